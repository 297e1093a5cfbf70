// Convolution kernels of the MTL-DAS hot path for sm_100a.
//
//  conv_umma_kernel : bf16 implicit GEMM on tcgen05 tensor cores.  One CTA owns MT*128 consecutive
//                     flattened output positions (see types.h for the P8 layout).  The input slab
//                     (all channels, tile + halo) is brought into shared memory ONCE by 1-D bulk
//                     async copies (TMA engine); every filter tap is then just a different start
//                     address of the A-operand descriptor into that slab, so a 3x3 conv reads its
//                     input once from L2/HBM, not nine times.  Weights stream through a ring of
//                     shared-memory stages (all resident when small).  Accumulators live in TMEM;
//                     the epilogue (bias/folded BN, residual add, ReLU, sigmoid * feature) is fused
//                     and writes 16-byte granules, 512 contiguous bytes per warp.
//  conv_fp32_kernel : the FP32 path of the same op (CUDA-core FFMA, fp32 activations): true fp32
//                     products as the reference's fp32 tolerance (1e-4) requires.
//
// Reference ops replaced: nn.Conv2d (+BatchNorm2d eval fold) + ReLU / residual add / Sigmoid * feature,
// /root/reference/model/modelA_MTL.py:12-32 (ResBlock), :42-50 (att_generator), :101-113 (output layers),
// :144,151,157,163 (mask multiply), :149,155,161 (torch.cat -> two sources).
#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cuda_bf16.h>

#include "gran.cuh"
#include "kernels.h"
#include "ptx.cuh"
#include "train_types.h"

namespace mtldas {

// Fused epilogue for 16 consecutive output channels [c0, c0+16) of flattened output position f.
template <typename T>
__device__ __forceinline__ void epilogue16(const ConvParams& p, const ConvGroupDev& g, int f, int c0, float* v) {
  if (f >= p.body) return;
  int n = f / p.Pimg;
  int rem = f - n * p.Pimg;
  int h = rem / p.Wp;
  int w = rem - h * p.Wp;
  bool valid = (h < p.H) && (w < p.W);
  if (valid) {
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] += __ldg(g.bias + c0 + j);
    if (p.epi == EPI_ADD_RELU || p.epi == EPI_SIGMOID_MUL) {
      float a[16];
      const int ac = (c0 >> 3) % g.aux_nchunks;
      load8<T>(view_addr<T>(g.aux, ac, f, n, h, w), a);
      load8<T>(view_addr<T>(g.aux, ac + 1, f, n, h, w), a + 8);
      if (p.epi == EPI_ADD_RELU) {
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = fmaxf(v[j] + a[j], 0.f);
      } else {
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] = a[j] / (1.f + __expf(-v[j]));
      }
    } else if (p.epi == EPI_RELU) {
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = fmaxf(v[j], 0.f);
    } else if (p.epi == EPI_ACCUM) {
      float a[16];
      load8<T>(view_addr<T>(g.out, c0 >> 3, f, n, h, w), a);
      load8<T>(view_addr<T>(g.out, (c0 >> 3) + 1, f, n, h, w), a + 8);
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] += a[j];
    }
  } else {
    if (g.out.layout != LAY_P8 || p.epi == EPI_ACCUM) return;  // pre-zeroed pads stay untouched
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] = 0.f;
  }
  store8<T>(view_addr<T>(g.out, c0 >> 3, f, n, h, w), v);
  store8<T>(view_addr<T>(g.out, (c0 >> 3) + 1, f, n, h, w), v + 8);
}

// ---------------------------------------------------------------------------------------------
// tcgen05 kernel
// ---------------------------------------------------------------------------------------------
// shared memory header: mbarriers [0,192), TMEM base slot at 192, bias (<= 128 floats... 256) at 256
constexpr int kSmemHeader = 1280;
constexpr int kBarSlab = 0, kBarAcc = 1, kBarWFull = 5, kBarWEmpty = 5 + kMaxRing;  // 8-byte slots

__device__ __forceinline__ long long gtime() {
  long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// slot 0: globaltimer (ns) at CTA start; slots 1..7: SM clock at each stage
#define DBG(slot)                                                                       \
  if (p.dbg && blockIdx.y == 0) {                                                       \
    if ((slot) == 0) p.dbg[(size_t)blockIdx.x * 16] = gtime();                           \
    p.dbg[(size_t)blockIdx.x * 16 + (slot) + 1] = clock64();                             \
  }

// relu + round-to-nearest bf16 of two floats in one instruction; lo lands in the low half (lower channel)
__device__ __forceinline__ uint32_t relu_pack_bf16x2(float lo, float hi) {
  uint32_t r;
  asm("cvt.rn.relu.bf16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi), "f"(lo));
  return r;
}

// Epilogue of 16 consecutive channels [c0, c0+16) of one output position held in registers.
// Everything it needs is passed in registers (the asm "memory" clobbers of the pipeline would otherwise make
// the compiler re-read the kernel parameter block for every use).
template <typename T, int EPI>
__device__ __forceinline__ void epilogue_cols(const float* bias_s, int c0, bool valid, float* v, const char* aux_addr0,
                                              long long aux_plane_bytes, int aux_nchunks, char* out_addr0,
                                              long long out_plane_bytes) {
  const int ch = c0 >> 3;
  if (valid) {
    float a[16];
    if (EPI == EPI_ADD_RELU || EPI == EPI_SIGMOID_MUL) {
      const int ac = ch >= aux_nchunks ? ch - aux_nchunks : ch;  // tasks stacked along Cout share one feature map
      load8<T>(aux_addr0 + ac * aux_plane_bytes, a);
      load8<T>(aux_addr0 + (ac + 1) * aux_plane_bytes, a + 8);
    }
    const float4* b4 = reinterpret_cast<const float4*>(bias_s + c0);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 bb = b4[q];
      v[4 * q + 0] += bb.x, v[4 * q + 1] += bb.y, v[4 * q + 2] += bb.z, v[4 * q + 3] += bb.w;
    }
    if ((EPI == EPI_ADD_RELU || EPI == EPI_RELU) && sizeof(T) == 2) {
      // ReLU folded into the bf16 conversion (cvt.rn.relu.bf16x2): 8 instructions instead of 16 max + 8 cvt
      if (EPI == EPI_ADD_RELU) {
#pragma unroll
        for (int j = 0; j < 16; ++j) v[j] += a[j];
      }
      uint4 q0, q1;
      q0.x = relu_pack_bf16x2(v[0], v[1]), q0.y = relu_pack_bf16x2(v[2], v[3]);
      q0.z = relu_pack_bf16x2(v[4], v[5]), q0.w = relu_pack_bf16x2(v[6], v[7]);
      q1.x = relu_pack_bf16x2(v[8], v[9]), q1.y = relu_pack_bf16x2(v[10], v[11]);
      q1.z = relu_pack_bf16x2(v[12], v[13]), q1.w = relu_pack_bf16x2(v[14], v[15]);
      *reinterpret_cast<uint4*>(out_addr0 + ch * out_plane_bytes) = q0;
      *reinterpret_cast<uint4*>(out_addr0 + (ch + 1) * out_plane_bytes) = q1;
      return;
    }
    if (EPI == EPI_ADD_RELU) {
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = fmaxf(v[j] + a[j], 0.f);
    } else if (EPI == EPI_SIGMOID_MUL) {
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = __fdividef(a[j], 1.f + __expf(-v[j]));
    } else if (EPI == EPI_RELU) {
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = fmaxf(v[j], 0.f);
    } else if (EPI == EPI_ACCUM) {  // gradient accumulation: out += conv
      load8<T>(out_addr0 + ch * out_plane_bytes, a);
      load8<T>(out_addr0 + (ch + 1) * out_plane_bytes, a + 8);
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] += a[j];
    }
  } else {
    if (EPI == EPI_ACCUM) return;  // pads of an accumulated tensor are already zero
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] = 0.f;
  }
  store8<T>(out_addr0 + ch * out_plane_bytes, v);
  store8<T>(out_addr0 + (ch + 1) * out_plane_bytes, v + 8);
}

// ---- epilogues with a second operand (residual / feature map): the aux granules are fetched kAuxDepth blocks of
// 16 channels ahead of the accumulator read, across accumulator-tile boundaries, and the first fetches are issued
// BEFORE the wait for the accumulator -- otherwise every block pays a full L2 round trip between tcgen05.ld and the
// store (measured: the residual layers ran 1.5-1.7x slower than their plain twins).
struct AuxPair {
  uint4 a0, a1;
};
struct EpiRow {            // one output position (TMEM lane) of one accumulator tile
  bool valid, do_store;
  char* out_addr0;
  const char* aux_addr0;
  uint32_t trow;
};
constexpr int kRingAuxDepth = 4, kPersistAuxDepth = 4;

template <int EPI>
__device__ __forceinline__ void aux_fetch(AuxPair& r, const EpiRow& row, long long aux_plane_bytes, int aux_nchunks, int c0) {
  if (row.valid) {
    const int ch = c0 >> 3;
    const int ac = ch >= aux_nchunks ? ch - aux_nchunks : ch;  // tasks stacked along Cout share one feature map
    r.a0 = *reinterpret_cast<const uint4*>(row.aux_addr0 + ac * aux_plane_bytes);
    r.a1 = *reinterpret_cast<const uint4*>(row.aux_addr0 + (ac + 1) * aux_plane_bytes);
  }
}

template <int EPI>
__device__ __forceinline__ void epilogue_cols_aux(const float* bias_s, int c0, const EpiRow& row, float* v, const AuxPair& ax,
                                                  long long out_plane_bytes) {
  const int ch = c0 >> 3;
  if (row.valid) {
    float a[16];
    const __nv_bfloat162* h0 = reinterpret_cast<const __nv_bfloat162*>(&ax.a0);
    const __nv_bfloat162* h1 = reinterpret_cast<const __nv_bfloat162*>(&ax.a1);
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float2 f0 = __bfloat1622float2(h0[i]), f1 = __bfloat1622float2(h1[i]);
      a[2 * i] = f0.x, a[2 * i + 1] = f0.y, a[8 + 2 * i] = f1.x, a[8 + 2 * i + 1] = f1.y;
    }
    const float4* b4 = reinterpret_cast<const float4*>(bias_s + c0);
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const float4 bb = b4[q];
      v[4 * q + 0] += bb.x, v[4 * q + 1] += bb.y, v[4 * q + 2] += bb.z, v[4 * q + 3] += bb.w;
    }
    if (EPI == EPI_ADD_RELU) {
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = fmaxf(v[j] + a[j], 0.f);
    } else {  // EPI_SIGMOID_MUL
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = __fdividef(a[j], 1.f + __expf(-v[j]));
    }
  } else {
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] = 0.f;
  }
  store8<__nv_bfloat16>(row.out_addr0 + ch * out_plane_bytes, v);
  store8<__nv_bfloat16>(row.out_addr0 + (ch + 1) * out_plane_bytes, v + 8);
}

// Drain `ntile` accumulator tiles (this warp's share) of Cout columns each.  row_of(i) gives the row of the i-th tile,
// ready(i) blocks until its accumulator is complete.
template <int EPI, int kAuxDepth, class RowFn, class ReadyFn>
__device__ __forceinline__ void drain_with_aux(int ntile, int Cout, const float* bias_s, long long aux_plane_bytes,
                                               int aux_nchunks, long long out_plane_bytes, RowFn row_of, ReadyFn ready) {
  const int bpt = Cout >> 4, nblk = ntile * bpt;  // blocks of 16 channels
  AuxPair buf[kAuxDepth];
  // fetch cursor (runs kAuxDepth blocks ahead of the consume cursor)
  int f_tile = 0, f_c0 = 0;
  EpiRow f_row = row_of(0);
#pragma unroll
  for (int i = 0; i < kAuxDepth; ++i) {
    if (f_tile < ntile) {
      aux_fetch<EPI>(buf[i], f_row, aux_plane_bytes, aux_nchunks, f_c0);
      f_c0 += 16;
      if (f_c0 >= Cout) {
        f_c0 = 0, ++f_tile;
        if (f_tile < ntile) f_row = row_of(f_tile);
      }
    }
  }
  int c_tile = 0, c_c0 = 0;
  EpiRow c_row = row_of(0);
  ready(0);
  for (int b0 = 0; b0 < nblk; b0 += kAuxDepth) {
#pragma unroll
    for (int i = 0; i < kAuxDepth; ++i) {
      if (b0 + i < nblk) {
        uint32_t r[16];
        ptx::tmem_ld16_issue(c_row.trow + (uint32_t)c_c0, r);
        ptx::tmem_wait16(r);
        if (c_row.do_store) {
          float v[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
          epilogue_cols_aux<EPI>(bias_s, c_c0, c_row, v, buf[i], out_plane_bytes);
        }
        c_c0 += 16;
        if (c_c0 >= Cout) {
          c_c0 = 0, ++c_tile;
          if (c_tile < ntile) {
            c_row = row_of(c_tile);
            ready(c_tile);
          }
        }
        if (f_tile < ntile) {  // refill the slot just consumed
          aux_fetch<EPI>(buf[i], f_row, aux_plane_bytes, aux_nchunks, f_c0);
          f_c0 += 16;
          if (f_c0 >= Cout) {
            f_c0 = 0, ++f_tile;
            if (f_tile < ntile) f_row = row_of(f_tile);
          }
        }
      }
    }
  }
}

template <int EPI>
__global__ void __launch_bounds__(256) conv_umma_kernel(const __grid_constant__ ConvParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  const ConvGroupDev& g = p.g[blockIdx.y];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 192);
  float* bias_s = reinterpret_cast<float*>(smem + 256);
  uint8_t* slab = smem + kSmemHeader;
  const uint32_t chunk_bytes = (uint32_t)p.SP * 16u;  // one channel chunk of the slab
  const uint32_t slab_bytes = (uint32_t)(p.nsrc * p.nchunks) * chunk_bytes;
  uint8_t* wring = slab + ((slab_bytes + 127u) & ~127u);
  const uint32_t stage_bytes = (uint32_t)(p.tps * p.tap_bytes);
  const bool resident = (p.nstages == 1);  // all taps in one stage: accumulators can finish one by one

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile_start = blockIdx.x * (p.MT * kTileM);

  ptx::pdl_launch_dependents();
  if (threadIdx.x == 0) {
    DBG(0)
    ptx::mbar_init(bars + kBarSlab, 1);
    for (int i = 0; i < 4; ++i) ptx::mbar_init(bars + kBarAcc + i, 1);
    for (int i = 0; i < p.NS; ++i) {
      ptx::mbar_init(bars + kBarWFull + i, 1);
      ptx::mbar_init(bars + kBarWEmpty + i, 1);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, (uint32_t)p.tmem_cols);
  for (int i = threadIdx.x; i < p.Cout; i += blockDim.x) bias_s[i] = g.bias[i];
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ---- producer warp: one elected thread issues the input slab, then the weight stages --------
    if (ptx::elect_one()) {
      // weights first (they do not depend on the previous kernel), then wait for it, then the input slab
      const int pre = min(p.nstages, p.NS);
      for (int st = 0; st < pre; ++st) {
        const int ntap = min(p.tps, p.ntaps - st * p.tps);
        const uint32_t bytes = (uint32_t)(ntap * p.tap_bytes);
        ptx::mbar_arrive_expect_tx(bars + kBarWFull + st, bytes);
        ptx::bulk_g2s(wring + (size_t)st * stage_bytes, g.w + (size_t)st * stage_bytes, bytes, bars + kBarWFull + st);
      }
      ptx::pdl_grid_wait();
      DBG(1)
      ptx::mbar_arrive_expect_tx(bars + kBarSlab, slab_bytes);
      for (int s = 0; s < p.nsrc; ++s) {
        const char* src = g.src[s].base + (long long)(tile_start - p.halo_lo) * 16;
        for (int c = 0; c < p.nchunks; ++c)
          ptx::bulk_g2s(slab + (size_t)(s * p.nchunks + c) * chunk_bytes, src + (long long)c * g.src[s].plane_bytes,
                        chunk_bytes, bars + kBarSlab);
      }
      for (int st = pre; st < p.nstages; ++st) {
        const int slot = st % p.NS;
        ptx::mbar_wait(bars + kBarWEmpty + slot, ((st / p.NS) - 1) & 1);
        const int ntap = min(p.tps, p.ntaps - st * p.tps);
        const uint32_t bytes = (uint32_t)(ntap * p.tap_bytes);
        ptx::mbar_arrive_expect_tx(bars + kBarWFull + slot, bytes);
        ptx::bulk_g2s(wring + (size_t)slot * stage_bytes, g.w + (size_t)st * stage_bytes, bytes,
                      bars + kBarWFull + slot);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ---- MMA warp: ONE elected thread runs the whole issue loop (ptxas keeps the descriptors in uniform
    // registers inside an elect.sync region).  Per MMA only the 14-bit start-address fields change, so the
    // loop adds small constants to the low descriptor words.
    if (ptx::elect_one()) {
      const uint32_t idesc = ptx::umma_idesc_bf16(kTileM, p.Cout, 0, 0);
      const uint32_t b_kstride = (uint32_t)p.Cout * 16u;  // bytes between 8-channel chunks of the weights
      const uint64_t a_hi = (ptx::umma_desc(0, p.swap_desc ? 128u : chunk_bytes, p.swap_desc ? chunk_bytes : 128u));
      const uint64_t b_hi = (ptx::umma_desc(0, p.swap_desc ? 128u : b_kstride, p.swap_desc ? b_kstride : 128u));
      const uint32_t slab16 = ptx::smem_u32(slab) >> 4;        // all addresses below in 16-byte units
      const uint32_t wring16 = ptx::smem_u32(wring) >> 4;
      const uint32_t chunk16 = chunk_bytes >> 4;
      const uint32_t a_kstep = 2u * chunk16, b_kstep = 2u * (uint32_t)p.Cout;
      const uint32_t tap16 = (uint32_t)p.tap_bytes >> 4, stage16 = stage_bytes >> 4;
      const int MT = p.MT, KC2 = p.nchunks >> 1, Cout = p.Cout, ntaps = p.ntaps;
      uint32_t tap_off[kMaxTaps];  // per-tap slab offsets in registers (no parameter-block loads in the issue loop)
#pragma unroll
      for (int t = 0; t < kMaxTaps; ++t)
        tap_off[t] = t < ntaps ? (uint32_t)(p.tap[t].src * p.nchunks) * chunk16 + (uint32_t)(p.halo_lo + p.tap[t].off) : 0u;
      ptx::mbar_wait(bars + kBarSlab, 0);
      ptx::tc_fence_after();
      DBG(2)
      if (resident) {
        ptx::mbar_wait(bars + kBarWFull, 0);
        ptx::tc_fence_after();
        DBG(3)
        for (int mt = 0; mt < MT; ++mt) {
          const uint32_t d_tmem = tmem_base + (uint32_t)(mt * Cout);
          uint32_t acc = 0u;
          uint32_t b_lo = wring16 | (uint32_t)b_hi;
#pragma unroll
          for (int t = 0; t < kMaxTaps; ++t) {
            if (t < ntaps) {
              uint32_t a_lo = (slab16 + tap_off[t] + (uint32_t)(mt * kTileM)) | (uint32_t)a_hi;
              for (int k = 0; k < KC2; ++k) {
                ptx::umma_bf16_lean(d_tmem, a_lo, (uint32_t)(a_hi >> 32), b_lo, (uint32_t)(b_hi >> 32), idesc, acc);
                acc = 1u;
                a_lo += a_kstep;
                b_lo += b_kstep;
              }
            }
          }
          ptx::umma_commit(bars + kBarAcc + mt);
        }
      } else {
        // streamed weights.  With p.ksplit the even / odd 16-channel K steps accumulate into two TMEM sets that the
        // epilogue adds: consecutive MMAs then never target the same accumulator tile twice within 2*MT issues
        // (back-to-back accumulation into one tile serialises on the accumulate latency).
        uint32_t acc0 = 0u, acc1 = 0u;
        const uint32_t set1 = p.ksplit ? (uint32_t)(MT * Cout) : 0u;
        const int nchunks_r = p.nchunks, halo_lo_r = p.halo_lo;
        const bool ksplit_r = p.ksplit != 0;
        for (int st = 0; st < p.nstages; ++st) {
          const int slot = st % p.NS;
          ptx::mbar_wait(bars + kBarWFull + slot, (st / p.NS) & 1);
          ptx::tc_fence_after();
          if (st == 0) DBG(3)
          const int ntap = min(p.tps, ntaps - st * p.tps);
          for (int ti = 0; ti < ntap; ++ti) {
            const ConvTapDev tp = p.tap[st * p.tps + ti];
            uint32_t a_lo = (slab16 + (uint32_t)(tp.src * nchunks_r) * chunk16 + (uint32_t)(halo_lo_r + tp.off)) | (uint32_t)a_hi;
            uint32_t b_lo = (wring16 + (uint32_t)slot * stage16 + (uint32_t)ti * tap16) | (uint32_t)b_hi;
            const uint32_t a_hh = (uint32_t)(a_hi >> 32), b_hh = (uint32_t)(b_hi >> 32);
            for (int k = 0; k < KC2; ++k) {
              const bool odd = ksplit_r && (k & 1);
              const uint32_t d0 = tmem_base + (odd ? set1 : 0u);
              const uint32_t acc = odd ? acc1 : acc0;
              if (MT == 4) {
#pragma unroll
                for (int mt = 0; mt < 4; ++mt)
                  ptx::umma_bf16_lean(d0 + (uint32_t)(mt * Cout), a_lo + (uint32_t)(mt * kTileM), a_hh, b_lo, b_hh, idesc, acc);
              } else {
                for (int mt = 0; mt < MT; ++mt)
                  ptx::umma_bf16_lean(d0 + (uint32_t)(mt * Cout), a_lo + (uint32_t)(mt * kTileM), a_hh, b_lo, b_hh, idesc, acc);
              }
              if (odd) acc1 = 1u; else acc0 = 1u;
              a_lo += a_kstep;
              b_lo += b_kstep;
            }
          }
          ptx::umma_commit(bars + kBarWEmpty + slot);  // slot reusable once these MMAs are done
        }
        for (int mt = 0; mt < MT; ++mt) ptx::umma_commit(bars + kBarAcc + mt);
      }
      DBG(4)
    }
    __syncwarp();
  }

  // ---- epilogue: warp w owns TMEM lanes [32w, 32w+32); accumulator tiles are drained as they finish --
  {
    const ActView outv = g.out, auxv = g.aux;  // register copies (see epilogue_cols)
    const long long out_plane_bytes = outv.plane * 16, aux_plane_bytes = auxv.plane * 16;
    const int aux_nchunks = g.aux_nchunks;
    const int Cout = p.Cout, MT = p.MT, H = p.H, W = p.W, Wp = p.Wp, Pimg = p.Pimg, body = (int)p.body;
    // launched with 256 threads (layers that own the SM alone): warps w and w+4 share TMEM lane quarter w % 4 and
    // take alternate accumulator tiles, halving the serial epilogue of the CTA
    const int quarter = warp & 3, mt0 = warp >> 2, mt_step = (int)(blockDim.x >> 7);
    if ((EPI == EPI_ADD_RELU || EPI == EPI_SIGMOID_MUL) && !p.ksplit) {
      // the aux fetches below run ahead of every barrier of this kernel: order them after the previous grid here
      // (only the producer thread waited so far)
      ptx::pdl_grid_wait();
      auto row_of = [&](int i) {
        const int mt = mt0 + i * mt_step;
        const int f = tile_start + mt * kTileM + quarter * 32 + lane;
        const bool in_body = f < body;
        int n = 0, h = 0, w = 0;
        if (in_body) {
          n = f / Pimg;
          const int rem = f - n * Pimg;
          h = rem / Wp;
          w = rem - h * Wp;
        }
        EpiRow r;
        r.valid = in_body && (h < H) && (w < W);
        r.do_store = in_body && (r.valid || outv.layout == LAY_P8);
        r.out_addr0 = view_addr<__nv_bfloat16>(outv, 0, f, n, h, w);
        r.aux_addr0 = view_addr<__nv_bfloat16>(auxv, 0, f, n, h, w);
        r.trow = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(mt * Cout);
        return r;
      };
      auto ready = [&](int i) {
        ptx::mbar_wait(bars + kBarAcc + mt0 + i * mt_step, 0);
        ptx::tc_fence_after();
      };
      const int nmine = mt0 < MT ? (MT - mt0 + mt_step - 1) / mt_step : 0;
      if (nmine > 0) drain_with_aux<EPI, kRingAuxDepth>(nmine, Cout, bias_s, aux_plane_bytes, aux_nchunks, out_plane_bytes, row_of, ready);
    } else
    for (int mt = mt0; mt < MT; mt += mt_step) {
      const int f = tile_start + mt * kTileM + quarter * 32 + lane;
      const bool in_body = f < body;
      int n = 0, h = 0, w = 0;
      if (in_body) {
        n = f / Pimg;
        const int rem = f - n * Pimg;
        h = rem / Wp;
        w = rem - h * Wp;
      }
      const bool valid = in_body && (h < H) && (w < W);
      const bool do_store = in_body && (valid || outv.layout == LAY_P8);
      char* out_addr0 = view_addr<__nv_bfloat16>(outv, 0, f, n, h, w);
      const char* aux_addr0 = view_addr<__nv_bfloat16>(auxv, 0, f, n, h, w);
      ptx::mbar_wait(bars + kBarAcc + mt, 0);
      ptx::tc_fence_after();
      if (mt == 0 && threadIdx.x == 64) DBG(5)
      const uint32_t trow = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(mt * Cout);
      if (Cout >= 32) {
        for (int c0 = 0; c0 < Cout; c0 += 32) {
          uint32_t r[32];
          ptx::tmem_ld32_issue(trow + (uint32_t)c0, r);
          ptx::tmem_wait32(r);
          if (p.ksplit && !resident) {  // second K-split accumulator set
            uint32_t r2[32];
            ptx::tmem_ld32_issue(trow + (uint32_t)(MT * Cout + c0), r2);
            ptx::tmem_wait32(r2);
#pragma unroll
            for (int j = 0; j < 32; ++j) r[j] = __float_as_uint(__uint_as_float(r[j]) + __uint_as_float(r2[j]));
          }
          if (mt == 0 && c0 == 0 && threadIdx.x == 64) DBG(7)
          if (do_store) {
#pragma unroll
            for (int hlf = 0; hlf < 2; ++hlf) {
              float v[16];
#pragma unroll
              for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[hlf * 16 + j]);
              epilogue_cols<__nv_bfloat16, EPI>(bias_s, c0 + hlf * 16, valid, v, aux_addr0, aux_plane_bytes,
                                                aux_nchunks, out_addr0, out_plane_bytes);
            }
          }
          if (mt == 0 && c0 == 0 && threadIdx.x == 64) DBG(8)
        }
      } else {
        uint32_t r[16];
        ptx::tmem_ld16_issue(trow, r);
        ptx::tmem_wait16(r);
        if (do_store) {
          float v[16];
#pragma unroll
          for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
          epilogue_cols<__nv_bfloat16, EPI>(bias_s, 0, valid, v, aux_addr0, aux_plane_bytes, aux_nchunks, out_addr0,
                                            out_plane_bytes);
        }
      }
    }
  }
  if (threadIdx.x == 64) DBG(9)
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 64) DBG(6)
  if (warp == 1) ptx::tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols);
}

// ---------------------------------------------------------------------------------------------
// persistent, pipelined variant for layers whose packed weights stay resident in shared memory.
// One CTA loops over tiles (tile = blockIdx.x + k*gridDim.x).  Warp 0: producer (weights once, then a ring of
// input slabs); warp 1: MMA issuer into one of two TMEM accumulator sets; warps 2..9: epilogue (two warps per
// TMEM lane quarter, accumulator tiles split between them).  The tensor pipe, the bulk-copy engine and the
// epilogue warps of different tiles overlap inside the CTA instead of relying on CTA-level concurrency.
// ---------------------------------------------------------------------------------------------
constexpr int kPersistThreads = 320;
constexpr int kPBarW = 0, kPBarSlabFull = 1, kPBarSlabEmpty = 4, kPBarAccFull = 7, kPBarAccEmpty = 9;  // 8-byte slots

__device__ __forceinline__ int fast_div(int f, unsigned long long magic) {
  return (int)(((unsigned long long)(unsigned)f * magic) >> 40);
}

// All MMAs of one tile set (MT accumulator tiles x taps x K steps), issued by ONE thread.  The issue loop used to cost
// ~100 cycles per MMA (kernel timeline, round 2): about forty uniform-datapath instructions per K step, parameter
// re-loads after every MMA (the asm's memory clobber) and R2UR moves of loop bounds -- twice the tensor pipe's own
// 45-64 cycles.  Here MT is a compile-time constant (the mt loop is fully unrolled), the descriptors are 32-bit
// running sums, and nothing is read from memory inside the loop.
template <int MT_, int KS_>
__device__ __forceinline__ void issue_tile_mmas(uint32_t d_base, uint32_t part_cols, uint32_t slab16,
                                                const uint32_t (&tap_off)[kMaxTaps], int ntaps, int KC2, uint32_t Cout,
                                                uint32_t a_hi_lo, uint32_t a_hi_hi, uint32_t b_hi_lo, uint32_t b_hi_hi,
                                                uint32_t w16, uint32_t a_kstep, uint32_t b_kstep, uint32_t idesc) {
  // KS_ > 1: the K steps rotate over KS_ accumulator partials (summed in the epilogue): back-to-back MMAs into ONE
  // accumulator serialise on the accumulate latency (~150 cycles measured for M128 N64 against ~45 of issue), so a
  // CTA wants MT_ * KS_ >= 4 independent chains
  uint32_t started = 0u, b_lo = w16 | b_hi_lo, step = 0u;
#pragma unroll
  for (int t = 0; t < kMaxTaps; ++t) {
    if (t < ntaps) {
      uint32_t a_lo = (slab16 + tap_off[t]) | a_hi_lo;   // start address < 2^14: the OR never carries into the LBO field
      for (int k = 0; k < KC2; ++k) {
        const uint32_t h = KS_ > 1 ? (step & (uint32_t)(KS_ - 1)) : 0u;
        const uint32_t acc = (started >> h) & 1u;
        const uint32_t d = d_base + h * part_cols;
#pragma unroll
        for (int mt = 0; mt < MT_; ++mt)
          ptx::umma_bf16_lean(d + (uint32_t)mt * Cout, a_lo + (uint32_t)(mt * kTileM), a_hi_hi, b_lo, b_hi_hi, idesc, acc);
        started |= 1u << h;
        ++step;
        a_lo += a_kstep;
        b_lo += b_kstep;
      }
    }
  }
}

// 32 / 16 accumulator columns of this thread's row, summed over the KS K-split partials
__device__ __forceinline__ void load_acc32(uint32_t taddr, int KS, uint32_t part_cols, uint32_t* r) {
  ptx::tmem_ld32_issue(taddr, r);
  ptx::tmem_wait32(r);
  for (int h = 1; h < KS; ++h) {
    uint32_t r2[32];
    ptx::tmem_ld32_issue(taddr + (uint32_t)h * part_cols, r2);
    ptx::tmem_wait32(r2);
#pragma unroll
    for (int j = 0; j < 32; ++j) r[j] = __float_as_uint(__uint_as_float(r[j]) + __uint_as_float(r2[j]));
  }
}
__device__ __forceinline__ void load_acc16(uint32_t taddr, int KS, uint32_t part_cols, uint32_t* r) {
  ptx::tmem_ld16_issue(taddr, r);
  ptx::tmem_wait16(r);
  for (int h = 1; h < KS; ++h) {
    uint32_t r2[16];
    ptx::tmem_ld16_issue(taddr + (uint32_t)h * part_cols, r2);
    ptx::tmem_wait16(r2);
#pragma unroll
    for (int j = 0; j < 16; ++j) r[j] = __float_as_uint(__uint_as_float(r[j]) + __uint_as_float(r2[j]));
  }
}

// WIDE = 1: layers that own the SM alone anyway (>= 64 output channels, large weight image): no 96-register cap, and the
// second-operand epilogues run the fetch-ahead drain; WIDE = 0: two CTAs per SM (16 / 32 output channels).
template <int EPI, int WIDE>
__global__ void __launch_bounds__(kPersistThreads, WIDE ? 1 : 2) conv_umma_persist_kernel(const __grid_constant__ ConvParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  const ConvGroupDev& g = p.g[blockIdx.y];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 192);
  float* bias_s = reinterpret_cast<float*>(smem + 256);
  uint8_t* wsm = smem + kSmemHeader;
  const uint32_t w_bytes = (uint32_t)(p.ntaps * p.tap_bytes);
  const uint32_t chunk_bytes = (uint32_t)p.SP * 16u;
  const uint32_t slab_bytes = (uint32_t)(p.nsrc * p.nchunks) * chunk_bytes;
  const uint32_t slab_stride = (slab_bytes + 127u) & ~127u;
  uint8_t* slabs = wsm + ((w_bytes + 127u) & ~127u);
  const int S = p.persist_stages;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile_rows = p.MT * kTileM;                                   // rows computed per tile
  const int tile_pos = p.pool_k > 0 ? 2 * p.Wp * p.pool_k : tile_rows;   // rows a tile advances by
  const int ntiles = (int)((p.body + tile_pos - 1) / tile_pos);
  uint8_t* staging = slabs + (size_t)p.persist_stages * slab_stride;     // fused max-pool: [chunk][tile_rows][16 B]
  const int KS = p.ksplit > 1 ? p.ksplit : 1;                      // K-split partials per accumulator tile (launcher)
  const uint32_t part_cols = (uint32_t)(p.MT * p.Cout);            // one partial of all MT tiles
  const uint32_t acc_cols = part_cols * (uint32_t)KS;              // one accumulator set
  const int nsets = p.acc_sets;                       // 1 or 2 TMEM accumulator sets
  const int nepi = (int)(blockDim.x >> 5) - 2;        // 4 or 8 epilogue warps
  uint32_t cols = 32;
  while (cols < (uint32_t)nsets * acc_cols) cols <<= 1;

  ptx::pdl_launch_dependents();
  if (threadIdx.x == 0) {
    ptx::mbar_init(bars + kPBarW, 1);
    for (int i = 0; i < 3; ++i) {
      ptx::mbar_init(bars + kPBarSlabFull + i, 1);
      ptx::mbar_init(bars + kPBarSlabEmpty + i, 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(bars + kPBarAccFull + i, 1);
      ptx::mbar_init(bars + kPBarAccEmpty + i, (uint32_t)nepi);  // one arrival per epilogue warp
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, cols);
  for (int i = threadIdx.x; i < p.Cout; i += blockDim.x) bias_s[i] = g.bias[i];
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (ptx::elect_one()) {
      ptx::mbar_arrive_expect_tx(bars + kPBarW, w_bytes);
      ptx::bulk_g2s(wsm, g.w, w_bytes, bars + kPBarW);
      ptx::pdl_grid_wait();
      int it = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int s = it % S;
        if (it >= S) ptx::mbar_wait(bars + kPBarSlabEmpty + s, ((it / S) - 1) & 1);
        ptx::mbar_arrive_expect_tx(bars + kPBarSlabFull + s, slab_bytes);
        uint8_t* slab = slabs + (size_t)s * slab_stride;
        const long long start = (long long)tile * tile_pos - p.halo_lo;
        for (int sr = 0; sr < p.nsrc; ++sr) {
          const char* src = g.src[sr].base + start * 16;
          for (int c = 0; c < p.nchunks; ++c)
            ptx::bulk_g2s(slab + (size_t)(sr * p.nchunks + c) * chunk_bytes, src + (long long)c * g.src[sr].plane_bytes,
                          chunk_bytes, bars + kPBarSlabFull + s);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (ptx::elect_one()) {
      const uint32_t idesc = ptx::umma_idesc_bf16(kTileM, p.Cout, 0, 0);
      const uint32_t b_kstride = (uint32_t)p.Cout * 16u;
      const uint64_t a_hi = ptx::umma_desc(0, chunk_bytes, 128u);
      const uint64_t b_hi = ptx::umma_desc(0, b_kstride, 128u);
      const uint32_t w16 = ptx::smem_u32(wsm) >> 4, slabs16 = ptx::smem_u32(slabs) >> 4;
      const uint32_t chunk16 = chunk_bytes >> 4, a_kstep = 2u * chunk16, b_kstep = 2u * (uint32_t)p.Cout;
      const int MT = p.MT, KC2 = p.nchunks >> 1, Cout = p.Cout, ntaps = p.ntaps;
      // per-tap slab offsets in registers (the tap table lives in the kernel parameter block; reading it inside
      // the issue loop would put a constant-bank load in front of every MMA)
      uint32_t tap_off[kMaxTaps];
#pragma unroll
      for (int t = 0; t < kMaxTaps; ++t)
        tap_off[t] = t < ntaps ? (uint32_t)(p.tap[t].src * p.nchunks) * chunk16 + (uint32_t)(p.halo_lo + p.tap[t].off) : 0u;
      ptx::mbar_wait(bars + kPBarW, 0);
      int it = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int s = it % S, a = nsets == 2 ? (it & 1) : 0;
        const bool dbg_on = p.dbg && blockIdx.x == 0 && blockIdx.y == 0 && it < 12;
        if (dbg_on) p.dbg[it * 8 + 0] = clock64();
        ptx::mbar_wait(bars + kPBarSlabFull + s, (it / S) & 1);
        if (dbg_on) p.dbg[it * 8 + 1] = clock64();
        if (it >= nsets) ptx::mbar_wait(bars + kPBarAccEmpty + a, ((it / nsets) - 1) & 1);
        if (dbg_on) p.dbg[it * 8 + 2] = clock64();
        ptx::tc_fence_after();
        const uint32_t slab16 = slabs16 + (uint32_t)s * (slab_stride >> 4);
        // consecutive MMAs go to DIFFERENT accumulator tiles (mt innermost): back-to-back accumulation into the same
        // TMEM tile serialises on the accumulate latency, interleaving the MT tiles keeps the tensor pipe busy
        const uint32_t d_base = tmem_base + (uint32_t)a * acc_cols;
        const uint32_t ah_lo = (uint32_t)a_hi, ah_hi = (uint32_t)(a_hi >> 32), bh_lo = (uint32_t)b_hi, bh_hi = (uint32_t)(b_hi >> 32);
#define MTLDAS_ISSUE(M_, K_)                                                                                           \
  issue_tile_mmas<M_, K_>(d_base, part_cols, slab16, tap_off, ntaps, KC2, (uint32_t)Cout, ah_lo, ah_hi, bh_lo, bh_hi, w16, \
                          a_kstep, b_kstep, idesc)
        if ((EPI == EPI_RELU && p.pool_k > 0) || (!WIDE && (EPI == EPI_ADD_RELU || EPI == EPI_SIGMOID_MUL))) {
          // (same for the narrow second-operand layers, whose epilogue is the critical path: 0.238 -> 0.248 ms otherwise)
          // fused-pool layers are bound by their epilogue's shared-memory traffic (staging tile); MMAs issued back to
          // back take the whole shared-memory bandwidth in bursts and delay exactly that critical path (measured:
          // 0.56 -> 0.62 ms on output_layer1).  The plain loop below paces itself at ~90 cycles per MMA.
          uint32_t acc = 0u, b_lo = w16;
#pragma unroll
          for (int t = 0; t < kMaxTaps; ++t) {
            if (t < ntaps) {
              uint32_t a_lo = slab16 + tap_off[t];
              for (int k = 0; k < KC2; ++k) {
                for (int mt = 0; mt < MT; ++mt)
                  ptx::umma_bf16(d_base + (uint32_t)(mt * Cout), a_hi | (a_lo + (uint32_t)(mt * kTileM)), b_hi | b_lo, idesc, acc);
                acc = 1u;
                a_lo += a_kstep;
                b_lo += b_kstep;
              }
            }
          }
        } else if (MT == 4) MTLDAS_ISSUE(4, 1);
        else if (MT == 2 && KS == 1) MTLDAS_ISSUE(2, 1);
        else if (MT == 2) MTLDAS_ISSUE(2, 2);
        else if (KS == 1) MTLDAS_ISSUE(1, 1);
        else if (KS == 2) MTLDAS_ISSUE(1, 2);
        else MTLDAS_ISSUE(1, 4);
#undef MTLDAS_ISSUE
        ptx::umma_commit(bars + kPBarSlabEmpty + s);
        ptx::umma_commit(bars + kPBarAccFull + a);
        if (dbg_on) p.dbg[it * 8 + 3] = clock64();
      }
    }
    __syncwarp();
  } else {
    // ---- epilogue warps: TMEM lane quarter = warp % 4; the two warps of a quarter take alternate accumulator tiles
    // (second-operand epilogues fetch ahead of every barrier of this kernel: order them after the previous grid here)
    if (WIDE && (EPI == EPI_ADD_RELU || EPI == EPI_SIGMOID_MUL)) ptx::pdl_grid_wait();
    const int quarter = warp & 3, half = (warp - 2) >> 2, mt_step = nepi >> 2;
    const ActView outv = g.out, auxv = g.aux;
    const long long out_plane_bytes = outv.plane * 16, aux_plane_bytes = auxv.plane * 16;
    const int aux_nchunks = g.aux_nchunks;
    const int Cout = p.Cout, MT = p.MT, H = p.H, W = p.W, Wp = p.Wp, Pimg = p.Pimg, body = (int)p.body;
    const unsigned long long m_pimg = p.magic_pimg, m_wp = p.magic_wp;
    // fused pool: exact multiply-shift divisors of the pooled index math (per-thread, once)
    unsigned long long m_per_chunk = 0, m_wph = 0, m_hph = 0;
    if (EPI == EPI_RELU && p.pool_k > 0) {
      m_per_chunk = (1ull << 40) / (unsigned long long)(p.pool_k * (Wp >> 1)) + 1;
      m_wph = (1ull << 40) / (unsigned long long)(Wp >> 1) + 1;
      m_hph = (1ull << 40) / (unsigned long long)((H + 1) >> 1) + 1;
    }
    int it = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const int a = nsets == 2 ? (it & 1) : 0;
      const bool dbg_on = p.dbg && blockIdx.x == 0 && blockIdx.y == 0 && it < 12 && warp == 2 && lane == 0;
      if (dbg_on) p.dbg[it * 8 + 4] = clock64();
      if (WIDE && (EPI == EPI_ADD_RELU || EPI == EPI_SIGMOID_MUL)) {   // (the launcher never K-splits these)
        // second-operand epilogues: the residual / feature granules of this warp's rows are fetched a few 16-channel
        // blocks ahead of the accumulator reads, the first ones BEFORE the wait for the accumulator -- one exposed L2
        // round trip per tile instead of one per block.  Measured: pays from 64 output channels on (0.13 -> 0.12 ms at
        // stage 3); with 16 / 32 channels a row is one or two blocks and the plain path is faster (0.25 vs 0.28 ms).
        auto row_of = [&](int i) {
          const int mt = half + i * mt_step;
          const int f = tile * tile_pos + mt * kTileM + quarter * 32 + lane;
          const bool in_body = f < body;
          int n = 0, h = 0, w = 0;
          if (in_body) {
            n = fast_div(f, m_pimg);
            const int rem = f - n * Pimg;
            h = fast_div(rem, m_wp);
            w = rem - h * Wp;
          }
          EpiRow r;
          r.valid = in_body && (h < H) && (w < W);
          r.do_store = in_body && (r.valid || outv.layout == LAY_P8);
          r.out_addr0 = view_addr<__nv_bfloat16>(outv, 0, f, n, h, w);
          r.aux_addr0 = view_addr<__nv_bfloat16>(auxv, 0, f, n, h, w);
          r.trow = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)a * acc_cols + (uint32_t)(mt * Cout);
          return r;
        };
        auto ready = [&](int i) {
          if (i == 0) {
            ptx::mbar_wait(bars + kPBarAccFull + a, (it / nsets) & 1);
            ptx::tc_fence_after();
          }
        };
        const int nmine = half < MT ? (MT - half + mt_step - 1) / mt_step : 0;
        if (nmine > 0)
          drain_with_aux<EPI, kPersistAuxDepth>(nmine, Cout, bias_s, aux_plane_bytes, aux_nchunks, out_plane_bytes, row_of, ready);
        else
          ready(0);
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(bars + kPBarAccEmpty + a);
        continue;
      }
      ptx::mbar_wait(bars + kPBarAccFull + a, (it / nsets) & 1);
      if (dbg_on) p.dbg[it * 8 + 5] = clock64();
      ptx::tc_fence_after();
      for (int mt = half; mt < MT; mt += mt_step) {
        const int f = tile * tile_pos + mt * kTileM + quarter * 32 + lane;
        const bool in_body = f < body;
        int n = 0, h = 0, w = 0;
        if (in_body) {
          n = fast_div(f, m_pimg);
          const int rem = f - n * Pimg;
          h = fast_div(rem, m_wp);
          w = rem - h * Wp;
        }
        const bool valid = in_body && (h < H) && (w < W);
        const bool do_store = in_body && (valid || outv.layout == LAY_P8);
        char* out_addr0 = view_addr<__nv_bfloat16>(outv, 0, f, n, h, w);
        const char* aux_addr0 = view_addr<__nv_bfloat16>(auxv, 0, f, n, h, w);
        const uint32_t trow = tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)a * acc_cols + (uint32_t)(mt * Cout);
        if (EPI == EPI_RELU && p.pool_k > 0) {
          // fused max-pool: relu(acc + bias) of this row goes to the shared-memory staging tile (zero at pads)
          const int row = mt * kTileM + quarter * 32 + lane;
          for (int c0 = 0; c0 < Cout; c0 += 32) {  // Cout is a multiple of 32 on the pooled layers (32 / 64 / 128)
            uint32_t r[32];
            load_acc32(trow + (uint32_t)c0, KS, part_cols, r);
#pragma unroll
            for (int hlf = 0; hlf < 2; ++hlf) {
              float v[16];
#pragma unroll
              for (int j = 0; j < 16; ++j)
                v[j] = valid ? fmaxf(__uint_as_float(r[hlf * 16 + j]) + bias_s[c0 + hlf * 16 + j], 0.f) : 0.f;
              const int ch = (c0 >> 3) + 2 * hlf;
              store8<__nv_bfloat16>(reinterpret_cast<char*>(staging) + ((size_t)ch * tile_rows + row) * 16, v);
              store8<__nv_bfloat16>(reinterpret_cast<char*>(staging) + ((size_t)(ch + 1) * tile_rows + row) * 16, v + 8);
            }
          }
        } else if (Cout >= 32) {
          for (int c0 = 0; c0 < Cout; c0 += 32) {
            uint32_t r[32];
            load_acc32(trow + (uint32_t)c0, KS, part_cols, r);
            if (do_store) {
#pragma unroll
              for (int hlf = 0; hlf < 2; ++hlf) {
                float v[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[hlf * 16 + j]);
                epilogue_cols<__nv_bfloat16, EPI>(bias_s, c0 + hlf * 16, valid, v, aux_addr0, aux_plane_bytes,
                                                  aux_nchunks, out_addr0, out_plane_bytes);
              }
            }
          }
        } else {
          uint32_t r[16];
          load_acc16(trow, KS, part_cols, r);
          if (do_store) {
            float v[16];
#pragma unroll
            for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
            epilogue_cols<__nv_bfloat16, EPI>(bias_s, 0, valid, v, aux_addr0, aux_plane_bytes, aux_nchunks, out_addr0,
                                              out_plane_bytes);
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (dbg_on) p.dbg[it * 8 + 6] = clock64();
      if (lane == 0) ptx::mbar_arrive(bars + kPBarAccEmpty + a);  // this warp is done reading accumulator set a
      if (EPI == EPI_RELU && p.pool_k > 0) {
        // ---- fused MaxPool2d(2,2,ceil): the tile is pool_k image-row pairs; every pooled granule is the max of four
        // staged granules (pads were staged as zero, and post-ReLU values are >= 0, so clipped windows are correct)
        const int nthr = nepi * 32, e = (warp - 2) * 32 + lane;
        asm volatile("bar.sync 1, %0;" ::"r"(nthr) : "memory");
        const int Wph = Wp >> 1, nch = Cout >> 3, per_chunk = p.pool_k * Wph;
        const int hp_half = (H + 1) >> 1;                      // row pairs per image (Hp is even)
        const ActView pv = g.pool_out;
        for (int o = e; o < per_chunk * nch; o += nthr) {
          const int c = fast_div(o, m_per_chunk), q = o - c * per_chunk;
          const int rp = fast_div(q, m_wph), cp = q - rp * Wph;
          const int r0 = 2 * rp * Wp + 2 * cp;                 // row in the staging tile
          if ((long long)tile * tile_pos + r0 >= p.body) continue;
          const uint4* sp = reinterpret_cast<const uint4*>(staging + ((size_t)c * tile_rows + r0) * 16);
          uint4 m = sp[0];
          const uint4 b1 = sp[1], b2 = sp[Wp], b3 = sp[Wp + 1];
          __nv_bfloat162* mm = reinterpret_cast<__nv_bfloat162*>(&m);
          const __nv_bfloat162* x1 = reinterpret_cast<const __nv_bfloat162*>(&b1);
          const __nv_bfloat162* x2 = reinterpret_cast<const __nv_bfloat162*>(&b2);
          const __nv_bfloat162* x3 = reinterpret_cast<const __nv_bfloat162*>(&b3);
#pragma unroll
          for (int j = 0; j < 4; ++j) mm[j] = __hmax2(__hmax2(mm[j], x1[j]), __hmax2(x2[j], x3[j]));
          const int R = tile * p.pool_k + rp;                  // global row-pair index
          const int n = fast_div(R, m_hph), ho = R - n * hp_half;
          const long long pos = (long long)c * pv.plane + pv.g_lo + (long long)n * p.pool_Pimgo + ho * p.pool_Wpo + cp;
          *reinterpret_cast<uint4*>(pv.base + pos * 16) = m;
        }
        asm volatile("bar.sync 1, %0;" ::"r"(nthr) : "memory");  // staging is reused by the next tile
        if (dbg_on) p.dbg[it * 8 + 7] = clock64();
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, cols);
}

size_t conv_persist_smem_bytes(const ConvParams& p) {
  size_t w = ((size_t)p.ntaps * p.tap_bytes + 127) & ~(size_t)127;
  size_t slab = ((size_t)p.nsrc * p.nchunks * p.SP * 16 + 127) & ~(size_t)127;
  size_t staging = p.pool_k > 0 ? (size_t)(p.Cout / 8) * p.MT * kTileM * 16 : 0;
  return kSmemHeader + w + (size_t)p.persist_stages * slab + staging;
}

// ---------------------------------------------------------------------------------------------
// FP32 path
// ---------------------------------------------------------------------------------------------
constexpr int kFp32CoutBlock = 32;

__global__ void __launch_bounds__(128) conv_fp32_kernel(const __grid_constant__ ConvParams p) {
  extern __shared__ __align__(16) float ws[];  // [K_tap][32]
  ptx::pdl_launch_dependents();
  ptx::pdl_grid_wait();
  const ConvGroupDev& g = p.g[blockIdx.y];
  const int f = blockIdx.x * 128 + threadIdx.x;
  const int cb = blockIdx.z * kFp32CoutBlock;
  const int ktap = p.nchunks * 8;
  const float* wg = reinterpret_cast<const float*>(g.w);
  float acc[kFp32CoutBlock];
#pragma unroll
  for (int j = 0; j < kFp32CoutBlock; ++j) acc[j] = 0.f;
  for (int t = 0; t < p.ntaps; ++t) {
    __syncthreads();
    for (int i = threadIdx.x; i < ktap * kFp32CoutBlock; i += 128) {
      int k = i / kFp32CoutBlock, co = i - k * kFp32CoutBlock;
      ws[i] = (cb + co < p.Cout) ? wg[((size_t)t * ktap + k) * p.Cout + cb + co] : 0.f;
    }
    __syncthreads();
    const ConvTapDev tp = p.tap[t];
    const char* src = g.src[tp.src].base + (long long)(f + tp.off) * 32;
    for (int c = 0; c < p.nchunks; ++c) {
      float x[8];
      load8<float>(src + (long long)c * g.src[tp.src].plane_bytes, x);
#pragma unroll
      for (int ci = 0; ci < 8; ++ci) {
        const float4* wr = reinterpret_cast<const float4*>(ws + (c * 8 + ci) * kFp32CoutBlock);
#pragma unroll
        for (int q = 0; q < kFp32CoutBlock / 4; ++q) {
          float4 w4 = wr[q];
          acc[4 * q + 0] = fmaf(x[ci], w4.x, acc[4 * q + 0]);
          acc[4 * q + 1] = fmaf(x[ci], w4.y, acc[4 * q + 1]);
          acc[4 * q + 2] = fmaf(x[ci], w4.z, acc[4 * q + 2]);
          acc[4 * q + 3] = fmaf(x[ci], w4.w, acc[4 * q + 3]);
        }
      }
    }
  }
  for (int c0 = 0; c0 < kFp32CoutBlock; c0 += 16)
    if (cb + c0 < p.Cout) epilogue16<float>(p, g, f, cb + c0, acc + c0);
}

// ---------------------------------------------------------------------------------------------
// launchers
// ---------------------------------------------------------------------------------------------
size_t conv_umma_smem_bytes(const ConvParams& p) {
  size_t slab = (size_t)p.nsrc * p.nchunks * p.SP * 16;
  slab = (slab + 127) & ~(size_t)127;
  return kSmemHeader + slab + (size_t)p.NS * p.tps * p.tap_bytes;
}

template <typename K, typename P>
static cudaError_t launch_ex(K kernel, dim3 grid, dim3 block, size_t smem, cudaStream_t stream, bool pdl, const P& params) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid, cfg.blockDim = block, cfg.dynamicSmemBytes = smem, cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr, cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, params);
}

cudaError_t launch_conv(const ConvParams& p, int precision, cudaStream_t stream, bool pdl) {
  if (precision == 0 && p.persist_stages > 0) {
    const size_t smem = conv_persist_smem_bytes(p);
    static PerDeviceSmem configured_p;
    {
      cudaError_t e = configured_p.ensure(smem, [](int bytes) {
        cudaError_t e = cudaSuccess;
#define MTLDAS_SET(E_, W_) \
  if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_umma_persist_kernel<E_, W_>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes)
        MTLDAS_SET(EPI_NONE, 0); MTLDAS_SET(EPI_RELU, 0); MTLDAS_SET(EPI_ADD_RELU, 0); MTLDAS_SET(EPI_SIGMOID_MUL, 0); MTLDAS_SET(EPI_ACCUM, 0);
        MTLDAS_SET(EPI_NONE, 1); MTLDAS_SET(EPI_RELU, 1); MTLDAS_SET(EPI_ADD_RELU, 1); MTLDAS_SET(EPI_SIGMOID_MUL, 1); MTLDAS_SET(EPI_ACCUM, 1);
#undef MTLDAS_SET
        return e;
      });
      if (e != cudaSuccess) return e;
    }
    const int tile = p.pool_k > 0 ? 2 * p.Wp * p.pool_k : p.MT * kTileM;
    const int ntiles = (int)((p.body + tile - 1) / tile);
    static int env_threads = -1, env_ctas = -1;
    if (env_threads < 0) {
      const char* a = getenv("MTLDAS_PERSIST_THREADS");
      const char* b = getenv("MTLDAS_PERSIST_CTAS");
      env_threads = a ? atoi(a) : 320;
      env_ctas = b ? atoi(b) : 2;
    }
    const int threads = env_threads == 320 ? 320 : 192;
    int per_sm = (int)std::min<size_t>((size_t)env_ctas, (220 * 1024) / smem);
    per_sm = std::min(per_sm, 65536 / (96 * threads));
    per_sm = std::max(per_sm, 1);
    ConvParams q = p;
    q.acc_sets = (per_sm * 2 * p.MT * p.Cout <= 512) ? 2 : 1;
    // K-split partials (see issue_tile_mmas): four independent accumulator chains for N <= 64 instructions, two for
    // N = 128; never more partials than K steps per tile; only where TMEM has the columns
    static int env_ks = -1;
    if (env_ks < 0) {
      const char* e = getenv("MTLDAS_PERSIST_KSPLIT");
      env_ks = e ? atoi(e) : 1;   // measured: no gain on B200 (the chains are not what binds), kept as an experiment switch
    }
    int KS = 1;
    const int want = p.Cout <= 64 ? 4 : 2, ksteps = p.ntaps * (p.nchunks >> 1);
    while (p.MT < 4 && p.MT * KS < want && KS * 2 <= env_ks && KS * 2 <= ksteps &&
           per_sm * q.acc_sets * p.MT * p.Cout * KS * 2 <= 512)
      KS *= 2;
    if (p.MT == 2 && KS > 2) KS = 2;
    if (p.epi == EPI_ADD_RELU || p.epi == EPI_SIGMOID_MUL) KS = 1;
    q.ksplit = KS;
    uint32_t cols = 32;
    while (cols < (uint32_t)(q.acc_sets * p.MT * p.Cout * KS)) cols <<= 1;
    per_sm = std::max(1, std::min(per_sm, (int)(512 / cols)));
    int ctas = std::min(ntiles, 148 * per_sm / std::max(1, p.ngroups));
    ctas = std::max(ctas, 1);
    dim3 grid(ctas, p.ngroups, 1);
    const bool wide = per_sm == 1 && p.Cout >= 64;
#define MTLDAS_GO(E_)                                                                                             \
  return wide ? launch_ex(conv_umma_persist_kernel<E_, 1>, grid, dim3(threads), smem, stream, pdl, q)             \
              : launch_ex(conv_umma_persist_kernel<E_, 0>, grid, dim3(threads), smem, stream, pdl, q)
    switch (p.epi) {
      case EPI_NONE: MTLDAS_GO(EPI_NONE);
      case EPI_RELU: MTLDAS_GO(EPI_RELU);
      case EPI_ADD_RELU: MTLDAS_GO(EPI_ADD_RELU);
      case EPI_ACCUM: MTLDAS_GO(EPI_ACCUM);
      default: MTLDAS_GO(EPI_SIGMOID_MUL);
    }
#undef MTLDAS_GO
  } else if (precision == 0) {
    size_t smem = conv_umma_smem_bytes(p);
    static PerDeviceSmem configured;
    {
      cudaError_t e = configured.ensure(smem, [](int bytes) {
        cudaError_t e = cudaFuncSetAttribute(conv_umma_kernel<EPI_NONE>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_umma_kernel<EPI_RELU>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_umma_kernel<EPI_ADD_RELU>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_umma_kernel<EPI_SIGMOID_MUL>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
        if (e == cudaSuccess) e = cudaFuncSetAttribute(conv_umma_kernel<EPI_ACCUM>, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
        return e;
      });
      if (e != cudaSuccess) return e;
    }
    int tile = p.MT * kTileM;
    dim3 grid((unsigned)((p.body + tile - 1) / tile), p.ngroups, 1);
    // one CTA per SM anyway (all of TMEM or most of shared memory): give it a second set of epilogue warps
    const dim3 block(p.tmem_cols > 256 && p.MT == 4 ? 256 : 128);  // (MT = 2 layers measured slower with 256)
    switch (p.epi) {
      case EPI_NONE: return launch_ex(conv_umma_kernel<EPI_NONE>, grid, block, smem, stream, pdl, p);
      case EPI_RELU: return launch_ex(conv_umma_kernel<EPI_RELU>, grid, block, smem, stream, pdl, p);
      case EPI_ADD_RELU: return launch_ex(conv_umma_kernel<EPI_ADD_RELU>, grid, block, smem, stream, pdl, p);
      case EPI_ACCUM: return launch_ex(conv_umma_kernel<EPI_ACCUM>, grid, block, smem, stream, pdl, p);
      default: return launch_ex(conv_umma_kernel<EPI_SIGMOID_MUL>, grid, block, smem, stream, pdl, p);
    }
  } else {
    size_t smem = (size_t)p.nchunks * 8 * kFp32CoutBlock * sizeof(float);
    dim3 grid((unsigned)((p.body + 127) / 128), p.ngroups, (p.Cout + kFp32CoutBlock - 1) / kFp32CoutBlock);
    return launch_ex(conv_fp32_kernel, grid, dim3(128), smem, stream, pdl, p);
  }
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// descriptor self test: D[128 x N] = A[128 x 16] * B[N x 16]^T with A rows taken at `row_shift`
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) umma_selftest_kernel(int N, int shift, int swap, float* max_err) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
  uint32_t* slot = reinterpret_cast<uint32_t*>(smem + 64);
  const int R = 128 + shift;
  __nv_bfloat16* A = reinterpret_cast<__nv_bfloat16*>(smem + 128);          // [2][R][8]
  __nv_bfloat16* B = A + 2 * R * 8;                                         // [2][N][8]
  auto aval = [](int r, int k) { return (float)(((r * 3 + k * 5) % 7) - 3); };
  auto bval = [](int n, int k) { return (float)(((n * 2 + k * 3) % 5) - 2); };
  for (int i = threadIdx.x; i < 2 * R * 8; i += 128) {
    int kc = i / (R * 8), r = (i / 8) % R, kk = i % 8;
    A[i] = __float2bfloat16(aval(r, kc * 8 + kk));
  }
  for (int i = threadIdx.x; i < 2 * N * 8; i += 128) {
    int kc = i / (N * 8), n = (i / 8) % N, kk = i % 8;
    B[i] = __float2bfloat16(bval(n, kc * 8 + kk));
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (threadIdx.x == 0) {
    ptx::mbar_init(bar, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc(slot, 256);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy smem writes -> async proxy (MMA)
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *slot;
  if (threadIdx.x == 0) {
    uint32_t a_chunk = (uint32_t)R * 16u, b_chunk = (uint32_t)N * 16u;
    uint64_t ad = ptx::umma_desc(ptx::smem_u32(A) + (uint32_t)shift * 16u, swap ? 128u : a_chunk, swap ? a_chunk : 128u);
    uint64_t bd = ptx::umma_desc(ptx::smem_u32(B), swap ? 128u : b_chunk, swap ? b_chunk : 128u);
    ptx::umma_bf16(tmem, ad, bd, ptx::umma_idesc_bf16(128, N, 0, 0), 0u);
    ptx::umma_commit(bar);
  }
  __syncwarp();
  ptx::mbar_wait(bar, 0);
  ptx::tc_fence_after();
  float err = 0.f;
  const int row = warp * 32 + lane;
  for (int c0 = 0; c0 < N; c0 += 16) {
    float v[16];
    ptx::tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + (uint32_t)c0, v);
    for (int j = 0; j < 16; ++j) {
      float ref = 0.f;
      for (int k = 0; k < 16; ++k) ref += aval(row + shift, k) * bval(c0 + j, k);
      err = fmaxf(err, fabsf(ref - v[j]));
    }
  }
  atomicMax(reinterpret_cast<int*>(max_err), __float_as_int(err));  // err >= 0: int order == float order
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem, 256);
}

float umma_selftest(int n_cols, int row_shift, int swap) {
  if (n_cols < 16 || n_cols > 256 || (n_cols % 16) || row_shift < 0 || row_shift > 256) return -1.f;
  float* d = nullptr;
  if (cudaMalloc(&d, sizeof(float)) != cudaSuccess) return -2.f;
  cudaMemset(d, 0, sizeof(float));
  size_t smem = 128 + (size_t)2 * (128 + row_shift) * 16 + (size_t)2 * n_cols * 16;
  umma_selftest_kernel<<<1, 128, smem>>>(n_cols, row_shift, swap, d);
  float h = -3.f;
  cudaError_t e = cudaDeviceSynchronize();
  if (e == cudaSuccess) e = cudaMemcpy(&h, d, sizeof(float), cudaMemcpyDeviceToHost);
  cudaFree(d);
  if (e != cudaSuccess) {
    fprintf(stderr, "mtldas selftest: %s\n", cudaGetErrorString(e));
    return -3.f;
  }
  return h;
}


// ---------------------------------------------------------------------------------------------
// tensor-pipe rate probe: `reps` x 16 back-to-back 128 x N x 16 MMAs on fixed operands.
// mode 0: interleaved (no swizzle) descriptors as used by the conv kernel; mode 1: 128B-swizzle descriptors;
// mode 2: interleaved, descriptor start address advanced every MMA (register arithmetic in the loop).
// Returns SM cycles per MMA as seen by the issuing thread (issue start -> commit observed).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) umma_rate_kernel(int N, int mode, int reps, float* out) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bar = reinterpret_cast<uint64_t*>(smem);
  uint32_t* slot = reinterpret_cast<uint32_t*>(smem + 64);
  uint8_t* A = smem + 1024;            // 16 KB region
  uint8_t* B = smem + 1024 + 16384;    // 16 KB region
  for (int i = threadIdx.x; i < 8192; i += 128) reinterpret_cast<uint32_t*>(A)[i] = 0x3c003c00u;  // both regions
  const int warp = threadIdx.x >> 5;
  if (threadIdx.x == 0) {
    ptx::mbar_init(bar, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc(slot, 256);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = *slot;
  if (threadIdx.x == 32) {
    const uint32_t idesc = ptx::umma_idesc_bf16(128, N, 0, 0);
    uint64_t ad, bd;
    if (mode == 1) {  // SWIZZLE_128B K-major: SBO = 1024, layout type 2
      ad = ptx::umma_desc(ptx::smem_u32(A), 16, 1024) | (2ull << 61);
      bd = ptx::umma_desc(ptx::smem_u32(B), 16, 1024) | (2ull << 61);
    } else {
      ad = ptx::umma_desc(ptx::smem_u32(A), 4096, 128);
      bd = ptx::umma_desc(ptx::smem_u32(B), (uint32_t)N * 16u, 128);
    }
    const long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
#pragma unroll
      for (int i = 0; i < 16; ++i) {
        uint64_t a2 = ad;
        if (mode == 2) a2 = ad + (uint64_t)((r * 16 + i) & 7);
        ptx::umma_bf16(tmem, a2, bd, idesc, 1u);
      }
    }
    const long long t1 = clock64();
    ptx::umma_commit(bar);
    ptx::mbar_wait(bar, 0);
    const long long t2 = clock64();
    out[0] = (float)(t2 - t0) / (float)(reps * 16);
    out[1] = (float)(t1 - t0) / (float)(reps * 16);
  }
  __syncthreads();
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem, 256);
}

int umma_rate(int n_cols, int mode, int reps, float* host_out2) {
  float* d = nullptr;
  if (cudaMalloc(&d, 2 * sizeof(float)) != cudaSuccess) return -2;
  size_t smem = 1024 + 32768;
  umma_rate_kernel<<<1, 128, smem>>>(n_cols, mode, reps, d);
  cudaError_t e = cudaDeviceSynchronize();
  if (e == cudaSuccess) e = cudaMemcpy(host_out2, d, 2 * sizeof(float), cudaMemcpyDeviceToHost);
  cudaFree(d);
  return e == cudaSuccess ? 0 : -2;
}

}  // namespace mtldas
