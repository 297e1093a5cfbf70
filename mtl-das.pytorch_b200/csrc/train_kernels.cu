// Training-mode kernels of the MTL-DAS hot path: weight gradients (tcgen05 and CUDA-core), joint NLL loss, head
// backward, gradient unpack.  The BatchNorm / activation kernels live in train_ew.cu.
//
// Reference semantics restated (autograd of /root/reference/model/modelA_MTL.py:127-174 under utils.py:359-373):
//   BatchNorm2d train: y_hat = (y - mean) * invstd (biased batch variance), out = gamma*y_hat + beta;
//     running_mean/var updated with momentum 0.1 (unbiased variance), num_batches_tracked += 1;
//   backward: dgamma = sum(g*y_hat), dbeta = sum(g), dy = gamma*invstd*(g - mean(g) - y_hat*mean(g*y_hat));
//   ReLU' = [out > 0]; sigmoid' = s(1-s); mask multiply: d mask = dz*feat, d feat = dz*mask;
//   MaxPool2d(2,2,ceil): gradient to the first maximal element of the window (scan order h, w).
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "gran.cuh"
#include "kernels.h"
#include "ptx.cuh"
#include "train_kernels.h"

namespace mtldas {


// ---------------------------------------------------------------------------------------------
// weight gradient on CUDA cores: dW[tap][k][cout] += sum_pos zin[pos+off][k] * dy[pos][cout]
// grid (position slices, taps); block 256 = 16 x 16 thread grid, each thread (K/16) x (Cout/16) outputs.
// ---------------------------------------------------------------------------------------------
constexpr int kWgPos = 32;  // positions staged per iteration

template <typename T, int KR, int CR>
__global__ void __launch_bounds__(256) wgrad_kernel(const __grid_constant__ WgradParams p) {
  extern __shared__ __align__(16) float wsm[];
  const int K = p.nchunks * 8, Cout = p.Cout;
  float* zs = wsm;                       // [kWgPos][K]
  float* ds = wsm + kWgPos * K;          // [kWgPos][Cout]
  const ConvTapDev tp = p.tap[blockIdx.y];
  const ConvSrcDev src = p.src[tp.src];
  const long long f0 = (long long)blockIdx.x * p.slice;
  const long long f1 = min(f0 + (long long)p.slice, p.body);
  const int tk = threadIdx.x >> 4, tc = threadIdx.x & 15;
  constexpr int G = Gran<T>::bytes;
  float acc[KR][CR];
#pragma unroll
  for (int i = 0; i < KR; ++i)
#pragma unroll
    for (int j = 0; j < CR; ++j) acc[i][j] = 0.f;
  for (long long fb = f0; fb < f1; fb += kWgPos) {
    __syncthreads();
    for (int i = threadIdx.x; i < kWgPos * p.nchunks; i += 256) {
      const int c = i / kWgPos, pp = i - c * kWgPos;
      float v[8];
      if (fb + pp < f1)
        load8<T>(src.base + (long long)c * src.plane_bytes + (fb + pp + tp.off) * G, v);
      else {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = 0.f;
      }
      float4* d = reinterpret_cast<float4*>(zs + pp * K + c * 8);
      d[0] = make_float4(v[0], v[1], v[2], v[3]);
      d[1] = make_float4(v[4], v[5], v[6], v[7]);
    }
    for (int i = threadIdx.x; i < kWgPos * (Cout / 8); i += 256) {
      const int c = i / kWgPos, pp = i - c * kWgPos;
      float v[8];
      if (fb + pp < f1)
        load8<T>(p.dy + (long long)c * p.dy_plane_bytes + (fb + pp) * G, v);
      else {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = 0.f;
      }
      float4* d = reinterpret_cast<float4*>(ds + pp * Cout + c * 8);
      d[0] = make_float4(v[0], v[1], v[2], v[3]);
      d[1] = make_float4(v[4], v[5], v[6], v[7]);
    }
    __syncthreads();
#pragma unroll 4
    for (int pp = 0; pp < kWgPos; ++pp) {
      float zv[KR], dv[CR];
#pragma unroll
      for (int i = 0; i < KR; ++i) zv[i] = zs[pp * K + tk + 16 * i];
#pragma unroll
      for (int j = 0; j < CR; ++j) dv[j] = ds[pp * Cout + tc + 16 * j];
#pragma unroll
      for (int i = 0; i < KR; ++i)
#pragma unroll
        for (int j = 0; j < CR; ++j) acc[i][j] = fmaf(zv[i], dv[j], acc[i][j]);
    }
  }
  float* dw = p.dw + (size_t)blockIdx.y * K * Cout;
#pragma unroll
  for (int i = 0; i < KR; ++i)
#pragma unroll
    for (int j = 0; j < CR; ++j) atomicAdd(dw + (size_t)(tk + 16 * i) * Cout + tc + 16 * j, acc[i][j]);
}

template <typename T>
static cudaError_t launch_wgrad_t(const WgradParams& p, cudaStream_t st) {
  const int K = p.nchunks * 8;
  dim3 grid((unsigned)((p.body + p.slice - 1) / p.slice), p.ntaps);
  const size_t smem = (size_t)kWgPos * (K + p.Cout) * sizeof(float);
  const int kr = K / 16, cr = p.Cout / 16;
#define WG(KR_, CR_)                                                                       \
  if (kr == KR_ && cr == CR_) {                                                            \
    wgrad_kernel<T, KR_, CR_><<<grid, 256, smem, st>>>(p);                                 \
    return cudaGetLastError();                                                             \
  }
  WG(1, 1) WG(1, 2) WG(1, 4) WG(1, 8) WG(2, 1) WG(2, 2) WG(2, 4) WG(2, 8)
  WG(4, 1) WG(4, 2) WG(4, 4) WG(4, 8) WG(8, 1) WG(8, 2) WG(8, 4) WG(8, 8)
#undef WG
  return cudaErrorInvalidValue;
}


// ---------------------------------------------------------------------------------------------
// weight gradient on tcgen05 (bf16 path):  D_t[co][ci] = sum_pos dy[pos][co] * zin[pos + off_t][ci]
// Both operands are read from shared memory in the MN-major interleaved layout, which is exactly the P8 slab
// [chunk][position][8 channels]: the MMA's K dimension is the POSITION axis (16 positions per instruction),
// M = output channels (128 rows; rows >= Cout are zero), N = input channels of one tap, and a tap is again a
// start-address shift of the B descriptor.  The accumulators of all taps of a tap group stay in TMEM for the whole
// position slice of the CTA (grid.x = slices, grid.y = tap groups) and are added to the packed fp32 gradient
// [tap][ci][co] with coalesced red.global.add (32 consecutive co per warp instruction).
// ---------------------------------------------------------------------------------------------
constexpr int kWgStage = 128;   // positions per pipeline stage
constexpr int kWgHeader = 256;

__global__ void __launch_bounds__(128) wgrad_umma_kernel(const __grid_constant__ WgradParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);  // [0,1] full, [2,3] empty, [4] acc
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 64);
  const int Cin = p.nchunks * 8, Cout = p.Cout;
  const int tap0 = blockIdx.y * p.taps_per_cta;
  const int ntap = min(p.taps_per_cta, p.ntaps - tap0);
  const uint32_t a_chunk = kWgStage * 16u;                          // dy slab: [16 chunk slots][128][16 B]
  const uint32_t a_stage = 16u * a_chunk;
  const int SPb = kWgStage + p.halo_lo + p.halo_hi;
  const uint32_t b_chunk = (uint32_t)SPb * 16u;                     // zin slab: [nsrc*nchunks][SPb][16 B]
  const uint32_t b_stage = ((uint32_t)(p.nsrc * p.nchunks) * b_chunk + 127u) & ~127u;
  uint8_t* a_slab = smem + kWgHeader;
  uint8_t* b_slab = a_slab + 2 * a_stage;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long f0 = (long long)blockIdx.x * p.slice;
  const long long f1 = min(f0 + (long long)p.slice, (p.body + 15) / 16 * 16);
  const int nstage = (int)((f1 - f0 + kWgStage - 1) / kWgStage);
  uint32_t cols = 32;
  while ((int)cols < ntap * Cin) cols <<= 1;

  if (threadIdx.x == 0) {
    for (int i = 0; i < 5; ++i) ptx::mbar_init(bars + i, 1);
    ptx::fence_mbar_init();
  }
  // rows >= Cout of the M = 128 operand: zero once (never overwritten by the bulk copies)
  for (int i = threadIdx.x; i < (int)(2 * a_stage / 16); i += 128) reinterpret_cast<uint4*>(a_slab)[i] = make_uint4(0, 0, 0, 0);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  if (warp == 1) ptx::tmem_alloc(tmem_slot, cols);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (ptx::elect_one()) {
      for (int s = 0; s < nstage; ++s) {
        const int b = s & 1;
        if (s >= 2) ptx::mbar_wait(bars + 2 + b, ((s >> 1) - 1) & 1);
        const long long fb = f0 + (long long)s * kWgStage;
        const uint32_t bytes = (uint32_t)(Cout / 8) * a_chunk + (uint32_t)(p.nsrc * p.nchunks) * b_chunk;
        ptx::mbar_arrive_expect_tx(bars + b, bytes);
        for (int c = 0; c < Cout / 8; ++c)
          ptx::bulk_g2s(a_slab + b * a_stage + c * a_chunk, p.dy + (long long)c * p.dy_plane_bytes + fb * 16, a_chunk, bars + b);
        for (int sr = 0; sr < p.nsrc; ++sr)
          for (int c = 0; c < p.nchunks; ++c)
            ptx::bulk_g2s(b_slab + b * b_stage + (size_t)(sr * p.nchunks + c) * b_chunk,
                          p.src[sr].base + (long long)c * p.src[sr].plane_bytes + (fb - p.halo_lo) * 16, b_chunk, bars + b);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (ptx::elect_one()) {
      const uint32_t idesc = ptx::umma_idesc_bf16(128, Cin, 1, 1);
      // MN-major interleaved: LBO = distance between 8-position groups (128 B), SBO = distance between 8-channel chunks
      const uint64_t a_hi = ptx::umma_desc(0, 128u, a_chunk);
      const uint64_t b_hi = ptx::umma_desc(0, 128u, b_chunk);
      const uint32_t a16 = ptx::smem_u32(a_slab) >> 4, b16 = ptx::smem_u32(b_slab) >> 4;
      uint32_t acc = 0u;
      uint32_t tap_off[kMaxTaps];  // per-tap B-operand offsets in registers (no parameter loads in the issue loop)
#pragma unroll
      for (int t = 0; t < kMaxTaps; ++t)
        tap_off[t] = t < ntap ? (uint32_t)(p.tap[tap0 + t].src * p.nchunks) * (b_chunk >> 4) +
                                    (uint32_t)(p.halo_lo + p.tap[tap0 + t].off)
                              : 0u;
      for (int s = 0; s < nstage; ++s) {
        const int b = s & 1;
        ptx::mbar_wait(bars + b, (s >> 1) & 1);
        ptx::tc_fence_after();
        const long long fb = f0 + (long long)s * kWgStage;
        const int ksteps = (int)min((long long)(kWgStage / 16), (f1 - fb + 15) / 16);
        for (int j = 0; j < ksteps; ++j) {
          const uint32_t a_lo = a16 + (uint32_t)b * (a_stage >> 4) + (uint32_t)j * 16u;
          const uint32_t b_j = b16 + (uint32_t)b * (b_stage >> 4) + (uint32_t)(j * 16);
#pragma unroll
          for (int t = 0; t < kMaxTaps; ++t)
            if (t < ntap) ptx::umma_bf16(tmem_base + (uint32_t)(t * Cin), a_hi | a_lo, b_hi | (b_j + tap_off[t]), idesc, acc);
          acc = 1u;
        }
        ptx::umma_commit(bars + 2 + b);
      }
      ptx::umma_commit(bars + 4);
    }
    __syncwarp();
  }

  // ---- epilogue: lane = output channel (row of D); columns = (tap, ci) ------------------------------
  ptx::mbar_wait(bars + 4, 0);
  ptx::tc_fence_after();
  const int co = warp * 32 + lane;
  const int K = Cin;
  for (int t = 0; t < ntap; ++t) {
    for (int c0 = 0; c0 < Cin; c0 += 16) {
      uint32_t r[16];
      ptx::tmem_ld16_issue(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(t * Cin + c0), r);
      ptx::tmem_wait16(r);
      if (co < Cout && nstage > 0) {
        float* dst = p.dw + ((size_t)(tap0 + t) * K + c0) * Cout + co;
#pragma unroll
        for (int j = 0; j < 16; ++j) atomicAdd(dst + (size_t)j * Cout, __uint_as_float(r[j]));
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, cols);
}

// ---------------------------------------------------------------------------------------------
// weight gradient, narrow layers (Cin <= 64, 3x3): the taps side by side along the MMA's N dimension.
// An M = 128, K = 16 instruction costs ~45 cycles of A-operand reads whatever N <= 64 is, so nine N = Cin instructions
// per 16 positions waste the tensor pipe; here the stage's input slab (128 positions + halo, one bulk copy per chunk
// plane as before) is re-laid by six copy warps into one slot per (tap, chunk) holding the positions ALREADY SHIFTED
// by the tap, and ONE instruction with N = taps*Cin (<= 256) covers them all.  The replication stays inside shared
// memory: doing it with shifted bulk copies would multiply the L2 traffic by the tap count.
// A (dy) needs only Cout/8 real chunk slots: the instruction reads 16, the rest alias whatever follows in shared
// memory and feed accumulator lanes >= Cout that nobody reads.
// ---------------------------------------------------------------------------------------------
constexpr int kWg2Threads = 256, kWg2Rep = 6;
constexpr int kWg2Ring = 4;  // max stages of the dy / raw-slab ring (the bulk-copy latency is ~4 stages of MMA work)
constexpr int kWg2AFull = 0, kWg2RawFull = 4, kWg2RawEmpty = 8, kWg2AEmpty = 12, kWg2BFull = 16, kWg2MmaDone = 18, kWg2Acc = 20;

__global__ void __launch_bounds__(kWg2Threads) wgrad_umma2_kernel(const __grid_constant__ WgradParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 200);
  const int Cin = p.nchunks * 8, Cout = p.Cout, NR = p.nring;
  const int tap0 = blockIdx.y * p.taps_per_cta;
  const int ntap = min(p.taps_per_cta, p.ntaps - tap0);
  const uint32_t chunk2k = kWgStage * 16u;
  const uint32_t a_real = (uint32_t)(Cout / 8) * chunk2k;
  const int SPb = kWgStage + p.halo_lo + p.halo_hi;
  const uint32_t raw_chunk = (uint32_t)SPb * 16u;
  const uint32_t raw_stage = ((uint32_t)(p.nsrc * p.nchunks) * raw_chunk + 127u) & ~127u;
  const uint32_t b_stage = (uint32_t)(p.taps_per_cta * p.nchunks) * chunk2k;
  uint8_t* a_slab = smem + kWgHeader;
  uint8_t* raw = a_slab + NR * a_real;
  uint8_t* bexp = raw + NR * raw_stage;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long f0 = (long long)blockIdx.x * p.slice;
  const long long f1 = min(f0 + (long long)p.slice, (p.body + 15) / 16 * 16);
  const int nstage = (int)((f1 - f0 + kWgStage - 1) / kWgStage);
  uint32_t cols = 32;
  while ((int)cols < ntap * Cin) cols <<= 1;

  if (threadIdx.x == 0) {
    for (int i = 0; i < kWg2Ring; ++i) {
      ptx::mbar_init(bars + kWg2AFull + i, 1), ptx::mbar_init(bars + kWg2RawFull + i, 1);
      ptx::mbar_init(bars + kWg2RawEmpty + i, kWg2Rep), ptx::mbar_init(bars + kWg2AEmpty + i, 1);
    }
    ptx::mbar_init(bars + kWg2BFull + 0, kWg2Rep), ptx::mbar_init(bars + kWg2BFull + 1, kWg2Rep);
    ptx::mbar_init(bars + kWg2MmaDone + 0, 1), ptx::mbar_init(bars + kWg2MmaDone + 1, 1);
    ptx::mbar_init(bars + kWg2Acc, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, cols);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ---- producer: dy chunks -> A stage, input slab (tile + halo) -> raw stage ---------------------------------
    if (ptx::elect_one()) {
      for (int s = 0; s < nstage; ++s) {
        const int b = s % NR;
        if (s >= NR) {
          ptx::mbar_wait(bars + kWg2AEmpty + b, ((s / NR) - 1) & 1);    // A stage b consumed
          ptx::mbar_wait(bars + kWg2RawEmpty + b, ((s / NR) - 1) & 1);  // raw stage b re-laid
        }
        const long long fb = f0 + (long long)s * kWgStage;
        ptx::mbar_arrive_expect_tx(bars + kWg2AFull + b, a_real);
        for (int c = 0; c < Cout / 8; ++c)
          ptx::bulk_g2s(a_slab + b * a_real + c * chunk2k, p.dy + (long long)c * p.dy_plane_bytes + fb * 16, chunk2k,
                        bars + kWg2AFull + b);
        ptx::mbar_arrive_expect_tx(bars + kWg2RawFull + b, (uint32_t)(p.nsrc * p.nchunks) * raw_chunk);
        for (int sr = 0; sr < p.nsrc; ++sr)
          for (int c = 0; c < p.nchunks; ++c)
            ptx::bulk_g2s(raw + b * raw_stage + (size_t)(sr * p.nchunks + c) * raw_chunk,
                          p.src[sr].base + (long long)c * p.src[sr].plane_bytes + (fb - p.halo_lo) * 16, raw_chunk,
                          bars + kWg2RawFull + b);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ---- MMA issuer --------------------------------------------------------------------------------------------
    if (ptx::elect_one()) {
      // MN-major interleaved: LBO = distance between 8-position groups (128 B), SBO = distance between 8-channel chunks
      const uint64_t ab_hi = ptx::umma_desc(0, 128u, chunk2k);
      const uint32_t a16 = ptx::smem_u32(a_slab) >> 4, b16 = ptx::smem_u32(bexp) >> 4;
      const int tpg = max(1, 256 / Cin), ngrp = (ntap + tpg - 1) / tpg;  // taps per instruction (N <= 256)
      const uint32_t idesc_full = ptx::umma_idesc_bf16(128, min(tpg, ntap) * Cin, 1, 1);
      const uint32_t idesc_last = ptx::umma_idesc_bf16(128, (ntap - (ngrp - 1) * tpg) * Cin, 1, 1);
      uint32_t acc = 0u;
      for (int s = 0; s < nstage; ++s) {
        const int b = s & 1, ra = s % NR;
        ptx::mbar_wait(bars + kWg2AFull + ra, (s / NR) & 1);
        ptx::mbar_wait(bars + kWg2BFull + b, (s >> 1) & 1);
        ptx::tc_fence_after();
        const long long fb = f0 + (long long)s * kWgStage;
        const int ksteps = (int)min((long long)(kWgStage / 16), (f1 - fb + 15) / 16);
        for (int j = 0; j < ksteps && !(p.dbg_flags & 2); ++j) {
          const uint32_t a_lo = a16 + (uint32_t)ra * (a_real >> 4) + (uint32_t)j * 16u;
          const uint32_t b_j = b16 + (uint32_t)b * (b_stage >> 4) + (uint32_t)(j * 16);
          for (int gq = 0; gq < ngrp; ++gq)
            ptx::umma_bf16(tmem_base + (uint32_t)(gq * tpg * Cin), ab_hi | a_lo,
                           ab_hi | (b_j + (uint32_t)(gq * tpg * p.nchunks) * (chunk2k >> 4)),
                           gq == ngrp - 1 ? idesc_last : idesc_full, acc);
          acc = 1u;
        }
        ptx::umma_commit(bars + kWg2AEmpty + ra);
        ptx::umma_commit(bars + kWg2MmaDone + b);
      }
      ptx::umma_commit(bars + kWg2Acc);
    }
    __syncwarp();
  } else {
    // ---- copy warps: slot (tap, chunk) = the raw chunk plane shifted by the tap --------------------------------
    const int tr = threadIdx.x - 64, nslot = ntap * p.nchunks;
    for (int s = 0; s < nstage; ++s) {
      const int b = s & 1, rr = s % NR;
      ptx::mbar_wait(bars + kWg2RawFull + rr, (s / NR) & 1);
      if (s >= 2) ptx::mbar_wait(bars + kWg2MmaDone + b, ((s >> 1) - 1) & 1);  // expanded stage b consumed
      const uint8_t* rsrc = raw + rr * raw_stage;
      uint8_t* dst = bexp + b * b_stage;
      for (int i = tr; i < nslot * (kWgStage / 2) && !(p.dbg_flags & 1); i += kWg2Rep * 32) {   // two granules per thread and step
        const int slot = i >> 6, pos = (i & 63) * 2;
        const int t = slot / p.nchunks, c = slot - t * p.nchunks;
        const ConvTapDev tp = p.tap[tap0 + t];
        const uint4* sp = reinterpret_cast<const uint4*>(rsrc + (size_t)(tp.src * p.nchunks + c) * raw_chunk +
                                                         (size_t)(p.halo_lo + tp.off + pos) * 16);
        uint4* dp = reinterpret_cast<uint4*>(dst + (size_t)slot * chunk2k + (size_t)pos * 16);
        const uint4 v0 = sp[0], v1 = sp[1];
        dp[0] = v0, dp[1] = v1;
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> tensor-core reads
      __syncwarp();
      if (lane == 0) {
        ptx::mbar_arrive(bars + kWg2BFull + b);
        ptx::mbar_arrive(bars + kWg2RawEmpty + rr);
      }
    }
  }

  // ---- epilogue: lane = output channel (row of D); columns = (tap, ci); the two warps of a lane quarter split the taps
  ptx::mbar_wait(bars + kWg2Acc, 0);
  ptx::tc_fence_after();
  const int quarter = warp & 3, co = quarter * 32 + lane;
  if (quarter * 32 < Cout) {
    for (int t = warp >> 2; t < ntap; t += 2) {
      for (int c0 = 0; c0 < Cin; c0 += 16) {
        uint32_t r[16];
        ptx::tmem_ld16_issue(tmem_base + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(t * Cin + c0), r);
        ptx::tmem_wait16(r);
        if (co < Cout && nstage > 0) {
          float* dst = p.dw + ((size_t)(tap0 + t) * Cin + c0) * Cout + co;
#pragma unroll
          for (int j = 0; j < 16; ++j) atomicAdd(dst + (size_t)j * Cout, __uint_as_float(r[j]));
        }
      }
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) ptx::tmem_dealloc(tmem_base, cols);
}

// shared memory of wgrad_umma2_kernel; the A operand reads 32 KB from each A stage base whatever Cout is
static size_t wgrad2_smem(const WgradParams& p, int taps_per_cta, int nring) {
  const size_t a_real = (size_t)(p.Cout / 8) * kWgStage * 16;
  const size_t raw_stage = ((size_t)(p.nsrc * p.nchunks) * (kWgStage + p.halo_lo + p.halo_hi) * 16 + 127) & ~(size_t)127;
  const size_t b_stage = (size_t)taps_per_cta * p.nchunks * kWgStage * 16;
  const size_t used = kWgHeader + nring * (a_real + raw_stage) + 2 * b_stage;
  return std::max(used, (size_t)kWgHeader + (size_t)(nring - 1) * a_real + 16 * (size_t)kWgStage * 16);
}

static cudaError_t launch_wgrad_umma2(WgradParams p, cudaStream_t st, bool* handled) {
  const int Cin = p.nchunks * 8;
  *handled = false;
  if (Cin > 64 || p.ntaps < 3) return cudaSuccess;
  int lo = 0, hi = 0;
  for (int t = 0; t < p.ntaps; ++t) lo = std::max(lo, -p.tap[t].off), hi = std::max(hi, p.tap[t].off);
  p.halo_lo = lo, p.halo_hi = hi;
  int tpc = std::min(p.ntaps, 512 / Cin);
  while (tpc > 1 && wgrad2_smem(p, tpc, 2) > 200 * 1024) --tpc;
  if (wgrad2_smem(p, tpc, 2) > 200 * 1024) return cudaSuccess;
  const int ngroups = (p.ntaps + tpc - 1) / tpc;
  p.taps_per_cta = (p.ntaps + ngroups - 1) / ngroups;
  {
    const char* d = getenv("MTLDAS_WG2_DEBUG");
    p.dbg_flags = d ? atoi(d) : 0;
  }
  // deepest ring that still lets two CTAs share an SM (measured: two CTAs with a 2-stage ring beat one CTA with
  // four stages, 50 vs 59 us on the 16-channel layers at batch 256); otherwise the deepest that fits
  p.nring = 0;
  for (int r = kWg2Ring; r >= 2 && !p.nring; --r)
    if (wgrad2_smem(p, p.taps_per_cta, r) <= 113 * 1024) p.nring = r;
  for (int r = kWg2Ring; r >= 2 && !p.nring; --r)
    if (wgrad2_smem(p, p.taps_per_cta, r) <= 220 * 1024) p.nring = r;
  if (!p.nring) return cudaSuccess;
  const size_t smem = wgrad2_smem(p, p.taps_per_cta, p.nring);
  int cols = 32;
  while (cols < p.taps_per_cta * Cin) cols <<= 1;
  const int resident = std::max(1, std::min({2, 512 / cols, (int)((size_t)227 * 1024 / (smem + 1024))}));
  const long long body16 = (p.body + 15) / 16 * 16;
  const long long want = std::max<long long>(1, (long long)148 * resident / ngroups);
  const long long slice = std::max<long long>(512, (body16 + want - 1) / want);
  p.slice = (int)((slice + kWgStage - 1) / kWgStage * kWgStage);
  static PerDeviceSmem configured;
  {
    cudaError_t e = configured.ensure(smem, [](int bytes) {
      return cudaFuncSetAttribute(wgrad_umma2_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    });
    if (e != cudaSuccess) return e;
  }
  dim3 grid((unsigned)((body16 + p.slice - 1) / p.slice), ngroups);
  wgrad_umma2_kernel<<<grid, kWg2Threads, smem, st>>>(p);
  *handled = true;
  return cudaGetLastError();
}

static cudaError_t launch_wgrad_umma(WgradParams p, cudaStream_t st) {
  const int Cin = p.nchunks * 8;
  const int tg_max = std::max(1, 512 / Cin);
  const int ngroups = (p.ntaps + tg_max - 1) / tg_max;
  p.taps_per_cta = (p.ntaps + ngroups - 1) / ngroups;
  int lo = 0, hi = 0;
  for (int t = 0; t < p.ntaps; ++t) lo = std::max(lo, -p.tap[t].off), hi = std::max(hi, p.tap[t].off);
  p.halo_lo = lo, p.halo_hi = hi;
  const long long body16 = (p.body + 15) / 16 * 16;
  const int SPb = kWgStage + lo + hi;
  const size_t b_stage = ((size_t)(p.nsrc * p.nchunks) * SPb * 16 + 127) & ~(size_t)127;
  const size_t smem = kWgHeader + 2 * (size_t)16 * kWgStage * 16 + 2 * b_stage;
  // one wave: CTAs resident per SM are bounded by the TMEM columns (ntaps*Cin rounded to a power of two) and by
  // shared memory; the position slices are sized so that the whole grid is resident at once
  int cols = 32;
  while (cols < p.taps_per_cta * Cin) cols <<= 1;
  const int resident = std::max(1, std::min({3, 512 / cols, (int)((size_t)227 * 1024 / (smem + 1024))}));
  long long want = std::max<long long>(1, (long long)148 * resident / ngroups);
  long long slice = std::max<long long>(512, (body16 + want - 1) / want);
  p.slice = (int)((slice + kWgStage - 1) / kWgStage * kWgStage);
  static PerDeviceSmem configured;
  {
    cudaError_t e = configured.ensure(smem, [](int bytes) {
      return cudaFuncSetAttribute(wgrad_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    });
    if (e != cudaSuccess) return e;
  }
  dim3 grid((unsigned)((body16 + p.slice - 1) / p.slice), ngroups);
  wgrad_umma_kernel<<<grid, 128, smem, st>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_wgrad(const WgradParams& p, int precision, cudaStream_t st) {
  if (precision == 0) {
    const char* e = getenv("MTLDAS_WGRAD_CUDA_CORES");
    if (!(e && e[0] == '1')) {
      static int wide = -1;
      if (wide < 0) {
        const char* w = getenv("MTLDAS_WGRAD_WIDE");
        wide = (w && w[0] == '0') ? 0 : 1;
      }
      if (wide) {
        bool handled = false;
        cudaError_t ce = launch_wgrad_umma2(p, st, &handled);
        if (ce != cudaSuccess || handled) return ce;
      }
      return launch_wgrad_umma(p, st);
    }
    return launch_wgrad_t<__nv_bfloat16>(p, st);
  }
  return launch_wgrad_t<float>(p, st);
}

// ---------------------------------------------------------------------------------------------
// joint NLL loss of the task heads and its gradient (utils.py:139,361,367: NLLLoss(mean) per task, summed):
//   L = -(1/n) sum_i logp_t[i, y_t[i]] summed over the tasks;  dL/dlogp_t[i, j] = -(1/n) [j == y_t[i]]
// One block: the loss is a deterministic tree sum; the gradients are written densely (they feed head_bwd).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) nll_loss_kernel(const float* logp0, const long long* y0, int c0, float* d0,
                                                       const float* logp1, const long long* y1, int c1, float* d1,
                                                       int n, float* loss, int* bad_label) {
  __shared__ float sh[8];
  const float k = 1.f / (float)n;
  float acc = 0.f;
  for (int t = 0; t < 2; ++t) {
    const float* lp = t ? logp1 : logp0;
    const long long* y = t ? y1 : y0;
    float* d = t ? d1 : d0;
    const int c = t ? c1 : c0;
    if (!lp) continue;
    for (int i = threadIdx.x; i < n * c; i += 256) {
      const int row = i / c, col = i - row * c;
      const long long lab = y[row];
      const bool hit = lab == col;
      if (d) d[i] = hit ? -k : 0.f;
      if (hit) acc -= lp[i] * k;
      if (bad_label && col == 0 && (lab < 0 || lab >= c)) *bad_label = 1;
    }
  }
  for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
  if ((threadIdx.x & 31) == 0) sh[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int w = 0; w < 8; ++w) s += sh[w];
    *loss = s;
  }
}

cudaError_t launch_nll_loss(const float* logp0, const long long* y0, int c0, float* d0, const float* logp1,
                            const long long* y1, int c1, float* d1, int n, float* loss, int* bad_label, cudaStream_t st) {
  nll_loss_kernel<<<1, 256, 0, st>>>(logp0, y0, c0, d0, logp1, y1, c1, d1, n, loss, bad_label);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// head backward: d a4[n,c,h,w] = dz[n, c / group] / (H*W*group), dz = dlogp - softmax * sum(dlogp)
// one block per sample, thread t: chunk t%16, positions t/16, t/16+8, ...
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(128) head_bwd_kernel(const __grid_constant__ HeadBwdParams p) {
  __shared__ float dzc[16];
  const int n = blockIdx.x;
  if (threadIdx.x == 0) {
    float s = 0.f;
    for (int j = 0; j < p.classes; ++j) s += p.dlogp[(size_t)n * p.classes + j];
    const int group = 128 / p.classes;
    const float k = 1.f / (float)(p.H * p.W * group);
    for (int j = 0; j < p.classes; ++j)
      dzc[j] = (p.dlogp[(size_t)n * p.classes + j] - expf(p.logp[(size_t)n * p.classes + j]) * s) * k;
  }
  __syncthreads();
  const int c = threadIdx.x & 15, pl = threadIdx.x >> 4;
  const int group = 128 / p.classes;
  float v[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = dzc[(c * 8 + j) / group];
  float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  for (int q = pl; q < p.Pimg; q += 8) {
    const int h = q / p.Wp, w = q - h * p.Wp;
    const bool valid = h < p.H && w < p.W;
    store8<T>(p.da4.base + ((long long)c * p.da4.plane + p.da4.g_lo + (long long)n * p.Pimg + q) * Gran<T>::bytes,
              valid ? v : z);
  }
}

cudaError_t launch_head_bwd(const HeadBwdParams& p, int precision, cudaStream_t st) {
  if (precision == 0)
    head_bwd_kernel<__nv_bfloat16><<<p.nb, 128, 0, st>>>(p);
  else
    head_bwd_kernel<float><<<p.nb, 128, 0, st>>>(p);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// packed fp32 weight gradient -> reference parameter layout (inverse of pack_kernel's placement)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) unpack_batch_kernel(const __grid_constant__ UnpackBatch b) {
  const UnpackParams& p = b.u[blockIdx.y];
  const int kk = p.ksize * p.ksize;
  const int total = p.cout * p.cin * kk;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < total; i += gridDim.x * 256) {
    const int co = i / (p.cin * kk);
    const int r1 = i - co * (p.cin * kk);
    const int ci = r1 / kk;
    int tap = r1 - ci * kk, k = ci;
    if (ci >= p.split) tap += 1, k -= p.split;
    k += p.k_off;
    p.dst[i] = p.src[((size_t)tap * p.K_tap + k) * p.Cout_p + p.cout_off + co];
  }
}

cudaError_t launch_unpack_batch(const UnpackParams* u, int n, cudaStream_t st) {
  static UnpackBatch b;
  for (int i0 = 0; i0 < n; i0 += kUnpackBatch) {
    const int m = std::min(kUnpackBatch, n - i0);
    for (int i = 0; i < m; ++i) b.u[i] = u[i0 + i];
    unpack_batch_kernel<<<dim3(48, m), 256, 0, st>>>(b);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}


}  // namespace mtldas
