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
#include <cstdio>
#include <cuda_bf16.h>

#include "gran.cuh"
#include "kernels.h"
#include "ptx.cuh"

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
    }
  } else {
    if (g.out.layout != LAY_P8) return;  // phase-split tensors keep their (pre-zeroed) pads untouched
#pragma unroll
    for (int j = 0; j < 16; ++j) v[j] = 0.f;
  }
  store8<T>(view_addr<T>(g.out, c0 >> 3, f, n, h, w), v);
  store8<T>(view_addr<T>(g.out, (c0 >> 3) + 1, f, n, h, w), v + 8);
}

// ---------------------------------------------------------------------------------------------
// tcgen05 kernel
// ---------------------------------------------------------------------------------------------
constexpr int kSmemHeader = 256;  // barriers + TMEM slot

__device__ __forceinline__ long long gtime() {
  long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// slot 0: globaltimer (ns) at CTA start; slots 1..7: SM clock at each stage
#define DBG(slot)                                                                       \
  if (p.dbg && blockIdx.y == 0) {                                                       \
    if ((slot) == 0) p.dbg[(size_t)blockIdx.x * 8] = gtime();                           \
    p.dbg[(size_t)blockIdx.x * 8 + (slot) + 1] = clock64();                             \
  }

__global__ void __launch_bounds__(128) conv_umma_kernel(const __grid_constant__ ConvParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  const ConvGroupDev& g = p.g[blockIdx.y];
  uint64_t* bar_slab = reinterpret_cast<uint64_t*>(smem);
  uint64_t* bar_acc = bar_slab + 1;
  uint64_t* bar_wfull = bar_slab + 2;
  uint64_t* bar_wempty = bar_slab + 2 + kMaxRing;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 192);
  uint8_t* slab = smem + kSmemHeader;
  const uint32_t chunk_bytes = (uint32_t)p.SP * 16u;                     // one channel chunk of the slab
  const uint32_t slab_bytes = (uint32_t)(p.nsrc * p.nchunks) * chunk_bytes;
  uint8_t* wring = slab + ((slab_bytes + 127u) & ~127u);
  const uint32_t stage_bytes = (uint32_t)(p.tps * p.tap_bytes);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile_start = blockIdx.x * (p.MT * kTileM);

  if (threadIdx.x == 0) {
    DBG(0)
    ptx::mbar_init(bar_slab, 1);
    ptx::mbar_init(bar_acc, 1);
    for (int i = 0; i < p.NS; ++i) {
      ptx::mbar_init(bar_wfull + i, 1);
      ptx::mbar_init(bar_wempty + i, 1);
    }
    ptx::fence_mbar_init();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, (uint32_t)p.tmem_cols);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0) {
    // ---- producer: input slab, then the weight stages -------------------------------------
    DBG(1)
    ptx::mbar_arrive_expect_tx(bar_slab, slab_bytes);
    for (int s = 0; s < p.nsrc; ++s) {
      const char* src = g.src[s].base + (long long)(tile_start - p.halo_lo) * 16;
      for (int c = 0; c < p.nchunks; ++c)
        ptx::bulk_g2s(slab + (size_t)(s * p.nchunks + c) * chunk_bytes, src + (long long)c * g.src[s].plane_bytes,
                      chunk_bytes, bar_slab);
    }
    for (int st = 0; st < p.nstages; ++st) {
      const int slot = st % p.NS;
      if (st >= p.NS) ptx::mbar_wait(bar_wempty + slot, ((st / p.NS) - 1) & 1);
      int ntap = min(p.tps, p.ntaps - st * p.tps);
      uint32_t bytes = (uint32_t)(ntap * p.tap_bytes);
      ptx::mbar_arrive_expect_tx(bar_wfull + slot, bytes);
      ptx::bulk_g2s(wring + (size_t)slot * stage_bytes, g.w + (size_t)st * stage_bytes, bytes, bar_wfull + slot);
    }
  } else if (warp == 1 && lane == 0) {
    // ---- MMA issuer -------------------------------------------------------------------------
    const uint32_t idesc = ptx::umma_idesc_bf16(kTileM, p.Cout, 0, 0);
    const uint32_t slab_a = ptx::smem_u32(slab);
    const uint32_t wring_a = ptx::smem_u32(wring);
    const uint32_t a_lbo = p.swap_desc ? 128u : chunk_bytes;
    const uint32_t a_sbo = p.swap_desc ? chunk_bytes : 128u;
    const uint32_t b_kstride = (uint32_t)p.Cout * 16u;  // bytes between 8-channel chunks of the weights
    const uint32_t b_lbo = p.swap_desc ? 128u : b_kstride;
    const uint32_t b_sbo = p.swap_desc ? b_kstride : 128u;
    ptx::mbar_wait(bar_slab, 0);
    ptx::tc_fence_after();
    DBG(2)
    bool first = true;
    for (int st = 0; st < p.nstages; ++st) {
      const int slot = st % p.NS;
      ptx::mbar_wait(bar_wfull + slot, (st / p.NS) & 1);
      ptx::tc_fence_after();
      if (st == 0) DBG(3)
      const int ntap = min(p.tps, p.ntaps - st * p.tps);
      for (int ti = 0; ti < ntap; ++ti) {
        const ConvTapDev tp = p.tap[st * p.tps + ti];
        const uint32_t a_tap = slab_a + (uint32_t)(tp.src * p.nchunks) * chunk_bytes +
                               (uint32_t)(p.halo_lo + tp.off) * 16u;
        const uint32_t b_tap = wring_a + (uint32_t)slot * stage_bytes + (uint32_t)(ti * p.tap_bytes);
        for (int kc = 0; kc < p.nchunks; kc += 2) {
          const uint64_t bdesc = ptx::umma_desc(b_tap + (uint32_t)kc * b_kstride, b_lbo, b_sbo);
          for (int mt = 0; mt < p.MT; ++mt) {
            const uint64_t adesc =
                ptx::umma_desc(a_tap + (uint32_t)kc * chunk_bytes + (uint32_t)(mt * kTileM) * 16u, a_lbo, a_sbo);
            ptx::umma_bf16(tmem_base + (uint32_t)(mt * p.Cout), adesc, bdesc, idesc, first ? 0u : 1u);
          }
          first = false;
        }
      }
      ptx::umma_commit(bar_wempty + slot);  // stage may be refilled once these MMAs have read it
    }
    ptx::umma_commit(bar_acc);
    DBG(4)
  }
  __syncwarp();

  // ---- epilogue: warp w owns TMEM lanes [32w, 32w+32) ----------------------------------------
  ptx::mbar_wait(bar_acc, 0);
  ptx::tc_fence_after();
  if (threadIdx.x == 64) DBG(5)
  for (int mt = 0; mt < p.MT; ++mt) {
    const int f = tile_start + mt * kTileM + warp * 32 + lane;
    for (int c0 = 0; c0 < p.Cout; c0 += 16) {
      float v[16];
      ptx::tmem_ld16(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(mt * p.Cout + c0), v);
      epilogue16<__nv_bfloat16>(p, g, f, c0, v);
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (threadIdx.x == 64) DBG(6)
  if (warp == 1) ptx::tmem_dealloc(tmem_base, (uint32_t)p.tmem_cols);
}

// ---------------------------------------------------------------------------------------------
// FP32 path
// ---------------------------------------------------------------------------------------------
constexpr int kFp32CoutBlock = 32;

__global__ void __launch_bounds__(128) conv_fp32_kernel(const __grid_constant__ ConvParams p) {
  extern __shared__ __align__(16) float ws[];  // [K_tap][32]
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

cudaError_t launch_conv(const ConvParams& p, int precision, cudaStream_t stream) {
  if (precision == 0) {
    size_t smem = conv_umma_smem_bytes(p);
    static size_t configured = 0;
    if (smem > configured) {
      cudaError_t e = cudaFuncSetAttribute(conv_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) return e;
      configured = smem;
    }
    int tile = p.MT * kTileM;
    dim3 grid((unsigned)((p.body + tile - 1) / tile), p.ngroups, 1);
    conv_umma_kernel<<<grid, 128, smem, stream>>>(p);
  } else {
    size_t smem = (size_t)p.nchunks * 8 * kFp32CoutBlock * sizeof(float);
    dim3 grid((unsigned)((p.body + 127) / 128), p.ngroups, (p.Cout + kFp32CoutBlock - 1) / kFp32CoutBlock);
    conv_fp32_kernel<<<grid, 128, smem, stream>>>(p);
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

}  // namespace mtldas
