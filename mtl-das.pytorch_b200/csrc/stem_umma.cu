// Stem convolution (Conv2d(1,16,7x7,stride 3,pad 2), /root/reference/model/modelA_MTL.py:73-77) on tcgen05.
//
// Cin = 1, so the GEMM K dimension is the 49 filter taps (padded to 64).  For a tile of 128 consecutive flattened
// output positions the worker warps (a) stage the <= 3 x 7 input rows the tile touches into shared memory with
// coalesced 8-byte loads, (b) build the im2col A operand [8 K-chunks][128 rows][8] in bf16 straight from that
// staging (stride-3 word reads: bank-conflict free) and (c) drain the previous tile's accumulator (bias, ReLU, bf16
// pack, P8 store) while the MMA warp issues the four K=16 MMAs of the current tile.  A tiles and TMEM accumulators are
// double buffered.  HBM traffic = the input once + the output once.
#include <algorithm>
#include <cstdio>

#include "gran.cuh"
#include "kernels.h"
#include "ptx.cuh"

namespace mtldas {

constexpr int kStemThreads = 160;  // warps 0-3: workers (thread = tile row), warp 4: MMA issuer
constexpr int kStemXRows = 21, kStemXCols = 256;

__global__ void __launch_bounds__(kStemThreads) stem_umma_kernel(const __grid_constant__ StemParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);  // [0,1] a_full (128 arrivals), [2,3] acc_full, [4] w_full
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 64);
  float* bias_s = reinterpret_cast<float*>(smem + 128);
  uint8_t* wsm = smem + 256;                               // 2 KB weight image
  uint8_t* asm_ = smem + 256 + 2048;                        // A[2][8][128][16 B] = 32 KB
  float* xs = reinterpret_cast<float*>(asm_ + 2 * 16384);   // [3][7][256] fp32 input staging
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int ntiles = (int)((p.body + 127) / 128);
  const int Hp = p.Pimg / p.Wp;

  ptx::pdl_launch_dependents();
  if (threadIdx.x == 0) {
    ptx::mbar_init(bars + 0, 128), ptx::mbar_init(bars + 1, 128);
    ptx::mbar_init(bars + 2, 1), ptx::mbar_init(bars + 3, 1), ptx::mbar_init(bars + 4, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 4) ptx::tmem_alloc(tmem_slot, 32);
  if (threadIdx.x < 16) bias_s[threadIdx.x] = p.bias[threadIdx.x];
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 4) {
    if (ptx::elect_one()) {
      ptx::mbar_arrive_expect_tx(bars + 4, 2048);
      ptx::bulk_g2s(wsm, p.w_umma, 2048, bars + 4);
      ptx::mbar_wait(bars + 4, 0);
      const uint32_t idesc = ptx::umma_idesc_bf16(128, 16, 0, 0);
      const uint64_t a_hi = ptx::umma_desc(0, 2048u, 128u);  // K chunks 2048 B apart, 8-row groups 128 B apart
      const uint64_t b_hi = ptx::umma_desc(0, 256u, 128u);
      const uint32_t a16 = ptx::smem_u32(asm_) >> 4, w16 = ptx::smem_u32(wsm) >> 4;
      int it = 0;
      for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
        const int b = it & 1;
        ptx::mbar_wait(bars + b, (it >> 1) & 1);
        ptx::tc_fence_after();
#pragma unroll
        for (int k = 0; k < 4; ++k)
          ptx::umma_bf16(tmem_base + (uint32_t)(b * 16), a_hi | (a16 + (uint32_t)b * 1024u + (uint32_t)k * 256u),
                         b_hi | (w16 + (uint32_t)k * 32u), idesc, k ? 1u : 0u);
        ptx::umma_commit(bars + 2 + b);
      }
    }
    __syncwarp();
  } else {
    const int r = threadIdx.x;  // row of the tile
    ptx::pdl_grid_wait();
    auto epilogue = [&](int tile, int it) {
      const int b = it & 1;
      ptx::mbar_wait(bars + 2 + b, (it >> 1) & 1);
      ptx::tc_fence_after();
      uint32_t q[16];
      ptx::tmem_ld16_issue(tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)(b * 16), q);
      ptx::tmem_wait16(q);
      ptx::tc_fence_before();
      const long long f = (long long)tile * 128 + r;
      if (f >= p.body) return;
      const int n = (int)(f / p.Pimg);
      const int rem = (int)(f - (long long)n * p.Pimg);
      const int h = rem / p.Wp, w = rem - h * p.Wp;
      const bool valid = h < p.H && w < p.W;
      float v[16];
      const float lo = p.raw ? -INFINITY : 0.f;
#pragma unroll
      for (int j = 0; j < 16; ++j) v[j] = valid ? fmaxf(__uint_as_float(q[j]) + bias_s[j], lo) : 0.f;
      store8<__nv_bfloat16>(view_addr<__nv_bfloat16>(p.out, 0, (int)f, n, h, w), v);
      store8<__nv_bfloat16>(view_addr<__nv_bfloat16>(p.out, 1, (int)f, n, h, w), v + 8);
    };
    int it = 0, prev_tile = -1;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      const int b = it & 1;
      const long long f0 = (long long)tile * 128;
      const int g0 = (int)(f0 / p.Wp);  // first flattened image row (n*Hp + h) touched by the tile
      asm volatile("bar.sync 1, 128;" ::: "memory");  // previous tile's reads of the staging are done
      // ---- (a) stage input rows: slot (j, kr) = input row 3h-2+kr of flattened output row g0+j, columns ix = c-2
      // one warp per staged row (row-uniform index math), lanes sweep the 128 column pairs
      for (int row = warp; row < kStemXRows; row += 4) {
        const int j = row / 7, kr = row - j * 7;
        const int g = g0 + j, n = g / Hp, h = g - n * Hp;
        const int iy = 3 * h - 2 + kr;
        const bool row_ok = h < p.H && (long long)g * p.Wp < p.body && iy >= 0 && iy < p.Hin;
        const float* src = p.x + ((size_t)n * p.Hin + (row_ok ? iy : 0)) * p.Win;
        float* dst = xs + row * kStemXCols;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const int cp = lane + 32 * q, ix = 2 * cp - 2;
          float2 v = make_float2(0.f, 0.f);
          if (row_ok && ix >= 0 && ix + 1 < p.Win) v = *reinterpret_cast<const float2*>(src + ix);
          *reinterpret_cast<float2*>(dst + 2 * cp) = v;
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");
      // ---- (b) im2col row r of A[b]
      {
        const long long f = f0 + r;
        const int g = (int)(f / p.Wp), w = (int)(f - (long long)g * p.Wp);
        const int n = g / Hp, h = g - n * Hp;
        const bool valid = f < p.body && h < p.H && w < p.W;
        const float* xr = xs + (g - g0) * 7 * kStemXCols + 3 * w;  // tap (kr, ks) -> xr[kr*256 + ks]
        uint8_t* arow = asm_ + (size_t)b * 16384 + (size_t)r * 16;
#pragma unroll
        for (int c = 0; c < 8; ++c) {
          float v[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const int k = c * 8 + j;  // compile-time after unrolling
            v[j] = (valid && k < 49) ? xr[(k / 7) * kStemXCols + (k % 7)] : 0.f;
          }
          store8<__nv_bfloat16>(reinterpret_cast<char*>(arow) + (size_t)c * 2048, v);
        }
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> tensor-core reads
      ptx::mbar_arrive(bars + b);
      // ---- (c) drain the previous tile while the MMA warp works on this one
      if (prev_tile >= 0) epilogue(prev_tile, it - 1);
      prev_tile = tile;
    }
    if (prev_tile >= 0) epilogue(prev_tile, it - 1);
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 4) ptx::tmem_dealloc(tmem_base, 32);
}

cudaError_t launch_stem_umma(const StemParams& p, cudaStream_t stream, bool pdl) {
  const size_t smem = 256 + 2048 + 2 * 16384 + (size_t)kStemXRows * kStemXCols * sizeof(float);
  static bool configured = false;
  if (!configured) {
    cudaError_t e = cudaFuncSetAttribute(stem_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    configured = true;
  }
  const int ntiles = (int)((p.body + 127) / 128);
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)std::min(ntiles, 148 * 3)), cfg.blockDim = dim3(kStemThreads);
  cfg.dynamicSmemBytes = smem, cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr, cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, stem_umma_kernel, p);
}

}  // namespace mtldas
