// Model C (Multi_Classifier = Inception-V3 with a 1-channel stem and 32 classes), eval-mode forward, for sm_100a.
//
// Reference: /root/reference/model/modelC_multiClassifier.py:63-86 (layers), :102-147 (_forward) and torchvision's
// InceptionA..E blocks (torchvision/models/inception.py:176-365) that it instantiates: 94 conv + BatchNorm(eps 1e-3) +
// ReLU units with kernels 1x1, 3x3 (pad 0 / 1, stride 1 / 2), 5x5 p2, 1x7, 7x1, 1x3, 3x1; F.avg_pool2d(3,1,1) (count
// includes the padding), F.max_pool2d(3,2) / nn.MaxPool2d(3,2); channel concatenation; AdaptiveAvgPool2d(1) + Dropout
// (identity in eval) + Linear(2048, 32).
//
// Layout "planar-8": a tensor [n, C, H, W] is C/8 planes of n*H*W positions, one position of one plane = 8 consecutive
// channels as bf16 (16 bytes).  A channel concat is a chunk offset into the destination tensor.
//
// convc_umma_kernel: implicit GEMM on tcgen05.  One CTA = 128 output positions x Nt <= 256 output channels.  The K loop
// runs over (tap, 32-channel unit); four gather warps build the A tile of a unit with 16-byte cp.async copies (zero fill
// outside the image: any kernel size / stride / padding is index arithmetic, no im2col tensor, no tensor maps), one
// producer thread brings the unit's weight slice with bulk copies, one thread issues the UMMAs into a TMEM accumulator;
// a ring of stages overlaps the three.  The epilogue (folded-BN bias + ReLU, bf16, 16-byte stores at the concat offset)
// is done by the gather warps.  This model is compute bound (1.1 GMAC per patch, Cout up to 448).
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>

#include "../../include/mtldas.h"
#include "kernels.h"
#include "plan.h"
#include "ptx.cuh"

namespace mtldas {
namespace {

constexpr int kCStages = 4;       // ring depth
constexpr int kCUnitSteps = 2;    // K steps (16 channels each) per ring stage
constexpr int kCThreads = 192;    // warps 0-3 gather + epilogue, warp 4 MMA issue, warp 5 weight producer
constexpr int kCHeader = 1024;    // mbarriers, TMEM slot, then bias at kCHeader

struct ConvCParams {
  const char* x;        // planar-8 bf16 input, cin/8 planes of m_in positions
  const char* w;        // packed weights [tap][cin/8][cout][8] bf16 (BatchNorm folded)
  const float* bias;    // [cout] folded BatchNorm shift
  char* out;            // planar-8 bf16 destination tensor (out_planes planes of m_out positions)
  long long m_in, m_out;
  int n, cin, H, W, Ho, Wo, cout, kh, kw, sh, sw, ph, pw;
  int out_chunk0;       // first destination plane (channel offset / 8 inside the concat tensor)
  int Nt;               // output channels per CTA (multiple of 16, <= 256)
  int relu;
};

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__global__ void __launch_bounds__(kCThreads, 2) convc_umma_kernel(const __grid_constant__ ConvCParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);              // [kCStages]
  uint64_t* empty = full + kCStages;                                // [kCStages]
  uint64_t* acc_full = empty + kCStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 128);
  float* bias_s = reinterpret_cast<float*>(smem + 256);             // Nt floats (<= 1 KB -> header is 256 + 1024)
  const int Nt = p.Nt;
  const uint32_t a_stage = 2u * kCUnitSteps * 128u * 16u;           // 8 KB: 2*steps chunks x 128 rows x 16 B
  const uint32_t b_stage = 2u * kCUnitSteps * (uint32_t)Nt * 16u;
  uint8_t* a_base = smem + 256 + 1024;
  uint8_t* b_base = a_base + (size_t)kCStages * a_stage;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.y * Nt;
  const int KC = p.cin >> 4;                                        // K steps per tap
  const int units_per_tap = (KC + kCUnitSteps - 1) / kCUnitSteps;
  const int ntaps = p.kh * p.kw;
  const int nunits = ntaps * units_per_tap;
  uint32_t cols = 32;
  while (cols < (uint32_t)Nt) cols <<= 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kCStages; ++s) {
      ptx::mbar_init(full + s, 128 + 1);
      ptx::mbar_init(empty + s, 1);
    }
    ptx::mbar_init(acc_full, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 4) ptx::tmem_alloc(tmem_slot, cols);
  for (int i = threadIdx.x; i < Nt; i += blockDim.x) bias_s[i] = p.bias[n0 + i];
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 4) {
    // ---------------- gather: thread = one output position (one row of the A tile) ----------------
    const long long m = (long long)blockIdx.x * 128 + threadIdx.x;
    const bool row_ok = m < p.m_out;
    int n = 0, ho = 0, wo = 0;
    if (row_ok) {
      const int hw = p.Ho * p.Wo;
      n = (int)(m / hw);
      const int rem = (int)(m - (long long)n * hw);
      ho = rem / p.Wo;
      wo = rem - ho * p.Wo;
    }
    const uint32_t a_row = ptx::smem_u32(a_base) + (uint32_t)threadIdx.x * 16u;
    int u = 0;
    for (int tap = 0; tap < ntaps; ++tap) {
      const int r = tap / p.kw, q = tap - r * p.kw;
      const int hi = ho * p.sh + r - p.ph, wi = wo * p.sw + q - p.pw;
      const bool inb = row_ok && hi >= 0 && hi < p.H && wi >= 0 && wi < p.W;
      const long long src_pos = inb ? ((long long)n * p.H + hi) * p.W + wi : 0;
      const char* src0 = p.x + src_pos * 16;
      const uint32_t nbytes = inb ? 16u : 0u;
      for (int kc0 = 0; kc0 < KC; kc0 += kCUnitSteps, ++u) {
        const int s = u % kCStages;
        if (u >= kCStages) ptx::mbar_wait(empty + s, ((u / kCStages) - 1) & 1);
        const int nch = 2 * min(kCUnitSteps, KC - kc0);
        const uint32_t dst = a_row + (uint32_t)s * a_stage;
#pragma unroll
        for (int c = 0; c < 2 * kCUnitSteps; ++c)
          if (c < nch) cp_async16(dst + (uint32_t)c * 2048u, src0 + (long long)(kc0 * 2 + c) * p.m_in * 16, nbytes);
        cp_async_commit();
        if (u >= 2) {                       // unit u-2 has landed: publish it to the MMA thread
          cp_async_wait<2>();
          fence_proxy_async();
          ptx::mbar_arrive(full + (u - 2) % kCStages);
        }
      }
    }
    if (nunits >= 2) {
      cp_async_wait<1>();
      fence_proxy_async();
      ptx::mbar_arrive(full + (nunits - 2) % kCStages);
    }
    cp_async_wait<0>();
    fence_proxy_async();
    ptx::mbar_arrive(full + (nunits - 1) % kCStages);

    // ---------------- epilogue: bias + ReLU, bf16, one 16-byte granule per 8 channels ----------------
    ptx::mbar_wait(acc_full, 0);
    ptx::tc_fence_after();
    const uint32_t trow = tmem_base + ((uint32_t)(warp * 32) << 16);
    char* out_row = p.out + ((long long)(p.out_chunk0 + (n0 >> 3)) * p.m_out + m) * 16;
    for (int c0 = 0; c0 < Nt; c0 += 16) {
      uint32_t r[16];
      ptx::tmem_ld16_issue(trow + (uint32_t)c0, r);
      ptx::tmem_wait16(r);
      if (row_ok) {
        float v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          v[j] = __uint_as_float(r[j]) + bias_s[c0 + j];
          if (p.relu) v[j] = fmaxf(v[j], 0.f);
        }
        uint4 q0, q1;
        __nv_bfloat162* h0 = reinterpret_cast<__nv_bfloat162*>(&q0);
        __nv_bfloat162* h1 = reinterpret_cast<__nv_bfloat162*>(&q1);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          h0[j] = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
          h1[j] = __floats2bfloat162_rn(v[8 + 2 * j], v[8 + 2 * j + 1]);
        }
        *reinterpret_cast<uint4*>(out_row + (long long)(c0 >> 3) * p.m_out * 16) = q0;
        *reinterpret_cast<uint4*>(out_row + (long long)((c0 >> 3) + 1) * p.m_out * 16) = q1;
      }
    }
  } else if (warp == 4) {
    // ---------------- MMA issue ----------------
    if (ptx::elect_one()) {
      const uint32_t idesc = ptx::umma_idesc_bf16(128, Nt, 0, 0);
      const uint64_t a_hi = ptx::umma_desc(0, 2048u, 128u);
      const uint64_t b_hi = ptx::umma_desc(0, (uint32_t)Nt * 16u, 128u);
      const uint32_t a16 = ptx::smem_u32(a_base) >> 4, b16 = ptx::smem_u32(b_base) >> 4;
      uint32_t acc = 0u;
      int u = 0;
      for (int tap = 0; tap < ntaps; ++tap)
        for (int kc0 = 0; kc0 < KC; kc0 += kCUnitSteps, ++u) {
          const int s = u % kCStages;
          ptx::mbar_wait(full + s, (u / kCStages) & 1);
          ptx::tc_fence_after();
          const int nk = min(kCUnitSteps, KC - kc0);
          for (int j = 0; j < nk; ++j) {
            const uint32_t a_lo = a16 + (uint32_t)s * (a_stage >> 4) + (uint32_t)j * (2u * 2048u >> 4);
            const uint32_t b_lo = b16 + (uint32_t)s * (b_stage >> 4) + (uint32_t)j * (2u * (uint32_t)Nt);
            ptx::umma_bf16_lean(tmem_base, a_lo | (uint32_t)a_hi, (uint32_t)(a_hi >> 32), b_lo | (uint32_t)b_hi,
                                (uint32_t)(b_hi >> 32), idesc, acc);
            acc = 1u;
          }
          ptx::umma_commit(empty + s);
        }
      ptx::umma_commit(acc_full);
    }
    __syncwarp();
  } else {
    // ---------------- weight producer: the unit's [2*steps chunks][Nt][8] slice, one bulk copy per chunk ----------------
    if (ptx::elect_one()) {
      const int cin_chunks = p.cin >> 3;
      int u = 0;
      for (int tap = 0; tap < ntaps; ++tap)
        for (int kc0 = 0; kc0 < KC; kc0 += kCUnitSteps, ++u) {
          const int s = u % kCStages;
          if (u >= kCStages) ptx::mbar_wait(empty + s, ((u / kCStages) - 1) & 1);
          const int nch = 2 * min(kCUnitSteps, KC - kc0);
          ptx::mbar_arrive_expect_tx(full + s, (uint32_t)(nch * Nt * 16));
          for (int c = 0; c < nch; ++c)
            ptx::bulk_g2s(b_base + (size_t)s * b_stage + (size_t)c * Nt * 16,
                          p.w + (((long long)tap * cin_chunks + kc0 * 2 + c) * p.cout + n0) * 16, (uint32_t)(Nt * 16), full + s);
        }
    }
    __syncwarp();
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 4) ptx::tmem_dealloc(tmem_base, cols);
}

// ---------------------------------------------------------------------------------------------
// Conv2d_1a_3x3 (modelC:63): 1 -> 32 channels, 3x3, stride 2, no padding, straight from the fp32 patch (K = 9: CUDA cores)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) convc_first_kernel(const float* __restrict__ x, const float* __restrict__ w9,
                                                          const float* __restrict__ bias, char* __restrict__ out, int n,
                                                          int H, int W, int Ho, int Wo, int cout) {
  __shared__ float ws[9 * 64], bs[64];
  for (int i = threadIdx.x; i < 9 * cout; i += 256) ws[i] = w9[i];       // [tap][cout]
  for (int i = threadIdx.x; i < cout; i += 256) bs[i] = bias[i];
  __syncthreads();
  const long long M = (long long)n * Ho * Wo;
  for (long long m = (long long)blockIdx.x * 256 + threadIdx.x; m < M; m += (long long)gridDim.x * 256) {
    const int img = (int)(m / (Ho * Wo));
    const int rem = (int)(m - (long long)img * Ho * Wo);
    const int ho = rem / Wo, wo = rem - ho * Wo;
    const float* src = x + ((long long)img * H + 2 * ho) * W + 2 * wo;
    float v[9];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int q = 0; q < 3; ++q) v[r * 3 + q] = src[r * W + q];
    for (int c0 = 0; c0 < cout; c0 += 8) {
      float acc[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] = bs[c0 + j];
#pragma unroll
      for (int t = 0; t < 9; ++t)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] = fmaf(v[t], ws[t * cout + c0 + j], acc[j]);
      uint4 q4;
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&q4);
#pragma unroll
      for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(fmaxf(acc[2 * j], 0.f), fmaxf(acc[2 * j + 1], 0.f));
      *reinterpret_cast<uint4*>(out + ((long long)(c0 >> 3) * M + m) * 16) = q4;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// 3x3 pools: max (no padding, stride 2: nn.MaxPool2d(3,2) / F.max_pool2d(x,3,2)) and average (stride 1, pad 1, the
// padding counts: F.avg_pool2d(x, 3, 1, 1) divides by 9 everywhere)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) poolc_kernel(const char* __restrict__ x, char* __restrict__ out, int n, int chunks,
                                                    int H, int W, int Ho, int Wo, int stride, int pad, int is_avg,
                                                    int out_chunk0) {
  // grid = (position blocks, channel chunks): 32-bit index arithmetic only (64-bit divisions dominated the first version)
  const unsigned Mo = (unsigned)n * Ho * Wo, hw = (unsigned)Ho * Wo;
  const long long Mi = (long long)n * H * W;
  const int c = blockIdx.y;
  const char* xin = x + (long long)c * Mi * 16;
  char* dst = out + (long long)(out_chunk0 + c) * Mo * 16;
  for (unsigned m = blockIdx.x * 256u + threadIdx.x; m < Mo; m += gridDim.x * 256u) {
    const unsigned img = m / hw, rem = m - img * hw;
    const int ho = (int)(rem / (unsigned)Wo), wo = (int)(rem - (unsigned)ho * Wo);
    const int h0 = ho * stride - pad, w0 = wo * stride - pad;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = is_avg ? 0.f : -INFINITY;
    const char* base = xin + ((long long)img * H * W) * 16;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int hi = h0 + r;
      if (hi < 0 || hi >= H) continue;
#pragma unroll
      for (int q = 0; q < 3; ++q) {
        const int wi = w0 + q;
        if (wi < 0 || wi >= W) continue;
        const uint4 g = *reinterpret_cast<const uint4*>(base + (hi * W + wi) * 16);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&g);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 f = __bfloat1622float2(h[j]);
          if (is_avg) acc[2 * j] += f.x, acc[2 * j + 1] += f.y;
          else acc[2 * j] = fmaxf(acc[2 * j], f.x), acc[2 * j + 1] = fmaxf(acc[2 * j + 1], f.y);
        }
      }
    }
    uint4 o;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&o);
    const float sc = is_avg ? (1.f / 9.f) : 1.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(acc[2 * j] * sc, acc[2 * j + 1] * sc);
    *reinterpret_cast<uint4*>(dst + (long long)m * 16) = o;
  }
}

// ---------------------------------------------------------------------------------------------
// AdaptiveAvgPool2d((1,1)) + flatten + Linear(C, classes)   (modelC:137-146; Dropout is the identity in eval mode)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) headc_kernel(const char* __restrict__ x, int C, int hw, long long M,
                                                    const float* __restrict__ fw, const float* __restrict__ fb, int classes,
                                                    float* __restrict__ logits) {
  extern __shared__ float gap[];                     // [C]
  const int img = blockIdx.x;
  for (int c = threadIdx.x; c < C / 8; c += 256) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int q = 0; q < hw; ++q) {
      const uint4 g = *reinterpret_cast<const uint4*>(x + ((long long)c * M + (long long)img * hw + q) * 16);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&g);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = __bfloat1622float2(h[j]);
        acc[2 * j] += f.x, acc[2 * j + 1] += f.y;
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) gap[c * 8 + j] = acc[j] / (float)hw;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int k = warp; k < classes; k += 8) {
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s = fmaf(gap[c], fw[(size_t)k * C + c], s);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) logits[(size_t)img * classes + k] = s + fb[k];
  }
}

// ---------------------------------------------------------------------------------------------
// weight packing with the eval-mode BatchNorm fold (eps 1e-3): w' = w*g/sqrt(rv+eps), b' = beta - rm*g/sqrt(rv+eps)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) packc_kernel(const float* __restrict__ w, const float* __restrict__ gamma,
                                                    const float* __restrict__ beta, const float* __restrict__ rmean,
                                                    const float* __restrict__ rvar, float eps, int cout, int cin, int taps,
                                                    __nv_bfloat16* __restrict__ dst, float* __restrict__ bias,
                                                    float* __restrict__ first_w) {
  const int total = cout * cin * taps;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < total; i += gridDim.x * 256) {
    const int co = i / (cin * taps), r1 = i - co * cin * taps, ci = r1 / taps, t = r1 - ci * taps;
    const float sc = gamma[co] * rsqrtf(rvar[co] + eps);
    if (r1 == 0) bias[co] = beta[co] - rmean[co] * sc;
    const float v = w[i] * sc;
    if (first_w) first_w[t * cout + co] = v;                            // [tap][cout] fp32 (cin == 1)
    else dst[(((size_t)t * (cin >> 3) + (ci >> 3)) * cout + co) * 8 + (ci & 7)] = __float2bfloat16(v);
  }
}

__global__ void __launch_bounds__(256) readc_kernel(const char* __restrict__ x, long long M, int hw, int C, int c0,
                                                    float* __restrict__ nchw) {
  const long long total = M * (C / 8);
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int c = (int)(i / M);
    const long long m = i - (long long)c * M;
    const long long img = m / hw, q = m - img * hw;
    const uint4 g = *reinterpret_cast<const uint4*>(x + ((long long)(c0 / 8 + c) * M + m) * 16);
    const __nv_bfloat16* h = reinterpret_cast<const __nv_bfloat16*>(&g);
    for (int j = 0; j < 8; ++j) nchw[((img * C) + c * 8 + j) * hw + q] = __bfloat162float(h[j]);
  }
}

int pick_nt(int cout) {   // output channels per CTA: the largest divisor of cout that is a multiple of 16 and <= 256
  for (int parts = 1; parts <= 8; ++parts)
    if (cout % parts == 0 && (cout / parts) % 16 == 0 && cout / parts <= 256) return cout / parts;
  return 0;
}

}  // namespace
}  // namespace mtldas

using namespace mtldas;

#define CKC(call, what)                                                     \
  do {                                                                      \
    cudaError_t e_ = (call);                                                \
    if (e_ != cudaSuccess) {                                                \
      set_error("%s: %s", what, cudaGetErrorString(e_));                    \
      return MTLDAS_ERR_CUDA;                                               \
    }                                                                       \
  } while (0)

extern "C" int mtldas_c_pack_conv(const float* dev_w, const float* dev_gamma, const float* dev_beta, const float* dev_rmean,
                                  const float* dev_rvar, float eps, int cout, int cin, int kh, int kw, void* dev_packed,
                                  float* dev_bias, void* stream) {
  if (!dev_w || !dev_gamma || !dev_beta || !dev_rmean || !dev_rvar || !dev_packed || !dev_bias || cout < 1 || kh < 1 || kw < 1 ||
      !(cin == 1 || cin % 16 == 0)) {
    set_error("c_pack_conv: bad arguments (cin must be 1 or a multiple of 16)");
    return MTLDAS_ERR_ARG;
  }
  const int total = cout * cin * kh * kw;
  packc_kernel<<<std::min(592, (total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      dev_w, dev_gamma, dev_beta, dev_rmean, dev_rvar, eps, cout, cin, kh * kw, cin == 1 ? nullptr : (__nv_bfloat16*)dev_packed,
      dev_bias, cin == 1 ? (float*)dev_packed : nullptr);
  CKC(cudaGetLastError(), "c_pack_conv");
  return 0;
}

extern "C" int mtldas_c_conv(const void* dev_x, int n, int cin, int h, int w, const void* dev_packed, const float* dev_bias,
                             void* dev_out, int out_channels_total, int out_channel0, int cout, int kh, int kw, int sh, int sw,
                             int ph, int pw, int relu, void* stream) {
  const int Ho = (h + 2 * ph - kh) / sh + 1, Wo = (w + 2 * pw - kw) / sw + 1;
  if (!dev_x || !dev_packed || !dev_bias || !dev_out || n < 1 || Ho < 1 || Wo < 1 || (out_channel0 % 8) || (out_channels_total % 8) ||
      out_channel0 + cout > out_channels_total) {
    set_error("c_conv: bad arguments");
    return MTLDAS_ERR_ARG;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const long long m_out = (long long)n * Ho * Wo;
  if (cin == 1) {   // Conv2d_1a_3x3: fp32 patches in, CUDA cores
    if (kh != 3 || kw != 3 || sh != 2 || sw != 2 || ph || pw || cout > 64 || (cout % 8) || out_channel0) {
      set_error("c_conv: the 1-channel entry convolution is 3x3 / stride 2 / no padding (modelC:63)");
      return MTLDAS_ERR_UNSUPPORTED;
    }
    const int blocks = (int)std::min<long long>(148 * 8, (m_out + 255) / 256);
    convc_first_kernel<<<blocks, 256, 0, st>>>((const float*)dev_x, (const float*)dev_packed, dev_bias, (char*)dev_out, n, h, w,
                                               Ho, Wo, cout);
    CKC(cudaGetLastError(), "c_conv (first)");
    return 0;
  }
  const int Nt = pick_nt(cout);
  if ((cin % 16) || !Nt) {
    set_error("c_conv: cin %d / cout %d not supported (multiples of 16)", cin, cout);
    return MTLDAS_ERR_UNSUPPORTED;
  }
  ConvCParams p{};
  p.x = (const char*)dev_x, p.w = (const char*)dev_packed, p.bias = dev_bias, p.out = (char*)dev_out;
  p.m_in = (long long)n * h * w, p.m_out = m_out;
  p.n = n, p.cin = cin, p.H = h, p.W = w, p.Ho = Ho, p.Wo = Wo, p.cout = cout, p.kh = kh, p.kw = kw, p.sh = sh, p.sw = sw;
  p.ph = ph, p.pw = pw, p.out_chunk0 = out_channel0 / 8, p.Nt = Nt, p.relu = relu;
  const size_t smem = 256 + 1024 + (size_t)kCStages * (2 * kCUnitSteps * 128 * 16 + 2 * kCUnitSteps * Nt * 16);
  static PerDeviceSmem configured;
  CKC(configured.ensure(smem, [](int bytes) {
        return cudaFuncSetAttribute(convc_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
      }), "c_conv (attribute)");
  dim3 grid((unsigned)((m_out + 127) / 128), (unsigned)(cout / Nt));
  convc_umma_kernel<<<grid, kCThreads, smem, st>>>(p);
  CKC(cudaGetLastError(), "c_conv");
  return 0;
}

extern "C" int mtldas_c_pool(const void* dev_x, int n, int channels, int h, int w, void* dev_out, int out_channels_total,
                             int out_channel0, int stride, int pad, int is_avg, void* stream) {
  const int Ho = (h + 2 * pad - 3) / stride + 1, Wo = (w + 2 * pad - 3) / stride + 1;
  if (!dev_x || !dev_out || n < 1 || (channels % 8) || (out_channel0 % 8) || out_channel0 + channels > out_channels_total ||
      Ho < 1 || Wo < 1) {
    set_error("c_pool: bad arguments");
    return MTLDAS_ERR_ARG;
  }
  const long long mo = (long long)n * Ho * Wo;
  if (mo >= (1ll << 31) || (long long)n * h * w >= (1ll << 27)) {
    set_error("c_pool: batch too large for the 32-bit index arithmetic of this kernel");
    return MTLDAS_ERR_UNSUPPORTED;
  }
  const int chunks = channels / 8;
  const unsigned bx = (unsigned)std::max<long long>(1, std::min<long long>((mo + 255) / 256, (148 * 16 + chunks - 1) / chunks));
  poolc_kernel<<<dim3(bx, (unsigned)chunks), 256, 0, (cudaStream_t)stream>>>(
      (const char*)dev_x, (char*)dev_out, n, channels / 8, h, w, Ho, Wo, stride, pad, is_avg, out_channel0 / 8);
  CKC(cudaGetLastError(), "c_pool");
  return 0;
}

extern "C" int mtldas_c_head(const void* dev_x, int n, int channels, int hw, const float* dev_fc_w, const float* dev_fc_b,
                             int classes, float* dev_logits, void* stream) {
  if (!dev_x || !dev_fc_w || !dev_fc_b || !dev_logits || n < 1 || (channels % 8) || hw < 1 || classes < 1 || channels > 8192) {
    set_error("c_head: bad arguments");
    return MTLDAS_ERR_ARG;
  }
  headc_kernel<<<n, 256, (size_t)channels * sizeof(float), (cudaStream_t)stream>>>((const char*)dev_x, channels, hw,
                                                                                  (long long)n * hw, dev_fc_w, dev_fc_b,
                                                                                  classes, dev_logits);
  CKC(cudaGetLastError(), "c_head");
  return 0;
}

extern "C" int mtldas_c_read(const void* dev_x, int n, int channels_total, int channel0, int channels, int hw,
                             float* dev_out_nchw, void* stream) {
  if (!dev_x || !dev_out_nchw || (channel0 % 8) || (channels % 8) || channel0 + channels > channels_total) {
    set_error("c_read: bad arguments");
    return MTLDAS_ERR_ARG;
  }
  const long long M = (long long)n * hw, total = M * (channels / 8);
  readc_kernel<<<(int)std::min<long long>(148 * 16, (total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      (const char*)dev_x, M, hw, channels, channel0, dev_out_nchw);
  CKC(cudaGetLastError(), "c_read");
  return 0;
}
