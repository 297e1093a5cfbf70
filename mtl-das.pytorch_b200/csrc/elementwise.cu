// Streaming (HBM-bound) kernels of the MTL-DAS hot path: stem convolution (Cin = 1), 2x2 ceil max-pool,
// classification heads, weight packing with eval-BatchNorm folding, layout conversion.
// All global accesses are 16/32-byte granules, consecutive threads -> consecutive granules.
#include <climits>
#include <cstdio>

#include "gran.cuh"
#include "kernels.h"

namespace mtldas {

// ---------------------------------------------------------------------------------------------
// stem: Conv2d(1,16,7x7,stride 3,pad 2,bias=False)+BN(eval, folded)+ReLU -- modelA_MTL.py:73-77,128
// thread = one flattened output position, 16 output channels in registers.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(128) stem_kernel(const __grid_constant__ StemParams p) {
  __shared__ float ws[49 * 16];
  __shared__ float bs[16];
  for (int i = threadIdx.x; i < 49 * 16; i += 128) ws[i] = p.w[i];
  if (threadIdx.x < 16) bs[threadIdx.x] = p.bias[threadIdx.x];
  __syncthreads();
  const int f = blockIdx.x * 128 + threadIdx.x;
  if (f >= p.body) return;
  const int n = f / p.Pimg;
  const int rem = f - n * p.Pimg;
  const int h = rem / p.Wp, w = rem - (rem / p.Wp) * p.Wp;
  float acc[16];
#pragma unroll
  for (int j = 0; j < 16; ++j) acc[j] = 0.f;
  if (h < p.H && w < p.W) {
    const float* xin = p.x + (size_t)n * p.Hin * p.Win;
#pragma unroll 1
    for (int r = 0; r < 7; ++r) {
      const int iy = 3 * h - 2 + r;
      if (iy < 0 || iy >= p.Hin) continue;
#pragma unroll
      for (int s = 0; s < 7; ++s) {
        const int ix = 3 * w - 2 + s;
        if (ix < 0 || ix >= p.Win) continue;
        const float xv = __ldg(xin + iy * p.Win + ix);
        const float4* wr = reinterpret_cast<const float4*>(ws + (r * 7 + s) * 16);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          float4 w4 = wr[q];
          acc[4 * q + 0] = fmaf(xv, w4.x, acc[4 * q + 0]);
          acc[4 * q + 1] = fmaf(xv, w4.y, acc[4 * q + 1]);
          acc[4 * q + 2] = fmaf(xv, w4.z, acc[4 * q + 2]);
          acc[4 * q + 3] = fmaf(xv, w4.w, acc[4 * q + 3]);
        }
      }
    }
#pragma unroll
    for (int j = 0; j < 16; ++j) acc[j] = fmaxf(acc[j] + bs[j], 0.f);
  }
  store8<T>(view_addr<T>(p.out, 0, f, n, h, w), acc);
  store8<T>(view_addr<T>(p.out, 1, f, n, h, w), acc + 8);
}

cudaError_t launch_stem(const StemParams& p, int precision, cudaStream_t stream) {
  unsigned grid = (unsigned)((p.body + 127) / 128);
  if (precision == 0)
    stem_kernel<__nv_bfloat16><<<grid, 128, 0, stream>>>(p);
  else
    stem_kernel<float><<<grid, 128, 0, stream>>>(p);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// MaxPool2d(kernel 2, stride 2, ceil_mode=True) -- modelA_MTL.py:116,145,152,158
// thread = one output granule (position x 8 channels); windows clipped at the image border.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) pool_kernel(const __grid_constant__ PoolParams p) {
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  const long long total = p.body_o * p.nchunks;
  if (idx >= total) return;
  const int c = (int)(idx / p.body_o);
  const int fo = (int)(idx - (long long)c * p.body_o);
  const int n = fo / p.Pimgo;
  const int rem = fo - n * p.Pimgo;
  const int ho = rem / p.Wpo, wo = rem - (rem / p.Wpo) * p.Wpo;
  float m[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) m[j] = 0.f;
  if (ho < p.Ho && wo < p.Wo) {
#pragma unroll
    for (int j = 0; j < 8; ++j) m[j] = -INFINITY;
#pragma unroll
    for (int dy = 0; dy < 2; ++dy) {
      const int h = 2 * ho + dy;
      if (h >= p.H) continue;
#pragma unroll
      for (int dx = 0; dx < 2; ++dx) {
        const int w = 2 * wo + dx;
        if (w >= p.W) continue;
        float v[8];
        load8<T>(view_addr<T>(p.in, c, n * p.Pimg + h * p.Wp + w, n, h, w), v);
#pragma unroll
        for (int j = 0; j < 8; ++j) m[j] = fmaxf(m[j], v[j]);
      }
    }
  }
  store8<T>(view_addr<T>(p.out, c, fo, n, ho, wo), m);
}

cudaError_t launch_pool(const PoolParams& p, int precision, cudaStream_t stream) {
  long long total = p.body_o * p.nchunks;
  unsigned grid = (unsigned)((total + 255) / 256);
  if (precision == 0)
    pool_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>(p);
  else
    pool_kernel<float><<<grid, 256, 0, stream>>>(p);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// heads: AdaptiveAvgPool2d(1,1) -> AvgPool1d(k = 128/classes) -> log_softmax -- modelA_MTL.py:165-172
// one CTA (128 threads) per sample; the padded 6x12 image block is summed whole (pads are zero).
// thread t: chunk t%16, positions t/16, t/16+8, ...
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(128) head_kernel(const __grid_constant__ HeadParams p) {
  __shared__ float part[8][128];
  __shared__ float z[16];
  const int n = blockIdx.x;
  const int c = threadIdx.x & 15, pl = threadIdx.x >> 4;
  float s[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = 0.f;
  for (int q = pl; q < p.npos; q += 8) {
    float v[8];
    load8<T>(view_addr<T>(p.in, c, n * p.npos + q, 0, 0, 0), v);
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] += v[j];
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) part[pl][c * 8 + j] = s[j];
  __syncthreads();
  // channel sums -> class means
  const int group = 128 / p.classes;
  if (threadIdx.x < p.classes) {
    float acc = 0.f;
    for (int ch = threadIdx.x * group; ch < (threadIdx.x + 1) * group; ++ch)
      for (int q = 0; q < 8; ++q) acc += part[q][ch];
    z[threadIdx.x] = acc * p.inv_count;
  }
  __syncthreads();
  if (threadIdx.x < p.classes) {
    float mx = -INFINITY;
    for (int j = 0; j < p.classes; ++j) mx = fmaxf(mx, z[j]);
    float se = 0.f;
    for (int j = 0; j < p.classes; ++j) se += expf(z[j] - mx);
    p.out[(size_t)n * p.classes + threadIdx.x] = z[threadIdx.x] - mx - logf(se);
  }
}

cudaError_t launch_head(const HeadParams& p, int precision, cudaStream_t stream) {
  if (precision == 0)
    head_kernel<__nv_bfloat16><<<p.nb, 128, 0, stream>>>(p);
  else
    head_kernel<float><<<p.nb, 128, 0, stream>>>(p);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// weight packing (+ eval BatchNorm fold: w' = w*g/sqrt(rv+eps), b' = beta + (b - rm)*g/sqrt(rv+eps))
// thread = one element of the source weight [cout, cin, k, k].
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) pack_kernel(const __grid_constant__ PackParams p) {
  const PackSeg& s = p.seg;
  const int kk = s.ksize * s.ksize;
  const int total = s.cout * s.cin * kk;
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  const int co = i / (s.cin * kk);
  const int r1 = i - co * (s.cin * kk);
  const int ci = r1 / kk;
  const int rs = r1 - ci * kk;
  float scale = 1.f;
  if (p.fold) scale = s.gamma[co] * rsqrtf(s.rvar[co] + p.eps);
  if (p.fold && r1 == 0) {
    float b = s.conv_bias ? s.conv_bias[co] : 0.f;
    p.bias_dst[s.cout_off + co] = s.beta[co] + (b - s.rmean[co]) * scale;
  }
  const float v = s.w[i] * scale;
  int tap = rs, k = ci;
  if (ci >= s.split) {
    tap += 1;
    k -= s.split;
  }
  k += s.k_off;
  const int cop = s.cout_off + co;
  if (p.fp32) {
    reinterpret_cast<float*>(p.dst)[((size_t)tap * p.K_tap + k) * p.Cout_p + cop] = v;
  } else {
    size_t e = (((size_t)tap * (p.K_tap / 8) + (k >> 3)) * p.Cout_p + cop) * 8 + (k & 7);
    reinterpret_cast<__nv_bfloat16*>(p.dst)[e] = __float2bfloat16(v);
  }
}

cudaError_t launch_pack(const PackParams& p, cudaStream_t stream) {
  int total = p.seg.cout * p.seg.cin * p.seg.ksize * p.seg.ksize;
  pack_kernel<<<(total + 255) / 256, 256, 0, stream>>>(p);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// fp32 NCHW <-> view
// ---------------------------------------------------------------------------------------------
template <typename T, bool ToView>
__global__ void __launch_bounds__(256) convert_kernel(float* nchw, ActView v, int nb, int C, int H, int W) {
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  const long long total = (long long)nb * (C / 8) * H * W;
  if (idx >= total) return;
  const int w = (int)(idx % W);
  const int h = (int)((idx / W) % H);
  const int c = (int)((idx / ((long long)W * H)) % (C / 8));
  const int n = (int)(idx / ((long long)W * H * (C / 8)));
  const int Wp = W + 1, Pimg = (H + 1) * Wp;
  char* g = view_addr<T>(v, c, n * Pimg + h * Wp + w, n, h, w);
  float t[8];
  if (ToView) {
#pragma unroll
    for (int j = 0; j < 8; ++j) t[j] = nchw[(((size_t)n * C + c * 8 + j) * H + h) * W + w];
    store8<T>(g, t);
  } else {
    load8<T>(g, t);
#pragma unroll
    for (int j = 0; j < 8; ++j) nchw[(((size_t)n * C + c * 8 + j) * H + h) * W + w] = t[j];
  }
}

cudaError_t launch_nchw_to_view(const float* nchw, ActView v, int nb, int C, int H, int W, int precision,
                                cudaStream_t stream) {
  long long total = (long long)nb * (C / 8) * H * W;
  unsigned grid = (unsigned)((total + 255) / 256);
  if (precision == 0)
    convert_kernel<__nv_bfloat16, true><<<grid, 256, 0, stream>>>(const_cast<float*>(nchw), v, nb, C, H, W);
  else
    convert_kernel<float, true><<<grid, 256, 0, stream>>>(const_cast<float*>(nchw), v, nb, C, H, W);
  return cudaGetLastError();
}
cudaError_t launch_view_to_nchw(ActView v, float* nchw, int nb, int C, int H, int W, int precision,
                                cudaStream_t stream) {
  long long total = (long long)nb * (C / 8) * H * W;
  unsigned grid = (unsigned)((total + 255) / 256);
  if (precision == 0)
    convert_kernel<__nv_bfloat16, false><<<grid, 256, 0, stream>>>(nchw, v, nb, C, H, W);
  else
    convert_kernel<float, false><<<grid, 256, 0, stream>>>(nchw, v, nb, C, H, W);
  return cudaGetLastError();
}

}  // namespace mtldas
