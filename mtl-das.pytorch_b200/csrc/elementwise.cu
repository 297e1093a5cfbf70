// Streaming (HBM-bound) kernels of the MTL-DAS hot path: stem convolution (Cin = 1), 2x2 ceil max-pool,
// classification heads, weight packing with eval-BatchNorm folding, layout conversion.
// All global accesses are 16/32-byte granules, consecutive threads -> consecutive granules.
#include <climits>
#include <cstdio>

#include "gran.cuh"
#include "kernels.h"
#include "ptx.cuh"

namespace mtldas {

template <typename K, typename P>
static cudaError_t launch_ex(K kernel, dim3 grid, dim3 block, cudaStream_t stream, bool pdl, const P& params) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid, cfg.blockDim = block, cfg.dynamicSmemBytes = 0, cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr, cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, params);
}

// ---------------------------------------------------------------------------------------------
// stem: Conv2d(1,16,7x7,stride 3,pad 2,bias=False)+BN(eval, folded)+ReLU -- modelA_MTL.py:73-77,128
// thread = one flattened output position, 16 output channels in registers.
// ---------------------------------------------------------------------------------------------
// thread = two horizontally adjacent output positions (Wp = 84 is even, so a pair never straddles a row):
// the 7x10 input window is read once for both, every weight float4 fetched from shared memory feeds two
// positions, and the FMAs are issued as packed fma.rn.f32x2 (two channels per instruction).
template <typename T>
__global__ void __launch_bounds__(128) stem_kernel(const __grid_constant__ StemParams p) {
  __shared__ __align__(16) float ws[49 * 16];
  __shared__ float bs[16];
  ptx::pdl_launch_dependents();
  for (int i = threadIdx.x; i < 49 * 16; i += 128) ws[i] = p.w[i];
  if (threadIdx.x < 16) bs[threadIdx.x] = p.bias[threadIdx.x];
  __syncthreads();
  ptx::pdl_grid_wait();  // the output tensor may still be read by the previous pass's kernels
  const int f = (blockIdx.x * 128 + threadIdx.x) * 2;
  if (f >= p.body) return;
  const int n = f / p.Pimg;
  const int rem = f - n * p.Pimg;
  const int h = rem / p.Wp, w = rem - (rem / p.Wp) * p.Wp;
  float2 acc[2][8];
#pragma unroll
  for (int q = 0; q < 8; ++q) acc[0][q] = acc[1][q] = make_float2(0.f, 0.f);
  const bool row_ok = h < p.H;
  const bool v0 = row_ok && w < p.W, v1 = row_ok && (w + 1) < p.W;
  if (v0) {
    const float* xin = p.x + (size_t)n * p.Hin * p.Win;
    const int ix0 = 3 * w - 2;
#pragma unroll 1
    for (int r = 0; r < 7; ++r) {
      const int iy = 3 * h - 2 + r;
      if (iy < 0 || iy >= p.Hin) continue;
      const float* row = xin + iy * p.Win;
      float xv[10];
#pragma unroll
      for (int s = 0; s < 10; ++s) {
        const int ix = ix0 + s;
        xv[s] = (ix >= 0 && ix < p.Win) ? __ldg(row + ix) : 0.f;
      }
#pragma unroll
      for (int s = 0; s < 7; ++s) {
        const float4* wr = reinterpret_cast<const float4*>(ws + (r * 7 + s) * 16);
        const float2 xa = make_float2(xv[s], xv[s]), xb = make_float2(xv[s + 3], xv[s + 3]);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          const float4 w4 = wr[q];
          const float2 wlo = make_float2(w4.x, w4.y), whi = make_float2(w4.z, w4.w);
          acc[0][2 * q] = __ffma2_rn(xa, wlo, acc[0][2 * q]);
          acc[0][2 * q + 1] = __ffma2_rn(xa, whi, acc[0][2 * q + 1]);
          acc[1][2 * q] = __ffma2_rn(xb, wlo, acc[1][2 * q]);
          acc[1][2 * q + 1] = __ffma2_rn(xb, whi, acc[1][2 * q + 1]);
        }
      }
    }
  }
#pragma unroll
  for (int pos = 0; pos < 2; ++pos) {
    const bool valid = pos == 0 ? v0 : v1;
    float o[16];
#pragma unroll
    for (int q = 0; q < 8; ++q) {
      const float lo = p.raw ? -INFINITY : 0.f;
      o[2 * q] = valid ? fmaxf(acc[pos][q].x + bs[2 * q], lo) : 0.f;
      o[2 * q + 1] = valid ? fmaxf(acc[pos][q].y + bs[2 * q + 1], lo) : 0.f;
    }
    if (f + pos < p.body) {
      store8<T>(view_addr<T>(p.out, 0, f + pos, n, h, w + pos), o);
      store8<T>(view_addr<T>(p.out, 1, f + pos, n, h, w + pos), o + 8);
    }
  }
}

cudaError_t launch_stem_umma(const StemParams& p, cudaStream_t stream, bool pdl);

cudaError_t launch_stem(const StemParams& p, int precision, cudaStream_t stream, bool pdl) {
  if (precision == 0 && p.w_umma) {
    // the tensor-core stem needs a 16-byte aligned patch tensor (bulk copies); anything else takes the FFMA kernel
    cudaError_t e = launch_stem_umma(p, stream, pdl);
    if (e != cudaErrorInvalidValue) return e;
    cudaGetLastError();
  }
  unsigned grid = (unsigned)((p.body + 255) / 256);
  if (precision == 0) return launch_ex(stem_kernel<__nv_bfloat16>, dim3(grid), dim3(128), stream, pdl, p);
  return launch_ex(stem_kernel<float>, dim3(grid), dim3(128), stream, pdl, p);
}

// ---------------------------------------------------------------------------------------------
// MaxPool2d(kernel 2, stride 2, ceil_mode=True) -- modelA_MTL.py:116,145,152,158
// thread = one output granule (position x 8 channels); windows clipped at the image border.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) pool_kernel(const __grid_constant__ PoolParams p) {
  ptx::pdl_launch_dependents();
  ptx::pdl_grid_wait();
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

cudaError_t launch_pool(const PoolParams& p, int precision, cudaStream_t stream, bool pdl) {
  long long total = p.body_o * p.nchunks;
  unsigned grid = (unsigned)((total + 255) / 256);
  if (precision == 0) return launch_ex(pool_kernel<__nv_bfloat16>, dim3(grid), dim3(256), stream, pdl, p);
  return launch_ex(pool_kernel<float>, dim3(grid), dim3(256), stream, pdl, p);
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
  ptx::pdl_launch_dependents();
  ptx::pdl_grid_wait();
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

cudaError_t launch_head(const HeadParams& p, int precision, cudaStream_t stream, bool pdl) {
  if (precision == 0) return launch_ex(head_kernel<__nv_bfloat16>, dim3(p.nb), dim3(128), stream, pdl, p);
  return launch_ex(head_kernel<float>, dim3(p.nb), dim3(128), stream, pdl, p);
}

// ---------------------------------------------------------------------------------------------
// weight packing (+ eval BatchNorm fold: w' = w*g/sqrt(rv+eps), b' = beta + (b - rm)*g/sqrt(rv+eps))
// thread = one element of the source weight [cout, cin, k, k].
// ---------------------------------------------------------------------------------------------
__device__ __forceinline__ void pack_element(const PackParams& p, int i) {
  const PackSeg& s = p.seg;
  const int kk = s.ksize * s.ksize;
  const int total = s.cout * s.cin * kk;
  if (i >= total) return;
  const int co = i / (s.cin * kk);
  const int r1 = i - co * (s.cin * kk);
  const int ci = r1 / kk;
  const int rs = r1 - ci * kk;
  int col = co;            // column inside this block's slice of the layer's output channels
  if (s.co_n > 0 && !s.transpose && !s.k_from_tap) {
    if (co < s.co_lo || co >= s.co_lo + s.co_n) return;
    col = co - s.co_lo;
  }
  float scale = 1.f;
  if (p.fold) scale = s.gamma[co] * rsqrtf(s.rvar[co] + p.eps);
  if (p.fold && r1 == 0) {
    float b = s.conv_bias ? s.conv_bias[co] : 0.f;
    p.bias_dst[s.cout_off + col] = s.beta[co] + (b - s.rmean[co]) * scale;
  }
  const float v = s.w[i] * scale;
  int tap = rs, k = ci, cop;
  if (s.k_from_tap) {  // 1: k = tap index; 2: banded stem image, k = filter row * 8 + filter column
    tap = 0, k = s.k_from_tap == 2 ? (rs / s.ksize) * 8 + rs % s.ksize : rs, cop = s.cout_off + co;
  } else if (s.transpose) {
    if (ci < s.ci_lo || ci >= s.ci_lo + s.ci_n) return;
    tap = s.tap_map[rs];
    if (tap < 0) return;
    k = s.k_off + co;
    cop = s.cout_off + (ci - s.ci_lo);
  } else {
    if (ci >= s.split) {
      tap += 1;
      k -= s.split;
    }
    k += s.k_off;
    cop = s.cout_off + col;
  }
  if (p.fp32) {
    reinterpret_cast<float*>(p.dst)[((size_t)tap * p.K_tap + k) * p.Cout_p + cop] = v;
  } else {
    size_t e = (((size_t)tap * (p.K_tap / 8) + (k >> 3)) * p.Cout_p + cop) * 8 + (k & 7);
    reinterpret_cast<__nv_bfloat16*>(p.dst)[e] = __float2bfloat16(v);
  }
}

__global__ void __launch_bounds__(256) pack_kernel(const __grid_constant__ PackParams p) {
  pack_element(p, blockIdx.x * 256 + threadIdx.x);
}

__global__ void __launch_bounds__(256) pack_batch_kernel(const __grid_constant__ PackBatch b) {
  const PackParams& p = b.p[blockIdx.y];
  const int total = p.seg.cout * p.seg.cin * p.seg.ksize * p.seg.ksize;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < total; i += gridDim.x * 256) pack_element(p, i);
}

cudaError_t launch_pack_batch(const PackParams* packs, int n, cudaStream_t stream) {
  static PackBatch b;  // host staging (32 KB kernel parameter block)
  for (int i0 = 0; i0 < n; i0 += kPackBatch) {
    const int m = n - i0 < kPackBatch ? n - i0 : kPackBatch;
    for (int i = 0; i < m; ++i) b.p[i] = packs[i0 + i];
    pack_batch_kernel<<<dim3(48, m), 256, 0, stream>>>(b);
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return e;
  }
  return cudaSuccess;
}

cudaError_t launch_pack(const PackParams& p, cudaStream_t stream) {
  int total = p.seg.cout * p.seg.cin * p.seg.ksize * p.seg.ksize;
  pack_kernel<<<(total + 255) / 256, 256, 0, stream>>>(p);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// streaming inference: cut [100 x 250] patches out of a large DAS array [n_channels, n_time] (fp32, row major).
// patch (wi, tb) = array[start0 + wi*stride : +100, tb*250 : +250]; output order: window-major, then time block.
// thread = one float4 of a patch row (250 floats = 62 float4 + 2 floats handled by the last lanes).
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) window_gather_kernel(const float* __restrict__ arr, long long n_time,
                                                            int start0, int stride, int n_windows, int n_tblocks,
                                                            float* __restrict__ out) {
  const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
  const long long total = (long long)n_windows * n_tblocks * 100 * 250;
  if (idx >= total) return;
  const int col = (int)(idx % 250);
  const int row = (int)((idx / 250) % 100);
  const long long pidx = idx / 25000;
  const int tb = (int)(pidx % n_tblocks);
  const int wi = (int)(pidx / n_tblocks);
  out[idx] = arr[(long long)(start0 + wi * stride + row) * n_time + (long long)tb * 250 + col];
}

cudaError_t launch_window_gather(const float* arr, long long n_time, int start0, int stride, int n_windows,
                                 int n_tblocks, float* out, cudaStream_t stream) {
  const long long total = (long long)n_windows * n_tblocks * 25000;
  if (total == 0) return cudaSuccess;
  window_gather_kernel<<<(unsigned)((total + 255) / 256), 256, 0, stream>>>(arr, n_time, start0, stride, n_windows,
                                                                             n_tblocks, out);
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

// ---------------------------------------------------------------------------------------------
// bf16 -> fp32 widening of host patches that came over PCIe as bf16 (mtldas_plan_forward_eval_host_bf16)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) widen_bf16_kernel(const uint4* __restrict__ in, float4* __restrict__ out, long long n8,
                                                         const uint16_t* in_tail, float* out_tail, int tail) {
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += stride) {
    const uint4 q = in[i];
    const uint32_t w[4] = {q.x, q.y, q.z, q.w};
    float f[8];
#pragma unroll
    for (int j = 0; j < 4; ++j) f[2 * j] = __uint_as_float(w[j] << 16), f[2 * j + 1] = __uint_as_float(w[j] & 0xffff0000u);
    out[2 * i] = make_float4(f[0], f[1], f[2], f[3]);
    out[2 * i + 1] = make_float4(f[4], f[5], f[6], f[7]);
  }
  if (blockIdx.x == 0 && (int)threadIdx.x < tail) out_tail[threadIdx.x] = __uint_as_float((uint32_t)in_tail[threadIdx.x] << 16);
}

cudaError_t launch_widen_bf16(const uint16_t* in, float* out, long long n, cudaStream_t stream) {
  if (n <= 0) return cudaSuccess;
  const long long n8 = n / 8;
  const int tail = (int)(n - n8 * 8);
  const int blocks = (int)std::min<long long>(148 * 8, std::max<long long>(1, (n8 + 255) / 256));
  widen_bf16_kernel<<<blocks, 256, 0, stream>>>(reinterpret_cast<const uint4*>(in), reinterpret_cast<float4*>(out), n8,
                                                in + n8 * 8, out + n8 * 8, tail);
  return cudaGetLastError();
}

}  // namespace mtldas
