// Training-mode kernels of the MTL-DAS hot path: batch-statistics BatchNorm (stats / finalize / fused apply),
// fused elementwise backward (ReLU / residual / sigmoid*feature / ReLU+maxpool with BN backward), weight gradients,
// head backward.  All streaming kernels move 16/32-byte granules with consecutive threads on consecutive positions.
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
#include "ptx.cuh"
#include "train_kernels.h"

namespace mtldas {

namespace {

struct Pos {
  int n, h, w;
  bool valid;
};
__device__ __forceinline__ Pos decode(int f, int Pimg, int Wp, int H, int W) {
  Pos p;
  p.n = f / Pimg;
  const int rem = f - p.n * Pimg;
  p.h = rem / Wp;
  p.w = rem - p.h * Wp;
  p.valid = p.h < H && p.w < W;
  return p;
}

__device__ __forceinline__ void load_affine(const float* a, int c0, float* v) {
  const float4 x = reinterpret_cast<const float4*>(a + c0)[0], y = reinterpret_cast<const float4*>(a + c0)[1];
  v[0] = x.x, v[1] = x.y, v[2] = x.z, v[3] = x.w, v[4] = y.x, v[5] = y.y, v[6] = y.z, v[7] = y.w;
}

__device__ __forceinline__ void load_affine_d(const double* a, int c0, float k, float* v) {
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = (float)(a[c0 + j] * (double)k);
}

// block-wide sum of 8 per-thread values -> atomicAdd(double) by warp 0; blockDim.x == 256
__device__ __forceinline__ void block_reduce8_atomic(float* v, double* dst, float (*sh)[8]) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[j] += __shfl_xor_sync(0xffffffffu, v[j], o);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (lane == 0) {
#pragma unroll
    for (int j = 0; j < 8; ++j) sh[warp][j] = v[j];
  }
  __syncthreads();
  if (threadIdx.x < 8) {
    float s = 0.f;
    for (int wv = 0; wv < 8; ++wv) s += sh[wv][threadIdx.x];
    atomicAdd(dst + threadIdx.x, (double)s);
  }
  __syncthreads();
}

}  // namespace

__device__ __forceinline__ void finalize_channel(const FinalizeParams& p, int c) {
  if (c >= p.C) return;
  const double mean = __ldcg(p.bn.sum + c) / p.n;  // L2 reads: the sums were produced by other blocks' atomics
  double var = __ldcg(p.bn.sumsq + c) / p.n - mean * mean;
  if (var < 0.0) var = 0.0;
  const double invstd = rsqrt(var + (double)p.eps);
  float scale = 0.f, shift = 0.f, fm = 0.f, fi = 0.f;
  for (int s = 0; s < p.nseg; ++s) {
    const FinalizeSeg& g = p.seg[s];
    if (c >= g.ch_off && c < g.ch_off + g.count) {
      const int i = c - g.ch_off;
      fm = (float)mean, fi = (float)invstd;
      scale = (float)((double)g.gamma[i] * invstd);
      shift = (float)((double)g.beta[i] - mean * (double)g.gamma[i] * invstd);
      const double bias = g.conv_bias ? (double)g.conv_bias[i] : 0.0;
      const double unbiased = p.n > 1.0 ? var * p.n / (p.n - 1.0) : var;
      g.running_mean[i] = (float)((1.0 - p.momentum) * (double)g.running_mean[i] + p.momentum * (mean + bias));
      g.running_var[i] = (float)((1.0 - p.momentum) * (double)g.running_var[i] + p.momentum * unbiased);
      if (i == 0) *g.num_batches_tracked += 1;
    }
  }
  p.bn.scale[c] = scale, p.bn.shift[c] = shift, p.bn.mean[c] = fm, p.bn.invstd[c] = fi;
}

__global__ void finalize_kernel(const __grid_constant__ FinalizeParams p) { finalize_channel(p, threadIdx.x); }

// ---------------------------------------------------------------------------------------------
// batch statistics: per-channel sum and sum of squares (pads and guards are zero -> planes are summed whole)
// grid (ceil(body/1024), nchunks), block 256
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) stats_kernel(const __grid_constant__ StatsParams p) {
  __shared__ float sh[8][8];
  const int c = blockIdx.y;
  float s[8], q[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
  const long long base = (long long)blockIdx.x * 1024 + threadIdx.x;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long f = base + i * 256;
    if (f < p.body) {
      float v[8];
      load8<T>(p.y.base + ((long long)c * p.y.plane + p.y.g_lo + f) * Gran<T>::bytes, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] += v[j], q[j] = fmaf(v[j], v[j], q[j]);
    }
  }
  block_reduce8_atomic(s, p.sum + c * 8, sh);
  block_reduce8_atomic(q, p.sumsq + c * 8, sh);
  // last block to finish turns the sums into scale / shift / running statistics (no separate finalize launch)
  __shared__ unsigned int last;
  __threadfence();
  if (threadIdx.x == 0) last = atomicAdd(p.done, 1u) == gridDim.x * gridDim.y - 1 ? 1u : 0u;
  __syncthreads();
  if (last) {
    __threadfence();
    finalize_channel(p.fin, threadIdx.x);
  }
}

cudaError_t launch_stats(const StatsParams& p, int precision, cudaStream_t st) {
  dim3 grid((unsigned)((p.body + 1023) / 1024), p.nchunks);
  if (precision == 0)
    stats_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(p);
  else
    stats_kernel<float><<<grid, 256, 0, st>>>(p);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// finalize: mean / invstd / scale / shift, running statistics (one block, one thread per channel)
// ---------------------------------------------------------------------------------------------
cudaError_t launch_finalize(const FinalizeParams& p, cudaStream_t st) {
  finalize_kernel<<<1, 256, 0, st>>>(p);
  return cudaGetLastError();
}

__global__ void bn_grad_kernel(const __grid_constant__ BnGradParams p) {
  const int c = threadIdx.x;
  for (int s = 0; s < p.nseg; ++s) {
    const BnGradSeg& g = p.seg[s];
    if (c >= g.ch_off && c < g.ch_off + g.count) {
      g.dgamma[c - g.ch_off] = (float)p.bn.red_gy[c];
      g.dbeta[c - g.ch_off] = (float)p.bn.red_g[c];
    }
  }
}
cudaError_t launch_bn_grad(const BnGradParams& p, cudaStream_t st) {
  bn_grad_kernel<<<1, 256, 0, st>>>(p);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// fused BatchNorm apply + activation (forward)
// thread = (position, chunk-of-a-task); EW_SIGMOID_MUL loops over the tasks that share the feature map;
// EW_RELU_POOL: thread = (pooled position, chunk).
// ---------------------------------------------------------------------------------------------
#ifndef MTLDAS_EW_PPT
#define MTLDAS_EW_PPT 1
#endif
constexpr int kEwPpt = MTLDAS_EW_PPT;  // positions per thread (per-channel BatchNorm constants are loaded once per thread)

template <typename T>
__global__ void __launch_bounds__(256) ew_fwd_kernel(const __grid_constant__ EwParams p) {
  constexpr int G = Gran<T>::bytes;
  if (p.mode == EW_RELU_POOL) {
    const long long idx = (long long)blockIdx.x * 256 + threadIdx.x;
    if (idx >= p.body * p.nchunks) return;
    const int c = (int)(idx / p.body);
    const int f = (int)(idx - (long long)c * p.body);
    const Pos o = decode(f, p.Pimgo, p.Wpo, p.Ho, p.Wo);
    float m[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) m[j] = 0.f;
    if (o.valid) {
      float sc[8], sh[8];
      load_affine(p.bn.scale, c * 8, sc);
      load_affine(p.bn.shift, c * 8, sh);
#pragma unroll
      for (int dy = 0; dy < 2; ++dy) {
        const int h = 2 * o.h + dy;
        if (h >= p.H) continue;
#pragma unroll
        for (int dx = 0; dx < 2; ++dx) {
          const int w = 2 * o.w + dx;
          if (w >= p.W) continue;
          float v[8];
          load8<T>(view_addr<T>(p.y, c, o.n * p.Pimg + h * p.Wp + w, o.n, h, w), v);
#pragma unroll
          for (int j = 0; j < 8; ++j) m[j] = fmaxf(m[j], fmaf(v[j], sc[j], sh[j]));  // relu folded into max with 0
        }
      }
    }
    store8<T>(view_addr<T>(p.out, c, f, o.n, o.h, o.w), m);
    return;
  }
  // thread = chunk c (of a task) x kEwPpt positions 256 apart; a block covers 1024 consecutive positions of one chunk
  const long long per_chunk = (p.body + 256 * kEwPpt - 1) / (256 * kEwPpt) * (256 * kEwPpt);
  const long long base = (long long)blockIdx.x * (256 * kEwPpt) + threadIdx.x;
  const int c = (int)(base / per_chunk);
  const int f0 = (int)(base - (long long)c * per_chunk);
  for (int t = 0; t < p.ntasks; ++t) {
    const int ch = t * p.nchunks + c;
    float sc[8], sh[8], s2[8], h2[8];
    load_affine(p.bn.scale, ch * 8, sc);
    load_affine(p.bn.shift, ch * 8, sh);
    if (p.mode == EW_ADD_RELU_SC) {
      load_affine(p.bn2.scale, ch * 8, s2);
      load_affine(p.bn2.shift, ch * 8, h2);
    }
#pragma unroll
    for (int i = 0; i < kEwPpt; ++i) {
      const int f = f0 + i * 256;
      if (f >= p.body) continue;
      const Pos q = decode(f, p.Pimg, p.Wp, p.H, p.W);
      if (!q.valid && p.out.layout != LAY_P8) continue;
      float z[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) z[j] = 0.f;
      if (q.valid) {
        float v[8];
        load8<T>(p.y.base + ((long long)ch * p.y.plane + p.y.g_lo + f) * G, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = fmaf(v[j], sc[j], sh[j]);
        if (p.mode == EW_RELU) {
#pragma unroll
          for (int j = 0; j < 8; ++j) z[j] = fmaxf(v[j], 0.f);
        } else if (p.mode == EW_ADD_RELU_SC) {
          float v2[8];
          load8<T>(p.y2.base + ((long long)ch * p.y2.plane + p.y2.g_lo + f) * G, v2);
#pragma unroll
          for (int j = 0; j < 8; ++j) z[j] = fmaxf(v[j] + fmaf(v2[j], s2[j], h2[j]), 0.f);
        } else {
          float aux[8];
          load8<T>(view_addr<T>(p.aux, c, f, q.n, q.h, q.w), aux);
          if (p.mode == EW_ADD_RELU_ID) {
#pragma unroll
            for (int j = 0; j < 8; ++j) z[j] = fmaxf(v[j] + aux[j], 0.f);
          } else {  // EW_SIGMOID_MUL
#pragma unroll
            for (int j = 0; j < 8; ++j) z[j] = aux[j] / (1.f + expf(-v[j]));
          }
        }
      }
      store8<T>(view_addr<T>(p.out, ch, f, q.n, q.h, q.w), z);
    }
  }
}

cudaError_t launch_ew_fwd(const EwParams& p, int precision, cudaStream_t st) {
  const long long total = p.body * p.nchunks;
  const unsigned grid = p.mode == EW_RELU_POOL ? (unsigned)((total + 255) / 256)
                                               : (unsigned)((p.body + 256 * kEwPpt - 1) / (256 * kEwPpt) * p.nchunks);
  if (precision == 0)
    ew_fwd_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(p);
  else
    ew_fwd_kernel<float><<<grid, 256, 0, st>>>(p);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// fused elementwise backward.  PASS 0: reductions sum(g), sum(g*y_hat); PASS 1: dy (and d aux).
// g = dz * activation'(.) is recomputed in both passes from dz and the saved raw tensors.
// ---------------------------------------------------------------------------------------------
// d gamma / d beta of chunk c (all tasks) from the completed reductions: written once, by the thread at f == 0
__device__ __forceinline__ void write_bn_grads(const EwParams& p, const BnArrays& bn, int c) {
  const int ntask = p.mode == EW_RELU_POOL ? 1 : p.ntasks;
  for (int t = 0; t < ntask; ++t)
    for (int j = 0; j < 8; ++j) {
      const int ch = (t * p.nchunks + c) * 8 + j;
      for (int s = 0; s < p.ngseg; ++s) {
        const BnGradSeg& g = p.gseg[s];
        if (ch >= g.ch_off && ch < g.ch_off + g.count) {
          g.dgamma[ch - g.ch_off] = (float)bn.red_gy[ch];
          g.dbeta[ch - g.ch_off] = (float)bn.red_g[ch];
        }
      }
    }
}

template <typename T, int PASS>
__global__ void __launch_bounds__(256) ew_bwd_kernel(const __grid_constant__ EwParams p) {
  __shared__ float sh[8][8];
  // a block never straddles two chunks: pool mode covers 256 pooled positions per block, the other modes 1024
  // positions (kEwPpt per thread); the launcher sizes the grid accordingly
  const int span = p.mode == EW_RELU_POOL ? 256 : 256 * kEwPpt;
  const long long per_chunk = (p.body + span - 1) / span * span;
  const long long idx2 = (long long)blockIdx.x * span + threadIdx.x;
  const int c = (int)(idx2 / per_chunk);
  const int f = (int)(idx2 - (long long)c * per_chunk);
  const bool in_range = f < p.body;
  constexpr int G = Gran<T>::bytes;
  const BnArrays& bn = p.second ? p.bn2 : p.bn;     // the BatchNorm whose input gradient this launch produces
  const ActView& ym = p.second ? p.y2 : p.y;

  if (p.mode == EW_RELU_POOL) {
    // thread = pooled position; the 2x2 window's four full-resolution positions belong to this thread alone
    float rg[8], rgy[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) rg[j] = rgy[j] = 0.f;
    Pos o = decode(in_range ? f : 0, p.Pimgo, p.Wpo, p.Ho, p.Wo);
    if (in_range && o.valid) {
      float sc[8], shf[8], mu[8], is[8], dzv[8];
      load_affine(bn.scale, c * 8, sc);
      load_affine(bn.shift, c * 8, shf);
      load_affine(bn.mean, c * 8, mu);
      load_affine(bn.invstd, c * 8, is);
      load8<T>(view_addr<T>(p.dz, c, f, o.n, o.h, o.w), dzv);
      float yv[4][8];
      bool inb[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int h = 2 * o.h + (k >> 1), w = 2 * o.w + (k & 1);
        inb[k] = h < p.H && w < p.W;
        if (inb[k]) load8<T>(view_addr<T>(ym, c, o.n * p.Pimg + h * p.Wp + w, o.n, h, w), yv[k]);
      }
      int arg[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float best = -INFINITY;
        arg[j] = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          if (!inb[k]) continue;
          const float r = fmaxf(fmaf(yv[k][j], sc[j], shf[j]), 0.f);
          if (r > best) best = r, arg[j] = k;   // strict '>' : first maximum wins (scan order h, w)
        }
      }
      float cg[8], cgy[8];
      if (PASS == 1) {
        load_affine_d(bn.red_g, c * 8, p.inv_n, cg);
        load_affine_d(bn.red_gy, c * 8, p.inv_n, cgy);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (!inb[k]) continue;
        float dyv[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float v = fmaf(yv[k][j], sc[j], shf[j]);
          const float g = (arg[j] == k && v > 0.f) ? dzv[j] : 0.f;
          const float yh = (yv[k][j] - mu[j]) * is[j];
          if (PASS == 0) {
            rg[j] += g, rgy[j] = fmaf(g, yh, rgy[j]);
          } else {
            dyv[j] = sc[j] * (g - cg[j] - yh * cgy[j]);
          }
        }
        if (PASS == 1) {
          const int h = 2 * o.h + (k >> 1), w = 2 * o.w + (k & 1);
          store8<T>(view_addr<T>(p.dy, c, o.n * p.Pimg + h * p.Wp + w, o.n, h, w), dyv);
        }
      }
    }
    if (PASS == 1 && f == 0) write_bn_grads(p, bn, c);
    if (PASS == 0) {
      block_reduce8_atomic(rg, bn.red_g + c * 8, sh);
      block_reduce8_atomic(rgy, bn.red_gy + c * 8, sh);
    } else if (f < 16) {
      // the tensor-core weight gradient reduces 16 positions per MMA: keep [body, round_up(body,16)) zero
      const long long full = (p.body / p.Pimgo) * p.Pimg;
      if (full + f < (full + 15) / 16 * 16) {
        float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        store8<T>(p.dy.base + ((long long)c * p.dy.plane + p.dy.g_lo + full + f) * G, z);
      }
    }
    return;
  }

  // non-pool modes: thread = chunk c x kEwPpt positions (f, f+256, ...) of a 1024-position block
  float dacc[kEwPpt][8];
#pragma unroll
  for (int i = 0; i < kEwPpt; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) dacc[i][j] = 0.f;
  const int body16 = (int)((p.body + 15) / 16 * 16);
  for (int t = 0; t < p.ntasks; ++t) {
    const int ch = t * p.nchunks + c;
    float sc[8], shf[8], mu[8], is[8], s2[8], h2[8], cg[8], cgy[8], rg[8], rgy[8];
    load_affine(bn.scale, ch * 8, sc);
    load_affine(bn.shift, ch * 8, shf);
    load_affine(bn.mean, ch * 8, mu);
    load_affine(bn.invstd, ch * 8, is);
    if (p.mode == EW_ADD_RELU_SC) {
      const BnArrays& bo = p.second ? p.bn : p.bn2;
      load_affine(bo.scale, ch * 8, s2);
      load_affine(bo.shift, ch * 8, h2);
    }
    if (PASS == 1) {
      load_affine_d(bn.red_g, ch * 8, p.inv_n, cg);
      load_affine_d(bn.red_gy, ch * 8, p.inv_n, cgy);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) rg[j] = rgy[j] = 0.f;
#pragma unroll
    for (int i = 0; i < kEwPpt; ++i) {
      const int fi = f + i * 256;
      const bool in_r = fi < p.body;
      const Pos q = decode(in_r ? fi : 0, p.Pimg, p.Wp, p.H, p.W);
      const bool valid = in_r && q.valid;
      float dyv[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) dyv[j] = 0.f;
      if (valid) {
        float y[8], dzv[8], g[8];
        load8<T>(ym.base + ((long long)ch * ym.plane + ym.g_lo + fi) * G, y);
        load8<T>(view_addr<T>(p.dz, ch, fi, q.n, q.h, q.w), dzv);
        if (p.mode == EW_RELU) {
#pragma unroll
          for (int j = 0; j < 8; ++j) g[j] = fmaf(y[j], sc[j], shf[j]) > 0.f ? dzv[j] : 0.f;
        } else if (p.mode == EW_ADD_RELU_SC) {
          const ActView& yo = p.second ? p.y : p.y2;
          float y2[8];
          load8<T>(yo.base + ((long long)ch * yo.plane + yo.g_lo + fi) * G, y2);
#pragma unroll
          for (int j = 0; j < 8; ++j) g[j] = (fmaf(y[j], sc[j], shf[j]) + fmaf(y2[j], s2[j], h2[j])) > 0.f ? dzv[j] : 0.f;
        } else {
          float aux[8];
          load8<T>(view_addr<T>(p.aux, c, fi, q.n, q.h, q.w), aux);
          if (p.mode == EW_ADD_RELU_ID) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              g[j] = (fmaf(y[j], sc[j], shf[j]) + aux[j]) > 0.f ? dzv[j] : 0.f;
              dacc[i][j] = g[j];
            }
          } else {  // EW_SIGMOID_MUL
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const float sg = 1.f / (1.f + expf(-fmaf(y[j], sc[j], shf[j])));
              g[j] = dzv[j] * aux[j] * sg * (1.f - sg);
              dacc[i][j] = fmaf(dzv[j], sg, dacc[i][j]);
            }
          }
        }
        if (PASS == 0) {
#pragma unroll
          for (int j = 0; j < 8; ++j) rg[j] += g[j], rgy[j] = fmaf(g[j], (y[j] - mu[j]) * is[j], rgy[j]);
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) dyv[j] = sc[j] * (g[j] - cg[j] - (y[j] - mu[j]) * is[j] * cgy[j]);
        }
      }
      // zero at pads, and zero up to the next multiple of 16 positions (tensor-core wgrad granularity)
      if (PASS == 1 && fi < body16) store8<T>(p.dy.base + ((long long)ch * p.dy.plane + p.dy.g_lo + fi) * G, dyv);
    }
    if (PASS == 0) {
      block_reduce8_atomic(rg, bn.red_g + ch * 8, sh);
      block_reduce8_atomic(rgy, bn.red_gy + ch * 8, sh);
    }
  }
  if (PASS == 1 && f == 0) write_bn_grads(p, bn, c);
  if (PASS == 1 && p.want_daux && !p.second) {
#pragma unroll
    for (int i = 0; i < kEwPpt; ++i) {
      const int fi = f + i * 256;
      if (fi >= p.body) continue;
      const Pos q = decode(fi, p.Pimg, p.Wp, p.H, p.W);
      if (!(q.valid || p.daux.layout == LAY_P8)) continue;
      char* a = view_addr<T>(p.daux, c, fi, q.n, q.h, q.w);
      if (p.daux_accum) {
        if (!q.valid) continue;
        float old[8];
        load8<T>(a, old);
#pragma unroll
        for (int j = 0; j < 8; ++j) dacc[i][j] += old[j];
      }
      store8<T>(a, dacc[i]);
    }
  }
}

cudaError_t launch_ew_bwd(const EwParams& p, int pass, int precision, cudaStream_t st) {
  const int span = p.mode == EW_RELU_POOL ? 256 : 256 * kEwPpt;
  const long long per_chunk = (p.body + span - 1) / span * span;
  const unsigned grid = (unsigned)(per_chunk * p.nchunks / span);
  if (precision == 0) {
    if (pass == 0)
      ew_bwd_kernel<__nv_bfloat16, 0><<<grid, 256, 0, st>>>(p);
    else
      ew_bwd_kernel<__nv_bfloat16, 1><<<grid, 256, 0, st>>>(p);
  } else {
    if (pass == 0)
      ew_bwd_kernel<float, 0><<<grid, 256, 0, st>>>(p);
    else
      ew_bwd_kernel<float, 1><<<grid, 256, 0, st>>>(p);
  }
  return cudaGetLastError();
}

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

static cudaError_t launch_wgrad_umma(WgradParams p, cudaStream_t st) {
  const int Cin = p.nchunks * 8;
  const int tg_max = std::max(1, 512 / Cin);
  const int ngroups = (p.ntaps + tg_max - 1) / tg_max;
  p.taps_per_cta = (p.ntaps + ngroups - 1) / ngroups;
  int lo = 0, hi = 0;
  for (int t = 0; t < p.ntaps; ++t) lo = std::max(lo, -p.tap[t].off), hi = std::max(hi, p.tap[t].off);
  p.halo_lo = lo, p.halo_hi = hi;
  const long long body16 = (p.body + 15) / 16 * 16;
  long long want = std::max<long long>(1, 444 / ngroups);              // about three CTAs per SM
  long long slice = std::max<long long>(1024, (body16 + want - 1) / want);
  p.slice = (int)((slice + kWgStage - 1) / kWgStage * kWgStage);
  const int SPb = kWgStage + lo + hi;
  const size_t b_stage = ((size_t)(p.nsrc * p.nchunks) * SPb * 16 + 127) & ~(size_t)127;
  const size_t smem = kWgHeader + 2 * (size_t)16 * kWgStage * 16 + 2 * b_stage;
  static size_t configured = 0;
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(wgrad_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    configured = smem;
  }
  dim3 grid((unsigned)((body16 + p.slice - 1) / p.slice), ngroups);
  wgrad_umma_kernel<<<grid, 128, smem, st>>>(p);
  return cudaGetLastError();
}

cudaError_t launch_wgrad(const WgradParams& p, int precision, cudaStream_t st) {
  if (precision == 0) {
    const char* e = getenv("MTLDAS_WGRAD_CUDA_CORES");
    if (!(e && e[0] == '1')) return launch_wgrad_umma(p, st);
    return launch_wgrad_t<__nv_bfloat16>(p, st);
  }
  return launch_wgrad_t<float>(p, st);
}

// ---------------------------------------------------------------------------------------------
// stem weight gradient: dW[tap][co] += sum_pos x[n, 3h-2+r, 3w-2+s] * dy[pos][co]
// block 256 handles `slice` positions, 64 at a time staged in shared memory; 784 outputs over 256 threads.
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) stem_wgrad_kernel(const __grid_constant__ StemWgradParams p, int slice) {
  __shared__ float xw[64][49];
  __shared__ float dys[64][16];
  const long long f0 = (long long)blockIdx.x * slice, f1 = min(f0 + (long long)slice, p.body);
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  for (long long fb = f0; fb < f1; fb += 64) {
    __syncthreads();
    for (int i = threadIdx.x; i < 64 * 49; i += 256) {
      const int pp = i / 49, t = i - pp * 49;
      float v = 0.f;
      if (fb + pp < f1) {
        const Pos q = decode((int)(fb + pp), p.Pimg, p.Wp, p.H, p.W);
        const int iy = 3 * q.h - 2 + t / 7, ix = 3 * q.w - 2 + t % 7;
        if (q.valid && iy >= 0 && iy < p.Hin && ix >= 0 && ix < p.Win) v = p.x[((size_t)q.n * p.Hin + iy) * p.Win + ix];
      }
      xw[pp][t] = v;
    }
    for (int i = threadIdx.x; i < 64 * 2; i += 256) {
      const int pp = i >> 1, c = i & 1;
      float v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) v[j] = 0.f;
      if (fb + pp < f1) load8<T>(p.dy.base + ((long long)c * p.dy.plane + p.dy.g_lo + fb + pp) * Gran<T>::bytes, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) dys[pp][c * 8 + j] = v[j];
    }
    __syncthreads();
#pragma unroll
    for (int r = 0; r < 4; ++r) {
      const int o = threadIdx.x + 256 * r;
      if (o < 784) {
        const int t = o >> 4, co = o & 15;
        float a = acc[r];
        for (int pp = 0; pp < 64; ++pp) a = fmaf(xw[pp][t], dys[pp][co], a);
        acc[r] = a;
      }
    }
  }
#pragma unroll
  for (int r = 0; r < 4; ++r) {
    const int o = threadIdx.x + 256 * r;
    if (o < 784) atomicAdd(p.dw + o, acc[r]);
  }
}

cudaError_t launch_stem_wgrad(const StemWgradParams& p, int precision, cudaStream_t st) {
  const int slice = 2048;
  const unsigned grid = (unsigned)((p.body + slice - 1) / slice);
  if (precision == 0)
    stem_wgrad_kernel<__nv_bfloat16><<<grid, 256, 0, st>>>(p, slice);
  else
    stem_wgrad_kernel<float><<<grid, 256, 0, st>>>(p, slice);
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
__global__ void __launch_bounds__(256) unpack_kernel(const __grid_constant__ UnpackParams p) {
  const int kk = p.ksize * p.ksize;
  const int total = p.cout * p.cin * kk;
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= total) return;
  const int co = i / (p.cin * kk);
  const int r1 = i - co * (p.cin * kk);
  const int ci = r1 / kk;
  int tap = r1 - ci * kk, k = ci;
  if (ci >= p.split) tap += 1, k -= p.split;
  k += p.k_off;
  p.dst[i] = p.src[((size_t)tap * p.K_tap + k) * p.Cout_p + p.cout_off + co];
}

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

cudaError_t launch_unpack(const UnpackParams& p, cudaStream_t st) {
  const int total = p.cout * p.cin * p.ksize * p.ksize;
  unpack_kernel<<<(total + 255) / 256, 256, 0, st>>>(p);
  return cudaGetLastError();
}

}  // namespace mtldas
