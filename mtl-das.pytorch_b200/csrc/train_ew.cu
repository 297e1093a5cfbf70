// Streaming training kernels of the MTL-DAS hot path, written as chunk-resident grid-stride loops:
//   stats_kernel        per-channel sum / sum of squares of a raw conv output (+ BatchNorm finalize in the last block)
//   ew_fwd_kernel       fused BatchNorm apply + activation (ReLU / residual / sigmoid*feature / ReLU+maxpool)
//   ew_bwd_kernel       fused activation' + BatchNorm backward: PASS 0 reductions, PASS 1 input gradient
//   stem_wgrad_kernel   weight gradient of the 7x7/s3 stem from the fp32 patches
// grid = (blocks per chunk, chunks): a block owns ONE 8-channel chunk, so the per-channel BatchNorm constants
// live in registers for the whole loop, the per-thread partial sums are reduced once per block (not once per
// position), and consecutive threads move consecutive 16/32-byte granules.
//
// Reference semantics restated (autograd of /root/reference/model/modelA_MTL.py:127-174 under utils.py:359-373):
//   BatchNorm2d train: y_hat = (y - mean) * invstd (biased batch variance), out = gamma*y_hat + beta;
//   backward: dgamma = sum(g*y_hat), dbeta = sum(g), dy = gamma*invstd*(g - mean(g) - y_hat*mean(g*y_hat));
//   ReLU' = [out > 0]; sigmoid' = s(1-s); mask multiply: d mask = dz*feat, d feat = dz*mask;
//   MaxPool2d(2,2,ceil): gradient to the first maximal element of the window (scan order h, w).
#include <algorithm>
#include <cstdio>
#include <cstdlib>

#include "gran.cuh"
#include "train_kernels.h"

namespace mtldas {

namespace {

constexpr int kEwThreads = 256;

__device__ __forceinline__ int fdiv(int f, unsigned long long magic) {
  return (int)(((unsigned long long)(unsigned)f * magic) >> 40);
}

struct Pos3 {
  int n, h, w;
};
__device__ __forceinline__ Pos3 decode3(int f, int Pimg, int Wp, unsigned long long m_pimg, unsigned long long m_wp) {
  Pos3 q;
  q.n = fdiv(f, m_pimg);
  const int rem = f - q.n * Pimg;
  q.h = fdiv(rem, m_wp);
  q.w = rem - q.h * Wp;
  return q;
}

__device__ __forceinline__ void ld_affine(const float* a, int c0, float* v) {
  const float4 x = reinterpret_cast<const float4*>(a + c0)[0], y = reinterpret_cast<const float4*>(a + c0)[1];
  v[0] = x.x, v[1] = x.y, v[2] = x.z, v[3] = x.w, v[4] = y.x, v[5] = y.y, v[6] = y.z, v[7] = y.w;
}

// block-wide sum of 8 per-thread values -> atomicAdd(double) by the first 8 threads; blockDim.x == 256
__device__ __forceinline__ void block_reduce8(float* v, double* dst, float (*sh)[8]) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v[j] += __shfl_xor_sync(0xffffffffu, v[j], o);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();
  if (lane == 0) {
#pragma unroll
    for (int j = 0; j < 8; ++j) sh[warp][j] = v[j];
  }
  __syncthreads();
  if (threadIdx.x < 8) {
    float s = 0.f;
#pragma unroll
    for (int wv = 0; wv < kEwThreads / 32; ++wv) s += sh[wv][threadIdx.x];
    atomicAdd(dst + threadIdx.x, (double)s);
  }
}

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

// blocks per chunk: about four blocks of 256 threads per SM over the whole grid, never more than the work
inline unsigned blocks_per_chunk(long long positions, int chunks) {
  static int sms = 0;
  if (!sms) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    if (sms <= 0) sms = 148;
  }
  const long long need = (positions + kEwThreads - 1) / kEwThreads;
  const long long want = std::max<long long>(1, ((long long)sms * 4 + chunks - 1) / chunks);
  return (unsigned)std::max<long long>(1, std::min(need, want));
}

}  // namespace

__global__ void finalize_kernel(const __grid_constant__ FinalizeParams p) { finalize_channel(p, threadIdx.x); }

cudaError_t launch_finalize(const FinalizeParams& p, cudaStream_t st) {
  finalize_kernel<<<1, 256, 0, st>>>(p);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// batch statistics (pads and guards are zero -> planes are summed whole, no position decoding)
// ---------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kEwThreads) stats_kernel(const __grid_constant__ StatsParams p) {
  __shared__ float sh[kEwThreads / 32][8];
  constexpr int G = Gran<T>::bytes;
  const int c = blockIdx.y;
  const char* base = p.y.base + ((long long)c * p.y.plane + p.y.g_lo) * G;
  float s[8], q[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
  const long long stride = (long long)gridDim.x * kEwThreads;
  long long f = (long long)blockIdx.x * kEwThreads + threadIdx.x;
  for (; f + 3 * stride < p.body; f += 4 * stride) {  // four independent loads in flight
    float v[4][8];
#pragma unroll
    for (int i = 0; i < 4; ++i) load8<T>(base + (f + i * stride) * G, v[i]);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] += v[i][j], q[j] = fmaf(v[i][j], v[i][j], q[j]);
  }
  for (; f < p.body; f += stride) {
    float v[8];
    load8<T>(base + f * G, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] += v[j], q[j] = fmaf(v[j], v[j], q[j]);
  }
  block_reduce8(s, p.sum + c * 8, sh);
  block_reduce8(q, p.sumsq + c * 8, sh);
  // last block to finish turns the sums into scale / shift / running statistics (no separate finalize launch)
  __shared__ unsigned int last;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) last = atomicAdd(p.done, 1u) == gridDim.x * gridDim.y - 1 ? 1u : 0u;
  __syncthreads();
  if (last) {
    __threadfence();
    finalize_channel(p.fin, threadIdx.x);
  }
}

cudaError_t launch_stats(const StatsParams& p, int precision, cudaStream_t st) {
  dim3 grid(blocks_per_chunk(p.body, p.nchunks), p.nchunks);
  if (precision == 0)
    stats_kernel<__nv_bfloat16><<<grid, kEwThreads, 0, st>>>(p);
  else
    stats_kernel<float><<<grid, kEwThreads, 0, st>>>(p);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------------------------
// fused BatchNorm apply + activation (forward).  EW_SIGMOID_MUL loops over the tasks that share the feature map;
// EW_RELU_POOL iterates pooled positions (four full-resolution loads per thread and chunk).
// ---------------------------------------------------------------------------------------------
template <typename T, int MODE>
__global__ void __launch_bounds__(kEwThreads) ew_fwd_kernel(const __grid_constant__ EwParams p) {
  constexpr int G = Gran<T>::bytes;
  const int c = blockIdx.y;
  const int stride = gridDim.x * kEwThreads;
  const int body = (int)p.body;
  if (MODE == EW_RELU_POOL) {
    float sc[8], sh[8];
    ld_affine(p.bn.scale, c * 8, sc);
    ld_affine(p.bn.shift, c * 8, sh);
    const char* ybase = p.y.base + ((long long)c * p.y.plane + p.y.g_lo) * G;
    for (int f = blockIdx.x * kEwThreads + threadIdx.x; f < body; f += stride) {
      const Pos3 o = decode3(f, p.Pimgo, p.Wpo, p.magic_pimgo, p.magic_wpo);
      float m[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) m[j] = 0.f;
      if (o.h < p.Ho && o.w < p.Wo) {
        const int h0 = 2 * o.h, w0 = 2 * o.w;
        const bool hb = h0 + 1 < p.H, wb = w0 + 1 < p.W;
        const char* a = ybase + ((long long)o.n * p.Pimg + h0 * p.Wp + w0) * G;
        float v[4][8];
        load8<T>(a, v[0]);
        if (wb) load8<T>(a + G, v[1]);
        if (hb) load8<T>(a + (long long)p.Wp * G, v[2]);
        if (hb && wb) load8<T>(a + (long long)(p.Wp + 1) * G, v[3]);
#pragma unroll
        for (int j = 0; j < 8; ++j) {  // relu folded into the max with 0
          m[j] = fmaxf(m[j], fmaf(v[0][j], sc[j], sh[j]));
          if (wb) m[j] = fmaxf(m[j], fmaf(v[1][j], sc[j], sh[j]));
          if (hb) m[j] = fmaxf(m[j], fmaf(v[2][j], sc[j], sh[j]));
          if (hb && wb) m[j] = fmaxf(m[j], fmaf(v[3][j], sc[j], sh[j]));
        }
      }
      store8<T>(view_addr<T>(p.out, c, f, o.n, o.h, o.w), m);
    }
    return;
  }
  for (int t = 0; t < p.ntasks; ++t) {
    const int ch = t * p.nchunks + c;
    float sc[8], sh[8], s2[8], h2[8];
    ld_affine(p.bn.scale, ch * 8, sc);
    ld_affine(p.bn.shift, ch * 8, sh);
    if (MODE == EW_ADD_RELU_SC) {
      ld_affine(p.bn2.scale, ch * 8, s2);
      ld_affine(p.bn2.shift, ch * 8, h2);
    }
    const char* ybase = p.y.base + ((long long)ch * p.y.plane + p.y.g_lo) * G;
    const char* y2base = MODE == EW_ADD_RELU_SC ? p.y2.base + ((long long)ch * p.y2.plane + p.y2.g_lo) * G : nullptr;
#pragma unroll 2
    for (int f = blockIdx.x * kEwThreads + threadIdx.x; f < body; f += stride) {
      const Pos3 q = decode3(f, p.Pimg, p.Wp, p.magic_pimg, p.magic_wp);
      const bool valid = q.h < p.H && q.w < p.W;
      if (!valid && p.out.layout != LAY_P8) continue;
      float z[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) z[j] = 0.f;
      if (valid) {
        float v[8];
        load8<T>(ybase + (long long)f * G, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[j] = fmaf(v[j], sc[j], sh[j]);
        if (MODE == EW_RELU) {
#pragma unroll
          for (int j = 0; j < 8; ++j) z[j] = fmaxf(v[j], 0.f);
        } else if (MODE == EW_ADD_RELU_SC) {
          float v2[8];
          load8<T>(y2base + (long long)f * G, v2);
#pragma unroll
          for (int j = 0; j < 8; ++j) z[j] = fmaxf(v[j] + fmaf(v2[j], s2[j], h2[j]), 0.f);
        } else {
          float aux[8];
          load8<T>(view_addr<T>(p.aux, c, f, q.n, q.h, q.w), aux);
          if (MODE == EW_ADD_RELU_ID) {
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

template <typename T>
static cudaError_t launch_ew_fwd_t(const EwParams& p, cudaStream_t st) {
  dim3 grid(blocks_per_chunk(p.body, p.nchunks), p.nchunks);
  switch (p.mode) {
    case EW_RELU: ew_fwd_kernel<T, EW_RELU><<<grid, kEwThreads, 0, st>>>(p); break;
    case EW_ADD_RELU_ID: ew_fwd_kernel<T, EW_ADD_RELU_ID><<<grid, kEwThreads, 0, st>>>(p); break;
    case EW_ADD_RELU_SC: ew_fwd_kernel<T, EW_ADD_RELU_SC><<<grid, kEwThreads, 0, st>>>(p); break;
    case EW_SIGMOID_MUL: ew_fwd_kernel<T, EW_SIGMOID_MUL><<<grid, kEwThreads, 0, st>>>(p); break;
    case EW_RELU_POOL: ew_fwd_kernel<T, EW_RELU_POOL><<<grid, kEwThreads, 0, st>>>(p); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

cudaError_t launch_ew_fwd(const EwParams& p, int precision, cudaStream_t st) {
  return precision == 0 ? launch_ew_fwd_t<__nv_bfloat16>(p, st) : launch_ew_fwd_t<float>(p, st);
}

// ---------------------------------------------------------------------------------------------
// fused elementwise backward.  PASS 0: reductions sum(g), sum(g*y_hat); PASS 1: dy (and d aux).
// g = dz * activation'(.) is recomputed in both passes from dz and the saved raw tensors.
// PASS 1 uses dy = scale*g + A + B*y with A = -scale*(mean(g) - mean(g*y_hat)*mean*invstd) and
// B = -scale*mean(g*y_hat)*invstd (formed in double per thread, once), i.e. scale*(g - mean(g) - y_hat*mean(g*y_hat)).
// ---------------------------------------------------------------------------------------------
struct BwdConst {
  float sc[8], shf[8], a[8], b[8];  // PASS 0: a = mean, b = invstd;  PASS 1: a = A, b = B
};

template <int PASS>
__device__ __forceinline__ void load_bwd_const(const BnArrays& bn, int ch, float inv_n, BwdConst& k) {
  ld_affine(bn.scale, ch * 8, k.sc);
  ld_affine(bn.shift, ch * 8, k.shf);
  ld_affine(bn.mean, ch * 8, k.a);
  ld_affine(bn.invstd, ch * 8, k.b);
  if (PASS == 1) {
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const double cg = bn.red_g[ch * 8 + j] * (double)inv_n, cgy = bn.red_gy[ch * 8 + j] * (double)inv_n;
      const double mu = k.a[j], is = k.b[j], sc = k.sc[j];
      k.a[j] = (float)(-sc * (cg - cgy * mu * is));
      k.b[j] = (float)(-sc * cgy * is);
    }
  }
}

// d gamma / d beta of chunk c (all tasks) from the completed reductions: written once per chunk
__device__ __forceinline__ void write_bn_grads(const EwParams& p, const BnArrays& bn, int c, int ntask) {
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

template <typename T, int MODE, int PASS>
__global__ void __launch_bounds__(kEwThreads) ew_bwd_kernel(const __grid_constant__ EwParams p) {
  __shared__ float sh[kEwThreads / 32][8];
  constexpr int G = Gran<T>::bytes;
  const int c = blockIdx.y;
  const int stride = gridDim.x * kEwThreads;
  const int body = (int)p.body;
  const int first = blockIdx.x * kEwThreads + threadIdx.x;
  const BnArrays& bn = p.second ? p.bn2 : p.bn;     // the BatchNorm whose input gradient this launch produces
  const ActView& ym = p.second ? p.y2 : p.y;

  if (MODE == EW_RELU_POOL) {
    // thread = pooled position; the 2x2 window's four full-resolution positions belong to this thread alone
    BwdConst k;
    load_bwd_const<PASS>(bn, c, p.inv_n, k);
    float rg[8], rgy[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) rg[j] = rgy[j] = 0.f;
    const char* ybase = ym.base + ((long long)c * ym.plane + ym.g_lo) * G;
    char* dybase = p.dy.base + ((long long)c * p.dy.plane + p.dy.g_lo) * G;
    for (int f = first; f < body; f += stride) {
      const Pos3 o = decode3(f, p.Pimgo, p.Wpo, p.magic_pimgo, p.magic_wpo);
      if (!(o.h < p.Ho && o.w < p.Wo)) continue;
      float dzv[8];
      load8<T>(view_addr<T>(p.dz, c, f, o.n, o.h, o.w), dzv);
      const int h0 = 2 * o.h, w0 = 2 * o.w;
      const long long pos0 = (long long)o.n * p.Pimg + h0 * p.Wp + w0;
      bool inb[4];
      inb[0] = true, inb[1] = w0 + 1 < p.W, inb[2] = h0 + 1 < p.H, inb[3] = inb[1] && inb[2];
      float yv[4][8];
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if (inb[q]) load8<T>(ybase + (pos0 + (q >> 1) * p.Wp + (q & 1)) * G, yv[q]);
      int arg[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float best = -INFINITY;
        arg[j] = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          if (!inb[q]) continue;
          const float r = fmaxf(fmaf(yv[q][j], k.sc[j], k.shf[j]), 0.f);
          if (r > best) best = r, arg[j] = q;   // strict '>' : first maximum wins (scan order h, w)
        }
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (!inb[q]) continue;
        float dyv[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float v = fmaf(yv[q][j], k.sc[j], k.shf[j]);
          const float g = (arg[j] == q && v > 0.f) ? dzv[j] : 0.f;
          if (PASS == 0) {
            rg[j] += g, rgy[j] = fmaf(g, (yv[q][j] - k.a[j]) * k.b[j], rgy[j]);
          } else {
            dyv[j] = fmaf(k.sc[j], g, fmaf(k.b[j], yv[q][j], k.a[j]));
          }
        }
        if (PASS == 1) store8<T>(dybase + (pos0 + (q >> 1) * p.Wp + (q & 1)) * G, dyv);
      }
    }
    if (PASS == 0) {
      block_reduce8(rg, bn.red_g + c * 8, sh);
      block_reduce8(rgy, bn.red_gy + c * 8, sh);
    } else if (blockIdx.x == 0) {
      if (threadIdx.x == 0) write_bn_grads(p, bn, c, 1);
      // the tensor-core weight gradient reduces 16 positions per MMA: keep [body, round_up(body,16)) zero
      const long long full = (p.body / p.Pimgo) * p.Pimg;
      if (threadIdx.x < 16 && full + threadIdx.x < (full + 15) / 16 * 16) {
        float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        store8<T>(dybase + (full + threadIdx.x) * G, z);
      }
    }
    return;
  }

  const int body16 = (body + 15) / 16 * 16;
  const bool dz_flat = p.dz.layout == LAY_P8;
  if (PASS == 0 || MODE != EW_SIGMOID_MUL) {
    // tasks outer, positions inner: the constants of (task, chunk) stay in registers
    for (int t = 0; t < p.ntasks; ++t) {
      const int ch = t * p.nchunks + c;
      BwdConst k;
      load_bwd_const<PASS>(bn, ch, p.inv_n, k);
      float s2[8], h2[8], rg[8], rgy[8];
      if (MODE == EW_ADD_RELU_SC) {
        const BnArrays& bo = p.second ? p.bn : p.bn2;
        ld_affine(bo.scale, ch * 8, s2);
        ld_affine(bo.shift, ch * 8, h2);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) rg[j] = rgy[j] = 0.f;
      const char* ybase = ym.base + ((long long)ch * ym.plane + ym.g_lo) * G;
      const ActView& yo = p.second ? p.y : p.y2;
      const char* y2base = MODE == EW_ADD_RELU_SC ? yo.base + ((long long)ch * yo.plane + yo.g_lo) * G : nullptr;
      char* dybase = p.dy.base + ((long long)ch * p.dy.plane + p.dy.g_lo) * G;
      const char* dzbase = p.dz.base + ((long long)ch * p.dz.plane + p.dz.g_lo) * G;
      const int limit = PASS == 0 ? body : body16;
#pragma unroll 2
      for (int f = first; f < limit; f += stride) {
        // PASS 0 over flat tensors needs no coordinates: dz, y and aux are zero at pads, so g is zero there
        bool valid = true;
        Pos3 q{0, 0, 0};
        if (PASS == 1 || !dz_flat || (MODE != EW_RELU && MODE != EW_ADD_RELU_SC && p.aux.layout != LAY_P8)) {
          q = decode3(f < body ? f : 0, p.Pimg, p.Wp, p.magic_pimg, p.magic_wp);
          valid = f < body && q.h < p.H && q.w < p.W;
        }
        float dyv[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) dyv[j] = 0.f;
        if (valid) {
          float y[8], dzv[8], g[8];
          load8<T>(ybase + (long long)f * G, y);
          load8<T>(dz_flat ? dzbase + (long long)f * G : view_addr<T>(p.dz, ch, f, q.n, q.h, q.w), dzv);
          if (MODE == EW_RELU) {
#pragma unroll
            for (int j = 0; j < 8; ++j) g[j] = fmaf(y[j], k.sc[j], k.shf[j]) > 0.f ? dzv[j] : 0.f;
          } else if (MODE == EW_ADD_RELU_SC) {
            float y2[8];
            load8<T>(y2base + (long long)f * G, y2);
#pragma unroll
            for (int j = 0; j < 8; ++j)
              g[j] = (fmaf(y[j], k.sc[j], k.shf[j]) + fmaf(y2[j], s2[j], h2[j])) > 0.f ? dzv[j] : 0.f;
          } else {
            float aux[8];
            load8<T>(view_addr<T>(p.aux, c, f, q.n, q.h, q.w), aux);
            if (MODE == EW_ADD_RELU_ID) {
#pragma unroll
              for (int j = 0; j < 8; ++j) g[j] = (fmaf(y[j], k.sc[j], k.shf[j]) + aux[j]) > 0.f ? dzv[j] : 0.f;
            } else {  // EW_SIGMOID_MUL (PASS 0 only on this path)
#pragma unroll
              for (int j = 0; j < 8; ++j) {
                const float sg = 1.f / (1.f + expf(-fmaf(y[j], k.sc[j], k.shf[j])));
                g[j] = dzv[j] * aux[j] * sg * (1.f - sg);
              }
            }
          }
          if (PASS == 0) {
#pragma unroll
            for (int j = 0; j < 8; ++j) rg[j] += g[j], rgy[j] = fmaf(g[j], (y[j] - k.a[j]) * k.b[j], rgy[j]);
          } else {
#pragma unroll
            for (int j = 0; j < 8; ++j) dyv[j] = fmaf(k.sc[j], g[j], fmaf(k.b[j], y[j], k.a[j]));
            if (MODE == EW_ADD_RELU_ID && p.want_daux) {  // d residual = g
              char* a = view_addr<T>(p.daux, c, f, q.n, q.h, q.w);
              if (p.daux_accum) {
                float old[8];
                load8<T>(a, old);
#pragma unroll
                for (int j = 0; j < 8; ++j) g[j] += old[j];
              }
              store8<T>(a, g);
            }
          }
        } else if (PASS == 1 && MODE == EW_ADD_RELU_ID && p.want_daux && !p.daux_accum && f < body &&
                   p.daux.layout == LAY_P8) {
          store8<T>(view_addr<T>(p.daux, c, f, q.n, q.h, q.w), dyv);  // zero at the pads of a freshly written tensor
        }
        // zero at pads, and zero up to the next multiple of 16 positions (tensor-core wgrad granularity)
        if (PASS == 1) store8<T>(dybase + (long long)f * G, dyv);
      }
      if (PASS == 0) {
        block_reduce8(rg, bn.red_g + ch * 8, sh);
        block_reduce8(rgy, bn.red_gy + ch * 8, sh);
      }
    }
    if (PASS == 1 && first == 0) write_bn_grads(p, bn, c, p.ntasks);
    return;
  }

  // EW_SIGMOID_MUL, PASS 1: d feat sums over the tasks at one position, so the tasks are the inner loop and
  // their constants sit in shared memory
  constexpr int kMaxTasks = 4;
  __shared__ float cst[kMaxTasks][4][8];
  if (threadIdx.x < p.ntasks * 8) {
    const int t = threadIdx.x >> 3, j = threadIdx.x & 7, chn = (t * p.nchunks + c) * 8 + j;
    const double cg = bn.red_g[chn] * (double)p.inv_n, cgy = bn.red_gy[chn] * (double)p.inv_n;
    const double mu = bn.mean[chn], is = bn.invstd[chn], sc = bn.scale[chn];
    cst[t][0][j] = bn.scale[chn], cst[t][1][j] = bn.shift[chn];
    cst[t][2][j] = (float)(-sc * (cg - cgy * mu * is));
    cst[t][3][j] = (float)(-sc * cgy * is);
  }
  __syncthreads();
  for (int f = first; f < body16; f += stride) {
    const Pos3 q = decode3(f < body ? f : 0, p.Pimg, p.Wp, p.magic_pimg, p.magic_wp);
    const bool valid = f < body && q.h < p.H && q.w < p.W;
    float dacc[8], aux[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) dacc[j] = 0.f;
    if (valid) load8<T>(view_addr<T>(p.aux, c, f, q.n, q.h, q.w), aux);
    for (int t = 0; t < p.ntasks; ++t) {
      const int ch = t * p.nchunks + c;
      float dyv[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) dyv[j] = 0.f;
      if (valid) {
        float y[8], dzv[8];
        load8<T>(ym.base + ((long long)ch * ym.plane + ym.g_lo + f) * G, y);
        load8<T>(view_addr<T>(p.dz, ch, f, q.n, q.h, q.w), dzv);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float sg = 1.f / (1.f + expf(-fmaf(y[j], cst[t][0][j], cst[t][1][j])));
          const float g = dzv[j] * aux[j] * sg * (1.f - sg);
          dacc[j] = fmaf(dzv[j], sg, dacc[j]);
          dyv[j] = fmaf(cst[t][0][j], g, fmaf(cst[t][3][j], y[j], cst[t][2][j]));
        }
      }
      store8<T>(p.dy.base + ((long long)ch * p.dy.plane + p.dy.g_lo + f) * G, dyv);
    }
    if (p.want_daux && f < body && (valid || p.daux.layout == LAY_P8)) {
      char* a = view_addr<T>(p.daux, c, f, q.n, q.h, q.w);
      if (p.daux_accum) {
        if (!valid) continue;
        float old[8];
        load8<T>(a, old);
#pragma unroll
        for (int j = 0; j < 8; ++j) dacc[j] += old[j];
      }
      store8<T>(a, dacc);
    }
  }
  if (first == 0) write_bn_grads(p, bn, c, p.ntasks);
}

template <typename T, int PASS>
static cudaError_t launch_ew_bwd_t(const EwParams& p, cudaStream_t st) {
  if (p.mode == EW_SIGMOID_MUL && p.ntasks > 4) return cudaErrorInvalidValue;
  dim3 grid(blocks_per_chunk(p.body, p.nchunks), p.nchunks);
  switch (p.mode) {
    case EW_RELU: ew_bwd_kernel<T, EW_RELU, PASS><<<grid, kEwThreads, 0, st>>>(p); break;
    case EW_ADD_RELU_ID: ew_bwd_kernel<T, EW_ADD_RELU_ID, PASS><<<grid, kEwThreads, 0, st>>>(p); break;
    case EW_ADD_RELU_SC: ew_bwd_kernel<T, EW_ADD_RELU_SC, PASS><<<grid, kEwThreads, 0, st>>>(p); break;
    case EW_SIGMOID_MUL: ew_bwd_kernel<T, EW_SIGMOID_MUL, PASS><<<grid, kEwThreads, 0, st>>>(p); break;
    case EW_RELU_POOL: ew_bwd_kernel<T, EW_RELU_POOL, PASS><<<grid, kEwThreads, 0, st>>>(p); break;
    default: return cudaErrorInvalidValue;
  }
  return cudaGetLastError();
}

cudaError_t launch_ew_bwd(const EwParams& p, int pass, int precision, cudaStream_t st) {
  if (precision == 0)
    return pass == 0 ? launch_ew_bwd_t<__nv_bfloat16, 0>(p, st) : launch_ew_bwd_t<__nv_bfloat16, 1>(p, st);
  return pass == 0 ? launch_ew_bwd_t<float, 0>(p, st) : launch_ew_bwd_t<float, 1>(p, st);
}

// ---------------------------------------------------------------------------------------------
// stem weight gradient: dW[r*7+s][co] += sum_{n,h,w} x[n, 3h-2+r, 3w-2+s] * dy[n,co,h,w]   (modelA_MTL.py:74)
// A block walks output rows (n,h): the seven fp32 input rows and the transposed dy row are staged in shared
// memory; thread (r, co) keeps dW[r][0..6][co] in registers and consumes four output columns per step
// (16 consecutive input samples = four aligned 128-bit loads, one 128-bit load of dy).
// ---------------------------------------------------------------------------------------------
constexpr int kSwMaxW = 96;                    // output columns (multiple of 4)
constexpr int kSwXPitch = 3 * kSwMaxW + 12;    // input samples per staged row (2 left pad + 3*W + 6 + slack), 16 B multiple
constexpr int kSwDyPitch = kSwMaxW + 4;        // 100: conflict-free 128-bit reads across the 16 channels

template <typename T>
__global__ void __launch_bounds__(128) stem_wgrad_kernel(const __grid_constant__ StemWgradParams p, int rows_per_block,
                                                         int total_rows) {
  __shared__ __align__(16) float xs[7][kSwXPitch];
  __shared__ __align__(16) float dyT[16][kSwDyPitch];
  constexpr int G = Gran<T>::bytes;
  const int tid = threadIdx.x;
  const int W4 = (p.W + 3) & ~3;
  for (int i = tid; i < 7 * kSwXPitch; i += 128) (&xs[0][0])[i] = 0.f;
  for (int i = tid; i < 16 * kSwDyPitch; i += 128) (&dyT[0][0])[i] = 0.f;
  const int r = tid >> 4, co = tid & 15;
  float acc[7];
#pragma unroll
  for (int s = 0; s < 7; ++s) acc[s] = 0.f;
  const int row0 = blockIdx.x * rows_per_block, row1 = min(row0 + rows_per_block, total_rows);
  for (int row = row0; row < row1; ++row) {
    const int n = row / p.H, h = row - n * p.H;
    __syncthreads();
    for (int rr = 0; rr < 7; ++rr) {
      const int iy = 3 * h - 2 + rr;
      const bool in = iy >= 0 && iy < p.Hin;
      const float* src = p.x + ((size_t)n * p.Hin + (in ? iy : 0)) * p.Win;
      for (int ix = tid; ix < p.Win; ix += 128) xs[rr][ix + 2] = in ? src[ix] : 0.f;
    }
    for (int i = tid; i < 2 * p.W; i += 128) {
      const int c = i >= p.W ? 1 : 0, w = i - c * p.W;
      float v[8];
      load8<T>(p.dy.base + ((long long)c * p.dy.plane + p.dy.g_lo + (long long)n * p.Pimg + h * p.Wp + w) * G, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) dyT[c * 8 + j][w] = v[j];
    }
    __syncthreads();
    if (tid < 112) {
      for (int w0 = 0; w0 < W4; w0 += 4) {
        const float4 d = *reinterpret_cast<const float4*>(&dyT[co][w0]);
        const float4* xp = reinterpret_cast<const float4*>(&xs[r][3 * w0]);
        const float4 x0 = xp[0], x1 = xp[1], x2 = xp[2], x3 = xp[3];
        const float xv[16] = {x0.x, x0.y, x0.z, x0.w, x1.x, x1.y, x1.z, x1.w,
                              x2.x, x2.y, x2.z, x2.w, x3.x, x3.y, x3.z, x3.w};
        const float dv[4] = {d.x, d.y, d.z, d.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int s = 0; s < 7; ++s) acc[s] = fmaf(xv[3 * i + s], dv[i], acc[s]);
      }
    }
  }
  if (tid < 112) {
#pragma unroll
    for (int s = 0; s < 7; ++s) atomicAdd(p.dw + (r * 7 + s) * 16 + co, acc[s]);
  }
}

cudaError_t launch_stem_wgrad(const StemWgradParams& p, int precision, cudaStream_t st) {
  if (p.W > kSwMaxW || p.Win + 2 > kSwXPitch || 3 * ((p.W + 3) & ~3) + 4 > kSwXPitch) return cudaErrorInvalidValue;
  const int total_rows = (int)(p.body / p.Pimg) * p.H;
  const int want = 148 * 8;
  const int rows_per_block = std::max(1, (total_rows + want - 1) / want);
  const unsigned grid = (unsigned)((total_rows + rows_per_block - 1) / rows_per_block);
  if (precision == 0)
    stem_wgrad_kernel<__nv_bfloat16><<<grid, 128, 0, st>>>(p, rows_per_block, total_rows);
  else
    stem_wgrad_kernel<float><<<grid, 128, 0, st>>>(p, rows_per_block, total_rows);
  return cudaGetLastError();
}

}  // namespace mtldas
