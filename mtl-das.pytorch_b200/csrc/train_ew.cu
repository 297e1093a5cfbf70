// Streaming training kernels of the MTL-DAS hot path, written as chunk-resident grid-stride loops:
//   ew_fwd_kernel       batch statistics -> chunk barrier -> BatchNorm finalize -> apply + activation (ReLU / residual /
//                       sigmoid*feature / ReLU+maxpool), one launch
//   ew_bwd_kernel       activation' + BatchNorm backward: reductions -> chunk barrier -> input gradient, one launch
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
#include "ptx.cuh"
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

// Barrier among the gridDim.x blocks that own one chunk (arrival counter zeroed by the host-side memset of the
// sums).  Every block of the grid is resident (the launchers size the grid from the occupancy query), so the
// spin cannot starve a block that has not started; the clock bound turns a broken invariant into a trap, not a hang.
__device__ __forceinline__ void chunk_barrier(unsigned int* ctr, unsigned int nblocks) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    atomicAdd(ctr, 1u);
    const long long t0 = clock64();
    while (*reinterpret_cast<volatile unsigned int*>(ctr) < nblocks) {
      __nanosleep(40);
      if (clock64() - t0 > 8000000000ll) __trap();
    }
    __threadfence();
  }
  __syncthreads();
}

// per-thread partial sum / sum of squares of one chunk plane (pads and guards are zero: no position decoding)
template <typename T>
__device__ __forceinline__ void plane_sums(const char* base, long long body, long long first, long long stride,
                                           float* s, float* q) {
  constexpr int G = Gran<T>::bytes;
  long long f = first;
  for (; f + 3 * stride < body; f += 4 * stride) {  // four independent loads in flight
    float v[4][8];
#pragma unroll
    for (int i = 0; i < 4; ++i) load8<T>(base + (f + i * stride) * G, v[i]);
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] += v[i][j], q[j] = fmaf(v[i][j], v[i][j], q[j]);
  }
  for (; f < body; f += stride) {
    float v[8];
    load8<T>(base + f * G, v);
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] += v[j], q[j] = fmaf(v[j], v[j], q[j]);
  }
}

// scale / shift of channel c from the completed sums (every block of the chunk computes its own copy)
__device__ __forceinline__ void affine_of(const FinalizeParams& p, int c, float& scale, float& shift) {
  const double mean = __ldcg(p.bn.sum + c) / p.n;
  double var = __ldcg(p.bn.sumsq + c) / p.n - mean * mean;
  if (var < 0.0) var = 0.0;
  const double invstd = rsqrt(var + (double)p.eps);
  scale = 0.f, shift = 0.f;
  for (int s = 0; s < p.nseg; ++s) {
    const FinalizeSeg& g = p.seg[s];
    if (c >= g.ch_off && c < g.ch_off + g.count) {
      const int i = c - g.ch_off;
      scale = (float)((double)g.gamma[i] * invstd);
      shift = (float)((double)g.beta[i] - mean * (double)g.gamma[i] * invstd);
    }
  }
}

// grid.x of a fused (barrier) kernel: all blocks resident at once
template <typename K>
inline unsigned resident_blocks_per_chunk(K kernel, long long positions, int chunks) {
  int sms = 148, dev = 0, occ = 1;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kernel, kEwThreads, 0) != cudaSuccess || occ < 1) occ = 1;
  const long long need = (positions + kEwThreads - 1) / kEwThreads;
  const long long cap = std::max<long long>(1, (long long)sms * std::min(occ, 4) / chunks);
  return (unsigned)std::max<long long>(1, std::min(need, cap));
}

}  // namespace

// ---------------------------------------------------------------------------------------------
// fused forward of one BatchNorm + activation: batch statistics -> per-chunk barrier -> finalize -> apply.
// The raw tensor is read twice (the second time out of L2); scale / shift never leave the SM between the phases.
// EW_SIGMOID_MUL loops over the tasks that share the feature map; EW_ADD_RELU_SC normalises the main and the
// shortcut branch in one go; EW_RELU_POOL applies over pooled positions (four full-resolution loads per thread).
// ---------------------------------------------------------------------------------------------
constexpr int kMaxTasks = 2;  // tasks stacked on one feature map (MTL_Net: distance + event)

template <typename T, int MODE>
__global__ void __launch_bounds__(kEwThreads) ew_fwd_kernel(const __grid_constant__ EwParams p) {
  constexpr int G = Gran<T>::bytes;
  __shared__ float sh[kEwThreads / 32][8];
  __shared__ float aff[kMaxTasks][4][8];  // per task: scale, shift, scale2, shift2 of the chunk's 8 channels
  // programmatic dependent launch: let the next kernel's CTAs get scheduled, then wait for everything before us
  ptx::pdl_launch_dependents();
  ptx::pdl_grid_wait();
  const int c = blockIdx.y;
  const int stride = gridDim.x * kEwThreads;
  const int first = blockIdx.x * kEwThreads + threadIdx.x;
  const int body = (int)p.body;
  const int ntasks = MODE == EW_RELU_POOL ? 1 : p.ntasks;

  // ---- phase 1: statistics of this chunk (all tasks; both branches) ------------------------------
  for (int t = 0; t < ntasks; ++t) {
    const int ch = t * p.nchunks + c;
    float s[8], q[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
    plane_sums<T>(p.y.base + ((long long)ch * p.y.plane + p.y.g_lo) * G, p.body_full, first, stride, s, q);
    block_reduce8(s, p.fin.bn.sum + ch * 8, sh);
    block_reduce8(q, p.fin.bn.sumsq + ch * 8, sh);
    if (MODE == EW_ADD_RELU_SC) {
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j] = q[j] = 0.f;
      plane_sums<T>(p.y2.base + ((long long)ch * p.y2.plane + p.y2.g_lo) * G, p.body_full, first, stride, s, q);
      block_reduce8(s, p.fin2.bn.sum + ch * 8, sh);
      block_reduce8(q, p.fin2.bn.sumsq + ch * 8, sh);
    }
  }
  chunk_barrier(p.sync + c, gridDim.x);

  // ---- phase 2: finalize (every block for itself; block 0 of the chunk also publishes) -------------
  if (threadIdx.x < ntasks * 8) {
    const int t = threadIdx.x >> 3, j = threadIdx.x & 7, chn = (t * p.nchunks + c) * 8 + j;
    affine_of(p.fin, chn, aff[t][0][j], aff[t][1][j]);
    if (MODE == EW_ADD_RELU_SC) affine_of(p.fin2, chn, aff[t][2][j], aff[t][3][j]);
    if (blockIdx.x == 0) {
      finalize_channel(p.fin, chn);
      if (MODE == EW_ADD_RELU_SC) finalize_channel(p.fin2, chn);
    }
  }
  __syncthreads();

  // ---- phase 3: apply ------------------------------------------------------------------------------
  if (MODE == EW_RELU_POOL) {
    float sc[8], shf[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) sc[j] = aff[0][0][j], shf[j] = aff[0][1][j];
    const char* ybase = p.y.base + ((long long)c * p.y.plane + p.y.g_lo) * G;
    for (int f = first; f < body; f += stride) {
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
          m[j] = fmaxf(m[j], fmaf(v[0][j], sc[j], shf[j]));
          if (wb) m[j] = fmaxf(m[j], fmaf(v[1][j], sc[j], shf[j]));
          if (hb) m[j] = fmaxf(m[j], fmaf(v[2][j], sc[j], shf[j]));
          if (hb && wb) m[j] = fmaxf(m[j], fmaf(v[3][j], sc[j], shf[j]));
        }
      }
      store8<T>(view_addr<T>(p.out, c, f, o.n, o.h, o.w), m);
    }
    return;
  }
  for (int t = 0; t < ntasks; ++t) {
    const int ch = t * p.nchunks + c;
    float sc[8], shf[8], s2[8], h2[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      sc[j] = aff[t][0][j], shf[j] = aff[t][1][j];
      if (MODE == EW_ADD_RELU_SC) s2[j] = aff[t][2][j], h2[j] = aff[t][3][j];
    }
    const char* ybase = p.y.base + ((long long)ch * p.y.plane + p.y.g_lo) * G;
    const char* y2base = MODE == EW_ADD_RELU_SC ? p.y2.base + ((long long)ch * p.y2.plane + p.y2.g_lo) * G : nullptr;
    constexpr int U = Batch<T>::n;
    for (int f0 = first; f0 < body; f0 += U * stride) {
      Raw<T> ry[U], ro[U];
      Pos3 q[U];
      bool valid[U];
#pragma unroll
      for (int i = 0; i < U; ++i) {  // issue every load of the batch before the first use
        const int f = f0 + i * stride;
        q[i] = decode3(f < body ? f : 0, p.Pimg, p.Wp, p.magic_pimg, p.magic_wp);
        valid[i] = f < body && q[i].h < p.H && q[i].w < p.W;
        if (valid[i]) {
          load_raw(ybase + (long long)f * G, ry[i]);
          if (MODE == EW_ADD_RELU_SC) load_raw(y2base + (long long)f * G, ro[i]);
          if (MODE == EW_ADD_RELU_ID || MODE == EW_SIGMOID_MUL)
            load_raw(view_addr<T>(p.aux, c, f, q[i].n, q[i].h, q[i].w), ro[i]);
        }
      }
#pragma unroll
      for (int i = 0; i < U; ++i) {
        const int f = f0 + i * stride;
        if (f >= body || (!valid[i] && p.out.layout != LAY_P8)) continue;
        float z[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) z[j] = 0.f;
        if (valid[i]) {
          float v[8], o[8];
          unpack(ry[i], v);
          if (MODE != EW_RELU) unpack(ro[i], o);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float bn = fmaf(v[j], sc[j], shf[j]);
            if (MODE == EW_RELU) z[j] = fmaxf(bn, 0.f);
            else if (MODE == EW_ADD_RELU_SC) z[j] = fmaxf(bn + fmaf(o[j], s2[j], h2[j]), 0.f);
            else if (MODE == EW_ADD_RELU_ID) z[j] = fmaxf(bn + o[j], 0.f);
            else z[j] = o[j] / (1.f + expf(-bn));  // EW_SIGMOID_MUL
          }
        }
        store8<T>(view_addr<T>(p.out, ch, f, q[i].n, q[i].h, q[i].w), z);
      }
    }
  }
}


// launch with the programmatic-stream-serialization attribute (the kernels above call griddepcontrol.wait first)
template <typename K>
static cudaError_t launch_pdl(K kernel, dim3 grid, const EwParams& p, cudaStream_t st) {
  static int use_pdl = -1;
  if (use_pdl < 0) {
    const char* e = getenv("MTLDAS_NO_PDL");
    use_pdl = (e && e[0] == '1') ? 0 : 1;
  }
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid, cfg.blockDim = dim3(kEwThreads), cfg.dynamicSmemBytes = 0, cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr, cfg.numAttrs = use_pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, p);
}

template <typename T, int MODE>
static cudaError_t launch_ew_fwd_m(const EwParams& p, cudaStream_t st) {
  static unsigned per_sm_cache = 0;  // occupancy of this instantiation
  (void)per_sm_cache;
  dim3 grid(resident_blocks_per_chunk(ew_fwd_kernel<T, MODE>, std::max(p.body, p.body_full), p.nchunks), p.nchunks);
  return launch_pdl(ew_fwd_kernel<T, MODE>, grid, p, st);
}

template <typename T>
static cudaError_t launch_ew_fwd_t(const EwParams& p, cudaStream_t st) {
  if (p.ntasks > kMaxTasks || !p.sync) return cudaErrorInvalidValue;
  switch (p.mode) {
    case EW_RELU: return launch_ew_fwd_m<T, EW_RELU>(p, st);
    case EW_ADD_RELU_ID: return launch_ew_fwd_m<T, EW_ADD_RELU_ID>(p, st);
    case EW_ADD_RELU_SC: return launch_ew_fwd_m<T, EW_ADD_RELU_SC>(p, st);
    case EW_SIGMOID_MUL: return launch_ew_fwd_m<T, EW_SIGMOID_MUL>(p, st);
    case EW_RELU_POOL: return launch_ew_fwd_m<T, EW_RELU_POOL>(p, st);
    default: return cudaErrorInvalidValue;
  }
}

cudaError_t launch_ew_fwd(const EwParams& p, int precision, cudaStream_t st) {
  return precision == 0 ? launch_ew_fwd_t<__nv_bfloat16>(p, st) : launch_ew_fwd_t<float>(p, st);
}

// ---------------------------------------------------------------------------------------------
// fused elementwise backward of one BatchNorm + activation:
//   phase 1  reductions sum(g), sum(g*y_hat)            (g = dz * activation'(.), recomputed from the saved raw tensors)
//   barrier  among the blocks of the chunk
//   phase 2  dy = scale*g + A + B*y with A = -scale*(mean(g) - mean(g*y_hat)*mean*invstd), B = -scale*mean(g*y_hat)*invstd
//            (formed in double, once per thread), i.e. scale*(g - mean(g) - y_hat*mean(g*y_hat)); d aux where the
//            activation has a second input; d gamma / d beta written by block 0 of the chunk.
// EW_ADD_RELU_SC differentiates the main and the shortcut BatchNorm together (they share g).
// ---------------------------------------------------------------------------------------------
struct BwdConst {
  float sc[8], shf[8], a[8], b[8];  // phase 1: a = mean, b = invstd;  phase 2: a = A, b = B
};

__device__ __forceinline__ void load_fwd_const(const BnArrays& bn, int ch, BwdConst& k) {
  ld_affine(bn.scale, ch * 8, k.sc);
  ld_affine(bn.shift, ch * 8, k.shf);
  ld_affine(bn.mean, ch * 8, k.a);
  ld_affine(bn.invstd, ch * 8, k.b);
}

// a = mean, b = invstd  ->  a = A, b = B from the completed reductions (L2 reads: other blocks' atomics)
__device__ __forceinline__ void to_grad_const(const BnArrays& bn, int ch, float inv_n, BwdConst& k) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const double cg = __ldcg(bn.red_g + ch * 8 + j) * (double)inv_n, cgy = __ldcg(bn.red_gy + ch * 8 + j) * (double)inv_n;
    const double mu = k.a[j], is = k.b[j], sc = k.sc[j];
    k.a[j] = (float)(-sc * (cg - cgy * mu * is));
    k.b[j] = (float)(-sc * cgy * is);
  }
}

// d gamma / d beta of chunk c (all tasks) from the completed reductions: written once per chunk
__device__ __forceinline__ void write_bn_grads(const BnArrays& bn, const BnGradSeg* gseg, int ngseg, int nchunks, int c,
                                               int ntask) {
  for (int t = 0; t < ntask; ++t)
    for (int j = 0; j < 8; ++j) {
      const int ch = (t * nchunks + c) * 8 + j;
      for (int s = 0; s < ngseg; ++s) {
        const BnGradSeg& g = gseg[s];
        if (ch >= g.ch_off && ch < g.ch_off + g.count) {
          g.dgamma[ch - g.ch_off] = (float)__ldcg(bn.red_gy + ch);
          g.dbeta[ch - g.ch_off] = (float)__ldcg(bn.red_g + ch);
        }
      }
    }
}

// g of one position for the non-pool modes (y, dz, and the mode's second input already loaded)
template <int MODE>
__device__ __forceinline__ void act_grad(const float* y, const float* dzv, const float* other, const BwdConst& k,
                                         const float* s2, const float* h2, float* g, float* sgm) {
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const float v = fmaf(y[j], k.sc[j], k.shf[j]);
    if (MODE == EW_RELU) {
      g[j] = v > 0.f ? dzv[j] : 0.f;
    } else if (MODE == EW_ADD_RELU_SC) {
      g[j] = (v + fmaf(other[j], s2[j], h2[j])) > 0.f ? dzv[j] : 0.f;
    } else if (MODE == EW_ADD_RELU_ID) {
      g[j] = (v + other[j]) > 0.f ? dzv[j] : 0.f;
    } else {  // EW_SIGMOID_MUL
      const float sg = 1.f / (1.f + expf(-v));
      g[j] = dzv[j] * other[j] * sg * (1.f - sg);
      sgm[j] = sg;
    }
  }
}

// One pass of the ReLU + max-pool backward over the pooled positions of chunk c.  thread = pooled position; the 2x2
// window's four full-resolution positions belong to this thread alone.  Two pooled positions (ten 16-byte loads)
// are in flight per thread.  PASS 0 accumulates sum(g), sum(g*y_hat); PASS 1 writes dy.
template <typename T, int PASS>
__device__ __forceinline__ void pool_bwd_pass(const EwParams& p, int c, const BwdConst& k, int first, int stride,
                                              const char* ybase, char* dybase, float* rg, float* rgy) {
  constexpr int G = Gran<T>::bytes;
  constexpr int U = Batch<T>::n / 2;  // bf16: two pooled positions per iteration, fp32: one (registers)
  const int body = (int)p.body;
  for (int f0 = first; f0 < body; f0 += U * stride) {
    Raw<T> rdz[U], ry[U][4];
    long long pos0[U];
    bool ok[U], inb[U][4];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const int f = f0 + u * stride;
      const Pos3 o = decode3(f < body ? f : 0, p.Pimgo, p.Wpo, p.magic_pimgo, p.magic_wpo);
      ok[u] = f < body && o.h < p.Ho && o.w < p.Wo;
      const int h0 = 2 * o.h, w0 = 2 * o.w;
      pos0[u] = (long long)o.n * p.Pimg + h0 * p.Wp + w0;
      inb[u][0] = ok[u], inb[u][1] = ok[u] && w0 + 1 < p.W, inb[u][2] = ok[u] && h0 + 1 < p.H;
      inb[u][3] = inb[u][1] && inb[u][2];
      if (ok[u]) load_raw(view_addr<T>(p.dz, c, f, o.n, o.h, o.w), rdz[u]);
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if (inb[u][q]) load_raw(ybase + (pos0[u] + (q >> 1) * p.Wp + (q & 1)) * G, ry[u][q]);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      if (!ok[u]) continue;
      float dzv[8], yv[4][8];
      unpack(rdz[u], dzv);
#pragma unroll
      for (int q = 0; q < 4; ++q)
        if (inb[u][q]) unpack(ry[u][q], yv[q]);
      int arg[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float best = -INFINITY;
        arg[j] = 0;
#pragma unroll
        for (int q = 0; q < 4; ++q) {
          if (!inb[u][q]) continue;
          const float r = fmaxf(fmaf(yv[q][j], k.sc[j], k.shf[j]), 0.f);
          if (r > best) best = r, arg[j] = q;   // strict '>' : first maximum wins (scan order h, w)
        }
      }
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        if (!inb[u][q]) continue;
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
        if (PASS == 1) store8<T>(dybase + (pos0[u] + (q >> 1) * p.Wp + (q & 1)) * G, dyv);
      }
    }
  }
}

template <typename T, int MODE>
__global__ void __launch_bounds__(kEwThreads, 2) ew_bwd_kernel(const __grid_constant__ EwParams p) {
  __shared__ float sh[kEwThreads / 32][8];
  constexpr int G = Gran<T>::bytes;
  ptx::pdl_launch_dependents();
  ptx::pdl_grid_wait();
  const int c = blockIdx.y;
  const int stride = gridDim.x * kEwThreads;
  const int body = (int)p.body;
  const int first = blockIdx.x * kEwThreads + threadIdx.x;

  if (MODE == EW_RELU_POOL) {
    BwdConst k;
    load_fwd_const(p.bn, c, k);
    const char* ybase = p.y.base + ((long long)c * p.y.plane + p.y.g_lo) * G;
    char* dybase = p.dy.base + ((long long)c * p.dy.plane + p.dy.g_lo) * G;
    float rg[8], rgy[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) rg[j] = rgy[j] = 0.f;
    pool_bwd_pass<T, 0>(p, c, k, first, stride, ybase, dybase, rg, rgy);
    block_reduce8(rg, p.bn.red_g + c * 8, sh);
    block_reduce8(rgy, p.bn.red_gy + c * 8, sh);
    chunk_barrier(p.sync + c, gridDim.x);
    to_grad_const(p.bn, c, p.inv_n, k);
    pool_bwd_pass<T, 1>(p, c, k, first, stride, ybase, dybase, rg, rgy);
    if (blockIdx.x == 0) {
      if (threadIdx.x == 0) write_bn_grads(p.bn, p.gseg, p.ngseg, p.nchunks, c, 1);
      // the tensor-core weight gradient reduces 16 positions per MMA: keep [body, round_up(body,16)) zero
      if (threadIdx.x < 16 && p.body_full + threadIdx.x < (p.body_full + 15) / 16 * 16) {
        float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        store8<T>(dybase + (p.body_full + threadIdx.x) * G, z);
      }
    }
    return;
  }

  const int body16 = (body + 15) / 16 * 16;
  const bool dz_flat = p.dz.layout == LAY_P8;
  const bool need_pos = !dz_flat || ((MODE == EW_ADD_RELU_ID || MODE == EW_SIGMOID_MUL) && p.aux.layout != LAY_P8);
  constexpr int U = MODE == EW_RELU ? Batch<T>::n : Batch<T>::n / 2;  // three-input modes: smaller batches (registers)
  constexpr bool kOther = MODE != EW_RELU;   // a second saved input (shortcut raw tensor / residual / feature map)

  // ---- phase 1: reductions (tasks outer; the constants of (task, chunk) stay in registers) -------------------
  for (int t = 0; t < p.ntasks; ++t) {
    const int ch = t * p.nchunks + c;
    BwdConst k;
    load_fwd_const(p.bn, ch, k);
    float s2[8], h2[8], m2[8], i2[8], rg[8], rgy[8], rgy2[8];
    if (MODE == EW_ADD_RELU_SC) {
      ld_affine(p.bn2.scale, ch * 8, s2);
      ld_affine(p.bn2.shift, ch * 8, h2);
      ld_affine(p.bn2.mean, ch * 8, m2);
      ld_affine(p.bn2.invstd, ch * 8, i2);
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) rg[j] = rgy[j] = rgy2[j] = 0.f;
    const char* ybase = p.y.base + ((long long)ch * p.y.plane + p.y.g_lo) * G;
    const char* y2base = MODE == EW_ADD_RELU_SC ? p.y2.base + ((long long)ch * p.y2.plane + p.y2.g_lo) * G : nullptr;
    const char* dzbase = p.dz.base + ((long long)ch * p.dz.plane + p.dz.g_lo) * G;
    for (int f0 = first; f0 < body; f0 += U * stride) {
      Raw<T> ry[U], rdz[U], ro[U];
      bool ok[U];
#pragma unroll
      for (int i = 0; i < U; ++i) {  // issue every load of the batch before the first use
        const int f = f0 + i * stride;
        ok[i] = f < body;
        // over flat tensors no coordinates are needed: dz, y and aux are zero at pads, so g is zero there
        Pos3 q{0, 0, 0};
        if (need_pos) {
          q = decode3(ok[i] ? f : 0, p.Pimg, p.Wp, p.magic_pimg, p.magic_wp);
          ok[i] = ok[i] && q.h < p.H && q.w < p.W;
        }
        if (ok[i]) {
          load_raw(ybase + (long long)f * G, ry[i]);
          load_raw(dz_flat ? dzbase + (long long)f * G : view_addr<T>(p.dz, ch, f, q.n, q.h, q.w), rdz[i]);
          if (MODE == EW_ADD_RELU_SC) load_raw(y2base + (long long)f * G, ro[i]);
          if (MODE == EW_ADD_RELU_ID || MODE == EW_SIGMOID_MUL) load_raw(view_addr<T>(p.aux, c, f, q.n, q.h, q.w), ro[i]);
        }
      }
#pragma unroll
      for (int i = 0; i < U; ++i) {
        if (!ok[i]) continue;
        float y[8], dzv[8], other[8], g[8], sg[8];
        unpack(ry[i], y);
        unpack(rdz[i], dzv);
        if (kOther) unpack(ro[i], other);
        act_grad<MODE>(y, dzv, other, k, s2, h2, g, sg);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          rg[j] += g[j], rgy[j] = fmaf(g[j], (y[j] - k.a[j]) * k.b[j], rgy[j]);
          if (MODE == EW_ADD_RELU_SC) rgy2[j] = fmaf(g[j], (other[j] - m2[j]) * i2[j], rgy2[j]);
        }
      }
    }
    if (MODE == EW_ADD_RELU_SC) {
      float rg2[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) rg2[j] = rg[j];
      block_reduce8(rg2, p.bn2.red_g + ch * 8, sh);
      block_reduce8(rgy2, p.bn2.red_gy + ch * 8, sh);
    }
    block_reduce8(rg, p.bn.red_g + ch * 8, sh);
    block_reduce8(rgy, p.bn.red_gy + ch * 8, sh);
  }
  chunk_barrier(p.sync + c, gridDim.x);
  if (first == 0) {
    write_bn_grads(p.bn, p.gseg, p.ngseg, p.nchunks, c, p.ntasks);
    if (MODE == EW_ADD_RELU_SC) write_bn_grads(p.bn2, p.gseg2, p.ngseg2, p.nchunks, c, p.ntasks);
  }

  // ---- phase 2: gradients ------------------------------------------------------------------------------------
  if (MODE != EW_SIGMOID_MUL) {
    for (int t = 0; t < p.ntasks; ++t) {
      const int ch = t * p.nchunks + c;
      BwdConst k, k2;
      load_fwd_const(p.bn, ch, k);
      to_grad_const(p.bn, ch, p.inv_n, k);
      if (MODE == EW_ADD_RELU_SC) {
        load_fwd_const(p.bn2, ch, k2);
        to_grad_const(p.bn2, ch, p.inv_n, k2);
      }
      const char* ybase = p.y.base + ((long long)ch * p.y.plane + p.y.g_lo) * G;
      const char* y2base = MODE == EW_ADD_RELU_SC ? p.y2.base + ((long long)ch * p.y2.plane + p.y2.g_lo) * G : nullptr;
      const char* dzbase = p.dz.base + ((long long)ch * p.dz.plane + p.dz.g_lo) * G;
      char* dybase = p.dy.base + ((long long)ch * p.dy.plane + p.dy.g_lo) * G;
      char* dy2base = MODE == EW_ADD_RELU_SC ? p.dy2.base + ((long long)ch * p.dy2.plane + p.dy2.g_lo) * G : nullptr;
      const bool rmw = MODE == EW_ADD_RELU_ID && p.want_daux && p.daux_accum;
      for (int f0 = first; f0 < body16; f0 += U * stride) {
        Raw<T> ry[U], rdz[U], ro[U], rold[U];
        Pos3 q[U];
        bool valid[U];
#pragma unroll
        for (int i = 0; i < U; ++i) {
          const int f = f0 + i * stride;
          q[i] = decode3(f < body ? f : 0, p.Pimg, p.Wp, p.magic_pimg, p.magic_wp);
          valid[i] = f < body && q[i].h < p.H && q[i].w < p.W;
          if (valid[i]) {
            load_raw(ybase + (long long)f * G, ry[i]);
            load_raw(dz_flat ? dzbase + (long long)f * G : view_addr<T>(p.dz, ch, f, q[i].n, q[i].h, q[i].w), rdz[i]);
            if (MODE == EW_ADD_RELU_SC) load_raw(y2base + (long long)f * G, ro[i]);
            if (MODE == EW_ADD_RELU_ID) load_raw(view_addr<T>(p.aux, c, f, q[i].n, q[i].h, q[i].w), ro[i]);
            if (rmw) load_raw(view_addr<T>(p.daux, c, f, q[i].n, q[i].h, q[i].w), rold[i]);
          }
        }
#pragma unroll
        for (int i = 0; i < U; ++i) {
          const int f = f0 + i * stride;
          if (f >= body16) continue;
          float dyv[8], dyv2[8];
#pragma unroll
          for (int j = 0; j < 8; ++j) dyv[j] = dyv2[j] = 0.f;
          if (valid[i]) {
            float y[8], dzv[8], other[8], g[8], sg[8];
            unpack(ry[i], y);
            unpack(rdz[i], dzv);
            if (kOther) unpack(ro[i], other);
            // the sign test needs the forward scale / shift: k keeps them (only a / b changed meaning)
            act_grad<MODE>(y, dzv, other, k, k2.sc, k2.shf, g, sg);
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              dyv[j] = fmaf(k.sc[j], g[j], fmaf(k.b[j], y[j], k.a[j]));
              if (MODE == EW_ADD_RELU_SC) dyv2[j] = fmaf(k2.sc[j], g[j], fmaf(k2.b[j], other[j], k2.a[j]));
            }
            if (MODE == EW_ADD_RELU_ID && p.want_daux) {  // d residual = g
              if (rmw) {
                float old[8];
                unpack(rold[i], old);
#pragma unroll
                for (int j = 0; j < 8; ++j) g[j] += old[j];
              }
              store8<T>(view_addr<T>(p.daux, c, f, q[i].n, q[i].h, q[i].w), g);
            }
          } else if (MODE == EW_ADD_RELU_ID && p.want_daux && !p.daux_accum && f < body && p.daux.layout == LAY_P8) {
            store8<T>(view_addr<T>(p.daux, c, f, q[i].n, q[i].h, q[i].w), dyv);  // zero at the pads of a fresh tensor
          }
          // zero at pads, and zero up to the next multiple of 16 positions (tensor-core wgrad granularity)
          store8<T>(dybase + (long long)f * G, dyv);
          if (MODE == EW_ADD_RELU_SC) store8<T>(dy2base + (long long)f * G, dyv2);
        }
      }
    }
    return;
  }

  // EW_SIGMOID_MUL: d feat sums over the tasks at one position, so the tasks are the inner loop and their
  // constants sit in shared memory
  __shared__ float cst[kMaxTasks][4][8];
  if (threadIdx.x < p.ntasks * 8) {
    const int t = threadIdx.x >> 3, j = threadIdx.x & 7, chn = (t * p.nchunks + c) * 8 + j;
    const double cg = __ldcg(p.bn.red_g + chn) * (double)p.inv_n, cgy = __ldcg(p.bn.red_gy + chn) * (double)p.inv_n;
    const double mu = p.bn.mean[chn], is = p.bn.invstd[chn], sc = p.bn.scale[chn];
    cst[t][0][j] = p.bn.scale[chn], cst[t][1][j] = p.bn.shift[chn];
    cst[t][2][j] = (float)(-sc * (cg - cgy * mu * is));
    cst[t][3][j] = (float)(-sc * cgy * is);
  }
  __syncthreads();
  constexpr int V = Batch<T>::n / 2;  // every position carries 1 + 2*ntasks (+1) loads already
  const bool rmw = p.want_daux && p.daux_accum;
  for (int f0 = first; f0 < body16; f0 += V * stride) {
    Raw<T> raux[V], rold[V], ry[V][kMaxTasks], rdz[V][kMaxTasks];
    Pos3 q[V];
    bool valid[V];
#pragma unroll
    for (int i = 0; i < V; ++i) {
      const int f = f0 + i * stride;
      q[i] = decode3(f < body ? f : 0, p.Pimg, p.Wp, p.magic_pimg, p.magic_wp);
      valid[i] = f < body && q[i].h < p.H && q[i].w < p.W;
      if (valid[i]) {
        load_raw(view_addr<T>(p.aux, c, f, q[i].n, q[i].h, q[i].w), raux[i]);
        if (rmw) load_raw(view_addr<T>(p.daux, c, f, q[i].n, q[i].h, q[i].w), rold[i]);
#pragma unroll
        for (int t = 0; t < kMaxTasks; ++t)
          if (t < p.ntasks) {
            const int ch = t * p.nchunks + c;
            load_raw(p.y.base + ((long long)ch * p.y.plane + p.y.g_lo + f) * G, ry[i][t]);
            load_raw(view_addr<T>(p.dz, ch, f, q[i].n, q[i].h, q[i].w), rdz[i][t]);
          }
      }
    }
#pragma unroll
    for (int i = 0; i < V; ++i) {
      const int f = f0 + i * stride;
      if (f >= body16) continue;
      float dacc[8], aux[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) dacc[j] = 0.f;
      if (valid[i]) unpack(raux[i], aux);
#pragma unroll
      for (int t = 0; t < kMaxTasks; ++t) {
        if (t >= p.ntasks) continue;
        const int ch = t * p.nchunks + c;
        float dyv[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) dyv[j] = 0.f;
        if (valid[i]) {
          float y[8], dzv[8];
          unpack(ry[i][t], y);
          unpack(rdz[i][t], dzv);
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
      if (p.want_daux && f < body && (valid[i] || p.daux.layout == LAY_P8)) {
        if (p.daux_accum) {
          if (!valid[i]) continue;
          float old[8];
          unpack(rold[i], old);
#pragma unroll
          for (int j = 0; j < 8; ++j) dacc[j] += old[j];
        }
        store8<T>(view_addr<T>(p.daux, c, f, q[i].n, q[i].h, q[i].w), dacc);
      }
    }
  }
}

template <typename T, int MODE>
static cudaError_t launch_ew_bwd_m(const EwParams& p, cudaStream_t st) {
  dim3 grid(resident_blocks_per_chunk(ew_bwd_kernel<T, MODE>, std::max(p.body, p.body_full), p.nchunks), p.nchunks);
  return launch_pdl(ew_bwd_kernel<T, MODE>, grid, p, st);
}

template <typename T>
static cudaError_t launch_ew_bwd_t(const EwParams& p, cudaStream_t st) {
  if (p.ntasks > kMaxTasks || !p.sync) return cudaErrorInvalidValue;
  switch (p.mode) {
    case EW_RELU: return launch_ew_bwd_m<T, EW_RELU>(p, st);
    case EW_ADD_RELU_ID: return launch_ew_bwd_m<T, EW_ADD_RELU_ID>(p, st);
    case EW_ADD_RELU_SC: return launch_ew_bwd_m<T, EW_ADD_RELU_SC>(p, st);
    case EW_SIGMOID_MUL: return launch_ew_bwd_m<T, EW_SIGMOID_MUL>(p, st);
    case EW_RELU_POOL: return launch_ew_bwd_m<T, EW_RELU_POOL>(p, st);
    default: return cudaErrorInvalidValue;
  }
}

cudaError_t launch_ew_bwd(const EwParams& p, int precision, cudaStream_t st) {
  return precision == 0 ? launch_ew_bwd_t<__nv_bfloat16>(p, st) : launch_ew_bwd_t<float>(p, st);
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
