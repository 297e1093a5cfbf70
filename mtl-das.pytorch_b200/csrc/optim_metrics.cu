// Neighbours of the hot path that the reference runs as torch library calls right after model(x):
//   adam_kernel     torch.optim.Adam(model.parameters(), lr, weight_decay) .step()   /root/reference/utils.py:133-134,374
//                   (L2 weight decay folded into the gradient, betas (0.9, 0.999), eps 1e-8, bias correction), one launch
//                   over the FLAT parameter buffer (1 136 224 fp32 for model A) instead of a sweep over 126 tensors;
//   metrics_kernel  the validation bookkeeping of utils.py:274-291,313: NLL sums, torch.max(out, 1)[1] predictions,
//                   correct counts and confusion matrices accumulated ON THE DEVICE (the reference converts every
//                   prediction to a Python int: one device sync per sample).
// Both are pure streaming kernels (HBM / latency bound): 128-bit accesses, no shared state beyond the counters.
#include <cuda_runtime.h>

#include "../../include/mtldas.h"
#include "plan.h"

namespace mtldas {

struct AdamArgs {
  float* p;
  const float* g;
  float* m;
  float* v;
  long long n;
  float lr_over_bc1;     // lr / (1 - beta1^t)
  float inv_sqrt_bc2;    // 1 / sqrt(1 - beta2^t)
  float beta1, beta2, eps, weight_decay;
};

__device__ __forceinline__ void adam_one(float& p, float g, float& m, float& v, const AdamArgs& a) {
  g = fmaf(a.weight_decay, p, g);                         // grad + wd * param   (L2 penalty, not AdamW)
  m = fmaf(1.f - a.beta1, g - m, m);                      // exp_avg.lerp_(grad, 1 - beta1)
  v = fmaf(1.f - a.beta2, g * g, a.beta2 * v);            // exp_avg_sq = beta2 * v + (1 - beta2) * g^2
  const float denom = sqrtf(v) * a.inv_sqrt_bc2 + a.eps;  // sqrt(v) / sqrt(bias_correction2) + eps
  p -= a.lr_over_bc1 * (m / denom);
}

__global__ void __launch_bounds__(256) adam_kernel(const AdamArgs a) {
  const long long n4 = a.n >> 2;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    float4 p = reinterpret_cast<float4*>(a.p)[i], m = reinterpret_cast<float4*>(a.m)[i], v = reinterpret_cast<float4*>(a.v)[i];
    const float4 g = reinterpret_cast<const float4*>(a.g)[i];
    adam_one(p.x, g.x, m.x, v.x, a);
    adam_one(p.y, g.y, m.y, v.y, a);
    adam_one(p.z, g.z, m.z, v.z, a);
    adam_one(p.w, g.w, m.w, v.w, a);
    reinterpret_cast<float4*>(a.p)[i] = p, reinterpret_cast<float4*>(a.m)[i] = m, reinterpret_cast<float4*>(a.v)[i] = v;
  }
  for (long long i = (n4 << 2) + (long long)blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += stride)
    adam_one(a.p[i], a.g[i], a.m[i], a.v[i], a);
}

// ---------------------------------------------------------------------------------------------------------------
// Gradient all-reduce + Adam + parameter broadcast in ONE kernel over NVLink / NVSwitch multicast:
// every rank owns the slice [rank * n / world, (rank + 1) * n / world) of the flat buffers.  For its slice it
//   g   = multimem.ld_reduce.add( grads of ALL ranks )   (summed inside the switch, one load per 4 floats) * 1 / world
//   p   = Adam(p, g)                                      (exp_avg / exp_avg_sq exist for the owned slice only)
//   multimem.st( params of ALL ranks ) = p                (the updated slice lands in every replica)
// replacing ncclAllReduce (4.54 MB) + the Adam launch.  Every element is updated once, by its owner, and broadcast, so
// the replicas stay bit-identical by construction.  No waiting inside the kernel: the host brackets it with the
// symmetric-memory barrier (all gradients written before, all slices landed after).
// ---------------------------------------------------------------------------------------------------------------
struct ReduceAdamArgs {
  const float* mc_grads;   // multicast address of the flat gradient buffer
  float* mc_params;        // multicast address of the flat parameter buffer
  const float* params;     // this rank's own (unicast) parameter buffer
  float* m;                // exp_avg of the owned slice   [n_slice]
  float* v;                // exp_avg_sq of the owned slice [n_slice]
  long long first, n_slice;  // owned slice, in floats (multiples of 4)
  float inv_world;
  AdamArgs a;              // p/g/m/v/n unused; the scalars are
};

__device__ __forceinline__ float4 multimem_ld_reduce_add(const float* mc_addr) {
  float4 r;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w)
               : "l"(mc_addr)
               : "memory");
  return r;
}
__device__ __forceinline__ void multimem_st(float* mc_addr, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "f"(v.x), "f"(v.y), "f"(v.z),
               "f"(v.w)
               : "memory");
}

__global__ void __launch_bounds__(256) reduce_adam_multimem_kernel(const ReduceAdamArgs q) {
  const long long n4 = q.n_slice >> 2, stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += stride) {
    const long long e = q.first + 4 * i;
    float4 g = multimem_ld_reduce_add(q.mc_grads + e);
    g.x *= q.inv_world, g.y *= q.inv_world, g.z *= q.inv_world, g.w *= q.inv_world;
    float4 p = *reinterpret_cast<const float4*>(q.params + e);
    float4 m = reinterpret_cast<float4*>(q.m)[i], v = reinterpret_cast<float4*>(q.v)[i];
    adam_one(p.x, g.x, m.x, v.x, q.a);
    adam_one(p.y, g.y, m.y, v.y, q.a);
    adam_one(p.z, g.z, m.z, v.z, q.a);
    adam_one(p.w, g.w, m.w, v.w, q.a);
    reinterpret_cast<float4*>(q.m)[i] = m, reinterpret_cast<float4*>(q.v)[i] = v;
    multimem_st(q.mc_params + e, p);
  }
  __threadfence_system();   // the broadcast stores are ordered before the caller's cross-rank barrier
}

struct MetricsArgs {
  const float* logp[2];
  const long long* y[2];
  int classes[2];
  long long* pred[2];     // optional per-sample predictions (int64 as torch.max(...)[1])
  int n, ntasks;
  long long* counts;      // [0..1] correct per task, [2] samples seen, [3] samples with an out-of-range label
  double* loss_sums;      // [2] sum over samples of -logp[y] per task
  long long* confusion[2];  // [classes x classes], row = label, column = prediction (sklearn.metrics.confusion_matrix)
};

constexpr int kMaxClasses = 16;

__global__ void __launch_bounds__(256) metrics_kernel(const MetricsArgs a) {
  __shared__ int conf[2][kMaxClasses * kMaxClasses];
  __shared__ int correct[2], bad;
  __shared__ float loss[2];
  for (int i = threadIdx.x; i < 2 * kMaxClasses * kMaxClasses; i += blockDim.x) (&conf[0][0])[i] = 0;
  if (threadIdx.x < 2) correct[threadIdx.x] = 0, loss[threadIdx.x] = 0.f;
  if (threadIdx.x == 0) bad = 0;
  __syncthreads();
  int seen = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < a.n; i += gridDim.x * blockDim.x) {
    ++seen;
    for (int t = 0; t < a.ntasks; ++t) {
      const int C = a.classes[t];
      const float* lp = a.logp[t] + (size_t)i * C;
      int best = 0;
      float bv = lp[0];
      for (int j = 1; j < C; ++j) {
        const float x = lp[j];
        if (x > bv) bv = x, best = j;   // strict '>' : the first maximum wins, as torch.max does
      }
      if (a.pred[t]) a.pred[t][i] = best;
      const long long y = a.y[t][i];
      if (y < 0 || y >= C) {
        atomicAdd(&bad, 1);
        continue;
      }
      if (best == (int)y) atomicAdd(&correct[t], 1);
      atomicAdd(&loss[t], -lp[y]);
      atomicAdd(&conf[t][(int)y * C + best], 1);
    }
  }
  // one atomic per thread for the sample count is wasteful: reduce through the first warp lane
  for (int o = 16; o > 0; o >>= 1) seen += __shfl_xor_sync(0xffffffffu, seen, o);
  if ((threadIdx.x & 31) == 0 && seen) atomicAdd(reinterpret_cast<unsigned long long*>(a.counts + 2), (unsigned long long)seen);
  __syncthreads();
  for (int t = 0; t < a.ntasks; ++t) {
    const int C = a.classes[t];
    for (int i = threadIdx.x; i < C * C; i += blockDim.x)
      if (conf[t][i]) atomicAdd(reinterpret_cast<unsigned long long*>(a.confusion[t] + i), (unsigned long long)conf[t][i]);
  }
  if (threadIdx.x < a.ntasks) {
    const int t = threadIdx.x;
    if (correct[t]) atomicAdd(reinterpret_cast<unsigned long long*>(a.counts + t), (unsigned long long)correct[t]);
    atomicAdd(a.loss_sums + t, (double)loss[t]);
  }
  if (threadIdx.x == 0 && bad) atomicAdd(reinterpret_cast<unsigned long long*>(a.counts + 3), (unsigned long long)bad);
}

}  // namespace mtldas

using namespace mtldas;

extern "C" int mtldas_adam_step(float* dev_params, const float* dev_grads, float* dev_exp_avg, float* dev_exp_avg_sq,
                                size_t n, int step, float lr, float beta1, float beta2, float eps, float weight_decay,
                                void* stream) {
  if (!dev_params || !dev_grads || !dev_exp_avg || !dev_exp_avg_sq || step < 1 || !(beta1 >= 0.f && beta1 < 1.f) ||
      !(beta2 >= 0.f && beta2 < 1.f) ||
      ((reinterpret_cast<uintptr_t>(dev_params) | reinterpret_cast<uintptr_t>(dev_grads) |
        reinterpret_cast<uintptr_t>(dev_exp_avg) | reinterpret_cast<uintptr_t>(dev_exp_avg_sq)) & 15)) {
    set_error("adam_step: bad arguments (non-null 16-byte aligned buffers, step >= 1, betas in [0,1))");
    return MTLDAS_ERR_ARG;
  }
  if (n == 0) return 0;
  AdamArgs a;
  a.p = dev_params, a.g = dev_grads, a.m = dev_exp_avg, a.v = dev_exp_avg_sq, a.n = (long long)n;
  const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
  a.lr_over_bc1 = (float)((double)lr / bc1);
  a.inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
  a.beta1 = beta1, a.beta2 = beta2, a.eps = eps, a.weight_decay = weight_decay;
  const long long n4 = ((long long)n + 3) / 4;
  const int blocks = (int)std::min<long long>(148 * 8, (n4 + 255) / 256);
  adam_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(a);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("adam_step: %s", cudaGetErrorString(e));
    return MTLDAS_ERR_CUDA;
  }
  return 0;
}

extern "C" int mtldas_reduce_adam_multimem(const float* mc_grads, float* mc_params, const float* dev_params,
                                           float* dev_exp_avg_slice, float* dev_exp_avg_sq_slice, size_t n_total, int rank,
                                           int world, int step, float lr, float beta1, float beta2, float eps,
                                           float weight_decay, void* stream) {
  if (!mc_grads || !mc_params || !dev_params || !dev_exp_avg_slice || !dev_exp_avg_sq_slice || world < 1 || rank < 0 ||
      rank >= world || step < 1 || n_total == 0 || (n_total % (size_t)(4 * world)) != 0 ||
      ((reinterpret_cast<uintptr_t>(mc_grads) | reinterpret_cast<uintptr_t>(mc_params) | reinterpret_cast<uintptr_t>(dev_params) |
        reinterpret_cast<uintptr_t>(dev_exp_avg_slice) | reinterpret_cast<uintptr_t>(dev_exp_avg_sq_slice)) & 15)) {
    set_error("reduce_adam_multimem: bad arguments (16-byte aligned buffers; n_total must be a multiple of 4 * world)");
    return MTLDAS_ERR_ARG;
  }
  ReduceAdamArgs q{};
  q.mc_grads = mc_grads, q.mc_params = mc_params, q.params = dev_params, q.m = dev_exp_avg_slice, q.v = dev_exp_avg_sq_slice;
  q.n_slice = (long long)(n_total / (size_t)world), q.first = q.n_slice * rank;
  q.inv_world = 1.f / (float)world;
  const double bc1 = 1.0 - pow((double)beta1, (double)step), bc2 = 1.0 - pow((double)beta2, (double)step);
  q.a.lr_over_bc1 = (float)((double)lr / bc1), q.a.inv_sqrt_bc2 = (float)(1.0 / sqrt(bc2));
  q.a.beta1 = beta1, q.a.beta2 = beta2, q.a.eps = eps, q.a.weight_decay = weight_decay;
  const long long n4 = q.n_slice / 4;
  const int blocks = (int)std::max<long long>(1, std::min<long long>(148 * 4, (n4 + 255) / 256));
  reduce_adam_multimem_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(q);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("reduce_adam_multimem: %s", cudaGetErrorString(e));
    return MTLDAS_ERR_CUDA;
  }
  return 0;
}

extern "C" int mtldas_eval_metrics(const float* dev_logp0, const long long* dev_y0, int classes0, long long* dev_pred0,
                                   const float* dev_logp1, const long long* dev_y1, int classes1, long long* dev_pred1,
                                   int n, long long* dev_counts4, double* dev_loss_sums2, long long* dev_confusion0,
                                   long long* dev_confusion1, void* stream) {
  const bool two = dev_logp1 != nullptr;
  if (!dev_logp0 || !dev_y0 || !dev_counts4 || !dev_loss_sums2 || !dev_confusion0 || n < 0 || classes0 < 1 ||
      classes0 > kMaxClasses || (two && (!dev_y1 || !dev_confusion1 || classes1 < 1 || classes1 > kMaxClasses))) {
    set_error("eval_metrics: bad arguments (classes <= %d, counters / loss sums / confusion buffers required)", kMaxClasses);
    return MTLDAS_ERR_ARG;
  }
  if (n == 0) return 0;
  MetricsArgs a{};
  a.logp[0] = dev_logp0, a.y[0] = dev_y0, a.classes[0] = classes0, a.pred[0] = dev_pred0, a.confusion[0] = dev_confusion0;
  a.logp[1] = dev_logp1, a.y[1] = dev_y1, a.classes[1] = classes1, a.pred[1] = dev_pred1, a.confusion[1] = dev_confusion1;
  a.n = n, a.ntasks = two ? 2 : 1, a.counts = dev_counts4, a.loss_sums = dev_loss_sums2;
  const int blocks = std::min(148, (n + 255) / 256);
  metrics_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(a);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    set_error("eval_metrics: %s", cudaGetErrorString(e));
    return MTLDAS_ERR_CUDA;
  }
  return 0;
}
