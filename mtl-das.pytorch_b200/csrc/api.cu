// extern "C" surface of libmtldas_b200.so (see include/mtldas.h).
#include <algorithm>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "../../include/mtldas.h"
#include "kernels.h"
#include "plan.h"
#include "train.h"
#include "train_kernels.h"

using namespace mtldas;

struct mtldas_plan {
  Plan* p;
};

extern "C" {

#ifndef MTLDAS_BUILD_ID
#define MTLDAS_BUILD_ID "dev"
#endif
const char* mtldas_version(void) { return "mtldas-b200 0.2 (sm_100a) build " MTLDAS_BUILD_ID; }
const char* mtldas_last_error(void) { return last_error_text(); }

int mtldas_device_ok(void) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, dev) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return prop.major == 10 ? 1 : 0;
}

mtldas_plan* mtldas_plan_create(int model_kind, int micro_batch, int precision) {
  if (model_kind < 0 || model_kind > 2 || micro_batch < 1 || micro_batch > 4096 || precision < 0 || precision > 1) {
    set_error("plan_create: bad arguments (kind %d, micro_batch %d, precision %d)", model_kind, micro_batch,
              precision);
    return nullptr;
  }
  mtldas_plan* h = new mtldas_plan();
  h->p = build_plan(model_kind, micro_batch, precision);
  const char* sw = getenv("MTLDAS_UMMA_SWAP_DESC");
  h->p->swap_desc = (sw && sw[0] == '1') ? 1 : 0;
  const char* ng = getenv("MTLDAS_NO_GRAPH");
  h->p->use_graphs = (ng && ng[0] == '1') ? 0 : 1;
  const char* np = getenv("MTLDAS_NO_PDL");
  h->p->use_pdl = (np && np[0] == '1') ? 0 : 1;
  // tensor-core stem (stem_umma.cu, one banded GEMM per output row): 0.25 vs 0.58 ms per 4096 patches for the FFMA2
  // kernel; MTLDAS_UMMA_STEM=0 falls back to the CUDA-core stem
  const char* nus = getenv("MTLDAS_UMMA_STEM");
  h->p->use_umma_stem = (nus && nus[0] == '0') ? 0 : 1;
  const char* npz = getenv("MTLDAS_NO_PERSIST");
  h->p->use_persist = (npz && npz[0] == '1') ? 0 : 1;
  return h;
}
void mtldas_plan_destroy(mtldas_plan* plan) {
  if (!plan) return;
  delete plan->p;
  delete plan;
}
int mtldas_plan_num_state_entries(const mtldas_plan* plan) { return plan ? (int)plan->p->state_keys.size() : -1; }
const char* mtldas_plan_state_key(const mtldas_plan* plan, int i) {
  if (!plan || i < 0 || i >= (int)plan->p->state_keys.size()) return nullptr;
  return plan->p->state_keys[i].c_str();
}
size_t mtldas_plan_workspace_bytes(const mtldas_plan* plan) { return plan ? plan->p->ws_bytes : 0; }

int mtldas_plan_bind_workspace(mtldas_plan* plan, void* dev_workspace, size_t bytes, void* stream) {
  if (!plan || !dev_workspace || bytes < plan->p->ws_bytes || ((uintptr_t)dev_workspace & 255)) {
    set_error("bind_workspace: need %zu bytes, 256-byte aligned", plan ? plan->p->ws_bytes : (size_t)0);
    return MTLDAS_ERR_ARG;
  }
  cudaError_t e = cudaMemsetAsync(dev_workspace, 0, plan->p->ws_bytes, (cudaStream_t)stream);
  if (e != cudaSuccess) {
    set_error("bind_workspace: %s", cudaGetErrorString(e));
    return MTLDAS_ERR_CUDA;
  }
  Plan& pp = *plan->p;
  if (!pp.copy_stream) {  // created once here so that the forward entry points never create anything
    bool ok = cudaStreamCreateWithFlags(&pp.copy_stream, cudaStreamNonBlocking) == cudaSuccess;
    ok = ok && cudaEventCreateWithFlags(&pp.ev_start, cudaEventDisableTiming) == cudaSuccess;
    for (int i = 0; i < 2 && ok; ++i)
      ok = cudaEventCreateWithFlags(&pp.ev_h2d[i], cudaEventDisableTiming) == cudaSuccess &&
           cudaEventCreateWithFlags(&pp.ev_comp[i], cudaEventDisableTiming) == cudaSuccess;
    if (!ok) {
      set_error("bind_workspace: could not create the copy stream / events");
      return MTLDAS_ERR_CUDA;
    }
  }
  pp.ws = (char*)dev_workspace;
  pp.packed = false;
  return 0;
}

size_t mtldas_plan_describe(const mtldas_plan* plan, char* buf, size_t cap) {
  if (!plan) return 0;
  std::string s = plan->p->describe();
  if (buf && cap) {
    size_t n = s.size() < cap - 1 ? s.size() : cap - 1;
    memcpy(buf, s.data(), n);
    buf[n] = 0;
  }
  return s.size() + 1;
}

int mtldas_plan_pack_eval(mtldas_plan* plan, const void* const* dev_state, int n_entries, void* stream) {
  if (!plan || !dev_state || n_entries != (int)plan->p->state_keys.size()) {
    set_error("pack_eval: expected %d state entries", plan ? (int)plan->p->state_keys.size() : -1);
    return MTLDAS_ERR_ARG;
  }
  for (int i = 0; i < n_entries; ++i)
    if (!dev_state[i]) {
      set_error("pack_eval: state entry %d (%s) is NULL", i, plan->p->state_keys[i].c_str());
      return MTLDAS_ERR_ARG;
    }
  return plan_pack_eval(*plan->p, dev_state, (cudaStream_t)stream);
}

int mtldas_plan_forward_eval(mtldas_plan* plan, const float* dev_x, int n, float* dev_out0, float* dev_out1,
                             void* stream) {
  if (!plan || !dev_x || !dev_out0 || n < 0) {
    set_error("forward_eval: bad arguments");
    return MTLDAS_ERR_ARG;
  }
  if (n == 0) return 0;
  return plan_forward_eval(*plan->p, dev_x, n, dev_out0, dev_out1, (cudaStream_t)stream);
}

// Chunk sizes of the host entry point (two staging slots: the copy of chunk i+1 runs under the compute of chunk i).
// Measured end to end at 4096 patches: fp32 patches are bound by the PCIe copy (55 GB/s = 554 k patches/s) and want
// short chunks (256 -> 418 k, 512 -> 495 k, 1024 -> 463 k patches/s: the exposed tail is the last chunk's compute);
// bf16 patches (half the bytes) are bound by compute and want larger ones (512 -> 540 k, 1024 -> 594 k, 2048 -> 561 k).
// Growing / shrinking schedules measured worse than the constants (more chunks, more per-chunk fixed cost).
static std::vector<int> host_chunk_schedule(int n, int max_chunk, bool x_is_bf16) {
  static int env_chunk = -1;
  if (env_chunk < 0) {
    const char* e = getenv("MTLDAS_HOST_CHUNK_SIZE");
    env_chunk = e ? atoi(e) : 0;
  }
  int c = env_chunk > 0 ? env_chunk : (x_is_bf16 ? 1024 : 512);
  c = std::max(1, std::min(c, max_chunk));
  std::vector<int> out;
  for (int rem = n; rem > 0; rem -= std::min(c, rem)) out.push_back(std::min(c, rem));
  return out;
}

static int forward_eval_host_impl(mtldas_plan* plan, const void* host_x_any, bool x_is_bf16, int n, float* host_out0,
                                  float* host_out1, void* stream) {
  if (!plan || !host_x_any || !host_out0 || n < 0) {
    set_error("forward_eval_host: bad arguments");
    return MTLDAS_ERR_ARG;
  }
  const float* host_x = static_cast<const float*>(host_x_any);
  const uint16_t* host_xh = static_cast<const uint16_t*>(host_x_any);
  Plan& p = *plan->p;
  if (x_is_bf16 && p.precision != 0) {
    set_error("forward_eval_host_bf16: bf16 patches are accepted by the bf16 plan only (the fp32 path keeps fp32 inputs)");
    return MTLDAS_ERR_UNSUPPORTED;
  }
  if (!p.ws || !p.packed) {
    set_error("forward_eval_host: bind a workspace and pack weights first");
    return MTLDAS_ERR_STATE;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const int c0 = p.heads[0].classes, c1 = p.heads.size() > 1 ? p.heads[1].classes : 0;
  const int chunk = p.host_chunk;
  const size_t xslot = (size_t)chunk * kPatchH * kPatchW;
  float* xs[2] = {reinterpret_cast<float*>(p.ws + p.x_stage_off), reinterpret_cast<float*>(p.ws + p.x_stage_off) + xslot};
  float* o0[2] = {reinterpret_cast<float*>(p.ws + p.out_stage_off),
                  reinterpret_cast<float*>(p.ws + p.out_stage_off) + (size_t)chunk * 18};
  float* o1[2] = {o0[0] + (size_t)chunk * 16, o0[1] + (size_t)chunk * 16};
  uint16_t* xh[2] = {reinterpret_cast<uint16_t*>(p.ws + p.xh_stage_off), reinterpret_cast<uint16_t*>(p.ws + p.xh_stage_off) + xslot};
#define CKH(call)                                                        \
  do {                                                                   \
    cudaError_t e_ = (call);                                             \
    if (e_ != cudaSuccess) {                                             \
      set_error("forward_eval_host: %s: %s", #call, cudaGetErrorString(e_)); \
      return MTLDAS_ERR_CUDA;                                            \
    }                                                                    \
  } while (0)
  // Two staging slots: the copy stream brings pass i+1 in while the caller's stream computes pass i.
  CKH(cudaEventRecord(p.ev_start, st));
  CKH(cudaStreamWaitEvent(p.copy_stream, p.ev_start, 0));
  int launches = 0, i = 0, s = 0;
  const std::vector<int> schedule = host_chunk_schedule(n, chunk, x_is_bf16);
  for (size_t ci = 0; ci < schedule.size(); s += schedule[ci], ++ci, ++i) {
    const int nb = schedule[ci];
    const int slot = i & 1;
    if (i >= 2) CKH(cudaStreamWaitEvent(p.copy_stream, p.ev_comp[slot], 0));  // slot free again
    if (x_is_bf16)   // half the bytes over PCIe; widened on the device (exact), the stem rounds to the same bf16 values
      CKH(cudaMemcpyAsync(xh[slot], host_xh + (size_t)s * kPatchH * kPatchW, (size_t)nb * kPatchH * kPatchW * 2,
                          cudaMemcpyHostToDevice, p.copy_stream));
    else
      CKH(cudaMemcpyAsync(xs[slot], host_x + (size_t)s * kPatchH * kPatchW, (size_t)nb * kPatchH * kPatchW * 4,
                          cudaMemcpyHostToDevice, p.copy_stream));
    CKH(cudaEventRecord(p.ev_h2d[slot], p.copy_stream));
    CKH(cudaStreamWaitEvent(st, p.ev_h2d[slot], 0));
    if (x_is_bf16) CKH(launch_widen_bf16(xh[slot], xs[slot], (long long)nb * kPatchH * kPatchW, st));
    int rc = plan_forward_eval(p, xs[slot], nb, o0[slot], c1 ? o1[slot] : nullptr, st);
    if (rc) return rc;
    launches += p.last_launches;
    CKH(cudaEventRecord(p.ev_comp[slot], st));
    CKH(cudaMemcpyAsync(host_out0 + (size_t)s * c0, o0[slot], (size_t)nb * c0 * 4, cudaMemcpyDeviceToHost, st));
    if (c1 && host_out1)
      CKH(cudaMemcpyAsync(host_out1 + (size_t)s * c1, o1[slot], (size_t)nb * c1 * 4, cudaMemcpyDeviceToHost, st));
  }
  p.last_launches = launches;
  CKH(cudaStreamSynchronize(st));
#undef CKH
  return 0;
}

int mtldas_plan_forward_eval_host(mtldas_plan* plan, const float* host_x, int n, float* host_out0, float* host_out1,
                                  void* stream) {
  return forward_eval_host_impl(plan, host_x, false, n, host_out0, host_out1, stream);
}

int mtldas_plan_forward_eval_host_bf16(mtldas_plan* plan, const void* host_x_bf16, int n, float* host_out0,
                                       float* host_out1, void* stream) {
  return forward_eval_host_impl(plan, host_x_bf16, true, n, host_out0, host_out1, stream);
}

int mtldas_plan_profile_eval(mtldas_plan* plan, const float* dev_x, int n, float* dev_out0, float* dev_out1,
                             void* stream, float* host_ms_per_op, int cap) {
  if (!plan || !dev_x || !dev_out0 || !host_ms_per_op || n < 1) {
    set_error("profile_eval: bad arguments");
    return MTLDAS_ERR_ARG;
  }
  return plan_profile_eval(*plan->p, dev_x, n, dev_out0, dev_out1, (cudaStream_t)stream, host_ms_per_op, cap);
}

int mtldas_plan_last_launch_count(const mtldas_plan* plan) { return plan ? plan->p->last_launches : -1; }

int mtldas_plan_read_tensor(mtldas_plan* plan, const char* name, int nb, float* dev_out_nchw, void* stream) {
  if (!plan || !name || !dev_out_nchw) return MTLDAS_ERR_ARG;
  return plan_read_tensor(*plan->p, name, nb, dev_out_nchw, (cudaStream_t)stream);
}

int mtldas_conv2d_nchw(const float* dev_x, const float* dev_x2, const float* dev_weight, const float* dev_bias,
                       const float* dev_aux, float* dev_out, int n, int cin, int cin2, int h, int w, int cout,
                       int ksize, int stride, int epilogue, int precision, int aux_phase_split) {
  const char* sw = getenv("MTLDAS_UMMA_SWAP_DESC");
  return single_conv(dev_x, dev_x2, dev_weight, dev_bias, dev_aux, dev_out, n, cin, cin2, h, w, cout, ksize, stride,
                     epilogue, precision, aux_phase_split, (sw && sw[0] == '1') ? 1 : 0);
}

int mtldas_nll_loss(const float* dev_logp0, const long long* dev_y0, int classes0, float* dev_dlogp0,
                    const float* dev_logp1, const long long* dev_y1, int classes1, float* dev_dlogp1, int n,
                    float* dev_loss, int* dev_status, void* stream) {
  if (!dev_logp0 || !dev_y0 || classes0 < 1 || n < 1 || !dev_loss || (dev_logp1 && (!dev_y1 || classes1 < 1))) {
    set_error("nll_loss: bad arguments (n %d, classes %d / %d)", n, classes0, classes1);
    return MTLDAS_ERR_ARG;
  }
  cudaError_t e = launch_nll_loss(dev_logp0, dev_y0, classes0, dev_dlogp0, dev_logp1, dev_y1, classes1, dev_dlogp1, n,
                                  dev_loss, dev_status, (cudaStream_t)stream);
  if (e != cudaSuccess) {
    set_error("nll_loss: %s", cudaGetErrorString(e));
    return MTLDAS_ERR_CUDA;
  }
  return 0;
}

int mtldas_plan_forward_eval_windows(mtldas_plan* plan, const float* dev_array, long long n_channels, long long n_time,
                                     int first_channel, int window_stride, int n_windows, float* dev_out0,
                                     float* dev_out1, void* stream) {
  if (!plan || !dev_array || !dev_out0 || window_stride < 1 || first_channel < 0) {
    set_error("forward_eval_windows: bad arguments");
    return MTLDAS_ERR_ARG;
  }
  if (n_windows == 0) return 0;
  return plan_forward_windows(*plan->p, dev_array, n_channels, n_time, first_channel, window_stride, n_windows, dev_out0,
                              dev_out1, (cudaStream_t)stream);
}

int mtldas_window_gather(const float* dev_array, long long n_channels, long long n_time, int first_channel,
                         int window_stride, int n_windows, float* dev_patches, void* stream) {
  if (!dev_array || !dev_patches || n_windows < 0 || window_stride < 1 || first_channel < 0 || n_time < kPatchW ||
      (long long)first_channel + (long long)(n_windows > 0 ? n_windows - 1 : 0) * window_stride + kPatchH > n_channels) {
    set_error("window_gather: windows do not fit the array");
    return MTLDAS_ERR_ARG;
  }
  cudaError_t e = launch_window_gather(dev_array, n_time, first_channel, window_stride, n_windows,
                                       (int)(n_time / kPatchW), dev_patches, (cudaStream_t)stream);
  if (e != cudaSuccess) {
    set_error("window_gather: %s", cudaGetErrorString(e));
    return MTLDAS_ERR_CUDA;
  }
  return 0;
}

float mtldas_umma_selftest(int n_cols, int row_shift, int swap_lbo_sbo) {
  return umma_selftest(n_cols, row_shift, swap_lbo_sbo);
}

int mtldas_umma_rate(int n_cols, int mode, int reps, float* host_out2) {
  return umma_rate(n_cols, mode, reps, host_out2);
}

struct mtldas_train {
  TrainPlan* tp;
};

mtldas_train* mtldas_train_create(int model_kind, int batch, int precision) {
  if (model_kind < 0 || model_kind > 2 || batch < 1 || batch > 1024 || precision < 0 || precision > 1) {
    set_error("train_create: bad arguments (kind %d, batch %d, precision %d)", model_kind, batch, precision);
    return nullptr;
  }
  mtldas_train* h = new mtldas_train();
  h->tp = build_train_plan(model_kind, batch, precision);
  const char* nus = getenv("MTLDAS_UMMA_STEM");
  if (h->tp && h->tp->p) h->tp->p->use_umma_stem = (nus && nus[0] == '0') ? 0 : 1;
  return h;
}
void mtldas_train_destroy(mtldas_train* t) {
  if (!t) return;
  delete t->tp;
  delete t;
}
int mtldas_train_num_state_entries(const mtldas_train* t) { return t ? (int)t->tp->p->state_keys.size() : -1; }
size_t mtldas_train_num_params(const mtldas_train* t) { return t ? t->tp->n_params : 0; }
size_t mtldas_train_workspace_bytes(const mtldas_train* t) { return t ? t->tp->p->ws_bytes : 0; }
int mtldas_train_bind_workspace(mtldas_train* t, void* dev_workspace, size_t bytes, void* stream) {
  if (!t || !dev_workspace || bytes < t->tp->p->ws_bytes || ((uintptr_t)dev_workspace & 255)) {
    set_error("train_bind_workspace: need %zu bytes, 256-byte aligned", t ? t->tp->p->ws_bytes : (size_t)0);
    return MTLDAS_ERR_ARG;
  }
  cudaError_t e = cudaMemsetAsync(dev_workspace, 0, t->tp->p->ws_bytes, (cudaStream_t)stream);
  if (e != cudaSuccess) {
    set_error("train_bind_workspace: %s", cudaGetErrorString(e));
    return MTLDAS_ERR_CUDA;
  }
  t->tp->p->ws = (char*)dev_workspace;
  t->tp->have_forward = false;
  return 0;
}
int mtldas_train_forward(mtldas_train* t, const void* const* dev_state, int n_entries, const float* dev_x, int n,
                         float* dev_out0, float* dev_out1, void* stream) {
  if (!t || !dev_state || !dev_x || !dev_out0 || n_entries != (int)t->tp->p->state_keys.size()) {
    set_error("train_forward: bad arguments (expected %d state entries)", t ? (int)t->tp->p->state_keys.size() : -1);
    return MTLDAS_ERR_ARG;
  }
  for (int i = 0; i < n_entries; ++i)
    if (!dev_state[i]) {
      set_error("train_forward: state entry %d (%s) is NULL", i, t->tp->p->state_keys[i].c_str());
      return MTLDAS_ERR_ARG;
    }
  return train_forward(*t->tp, dev_state, dev_x, n, dev_out0, dev_out1, (cudaStream_t)stream);
}
int mtldas_train_backward(mtldas_train* t, const void* const* dev_state, int n_entries, const float* dev_dout0,
                          const float* dev_dout1, float* dev_flat_grads, void* stream) {
  if (!t || !dev_state || !dev_dout0 || !dev_flat_grads || n_entries != (int)t->tp->p->state_keys.size()) {
    set_error("train_backward: bad arguments");
    return MTLDAS_ERR_ARG;
  }
  return train_backward(*t->tp, dev_state, dev_dout0, dev_dout1, dev_flat_grads, (cudaStream_t)stream);
}
int mtldas_train_last_launch_count(const mtldas_train* t) { return t ? t->tp->launches : -1; }
int mtldas_train_read_tensor(mtldas_train* t, const char* name, int nb, float* dev_out_nchw, void* stream) {
  if (!t || !name || !dev_out_nchw) return MTLDAS_ERR_ARG;
  return plan_read_tensor(*t->tp->p, name, nb, dev_out_nchw, (cudaStream_t)stream);
}
size_t mtldas_train_describe(const mtldas_train* t, char* buf, size_t cap) {
  if (!t) return 0;
  std::string s = t->tp->p->describe();
  if (buf && cap) {
    size_t n = s.size() < cap - 1 ? s.size() : cap - 1;
    memcpy(buf, s.data(), n);
    buf[n] = 0;
  }
  return s.size() + 1;
}

}  // extern "C"
