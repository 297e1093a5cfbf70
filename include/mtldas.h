/*
 * mtldas.h -- C ABI of libmtldas_b200.so: the B200 (sm_100a) implementation of the
 * MTL-DAS hot path (forward / backward of the two-level multi-task CNN).
 *
 * The reference (sunmin123456/MTL-DAS.PyTorch) has no FFI of its own: its "plugin
 * boundary" is the Python nn.Module protocol used by utils.py.  Every entry point below
 * therefore cites the reference *Python* interface it serves; the ctypes binding that a
 * maintainer adds on the reference side is shown in INTEGRATION.md and implemented in
 * mtl-das.pytorch_b200/mtldas_b200/_lib.py.
 *
 * Conventions
 *   - plain pointers and sizes only; all `dev` pointers are device pointers owned by the
 *     caller, all `host` pointers are host pointers owned by the caller;
 *   - every call is asynchronous on `stream` (a cudaStream_t passed as void*) unless the
 *     name ends in `_host` (those synchronise the stream before returning);
 *   - return value 0 = OK, negative = error; text via mtldas_last_error() (thread local);
 *   - the library keeps no pointer beyond the call except the workspace bound to a plan;
 *   - current-device semantics: the caller makes the plan's device current;
 *   - there is NO CPU fallback anywhere: without a B200 the compute calls fail.
 */
#ifndef MTLDAS_H_
#define MTLDAS_H_

#include <stddef.h>
#include <stdint.h>

#if defined(__GNUC__)
#define MTLDAS_API __attribute__((visibility("default")))
#else
#define MTLDAS_API
#endif

#ifdef __cplusplus
extern "C" {
#endif

#define MTLDAS_OK 0
#define MTLDAS_ERR_ARG -1
#define MTLDAS_ERR_CUDA -2
#define MTLDAS_ERR_STATE -3
#define MTLDAS_ERR_UNSUPPORTED -4

/* model kinds: utils.py:85-96 (`--model MTL | single_distance | single_event`) */
#define MTLDAS_MODEL_MTL 0             /* model/modelA_MTL.py:53  MTL_Net()                        */
#define MTLDAS_MODEL_SINGLE_DISTANCE 1 /* model/modelB_singleTask.py:53 Single_Task_Net('distance') */
#define MTLDAS_MODEL_SINGLE_EVENT 2    /* model/modelB_singleTask.py:53 Single_Task_Net('event')    */

/* arithmetic of the convolutions (BASELINE.json north_star: "FP32 path" / "BF16 tensor-core path") */
#define MTLDAS_PREC_BF16 0 /* tcgen05 (UMMA) bf16 x bf16 -> fp32 TMEM accumulators, bf16 activations in HBM */
#define MTLDAS_PREC_FP32 1 /* CUDA-core FFMA, fp32 activations in HBM                                       */

/* epilogues of the single-conv entry point */
#define MTLDAS_EPI_NONE 0        /* y = conv + bias                                   */
#define MTLDAS_EPI_RELU 1        /* relu(y)                 modelA_MTL.py:16,47,102   */
#define MTLDAS_EPI_ADD_RELU 2    /* relu(y + aux)           modelA_MTL.py:29-30       */
#define MTLDAS_EPI_SIGMOID_MUL 3 /* sigmoid(y) * aux        modelA_MTL.py:50,144      */
#define MTLDAS_EPI_RELU_POOL 5   /* maxpool2x2_ceil(relu(y)) modelA_MTL.py:102,116,145 (bf16 3x3/s1 only; out is
                                    [n, cout, ceil(h/2), ceil(w/2)]; the full-resolution tensor is never written)     */

typedef struct mtldas_plan mtldas_plan; /* opaque; not thread safe; distinct plans may run concurrently */

/* library identity / error text ------------------------------------------------------------ */
MTLDAS_API const char* mtldas_version(void);
MTLDAS_API const char* mtldas_last_error(void);
/* 1 if the current device is compute capability 10.x, 0 otherwise (no device: 0). */
MTLDAS_API int mtldas_device_ok(void);

/* plan life cycle ---------------------------------------------------------------------------
 * Replaces: constructing the module and moving it to the GPU
 *   utils.py:85-96 (model = MTL_Net() / Single_Task_Net(task=...)), utils.py:124-125 (model.cuda()).
 * `micro_batch` is the number of patches processed per pass over the layer program (the
 * workspace is sized for it; any batch size is accepted by forward and is cut into passes). */
MTLDAS_API mtldas_plan* mtldas_plan_create(int model_kind, int micro_batch, int precision);
MTLDAS_API void mtldas_plan_destroy(mtldas_plan* plan);
/* number of state_dict entries the plan expects (268 for MTL, 194 for single task: SURVEY.md App. C) */
MTLDAS_API int mtldas_plan_num_state_entries(const mtldas_plan* plan);
/* key of entry i in state_dict order, e.g. "att_mask_generato2.0.0.weight" (modelA_MTL.py:93, typo kept) */
MTLDAS_API const char* mtldas_plan_state_key(const mtldas_plan* plan, int i);
MTLDAS_API size_t mtldas_plan_workspace_bytes(const mtldas_plan* plan);
/* Bind caller-owned device scratch (>= workspace_bytes, 256-byte aligned).  Zero-fills it once
 * (the zero padding of every activation tensor lives there) on `stream`. */
MTLDAS_API int mtldas_plan_bind_workspace(mtldas_plan* plan, void* dev_workspace, size_t bytes, void* stream);
/* JSON description of tensors / ops / weight packing, for tests and debugging. Returns needed size. */
MTLDAS_API size_t mtldas_plan_describe(const mtldas_plan* plan, char* buf, size_t cap);

/* weights ------------------------------------------------------------------------------------
 * Replaces: model.load_state_dict(sd, strict=True) + model.eval()  (utils.py:122-123,249-250).
 * `dev_state[i]` is the device pointer of state_dict entry i (fp32 tensors, int64 counters) in the
 * reference's state_dict order.  Folds eval-mode BatchNorm (running statistics, eps 1e-5) into the
 * packed convolution weights. Must be called again whenever the parameters change. */
MTLDAS_API int mtldas_plan_pack_eval(mtldas_plan* plan, const void* const* dev_state, int n_entries, void* stream);

/* forward ------------------------------------------------------------------------------------
 * Replaces: out1, out2 = model(x) in eval mode  (utils.py:272; modelA_MTL.py:127-174,
 * modelB_singleTask.py:130-178).  dev_x: fp32 [n,1,100,250] contiguous.
 * MTL: dev_out0 = log-probs distance [n,16], dev_out1 = log-probs event [n,2].
 * single task: dev_out0 = [n,16] or [n,2]; dev_out1 ignored (may be NULL). */
MTLDAS_API int mtldas_plan_forward_eval(mtldas_plan* plan, const float* dev_x, int n, float* dev_out0, float* dev_out1,
                             void* stream);
/* Same through HOST buffers (pinned or pageable): H2D of x, forward, D2H of the log-probs, then a
 * stream synchronise.  Replaces the `valX.cuda()` + `model(valX)` + `.cpu()` sequence of
 * utils.py:264-277.  Uses staging space inside the bound workspace; allocates nothing. */
MTLDAS_API int mtldas_plan_forward_eval_host(mtldas_plan* plan, const float* host_x, int n, float* host_out0,
                                  float* host_out1, void* stream);
/* The same with the host patches already in bfloat16 (raw 16-bit words, [n,1,100,250]): HALF the bytes over PCIe.
 * The bf16 path rounds the patches to bf16 before its first MMA anyway, so for patches that are bf16-representable the
 * results are bit-identical to mtldas_plan_forward_eval_host on the same values.  bf16 plans only. */
MTLDAS_API int mtldas_plan_forward_eval_host_bf16(mtldas_plan* plan, const void* host_x_bf16, int n, float* host_out0,
                                                  float* host_out1, void* stream);
/* Instrumented forward for bench.py: same work as forward_eval with a CUDA-event pair around every launch;
 * host_ms_per_op[i] receives the device time of program op i (order of describe()'s "program") summed over
 * the micro-batch passes. Synchronises the stream. No reference counterpart (measurement only). */
MTLDAS_API int mtldas_plan_profile_eval(mtldas_plan* plan, const float* dev_x, int n, float* dev_out0, float* dev_out1,
                                        void* stream, float* host_ms_per_op, int cap);
/* number of kernels the last forward launched (for bench.py's gpu_launches) */
MTLDAS_API int mtldas_plan_last_launch_count(const mtldas_plan* plan);
/* Debug/test: copy an internal activation tensor of the LAST micro-batch pass out as fp32 NCHW
 * [nb, C, H, W] (C,H,W as listed by describe). */
MTLDAS_API int mtldas_plan_read_tensor(mtldas_plan* plan, const char* name, int nb, float* dev_out_nchw, void* stream);

/* single fused convolution (unit-parity entry point) -----------------------------------------
 * Replaces one nn.Conv2d (+ folded eval BatchNorm) + activation of the reference, e.g.
 * modelA_MTL.py:13-18 (ResBlock.left), :23 (shortcut), :45-50 (att_generator), :37 (conv3x3).
 * x: fp32 NCHW [n, cin, h, w]; x2 (optional, may be NULL): second source concatenated after x along
 * channels [n, cin2, h, w] (torch.cat, modelA_MTL.py:149); weight fp32 [cout, cin+cin2, k, k];
 * bias fp32 [cout] or NULL; aux fp32 NCHW [n, cout, ho, wo] for ADD_RELU / SIGMOID_MUL;
 * out fp32 NCHW [n, cout, ho, wo].  ksize in {1,3}, stride in {1,2}, padding = ksize/2 for 3, 0 for 1.
 * Allocates temporaries with cudaMalloc (test entry point, not for the hot loop). Synchronous. */
MTLDAS_API int mtldas_conv2d_nchw(const float* dev_x, const float* dev_x2, const float* dev_weight, const float* dev_bias,
                       const float* dev_aux, float* dev_out, int n, int cin, int cin2, int h, int w, int cout,
                       int ksize, int stride, int epilogue, int precision, int aux_phase_split);

/* single fused convolution WITHOUT any allocation -----------------------------------------------
 * No counterpart in the reference: this exists for hosts of this library that cannot allocate device memory --
 * a DAS acquisition / streaming loop that must never hit the device-wide synchronisation of cudaMalloc/cudaFree,
 * or a process whose GPU memory already belongs to one pool (e.g. PyTorch's caching allocator or a start-up arena).
 * Same arguments and results as mtldas_conv2d_nchw plus a stream; asynchronous on `stream` (no synchronise).
 * Its temporaries live in a fixed 32 MiB array inside the library image (held by every GPU that loads the
 * library, until exit); a call that needs more returns MTLDAS_ERR_UNSUPPORTED with the required size in
 * mtldas_last_error() and allocates nothing.  The first call on a device looks the array's device address up
 * (cudaGetSymbolAddress) -- make that first call before any stream capture; the address is cached per device ordinal
 * for the life of the process, so after cudaDeviceReset the entry point must not be used again.  All calls share the array: the part a
 * call uses is cleared first, and calls whose GPU work overlaps (other streams / threads) race and must be serialised. */
MTLDAS_API int mtldas_conv2d_nchw_static(const float* dev_x, const float* dev_x2, const float* dev_weight,
                                         const float* dev_bias, const float* dev_aux, float* dev_out, int n, int cin,
                                         int cin2, int h, int w, int cout, int ksize, int stride, int epilogue,
                                         int precision, int aux_phase_split, void* stream);

/* training: forward with batch-statistics BatchNorm + backward -------------------------------------
 * Replaces the reference training step's model calls (utils.py:344,359,372-373):
 *     model.train(); out1, out2 = model(x); ...; loss_balance.backward()
 * i.e. MTL_Net.forward in train mode (modelA_MTL.py:127-174, nn.BatchNorm2d batch statistics, running-stat update
 * with momentum 0.1) and the autograd backward of it.  The loss (NLLLoss per task, utils.py:139,361,367) stays
 * outside: backward takes d(loss)/d(log-probs).
 * `batch` is the largest batch the plan accepts (BatchNorm couples the batch, so one pass holds it all).
 * EXCLUSIVITY: the fused BatchNorm kernels synchronise the thread blocks of one channel chunk through a device-side
 * counter and size their grids so that every block is resident on an otherwise idle GPU (the plan's own side stream is
 * accounted for).  Keep at most ONE training plan with work in flight per device -- two training steps overlapping on
 * one GPU (other streams / host threads, or another process under MPS) can starve each other's blocks; the kernels
 * then trap after a bounded wait instead of hanging.  Inference plans have no such restriction. */
typedef struct mtldas_train mtldas_train;
MTLDAS_API mtldas_train* mtldas_train_create(int model_kind, int batch, int precision);
MTLDAS_API void mtldas_train_destroy(mtldas_train* t);
MTLDAS_API int mtldas_train_num_state_entries(const mtldas_train* t);
/* number of fp32 elements of the flat gradient buffer = sum of parameter sizes in named_parameters() order */
MTLDAS_API size_t mtldas_train_num_params(const mtldas_train* t);
MTLDAS_API size_t mtldas_train_workspace_bytes(const mtldas_train* t);
MTLDAS_API int mtldas_train_bind_workspace(mtldas_train* t, void* dev_workspace, size_t bytes, void* stream);
/* dev_state: the module's state_dict entries (reference order). Parameters are read; running_mean / running_var /
 * num_batches_tracked are UPDATED IN PLACE as nn.BatchNorm2d does in train mode. Outputs as forward_eval. */
MTLDAS_API int mtldas_train_forward(mtldas_train* t, const void* const* dev_state, int n_entries, const float* dev_x,
                                    int n, float* dev_out0, float* dev_out1, void* stream);
/* Gradients of the LAST train_forward: dev_dout0/1 = d loss / d log-probs ([n,16] / [n,2]); writes every parameter
 * gradient into dev_flat_grads (fp32, named_parameters() order: conv weight, [conv bias], BN weight, BN bias per
 * layer).  Pre-BatchNorm conv biases get exactly 0 (their gradient cancels). */
MTLDAS_API int mtldas_train_backward(mtldas_train* t, const void* const* dev_state, int n_entries,
                                     const float* dev_dout0, const float* dev_dout1, float* dev_flat_grads, void* stream);
MTLDAS_API int mtldas_train_last_launch_count(const mtldas_train* t);
MTLDAS_API int mtldas_train_read_tensor(mtldas_train* t, const char* name, int nb, float* dev_out_nchw, void* stream);
MTLDAS_API size_t mtldas_train_describe(const mtldas_train* t, char* buf, size_t cap);

/* joint loss of the task heads and its gradient ----------------------------------------------------
 * Replaces `criterion(pred1, y1) + criterion(pred2, y2)` with criterion = nn.NLLLoss() (utils.py:139,361,367) and
 * its autograd: *dev_loss = -(1/n) sum_i logp0[i, y0[i]] - (1/n) sum_i logp1[i, y1[i]];
 * dev_dlogp_t[i, j] = -(1/n) [j == y_t[i]] (dense [n, classes_t] fp32, the upstream gradients mtldas_train_backward
 * takes).  Labels are int64 as after `.long()`.  The second task (logp1 / y1 / dlogp1) may be NULL for the
 * single-task nets; either dlogp may be NULL (loss only).  dev_status (int, may be NULL) is set to 1 when a label
 * lies outside [0, classes) -- torch raises there; the caller checks it where it would have synchronised anyway. */
MTLDAS_API int mtldas_nll_loss(const float* dev_logp0, const long long* dev_y0, int classes0, float* dev_dlogp0,
                               const float* dev_logp1, const long long* dev_y1, int classes1, float* dev_dlogp1,
                               int n, float* dev_loss, int* dev_status, void* stream);

/* optimizer step over the flat parameter buffer -------------------------------------------------------
 * Replaces `optimizer.step()` with optimizer = optim.Adam(model.parameters(), lr=1e-3, weight_decay=1e-5)
 * (utils.py:133-134,374): torch's Adam with L2 weight decay folded into the gradient (not AdamW), bias-corrected,
 * no amsgrad.  One launch over n contiguous fp32 elements (all four buffers 16-byte aligned); `step` is the 1-based
 * step count (the bias corrections 1 - beta^step are formed in double on the host).  The learning-rate schedule
 * (lr /= 1.5 every 5 epochs, utils.py:245-247) stays with the caller: pass the current lr. */
MTLDAS_API int mtldas_adam_step(float* dev_params, const float* dev_grads, float* dev_exp_avg, float* dev_exp_avg_sq,
                                size_t n, int step, float lr, float beta1, float beta2, float eps, float weight_decay,
                                void* stream);

/* gradient all-reduce + Adam + parameter broadcast in one kernel (multi-GPU, NVLink / NVSwitch multicast) -----------
 * Replaces, for data-parallel training, `all-reduce(grads); optimizer.step()` (the reference has no distributed code;
 * its step is utils.py:372-374): rank r owns the slice [r*n/world, (r+1)*n/world) of the flat buffers, sums that slice
 * of every rank's gradient inside the switch (multimem.ld_reduce), divides by world, applies the Adam rule of
 * mtldas_adam_step and stores the updated slice into every rank's parameter buffer (multimem.st).
 * mc_grads / mc_params are MULTICAST addresses of symmetric allocations (e.g. torch.distributed._symmetric_memory:
 * handle.multicast_ptr) holding the flat gradient / parameter buffers; dev_params is this rank's ordinary address of
 * the same parameter buffer; the two moment buffers hold the owned slice only (n_total / world floats).
 * The kernel waits for nobody: the caller runs a cross-rank barrier before it (every gradient written) and after it
 * (every slice landed) -- mtldas_b200.ddp.DataParallelTrainer uses the symmetric-memory handle's barrier. */
MTLDAS_API int mtldas_reduce_adam_multimem(const float* mc_grads, float* mc_params, const float* dev_params,
                                           float* dev_exp_avg_slice, float* dev_exp_avg_sq_slice, size_t n_total, int rank,
                                           int world, int step, float lr, float beta1, float beta2, float eps,
                                           float weight_decay, void* stream);

/* validation bookkeeping on the device -----------------------------------------------------------------
 * Replaces, per validation batch, utils.py:274-291: criterion(out_t, y_t) sums, pred_t = torch.max(out_t, 1)[1]
 * (first maximum wins), (pred_t == y_t).sum(), and the label / prediction lists that become
 * sklearn.metrics.confusion_matrix at utils.py:313 -- without the per-sample `int(x)` device syncs of
 * utils.py:308-311.  ACCUMULATES into caller-zeroed buffers, so one read-back per epoch suffices:
 *   dev_counts4[0..1] correct predictions per task, [2] samples seen, [3] labels outside [0, classes);
 *   dev_loss_sums2[t] = sum_i -logp_t[i, y_t[i]]  (NLLLoss(mean) of a batch = its share / n);
 *   dev_confusion_t[label * classes_t + prediction] (int64).
 * dev_pred_t (int64 [n]) may be NULL.  The second task may be NULL (single-task nets).  classes <= 16. */
MTLDAS_API int mtldas_eval_metrics(const float* dev_logp0, const long long* dev_y0, int classes0, long long* dev_pred0,
                                   const float* dev_logp1, const long long* dev_y1, int classes1, long long* dev_pred1,
                                   int n, long long* dev_counts4, double* dev_loss_sums2, long long* dev_confusion0,
                                   long long* dev_confusion1, void* stream);

/* model C: Multi_Classifier (Inception-V3, 1-channel stem, 32 classes), eval-mode forward -------------------------
 * Replaces, op by op, what /root/reference/model/modelC_multiClassifier.py:102-147 executes in eval mode (utils.py:648):
 * BasicConv2d = Conv2d(bias=False) + BatchNorm2d(eps=1e-3) + ReLU (modelC:10-25 and torchvision's Inception blocks),
 * the 3x3 pools, the channel concatenations (a channel offset into the destination tensor) and the classifier head.
 * Activations use the "planar-8" layout: [channels/8][n*H*W][8] bf16; the host side (mtldas_b200/nets_c.py) owns the
 * tensors and the layer program.  bf16 tensor-core path only.  All calls are asynchronous on `stream`.
 *
 * c_pack_conv: folds BatchNorm (running statistics, eps as given) into the conv weight [cout, cin, kh, kw] and writes
 *   the packed image (kh*kw*cin*cout bf16 in the kernels' own order [output-channel tile][K unit][tap][8-channel
 *   chunk][channel of the tile][8]: what one CTA consumes at a time is contiguous; cin == 1: [kh*kw][cout] fp32) and
 *   the bias [cout] fp32.  The image is only meaningful to c_conv called with the same cout, cin, kh, kw.
 * c_conv: out[:, out_channel0 : +cout] = relu?(conv(x) + bias); x planar-8 [n, cin, h, w] (cin == 1: the fp32 patches
 *   [n,1,h,w], 3x3 / stride 2 / no padding only); any kernel size / stride / padding; cin, cout multiples of 16.
 *   Stride-1 layers run the window kernel (the input window of a tile is staged once in shared memory, every tap is
 *   a shifted tcgen05 operand descriptor), the others the gather kernel; the choice is internal.
 * c_pool: 3x3 max pool (is_avg 0; F.max_pool2d(x, 3, stride), no padding) or 3x3 average pool that counts the padding
 *   (is_avg 1; F.avg_pool2d(x, 3, stride, pad)), written at out_channel0 of the destination tensor.
 * c_head: logits = Linear(mean over the hw positions)  (AdaptiveAvgPool2d(1) + flatten + fc, modelC:137-146).
 * c_read: debug / tests: planar-8 channels [channel0, +channels) -> fp32 [n, channels, hw]. */
MTLDAS_API int mtldas_c_pack_conv(const float* dev_w, const float* dev_gamma, const float* dev_beta, const float* dev_rmean,
                                  const float* dev_rvar, float eps, int cout, int cin, int kh, int kw, void* dev_packed,
                                  float* dev_bias, void* stream);
MTLDAS_API int mtldas_c_conv(const void* dev_x, int n, int cin, int h, int w, const void* dev_packed, const float* dev_bias,
                             void* dev_out, int out_channels_total, int out_channel0, int cout, int kh, int kw, int sh,
                             int sw, int ph, int pw, int relu, void* stream);
MTLDAS_API int mtldas_c_pool(const void* dev_x, int n, int channels, int h, int w, void* dev_out, int out_channels_total,
                             int out_channel0, int stride, int pad, int is_avg, void* stream);
MTLDAS_API int mtldas_c_head(const void* dev_x, int n, int channels, int hw, const float* dev_fc_w, const float* dev_fc_b,
                             int classes, float* dev_logits, void* stream);
MTLDAS_API int mtldas_c_read(const void* dev_x, int n, int channels_total, int channel0, int channels, int hw,
                             float* dev_out_nchw, void* stream);

/* streaming inference over a large DAS array --------------------------------------------------------
 * No counterpart in the reference (its README only motivates it): cuts the [100 x 250] patches the
 * network expects (utils.py:128) out of dev_array [n_channels, n_time] fp32 row-major: window wi covers fibre
 * channels first_channel + wi*window_stride .. +100, time block tb covers samples tb*250 .. +250
 * (n_time / 250 full blocks); dev_patches receives [n_windows * (n_time/250), 1, 100, 250], window-major.
 * Shard windows across GPUs with no collective (mtldas_b200.ddp.shard_windows). */
/* The same forward WITHOUT materialising the patches: the stem reads every window's 100 rows x 250 samples straight
 * out of dev_array (one bulk copy per input row, row pitch n_time), outputs [n_windows * (n_time/250), classes],
 * window-major.  Each array element is read once per window that covers it instead of being copied into a patch
 * tensor and read again.  bf16 plan only; dev_array 16-byte aligned and n_time a multiple of 4 (else
 * MTLDAS_ERR_UNSUPPORTED: gather with mtldas_window_gather and call mtldas_plan_forward_eval). */
MTLDAS_API int mtldas_plan_forward_eval_windows(mtldas_plan* plan, const float* dev_array, long long n_channels,
                                                long long n_time, int first_channel, int window_stride, int n_windows,
                                                float* dev_out0, float* dev_out1, void* stream);
MTLDAS_API int mtldas_window_gather(const float* dev_array, long long n_channels, long long n_time, int first_channel,
                                    int window_stride, int n_windows, float* dev_patches, void* stream);

/* tcgen05 descriptor self-test: runs one 128xNx16 UMMA on known operands under the descriptor
 * conventions the library uses and returns the max abs error (or <0 on error). */
MTLDAS_API float mtldas_umma_selftest(int n_cols, int row_shift, int swap_lbo_sbo);

/* tensor-pipe rate probe (diagnostics): cycles per 128 x n_cols x 16 MMA; host_out2[0] = to completion,
 * host_out2[1] = issue only. mode 0 interleaved layout, 1 swizzle-128B, 2 interleaved + per-MMA address math. */
MTLDAS_API int mtldas_umma_rate(int n_cols, int mode, int reps, float* host_out2);

#ifdef __cplusplus
}
#endif
#endif /* MTLDAS_H_ */
