// Parameter blocks of the training-mode kernels (batch-statistics BatchNorm forward, fused elementwise
// backward, weight gradients, heads).  See types.h for the P8 / P8S2 activation layouts.
#pragma once
#include "types.h"

namespace mtldas {

enum : int { EPI_ACCUM = 4 };  // conv epilogue: out += conv (gradient accumulation, no bias)

// elementwise modes (what follows the BatchNorm of a raw conv output y)
enum : int {
  EW_RELU = 0,         // z = relu(bn(y))                                   modelA_MTL.py:16,47,76
  EW_ADD_RELU_ID = 1,  // z = relu(bn(y) + r), r an activation tensor       modelA_MTL.py:28-30 (identity shortcut)
  EW_ADD_RELU_SC = 2,  // z = relu(bn(y) + bn2(y2)), y2 the raw shortcut conv  modelA_MTL.py:22-25,28-30
  EW_SIGMOID_MUL = 3,  // a = sigmoid(bn(y)) * feat                         modelA_MTL.py:50,144,151,157,163
  EW_RELU_POOL = 4,    // p = maxpool2x2_ceil(relu(bn(y)))                  modelA_MTL.py:102,116,145
};

// per-channel BatchNorm work arrays of one raw tensor (all its channels, tasks stacked)
struct BnArrays {
  double* sum;     // [C] sum of y over valid positions
  double* sumsq;   // [C]
  double* red_g;   // [C] sum of g            (backward)   -> d beta
  double* red_gy;  // [C] sum of g * yhat     (backward)   -> d gamma
  float* scale;    // [C] gamma * invstd
  float* shift;    // [C] beta - mean * scale
  float* mean;     // [C]
  float* invstd;   // [C]
};

struct BnGradSeg {       // d gamma / d beta of one reference BatchNorm out of a tensor's reduction arrays
  float* dgamma;
  float* dbeta;
  int ch_off, count;
};

struct FinalizeSeg {     // channels [ch_off, ch_off+count) of the tensor belong to one reference BatchNorm
  const float* gamma;
  const float* beta;
  float* running_mean;   // updated in place (momentum 0.1, unbiased variance)
  float* running_var;
  long long* num_batches_tracked;
  const float* conv_bias;  // pre-BN conv bias (cancels in the output; only shifts running_mean) or null
  int ch_off, count;
};
struct FinalizeParams {
  BnArrays bn;
  int C;
  double n;              // nb*H*W
  float eps, momentum;
  int nseg;
  FinalizeSeg seg[2];
};

struct EwParams {
  int mode;
  int ntasks;            // tasks stacked along channels that share `aux` (EW_SIGMOID_MUL), else 1
  int nchunks;           // chunks per task of y / out / dz / dy
  int H, W, Wp, Pimg;    // geometry of y (full resolution)
  int Ho, Wo, Wpo, Pimgo;  // pooled geometry (EW_RELU_POOL)
  long long body;        // nb * Pimg   (EW_RELU_POOL: nb * Pimgo, threads iterate pooled positions)
  float inv_n;           // 1 / (nb*H*W)
  ActView y;             // raw conv output
  BnArrays bn;
  ActView y2;            // EW_ADD_RELU_SC: raw shortcut conv output
  BnArrays bn2;
  ActView aux;           // identity residual / feature map (EW_ADD_RELU_ID, EW_SIGMOID_MUL)
  ActView out;           // forward result
  ActView dz;            // backward: gradient w.r.t. the forward result
  ActView dy;            // backward: gradient w.r.t. y (written, zero at pads)
  ActView dy2;           // backward, EW_ADD_RELU_SC: gradient w.r.t. y2
  ActView daux;          // backward: gradient w.r.t. aux
  int daux_accum;        // 1: daux += , 0: daux =
  int want_daux;
  int ngseg;             // the backward kernel also writes d gamma / d beta of the BatchNorm it differentiates
  BnGradSeg gseg[2];
  int ngseg2;            // EW_ADD_RELU_SC: the shortcut BatchNorm's d gamma / d beta
  BnGradSeg gseg2[2];
  unsigned long long magic_pimg, magic_wp, magic_pimgo, magic_wpo;  // floor(2^40/d)+1 (exact f/d by multiply-shift)
  // fused kernels: statistics -> per-chunk barrier -> apply (forward), reductions -> barrier -> gradients (backward)
  FinalizeParams fin, fin2;  // forward: BatchNorm finalize of bn (and bn2, EW_ADD_RELU_SC) done inside the kernel
  unsigned int* sync;        // [chunks] arrival counters, zeroed with the sums (forward) / reductions (backward)
  long long body_full;       // nb * Pimg (== body except EW_RELU_POOL, where body counts pooled positions)
};


struct WgradParams {     // dW[tap][k][cout] += sum_pos zin[pos + off_tap][k] * dy[pos][cout]   (fp32 result)
  ConvSrcDev src[kMaxSrc];   // forward input sources (as in ConvParams, body position 0)
  ConvTapDev tap[kMaxTaps];
  const char* dy;            // chunk 0, body position 0
  long long dy_plane_bytes;
  float* dw;                 // [ntaps][K][Cout] fp32, zeroed before the launch
  int ntaps, nchunks, Cout;  // K = 8*nchunks
  long long body;
  int slice;                 // positions per CTA
  int nsrc;                  // sources referenced by the taps
  int taps_per_cta, halo_lo, halo_hi;  // tensor-core kernel only (filled by the launcher)
  int nring;                 // wide-N kernel: stages of the dy / input-slab ring (2..4)
  int dbg_flags;             // diagnostics (MTLDAS_WG2_DEBUG): 1 = skip the re-lay copies, 2 = skip the MMAs (results invalid)
};

struct HeadBwdParams {   // d(log_softmax(group mean(GAP(a4)))) -> d a4
  const float* logp;     // [nb, classes] forward output
  const float* dlogp;    // [nb, classes] upstream
  ActView da4;           // [nb,128,5,11] view (task chunk offset applied)
  int nb, classes, H, W, Wp, Pimg;
};

struct StemWgradParams { // dW[49][16] of the stem from the fp32 input and dy
  const float* x;        // [nb,1,Hin,Win]
  ActView dy;            // [nb,16,H,W]
  float* dw;             // [49][16] zeroed before the launch
  int Hin, Win, H, W, Wp, Pimg;
  long long body;
};

struct UnpackParams {    // packed fp32 gradient [tap][K][Cout_p] -> reference layout [cout][cin][k][k]
  const float* src;
  float* dst;
  int cout, cin, ksize, split, k_off, cout_off, Cout_p, K_tap;
};

constexpr int kUnpackBatch = 96;
struct UnpackBatch {
  UnpackParams u[kUnpackBatch];
};

}  // namespace mtldas
