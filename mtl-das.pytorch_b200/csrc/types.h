// Host/device shared parameter blocks of the mtldas kernels.
//
// Activation layout "P8" (planar, 8-channel granules, flattened zero-padded positions):
//   a tensor [nb, C, H, W] is stored as C/8 planes; a plane is a 1-D array of `plane` positions;
//   position index = g_lo + n*Pimg + h*Wp + w with Wp = W+1, Pimg = (H+1)*Wp; one position of one
//   plane is a granule of 8 consecutive channels (16 bytes bf16 / 32 bytes fp32).  Row h==H and
//   column w==W of every image, and the guard positions before/after the body, are ZERO: they are
//   the convolution's zero padding (right pad of row h == left pad of row h+1, bottom pad of image
//   n == top pad of image n+1).  A 3x3/s1/p1 convolution is then nine GEMMs whose A operands are the
//   same plane shifted by dr*Wp+ds positions -- i.e. nine start addresses into ONE shared-memory
//   slab -- and a channel concat is just two sources.
// Layout "P8S2" (input of a stride-2 convolution, phase split):
//   four P8-like plane sets, one per (h&1, w&1) phase, each with the geometry of the stride-2
//   OUTPUT (Ho = ceil(H/2), Wo = ceil(W/2)); element (n,c,h,w) lives in phase (h&1)*2+(w&1) at
//   position g_lo + n*Pimg2 + (h>>1)*Wp2 + (w>>1).  A 3x3/s2/p1 conv becomes nine shifted GEMMs
//   over the four phase planes with shifts in {-Wp2-1, -Wp2, -1, 0}.
#pragma once
#include <cstdint>

namespace mtldas {

constexpr int kMaxSrc = 4;
constexpr int kMaxTaps = 9;
constexpr int kMaxGroups = 4;  // tasks x output-channel halves (N-split of layers whose weights would not stay resident)
constexpr int kTileM = 128;
constexpr int kMaxRing = 9;

enum : int { LAY_P8 = 0, LAY_P8S2 = 1 };
enum : int { EPI_NONE = 0, EPI_RELU = 1, EPI_ADD_RELU = 2, EPI_SIGMOID_MUL = 3 };

struct ActView {
  char* base;              // chunk 0 (of this view), phase 0, position index 0 (first guard position)
  long long plane;         // positions per chunk plane
  long long phase_stride;  // positions between phase plane sets (P8S2), else 0
  int g_lo;                // guard positions before the body
  int layout;              // LAY_P8 / LAY_P8S2
  int Wp2, Pimg2;          // phase plane geometry (P8S2 only)
};

struct ConvSrcDev {
  const char* base;        // first chunk of the source view, body position 0 (guard and phase applied)
  long long plane_bytes;   // bytes between chunk planes
};
struct ConvTapDev {
  int src;                 // source slot
  int off;                 // position shift of this tap
};
struct ConvGroupDev {
  ConvSrcDev src[kMaxSrc];
  const char* w;           // packed weights of this group
  const float* bias;       // [Cout] fp32 (folded BN shift / conv bias)
  ActView out;
  ActView aux;
  int aux_nchunks;         // aux chunk = (output chunk) % aux_nchunks (tasks sharing one feature map)
  int pad_;
  ActView pool_out;        // fused ReLU + 2x2 ceil max-pool: pooled tensor (persistent kernel, pool_k > 0)
};
struct ConvParams {
  int H, W, Wp, Pimg;      // output geometry
  long long body;          // nb * Pimg
  int Cout;                // multiple of 16
  int nchunks;             // 8-channel chunks per source (K per tap = 8*nchunks)
  int nsrc, ntaps;
  int halo_lo, halo_hi;    // slab halo in positions
  int MT;                  // 128-row accumulator tiles per CTA
  int SP;                  // slab positions = MT*128 + halo_lo + halo_hi
  int tps;                 // taps per weight stage
  int NS;                  // weight ring slots
  int nstages;             // ceil(ntaps / tps)
  int tap_bytes;           // 8*nchunks*Cout*2 (bf16)
  int tmem_cols;           // power of two >= max(32, MT*Cout)
  int epi;
  int swap_desc;           // debug: swap the LBO/SBO roles in the UMMA descriptors
  int ngroups;
  unsigned long long magic_pimg, magic_wp;  // floor(2^40/d)+1: exact f/d for f < 2^22 by multiply-shift
  int pool_k;              // > 0: fused max-pool, the CTA tile is pool_k image-row pairs (2*Wp*pool_k positions)
  int pool_Wpo, pool_Pimgo; // pooled geometry
  int ksplit;              // ring kernel: even/odd K steps accumulate into two TMEM sets (summed in the epilogue)
  int acc_sets;            // persistent kernel: TMEM accumulator sets (1 or 2), chosen by the launcher
  int persist_stages;      // > 0: resident weights, persistent pipelined kernel with this many slab stages
  long long* dbg;          // optional [grid.x][8] timeline (globaltimer ns) for diagnostics, else null
  ConvTapDev tap[kMaxTaps];
  ConvGroupDev g[kMaxGroups];
};

struct StemParams {        // 7x7 / stride 3 / pad 2, Cin = 1 (modelA_MTL.py:74)
  const float* x;          // [nb,1,Hin,Win] fp32
  const float* w;          // [49][16] fp32 (folded)
  const float* bias;       // [16]
  ActView out;
  int Hin, Win, H, W, Wp, Pimg;
  long long body;
  int raw;                 // 1: store conv + bias without the ReLU (training: BatchNorm follows as its own pass)
  const char* w_umma;      // bf16 UMMA image [8 K-chunks][16 cout][8] (K = 49 taps padded to 64) or null
  // Where image n's samples live, in floats from x (all zero = a contiguous patch tensor [nb,1,Hin,Win]):
  //   first sample of image n = ((n + x_n0) / x_nt) * x_stride_a + ((n + x_n0) % x_nt) * x_stride_b, rows x_row_pitch apart.
  // A streaming array [channels, n_time] read in place (no patch materialisation): x_nt = time blocks per window,
  // x_stride_a = window_stride * n_time, x_stride_b = Win, x_row_pitch = n_time, x pointing at the first window's row.
  long long x_stride_a, x_stride_b, x_row_pitch;
  int x_nt, x_n0;
  int raw_stage_bytes;     // tensor-core stem: bytes of one fp32 staging stage (set by its launcher)
};

struct PoolParams {        // MaxPool2d(2,2,ceil_mode=True) (modelA_MTL.py:116)
  ActView in, out;
  int nchunks;
  int H, W, Wp, Pimg;      // input geometry
  int Ho, Wo, Wpo, Pimgo;  // output geometry
  long long body_o;
};

struct HeadParams {        // GAP + AvgPool1d + log_softmax (modelA_MTL.py:165-172)
  ActView in;              // [nb,128,5,11]
  float* out;              // [nb, classes]
  int nb, classes, npos;   // npos = Pimg (pads are zero so the whole image block is summed)
  float inv_count;         // 1 / (H*W*group)
};

struct PackSeg {           // one source layer scattered into a packed weight block
  const float* w;          // [cout, cin, k, k]
  const float* conv_bias;  // [cout] or null
  const float* gamma;      // BN weight
  const float* beta;       // BN bias
  const float* rmean;      // BN running_mean
  const float* rvar;       // BN running_var
  int cout, cin, ksize;
  int split;               // cin >= split go to tap+1 (dual-source 1x1), else INT_MAX
  int k_off, cout_off;     // placement inside the packed [K_tap x Cout_p] tap matrix
  // dgrad packing (transpose = 1): packed K runs over the layer's OUTPUT channels (at k_off), packed N over its
  // input channels [ci_lo, ci_lo+ci_n) (at cout_off); tap_map selects / reorders taps (-1 = tap not used)
  int transpose, ci_lo, ci_n;
  int tap_map[9];
  int k_from_tap;          // stem: the filter taps ARE the K dimension (k = r*ksize+s, single packed tap)
  int co_lo, co_n;         // co_n > 0: only output channels [co_lo, co_lo+co_n) of the layer go into this block (N-split)
};
struct PackParams {
  PackSeg seg;
  char* dst;               // packed block
  float* bias_dst;         // [Cout_p]
  int Cout_p, K_tap;
  int fp32;                // 1: fp32 [tap][K][Cout] ; 0: bf16 UMMA image [tap][K/8][Cout][8]
  int fold;                // 1: fold eval BN
  float eps;
};

constexpr int kPackBatch = 128;
struct PackBatch {          // many independent packs in one launch (blockIdx.y selects the entry)
  PackParams p[kPackBatch];
};

}  // namespace mtldas
