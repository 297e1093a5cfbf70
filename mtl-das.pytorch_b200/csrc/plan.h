// Host-side execution plan: the layer program of MTL_Net / Single_Task_Net expressed over the
// P8 / P8S2 activation layouts (types.h), the workspace map, and the launch sequence.
#pragma once
#include <cuda_runtime.h>

#include <string>
#include <vector>

#include "types.h"

namespace mtldas {

constexpr int kPatchH = 100, kPatchW = 250;  // utils.py:128 (input patch shape)
constexpr int kMaxTile = 512;                // largest CTA tile in positions (guard sizing)

struct LayerSpec {  // one (conv, bn) pair of the reference, in state_dict order
  std::string conv, bn;
  int cin, cout, ksize, stride, pad;
  bool bias;
  int i_w, i_b, i_gamma, i_beta, i_rm, i_rv, i_nbt;  // indices into the state_dict entry list
};

struct TensorInfo {
  std::string name;
  int C, H, W;     // logical (full resolution) shape per sample; C multiple of 16
  int layout;      // LAY_P8 / LAY_P8S2
  int Hs, Ws;      // stored plane geometry (== H,W for P8; ceil(H/2), ceil(W/2) for P8S2)
  int Wp, Pimg;    // Ws+1, (Hs+1)*Wp
  int g_lo;
  long long plane; // positions per chunk plane
  int nchunks, nphase;
  size_t offset, bytes;
};

struct SrcRef {
  int tensor, chunk0, phase;  // phase -1: P8 source
};
struct PackSegHost {
  int layer;                  // index into layers
  int split, k_off, cout_off;
  int co_lo = 0, co_n = 0;    // co_n > 0: only output channels [co_lo, co_lo+co_n) of the layer (N-split group)
};
struct ConvGroupHost {
  SrcRef src[kMaxSrc];
  int out_tensor, out_chunk0;
  int pool_tensor = -1;       // >= 0: fused ReLU + max-pool epilogue writes this pooled tensor (out_tensor unused)
  int out_phase = -1;         // >= 0: write one phase plane of a P8S2 tensor (stride-2 dgrad)
  int aux_tensor, aux_chunk0, aux_nchunks;
  size_t w_off, bias_off;     // into the workspace
  std::vector<PackSegHost> segs;
};
struct ConvOpHost {
  std::string name;
  int H, W, Cout, nchunks, nsrc, ntaps, halo_lo, halo_hi, epi, ngroups;
  int Hv = 0, Wv = 0;         // valid extents when they differ from the plane geometry (0 = H, W)
  ConvTapDev tap[kMaxTaps];
  ConvGroupHost g[kMaxGroups];
  int MT, tps, NS, nstages, tap_bytes, tmem_cols, ksplit;
};
struct PoolOpHost {
  std::string name;
  int in_tensor, out_tensor;
};
struct HeadOpHost {
  int in_tensor, chunk0, classes, out_index;
};
struct Op {
  int kind;  // 0 stem, 1 conv, 2 pool, 3 head
  int index;
};

struct StemSource {   // see StemParams::x_stride_a ... (types.h); nt == 0: contiguous patches
  long long stride_a = 0, stride_b = 0, row_pitch = 0;
  int nt = 0, n0 = 0;
};

struct Plan {
  int kind, mb, precision, ntasks;
  bool training = false;   // plan built for the training program (no inference-only re-groupings)
  bool use_nsplit = true;
  int nsplit_mode = 2;
  bool pool_split = true;
  int gran;  // bytes per granule (16 bf16 / 32 fp32)
  std::vector<std::string> state_keys;
  std::vector<LayerSpec> layers;
  std::vector<TensorInfo> tensors;
  std::vector<ConvOpHost> convs;
  std::vector<PoolOpHost> pools;
  std::vector<HeadOpHost> heads;
  std::vector<Op> program;
  int stem_layer, stem_out;
  size_t stem_w_off, stem_b_off, stem_wu_off = 0;
  int use_umma_stem = 0;
  size_t x_stage_off, out_stage_off, xh_stage_off = 0;
  int host_chunk = 0;
  size_t ws_bytes;
  char* ws = nullptr;
  bool packed = false;
  int last_launches = 0;
  int swap_desc = 0;
  // CUDA graphs of whole forwards, keyed by the caller's buffers (replayed when the same buffers come back)
  struct GraphEntry {
    const float* x;
    int n;
    float *o0, *o1;
    cudaGraphExec_t exec;
    int uses, launches;
    unsigned long long stamp;
  };
  std::vector<GraphEntry> graphs;
  unsigned long long graph_clock = 0;
  int use_graphs = 1;
  int use_pdl = 1;
  int use_persist = 1;
  cudaStream_t cap_stream = nullptr;
  // host entry point: private copy stream and events so that the H2D of pass i+1 overlaps the compute of pass i
  cudaStream_t copy_stream = nullptr;
  cudaEvent_t ev_start = nullptr, ev_h2d[2] = {nullptr, nullptr}, ev_comp[2] = {nullptr, nullptr};
  ~Plan();

  int find_tensor(const std::string& n) const;
  ActView view(int tensor, int chunk0) const;
  ConvParams conv_params(const ConvOpHost& op, int nb) const;
  std::string describe() const;
};

Plan* build_plan(int kind, int mb, int precision, bool training = false);
int add_tensor(Plan& p, const std::string& name, int C, int H, int W, int layout);
PackSeg make_seg(const Plan& p, const PackSegHost& h, const void* const* state);
void choose_tiling(ConvOpHost& op, long long positions = 0, bool small_resident = false);  // positions > 0: batch-aware tile count
bool split_output_channels(ConvOpHost& op, long long positions, bool even_if_resident = false);  // N-split into two groups when that makes the weights resident
int persist_stages_for(const ConvOpHost& op, size_t extra);
int plan_pack_eval(Plan& p, const void* const* state, cudaStream_t stream);
int plan_forward_eval(Plan& p, const float* x, int n, float* out0, float* out1, cudaStream_t stream);

// Forward over the windows of a streaming array read IN PLACE by the stem (no patch tensor): window wi = fibre channels
// first_channel + wi*window_stride .. +100, time block tb = samples tb*250 .. +250; outputs window-major.
int plan_forward_windows(Plan& p, const float* array, long long n_channels, long long n_time, int first_channel,
                         int window_stride, int n_windows, float* out0, float* out1, cudaStream_t stream);

int plan_profile_eval(Plan& p, const float* x, int n, float* out0, float* out1, cudaStream_t stream,
                      float* ms_per_op, int cap);
int plan_read_tensor(Plan& p, const char* name, int nb, float* out_nchw, cudaStream_t stream);
int single_conv(const float* x, const float* x2, const float* weight, const float* bias, const float* aux,
                float* out, int n, int cin, int cin2, int h, int w, int cout, int ksize, int stride, int epi,
                int precision, int aux_phase_split, int swap_desc);

void set_error(const char* fmt, ...);
const char* last_error_text();

}  // namespace mtldas
