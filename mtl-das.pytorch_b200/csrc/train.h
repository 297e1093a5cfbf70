// Host-side training plan (see train.cu).
#pragma once
#include <map>
#include <set>
#include <string>
#include <vector>

#include "plan.h"
#include "train_types.h"

namespace mtldas {

enum : int {
  TOP_STEM = 0, TOP_CONV, TOP_EW, TOP_HEAD,              // forward (BatchNorm statistics + finalize live inside TOP_EW)
  TOP_HEADBWD, TOP_EWB, TOP_WGRAD, TOP_DGRAD, TOP_STEMWG  // backward (BatchNorm gradients are written by TOP_EWB)
};

struct TrainPlan {
  struct TOp {
    int kind, a, b;
  };
  struct ParamOff {
    size_t w, b, gamma, beta;  // offsets (floats) into the flat gradient buffer, named_parameters order
  };
  struct BnSeg {
    int layer, ch_off, count;
  };
  struct BnHost {
    int tensor, C, H, W;
    std::vector<BnSeg> segs;
    size_t sum_off, red_off, f_off;
  };
  struct EwHost {
    int mode, ntasks, nchunks, conv, sc_conv;
    int y, bn, y2, bn2, aux, out;
    int dz, dy, dy2, daux;
    bool daux_accum;
  };
  struct WgHost {
    int conv, group, dy, inf_conv;
    size_t dw_off;
  };
  struct TSeg {
    int layer, ci_lo, ci_n, k_off, n_off;
    int group = 0;           // group of the input-gradient conv this block of transposed weights belongs to
  };
  struct DgradPack {
    int conv;
    int tap_map[9];
    std::vector<TSeg> segs;
  };

  Plan* p = nullptr;
  std::vector<ParamOff> param_off;
  size_t n_params = 0;
  std::vector<BnHost> bns;
  std::vector<EwHost> ews;
  std::vector<WgHost> wgs;
  std::vector<DgradPack> dgrads;
  std::vector<TOp> fwd, bwd;
  std::vector<int> y_of, bn_of, ew_of, fconv_of, dy_of;  // indexed by inference conv op
  std::map<int, int> grad_tensor;                         // activation tensor -> its gradient tensor
  std::set<long long> written;
  int y_stem = -1, bn_stem = -1, ew_stem = -1, dy_stem = -1, n_inf_convs = 0;
  size_t zero_fwd_off = 0, zero_fwd_end = 0, zero_bwd_off = 0, zero_bwd_end = 0;
  size_t stem_dw_off = 0, zero_bias_off = 0, logp_off = 0, pack_off = 0, pack_end = 0;
  struct GraphEntry {
    std::vector<const void*> key;
    int nb;
    cudaGraphExec_t exec;
    int uses, launches;
  };
  std::vector<GraphEntry> fwd_graphs, bwd_graphs;
  cudaStream_t cap_stream = nullptr;
  // weight gradients run on a side stream (forked after the layer's dy is complete, joined before the unpack):
  // they are off the dgrad -> elementwise chain, so the small latency-bound launches overlap
  cudaStream_t side_stream = nullptr;
  std::vector<cudaEvent_t> fork_events;
  cudaEvent_t join_event = nullptr;
  // state of the last forward
  int nb = 0, launches = 0;
  const float* x = nullptr;
  bool have_forward = false;

  ~TrainPlan();
};

TrainPlan* build_train_plan(int kind, int batch, int precision);
int train_forward(TrainPlan& tp, const void* const* state, const float* x, int nb, float* out0, float* out1,
                  cudaStream_t st);
int train_backward(TrainPlan& tp, const void* const* state, const float* dout0, const float* dout1, float* grads,
                   cudaStream_t st);

}  // namespace mtldas
