// Training-mode plan: forward with batch-statistics BatchNorm and the hand-scheduled backward pass of
// MTL_Net / Single_Task_Net (what autograd does for /root/reference/model/modelA_MTL.py:127-174 when
// utils.py:359-373 calls model(x), NLLLoss + NLLLoss, loss.backward()).
//
// The training program is derived mechanically from the inference plan's layer program (plan.cu), which fixes
// sources, taps, channel slices and weight placement:
//   forward, per conv op   : conv (raw output y, no bias) -> batch statistics -> finalize -> fused BN+activation
//   backward, reverse order: fused activation/BN backward (two passes: reductions, then dy) -> d gamma/d beta
//                            -> weight gradient -> input gradients (the same conv kernels with transposed weights
//                            and negated tap shifts; stride-2 convs scatter into the four phase planes)
// A tensor's first gradient contribution overwrites, later ones accumulate (EPI_ACCUM / daux_accum).
#include "train.h"

#include <algorithm>
#include <climits>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sstream>

#include "kernels.h"
#include "train_kernels.h"

namespace mtldas {

#define CK(call)                                                                             \
  do {                                                                                       \
    cudaError_t e_ = (call);                                                                 \
    if (e_ != cudaSuccess) {                                                                 \
      set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
      return -2;                                                                             \
    }                                                                                        \
  } while (0)

static inline size_t round_up_sz(size_t v, size_t a) { return (v + a - 1) / a * a; }

// MTLDAS_TRAIN_PROFILE=1: CUDA-event pair around every op, table of device milliseconds per op kind on stderr
struct OpTimer {
  bool on;
  cudaStream_t st;
  std::vector<cudaEvent_t> ev;
  std::vector<int> kind;
  explicit OpTimer(cudaStream_t s) : st(s) {
    const char* e = getenv("MTLDAS_TRAIN_PROFILE");
    on = e && e[0] == '1';
  }
  void begin(int k) {
    if (!on) return;
    cudaEvent_t a;
    cudaEventCreate(&a);
    cudaEventRecord(a, st);
    ev.push_back(a);
    kind.push_back(k);
  }
  void end() {
    if (!on) return;
    cudaEvent_t a;
    cudaEventCreate(&a);
    cudaEventRecord(a, st);
    ev.push_back(a);
  }
  void report(const char* what) {
    if (!on) return;
    cudaStreamSynchronize(st);
    static const char* names[] = {"stem", "conv", "bn+act", "head", "head_bwd", "bn+act_bwd", "wgrad", "dgrad",
                                  "stem_wgrad", "", "", ""};
    float tot[12] = {0};
    int cnt[12] = {0};
    for (size_t i = 0; i < kind.size(); ++i) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, ev[2 * i], ev[2 * i + 1]);
      tot[kind[i]] += ms, cnt[kind[i]]++;
    }
    float all = 0.f;
    for (float t : tot) all += t;
    fprintf(stderr, "[trainprof] %s total %.3f ms:", what, all);
    for (int k = 0; k < 12; ++k)
      if (cnt[k]) fprintf(stderr, " %s x%d %.3f;", names[k], cnt[k], tot[k]);
    fprintf(stderr, "\n");
    for (auto e : ev) cudaEventDestroy(e);
  }
};

static int clone_tensor(Plan& p, const std::string& name, int src, int layout) {
  const TensorInfo t = p.tensors[src];
  return add_tensor(p, name, t.C, t.H, t.W, layout);
}

// gradient tensor of an activation tensor (created on demand, same shape and layout)
static int grad_of(TrainPlan& tp, int tensor) {
  if (tp.grad_tensor.count(tensor)) return tp.grad_tensor[tensor];
  Plan& p = *tp.p;
  int g = clone_tensor(p, "d:" + p.tensors[tensor].name, tensor, p.tensors[tensor].layout);
  tp.grad_tensor[tensor] = g;
  return g;
}

// first contribution to (tensor, phase) overwrites, later ones accumulate; phase -1 = whole tensor
static bool claim(TrainPlan& tp, int tensor, int chunk0, int nch, int phase) {
  bool seen = false;
  for (int c = chunk0; c < chunk0 + nch; ++c)
    for (int ph = 0; ph < 4; ++ph) {
      if (phase >= 0 && ph != phase) continue;
      long long key = ((long long)tensor << 16) | (c << 2) | ph;
      if (tp.written.count(key)) seen = true;
      tp.written.insert(key);
    }
  return seen;
}

TrainPlan* build_train_plan(int kind, int batch, int precision) {
  TrainPlan* tpp = new TrainPlan();
  TrainPlan& tp = *tpp;
  tp.p = build_plan(kind, batch, precision, /*training=*/true);
  Plan& p = *tp.p;
  const int T = p.ntasks;
  const std::vector<Op> inf_prog = p.program;
  const size_t n_inf_convs = p.convs.size();

  // parameter offsets in the flat gradient buffer (named_parameters order)
  size_t off = 0;
  for (auto& L : p.layers) {
    TrainPlan::ParamOff po{};
    po.w = off, off += (size_t)L.cout * L.cin * L.ksize * L.ksize;
    po.b = (size_t)-1;
    if (L.bias) po.b = off, off += L.cout;
    po.gamma = off, off += L.cout;
    po.beta = off, off += L.cout;
    tp.param_off.push_back(po);
  }
  tp.n_params = off;

  auto add_bn = [&](int tensor, int H, int W) {
    TrainPlan::BnHost b{};
    b.tensor = tensor, b.C = p.tensors[tensor].C, b.H = H, b.W = W;
    tp.bns.push_back(b);
    return (int)tp.bns.size() - 1;
  };

  // ---- stem ------------------------------------------------------------------------------------------
  tp.y_stem = clone_tensor(p, "y:stem", p.stem_out, LAY_P8);
  tp.bn_stem = add_bn(tp.y_stem, p.tensors[p.stem_out].H, p.tensors[p.stem_out].W);
  tp.bns[tp.bn_stem].segs.push_back(TrainPlan::BnSeg{p.stem_layer, 0, 16});
  {
    TrainPlan::EwHost e{};
    e.mode = EW_RELU, e.ntasks = 1, e.nchunks = 2, e.y = tp.y_stem, e.bn = tp.bn_stem, e.out = p.stem_out;
    e.y2 = e.aux = e.bn2 = -1;
    tp.ews.push_back(e);
    tp.ew_stem = 0;
  }
  tp.fwd.push_back({TOP_STEM, 0, 0});
  tp.fwd.push_back({TOP_EW, tp.ew_stem, 0});

  // ---- forward program from the inference program -----------------------------------------------------
  tp.y_of.assign(n_inf_convs, -1), tp.bn_of.assign(n_inf_convs, -1), tp.ew_of.assign(n_inf_convs, -1);
  tp.fconv_of.assign(n_inf_convs, -1);
  std::vector<int> conv_writing(p.tensors.size() + 64, -1);  // activation tensor -> inference conv producing it
  for (size_t oi = 0; oi < inf_prog.size(); ++oi) {
    const Op op = inf_prog[oi];
    if (op.kind == 3) {
      tp.fwd.push_back({TOP_HEAD, op.index, 0});
      continue;
    }
    if (op.kind != 1) continue;
    const int ci = op.index;
    const ConvOpHost inf = p.convs[ci];
    const int out_t = inf.g[0].out_tensor;
    conv_writing[out_t] = ci;
    const int y = clone_tensor(p, "y:" + inf.name, out_t, LAY_P8);
    tp.y_of[ci] = y;
    const int bn = add_bn(y, inf.H, inf.W);
    tp.bn_of[ci] = bn;
    for (int g = 0; g < inf.ngroups; ++g)
      for (auto& s : inf.g[g].segs)
        tp.bns[bn].segs.push_back(TrainPlan::BnSeg{s.layer, inf.g[g].out_chunk0 * 8 + s.cout_off, p.layers[s.layer].cout});
    ConvOpHost f = inf;  // same sources / taps / tiling; raw output, no bias, no fused activation
    f.name = "fwd:" + inf.name, f.epi = EPI_NONE;
    for (int g = 0; g < f.ngroups; ++g) f.g[g].out_tensor = y, f.g[g].aux_tensor = -1, f.g[g].pool_tensor = -1;
    p.convs.push_back(f);
    tp.fconv_of[ci] = (int)p.convs.size() - 1;
    tp.fwd.push_back({TOP_CONV, tp.fconv_of[ci], 0});
    if (inf.epi == EPI_NONE) continue;  // shortcut conv: its BatchNorm is applied by the consumer (EW_ADD_RELU_SC)
    TrainPlan::EwHost e{};
    e.conv = ci, e.y = y, e.bn = bn, e.y2 = e.bn2 = e.aux = -1, e.out = out_t;
    e.ntasks = 1, e.nchunks = p.tensors[out_t].nchunks;
    if (inf.epi == EPI_RELU) {
      e.mode = EW_RELU;
      if (inf.g[0].pool_tensor >= 0) {  // inference fuses the pool into this conv's epilogue
        e.mode = EW_RELU_POOL;
        e.out = inf.g[0].pool_tensor;
      } else if (oi + 1 < inf_prog.size() && inf_prog[oi + 1].kind == 2 && p.pools[inf_prog[oi + 1].index].in_tensor == out_t) {
        e.mode = EW_RELU_POOL;
        e.out = p.pools[inf_prog[oi + 1].index].out_tensor;
      }
    } else if (inf.epi == EPI_ADD_RELU) {
      const int a = inf.g[0].aux_tensor;
      const int sc_conv = conv_writing[a];
      if (sc_conv >= 0 && p.convs[sc_conv].epi == EPI_NONE) {
        e.mode = EW_ADD_RELU_SC, e.y2 = tp.y_of[sc_conv], e.bn2 = tp.bn_of[sc_conv], e.sc_conv = sc_conv;
      } else {
        e.mode = EW_ADD_RELU_ID, e.aux = a;
      }
    } else {
      e.mode = EW_SIGMOID_MUL, e.aux = inf.g[0].aux_tensor;
      e.nchunks = inf.g[0].aux_nchunks;
      e.ntasks = p.tensors[out_t].nchunks / e.nchunks;
    }
    tp.ews.push_back(e);
    tp.ew_of[ci] = (int)tp.ews.size() - 1;
    tp.fwd.push_back({TOP_EW, tp.ew_of[ci], 0});
  }

  // ---- backward program --------------------------------------------------------------------------------
  for (size_t hi = 0; hi < p.heads.size(); ++hi) {
    const HeadOpHost& h = p.heads[hi];
    const int d = grad_of(tp, h.in_tensor);
    claim(tp, d, h.chunk0, 16, -1);
    tp.bwd.push_back({TOP_HEADBWD, (int)hi, 0});
  }
  auto add_dgrad = [&](const ConvOpHost& inf, int ci, int g, int src_slot_lo, int src_slot_hi, bool stack_tasks,
                       bool all_groups = false) {
    // gradient w.r.t. source tensor(s) of fprop group g read through slots [lo,hi) (one tensor, maybe 4 phases).
    // all_groups: the tasks' input gradients (same layer shape, different channel slices of one tensor) go out as the
    // groups of ONE launch instead of one launch per task.
    const int G = all_groups ? inf.ngroups : 1;
    const SrcRef sr0 = inf.g[all_groups ? 0 : g].src[src_slot_lo];
    const int dsrc = grad_of(tp, sr0.tensor);
    const int nphase = src_slot_hi - src_slot_lo > 1 ? 4 : 1;
    for (int ph = 0; ph < nphase; ++ph) {
      ConvOpHost d{};
      d.name = "dgrad:" + inf.name + (nphase > 1 ? "/ph" + std::to_string(ph) : "") + "/s" + std::to_string(src_slot_lo);
      d.H = inf.H, d.W = inf.W, d.ngroups = G, d.nsrc = 1;
      d.Cout = inf.nchunks * 8;                                  // channels of the source
      d.nchunks = (stack_tasks ? inf.ngroups : 1) * inf.Cout / 8;  // K = output channels of the fprop conv
      TrainPlan::DgradPack dp{};
      int nt = 0, lo = 0, hi = 0;
      for (int t = 0; t < 9; ++t) dp.tap_map[t] = -1;
      for (int t = 0; t < inf.ntaps; ++t) {
        const int slot = inf.tap[t].src;
        if (slot < src_slot_lo || slot >= src_slot_hi) continue;
        if (nphase > 1 && slot - src_slot_lo != ph) continue;
        d.tap[nt] = ConvTapDev{0, -inf.tap[t].off};
        lo = std::max(lo, inf.tap[t].off), hi = std::max(hi, -inf.tap[t].off);
        // packed tap index of the ORIGINAL filter tap: dual-source 1x1 convs use tap 1 for the second source
        dp.tap_map[inf.ntaps == inf.nsrc && inf.nsrc <= 2 && inf.halo_lo == 0 ? 0 : t] = nt;
        ++nt;
      }
      if (nt == 0) continue;
      d.ntaps = nt, d.halo_lo = lo, d.halo_hi = hi;  // reads at f - off: positive original shifts reach backwards
      for (auto& gg : d.g) {
        gg.aux_tensor = -1, gg.aux_chunk0 = 0, gg.aux_nchunks = d.Cout / 8, gg.out_tensor = -1, gg.out_chunk0 = 0;
        for (auto& s : gg.src) s = SrcRef{-1, 0, -1};
      }
      bool accum = false;
      for (int gi = 0; gi < G; ++gi) {
        const int fgi = all_groups ? gi : g;
        const ConvGroupHost& fg = inf.g[fgi];
        const SrcRef sr = fg.src[src_slot_lo];
        d.g[gi].src[0] = SrcRef{tp.dy_of[ci], stack_tasks ? 0 : fg.out_chunk0, -1};
        d.g[gi].out_tensor = dsrc, d.g[gi].out_chunk0 = sr.chunk0;
        if (p.tensors[sr.tensor].layout == LAY_P8S2) {
          const int phase = nphase > 1 ? ph : sr.phase;
          d.g[gi].out_phase = phase;
          const TensorInfo& st = p.tensors[sr.tensor];
          d.Hv = (st.H - (phase >> 1) + 1) / 2, d.Wv = (st.W - (phase & 1) + 1) / 2;
        }
        const bool seen = claim(tp, dsrc, sr.chunk0, d.Cout / 8, d.g[gi].out_phase);
        if (gi == 0) accum = seen;
        // (the groups of a merged launch are the tasks' slices of one tensor at one point of the schedule: same state)
      }
      d.epi = accum ? (int)EPI_ACCUM : (int)EPI_NONE;
      choose_tiling(d, (long long)p.mb * (d.H + 1) * (d.W + 1));
      p.convs.push_back(d);
      dp.conv = (int)p.convs.size() - 1;
      // transposed weight placement
      for (int gi = 0; gi < G; ++gi) {
        const int fgi = all_groups ? gi : g;
        const int g_lo = stack_tasks ? 0 : fgi, g_hi = stack_tasks ? inf.ngroups : fgi + 1;
        for (int gj = g_lo; gj < g_hi; ++gj)
          for (auto& s : inf.g[gj].segs) {
            TrainPlan::TSeg ts{};
            ts.layer = s.layer;
            const bool dual = s.split != INT_MAX;
            ts.ci_lo = dual ? (src_slot_lo == 0 ? 0 : s.split) : 0;
            ts.ci_n = dual ? s.split : p.layers[s.layer].cin;
            ts.k_off = s.cout_off + (stack_tasks ? gj * inf.Cout : 0);
            ts.n_off = dual ? 0 : s.k_off;
            ts.group = gi;
            dp.segs.push_back(ts);
          }
      }
      tp.dgrads.push_back(dp);
      tp.bwd.push_back({TOP_DGRAD, (int)tp.dgrads.size() - 1, 0});
    }
  };

  const char* nomerge = getenv("MTLDAS_NO_DGRAD_MERGE");
  const bool merge_task_dgrads = !(nomerge && nomerge[0] == '1');
  auto same_source_tensor = [&](const ConvOpHost& inf, int slot) {
    for (int g = 1; g < inf.ngroups; ++g)
      if (inf.g[g].src[slot].tensor != inf.g[0].src[slot].tensor || inf.g[g].src[slot].phase != inf.g[0].src[slot].phase) return false;
    return true;
  };
  tp.dy_of.assign(n_inf_convs, -1);
  for (size_t ci = 0; ci < n_inf_convs; ++ci) tp.dy_of[ci] = clone_tensor(p, "dy:" + p.convs[ci].name, tp.y_of[ci], LAY_P8);
  tp.dy_stem = clone_tensor(p, "dy:stem", tp.y_stem, LAY_P8);

  for (int oi = (int)inf_prog.size() - 1; oi >= 0; --oi) {
    const Op op = inf_prog[oi];
    if (op.kind != 1) continue;
    const int ci = op.index;
    const ConvOpHost inf = p.convs[ci];
    if (tp.ew_of[ci] >= 0) {
      TrainPlan::EwHost& e = tp.ews[tp.ew_of[ci]];
      e.dz = grad_of(tp, e.out), e.dy = tp.dy_of[ci], e.dy2 = -1, e.daux = -1;
      if (e.mode == EW_ADD_RELU_ID || e.mode == EW_SIGMOID_MUL) {
        e.daux = grad_of(tp, e.aux);
        e.daux_accum = claim(tp, e.daux, 0, p.tensors[e.aux].nchunks, -1);
      }
      if (e.mode == EW_ADD_RELU_SC) e.dy2 = tp.dy_of[e.sc_conv];
      tp.bwd.push_back({TOP_EWB, tp.ew_of[ci], 0});
    }
    // weight gradients (one per group)
    for (int g = 0; g < inf.ngroups; ++g) {
      TrainPlan::WgHost w{};
      w.conv = tp.fconv_of[ci], w.group = g, w.dy = tp.dy_of[ci], w.inf_conv = ci;
      tp.wgs.push_back(w);
      tp.bwd.push_back({TOP_WGRAD, (int)tp.wgs.size() - 1, 0});
    }
    // input gradients
    if (inf.nsrc == 4) {
      add_dgrad(inf, ci, 0, 0, 4, false);
    } else if (inf.nsrc == 2) {
      add_dgrad(inf, ci, 0, 0, 1, true);                       // shared source: tasks stacked along K
      if (merge_task_dgrads && same_source_tensor(inf, 1)) add_dgrad(inf, ci, 0, 1, 2, false, true);
      else for (int g = 0; g < inf.ngroups; ++g) add_dgrad(inf, ci, g, 1, 2, false);
    } else {
      if (merge_task_dgrads && inf.ngroups > 1 && same_source_tensor(inf, 0)) add_dgrad(inf, ci, 0, 0, 1, false, true);
      else for (int g = 0; g < inf.ngroups; ++g) add_dgrad(inf, ci, g, 0, 1, false);
    }
  }
  // stem
  tp.ews[tp.ew_stem].dz = grad_of(tp, p.stem_out), tp.ews[tp.ew_stem].dy = tp.dy_stem;
  tp.ews[tp.ew_stem].dy2 = tp.ews[tp.ew_stem].daux = -1;
  tp.bwd.push_back({TOP_EWB, tp.ew_stem, 0});
  tp.bwd.push_back({TOP_STEMWG, 0, 0});
  (void)T;

  // ---- workspace map ------------------------------------------------------------------------------------
  size_t o = 0;
  auto take = [&](size_t bytes) {
    size_t r = o;
    o = round_up_sz(o + bytes, 256);
    return r;
  };
  for (auto& t : p.tensors) t.offset = take(t.bytes);
  tp.zero_fwd_off = o;  // [zero_fwd_off, zero_fwd_end): BN sums, cleared at every forward
  // per BatchNorm: sums [2C] double + one arrival counter per chunk (the fused kernels' barrier)
  for (auto& b : tp.bns) b.sum_off = take((size_t)b.C * 2 * sizeof(double) + round_up_sz((size_t)b.C / 8 * 4, 16));
  tp.zero_fwd_end = o;
  tp.zero_bwd_off = o;  // cleared at every backward: BN reductions and packed weight-gradient buffers
  for (auto& b : tp.bns) b.red_off = take((size_t)b.C * 2 * sizeof(double) + round_up_sz((size_t)b.C / 8 * 4, 16));
  for (auto& w : tp.wgs) {
    const ConvOpHost& f = p.convs[w.conv];
    w.dw_off = take((size_t)f.ntaps * f.nchunks * 8 * f.Cout * sizeof(float));
  }
  tp.stem_dw_off = take(49 * 16 * sizeof(float));
  tp.zero_bwd_end = o;
  for (auto& b : tp.bns) b.f_off = take((size_t)b.C * 4 * sizeof(float));
  tp.pack_off = o;
  for (size_t i = 0; i < p.convs.size(); ++i) {
    if (i < n_inf_convs) continue;  // inference packs are not used by the training plan
    ConvOpHost& c = p.convs[i];
    for (int gi = 0; gi < c.ngroups; ++gi) {
      c.g[gi].w_off = take((size_t)c.ntaps * c.nchunks * 8 * c.Cout * (precision == 0 ? 2 : 4));
      c.g[gi].bias_off = (size_t)-1;
    }
  }
  p.stem_w_off = take(49 * 16 * 4);
  p.stem_wu_off = take(2048);  // banded bf16 image of the tensor-core stem (zeroed with the other packs, then packed)
  tp.pack_end = o;
  tp.zero_bias_off = take(256 * sizeof(float));
  p.stem_b_off = tp.zero_bias_off;
  tp.logp_off = take((size_t)batch * 18 * sizeof(float));
  for (size_t i = n_inf_convs; i < p.convs.size(); ++i)
    for (int gi = 0; gi < p.convs[i].ngroups; ++gi) p.convs[i].g[gi].bias_off = tp.zero_bias_off;
  p.ws_bytes = o;
  tp.n_inf_convs = (int)n_inf_convs;
  return tpp;
}

// ---------------------------------------------------------------------------------------------------------
static BnArrays bn_arrays(const TrainPlan& tp, int bi) {
  const TrainPlan::BnHost& b = tp.bns[bi];
  char* ws = tp.p->ws;
  BnArrays a;
  a.sum = reinterpret_cast<double*>(ws + b.sum_off);
  a.sumsq = a.sum + b.C;
  a.red_g = reinterpret_cast<double*>(ws + b.red_off);
  a.red_gy = a.red_g + b.C;
  a.scale = reinterpret_cast<float*>(ws + b.f_off);
  a.shift = a.scale + b.C, a.mean = a.scale + 2 * b.C, a.invstd = a.scale + 3 * b.C;
  return a;
}

static FinalizeParams finalize_params(const TrainPlan& tp, int bi, const void* const* state, int nb) {
  const Plan& p = *tp.p;
  const TrainPlan::BnHost& b = tp.bns[bi];
  FinalizeParams fq{};
  fq.bn = bn_arrays(tp, bi), fq.C = b.C, fq.n = (double)nb * b.H * b.W, fq.eps = 1e-5f, fq.momentum = 0.1f;
  fq.nseg = (int)b.segs.size();
  for (int s = 0; s < fq.nseg; ++s) {
    const LayerSpec& L = p.layers[b.segs[s].layer];
    FinalizeSeg& fs = fq.seg[s];
    fs.gamma = (const float*)state[L.i_gamma], fs.beta = (const float*)state[L.i_beta];
    fs.running_mean = (float*)state[L.i_rm], fs.running_var = (float*)state[L.i_rv];
    fs.num_batches_tracked = (long long*)state[L.i_nbt];
    fs.conv_bias = L.i_b >= 0 ? (const float*)state[L.i_b] : nullptr;
    fs.ch_off = b.segs[s].ch_off, fs.count = b.segs[s].count;
  }
  return fq;
}

static EwParams ew_params(const TrainPlan& tp, const TrainPlan::EwHost& e, int nb) {
  const Plan& p = *tp.p;
  EwParams q{};
  const TensorInfo& ty = p.tensors[e.y];
  q.mode = e.mode, q.ntasks = e.ntasks, q.nchunks = e.nchunks;
  q.H = ty.H, q.W = ty.W, q.Wp = ty.Wp, q.Pimg = ty.Pimg;
  q.body = (long long)nb * ty.Pimg, q.body_full = q.body;
  q.inv_n = 1.f / (float)((double)nb * ty.H * ty.W);
  q.y = p.view(e.y, 0), q.bn = bn_arrays(tp, e.bn);
  if (e.y2 >= 0) q.y2 = p.view(e.y2, 0), q.bn2 = bn_arrays(tp, e.bn2);
  if (e.aux >= 0) q.aux = p.view(e.aux, 0);
  q.out = p.view(e.out, 0);
  if (e.mode == EW_RELU_POOL) {
    const TensorInfo& to = p.tensors[e.out];
    q.Ho = to.H, q.Wo = to.W, q.Wpo = to.Wp, q.Pimgo = to.Pimg;
    q.body = (long long)nb * to.Pimg;
    q.magic_pimgo = (1ull << 40) / (unsigned long long)q.Pimgo + 1, q.magic_wpo = (1ull << 40) / (unsigned long long)q.Wpo + 1;
  }
  q.magic_pimg = (1ull << 40) / (unsigned long long)q.Pimg + 1, q.magic_wp = (1ull << 40) / (unsigned long long)q.Wp + 1;
  return q;
}

static int pack_train(TrainPlan& tp, const void* const* state, cudaStream_t st) {
  Plan& p = *tp.p;
  // one memset over all packed weight images (merged / channel-padded blocks keep zeros), one batched pack launch
  CK(cudaMemsetAsync(p.ws + tp.pack_off, 0, tp.pack_end - tp.pack_off, st));
  std::vector<PackParams> packs;
  for (size_t ci = 0; ci < (size_t)tp.n_inf_convs; ++ci) {  // forward packs: same placement as inference, no fold
    const ConvOpHost& c = p.convs[tp.fconv_of[ci]];
    for (int gi = 0; gi < c.ngroups; ++gi)
      for (auto& sh : c.g[gi].segs) {
        PackParams q{};
        q.seg = make_seg(p, sh, state);
        q.dst = p.ws + c.g[gi].w_off, q.bias_dst = nullptr;
        q.Cout_p = c.Cout, q.K_tap = c.nchunks * 8, q.fp32 = p.precision == 0 ? 0 : 1, q.fold = 0;
        packs.push_back(q);
      }
  }
  for (auto& dp : tp.dgrads) {
    const ConvOpHost& c = p.convs[dp.conv];
    for (auto& ts : dp.segs) {
      PackParams q{};
      q.seg = make_seg(p, PackSegHost{ts.layer, INT_MAX, 0, 0}, state);
      q.seg.transpose = 1, q.seg.ci_lo = ts.ci_lo, q.seg.ci_n = ts.ci_n, q.seg.k_off = ts.k_off, q.seg.cout_off = ts.n_off;
      for (int t = 0; t < 9; ++t) q.seg.tap_map[t] = dp.tap_map[t];
      q.dst = p.ws + c.g[ts.group].w_off, q.bias_dst = nullptr;
      q.Cout_p = c.Cout, q.K_tap = c.nchunks * 8, q.fp32 = p.precision == 0 ? 0 : 1, q.fold = 0;
      packs.push_back(q);
    }
  }
  {
    PackParams q{};
    q.seg = make_seg(p, PackSegHost{p.stem_layer, INT_MAX, 0, 0}, state);
    q.dst = p.ws + p.stem_w_off, q.bias_dst = nullptr;
    q.Cout_p = 16, q.K_tap = 1, q.fp32 = 1, q.fold = 0;
    packs.push_back(q);
    if (p.precision == 0 && p.use_umma_stem) {  // banded bf16 image of the tensor-core stem (its pads were zeroed by the memset above)
      q.seg.k_from_tap = 2, q.dst = p.ws + p.stem_wu_off, q.K_tap = 64, q.fp32 = 0;
      packs.push_back(q);
    }
  }
  CK(launch_pack_batch(packs.data(), (int)packs.size(), st));
  return 0;
}

static int train_forward_eager(TrainPlan& tp, const void* const* state, const float* x, int nb, float* out0,
                               float* out1, cudaStream_t st) {
  Plan& p = *tp.p;
  if (!p.ws) {
    set_error("train_forward: no workspace bound");
    return -3;
  }
  if (nb < 1 || nb > p.mb) {
    set_error("train_forward: batch %d outside [1, %d] (BatchNorm statistics need the whole batch in one pass)", nb, p.mb);
    return -1;
  }
  int rc = pack_train(tp, state, st);
  if (rc) return rc;
  CK(cudaMemsetAsync(p.ws + tp.zero_fwd_off, 0, tp.zero_fwd_end - tp.zero_fwd_off, st));
  tp.nb = nb, tp.x = x, tp.launches = 0;
  float* logp0 = reinterpret_cast<float*>(p.ws + tp.logp_off);
  float* logp1 = logp0 + (size_t)p.mb * 16;
  OpTimer timer(st);
  for (auto& op : tp.fwd) {
    timer.begin(op.kind);
    if (op.kind == TOP_STEM) {
      const TensorInfo& t = p.tensors[tp.y_stem];
      StemParams q{};
      q.x = x, q.w = reinterpret_cast<const float*>(p.ws + p.stem_w_off);
      q.bias = reinterpret_cast<const float*>(p.ws + tp.zero_bias_off);
      q.out = p.view(tp.y_stem, 0);
      q.Hin = kPatchH, q.Win = kPatchW, q.H = t.H, q.W = t.W, q.Wp = t.Wp, q.Pimg = t.Pimg;
      q.body = (long long)nb * t.Pimg, q.raw = 1;
      q.w_umma = (p.precision == 0 && p.use_umma_stem) ? p.ws + p.stem_wu_off : nullptr;
      CK(launch_stem(q, p.precision, st));
    } else if (op.kind == TOP_CONV) {
      ConvParams q = p.conv_params(p.convs[op.a], nb);
      CK(launch_conv(q, p.precision, st, p.use_pdl != 0));
    } else if (op.kind == TOP_EW) {
      const TrainPlan::EwHost& e = tp.ews[op.a];
      EwParams q = ew_params(tp, e, nb);
      q.fin = finalize_params(tp, e.bn, state, nb);
      if (e.y2 >= 0) q.fin2 = finalize_params(tp, e.bn2, state, nb);
      q.sync = reinterpret_cast<unsigned int*>(p.ws + tp.bns[e.bn].sum_off + (size_t)tp.bns[e.bn].C * 2 * sizeof(double));
      CK(launch_ew_fwd(q, p.precision, st));
    } else if (op.kind == TOP_HEAD) {
      const HeadOpHost& h = p.heads[op.a];
      const TensorInfo& t = p.tensors[h.in_tensor];
      HeadParams q{};
      q.in = p.view(h.in_tensor, h.chunk0);
      q.out = h.out_index == 0 ? logp0 : logp1;
      q.nb = nb, q.classes = h.classes, q.npos = t.Pimg;
      q.inv_count = 1.f / (float)(t.H * t.W * (128 / h.classes));
      CK(launch_head(q, p.precision, st));
      float* dst = h.out_index == 0 ? out0 : out1;
      if (!dst) {
        set_error("train_forward: output %d is NULL", h.out_index);
        return -1;
      }
      CK(cudaMemcpyAsync(dst, q.out, (size_t)nb * h.classes * sizeof(float), cudaMemcpyDeviceToDevice, st));
    }
    timer.end();
    ++tp.launches;
  }
  timer.report("forward");
  tp.have_forward = true;
  return 0;
}

static int train_backward_eager(TrainPlan& tp, const void* const* state, const float* dout0, const float* dout1,
                                float* grads, cudaStream_t st) {
  Plan& p = *tp.p;
  if (!tp.have_forward) {
    set_error("train_backward: no forward pass to differentiate");
    return -3;
  }
  const int nb = tp.nb;
  CK(cudaMemsetAsync(p.ws + tp.zero_bwd_off, 0, tp.zero_bwd_end - tp.zero_bwd_off, st));
  CK(cudaMemsetAsync(grads, 0, tp.n_params * sizeof(float), st));
  float* logp0 = reinterpret_cast<float*>(p.ws + tp.logp_off);
  float* logp1 = logp0 + (size_t)p.mb * 16;
  OpTimer timer(st);
  std::vector<UnpackParams> unpacks;  // packed weight gradients -> parameter layout, one batched launch at the end
  static int side_env = -1;
  if (side_env < 0) {
    const char* e = getenv("MTLDAS_WGRAD_SIDE");
    side_env = (e && e[0] == '0') ? 0 : 1;
  }
  const bool use_side = side_env && !timer.on;
  size_t n_fork = 0;
  if (use_side && !tp.side_stream) {
    CK(cudaStreamCreateWithFlags(&tp.side_stream, cudaStreamNonBlocking));
    CK(cudaEventCreateWithFlags(&tp.join_event, cudaEventDisableTiming));
  }
  // the side stream picks up after everything enqueued on `st` so far (the layer's dy, the zeroed dW buffers)
  auto fork = [&](cudaStream_t& out) -> bool {
    out = st;
    if (!use_side) return true;
    if (n_fork == tp.fork_events.size()) {
      cudaEvent_t e = nullptr;
      if (cudaEventCreateWithFlags(&e, cudaEventDisableTiming) != cudaSuccess) return false;
      tp.fork_events.push_back(e);
    }
    cudaEvent_t e = tp.fork_events[n_fork++];
    if (cudaEventRecord(e, st) != cudaSuccess || cudaStreamWaitEvent(tp.side_stream, e, 0) != cudaSuccess) return false;
    out = tp.side_stream;
    return true;
  };
  for (auto& op : tp.bwd) {
    timer.begin(op.kind);
    if (op.kind == TOP_HEADBWD) {
      const HeadOpHost& h = p.heads[op.a];
      const TensorInfo& t = p.tensors[h.in_tensor];
      HeadBwdParams q{};
      q.logp = h.out_index == 0 ? logp0 : logp1;
      q.dlogp = h.out_index == 0 ? dout0 : dout1;
      if (!q.dlogp) {
        set_error("train_backward: upstream gradient %d is NULL", h.out_index);
        return -1;
      }
      q.da4 = p.view(tp.grad_tensor.at(h.in_tensor), h.chunk0);
      q.nb = nb, q.classes = h.classes, q.H = t.H, q.W = t.W, q.Wp = t.Wp, q.Pimg = t.Pimg;
      CK(launch_head_bwd(q, p.precision, st));
    } else if (op.kind == TOP_EWB) {
      const TrainPlan::EwHost& e = tp.ews[op.a];
      EwParams q = ew_params(tp, e, nb);
      q.dz = p.view(e.dz, 0), q.dy = p.view(e.dy, 0);
      q.want_daux = e.daux >= 0, q.daux_accum = e.daux_accum;
      if (e.daux >= 0) q.daux = p.view(e.daux, 0);
      auto set_gsegs = [&](int bi, int& n, BnGradSeg* out) {
        const TrainPlan::BnHost& b = tp.bns[bi];
        n = (int)b.segs.size();
        for (int s = 0; s < n; ++s) {
          const TrainPlan::ParamOff& po = tp.param_off[b.segs[s].layer];
          out[s] = BnGradSeg{grads + po.gamma, grads + po.beta, b.segs[s].ch_off, b.segs[s].count};
        }
      };
      set_gsegs(e.bn, q.ngseg, q.gseg);
      if (e.mode == EW_ADD_RELU_SC) {  // the shortcut branch shares g; its own BatchNorm gives its own dy
        q.dy2 = p.view(e.dy2, 0);
        set_gsegs(e.bn2, q.ngseg2, q.gseg2);
      }
      q.sync = reinterpret_cast<unsigned int*>(p.ws + tp.bns[e.bn].red_off + (size_t)tp.bns[e.bn].C * 2 * sizeof(double));
      CK(launch_ew_bwd(q, p.precision, st));
    } else if (op.kind == TOP_WGRAD) {
      const TrainPlan::WgHost& w = tp.wgs[op.a];
      const ConvOpHost& f = p.convs[w.conv];
      ConvParams cp = p.conv_params(f, nb);
      WgradParams q{};
      for (int s = 0; s < f.nsrc; ++s) q.src[s] = cp.g[w.group].src[s];
      for (int t = 0; t < f.ntaps; ++t) q.tap[t] = f.tap[t];
      const TensorInfo& td = p.tensors[w.dy];
      q.dy = p.ws + td.offset + ((size_t)f.g[w.group].out_chunk0 * td.plane + td.g_lo) * p.gran;
      q.dy_plane_bytes = td.plane * p.gran;
      q.dw = reinterpret_cast<float*>(p.ws + w.dw_off);
      q.ntaps = f.ntaps, q.nchunks = f.nchunks, q.Cout = f.Cout, q.body = cp.body, q.nsrc = f.nsrc;
      long long want = std::max<long long>(1, 592 / f.ntaps);
      q.slice = (int)std::max<long long>(256, (cp.body + want - 1) / want);
      q.slice = (q.slice + 31) / 32 * 32;
      cudaStream_t ws = st;
      if (!fork(ws)) {
        set_error("train_backward: side-stream fork failed: %s", cudaGetErrorString(cudaGetLastError()));
        return -4;
      }
      CK(launch_wgrad(q, p.precision, ws));
      for (auto& sh : f.g[w.group].segs) {
        const LayerSpec& L = p.layers[sh.layer];
        UnpackParams u{};
        u.src = q.dw, u.dst = grads + tp.param_off[sh.layer].w;
        u.cout = L.cout, u.cin = L.cin, u.ksize = L.ksize, u.split = sh.split, u.k_off = sh.k_off, u.cout_off = sh.cout_off;
        u.Cout_p = f.Cout, u.K_tap = f.nchunks * 8;
        unpacks.push_back(u);
      }
    } else if (op.kind == TOP_DGRAD) {
      ConvParams q = p.conv_params(p.convs[tp.dgrads[op.a].conv], nb);
      CK(launch_conv(q, p.precision, st, p.use_pdl != 0));
    } else if (op.kind == TOP_STEMWG) {
      const TensorInfo& t = p.tensors[tp.dy_stem];
      StemWgradParams q{};
      q.x = tp.x, q.dy = p.view(tp.dy_stem, 0), q.dw = reinterpret_cast<float*>(p.ws + tp.stem_dw_off);
      q.Hin = kPatchH, q.Win = kPatchW, q.H = t.H, q.W = t.W, q.Wp = t.Wp, q.Pimg = t.Pimg;
      q.body = (long long)nb * t.Pimg;
      cudaStream_t ws = st;
      if (!fork(ws)) {
        set_error("train_backward: side-stream fork failed: %s", cudaGetErrorString(cudaGetLastError()));
        return -4;
      }
      CK(launch_stem_wgrad(q, p.precision, ws));
      UnpackParams u{};
      u.src = q.dw, u.dst = grads + tp.param_off[p.stem_layer].w;
      u.cout = 16, u.cin = 1, u.ksize = 7, u.split = INT_MAX, u.k_off = 0, u.cout_off = 0, u.Cout_p = 16, u.K_tap = 1;
      unpacks.push_back(u);
    }
    timer.end();
    ++tp.launches;
  }
  if (use_side && n_fork) {  // join: the unpack reads every dW buffer
    CK(cudaEventRecord(tp.join_event, tp.side_stream));
    CK(cudaStreamWaitEvent(st, tp.join_event, 0));
  }
  CK(launch_unpack_batch(unpacks.data(), (int)unpacks.size(), st));
  ++tp.launches;
  timer.report("backward");
  (void)state;
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// CUDA-graph replay: a training step is ~125 forward + ~235 backward launches; when the same buffers come back
// (steady-state training loop) the captured graph is replayed.  First use eager, second use captured.
// ---------------------------------------------------------------------------------------------------------
template <typename F>
static int run_with_graph(TrainPlan& tp, std::vector<TrainPlan::GraphEntry>& cache, const std::vector<const void*>& key,
                          int nb, cudaStream_t st, F&& eager) {
  const char* ng = getenv("MTLDAS_NO_GRAPH");
  const char* prof = getenv("MTLDAS_TRAIN_PROFILE");
  if ((ng && ng[0] == '1') || (prof && prof[0] == '1')) return eager(st);
  TrainPlan::GraphEntry* e = nullptr;
  for (auto& g : cache)
    if (g.nb == nb && g.key == key) e = &g;
  if (!e) {
    if (cache.size() >= 4) {
      if (cache.front().exec) cudaGraphExecDestroy(cache.front().exec);
      cache.erase(cache.begin());
    }
    cache.push_back(TrainPlan::GraphEntry{key, nb, nullptr, 0, 0});
    e = &cache.back();
  }
  e->uses++;
  if (e->exec) {
    CK(cudaGraphLaunch(e->exec, st));
    tp.launches = e->launches;
    return 0;
  }
  if (e->uses < 2) return eager(st);
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(st, &cs) != cudaSuccess) cudaGetLastError();
  if (cs != cudaStreamCaptureStatusNone) return eager(st);
  if (!tp.cap_stream) CK(cudaStreamCreateWithFlags(&tp.cap_stream, cudaStreamNonBlocking));
  CK(cudaStreamBeginCapture(tp.cap_stream, cudaStreamCaptureModeThreadLocal));
  int rc = eager(tp.cap_stream);
  cudaGraph_t graph = nullptr;
  cudaError_t ce = cudaStreamEndCapture(tp.cap_stream, &graph);
  if (rc || ce != cudaSuccess) {
    if (graph) cudaGraphDestroy(graph);
    cudaGetLastError();
    if (!rc) set_error("train graph capture failed: %s", cudaGetErrorString(ce));
    return rc ? rc : -2;
  }
  e->launches = tp.launches;
  ce = cudaGraphInstantiate(&e->exec, graph, 0);
  cudaGraphDestroy(graph);
  if (ce != cudaSuccess) {
    e->exec = nullptr;
    set_error("train graph instantiate failed: %s", cudaGetErrorString(ce));
    return -2;
  }
  CK(cudaGraphLaunch(e->exec, st));
  return 0;
}

int train_forward(TrainPlan& tp, const void* const* state, const float* x, int nb, float* out0, float* out1,
                  cudaStream_t st) {
  Plan& p = *tp.p;
  if (!p.ws) {
    set_error("train_forward: no workspace bound");
    return -3;
  }
  if (nb < 1 || nb > p.mb) {
    set_error("train_forward: batch %d outside [1, %d] (BatchNorm statistics need the whole batch in one pass)", nb, p.mb);
    return -1;
  }
  std::vector<const void*> key{x, out0, out1};
  for (size_t i = 0; i < p.state_keys.size(); ++i) key.push_back(state[i]);
  int rc = run_with_graph(tp, tp.fwd_graphs, key, nb, st,
                          [&](cudaStream_t s) { return train_forward_eager(tp, state, x, nb, out0, out1, s); });
  if (!rc) tp.nb = nb, tp.x = x, tp.have_forward = true;
  return rc;
}

int train_backward(TrainPlan& tp, const void* const* state, const float* dout0, const float* dout1, float* grads,
                   cudaStream_t st) {
  if (!tp.have_forward) {
    set_error("train_backward: no forward pass to differentiate");
    return -3;
  }
  std::vector<const void*> key{dout0, dout1, grads, tp.x};
  return run_with_graph(tp, tp.bwd_graphs, key, tp.nb, st,
                        [&](cudaStream_t s) { return train_backward_eager(tp, state, dout0, dout1, grads, s); });
}

TrainPlan::~TrainPlan() {
  for (auto* c : {&fwd_graphs, &bwd_graphs})
    for (auto& g : *c)
      if (g.exec) cudaGraphExecDestroy(g.exec);
  if (cap_stream) cudaStreamDestroy(cap_stream);
  if (side_stream) cudaStreamDestroy(side_stream);
  for (auto e : fork_events) cudaEventDestroy(e);
  if (join_event) cudaEventDestroy(join_event);
  delete p;
}

}  // namespace mtldas
