// Builds the layer program of MTL_Net (modelA_MTL.py:59-174) / Single_Task_Net
// (modelB_singleTask.py:58-178) over the P8 layouts and runs it.
#include "plan.h"

#include <algorithm>
#include <climits>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <sstream>

#include "kernels.h"

namespace mtldas {

static thread_local char g_err[512] = "";
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
const char* last_error_text() { return g_err; }

static inline long long round_up(long long v, long long a) { return (v + a - 1) / a * a; }

// ---------------------------------------------------------------------------------------------
// layer table in state_dict registration order (modelA_MTL.py:73-113)
// ---------------------------------------------------------------------------------------------
static void add_layer(Plan& p, const std::string& conv, const std::string& bn, int cin, int cout, int k, int s,
                      int pad, bool bias) {
  LayerSpec L;
  L.conv = conv, L.bn = bn, L.cin = cin, L.cout = cout, L.ksize = k, L.stride = s, L.pad = pad, L.bias = bias;
  auto push = [&](const std::string& key) {
    p.state_keys.push_back(key);
    return (int)p.state_keys.size() - 1;
  };
  L.i_w = push(conv + ".weight");
  L.i_b = bias ? push(conv + ".bias") : -1;
  L.i_gamma = push(bn + ".weight");
  L.i_beta = push(bn + ".bias");
  L.i_rm = push(bn + ".running_mean");
  L.i_rv = push(bn + ".running_var");
  L.i_nbt = push(bn + ".num_batches_tracked");
  p.layers.push_back(L);
}

static int layer_index(const Plan& p, const std::string& conv) {
  for (size_t i = 0; i < p.layers.size(); ++i)
    if (p.layers[i].conv == conv) return (int)i;
  return -1;
}

static const int kResCfg[8][3] = {{16, 16, 1}, {16, 16, 1}, {16, 32, 2}, {32, 32, 1},
                                  {32, 64, 2}, {64, 64, 1}, {64, 128, 2}, {128, 128, 1}};
static const char* kAttNames[4] = {"att_mask_generator1", "att_mask_generato2", "att_mask_generator3",
                                   "att_mask_generator4"};  // typo is part of the contract (modelA_MTL.py:93)
static const int kAttCfg[4][3] = {{16, 8, 16}, {64, 16, 32}, {128, 32, 64}, {256, 64, 128}};
static const int kOutCfg[3][2] = {{16, 32}, {32, 64}, {64, 128}};

static void build_layers(Plan& p) {
  add_layer(p, "conv1.0", "conv1.1", 1, 16, 7, 3, 2, false);
  for (int i = 0; i < 8; ++i) {
    std::string rb = "resblock" + std::to_string(i + 1);
    int ci = kResCfg[i][0], co = kResCfg[i][1], s = kResCfg[i][2];
    add_layer(p, rb + ".left.0", rb + ".left.1", ci, co, 3, s, 1, false);
    add_layer(p, rb + ".left.3", rb + ".left.4", co, co, 3, 1, 1, false);
    if (s != 1 || ci != co) add_layer(p, rb + ".shortcut.0", rb + ".shortcut.1", ci, co, 1, s, 0, false);
  }
  for (int l = 0; l < 4; ++l)
    for (int t = 0; t < p.ntasks; ++t) {
      std::string a = std::string(kAttNames[l]) + "." + std::to_string(t);
      add_layer(p, a + ".0", a + ".1", kAttCfg[l][0], kAttCfg[l][1], 1, 1, 0, true);
      add_layer(p, a + ".3", a + ".4", kAttCfg[l][1], kAttCfg[l][2], 3, 1, 1, true);
    }
  for (int l = 0; l < 3; ++l)
    for (int t = 0; t < p.ntasks; ++t) {
      std::string o = "output_layer" + std::to_string(l + 1) + "." + std::to_string(t);
      add_layer(p, o + ".0", o + ".1", kOutCfg[l][0], kOutCfg[l][1], 3, 1, 1, false);
    }
}

// ---------------------------------------------------------------------------------------------
// tensors
// ---------------------------------------------------------------------------------------------
int add_tensor(Plan& p, const std::string& name, int C, int H, int W, int layout) {
  TensorInfo t;
  t.name = name, t.C = C, t.H = H, t.W = W, t.layout = layout;
  t.Hs = layout == LAY_P8 ? H : (H + 1) / 2;
  t.Ws = layout == LAY_P8 ? W : (W + 1) / 2;
  t.Wp = t.Ws + 1;
  t.Pimg = (t.Hs + 1) * t.Wp;
  t.g_lo = (int)round_up(t.Wp + 1, 8);
  t.plane = round_up((long long)t.g_lo + (long long)p.mb * t.Pimg + kMaxTile + t.Wp + 1, 8);
  t.nchunks = C / 8;
  t.nphase = layout == LAY_P8 ? 1 : 4;
  t.bytes = (size_t)t.nphase * t.nchunks * t.plane * p.gran;
  t.offset = 0;
  p.tensors.push_back(t);
  return (int)p.tensors.size() - 1;
}

int Plan::find_tensor(const std::string& n) const {
  for (size_t i = 0; i < tensors.size(); ++i)
    if (tensors[i].name == n) return (int)i;
  return -1;
}

ActView Plan::view(int tensor, int chunk0) const {
  const TensorInfo& t = tensors[tensor];
  ActView v;
  v.base = ws + t.offset + (size_t)chunk0 * t.plane * gran;
  v.plane = t.plane;
  v.phase_stride = t.layout == LAY_P8 ? 0 : (long long)t.nchunks * t.plane;
  v.g_lo = t.g_lo;
  v.layout = t.layout;
  v.Wp2 = t.Wp;
  v.Pimg2 = t.Pimg;
  return v;
}

// ---------------------------------------------------------------------------------------------
// conv op construction
// ---------------------------------------------------------------------------------------------
// shared-memory stages of the persistent kernel for this op (0 = not eligible); `extra` = bytes of other buffers
// dynamic shared memory one CTA may own on sm_100 (227 KB) minus the kernel's header and alignment slack
constexpr size_t kSmemOneCta = 232448 - 1280 - 384;

int persist_stages_for(const ConvOpHost& op, size_t extra) {
  if (op.nstages != 1) return 0;
  const size_t SP = (size_t)op.MT * kTileM + op.halo_lo + op.halo_hi;
  const size_t slab = (size_t)op.nsrc * op.nchunks * SP * 16, wb = (size_t)op.ntaps * op.tap_bytes;
  static int env_stages = -1;
  if (env_stages < 0) {
    const char* e = getenv("MTLDAS_PERSIST_STAGES");
    env_stages = e ? atoi(e) : 3;
  }
  if (wb + (size_t)env_stages * slab + extra <= 108 * 1024) return env_stages;
  if (wb + 2 * ((slab + 127) & ~(size_t)127) + extra <= kSmemOneCta) return 2;
  return 0;
}

void choose_tiling(ConvOpHost& op, long long positions, bool small_resident) {
  op.tap_bytes = op.nchunks * 8 * op.Cout * 2;
  const int total_w = op.ntaps * op.tap_bytes;
  int MT = 4;
  // small batches: prefer more, smaller CTAs over multi-tile CTAs until every SM has about two
  if (positions > 0)
    while (MT > 1 && (positions + kTileM - 1) / kTileM / MT < 296) MT >>= 1;
  size_t slab = 0;
  for (;; MT >>= 1) {
    int SP = MT * kTileM + op.halo_lo + op.halo_hi;
    slab = (size_t)op.nsrc * op.nchunks * SP * 16;
    if (MT == 1) break;
    if (MT * op.Cout <= 256 && slab <= 80 * 1024) break;
    // compute-heavy layers (>= 128 output channels): four independent accumulator tiles keep the tensor pipe busy
    // (two serialise on the accumulate latency); needs the slab plus a two-slot weight ring in shared memory
    if (MT * op.Cout <= 512 && op.Cout >= 128 && op.nchunks >= 16 && slab + 2 * (size_t)op.tap_bytes <= 216 * 1024) break;
  }
  op.MT = MT;
  // weights too large for two CTAs per SM but small enough to stay resident next to two input slabs: one persistent
  // CTA per SM with two TMEM accumulator sets (epilogue of tile i under the MMAs of tile i+1) instead of the
  // one-tile-per-CTA ring kernel whose phases run back to back
  static int big_resident = -1;
  if (big_resident < 0) {
    const char* e = getenv("MTLDAS_BIG_RESIDENT");
    big_resident = (e && e[0] == '0') ? 0 : 1;
  }
  if (big_resident && slab + total_w > 96 * 1024) {
    // MT = 1 with full-width (128-column) weights measured slower than the ring kernel; the N-split halves
    // (small_resident) are where one tile per iteration pays: their weights only fit next to single-tile slabs
    for (int m = 2; m >= (small_resident ? 1 : 2); m >>= 1) {
      if (m > MT) continue;
      const size_t sl = ((size_t)op.nsrc * op.nchunks * (m * kTileM + op.halo_lo + op.halo_hi) * 16 + 127) & ~(size_t)127;
      if ((size_t)total_w + 2 * sl <= kSmemOneCta && 2 * m * op.Cout <= 512) {
        MT = m, slab = sl;
        op.MT = MT, op.tps = op.ntaps, op.NS = 1, op.nstages = 1, op.ksplit = 0;
        int c = 32;
        while (c < MT * op.Cout) c <<= 1;
        op.tmem_cols = c;
        return;
      }
    }
  }
  if (slab + total_w <= 96 * 1024) {
    op.tps = op.ntaps;
    op.NS = 1;
  } else {
    op.tps = 1;
    long long room = 110 * 1024 - (long long)slab;
    int ns = (int)(room / op.tap_bytes);
    if (ns < 3) ns = (int)((220 * 1024 - (long long)slab) / op.tap_bytes);
    op.NS = std::max(1, std::min(std::min(ns, op.ntaps), kMaxRing));
  }
  op.nstages = (op.ntaps + op.tps - 1) / op.tps;
  // K-split accumulation (two TMEM sets) is implemented in the kernel but measured slower on B200 for these
  // layers (TMEM-limited CTA concurrency, doubled epilogue loads): left off.
  op.ksplit = 0;
  int cols = 32;
  while (cols < (op.ksplit ? 2 : 1) * MT * op.Cout) cols <<= 1;
  op.tmem_cols = cols;
}

// Layers whose packed weights cannot stay in shared memory (64->128 / 128->128: 147 / 295 KB) ran one tile per CTA
// with the weights re-streamed from L2 per tile and load -> MMA -> epilogue back to back (0.13-0.37 of their roofline).
// Splitting the OUTPUT channels into two groups halves the weight image per CTA, so each half stays resident and runs
// the persistent pipelined kernel (epilogue of tile i under the MMAs of tile i+1).  The two halves of a tile run on
// different CTAs at about the same time, so the second read of the input slab hits L2.  Cost: N = 64 instructions
// are bound by the A-operand read (~45 cycles against 32 ideal), ~1.3x the tensor time of N = 128.
bool split_output_channels(ConvOpHost& op, long long positions, bool even_if_resident) {
  if ((op.nstages == 1 && !even_if_resident) || op.Cout < 128 || (op.Cout % 32) || op.ngroups * 2 > kMaxGroups) return false;
  const int Cs = op.Cout / 2;
  ConvOpHost s = op;
  s.Cout = Cs, s.ngroups = op.ngroups * 2;
  for (int g = 0; g < op.ngroups; ++g) {
    if (op.g[g].aux_tensor >= 0 && op.g[g].aux_nchunks != op.Cout / 8) return false;
    for (int half = 0; half < 2; ++half) {
      ConvGroupHost ng = op.g[g];
      ng.out_chunk0 += half * (Cs / 8);
      if (ng.aux_tensor >= 0) ng.aux_chunk0 += half * (Cs / 8), ng.aux_nchunks = Cs / 8;
      ng.segs.clear();
      for (const PackSegHost& sg : op.g[g].segs) {
        if (sg.layer < 0) continue;
        // the caller's segs place whole layers; single-layer groups only (all candidates are)
        PackSegHost h = sg;
        h.co_lo = sg.co_lo + half * Cs, h.co_n = Cs;
        ng.segs.push_back(h);
      }
      s.g[g * 2 + half] = ng;
    }
  }
  choose_tiling(s, positions, /*small_resident=*/true);
  if (s.nstages != 1 || persist_stages_for(s, 0) == 0) return false;
  op = s;
  return true;
}

namespace {
struct Builder {
  Plan& p;
  explicit Builder(Plan& pl) : p(pl) {}

  ConvOpHost base(const std::string& name, int H, int W, int Cout, int nchunks, int epi, int ngroups) {
    ConvOpHost op{};
    op.name = name, op.H = H, op.W = W, op.Cout = Cout, op.nchunks = nchunks, op.epi = epi, op.ngroups = ngroups;
    for (auto& g : op.g) {
      g.aux_tensor = -1, g.aux_chunk0 = 0, g.aux_nchunks = Cout / 8;
      g.out_tensor = -1, g.out_chunk0 = 0;
      for (auto& s : g.src) s = SrcRef{-1, 0, -1};
    }
    return op;
  }
  // 3x3 stride 1 pad 1 over one P8 source
  void taps_3x3_s1(ConvOpHost& op) {
    int Wp = op.W + 1;
    op.nsrc = 1, op.ntaps = 9;
    for (int r = 0; r < 3; ++r)
      for (int s = 0; s < 3; ++s) op.tap[r * 3 + s] = ConvTapDev{0, (r - 1) * Wp + (s - 1)};
    op.halo_lo = Wp + 1, op.halo_hi = Wp + 1;
  }
  // 3x3 stride 2 pad 1 over the four phase planes of a P8S2 source: input row 2i-1+r
  //   r=0 -> odd rows, index i-1 ; r=1 -> even rows, index i ; r=2 -> odd rows, index i
  void taps_3x3_s2(ConvOpHost& op) {
    int Wp = op.W + 1;
    op.nsrc = 4, op.ntaps = 9;
    for (int r = 0; r < 3; ++r)
      for (int s = 0; s < 3; ++s) {
        int ph = (r == 1) ? 0 : 1, pw = (s == 1) ? 0 : 1;
        int dh = (r == 0) ? -1 : 0, dw = (s == 0) ? -1 : 0;
        op.tap[r * 3 + s] = ConvTapDev{ph * 2 + pw, dh * Wp + dw};
      }
    op.halo_lo = Wp + 1, op.halo_hi = 0;
  }
  void taps_1x1(ConvOpHost& op, int nsrc) {
    op.nsrc = nsrc, op.ntaps = nsrc;
    for (int s = 0; s < nsrc; ++s) op.tap[s] = ConvTapDev{s, 0};
    op.halo_lo = op.halo_hi = 0;
  }
  void finish(ConvOpHost& op, bool tiled = false) {
    const long long positions = (long long)p.mb * (op.H + 1) * (op.W + 1);
    if (!tiled) {
      choose_tiling(op, positions);
      // mode 1: only the stride-2 layers (four phase sources); mode 2 (default): every layer whose weights stream
      if (p.precision == 0 && !p.training && p.use_nsplit && (p.nsplit_mode >= 2 || op.nsrc == 4))
        split_output_channels(op, positions);
    }
    p.convs.push_back(op);
    p.program.push_back(Op{1, (int)p.convs.size() - 1});
  }
  PackSegHost seg(const std::string& conv, int split = INT_MAX, int k_off = 0, int cout_off = 0) {
    int li = layer_index(p, conv);
    return PackSegHost{li, split, k_off, cout_off};
  }
};
}  // namespace

Plan* build_plan(int kind, int mb, int precision, bool training) {
  Plan* pp = new Plan();
  Plan& p = *pp;
  p.kind = kind, p.mb = mb, p.precision = precision, p.training = training;
  {
    const char* e = getenv("MTLDAS_NO_NSPLIT");
    p.use_nsplit = !(e && e[0] == '1');
    const char* m = getenv("MTLDAS_NSPLIT_MODE");
    p.nsplit_mode = m ? atoi(m) : 1;   // measured: only the stride-2 64->128 layer gains (0.19 -> 0.07 ms); N = 64 halves cost the
                                         // same ~60 cycles per MMA as N = 128 (operand reads bind), so splitting 128->128 only adds work
    const char* o3 = getenv("MTLDAS_POOL_SPLIT");
    p.pool_split = (o3 && o3[0] == '1');   // measured slower than resident full-width + separate pool (0.33 vs 0.30 ms)
  }
  p.ntasks = kind == 0 ? 2 : 1;
  p.gran = precision == 0 ? 16 : 32;
  build_layers(p);
  Builder b(p);
  const int T = p.ntasks;
  const int H[4] = {33, 17, 9, 5}, W[4] = {83, 42, 21, 11}, C[4] = {16, 32, 64, 128};

  // ---- shared trunk ---------------------------------------------------------------------------
  p.stem_layer = 0;
  int stem = add_tensor(p, "stem", 16, H[0], W[0], LAY_P8);
  p.stem_out = stem;
  p.program.push_back(Op{0, 0});
  int sb[8];
  int prev = stem;
  for (int i = 0; i < 8; ++i) {
    const int st = i / 2;  // stage of the block output
    const std::string rb = "resblock" + std::to_string(i + 1);
    const bool down = kResCfg[i][2] == 2;
    const bool next_down = (i + 1 < 8) && kResCfg[i + 1][2] == 2;  // output feeds stride-2 convs -> phase split
    int mid = add_tensor(p, rb + ".mid", C[st], H[st], W[st], LAY_P8);
    int sc = -1;
    {
      ConvOpHost op = b.base(rb + ".left.0", H[st], W[st], C[st], p.tensors[prev].nchunks, EPI_RELU, 1);
      if (down) {
        b.taps_3x3_s2(op);
        for (int ph = 0; ph < 4; ++ph) op.g[0].src[ph] = SrcRef{prev, 0, ph};
      } else {
        b.taps_3x3_s1(op);
        op.g[0].src[0] = SrcRef{prev, 0, -1};
      }
      op.g[0].out_tensor = mid;
      op.g[0].segs.push_back(b.seg(rb + ".left.0"));
      b.finish(op);
    }
    if (down) {
      sc = add_tensor(p, rb + ".sc", C[st], H[st], W[st], LAY_P8);
      ConvOpHost op = b.base(rb + ".shortcut.0", H[st], W[st], C[st], p.tensors[prev].nchunks, EPI_NONE, 1);
      b.taps_1x1(op, 1);
      op.g[0].src[0] = SrcRef{prev, 0, 0};  // 1x1 stride 2 reads x[2i,2j] = the (even,even) phase plane
      op.g[0].out_tensor = sc;
      op.g[0].segs.push_back(b.seg(rb + ".shortcut.0"));
      b.finish(op);
    }
    sb[i] = add_tensor(p, "sb" + std::to_string(i), C[st], H[st], W[st], next_down ? LAY_P8S2 : LAY_P8);
    {
      ConvOpHost op = b.base(rb + ".left.3", H[st], W[st], C[st], C[st] / 8, EPI_ADD_RELU, 1);
      b.taps_3x3_s1(op);
      op.g[0].src[0] = SrcRef{mid, 0, -1};
      op.g[0].out_tensor = sb[i];
      op.g[0].aux_tensor = down ? sc : prev;
      op.g[0].segs.push_back(b.seg(rb + ".left.3"));
      b.finish(op);
    }
    prev = sb[i];
  }

  // ---- task-specific sub-networks (modelA_MTL.py:142-163) --------------------------------------
  int a_prev = -1;  // pooled task feature of the previous level, channels = C[l] * T (task t at chunk C[l]/8*t)
  for (int l = 0; l < 4; ++l) {
    const std::string an = kAttNames[l];
    const int cm = kAttCfg[l][1], co = kAttCfg[l][2];
    const int cm_p = std::max(cm, 16);  // level 1 has 8 mid channels per task: padded / merged to 16
    int tmid, amask;
    if (l == 0) {
      // both tasks read the same sb0: one 1x1 conv with the tasks' filters stacked along Cout (8+8),
      // then one 3x3 conv with block-diagonal weights producing [mask_0 | mask_1].
      tmid = add_tensor(p, "att1.mid", 16, H[0], W[0], LAY_P8);
      amask = add_tensor(p, "a1", co * T, H[0], W[0], LAY_P8);
      ConvOpHost op = b.base(an + ".0", H[0], W[0], 16, 2, EPI_RELU, 1);
      b.taps_1x1(op, 1);
      op.g[0].src[0] = SrcRef{sb[0], 0, -1};
      op.g[0].out_tensor = tmid;
      for (int t = 0; t < T; ++t) op.g[0].segs.push_back(b.seg(an + "." + std::to_string(t) + ".0", INT_MAX, 0, 8 * t));
      b.finish(op);
      ConvOpHost op2 = b.base(an + ".3", H[0], W[0], co * T, 2, EPI_SIGMOID_MUL, 1);
      b.taps_3x3_s1(op2);
      op2.g[0].src[0] = SrcRef{tmid, 0, -1};
      op2.g[0].out_tensor = amask;
      op2.g[0].aux_tensor = sb[1];
      op2.g[0].aux_nchunks = co / 8;  // both tasks multiply the same 16-channel sb1
      for (int t = 0; t < T; ++t)
        op2.g[0].segs.push_back(b.seg(an + "." + std::to_string(t) + ".3", INT_MAX, 8 * t, co * t));
      b.finish(op2);
    } else {
      tmid = add_tensor(p, "att" + std::to_string(l + 1) + ".mid", cm_p * T, H[l], W[l], LAY_P8);
      amask = add_tensor(p, "a" + std::to_string(l + 1), co * T, H[l], W[l], LAY_P8);
      ConvOpHost op = b.base(an + ".0", H[l], W[l], cm_p, C[l] / 8, EPI_RELU, T);
      b.taps_1x1(op, 2);  // torch.cat((shared, task), dim=1): shared channels first
      for (int t = 0; t < T; ++t) {
        op.g[t].src[0] = SrcRef{sb[2 * l], 0, -1};
        op.g[t].src[1] = SrcRef{a_prev, (C[l] / 8) * t, -1};
        op.g[t].out_tensor = tmid, op.g[t].out_chunk0 = (cm_p / 8) * t;
        op.g[t].segs.push_back(b.seg(an + "." + std::to_string(t) + ".0", C[l], 0, 0));
      }
      b.finish(op);
      ConvOpHost op2 = b.base(an + ".3", H[l], W[l], co, cm_p / 8, EPI_SIGMOID_MUL, T);
      b.taps_3x3_s1(op2);
      for (int t = 0; t < T; ++t) {
        op2.g[t].src[0] = SrcRef{tmid, (cm_p / 8) * t, -1};
        op2.g[t].out_tensor = amask, op2.g[t].out_chunk0 = (co / 8) * t;
        op2.g[t].aux_tensor = sb[2 * l + 1];
        op2.g[t].segs.push_back(b.seg(an + "." + std::to_string(t) + ".3"));
      }
      b.finish(op2);
    }
    if (l == 3) {
      for (int t = 0; t < T; ++t) {
        int classes = (p.kind == 2 || (p.kind == 0 && t == 1)) ? 2 : 16;
        p.heads.push_back(HeadOpHost{amask, (co / 8) * t, classes, t});
        p.program.push_back(Op{3, (int)p.heads.size() - 1});
      }
      break;
    }
    // output layer l+1 and down-sampling
    const int oc = kOutCfg[l][1];
    int o = add_tensor(p, "o" + std::to_string(l + 1), oc * T, H[l], W[l], LAY_P8);
    int pooled = add_tensor(p, "p" + std::to_string(l + 1), oc * T, H[l + 1], W[l + 1], LAY_P8);
    ConvOpHost op3 = b.base("output_layer" + std::to_string(l + 1), H[l], W[l], oc, co / 8, EPI_RELU, T);
    b.taps_3x3_s1(op3);
    for (int t = 0; t < T; ++t) {
      op3.g[t].src[0] = SrcRef{amask, (co / 8) * t, -1};
      op3.g[t].out_tensor = o, op3.g[t].out_chunk0 = (oc / 8) * t;
      op3.g[t].segs.push_back(b.seg("output_layer" + std::to_string(l + 1) + "." + std::to_string(t) + ".0"));
    }
    choose_tiling(op3);
    // bf16 inference: ReLU + 2x2 max-pool are fused into the conv epilogue (shared-memory staging tile) when the layer
    // qualifies for the persistent kernel; the full-resolution tensor `o` is then never written.  output_layer3 (64->128,
    // 147 KB of weights per task) qualifies once its output channels are split into two groups (split_output_channels).
    const char* nofuse = getenv("MTLDAS_NO_POOL_FUSE");
    auto fusable = [&](const ConvOpHost& o) {
      const size_t staging = (size_t)(o.Cout / 8) * o.MT * kTileM * 16;
      return precision == 0 && !(nofuse && nofuse[0] == '1') && persist_stages_for(o, staging) > 0 &&
             (o.MT * kTileM) / (2 * (W[l] + 1)) >= 1;
    };
    bool fuse = fusable(op3);
    if (!fuse && precision == 0 && !training && p.use_nsplit && p.pool_split) {
      ConvOpHost sp = op3;
      if (split_output_channels(sp, 0, /*even_if_resident=*/true) && fusable(sp)) op3 = sp, fuse = true;
    }
    if (fuse)
      for (int g = 0; g < op3.ngroups; ++g) op3.g[g].pool_tensor = pooled;
    b.finish(op3, /*tiled=*/fuse);
    if (!fuse) {
      // (finish() tiles batch-aware and may still split the layer for the persistent kernel without the pool)
    }
    if (!fuse) {
      p.pools.push_back(PoolOpHost{"pool" + std::to_string(l + 1), o, pooled});
      p.program.push_back(Op{2, (int)p.pools.size() - 1});
    }
    a_prev = pooled;
  }

  // ---- workspace map ---------------------------------------------------------------------------
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off = (size_t)round_up((long long)(off + bytes), 256);
    return o;
  };
  for (auto& t : p.tensors) t.offset = take(t.bytes);
  for (auto& c : p.convs)
    for (int gi = 0; gi < c.ngroups; ++gi) {
      size_t wbytes = (size_t)c.ntaps * c.nchunks * 8 * c.Cout * (precision == 0 ? 2 : 4);
      c.g[gi].w_off = take(wbytes);
      c.g[gi].bias_off = take((size_t)c.Cout * 4);
    }
  p.stem_w_off = take(49 * 16 * 4);
  p.stem_b_off = take(16 * 4);
  p.stem_wu_off = take(2048);
  // host entry point: the batch goes through two staging slots in chunks (H2D of chunk i+1 under the compute of chunk
  // i); the slots hold the largest chunk of the schedule (api.cu::host_chunk_schedule)
  {
    const char* e = getenv("MTLDAS_HOST_CHUNK");
    int hc = e ? atoi(e) : 2048;
    if (hc < 64) hc = 64;
    p.host_chunk = std::min(mb, hc);
  }
  p.x_stage_off = take((size_t)2 * p.host_chunk * kPatchH * kPatchW * 4);
  p.out_stage_off = take((size_t)2 * p.host_chunk * 18 * 4);
  p.xh_stage_off = precision == 0 ? take((size_t)2 * p.host_chunk * kPatchH * kPatchW * 2) : 0;  // bf16 host patches
  p.ws_bytes = off;
  return pp;
}

// ---------------------------------------------------------------------------------------------
// device parameter blocks
// ---------------------------------------------------------------------------------------------
ConvParams Plan::conv_params(const ConvOpHost& op, int nb) const {
  ConvParams q{};
  q.H = op.Hv ? op.Hv : op.H, q.W = op.Wv ? op.Wv : op.W, q.Wp = op.W + 1, q.Pimg = (op.H + 1) * (op.W + 1);
  q.body = (long long)nb * q.Pimg;
  q.Cout = op.Cout, q.nchunks = op.nchunks, q.nsrc = op.nsrc, q.ntaps = op.ntaps;
  q.halo_lo = op.halo_lo, q.halo_hi = op.halo_hi;
  q.MT = op.MT, q.SP = op.MT * kTileM + op.halo_lo + op.halo_hi;
  q.tps = op.tps, q.NS = op.NS, q.nstages = op.nstages, q.tap_bytes = op.tap_bytes, q.tmem_cols = op.tmem_cols;
  q.epi = op.epi, q.swap_desc = swap_desc, q.ngroups = op.ngroups, q.ksplit = op.ksplit;
  q.magic_pimg = (1ull << 40) / (unsigned long long)q.Pimg + 1, q.magic_wp = (1ull << 40) / (unsigned long long)q.Wp + 1;
  q.pool_k = 0;
  const bool pool = op.g[0].pool_tensor >= 0;
  const size_t staging = pool ? (size_t)(op.Cout / 8) * op.MT * kTileM * 16 : 0;
  q.persist_stages = (use_persist || pool) ? persist_stages_for(op, staging) : 0;
  if (pool) {
    const TensorInfo& tp = tensors[op.g[0].pool_tensor];
    q.pool_k = (op.MT * kTileM) / (2 * q.Wp);
    q.pool_Wpo = tp.Wp, q.pool_Pimgo = tp.Pimg;
  }
  for (int t = 0; t < op.ntaps; ++t) q.tap[t] = op.tap[t];
  for (int gi = 0; gi < op.ngroups; ++gi) {
    const ConvGroupHost& g = op.g[gi];
    ConvGroupDev& d = q.g[gi];
    for (int s = 0; s < op.nsrc; ++s) {
      const TensorInfo& t = tensors[g.src[s].tensor];
      long long pos = (long long)g.src[s].chunk0 * t.plane + t.g_lo;
      if (g.src[s].phase >= 0) pos += (long long)g.src[s].phase * t.nchunks * t.plane;
      d.src[s].base = ws + t.offset + (size_t)pos * gran;
      d.src[s].plane_bytes = t.plane * gran;
    }
    d.w = ws + g.w_off;
    d.bias = reinterpret_cast<const float*>(ws + g.bias_off);
    d.out = view(g.out_tensor, g.out_chunk0);
    if (g.out_phase >= 0) {  // one phase plane of a phase-split tensor, addressed as a plain P8 tensor
      d.out.base += (size_t)g.out_phase * d.out.phase_stride * gran;
      d.out.layout = LAY_P8;
    }
    if (g.aux_tensor >= 0) d.aux = view(g.aux_tensor, g.aux_chunk0);
    d.aux_nchunks = g.aux_nchunks;
    if (g.pool_tensor >= 0) d.pool_out = view(g.pool_tensor, g.out_chunk0);
  }
  return q;
}

// ---------------------------------------------------------------------------------------------
// weight packing (eval)
// ---------------------------------------------------------------------------------------------
#define CK(call)                                                                      \
  do {                                                                                \
    cudaError_t e_ = (call);                                                          \
    if (e_ != cudaSuccess) {                                                          \
      set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
      return -2;                                                                      \
    }                                                                                 \
  } while (0)

PackSeg make_seg(const Plan& p, const PackSegHost& h, const void* const* state) {
  const LayerSpec& L = p.layers[h.layer];
  PackSeg s{};
  s.w = (const float*)state[L.i_w];
  s.conv_bias = L.i_b >= 0 ? (const float*)state[L.i_b] : nullptr;
  s.gamma = (const float*)state[L.i_gamma];
  s.beta = (const float*)state[L.i_beta];
  s.rmean = (const float*)state[L.i_rm];
  s.rvar = (const float*)state[L.i_rv];
  s.cout = L.cout, s.cin = L.cin, s.ksize = L.ksize;
  s.split = h.split, s.k_off = h.k_off, s.cout_off = h.cout_off;
  s.co_lo = h.co_lo, s.co_n = h.co_n;
  return s;
}

int plan_pack_eval(Plan& p, const void* const* state, cudaStream_t stream) {
  if (!p.ws) {
    set_error("pack_eval: no workspace bound");
    return -3;
  }
  for (auto& c : p.convs)
    for (int gi = 0; gi < c.ngroups; ++gi) {
      size_t wbytes = (size_t)c.ntaps * c.nchunks * 8 * c.Cout * (p.precision == 0 ? 2 : 4);
      CK(cudaMemsetAsync(p.ws + c.g[gi].w_off, 0, wbytes, stream));
      CK(cudaMemsetAsync(p.ws + c.g[gi].bias_off, 0, (size_t)c.Cout * 4, stream));
      for (auto& sh : c.g[gi].segs) {
        PackParams q{};
        q.seg = make_seg(p, sh, state);
        q.dst = p.ws + c.g[gi].w_off;
        q.bias_dst = reinterpret_cast<float*>(p.ws + c.g[gi].bias_off);
        q.Cout_p = c.Cout, q.K_tap = c.nchunks * 8;
        q.fp32 = p.precision == 0 ? 0 : 1;
        q.fold = 1, q.eps = 1e-5f;
        CK(launch_pack(q, stream));
      }
    }
  {
    PackParams q{};
    q.seg = make_seg(p, PackSegHost{p.stem_layer, INT_MAX, 0, 0}, state);
    q.dst = p.ws + p.stem_w_off;
    q.bias_dst = reinterpret_cast<float*>(p.ws + p.stem_b_off);
    q.Cout_p = 16, q.K_tap = 1, q.fp32 = 1, q.fold = 1, q.eps = 1e-5f;
    CK(launch_pack(q, stream));
    if (p.precision == 0) {  // bf16 tensor-core stem: taps are the K dimension (49 -> 64)
      CK(cudaMemsetAsync(p.ws + p.stem_wu_off, 0, 2048, stream));
      q.seg.k_from_tap = 2, q.dst = p.ws + p.stem_wu_off, q.K_tap = 64, q.fp32 = 0;
      CK(launch_pack(q, stream));
    }
  }
  p.packed = true;
  return 0;
}

// ---------------------------------------------------------------------------------------------
// forward (eval)
// ---------------------------------------------------------------------------------------------
static int run_pass(Plan& p, const float* x, int nb, float* out0, float* out1, cudaStream_t stream,
                    std::vector<cudaEvent_t>* evs = nullptr, bool first_pass = true, const StemSource* src = nullptr) {
  bool first_op = true;
  for (const Op& op : p.program) {
    // programmatic dependent launch for everything but the first kernel of a forward (what runs before it on the
    // stream is not ours); not under the per-launch event instrumentation
    const bool pdl = p.use_pdl && !evs && !(first_pass && first_op);
    first_op = false;
    if (evs) {
      cudaEvent_t e;
      CK(cudaEventCreate(&e));
      CK(cudaEventRecord(e, stream));
      evs->push_back(e);
    }
    if (op.kind == 0) {
      const TensorInfo& t = p.tensors[p.stem_out];
      StemParams q{};
      q.x = x;
      q.w = reinterpret_cast<const float*>(p.ws + p.stem_w_off);
      q.bias = reinterpret_cast<const float*>(p.ws + p.stem_b_off);
      q.out = p.view(p.stem_out, 0);
      q.Hin = kPatchH, q.Win = kPatchW, q.H = t.H, q.W = t.W, q.Wp = t.Wp, q.Pimg = t.Pimg;
      q.body = (long long)nb * t.Pimg;
      q.w_umma = (p.precision == 0 && p.use_umma_stem) ? p.ws + p.stem_wu_off : nullptr;
      if (src && src->nt > 0) {   // windows of a streaming array, read in place (tensor-core stem only)
        q.x_stride_a = src->stride_a, q.x_stride_b = src->stride_b, q.x_row_pitch = src->row_pitch;
        q.x_nt = src->nt, q.x_n0 = src->n0;
        if (!q.w_umma) {
          set_error("forward_windows: needs the bf16 tensor-core stem");
          return -4;
        }
        cudaError_t se = launch_stem_umma(q, stream, pdl);
        if (se == cudaErrorInvalidValue) {
          cudaGetLastError();
          set_error("forward_windows: the array must be 16-byte aligned with n_time a multiple of 4");
          return -4;
        }
        CK(se);
      } else {
        CK(launch_stem(q, p.precision, stream, pdl));
      }
    } else if (op.kind == 1) {
      ConvParams q = p.conv_params(p.convs[op.index], nb);
      static int pool_dbg = -1;
      if (pool_dbg < 0) {
        const char* e = getenv("MTLDAS_POOL_DEBUG");
        pool_dbg = (e && e[0] == '1') ? 1 : 0;
      }
      if (pool_dbg && q.pool_k > 0 && q.persist_stages > 0) {  // diagnostic: per-iteration timeline of CTA 0
        long long* dbg = nullptr;
        CK(cudaMalloc(&dbg, 16 * 8 * sizeof(long long)));
        CK(cudaMemset(dbg, 0, 16 * 8 * sizeof(long long)));
        q.dbg = dbg;
        CK(launch_conv(q, p.precision, stream, false));
        CK(cudaStreamSynchronize(stream));
        long long h[16 * 8];
        CK(cudaMemcpy(h, dbg, sizeof(h), cudaMemcpyDeviceToHost));
        cudaFree(dbg);
        fprintf(stderr, "[pooldbg] %s MT %d pool_k %d Cout %d:\n", p.convs[op.index].name.c_str(), q.MT, q.pool_k, q.Cout);
        for (int it = 0; it < 12; ++it) {
          const long long* r = &h[(size_t)it * 8];
          if (!r[0]) break;
          fprintf(stderr, "[pooldbg]  it %2d mma: start %lld slab-ready +%lld acc-free +%lld issued +%lld | epi(w2): wait-start %lld acc-ready +%lld drained +%lld pooled +%lld\n",
                  it, r[0] - h[0], r[1] - r[0], r[2] - r[1], r[3] - r[2], r[4] - h[0], r[5] - r[4], r[6] - r[5], r[7] - r[6]);
        }
        continue;
      }
      CK(launch_conv(q, p.precision, stream, pdl));
    } else if (op.kind == 2) {
      const PoolOpHost& po = p.pools[op.index];
      const TensorInfo& ti = p.tensors[po.in_tensor];
      const TensorInfo& to = p.tensors[po.out_tensor];
      PoolParams q{};
      q.in = p.view(po.in_tensor, 0), q.out = p.view(po.out_tensor, 0);
      q.nchunks = ti.nchunks;
      q.H = ti.H, q.W = ti.W, q.Wp = ti.Wp, q.Pimg = ti.Pimg;
      q.Ho = to.H, q.Wo = to.W, q.Wpo = to.Wp, q.Pimgo = to.Pimg;
      q.body_o = (long long)nb * to.Pimg;
      CK(launch_pool(q, p.precision, stream, pdl));
    } else {
      const HeadOpHost& h = p.heads[op.index];
      const TensorInfo& t = p.tensors[h.in_tensor];
      HeadParams q{};
      q.in = p.view(h.in_tensor, h.chunk0);
      q.out = h.out_index == 0 ? out0 : out1;
      q.nb = nb, q.classes = h.classes, q.npos = t.Pimg;
      q.inv_count = 1.f / (float)(t.H * t.W * (128 / h.classes));
      if (!q.out) {
        set_error("forward: output %d is NULL", h.out_index);
        return -1;
      }
      CK(launch_head(q, p.precision, stream, pdl));
    }
    ++p.last_launches;
    if (evs) {
      cudaEvent_t e;
      CK(cudaEventCreate(&e));
      CK(cudaEventRecord(e, stream));
      evs->push_back(e);
    }
  }
  return 0;
}

// One forward with a CUDA-event pair around every launch: ms_per_op[i] = device time of program op i summed
// over the micro-batch passes (bench.py's live per-kernel durations for the roofline).
int plan_profile_eval(Plan& p, const float* x, int n, float* out0, float* out1, cudaStream_t stream,
                      float* ms_per_op, int cap) {
  if (!p.ws || !p.packed) {
    set_error("profile_eval: bind a workspace and pack weights first");
    return -3;
  }
  const int nops = (int)p.program.size();
  if (cap < nops) {
    set_error("profile_eval: need room for %d ops", nops);
    return -1;
  }
  for (int i = 0; i < nops; ++i) ms_per_op[i] = 0.f;
  std::vector<cudaEvent_t> evs;
  int c0 = p.heads[0].classes, c1 = p.heads.size() > 1 ? p.heads[1].classes : 0;
  p.last_launches = 0;
  int rc = 0;
  for (int s = 0; s < n && !rc; s += p.mb) {
    int nb = std::min(p.mb, n - s);
    rc = run_pass(p, x + (size_t)s * kPatchH * kPatchW, nb, out0 + (size_t)s * c0,
                  out1 ? out1 + (size_t)s * c1 : nullptr, stream, &evs);
  }
  if (!rc && cudaStreamSynchronize(stream) != cudaSuccess) rc = -2;
  if (!rc)
    for (size_t i = 0; i + 1 < evs.size(); i += 2) {
      float ms = 0.f;
      cudaEventElapsedTime(&ms, evs[i], evs[i + 1]);
      ms_per_op[(i / 2) % nops] += ms;
    }
  for (cudaEvent_t e : evs) cudaEventDestroy(e);
  return rc;
}

static int forward_eager(Plan& p, const float* x, int n, float* out0, float* out1, cudaStream_t stream) {
  p.last_launches = 0;
  int c0 = p.heads[0].classes, c1 = p.heads.size() > 1 ? p.heads[1].classes : 0;
  for (int s = 0; s < n; s += p.mb) {
    int nb = std::min(p.mb, n - s);
    int rc = run_pass(p, x + (size_t)s * kPatchH * kPatchW, nb, out0 + (size_t)s * c0,
                      out1 ? out1 + (size_t)s * c1 : nullptr, stream, nullptr, s == 0);
    if (rc) return rc;
  }
  return 0;
}

int plan_forward_windows(Plan& p, const float* array, long long n_channels, long long n_time, int first_channel,
                         int window_stride, int n_windows, float* out0, float* out1, cudaStream_t stream) {
  if (!p.ws || !p.packed) {
    set_error("forward_windows: bind a workspace and pack weights first");
    return -3;
  }
  if (p.precision != 0 || !p.use_umma_stem || (n_time & 3) || (reinterpret_cast<uintptr_t>(array) & 15)) {
    set_error("forward_windows: needs the bf16 plan with the tensor-core stem, a 16-byte aligned array and n_time %% 4 == 0 "
              "(otherwise gather the windows with mtldas_window_gather)");
    return -4;
  }
  const int nt = (int)(n_time / kPatchW);
  const long long n = (long long)n_windows * nt;
  if (nt < 1 || n_windows < 0 || n > (1ll << 30) ||
      (long long)first_channel + (long long)(n_windows > 0 ? n_windows - 1 : 0) * window_stride + kPatchH > n_channels) {
    set_error("forward_windows: windows do not fit the array");
    return -1;
  }
  StemSource src;
  src.stride_a = (long long)window_stride * n_time, src.stride_b = kPatchW, src.row_pitch = n_time, src.nt = nt;
  const float* base = array + (long long)first_channel * n_time;
  const int c0 = p.heads[0].classes, c1 = p.heads.size() > 1 ? p.heads[1].classes : 0;
  p.last_launches = 0;
  for (long long s = 0; s < n; s += p.mb) {
    const int nb = (int)std::min<long long>(p.mb, n - s);
    src.n0 = (int)s;
    int rc = run_pass(p, base, nb, out0 + (size_t)s * c0, out1 ? out1 + (size_t)s * c1 : nullptr, stream, nullptr, s == 0, &src);
    if (rc) return rc;
  }
  return 0;
}

Plan::~Plan() {
  for (auto& g : graphs)
    if (g.exec) cudaGraphExecDestroy(g.exec);
  if (cap_stream) cudaStreamDestroy(cap_stream);
  if (copy_stream) cudaStreamDestroy(copy_stream);
  for (auto e : {ev_start, ev_h2d[0], ev_h2d[1], ev_comp[0], ev_comp[1]})
    if (e) cudaEventDestroy(e);
}

// The layer program is ~36 launches per micro-batch pass (576 for a 4096 batch): replaying it as one CUDA graph
// removes the per-launch host cost.  First use of a buffer set runs eagerly, the second is captured, later ones replay.
int plan_forward_eval(Plan& p, const float* x, int n, float* out0, float* out1, cudaStream_t stream) {
  if (!p.ws || !p.packed) {
    set_error("forward_eval: bind a workspace and pack weights first");
    return -3;
  }
  if (!p.use_graphs) return forward_eager(p, x, n, out0, out1, stream);
  Plan::GraphEntry* e = nullptr;
  for (auto& g : p.graphs)
    if (g.x == x && g.n == n && g.o0 == out0 && g.o1 == out1) e = &g;
  if (!e) {
    if (p.graphs.size() >= 24) {  // evict the least recently used entry (the host schedule uses 2 slots x 5 chunk sizes)
      size_t victim = 0;
      for (size_t i = 1; i < p.graphs.size(); ++i)
        if (p.graphs[i].stamp < p.graphs[victim].stamp) victim = i;
      if (p.graphs[victim].exec) cudaGraphExecDestroy(p.graphs[victim].exec);
      p.graphs.erase(p.graphs.begin() + victim);
    }
    p.graphs.push_back(Plan::GraphEntry{x, n, out0, out1, nullptr, 0, 0, 0});
    e = &p.graphs.back();
  }
  e->stamp = ++p.graph_clock;
  e->uses++;
  if (e->exec) {
    CK(cudaGraphLaunch(e->exec, stream));
    p.last_launches = e->launches;
    return 0;
  }
  if (e->uses < 2) return forward_eager(p, x, n, out0, out1, stream);
  cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream, &st) != cudaSuccess) cudaGetLastError();
  if (st != cudaStreamCaptureStatusNone) return forward_eager(p, x, n, out0, out1, stream);  // caller is capturing
  // capture on a private stream (the caller's may be the legacy default stream, which cannot be captured);
  // the instantiated graph is then launched on the caller's stream.
  if (!p.cap_stream) CK(cudaStreamCreateWithFlags(&p.cap_stream, cudaStreamNonBlocking));
  CK(cudaStreamBeginCapture(p.cap_stream, cudaStreamCaptureModeThreadLocal));
  int rc = forward_eager(p, x, n, out0, out1, p.cap_stream);
  cudaGraph_t graph = nullptr;
  cudaError_t ce = cudaStreamEndCapture(p.cap_stream, &graph);
  if (rc || ce != cudaSuccess) {
    if (graph) cudaGraphDestroy(graph);
    cudaGetLastError();
    if (!rc) set_error("graph capture failed: %s", cudaGetErrorString(ce));
    return rc ? rc : -2;
  }
  e->launches = p.last_launches;
  ce = cudaGraphInstantiate(&e->exec, graph, 0);
  cudaGraphDestroy(graph);
  if (ce != cudaSuccess) {
    e->exec = nullptr;
    set_error("graph instantiate failed: %s", cudaGetErrorString(ce));
    return -2;
  }
  CK(cudaGraphLaunch(e->exec, stream));
  return 0;
}

// ---------------------------------------------------------------------------------------------
// debug read-back and the single-convolution entry point (unit parity tests)
// ---------------------------------------------------------------------------------------------
int plan_read_tensor(Plan& p, const char* name, int nb, float* out_nchw, cudaStream_t stream) {
  int ti = p.find_tensor(name);
  if (ti < 0 || !p.ws) {
    set_error("read_tensor: unknown tensor '%s' or no workspace", name);
    return -1;
  }
  const TensorInfo& t = p.tensors[ti];
  CK(launch_view_to_nchw(p.view(ti, 0), out_nchw, nb, t.C, t.H, t.W, p.precision, stream));
  return 0;
}

int single_conv(const float* x, const float* x2, const float* weight, const float* bias, const float* aux,
                float* out, int n, int cin, int cin2, int h, int w, int cout, int ksize, int stride, int epi,
                int precision, int aux_phase_split, int swap_desc) {
  if (n < 1 || (cin % 16) || (cin2 % 16) || (cout % 16) || cout > 256 || cin < 16 || (ksize != 1 && ksize != 3) ||
      (stride != 1 && stride != 2) || (x2 && !(ksize == 1 && stride == 1 && cin2 == cin)) || (!x2 && cin2) ||
      ((epi == EPI_ADD_RELU || epi == EPI_SIGMOID_MUL) && !aux) || epi < 0 || epi > 5 || epi == 4 ||
      (epi == 5 && (precision != 0 || ksize != 3 || stride != 1))) {
    set_error("conv2d_nchw: unsupported arguments");
    return -1;
  }
  const bool want_pool = epi == 5;   // MTLDAS_EPI_RELU_POOL: ReLU + MaxPool2d(2,2,ceil) fused into the epilogue
  if (want_pool) epi = EPI_RELU;
  Plan p;
  p.kind = -1, p.mb = n, p.precision = precision, p.ntasks = 1, p.gran = precision == 0 ? 16 : 32;
  p.swap_desc = swap_desc;
  Builder b(p);
  const int Ho = stride == 1 ? h : (h + 1) / 2, Wo = stride == 1 ? w : (w + 1) / 2;
  int tin = add_tensor(p, "x", cin, h, w, stride == 2 ? LAY_P8S2 : LAY_P8);
  int tin2 = x2 ? add_tensor(p, "x2", cin2, h, w, LAY_P8) : -1;
  int tout = add_tensor(p, "y", cout, Ho, Wo, LAY_P8);
  int taux = aux ? add_tensor(p, "aux", cout, Ho, Wo, aux_phase_split ? LAY_P8S2 : LAY_P8) : -1;
  const int Hq = (Ho + 1) / 2, Wq = (Wo + 1) / 2;
  int tpool = want_pool ? add_tensor(p, "yp", cout, Hq, Wq, LAY_P8) : -1;
  ConvOpHost op = b.base("conv", Ho, Wo, cout, cin / 8, epi, 1);
  if (ksize == 3 && stride == 1) {
    b.taps_3x3_s1(op);
    op.g[0].src[0] = SrcRef{tin, 0, -1};
  } else if (ksize == 3) {
    b.taps_3x3_s2(op);
    for (int ph = 0; ph < 4; ++ph) op.g[0].src[ph] = SrcRef{tin, 0, ph};
  } else if (stride == 2) {
    b.taps_1x1(op, 1);
    op.g[0].src[0] = SrcRef{tin, 0, 0};
  } else {
    b.taps_1x1(op, x2 ? 2 : 1);
    op.g[0].src[0] = SrcRef{tin, 0, -1};
    if (x2) op.g[0].src[1] = SrcRef{tin2, 0, -1};
  }
  op.g[0].out_tensor = tout;
  op.g[0].aux_tensor = taux;
  choose_tiling(op);
  // the same re-grouping the inference plan applies (layers whose weights would not stay resident: N-split halves)
  const char* nosplit = getenv("MTLDAS_NO_NSPLIT");
  if (precision == 0 && !(nosplit && nosplit[0] == '1')) split_output_channels(op, 0, /*even_if_resident=*/want_pool && cout >= 128);
  if (want_pool) {   // same eligibility rule as build_plan's output layers
    const size_t staging = (size_t)(op.Cout / 8) * op.MT * kTileM * 16;
    if (persist_stages_for(op, staging) == 0 || (op.MT * kTileM) / (2 * (Wo + 1)) < 1) {
      set_error("conv2d_nchw: this shape does not qualify for the fused ReLU + max-pool epilogue");
      return -4;
    }
    for (int gi = 0; gi < op.ngroups; ++gi) op.g[gi].pool_tensor = tpool;
  }
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off = (size_t)round_up((long long)(off + bytes), 256);
    return o;
  };
  for (auto& t : p.tensors) t.offset = take(t.bytes);
  const size_t wbytes = (size_t)op.ntaps * op.nchunks * 8 * op.Cout * (precision == 0 ? 2 : 4);
  for (int gi = 0; gi < op.ngroups; ++gi) {
    op.g[gi].w_off = take(wbytes);
    op.g[gi].bias_off = take((size_t)op.Cout * 4);
  }
  CK(cudaMalloc(&p.ws, off));
  int rc = 0;
  do {
    cudaError_t e;
#define CKB(call)                                                                          \
  if ((e = (call)) != cudaSuccess) {                                                      \
    set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e), __FILE__, __LINE__); \
    rc = -2;                                                                              \
    break;                                                                                \
  }
    CKB(cudaMemset(p.ws, 0, off));
    CKB(launch_nchw_to_view(x, p.view(tin, 0), n, cin, h, w, precision, 0));
    if (x2) CKB(launch_nchw_to_view(x2, p.view(tin2, 0), n, cin2, h, w, precision, 0));
    if (aux) CKB(launch_nchw_to_view(aux, p.view(taux, 0), n, cout, Ho, Wo, precision, 0));
    for (int gi = 0; gi < op.ngroups && !rc; ++gi) {
      PackParams q{};
      q.seg.w = weight;
      q.seg.cout = cout, q.seg.cin = cin + cin2, q.seg.ksize = ksize;
      q.seg.split = x2 ? cin : INT_MAX;
      if (op.ngroups > 1) q.seg.co_lo = gi * op.Cout, q.seg.co_n = op.Cout;
      q.dst = p.ws + op.g[gi].w_off;
      q.bias_dst = reinterpret_cast<float*>(p.ws + op.g[gi].bias_off);
      q.Cout_p = op.Cout, q.K_tap = op.nchunks * 8, q.fp32 = precision == 0 ? 0 : 1, q.fold = 0, q.eps = 0.f;
      CKB(launch_pack(q, 0));
      if (bias)
        CKB(cudaMemcpy(p.ws + op.g[gi].bias_off, bias + (size_t)gi * op.Cout, (size_t)op.Cout * 4, cudaMemcpyDeviceToDevice));
    }
    if (rc) break;
    ConvParams cp = p.conv_params(op, n);
    const char* dbg_env = getenv("MTLDAS_CONV_DEBUG");
    if (dbg_env && precision == 0) {
      // diagnostics: per-CTA timeline of a warm (second) launch
      int ntile = (int)((cp.body + cp.MT * kTileM - 1) / (cp.MT * kTileM));
      long long* dbg = nullptr;
      CKB(cudaMalloc(&dbg, (size_t)ntile * 16 * sizeof(long long)));
      CKB(cudaMemset(dbg, 0, (size_t)ntile * 16 * sizeof(long long)));
      CKB(launch_conv(cp, precision, 0));
      CKB(cudaDeviceSynchronize());
      cp.dbg = dbg;
      CKB(launch_conv(cp, precision, 0));
      CKB(cudaDeviceSynchronize());
      std::vector<long long> h((size_t)ntile * 16);
      CKB(cudaMemcpy(h.data(), dbg, h.size() * sizeof(long long), cudaMemcpyDeviceToHost));
      cudaFree(dbg);
      long long t0 = h[0], t1 = 0;
      for (int i = 0; i < ntile; ++i) {
        t0 = std::min(t0, h[(size_t)i * 16]);
        t1 = std::max(t1, h[(size_t)i * 16]);
      }
      if (cp.persist_stages > 0) {
        fprintf(stderr, "[convdbg] persistent kernel, stages %d, tiles %d; CTA 0 per iteration (SM clocks relative to it 0):\n",
                cp.persist_stages, ntile);
        for (int it = 0; it < 12; ++it) {
          const long long* r = &h[(size_t)it * 8];
          if (!r[0]) break;
          fprintf(stderr, "[convdbg]  it %2d mma: start %lld slab-ready +%lld acc-free +%lld issued +%lld | epi(w2): wait-start %lld acc-ready +%lld done +%lld\n",
                  it, r[0] - h[0], r[1] - r[0], r[2] - r[1], r[3] - r[2], r[4] - h[0], r[5] - r[4], r[6] - r[5]);
        }
        break;
      }
      fprintf(stderr, "[convdbg] tiles=%d MT=%d NS=%d tps=%d smem=%zu last CTA start +%lld ns\n", ntile, cp.MT, cp.NS,
              cp.tps, conv_umma_smem_bytes(cp), t1 - t0);
      for (int i = 0; i < ntile; i += std::max(1, ntile / 6)) {
        const long long* r = &h[(size_t)i * 16];
        fprintf(stderr,
                "[convdbg] cta %4d start+%lld ns | clk: setup %lld slab %lld w0 %lld issued %lld acc %lld epi %lld | "
                "detail: first-ld %lld first-cols %lld loop-end %lld final-sync %lld\n",
                i, r[0] - t0, r[2] - r[1], r[3] - r[2], r[4] - r[3], r[5] - r[4], r[6] - r[5], r[7] - r[6], r[8] - r[6],
                r[9] - r[8], r[10] - r[6], r[7] - r[10]);
      }
      cp.dbg = nullptr;
    }
    CKB(launch_conv(cp, precision, 0));
    if (want_pool) {
      CKB(launch_view_to_nchw(p.view(tpool, 0), out, n, cout, Hq, Wq, precision, 0));
    } else {
      CKB(launch_view_to_nchw(p.view(tout, 0), out, n, cout, Ho, Wo, precision, 0));
    }
    CKB(cudaDeviceSynchronize());
#undef CKB
  } while (0);
  cudaFree(p.ws);
  p.ws = nullptr;
  return rc;
}

// ---------------------------------------------------------------------------------------------
// JSON description
// ---------------------------------------------------------------------------------------------
std::string Plan::describe() const {
  std::ostringstream o;
  o << "{\"kind\":" << kind << ",\"micro_batch\":" << mb << ",\"precision\":" << precision << ",\"gran\":" << gran
    << ",\"workspace_bytes\":" << ws_bytes << ",\"stem_out\":\"" << tensors[stem_out].name << "\",\"layers\":[";
  for (size_t i = 0; i < layers.size(); ++i) {
    const LayerSpec& L = layers[i];
    o << (i ? "," : "") << "{\"conv\":\"" << L.conv << "\",\"bn\":\"" << L.bn << "\",\"cin\":" << L.cin
      << ",\"cout\":" << L.cout << ",\"k\":" << L.ksize << ",\"stride\":" << L.stride << ",\"pad\":" << L.pad
      << ",\"bias\":" << (L.bias ? 1 : 0) << "}";
  }
  o << "],\"tensors\":[";
  for (size_t i = 0; i < tensors.size(); ++i) {
    const TensorInfo& t = tensors[i];
    o << (i ? "," : "") << "{\"name\":\"" << t.name << "\",\"C\":" << t.C << ",\"H\":" << t.H << ",\"W\":" << t.W
      << ",\"layout\":" << t.layout << ",\"Hs\":" << t.Hs << ",\"Ws\":" << t.Ws << ",\"Wp\":" << t.Wp
      << ",\"Pimg\":" << t.Pimg << ",\"g_lo\":" << t.g_lo << ",\"plane\":" << t.plane << ",\"nchunks\":" << t.nchunks
      << ",\"nphase\":" << t.nphase << ",\"offset\":" << t.offset << ",\"bytes\":" << t.bytes << "}";
  }
  o << "],\"program\":[";
  for (size_t i = 0; i < program.size(); ++i) {
    const Op& op = program[i];
    o << (i ? "," : "");
    if (op.kind == 0) {
      o << "{\"op\":\"stem\",\"layer\":\"" << layers[stem_layer].conv << "\",\"out\":\"" << tensors[stem_out].name
        << "\"}";
    } else if (op.kind == 1) {
      const ConvOpHost& c = convs[op.index];
      o << "{\"op\":\"conv\",\"name\":\"" << c.name << "\",\"H\":" << c.H << ",\"W\":" << c.W << ",\"Cout\":" << c.Cout
        << ",\"nchunks\":" << c.nchunks << ",\"nsrc\":" << c.nsrc << ",\"halo_lo\":" << c.halo_lo
        << ",\"halo_hi\":" << c.halo_hi << ",\"epi\":" << c.epi << ",\"MT\":" << c.MT << ",\"tps\":" << c.tps
        << ",\"NS\":" << c.NS << ",\"nstages\":" << c.nstages << ",\"tap_bytes\":" << c.tap_bytes
        << ",\"tmem_cols\":" << c.tmem_cols << ",\"taps\":[";
      for (int t = 0; t < c.ntaps; ++t) o << (t ? "," : "") << "[" << c.tap[t].src << "," << c.tap[t].off << "]";
      o << "],\"groups\":[";
      for (int gi = 0; gi < c.ngroups; ++gi) {
        const ConvGroupHost& g = c.g[gi];
        o << (gi ? "," : "") << "{\"src\":[";
        for (int s = 0; s < c.nsrc; ++s)
          o << (s ? "," : "") << "[\"" << tensors[g.src[s].tensor].name << "\"," << g.src[s].chunk0 << ","
            << g.src[s].phase << "]";
        o << "],\"out\":[\"" << tensors[g.out_tensor].name << "\"," << g.out_chunk0 << "],\"aux\":";
        if (g.aux_tensor >= 0)
          o << "[\"" << tensors[g.aux_tensor].name << "\"," << g.aux_chunk0 << "," << g.aux_nchunks << "]";
        else
          o << "null";
        if (g.pool_tensor >= 0) o << ",\"pool\":\"" << tensors[g.pool_tensor].name << "\"";
        o << ",\"w_off\":" << g.w_off << ",\"bias_off\":" << g.bias_off << ",\"segs\":[";
        for (size_t si = 0; si < g.segs.size(); ++si) {
          const PackSegHost& s = g.segs[si];
          o << (si ? "," : "") << "{\"layer\":\"" << layers[s.layer].conv << "\",\"split\":"
            << (s.split == INT_MAX ? -1 : s.split) << ",\"k_off\":" << s.k_off << ",\"cout_off\":" << s.cout_off
            << ",\"co_lo\":" << s.co_lo << ",\"co_n\":" << s.co_n << "}";
        }
        o << "]}";
      }
      o << "]}";
    } else if (op.kind == 2) {
      const PoolOpHost& po = pools[op.index];
      o << "{\"op\":\"pool\",\"in\":\"" << tensors[po.in_tensor].name << "\",\"out\":\""
        << tensors[po.out_tensor].name << "\"}";
    } else {
      const HeadOpHost& h = heads[op.index];
      o << "{\"op\":\"head\",\"in\":\"" << tensors[h.in_tensor].name << "\",\"chunk0\":" << h.chunk0
        << ",\"classes\":" << h.classes << ",\"out\":" << h.out_index << "}";
    }
  }
  o << "]}";
  return o.str();
}

}  // namespace mtldas
