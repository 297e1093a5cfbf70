// mtldas_conv2d_nchw_static: the single fused convolution for callers that must not allocate device memory.
//
// Who cannot allocate: a DAS acquisition / streaming-inference loop that keeps copies and kernels in flight on
// several streams cannot afford cudaMalloc or cudaFree (both synchronise the whole device and would stall the
// interrogator's transfers), and a process whose GPU memory has already been handed to one pool (PyTorch's caching
// allocator, or a fixed arena sized at start-up) may simply get an out-of-memory error from cudaMalloc even though
// "its" memory is free.  mtldas_conv2d_nchw (plan.cu) allocates its temporaries per call and is therefore unusable
// from such code.  This variant keeps the temporaries -- the P8 / phase-split copies of the inputs, the packed
// weight image, the bias, the P8 output -- in ONE fixed array that is part of the library image.
//
//   * the array is kStaticScratchBytes (32 MiB) and every GPU that loads the library carries it until the process
//     exits; a call that needs more is refused with an error, nothing is allocated instead;
//   * in host code the array's NAME is only a host-side handle, not a device address, so the device address is
//     looked up with cudaGetSymbolAddress the first time the entry point runs on a device (with that device
//     current) and cached per device.  Make that first call before starting any stream capture;
//   * every call uses the same array and finds the previous call's bytes in it: the part a call uses is cleared
//     first (the zero padding of the P8 layout lives there), and two calls whose GPU work overlaps -- different
//     streams, or different host threads -- race on it.  Serialise them.
#include <mutex>

#include "../../include/mtldas.h"
#include "kernels.h"
#include "plan.h"

namespace {

using namespace mtldas;

constexpr size_t kStaticScratchBytes = 32u << 20;
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
char* g_scratch_addr[kMaxDevices];  // device address of g_static_scratch, per device ordinal (host-side cache)
std::mutex g_scratch_mutex;

// Device address of the scratch array on the current device (looked up once per device).
int scratch_for_current_device(char** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) {
    set_error("conv2d_nchw_static: cudaGetDevice failed: %s", cudaGetErrorString(e));
    return MTLDAS_ERR_CUDA;
  }
  if (dev < 0 || dev >= kMaxDevices) {
    set_error("conv2d_nchw_static: device ordinal %d is outside the %d ordinals this entry point keeps an address for", dev,
              kMaxDevices);
    return MTLDAS_ERR_UNSUPPORTED;
  }
  std::lock_guard<std::mutex> lock(g_scratch_mutex);
  if (!g_scratch_addr[dev]) {
    void* ptr = nullptr;
    e = cudaGetSymbolAddress(&ptr, g_static_scratch);
    if (e != cudaSuccess || !ptr) {
      set_error("conv2d_nchw_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return MTLDAS_ERR_CUDA;
    }
    g_scratch_addr[dev] = static_cast<char*>(ptr);
  }
  *out = g_scratch_addr[dev];
  return 0;
}

// Geometry of one P8 / P8S2 tensor placed in the scratch (same rules as the plan's tensors, types.h).
struct Geo {
  int C, H, W, layout, Wp, Pimg, g_lo, nchunks, nphase;
  long long plane;
  size_t offset, bytes;
};

long long round_up_ll(long long v, long long a) { return (v + a - 1) / a * a; }

Geo make_geo(int nb, int C, int H, int W, int layout, int gran) {
  Geo t{};
  t.C = C, t.H = H, t.W = W, t.layout = layout;
  const int Hs = layout == LAY_P8 ? H : (H + 1) / 2, Ws = layout == LAY_P8 ? W : (W + 1) / 2;
  t.Wp = Ws + 1;
  t.Pimg = (Hs + 1) * t.Wp;
  t.g_lo = (int)round_up_ll(t.Wp + 1, 8);
  t.plane = round_up_ll((long long)t.g_lo + (long long)nb * t.Pimg + kMaxTile + t.Wp + 1, 8);
  t.nchunks = C / 8;
  t.nphase = layout == LAY_P8 ? 1 : 4;
  t.bytes = (size_t)t.nphase * t.nchunks * t.plane * gran;
  return t;
}

ActView make_view(char* scratch, const Geo& t, int gran) {
  ActView v{};
  v.base = scratch + t.offset;
  v.plane = t.plane;
  v.phase_stride = t.layout == LAY_P8 ? 0 : (long long)t.nchunks * t.plane;
  v.g_lo = t.g_lo;
  v.layout = t.layout;
  v.Wp2 = t.Wp;
  v.Pimg2 = t.Pimg;
  (void)gran;
  return v;
}

ConvSrcDev make_src(char* scratch, const Geo& t, int phase, int gran) {
  long long pos = t.g_lo;
  if (phase >= 0) pos += (long long)phase * t.nchunks * t.plane;
  ConvSrcDev s{};
  s.base = scratch + t.offset + (size_t)pos * gran;
  s.plane_bytes = t.plane * gran;
  return s;
}

}  // namespace

extern "C" int mtldas_conv2d_nchw_static(const float* dev_x, const float* dev_x2, const float* dev_weight,
                                         const float* dev_bias, const float* dev_aux, float* dev_out, int n, int cin,
                                         int cin2, int h, int w, int cout, int ksize, int stride, int epilogue,
                                         int precision, int aux_phase_split, void* stream) {
  const bool has_aux = (epilogue == EPI_ADD_RELU || epilogue == EPI_SIGMOID_MUL);
  if (!dev_x || !dev_weight || !dev_out || n < 1 || h < 1 || w < 1 || cin < 16 || (cin % 16) || (cin2 % 16) || (cout % 16) ||
      cout < 16 || cout > 256 || (ksize != 1 && ksize != 3) || (stride != 1 && stride != 2) || epilogue < 0 ||
      epilogue > 3 || precision < 0 || precision > 1 || (dev_x2 && !(ksize == 1 && stride == 1 && cin2 == cin)) ||
      (!dev_x2 && cin2) || (has_aux && !dev_aux)) {
    set_error("conv2d_nchw_static: unsupported arguments");
    return MTLDAS_ERR_ARG;
  }
  char* scratch = nullptr;
  int rc = scratch_for_current_device(&scratch);
  if (rc) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const int gran = precision == 0 ? 16 : 32;
  const int Ho = stride == 1 ? h : (h + 1) / 2, Wo = stride == 1 ? w : (w + 1) / 2;

  // ---- lay the temporaries out in the scratch -------------------------------------------------------
  Geo gx = make_geo(n, cin, h, w, stride == 2 ? LAY_P8S2 : LAY_P8, gran);
  Geo gx2 = make_geo(n, dev_x2 ? cin2 : 16, h, w, LAY_P8, gran);
  Geo gy = make_geo(n, cout, Ho, Wo, LAY_P8, gran);
  Geo ga = make_geo(n, cout, Ho, Wo, aux_phase_split ? LAY_P8S2 : LAY_P8, gran);
  ConvOpHost op{};  // only its plain fields are used here (tiling heuristics shared with the plan)
  op.H = Ho, op.W = Wo, op.Cout = cout, op.nchunks = cin / 8, op.epi = epilogue, op.ngroups = 1;
  const int Wp = Wo + 1;
  if (ksize == 3 && stride == 1) {
    op.nsrc = 1, op.ntaps = 9;
    for (int r = 0; r < 3; ++r)
      for (int s = 0; s < 3; ++s) op.tap[r * 3 + s] = ConvTapDev{0, (r - 1) * Wp + (s - 1)};
    op.halo_lo = op.halo_hi = Wp + 1;
  } else if (ksize == 3) {
    op.nsrc = 4, op.ntaps = 9;
    for (int r = 0; r < 3; ++r)
      for (int s = 0; s < 3; ++s)
        op.tap[r * 3 + s] = ConvTapDev{((r == 1) ? 0 : 1) * 2 + ((s == 1) ? 0 : 1),
                                       ((r == 0) ? -1 : 0) * Wp + ((s == 0) ? -1 : 0)};
    op.halo_lo = Wp + 1, op.halo_hi = 0;
  } else {
    op.nsrc = op.ntaps = dev_x2 ? 2 : 1;
    for (int s = 0; s < op.nsrc; ++s) op.tap[s] = ConvTapDev{s, 0};
    op.halo_lo = op.halo_hi = 0;
  }
  choose_tiling(op);
  size_t off = 0;
  auto take = [&](size_t bytes) {
    size_t o = off;
    off = (size_t)round_up_ll((long long)(off + bytes), 256);
    return o;
  };
  gx.offset = take(gx.bytes);
  if (dev_x2) gx2.offset = take(gx2.bytes);
  gy.offset = take(gy.bytes);
  if (has_aux) ga.offset = take(ga.bytes);
  const size_t wbytes = (size_t)op.ntaps * op.nchunks * 8 * cout * (precision == 0 ? 2 : 4);
  const size_t w_off = take(wbytes);
  const size_t b_off = take((size_t)cout * 4);
  if (off > kStaticScratchBytes) {
    set_error("conv2d_nchw_static: this call needs %zu bytes of scratch but the built-in scratch holds %zu; "
              "use mtldas_conv2d_nchw (which allocates) or a smaller batch",
              off, kStaticScratchBytes);
    return MTLDAS_ERR_UNSUPPORTED;
  }

#define CKS(call)                                                                                       \
  do {                                                                                                  \
    cudaError_t e_ = (call);                                                                            \
    if (e_ != cudaSuccess) {                                                                            \
      set_error("conv2d_nchw_static: %s failed: %s", #call, cudaGetErrorString(e_));                    \
      return MTLDAS_ERR_CUDA;                                                                           \
    }                                                                                                   \
  } while (0)

  // ---- previous contents are stale: everything this call touches starts from zero ---------------------
  CKS(cudaMemsetAsync(scratch, 0, off, st));
  CKS(launch_nchw_to_view(dev_x, make_view(scratch, gx, gran), n, cin, h, w, precision, st));
  if (dev_x2) CKS(launch_nchw_to_view(dev_x2, make_view(scratch, gx2, gran), n, cin2, h, w, precision, st));
  if (has_aux) CKS(launch_nchw_to_view(dev_aux, make_view(scratch, ga, gran), n, cout, Ho, Wo, precision, st));
  PackParams pk{};
  pk.seg.w = dev_weight;
  pk.seg.cout = cout, pk.seg.cin = cin + cin2, pk.seg.ksize = ksize;
  pk.seg.split = dev_x2 ? cin : 0x7fffffff;
  pk.dst = scratch + w_off;
  pk.bias_dst = reinterpret_cast<float*>(scratch + b_off);
  pk.Cout_p = cout, pk.K_tap = op.nchunks * 8, pk.fp32 = precision == 0 ? 0 : 1, pk.fold = 0, pk.eps = 0.f;
  CKS(launch_pack(pk, st));
  if (dev_bias) CKS(cudaMemcpyAsync(scratch + b_off, dev_bias, (size_t)cout * 4, cudaMemcpyDeviceToDevice, st));

  ConvParams q{};
  q.H = Ho, q.W = Wo, q.Wp = Wp, q.Pimg = (Ho + 1) * Wp;
  q.body = (long long)n * q.Pimg;
  q.Cout = cout, q.nchunks = op.nchunks, q.nsrc = op.nsrc, q.ntaps = op.ntaps;
  q.halo_lo = op.halo_lo, q.halo_hi = op.halo_hi;
  q.MT = op.MT, q.SP = op.MT * kTileM + op.halo_lo + op.halo_hi;
  q.tps = op.tps, q.NS = op.NS, q.nstages = op.nstages, q.tap_bytes = op.tap_bytes, q.tmem_cols = op.tmem_cols;
  q.epi = epilogue, q.swap_desc = 0, q.ngroups = 1, q.dbg = nullptr, q.ksplit = op.ksplit;
  q.persist_stages = persist_stages_for(op, 0);  // the kernel mtldas_conv2d_nchw picks for the same arguments
  q.magic_pimg = (1ull << 40) / (unsigned long long)q.Pimg + 1, q.magic_wp = (1ull << 40) / (unsigned long long)q.Wp + 1;
  for (int t = 0; t < op.ntaps; ++t) q.tap[t] = op.tap[t];
  ConvGroupDev& d = q.g[0];
  if (stride == 2 && ksize == 3) {
    for (int ph = 0; ph < 4; ++ph) d.src[ph] = make_src(scratch, gx, ph, gran);
  } else if (stride == 2) {
    d.src[0] = make_src(scratch, gx, 0, gran);
  } else {
    d.src[0] = make_src(scratch, gx, -1, gran);
    if (dev_x2) d.src[1] = make_src(scratch, gx2, -1, gran);
  }
  d.w = scratch + w_off;
  d.bias = reinterpret_cast<const float*>(scratch + b_off);
  d.out = make_view(scratch, gy, gran);
  if (has_aux) d.aux = make_view(scratch, ga, gran);
  d.aux_nchunks = cout / 8;
  CKS(launch_conv(q, precision, st));
  CKS(launch_view_to_nchw(make_view(scratch, gy, gran), dev_out, n, cout, Ho, Wo, precision, st));
#undef CKS
  return 0;
}
