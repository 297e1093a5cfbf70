// Granule (8-channel) load/store helpers and view addressing shared by all kernels.
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include "types.h"

namespace mtldas {

// ---------------------------------------------------------------------------------------------
// granule helpers
// ---------------------------------------------------------------------------------------------
template <typename T>
struct Gran;
template <>
struct Gran<__nv_bfloat16> {
  static constexpr int bytes = 16;
};
template <>
struct Gran<float> {
  static constexpr int bytes = 32;
};

template <typename T>
__device__ __forceinline__ void load8(const char* p, float* v);
template <>
__device__ __forceinline__ void load8<__nv_bfloat16>(const char* p, float* v) {
  uint4 q = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 f = __bfloat1622float2(h[i]);
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
}
template <>
__device__ __forceinline__ void load8<float>(const char* p, float* v) {
  float4 a = reinterpret_cast<const float4*>(p)[0];
  float4 b = reinterpret_cast<const float4*>(p)[1];
  v[0] = a.x, v[1] = a.y, v[2] = a.z, v[3] = a.w, v[4] = b.x, v[5] = b.y, v[6] = b.z, v[7] = b.w;
}
template <typename T>
__device__ __forceinline__ void store8(char* p, const float* v);
template <>
__device__ __forceinline__ void store8<__nv_bfloat16>(char* p, const float* v) {
  uint4 q;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&q);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
  *reinterpret_cast<uint4*>(p) = q;
}
template <>
__device__ __forceinline__ void store8<float>(char* p, const float* v) {
  reinterpret_cast<float4*>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
  reinterpret_cast<float4*>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
}

// raw granules: loads are issued for a batch of positions first and unpacked afterwards, so one thread keeps
// several independent 16-byte requests in flight (the streaming kernels are latency-bound otherwise)
template <typename T>
struct Raw;
template <>
struct Raw<__nv_bfloat16> {
  uint4 q;
};
template <>
struct Raw<float> {
  float4 a, b;
};
__device__ __forceinline__ void load_raw(const char* p, Raw<__nv_bfloat16>& r) { r.q = *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ void load_raw(const char* p, Raw<float>& r) {
  r.a = reinterpret_cast<const float4*>(p)[0];
  r.b = reinterpret_cast<const float4*>(p)[1];
}
__device__ __forceinline__ void unpack(const Raw<__nv_bfloat16>& r, float* v) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r.q);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float2 f = __bfloat1622float2(h[i]);
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
}
__device__ __forceinline__ void unpack(const Raw<float>& r, float* v) {
  v[0] = r.a.x, v[1] = r.a.y, v[2] = r.a.z, v[3] = r.a.w, v[4] = r.b.x, v[5] = r.b.y, v[6] = r.b.z, v[7] = r.b.w;
}
template <typename T>
struct Batch {  // positions per thread and loop iteration
  static constexpr int n = sizeof(T) == 2 ? 4 : 2;
};

// byte address of granule (chunk c, output-space coordinates) inside a view
template <typename T>
__device__ __forceinline__ char* view_addr(const ActView& a, int c, int f, int n, int h, int w) {
  long long pos;
  if (a.layout == LAY_P8) {
    pos = (long long)c * a.plane + a.g_lo + f;
  } else {
    int phase = ((h & 1) << 1) | (w & 1);
    pos = (long long)phase * a.phase_stride + (long long)c * a.plane + a.g_lo + (long long)n * a.Pimg2 +
          (h >> 1) * a.Wp2 + (w >> 1);
  }
  return a.base + pos * Gran<T>::bytes;
}

}  // namespace mtldas
