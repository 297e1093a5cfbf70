// Stem convolution (Conv2d(1,16,7x7,stride 3,pad 2), /root/reference/model/modelA_MTL.py:73-77) on tcgen05.
//
// Cin = 1.  One GEMM per OUTPUT ROW (n,h):  M = the row's W+1 positions (128-row tile), N = 16 output channels,
// K = 7 input rows x 8-sample windows = 64 (the 8th sample of a window and the 8th row carry zero weights):
//     out[n,h,j][co] = sum_{r<7} sum_{s<8} Y[3h-2+r][j][s] * Wb[r*8+s][co],   Y[y][j][s] = x[n, y, 3j-2+s] (bf16)
// Y is the input row re-windowed into 16-byte granules (2.7x the row's bytes, not the 49x of an im2col).  In the
// no-swizzle K-major operand layout the granule of (row j, K chunk r) sits at  base(y0) + r*LBO + j*16  with
// LBO = the pitch of one windowed input row, so the A operand of output row h is a START ADDRESS into the
// windowed image (input row 3h-2) -- consecutive output rows reuse the same shared-memory data shifted by three
// rows, exactly like the taps of the 3x3 layers.
//
// CTA = 14 warps, persistent over units (image, block of up to kStemRows output rows):
//   warps 0-3  epilogue   TMEM -> bias / ReLU -> bf16 -> two 16-byte stores per position
//   warp  4    MMA issuer four K=16 instructions per output row, rows interleaved over the accumulator tiles
//   warps 5-12 expanders  staged fp32 patch rows -> bf16 windows in shared memory (double buffered)
//   warp  13   producer   the unit's input rows are ONE contiguous span of the patch: a single bulk copy per unit
//                         (from the 16-byte-aligned address below it) into a two-stage fp32 staging ring
// HBM traffic = the patch once + the output once.
#include <algorithm>
#include <cstdio>

#include "gran.cuh"
#include "kernels.h"
#include "ptx.cuh"

namespace mtldas {

constexpr int kStemThreads = 14 * 32;
constexpr int kStemRows = 12;                       // output rows per unit (12 x 16 columns x 2 sets <= 512 TMEM columns)
constexpr int kStemExpanders = 8;                   // warps
constexpr int kStemInRows = 3 * kStemRows + 5;      // input rows a unit touches + one zero row (8th K chunk)
constexpr int kStemBarYFull = 0, kStemBarYEmpty = 2, kStemBarAccFull = 4, kStemBarAccEmpty = 6, kStemBarW = 8,
              kStemBarRawFull = 9, kStemBarRawEmpty = 11;

__device__ __forceinline__ int stem_fdiv(int f, unsigned long long magic) {
  return (int)(((unsigned long long)(unsigned)f * magic) >> 40);
}

__global__ void __launch_bounds__(kStemThreads, 1) stem_umma_kernel(const __grid_constant__ StemParams p, int units_per_image,
                                                                    int nunits, unsigned long long magic_wp,
                                                                    unsigned long long magic_upi) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 96);
  float* bias_s = reinterpret_cast<float*>(smem + 128);
  uint8_t* wsm = smem + 256;                                   // 2 KB banded weight image
  const uint32_t row_pitch = (uint32_t)p.Wp * 16u;             // one windowed input row
  const uint32_t ybuf_bytes = ((uint32_t)kStemInRows * row_pitch + (128u - (uint32_t)p.Wp) * 16u + 127u) & ~127u;
  uint8_t* ybuf = smem + 256 + 2048;                           // [2][kStemInRows][Wp][16 B] (+ slack for rows >= Wp)
  const uint32_t raw_bytes = (uint32_t)p.raw_stage_bytes;
  uint8_t* rawbuf = ybuf + 2 * ybuf_bytes;                     // [2] fp32 staging of a unit's input rows
  // windowed source (rows of an image are x_row_pitch floats apart, e.g. windows of a streaming array): one bulk copy
  // per input row into a staging row of raw_pitch bytes; contiguous patches: ONE bulk copy per unit
  const bool windowed = p.x_row_pitch != (long long)p.Win;
  const uint32_t raw_pitch = ((uint32_t)p.Win * 4u + 16u + 15u) & ~15u;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  ptx::pdl_launch_dependents();
  if (threadIdx.x == 0) {
    ptx::mbar_init(bars + kStemBarYFull + 0, kStemExpanders), ptx::mbar_init(bars + kStemBarYFull + 1, kStemExpanders);
    ptx::mbar_init(bars + kStemBarYEmpty + 0, 1), ptx::mbar_init(bars + kStemBarYEmpty + 1, 1);
    ptx::mbar_init(bars + kStemBarAccFull + 0, 1), ptx::mbar_init(bars + kStemBarAccFull + 1, 1);
    ptx::mbar_init(bars + kStemBarAccEmpty + 0, 4), ptx::mbar_init(bars + kStemBarAccEmpty + 1, 4);
    ptx::mbar_init(bars + kStemBarW, 1);
    ptx::mbar_init(bars + kStemBarRawFull + 0, 1), ptx::mbar_init(bars + kStemBarRawFull + 1, 1);
    ptx::mbar_init(bars + kStemBarRawEmpty + 0, kStemExpanders), ptx::mbar_init(bars + kStemBarRawEmpty + 1, kStemExpanders);
    ptx::fence_mbar_init();
  }
  if (warp == 4) ptx::tmem_alloc(tmem_slot, 512);
  if (threadIdx.x < 16) bias_s[threadIdx.x] = p.bias[threadIdx.x];
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  ptx::pdl_grid_wait();

  if (warp == 4) {
    // ---- MMA issuer ------------------------------------------------------------------------------------------
    if (ptx::elect_one()) {
      ptx::mbar_arrive_expect_tx(bars + kStemBarW, 2048);
      ptx::bulk_g2s(wsm, p.w_umma, 2048, bars + kStemBarW);
      ptx::mbar_wait(bars + kStemBarW, 0);
      const uint32_t idesc = ptx::umma_idesc_bf16(128, 16, 0, 0);
      const uint64_t a_hi = ptx::umma_desc(0, row_pitch, 128u);  // K chunks = input rows, 8-position groups 128 B apart
      const uint64_t b_hi = ptx::umma_desc(0, 256u, 128u);
      const uint32_t y16 = ptx::smem_u32(ybuf) >> 4, w16 = ptx::smem_u32(wsm) >> 4, pitch16 = row_pitch >> 4;
      int it = 0;
      for (int u = blockIdx.x; u < nunits; u += gridDim.x, ++it) {
        const int b = it & 1;
        const int n = stem_fdiv(u, magic_upi), q = u - n * units_per_image;
        const int rows = min(kStemRows, p.H - q * kStemRows);
        ptx::mbar_wait(bars + kStemBarYFull + b, (it >> 1) & 1);
        if (it >= 2) ptx::mbar_wait(bars + kStemBarAccEmpty + b, ((it >> 1) - 1) & 1);
        ptx::tc_fence_after();
        const uint32_t a0 = y16 + (uint32_t)b * (ybuf_bytes >> 4);
        const uint32_t d0 = tmem_base + (uint32_t)b * 256u;
        // rows innermost: consecutive instructions target different accumulator tiles (no accumulate-latency stall)
#pragma unroll
        for (int k = 0; k < 4; ++k)
          for (int i = 0; i < rows; ++i)
            ptx::umma_bf16(d0 + (uint32_t)(i * 16), a_hi | (a0 + (uint32_t)(3 * i + 2 * k) * pitch16),
                           b_hi | (w16 + (uint32_t)k * 32u), idesc, k ? 1u : 0u);
        ptx::umma_commit(bars + kStemBarYEmpty + b);
        ptx::umma_commit(bars + kStemBarAccFull + b);
      }
    }
    __syncwarp();
  } else if (warp == 13) {
    // ---- producer: rows [ya, yb) of image n, copied from the aligned address at or below the first sample -------
    if (ptx::elect_one()) {
      int it = 0;
      for (int u = blockIdx.x; u < nunits; u += gridDim.x, ++it) {
        const int b = it & 1;
        if (it >= 2) ptx::mbar_wait(bars + kStemBarRawEmpty + b, ((it >> 1) - 1) & 1);
        const int n = stem_fdiv(u, magic_upi), q = u - n * units_per_image;
        const int h0 = q * kStemRows, rows = min(kStemRows, p.H - h0);
        const int ya = max(3 * h0 - 2, 0), yb = min(3 * h0 - 2 + 3 * rows + 4, p.Hin);
        const long long gi = (long long)n + p.x_n0;
        const unsigned long long img = (unsigned long long)((gi / p.x_nt) * p.x_stride_a + (gi % p.x_nt) * p.x_stride_b);
        if (!windowed) {
          const unsigned long long lo = (img + (unsigned long long)ya * p.Win) * 4ull;
          const unsigned long long hi = (img + (unsigned long long)yb * p.Win) * 4ull;
          const unsigned long long alo = lo & ~15ull, ahi = (hi + 15ull) & ~15ull;
          const uint32_t bytes = (uint32_t)(ahi - alo);
          ptx::mbar_arrive_expect_tx(bars + kStemBarRawFull + b, bytes);
          ptx::bulk_g2s(rawbuf + (size_t)b * raw_bytes, reinterpret_cast<const char*>(p.x) + alo, bytes,
                        bars + kStemBarRawFull + b);
        } else {
          // every row starts at the same offset inside its 16-byte line (x_row_pitch is a multiple of 4 floats)
          const unsigned long long a0 = (img + (unsigned long long)ya * (unsigned long long)p.x_row_pitch) * 4ull;
          const uint32_t row_bytes = (uint32_t)(((a0 & 15ull) + (unsigned long long)p.Win * 4ull + 15ull) & ~15ull);
          ptx::mbar_arrive_expect_tx(bars + kStemBarRawFull + b, row_bytes * (uint32_t)(yb - ya));
          for (int y = ya; y < yb; ++y) {
            const unsigned long long a = (img + (unsigned long long)y * (unsigned long long)p.x_row_pitch) * 4ull;
            ptx::bulk_g2s(rawbuf + (size_t)b * raw_bytes + (size_t)(y - ya) * raw_pitch,
                          reinterpret_cast<const char*>(p.x) + (a & ~15ull), row_bytes, bars + kStemBarRawFull + b);
          }
        }
      }
    }
    __syncwarp();
  } else if (warp > 4) {
    // ---- expanders: Y[r][j] = bf16(x[n, y0+r, 3j-2 .. 3j+5]); one extra zero row (the 8th K chunk has zero weights,
    // but 0 * garbage must stay 0) -----------------------------------------------------------------------------
    const int te = threadIdx.x - 5 * 32;
    int it = 0;
    for (int u = blockIdx.x; u < nunits; u += gridDim.x, ++it) {
      const int b = it & 1;
      const int n = stem_fdiv(u, magic_upi), q = u - n * units_per_image;
      const int h0 = q * kStemRows, rows = min(kStemRows, p.H - h0);
      const int y0 = 3 * h0 - 2, total = (3 * rows + 5) * p.Wp;
      const int ya = max(y0, 0), yb = min(y0 + 3 * rows + 4, p.Hin);
      const long long gi = (long long)n + p.x_n0;
      const unsigned long long img = (unsigned long long)((gi / p.x_nt) * p.x_stride_a + (gi % p.x_nt) * p.x_stride_b);
      const unsigned long long first = img + (unsigned long long)ya * (unsigned long long)(windowed ? p.x_row_pitch : (long long)p.Win);
      const int mis = (int)((first * 4ull & 15ull) >> 2);
      const int src_pitch = windowed ? (int)(raw_pitch >> 2) : p.Win;   // floats between staged rows
      uint8_t* dst = ybuf + (size_t)b * ybuf_bytes;
      const float* raw = reinterpret_cast<const float*>(rawbuf + (size_t)b * raw_bytes) + mis;
      ptx::mbar_wait(bars + kStemBarRawFull + b, (it >> 1) & 1);
      if (it >= 2) ptx::mbar_wait(bars + kStemBarYEmpty + b, ((it >> 1) - 1) & 1);
      for (int g = te; g < total; g += kStemExpanders * 32) {
        const int r = stem_fdiv(g, magic_wp), j = g - r * p.Wp;
        const int iy = y0 + r, ix0 = 3 * j - 2;
        const bool row_ok = iy >= ya && iy < yb;
        const float* src = raw + (row_ok ? iy - ya : 0) * src_pitch;
        float v[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const int ix = ix0 + i;
          v[i] = (row_ok && ix >= 0 && ix < p.Win) ? src[ix] : 0.f;
        }
        store8<__nv_bfloat16>(reinterpret_cast<char*>(dst) + (size_t)g * 16, v);
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic-proxy writes -> tensor-core reads
      __syncwarp();
      if (lane == 0) {
        ptx::mbar_arrive(bars + kStemBarYFull + b);
        ptx::mbar_arrive(bars + kStemBarRawEmpty + b);
      }
    }
  } else {
    // ---- epilogue: warp w owns TMEM lanes (= positions of the row) [32w, 32w+32) --------------------------------
    const int j = warp * 32 + lane;
    const bool warp_on = warp * 32 < p.Wp, active = j < p.Wp, valid = j < p.W;
    const float lo = p.raw ? -INFINITY : 0.f;
    float bias[16];
#pragma unroll
    for (int c = 0; c < 16; ++c) bias[c] = bias_s[c];
    int it = 0;
    for (int u = blockIdx.x; u < nunits; u += gridDim.x, ++it) {
      const int b = it & 1;
      const int n = stem_fdiv(u, magic_upi), q = u - n * units_per_image;
      const int h0 = q * kStemRows, rows = min(kStemRows, p.H - h0);
      ptx::mbar_wait(bars + kStemBarAccFull + b, (it >> 1) & 1);
      ptx::tc_fence_after();
      if (warp_on) {
        const uint32_t trow = tmem_base + ((uint32_t)(warp * 32) << 16) + (uint32_t)b * 256u;
        for (int i0 = 0; i0 < rows; i0 += 2) {  // two rows of TMEM loads in flight
          uint32_t r0[16], r1[16];
          const bool two = i0 + 1 < rows;
          ptx::tmem_ld16_issue(trow + (uint32_t)(i0 * 16), r0);
          if (two) ptx::tmem_ld16_issue(trow + (uint32_t)((i0 + 1) * 16), r1);
          ptx::tmem_wait16(r0);
          if (two) ptx::tmem_wait16(r1);
          if (active) {
#pragma unroll
            for (int t = 0; t < 2; ++t) {
              if (t == 1 && !two) break;
              const uint32_t* r = t ? r1 : r0;
              float v[16];
#pragma unroll
              for (int c = 0; c < 16; ++c) v[c] = valid ? fmaxf(__uint_as_float(r[c]) + bias[c], lo) : 0.f;
              const int h = h0 + i0 + t;
              const int f = n * p.Pimg + h * p.Wp + j;
              store8<__nv_bfloat16>(view_addr<__nv_bfloat16>(p.out, 0, f, n, h, j), v);
              store8<__nv_bfloat16>(view_addr<__nv_bfloat16>(p.out, 1, f, n, h, j), v + 8);
            }
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(bars + kStemBarAccEmpty + b);
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 4) ptx::tmem_dealloc(tmem_base, 512);
}

cudaError_t launch_stem_umma(const StemParams& p_in, cudaStream_t stream, bool pdl) {
  StemParams p = p_in;
  if (p.x_nt == 0) p.x_nt = 1, p.x_n0 = 0, p.x_stride_a = (long long)p.Hin * p.Win, p.x_stride_b = 0, p.x_row_pitch = p.Win;
  const bool windowed = p.x_row_pitch != (long long)p.Win;
  const size_t row_pitch = (size_t)p.Wp * 16;
  const size_t ybuf = (((size_t)(3 * kStemRows + 5) * row_pitch + (size_t)(128 - p.Wp) * 16) + 127) & ~(size_t)127;
  const size_t raw_contig = (size_t)(3 * kStemRows + 4) * p.Win * 4 + 32;
  const size_t raw_rows = (size_t)(3 * kStemRows + 4) * ((((size_t)p.Win * 4 + 16 + 15) & ~(size_t)15));
  const size_t raw = ((windowed ? raw_rows : raw_contig) + 127) & ~(size_t)127;
  p.raw_stage_bytes = (int)raw;
  const size_t smem = 256 + 2048 + 2 * ybuf + 2 * raw;
  // bulk copies need 16-byte aligned spans: the source must start aligned; contiguous patches: an image must be a
  // multiple of 16 B; windowed rows: the row pitch and the image strides must be multiples of 4 floats... except the
  // time-block stride (Win = 250 floats): blocks alternate between offsets 0 and 8 inside a 16-byte line, handled
  // per image by the kernel
  if (p.Wp > 128 || smem > 225 * 1024 || (reinterpret_cast<uintptr_t>(p.x) & 15)) return cudaErrorInvalidValue;
  if (!windowed && ((size_t)p.Hin * p.Win * 4) % 16) return cudaErrorInvalidValue;
  if (windowed && ((p.x_row_pitch & 3) || (p.x_stride_a & 1) || (p.x_stride_b & 1))) return cudaErrorInvalidValue;
  static PerDeviceSmem configured;
  {
    cudaError_t e = configured.ensure(smem, [](int bytes) {
      return cudaFuncSetAttribute(stem_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
    });
    if (e != cudaSuccess) return e;
  }
  const int nb = (int)(p.body / p.Pimg);
  const int upi = (p.H + kStemRows - 1) / kStemRows, nunits = nb * upi;
  const unsigned long long magic_wp = (1ull << 40) / (unsigned long long)p.Wp + 1;
  const unsigned long long magic_upi = (1ull << 40) / (unsigned long long)upi + 1;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3((unsigned)std::min(nunits, 148)), cfg.blockDim = dim3(kStemThreads);
  cfg.dynamicSmemBytes = smem, cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr, cfg.numAttrs = pdl ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, stem_umma_kernel, p, upi, nunits, magic_wp, magic_upi);
}

}  // namespace mtldas
