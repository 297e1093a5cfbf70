// Model C (Multi_Classifier = Inception-V3 with a 1-channel stem and 32 classes), eval-mode forward, for sm_100a.
//
// Reference: /root/reference/model/modelC_multiClassifier.py:63-86 (layers), :102-147 (_forward) and torchvision's
// InceptionA..E blocks (torchvision/models/inception.py:176-365) that it instantiates: 94 conv + BatchNorm(eps 1e-3) +
// ReLU units with kernels 1x1, 3x3 (pad 0 / 1, stride 1 / 2), 5x5 p2, 1x7, 7x1, 1x3, 3x1; F.avg_pool2d(3,1,1) (count
// includes the padding), F.max_pool2d(3,2) / nn.MaxPool2d(3,2); channel concatenation; AdaptiveAvgPool2d(1) + Dropout
// (identity in eval) + Linear(2048, 32).
//
// Layout "planar-8": a tensor [n, C, H, W] is C/8 planes of n*H*W positions, one position of one plane = 8 consecutive
// channels as bf16 (16 bytes).  A channel concat is a chunk offset into the destination tensor.
//
// Two tcgen05 implicit-GEMM convolution kernels share one weight layout (weight_layout() below):
//   convc_window_kernel  stride-1 layers: the window of input positions a tile needs is staged once in shared memory and
//                        every tap is a shifted operand descriptor (see the comment in front of it);
//   convc_umma_kernel    stride-2 layers and layers whose padded enumeration would waste > 45 %: one CTA = 128 output
//                        positions x Nt <= 256 output channels; the K loop runs over (tap, 32-channel unit); four gather
//                        warps build the A tile of a unit with 16-byte cp.async copies (zero fill outside the image: any
//                        kernel size / stride / padding is index arithmetic), one producer thread brings the unit's
//                        weights with bulk copies, one thread issues the UMMAs into a TMEM accumulator; a ring of stages
//                        overlaps the three; the gather warps then run the epilogue.
// Epilogue of both: folded-BN bias + ReLU, bf16, 16-byte stores at the concat offset.  No im2col tensor, no tensor maps.
// This model is compute bound (1.1 GMAC per patch, Cout up to 448).
#include <cuda_bf16.h>
#include <cuda_runtime.h>

#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <cstdlib>

#include "../../include/mtldas.h"
#include "kernels.h"
#include "plan.h"
#include "ptx.cuh"

namespace mtldas {
namespace {

constexpr int kCStages = 4;       // ring depth
constexpr int kCUnitSteps = 2;    // K steps (16 channels each) per ring stage
constexpr int kCThreads = 192;    // warps 0-3 gather + epilogue, warp 4 MMA issue, warp 5 weight producer

struct ConvCParams {
  const char* x;        // planar-8 bf16 input, cin/8 planes of m_in positions
  const char* w;        // packed weights (BatchNorm folded), layout of weight_layout(): [n tile][K unit][tap][chunk][Nt][8] bf16
  const float* bias;    // [cout] folded BatchNorm shift
  char* out;            // planar-8 bf16 destination tensor (out_planes planes of m_out positions)
  long long m_in, m_out;
  int n, cin, H, W, Ho, Wo, cout, kh, kw, sh, sw, ph, pw;
  int out_chunk0;       // first destination plane (channel offset / 8 inside the concat tensor)
  int Nt;               // output channels per CTA (multiple of 16, <= 256)
  int relu;
  int ks;               // K steps per K unit of the weight layout (weight_layout())
};

__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void cp_async_arrive_noinc(uint64_t* bar) {   // this thread's arrival, deferred until its copies landed
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(ptx::smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__global__ void __launch_bounds__(kCThreads, 2) convc_umma_kernel(const __grid_constant__ ConvCParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* full = reinterpret_cast<uint64_t*>(smem);              // [kCStages]
  uint64_t* empty = full + kCStages;                                // [kCStages]
  uint64_t* acc_full = empty + kCStages;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 128);
  float* bias_s = reinterpret_cast<float*>(smem + 256);             // Nt floats (<= 1 KB -> header is 256 + 1024)
  const int Nt = p.Nt;
  const uint32_t a_stage = 2u * kCUnitSteps * 128u * 16u;           // 8 KB: 2*steps chunks x 128 rows x 16 B
  const uint32_t b_stage = 2u * kCUnitSteps * (uint32_t)Nt * 16u;
  uint8_t* a_base = smem + 256 + 1024;
  uint8_t* b_base = a_base + (size_t)kCStages * a_stage;
  const int warp = threadIdx.x >> 5;
  const int n0 = blockIdx.y * Nt;
  const int KC = p.cin >> 4;                                        // K steps per tap
  const int units_per_tap = (KC + kCUnitSteps - 1) / kCUnitSteps;
  const int ntaps = p.kh * p.kw;
  const int nunits = ntaps * units_per_tap;
  uint32_t cols = 32;
  while (cols < (uint32_t)Nt) cols <<= 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kCStages; ++s) {
      ptx::mbar_init(full + s, 128 + 1);
      ptx::mbar_init(empty + s, 1);
    }
    ptx::mbar_init(acc_full, 1);
    ptx::fence_mbar_init();
  }
  if (warp == 4) ptx::tmem_alloc(tmem_slot, cols);
  for (int i = threadIdx.x; i < Nt; i += blockDim.x) bias_s[i] = p.bias[n0 + i];
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp < 4) {
    // ---------------- gather: thread = one output position (one row of the A tile) ----------------
    const long long m = (long long)blockIdx.x * 128 + threadIdx.x;
    const bool row_ok = m < p.m_out;
    int n = 0, ho = 0, wo = 0;
    if (row_ok) {
      const int hw = p.Ho * p.Wo;
      n = (int)(m / hw);
      const int rem = (int)(m - (long long)n * hw);
      ho = rem / p.Wo;
      wo = rem - ho * p.Wo;
    }
    const uint32_t a_row = ptx::smem_u32(a_base) + (uint32_t)threadIdx.x * 16u;
    int u = 0;
    for (int tap = 0; tap < ntaps; ++tap) {
      const int r = tap / p.kw, q = tap - r * p.kw;
      const int hi = ho * p.sh + r - p.ph, wi = wo * p.sw + q - p.pw;
      const bool inb = row_ok && hi >= 0 && hi < p.H && wi >= 0 && wi < p.W;
      const long long src_pos = inb ? ((long long)n * p.H + hi) * p.W + wi : 0;
      const char* src0 = p.x + src_pos * 16;
      const uint32_t nbytes = inb ? 16u : 0u;
      for (int kc0 = 0; kc0 < KC; kc0 += kCUnitSteps, ++u) {
        const int s = u % kCStages;
        if (u >= kCStages) ptx::mbar_wait(empty + s, ((u / kCStages) - 1) & 1);
        const int nch = 2 * min(kCUnitSteps, KC - kc0);
        const uint32_t dst = a_row + (uint32_t)s * a_stage;
#pragma unroll
        for (int c = 0; c < 2 * kCUnitSteps; ++c)
          if (c < nch) cp_async16(dst + (uint32_t)c * 2048u, src0 + (long long)(kc0 * 2 + c) * p.m_in * 16, nbytes);
        cp_async_commit();
        if (u >= 2) {                       // unit u-2 has landed: publish it to the MMA thread
          cp_async_wait<2>();
          fence_proxy_async();
          ptx::mbar_arrive(full + (u - 2) % kCStages);
        }
      }
    }
    if (nunits >= 2) {
      cp_async_wait<1>();
      fence_proxy_async();
      ptx::mbar_arrive(full + (nunits - 2) % kCStages);
    }
    cp_async_wait<0>();
    fence_proxy_async();
    ptx::mbar_arrive(full + (nunits - 1) % kCStages);

    // ---------------- epilogue: bias + ReLU, bf16, one 16-byte granule per 8 channels ----------------
    ptx::mbar_wait(acc_full, 0);
    ptx::tc_fence_after();
    const uint32_t trow = tmem_base + ((uint32_t)(warp * 32) << 16);
    char* out_row = p.out + ((long long)(p.out_chunk0 + (n0 >> 3)) * p.m_out + m) * 16;
    for (int c0 = 0; c0 < Nt; c0 += 16) {
      uint32_t r[16];
      ptx::tmem_ld16_issue(trow + (uint32_t)c0, r);
      ptx::tmem_wait16(r);
      if (row_ok) {
        float v[16];
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          v[j] = __uint_as_float(r[j]) + bias_s[c0 + j];
          if (p.relu) v[j] = fmaxf(v[j], 0.f);
        }
        uint4 q0, q1;
        __nv_bfloat162* h0 = reinterpret_cast<__nv_bfloat162*>(&q0);
        __nv_bfloat162* h1 = reinterpret_cast<__nv_bfloat162*>(&q1);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          h0[j] = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
          h1[j] = __floats2bfloat162_rn(v[8 + 2 * j], v[8 + 2 * j + 1]);
        }
        *reinterpret_cast<uint4*>(out_row + (long long)(c0 >> 3) * p.m_out * 16) = q0;
        *reinterpret_cast<uint4*>(out_row + (long long)((c0 >> 3) + 1) * p.m_out * 16) = q1;
      }
    }
  } else if (warp == 4) {
    // ---------------- MMA issue ----------------
    if (ptx::elect_one()) {
      const uint32_t idesc = ptx::umma_idesc_bf16(128, Nt, 0, 0);
      const uint64_t a_hi = ptx::umma_desc(0, 2048u, 128u);
      const uint64_t b_hi = ptx::umma_desc(0, (uint32_t)Nt * 16u, 128u);
      const uint32_t a16 = ptx::smem_u32(a_base) >> 4, b16 = ptx::smem_u32(b_base) >> 4;
      uint32_t acc = 0u;
      int u = 0;
      for (int tap = 0; tap < ntaps; ++tap)
        for (int kc0 = 0; kc0 < KC; kc0 += kCUnitSteps, ++u) {
          const int s = u % kCStages;
          ptx::mbar_wait(full + s, (u / kCStages) & 1);
          ptx::tc_fence_after();
          const int nk = min(kCUnitSteps, KC - kc0);
          for (int j = 0; j < nk; ++j) {
            const uint32_t a_lo = a16 + (uint32_t)s * (a_stage >> 4) + (uint32_t)j * (2u * 2048u >> 4);
            const uint32_t b_lo = b16 + (uint32_t)s * (b_stage >> 4) + (uint32_t)j * (2u * (uint32_t)Nt);
            ptx::umma_bf16_lean(tmem_base, a_lo | (uint32_t)a_hi, (uint32_t)(a_hi >> 32), b_lo | (uint32_t)b_hi,
                                (uint32_t)(b_hi >> 32), idesc, acc);
            acc = 1u;
          }
          ptx::umma_commit(empty + s);
        }
      ptx::umma_commit(acc_full);
    }
    __syncwarp();
  } else {
    // ---------------- weight producer: the unit's [2*steps chunks][Nt][8] slice, one bulk copy per chunk ----------------
    if (ptx::elect_one()) {
      const int cin_chunks = p.cin >> 3;
      int u = 0;
      for (int tap = 0; tap < ntaps; ++tap)
        for (int kc0 = 0; kc0 < KC; kc0 += kCUnitSteps, ++u) {
          const int s = u % kCStages;
          if (u >= kCStages) ptx::mbar_wait(empty + s, ((u / kCStages) - 1) & 1);
          const int nch = 2 * min(kCUnitSteps, KC - kc0);
          ptx::mbar_arrive_expect_tx(full + s, (uint32_t)(nch * Nt * 16));
          for (int c = 0; c < nch;) {                     // layout [n tile][K unit][tap][chunk][Nt][8]: the chunks of
            const int gc = kc0 * 2 + c, ku = gc / (2 * p.ks), cl = gc - ku * 2 * p.ks;   // one K unit are contiguous
            const int nchu = min(2 * p.ks, cin_chunks - ku * 2 * p.ks);
            const int run = min(nch - c, nchu - cl);      // one bulk copy per run (a copy costs ~70 cycles of issue)
            ptx::bulk_g2s(b_base + (size_t)s * b_stage + (size_t)c * Nt * 16,
                          p.w + (((long long)blockIdx.y * ntaps * cin_chunks + (long long)ku * ntaps * 2 * p.ks + tap * nchu + cl) * Nt) * 16,
                          (uint32_t)(run * Nt * 16), full + s);
            c += run;
          }
        }
    }
    __syncwarp();
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 4) ptx::tmem_dealloc(tmem_base, cols);
}

// ---------------------------------------------------------------------------------------------
// convc_window_kernel: the stride-1 convolutions (84 of the 94).  Positions are enumerated in a PADDED image whose
// padding is shared between neighbours: Wp = W + pw positions per row (the right padding of a row is the left padding
// of the next one), Hp = H + ph rows per image (the bottom padding of an image is the top padding of the next), pixel
// (0, 0) of image 0 at position off = ph * Wp + pw.  Output (img, ho, wo) is position u = (img * Hp + ho) * Wp + wo and
// reads, for tap (r, q), position u + r * Wp + q.  A work item is MT x 128 consecutive positions; ONE window of
// MT * 128 + (kh-1) * Wp + (kw-1) positions per K unit is brought into shared memory (bulk copies of whole row
// segments; the zero padding is written there, not in HBM) and every tap is a UMMA whose A descriptor start address is
// shifted: the input is read from L2 about once instead of kh * kw times, by one warp issuing a few bulk copies instead
// of 128 threads issuing 16-byte copies per tap.  Positions that are no valid output (margins, padding rows) are
// computed and dropped: layers where they are more than ~45 % extra go to the gather kernel instead.
// Persistent: the CTAs walk over the (m group, n tile) work items; the accumulators live in TMEM (MT per set, two sets
// when they fit: four epilogue warps store item t while the MMA warp is on item t + 1).
// ---------------------------------------------------------------------------------------------
struct ConvWParams {
  const char* x;        // planar-8 bf16 input (unpadded), cin/8 planes of m_in positions
  const char* w;        // packed weights, layout of weight_layout(): [n tile][K unit][tap][chunk][Nt][8] bf16
  const float* bias;
  char* out;            // planar-8 bf16 destination (unpadded)
  long long m_in, m_out;
  int pp_total;         // n * Hp * Wp padded positions
  int n, cin, H, W, Ho, Wo, cout;
  int Hp, Wp;           // H + ph rows per image, W + pw positions per row (see the kernel comment)
  int kh, kw;           // taps used: rows [tap_r0, tap_r0 + kh) of the kernel, all kw columns
  int ph, pw;           // padding seen by those taps
  int off;              // ph * Wp + pw: position of pixel (0, 0) of image 0 in the padded enumeration
  int cells;            // padded window built cell by cell (cp.async) instead of row by row (bulk copies): narrow images
  int tap_r0;           // H == 1: the kernel rows above / below the centre only ever meet padding and are skipped
  int ntaps_full;       // taps of the whole kernel (the weight layout's)
  int out_chunk0, Nt, relu;
  int m_tiles, n_tiles;
  int mt;               // m tiles (128 positions each) per work item: 1, 2 or 4 independent accumulators
  int sets;             // accumulator sets in TMEM: 2 = the epilogue of an item overlaps the UMMAs of the next one
  int win;              // positions per window
  int a_stages, b_stages;
  int ks;               // K steps (16 channels) per window stage / weight slice
  int tps;              // taps per weight slice (one slice = one B stage)
  long long* trace;     // MTLDAS_C_TRACE: clock64 stamps of CTA 0, [role 0..3][event < 64][4]
};

constexpr int kWThreads = 256;                  // warp 0 windows, warp 1 weights, warp 2 MMA issue, warps 4-7 epilogue
constexpr int kWMaxA = 8, kWMaxB = 12;
constexpr uint32_t kWHeader = 512 + 2048;       // mbarriers + TMEM slot, then the layer's bias (<= 512 floats)
constexpr uint32_t kWSmemBudget = 225 * 1024;   // one CTA per SM (<= 512 TMEM columns); two CTAs per SM get half each (<= 256)

// Waits of the producer / epilogue warps: they back off between polls so that the MMA-issue thread (one thread, bound by
// its own instruction latency) keeps the issue slots and the mbarrier unit.  Bounded like ptx::mbar_wait.
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t* bar, uint32_t parity) {
  if (ptx::mbar_try_wait(bar, parity)) return;
  const long long t0 = clock64();
  uint32_t spins = 0;
  while (!ptx::mbar_try_wait(bar, parity)) {
    __nanosleep(32);
    if ((++spins & 255u) == 0 && clock64() - t0 > 4000000000ll) {
      printf("mtldas: mbarrier wait timed out (block %d thread %d)\n", blockIdx.x, threadIdx.x);
      __trap();
    }
  }
}

__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.expect_tx.relaxed.cta.shared::cta.b64 [%0], %1;" ::"r"(ptx::smem_u32(bar)), "r"(bytes) : "memory");
}

// UMMA / commit executed by the whole warp on warp-uniform operands; one elected lane issues (always the same one: the
// commit tracks the UMMAs of the thread that issued them)
__device__ __forceinline__ void umma_bf16_elect(uint32_t tmem_d, uint32_t a_lo, uint32_t a_hi, uint32_t b_lo, uint32_t b_hi,
                                                uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, e;\n\t.reg .b64 da, db;\n\t"
      "setp.ne.b32 p, %6, 0;\n\t"
      "mov.b64 da, {%1, %2};\n\t"
      "mov.b64 db, {%3, %4};\n\t"
      "elect.sync _|e, 0xffffffff;\n\t"
      "@e tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %5, p;\n\t}"
      ::"r"(tmem_d), "r"(a_lo), "r"(a_hi), "r"(b_lo), "r"(b_hi), "r"(idesc), "r"(accumulate));
}
__device__ __forceinline__ void umma_commit_elect(uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred e;\n\t"
      "elect.sync _|e, 0xffffffff;\n\t"
      "@e tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}"
      ::"r"(ptx::smem_u32(bar)) : "memory");
}

// eight accumulator columns + bias, lower clamp (0 = ReLU), rounded to one 16-byte granule of bf16
__device__ __forceinline__ uint4 bias_act_pack8(const uint32_t* r, const float* bias, float lo) {
  const float4 b0 = *reinterpret_cast<const float4*>(bias), b1 = *reinterpret_cast<const float4*>(bias + 4);
  uint4 o;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&o);
  h[0] = __floats2bfloat162_rn(fmaxf(__uint_as_float(r[0]) + b0.x, lo), fmaxf(__uint_as_float(r[1]) + b0.y, lo));
  h[1] = __floats2bfloat162_rn(fmaxf(__uint_as_float(r[2]) + b0.z, lo), fmaxf(__uint_as_float(r[3]) + b0.w, lo));
  h[2] = __floats2bfloat162_rn(fmaxf(__uint_as_float(r[4]) + b1.x, lo), fmaxf(__uint_as_float(r[5]) + b1.y, lo));
  h[3] = __floats2bfloat162_rn(fmaxf(__uint_as_float(r[6]) + b1.z, lo), fmaxf(__uint_as_float(r[7]) + b1.w, lo));
  return o;
}

#define WTRACE(role, ev, slot) \
  do { if (p.trace && blockIdx.x == 0 && (ev) < 64) p.trace[((role) * 64 + (ev)) * 4 + (slot)] = clock64(); } while (0)

__global__ void __launch_bounds__(kWThreads, 2) convc_window_kernel(const __grid_constant__ ConvWParams p) {
  extern __shared__ __align__(128) uint8_t smem[];
  uint64_t* a_full = reinterpret_cast<uint64_t*>(smem);            // [kWMaxA] one arrival + the window's bytes
  uint64_t* a_empty = a_full + kWMaxA;                              // [kWMaxA] one tcgen05.commit
  uint64_t* b_full = a_empty + kWMaxA;                              // [kWMaxB] one arrival + the slice's bytes
  uint64_t* b_empty = b_full + kWMaxB;                              // [kWMaxB] one tcgen05.commit
  uint64_t* acc_full = b_empty + kWMaxB;                            // [2] one tcgen05.commit per tile
  uint64_t* acc_empty = acc_full + 2;                               // [2] four epilogue warps
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(smem + 384);
  float* bias_s = reinterpret_cast<float*>(smem + 512);             // [cout]
  const int Nt = p.Nt, SA = p.a_stages, SB = p.b_stages, win = p.win, ks = p.ks, MT = p.mt, SETS = p.sets;
  const uint32_t a_stage = 2u * (uint32_t)ks * (uint32_t)win * 16u;  // 2 ks chunk planes of `win` positions
  const uint32_t b_stage = (uint32_t)p.tps * 2u * (uint32_t)ks * (uint32_t)Nt * 16u;
  uint8_t* a_base = smem + kWHeader;
  uint8_t* b_base = a_base + (size_t)SA * a_stage;
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);      // provably warp-uniform: role code stays in
  const int lane = threadIdx.x & 31;                                          // uniform registers
  const int KC = p.cin >> 4;                                        // K steps (16 channels)
  const int KU = (KC + ks - 1) / ks;                                // units of ks K steps
  const int ntaps = p.kh * p.kw;
  const int nslices = (ntaps + p.tps - 1) / p.tps;
  const int total = p.m_tiles * p.n_tiles;
  const uint32_t set_cols = ((uint32_t)Nt + 31u) & ~31u;            // TMEM columns of one accumulator set
  uint32_t cols = 32;
  while (cols < (uint32_t)(SETS * MT) * set_cols) cols <<= 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < SA; ++s) ptx::mbar_init(a_full + s, p.cells ? 33 : 1), ptx::mbar_init(a_empty + s, 1);
    for (int s = 0; s < SB; ++s) ptx::mbar_init(b_full + s, 1), ptx::mbar_init(b_empty + s, 1);
    for (int s = 0; s < 2; ++s) ptx::mbar_init(acc_full + s, 1), ptx::mbar_init(acc_empty + s, 4);
    ptx::fence_mbar_init();
  }
  if (warp == 2) ptx::tmem_alloc(tmem_slot, cols);
  for (int i = threadIdx.x; i < p.cout; i += blockDim.x) bias_s[i] = p.bias[i];
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_slot, 0);

  if (warp == 0) {
    // ---------------- window producer: the whole warp; one stage = one tile's window of one 32-channel unit ----------------
    const bool padded = (p.ph | p.pw) != 0;
    int a = 0, round = 0, ev = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x) {
      const int pp0 = (w / p.n_tiles) * 128 * MT;
      for (int ku = 0; ku < KU; ++ku) {
        if (lane == 0) WTRACE(0, ev, 0);
        if (round) mbar_wait_relaxed(a_empty + a, (uint32_t)(round - 1) & 1u);
        if (lane == 0) WTRACE(0, ev, 1);
        uint8_t* stage = a_base + (size_t)a * a_stage;
        const int nch = 2 * min(ks, KC - ku * ks);
        const char* xk = p.x + (long long)(ku * 2 * ks) * p.m_in * 16;
        if (!padded) {
          // padded position == input position: one copy per chunk plane, across rows and images
          const int npos = min(win, p.pp_total - pp0);
          if (lane < nch) {
            mbar_expect_tx(a_full + a, (uint32_t)npos * 16u);
            ptx::bulk_g2s(stage + (size_t)lane * win * 16, xk + ((long long)lane * p.m_in + pp0) * 16, (uint32_t)npos * 16u,
                          a_full + a);
          }
        } else if (p.cells) {
          // narrow images: a row segment is a few cells, and a bulk copy costs ~70 cycles of issue whatever its size.
          // Cell by cell instead: 16-byte cp.async with zero fill for the padding, one position (all chunk planes) per
          // lane and step; each lane's arrival on the stage's barrier is deferred until its copies have landed.
          const int s0 = pp0 - p.off + p.Hp * p.Wp;
          for (int pos = lane; pos < win; pos += 32) {
            const int v = s0 + pos, row = v / p.Wp, wp = v - row * p.Wp;
            const int img = row / p.Hp - 1, hp = row - (img + 1) * p.Hp;
            const bool ok = img >= 0 && img < p.n && hp < p.H && wp < p.W;
            const char* src = ok ? xk + (((long long)img * p.H + hp) * p.W + wp) * 16 : p.x;
            const long long plane = ok ? p.m_in * 16 : 0;
            const uint32_t dst = ptx::smem_u32(stage) + (uint32_t)pos * 16u;
            for (int c = 0; c < nch; ++c) cp_async16(dst + (uint32_t)(c * win) * 16u, src + c * plane, ok ? 16u : 0u);
          }
          cp_async_arrive_noinc(a_full + a);
        } else {
          // one task per (row of the window, chunk plane): bulk copy of the row's pixels, zeros in its padding.  Rows are
          // counted from one image period before the first image (the window may start in the top padding of image 0).
          const uint4 z = make_uint4(0u, 0u, 0u, 0u);
          const int s0 = pp0 - p.off + p.Hp * p.Wp;
          const int row0 = s0 / p.Wp, row1 = min((s0 + win - 1) / p.Wp, (p.n + 1) * p.Hp - 1);
          const int ncopies = (row1 - row0 + 1) * nch;
          for (int i = lane; i < ncopies; i += 32) {
            const int row = row0 + i / nch, c = i - (i / nch) * nch;
            const int img = row / p.Hp - 1, hp = row - (img + 1) * p.Hp;
            const int rel = s0 - row * p.Wp;                          // window start relative to this row
            const int w_lo = max(0, rel), w_hi = min(p.Wp, rel + win);  // the row's cells inside the window
            uint4* cells = reinterpret_cast<uint4*>(stage + ((long long)c * win - rel) * 16);   // cells[wp]
            if (img < 0 || hp >= p.H) {                               // a padding row (shared by two images)
              for (int x = w_lo; x < w_hi; ++x) cells[x] = z;
              continue;
            }
            for (int x = max(p.W, w_lo); x < w_hi; ++x) cells[x] = z;   // right padding = left padding of the next row
            const int hi = min(p.W, w_hi);
            if (w_lo >= hi) continue;
            const uint32_t bytes = (uint32_t)(hi - w_lo) * 16u;
            mbar_expect_tx(a_full + a, bytes);
            ptx::bulk_g2s(stage + ((size_t)c * win + (w_lo - rel)) * 16,
                          xk + ((long long)c * p.m_in + ((long long)img * p.H + hp) * p.W + w_lo) * 16, bytes, a_full + a);
          }
          fence_proxy_async();               // the zeros (generic proxy) are read by the UMMAs (async proxy)
        }
        __syncwarp();
        if (lane == 0) ptx::mbar_arrive(a_full + a);
        if (lane == 0) WTRACE(0, ev, 2);
        ++ev;
        if (++a == SA) a = 0, ++round;
      }
    }
  } else if (warp == 1) {
    // ---------------- weight producer: one stage = `tps` taps of one K unit for this tile's output channels: contiguous in
    // the packed layout [n tile][K unit][tap][chunk][Nt][8], one bulk copy ----------------
    const int cin_chunks = p.cin >> 3;
    int b = 0, round = 0, ev = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x) {
      const char* wn = p.w + (long long)(w % p.n_tiles) * p.ntaps_full * cin_chunks * Nt * 16;
      for (int ku = 0; ku < KU; ++ku) {
        const int nch = 2 * min(ks, KC - ku * ks);
        const char* wu = wn + ((long long)ku * p.ntaps_full * 2 * ks + (long long)p.tap_r0 * p.kw * nch) * Nt * 16;
        for (int sl = 0; sl < nslices; ++sl) {
          const int t0 = sl * p.tps, nt = min(p.tps, ntaps - t0);
          if (lane == 0) WTRACE(1, ev, 0);
          if (round) mbar_wait_relaxed(b_empty + b, (uint32_t)(round - 1) & 1u);
          if (lane == 0) {
            WTRACE(1, ev, 1);
            const uint32_t bytes = (uint32_t)(nt * nch * Nt * 16);
            ptx::mbar_arrive_expect_tx(b_full + b, bytes);
            ptx::bulk_g2s(b_base + (size_t)b * b_stage, wu + (long long)t0 * nch * Nt * 16, bytes, b_full + b);
            WTRACE(1, ev, 2);
          }
          ++ev;
          if (++b == SB) b = 0, ++round;
        }
      }
    }
  }
  else if (warp == 2) {
    // ---------------- MMA issue: tile t accumulates into TMEM set t & 1; a tap is a shifted A descriptor.  All 32 lanes
    // run this code on warp-uniform values only (kernel parameters and loop counters: the descriptors then live in
    // uniform registers, which is what UTCHMMA reads) and one elected lane issues; consecutive UMMAs go to the MT
    // independent accumulators of the work item ----------------
    const uint32_t idesc = ptx::umma_idesc_bf16(128, Nt, 0, 0);
    const uint64_t a_hi = ptx::umma_desc(0, (uint32_t)win * 16u, 128u);        // LBO: next chunk plane; SBO: next 8 rows
    const uint64_t b_hi = ptx::umma_desc(0, (uint32_t)Nt * 16u, 128u);
    const uint32_t a_hl = (uint32_t)a_hi, a_hh = (uint32_t)(a_hi >> 32), b_hl = (uint32_t)b_hi, b_hh = (uint32_t)(b_hi >> 32);
    const uint32_t a16 = ptx::smem_u32(a_base) >> 4, b16 = ptx::smem_u32(b_base) >> 4;
    int a = 0, around = 0, b = 0, bround = 0, t = 0, ev = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x, ++t) {
      // one accumulator set: the epilogue of item t - 1 must have drained it (a second CTA on the SM fills the gap);
      // two sets: item t uses set t & 1, drained by the epilogue of item t - 2
      const int set = SETS == 2 ? (t & 1) : 0, use = SETS == 2 ? (t >> 1) : t;
      if (use >= 1) {
        ptx::mbar_wait(acc_empty + set, (uint32_t)(use - 1) & 1u);
        ptx::tc_fence_after();
      }
      const uint32_t d = tmem_base + (uint32_t)(set * MT) * set_cols;
      uint32_t acc = 0u;
      for (int ku = 0; ku < KU; ++ku) {
        const int nk = min(ks, KC - ku * ks);
        if (lane == 0) WTRACE(2, ev, 0);
        ptx::mbar_wait(a_full + a, (uint32_t)around & 1u);
        if (p.cells) fence_proxy_async();      // the cells were written through the generic proxy (cp.async)
        const uint32_t a_st = (a16 + (uint32_t)a * (a_stage >> 4)) | a_hl;
        int r = 0, q = 0;
        for (int sl = 0; sl < nslices; ++sl) {
          const int nt = min(p.tps, ntaps - sl * p.tps);
          if (lane == 0) WTRACE(2, ev, 1);
          ptx::mbar_wait(b_full + b, (uint32_t)bround & 1u);
          ptx::tc_fence_after();
          if (lane == 0) WTRACE(2, ev, 2);
          uint32_t b_lo = (b16 + (uint32_t)b * (b_stage >> 4)) | b_hl;
          for (int tl = 0; tl < nt; ++tl) {
            uint32_t a_lo = a_st + (uint32_t)(r * p.Wp + q);                   // 16-byte units: one position
            for (int j = 0; j < nk; ++j, a_lo += (uint32_t)(2 * win), b_lo += (uint32_t)(2 * Nt)) {
              uint32_t dm = d, am = a_lo;
              for (int mt = 0; mt < MT; ++mt, dm += set_cols, am += 128u) umma_bf16_elect(dm, am, a_hh, b_lo, b_hh, idesc, acc);
              acc = 1u;
            }
            if (++q == p.kw) q = 0, ++r;
          }
          umma_commit_elect(b_empty + b);
          if (lane == 0) WTRACE(2, ev, 3);
          ++ev;
          if (++b == SB) b = 0, ++bround;
        }
        umma_commit_elect(a_empty + a);
        if (++a == SA) a = 0, ++around;
      }
      umma_commit_elect(acc_full + set);
    }
  } else if (warp >= 4) {
    // ---------------- epilogue: bias + ReLU, bf16, 16-byte granules at the concat offset ----------------
    const int q4 = warp & 3;                                        // the TMEM lane quarter this warp may read
    const int hwp = p.Hp * p.Wp;
    const long long oplane = p.m_out * 16;
    const float lo = p.relu ? 0.f : -INFINITY;
    int t = 0;
    for (int w = blockIdx.x; w < total; w += gridDim.x, ++t) {
      const int set = SETS == 2 ? (t & 1) : 0, use = SETS == 2 ? (t >> 1) : t;
      const int n0 = (w % p.n_tiles) * Nt;
      const float* bias = bias_s + n0;
      if (threadIdx.x == 128) WTRACE(3, t, 0);
      mbar_wait_relaxed(acc_full + set, (uint32_t)use & 1u);
      ptx::tc_fence_after();
      if (threadIdx.x == 128) WTRACE(3, t, 1);
      for (int mt = 0; mt < MT; ++mt) {
        const int pp = ((w / p.n_tiles) * MT + mt) * 128 + q4 * 32 + lane;       // padded position of this thread's row
        const int img = pp / hwp, rem = pp - img * hwp;
        const int ho = rem / p.Wp, wo = rem - ho * p.Wp;
        const bool store = pp < p.pp_total && ho < p.Ho && wo < p.Wo;
        const long long m = ((long long)img * p.Ho + ho) * p.Wo + wo;
        char* out_row = p.out + ((long long)(p.out_chunk0 + (n0 >> 3)) * p.m_out + m) * 16;
        const uint32_t trow = tmem_base + ((uint32_t)(q4 * 32) << 16) + (uint32_t)(set * MT + mt) * set_cols;
        int c0 = 0;
        for (; c0 + 32 <= Nt; c0 += 32) {
          uint32_t r[32];
          ptx::tmem_ld32_issue(trow + (uint32_t)c0, r);
          ptx::tmem_wait32(r);
          if (store) {
#pragma unroll
            for (int k = 0; k < 4; ++k)
              *reinterpret_cast<uint4*>(out_row + ((c0 >> 3) + k) * oplane) = bias_act_pack8(r + 8 * k, bias + c0 + 8 * k, lo);
          }
        }
        if (c0 < Nt) {                                              // Nt is a multiple of 16
          uint32_t r[16];
          ptx::tmem_ld16_issue(trow + (uint32_t)c0, r);
          ptx::tmem_wait16(r);
          if (store) {
#pragma unroll
            for (int k = 0; k < 2; ++k)
              *reinterpret_cast<uint4*>(out_row + ((c0 >> 3) + k) * oplane) = bias_act_pack8(r + 8 * k, bias + c0 + 8 * k, lo);
          }
        }
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(acc_empty + set);
      if (threadIdx.x == 128) WTRACE(3, t, 2);
    }
  }
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 2) ptx::tmem_dealloc(tmem_base, cols);
}

// ---------------------------------------------------------------------------------------------
// Conv2d_1a_3x3 (modelC:63): 1 -> 32 channels, 3x3, stride 2, no padding, straight from the fp32 patch (K = 9: CUDA cores)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) convc_first_kernel(const float* __restrict__ x, const float* __restrict__ w9,
                                                          const float* __restrict__ bias, char* __restrict__ out, int n,
                                                          int H, int W, int Ho, int Wo, int cout) {
  __shared__ __align__(16) float ws[9 * 64], bs[64];
  for (int i = threadIdx.x; i < 9 * cout; i += 256) ws[i] = w9[i];       // [tap][cout]
  for (int i = threadIdx.x; i < cout; i += 256) bs[i] = bias[i];
  __syncthreads();
  const long long M = (long long)n * Ho * Wo;
  for (long long m = (long long)blockIdx.x * 256 + threadIdx.x; m < M; m += (long long)gridDim.x * 256) {
    const int img = (int)(m / (Ho * Wo));
    const int rem = (int)(m - (long long)img * Ho * Wo);
    const int ho = rem / Wo, wo = rem - ho * Wo;
    const float* src = x + ((long long)img * H + 2 * ho) * W + 2 * wo;
    float v[9];
#pragma unroll
    for (int r = 0; r < 3; ++r)
#pragma unroll
      for (int q = 0; q < 3; ++q) v[r * 3 + q] = src[r * W + q];
    for (int c0 = 0; c0 < cout; c0 += 8) {
      float acc[8];
      *reinterpret_cast<float4*>(acc) = *reinterpret_cast<const float4*>(bs + c0);
      *reinterpret_cast<float4*>(acc + 4) = *reinterpret_cast<const float4*>(bs + c0 + 4);
#pragma unroll
      for (int t = 0; t < 9; ++t) {               // 128-bit broadcast reads: the kernel was bound by 288 scalar LDS per thread
        const float4 w0 = *reinterpret_cast<const float4*>(ws + t * cout + c0);
        const float4 w1 = *reinterpret_cast<const float4*>(ws + t * cout + c0 + 4);
        acc[0] = fmaf(v[t], w0.x, acc[0]), acc[1] = fmaf(v[t], w0.y, acc[1]);
        acc[2] = fmaf(v[t], w0.z, acc[2]), acc[3] = fmaf(v[t], w0.w, acc[3]);
        acc[4] = fmaf(v[t], w1.x, acc[4]), acc[5] = fmaf(v[t], w1.y, acc[5]);
        acc[6] = fmaf(v[t], w1.z, acc[6]), acc[7] = fmaf(v[t], w1.w, acc[7]);
      }
      uint4 q4;
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&q4);
#pragma unroll
      for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(fmaxf(acc[2 * j], 0.f), fmaxf(acc[2 * j + 1], 0.f));
      *reinterpret_cast<uint4*>(out + ((long long)(c0 >> 3) * M + m) * 16) = q4;
    }
  }
}

// ---------------------------------------------------------------------------------------------
// 3x3 pools: max (no padding, stride 2: nn.MaxPool2d(3,2) / F.max_pool2d(x,3,2)) and average (stride 1, pad 1, the
// padding counts: F.avg_pool2d(x, 3, 1, 1) divides by 9 everywhere)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) poolc_kernel(const char* __restrict__ x, char* __restrict__ out, int n, int chunks,
                                                    int H, int W, int Ho, int Wo, int stride, int pad, int is_avg,
                                                    int out_chunk0) {
  // grid = (position blocks, channel chunks): 32-bit index arithmetic only (64-bit divisions dominated the first version)
  const unsigned Mo = (unsigned)n * Ho * Wo, hw = (unsigned)Ho * Wo;
  const long long Mi = (long long)n * H * W;
  const int c = blockIdx.y;
  const char* xin = x + (long long)c * Mi * 16;
  char* dst = out + (long long)(out_chunk0 + c) * Mo * 16;
  for (unsigned m = blockIdx.x * 256u + threadIdx.x; m < Mo; m += gridDim.x * 256u) {
    const unsigned img = m / hw, rem = m - img * hw;
    const int ho = (int)(rem / (unsigned)Wo), wo = (int)(rem - (unsigned)ho * Wo);
    const int h0 = ho * stride - pad, w0 = wo * stride - pad;
    float acc[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) acc[j] = is_avg ? 0.f : -INFINITY;
    const char* base = xin + ((long long)img * H * W) * 16;
#pragma unroll
    for (int r = 0; r < 3; ++r) {
      const int hi = h0 + r;
      if (hi < 0 || hi >= H) continue;
#pragma unroll
      for (int q = 0; q < 3; ++q) {
        const int wi = w0 + q;
        if (wi < 0 || wi >= W) continue;
        const uint4 g = *reinterpret_cast<const uint4*>(base + (hi * W + wi) * 16);
        const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&g);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float2 f = __bfloat1622float2(h[j]);
          if (is_avg) acc[2 * j] += f.x, acc[2 * j + 1] += f.y;
          else acc[2 * j] = fmaxf(acc[2 * j], f.x), acc[2 * j + 1] = fmaxf(acc[2 * j + 1], f.y);
        }
      }
    }
    uint4 o;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&o);
    const float sc = is_avg ? (1.f / 9.f) : 1.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(acc[2 * j] * sc, acc[2 * j + 1] * sc);
    *reinterpret_cast<uint4*>(dst + (long long)m * 16) = o;
  }
}

// F.avg_pool2d(x, 3, 1, 1) as a separable rolling sum: a thread owns one (image, column, 8-channel chunk) and walks down
// the rows, three 16-byte loads and ~50 arithmetic instructions per output instead of nine loads and ~250 (the generic
// kernel above is instruction bound on these layers, three times off the HBM time).
__device__ __forceinline__ void add_cell(float* acc, const char* p) {
  const uint4 g = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    acc[2 * j] += __uint_as_float(w[j] << 16);              // bf16 -> fp32 is a shift
    acc[2 * j + 1] += __uint_as_float(w[j] & 0xffff0000u);
  }
}
__global__ void __launch_bounds__(256) avgpoolc_kernel(const char* __restrict__ x, char* __restrict__ out, int n, int H, int W,
                                                       int out_chunk0) {
  const unsigned idx = blockIdx.x * 256u + threadIdx.x;
  if (idx >= (unsigned)n * W) return;
  const unsigned img = idx / (unsigned)W;
  const int wo = (int)(idx - img * (unsigned)W);
  const long long M = (long long)n * H * W;
  const char* col = x + ((long long)blockIdx.y * M + (long long)img * H * W + wo) * 16;
  char* dst = out + ((long long)(out_chunk0 + blockIdx.y) * M + (long long)img * H * W + wo) * 16;
  const bool left = wo > 0, right = wo + 1 < W;
  float r0[8], r1[8], r2[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) r0[j] = 0.f, r1[j] = 0.f;
  add_cell(r1, col);
  if (left) add_cell(r1, col - 16);
  if (right) add_cell(r1, col + 16);
  for (int h = 0; h < H; ++h) {
#pragma unroll
    for (int j = 0; j < 8; ++j) r2[j] = 0.f;
    if (h + 1 < H) {
      const char* p = col + (long long)(h + 1) * W * 16;
      add_cell(r2, p);
      if (left) add_cell(r2, p - 16);
      if (right) add_cell(r2, p + 16);
    }
    uint4 o;
    __nv_bfloat162* hh = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
    for (int j = 0; j < 4; ++j)
      hh[j] = __floats2bfloat162_rn((r0[2 * j] + r1[2 * j] + r2[2 * j]) * (1.f / 9.f),
                                    (r0[2 * j + 1] + r1[2 * j + 1] + r2[2 * j + 1]) * (1.f / 9.f));
    *reinterpret_cast<uint4*>(dst + (long long)h * W * 16) = o;
#pragma unroll
    for (int j = 0; j < 8; ++j) r0[j] = r1[j], r1[j] = r2[j];
  }
}

// ---------------------------------------------------------------------------------------------
// AdaptiveAvgPool2d((1,1)) + flatten + Linear(C, classes)   (modelC:137-146; Dropout is the identity in eval mode)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) headc_kernel(const char* __restrict__ x, int C, int hw, long long M,
                                                    const float* __restrict__ fw, const float* __restrict__ fb, int classes,
                                                    float* __restrict__ logits) {
  extern __shared__ float gap[];                     // [C]
  const int img = blockIdx.x;
  for (int c = threadIdx.x; c < C / 8; c += 256) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int q = 0; q < hw; ++q) {
      const uint4 g = *reinterpret_cast<const uint4*>(x + ((long long)c * M + (long long)img * hw + q) * 16);
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&g);
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const float2 f = __bfloat1622float2(h[j]);
        acc[2 * j] += f.x, acc[2 * j + 1] += f.y;
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) gap[c * 8 + j] = acc[j] / (float)hw;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int k = warp; k < classes; k += 8) {
    float s = 0.f;
    for (int c = lane; c < C; c += 32) s = fmaf(gap[c], fw[(size_t)k * C + c], s);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) logits[(size_t)img * classes + k] = s + fb[k];
  }
}

// ---------------------------------------------------------------------------------------------
// weight packing with the eval-mode BatchNorm fold (eps 1e-3): w' = w*g/sqrt(rv+eps), b' = beta - rm*g/sqrt(rv+eps)
// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) packc_kernel(const float* __restrict__ w, const float* __restrict__ gamma,
                                                    const float* __restrict__ beta, const float* __restrict__ rmean,
                                                    const float* __restrict__ rvar, float eps, int cout, int cin, int taps,
                                                    __nv_bfloat16* __restrict__ dst, float* __restrict__ bias,
                                                    float* __restrict__ first_w, int Nt, int ks) {
  const int total = cout * cin * taps;
  for (int i = blockIdx.x * 256 + threadIdx.x; i < total; i += gridDim.x * 256) {
    const int co = i / (cin * taps), r1 = i - co * cin * taps, ci = r1 / taps, t = r1 - ci * taps;
    const float sc = gamma[co] * rsqrtf(rvar[co] + eps);
    if (r1 == 0) bias[co] = beta[co] - rmean[co] * sc;
    const float v = w[i] * sc;
    if (first_w) first_w[t * cout + co] = v;                            // [tap][cout] fp32 (cin == 1)
    else {                                                              // [n tile][K unit][tap][chunk][Nt][8]
      const int CC = cin >> 3, gc = ci >> 3, ku = gc / (2 * ks), cl = gc - ku * 2 * ks, nch = min(2 * ks, CC - ku * 2 * ks);
      const int nt = co / Nt, col = co - nt * Nt;
      dst[((((size_t)nt * taps * CC + (size_t)ku * taps * 2 * ks + (size_t)t * nch + cl) * Nt) + col) * 8 + (ci & 7)] =
          __float2bfloat16(v);
    }
  }
}

__global__ void __launch_bounds__(256) readc_kernel(const char* __restrict__ x, long long M, int hw, int C, int c0,
                                                    float* __restrict__ nchw) {
  const long long total = M * (C / 8);
  for (long long i = (long long)blockIdx.x * 256 + threadIdx.x; i < total; i += (long long)gridDim.x * 256) {
    const int c = (int)(i / M);
    const long long m = i - (long long)c * M;
    const long long img = m / hw, q = m - img * hw;
    const uint4 g = *reinterpret_cast<const uint4*>(x + ((long long)(c0 / 8 + c) * M + m) * 16);
    const __nv_bfloat16* h = reinterpret_cast<const __nv_bfloat16*>(&g);
    for (int j = 0; j < 8; ++j) nchw[((img * C) + c * 8 + j) * hw + q] = __bfloat162float(h[j]);
  }
}

int pick_nt(int cout) {   // output channels per CTA: the largest divisor of cout that is a multiple of 16 and <= 256
  for (int parts = 1; parts <= 8; ++parts)
    if (cout % parts == 0 && (cout / parts) % 16 == 0 && cout / parts <= 256) return cout / parts;
  return 0;
}


// Weight layout shared by both convolution kernels: [n tile][K unit][tap][chunk of the unit][Nt][8] bf16, a K unit being
// `ks` K steps (2 ks chunks of 8 channels; the last unit may be shorter).  Every (n tile, K unit, run of taps) is one
// contiguous block: the window kernel brings a weight slice with ONE bulk copy (a bulk copy costs ~70 cycles of issue
// whatever its size).  ks is chosen so that a slice (one synchronisation step of the MMA warp) carries >= ~8 UMMAs.
struct WeightLayout {
  int Nt, ks, m_max;
};
WeightLayout weight_layout(int cin, int cout, int kh, int kw) {
  WeightLayout l{};
  l.Nt = pick_nt(cout);
  if (!l.Nt) return l;
  const size_t tap_step = 32 * (size_t)l.Nt;                       // bytes of one tap, one K step
  l.m_max = (int)std::max<size_t>(1, (l.Nt >= 128 ? 48 * 1024 : 32 * 1024) / tap_step);
  const int ntaps = kh * kw, KC = cin / 16;
  l.ks = 1;
  if (ntaps < l.m_max)
    while (l.ks * 2 <= KC && l.ks * 2 * ntaps <= l.m_max && l.ks < 8) l.ks *= 2;
  // 1x1 layers: two window stages (128 positions x 2 ks chunks) and two weight slices must fit HALF an SM's shared
  // memory, so that two CTAs share the SM -- two issuing warps keep the tensor pipe twice as busy as one (measured on
  // 768 -> 192: one CTA with ks = 8 spends 1500 cycles per 8 UMMAs of 98 cycles each)
  if (ntaps == 1)
    while (l.ks > 1 && 2 * (size_t)l.ks * (32 * 128 + tap_step) > 104 * 1024) l.ks /= 2;
  // At least two K steps per unit: the gather kernel's ring unit (two K steps) then finds its weights in one contiguous
  // block = one bulk copy (7x1 layers 750 -> 810 TFLOP/s), and the window kernel does twice the UMMAs per window stage
  // (3x3 / 5x5 layers +7 %).  Not for wide kernels (1x7): their window rows are short and every extra chunk plane costs
  // a bulk copy per row (0.080 -> 0.107 ms).
  if (kw <= kh && l.ks == 1 && KC >= 2) l.ks = 2;
  return l;
}

}  // namespace
}  // namespace mtldas

using namespace mtldas;

#define CKC(call, what)                                                     \
  do {                                                                      \
    cudaError_t e_ = (call);                                                \
    if (e_ != cudaSuccess) {                                                \
      set_error("%s: %s", what, cudaGetErrorString(e_));                    \
      return MTLDAS_ERR_CUDA;                                               \
    }                                                                       \
  } while (0)

extern "C" int mtldas_c_pack_conv(const float* dev_w, const float* dev_gamma, const float* dev_beta, const float* dev_rmean,
                                  const float* dev_rvar, float eps, int cout, int cin, int kh, int kw, void* dev_packed,
                                  float* dev_bias, void* stream) {
  if (!dev_w || !dev_gamma || !dev_beta || !dev_rmean || !dev_rvar || !dev_packed || !dev_bias || cout < 1 || kh < 1 || kw < 1 ||
      !(cin == 1 || cin % 16 == 0)) {
    set_error("c_pack_conv: bad arguments (cin must be 1 or a multiple of 16)");
    return MTLDAS_ERR_ARG;
  }
  const int total = cout * cin * kh * kw;
  WeightLayout lay{1, 1, 1};
  if (cin != 1) {
    lay = weight_layout(cin, cout, kh, kw);
    if (!lay.Nt) {
      set_error("c_pack_conv: cout %d not supported (a multiple of 16 with a divisor <= 256)", cout);
      return MTLDAS_ERR_UNSUPPORTED;
    }
  }
  packc_kernel<<<std::min(592, (total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      dev_w, dev_gamma, dev_beta, dev_rmean, dev_rvar, eps, cout, cin, kh * kw, cin == 1 ? nullptr : (__nv_bfloat16*)dev_packed,
      dev_bias, cin == 1 ? (float*)dev_packed : nullptr, lay.Nt, lay.ks);
  CKC(cudaGetLastError(), "c_pack_conv");
  return 0;
}

extern "C" int mtldas_c_conv(const void* dev_x, int n, int cin, int h, int w, const void* dev_packed, const float* dev_bias,
                             void* dev_out, int out_channels_total, int out_channel0, int cout, int kh, int kw, int sh, int sw,
                             int ph, int pw, int relu, void* stream) {
  const int Ho = (h + 2 * ph - kh) / sh + 1, Wo = (w + 2 * pw - kw) / sw + 1;
  if (!dev_x || !dev_packed || !dev_bias || !dev_out || n < 1 || Ho < 1 || Wo < 1 || (out_channel0 % 8) || (out_channels_total % 8) ||
      out_channel0 + cout > out_channels_total) {
    set_error("c_conv: bad arguments");
    return MTLDAS_ERR_ARG;
  }
  cudaStream_t st = (cudaStream_t)stream;
  const long long m_out = (long long)n * Ho * Wo;
  if (cin == 1) {   // Conv2d_1a_3x3: fp32 patches in, CUDA cores
    if (kh != 3 || kw != 3 || sh != 2 || sw != 2 || ph || pw || cout > 64 || (cout % 8) || out_channel0) {
      set_error("c_conv: the 1-channel entry convolution is 3x3 / stride 2 / no padding (modelC:63)");
      return MTLDAS_ERR_UNSUPPORTED;
    }
    const int blocks = (int)std::min<long long>(148 * 8, (m_out + 255) / 256);
    convc_first_kernel<<<blocks, 256, 0, st>>>((const float*)dev_x, (const float*)dev_packed, dev_bias, (char*)dev_out, n, h, w,
                                               Ho, Wo, cout);
    CKC(cudaGetLastError(), "c_conv (first)");
    return 0;
  }
  const WeightLayout lay = weight_layout(cin, cout, kh, kw);
  const int Nt = lay.Nt;
  if ((cin % 16) || !Nt) {
    set_error("c_conv: cin %d / cout %d not supported (multiples of 16)", cin, cout);
    return MTLDAS_ERR_UNSUPPORTED;
  }
  static const bool use_window = [] {
    const char* e = getenv("MTLDAS_C_WINDOW");       // 0: the gather kernel for every layer (A/B timing)
    return !(e && e[0] == '0');
  }();
  if (use_window && sh == 1 && sw == 1 && cout <= 512) {
    ConvWParams q{};
    q.x = (const char*)dev_x, q.w = (const char*)dev_packed, q.bias = dev_bias, q.out = (char*)dev_out;
    q.m_in = (long long)n * h * w, q.m_out = m_out;
    q.n = n, q.cin = cin, q.H = h, q.W = w, q.Ho = Ho, q.Wo = Wo, q.cout = cout;
    q.kh = kh, q.kw = kw, q.ph = ph, q.pw = pw, q.tap_r0 = 0;
    if (h == 1 && kh == 2 * ph + 1) q.tap_r0 = ph, q.kh = 1, q.ph = 0;      // one-row images: only the centre kernel row
    q.Hp = h + q.ph, q.Wp = w + q.pw, q.off = q.ph * q.Wp + q.pw;     // padding shared by neighbouring rows / images
    static const int cells_w = [] { const char* e = getenv("MTLDAS_C_CELLS_W"); return e ? atoi(e) : 8; }();
    q.cells = (q.ph | q.pw) != 0 && w <= cells_w;
    q.out_chunk0 = out_channel0 / 8, q.Nt = Nt, q.relu = relu;
    const long long pp_total = (long long)n * q.Hp * q.Wp;
    static const int max_waste = [] { const char* e = getenv("MTLDAS_C_WASTE"); return e ? atoi(e) : 145; }();
    const bool window_ok = q.kh > q.ph && q.kw > q.pw && pp_total * 100 <= m_out * max_waste;   // else: the gather kernel
    const int ntaps = q.kh * q.kw;
    // One synchronisation step of the MMA warp (a weight slice) carries up to m_max UMMAs: `tps` taps x `ks` K steps,
    // ks being fixed by the weight layout.
    const size_t tap_step = 32 * (size_t)Nt;                                 // bytes of one tap, one K step
    q.ks = lay.ks, q.ntaps_full = kh * kw;
    const int per_slice = std::max(1, lay.m_max / lay.ks), nsl = (ntaps + per_slice - 1) / per_slice;
    q.tps = (ntaps + nsl - 1) / nsl;
    const size_t b_stage = (size_t)q.tps * q.ks * tap_step;
    const int set_cols = (Nt + 31) & ~31, halo = (q.kh - 1) * q.Wp + (q.kw - 1);
    int dev = 0, sms = 0;
    CKC(cudaGetDevice(&dev), "c_conv (device)");
    CKC(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev), "c_conv (device)");
    // Configurations, best first.  Two CTAs per SM double the UMMA issue rate (one issuing warp per CTA, bound by its own
    // instruction latency: ~70 cycles per UMMA in the unrolled m-tile loop, ~220 in the tap loop), so they come first;
    // each CTA then gets half of the shared memory and at most 256 TMEM columns.
    static const int force_ctas = [] { const char* e = getenv("MTLDAS_C_CTAS"); return e ? atoi(e) : 0; }();
    static const int mt_cap = [] { const char* e = getenv("MTLDAS_C_MT"); return e ? atoi(e) : 4; }();
    struct Cfg { int ctas, sets, mt; };
    static const Cfg cfgs[] = {{2, 2, 4}, {2, 2, 2}, {2, 1, 2}, {2, 2, 1}, {2, 1, 1}, {1, 2, 4}, {1, 2, 2}, {1, 2, 1}, {1, 1, 2}, {1, 1, 1}};
    size_t a_stage = 0, smem_w = 0;
    int ctas = 0;
    static const Cfg force_cfg = [] {          // MTLDAS_C_CFG="ctas,sets,mt": experiments
      Cfg f{0, 0, 0};
      const char* e = getenv("MTLDAS_C_CFG");
      if (e) sscanf(e, "%d,%d,%d", &f.ctas, &f.sets, &f.mt);
      return f;
    }();
    for (const Cfg& c : cfgs) {
      if ((force_ctas && c.ctas != force_ctas) || c.mt > mt_cap) continue;
      if (force_cfg.ctas && (c.ctas != force_cfg.ctas || c.sets != force_cfg.sets || c.mt != force_cfg.mt)) continue;
      if (c.sets * c.mt * set_cols > (c.ctas == 2 ? 256 : 512)) continue;
      const long long items_c = (pp_total + 128 * c.mt - 1) / (128 * c.mt) * (cout / Nt);
      if (c.mt > 1 && items_c < 2ll * sms * c.ctas && !force_cfg.ctas) continue;   // keep every CTA busy for a few items
      if (c.ctas == 2 && items_c <= sms && !force_ctas && !force_cfg.ctas) continue;          // one item per SM at most: deeper rings instead
      const size_t bud = (c.ctas == 2 ? kWSmemBudget / 2 - 1024 : kWSmemBudget) - kWHeader;
      q.win = (128 * c.mt + halo + 7) & ~7;
      a_stage = 32 * (size_t)q.ks * q.win;
      if (2 * a_stage + 2 * b_stage > bud) continue;
      q.b_stages = 3 * b_stage <= bud * 45 / 100 ? (4 * b_stage <= bud * 30 / 100 ? 4 : 3) : 2;
      while (q.b_stages > 2 && (bud - q.b_stages * b_stage) / a_stage < 2) --q.b_stages;   // two window stages come first
      q.a_stages = (int)std::min<size_t>(kWMaxA, (bud - q.b_stages * b_stage) / a_stage);
      if (q.a_stages < 2) continue;
      q.mt = c.mt, q.sets = c.sets, ctas = c.ctas, smem_w = bud + kWHeader;
      break;
    }
    if (pp_total < (1ll << 31) - 65536 && ctas && window_ok) {
      q.pp_total = (int)pp_total;
      q.m_tiles = (int)((pp_total + 128 * q.mt - 1) / (128 * q.mt)), q.n_tiles = cout / Nt;
      static PerDeviceSmem configured_w;
      CKC(configured_w.ensure(kWSmemBudget, [](int bytes) {
            return cudaFuncSetAttribute(convc_window_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
          }), "c_conv (attribute)");
      const long long items = (long long)q.m_tiles * q.n_tiles;
      static const bool trace = getenv("MTLDAS_C_TRACE") != nullptr;      // timeline of CTA 0 on stdout (tools/modelc_layer.py)
      if (trace) cudaMalloc(&q.trace, 1024 * 8), cudaMemset(q.trace, 0, 1024 * 8);
      convc_window_kernel<<<(unsigned)std::min<long long>(items, (long long)sms * ctas), kWThreads, smem_w, st>>>(q);
      CKC(cudaGetLastError(), "c_conv (window)");
      if (trace) {
        static long long h[1024];
        cudaDeviceSynchronize();
        cudaMemcpy(h, q.trace, sizeof(h), cudaMemcpyDeviceToHost);
        cudaFree(q.trace);
        const long long z = h[2 * 256];
        printf("window kernel: %d CTA/SM, A stages %d x %zu B, B stages %d x %zu B, ks %d tps %d mt %d sets %d win %d Nt %d\n", ctas, q.a_stages,
               a_stage, q.b_stages, b_stage, q.ks, q.tps, q.mt, q.sets, q.win, Nt);
        for (int i = 0; i < 40; ++i) {
          const long long *a = h + i * 4, *b = h + 256 + i * 4, *m = h + 512 + i * 4, *e = h + 768 + i * 4;
          printf("%2d | A wait %6lld got %6lld issued %6lld | B wait %6lld got %6lld issued %6lld | MMA a-wait %6lld b-wait %6lld got %6lld "
                 "committed %6lld | EPI wait %6lld got %6lld done %6lld\n", i, a[0] - z, a[1] - z, a[2] - z, b[0] - z, b[1] - z, b[2] - z,
                 m[0] - z, m[1] - z, m[2] - z, m[3] - z, e[0] - z, e[1] - z, e[2] - z);
        }
      }
      return 0;
    }
  }
  ConvCParams p{};
  p.x = (const char*)dev_x, p.w = (const char*)dev_packed, p.bias = dev_bias, p.out = (char*)dev_out;
  p.m_in = (long long)n * h * w, p.m_out = m_out;
  p.n = n, p.cin = cin, p.H = h, p.W = w, p.Ho = Ho, p.Wo = Wo, p.cout = cout, p.kh = kh, p.kw = kw, p.sh = sh, p.sw = sw;
  p.ph = ph, p.pw = pw, p.out_chunk0 = out_channel0 / 8, p.Nt = Nt, p.relu = relu, p.ks = lay.ks;
  const size_t smem = 256 + 1024 + (size_t)kCStages * (2 * kCUnitSteps * 128 * 16 + 2 * kCUnitSteps * Nt * 16);
  static PerDeviceSmem configured;
  CKC(configured.ensure(smem, [](int bytes) {
        return cudaFuncSetAttribute(convc_umma_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes);
      }), "c_conv (attribute)");
  dim3 grid((unsigned)((m_out + 127) / 128), (unsigned)(cout / Nt));
  convc_umma_kernel<<<grid, kCThreads, smem, st>>>(p);
  CKC(cudaGetLastError(), "c_conv");
  return 0;
}

extern "C" int mtldas_c_pool(const void* dev_x, int n, int channels, int h, int w, void* dev_out, int out_channels_total,
                             int out_channel0, int stride, int pad, int is_avg, void* stream) {
  const int Ho = (h + 2 * pad - 3) / stride + 1, Wo = (w + 2 * pad - 3) / stride + 1;
  if (!dev_x || !dev_out || n < 1 || (channels % 8) || (out_channel0 % 8) || out_channel0 + channels > out_channels_total ||
      Ho < 1 || Wo < 1) {
    set_error("c_pool: bad arguments");
    return MTLDAS_ERR_ARG;
  }
  const long long mo = (long long)n * Ho * Wo;
  if (mo >= (1ll << 31) || (long long)n * h * w >= (1ll << 27)) {
    set_error("c_pool: batch too large for the 32-bit index arithmetic of this kernel");
    return MTLDAS_ERR_UNSUPPORTED;
  }
  const int chunks = channels / 8;
  if (is_avg && stride == 1 && pad == 1) {
    avgpoolc_kernel<<<dim3((unsigned)(((long long)n * w + 255) / 256), (unsigned)chunks), 256, 0, (cudaStream_t)stream>>>(
        (const char*)dev_x, (char*)dev_out, n, h, w, out_channel0 / 8);
    CKC(cudaGetLastError(), "c_pool (avg)");
    return 0;
  }
  const unsigned bx = (unsigned)std::max<long long>(1, std::min<long long>((mo + 255) / 256, (148 * 16 + chunks - 1) / chunks));
  poolc_kernel<<<dim3(bx, (unsigned)chunks), 256, 0, (cudaStream_t)stream>>>(
      (const char*)dev_x, (char*)dev_out, n, channels / 8, h, w, Ho, Wo, stride, pad, is_avg, out_channel0 / 8);
  CKC(cudaGetLastError(), "c_pool");
  return 0;
}

extern "C" int mtldas_c_head(const void* dev_x, int n, int channels, int hw, const float* dev_fc_w, const float* dev_fc_b,
                             int classes, float* dev_logits, void* stream) {
  if (!dev_x || !dev_fc_w || !dev_fc_b || !dev_logits || n < 1 || (channels % 8) || hw < 1 || classes < 1 || channels > 8192) {
    set_error("c_head: bad arguments");
    return MTLDAS_ERR_ARG;
  }
  headc_kernel<<<n, 256, (size_t)channels * sizeof(float), (cudaStream_t)stream>>>((const char*)dev_x, channels, hw,
                                                                                  (long long)n * hw, dev_fc_w, dev_fc_b,
                                                                                  classes, dev_logits);
  CKC(cudaGetLastError(), "c_head");
  return 0;
}

extern "C" int mtldas_c_read(const void* dev_x, int n, int channels_total, int channel0, int channels, int hw,
                             float* dev_out_nchw, void* stream) {
  if (!dev_x || !dev_out_nchw || (channel0 % 8) || (channels % 8) || channel0 + channels > channels_total) {
    set_error("c_read: bad arguments");
    return MTLDAS_ERR_ARG;
  }
  const long long M = (long long)n * hw, total = M * (channels / 8);
  readc_kernel<<<(int)std::min<long long>(148 * 16, (total + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      (const char*)dev_x, M, hw, channels, channel0, dev_out_nchw);
  CKC(cudaGetLastError(), "c_read");
  return 0;
}
