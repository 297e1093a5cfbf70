// Launch functions of the mtldas kernels (host side).
#pragma once
#include <cuda_runtime.h>

#include <mutex>

#include "types.h"

namespace mtldas {

// cudaFuncAttributeMaxDynamicSharedMemorySize is a property of (kernel, DEVICE): every launcher keeps one of these
// tables per kernel family and raises the limit on the device that is current at the launch.  Safe under concurrent
// launches from several host threads (distinct plans may run concurrently, include/mtldas.h).
class PerDeviceSmem {
 public:
  template <class F>
  cudaError_t ensure(size_t smem, F&& set_attr) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= kMaxDev) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu_);
    if (smem <= bytes_[dev]) return cudaSuccess;
    e = set_attr((int)smem);
    if (e == cudaSuccess) bytes_[dev] = smem;
    return e;
  }

 private:
  static constexpr int kMaxDev = 64;
  std::mutex mu_;
  size_t bytes_[kMaxDev] = {};
};

// precision: 0 = bf16 (tcgen05), 1 = fp32 (CUDA cores)
cudaError_t launch_conv(const ConvParams& p, int precision, cudaStream_t stream, bool pdl = false);
size_t conv_umma_smem_bytes(const ConvParams& p);
float umma_selftest(int n_cols, int row_shift, int swap);
int umma_rate(int n_cols, int mode, int reps, float* host_out2);

cudaError_t launch_stem(const StemParams& p, int precision, cudaStream_t stream, bool pdl = false);
cudaError_t launch_stem_umma(const StemParams& p, cudaStream_t stream, bool pdl);   // cudaErrorInvalidValue: not eligible
cudaError_t launch_pool(const PoolParams& p, int precision, cudaStream_t stream, bool pdl = false);
cudaError_t launch_head(const HeadParams& p, int precision, cudaStream_t stream, bool pdl = false);
cudaError_t launch_pack(const PackParams& p, cudaStream_t stream);
cudaError_t launch_pack_batch(const PackParams* packs, int n, cudaStream_t stream);

cudaError_t launch_window_gather(const float* arr, long long n_time, int start0, int stride, int n_windows,
                                 int n_tblocks, float* out, cudaStream_t stream);

// bf16 -> fp32 (exact) for host patches uploaded in half the bytes
cudaError_t launch_widen_bf16(const uint16_t* in, float* out, long long n, cudaStream_t stream);

// layout conversion (tests, debugging, and the autograd boundary): fp32 NCHW <-> P8 / P8S2 views
cudaError_t launch_nchw_to_view(const float* nchw, ActView v, int nb, int C, int H, int W, int precision,
                                cudaStream_t stream);
cudaError_t launch_view_to_nchw(ActView v, float* nchw, int nb, int C, int H, int W, int precision,
                                cudaStream_t stream);

}  // namespace mtldas
