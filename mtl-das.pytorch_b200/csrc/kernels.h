// Launch functions of the mtldas kernels (host side).
#pragma once
#include <cuda_runtime.h>

#include "types.h"

namespace mtldas {

// precision: 0 = bf16 (tcgen05), 1 = fp32 (CUDA cores)
cudaError_t launch_conv(const ConvParams& p, int precision, cudaStream_t stream, bool pdl = false);
size_t conv_umma_smem_bytes(const ConvParams& p);
float umma_selftest(int n_cols, int row_shift, int swap);
int umma_rate(int n_cols, int mode, int reps, float* host_out2);

cudaError_t launch_stem(const StemParams& p, int precision, cudaStream_t stream, bool pdl = false);
cudaError_t launch_pool(const PoolParams& p, int precision, cudaStream_t stream, bool pdl = false);
cudaError_t launch_head(const HeadParams& p, int precision, cudaStream_t stream, bool pdl = false);
cudaError_t launch_pack(const PackParams& p, cudaStream_t stream);
cudaError_t launch_pack_batch(const PackParams* packs, int n, cudaStream_t stream);

cudaError_t launch_window_gather(const float* arr, long long n_time, int start0, int stride, int n_windows,
                                 int n_tblocks, float* out, cudaStream_t stream);

// layout conversion (tests, debugging, and the autograd boundary): fp32 NCHW <-> P8 / P8S2 views
cudaError_t launch_nchw_to_view(const float* nchw, ActView v, int nb, int C, int H, int W, int precision,
                                cudaStream_t stream);
cudaError_t launch_view_to_nchw(ActView v, float* nchw, int nb, int C, int H, int W, int precision,
                                cudaStream_t stream);

}  // namespace mtldas
