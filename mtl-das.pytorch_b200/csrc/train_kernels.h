// Launch functions of the training-mode kernels.
#pragma once
#include <cuda_runtime.h>

#include "train_types.h"

namespace mtldas {

cudaError_t launch_ew_fwd(const EwParams& p, int precision, cudaStream_t st);
cudaError_t launch_ew_bwd(const EwParams& p, int precision, cudaStream_t st);
cudaError_t launch_wgrad(const WgradParams& p, int precision, cudaStream_t st);
cudaError_t launch_stem_wgrad(const StemWgradParams& p, int precision, cudaStream_t st);
cudaError_t launch_nll_loss(const float* logp0, const long long* y0, int c0, float* d0, const float* logp1,
                            const long long* y1, int c1, float* d1, int n, float* loss, int* bad_label, cudaStream_t st);
cudaError_t launch_head_bwd(const HeadBwdParams& p, int precision, cudaStream_t st);
cudaError_t launch_unpack_batch(const UnpackParams* u, int n, cudaStream_t st);

}  // namespace mtldas
