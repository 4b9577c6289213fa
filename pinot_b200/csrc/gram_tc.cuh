// Tensor-core Gram matrix (gram_tc.cu): interface used by the C-ABI entry points in gram.cu.
#pragma once
#include "common.cuh"

namespace pb {

bool gram_tc_applicable(int64_t n, int64_t n2, int64_t d, bool backward);
int64_t gram_tc_workspace_floats(int64_t n, int64_t n2, int64_t d, bool same, bool backward, bool need_y_grad);
int gram_fwd_tc(int kind, const float* x, const float* y, const float* scale, int scale_stride, const float* variance,
                float* out, int64_t ldo, int64_t n, int64_t n2, int d, float diag_add, const float* diag_add_log,
                int lower_only, float* workspace, cudaStream_t stream);
int gram_bwd_tc(int kind, const float* x, const float* y, const float* scale, int scale_stride, const float* variance,
                const float* dk, int64_t ldk, int64_t n, int64_t n2, int d, float* dx, float* dy, float* dscale,
                float* dvariance, int symmetric, float* workspace, cudaStream_t stream);

}  // namespace pb
