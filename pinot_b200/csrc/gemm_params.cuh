// Kernel-side parameter block shared by the FFMA GEMM (gemm.cu) and the tcgen05 GEMM (gemm_tc.cu).
#pragma once
#include "gemm.cuh"

namespace pb {

struct KParams {
    int M, N, K;
    const float* A; int64_t lda;
    const float* B; int64_t ldb;
    float* C; int64_t ldc;
    float alpha, beta;
    const float* alpha_dev;
    const float* bias;
    int act;
    const float* kscale;
    const float* r1u; const float* r1v; float r1coef;
    int lower, mirror, ktri, splitk;
    int vecA, vecB, vecC;
    float* post_wdiag; float* post_winv; int* post_info; int post_nb, post_c0;
    int* post_flag; int post_solve;
    int pdl;                                // launched as a programmatic dependent of the kernel before it (sweep steps)
};

// gemm_tc.cu: 3xTF32 tensor-core (tcgen05 / TMEM) implementation of the same contract.
int launch_gemm_tc(const KParams& p, bool transA, bool transB, dim3 grid, cudaStream_t stream);

// gemm_tma.cu: TMA-fed NT variant working from pre-split (hi / lo) operands in caller-provided scratch.
int64_t gemm_tma_scratch_floats(int64_t M, int64_t N, int64_t K, bool same_operand);
int launch_gemm_tma(const KParams& p, float* scratch, bool same_operand, cudaStream_t stream);
bool gemm_tma_enabled();   // tensor-core GEMMs selected and PINOT_B200_TMA != 0
// one operand (rows x K, row-major) -> hi / lo arrays with leading dimension ceil4(K); scale_mode 0 none,
// 1 = columns * colscale, 2 = columns * sqrt(colscale); tri: upper-triangular operand of a ktri product
int tma_split_operand(const float* x, int64_t ldx, int64_t rows, int64_t K, float* hi, float* lo, const float* colscale,
                      int scale_mode, int tri, cudaStream_t stream);
// tile_counter: one int of device scratch for the kernel's dynamic tile queue (zeroed here, on the stream).
// ld_a / ld_b: leading dimensions of the pre-split arrays (0 = ceil4(K): compact); a block of a larger array is passed
// as a pointer to its first element plus the array's leading dimension.  accurate: accumulators flushed into registers
// every 128 K-columns (plain alpha / beta products only).
int launch_gemm_tma_presplit(const KParams& p, const float* a_hi, const float* a_lo, const float* b_hi, const float* b_lo,
                             int* tile_counter, cudaStream_t stream, int64_t ld_a = 0, int64_t ld_b = 0, bool accurate = false);
constexpr int64_t TMA_COUNTER_FLOATS = 64;   // callers reserve 256 bytes for that counter

}  // namespace pb
