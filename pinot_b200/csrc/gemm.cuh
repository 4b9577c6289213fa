// Internal fp32 GEMM used by the node-update layers (kernel 2) and by the trailing updates of the
// blocked Cholesky sweep (kernel 4).  Row-major everywhere.
#pragma once
#include "common.cuh"

namespace pb {

struct GemmParams {
    int M = 0, N = 0, K = 0;
    bool transA = false;  // false: A is (M,K) row-major; true: A is stored (K,M)
    bool transB = false;  // false: B is (K,N) row-major; true: B is stored (N,K)
    const float* A = nullptr; int64_t lda = 0;
    const float* B = nullptr; int64_t ldb = 0;
    float* C = nullptr;       int64_t ldc = 0;
    float alpha = 1.f, beta = 0.f;
    const float* alpha_dev = nullptr;  // optional device scalar multiplied into alpha and r1coef
    const float* bias = nullptr;       // optional [N], added before the activation
    int act = 0;                       // 0 none, 1 tanh
    const float* kscale = nullptr;     // optional [K]: op(A)[m,k] *= kscale[k]
    const float* r1u = nullptr;        // optional rank-1 epilogue: C += r1coef * r1u[m] * r1v[n]
    const float* r1v = nullptr;
    float r1coef = 0.f;
    int lower = 0;   // 1: square-origin C, only elements with n <= m are computed/written
    int mirror = 0;  // with lower: element (m,n) is also written to (n,m)
    int ktri = 0;    // 1: op(A)[m,k] == 0 for k < m (and op(B)[k,n] == 0 for k < n): skip k < tile row origin
    int splitk = 1;  // >1: K is split over grid.z and partial sums are atomically added into C
                     //     (beta must be 0 or 1; for 0 the caller's C is zeroed by gemm())
    // Cholesky look-ahead: after storing its tile, the CTA that owns tile (0, 0) factors the
    // post_nb x post_nb block at C's origin (the next diagonal block of the sweep) and writes
    // L11 / L11^{-T} to post_wdiag / post_winv (64 x 64 floats each); a non-positive pivot sets
    // *post_info to post_c0 + pivot + 1 (if still 0).
    float* post_wdiag = nullptr;
    float* post_winv = nullptr;
    int* post_info = nullptr;
    int post_nb = 0;
    int post_c0 = 0;
    // Fused panel solve (tensor-core kernel only): once that factor is published through
    // *post_flag, every CTA of tile column 0 multiplies its rows of the next panel (columns
    // [0, post_nb) of C, rows >= post_nb) by L11^{-T} in place, which replaces the separate
    // panel-TRSM launch of the next sweep step.
    int* post_flag = nullptr;
    int post_solve = 0;
    int pdl = 0;                // the kernel before this one in the stream is the step before it in a sweep: launch as its programmatic dependent
};

// True when gemm() dispatches to the tcgen05 kernel (PINOT_B200_GEMM != "ffma").
bool gemm_uses_tensor_cores();

// Launches the GEMM on `stream`.  Returns PB_OK or an error code (message via set_error).
int gemm(const GemmParams& p, cudaStream_t stream);

// gemm_node.cu: persistent TMA-fed kernel for M >> N, K with N, K <= 128 (node-update layers); gemm() routes
// to it when applicable (PINOT_B200_NODE_GEMM=0 switches it off).
bool node_gemm_applicable(const GemmParams& g);
int launch_node_gemm(const GemmParams& g, cudaStream_t stream);
bool gemm_tma_enabled();

}  // namespace pb
