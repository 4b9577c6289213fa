// Cycles of one 64 x 64 diagonal-block factorisation (pinot_b200/csrc/chol_diag.cuh) inside a CTA of 256 threads,
// timed with clock64 by thread 0.  Build (from the repository root):
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a --expt-relaxed-constexpr -I pinot_b200/csrc \
//        -o scripts/micro/factor_time scripts/micro/factor_time.cu
// Measured on B200 (round 2): 15 823 cycles = 247 per column for the committed formulation (128 threads, one named
// barrier per column); 17 056 = 266 per column for a formulation with the pivot chain inside one warp (rows in
// registers, column vector broadcast from shared memory, next pivot by one shuffle, identity rows following through
// mbarriers) -- the 64-fold unrolled code runs once per call, straight from the instruction fetch path, and both
// come out at about the same ~6 bytes of code per cycle.
#include <cstdio>
#include <vector>
#include <cuda_runtime.h>
#include "chol_diag.cuh"
__global__ void __launch_bounds__(256) k(const float* E, float* wdiag, float* winv, long long* cyc, int* failout) {
    __shared__ __align__(16) float S[pb::DIAG_SMEM_FLOATS];
    __syncthreads();
    const long long t0 = clock64();
    const int fail = pb::factor_diag_block_regs(E, 64, 0, 64, false, S, wdiag, winv);
    __syncthreads();
    const long long t1 = clock64();
    if (threadIdx.x == 0) { cyc[0] = t1 - t0; failout[0] = fail; }
}
int main() {
    const int n = 64;
    std::vector<float> a(n * n), h(n * n);
    for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) a[i * n + j] = (float)((i * 37 + j * 11) % 17 - 8) / 16.f;
    for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) { float s = 0; for (int k = 0; k < n; ++k) s += a[i * n + k] * a[j * n + k]; h[i * n + j] = s / n + (i == j ? 1.f : 0.f); }
    float *E, *wd, *wi; long long* cyc; int* fail;
    cudaMalloc(&E, n * n * 4); cudaMalloc(&wd, n * n * 4); cudaMalloc(&wi, n * n * 4); cudaMalloc(&cyc, 8); cudaMalloc(&fail, 4);
    cudaMemcpy(E, h.data(), n * n * 4, cudaMemcpyHostToDevice);
    long long best = 1ll << 60; int f = 0;
    for (int r = 0; r < 10; ++r) {
        k<<<1, 256>>>(E, wd, wi, cyc, fail);
        cudaDeviceSynchronize();
        long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&f, fail, 4, cudaMemcpyDeviceToHost);
        if (c < best) best = c;
    }
    std::vector<float> L(n * n); cudaMemcpy(L.data(), wd, n * n * 4, cudaMemcpyDeviceToHost);
    double err = 0; for (int i = 0; i < n; ++i) for (int j = 0; j <= i; ++j) { double s = 0; for (int k = 0; k <= j; ++k) s += (double)L[i * n + k] * L[j * n + k]; err = fmax(err, fabs(s - h[i * n + j])); }
    printf("best of 10: %lld cycles (%.1f per column), fail %d, max |L L^T - A| %.2e\n", best, best / 64.0, f, err);
    return 0;
}
