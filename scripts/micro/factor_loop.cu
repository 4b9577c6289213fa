// Experiment: the 64 x 64 diagonal-block factor as a LOOP (code that stays in the instruction cache) instead of
// 64-fold unrolled code.  Rows live in registers and are ROTATED by the rank-1 update itself (a[i] <- a[i+1] - y l[i]),
// so that "my entry of the current column" is always a[0] and the body is the same for every column.
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -o scripts/micro/factor_loop scripts/micro/factor_loop.cu
// Measured on B200 (round 2), cycles per column of the loop alone, against 247 for the committed unrolled version
// (whole call, scripts/micro/factor_time.cu):
//   one bar.sync per column, publish at the top of the body                    391
//   split phase: publish + mbarrier arrive (128 arrivals) before the bulk      562
//   the same with one arrival per warp                                         617   (this file)
// A loop does not beat the unrolled code: its body has to update the full width of the phase for every column, the
// next column's publish sits behind that in program order (or behind an mbarrier round trip of several hundred
// cycles), and the branches around the predicated stores add their own latency.  Kept as a record.
#include <cstdio>
#include <cstdint>
#include <cmath>
#include <vector>
#include <cuda_runtime.h>

constexpr int NB = 64;
constexpr int OUT_LD = NB + 1;
constexpr int COL_LD = 80;                 // one published column: [0..63] relative entries, [64] the pivot

__device__ __forceinline__ void bar_init(uint64_t* bar, int count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void bar_arrive(uint64_t* bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"((uint32_t)__cvta_generic_to_shared(bar)) : "memory");
}
__device__ __forceinline__ void bar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "W: mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@!p bra W;\n"
        "}\n" ::"r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(parity)
        : "memory");
}

// my entry of column jn goes into buffer jn % 3 (relative position = row - jn - 1; the pivot row into slot NB)
__device__ __forceinline__ void publish(float v, int jn, int r, bool is_a, float* col, uint64_t* bars) {
    if (is_a) {
        float* nbuf = col + (jn % 3) * COL_LD;
        const int rel = r - jn - 1;
        if (rel >= 0) nbuf[rel] = v;
        if (rel == -1) nbuf[NB] = v;
    }
    __syncwarp();                                   // the warp's stores, then ONE arrival per warp (an arrival is a
    if ((threadIdx.x & 31) == 0) bar_arrive(bars + jn % 3);   // shared-memory atomic: 128 of them serialise)
}

template <int W>   // W = number of trailing entries updated per column in this phase (multiple of 4, <= 64)
__device__ __forceinline__ void phase(float (&a)[NB], int j0, int j1, int r, bool is_a, float* col, uint64_t* bars, float* out, int& fail) {
#pragma unroll 1
    for (int j = j0; j < j1; ++j) {
        const float* buf = col + (j % 3) * COL_LD;
        bar_wait(bars + j % 3, (uint32_t)((j / 3) & 1));
        float d = buf[NB];
        if (!(d > 0.f)) { if (fail < 0) fail = j; d = 1.f; }
        const float inv = rsqrtf(d);
        const float mine = a[0] * inv;
        const bool live = is_a ? (r >= j) : (j >= r);
        const float y = (is_a ? (r > j) : true) ? mine * inv : 0.f;
        {
            const float4 l = *reinterpret_cast<const float4*>(buf);
            a[0] = fmaf(-y, l.x, a[1]);
            a[1] = fmaf(-y, l.y, a[2]);
            a[2] = fmaf(-y, l.z, a[3]);
            a[3] = fmaf(-y, l.w, a[4]);
        }
        if (j + 1 < NB) publish(a[0], j + 1, r, is_a, col, bars);      // the next column is on its way before the bulk
        out[r * OUT_LD + j] = live ? mine : 0.f;
#pragma unroll
        for (int i = 4; i < W; i += 4) {
            const float4 l = *reinterpret_cast<const float4*>(buf + i);
            a[i + 0] = fmaf(-y, l.x, a[i + 1]);
            a[i + 1] = fmaf(-y, l.y, a[i + 2]);
            a[i + 2] = fmaf(-y, l.z, a[i + 3]);
            a[i + 3] = (i + 4 < NB) ? fmaf(-y, l.w, a[(i + 4) & (NB - 1)]) : 0.f;
        }
    }
}

__global__ void __launch_bounds__(256) k(const float* E, float* wdiag, float* winv, long long* cyc, int* failout) {
    __shared__ __align__(16) float col[3 * COL_LD];
    __shared__ __align__(8) uint64_t bars[3];
    __shared__ float outs[2 * NB * OUT_LD];
    const int tid = threadIdx.x;
    if (tid < 3) bar_init(bars + tid, 4);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    __syncthreads();
    const long long t0 = clock64();
    int fail = -1;
    if (tid < 2 * NB) {
        const bool is_a = tid < NB;
        const int r = is_a ? tid : tid - NB;
        float a[NB];
#pragma unroll
        for (int c = 0; c < NB; c += 4) {
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (is_a && c <= r) v = __ldg(reinterpret_cast<const float4*>(E + r * NB + c));
            a[c + 0] = is_a ? ((c + 0 <= r) ? v.x : 0.f) : ((c + 0 == r) ? 1.f : 0.f);
            a[c + 1] = is_a ? ((c + 1 <= r) ? v.y : 0.f) : ((c + 1 == r) ? 1.f : 0.f);
            a[c + 2] = is_a ? ((c + 2 <= r) ? v.z : 0.f) : ((c + 2 == r) ? 1.f : 0.f);
            a[c + 3] = is_a ? ((c + 3 <= r) ? v.w : 0.f) : ((c + 3 == r) ? 1.f : 0.f);
        }
        // published columns beyond the rows that exist must read as zero
        for (int i = tid; i < 3 * COL_LD; i += 128) col[i] = 0.f;
        asm volatile("bar.sync 1, 128;" ::: "memory");
        float* out = outs + (is_a ? 0 : NB * OUT_LD);
        const long long ta = clock64();
        publish(a[0], 0, r, is_a, col, bars);
        phase<64>(a, 0, 16, r, is_a, col, bars, out, fail);
        phase<48>(a, 16, 32, r, is_a, col, bars, out, fail);
        phase<32>(a, 32, 48, r, is_a, col, bars, out, fail);
        phase<16>(a, 48, 64, r, is_a, col, bars, out, fail);
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (tid == 0) cyc[1] = clock64() - ta;
        float* dst = is_a ? wdiag : winv;
        for (int i = tid & 63; i < NB * NB; i += 64) dst[i] = out[(i >> 6) * OUT_LD + (i & 63)];
    }
    __syncthreads();
    const long long t1 = clock64();
    if (tid == 0) { cyc[0] = t1 - t0; failout[0] = fail; }
}

int main() {
    const int n = NB;
    std::vector<float> a(n * n), h(n * n);
    for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) a[i * n + j] = (float)((i * 37 + j * 11) % 17 - 8) / 16.f;
    for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) { float s = 0; for (int k = 0; k < n; ++k) s += a[i * n + k] * a[j * n + k]; h[i * n + j] = s / n + (i == j ? 1.f : 0.f); }
    float *E, *wd, *wi; long long* cyc; int* fail;
    cudaMalloc(&E, n * n * 4); cudaMalloc(&wd, n * n * 4); cudaMalloc(&wi, n * n * 4); cudaMalloc(&cyc, 16); cudaMalloc(&fail, 4);
    cudaMemcpy(E, h.data(), n * n * 4, cudaMemcpyHostToDevice);
    long long best = 1ll << 60, bestloop = 1ll << 60; int f = 0;
    for (int r = 0; r < 10; ++r) {
        k<<<1, 256>>>(E, wd, wi, cyc, fail);
        cudaDeviceSynchronize();
        long long c; cudaMemcpy(&c, cyc, 8, cudaMemcpyDeviceToHost); cudaMemcpy(&f, fail, 4, cudaMemcpyDeviceToHost);
        if (c < best) best = c;
        long long c2; cudaMemcpy(&c2, cyc + 1, 8, cudaMemcpyDeviceToHost); if (c2 < bestloop) bestloop = c2;
    }
    std::vector<float> L(n * n), M(n * n);
    cudaMemcpy(L.data(), wd, n * n * 4, cudaMemcpyDeviceToHost); cudaMemcpy(M.data(), wi, n * n * 4, cudaMemcpyDeviceToHost);
    double err = 0, errm = 0, up = 0;
    for (int i = 0; i < n; ++i) for (int j = 0; j < n; ++j) {
        if (j <= i) { double s = 0; for (int k = 0; k <= j; ++k) s += (double)L[i * n + k] * L[j * n + k]; err = fmax(err, fabs(s - h[i * n + j])); }
        else up = fmax(up, fabs(L[i * n + j]));
        double s = 0; for (int k = 0; k < n; ++k) s += (double)L[i * n + k] * M[j * n + k];     // L M^T = I
        errm = fmax(errm, fabs(s - (i == j ? 1.0 : 0.0)));
    }
    printf("loop only: %lld cycles (%.1f per column)\n", bestloop, bestloop / 64.0);
    printf("loop factor: best of 10 %lld cycles (%.1f per column), fail %d, |L L^T - A| %.2e, |L M^T - I| %.2e, upper(L) %.1e\n", best, best / 64.0, f, err, errm, up);
    return 0;
}
