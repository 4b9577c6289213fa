// Issue rate of SHFL.IDX inside ONE warp.  Written the obvious way -- shuffle, use, shuffle, use -- ptxas gives
// every result the same register and the pairs serialise at the shuffle latency; issued in batches of B shuffles
// before their B uses, they pipeline.  nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o shfl_rate shfl_rate.cu
#include <cstdio>
#include <cuda_runtime.h>
template <int B>
__global__ void k(float* out, long long* cyc, int reps) {
    float a[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) a[i] = threadIdx.x * 0.001f + i;
    float l = threadIdx.x * 0.5f + 1.f;
    __syncthreads();
    long long t0 = clock64();
    for (int r = 0; r < reps; ++r) {
#pragma unroll
        for (int c0 = 0; c0 < 32; c0 += B) {
            float lc[B];
#pragma unroll
            for (int i = 0; i < B; ++i) lc[i] = __shfl_sync(0xffffffffu, l, c0 + i);
#pragma unroll
            for (int i = 0; i < B; ++i) a[c0 + i] = fmaf(-l, lc[i], a[c0 + i]);
        }
        l += 1e-6f;
    }
    long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) s += a[i];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if ((threadIdx.x & 31) == 0) cyc[threadIdx.x / 32] = t1 - t0;
}
template <int B>
void run(float* out, long long* cyc) {
    for (int warps : {1, 4}) {
        k<B><<<1, 32 * warps>>>(out, cyc, 8);
        cudaDeviceSynchronize();
        k<B><<<1, 32 * warps>>>(out, cyc, 64);
        cudaDeviceSynchronize();
        long long h[4]; cudaMemcpy(h, cyc, 8 * warps, cudaMemcpyDeviceToHost);
        printf("batch %2d, %d warps per CTA: %.2f cycles per (SHFL + FFMA) pair\n", B, warps, (double)h[0] / (64.0 * 32.0));
    }
}
int main() {
    float* out; long long* cyc;
    cudaMalloc(&out, 1024 * 4); cudaMalloc(&cyc, 64 * 8);
    run<1>(out, cyc); run<4>(out, cyc); run<8>(out, cyc); run<16>(out, cyc); run<32>(out, cyc);
    return 0;
}
