// Epilogue shared by the tcgen05 GEMM kernels (gemm_tc.cu, gemm_tma.cu).
#pragma once
#include "gemm_params.cuh"

namespace pb {

constexpr int EPI_LD = 36;   // floats per row of a warp's 32x32 epilogue staging tile

// Writes one warp's 32 x 32 block (staged row-major in `tile`) to C.  EPI selects how much code is
// instantiated -- the epilogue runs once per tile, straight-line, and its instruction footprint is
// what it costs: 0 = alpha/beta/bias (+ lower mask); 1 = 0 + tanh; 2 = 0 + rank-1 term + mirror;
// 3 = everything incl. split-K atomics and unaligned C, as one compact scalar loop.
template <int EPI>
__device__ __forceinline__ void epilogue_rows(const KParams& p, const float* tile, float alpha, float beta, float r1c,
                                              bool last_chunk, int m0, int n0, int q, int c0, int lane) {
    const int c4 = (lane & 7) * 4;
    const int n = n0 + c0 + c4;
    if constexpr (EPI != 3) {
        // all C reads first (one L2 round trip), then 8 x {combine, store}
        float4 cold[8];
#pragma unroll
        for (int t8 = 0; t8 < 8; ++t8) {
            const int m = m0 + q * 32 + t8 * 4 + (lane >> 3);
            cold[t8] = make_float4(0.f, 0.f, 0.f, 0.f);
            // (on diagonal tiles of a `lower` update this also reads a few elements above the
            //  diagonal: harmless, they are never stored)
            if (beta != 0.f && m < p.M && n + 3 < p.N)
                cold[t8] = *reinterpret_cast<const float4*>(p.C + (int64_t)m * p.ldc + n);
        }
        float4 bv = make_float4(0.f, 0.f, 0.f, 0.f), rv = make_float4(0.f, 0.f, 0.f, 0.f);
        if (last_chunk && n + 3 < p.N) {
            if (p.bias) bv = make_float4(__ldg(p.bias + n), __ldg(p.bias + n + 1), __ldg(p.bias + n + 2), __ldg(p.bias + n + 3));
            if (EPI == 2 && p.r1u) rv = make_float4(__ldg(p.r1v + n), __ldg(p.r1v + n + 1), __ldg(p.r1v + n + 2), __ldg(p.r1v + n + 3));
        }
#pragma unroll
        for (int t8 = 0; t8 < 8; ++t8) {
            const int rl = t8 * 4 + (lane >> 3);
            const int m = m0 + q * 32 + rl;
            if (m >= p.M || n >= p.N) continue;
            const float4 acc = *reinterpret_cast<const float4*>(tile + rl * EPI_LD + c4);
            float* cp = p.C + (int64_t)m * p.ldc + n;
            if (n + 3 < p.N) {
                float4 o;
                o.x = fmaf(alpha, acc.x, fmaf(beta, cold[t8].x, bv.x));
                o.y = fmaf(alpha, acc.y, fmaf(beta, cold[t8].y, bv.y));
                o.z = fmaf(alpha, acc.z, fmaf(beta, cold[t8].z, bv.z));
                o.w = fmaf(alpha, acc.w, fmaf(beta, cold[t8].w, bv.w));
                if (EPI == 2 && last_chunk && p.r1u) {
                    const float um = __ldg(p.r1u + m) * r1c;
                    o.x = fmaf(um, rv.x, o.x); o.y = fmaf(um, rv.y, o.y);
                    o.z = fmaf(um, rv.z, o.z); o.w = fmaf(um, rv.w, o.w);
                }
                if (EPI == 1 && last_chunk && p.act == 1) {
                    o.x = tanhf(o.x); o.y = tanhf(o.y); o.z = tanhf(o.z); o.w = tanhf(o.w);
                }
                if (EPI == 2 && last_chunk && p.mirror) {
                    if (n + 0 < m) p.C[(int64_t)(n + 0) * p.ldc + m] = o.x;
                    if (n + 1 < m) p.C[(int64_t)(n + 1) * p.ldc + m] = o.y;
                    if (n + 2 < m) p.C[(int64_t)(n + 2) * p.ldc + m] = o.z;
                    if (n + 3 < m) p.C[(int64_t)(n + 3) * p.ldc + m] = o.w;
                }
                if (!p.lower || n + 3 <= m) {
                    *reinterpret_cast<float4*>(cp) = o;
                } else {
                    if (n + 0 <= m) cp[0] = o.x;
                    if (n + 1 <= m) cp[1] = o.y;
                    if (n + 2 <= m) cp[2] = o.z;
                }
            } else {
                const float a4[4] = {acc.x, acc.y, acc.z, acc.w};
#pragma unroll 1
                for (int j = 0; j < 4; ++j) {
                    if (n + j < p.N && (!p.lower || n + j <= m)) {
                        float o = alpha * a4[j];
                        if (beta != 0.f) o = fmaf(beta, cp[j], o);
                        if (last_chunk) {
                            if (p.bias) o += __ldg(p.bias + n + j);
                            if (EPI == 2 && p.r1u) o = fmaf(__ldg(p.r1u + m) * r1c, __ldg(p.r1v + n + j), o);
                            if (EPI == 1 && p.act == 1) o = tanhf(o);
                        }
                        cp[j] = o;
                        if (EPI == 2 && last_chunk && p.mirror && n + j < m) p.C[(int64_t)(n + j) * p.ldc + m] = o;
                    }
                }
            }
        }
    } else {
#pragma unroll 1
        for (int t8 = 0; t8 < 8; ++t8) {
            const int rl = t8 * 4 + (lane >> 3);
            const int m = m0 + q * 32 + rl;
            if (m >= p.M || n >= p.N) continue;
            const float4 acc = *reinterpret_cast<const float4*>(tile + rl * EPI_LD + c4);
            const float a4[4] = {acc.x, acc.y, acc.z, acc.w};
            float* cp = p.C + (int64_t)m * p.ldc + n;
            if (p.splitk > 1 && p.vecC && n + 3 < p.N && (!p.lower || n + 3 <= m)) {
                // split-K partial sums: one 16-byte vector reduction instead of four scalar atomics
                atomicAdd(reinterpret_cast<float4*>(cp), make_float4(alpha * a4[0], alpha * a4[1], alpha * a4[2], alpha * a4[3]));
                continue;
            }
            const float um = (last_chunk && p.r1u) ? __ldg(p.r1u + m) * r1c : 0.f;
#pragma unroll 1
            for (int j = 0; j < 4; ++j) {
                if (n + j >= p.N || (p.lower && n + j > m)) continue;
                float o = alpha * a4[j];
                if (p.splitk > 1) {
                    atomicAdd(cp + j, o);
                    continue;
                }
                if (beta != 0.f) o = fmaf(beta, cp[j], o);
                if (last_chunk) {
                    if (p.bias) o += __ldg(p.bias + n + j);
                    if (p.r1u) o = fmaf(um, __ldg(p.r1v + n + j), o);
                    if (p.act == 1) o = tanhf(o);
                }
                cp[j] = o;
                if (p.mirror && last_chunk && n + j < m) p.C[(int64_t)(n + j) * p.ldc + m] = o;
            }
        }
    }
}

}  // namespace pb
