// In-CTA factorisation of one 64x64 diagonal block of the Cholesky sweep, together with the
// transposed inverse of its factor.  Shared by the stand-alone diagonal kernel (chol.cu) and by the
// look-ahead epilogue of the trailing-update GEMMs (gemm.cu, gemm_tc.cu).
//
// The block is factored "augmented": rows 0..63 are the block A11, rows 64..127 an identity.
// Right-looking elimination of the stacked (128 x 64) matrix gives [ L11 ; L11^{-T} ] -- the same
// trick the sweep itself plays with its extension rows -- so the panel TRSM that follows becomes a
// plain GEMM with M = L11^{-T}.
//
// Mapping: thread r (r < 128) keeps row r of the stacked matrix in 64 registers.  Step j:
//   every thread publishes its (unscaled) column-j entry to shared memory; after ONE barrier it
//   reads the pivot and the column back (broadcast, 16-byte vector loads) and applies the scaled
//   rank-1 update to its remaining entries with 63-j independent FMAs.  No divergence, ~2000 FMAs
//   per thread in total: a few microseconds instead of the tens a shared-memory read-modify-write
//   formulation needs.  Only warps 0-3 take part (named barrier 1), other warps return at once.
#pragma once
#include "common.cuh"
#ifndef PB_STAMP
#define PB_STAMP(i)
#endif

namespace pb {

constexpr int DIAG_NB = 64;                 // panel width
constexpr int DIAG_LT = DIAG_NB;            // Lt[j][c] = column j of the stacked factor, c < 64 (A rows only)
constexpr int DIAG_SMEM_FLOATS = DIAG_NB * DIAG_LT;   // Lt (must be 16-byte aligned)

// Factor the nb x nb block at (c0, c0) of `src` (lower triangle read) and write L11 (zero above the
// diagonal) to wdiag and M = L11^{-T} (zero below it) to winv, both dense 64 x 64 (identity
// padding beyond nb).  smem: DIAG_SMEM_FLOATS floats.  All threads of the CTA must call this
// (blockDim.x >= 128) after a CTA-wide barrier; threads >= 128 return immediately.  `coherent` avoids the
// read-only path (the look-ahead epilogue reads values the same CTA has just written).
// Returns the first failing pivot (0-based) or -1 (valid in threads < 128).
__device__ __forceinline__ int factor_diag_block_regs(const float* src, int64_t ld, int c0, int nb, bool coherent,
                                                      float* smem, float* __restrict__ wdiag,
                                                      float* __restrict__ winv) {
    // NB `smem` must NOT be __restrict__: it carries data between threads across the barriers, and a
    // restrict-qualified pointer lets the compiler hoist those loads above __syncthreads().
    float* Lt = smem;                       // [64][64]: Lt[j][c] = L[c][j]
    const int tid = threadIdx.x;
    const bool active = tid < 2 * DIAG_NB;
    const bool is_a = tid < DIAG_NB;        // rows of A11; the rest are identity rows
    const int r = is_a ? tid : tid - DIAG_NB;

    float a[DIAG_NB];
#pragma unroll
    for (int c = 0; c < DIAG_NB; ++c) a[c] = 0.f;
    if (active) {
        if (is_a && r < nb) {
            const float* row = src + (int64_t)(c0 + r) * ld + c0;
            const bool vec = ((ld & 3) == 0) && ((reinterpret_cast<uintptr_t>(row) & 15u) == 0) && (nb == DIAG_NB);
            if (vec) {
#pragma unroll
                for (int c = 0; c < DIAG_NB; c += 4) {
                    const float4 v = coherent ? __ldcv(reinterpret_cast<const float4*>(row + c))
                                              : __ldg(reinterpret_cast<const float4*>(row + c));
                    a[c] = v.x; a[c + 1] = v.y; a[c + 2] = v.z; a[c + 3] = v.w;
                }
#pragma unroll
                for (int c = 0; c < DIAG_NB; ++c)
                    if (c > r) a[c] = 0.f;                                  // only the lower triangle is valid
            } else {
#pragma unroll
                for (int c = 0; c < DIAG_NB; ++c)
                    if (c <= r) a[c] = coherent ? __ldcv(row + c) : __ldg(row + c);
            }
        } else {
#pragma unroll
            for (int c = 0; c < DIAG_NB; ++c) a[c] = (c == r) ? 1.f : 0.f;   // identity rows / padding
        }
    }
    int fail = -1;
    if (!active) return fail;   // NB the caller may only use the return value of thread 0
#pragma unroll
    for (int j = 0; j < DIAG_NB; ++j) {
        // publish the unscaled column j (entry j of it is the pivot S[j][j])
        if (is_a) Lt[j * DIAG_LT + tid] = a[j];
        asm volatile("bar.sync 1, 128;" ::: "memory");
        float d = Lt[j * DIAG_LT + j];
        if (!(d > 0.f)) {   // also catches NaN
            if (fail < 0 && j < nb) fail = j;
            d = 1.f;
        }
        float inv;                                 // MUFU.RSQ alone: rsqrtf() wraps it in a denormal rescue (a pivot is
        asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(inv) : "f"(d));   // never denormal; same value for normal inputs)
        const float y = a[j] * inv * inv;          // S[r][j] / d
        a[j] *= inv;                               // L[r][j] (row j: sqrt(d)), or M[r][j]
        // S[r][c] -= S[r][j] S[c][j] / d for c > j; column read in aligned 16-byte pieces.  Pairs of columns go
        // through the packed FFMA2 (fma.rn.f32x2, same rounding as two FFMAs): this code is unrolled 64-fold and runs
        // once per call, at the pace of the instruction fetch path (scripts/micro/factor_time.cu) -- half the FMA
        // instructions is half of most of its bytes.
        const float2 ny2 = make_float2(-y, -y);
#pragma unroll
        for (int c4 = ((j + 1) / 4) * 4; c4 < DIAG_NB; c4 += 4) {
            const float4 l = *reinterpret_cast<const float4*>(Lt + j * DIAG_LT + c4);
            if (c4 + 0 > j) {
                const float2 t = __ffma2_rn(ny2, make_float2(l.x, l.y), make_float2(a[c4 + 0], a[c4 + 1]));
                a[c4 + 0] = t.x; a[c4 + 1] = t.y;
            } else if (c4 + 1 > j) {
                a[c4 + 1] = fmaf(-y, l.y, a[c4 + 1]);
            }
            if (c4 + 2 > j) {
                const float2 t = __ffma2_rn(ny2, make_float2(l.z, l.w), make_float2(a[c4 + 2], a[c4 + 3]));
                a[c4 + 2] = t.x; a[c4 + 3] = t.y;
            } else if (c4 + 3 > j) {
                a[c4 + 3] = fmaf(-y, l.w, a[c4 + 3]);
            }
        }
    }
    // results: thread r < 64 holds row r of L11 (entries c <= r), thread 64 + r row r of M (c >= r)
    if (active) {
        float* dst = (is_a ? wdiag : winv) + r * DIAG_NB;
#pragma unroll
        for (int c = 0; c < DIAG_NB; c += 4) {
            float4 v;
            v.x = (is_a ? (c + 0 <= r) : (c + 0 >= r)) ? a[c + 0] : 0.f;
            v.y = (is_a ? (c + 1 <= r) : (c + 1 >= r)) ? a[c + 1] : 0.f;
            v.z = (is_a ? (c + 2 <= r) : (c + 2 >= r)) ? a[c + 2] : 0.f;
            v.w = (is_a ? (c + 3 <= r) : (c + 3 >= r)) ? a[c + 3] : 0.f;
            *reinterpret_cast<float4*>(dst + c) = v;
        }
    }
    return fail;
}

// ---- fused panel solve ------------------------------------------------------------------------
constexpr int SOLVE_LD = DIAG_NB + 1;                                       // padded row of the staged X tile
constexpr int SOLVE_SMEM_FLOATS = 128 * SOLVE_LD + 12 + DIAG_NB * DIAG_NB;  // X tile (128 x 64) + M (64 x 64)

__device__ __forceinline__ void publish_flag(int* flag) {
    __threadfence();          // this thread's stores (L11 / M) are visible device-wide ...
    __syncthreads();
    if (threadIdx.x == 0) atomicExch(flag, 1);   // ... before the flag is
}
__device__ __forceinline__ void wait_flag(const int* flag) {
    // The flag is published by CTA (0,0) of the SAME grid.  Blocks are dispatched in ascending index order,
    // so that CTA is resident (or finished) whenever a waiter runs; CUDA's forward-progress model does not
    // promise it, hence the bounded spin: after ~2 s without the flag the kernel traps (a launch error the
    // host sees) instead of hanging the GPU.
    if (threadIdx.x == 0) {
        int v;
        unsigned spins = 0;
        do {
            asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(flag) : "memory");
            if (v == 0) {
                __nanosleep(64);
                if (++spins > (1u << 25)) __trap();
            }
        } while (v == 0);
    }
    __syncthreads();
}

// C[m][0:nb] <- C[m][0:nb] * M  for the rows m in [m0, m0 + 128) of C with nb <= m < rows,
// M = winv (64 x 64, upper triangular, dense).  256 threads; smem: SOLVE_SMEM_FLOATS floats.
// The rows were written by this CTA earlier in the same kernel: they are re-read through L2.
__device__ __forceinline__ void panel_solve_tile(float* C, int64_t ldc, int m0, int rows, int nb,
                                                 const float* winv, float* smem) {
    float* Xs = smem;                            // [128][SOLVE_LD]
    float* Ms = smem + 128 * SOLVE_LD + 12;      // [64][64], 16-byte aligned (128 * 65 + 12 = 8332 floats)
    const int tid = threadIdx.x;
    const bool vec = ((ldc & 3) == 0) && ((reinterpret_cast<uintptr_t>(C) & 15u) == 0);
    {
        // all loads first (they are independent), then the shared-memory stores: one L2 round trip
        float4 mreg[4], xreg[8];
#pragma unroll
        for (int i = 0; i < 4; ++i) mreg[i] = __ldcg(reinterpret_cast<const float4*>(winv) + tid + i * 256);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int idx = tid + i * 256;          // float4 index in the 128 x 16 tile
            const int r = idx >> 4, c = (idx & 15) * 4;
            const int m = m0 + r;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (m >= nb && m < rows && c < nb) {
                const float* src = C + (int64_t)m * ldc + c;
                if (vec && c + 3 < nb) {
                    v = __ldcg(reinterpret_cast<const float4*>(src));
                } else {
                    v.x = __ldcg(src);
                    if (c + 1 < nb) v.y = __ldcg(src + 1);
                    if (c + 2 < nb) v.z = __ldcg(src + 2);
                    if (c + 3 < nb) v.w = __ldcg(src + 3);
                }
            }
            xreg[i] = v;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) reinterpret_cast<float4*>(Ms)[tid + i * 256] = mreg[i];
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int idx = tid + i * 256;
            const int r = idx >> 4, c = (idx & 15) * 4;
            Xs[r * SOLVE_LD + c + 0] = xreg[i].x;
            Xs[r * SOLVE_LD + c + 1] = xreg[i].y;
            Xs[r * SOLVE_LD + c + 2] = xreg[i].z;
            Xs[r * SOLVE_LD + c + 3] = xreg[i].w;
        }
    }
    __syncthreads();
    PB_STAMP(8);
    // warp w: column group (w & 3) of 16 columns, row block (w >> 2) of 64 rows; lane l handles rows
    // l and l + 32 of that block.  Every lane of a warp reads the same M addresses (broadcast) and
    // its own conflict-free X rows: 2 scalar + 4 vector LDS per 32 FMAs.
    const int warp = tid >> 5, lane = tid & 31;
    const int cg = (warp & 3) * 16;
    const int r0 = (warp >> 2) * 64 + lane, r1 = r0 + 32;
    float acc0[16], acc1[16];
#pragma unroll
    for (int j = 0; j < 16; ++j) { acc0[j] = 0.f; acc1[j] = 0.f; }
#pragma unroll 4
    for (int k = 0; k < DIAG_NB; ++k) {
        const float x0 = Xs[r0 * SOLVE_LD + k];
        const float x1 = Xs[r1 * SOLVE_LD + k];
#pragma unroll
        for (int j4 = 0; j4 < 16; j4 += 4) {
            const float4 mv = *reinterpret_cast<const float4*>(Ms + k * DIAG_NB + cg + j4);
            acc0[j4 + 0] = fmaf(x0, mv.x, acc0[j4 + 0]);
            acc0[j4 + 1] = fmaf(x0, mv.y, acc0[j4 + 1]);
            acc0[j4 + 2] = fmaf(x0, mv.z, acc0[j4 + 2]);
            acc0[j4 + 3] = fmaf(x0, mv.w, acc0[j4 + 3]);
            acc1[j4 + 0] = fmaf(x1, mv.x, acc1[j4 + 0]);
            acc1[j4 + 1] = fmaf(x1, mv.y, acc1[j4 + 1]);
            acc1[j4 + 2] = fmaf(x1, mv.z, acc1[j4 + 2]);
            acc1[j4 + 3] = fmaf(x1, mv.w, acc1[j4 + 3]);
        }
    }
    __syncthreads();
    PB_STAMP(9);
#pragma unroll
    for (int j = 0; j < 16; ++j) {
        Xs[r0 * SOLVE_LD + cg + j] = acc0[j];
        Xs[r1 * SOLVE_LD + cg + j] = acc1[j];
    }
    __syncthreads();
    PB_STAMP(10);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int idx = tid + i * 256;
        const int rr = idx >> 4, c = (idx & 15) * 4;
        const int m = m0 + rr;
        if (m >= nb && m < rows && c < nb) {
            float* dst = C + (int64_t)m * ldc + c;
            const float4 v = make_float4(Xs[rr * SOLVE_LD + c], Xs[rr * SOLVE_LD + c + 1], Xs[rr * SOLVE_LD + c + 2],
                                         Xs[rr * SOLVE_LD + c + 3]);
            if (vec && c + 3 < nb) {
                *reinterpret_cast<float4*>(dst) = v;
            } else {
                dst[0] = v.x;
                if (c + 1 < nb) dst[1] = v.y;
                if (c + 2 < nb) dst[2] = v.z;
                if (c + 3 < nb) dst[3] = v.w;
            }
        }
    }
}

}  // namespace pb
