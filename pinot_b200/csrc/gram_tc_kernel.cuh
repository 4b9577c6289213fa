// Device side of the tensor-core Gram matrix (gram_tc.cu): a TMA-fed, warp-specialised tcgen05 GEMM over the
// pre-split (hi / lo TF32) centred rows whose epilogue turns the dot products into kernel values (MODE 1) or
// into the backward coefficients c_mn = dK_mn k'(d_mn) / d_mn (MODE 2); MODE 3 runs the diagonal tiles of x x^T
// only and writes the squared norms (see gram_tc.cu for why they come from here).
//
//   warp 0      one thread: TMA loads of 128 x 32 tiles of x_hi, x_lo, y_hi, y_lo into a 3-stage ring (64 KB / stage)
//   warp 1      one thread: 12 tcgen05.mma.kind::tf32 per stage (lo.hi + hi.lo -> cross accumulator, hi.hi -> main)
//   warps 2-9   epilogue, two warps per TMEM lane quadrant: thread = one output row, 32 consecutive columns per
//               tcgen05.ld; nothing is staged through shared memory, the row norm is one register, results leave
//               as 32-byte stores (whole L2 sectors).  Fast intrinsics only (MUFU.RSQ / MUFU.EX2): the epilogue
//               issues ~20 instructions per element on 8 warps, which is what keeps it off the critical path.
// The accumulators are double-buffered in TMEM so that the epilogue of one tile overlaps the MMAs of the next.
#pragma once
#include "tma_common.cuh"

namespace pb {
namespace gram {

using namespace tc;

constexpr int BM = 128, BN = 128, BK = 32, NSTAGE = 3;
constexpr int TILE_BYTES = BM * BK * 4;
constexpr int STAGE_BYTES = 4 * TILE_BYTES;
constexpr int SMEM_BYTES = NSTAGE * STAGE_BYTES + 1024;
constexpr int NTHREADS = 320;                      // producer, MMA, 8 epilogue warps
constexpr int NEPI = 256;

struct GramParams {
    int M, N, K;                            // rows of x, rows of y, features (K = d)
    int mirror;                             // fwd, x is y: lower tiles only, each element also stored transposed
    int mode, kind, same, lower, trans;     // 1 fwd / 2 bwd; 0 RBF / 1 Matern52; x is y; skip tiles above the diagonal;
                                            // bwd: dK is read transposed (column-side pass)
    float* C; int64_t ldc;                  // fwd: K matrix; bwd: coefficient matrix
    const float* nx; const float* ny;       // squared norms of the centred, scaled rows
    const float* x; const float* y; int d;  // raw inputs (ld = d): exact difference-form distance of flagged pairs
    const float* es;                        // exp(scale) per column
    const float* var; const float* dlog; float diag;
    float tau;                              // pairs with d^2 < (tau (|x|^2 + |y|^2))^2 are recomputed exactly
    const float* dk; int64_t lddk;          // bwd: upstream gradient
    float* rs; float* dvar; float* fix;     // bwd: row sums of c, sum(dK * K) (may be NULL), exact near-pair terms (ld = d)
    float* dbg;                             // development aid: count of exact recomputations (may be NULL)
    float* nrm_out;                         // MODE 3: the squared norms, taken from the diagonal of x x^T
};

__device__ __forceinline__ void tmem_ld_x32(uint32_t (&v)[32], uint32_t taddr) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]),
          "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]),
          "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}

// kernel value and (dK/dd)/d from the squared distance; fast intrinsics (relative error ~1e-6, far inside 1e-4)
__device__ __forceinline__ float ex2_fast(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
template <int KIND>
__device__ __forceinline__ void kernel_terms(float amp, float d2, float& kv, float& c_over) {
    const float inv = rsqrtf(fmaxf(d2, 1e-30f));
    const float d = d2 * inv;                              // 0 at d2 = 0
    if (KIND == 0) {
        kv = amp * ex2_fast(-0.72134752f * d);             // exp(-d / 2)
        c_over = d2 > 0.f ? -0.5f * kv * inv : 0.f;        // 0 at d = 0: torch.norm's sub-gradient
    } else {
        const float r = 2.2360680f * d;                    // sqrt(5 d^2)
        const float e = amp * ex2_fast(-1.4426950f * r);
        kv = fmaf(1.6666666f, d2, 1.f + r) * e;
        c_over = d2 > 0.f ? -1.6666666f * (1.f + r) * e : 0.f;
    }
}

// The pairs whose GEMM-form distance is unusable.  An epilogue warp owns a 32-row x 64-column piece of the tile;
// while it walks that piece it only LISTS the flagged pairs in shared memory (code = column_in_piece * 32 + lane;
// the placeholder it stores for them is overwritten here) and, once the piece is done and the accumulator has been
// handed back to the MMA warp, works through the list as a warp: eight pairs per pass, the lanes splitting the
// feature axis (coalesced loads, all eight pairs' loads in flight together, one warp reduction per pair).  Nothing
// of the hot loop is live across this call.  Forward: the exact kernel value replaces the stored one (and its
// mirror image).  Backward: the coefficient c from the exact distance replaces the zero in the coefficient
// matrix; only for true near-duplicates (d below 1e-3 of the norms, where xs_m rowsum(c) - (C Ys)_m would cancel)
// does the pair's whole contribution c (xs_m - ys_n) to row m go out in difference form (coalesced atomics into
// p.fix) with its entry of the coefficient matrix left zero.
constexpr int NEAR_SLOTS = 32 * (BN / 2);           // every pair of the piece
constexpr int NEAR_BATCH = 8;
constexpr float NEAR_GEMM_OK = 1e-6f;               // d^2 above this fraction of |x|^2 + |y|^2: the backward term may use the GEMM
__device__ __forceinline__ float diff_rn(float x, float y, float e) {
    // x * e - y * e as the reference computes it (kernels/rbf.py:31-33): two rounded products, then the difference
    // (no FMA contraction), so equal inputs give exactly zero
    return __fsub_rn(__fmul_rn(x, e), __fmul_rn(y, e));
}

template <int MODE>
__device__ __noinline__ void near_flush(const GramParams& p, int m_base, int n_base, int count, const uint16_t* list,
                                        float amp, float dadd, int lane) {
    if (p.dbg && lane == 0) atomicAdd(p.dbg, (float)count);
    // d <= 128 in whole float4s (the usual 128-wide embeddings): one 16-byte load per lane covers a row, a batch is
    // a single round trip to L2, and the differences stay in registers for the backward atomics
    const bool vec = (p.d % 4 == 0) && p.d <= 128 &&
                     ((reinterpret_cast<uintptr_t>(p.x) | reinterpret_cast<uintptr_t>(p.y) | reinterpret_cast<uintptr_t>(p.es)) & 15) == 0;
    for (int i0 = 0; i0 < count; i0 += NEAR_BATCH) {
        int code[NEAR_BATCH];
        int64_t xo[NEAR_BATCH], yo[NEAR_BATCH];
        float s[NEAR_BATCH];
        float4 df[NEAR_BATCH];
#pragma unroll
        for (int j = 0; j < NEAR_BATCH; ++j) {
            code[j] = i0 + j < count ? (int)list[i0 + j] : -1;
        }
#pragma unroll
        for (int j = 0; j < NEAR_BATCH; ++j) {
            // an absent pair of the last batch re-reads pair 0, which always exists: no branch between the loads,
            // all of them are in flight before the first use
            const int c = code[j] >= 0 ? code[j] : code[0];
            xo[j] = (int64_t)(m_base + (c & 31)) * p.d;
            yo[j] = (int64_t)(n_base + (c >> 5)) * p.d;
            s[j] = 0.f;
            df[j] = make_float4(0.f, 0.f, 0.f, 0.f);
        }
        if (vec) {
            const int k = 4 * lane;
            if (k < p.d) {
                const float4 e = __ldg(reinterpret_cast<const float4*>(p.es + k));
                float4 xv[NEAR_BATCH], yv[NEAR_BATCH];
#pragma unroll
                for (int j = 0; j < NEAR_BATCH; ++j) {
                    xv[j] = __ldg(reinterpret_cast<const float4*>(p.x + xo[j] + k));
                    yv[j] = __ldg(reinterpret_cast<const float4*>(p.y + yo[j] + k));
                }
#pragma unroll
                for (int j = 0; j < NEAR_BATCH; ++j) {
                    df[j] = make_float4(diff_rn(xv[j].x, yv[j].x, e.x), diff_rn(xv[j].y, yv[j].y, e.y),
                                        diff_rn(xv[j].z, yv[j].z, e.z), diff_rn(xv[j].w, yv[j].w, e.w));
                    s[j] = fmaf(df[j].w, df[j].w, fmaf(df[j].z, df[j].z, fmaf(df[j].y, df[j].y, df[j].x * df[j].x)));
                }
            }
        } else {
            for (int k = lane; k < p.d; k += 32) {
                const float e = __ldg(p.es + k);
                float xv[NEAR_BATCH], yv[NEAR_BATCH];
#pragma unroll
                for (int j = 0; j < NEAR_BATCH; ++j) {
                    xv[j] = __ldg(p.x + xo[j] + k);
                    yv[j] = __ldg(p.y + yo[j] + k);
                }
#pragma unroll
                for (int j = 0; j < NEAR_BATCH; ++j) {
                    const float t = diff_rn(xv[j], yv[j], e);
                    s[j] = fmaf(t, t, s[j]);
                }
            }
        }
        int my_code = -1;
        float my_d2 = 0.f;
#pragma unroll
        for (int j = 0; j < NEAR_BATCH; ++j) {
            float t = s[j];                                // (reduced for absent pairs too: no branch, no harm)
#pragma unroll
            for (int off = 16; off > 0; off >>= 1) t += __shfl_xor_sync(0xffffffffu, t, off);
            s[j] = t;
            if (lane == j) {
                my_code = code[j];
                my_d2 = t;
            }
        }
        if (MODE == 1) {
            if (my_code >= 0) {                            // lane j finishes pair j
                float kv, c_over;
                if (p.kind == 0) kernel_terms<0>(amp, my_d2, kv, c_over);
                else kernel_terms<1>(amp, my_d2, kv, c_over);
                const int64_t m = m_base + (my_code & 31), n = n_base + (my_code >> 5);
                if (m == n) kv += dadd;                    // K(x, y) with y != x: a duplicate pair on the main diagonal
                p.C[m * p.ldc + n] = kv;
                if (p.mirror) p.C[n * p.ldc + m] = kv;     // only pairs below the diagonal are listed in that mode
            }
        } else {
            // upstream gradient of pair j, fetched by lane j (one round trip for the batch), then its terms
            float my_g = 0.f;
            if (my_code >= 0) {
                const int64_t m = m_base + (my_code & 31), n = n_base + (my_code >> 5);
                my_g = p.trans ? __ldg(p.dk + n * p.lddk + m) : __ldg(p.dk + m * p.lddk + n);
            }
            float kv, c_over;
            if (p.kind == 0) kernel_terms<0>(amp, my_d2, kv, c_over);
            else kernel_terms<1>(amp, my_d2, kv, c_over);
            float my_dv = my_g * kv;                       // this pair's share of sum(dK * K)
            float my_cv = my_g * c_over;
            // With the exact distance the coefficient is good; whether the pair may then ride the GEMM
            // (dxs_m = xs_m rowsum(c) - (C Ys)_m, error ~eps |xs| / d of the pair's own term) depends on d against the
            // norms: beyond 1e-3 of them it may (one store, one atomic), true near-duplicates take the
            // difference form below (d atomics each)
            if (my_code >= 0 && my_cv != 0.f) {
                const int64_t m = m_base + (my_code & 31), n = n_base + (my_code >> 5);
                if (my_d2 > NEAR_GEMM_OK * (__ldg(p.nx + m) + __ldg(p.ny + n))) {
                    p.C[m * p.ldc + n] = my_cv;
                    atomicAdd(p.rs + m, my_cv);
                    my_cv = 0.f;
                }
            }
            if (p.dvar) {
#pragma unroll
                for (int off = NEAR_BATCH / 2; off > 0; off >>= 1) my_dv += __shfl_xor_sync(0xffffffffu, my_dv, off);
                if (lane == 0 && my_dv != 0.f) atomicAdd(p.dvar, my_dv);
            }
#pragma unroll
            for (int j = 0; j < NEAR_BATCH; ++j) {
                const float cv = __shfl_sync(0xffffffffu, my_cv, j);
                if (cv != 0.f) {                           // uniform; 0 for absent pairs (their gradient was not read)
                    float* fr = p.fix + xo[j];
                    if (vec) {
                        const int k = 4 * lane;
                        if (k < p.d) {
                            atomicAdd(fr + k, cv * df[j].x);
                            atomicAdd(fr + k + 1, cv * df[j].y);
                            atomicAdd(fr + k + 2, cv * df[j].z);
                            atomicAdd(fr + k + 3, cv * df[j].w);
                        }
                    } else {
                        for (int k = lane; k < p.d; k += 32)
                            atomicAdd(fr + k, cv * diff_rn(__ldg(p.x + xo[j] + k), __ldg(p.y + yo[j] + k), __ldg(p.es + k)));
                    }
                }
            }
        }
    }
}

template <int MODE, int KIND>
__global__ void __launch_bounds__(NTHREADS, 1)
gram_tma_kernel(const __grid_constant__ GramParams p, const __grid_constant__ CUtensorMap map_xh, const __grid_constant__ CUtensorMap map_xl,
                const __grid_constant__ CUtensorMap map_yh, const __grid_constant__ CUtensorMap map_yl) {
    extern __shared__ __align__(1024) uint8_t gram_smem[];
    __shared__ __align__(8) uint64_t full_bar[NSTAGE], empty_bar[NSTAGE], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_base_holder;
    // per epilogue warp: the flagged pairs of the piece it is walking (near_flush)
    __shared__ uint16_t near_list[NEPI / 32][NEAR_SLOTS];

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    uint8_t* smem = gram_smem + ((1024u - (s32(gram_smem) & 1023u)) & 1023u);
    const int gm = (p.M + BM - 1) / BM, gn = (p.N + BN - 1) / BN;
    const int ntiles = MODE == 3 ? gm : gm * gn;      // MODE 3: the diagonal tiles only
    const int num_k = (p.K + BK - 1) / BK;
    auto live = [&](int t, int& i, int& j) {
        if (MODE == 3) {
            i = j = t;
            return true;
        }
        i = t / gn;
        j = t - i * gn;
        return !(p.lower && j > i);
    };

    if (tid == 0) {
        for (int s = 0; s < NSTAGE; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&acc_full[s], 1);
            mbar_init(&acc_empty[s], NEPI);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_xh) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_xl) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_yh) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_yl) : "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&tmem_base_holder)), "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base_holder;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            uint32_t it = 0;
            for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
                int i, j;
                if (!live(t, i, j)) continue;
                for (int kt = 0; kt < num_k; ++kt, ++it) {
                    const uint32_t s = it % NSTAGE;
                    if (it >= NSTAGE) mbar_wait(&empty_bar[s], (it / NSTAGE - 1) & 1);
                    uint8_t* st = smem + s * STAGE_BYTES;
                    mbar_expect_tx(&full_bar[s], STAGE_BYTES);
                    tma_load_2d(st, &map_xh, kt * BK, i * BM, &full_bar[s]);
                    tma_load_2d(st + TILE_BYTES, &map_xl, kt * BK, i * BM, &full_bar[s]);
                    tma_load_2d(st + 2 * TILE_BYTES, &map_yh, kt * BK, j * BN, &full_bar[s]);
                    tma_load_2d(st + 3 * TILE_BYTES, &map_yl, kt * BK, j * BN, &full_bar[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
            uint32_t it = 0, gc = 0;
            for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
                int i, j;
                if (!live(t, i, j)) continue;
                const uint32_t set = gc & 1;
                if (gc >= 2) mbar_wait(&acc_empty[set], (gc / 2 - 1) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t acc_hi = tmem_d + set * 256u, acc_x = acc_hi + 128u;
                for (int kt = 0; kt < num_k; ++kt, ++it) {
                    const uint32_t s = it % NSTAGE;
                    mbar_wait(&full_bar[s], (it / NSTAGE) & 1);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t sb = s32(smem + s * STAGE_BYTES);
                    const uint64_t d_ah = sw128_desc(sb), d_al = sw128_desc(sb + TILE_BYTES);
                    const uint64_t d_bh = sw128_desc(sb + 2 * TILE_BYTES), d_bl = sw128_desc(sb + 3 * TILE_BYTES);
#pragma unroll
                    for (int k8 = 0; k8 < BK / 8; ++k8) {
                        const uint64_t adv = (uint64_t)((k8 * 32) >> 4);
                        const uint32_t acc = (kt > 0 || k8 > 0) ? 1u : 0u;
                        mma_tf32(acc_x, d_al + adv, d_bh + adv, idesc, acc);
                        mma_tf32(acc_x, d_ah + adv, d_bl + adv, idesc, 1u);
                        mma_tf32(acc_hi, d_ah + adv, d_bh + adv, idesc, acc);
                    }
                    mma_commit(&empty_bar[s]);
                }
                mma_commit(&acc_full[set]);
                ++gc;
            }
        }
    } else {
        // ===================== epilogue warps =====================
        if constexpr (MODE == 3) {
            // |x_m|^2 := (x x^T)_mm as THIS pipeline computes it (hi.hi + hi.lo + lo.hi, accumulated by the same MMA
            // sequence): the tensor core truncates when it accumulates (measured: -4.7e-7 relative over 16 MMAs,
            // scripts/tc_accumulate_bias.py), and norms that carry the same bias as the dot products cancel it in
            // |x|^2 + |y|^2 - 2 x.y where it matters, for x close to y.
            const int q = warp & 3, half = (warp - 2) >> 2;
            uint32_t gc = 0;
            for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++gc) {
                const uint32_t set = gc & 1;
                mbar_wait(&acc_full[set], (gc / 2) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                if (half == (q >> 1)) {               // columns q*32 .. q*32+31 of the tile lie in this warp's half
                    uint32_t v[32], w[32];
                    const uint32_t taddr = tmem_d + ((uint32_t)(q * 32) << 16) + set * 256u + (uint32_t)(q * 32);
                    tmem_ld_x32(v, taddr);
                    tmem_ld_x32(w, taddr + 128u);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    float a = 0.f;
#pragma unroll
                    for (int e = 0; e < 32; ++e)
                        if (lane == e) a = __uint_as_float(v[e]) + __uint_as_float(w[e]);
                    const int m = t * BM + q * 32 + lane;
                    if (m < p.M) p.nrm_out[m] = a;
                }
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                mbar_arrive(&acc_empty[set]);
            }
        } else {
#include "gram_tc_epilogue.inc"
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(512u) : "memory");
    }
}

}  // namespace gram
}  // namespace pb
