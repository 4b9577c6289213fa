// Tensor-core implementation of the library's fp32 GEMM contract: tcgen05.mma (kind::tf32) with the
// accumulator in TMEM, fp32 accuracy recovered by the 3xTF32 split
//     a = a_hi + a_lo,  b = b_hi + b_lo   (each part exactly representable in TF32)
//     a.b ~= a_hi.b_hi + a_hi.b_lo + a_lo.b_hi        (the dropped a_lo.b_lo term is ~2^-22 |a||b|)
// so that the embedding / Cholesky parity bars (1e-4 relative) keep holding where a single TF32 or
// BF16 pass would not.
//
// One CTA (256 threads) owns a 128 x 128 output tile:
//   * all 8 warps stream op(A) / op(B) K-slabs (16 columns at a time) from global memory, split them
//     and store hi / lo parts to shared memory in the canonical K-major, un-swizzled UMMA layout
//     (8-row x 16-byte core matrices; SBO = 128 B between 8-row groups, LBO = 2080 B between
//     K-adjacent core matrices -- the 32-byte pad keeps the coalesced stores bank-conflict free);
//   * one thread issues the tcgen05.mma instructions (M = 128, N <= 128, K = 8 each, 3 per K-step)
//     and commits them to an mbarrier per pipeline stage, which is what releases the stage's shared
//     memory for the next slab (3-stage ring, loads of slab k+1 overlap the MMAs of slab k);
//   * fp32 accumulation in TMEM truncates, so the error grows with the number of MMAs: the small cross
//     terms get their own accumulator (TMEM columns 128-255) and both are flushed into C every 512
//     K-columns, which bounds the accumulation error independently of K;
//   * the epilogue reads the accumulator back with tcgen05.ld (each warp its 32 TMEM lanes, i.e. 32
//     rows; warps 0-3 the left 64 columns, warps 4-7 the right ones), transposes 32x32 blocks through
//     shared memory so that C is read and written in row-contiguous 128-byte pieces, and applies the same fused
//     epilogue as the FFMA kernel (alpha/beta, bias, tanh, rank-1 term, lower/mirror masks,
//     split-K atomics, Cholesky look-ahead).
#include "gemm_params.cuh"
#include <stdlib.h>

namespace pb {

#ifdef PB_DEBUG_TIMING
__device__ unsigned long long g_dbg_times[16];
// per sweep step (index = first column of the block being factored / 64): the same stamps of tile (0,0), plus
// stamp 15 = the last tile-column-0 CTA to finish its panel solve (atomicMax).  Use with PINOT_B200_PDL=0: a
// dependent launch starts before the step before it has finished and would overwrite the step index.
__device__ unsigned long long g_dbg_steps[512][16];
__device__ int g_dbg_step_index;
#define PB_STAMP(i) do { if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && threadIdx.x == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); g_dbg_times[i] = t_; if (g_dbg_step_index >= 0 && g_dbg_step_index < 512) g_dbg_steps[g_dbg_step_index][i] = t_; } } while (0)
#define PB_STAMP_COL0_DONE() do { if (threadIdx.x == 0 && g_dbg_step_index >= 0 && g_dbg_step_index < 512) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_)); atomicMax(&g_dbg_steps[g_dbg_step_index][15], t_); } } while (0)
#else
#define PB_STAMP(i)
#define PB_STAMP_COL0_DONE()
#endif
}  // namespace pb
#include "chol_diag.cuh"
#include "gemm_epilogue.cuh"
namespace pb {

namespace {

constexpr int TM = 128, TK = 16;               // CTA tile rows; TK fp32 columns per pipeline stage
constexpr int STAGES = 3;
constexpr int NTHREADS = 256;
constexpr int SBO_BYTES = 128;                 // between 8-row groups
constexpr int KC_PER_STAGE = TK / 4;           // 16-byte K chunks per stage
constexpr int KCHUNK = 512;                    // accumulator is flushed to C every KCHUNK columns
constexpr int DIAG_BYTES = (DIAG_SMEM_FLOATS > SOLVE_SMEM_FLOATS ? DIAG_SMEM_FLOATS : SOLVE_SMEM_FLOATS) * 4;

// Geometry of a CTA tile of 128 x BN outputs (BN = 128: two CTAs per SM; BN = 256: one CTA per SM, all
// 512 TMEM columns, half the shared-memory operand traffic per flop -- for large products).
template <int BN>
struct Geo {
    static constexpr int LBO_A = (TM / 8) * 128 + 32;   // between K-adjacent core matrices (+32 B pad)
    static constexpr int LBO_B = (BN / 8) * 128 + 32;
    static constexpr int OP_A = KC_PER_STAGE * LBO_A;    // one of A_hi, A_lo
    static constexpr int OP_B = KC_PER_STAGE * LBO_B;    // one of B_hi, B_lo
    static constexpr int STAGE = 2 * OP_A + 2 * OP_B;
    static constexpr int PIPE = STAGES * STAGE;
    // tensor-core panel solve (look-ahead kernel): X tile 128 x 64 and M^T 64 x 64, hi | lo, whole K resident
    static constexpr int SOLVE_LBO_B = (DIAG_NB / 8) * 128 + 32;
    static constexpr int SOLVE_A = (DIAG_NB / 4) * LBO_A;          // one of X_hi, X_lo
    static constexpr int SOLVE_B = (DIAG_NB / 4) * SOLVE_LBO_B;    // one of M_hi, M_lo
    static constexpr int SOLVE_TC = 2 * SOLVE_A + 2 * SOLVE_B;
    static constexpr int NEED0 = PIPE > DIAG_BYTES ? PIPE : DIAG_BYTES;
    static constexpr int DYN_SMEM = (NEED0 > SOLVE_TC ? NEED0 : SOLVE_TC) + 128;
    static constexpr int TMEM_COLS = 2 * BN;             // [0,BN): hi.hi accumulator, [BN,2BN): cross terms
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_LOOP:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE;\n\t"
        "bra WAIT_LOOP;\n\t"
        "DONE:\n\t"
        "}\n" ::"r"(addr), "r"(parity)
        : "memory");
}

// 64-bit shared-memory matrix descriptor: K-major, SWIZZLE_NONE, version 1 (sm_100).
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, int lbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);                    // start address      bits [0,14)
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;          // leading byte off.  bits [16,30)
    d |= (uint64_t)((SBO_BYTES >> 4) & 0x3FFF) << 32;          // stride byte off.   bits [32,46)
    d |= (uint64_t)1 << 46;                                    // descriptor version bits [46,48)
    return d;                                                  // layout type 0 = no swizzle
}

// Instruction descriptor: D = F32, A = B = TF32, both K-major, M = 128, N = n.
__device__ __forceinline__ uint32_t make_idesc(int n) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(TM >> 4) << 24);
}

__device__ __forceinline__ void mma_tf32(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "setp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t"
        "}\n" ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
        : "memory");
}
__device__ __forceinline__ void mma_commit(uint64_t* bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
                 : "memory");
}

__device__ __forceinline__ void split_tf32(float x, float& hi, float& lo) {
    uint32_t h, l;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(h) : "f"(x));
    hi = __uint_as_float(h);
    const float r = x - hi;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(l) : "f"(r));
    lo = __uint_as_float(l);
}

// ---- operand staging -------------------------------------------------------------------------
// Register image of one K-slab of an operand tile of 128 * H rows (H = 1 or 2): 2 * H float4 per
// thread.  The two layouts below are applied to each 128-row half in turn (images [2h], [2h+1]).
//   k-contiguous storage ([row][k]):   thread -> row = tid/4 + 64*i, kc = tid%4   (float4 along k)
//   row-contiguous storage ([k][row]): thread -> kc = tid/64, row pair rp = tid%64; four float2
//     loads (k = 4 kc + i, rows 2 rp, 2 rp + 1) are transposed in registers into one float4 per
//     row, so the shared-memory stores stay 16-byte vectors along k.
template <int H>
__device__ __forceinline__ void tc_load_kcontig(const float* __restrict__ P, int64_t ld, int rows, int kend,
                                                int row0, int k0, int vec, const float* __restrict__ kscale,
                                                float4 (&r)[2 * H]) {
    const int tid = threadIdx.x;
    const int k = k0 + (tid & 3) * 4;
#pragma unroll
    for (int i = 0; i < 2 * H; ++i) {
        const int row = row0 + (tid >> 2) + 64 * i;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row < rows && k < kend) {
            const float* src = P + (int64_t)row * ld + k;
            if (vec && k + 3 < kend) {
                v = ldg4(src);
            } else {
                v.x = __ldg(src);
                if (k + 1 < kend) v.y = __ldg(src + 1);
                if (k + 2 < kend) v.z = __ldg(src + 2);
                if (k + 3 < kend) v.w = __ldg(src + 3);
            }
            if (kscale) {
                v.x *= __ldg(kscale + k);
                if (k + 1 < kend) v.y *= __ldg(kscale + k + 1);
                if (k + 2 < kend) v.z *= __ldg(kscale + k + 2);
                if (k + 3 < kend) v.w *= __ldg(kscale + k + 3);
            }
        }
        r[i] = v;
    }
}
template <int H>
__device__ __forceinline__ void tc_store_kcontig(uint8_t* hi, uint8_t* lo, int lbo, const float4 (&r)[2 * H]) {
    const int tid = threadIdx.x;
    const int kc = tid & 3;
#pragma unroll
    for (int i = 0; i < 2 * H; ++i) {
        const int row = (tid >> 2) + 64 * i;
        float4 h, l;
        split_tf32(r[i].x, h.x, l.x);
        split_tf32(r[i].y, h.y, l.y);
        split_tf32(r[i].z, h.z, l.z);
        split_tf32(r[i].w, h.w, l.w);
        const int off = kc * lbo + row * 16;   // (row/8)*128 + (row%8)*16 == row*16
        *reinterpret_cast<float4*>(hi + off) = h;
        *reinterpret_cast<float4*>(lo + off) = l;
    }
}

template <int H>
__device__ __forceinline__ void tc_load_rowcontig(const float* __restrict__ P, int64_t ld, int rows, int kend,
                                                  int row0, int k0, int vec, const float* __restrict__ kscale,
                                                  float4 (&r)[2 * H]) {
    const int tid = threadIdx.x;
    const int kbase = k0 + (tid >> 6) * 4;
#pragma unroll
    for (int h = 0; h < H; ++h) {
        const int row = row0 + 128 * h + (tid & 63) * 2;
        float a[4], b[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int k = kbase + i;
            float x = 0.f, y = 0.f;
            if (k < kend && row < rows) {
                const float* src = P + (int64_t)k * ld + row;
                if (vec && row + 1 < rows) {
                    const float2 t = __ldg(reinterpret_cast<const float2*>(src));
                    x = t.x; y = t.y;
                } else {
                    x = __ldg(src);
                    if (row + 1 < rows) y = __ldg(src + 1);
                }
                if (kscale) {
                    const float s = __ldg(kscale + k);
                    x *= s; y *= s;
                }
            }
            a[i] = x; b[i] = y;
        }
        r[2 * h] = make_float4(a[0], a[1], a[2], a[3]);
        r[2 * h + 1] = make_float4(b[0], b[1], b[2], b[3]);
    }
}
template <int H>
__device__ __forceinline__ void tc_store_rowcontig(uint8_t* hi, uint8_t* lo, int lbo, const float4 (&r)[2 * H]) {
    const int tid = threadIdx.x;
    const int kc = tid >> 6;
#pragma unroll
    for (int i = 0; i < 2 * H; ++i) {
        const int row = 128 * (i >> 1) + (tid & 63) * 2 + (i & 1);
        float4 h, l;
        split_tf32(r[i].x, h.x, l.x);
        split_tf32(r[i].y, h.y, l.y);
        split_tf32(r[i].z, h.z, l.z);
        split_tf32(r[i].w, h.w, l.w);
        const int off = kc * lbo + row * 16;
        *reinterpret_cast<float4*>(hi + off) = h;
        *reinterpret_cast<float4*>(lo + off) = l;
    }
}

template <bool TA, bool TB, int EPI, bool POST, int BN>
__global__ void __launch_bounds__(NTHREADS, BN == 128 ? 2 : 1) tgemm_kernel(const KParams p) {
    using G = Geo<BN>;
    constexpr int HB = BN / 128;
    extern __shared__ __align__(128) uint8_t dyn_smem[];
    __shared__ __align__(8) uint64_t mma_bar[STAGES];
    __shared__ __align__(8) uint64_t solve_bar;       // look-ahead kernel: completion of the panel-solve MMAs
    __shared__ uint32_t tmem_base_holder;

    // Sweep steps are launched as programmatic dependents of the step before them: every CTA lets the next step's
    // grid be scheduled at once (it takes SMs as this grid drains and does its set-up there), and no CTA touches
    // global memory before the grid before it has completed (griddepcontrol.wait below).  Both are no-ops in a
    // plain launch.
    if (POST) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    const int m0 = blockIdx.y * TM, n0 = blockIdx.x * BN;
    if (p.lower && n0 > m0 + TM - 1) return;

    int kb = 0, ke = p.K;
    if (p.ktri) kb = (m0 / TK) * TK;
    if (p.splitk > 1) {
        const int ktiles = (ke - kb + TK - 1) / TK;
        const int per = (ktiles + p.splitk - 1) / p.splitk;
        const int b = kb + (int)blockIdx.z * per * TK;
        ke = min(ke, b + per * TK);
        kb = b;
        if (kb >= ke) return;
    }
    const int num_k = (ke > kb) ? (ke - kb + TK - 1) / TK : 0;

#ifdef PB_DEBUG_TIMING
    if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && threadIdx.x == 0) g_dbg_step_index = (POST && p.post_wdiag) ? p.post_c0 / 64 : -1;
#endif
    PB_STAMP(0);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    // aligned by OFFSET, not by rounding the pointer through an integer: an integer round trip drops the shared
    // address space and every access to the operand / staging buffers becomes a generic LD / ST instead of LDS / STS
    uint8_t* smem = dyn_smem + ((128u - (smem_u32(dyn_smem) & 127u)) & 127u);

    // ---- one-time setup: mbarriers + TMEM allocation ----
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < STAGES; ++s) mbar_init(&mma_bar[s], 1);
        if (POST) mbar_init(&solve_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_holder)),
                     "r"((uint32_t)G::TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base_holder;
    if (POST) asm volatile("griddepcontrol.wait;" ::: "memory");
    PB_STAMP(1);

    // MMA N: the valid columns of this tile rounded up to a legal UMMA N (multiple of 16)
    const int n_valid = min(BN, p.N - n0);
    const int n_mma = min(BN, ((n_valid + 15) / 16) * 16);
    const uint32_t idesc = make_idesc(n_mma);
    const uint64_t desc_a = make_desc(0, G::LBO_A), desc_b = make_desc(0, G::LBO_B);

    // two register images per operand: two K-slabs are in flight while a third is being split
    float4 ra0[2], rb0[2 * HB], ra1[2], rb1[2 * HB];
    auto gload0 = [&](int k0) {
        if (TA) tc_load_rowcontig<1>(p.A, p.lda, p.M, ke, m0, k0, p.vecA, p.kscale, ra0);
        else    tc_load_kcontig<1>(p.A, p.lda, p.M, ke, m0, k0, p.vecA, p.kscale, ra0);
        if (TB) tc_load_kcontig<HB>(p.B, p.ldb, p.N, ke, n0, k0, p.vecB, nullptr, rb0);
        else    tc_load_rowcontig<HB>(p.B, p.ldb, p.N, ke, n0, k0, p.vecB, nullptr, rb0);
    };
    auto gload1 = [&](int k0) {
        if (TA) tc_load_rowcontig<1>(p.A, p.lda, p.M, ke, m0, k0, p.vecA, p.kscale, ra1);
        else    tc_load_kcontig<1>(p.A, p.lda, p.M, ke, m0, k0, p.vecA, p.kscale, ra1);
        if (TB) tc_load_kcontig<HB>(p.B, p.ldb, p.N, ke, n0, k0, p.vecB, nullptr, rb1);
        else    tc_load_rowcontig<HB>(p.B, p.ldb, p.N, ke, n0, k0, p.vecB, nullptr, rb1);
    };
    // stage layout: A_hi | A_lo | B_hi | B_lo
    auto sstore0 = [&](int stage) {
        uint8_t* base = smem + stage * G::STAGE;
        if (TA) tc_store_rowcontig<1>(base, base + G::OP_A, G::LBO_A, ra0);
        else    tc_store_kcontig<1>(base, base + G::OP_A, G::LBO_A, ra0);
        if (TB) tc_store_kcontig<HB>(base + 2 * G::OP_A, base + 2 * G::OP_A + G::OP_B, G::LBO_B, rb0);
        else    tc_store_rowcontig<HB>(base + 2 * G::OP_A, base + 2 * G::OP_A + G::OP_B, G::LBO_B, rb0);
    };
    auto sstore1 = [&](int stage) {
        uint8_t* base = smem + stage * G::STAGE;
        if (TA) tc_store_rowcontig<1>(base, base + G::OP_A, G::LBO_A, ra1);
        else    tc_store_kcontig<1>(base, base + G::OP_A, G::LBO_A, ra1);
        if (TB) tc_store_kcontig<HB>(base + 2 * G::OP_A, base + 2 * G::OP_A + G::OP_B, G::LBO_B, rb1);
        else    tc_store_rowcontig<HB>(base + 2 * G::OP_A, base + 2 * G::OP_A + G::OP_B, G::LBO_B, rb1);
    };

    // epilogue constants
    float alpha = p.alpha, r1c = p.r1coef;
    if (p.alpha_dev) {
        const float s = __ldg(p.alpha_dev);
        alpha *= s;
        r1c *= s;
    }
    const int q = warp & 3;                  // TMEM lane quadrant of this warp
    const int col_half = (warp >> 2) * (BN / 2);   // warps 0-3: left half of the columns, warps 4-7: right half

    const int tiles_per_chunk = KCHUNK / TK;
    const int num_chunks = num_k > 0 ? (num_k + tiles_per_chunk - 1) / tiles_per_chunk : 1;
    if (num_k > 0) gload0(kb);
    if (num_k > 1) gload1(kb + TK);
    for (int chunk = 0; chunk < num_chunks; ++chunk) {
        const int kt_beg = chunk * tiles_per_chunk;
        const int kt_end = min(num_k, kt_beg + tiles_per_chunk);
        const bool first_chunk = (chunk == 0), last_chunk = (chunk == num_chunks - 1);
        for (int kt = kt_beg; kt < kt_end; ++kt) {
            const int stage = kt % STAGES;
            if (kt >= STAGES) {
                // the MMAs that read this stage (slab kt - STAGES) must have completed
                mbar_wait(&mma_bar[stage], (uint32_t)((kt / STAGES - 1) & 1));
            }
            // slab kt lives in register image kt & 1; once it is in shared memory the image is refilled
            // with slab kt + 2
            if ((kt & 1) == 0) {
                sstore0(stage);
                if (kt + 2 < num_k) gload0(kb + (kt + 2) * TK);
            } else {
                sstore1(stage);
                if (kt + 2 < num_k) gload1(kb + (kt + 2) * TK);
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> tensor core
            __syncthreads();
            if (tid == 0) {
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                // descriptors differ only in their 14-bit start-address field: one add per operand
                const uint32_t s16 = smem_u32(smem + stage * G::STAGE) >> 4;
#pragma unroll
                for (int k8 = 0; k8 < TK / 8; ++k8) {
                    const uint64_t a_hi = desc_a + (uint64_t)(s16 + ((2 * k8 * G::LBO_A) >> 4));
                    const uint64_t a_lo = a_hi + (uint64_t)(G::OP_A >> 4);
                    const uint64_t b_hi = desc_b + (uint64_t)(s16 + ((2 * G::OP_A + 2 * k8 * G::LBO_B) >> 4));
                    const uint64_t b_lo = b_hi + (uint64_t)(G::OP_B >> 4);
                    const uint32_t acc = (kt > kt_beg || k8 > 0) ? 1u : 0u;
                    mma_tf32(tmem_d + BN, a_lo, b_hi, idesc, acc);   // cross terms: own accumulator
                    mma_tf32(tmem_d + BN, a_hi, b_lo, idesc, 1u);
                    mma_tf32(tmem_d, a_hi, b_hi, idesc, acc);
                }
                mma_commit(&mma_bar[stage]);
            }
        }
        if (kt_end > kt_beg) {
            const int last = kt_end - 1;
            mbar_wait(&mma_bar[last % STAGES], (uint32_t)((last / STAGES) & 1));
        }
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        PB_STAMP(2);

        // ------------- epilogue / flush: TMEM -> registers -> global -------------
        // chunk 0 applies beta; later chunks accumulate onto C; bias / rank-1 / activation / mirror
        // belong to the last chunk only.
        const float beta = first_chunk ? p.beta : 1.f;
#pragma unroll 1
        for (int cc = 0; cc < BN / 2; cc += 32) {
            const int c0 = col_half + cc;
            uint32_t v[32], w[32];
            if (num_k > 0 && c0 < n_mma) {
                const uint32_t taddr = tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)c0;
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
                asm volatile(
                    "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                    "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                    "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
                    : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]),
                      "=r"(w[8]), "=r"(w[9]), "=r"(w[10]), "=r"(w[11]), "=r"(w[12]), "=r"(w[13]), "=r"(w[14]), "=r"(w[15]),
                      "=r"(w[16]), "=r"(w[17]), "=r"(w[18]), "=r"(w[19]), "=r"(w[20]), "=r"(w[21]), "=r"(w[22]), "=r"(w[23]),
                      "=r"(w[24]), "=r"(w[25]), "=r"(w[26]), "=r"(w[27]), "=r"(w[28]), "=r"(w[29]), "=r"(w[30]), "=r"(w[31])
                    : "r"(taddr + (uint32_t)BN)
                    : "memory");
                asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                if (cc == 0) PB_STAMP(11);
            } else {
#pragma unroll
                for (int j = 0; j < 32; ++j) { v[j] = 0u; w[j] = 0u; }
            }
            // transpose through shared memory (the pipeline stages are idle now) so that global
            // accesses are row-contiguous: 8 lanes x 16 B cover one row's 32 columns, 4 rows per access
            float* tile = reinterpret_cast<float*>(smem) + warp * (32 * EPI_LD);
#pragma unroll
            for (int i = 0; i < 8; ++i) {
                float4 t;
                t.x = __uint_as_float(v[4 * i + 0]) + __uint_as_float(w[4 * i + 0]);
                t.y = __uint_as_float(v[4 * i + 1]) + __uint_as_float(w[4 * i + 1]);
                t.z = __uint_as_float(v[4 * i + 2]) + __uint_as_float(w[4 * i + 2]);
                t.w = __uint_as_float(v[4 * i + 3]) + __uint_as_float(w[4 * i + 3]);
                *reinterpret_cast<float4*>(tile + lane * EPI_LD + 4 * i) = t;
            }
            __syncwarp();
            if (cc == 0) PB_STAMP(12);
            epilogue_rows<EPI>(p, tile, alpha, beta, r1c, last_chunk, m0, n0, q, c0, lane);
            __syncwarp();
            if (cc == 0) PB_STAMP(13);
            if (cc == BN / 2 - 32) PB_STAMP(14);
        }
        // all warps have drained the accumulators before the next chunk overwrites them
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncthreads();
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    }

    PB_STAMP(3);
    // ---- teardown: every warp is done reading TMEM (synchronised above) before warp 0 frees it; the tiles
    //      that go on to a tensor-core panel solve keep their columns until that is done ----
    const bool tc_solve = POST && p.post_wdiag != nullptr && blockIdx.x == 0 && blockIdx.z == 0 && p.post_solve == 2;
    if (warp == 0 && !tc_solve) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"((uint32_t)G::TMEM_COLS)
                     : "memory");
    }

    // ---------------- Cholesky look-ahead and fused panel solve (tile column 0) ----------------
    if (POST && p.post_wdiag != nullptr && blockIdx.x == 0 && blockIdx.z == 0) {
        float* fsmem = reinterpret_cast<float*>(smem);
        const bool owner = (blockIdx.y == 0);
        PB_STAMP(4);
        if (owner) {
            // tile (0,0): factor the next diagonal block while the other tiles are still running
            const int fail = factor_diag_block_regs(p.C, p.ldc, 0, p.post_nb, /*coherent=*/true, fsmem,
                                                    p.post_wdiag, p.post_winv);
            if (fail >= 0 && threadIdx.x == 0 && p.post_info) atomicCAS(p.post_info, 0, p.post_c0 + fail + 1);
            PB_STAMP(5);
            if (p.post_flag) publish_flag(p.post_flag);
            PB_STAMP(6);
        }
        if (p.post_solve) {
            const bool tc = POST && p.post_solve == 2;
            // The rows of the panel this tile solves are its own epilogue's output: a tile that has to wait for the
            // factor (every tile of column 0 but the owner) loads, splits and stages them BEFORE it waits, so that
            // only M is on the path after the flag.  (One copy of the staging code: a two-pass loop.)
            constexpr int LBOB = G::SOLVE_LBO_B;
            uint8_t* x_hi = smem;
            uint8_t* x_lo = smem + G::SOLVE_A;
            uint8_t* m_hi = smem + 2 * G::SOLVE_A;
            uint8_t* m_lo = m_hi + G::SOLVE_B;
#pragma unroll 1
            for (int pass = 0; pass < 2; ++pass) {
                if (pass == 1) {
                    if (!owner) wait_flag(p.post_flag);
                    else __syncthreads();
                }
                if (tc && (pass == 0) == !owner) {
                    float4 xr[8];
#pragma unroll
                    for (int i = 0; i < 8; ++i) {          // all loads first: one L2 round trip
                        const int idx = tid + i * 256;
                        const int row = idx >> 4, kc = idx & 15;
                        const int m = m0 + row;
                        xr[i] = make_float4(0.f, 0.f, 0.f, 0.f);
                        if (m >= p.post_nb && m < p.M)
                            xr[i] = __ldcg(reinterpret_cast<const float4*>(p.C + (int64_t)m * p.ldc + kc * 4));
                    }
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        const int idx = tid + i * 256;
                        const int row = idx >> 4, kc = idx & 15;
                        float4 h, l;
                        split_tf32(xr[i].x, h.x, l.x); split_tf32(xr[i].y, h.y, l.y);
                        split_tf32(xr[i].z, h.z, l.z); split_tf32(xr[i].w, h.w, l.w);
                        const int off = kc * G::LBO_A + row * 16;
                        *reinterpret_cast<float4*>(x_hi + off) = h;
                        *reinterpret_cast<float4*>(x_lo + off) = l;
                    }
                }
            }
            if (tc) {
                // ---- panel solve X <- X M on the tensor cores (nb == 64, 16-byte aligned rows) ----
                // operands: A = this tile's 128 rows of the panel (K = 64, staged above), B = M^T (N = 64), both split
                // to hi / lo over the whole K in the idle pipeline memory; 24 MMAs; the result goes back in place.
                {
                    float4 mr[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int idx = tid + i * 256;
                        const int n = idx & 63, kc = idx >> 6;     // B[n][k] = M[k][n]
                        const float* src = p.post_winv + (kc * 4) * DIAG_NB + n;
                        mr[i] = make_float4(__ldcg(src), __ldcg(src + DIAG_NB), __ldcg(src + 2 * DIAG_NB), __ldcg(src + 3 * DIAG_NB));
                    }
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int idx = tid + i * 256;
                        const int n = idx & 63, kc = idx >> 6;
                        float4 h, l;
                        split_tf32(mr[i].x, h.x, l.x); split_tf32(mr[i].y, h.y, l.y);
                        split_tf32(mr[i].z, h.z, l.z); split_tf32(mr[i].w, h.w, l.w);
                        const int off = kc * LBOB + n * 16;
                        *reinterpret_cast<float4*>(m_hi + off) = h;
                        *reinterpret_cast<float4*>(m_lo + off) = l;
                    }
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncthreads();
                if (tid == 0) {
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t id64 = make_idesc(DIAG_NB);
                    const uint64_t dxa = make_desc(0, G::LBO_A), dmb = make_desc(0, LBOB);
                    const uint32_t xa = smem_u32(x_hi) >> 4, mb = smem_u32(m_hi) >> 4;
#pragma unroll
                    for (int k8 = 0; k8 < DIAG_NB / 8; ++k8) {
                        const uint64_t a_hi = dxa + (uint64_t)(xa + ((2 * k8 * G::LBO_A) >> 4));
                        const uint64_t a_lo = a_hi + (uint64_t)(G::SOLVE_A >> 4);
                        const uint64_t b_hi = dmb + (uint64_t)(mb + ((2 * k8 * LBOB) >> 4));
                        const uint64_t b_lo = b_hi + (uint64_t)(G::SOLVE_B >> 4);
                        const uint32_t acc = k8 > 0 ? 1u : 0u;
                        mma_tf32(tmem_d + BN, a_lo, b_hi, id64, acc);
                        mma_tf32(tmem_d + BN, a_hi, b_lo, id64, 1u);
                        mma_tf32(tmem_d, a_hi, b_hi, id64, acc);
                    }
                    mma_commit(&solve_bar);
                }
                mbar_wait(&solve_bar, 0u);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                {
                    KParams ps = p;
                    ps.N = DIAG_NB; ps.beta = 0.f; ps.bias = nullptr; ps.lower = 0; ps.act = 0;
                    const int c0 = (warp >> 2) * 32;       // warps 0-3: columns 0..31, warps 4-7: 32..63
                    uint32_t v[32], w[32];
                    const uint32_t taddr = tmem_d + ((uint32_t)(q * 32) << 16) + (uint32_t)c0;
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
                    asm volatile(
                        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
                        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
                        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
                        : "=r"(w[0]), "=r"(w[1]), "=r"(w[2]), "=r"(w[3]), "=r"(w[4]), "=r"(w[5]), "=r"(w[6]), "=r"(w[7]),
                          "=r"(w[8]), "=r"(w[9]), "=r"(w[10]), "=r"(w[11]), "=r"(w[12]), "=r"(w[13]), "=r"(w[14]), "=r"(w[15]),
                          "=r"(w[16]), "=r"(w[17]), "=r"(w[18]), "=r"(w[19]), "=r"(w[20]), "=r"(w[21]), "=r"(w[22]), "=r"(w[23]),
                          "=r"(w[24]), "=r"(w[25]), "=r"(w[26]), "=r"(w[27]), "=r"(w[28]), "=r"(w[29]), "=r"(w[30]), "=r"(w[31])
                        : "r"(taddr + (uint32_t)BN)
                        : "memory");
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    // the operands are dead (MMAs complete): the staging tiles may overwrite them, but only once
                    // every warp has its accumulators in registers
                    __syncthreads();
                    float* tile = reinterpret_cast<float*>(smem) + warp * (32 * EPI_LD);
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        float4 t;
                        t.x = __uint_as_float(v[4 * i + 0]) + __uint_as_float(w[4 * i + 0]);
                        t.y = __uint_as_float(v[4 * i + 1]) + __uint_as_float(w[4 * i + 1]);
                        t.z = __uint_as_float(v[4 * i + 2]) + __uint_as_float(w[4 * i + 2]);
                        t.w = __uint_as_float(v[4 * i + 3]) + __uint_as_float(w[4 * i + 3]);
                        *reinterpret_cast<float4*>(tile + lane * EPI_LD + 4 * i) = t;
                    }
                    __syncwarp();
                    // rows of the diagonal block itself (first 64 rows of tile (0,0)) are not panel rows
                    if (m0 + q * 32 >= p.post_nb)
                        epilogue_rows<0>(ps, tile, 1.f, 0.f, 0.f, true, m0, 0, q, c0, lane);
                }
                asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                __syncthreads();
                if (warp == 0) {
                    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d),
                                 "r"((uint32_t)G::TMEM_COLS)
                                 : "memory");
                }
            } else {
                panel_solve_tile(p.C, p.ldc, m0, p.M, p.post_nb, p.post_winv, fsmem);
            }
            PB_STAMP(7);
            PB_STAMP_COL0_DONE();
        }
    }
}

}  // namespace

#ifdef PB_DEBUG_TIMING
extern "C" __attribute__((visibility("default"))) int pinot_b200_debug_times(unsigned long long* out16) {
    return (int)cudaMemcpyFromSymbol(out16, g_dbg_times, sizeof(unsigned long long) * 16);
}
extern "C" __attribute__((visibility("default"))) int pinot_b200_debug_steps(unsigned long long* out512x16) {
    return (int)cudaMemcpyFromSymbol(out512x16, g_dbg_steps, sizeof(unsigned long long) * 512 * 16);
}
#endif

static bool pdl_enabled() {
    const char* v = getenv("PINOT_B200_PDL");          // =0: plain launches (read on every call)
    return !(v && v[0] == '0');
}

int launch_gemm_tc(const KParams& p, bool transA, bool transB, dim3 grid, cudaStream_t stream) {
    // epilogue variant (see epilogue_rows): the generic one whenever a specialised one is not built
    int epi = 0;
    if (p.splitk > 1 || !p.vecC) epi = 3;
    else if (p.r1u || p.mirror) epi = (!transA && transB) ? 2 : 3;
    else if (p.act) epi = (!transA) ? 1 : 3;
    // wide tiles (128 x 256, one CTA per SM) for large products: half the shared-memory operand
    // traffic per flop; PINOT_B200_WIDE=0 switches them off
    static const bool wide_ok = [] { const char* v = getenv("PINOT_B200_WIDE"); return !(v && v[0] == '0'); }();
    const bool wide = wide_ok && !p.post_wdiag && p.splitk == 1 && (epi == 0 || epi == 2) && p.M >= 1024 &&
                      p.N >= 1024 && p.K >= 256;
    auto launch = [&](auto kern, int smem_bytes, dim3 g) -> int {
        PB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_bytes));
        if (p.pdl && pdl_enabled()) {
            cudaLaunchConfig_t cfg = {};
            cfg.gridDim = g; cfg.blockDim = dim3(NTHREADS, 1, 1); cfg.dynamicSmemBytes = (size_t)smem_bytes; cfg.stream = stream;
            cudaLaunchAttribute attr[1];
            attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
            attr[0].val.programmaticStreamSerializationAllowed = 1;
            cfg.attrs = attr; cfg.numAttrs = 1;
            PB_CUDA(cudaLaunchKernelEx(&cfg, kern, p));
        } else {
            kern<<<g, NTHREADS, smem_bytes, stream>>>(p);
        }
        return check_launch("gemm (tcgen05)");
    };
    if (wide) {
        dim3 g2((unsigned)ceil_div(p.N, 256), grid.y, 1);
#define PB_TCW(TA_, TB_, E_) return launch(tgemm_kernel<TA_, TB_, E_, false, 256>, Geo<256>::DYN_SMEM, g2)
        if (epi == 2) PB_TCW(false, true, 2);
        if (!transA && !transB) PB_TCW(false, false, 0);
        if (!transA && transB) PB_TCW(false, true, 0);
        if (transA && !transB) PB_TCW(true, false, 0);
        PB_TCW(true, true, 0);
#undef PB_TCW
    }
    if (p.post_wdiag) {
        PB_REQUIRE(!transA && transB && epi == 0, "gemm: the Cholesky look-ahead is built for the NT update only");
        return launch(tgemm_kernel<false, true, 0, true, 128>, Geo<128>::DYN_SMEM, grid);
    }
#define PB_TC(TA_, TB_, E_) return launch(tgemm_kernel<TA_, TB_, E_, false, 128>, Geo<128>::DYN_SMEM, grid)
    if (!transA && !transB) { if (epi == 0) PB_TC(false, false, 0); if (epi == 1) PB_TC(false, false, 1); PB_TC(false, false, 3); }
    if (!transA && transB)  { if (epi == 0) PB_TC(false, true, 0);  if (epi == 1) PB_TC(false, true, 1);
                              if (epi == 2) PB_TC(false, true, 2);  PB_TC(false, true, 3); }
    if (transA && !transB)  { if (epi == 0) PB_TC(true, false, 0);  PB_TC(true, false, 3); }
    if (epi == 0) PB_TC(true, true, 0);
    PB_TC(true, true, 3);
#undef PB_TC
}

}  // namespace pb
