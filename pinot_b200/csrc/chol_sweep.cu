// Kernel (4), single-launch form: the whole blocked Cholesky sweep of the stacked matrix [A ; X]
// (see chol.cu for what the sweep computes) as ONE persistent cooperative kernel.
//
// chol.cu runs the sweep as one launch per 64-column panel; at n <= ~6000 that chain of ~n/64
// dependent launches (each: update -> factor the next diagonal block -> panel solve) is all that
// bounds it.  Here the launch boundaries become flags in global memory and every CTA works through
// its own, statically assigned share of a tile-level task graph:
//
//   * the matrix is cut into absolute 128 x 128 tiles; tile (I, J) belongs to CTA (I + 67 J) mod G for
//     the whole sweep, so every update of a tile is made by the same CTA, in program order;
//   * column tile J holds the panels 2J and 2J+1 (64 columns each).  Per tile of that column, in order:
//       S0  left panel:  [diagonal tile: factor D_2J, publish L11 and M = L11^-T]  X <- X M   (tcgen05)
//       U   right half -= X_left X_left[rows of the right half]^T                  (rank 64, tcgen05)
//       S1  right panel: [diagonal tile: factor D_2J+1]                            X <- X M
//     then progress[I] = 2J + 2 tells everyone that tile row I is final up to column tile J;
//   * every tile (I, J') to the right, J' > J, receives ONE rank-128 update per column tile J ("far"
//     task) as soon as progress[I] and progress[J'] allow -- left-looking in units of column tiles, so
//     the trailing matrix is read and written n/128 times, not n/64 times;
//   * a CTA keeps two cursors, one over its near tasks (S0 / U / S1, the critical chain) and one over
//     its far tasks, always prefers a ready near task, and does not start a far task when its next
//     near task is about to become ready (its panel is the one being factored or the next one);
//   * C tiles are updated with vector reductions (red.global.add.v4.f32) straight from the TMEM
//     accumulators -- each element has exactly one writer at a time, the reduction only saves the
//     read-modify-write round trip through L2; operands written by other CTAs are read with ld.global.cg.
//
// All CTAs must be co-resident (they wait on one another's flags): the kernel is launched with
// cudaLaunchCooperativeKernel and a grid of at most 2 CTAs per SM.
#include "gemm.cuh"
#include "chol_diag.cuh"
#include "chol_sweep.cuh"
#include <stdlib.h>

namespace pb {

namespace {

constexpr int TM = 128, TK = 16, STAGES = 3, NTH = 256;
constexpr int SBO_BYTES = 128;
constexpr int LBO_A = (TM / 8) * 128 + 32;           // between K-adjacent core matrices of a 128-row operand
constexpr int OP = (TK / 4) * LBO_A;                 // one of A_hi, A_lo, B_hi, B_lo per stage
constexpr int STAGE_BYTES = 4 * OP;
constexpr int PIPE_BYTES = STAGES * STAGE_BYTES;
constexpr int SOLVE_LBO_B = (DIAG_NB / 8) * 128 + 32;
constexpr int SOLVE_A = (DIAG_NB / 4) * LBO_A;       // X tile (128 x 64), one of hi / lo
constexpr int SOLVE_B = (DIAG_NB / 4) * SOLVE_LBO_B; // M^T (64 x 64), one of hi / lo
constexpr int SOLVE_BYTES = 2 * SOLVE_A + 2 * SOLVE_B;
constexpr int FFMA_BYTES = SOLVE_SMEM_FLOATS * 4;
constexpr int DIAG_BYTES = DIAG_SMEM_FLOATS * 4;
constexpr int cmax(int a, int b) { return a > b ? a : b; }
// fused diagonal-tile routine (factor / solve / update / factor on chip), offsets in floats
constexpr int DT_LD = 68;                            // padded, 16-byte aligned rows of the transposed blocks
constexpr int XS_LD = 65;
constexpr int F_LT = 0;                              // Lt[t][r] = L[r][t]
constexpr int F_YT = F_LT + 64 * 64;                 // Yt[t][r] = Y[r][t], Y = L^-T (rows of the identity part)
constexpr int F_DA = F_YT + 64 * 64;                 // D11 transposed (DT_LD), later M11 row-major (DT_LD)
constexpr int F_PC = F_DA + 64 * DT_LD;              // pivot columns of the current 16-block, 1/sqrt(pivot), fail flag
constexpr int F_XS = F_PC + 320;                     // X21 / L21 row-major (XS_LD)
constexpr int F_LT21 = F_XS + 64 * XS_LD;            // L21 transposed [k][r]
constexpr int F_DB = F_LT21 + 64 * 64;               // D22 transposed (DT_LD)
constexpr int FUSED_BYTES = (F_DB + 64 * DT_LD) * 4;
constexpr int DYN_SMEM = cmax(cmax(cmax(PIPE_BYTES, SOLVE_BYTES), cmax(FFMA_BYTES, DIAG_BYTES)), FUSED_BYTES) + 128;
constexpr int TMEM_COLS = 256;                       // [0,128): hi.hi accumulator, [128,256): cross terms
constexpr int EPI_LD = 36;
constexpr int OWNER_STRIDE = 67;
constexpr int MAX_RT = SWEEP_MAX_ROW_TILES;

// Dynamic shared memory, declared at file scope and used directly by every device function: a pointer to it
// that travels through a struct or an integer cast loses its address space, and the compiler then emits
// generic LD / ST in place of LDS / STS (measured: the shared-memory-bound routines ran ~2x slower).
extern __shared__ __align__(128) uint8_t dyn_smem[];

enum : int { T_DONE = 0, T_S0 = 1, T_U = 2, T_S1 = 3, T_FAR = 4 };

struct Task { int type, I, J, Jsrc; };

#ifdef PB_SWEEP_TIMING
// development aid: per-task timeline (cta, type, I, J, picked, started, finished in globaltimer ns)
__device__ __forceinline__ unsigned long long gtime() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
struct LogRec { int cta, type, I, J; unsigned long long t_pick, t_start, t_end, ph[8], clk_start, clk_end; };
__shared__ unsigned long long s_ph[8];
#define PB_PH(i) do { if (threadIdx.x == 0) s_ph[i] = gtime(); } while (0)
#else
#define PB_PH(i)
#endif
#ifdef PB_SWEEP_TIMING
constexpr int LOG_CAP = 1 << 16;
__device__ LogRec g_log[LOG_CAP];
__device__ int g_log_n;
__device__ unsigned long long g_fphase[4];
__device__ long long g_wstamp[4][4][6];   // [jb][warp][stamp]
#endif

struct Sched {
    int nJ, nPh, nI, near_done;     // near cursor: column tile, phase (0 S0, 1 U, 2 S1), tile row
    int fS, fP, fI, far_done;       // far cursor: source column tile, target column tile, tile row
    int nfact;                      // last value seen of the factored-panel counter
};

// ---- small PTX wrappers --------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    const uint32_t addr = smem_u32(bar);
    // bounded: a phase-bookkeeping error must end in a launch failure, not in a hung GPU
    for (unsigned spins = 0;; ++spins) {
        uint32_t done;
        asm volatile(
            "{\n\t"
            ".reg .pred p;\n\t"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
            "selp.u32 %0, 1, 0, p;\n\t"
            "}\n"
            : "=r"(done)
            : "r"(addr), "r"(parity)
            : "memory");
        if (done) return;
        if (spins > (1u << 26)) __trap();
    }
}
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, int lbo_bytes) {
    uint64_t d = 0;
    d |= (uint64_t)((saddr >> 4) & 0x3FFF);
    d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
    d |= (uint64_t)((SBO_BYTES >> 4) & 0x3FFF) << 32;
    d |= (uint64_t)1 << 46;
    return d;
}
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
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
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
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
          "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ int ld_acquire(const int* p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release(int* p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

// ---- task graph ------------------------------------------------------------------------------------
__device__ __forceinline__ bool tile_active(const SweepParams& p, int I, int J) {
    // does tile row I hold rows that take part in column tile J (are solved there / carry its updates)?
    const int r0 = I * TM, r1 = min(r0 + TM, p.rows);
    if (r0 < p.n && min(r1, p.n) > TM * J) return true;                 // rows of A at or below the column tile
    if (r1 > p.n && r0 < p.n + p.nd) return true;                       // dense extension rows
    if (p.ident) {
        const int i0 = p.n + p.nd, cend = min(TM * J + TM, p.n);         // identity row c is zero before column c
        if (r1 > i0 && r0 - i0 < cend) return true;
    }
    return false;
}
__device__ __forceinline__ int first_owned(int g, int G, int J, int lo) {
    // smallest I >= lo with (I + OWNER_STRIDE * J) mod G == g
    int i0 = (g - OWNER_STRIDE * J) % G;
    if (i0 < 0) i0 += G;
    if (lo > i0) i0 += ((lo - i0 + G - 1) / G) * G;
    return i0;
}
__device__ bool near_advance(const SweepParams& p, int g, int G, Sched& s) {
    while (s.nJ < p.CT) {
        const int phases = (TM * s.nJ + DIAG_NB < p.n) ? 3 : 1;
        while (s.nPh < phases) {
            for (int I = first_owned(g, G, s.nJ, max(s.nI, s.nJ)); I < p.RT; I += G)
                if (tile_active(p, I, s.nJ)) { s.nI = I; return true; }
            s.nPh++; s.nI = 0;
        }
        s.nJ++; s.nPh = 0; s.nI = 0;
    }
    s.near_done = 1;
    return false;
}
__device__ bool far_advance(const SweepParams& p, int g, int G, Sched& s) {
    while (s.fS < p.CT - 1) {
        while (s.fP < p.CT) {
            for (int I = first_owned(g, G, s.fP, max(s.fI, s.fP)); I < p.RT; I += G)
                if (tile_active(p, I, s.fS)) { s.fI = I; return true; }
            s.fP++; s.fI = 0;
        }
        s.fS++; s.fP = s.fS + 1; s.fI = 0;
    }
    s.far_done = 1;
    return false;
}

// ---- operand staging (k-contiguous rows of E -> hi / lo core matrices) -----------------------------
// thread -> rows tid/4 and tid/4 + 64 of the 128-row operand tile, 16-byte chunk tid%4 of the 16-column slab
__device__ __forceinline__ void load_slab(const float* __restrict__ E, int64_t ld, int row0, int row_end, int k0,
                                          float4 (&r)[2]) {
    const int tid = threadIdx.x;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int row = row0 + (tid >> 2) + 64 * i;
        r[i] = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row < row_end) r[i] = __ldcg(reinterpret_cast<const float4*>(E + (int64_t)row * ld + k0 + (tid & 3) * 4));
    }
}
__device__ __forceinline__ void store_slab(uint8_t* hi, uint8_t* lo, const float4 (&r)[2]) {
    const int tid = threadIdx.x;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int row = (tid >> 2) + 64 * i;
        float4 h, l;
        split_tf32(r[i].x, h.x, l.x);
        split_tf32(r[i].y, h.y, l.y);
        split_tf32(r[i].z, h.z, l.z);
        split_tf32(r[i].w, h.w, l.w);
        const int off = (tid & 3) * LBO_A + row * 16;
        *reinterpret_cast<float4*>(hi + off) = h;
        *reinterpret_cast<float4*>(lo + off) = l;
    }
}

struct Cta {
    uint64_t* mma_bar;      // [STAGES]
    uint64_t* solve_bar;
    uint32_t tmem;
    uint32_t slabs;         // slabs issued so far (stage ring position / mbarrier phase bookkeeping)
    uint32_t solves;        // solves issued so far (phase of solve_bar)
};

// C[arow0 + r][ccol0 + j] -= sum_t E[arow0 + r][kcol0 + t] * E[ccol0 + j][kcol0 + t],  t < K (64 or 128),
// j < ncols, stored only where  arow0 + r >= mask_lo,  < rows,  and  ccol0 + j <= arow0 + r  (lower part;
// extension rows lie below every column).
__device__ void tile_update(const SweepParams& p, Cta& c, int arow0, int kcol0, int K, int ccol0, int ncols,
                            int mask_lo) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int n_mma = min(TM, ((ncols + 15) / 16) * 16);
    const uint32_t idesc = make_idesc(n_mma);
    const uint64_t desc = make_desc(0, LBO_A);
    const int a_end = p.rows, b_end = ccol0 + ncols;
    const int nslab = K / TK;
    float4 ra0[2], rb0[2], ra1[2], rb1[2];
    load_slab(p.E, p.ld, arow0, a_end, kcol0, ra0);
    load_slab(p.E, p.ld, ccol0, b_end, kcol0, rb0);
    load_slab(p.E, p.ld, arow0, a_end, kcol0 + TK, ra1);
    load_slab(p.E, p.ld, ccol0, b_end, kcol0 + TK, rb1);
    PB_PH(0);
#pragma unroll 1
    for (int s = 0; s < nslab; s += 2) {
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const uint32_t g = c.slabs;
            const int stage = (int)(g % STAGES);
            if (g >= STAGES) mbar_wait(&c.mma_bar[stage], ((g / STAGES) - 1) & 1);
            uint8_t* base = dyn_smem + stage * STAGE_BYTES;
            if (h == 0) {
                store_slab(base, base + OP, ra0);
                store_slab(base + 2 * OP, base + 3 * OP, rb0);
                if (s + 2 < nslab) {
                    load_slab(p.E, p.ld, arow0, a_end, kcol0 + (s + 2) * TK, ra0);
                    load_slab(p.E, p.ld, ccol0, b_end, kcol0 + (s + 2) * TK, rb0);
                }
            } else {
                store_slab(base, base + OP, ra1);
                store_slab(base + 2 * OP, base + 3 * OP, rb1);
                if (s + 3 < nslab) {
                    load_slab(p.E, p.ld, arow0, a_end, kcol0 + (s + 3) * TK, ra1);
                    load_slab(p.E, p.ld, ccol0, b_end, kcol0 + (s + 3) * TK, rb1);
                }
            }
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
            if (tid == 0) {
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t s16 = smem_u32(base) >> 4;
#pragma unroll
                for (int k8 = 0; k8 < TK / 8; ++k8) {
                    const uint64_t a_hi = desc + (uint64_t)(s16 + ((2 * k8 * LBO_A) >> 4));
                    const uint64_t a_lo = a_hi + (uint64_t)(OP >> 4);
                    const uint64_t b_hi = a_hi + (uint64_t)((2 * OP) >> 4);
                    const uint64_t b_lo = b_hi + (uint64_t)(OP >> 4);
                    const uint32_t acc = (s + h > 0 || k8 > 0) ? 1u : 0u;
                    mma_tf32(c.tmem + TM, a_lo, b_hi, idesc, acc);
                    mma_tf32(c.tmem + TM, a_hi, b_lo, idesc, 1u);
                    mma_tf32(c.tmem, a_hi, b_hi, idesc, acc);
                }
                mma_commit(&c.mma_bar[stage]);
            }
            c.slabs = g + 1;
        }
    }
    PB_PH(1);
    // epilogue: C <- C - (hi.hi + cross) as a plain read-modify-write (one writer per element; vector reductions
    // were tried first and are limited to ~4 B/cycle per SM).  The C reads of a 32-column block are issued before
    // the accumulators are fetched (for the first block: before the wait for the last MMAs).  Tiles that lie
    // wholly inside the updated region take a mask-free path.
    const int q = warp & 3, half = warp >> 2;
    const int col4 = (lane & 7) * 4;
    const bool warp_on = half * (TM / 2) < n_mma;
    const bool interior = arow0 >= mask_lo && arow0 + TM <= p.rows && ncols == n_mma && ccol0 + ncols - 1 <= arow0;
    float* tile = reinterpret_cast<float*>(dyn_smem) + warp * (32 * EPI_LD);
    const int row_base = arow0 + q * 32 + (lane >> 3);
    auto elem_ok = [&](int m, int col) { return m >= mask_lo && m < p.rows && col < b_end && col <= m; };
    auto load_c = [&](int c0, float4 (&cv)[8]) {
        const int col = ccol0 + c0 + col4;
        if (interior) {
#pragma unroll
            for (int t8 = 0; t8 < 8; ++t8)
                cv[t8] = __ldcg(reinterpret_cast<const float4*>(p.E + (int64_t)(row_base + t8 * 4) * p.ld + col));
        } else {
#pragma unroll 1
            for (int t8 = 0; t8 < 8; ++t8) {
                const int m = row_base + t8 * 4;
                const float* cp = p.E + (int64_t)m * p.ld + col;
                float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
                if (elem_ok(m, col)) v.x = __ldcg(cp);
                if (elem_ok(m, col + 1)) v.y = __ldcg(cp + 1);
                if (elem_ok(m, col + 2)) v.z = __ldcg(cp + 2);
                if (elem_ok(m, col + 3)) v.w = __ldcg(cp + 3);
                cv[t8] = v;
            }
        }
    };
    auto stage_acc = [&](int c0) {
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {                 // 16 columns at a time keeps the register footprint small
            uint32_t v[16], w[16];
            const uint32_t taddr = c.tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)(c0 + 16 * hh);
            tmem_ld16(taddr, v);
            tmem_ld16(taddr + (uint32_t)TM, w);
            asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                float4 t;
                t.x = __uint_as_float(v[4 * i + 0]) + __uint_as_float(w[4 * i + 0]);
                t.y = __uint_as_float(v[4 * i + 1]) + __uint_as_float(w[4 * i + 1]);
                t.z = __uint_as_float(v[4 * i + 2]) + __uint_as_float(w[4 * i + 2]);
                t.w = __uint_as_float(v[4 * i + 3]) + __uint_as_float(w[4 * i + 3]);
                *reinterpret_cast<float4*>(tile + lane * EPI_LD + 16 * hh + 4 * i) = t;
            }
        }
        __syncwarp();
    };
    auto combine = [&](int c0, const float4 (&cv)[8]) {
        const int col = ccol0 + c0 + col4;
        if (interior) {
#pragma unroll
            for (int t8 = 0; t8 < 8; ++t8) {
                const float4 a = *reinterpret_cast<const float4*>(tile + (t8 * 4 + (lane >> 3)) * EPI_LD + col4);
                *reinterpret_cast<float4*>(p.E + (int64_t)(row_base + t8 * 4) * p.ld + col) =
                    make_float4(cv[t8].x - a.x, cv[t8].y - a.y, cv[t8].z - a.z, cv[t8].w - a.w);
            }
        } else {
#pragma unroll 1
            for (int t8 = 0; t8 < 8; ++t8) {
                const int m = row_base + t8 * 4;
                const float4 a = *reinterpret_cast<const float4*>(tile + (t8 * 4 + (lane >> 3)) * EPI_LD + col4);
                float* cp = p.E + (int64_t)m * p.ld + col;
                if (elem_ok(m, col)) cp[0] = cv[t8].x - a.x;
                if (elem_ok(m, col + 1)) cp[1] = cv[t8].y - a.y;
                if (elem_ok(m, col + 2)) cp[2] = cv[t8].z - a.z;
                if (elem_ok(m, col + 3)) cp[3] = cv[t8].w - a.w;
            }
        }
        __syncwarp();
    };
    float4 c_a[8];
    if (warp_on) load_c(half * (TM / 2), c_a);
    {
        const uint32_t last = c.slabs - 1;
        mbar_wait(&c.mma_bar[last % STAGES], (last / STAGES) & 1);
    }
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    PB_PH(2);
    if (warp_on) {
        const int c0 = half * (TM / 2);
        stage_acc(c0);
        combine(c0, c_a);
        if (c0 + 32 < n_mma) {
            load_c(c0 + 32, c_a);
            stage_acc(c0 + 32);
            combine(c0 + 32, c_a);
        }
    }
    PB_PH(3);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __threadfence();
    PB_PH(4);
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    PB_PH(5);
}

// X <- X M for the rows [max(row_lo, 128 I), min(128 I + 128, rows)) of tile row I, columns [c0, c0 + 64):
// one 128 x 64 x 64 product on the tensor cores (3xTF32), operands resident for the whole K.
__device__ void solve_tile_tc(const SweepParams& p, Cta& c, int I, int c0, int row_lo, const float* __restrict__ winv) {
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m0 = I * TM;
    uint8_t* x_hi = dyn_smem;
    uint8_t* x_lo = dyn_smem + SOLVE_A;
    uint8_t* m_hi = dyn_smem + 2 * SOLVE_A;
    uint8_t* m_lo = m_hi + SOLVE_B;
    {
        float4 xr[8], mr[4];
#pragma unroll
        for (int i = 0; i < 8; ++i) {                  // all loads first: one L2 round trip
            const int idx = tid + i * NTH;
            const int row = idx >> 4, kc = idx & 15;
            const int m = m0 + row;
            xr[i] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (m >= row_lo && m < p.rows)
                xr[i] = __ldcg(reinterpret_cast<const float4*>(p.E + (int64_t)m * p.ld + c0 + kc * 4));
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int idx = tid + i * NTH;
            const int n = idx & 63, kc = idx >> 6;     // B[n][k] = M[k][n]
            const float* src = winv + (kc * 4) * DIAG_NB + n;
            mr[i] = make_float4(__ldcg(src), __ldcg(src + DIAG_NB), __ldcg(src + 2 * DIAG_NB), __ldcg(src + 3 * DIAG_NB));
        }
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            const int idx = tid + i * NTH;
            const int row = idx >> 4, kc = idx & 15;
            float4 h, l;
            split_tf32(xr[i].x, h.x, l.x); split_tf32(xr[i].y, h.y, l.y);
            split_tf32(xr[i].z, h.z, l.z); split_tf32(xr[i].w, h.w, l.w);
            const int off = kc * LBO_A + row * 16;
            *reinterpret_cast<float4*>(x_hi + off) = h;
            *reinterpret_cast<float4*>(x_lo + off) = l;
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int idx = tid + i * NTH;
            const int n = idx & 63, kc = idx >> 6;
            float4 h, l;
            split_tf32(mr[i].x, h.x, l.x); split_tf32(mr[i].y, h.y, l.y);
            split_tf32(mr[i].z, h.z, l.z); split_tf32(mr[i].w, h.w, l.w);
            const int off = kc * SOLVE_LBO_B + n * 16;
            *reinterpret_cast<float4*>(m_hi + off) = h;
            *reinterpret_cast<float4*>(m_lo + off) = l;
        }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    PB_PH(4);
    if (tid == 0) {
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint32_t id64 = make_idesc(DIAG_NB);
        const uint64_t dxa = make_desc(0, LBO_A), dmb = make_desc(0, SOLVE_LBO_B);
        const uint32_t xa = smem_u32(x_hi) >> 4, mb = smem_u32(m_hi) >> 4;
#pragma unroll
        for (int k8 = 0; k8 < DIAG_NB / 8; ++k8) {
            const uint64_t a_hi = dxa + (uint64_t)(xa + ((2 * k8 * LBO_A) >> 4));
            const uint64_t a_lo = a_hi + (uint64_t)(SOLVE_A >> 4);
            const uint64_t b_hi = dmb + (uint64_t)(mb + ((2 * k8 * SOLVE_LBO_B) >> 4));
            const uint64_t b_lo = b_hi + (uint64_t)(SOLVE_B >> 4);
            const uint32_t acc = k8 > 0 ? 1u : 0u;
            mma_tf32(c.tmem + TM, a_lo, b_hi, id64, acc);
            mma_tf32(c.tmem + TM, a_hi, b_lo, id64, 1u);
            mma_tf32(c.tmem, a_hi, b_hi, id64, acc);
        }
        mma_commit(c.solve_bar);
    }
    mbar_wait(c.solve_bar, c.solves & 1);
    c.solves += 1;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    PB_PH(5);
    {
        const int q = warp & 3;
        const int cb = (warp >> 2) * 32;               // warps 0-3: columns 0..31, warps 4-7: 32..63
        uint32_t v[32], w[32];
        const uint32_t taddr = c.tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)cb;
        tmem_ld32(taddr, v);
        tmem_ld32(taddr + (uint32_t)TM, w);
        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
        __syncthreads();                               // operands are dead everywhere before the staging tiles reuse them
        float* tile = reinterpret_cast<float*>(dyn_smem) + warp * (32 * EPI_LD);
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
#pragma unroll
        for (int t8 = 0; t8 < 8; ++t8) {
            const int rl = t8 * 4 + (lane >> 3);
            const int m = m0 + q * 32 + rl;
            if (m < row_lo || m >= p.rows) continue;
            const float4 a = *reinterpret_cast<const float4*>(tile + rl * EPI_LD + (lane & 7) * 4);
            *reinterpret_cast<float4*>(p.E + (int64_t)m * p.ld + c0 + cb + (lane & 7) * 4) = a;
        }
    }
    PB_PH(6);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __threadfence();
    __syncthreads();
    PB_PH(7);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}

// Factor the diagonal block of panel k (by the CTA that owns the diagonal tile), store L11 into E and
// diag(L) into diag_out, publish L11 / M through the factored-panel counter.
__device__ void factor_panel(const SweepParams& p, Cta& c, int k) {
    const int c0 = k * DIAG_NB, nb = min(DIAG_NB, p.n - c0);
    float* wd = p.wdiag + (size_t)k * DIAG_NB * DIAG_NB;
    float* wi = p.winv + (size_t)k * DIAG_NB * DIAG_NB;
    __syncthreads();
    PB_PH(0);
    const int fail = factor_diag_block_regs(p.E, p.ld, c0, nb, /*coherent=*/true, reinterpret_cast<float*>(dyn_smem), wd, wi);
    PB_PH(1);
    if (threadIdx.x == 0 && fail >= 0) atomicCAS(p.info, 0, c0 + fail + 1);
    __threadfence();
    __syncthreads();
    PB_PH(2);
    {
        float v[16];                                  // all loads first: one L2 round trip, not sixteen
#pragma unroll
        for (int t = 0; t < 16; ++t) v[t] = __ldcg(wd + threadIdx.x + t * NTH);
#pragma unroll
        for (int t = 0; t < 16; ++t) {
            const int idx = threadIdx.x + t * NTH;
            const int i = idx >> 6, j = idx & 63;
            if (i < nb && j <= i) {
                p.E[(int64_t)(c0 + i) * p.ld + c0 + j] = v[t];
                if (i == j && p.diag_out) p.diag_out[c0 + i] = v[t];
            }
        }
    }
    __threadfence();
    __syncthreads();
    PB_PH(3);
    if (threadIdx.x == 0) st_release(p.nfact, k + 1);
}

// ---- on-chip factorisation of a 64 x 64 block and of its inverse transpose -----------------------------
// Input Dt[c * DT_LD + r] = D[r][c] (c <= r).  Output Lt[t * 64 + r] = L[r][t] and Yt[t * 64 + r] = (L^-T)[r][t].
// Threads 0..63 own the rows of D, threads 64..127 the rows of an identity block stacked below it (the same
// augmented elimination as chol_diag.cuh, which turns the identity into L^-T).  Blocked by 16 columns,
// left-looking: (a) every live row applies the finished columns to its 16 entries of the block, (b) the warp
// that holds the block's 16 rows factors the 16 x 16 diagonal block with shuffles -- no barrier inside the
// pivot chain -- and publishes the pivot columns, (c) every other row replays those 16 eliminations on its own.
// Three named barriers per block instead of one per pivot, and a loop body of ~1k instructions instead of a
// 64-fold unrolled one (the instruction stream of the unrolled version does not fit the instruction caches).
// Called by threads 0..127 (warps 0-3) only; returns the first failing pivot or -1 through sm[F_PC + 300].
#ifdef PB_SWEEP_TIMING
__shared__ unsigned long long s_fphase[4];
#define PB_FCLK(var) const long long var = clock64()
#define PB_FACC(i, a, b) do { if (threadIdx.x == 0) s_fphase[i] += (unsigned long long)((b) - (a)); } while (0)
#else
#define PB_FCLK(var)
#define PB_FACC(i, a, b)
#endif
__device__ void factor64(float* sm, const float* Dt) {
    float* Lt = sm + F_LT;
    float* Yt = sm + F_YT;
    float* Pc = sm + F_PC;
    float* dinv = Pc + 256;
    const int tid = threadIdx.x, lane = tid & 31;
    const bool is_a = tid < 64;
    const int r = tid & 63;
    if (tid == 0) reinterpret_cast<int*>(Pc)[300] = -1;
    // rows of L^-T are zero left of their diagonal entry: the left-looking update reads those entries
    for (int i = tid; i < 64 * 64; i += 128) Yt[i] = 0.f;
    asm volatile("bar.sync 1, 128;" ::: "memory");
#pragma unroll 1
    for (int jb = 0; jb < 4; ++jb) {
        const int c0 = 16 * jb;
        const bool live = is_a ? (r >= c0) : (r < c0 + 16);
        const bool in_blk = r >= c0 && r < c0 + 16;
        PB_FCLK(tk0);
        float x[16];
#pragma unroll
        for (int c = 0; c < 16; ++c) {
            if (is_a) x[c] = (live && c0 + c <= r) ? Dt[(c0 + c) * DT_LD + r] : 0.f;
            else x[c] = (r == c0 + c) ? 1.f : 0.f;
        }
        if (live && !(in_blk && !is_a)) {
            const float* mine = (is_a ? Lt : Yt) + r;
#pragma unroll 4
            for (int t = 0; t < c0; ++t) {
                const float m = mine[t * 64];
                const float4* lv = reinterpret_cast<const float4*>(Lt + t * 64 + c0);
                const float4 l0 = lv[0], l1 = lv[1], l2 = lv[2], l3 = lv[3];
                x[0] = fmaf(-m, l0.x, x[0]);   x[1] = fmaf(-m, l0.y, x[1]);   x[2] = fmaf(-m, l0.z, x[2]);   x[3] = fmaf(-m, l0.w, x[3]);
                x[4] = fmaf(-m, l1.x, x[4]);   x[5] = fmaf(-m, l1.y, x[5]);   x[6] = fmaf(-m, l1.z, x[6]);   x[7] = fmaf(-m, l1.w, x[7]);
                x[8] = fmaf(-m, l2.x, x[8]);   x[9] = fmaf(-m, l2.y, x[9]);   x[10] = fmaf(-m, l2.z, x[10]); x[11] = fmaf(-m, l2.w, x[11]);
                x[12] = fmaf(-m, l3.x, x[12]); x[13] = fmaf(-m, l3.y, x[13]); x[14] = fmaf(-m, l3.z, x[14]); x[15] = fmaf(-m, l3.w, x[15]);
            }
        }
        PB_FCLK(tk1);
        PB_FACC(0, tk0, tk1);
        // (b) the warp holding rows c0 .. c0+15 of D factors the diagonal block
        if ((tid >> 5) == (c0 >> 5)) {
            const int l0 = c0 & 31;
            const bool mine_blk = lane >= l0 && lane < l0 + 16;
            float xa[16];
#pragma unroll
            for (int c = 0; c < 16; ++c) xa[c] = x[c];
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                float d = __shfl_sync(0xffffffffu, xa[j], l0 + j);
                if (!(d > 0.f)) {                       // also catches NaN
                    if (lane == 0) atomicMax(reinterpret_cast<int*>(Pc) + 300, 0x7fffffff - (c0 + j));
                    d = 1.f;
                }
                const float inv = rsqrtf(d);
                if (mine_blk) Pc[j * 16 + (lane - l0)] = xa[j];          // unscaled column j of the block
                if (lane == l0 + j) dinv[j] = inv;
                const float y = xa[j] * inv * inv;
#pragma unroll
                for (int c = j + 1; c < 16; ++c) {
                    const float lc = __shfl_sync(0xffffffffu, xa[j], l0 + c);
                    xa[c] = fmaf(-y, lc, xa[c]);
                }
                xa[j] *= inv;
            }
            if (mine_blk) {
#pragma unroll
                for (int c = 0; c < 16; ++c) x[c] = xa[c];
            }
        }
        asm volatile("bar.sync 1, 128;" ::: "memory");
        PB_FCLK(tk2);
        PB_FACC(1, tk1, tk2);
#ifdef PB_SWEEP_TIMING
        if (lane == 0) { g_wstamp[jb][tid >> 5][0] = tk0; g_wstamp[jb][tid >> 5][1] = tk1; g_wstamp[jb][tid >> 5][2] = tk2; }
#endif
        // (c) all other live rows replay the 16 eliminations with the published columns
        if (live && !(is_a && in_blk)) {
#pragma unroll
            for (int j = 0; j < 16; ++j) {
                const float inv = dinv[j];
                const float y = x[j] * inv * inv;
                x[j] *= inv;
#pragma unroll
                for (int c = j + 1; c < 16; ++c) x[c] = fmaf(-y, Pc[j * 16 + c], x[c]);
            }
        }
        if (live) {
            float* dst = (is_a ? Lt : Yt) + r;
#pragma unroll
            for (int c = 0; c < 16; ++c) dst[(c0 + c) * 64] = x[c];
        }
#ifdef PB_SWEEP_TIMING
        if (lane == 0) g_wstamp[jb][tid >> 5][3] = clock64();
#endif
        asm volatile("bar.sync 1, 128;" ::: "memory");
        PB_FCLK(tk3);
#ifdef PB_SWEEP_TIMING
        if (lane == 0) g_wstamp[jb][tid >> 5][4] = tk3;
#endif
        PB_FACC(2, tk2, tk3);
    }
}

// Everything column tile J asks of its diagonal tile, on chip: factor D_2J, solve the 64 rows below it, update
// D_2J+1, factor it -- one load of the 128 x 128 lower tile, no global round trip between the four steps.
// Only for regular tiles (128 J + 128 <= n: two full panels, no extension rows inside).
__device__ void diag_tile_fused(const SweepParams& p, Cta& c, int J) {
    float* sm = reinterpret_cast<float*>(dyn_smem);
    float* Lt = sm + F_LT;
    float* Yt = sm + F_YT;
    float* DA = sm + F_DA;
    float* Xs = sm + F_XS;
    float* Lt21 = sm + F_LT21;
    float* DB = sm + F_DB;
    const int tid = threadIdx.x;
    const int c0 = J * TM;
    const int k0 = 2 * J;
    __syncthreads();
    PB_PH(0);
    {   // one batch of loads for all three blocks
        float4 d11[4], d22[4], x21[4];
        const int rt = tid & 63, qt = tid >> 6;              // transposed blocks: lane <-> row (conflict-free stores)
        const int rx = tid >> 4, qx = tid & 15;              // row-major block: lane <-> 16-byte chunk (coalesced)
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int cc = 4 * (qt + 4 * i);
            d11[i] = __ldcg(reinterpret_cast<const float4*>(p.E + (int64_t)(c0 + rt) * p.ld + c0 + cc));
            d22[i] = __ldcg(reinterpret_cast<const float4*>(p.E + (int64_t)(c0 + 64 + rt) * p.ld + c0 + 64 + cc));
            x21[i] = __ldcg(reinterpret_cast<const float4*>(p.E + (int64_t)(c0 + 64 + rx + 16 * i) * p.ld + c0 + 4 * qx));
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int cc = 4 * (qt + 4 * i);
            DA[(cc + 0) * DT_LD + rt] = d11[i].x; DA[(cc + 1) * DT_LD + rt] = d11[i].y;
            DA[(cc + 2) * DT_LD + rt] = d11[i].z; DA[(cc + 3) * DT_LD + rt] = d11[i].w;
            DB[(cc + 0) * DT_LD + rt] = d22[i].x; DB[(cc + 1) * DT_LD + rt] = d22[i].y;
            DB[(cc + 2) * DT_LD + rt] = d22[i].z; DB[(cc + 3) * DT_LD + rt] = d22[i].w;
            float* xr = Xs + (rx + 16 * i) * XS_LD + 4 * qx;
            xr[0] = x21[i].x; xr[1] = x21[i].y; xr[2] = x21[i].z; xr[3] = x21[i].w;
        }
    }
    __syncthreads();
    PB_PH(1);
    const int r = tid & 63, cg = tid >> 6;                   // write-out / product mapping: row r, 16-column group cg
    auto fail_check = [&](int col0) {
        if (tid == 0) {
            const int f = reinterpret_cast<int*>(sm + F_PC)[300];
            if (f >= 0) atomicCAS(p.info, 0, col0 + (0x7fffffff - f) + 1);
        }
    };
    auto write_out = [&](int col0, int k, bool keep_m) {
        // L (lower) -> E, diag(L) -> diag_out, M = L^-T (upper, zeros below) -> winv[k] (+ row-major copy in DA)
        float* wi = p.winv + (size_t)k * DIAG_NB * DIAG_NB;
#pragma unroll
        for (int c4 = 0; c4 < 16; c4 += 4) {
            const int cc = 16 * cg + c4;
            float l[4], m[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                l[i] = Lt[(cc + i) * 64 + r];
                m[i] = (cc + i >= r) ? Yt[(cc + i) * 64 + r] : 0.f;
            }
            float* ep = p.E + (int64_t)(col0 + r) * p.ld + col0 + cc;
            if (cc + 3 <= r) {
                *reinterpret_cast<float4*>(ep) = make_float4(l[0], l[1], l[2], l[3]);
            } else {
#pragma unroll
                for (int i = 0; i < 4; ++i)
                    if (cc + i <= r) ep[i] = l[i];
            }
            if (p.diag_out && r >= cc && r < cc + 4) p.diag_out[col0 + r] = l[r - cc];
            *reinterpret_cast<float4*>(wi + r * DIAG_NB + cc) = make_float4(m[0], m[1], m[2], m[3]);
            if (keep_m) *reinterpret_cast<float4*>(DA + r * DT_LD + cc) = make_float4(m[0], m[1], m[2], m[3]);
        }
    };
    // ---- panel 2J ----
#ifdef PB_SWEEP_TIMING
    if (tid == 0) { s_fphase[0] = s_fphase[1] = s_fphase[2] = 0; }
#endif
    if (tid < 128) factor64(sm, DA);
    __syncthreads();
#ifdef PB_SWEEP_TIMING
    if (tid == 0 && J == 2) { g_fphase[0] = s_fphase[0]; g_fphase[1] = s_fphase[1]; g_fphase[2] = s_fphase[2]; }
#endif
    PB_PH(2);
    fail_check(c0);
    write_out(c0, k0, true);
    __threadfence();
    __syncthreads();
    if (tid == 0) st_release(p.nfact, k0 + 1);
    PB_PH(3);
    // ---- L21 = X21 M11 (M11 upper triangular: k <= column) ----
    float acc[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) acc[i] = 0.f;
    {
        const int kend = 16 * cg + 16;
#pragma unroll 4
        for (int k = 0; k < kend; ++k) {
            const float xv = Xs[r * XS_LD + k];
            const float4* mv = reinterpret_cast<const float4*>(DA + k * DT_LD + 16 * cg);
            const float4 m0 = mv[0], m1 = mv[1], m2 = mv[2], m3 = mv[3];
            acc[0] = fmaf(xv, m0.x, acc[0]);   acc[1] = fmaf(xv, m0.y, acc[1]);   acc[2] = fmaf(xv, m0.z, acc[2]);   acc[3] = fmaf(xv, m0.w, acc[3]);
            acc[4] = fmaf(xv, m1.x, acc[4]);   acc[5] = fmaf(xv, m1.y, acc[5]);   acc[6] = fmaf(xv, m1.z, acc[6]);   acc[7] = fmaf(xv, m1.w, acc[7]);
            acc[8] = fmaf(xv, m2.x, acc[8]);   acc[9] = fmaf(xv, m2.y, acc[9]);   acc[10] = fmaf(xv, m2.z, acc[10]); acc[11] = fmaf(xv, m2.w, acc[11]);
            acc[12] = fmaf(xv, m3.x, acc[12]); acc[13] = fmaf(xv, m3.y, acc[13]); acc[14] = fmaf(xv, m3.z, acc[14]); acc[15] = fmaf(xv, m3.w, acc[15]);
        }
    }
    __syncthreads();                                         // every X21 entry has been read
    {
        float* ep = p.E + (int64_t)(c0 + 64 + r) * p.ld + c0 + 16 * cg;
#pragma unroll
        for (int c4 = 0; c4 < 16; c4 += 4) {
            *reinterpret_cast<float4*>(ep + c4) = make_float4(acc[c4], acc[c4 + 1], acc[c4 + 2], acc[c4 + 3]);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                Xs[r * XS_LD + 16 * cg + c4 + i] = acc[c4 + i];
                Lt21[(16 * cg + c4 + i) * 64 + r] = acc[c4 + i];
            }
        }
    }
    __threadfence();
    __syncthreads();
    if (tid == 0) st_release(p.progress + J, k0 + 1);         // rows 64..127: B rows of the other tiles' U tasks
    PB_PH(4);
    // ---- D22 -= L21 L21^T (lower part), in its transposed shared-memory copy ----
    if (16 * cg <= r) {
#pragma unroll
        for (int i = 0; i < 16; ++i) acc[i] = 0.f;
#pragma unroll 4
        for (int k = 0; k < 64; ++k) {
            const float xv = Xs[r * XS_LD + k];
            const float4* lv = reinterpret_cast<const float4*>(Lt21 + k * 64 + 16 * cg);
            const float4 l0 = lv[0], l1 = lv[1], l2 = lv[2], l3 = lv[3];
            acc[0] = fmaf(xv, l0.x, acc[0]);   acc[1] = fmaf(xv, l0.y, acc[1]);   acc[2] = fmaf(xv, l0.z, acc[2]);   acc[3] = fmaf(xv, l0.w, acc[3]);
            acc[4] = fmaf(xv, l1.x, acc[4]);   acc[5] = fmaf(xv, l1.y, acc[5]);   acc[6] = fmaf(xv, l1.z, acc[6]);   acc[7] = fmaf(xv, l1.w, acc[7]);
            acc[8] = fmaf(xv, l2.x, acc[8]);   acc[9] = fmaf(xv, l2.y, acc[9]);   acc[10] = fmaf(xv, l2.z, acc[10]); acc[11] = fmaf(xv, l2.w, acc[11]);
            acc[12] = fmaf(xv, l3.x, acc[12]); acc[13] = fmaf(xv, l3.y, acc[13]); acc[14] = fmaf(xv, l3.z, acc[14]); acc[15] = fmaf(xv, l3.w, acc[15]);
        }
#pragma unroll
        for (int i = 0; i < 16; ++i) DB[(16 * cg + i) * DT_LD + r] -= acc[i];
    }
    __syncthreads();
    PB_PH(5);
    // ---- panel 2J+1 ----
    if (tid < 128) factor64(sm, DB);
    __syncthreads();
    PB_PH(6);
    fail_check(c0 + 64);
    write_out(c0 + 64, k0 + 1, false);
    __threadfence();
    __syncthreads();
    if (tid == 0) {
        st_release(p.nfact, k0 + 2);
        st_release(p.progress + J, min(k0 + 2, p.NP));
    }
    PB_PH(7);
}

__device__ __forceinline__ bool fused_diag_enabled(const SweepParams& p) { return p.fused != 0; }

__device__ void solve_panel_tile(const SweepParams& p, Cta& c, int I, int k) {
    const int c0 = k * DIAG_NB, nb = min(DIAG_NB, p.n - c0);
    const int row_lo = c0 + nb;
    if (max(row_lo, I * TM) >= min(I * TM + TM, p.rows)) return;       // no rows below the diagonal block here
    const float* wi = p.winv + (size_t)k * DIAG_NB * DIAG_NB;
    if (nb == DIAG_NB) {
        solve_tile_tc(p, c, I, c0, row_lo, wi);
    } else {
        // ragged last panel: CUDA-core version, coordinates relative to the diagonal block's origin
        panel_solve_tile(p.E + (int64_t)c0 * p.ld + c0, p.ld, I * TM - c0, p.rows - c0, nb, wi,
                         reinterpret_cast<float*>(dyn_smem));
        __threadfence();
        __syncthreads();
    }
}

__global__ void __launch_bounds__(NTH, 2) chol_sweep_kernel(const SweepParams p) {
    __shared__ __align__(8) uint64_t mma_bar[STAGES];
    __shared__ __align__(8) uint64_t solve_bar;
    __shared__ uint32_t tmem_base_holder;
    __shared__ Task s_task;
    __shared__ Sched s_sched;
    __shared__ int s_prog[MAX_RT];

    const int tid = threadIdx.x, warp = tid >> 5;
    const int g = blockIdx.x, G = gridDim.x;
    Cta c;
    c.mma_bar = mma_bar;
    c.solve_bar = &solve_bar;
    c.slabs = 0;
    c.solves = 0;

    if (tid == 0) {
        for (int s = 0; s < STAGES; ++s) mbar_init(&mma_bar[s], 1);
        mbar_init(&solve_bar, 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        Sched s0;
        s0.nJ = 0; s0.nPh = 0; s0.nI = 0; s0.near_done = 0;
        s0.fS = 0; s0.fP = 1; s0.fI = 0; s0.far_done = 0; s0.nfact = 0;
        s_sched = s0;
    }
    for (int i = tid; i < p.RT; i += NTH) s_prog[i] = 0;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_holder)),
                     "r"((uint32_t)TMEM_COLS)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    c.tmem = tmem_base_holder;

    while (true) {
#ifdef PB_SWEEP_TIMING
        unsigned long long t_pick = 0, t_start = 0;
        if (tid == 0) t_pick = gtime();
#endif
        if (tid == 0) {
            Sched s = s_sched;
            Task t;
            t.type = T_DONE; t.I = 0; t.J = 0; t.Jsrc = 0;
            auto prog_ge = [&](int I, int want) {
                if (s_prog[I] >= want) return true;
                const int v = ld_acquire(p.progress + I);
                s_prog[I] = v;
                return v >= want;
            };
            auto done_val = [&](int J) { return min(2 * J + 2, p.NP); };
            for (unsigned spins = 0;; ++spins) {
                const bool nv = !s.near_done && near_advance(p, g, G, s);
                const bool fv = !s.far_done && far_advance(p, g, G, s);
                if (!nv && !fv) break;
                bool blocked = false;
                int kneed = 0;
                if (nv) {
                    const int J = s.nJ, I = s.nI, ph = s.nPh;
                    kneed = 2 * J + (ph == 2 ? 1 : 0);
                    // this tile's own rank-128 updates come first: far cursor <= (J-1, J, I) means one is pending
                    blocked = fv && (s.fS < J - 1 || (s.fS == J - 1 && (s.fP < J || (s.fP == J && s.fI <= I))));
                    if (!blocked) {
                        bool ready = true;
                        if (I != J) {
                            s.nfact = max(s.nfact, ld_acquire(p.nfact));
                            if (ph == 0) ready = s.nfact >= 2 * J + 1;
                            else if (ph == 1) ready = prog_ge(J, 2 * J + 1);
                            else ready = s.nfact >= 2 * J + 2;
                        }
                        if (ready) {
                            t.type = T_S0 + ph; t.I = I; t.J = J;
                            s.nI = I + 1;
                            break;
                        }
                    }
                }
                if (fv) {
                    const bool allow = !nv || blocked || kneed > s.nfact + 1;
                    if (allow && prog_ge(s.fI, done_val(s.fS)) && prog_ge(s.fP, done_val(s.fS))) {
                        t.type = T_FAR; t.I = s.fI; t.J = s.fP; t.Jsrc = s.fS;
                        s.fI = s.fI + 1;
                        break;
                    }
                }
                __nanosleep(100);
                if (spins > (1u << 24)) __trap();        // ~2 s without progress: fail loudly instead of hanging the GPU
            }
            s_sched = s;
            s_task = t;
        }
        __syncthreads();
        const Task t = s_task;
        if (t.type == T_DONE) break;
#ifdef PB_SWEEP_TIMING
        unsigned long long clk_start = 0;
        if (tid == 0) { t_start = gtime(); clk_start = clock64(); }
#endif
        if (t.type == T_FAR) {
            tile_update(p, c, t.I * TM, t.Jsrc * TM, TM, t.J * TM, min(TM, p.n - t.J * TM), (t.Jsrc + 1) * TM);
        } else if (t.type == T_U && !(t.I == t.J && TM * t.J + TM <= p.n && fused_diag_enabled(p))) {
            const int cc0 = t.J * TM + DIAG_NB;
            tile_update(p, c, t.I * TM, t.J * TM, DIAG_NB, cc0, min(DIAG_NB, p.n - cc0), cc0);
        } else if (t.I == t.J && TM * t.J + TM <= p.n && fused_diag_enabled(p)) {
            // regular diagonal tile: S0 runs the whole fused routine, its U and S1 entries have nothing left to do
            if (t.type == T_S0) diag_tile_fused(p, c, t.J);
        } else {
            const int k = 2 * t.J + (t.type == T_S1 ? 1 : 0);
            if (t.I == t.J) factor_panel(p, c, k);
            solve_panel_tile(p, c, t.I, k);
            const bool last_of_column = (t.type == T_S1) || (k + 1 >= p.NP) || (TM * t.J + DIAG_NB >= p.n);
            if (tid == 0) {
                if (last_of_column) st_release(p.progress + t.I, min(2 * t.J + 2, p.NP));
                else if (t.I == t.J) st_release(p.progress + t.I, k + 1);   // rows 64..127 of the diagonal tile: B rows of U
            }
        }
#ifdef PB_SWEEP_TIMING
        if (tid == 0) {
            const int slot = atomicAdd(&g_log_n, 1);
            if (slot < LOG_CAP) {
                LogRec r;
                r.cta = g; r.type = t.type; r.I = t.I; r.J = t.type == T_FAR ? t.J * 1000 + t.Jsrc : t.J;
                r.t_pick = t_pick; r.t_start = t_start; r.t_end = gtime();
                r.clk_start = clk_start; r.clk_end = clock64();
                for (int i = 0; i < 8; ++i) r.ph[i] = s_ph[i];
                g_log[slot] = r;
            }
        }
#endif
        __syncthreads();
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) {
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(c.tmem), "r"((uint32_t)TMEM_COLS)
                     : "memory");
    }
}

}  // namespace

#ifdef PB_SWEEP_TIMING
extern "C" __attribute__((visibility("default"))) int pinot_b200_debug_sweep_log(void* out, int cap_records, int reset) {
    int n = 0;
    cudaMemcpyFromSymbol(&n, g_log_n, sizeof(int));
    if (n > LOG_CAP) n = LOG_CAP;
    if (n > cap_records) n = cap_records;
    if (out && n > 0) cudaMemcpyFromSymbol(out, g_log, sizeof(LogRec) * (size_t)n);
    if (reset) { int z = 0; cudaMemcpyToSymbol(g_log_n, &z, sizeof(int)); }
    if (out && cap_records == -1) cudaMemcpyFromSymbol(out, g_fphase, sizeof(g_fphase));
    if (out && cap_records == -2) cudaMemcpyFromSymbol(out, g_wstamp, sizeof(g_wstamp));
    return n;
}
#endif

bool sweep_persistent_enabled() {
    // Opt-in: PINOT_B200_SWEEP=persistent.  Measured on B200 (profiles/r02/sweep_single_launch.md) the single launch
    // is correct but not yet faster than the launch-per-panel sweep (n = 4200: 2.75 vs 2.00 ms), so the default
    // stays the launch chain.  Read on every call so that tests can switch within one process.
    const char* v = getenv("PINOT_B200_SWEEP");
    return v && v[0] == 'p';
}

int64_t sweep_persistent_flag_bytes(int64_t n, int64_t n_ext_dense) {
    // progress counter per 128-row tile (worst case: identity rows present) + the factored-panel counter
    const int64_t rt = ceil_div(2 * n + n_ext_dense, TM);
    return (int64_t)sizeof(int) * (rt + 64);
}

bool sweep_persistent_applicable(int64_t n, int64_t rows_total) {
    // size test only (the workspace query must not depend on the environment); the dispatch also asks enabled()
    return gemm_uses_tensor_cores() && n >= 1 && n < SWEEP_PERSISTENT_MAX_N && ceil_div(rows_total, TM) <= MAX_RT;
}

int launch_sweep_persistent(float* e, int64_t ld, int64_t n, int64_t n_ext_dense, int ext_identity, float* diag_out,
                            float* wdiag, float* winv, int* flags, int* info, cudaStream_t stream) {
    SweepParams p;
    p.E = e; p.ld = ld; p.n = (int)n; p.nd = (int)n_ext_dense; p.ident = ext_identity ? 1 : 0;
    p.rows = (int)(n + n_ext_dense + (ext_identity ? n : 0));
    p.NP = (int)ceil_div(n, DIAG_NB);
    p.CT = (int)ceil_div(n, TM);
    p.RT = (int)ceil_div(p.rows, TM);
    p.wdiag = wdiag; p.winv = winv; p.nfact = flags; p.progress = flags + 32;
    p.info = info; p.diag_out = diag_out;
    static const int fused = [] { const char* v = getenv("PINOT_B200_SWEEP_FUSED_DIAG"); return (v && v[0] == '0') ? 0 : 1; }();
    p.fused = fused;
    PB_CUDA(cudaMemsetAsync(flags, 0, sizeof(int) * (size_t)(p.RT + 32), stream));
    static int grid_cache[64] = {0};
    int dev = 0;
    PB_CUDA(cudaGetDevice(&dev));
    if (dev < 0 || dev >= 64 || grid_cache[dev] == 0) {
        PB_CUDA(cudaFuncSetAttribute(chol_sweep_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, DYN_SMEM));
        int per_sm = 0, sms = 0;
        PB_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, chol_sweep_kernel, NTH, DYN_SMEM));
        PB_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
        PB_REQUIRE(per_sm >= 1 && sms >= 1, "potrf_ext: the persistent sweep kernel does not fit this device");
        const int gsz = (per_sm > 2 ? 2 : per_sm) * sms;
        if (dev >= 0 && dev < 64) grid_cache[dev] = gsz;
        else return PB_ERR_ARG;
    }
    // no more CTAs than tiles: the rest would only allocate tensor memory and leave
    int64_t tiles = (int64_t)p.RT * p.CT;
    int grid = grid_cache[dev];
    if (tiles < grid) grid = (int)(tiles < 1 ? 1 : tiles);
    void* args[] = {(void*)&p};
    PB_CUDA(cudaLaunchCooperativeKernel((const void*)chol_sweep_kernel, dim3((unsigned)grid), dim3(NTH), args,
                                        (size_t)DYN_SMEM, stream));
    return check_launch("chol_sweep (persistent)");
}

}  // namespace pb
