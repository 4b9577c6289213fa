// TMA-fed, warp-specialised tcgen05 GEMM for LARGE products  C = alpha A B^T (+ beta C ...)  with A (M x K)
// and B (N x K) both K-contiguous ("NT").
//
// The register-staged kernel in gemm_tc.cu spends its instruction-issue budget on loading,
// splitting (3xTF32) and storing operands -- every CTA of a tile row re-splits the same A slab.
// Here the split happens ONCE per operand, in a memory-bound pre-pass into caller-provided scratch
// (a_hi, a_lo, b_hi, b_lo: tf32-valued fp32 arrays), and the GEMM itself contains no loader code:
//   warp 0   one thread issues cp.async.bulk.tensor (TMA) loads of 128 x 32 tiles of the four arrays
//            into a 3-stage ring of 128B-swizzled shared-memory tiles (64 KB per stage);
//   warp 1   one thread issues tcgen05.mma.kind::tf32 (M = 128, N = 128, K = 8; 12 per stage:
//            lo.hi + hi.lo into the cross-term accumulator, hi.hi into the main one) and releases
//            each stage with tcgen05.commit;
//   warps 2-5 drain the accumulators: tcgen05.ld, transpose through shared memory, fused epilogue.
// The accumulators are double-buffered in TMEM (2 x 256 columns) and swapped every 512 K-columns
// (TMEM accumulation truncates, see gemm_tc.cu), so flushing chunk c overlaps the MMAs of chunk c+1.
#include "gemm_params.cuh"
#include "gemm_epilogue.cuh"
#include "tma_common.cuh"
#include <cuda.h>
#include "../../include/pinot_b200.h"
#include <stdlib.h>

namespace pb {

namespace {

using namespace tc;

constexpr int BM = 128, BN = 128, BK = 32, NSTAGE = 3;
constexpr int TILE_BYTES = BM * BK * 4;            // 16 KB: 128 rows x 128 bytes, SWIZZLE_128B
constexpr int STAGE_BYTES = 4 * TILE_BYTES;        // a_hi | a_lo | b_hi | b_lo
constexpr int EPI_BYTES = 4 * 32 * EPI_LD * 4;     // one 32x32 staging tile per epilogue warp
constexpr int SMEM_BYTES = NSTAGE * STAGE_BYTES + EPI_BYTES + 1024;
constexpr int NTHREADS = 192;                      // producer, MMA, 4 epilogue warps
constexpr int KCHUNK = 512;
constexpr int STAGES_PER_CHUNK = KCHUNK / BK;
constexpr int NSCHED = 4;                          // depth of the tile hand-off ring

// Live tiles in dispatch order.  Without `lower` the gm x gn rectangle row by row; with it only tiles on or
// below the diagonal (row i has min(i + 1, gn) of them), still row by row -- for a ktri product that is
// longest-K first.
struct TileGrid {
    int gm, gn, lower;
    __host__ __device__ int count() const {
        if (!lower) return gm * gn;
        const int t = gm < gn ? gm : gn;
        return t * (t + 1) / 2 + (gm > gn ? (gm - gn) * gn : 0);
    }
    __device__ void locate(int t, int& i, int& j) const {
        if (!lower) { i = t / gn; j = t - i * gn; return; }
        const int tri = gn * (gn + 1) / 2;
        if (t < tri) {
            i = (int)((sqrtf(8.f * (float)t + 1.f) - 1.f) * 0.5f);
            while ((i + 1) * (i + 2) / 2 <= t) ++i;
            while (i * (i + 1) / 2 > t) --i;
            j = t - i * (i + 1) / 2;
        } else {
            const int r = t - tri;
            i = gn + r / gn;
            j = r - (r / gn) * gn;
        }
    }
};

struct TileWork {
    int m0, n0, kb, num_k, num_chunks;
};
__device__ __forceinline__ TileWork tile_work(const KParams& p, const TileGrid& tg, int t, int spc = STAGES_PER_CHUNK) {
    int i, j;
    tg.locate(t, i, j);
    TileWork w;
    w.m0 = i * BM;
    w.n0 = j * BN;
    w.kb = p.ktri ? (w.m0 / BK) * BK : 0;
    w.num_k = (p.K > w.kb) ? (p.K - w.kb + BK - 1) / BK : 0;
    w.num_chunks = (w.num_k + spc - 1) / spc;   // 0 when there is nothing to multiply
    return w;
}

// Persistent with a dynamic tile queue: the producer thread draws the next live tile from a global counter
// (zeroed by the host before the launch) and hands it to the MMA and epilogue warps through a 4-slot ring in
// shared memory -- a ktri product has tiles of very different length, which a static assignment balances
// badly.  The three roles keep their own running counters (stage ring position, accumulator-set use count),
// so the drain of one tile's last K-chunk overlaps the MMAs of the next tile's first one.
// ACC: the accurate variant.  The tensor core truncates when it accumulates, so the error of a product grows with
// the number of MMAs between two flushes of the accumulators (scripts/tc_accumulate_bias.py).  Here a flush happens
// every 128 K-columns (16 MMAs) and goes into REGISTERS -- each epilogue thread keeps the 128 running sums of its
// row, added in fp32 round-to-nearest -- so that C is read and written once per tile, as in the 512-column variant,
// but with the accumulation error of 128-column products (profiles/r02/trsm_outer_block.md).
template <int EPI, bool ACC = false>
__global__ void __launch_bounds__(NTHREADS, 1)
tgemm_tma_kernel(const KParams p, const TileGrid tg, int* __restrict__ tile_counter,
                 const __grid_constant__ CUtensorMap map_ah,
                 const __grid_constant__ CUtensorMap map_al, const __grid_constant__ CUtensorMap map_bh,
                 const __grid_constant__ CUtensorMap map_bl) {
    extern __shared__ __align__(1024) uint8_t dyn_smem[];
    __shared__ __align__(8) uint64_t full_bar[NSTAGE], empty_bar[NSTAGE], acc_full[2], acc_empty[2];
    __shared__ __align__(8) uint64_t sched_full[NSCHED], sched_empty[NSCHED];
    __shared__ int sched_tile[NSCHED];
    __shared__ uint32_t tmem_base_holder;

    constexpr int SPC = ACC ? 128 / BK : STAGES_PER_CHUNK;      // stages between two flushes of the accumulators
    const int ntiles = tg.count();
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    // aligned by offset so that the pointer keeps its shared address space (LDS / STS instead of generic LD / ST)
    uint8_t* smem = dyn_smem + ((1024u - (tc::s32(dyn_smem) & 1023u)) & 1023u);
    float* epi_smem = reinterpret_cast<float*>(smem + NSTAGE * STAGE_BYTES);

    if (tid == 0) {
        for (int s = 0; s < NSTAGE; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&acc_full[s], 1);
            mbar_init(&acc_empty[s], 128);
        }
        for (int s = 0; s < NSCHED; ++s) {
            mbar_init(&sched_full[s], 1);
            mbar_init(&sched_empty[s], 129);   // the MMA thread and the 128 epilogue threads
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_ah) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_al) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_bh) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_bl) : "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&tmem_base_holder)),
                     "r"(512u)
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
            uint32_t it = 0;   // stages issued so far
            for (uint32_t ns = 0;; ++ns) {
                const uint32_t slot = ns % NSCHED;
                if (ns >= NSCHED) mbar_wait(&sched_empty[slot], (ns / NSCHED - 1) & 1);
                const int t = atomicAdd(tile_counter, 1);
                *reinterpret_cast<volatile int*>(&sched_tile[slot]) = t;
                mbar_arrive(&sched_full[slot]);
                if (t >= ntiles) break;
                const TileWork w = tile_work(p, tg, t, SPC);
                for (int kt = 0; kt < w.num_k; ++kt, ++it) {
                    const uint32_t s = it % NSTAGE;
                    if (it >= NSTAGE) mbar_wait(&empty_bar[s], (it / NSTAGE - 1) & 1);
                    uint8_t* st = smem + s * STAGE_BYTES;
                    mbar_expect_tx(&full_bar[s], STAGE_BYTES);
                    const int kc = w.kb + kt * BK;
                    tma_load_2d(st, &map_ah, kc, w.m0, &full_bar[s]);
                    tma_load_2d(st + TILE_BYTES, &map_al, kc, w.m0, &full_bar[s]);
                    tma_load_2d(st + 2 * TILE_BYTES, &map_bh, kc, w.n0, &full_bar[s]);
                    tma_load_2d(st + 3 * TILE_BYTES, &map_bl, kc, w.n0, &full_bar[s]);
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
            uint32_t it = 0, gc = 0;   // stages / accumulator sets consumed so far
            for (uint32_t ns = 0;; ++ns) {
                const uint32_t slot = ns % NSCHED;
                mbar_wait(&sched_full[slot], (ns / NSCHED) & 1);
                const int t = *reinterpret_cast<volatile int*>(&sched_tile[slot]);
                mbar_arrive(&sched_empty[slot]);
                if (t >= ntiles) break;
                const TileWork w = tile_work(p, tg, t, SPC);
                for (int chunk = 0; chunk < w.num_chunks; ++chunk, ++gc) {
                    const uint32_t set = gc & 1;
                    if (gc >= 2) mbar_wait(&acc_empty[set], (gc / 2 - 1) & 1);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t acc_hi = tmem_d + set * 256u, acc_x = acc_hi + 128u;
                    const int kt_beg = chunk * SPC, kt_end = min(w.num_k, kt_beg + SPC);
                    for (int kt = kt_beg; kt < kt_end; ++kt, ++it) {
                        const uint32_t s = it % NSTAGE;
                        mbar_wait(&full_bar[s], (it / NSTAGE) & 1);
                        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                        const uint32_t sb = s32(smem + s * STAGE_BYTES);
                        const uint64_t d_ah = sw128_desc(sb), d_al = sw128_desc(sb + TILE_BYTES);
                        const uint64_t d_bh = sw128_desc(sb + 2 * TILE_BYTES), d_bl = sw128_desc(sb + 3 * TILE_BYTES);
#pragma unroll
                        for (int k8 = 0; k8 < BK / 8; ++k8) {
                            const uint64_t adv = (uint64_t)((k8 * 32) >> 4);   // +32 bytes inside the swizzle atom
                            const uint32_t acc = (kt > kt_beg || k8 > 0) ? 1u : 0u;
                            mma_tf32(acc_x, d_al + adv, d_bh + adv, idesc, acc);
                            mma_tf32(acc_x, d_ah + adv, d_bl + adv, idesc, 1u);
                            mma_tf32(acc_hi, d_ah + adv, d_bh + adv, idesc, acc);
                        }
                        mma_commit(&empty_bar[s]);     // stage free once these MMAs have read it
                    }
                    mma_commit(&acc_full[set]);        // accumulators of this chunk complete
                }
            }
        }
    } else {
        // ===================== epilogue warps =====================
        float alpha = p.alpha, r1c = p.r1coef;
        if (p.alpha_dev) {
            const float s = __ldg(p.alpha_dev);
            alpha *= s;
            r1c *= s;
        }
        const int q = warp & 3;                       // TMEM lane quadrant this warp may access
        float* tile = epi_smem + (warp - 2) * (32 * EPI_LD);
        uint32_t gc = 0;
        for (uint32_t ns = 0;; ++ns) {
            const uint32_t slot = ns % NSCHED;
            mbar_wait(&sched_full[slot], (ns / NSCHED) & 1);
            const int t = *reinterpret_cast<volatile int*>(&sched_tile[slot]);
            mbar_arrive(&sched_empty[slot]);
            if (t >= ntiles) break;
            const TileWork w = tile_work(p, tg, t, SPC);
            if constexpr (ACC) {
                float run[BN];                        // this row's running sums, one per column of the tile
#pragma unroll
                for (int c = 0; c < BN; ++c) run[c] = 0.f;
                for (int chunk = 0; chunk < w.num_chunks; ++chunk, ++gc) {
                    const uint32_t set = gc & 1;
                    mbar_wait(&acc_full[set], (gc / 2) & 1);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
#pragma unroll
                    for (int c0 = 0; c0 < BN; c0 += 32) {
                        uint32_t v[32], x[32];
                        const uint32_t taddr = tmem_d + ((uint32_t)(q * 32) << 16) + set * 256u + (uint32_t)c0;
                        tc::tmem_ld_32x32b_x32(v, taddr);
                        tc::tmem_ld_32x32b_x32(x, taddr + 128u);
                        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
                        for (int i = 0; i < 32; ++i) run[c0 + i] += __uint_as_float(v[i]) + __uint_as_float(x[i]);
                    }
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    mbar_arrive(&acc_empty[set]);
                }
                // one pass over C: stage 32 columns at a time (this thread's row of the warp's 32 x 32 block)
#pragma unroll
                for (int c0 = 0; c0 < BN; c0 += 32) {
#pragma unroll
                    for (int i = 0; i < 8; ++i)
                        *reinterpret_cast<float4*>(tile + lane * EPI_LD + 4 * i) =
                            make_float4(run[c0 + 4 * i], run[c0 + 4 * i + 1], run[c0 + 4 * i + 2], run[c0 + 4 * i + 3]);
                    __syncwarp();
                    epilogue_rows<EPI>(p, tile, alpha, p.beta, r1c, true, w.m0, w.n0, q, c0, lane);
                    __syncwarp();
                }
                continue;
            }
            const int passes = w.num_chunks > 0 ? w.num_chunks : 1;   // an empty product still applies beta / bias
            for (int chunk = 0; chunk < passes; ++chunk) {
                const bool have_acc = w.num_chunks > 0;
                const uint32_t set = gc & 1;
                const bool last_chunk = (chunk == passes - 1);
                const float beta = (chunk == 0) ? p.beta : 1.f;
                if (have_acc) {
                    mbar_wait(&acc_full[set], (gc / 2) & 1);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                }
#pragma unroll 1
                for (int c0 = 0; c0 < BN; c0 += 32) {
                    uint32_t v[32], x[32];
                    if (have_acc) {
                        const uint32_t taddr = tmem_d + ((uint32_t)(q * 32) << 16) + set * 256u + (uint32_t)c0;
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
                            : "=r"(x[0]), "=r"(x[1]), "=r"(x[2]), "=r"(x[3]), "=r"(x[4]), "=r"(x[5]), "=r"(x[6]), "=r"(x[7]),
                              "=r"(x[8]), "=r"(x[9]), "=r"(x[10]), "=r"(x[11]), "=r"(x[12]), "=r"(x[13]), "=r"(x[14]), "=r"(x[15]),
                              "=r"(x[16]), "=r"(x[17]), "=r"(x[18]), "=r"(x[19]), "=r"(x[20]), "=r"(x[21]), "=r"(x[22]), "=r"(x[23]),
                              "=r"(x[24]), "=r"(x[25]), "=r"(x[26]), "=r"(x[27]), "=r"(x[28]), "=r"(x[29]), "=r"(x[30]), "=r"(x[31])
                            : "r"(taddr + 128u)
                            : "memory");
                        asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    } else {
#pragma unroll
                        for (int jj = 0; jj < 32; ++jj) { v[jj] = 0u; x[jj] = 0u; }
                    }
#pragma unroll
                    for (int i = 0; i < 8; ++i) {
                        float4 o;
                        o.x = __uint_as_float(v[4 * i + 0]) + __uint_as_float(x[4 * i + 0]);
                        o.y = __uint_as_float(v[4 * i + 1]) + __uint_as_float(x[4 * i + 1]);
                        o.z = __uint_as_float(v[4 * i + 2]) + __uint_as_float(x[4 * i + 2]);
                        o.w = __uint_as_float(v[4 * i + 3]) + __uint_as_float(x[4 * i + 3]);
                        *reinterpret_cast<float4*>(tile + lane * EPI_LD + 4 * i) = o;
                    }
                    __syncwarp();
                    epilogue_rows<EPI>(p, tile, alpha, beta, r1c, last_chunk, w.m0, w.n0, q, c0, lane);
                    __syncwarp();
                }
                if (have_acc) {
                    // this accumulator set may be overwritten by the chunk after next
                    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
                    mbar_arrive(&acc_empty[set]);
                    ++gc;
                }
            }
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(512u) : "memory");
    }
}

// hi = rna_tf32(v), lo = rna_tf32(v - hi) with v = x * f(scale[col]); f = identity (mode 1) or sqrt (mode 2: the
// same-operand form of A diag(s) A^T).  tri != 0: x is upper triangular and only columns >= the origin of the
// row's 128-row tile are ever loaded by the GEMM (ktri), so the rest is neither read nor written.
__global__ void __launch_bounds__(256)
split_tf32_kernel(const float* __restrict__ x, int64_t ldx, float* __restrict__ hi, float* __restrict__ lo, int64_t ldo,
                  int64_t rows, int cols4, const float* __restrict__ colscale, int scale_mode, int tri) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t r = t / cols4;
    if (r >= rows) return;
    const int c = (int)(t - r * cols4) * 4;
    if (tri && c + 4 <= (int)((r / BM) * BM)) return;
    float4 v = __ldg(reinterpret_cast<const float4*>(x + r * ldx + c));
    if (scale_mode) {
        float4 s = __ldg(reinterpret_cast<const float4*>(colscale + c));
        if (scale_mode == 2) { s.x = sqrtf(s.x); s.y = sqrtf(s.y); s.z = sqrtf(s.z); s.w = sqrtf(s.w); }
        v.x *= s.x; v.y *= s.y; v.z *= s.z; v.w *= s.w;
    }
    float4 h, l;
    uint32_t u;
#define PB_SPLIT(F)                                                       \
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v.F));                 \
    h.F = __uint_as_float(u);                                             \
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v.F - h.F));           \
    l.F = __uint_as_float(u);
    PB_SPLIT(x) PB_SPLIT(y) PB_SPLIT(z) PB_SPLIT(w)
#undef PB_SPLIT
    *reinterpret_cast<float4*>(hi + r * ldo + c) = h;
    *reinterpret_cast<float4*>(lo + r * ldo + c) = l;
}

}  // namespace

// hi / lo copies plus 64 floats (256 bytes) for the tile counter, which sits FIRST
int64_t gemm_tma_scratch_floats(int64_t M, int64_t N, int64_t K, bool same_operand) {
    const int64_t ldk = (K + 3) / 4 * 4;
    return TMA_COUNTER_FLOATS + 2 * M * ldk + (same_operand ? 0 : 2 * N * ldk);
}

bool gemm_tma_enabled() {
    static const bool on = [] {
        const char* e = getenv("PINOT_B200_TMA");
        return !(e && e[0] == '0');
    }();
    return on && gemm_uses_tensor_cores();
}

int tma_split_operand(const float* x, int64_t ldx, int64_t rows, int64_t K, float* hi, float* lo, const float* colscale,
                      int scale_mode, int tri, cudaStream_t stream) {
    const int64_t ldk = (K + 3) / 4 * 4;
    PB_REQUIRE(ldx % 4 == 0 && ldx >= ldk && aligned16(x) && aligned16(hi) && aligned16(lo),
               "gemm (TMA): operand rows must be 16-byte aligned and padded to a multiple of 4 columns");
    PB_REQUIRE(scale_mode == 0 || (colscale && aligned16(colscale)), "gemm (TMA): column scale must be 16-byte aligned");
    const int cols4 = (int)(ldk / 4);
    split_tf32_kernel<<<(unsigned)ceil_div(rows * cols4, 256), 256, 0, stream>>>(x, ldx, hi, lo, ldk, rows, cols4, colscale,
                                                                                 scale_mode, tri);
    return check_launch("split_tf32");
}

// C (M x N) = alpha A B^T (+ ...) from pre-split operands: a_* (M x K), b_* (N x K), leading dimension ceil4(K).
int launch_gemm_tma_presplit(const KParams& p, const float* a_hi, const float* a_lo, const float* b_hi, const float* b_lo,
                             int* tile_counter, cudaStream_t stream, int64_t ld_a, int64_t ld_b, bool accurate) {
    const int64_t M = p.M, N = p.N, K = p.K;
    const int64_t ldk = (K + 3) / 4 * 4;
    if (ld_a <= 0) ld_a = ldk;
    if (ld_b <= 0) ld_b = ldk;
    PB_REQUIRE(ld_a % 4 == 0 && ld_b % 4 == 0 && ld_a >= ldk && ld_b >= ldk && aligned16(a_hi) && aligned16(a_lo) && aligned16(b_hi) &&
                   aligned16(b_lo),
               "gemm (TMA): pre-split operands must be 16-byte aligned with leading dimensions that are multiples of 4");
    PB_REQUIRE(p.vecC && p.ldc % 4 == 0 && aligned16(p.C), "gemm (TMA): C must be 16-byte aligned with ldc % 4 == 0");
    PB_REQUIRE(p.splitk == 1 && !p.post_wdiag, "gemm (TMA): split-K / look-ahead are not supported on this path");
    PB_REQUIRE(M > 0 && N > 0 && K > 0, "gemm (TMA): empty product");
    alignas(64) CUtensorMap mah, mal, mbh, mbl;
    int rc;
    if ((rc = make_map_box(&mah, a_hi, M, K, ld_a))) return rc;
    if ((rc = make_map_box(&mal, a_lo, M, K, ld_a))) return rc;
    if ((rc = make_map_box(&mbh, b_hi, N, K, ld_b))) return rc;
    if ((rc = make_map_box(&mbl, b_lo, N, K, ld_b))) return rc;
    TileGrid tg;
    tg.gm = (int)ceil_div(M, BM); tg.gn = (int)ceil_div(N, BN); tg.lower = p.lower;
    const int ntiles = tg.count();
    const unsigned grid = (unsigned)(ntiles < num_sms() ? ntiles : num_sms());
    PB_REQUIRE(tile_counter != nullptr, "gemm (TMA): tile counter missing");
    PB_CUDA(cudaMemsetAsync(tile_counter, 0, sizeof(int), stream));
    auto launch = [&](auto kern) -> int {
        PB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        kern<<<grid, NTHREADS, SMEM_BYTES, stream>>>(p, tg, tile_counter, mah, mal, mbh, mbl);
        return check_launch("gemm (tcgen05 + TMA)");
    };
    if (accurate) {
        PB_REQUIRE(!p.r1u && !p.mirror && !p.act && !p.ktri, "gemm (TMA): the accurate variant is the plain alpha / beta product");
        return launch(tgemm_tma_kernel<0, true>);
    }
    if (p.r1u || p.mirror) return launch(tgemm_tma_kernel<2>);
    if (p.act) return launch(tgemm_tma_kernel<1>);
    return launch(tgemm_tma_kernel<0>);
}

int launch_gemm_tma(const KParams& p, float* scratch, bool same_operand, cudaStream_t stream) {
    const int64_t M = p.M, N = p.N, K = p.K;
    const int64_t ldk = (K + 3) / 4 * 4;
    float* a_hi = scratch + TMA_COUNTER_FLOATS;
    float* a_lo = a_hi + M * ldk;
    float* b_hi = same_operand ? a_hi : a_lo + M * ldk;
    float* b_lo = same_operand ? a_lo : b_hi + N * ldk;
    int rc = tma_split_operand(p.A, p.lda, M, K, a_hi, a_lo, p.kscale, p.kscale ? 1 : 0, 0, stream);
    if (rc) return rc;
    if (!same_operand) {
        rc = tma_split_operand(p.B, p.ldb, N, K, b_hi, b_lo, nullptr, 0, 0, stream);
        if (rc) return rc;
    }
    return launch_gemm_tma_presplit(p, a_hi, a_lo, b_hi, b_lo, reinterpret_cast<int*>(scratch), stream);
}

}  // namespace pb

extern "C" {

int64_t pinot_b200_sgemm_nt_tma_workspace_bytes(int64_t M, int64_t N, int64_t K) {
    if (M < 0 || N < 0 || K < 0) return -1;
    return pb::gemm_tma_scratch_floats(M, N, K, false) * (int64_t)sizeof(float);
}

int pinot_b200_sgemm_nt_tma(int64_t M, int64_t N, int64_t K, float alpha, const float* A, int64_t lda, const float* B,
                            int64_t ldb, float beta, float* C, int64_t ldc, void* workspace, int64_t workspace_bytes,
                            cudaStream_t stream) {
    PB_REQUIRE(M > 0 && N > 0 && K > 0, "sgemm_nt_tma: empty product");
    PB_REQUIRE(M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31), "sgemm_nt_tma: size exceeds int32");
    PB_REQUIRE(A && B && C && workspace, "sgemm_nt_tma: null pointer");
    PB_REQUIRE(workspace_bytes >= pinot_b200_sgemm_nt_tma_workspace_bytes(M, N, K),
               "sgemm_nt_tma: workspace too small (%lld bytes given)", (long long)workspace_bytes);
    auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    const int64_t k4 = (K + 3) / 4 * 4;
    PB_REQUIRE(lda % 4 == 0 && ldb % 4 == 0 && ldc % 4 == 0 && lda >= k4 && ldb >= k4 && ldc >= N && al16(A) && al16(B) &&
                   al16(C) && al16(workspace),
               "sgemm_nt_tma: operands must be 16-byte aligned with leading dimensions that are multiples of 4 (>= K rounded up)");
    pb::KParams p{};
    p.M = (int)M; p.N = (int)N; p.K = (int)K;
    p.A = A; p.lda = lda; p.B = B; p.ldb = ldb; p.C = C; p.ldc = ldc;
    p.alpha = alpha; p.beta = beta; p.splitk = 1; p.vecA = p.vecB = p.vecC = 1;
    return pb::launch_gemm_tma(p, static_cast<float*>(workspace), false, stream);
}

int pinot_b200_tf32_split(const float* x, int64_t ldx, int64_t rows, int64_t K, const float* colscale, int scale_mode,
                          float* hi, float* lo, cudaStream_t stream) {
    PB_REQUIRE(rows >= 0 && K > 0 && rows < (1ll << 31) && K < (1ll << 31), "tf32_split: invalid sizes");
    PB_REQUIRE(scale_mode >= 0 && scale_mode <= 2, "tf32_split: scale_mode must be 0, 1 or 2");
    if (rows == 0) return pb::PB_OK;
    PB_REQUIRE(x && hi && lo, "tf32_split: null pointer");
    return pb::tma_split_operand(x, ldx, rows, K, hi, lo, colscale, scale_mode, 0, stream);
}

int pinot_b200_sgemm_nt_presplit(int64_t M, int64_t N, int64_t K, float alpha, const float* a_hi, const float* a_lo,
                                 const float* b_hi, const float* b_lo, float beta, float* C, int64_t ldc, int lower,
                                 int* tile_counter, cudaStream_t stream) {
    PB_REQUIRE(M > 0 && N > 0 && K > 0, "sgemm_nt_presplit: empty product");
    PB_REQUIRE(M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31), "sgemm_nt_presplit: size exceeds int32");
    PB_REQUIRE(a_hi && a_lo && b_hi && b_lo && C && tile_counter, "sgemm_nt_presplit: null pointer");
    auto al16 = [](const void* q) { return (reinterpret_cast<uintptr_t>(q) & 15) == 0; };
    PB_REQUIRE(al16(a_hi) && al16(a_lo) && al16(b_hi) && al16(b_lo) && al16(C) && ldc % 4 == 0 && ldc >= N,
               "sgemm_nt_presplit: operands must be 16-byte aligned with ldc a multiple of 4");
    pb::KParams p{};
    p.M = (int)M; p.N = (int)N; p.K = (int)K;
    p.C = C; p.ldc = ldc;
    p.alpha = alpha; p.beta = beta; p.lower = lower ? 1 : 0; p.splitk = 1; p.vecC = 1;
    return pb::launch_gemm_tma_presplit(p, a_hi, a_lo, b_hi, b_lo, tile_counter, stream);
}

}  // extern "C"
