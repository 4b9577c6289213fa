// Tall-and-skinny fp32-accurate GEMM for the node-update layers:  C (M x N) = act(A (M x K) op(B) + bias) (+ beta C)
// with N, K <= 128 and M in the 10^5 .. 10^7 range (one row per atom).  HBM-bound by design: A is read once, C
// written once, B (the layer weight, <= 64 KB) lives in shared memory for the whole kernel.
//
// Persistent CTA per SM, five roles, no thread ever waits on global memory latency:
//   warp 0      TMA producer: raw fp32 slabs of A (128 rows x 32 columns, SWIZZLE_128B) into a 2-slot landing ring
//               (B leaves room for no more; a slot is handed back as soon as it has been read, not when consumed);
//   warps 18-21 converters: split a landed slab into tf32 hi / lo operand buffers -- the swizzled layout is
//               position-preserving, so the split is a flat 16-byte-chunk loop -- and hand the landing slot back
//               to the producer at once;
//   warp 1      MMA issuer: 3 tcgen05.mma per k-step (lo.hi + hi.lo -> cross accumulator, hi.hi -> main one)
//               against the resident pre-split B, tcgen05.commit frees the operand buffer;
//   warps 2-17  epilogue: the product is accumulated TRANSPOSED (weights on the M side), so tcgen05.ld gives each
//               thread one output column and 32 consecutive rows; bias / tanh / beta, then one 4-byte store per row:
//               a warp's request is 128 contiguous bytes of one row (one L1 wavefront); four warps per lane quadrant;
// accumulators double-buffered in TMEM so that the epilogue of row tile t overlaps the MMAs of tile t + 1.
#include "gemm_params.cuh"
#include "gemm_epilogue.cuh"
#include "tma_common.cuh"

namespace pb {

namespace {

using namespace tc;

constexpr int BM = 128, BK = 32;
constexpr int NRAW = 2;                             // landing slots for raw slabs
constexpr int NOP = 2;                              // operand (hi | lo) buffers: conversion of slab i + 1 overlaps the MMAs of slab i
constexpr int SLAB_BYTES = BM * BK * 4;             // 16 KB
constexpr int MAX_SLABS = 4;                        // K <= 128
constexpr int B_BYTES = 2 * MAX_SLABS * SLAB_BYTES; // hi | lo, 128 KB
constexpr int RAW_BYTES = NRAW * SLAB_BYTES;        // 32 KB
constexpr int OP_BYTES = NOP * 2 * SLAB_BYTES;      // per buffer: hi | lo
constexpr int SMEM_BYTES = B_BYTES + RAW_BYTES + OP_BYTES;   // 224 KB (+ 1 KB static) of the 227 KB a CTA may have
constexpr int NTHREADS = 704;                       // producer, MMA, 16 epilogue warps, 4 converter warps
constexpr int EPI_WARPS = 16, CONV_WARP0 = 2 + EPI_WARPS;
constexpr int BTASKS = (BM * MAX_SLABS * 8 + NTHREADS - 1) / NTHREADS;   // 16-byte chunks of B per thread

__device__ __forceinline__ void split4(const float4 v, float4& h, float4& l) {
    uint32_t u;
#define PB_SPLIT(F)                                               \
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v.F));         \
    h.F = __uint_as_float(u);                                     \
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v.F - h.F));   \
    l.F = __uint_as_float(u);
    PB_SPLIT(x) PB_SPLIT(y) PB_SPLIT(z) PB_SPLIT(w)
#undef PB_SPLIT
}

template <int EPI>
__global__ void __launch_bounds__(NTHREADS, 1)
node_gemm_kernel(const KParams p, const int transB, const __grid_constant__ CUtensorMap map_a, const int l2_ahead) {
    extern __shared__ __align__(1024) uint8_t dyn_smem[];
    __shared__ __align__(8) uint64_t raw_full[NRAW], raw_empty[NRAW], op_full[NOP], op_empty[NOP], acc_full[2], acc_empty[2];
    __shared__ uint32_t tmem_base_holder;

    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    uint8_t* smem = dyn_smem;                      // 1024-byte aligned by declaration (checked below)
    if ((s32(smem) & 1023u) != 0u) __trap();
    uint8_t* b_hi = smem;
    uint8_t* b_lo = smem + MAX_SLABS * SLAB_BYTES;
    uint8_t* raw_ring = smem + B_BYTES;
    uint8_t* op_ring = raw_ring + RAW_BYTES;      // buffer o: hi at o * 2 slabs, lo one slab later

    const int nslab = (p.K + BK - 1) / BK;
    const int ntiles = (p.M + BM - 1) / BM;

    if (tid == 0) {
        for (int s = 0; s < NRAW; ++s) {
            mbar_init(&raw_full[s], 1);
            mbar_init(&raw_empty[s], 128);
        }
        for (int s = 0; s < NOP; ++s) {
            mbar_init(&op_full[s], 128);
            mbar_init(&op_empty[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            mbar_init(&acc_full[s], 1);
            mbar_init(&acc_empty[s], EPI_WARPS * 32);
        }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&map_a) : "memory");
    }
    if (warp == 1) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&tmem_base_holder)),
                     "r"(512u)
                     : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    // ---- B (weights): split once, resident for the whole kernel, SWIZZLE_128B K-major slabs ----
    // element (n, k) -> slab k / 32, row n, 16-byte chunk c = (k % 32) / 4 stored at chunk c ^ (n & 7)
    {
        float4 bv[BTASKS];
#pragma unroll
        for (int u = 0; u < BTASKS; ++u) {           // all loads first: one round trip to L2, not BTASKS of them
            const int t = tid + u * NTHREADS;
            const int n = t & (BM - 1);
            const int k = (t >> 7) * 4;
            float e[4] = {0.f, 0.f, 0.f, 0.f};
            if (t < BM * nslab * 8 && n < p.N) {
                if (transB && p.vecB && k + 3 < p.K) {      // (N x K) weight: one 16-byte load along k
                    const float4 q4 = __ldg(reinterpret_cast<const float4*>(p.B + (int64_t)n * p.ldb + k));
                    e[0] = q4.x; e[1] = q4.y; e[2] = q4.z; e[3] = q4.w;
                } else {
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (k + j < p.K)
                            e[j] = transB ? __ldg(p.B + (int64_t)n * p.ldb + k + j) : __ldg(p.B + (int64_t)(k + j) * p.ldb + n);
                }
            }
            bv[u] = make_float4(e[0], e[1], e[2], e[3]);
        }
#pragma unroll
        for (int u = 0; u < BTASKS; ++u) {
            const int t = tid + u * NTHREADS;
            if (t >= BM * nslab * 8) continue;
            const int n = t & (BM - 1);
            const int kc = t >> 7;                    // global 16-byte chunk index along k
            float4 h, l;
            split4(bv[u], h, l);
            const int off = (kc >> 3) * SLAB_BYTES + n * 128 + (((kc & 7) ^ (n & 7)) << 4);
            *reinterpret_cast<float4*>(b_hi + off) = h;
            *reinterpret_cast<float4*>(b_lo + off) = l;
        }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> visible to tcgen05.mma
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tmem_d = tmem_base_holder;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            uint32_t it = 0;
            for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
                for (int ks = 0; ks < nslab; ++ks, ++it) {
                    const uint32_t s = it % NRAW;
                    if (it >= NRAW) mbar_wait(&raw_empty[s], (it / NRAW - 1) & 1);
                    mbar_expect_tx(&raw_full[s], SLAB_BYTES);
                    tma_load_2d(raw_ring + s * SLAB_BYTES, &map_a, ks * BK, t * BM, &raw_full[s]);
                    // Two 16 KB landing slots are too few bytes in flight to cover the HBM latency; the same slab of
                    // this CTA's NEXT row tile is pulled into L2 now, so that its TMA load later is an L2 hit.
                    if (l2_ahead && t + (int)gridDim.x < ntiles)
                        asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];" ::"l"(&map_a),
                                     "r"(ks * BK), "r"((t + (int)gridDim.x) * BM)
                                     : "memory");
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            // The product is accumulated TRANSPOSED: the weight slabs are the M-side operand (output column = TMEM lane,
            // zero rows beyond p.N) and the row slab is the N-side one (row of the tile = TMEM column).  Both are
            // 128-row K-major SWIZZLE_128B slabs, so only the descriptors trade places.  The epilogue then holds one
            // output COLUMN per thread and a warp's store covers 32 consecutive floats of one row: one 128-byte
            // wavefront per request instead of ~46 (thread = row: 32 rows x 32 bytes), which had the LSU data path
            // at 83 % (profiles/r02/node_gemm_data_pipe.md).
            const uint32_t idesc = (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(BM >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
            uint32_t it = 0, gc = 0;
            for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++gc) {
                const uint32_t set = gc & 1;
                if (gc >= 2) mbar_wait(&acc_empty[set], (gc / 2 - 1) & 1);
                asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                const uint32_t acc_hi = tmem_d + set * 256u, acc_x = acc_hi + 128u;
                for (int ks = 0; ks < nslab; ++ks, ++it) {
                    const uint32_t o = it % NOP;
                    mbar_wait(&op_full[o], (it / NOP) & 1);
                    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
                    const uint32_t sa = s32(op_ring + o * 2 * SLAB_BYTES);
                    const uint64_t d_ah = sw128_desc(sa), d_al = sw128_desc(sa + SLAB_BYTES);
                    const uint64_t d_bh = sw128_desc(s32(b_hi + ks * SLAB_BYTES)), d_bl = sw128_desc(s32(b_lo + ks * SLAB_BYTES));
#pragma unroll
                    for (int k8 = 0; k8 < BK / 8; ++k8) {
                        const uint64_t adv = (uint64_t)((k8 * 32) >> 4);
                        const uint32_t acc = (ks > 0 || k8 > 0) ? 1u : 0u;
                        mma_tf32(acc_x, d_bh + adv, d_al + adv, idesc, acc);
                        mma_tf32(acc_x, d_bl + adv, d_ah + adv, idesc, 1u);
                        mma_tf32(acc_hi, d_bh + adv, d_ah + adv, idesc, acc);
                    }
                    mma_commit(&op_empty[o]);
                }
                mma_commit(&acc_full[set]);
            }
        }
    } else if (warp >= CONV_WARP0) {
        // ===================== converters =====================
        // (A second converter group on alternate slabs was measured: +3 % before the transposed accumulator, nothing
        //  after it, and its 576 threads cap the epilogue at 96 registers.)
        const int ct = tid - CONV_WARP0 * 32;      // 0 .. 127
        uint32_t it = 0;
        for (int t = blockIdx.x; t < ntiles; t += gridDim.x) {
            for (int ks = 0; ks < nslab; ++ks, ++it) {
                const uint32_t s = it % NRAW;
                mbar_wait(&raw_full[s], (it / NRAW) & 1);
                const uint8_t* raw = raw_ring + s * SLAB_BYTES;
                float4 h[8], l[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    const float4 v = *reinterpret_cast<const float4*>(raw + (i * 128 + ct) * 16);
                    split4(v, h[i], l[i]);
                }
                // The slot was read through the generic proxy and will be rewritten by TMA (async proxy): without
                // this fence the refill can overtake the reads (seen as 1e-2 errors with two converter groups).
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_arrive(&raw_empty[s]);                      // the landing slot may be refilled
                const uint32_t o = it % NOP;
                if (it >= NOP) mbar_wait(&op_empty[o], (it / NOP - 1) & 1);   // the MMAs that read this buffer are done
                uint8_t* op_hi = op_ring + o * 2 * SLAB_BYTES;
                uint8_t* op_lo = op_hi + SLAB_BYTES;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    *reinterpret_cast<float4*>(op_hi + (i * 128 + ct) * 16) = h[i];
                    *reinterpret_cast<float4*>(op_lo + (i * 128 + ct) * 16) = l[i];
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                mbar_arrive(&op_full[o]);
            }
        }
    } else {
        // ===================== epilogue warps (2..17) =====================
        // thread = output column n (TMEM lane), registers = consecutive rows of the tile.  Sixteen warps (one per lane
        // quadrant and 32-row block): the tanh of a row block is a chain of dependent MUFU / FMA instructions that
        // ptxas does not interleave across elements, and with eight warps the tanh variant ran at the pace of that
        // chain (5.6 us per row tile against 3.8 without the activation).
        const float alpha = p.alpha, beta = p.beta;
        const int q = warp & 3;                 // TMEM lane quadrant of this warp: columns 32 q .. 32 q + 31
        const int r0 = ((warp - 2) >> 2) * 32;  // its rows of the tile
        const int n = q * 32 + lane;
        const bool n_ok = n < p.N;
        const float bias = (p.bias && n_ok) ? __ldg(p.bias + n) : 0.f;
        uint32_t gc = 0;
        for (int t = blockIdx.x; t < ntiles; t += gridDim.x, ++gc) {
            const uint32_t set = gc & 1;
            mbar_wait(&acc_full[set], (gc / 2) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
            const int rows_left = p.M - (t * BM + r0);
            if (rows_left > 0 && q * 32 < p.N) {                  // warp-uniform
#pragma unroll 1
                for (int h0 = 0; h0 < 32; h0 += 16) {
                    uint32_t v[16], x[16];
                    const uint32_t taddr = tmem_d + ((uint32_t)(q * 32) << 16) + set * 256u + (uint32_t)(r0 + h0);
                    tmem_ld_32x32b_x16(v, taddr);
                    tmem_ld_32x32b_x16(x, taddr + 128u);
                    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
                    if (n_ok) {
                        float* cp = p.C + (int64_t)(t * BM + r0 + h0) * p.ldc + n;
                        const int left = rows_left - h0;
                        // same arithmetic order as epilogue_rows: alpha * acc + (beta * c + bias)
#pragma unroll
                        for (int i0 = 0; i0 < 16; i0 += 8) {
                            float c[8], o[8];
#pragma unroll
                            for (int i = 0; i < 8; ++i) c[i] = (beta != 0.f && i0 + i < left) ? cp[(int64_t)(i0 + i) * p.ldc] : 0.f;
#pragma unroll
                            for (int i = 0; i < 8; ++i)
                                o[i] = fmaf(alpha, __uint_as_float(v[i0 + i]) + __uint_as_float(x[i0 + i]), fmaf(beta, c[i], bias));
                            if (EPI == 1) {
#pragma unroll
                                for (int i = 0; i < 8; ++i) o[i] = tanhf(o[i]);
                            }
#pragma unroll
                            for (int i = 0; i < 8; ++i)
                                if (i0 + i < left) cp[(int64_t)(i0 + i) * p.ldc] = o[i];
                        }
                    }
                }
            }
            asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
            mbar_arrive(&acc_empty[set]);
        }
    }

    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 1) {
        __syncwarp();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(512u) : "memory");
    }
}

}  // namespace

bool node_gemm_applicable(const GemmParams& g) {
    static const bool on = [] {
        const char* e = getenv("PINOT_B200_NODE_GEMM");
        return !(e && e[0] == '0');
    }();
    if (!on || !gemm_tma_enabled()) return false;
    if (g.transA || g.M < 8192 || g.N > 128 || g.K > 128 || g.N < 1 || g.K < 1) return false;
    if (g.splitk > 1 || g.kscale || g.r1u || g.mirror || g.lower || g.ktri || g.alpha_dev || g.post_wdiag) return false;
    if (g.lda % 4 != 0 || g.ldc % 4 != 0 || g.N % 4 != 0 || !aligned16(g.A) || !aligned16(g.C)) return false;
    if (g.bias && !aligned16(g.bias)) return false;
    if (g.lda < (g.K + 3) / 4 * 4) return false;
    return true;
}

int launch_node_gemm(const GemmParams& g, cudaStream_t stream) {
    KParams p{};
    p.M = g.M; p.N = g.N; p.K = g.K;
    p.A = g.A; p.lda = g.lda; p.B = g.B; p.ldb = g.ldb; p.C = g.C; p.ldc = g.ldc;
    p.alpha = g.alpha; p.beta = g.beta; p.bias = g.bias; p.act = g.act;
    p.splitk = 1; p.vecA = p.vecC = 1;
    p.vecB = aligned16(g.B) && (g.ldb % 4 == 0);
    alignas(64) CUtensorMap map_a;
    int rc = make_map_box(&map_a, g.A, g.M, g.K, g.lda, BK, BM);
    if (rc) return rc;
    const int ntiles = (int)ceil_div(g.M, BM);
    const unsigned grid = (unsigned)(ntiles < num_sms() ? ntiles : num_sms());
    const int transB = g.transB ? 1 : 0;
    const char* e = getenv("PINOT_B200_NODE_L2_AHEAD");
    const int l2_ahead = e ? (e[0] != '0') : 1;
    auto launch = [&](auto kern) -> int {
        PB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES));
        kern<<<grid, NTHREADS, SMEM_BYTES, stream>>>(p, transB, map_a, l2_ahead);
        return check_launch("node gemm (tcgen05 + TMA)");
    };
    if (g.act) return launch(node_gemm_kernel<1>);
    return launch(node_gemm_kernel<0>);
}

}  // namespace pb
