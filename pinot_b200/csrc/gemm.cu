// Kernel (2): fp32 GEMM with fused bias / tanh / rank-1 / triangular epilogues, and the node-update
// entry points built on it.  Replaces the cuBLAS sgemm that torch.nn.Linear (f_in,
// pinot/representation/sequential.py:41-43) and DGL GraphConv/SAGEConv weights
// (pinot/representation/dgl_legacy.py:40-50) dispatch to.
//
// fp32 FFMA path: embeddings must match the reference within 1e-4 relative, which rules out
// single-pass TF32/BF16 tensor-core products ("hard parts").
// Tile 128x128x16, 256 threads, 8x8 outputs per thread split in two 4-wide halves so that every
// shared-memory fragment read is a conflict-free (or broadcast) LDS.128.
#include "gemm_params.cuh"
#include "chol_diag.cuh"
#include <stdlib.h>
#include "../../include/pinot_b200.h"

namespace pb {

namespace {

constexpr int BM = 128, BN = 128, BK = 16, PAD = 4, NT = 256;

constexpr int TILE_FLOATS = 2 * BK * (BM + PAD);   // one operand, both pipeline stages
constexpr int SMEM_FLOATS = (2 * TILE_FLOATS > DIAG_SMEM_FLOATS) ? 2 * TILE_FLOATS : DIAG_SMEM_FLOATS;

// Tile of a matrix stored [row][k] (k contiguous): thread -> (row = tid/4 + 64*i, k = (tid%4)*4..+3)
__device__ __forceinline__ void load_kcontig(const float* __restrict__ P, int64_t ld, int rows, int kend,
                                             int row0, int k0, int vec, const float* __restrict__ kscale,
                                             float4 (&r)[2]) {
    const int tid = threadIdx.x;
    const int kq = (tid & 3) * 4;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int row = row0 + (tid >> 2) + 64 * i;
        const int k = k0 + kq;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row < rows) {
            const float* src = P + (int64_t)row * ld + k;
            if (vec && k + 3 < kend) {
                v = ldg4(src);
            } else {
                if (k + 0 < kend) v.x = __ldg(src + 0);
                if (k + 1 < kend) v.y = __ldg(src + 1);
                if (k + 2 < kend) v.z = __ldg(src + 2);
                if (k + 3 < kend) v.w = __ldg(src + 3);
            }
            if (kscale) {
                if (k + 0 < kend) v.x *= __ldg(kscale + k + 0);
                if (k + 1 < kend) v.y *= __ldg(kscale + k + 1);
                if (k + 2 < kend) v.z *= __ldg(kscale + k + 2);
                if (k + 3 < kend) v.w *= __ldg(kscale + k + 3);
            }
        }
        r[i] = v;
    }
}
__device__ __forceinline__ void store_kcontig(float (*S)[BM + PAD], const float4 (&r)[2]) {
    const int tid = threadIdx.x;
    const int kq = (tid & 3) * 4;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int row = (tid >> 2) + 64 * i;
        S[kq + 0][row] = r[i].x;
        S[kq + 1][row] = r[i].y;
        S[kq + 2][row] = r[i].z;
        S[kq + 3][row] = r[i].w;
    }
}

// Tile of a matrix stored [k][row] (row contiguous): thread -> (k = tid/32 + 8*i, row = (tid%32)*4..+3)
__device__ __forceinline__ void load_rowcontig(const float* __restrict__ P, int64_t ld, int rows, int kend,
                                               int row0, int k0, int vec, const float* __restrict__ kscale,
                                               float4 (&r)[2]) {
    const int tid = threadIdx.x;
    const int rq = (tid & 31) * 4;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int k = k0 + (tid >> 5) + 8 * i;
        const int row = row0 + rq;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (k < kend) {
            const float* src = P + (int64_t)k * ld + row;
            if (vec && row + 3 < rows) {
                v = ldg4(src);
            } else {
                if (row + 0 < rows) v.x = __ldg(src + 0);
                if (row + 1 < rows) v.y = __ldg(src + 1);
                if (row + 2 < rows) v.z = __ldg(src + 2);
                if (row + 3 < rows) v.w = __ldg(src + 3);
            }
            if (kscale) {
                const float s = __ldg(kscale + k);
                v.x *= s; v.y *= s; v.z *= s; v.w *= s;
            }
        }
        r[i] = v;
    }
}
__device__ __forceinline__ void store_rowcontig(float (*S)[BM + PAD], const float4 (&r)[2]) {
    const int tid = threadIdx.x;
    const int rq = (tid & 31) * 4;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
        const int k = (tid >> 5) + 8 * i;
        *reinterpret_cast<float4*>(&S[k][rq]) = r[i];
    }
}

template <bool TA, bool TB>
__global__ void __launch_bounds__(NT, 2) sgemm_kernel(const KParams p) {
    // one buffer: the A/B pipeline stages, reused by the look-ahead factorisation after the epilogue
    __shared__ __align__(16) float smem_all[SMEM_FLOATS];
    float (*As)[BK][BM + PAD] = reinterpret_cast<float (*)[BK][BM + PAD]>(smem_all);
    float (*Bs)[BK][BN + PAD] = reinterpret_cast<float (*)[BK][BN + PAD]>(smem_all + TILE_FLOATS);

    const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
    if (p.lower && n0 > m0 + BM - 1) return;

    int kb = 0, ke = p.K;
    if (p.ktri) kb = (m0 / BK) * BK;
    if (p.splitk > 1) {
        const int ktiles = (ke - kb + BK - 1) / BK;
        const int per = (ktiles + p.splitk - 1) / p.splitk;
        const int z = blockIdx.z;
        const int b = kb + z * per * BK;
        ke = min(ke, b + per * BK);
        kb = b;
        if (kb >= ke) return;
    }

    const int tid = threadIdx.x;
    const int ty = tid >> 4, tx = tid & 15;

    float acc[8][8];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[i][j] = 0.f;

    float4 ra[2], rb[2];
    auto gload = [&](int k0) {
        if (TA) load_rowcontig(p.A, p.lda, p.M, ke, m0, k0, p.vecA, p.kscale, ra);
        else    load_kcontig(p.A, p.lda, p.M, ke, m0, k0, p.vecA, p.kscale, ra);
        if (TB) load_kcontig(p.B, p.ldb, p.N, ke, n0, k0, p.vecB, nullptr, rb);
        else    load_rowcontig(p.B, p.ldb, p.N, ke, n0, k0, p.vecB, nullptr, rb);
    };
    auto sstore = [&](int buf) {
        if (TA) store_rowcontig(As[buf], ra); else store_kcontig(As[buf], ra);
        if (TB) store_kcontig(Bs[buf], rb);   else store_rowcontig(Bs[buf], rb);
    };

    int buf = 0;
    if (kb < ke) {
        gload(kb);
        sstore(0);
    }
    __syncthreads();

    for (int k0 = kb; k0 < ke; k0 += BK) {
        const bool has_next = (k0 + BK) < ke;
        if (has_next) gload(k0 + BK);
#pragma unroll
        for (int k = 0; k < BK; ++k) {
            const float4 a0 = *reinterpret_cast<const float4*>(&As[buf][k][ty * 4]);
            const float4 a1 = *reinterpret_cast<const float4*>(&As[buf][k][64 + ty * 4]);
            const float4 b0 = *reinterpret_cast<const float4*>(&Bs[buf][k][tx * 4]);
            const float4 b1 = *reinterpret_cast<const float4*>(&Bs[buf][k][64 + tx * 4]);
            const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
            const float b[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
            for (int i = 0; i < 8; ++i)
#pragma unroll
                for (int j = 0; j < 8; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
        }
        if (has_next) sstore(buf ^ 1);
        __syncthreads();
        buf ^= 1;
    }

    // ---------------- epilogue ----------------
    float alpha = p.alpha, r1c = p.r1coef;
    if (p.alpha_dev) {
        const float s = __ldg(p.alpha_dev);
        alpha *= s;
        r1c *= s;
    }
#pragma unroll
    for (int i = 0; i < 8; ++i) {
        const int m = m0 + (i < 4 ? ty * 4 + i : 64 + ty * 4 + (i - 4));
        if (m >= p.M) continue;
        const float um = p.r1u ? __ldg(p.r1u + m) * r1c : 0.f;
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const int n = n0 + h * 64 + tx * 4;
            if (n >= p.N) continue;
            float v[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = alpha * acc[i][h * 4 + j];
            float* cp = p.C + (int64_t)m * p.ldc + n;
            if (p.splitk > 1) {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (n + j < p.N && (!p.lower || n + j <= m)) atomicAdd(cp + j, v[j]);
                continue;
            }
            const bool full = (n + 3 < p.N) && (!p.lower || n + 3 <= m);
            if (p.beta != 0.f) {
                if (full && p.vecC) {
                    const float4 c = *reinterpret_cast<const float4*>(cp);
                    v[0] = fmaf(p.beta, c.x, v[0]); v[1] = fmaf(p.beta, c.y, v[1]);
                    v[2] = fmaf(p.beta, c.z, v[2]); v[3] = fmaf(p.beta, c.w, v[3]);
                } else {
#pragma unroll
                    for (int j = 0; j < 4; ++j)
                        if (n + j < p.N && (!p.lower || n + j <= m)) v[j] = fmaf(p.beta, cp[j], v[j]);
                }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (n + j < p.N) {
                    if (p.bias) v[j] += __ldg(p.bias + n + j);
                    if (p.r1u) v[j] = fmaf(um, __ldg(p.r1v + n + j), v[j]);
                    if (p.act == 1) v[j] = tanhf(v[j]);
                }
            }
            if (full && p.vecC) {
                *reinterpret_cast<float4*>(cp) = make_float4(v[0], v[1], v[2], v[3]);
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (n + j < p.N && (!p.lower || n + j <= m)) cp[j] = v[j];
            }
            if (p.mirror) {
#pragma unroll
                for (int j = 0; j < 4; ++j)
                    if (n + j < p.N && n + j < m) p.C[(int64_t)(n + j) * p.ldc + m] = v[j];
            }
        }
    }

    // ---------------- Cholesky look-ahead (tile (0,0) only) ----------------
    if (p.post_wdiag != nullptr && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0) {
        __syncthreads();   // this CTA's stores to C are visible to all of its threads
        const int fail = factor_diag_block_regs(p.C, p.ldc, 0, p.post_nb, /*coherent=*/true, smem_all,
                                                p.post_wdiag, p.post_winv);
        if (fail >= 0 && threadIdx.x == 0 && p.post_info) atomicCAS(p.post_info, 0, p.post_c0 + fail + 1);
    }
}

}  // namespace

// PINOT_B200_GEMM=ffma selects the CUDA-core kernel; the default is the tcgen05 3xTF32 kernel.
static bool use_tensor_cores() {
    static const bool on = [] {
        const char* v = getenv("PINOT_B200_GEMM");
        return !(v && (v[0] == 'f' || v[0] == 'F'));
    }();
    return on;
}

bool gemm_uses_tensor_cores() { return use_tensor_cores(); }

int gemm(const GemmParams& g, cudaStream_t stream) {
    PB_REQUIRE(g.M >= 0 && g.N >= 0 && g.K >= 0, "gemm: negative size");
    if (g.M == 0 || g.N == 0) return PB_OK;
    PB_REQUIRE(g.C && (g.K == 0 || (g.A && g.B)), "gemm: null pointer");
    PB_REQUIRE(!(g.mirror && !g.lower), "gemm: mirror requires lower");
    PB_REQUIRE(g.splitk >= 1 && g.splitk <= 65535, "gemm: bad splitk");
    // tall-and-skinny products (one row per atom, N and K <= 128): the persistent TMA-fed kernel
    if (g.K > 0 && node_gemm_applicable(g)) return launch_node_gemm(g, stream);
    KParams p{};
    p.M = g.M; p.N = g.N; p.K = g.K;
    p.A = g.A; p.lda = g.lda; p.B = g.B; p.ldb = g.ldb; p.C = g.C; p.ldc = g.ldc;
    p.alpha = g.alpha; p.beta = g.beta; p.alpha_dev = g.alpha_dev; p.bias = g.bias; p.act = g.act;
    p.kscale = g.kscale; p.r1u = g.r1u; p.r1v = g.r1v; p.r1coef = g.r1coef;
    p.lower = g.lower; p.mirror = g.mirror; p.ktri = g.ktri; p.splitk = g.splitk;
    p.vecA = aligned16(g.A) && (g.lda % 4 == 0);
    p.vecB = aligned16(g.B) && (g.ldb % 4 == 0);
    p.vecC = aligned16(g.C) && (g.ldc % 4 == 0);
    p.post_wdiag = g.post_wdiag; p.post_winv = g.post_winv; p.post_info = g.post_info;
    p.post_nb = g.post_nb; p.post_c0 = g.post_c0;
    p.post_flag = g.post_flag; p.post_solve = g.post_solve; p.pdl = g.pdl;
    PB_REQUIRE(!g.post_solve || (g.post_flag && g.post_wdiag && use_tensor_cores()),
               "gemm: the fused panel solve needs the look-ahead factor and the tensor-core kernel");
    PB_REQUIRE(!g.post_wdiag || (g.post_winv && g.splitk == 1 && g.post_nb > 0 && g.post_nb <= DIAG_NB),
               "gemm: invalid look-ahead parameters");
    if (g.splitk > 1) {
        PB_REQUIRE(g.beta == 0.f || g.beta == 1.f, "gemm: split-K needs beta in {0,1}");
        PB_REQUIRE(!g.bias && !g.act && !g.r1u && !g.mirror, "gemm: split-K supports a plain epilogue only");
        if (g.beta == 0.f) {
            if (g.ldc == g.N) {
                PB_CUDA(cudaMemsetAsync(g.C, 0, sizeof(float) * (size_t)g.M * g.N, stream));
            } else {
                PB_CUDA(cudaMemset2DAsync(g.C, sizeof(float) * g.ldc, 0, sizeof(float) * g.N, g.M, stream));
            }
        }
    }
    dim3 grid((unsigned)ceil_div(g.N, BN), (unsigned)ceil_div(g.M, BM), (unsigned)g.splitk);
    PB_REQUIRE(grid.y <= 65535, "gemm: M too large for one launch (%d rows)", g.M);
    if (use_tensor_cores()) return launch_gemm_tc(p, g.transA, g.transB, grid, stream);
    if (!g.transA && !g.transB) sgemm_kernel<false, false><<<grid, NT, 0, stream>>>(p);
    else if (!g.transA && g.transB) sgemm_kernel<false, true><<<grid, NT, 0, stream>>>(p);
    else if (g.transA && !g.transB) sgemm_kernel<true, false><<<grid, NT, 0, stream>>>(p);
    else sgemm_kernel<true, true><<<grid, NT, 0, stream>>>(p);
    return check_launch("gemm");
}

namespace {

// dz = dy * act'(y) (tanh: 1 - y^2), written to dz, plus db[n] += sum_m dz[m, n].
// Thread = 4 consecutive columns (16-byte accesses) of `rows_per_block` rows; column sums leave the
// block as one atomic per column.  N % 4 == 0 and 16-byte aligned pointers (vec) or the scalar path.
__global__ void __launch_bounds__(256)
act_bwd_bias_kernel(const float* __restrict__ dy, const float* __restrict__ y, float* __restrict__ dz,
                    float* __restrict__ db, int64_t M, int N, int act, int rows_per_block, int vec) {
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
    const int64_t r1 = min(M, r0 + rows_per_block);
    if (vec) {
        const int nv = N >> 2;                       // float4 columns
        const int cpr = min(nv, (int)blockDim.x);     // threads along a row
        const int tr = threadIdx.x / cpr, tc = threadIdx.x - tr * cpr;
        const int row_step = blockDim.x / cpr;
        const bool active = tr < row_step;   // leftover threads when cpr does not divide the block
        for (int cbase = blockIdx.y * cpr; cbase < nv; cbase += cpr * gridDim.y) {   // uniform trip count
            const int c = cbase + tc;
            const bool valid = active && c < nv;
            float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int64_t r = r0 + tr; valid && r < r1; r += row_step) {
                float4 g = __ldg(reinterpret_cast<const float4*>(dy + r * N) + c);
                if (act == 1) {
                    const float4 t = __ldg(reinterpret_cast<const float4*>(y + r * N) + c);
                    g.x *= (1.f - t.x * t.x); g.y *= (1.f - t.y * t.y);
                    g.z *= (1.f - t.z * t.z); g.w *= (1.f - t.w * t.w);
                }
                if (dz) reinterpret_cast<float4*>(dz + r * N)[c] = g;
                s.x += g.x; s.y += g.y; s.z += g.z; s.w += g.w;
            }
            if (db) {
                // combine the row lanes of this block in shared memory: one atomic per column
                __shared__ float4 part[256];
                part[threadIdx.x] = s;
                __syncthreads();
                if (tr == 0 && valid) {
                    for (int k = 1; k < row_step; ++k) {
                        const float4 o = part[k * cpr + tc];
                        s.x += o.x; s.y += o.y; s.z += o.z; s.w += o.w;
                    }
                    atomicAdd(db + 4 * c + 0, s.x); atomicAdd(db + 4 * c + 1, s.y);
                    atomicAdd(db + 4 * c + 2, s.z); atomicAdd(db + 4 * c + 3, s.w);
                }
                __syncthreads();
            }
        }
        return;
    }
    for (int n = threadIdx.x + blockIdx.y * blockDim.x; n < N; n += blockDim.x * gridDim.y) {
        float s = 0.f;
        for (int64_t r = r0; r < r1; ++r) {
            float g = __ldg(dy + r * N + n);
            if (act == 1) {
                const float t = __ldg(y + r * N + n);
                g *= (1.f - t * t);
            }
            if (dz) dz[r * N + n] = g;
            s += g;
        }
        if (db) atomicAdd(db + n, s);
    }
}

}  // namespace

}  // namespace pb

using pb::GemmParams;

extern "C" {

int pinot_b200_sgemm(int transA, int transB, int64_t M, int64_t N, int64_t K, float alpha,
                     const float* A, int64_t lda, const float* B, int64_t ldb, float beta, float* C,
                     int64_t ldc, cudaStream_t stream) {
    PB_REQUIRE(M < (1ll << 31) && N < (1ll << 31) && K < (1ll << 31), "sgemm: size exceeds int32");
    GemmParams g;
    g.M = (int)M; g.N = (int)N; g.K = (int)K; g.transA = transA != 0; g.transB = transB != 0;
    g.A = A; g.lda = lda; g.B = B; g.ldb = ldb; g.C = C; g.ldc = ldc; g.alpha = alpha; g.beta = beta;
    // Short-and-wide products (weight gradients) get split-K so that all SMs participate.
    const int64_t tiles = pb::ceil_div(M, 128) * pb::ceil_div(N, 128);
    if ((beta == 0.f || beta == 1.f) && tiles < 148 && K >= 2048) {
        int64_t s = 296 / tiles;   // two CTAs per SM
        const int64_t ktiles = pb::ceil_div(K, 16);
        if (s > ktiles / 8) s = ktiles / 8;
        if (s > 1) g.splitk = (int)s;
    }
    return pb::gemm(g, stream);
}

int pinot_b200_node_update_fwd(const float* x, const float* w, const float* bias, float* y,
                               int64_t n_rows, int64_t in_feats, int64_t out_feats, int w_is_out_in,
                               int act, int accumulate, cudaStream_t stream) {
    PB_REQUIRE(n_rows >= 0 && in_feats > 0 && out_feats > 0, "node_update_fwd: invalid sizes");
    PB_REQUIRE(n_rows < (1ll << 31) && in_feats < (1 << 24) && out_feats < (1 << 24), "node_update_fwd: size too large");
    PB_REQUIRE(act == 0 || act == 1, "node_update_fwd: act must be 0 (none) or 1 (tanh)");
    GemmParams g;
    g.M = (int)n_rows; g.N = (int)out_feats; g.K = (int)in_feats;
    g.A = x; g.lda = in_feats;
    g.B = w;
    if (w_is_out_in) { g.transB = true; g.ldb = in_feats; }   // torch.nn.Linear weight (out, in)
    else             { g.transB = false; g.ldb = out_feats; } // DGL GraphConv weight (in, out)
    g.C = y; g.ldc = out_feats; g.bias = bias; g.act = act;
    g.beta = accumulate ? 1.f : 0.f;
    return pb::gemm(g, stream);
}

int pinot_b200_node_update_bwd(const float* dy, const float* y, const float* x, const float* w,
                               float* dz_scratch, float* dx, float* dw, float* db, int64_t n_rows,
                               int64_t in_feats, int64_t out_feats, int w_is_out_in, int act,
                               cudaStream_t stream) {
    PB_REQUIRE(n_rows < (1ll << 31) && in_feats < (1 << 24) && out_feats < (1 << 24), "node_update_bwd: size too large");
    PB_REQUIRE(act == 0 || act == 1, "node_update_bwd: act must be 0 or 1");
    PB_REQUIRE(dy && w, "node_update_bwd: null pointer");
    PB_REQUIRE(act == 0 || (y && dz_scratch), "node_update_bwd: tanh needs y and dz_scratch");
    if (n_rows == 0) {
        if (dw) PB_CUDA(cudaMemsetAsync(dw, 0, sizeof(float) * in_feats * out_feats, stream));
        if (db) PB_CUDA(cudaMemsetAsync(db, 0, sizeof(float) * out_feats, stream));
        return pb::PB_OK;
    }
    const float* dz = dy;
    if (act == 1 || db) {
        if (db) PB_CUDA(cudaMemsetAsync(db, 0, sizeof(float) * out_feats, stream));
        const int vec = (out_feats % 4 == 0) && pb::aligned16(dy) && (!y || pb::aligned16(y)) &&
                        (act != 1 || pb::aligned16(dz_scratch));
        if (vec) {
            const int rows_per_block = 128;
            dim3 grid((unsigned)pb::ceil_div(n_rows, rows_per_block), (unsigned)pb::ceil_div(out_feats / 4, 256));
            pb::act_bwd_bias_kernel<<<grid, 256, 0, stream>>>(dy, y, act == 1 ? dz_scratch : nullptr, db, n_rows,
                                                             (int)out_feats, act, rows_per_block, 1);
        } else {
            const int rows_per_block = 64;
            const int tx = (int)(out_feats >= 256 ? 256 : ((out_feats + 31) / 32) * 32);
            dim3 grid((unsigned)pb::ceil_div(n_rows, rows_per_block), (unsigned)pb::ceil_div(out_feats, tx));
            pb::act_bwd_bias_kernel<<<grid, tx, 0, stream>>>(dy, y, act == 1 ? dz_scratch : nullptr, db, n_rows,
                                                            (int)out_feats, act, rows_per_block, 0);
        }
        int rc = pb::check_launch("node_update_bwd/act_bwd_bias");
        if (rc) return rc;
        if (act == 1) dz = dz_scratch;
    }
    if (dx) {
        // dx (n, in) = dz (n, out) * W^T
        GemmParams g;
        g.M = (int)n_rows; g.N = (int)in_feats; g.K = (int)out_feats;
        g.A = dz; g.lda = out_feats; g.B = w;
        if (w_is_out_in) { g.transB = false; g.ldb = in_feats; }  // W (out,in) is already (K,N)
        else             { g.transB = true; g.ldb = out_feats; }  // W (in,out) stored (N,K)
        g.C = dx; g.ldc = in_feats;
        int rc = pb::gemm(g, stream);
        if (rc) return rc;
    }
    if (dw) {
        PB_REQUIRE(x, "node_update_bwd: dw needs x");
        int rc;
        if (w_is_out_in) {
            // dW (out, in) = dz^T (out, n) * x (n, in)
            rc = pinot_b200_sgemm(1, 0, out_feats, in_feats, n_rows, 1.f, dz, out_feats, x, in_feats, 0.f,
                                  dw, in_feats, stream);
        } else {
            // dW (in, out) = x^T (in, n) * dz (n, out)
            rc = pinot_b200_sgemm(1, 0, in_feats, out_feats, n_rows, 1.f, x, in_feats, dz, out_feats, 0.f,
                                  dw, out_feats, stream);
        }
        if (rc) return rc;
    }
    return pb::PB_OK;
}

}  // extern "C"
