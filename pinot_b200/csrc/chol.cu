// Kernel (4): blocked right-looking Cholesky "sweep" over an augmented matrix, plus the small
// kernels that turn its outputs into the exact-GP loss, its closed-form gradient and the
// predictive mean / variance.  Replaces torch.cholesky + torch.triangular_solve (cuSOLVER potrf,
// cuBLAS trsm) and autograd's cholesky_backward on the path
// pinot/regressors/gaussian_process_regressor.py:85-121 (factor + alpha), :123-166 (loss),
// :168-233 (condition).
//
// One sweep factors the stacked matrix E = [ A ; X ] ((n + r) x n, row-major):
//     A (n x n, lower triangle used)  ->  L            with  A = L L^T
//     X (r x n, "extension rows")     ->  X L^{-T}
// because the panel TRSM and the trailing rank-nb update treat the extension rows exactly like
// the rows of A below the panel.  With X = [ y^T ; I ] the sweep yields z^T = (L^{-1} y)^T and
// L^{-T} (all that loss + gradient need); with X = [ y^T ; K_te,tr ] it yields z and
// V^T = K_te,tr L^{-T} (all that the predictive distribution needs).  Identity extension rows are
// only touched from the column block where their 1 sits, so L^{-T} costs n^3/3 extra flops.
#include "gemm.cuh"
#include "gemm_params.cuh"
#include "chol_diag.cuh"
#include "chol_sweep.cuh"
#include <stdlib.h>
#include "../../include/pinot_b200.h"
#include <math.h>
#include <mutex>

namespace {

constexpr int NB = 64;        // panel width
constexpr int64_t TMA_SPLIT_COLS = 128;   // tile column 0 of a rank-W update stays on the look-ahead kernel
constexpr int64_t TMA_MIN_N = 1024;       // below this the pre-split pass costs more than it saves
constexpr int LDL = NB + 4;   // shared stride of the diagonal block
constexpr int PR = 256;       // panel rows per CTA (one thread per row)
constexpr int LDT = NB + 1;   // shared stride of the staged panel rows

// Factor the nb x nb diagonal block held (unfactored, lower part valid) in Ls.  On exit the lower
// triangle of Ls holds L11 and invd[j] = 1 / L11[j][j].  Returns the first failing pivot (0-based)
// or -1.  All 256 threads take part.
__device__ int factor_diag_block(float* Ls, float* dj, int nb) {
    const int tid = threadIdx.x;
    const int ti = tid >> 4, tc = tid & 15;
    int fail = -1;
    for (int j = 0; j < nb; ++j) {
        __syncthreads();
        float d = Ls[j * LDL + j];
        if (!(d > 0.f)) {  // also catches NaN
            if (fail < 0) fail = j;
            d = 1.f;
        }
        if (tid == 0) dj[j] = d;
        const float inv = 1.f / d;
        for (int i = j + 1 + ti; i < nb; i += 16) {
            const float lij = Ls[i * LDL + j] * inv;
            for (int c = j + 1 + tc; c <= i; c += 16) Ls[i * LDL + c] = fmaf(-lij, Ls[c * LDL + j], Ls[i * LDL + c]);
        }
    }
    __syncthreads();
    // L[j][j] = sqrt(d_j); dj[j] <- 1 / L[j][j]; then scale the columns: L[i][j] = S[i][j] / L[j][j]
    if (tid < nb) {
        const float l = sqrtf(dj[tid]);
        Ls[tid * LDL + tid] = l;
        dj[tid] = 1.f / l;
    }
    __syncthreads();
    for (int idx = tid; idx < nb * nb; idx += blockDim.x) {
        const int i = idx / nb, j = idx - i * nb;
        if (j < i) Ls[i * LDL + j] *= dj[j];
    }
    __syncthreads();
    return fail;
}

// Panel step of the sweep for column block [c0, c0+nb):
//   * every CTA loads the diagonal block (from `D`, leading dimension ldd) and, unless it is
//     already factored, factors it in shared memory; CTA 0 stores L11 into `wdiag` (the deferred
//     diagonal write-back keeps late CTAs from reading a half-written block);
//   * each CTA then solves X <- X L11^{-T} for its PR rows of the panel (rows [row_beg,row_end)
//     of `P`, leading dimension ldp, columns [c0, c0+nb)).
__global__ void __launch_bounds__(PR)
chol_panel_kernel(const float* __restrict__ D, int64_t ldd, float* __restrict__ P, int64_t ldp,
                  int c0, int nb, int64_t row_beg, int64_t row_end, int factored,
                  float* __restrict__ wdiag, int* __restrict__ info) {
    extern __shared__ __align__(16) float smem[];
    float* Ls = smem;                 // [NB][LDL]
    float* dj = Ls + NB * LDL;        // [NB]
    float* Ts = dj + NB;              // [PR][LDT]
    const int tid = threadIdx.x;

    for (int idx = tid; idx < NB * NB; idx += blockDim.x) {
        const int i = idx >> 6, j = idx & 63;
        float v = (i == j) ? 1.f : 0.f;
        if (i < nb && j <= i) v = __ldg(D + (int64_t)(c0 + i) * ldd + c0 + j);
        Ls[i * LDL + j] = v;
    }
    for (int i = tid; i < NB; i += blockDim.x) dj[i] = 1.f;
    __syncthreads();
    if (!factored) {
        const int fail = factor_diag_block(Ls, dj, nb);
        if (blockIdx.x == 0) {
            if (fail >= 0 && tid == 0) atomicCAS(info, 0, c0 + fail + 1);
            for (int idx = tid; idx < NB * NB; idx += blockDim.x) {
                const int i = idx >> 6, j = idx & 63;
                wdiag[idx] = (j <= i) ? Ls[i * LDL + j] : 0.f;
            }
        }
    } else {
        if (tid < nb) dj[tid] = 1.f / Ls[tid * LDL + tid];
        __syncthreads();
    }

    const int64_t r0 = row_beg + (int64_t)blockIdx.x * PR;
    if (r0 >= row_end) return;
    const int rows_here = (int)min((int64_t)PR, row_end - r0);
    const bool vec = ((ldp & 3) == 0) && ((reinterpret_cast<uintptr_t>(P) & 15u) == 0) && ((c0 & 3) == 0);

    // stage rows (coalesced) -> Ts
    for (int idx = tid; idx < PR * (NB / 4); idx += blockDim.x) {
        const int r = idx >> 4, q = (idx & 15) * 4;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (r < rows_here) {
            const float* src = P + (r0 + r) * ldp + c0 + q;
            if (vec && q + 3 < nb) {
                v = *reinterpret_cast<const float4*>(src);
            } else {
                if (q + 0 < nb) v.x = src[0];
                if (q + 1 < nb) v.y = src[1];
                if (q + 2 < nb) v.z = src[2];
                if (q + 3 < nb) v.w = src[3];
            }
        }
        Ts[r * LDT + q + 0] = v.x;
        Ts[r * LDT + q + 1] = v.y;
        Ts[r * LDT + q + 2] = v.z;
        Ts[r * LDT + q + 3] = v.w;
    }
    __syncthreads();

    // forward substitution, one row per thread, the row lives in registers
    float xr[NB];
#pragma unroll
    for (int j = 0; j < NB; ++j) xr[j] = Ts[tid * LDT + j];
#pragma unroll
    for (int j = 0; j < NB; ++j) {
        float s0 = xr[j], s1 = 0.f, s2 = 0.f, s3 = 0.f;
#pragma unroll
        for (int t = 0; t < j; ++t) {
            const float l = Ls[j * LDL + t];
            if ((t & 3) == 0) s0 = fmaf(-xr[t], l, s0);
            else if ((t & 3) == 1) s1 = fmaf(-xr[t], l, s1);
            else if ((t & 3) == 2) s2 = fmaf(-xr[t], l, s2);
            else s3 = fmaf(-xr[t], l, s3);
        }
        xr[j] = ((s0 + s1) + (s2 + s3)) * dj[j];
    }
#pragma unroll
    for (int j = 0; j < NB; ++j) Ts[tid * LDT + j] = xr[j];
    __syncthreads();

    for (int idx = tid; idx < PR * (NB / 4); idx += blockDim.x) {
        const int r = idx >> 4, q = (idx & 15) * 4;
        if (r >= rows_here) continue;
        float* dst = P + (r0 + r) * ldp + c0 + q;
        const float4 v = make_float4(Ts[r * LDT + q], Ts[r * LDT + q + 1], Ts[r * LDT + q + 2], Ts[r * LDT + q + 3]);
        if (vec && q + 3 < nb) {
            *reinterpret_cast<float4*>(dst) = v;
        } else {
            if (q + 0 < nb) dst[0] = v.x;
            if (q + 1 < nb) dst[1] = v.y;
            if (q + 2 < nb) dst[2] = v.z;
            if (q + 3 < nb) dst[3] = v.w;
        }
    }
}

// Deferred write-back of the factored diagonal blocks, optional zeroing of the strict upper
// triangle of the n x n block (torch.cholesky returns zeros there), and extraction of diag(L).
__global__ void __launch_bounds__(256)
chol_writeback_kernel(float* __restrict__ E, int64_t ld, int n, const float* __restrict__ wdiag,
                      float* __restrict__ diag_out) {
    const int b = blockIdx.x, c0 = b * NB;
    const int nb = min(NB, n - c0);
    const float* W = wdiag + (size_t)b * NB * NB;
    for (int idx = threadIdx.x; idx < NB * NB; idx += blockDim.x) {
        const int i = idx >> 6, j = idx & 63;
        if (i < nb && j <= i) {
            const float v = W[idx];
            E[(int64_t)(c0 + i) * ld + c0 + j] = v;
            if (i == j && diag_out) diag_out[c0 + i] = v;
        }
    }
}

__global__ void __launch_bounds__(256)
zero_upper_kernel(float* __restrict__ A, int64_t ld, int64_t n) {
    for (int64_t i = blockIdx.y; i < n; i += gridDim.y)
        for (int64_t j = i + 1 + (int64_t)blockIdx.x * blockDim.x + threadIdx.x; j < n;
             j += (int64_t)gridDim.x * blockDim.x)
            A[i * ld + j] = 0.f;
}

__device__ __forceinline__ float block_sum(float v, float* red) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    __syncthreads();
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    float s = 0.f;
    const int nw = (blockDim.x + 31) >> 5;
    for (int w = 0; w < nw; ++w) s += red[w];
    return s;
}

// out[0] = 0.5 z.z + trace(L) + 0.5 n log(2 pi)        (the reference's objective, :160-164)
// out[1] = 0.5 z.z + sum log L_ii + 0.5 n log(2 pi)    (the exact negative log marginal likelihood)
// out[2] = z.z
__global__ void __launch_bounds__(1024)
gp_nll_kernel(const float* __restrict__ diagL, const float* __restrict__ z, int n, float* __restrict__ out) {
    __shared__ float red[32];
    float tr = 0.f, ld = 0.f, q = 0.f;
    for (int i = threadIdx.x; i < n; i += blockDim.x) {
        const float l = diagL[i];
        tr += l;
        ld += logf(l);
        const float zi = z[i];
        q = fmaf(zi, zi, q);
    }
    tr = block_sum(tr, red);
    ld = block_sum(ld, red);
    q = block_sum(q, red);
    if (threadIdx.x == 0) {
        const float c = 0.5f * (float)n * 1.8378770664093453f;  // log(2 pi)
        out[0] = 0.5f * q + tr + c;
        out[1] = 0.5f * q + ld + c;
        out[2] = q;
    }
}

// y[i] = sum_{k >= (upper ? i : 0)} M[i][k] * v[k]; one warp per row.
__global__ void __launch_bounds__(256)
gemv_rows_kernel(const float* __restrict__ M, int64_t ld, int64_t rows, int64_t cols, int upper,
                 const float* __restrict__ v, float* __restrict__ y) {
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    const int64_t kb = upper ? (row & ~(int64_t)31) : 0;
    float s = 0.f;
    for (int64_t k = kb + lane; k < cols; k += 32) {
        if (!upper || k >= row) s = fmaf(__ldg(M + row * ld + k), __ldg(v + k), s);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) y[row] = s;
}

// mean[i] = Vt[i,:] . z ; var[i] = kdiag - |Vt[i,:]|^2 ; one warp per test point.
__global__ void __launch_bounds__(256)
gp_predict_kernel(const float* __restrict__ Vt, int64_t ld, int64_t n_te, int64_t n, const float* __restrict__ z,
                  const float* __restrict__ variance, float eps, float* __restrict__ mean, float* __restrict__ var) {
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= n_te) return;
    const int lane = threadIdx.x & 31;
    float m = 0.f, q = 0.f;
    for (int64_t k = lane; k < n; k += 32) {
        const float v = __ldg(Vt + row * ld + k);
        m = fmaf(v, __ldg(z + k), m);
        q = fmaf(v, v, q);
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        m += __shfl_xor_sync(0xffffffffu, m, o);
        q += __shfl_xor_sync(0xffffffffu, q, o);
    }
    if (lane == 0) {
        mean[row] = m;
        if (var) var[row] = expf(__ldg(variance)) + eps - q;
    }
}

// Stand-alone factorisation of one diagonal block (first block of a sweep, or every block when the
// look-ahead is switched off): L11 -> wdiag, L11^{-T} -> winv.
__global__ void __launch_bounds__(256)
chol_diag_kernel(const float* __restrict__ E, int64_t ld, int c0, int nb, float* __restrict__ wdiag,
                 float* __restrict__ winv, int* __restrict__ info) {
    __shared__ __align__(16) float S[pb::DIAG_SMEM_FLOATS];
    const int fail = pb::factor_diag_block_regs(E, ld, c0, nb, /*coherent=*/false, S, wdiag, winv);
    if (fail >= 0 && threadIdx.x == 0) atomicCAS(info, 0, c0 + fail + 1);
}

// Outer block width of the two-level sweep (multiple of NB).  PINOT_B200_OUTER_BLOCK overrides.
int64_t outer_block_width(int64_t n) {
    static const int64_t forced = [] {
        const char* v = getenv("PINOT_B200_OUTER_BLOCK");
        const int64_t w = v ? atoll(v) : 0;
        return (w >= NB && w % NB == 0) ? w : (int64_t)0;
    }();
    if (forced) return forced;
    if (n >= 6144) return 512;
    return NB;
}

bool tc_solve_enabled() {
    static const bool on = [] {
        const char* v = getenv("PINOT_B200_SOLVE");
        return !(v && v[0] == 'f');      // PINOT_B200_SOLVE=ffma selects the CUDA-core panel solve
    }();
    return on;
}

bool lookahead_enabled() {
    static const bool on = [] {
        const char* v = getenv("PINOT_B200_LOOKAHEAD");
        return !(v && v[0] == '0');
    }();
    return on;
}

size_t panel_smem_bytes() { return sizeof(float) * (size_t)(NB * LDL + NB + PR * LDT); }

int launch_panel(const float* D, int64_t ldd, float* P, int64_t ldp, int c0, int nb, int64_t row_beg,
                 int64_t row_end, int factored, float* wdiag, int* info, cudaStream_t stream) {
    const size_t smem = panel_smem_bytes();
    PB_CUDA(cudaFuncSetAttribute(chol_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int64_t nblk = pb::ceil_div(row_end > row_beg ? row_end - row_beg : 0, PR);
    if (nblk < 1) nblk = 1;
    chol_panel_kernel<<<(unsigned)nblk, PR, smem, stream>>>(D, ldd, P, ldp, c0, nb, row_beg, row_end, factored,
                                                            wdiag, info);
    return pb::check_launch("chol_panel");
}

}  // namespace

extern "C" {

int64_t pinot_b200_potrf_workspace_bytes(int64_t n) {
    if (n < 0) return -1;
    // per column block: L11 (deferred diagonal write-back), L11^{-T} (panel TRSM as a GEMM) and a
    // "factor published" flag for the fused panel solve
    const int64_t nblk = pb::ceil_div(n, NB);
    return (int64_t)sizeof(float) * nblk * NB * NB * 2 + pb::ceil_div(nblk * (int64_t)sizeof(int), 256) * 256;
}

}  // extern "C"

namespace {
// hi / lo panel copies of the rank-W trailing updates (two-level sweep only: n >= 6144)
int64_t tma_panel_scratch_bytes(int64_t n, int64_t n_ext_dense) {
    const int64_t W = outer_block_width(n);
    if (W <= NB || !pb::gemm_tma_enabled()) return 0;   // single-level sweep: no rank-W updates
    // tile counter + hi / lo copies of one (n + n_ext_dense) x W panel
    return (int64_t)sizeof(float) * (pb::TMA_COUNTER_FLOATS + 2 * (n + n_ext_dense) * W);
}
}  // namespace

extern "C" {

int64_t pinot_b200_potrf_tma_scratch_bytes(int64_t n, int64_t n_ext_dense) {
    if (n < 0 || n_ext_dense < 0) return -1;
    // small and medium n: flags of the single-launch sweep (chol_sweep.cu); large n: the panel copies
    if (pb::sweep_persistent_applicable(n, 2 * n + n_ext_dense)) return pb::sweep_persistent_flag_bytes(n, n_ext_dense);
    return tma_panel_scratch_bytes(n, n_ext_dense);
}

int pinot_b200_potrf_ext(float* e, int64_t ld, int64_t n, int64_t n_ext_dense, int ext_identity,
                         float* diag_out, void* workspace, int64_t workspace_bytes, int* info,
                         cudaStream_t stream) {
    PB_REQUIRE(n >= 0 && n < (1ll << 30) && n_ext_dense >= 0 && ld >= n, "potrf_ext: invalid sizes");
    PB_REQUIRE(n_ext_dense < (1ll << 30) && 2 * n + n_ext_dense < (1ll << 30),
               "potrf_ext: %lld rows exceed the 2^30 row limit", (long long)(2 * n + n_ext_dense));
    PB_REQUIRE(info, "potrf_ext: info pointer is required");
    PB_CUDA(cudaMemsetAsync(info, 0, sizeof(int), stream));
    if (n == 0) return pb::PB_OK;
    PB_REQUIRE(e && workspace, "potrf_ext: null pointer");
    const int64_t base_bytes = pinot_b200_potrf_workspace_bytes(n);
    PB_REQUIRE(workspace_bytes >= base_bytes, "potrf_ext: workspace of %lld bytes is smaller than the %lld required",
               (long long)workspace_bytes, (long long)base_bytes);
    const int64_t nblk = pb::ceil_div(n, NB);
    float* wdiag = static_cast<float*>(workspace);
    float* winv = wdiag + nblk * NB * NB;
    int* flags = reinterpret_cast<int*>(winv + nblk * NB * NB);
    // optional tail of the workspace: hi / lo panel copies for the TMA-fed rank-W updates
    float* tma_scratch = nullptr;
    {
        const int64_t extra = tma_panel_scratch_bytes(n, n_ext_dense);     // NOT the flag tail of the single-launch sweep
        if (extra > 0 && workspace_bytes >= base_bytes + extra && (ld % 4 == 0) && pb::aligned16(e) &&
            pb::aligned16(workspace))
            tma_scratch = reinterpret_cast<float*>(static_cast<char*>(workspace) + base_bytes);
    }
    const int64_t r_total = n_ext_dense + (ext_identity ? n : 0);
    // One persistent cooperative kernel for the whole sweep when the caller's workspace has room for its flags
    // (the tail that pinot_b200_potrf_tma_scratch_bytes reports) and the rows are 16-byte aligned.
    if (pb::sweep_persistent_enabled() && pb::sweep_persistent_applicable(n, n + r_total) && (ld % 4 == 0) && pb::aligned16(e) && pb::aligned16(workspace) &&
        workspace_bytes >= base_bytes + pb::sweep_persistent_flag_bytes(n, n_ext_dense)) {
        int* sflags = reinterpret_cast<int*>(static_cast<char*>(workspace) + base_bytes);
        return pb::launch_sweep_persistent(e, ld, n, n_ext_dense, ext_identity, diag_out, wdiag, winv, sflags, info,
                                           stream);
    }
    PB_CUDA(cudaMemsetAsync(flags, 0, sizeof(int) * nblk, stream));
    // the fused look-ahead kernel is built for 16-byte aligned rows only (ld % 4 == 0)
    const bool lookahead = lookahead_enabled() && (ld % 4 == 0) && pb::aligned16(e);
    // With the tensor-core GEMM the panel solve of block b+1 is fused into the update that also
    // factors its diagonal block (one launch per sweep step instead of two).
    const bool fused = lookahead && pb::gemm_uses_tensor_cores();
    // Two-level blocking.  Column blocks of width NB are grouped into outer blocks of width W; the
    // rank-NB updates inside an outer block only touch that block's remaining columns, and
    // everything to the right of it receives ONE rank-W update per outer block.  The trailing
    // matrix is then read and written n/W times instead of n/NB times, which is what bounds large
    // factorisations (n = 16384: 275 GB of HBM traffic with W = NB = 64); small ones are bound by
    // the length of the dependent launch chain and keep W = NB.
    const int64_t W = outer_block_width(n);
    auto active_rows_end = [&](int64_t c1) { return n + (ext_identity ? (n_ext_dense + c1) : r_total); };
    auto attach_lookahead = [&](pb::GemmParams& g, int64_t c1) {
        if (!lookahead) return;
        const int64_t b1 = c1 / NB;
        g.post_wdiag = wdiag + b1 * NB * NB;
        g.post_winv = winv + b1 * NB * NB;
        g.post_info = info;
        g.post_nb = (int)((n - c1) < NB ? (n - c1) : NB);
        g.post_c0 = (int)c1;
        g.pdl = 1;      // every launch before this one in the sweep is a kernel: the step may be scheduled while it drains
        if (fused) {
            g.post_flag = flags + b1;
            // 2: the solve runs on the tensor cores (full 64-column blocks); 1: FFMA version
            g.post_solve = (tc_solve_enabled() && g.post_nb == NB) ? 2 : 1;
            // the identity rows that enter with block b1 ride along (their update is a no-op) so
            // that tile column 0 solves them too
            g.M = (int)(active_rows_end(c1 + g.post_nb) - c1);
        }
    };
    for (int64_t J0 = 0; J0 < n; J0 += W) {
        const int64_t J1 = (J0 + W < n) ? J0 + W : n;
        for (int64_t c0 = J0; c0 < J1; c0 += NB) {
            const int nb = (int)((n - c0) < NB ? (n - c0) : NB);
            const int64_t c1 = c0 + nb, b = c0 / NB;
            const int64_t row_end = active_rows_end(c1);
            // (a) factor of the diagonal block: first block, or no look-ahead
            if (c0 == 0 || !lookahead) {
                chol_diag_kernel<<<1, 256, 0, stream>>>(e, ld, (int)c0, nb, wdiag + b * NB * NB, winv + b * NB * NB, info);
                int rc = pb::check_launch("chol_diag");
                if (rc) return rc;
            }
            // (b) panel "TRSM" as a GEMM with M = L11^{-T}, in place (each CTA owns whole rows); in the
            //     fused scheme only the very first panel needs it
            if (row_end > c1 && (c0 == 0 || !fused)) {
                pb::GemmParams g;
                g.M = (int)(row_end - c1); g.N = nb; g.K = nb;
                g.A = e + c1 * ld + c0; g.lda = ld;
                g.B = winv + b * NB * NB; g.ldb = NB;
                g.C = e + c1 * ld + c0; g.ldc = ld;
                int rc = pb::gemm(g, stream);
                if (rc) return rc;
            }
            // (c) rank-nb update of the rest of this outer block (lower triangle + extension rows)
            if (c1 < J1) {
                pb::GemmParams g;
                g.M = (int)(row_end - c1); g.N = (int)(J1 - c1); g.K = nb;
                g.A = e + c1 * ld + c0; g.lda = ld; g.transA = false;
                g.B = e + c1 * ld + c0; g.ldb = ld; g.transB = true;
                g.C = e + c1 * ld + c1; g.ldc = ld;
                g.alpha = -1.f; g.beta = 1.f; g.lower = 1;
                attach_lookahead(g, c1);
                int rc = pb::gemm(g, stream);
                if (rc) return rc;
            }
        }
        // (d) rank-(J1 - J0) update of everything to the right of the outer block
        if (J1 < n && tma_scratch && lookahead && (n - J1) > 2 * TMA_SPLIT_COLS) {
            // Same update in two launches that share the panel: tile column 0 (which carries the
            // look-ahead factor + panel solve) with the register-staged kernel, everything to its
            // right with the TMA-fed kernel working from one hi / lo copy of the panel.
            const int64_t row_end = active_rows_end(J1);
            const int64_t rows = row_end - J1, K = J1 - J0, ldk = (K + 3) / 4 * 4;
            PB_REQUIRE(rows <= n + n_ext_dense, "potrf_ext: internal error (panel of %lld rows)", (long long)rows);
            float* hi = tma_scratch + pb::TMA_COUNTER_FLOATS;
            float* lo = hi + rows * ldk;
            int rc = pb::tma_split_operand(e + J1 * ld + J0, ld, rows, K, hi, lo, nullptr, 0, 0, stream);
            if (rc) return rc;
            pb::GemmParams g;
            g.M = (int)rows; g.N = (int)TMA_SPLIT_COLS; g.K = (int)K;
            g.A = e + J1 * ld + J0; g.lda = ld; g.transA = false;
            g.B = e + J1 * ld + J0; g.ldb = ld; g.transB = true;
            g.C = e + J1 * ld + J1; g.ldc = ld;
            g.alpha = -1.f; g.beta = 1.f; g.lower = 1;
            attach_lookahead(g, J1);
            rc = pb::gemm(g, stream);
            if (rc) return rc;
            const int64_t o = TMA_SPLIT_COLS;
            pb::KParams p{};
            p.M = (int)(rows - o); p.N = (int)(n - J1 - o); p.K = (int)K;
            p.C = e + (J1 + o) * ld + (J1 + o); p.ldc = ld;
            p.alpha = -1.f; p.beta = 1.f; p.lower = 1; p.splitk = 1; p.vecC = 1;
            rc = pb::launch_gemm_tma_presplit(p, hi + o * ldk, lo + o * ldk, hi + o * ldk, lo + o * ldk,
                                              reinterpret_cast<int*>(tma_scratch), stream);
            if (rc) return rc;
        } else if (J1 < n) {
            const int64_t row_end = active_rows_end(J1);
            pb::GemmParams g;
            g.M = (int)(row_end - J1); g.N = (int)(n - J1); g.K = (int)(J1 - J0);
            g.A = e + J1 * ld + J0; g.lda = ld; g.transA = false;
            g.B = e + J1 * ld + J0; g.ldb = ld; g.transB = true;
            g.C = e + J1 * ld + J1; g.ldc = ld;
            g.alpha = -1.f; g.beta = 1.f; g.lower = 1;
            attach_lookahead(g, J1);
            int rc = pb::gemm(g, stream);
            if (rc) return rc;
        }
    }
    chol_writeback_kernel<<<(unsigned)nblk, 256, 0, stream>>>(e, ld, (int)n, wdiag, diag_out);
    return pb::check_launch("chol_writeback");
}

// ---------------------------------------------------------------------------------------------
// Sweep with built-in scratch.
//
// Some hosts of this library may not allocate device memory at the moment they need a
// factorisation: a plug-in that is handed raw tensors by a framework which owns the whole GPU
// through its own memory pool and exposes no allocator to foreign code; a serving loop that must
// not call cudaMalloc because that call synchronises the device and stalls every other stream.
// For them the deferred diagonal blocks of the sweep live in this file-scope device array instead
// of a caller-provided workspace.  It costs PINOT_B200_POTRF_NOALLOC_MAX_N * 64 * 4 bytes (2 MiB) of
// device memory on every GPU that loads the library, for the life of the process.
// ---------------------------------------------------------------------------------------------
constexpr int64_t kNoAllocMaxN = PINOT_B200_POTRF_NOALLOC_MAX_N;
static_assert(kNoAllocMaxN % NB == 0, "the limit must be a whole number of panels");
static __device__ __align__(256) float g_noalloc_diag_scratch[(kNoAllocMaxN / NB) * NB * NB];

namespace {
constexpr int kMaxDevices = 64;
std::mutex g_noalloc_mu;
float* g_noalloc_ptr[kMaxDevices] = {nullptr};

// The array's name is only a host-side handle; kernels need its device address, which differs per
// GPU.  Looked up once per device, with that device current, and cached.
int noalloc_scratch_for_current_device(float** out) {
    int dev = -1;
    PB_CUDA(cudaGetDevice(&dev));
    PB_REQUIRE(dev >= 0 && dev < kMaxDevices, "potrf_ext_noalloc: device ordinal %d not supported", dev);
    std::lock_guard<std::mutex> lock(g_noalloc_mu);
    if (g_noalloc_ptr[dev] == nullptr) {
        float* p = nullptr;
        cudaError_t err = cudaGetSymbolAddress((void**)&p, g_noalloc_diag_scratch);
        if (err != cudaSuccess || p == nullptr) {
            pb::set_error("potrf_ext_noalloc: cudaGetSymbolAddress failed on device %d: %d (%s)", dev, (int)err,
                          cudaGetErrorString(err));
            return pb::PB_ERR_CUDA;
        }
        g_noalloc_ptr[dev] = p;
    }
    *out = g_noalloc_ptr[dev];
    return pb::PB_OK;
}
}  // namespace

int pinot_b200_potrf_ext_noalloc(float* e, int64_t ld, int64_t n, int64_t n_ext_dense, int ext_identity,
                                 float* diag_out, int* info, cudaStream_t stream) {
    PB_REQUIRE(n >= 0 && n_ext_dense >= 0 && ld >= n, "potrf_ext_noalloc: invalid sizes");
    PB_REQUIRE(n_ext_dense < (1ll << 30) && 2 * n + n_ext_dense < (1ll << 30),
               "potrf_ext_noalloc: %lld rows exceed the 2^30 row limit", (long long)(2 * n + n_ext_dense));
    PB_REQUIRE(info, "potrf_ext_noalloc: info pointer is required");
    if (n > kNoAllocMaxN) {
        pb::set_error("potrf_ext_noalloc: n = %lld exceeds the built-in scratch limit of %lld rows; call "
                      "pinot_b200_potrf_ext with a caller-provided workspace instead",
                      (long long)n, (long long)kNoAllocMaxN);
        return pb::PB_ERR_LIMIT;
    }
    float* wdiag = nullptr;
    int rc = noalloc_scratch_for_current_device(&wdiag);
    if (rc) return rc;
    if (e) {
        // the built-in array belongs to the current device: refuse a matrix that lives on another GPU
        cudaPointerAttributes attr;
        int cur = -1;
        if (cudaPointerGetAttributes(&attr, e) == cudaSuccess && cudaGetDevice(&cur) == cudaSuccess &&
            attr.type == cudaMemoryTypeDevice && attr.device != cur) {
            pb::set_error("potrf_ext_noalloc: the matrix lives on device %d but device %d is current; make the "
                          "matrix's device current before the call", attr.device, cur);
            return pb::PB_ERR_ARG;
        }
        (void)cudaGetLastError();
    }
    PB_CUDA(cudaMemsetAsync(info, 0, sizeof(int), stream));
    if (n == 0) return pb::PB_OK;
    PB_REQUIRE(e, "potrf_ext_noalloc: null pointer");
    // Nothing in the scratch has to start at zero: the panel step of column block b writes all
    // 64 x 64 entries of slot b before the final write-back reads them, so whatever an earlier
    // call left behind is never observed.
    const int64_t r_total = n_ext_dense + (ext_identity ? n : 0);
    for (int64_t c0 = 0; c0 < n; c0 += NB) {
        const int nb = (int)((n - c0) < NB ? (n - c0) : NB);
        const int64_t c1 = c0 + nb;
        const int64_t r_act = ext_identity ? (n_ext_dense + c1) : r_total;
        const int64_t row_end = n + r_act;
        rc = launch_panel(e, ld, e, ld, (int)c0, nb, c1, row_end, 0, wdiag + (c0 / NB) * NB * NB, info, stream);
        if (rc) return rc;
        if (c1 < n) {
            pb::GemmParams g;
            g.M = (int)(row_end - c1); g.N = (int)(n - c1); g.K = nb;
            g.A = e + c1 * ld + c0; g.lda = ld; g.transA = false;
            g.B = e + c1 * ld + c0; g.ldb = ld; g.transB = true;
            g.C = e + c1 * ld + c1; g.ldc = ld;
            g.alpha = -1.f; g.beta = 1.f; g.lower = 1;
            rc = pb::gemm(g, stream);
            if (rc) return rc;
        }
    }
    chol_writeback_kernel<<<(unsigned)pb::ceil_div(n, NB), 256, 0, stream>>>(e, ld, (int)n, wdiag, diag_out);
    return pb::check_launch("chol_writeback (noalloc)");
}

}  // extern "C"

namespace {

// X[:, c_beg:c_end) <- solved against L (columns before c_beg are solved and already applied to this range): panels of
// NB columns inside outer blocks of `ob` columns; updates stay inside [c_beg, c_end).
int trsm_range(const float* l, int64_t ldl, float* x, int64_t ldx, int64_t r, int64_t c_beg, int64_t c_end, int64_t ob,
               cudaStream_t stream) {
    auto update = [&](int64_t k0, int64_t k1, int64_t n0, int64_t n1) -> int {
        // X[:, n0:n1] -= X[:, k0:k1] L[n0:n1, k0:k1]^T
        if (n1 <= n0) return pb::PB_OK;
        pb::GemmParams g;
        g.M = (int)r; g.N = (int)(n1 - n0); g.K = (int)(k1 - k0);
        g.A = x + k0; g.lda = ldx; g.transA = false;
        g.B = l + n0 * ldl + k0; g.ldb = ldl; g.transB = true;
        g.C = x + n0; g.ldc = ldx; g.alpha = -1.f; g.beta = 1.f;
        return pb::gemm(g, stream);
    };
    for (int64_t o0 = c_beg; o0 < c_end; o0 += ob) {
        const int64_t o1 = (o0 + ob < c_end) ? o0 + ob : c_end;
        for (int64_t c0 = o0; c0 < o1; c0 += NB) {
            const int nb = (int)((o1 - c0) < NB ? (o1 - c0) : NB);
            int rc = launch_panel(l, ldl, x, ldx, (int)c0, nb, 0, r, 1, nullptr, nullptr, stream);
            if (rc) return rc;
            if ((rc = update(c0, c0 + nb, c0 + nb, o1))) return rc;
        }
        int rc = update(o0, o1, o1, c_end);
        if (rc) return rc;
    }
    return pb::PB_OK;
}

// Width of the rank-K updates of the triangular solve.  Bounded by ACCURACY, not by traffic: the tensor core
// truncates when it accumulates, so the error of a K-long product grows with K / 8 MMAs between two flushes of the
// accumulators (measured on the C1 posterior, worst element of the marginal variance against fp64, fp32 reference
// 3.6e-5: width 128 -> 5.2e-5, 256 -> 1.1e-4, 512 -> 1.8e-4; profiles/r02/trsm_outer_block.md).  128 keeps the
// posterior inside the parity bar; PINOT_B200_TRSM_OB overrides (development switch, read on every call).
int64_t trsm_outer_block() {
    const char* obv = getenv("PINOT_B200_TRSM_OB");
    const int64_t ob_req = obv ? atoll(obv) : 128;
    return (ob_req >= NB && ob_req % NB == 0) ? ob_req : 128;
}

constexpr int64_t TRSM_ACC_BLOCK = 512;     // columns per accurate rank-K update (flushed into registers every 128)

}  // namespace

extern "C" {

int pinot_b200_trsm_rlt(const float* l, int64_t ldl, int64_t n, float* x, int64_t ldx, int64_t r,
                        cudaStream_t stream) {
    // X (r x n) <- X L^{-T}, i.e. X^T <- L^{-1} X^T, with L already factored (lower, n x n).
    // Two levels: panels of NB columns are solved and applied inside an outer block; the columns to the right of the
    // outer block are updated once per outer block.  With r in the tens of thousands X does not fit in L2, and a
    // rank-NB update of everything to the right per panel would stream it n / NB times.
    PB_REQUIRE(n >= 0 && r >= 0 && ldl >= n && ldx >= n && n < (1ll << 30), "trsm_rlt: invalid sizes");
    if (n == 0 || r == 0) return pb::PB_OK;
    PB_REQUIRE(l && x, "trsm_rlt: null pointer");
    return trsm_range(l, ldl, x, ldx, r, 0, n, trsm_outer_block(), stream);
}

int64_t pinot_b200_trsm_rlt_workspace_bytes(int64_t n, int64_t r) {
    if (n < 0 || r < 0) return -1;
    if (n < 2 * TRSM_ACC_BLOCK || r < 1024 || !pb::gemm_tma_enabled()) return 0;      // the plain schedule is used
    const int64_t n4 = (n + 3) / 4 * 4;
    return (int64_t)sizeof(float) * (pb::TMA_COUNTER_FLOATS + 2 * n * n4 + 2 * r * TRSM_ACC_BLOCK);
}

int pinot_b200_trsm_rlt_ws(const float* l, int64_t ldl, int64_t n, float* x, int64_t ldx, int64_t r, void* workspace,
                           int64_t workspace_bytes, cudaStream_t stream) {
    // The same solve with the wide updates on the TMA-fed GEMM: inside 512-column blocks the schedule above; the
    // columns to the right of a block get ONE rank-512 update from pre-split operands, accumulated in the accurate
    // way (accumulators flushed into registers every 128 K-columns: the error of 128-wide updates, one pass over C).
    PB_REQUIRE(n >= 0 && r >= 0 && ldl >= n && ldx >= n && n < (1ll << 30), "trsm_rlt_ws: invalid sizes");
    if (n == 0 || r == 0) return pb::PB_OK;
    PB_REQUIRE(l && x, "trsm_rlt_ws: null pointer");
    const int64_t need = pinot_b200_trsm_rlt_workspace_bytes(n, r);
    const bool fast = need > 0 && workspace && workspace_bytes >= need && ldl % 4 == 0 && ldx % 4 == 0 && pb::aligned16(l) &&
                      pb::aligned16(x) && pb::aligned16(workspace);
    if (!fast) return trsm_range(l, ldl, x, ldx, r, 0, n, trsm_outer_block(), stream);
    const int64_t n4 = (n + 3) / 4 * 4;
    float* base = static_cast<float*>(workspace);
    float* lhi = base + pb::TMA_COUNTER_FLOATS;
    float* llo = lhi + n * n4;
    float* xhi = llo + n * n4;
    float* xlo = xhi + r * TRSM_ACC_BLOCK;
    int rc = pb::tma_split_operand(l, ldl, n, n, lhi, llo, nullptr, 0, 0, stream);
    if (rc) return rc;
    for (int64_t o0 = 0; o0 < n; o0 += TRSM_ACC_BLOCK) {
        const int64_t o1 = (o0 + TRSM_ACC_BLOCK < n) ? o0 + TRSM_ACC_BLOCK : n;
        if ((rc = trsm_range(l, ldl, x, ldx, r, o0, o1, trsm_outer_block(), stream))) return rc;
        if (o1 >= n) break;
        const int64_t K = o1 - o0;                        // a multiple of 4 (o1 < n here: K == TRSM_ACC_BLOCK)
        if ((rc = pb::tma_split_operand(x + o0, ldx, r, K, xhi, xlo, nullptr, 0, 0, stream))) return rc;
        pb::KParams p{};
        p.M = (int)r; p.N = (int)(n - o1); p.K = (int)K;
        p.C = x + o1; p.ldc = ldx;
        p.alpha = -1.f; p.beta = 1.f; p.splitk = 1; p.vecC = 1;
        rc = pb::launch_gemm_tma_presplit(p, xhi, xlo, lhi + o1 * n4 + o0, llo + o1 * n4 + o0, reinterpret_cast<int*>(base), stream,
                                          K, n4, /*accurate=*/true);
        if (rc) return rc;
    }
    return pb::PB_OK;
}

int pinot_b200_zero_upper(float* a, int64_t ld, int64_t n, cudaStream_t stream) {
    PB_REQUIRE(n >= 0 && ld >= n, "zero_upper: invalid sizes");
    if (n <= 1) return pb::PB_OK;
    PB_REQUIRE(a, "zero_upper: null pointer");
    const int64_t gx = pb::ceil_div(n, 256) < 64 ? pb::ceil_div(n, 256) : 64;
    dim3 grid((unsigned)gx, (unsigned)(n < 65535 ? n : 65535));
    zero_upper_kernel<<<grid, 256, 0, stream>>>(a, ld, n);
    return pb::check_launch("zero_upper");
}

int pinot_b200_gp_nll(const float* diag_l, const float* z, int64_t n, float* out3, cudaStream_t stream) {
    PB_REQUIRE(n >= 0 && n < (1ll << 31) && out3 && (n == 0 || (diag_l && z)), "gp_nll: invalid arguments");
    gp_nll_kernel<<<1, 1024, 0, stream>>>(diag_l, z, (int)n, out3);
    return pb::check_launch("gp_nll");
}

int pinot_b200_gemv_rows(const float* m, int64_t ld, int64_t rows, int64_t cols, int upper,
                         const float* v, float* y, cudaStream_t stream) {
    PB_REQUIRE(rows >= 0 && cols >= 0 && ld >= cols, "gemv_rows: invalid sizes");
    if (rows == 0) return pb::PB_OK;
    PB_REQUIRE(m && v && y, "gemv_rows: null pointer");
    gemv_rows_kernel<<<(unsigned)pb::ceil_div(rows, 8), 256, 0, stream>>>(m, ld, rows, cols, upper, v, y);
    return pb::check_launch("gemv_rows");
}

int64_t pinot_b200_gp_nll_grad_workspace_bytes(int64_t n) {
    if (n < 0) return -1;
    if (n < TMA_MIN_N || !pb::gemm_tma_enabled()) return 0;
    // tile counter + hi / lo copies of L^{-T} diag(L)^{1/2}
    return (int64_t)sizeof(float) * (pb::TMA_COUNTER_FLOATS + 2 * n * ((n + 3) / 4 * 4));
}

int pinot_b200_gp_nll_grad(const float* linv_t, int64_t ld, int64_t n, const float* diag_l,
                           const float* alpha, const float* upstream, int exact_logdet, float* dk,
                           int64_t lddk, void* workspace, int64_t workspace_bytes, cudaStream_t stream) {
    // dNLL/dK = 0.5 L^{-T} diag(L) L^{-1} - 0.5 alpha alpha^T      (trace(L) objective)
    //         = 0.5 L^{-T} L^{-1}         - 0.5 alpha alpha^T      (exact log-det objective)
    PB_REQUIRE(n >= 0 && n < (1ll << 31) && ld >= n && lddk >= n, "gp_nll_grad: invalid sizes");
    if (n == 0) return pb::PB_OK;
    PB_REQUIRE(linv_t && alpha && dk && (exact_logdet || diag_l), "gp_nll_grad: null pointer");
    {
        // Large n with a workspace: split W = L^{-T} diag(L)^{1/2} once (same operand on both sides of
        // W W^T) and run the TMA-fed tensor-core kernel on it.
        const int64_t need = pinot_b200_gp_nll_grad_workspace_bytes(n);
        const int64_t n4 = (n + 3) / 4 * 4;
        if (need > 0 && workspace && workspace_bytes >= need && ld % 4 == 0 && ld >= n4 && lddk % 4 == 0 &&
            pb::aligned16(linv_t) && pb::aligned16(dk) && pb::aligned16(workspace) &&
            (exact_logdet || pb::aligned16(diag_l))) {
            float* hi = static_cast<float*>(workspace) + pb::TMA_COUNTER_FLOATS;
            float* lo = hi + n * n4;
            int rc = pb::tma_split_operand(linv_t, ld, n, n, hi, lo, diag_l, exact_logdet ? 0 : 2, 1, stream);
            if (rc) return rc;
            pb::KParams p{};
            p.M = p.N = p.K = (int)n;
            p.C = dk; p.ldc = lddk;
            p.alpha = 0.5f; p.beta = 0.f; p.alpha_dev = upstream;
            p.r1u = alpha; p.r1v = alpha; p.r1coef = -0.5f;
            p.lower = 1; p.mirror = 1; p.ktri = 1; p.splitk = 1; p.vecC = 1;
            return pb::launch_gemm_tma_presplit(p, hi, lo, hi, lo, static_cast<int*>(workspace), stream);
        }
    }
    pb::GemmParams g;
    g.M = (int)n; g.N = (int)n; g.K = (int)n;
    g.A = linv_t; g.lda = ld; g.transA = false;
    g.B = linv_t; g.ldb = ld; g.transB = true;
    g.C = dk; g.ldc = lddk;
    g.alpha = 0.5f; g.beta = 0.f; g.alpha_dev = upstream;
    g.kscale = exact_logdet ? nullptr : diag_l;
    g.r1u = alpha; g.r1v = alpha; g.r1coef = -0.5f;
    g.lower = 1; g.mirror = 1; g.ktri = 1;
    return pb::gemm(g, stream);
}

int pinot_b200_gp_predict_mean_var(const float* vt, int64_t ld, int64_t n_te, int64_t n,
                                   const float* z, const float* variance, float eps, float* mean,
                                   float* var, cudaStream_t stream) {
    PB_REQUIRE(n_te >= 0 && n >= 0 && ld >= n, "gp_predict_mean_var: invalid sizes");
    if (n_te == 0) return pb::PB_OK;
    PB_REQUIRE(vt && z && mean && (!var || variance), "gp_predict_mean_var: null pointer");
    gp_predict_kernel<<<(unsigned)pb::ceil_div(n_te, 8), 256, 0, stream>>>(vt, ld, n_te, n, z, variance, eps, mean, var);
    return pb::check_launch("gp_predict_mean_var");
}

}  // extern "C"
