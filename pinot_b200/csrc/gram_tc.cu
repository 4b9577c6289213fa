// Kernel (3) on the tensor cores: the Gram matrix as a TMA-fed tcgen05 GEMM with the distance-to-kernel
// epilogue fused, forward and backward (pinot/regressors/kernels/rbf.py:31-55, matern.py:30-63).
//
// d_ij^2 = |xs_i|^2 + |ys_j|^2 - 2 xs_i . ys_j is exact enough for K = exp(-d/2) only while d is not small
// against the norms (pinot's "RBF" uses the UN-squared distance: dK/K = -1/2 dd, and d = sqrt(d^2) amplifies
// the cancellation error of d^2 near 0).  Four measures keep the 1e-4 parity bar:
//   * rows are centred by the column mean of y before scaling (distances do not change, the norms shrink);
//   * the dot products run as 3xTF32 (hi.hi + hi.lo + lo.hi);
//   * the norms are the diagonal of the rows' own product, taken from the SAME tensor-core pipeline (MODE 3 of
//     the kernel).  The tensor core truncates when it accumulates: -4.7e-7 relative over the 16 MMAs of d = 128
//     (scripts/tc_accumulate_bias.py), ten times the error of the formula itself; norms with the same bias cancel
//     it in |x|^2 + |y|^2 - 2 x.y for x near y (measured: K error 6x smaller, scripts/gram_gemm_form_error.py);
//   * every pair whose d comes out below tau (|xs_i|^2 + |ys_j|^2), tau = 0.03, is recomputed in the reference's
//     difference form (listed by the epilogue, worked through warp-wide after the tile: near_flush), and the
//     diagonal of K(x, x) is exactly zero distance.
// Backward: c_ij = dK_ij k'(d_ij) / d_ij is written as a matrix by the same GEMM + epilogue, and
//   dxs_i = sum_j c_ij (xs_i - ys_j) = xs_i rowsum_i(c) - (C Ys)_i
// is one more GEMM; flagged pairs get their coefficient from the exact distance, and the true near-duplicates
// among them (where that difference of products would cancel) contribute in difference form through atomics and
// are left out of C.
// Below ~2000^2 (forward) / ~1200^2 (backward) pairs the CUDA-core kernels of gram.cu are faster (this route's
// fixed cost is ~50 us) and are used instead: gram_tc_applicable.
#include "gemm.cuh"
#include "gram_tc.cuh"
#include "gram_tc_kernel.cuh"
#include "../../include/pinot_b200.h"
#include <stdlib.h>

namespace pb {

namespace {

constexpr int MEAN_BLOCKS = 64;
constexpr float GRAM_TAU = 0.03f;

// partial column sums of the first `rows` rows of y (row-major, ld = d): block b sums rows b, b + 64, ...
__global__ void __launch_bounds__(256) colsum_partial_kernel(const float* __restrict__ y, int64_t rows, int d,
                                                            float* __restrict__ partial /*[MEAN_BLOCKS][d]*/) {
    for (int k = threadIdx.x; k < d; k += blockDim.x) {
        float s = 0.f;
        for (int64_t r = blockIdx.x; r < rows; r += gridDim.x) s += __ldg(y + r * d + k);
        partial[blockIdx.x * d + k] = s;
    }
}
// shift[k] = mean_k, es[k] = exp(scale[k * stride])
__global__ void __launch_bounds__(256) colsum_final_kernel(const float* __restrict__ partial, int nblocks, int64_t rows, int d,
                                                          const float* __restrict__ scale, int scale_stride,
                                                          float* __restrict__ shift, float* __restrict__ es) {
    for (int k = threadIdx.x; k < d; k += blockDim.x) {
        float s = 0.f;
        for (int b = 0; b < nblocks; ++b) s += partial[b * d + k];
        shift[k] = s / (float)rows;
        es[k] = expf(__ldg(scale + (int64_t)k * scale_stride));
    }
}

// One warp per row: v = (x - shift) * es; hi / lo TF32 parts (ld = ldk, zero padded), |v|^2, optionally v itself.
__global__ void __launch_bounds__(256) gram_prep_kernel(const float* __restrict__ x, int64_t rows, int d, int ldk,
                                                       const float* __restrict__ shift, const float* __restrict__ es,
                                                       float* __restrict__ hi, float* __restrict__ lo,
                                                       float* __restrict__ plain, float* __restrict__ nrm) {
    const int64_t row = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    float s = 0.f;
    for (int k = lane; k < ldk; k += 32) {
        float v = 0.f;
        if (k < d) v = (__ldg(x + row * d + k) - __ldg(shift + k)) * __ldg(es + k);
        s = fmaf(v, v, s);
        uint32_t u;
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v));
        const float h = __uint_as_float(u);
        asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(v - h));
        hi[row * ldk + k] = h;
        lo[row * ldk + k] = __uint_as_float(u);
        if (plain) plain[row * ldk + k] = v;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
    if (lane == 0) nrm[row] = s;
}

// dxs = xc * rs - G + fix ;  dx = dxs * es ;  dscale_k += sum_i (x_ik es_k) dxs_ik  (uncentred xs: d xs / d scale)
__global__ void __launch_bounds__(256) gram_bwd_finish_kernel(const float* __restrict__ x, const float* __restrict__ xc,
                                                             int ldk, const float* __restrict__ rs,
                                                             const float* __restrict__ G, const float* __restrict__ fix,
                                                             const float* __restrict__ es, int64_t rows, int d,
                                                             float* __restrict__ dx, float* __restrict__ dscale,
                                                             int scale_scalar, int rows_per_block) {
    __shared__ float red[256];
    const int lanes_per_row = d < (int)blockDim.x ? d : (int)blockDim.x;   // threads along a row
    const int row_step = (int)blockDim.x / lanes_per_row;
    const int tr = threadIdx.x / lanes_per_row, tcol = threadIdx.x % lanes_per_row;
    const int64_t r0 = (int64_t)blockIdx.x * rows_per_block;
    const int64_t r1 = r0 + rows_per_block < rows ? r0 + rows_per_block : rows;
    float ds_total = 0.f;
    if (tr < row_step) {
        for (int kc = tcol; kc < d; kc += lanes_per_row) {
            const float e = __ldg(es + kc);
            float ds = 0.f;
            for (int64_t r = r0 + tr; r < r1; r += row_step) {
                const float v = xc[r * ldk + kc] * rs[r] - G[r * d + kc] + fix[r * d + kc];
                if (dx) dx[r * d + kc] = v * e;
                ds = fmaf(__ldg(x + r * d + kc) * e, v, ds);
            }
            if (dscale) {
                if (scale_scalar) ds_total += ds;
                else if (ds != 0.f) atomicAdd(dscale + kc, ds);
            }
        }
    }
    if (dscale && scale_scalar) {
        red[threadIdx.x] = ds_total;
        __syncthreads();
        for (int o = blockDim.x / 2; o > 0; o >>= 1) {
            if ((int)threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
            __syncthreads();
        }
        if (threadIdx.x == 0 && red[0] != 0.f) atomicAdd(dscale, red[0]);
    }
}

struct Scratch {
    float* dbg;        // development counter (64 floats)
    float* partial;    // MEAN_BLOCKS x d
    float* shift;      // d
    float* es;         // d
    float* nx; float* ny;
    float* x_hi; float* x_lo; float* y_hi; float* y_lo;
    float* xc; float* yc;          // backward only: centred, scaled rows (ld = ldk)
    float* cmat;                   // backward only: coefficient matrix, max(n x ceil4(n2), n2 x ceil4(n)) floats
    float* rs; float* fix;         // backward only: row sums (max(n, n2)) and near-pair terms (max(n, n2) x d), zeroed per pass
    float* g;                      // backward only: C Ys (max(n, n2) x d)
    int64_t zero_floats;           // rs .. end of fix (adjacent: one memset)
};

int64_t a64(int64_t v) { return (v + 63) / 64 * 64; }   // keep every array 256-byte aligned

int64_t carve(Scratch& s, float* base, int64_t n, int64_t n2, int64_t d, bool same, bool backward, bool need_y_grad) {
    const int64_t ldk = (d + 3) / 4 * 4;
    int64_t off = 0;
    auto take = [&](int64_t floats) { float* p = base ? base + off : nullptr; off += a64(floats); return p; };
    s.dbg = take(64);
    s.partial = take(MEAN_BLOCKS * d);
    s.shift = take(d);
    s.es = take(d);
    s.nx = take(n);
    s.x_hi = take(n * ldk);
    s.x_lo = take(n * ldk);
    if (same) {
        s.ny = s.nx; s.y_hi = s.x_hi; s.y_lo = s.x_lo;
    } else {
        s.ny = take(n2);
        s.y_hi = take(n2 * ldk);
        s.y_lo = take(n2 * ldk);
    }
    s.xc = s.yc = s.cmat = s.rs = s.fix = s.g = nullptr;
    s.zero_floats = 0;
    if (backward) {
        const int64_t rmax = (need_y_grad && n2 > n) ? n2 : n;
        const int64_t c1 = n * ((n2 + 3) / 4 * 4), c2 = need_y_grad ? n2 * ((n + 3) / 4 * 4) : 0;
        s.xc = take(n * ldk);
        s.yc = same ? s.xc : take(n2 * ldk);
        s.cmat = take(c1 > c2 ? c1 : c2);
        s.rs = take(rmax);
        s.fix = take(rmax * d);
        s.zero_floats = a64(rmax) + a64(rmax * d);
        s.g = take(rmax * d);
    }
    return off;
}

int prep_operands(const Scratch& s, const float* x, const float* y, const float* scale, int scale_stride, int64_t n,
                  int64_t n2, int d, bool same, bool backward, cudaStream_t stream) {
    const int ldk = (d + 3) / 4 * 4;
    const int64_t mean_rows = n2 < 16384 ? n2 : 16384;      // any common shift is valid; a sample of y is enough
    colsum_partial_kernel<<<MEAN_BLOCKS, 256, 0, stream>>>(y, mean_rows, d, s.partial);
    int rc = check_launch("gram (colsum)");
    if (rc) return rc;
    colsum_final_kernel<<<1, 256, 0, stream>>>(s.partial, MEAN_BLOCKS, mean_rows, d, scale, scale_stride, s.shift, s.es);
    if ((rc = check_launch("gram (mean)"))) return rc;
    gram_prep_kernel<<<(unsigned)ceil_div(n, 8), 256, 0, stream>>>(x, n, d, ldk, s.shift, s.es, s.x_hi, s.x_lo,
                                                                  backward ? s.xc : nullptr, s.nx);
    if ((rc = check_launch("gram (prep x)"))) return rc;
    if (!same) {
        gram_prep_kernel<<<(unsigned)ceil_div(n2, 8), 256, 0, stream>>>(y, n2, d, ldk, s.shift, s.es, s.y_hi, s.y_lo,
                                                                       backward ? s.yc : nullptr, s.ny);
        if ((rc = check_launch("gram (prep y)"))) return rc;
    }
    return PB_OK;
}

float gram_tau() {
    const char* v = getenv("PINOT_B200_GRAM_TAU");     // read on every call: a development switch
    return v ? (float)atof(v) : GRAM_TAU;
}

// one GEMM + epilogue pass: rows `a` (m of them) against rows `b` (nn of them)
template <int MODE>
int launch_pass(gram::GramParams& p, const float* a_hi, const float* a_lo, const float* b_hi, const float* b_lo,
                cudaStream_t stream) {
    const int64_t ldk = (p.K + 3) / 4 * 4;
    alignas(64) CUtensorMap mah, mal, mbh, mbl;
    int rc;
    if ((rc = tc::make_map_box(&mah, a_hi, p.M, p.K, ldk))) return rc;
    if ((rc = tc::make_map_box(&mal, a_lo, p.M, p.K, ldk))) return rc;
    if ((rc = tc::make_map_box(&mbh, b_hi, p.N, p.K, ldk))) return rc;
    if ((rc = tc::make_map_box(&mbl, b_lo, p.N, p.K, ldk))) return rc;
    const int64_t gm = ceil_div(p.M, gram::BM), gn = ceil_div(p.N, gram::BN);
    const int64_t live = MODE == 3 ? gm : p.lower ? (gm < gn ? gm * (gm + 1) / 2 : gn * (gn + 1) / 2 + (gm - gn) * gn) : gm * gn;
    const unsigned grid = (unsigned)(live < tc::num_sms() ? live : tc::num_sms());
    auto go = [&](auto kern) -> int {
        PB_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, gram::SMEM_BYTES));
        kern<<<grid, gram::NTHREADS, gram::SMEM_BYTES, stream>>>(p, mah, mal, mbh, mbl);
        return check_launch("gram (tcgen05 + TMA)");
    };
    return p.kind == 0 ? go(gram::gram_tma_kernel<MODE, 0>) : go(gram::gram_tma_kernel<MODE, 1>);
}

// |rows|^2 from the tensor core itself (MODE 3 of the kernel: the diagonal of the rows' own 3xTF32 product)
int norms_pass(const float* hi, const float* lo, int64_t rows, int d, float* nrm, cudaStream_t stream) {
    gram::GramParams p{};
    p.M = p.N = (int)rows; p.K = d;
    p.mode = 3; p.kind = 0; p.same = 1;
    p.nrm_out = nrm;
    return launch_pass<3>(p, hi, lo, hi, lo, stream);
}

}  // namespace

bool gram_tc_applicable(int64_t n, int64_t n2, int64_t d, bool backward) {
    const char* v = getenv("PINOT_B200_GRAM");         // =ffma: CUDA cores only; =tc: no lower size limit (tests).  Read on every call
    const bool on = !(v && v[0] == 'f'), forced = v && v[0] == 't';
    // Below these sizes the CUDA-core kernels win: this route has a fixed cost of ~50 us (operand preparation, the
    // norm pass, a persistent kernel's set-up); measured crossovers on B200 at d = 128: forward ~2000^2 pairs,
    // backward ~1200^2 (profiles/r02/gram_tensor_core.md)
    const int64_t min_pairs = forced ? 0 : backward ? 1500000 : 4000000;
    return on && gemm_tma_enabled() && d >= 8 && d <= 256 && n >= 256 && n2 >= 256 && n * n2 >= min_pairs && n < (1ll << 31) &&
           n2 < (1ll << 31);
}

int64_t gram_tc_workspace_floats(int64_t n, int64_t n2, int64_t d, bool same, bool backward, bool need_y_grad) {
    Scratch s;
    return carve(s, nullptr, n, n2, d, same, backward, need_y_grad);
}

int gram_fwd_tc(int kind, const float* x, const float* y, const float* scale, int scale_stride, const float* variance,
                float* out, int64_t ldo, int64_t n, int64_t n2, int d, float diag_add, const float* diag_add_log,
                int lower_only, float* workspace, cudaStream_t stream) {
    const bool same = (x == y) && (n == n2);
    Scratch s;
    carve(s, workspace, n, n2, d, same, false, false);
    int rc = prep_operands(s, x, y, scale, scale_stride, n, n2, d, same, false, stream);
    if (rc) return rc;
    if ((rc = norms_pass(s.x_hi, s.x_lo, n, d, s.nx, stream))) return rc;
    if (!same) rc = norms_pass(s.y_hi, s.y_lo, n2, d, s.ny, stream);
    if (rc) return rc;
    gram::GramParams p{};
    p.M = (int)n; p.N = (int)n2; p.K = d;
    p.mode = 1; p.kind = kind; p.same = same ? 1 : 0;
    p.mirror = (same && !lower_only) ? 1 : 0;
    p.lower = (lower_only || p.mirror) ? 1 : 0;
    p.C = out; p.ldc = ldo;
    p.nx = s.nx; p.ny = s.ny; p.x = x; p.y = y; p.d = d; p.es = s.es;
    p.var = variance; p.dlog = diag_add_log; p.diag = diag_add; p.tau = gram_tau();
    if (getenv("PINOT_B200_GRAM_DEBUG")) p.dbg = s.dbg;
    return launch_pass<1>(p, s.x_hi, s.x_lo, s.y_hi, s.y_lo, stream);
}

int gram_bwd_tc(int kind, const float* x, const float* y, const float* scale, int scale_stride, const float* variance,
                const float* dk, int64_t lddk, int64_t n, int64_t n2, int d, float* dx, float* dy, float* dscale,
                float* dvariance, int symmetric, float* workspace, cudaStream_t stream) {
    const bool same = (x == y) && (n == n2);
    const bool need_y = (symmetric == 0) && (dy != nullptr || dscale != nullptr);
    Scratch s;
    carve(s, workspace, n, n2, d, same, true, need_y);
    int rc = prep_operands(s, x, y, scale, scale_stride, n, n2, d, same, true, stream);
    if (rc) return rc;
    if ((rc = norms_pass(s.x_hi, s.x_lo, n, d, s.nx, stream))) return rc;
    if (!same) rc = norms_pass(s.y_hi, s.y_lo, n2, d, s.ny, stream);
    if (rc) return rc;
    const int ldkk = (d + 3) / 4 * 4;
    // One pass = coefficient matrix (GEMM + epilogue), G = C B (second GEMM), finish.  The row-side pass gives dx;
    // the column-side pass is the same with the roles of x and y swapped and dK read transposed.
    auto pass = [&](bool col_side) -> int {
        const int64_t m = col_side ? n2 : n, nn = col_side ? n : n2;
        const int64_t ldc = (nn + 3) / 4 * 4;
        PB_CUDA(cudaMemsetAsync(s.rs, 0, sizeof(float) * (size_t)s.zero_floats, stream));
        gram::GramParams p{};
        p.M = (int)m; p.N = (int)nn; p.K = d;
        p.mode = 2; p.kind = kind; p.same = same ? 1 : 0; p.lower = 0; p.trans = col_side ? 1 : 0;
        p.C = s.cmat; p.ldc = ldc;
        p.nx = col_side ? s.ny : s.nx; p.ny = col_side ? s.nx : s.ny;
        p.x = col_side ? y : x; p.y = col_side ? x : y; p.d = d; p.es = s.es;
        p.var = variance; p.tau = gram_tau();
        p.dk = dk; p.lddk = lddk;
        p.rs = s.rs; p.fix = s.fix; p.dvar = col_side ? nullptr : dvariance;     // sum(dK * K) once
        int rc2 = col_side ? launch_pass<2>(p, s.y_hi, s.y_lo, s.x_hi, s.x_lo, stream)
                           : launch_pass<2>(p, s.x_hi, s.x_lo, s.y_hi, s.y_lo, stream);
        if (rc2) return rc2;
        // G = C B_c  (m x d, K = nn): split-K so that the few output tiles still fill the GPU
        GemmParams gp;
        gp.M = (int)m; gp.N = d; gp.K = (int)nn;
        gp.A = s.cmat; gp.lda = ldc; gp.transA = false;
        gp.B = col_side ? s.xc : s.yc; gp.ldb = ldkk; gp.transB = false;
        gp.C = s.g; gp.ldc = d;
        gp.alpha = 1.f; gp.beta = 0.f;
        const int64_t tiles = ceil_div(m, 128) * ceil_div(d, 128);
        int64_t sk = tiles >= 148 ? 1 : (296 + tiles - 1) / tiles;
        const int64_t kmax = ceil_div(nn, 256);
        if (sk > kmax) sk = kmax;
        gp.splitk = (int)(sk < 1 ? 1 : sk);
        if ((rc2 = gemm(gp, stream))) return rc2;
        gram_bwd_finish_kernel<<<(unsigned)ceil_div(m, 32), 256, 0, stream>>>(
            col_side ? y : x, col_side ? s.yc : s.xc, ldkk, s.rs, s.g, s.fix, s.es, m, d, col_side ? dy : dx, dscale,
            scale_stride == 0 ? 1 : 0, 32);
        return check_launch("gram_bwd (finish)");
    };
    if (dvariance == nullptr) { /* nothing to zero */ }
    if ((rc = pass(false))) return rc;
    if (need_y && (rc = pass(true))) return rc;
    return PB_OK;
}

}  // namespace pb
