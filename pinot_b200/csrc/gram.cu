// Kernel (3): RBF / Matern-5/2 Gram matrix, forward and backward.
// Follows pinot/regressors/kernels/rbf.py:31-55 and pinot/regressors/kernels/matern.py:30-63:
//   xs = x * exp(scale);  d_ij = || xs_i - xs'_j ||_2  (difference form, NOT the squared distance);
//   RBF      K = exp(variance) * exp(-0.5 * d)
//   Matern52 K = exp(variance) * (1 + sqrt(5 d^2) + 5/3 d^2) * exp(-sqrt(5 d^2))
// plus the fused `_perturb` / noise diagonal (gaussian_process_regressor.py:29-46, :103-105).
//
// The distance is accumulated from explicit differences: the GEMM form |x|^2+|x'|^2-2x.x' loses
// 1e-2 relative accuracy in K after the square root, which breaks the 1e-4
// parity bar.  The reference materialises an (N, N', D) tensor; here each 64x64 output tile streams
// D through shared memory once and K is written exactly once.
#include "common.cuh"
#include "gram_tc.cuh"
#include "../../include/pinot_b200.h"

namespace {

constexpr int TS = 64;        // output tile edge
constexpr int LDS_ = TS + 4;  // shared row stride (floats), keeps float4 alignment

__device__ __forceinline__ float kernel_value(int kind, float amp, float d2) {
    const float d = sqrtf(d2);
    if (kind == 0) return amp * expf(-0.5f * d);
    const float dsq = d * d;
    const float r = sqrtf(5.f * dsq);
    return amp * (1.f + r + (5.f / 3.f) * dsq) * expf(-r);
}

// Loads a (TS rows x kc) slab of `p` (row-major, ld = D), scales column k by exp(scale[k*stride]),
// and stores it k-major: S[k][row].  Rows >= rows and k >= D are zero-filled.
__device__ __forceinline__ void load_scaled_tile(const float* __restrict__ p, int64_t rows, int D,
                                                 int64_t row0, int k0, int kc,
                                                 const float* __restrict__ scale, int scale_stride,
                                                 int vec, float* __restrict__ S /*[kc][LDS_]*/) {
    const int tid = threadIdx.x;
    const int kq4 = kc / 4;  // float4 groups per row
    for (int idx = tid; idx < TS * kq4; idx += blockDim.x) {
        const int r = idx / kq4;
        const int kk = (idx - r * kq4) * 4;
        const int64_t row = row0 + r;
        const int k = k0 + kk;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row < rows && k < D) {
            const float* src = p + row * (int64_t)D + k;
            if (vec && k + 3 < D) {
                v = pb::ldg4(src);
            } else {
                v.x = __ldg(src);
                if (k + 1 < D) v.y = __ldg(src + 1);
                if (k + 2 < D) v.z = __ldg(src + 2);
                if (k + 3 < D) v.w = __ldg(src + 3);
            }
            v.x *= expf(__ldg(scale + (int64_t)(k + 0) * scale_stride));
            if (k + 1 < D) v.y *= expf(__ldg(scale + (int64_t)(k + 1) * scale_stride));
            if (k + 2 < D) v.z *= expf(__ldg(scale + (int64_t)(k + 2) * scale_stride));
            if (k + 3 < D) v.w *= expf(__ldg(scale + (int64_t)(k + 3) * scale_stride));
        }
        S[(kk + 0) * LDS_ + r] = v.x;
        S[(kk + 1) * LDS_ + r] = v.y;
        S[(kk + 2) * LDS_ + r] = v.z;
        S[(kk + 3) * LDS_ + r] = v.w;
    }
}

// acc[a][b] += sum_k (Xs[k][ty*4+a] - Ys[k][tx*4+b])^2 for k in [0, kc)
__device__ __forceinline__ void tile_dist2(const float* __restrict__ Xs, const float* __restrict__ Ys,
                                           int kc, int ty, int tx, float (&acc)[4][4]) {
#pragma unroll 8
    for (int k = 0; k < kc; ++k) {
        const float4 a4 = *reinterpret_cast<const float4*>(Xs + k * LDS_ + ty * 4);
        const float4 b4 = *reinterpret_cast<const float4*>(Ys + k * LDS_ + tx * 4);
        const float a[4] = {a4.x, a4.y, a4.z, a4.w};
        const float b[4] = {b4.x, b4.y, b4.z, b4.w};
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float d = a[i] - b[j];
                acc[i][j] = fmaf(d, d, acc[i][j]);
            }
    }
}

constexpr int FWD_KC = 32;

__global__ void __launch_bounds__(256)
gram_fwd_kernel(int kind, const float* __restrict__ x, const float* __restrict__ y,
                const float* __restrict__ scale, int scale_stride, const float* __restrict__ variance,
                float* __restrict__ out, int64_t ldo, int64_t N, int64_t Np, int D, float diag_add,
                const float* __restrict__ diag_add_log_dev, int vec_in, int vec_out, int lower_only) {
    __shared__ __align__(16) float Xs[FWD_KC * LDS_];
    __shared__ __align__(16) float Ys[FWD_KC * LDS_];
    if (lower_only && blockIdx.x > blockIdx.y) return;   // tile strictly above the diagonal: the consumer never reads it
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    const int64_t i0 = (int64_t)blockIdx.y * TS, j0 = (int64_t)blockIdx.x * TS;
    float acc[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int k0 = 0; k0 < D; k0 += FWD_KC) {
        load_scaled_tile(x, N, D, i0, k0, FWD_KC, scale, scale_stride, vec_in, Xs);
        load_scaled_tile(y, Np, D, j0, k0, FWD_KC, scale, scale_stride, vec_in, Ys);
        __syncthreads();
        tile_dist2(Xs, Ys, FWD_KC, ty, tx, acc);
        __syncthreads();
    }
    const float amp = expf(__ldg(variance));
    float dadd = diag_add;
    if (diag_add_log_dev) dadd += expf(__ldg(diag_add_log_dev));
#pragma unroll
    for (int a = 0; a < 4; ++a) {
        const int64_t i = i0 + ty * 4 + a;
        if (i >= N) continue;
        float v[4];
#pragma unroll
        for (int b = 0; b < 4; ++b) {
            const int64_t j = j0 + tx * 4 + b;
            v[b] = kernel_value(kind, amp, acc[a][b]);
            if (i == j) v[b] += dadd;
        }
        const int64_t j = j0 + tx * 4;
        float* dst = out + i * ldo + j;
        if (vec_out && j + 3 < Np) {
            *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
        } else {
#pragma unroll
            for (int b = 0; b < 4; ++b)
                if (j + b < Np) dst[b] = v[b];
        }
    }
}

// Backward of the Gram matrix with respect to the "row" argument x (and the hyper-parameters):
//   c_ij   = dK_ij * dK/dd (d_ij) / d_ij          (0 where d_ij == 0: torch.norm's sub-gradient)
//   dxs_ik = sum_j c_ij (xs_ik - ys_jk)           (difference form again: no cancellation)
//   dx_ik += exp(scale_k) dxs_ik ;  dscale_k += sum_i xs_ik dxs_ik ;  dvariance += sum_ij dK_ij K_ij
// The gradient w.r.t. the "column" argument is the same kernel with the roles swapped (`trans`).
// grid = (row tiles, column splits); results are atomically accumulated into zeroed outputs.
template <int DT>
__global__ void __launch_bounds__(256)
gram_bwd_kernel(int kind, const float* __restrict__ x, const float* __restrict__ y,
                const float* __restrict__ scale, int scale_stride, const float* __restrict__ variance,
                const float* __restrict__ dK, int64_t ldk, int trans, int64_t N, int64_t Np, int D,
                float* __restrict__ dx, float* __restrict__ dscale, float* __restrict__ dvariance,
                int vec_in) {
    extern __shared__ __align__(16) float smem[];
    float* Xs = smem;                  // [DT][LDS_]
    float* Ys = Xs + DT * LDS_;        // [DT][LDS_]
    float* Cs = Ys + DT * LDS_;        // [TS][LDS_]
    float* dsc = Cs + TS * LDS_;       // [DT]
    __shared__ float red[8];

    constexpr int KQ = DT / 16;  // k's per thread in the accumulation phase
    const int tid = threadIdx.x, ty = tid >> 4, tx = tid & 15;
    const int64_t i0 = (int64_t)blockIdx.x * TS;
    const int64_t jtiles = (Np + TS - 1) / TS;
    const int64_t per = (jtiles + gridDim.y - 1) / gridDim.y;
    const int64_t jt_beg = (int64_t)blockIdx.y * per;
    const int64_t jt_end = min(jtiles, jt_beg + per);
    if (jt_beg >= jt_end) return;

    load_scaled_tile(x, N, D, i0, 0, DT, scale, scale_stride, vec_in, Xs);
    for (int k = tid; k < DT; k += blockDim.x) dsc[k] = 0.f;
    __syncthreads();

    // this thread's xs[i][k] for i = ty*4+a, k = tx + 16*q
    float xr[4][KQ], g[4][KQ];
#pragma unroll
    for (int a = 0; a < 4; ++a)
#pragma unroll
        for (int q = 0; q < KQ; ++q) {
            xr[a][q] = Xs[(tx + 16 * q) * LDS_ + ty * 4 + a];
            g[a][q] = 0.f;
        }
    const float amp = expf(__ldg(variance));
    float dvar_local = 0.f;

    for (int64_t jt = jt_beg; jt < jt_end; ++jt) {
        const int64_t j0 = jt * TS;
        load_scaled_tile(y, Np, D, j0, 0, DT, scale, scale_stride, vec_in, Ys);
        __syncthreads();
        float acc[4][4];
#pragma unroll
        for (int a = 0; a < 4; ++a)
#pragma unroll
            for (int b = 0; b < 4; ++b) acc[a][b] = 0.f;
        tile_dist2(Xs, Ys, DT, ty, tx, acc);
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            const int64_t i = i0 + ty * 4 + a;
            float c[4];
#pragma unroll
            for (int b = 0; b < 4; ++b) {
                const int64_t j = j0 + tx * 4 + b;
                float cv = 0.f;
                if (i < N && j < Np) {
                    const float gk = trans ? __ldg(dK + j * ldk + i) : __ldg(dK + i * ldk + j);
                    const float d = sqrtf(acc[a][b]);
                    if (kind == 0) {
                        const float kv = amp * expf(-0.5f * d);
                        dvar_local = fmaf(gk, kv, dvar_local);
                        cv = d > 0.f ? gk * (-0.5f * kv) / d : 0.f;
                    } else {
                        const float dsq = d * d;
                        const float r = sqrtf(5.f * dsq);
                        const float e = amp * expf(-r);
                        dvar_local = fmaf(gk, (1.f + r + (5.f / 3.f) * dsq) * e, dvar_local);
                        cv = d > 0.f ? gk * (-(5.f / 3.f) * (1.f + r) * e) : 0.f;
                    }
                }
                c[b] = cv;
            }
            *reinterpret_cast<float4*>(Cs + (ty * 4 + a) * LDS_ + tx * 4) = make_float4(c[0], c[1], c[2], c[3]);
        }
        __syncthreads();
        // g[a][q] += sum_j Cs[i][j] * (xs[i][k] - Ys[k][j])
#pragma unroll 2
        for (int j4 = 0; j4 < TS; j4 += 4) {
            float4 c4[4];
#pragma unroll
            for (int a = 0; a < 4; ++a)
                c4[a] = *reinterpret_cast<const float4*>(Cs + (ty * 4 + a) * LDS_ + j4);
#pragma unroll
            for (int q = 0; q < KQ; ++q) {
                const float4 y4 = *reinterpret_cast<const float4*>(Ys + (tx + 16 * q) * LDS_ + j4);
#pragma unroll
                for (int a = 0; a < 4; ++a) {
                    const float xv = xr[a][q];
                    g[a][q] = fmaf(c4[a].x, xv - y4.x, g[a][q]);
                    g[a][q] = fmaf(c4[a].y, xv - y4.y, g[a][q]);
                    g[a][q] = fmaf(c4[a].z, xv - y4.z, g[a][q]);
                    g[a][q] = fmaf(c4[a].w, xv - y4.w, g[a][q]);
                }
            }
        }
        __syncthreads();
    }

    // outputs
#pragma unroll
    for (int q = 0; q < KQ; ++q) {
        const int k = tx + 16 * q;
        if (k >= D) continue;
        const float s = expf(__ldg(scale + (int64_t)k * scale_stride));
        float ds = 0.f;
#pragma unroll
        for (int a = 0; a < 4; ++a) {
            const int64_t i = i0 + ty * 4 + a;
            if (i < N) {
                if (dx) atomicAdd(dx + i * D + k, s * g[a][q]);
                ds = fmaf(xr[a][q], g[a][q], ds);
            }
        }
        if (dscale) atomicAdd(dsc + k, ds);
    }
    // block reduction of dvariance
    if (dvariance) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) dvar_local += __shfl_xor_sync(0xffffffffu, dvar_local, o);
        if ((tid & 31) == 0) red[tid >> 5] = dvar_local;
    }
    __syncthreads();
    if (dvariance && tid == 0) {
        float s = 0.f;
        for (int w = 0; w < 8; ++w) s += red[w];
        atomicAdd(dvariance, s);
    }
    if (dscale) {
        if (scale_stride == 0) {
            // scalar length-scale (Matern52): every column contributes to the single parameter
            float s = 0.f;
            for (int k = tid; k < D; k += blockDim.x) s += dsc[k];
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
            if ((tid & 31) == 0 && s != 0.f) atomicAdd(dscale, s);
        } else {
            for (int k = tid; k < D; k += blockDim.x) atomicAdd(dscale + k, dsc[k]);
        }
    }
}

template <int DT>
int launch_bwd(int kind, const float* x, const float* y, const float* scale, int scale_stride,
               const float* variance, const float* dK, int64_t ldk, int trans, int64_t N, int64_t Np,
               int D, float* dx, float* dscale, float* dvariance, int vec_in, cudaStream_t stream) {
    const size_t smem = sizeof(float) * (size_t)(2 * DT * LDS_ + TS * LDS_ + DT);
    // set on every launch: the attribute is per device, and a process may drive several
    PB_CUDA(cudaFuncSetAttribute(gram_bwd_kernel<DT>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int64_t itiles = pb::ceil_div(N, TS), jtiles = pb::ceil_div(Np, TS);
    int64_t js = pb::ceil_div(592, itiles);  // aim for ~4 CTAs per SM worth of blocks
    if (js > jtiles) js = jtiles;
    if (js < 1) js = 1;
    dim3 grid((unsigned)itiles, (unsigned)js);
    gram_bwd_kernel<DT><<<grid, 256, smem, stream>>>(kind, x, y, scale, scale_stride, variance, dK, ldk,
                                                     trans, N, Np, D, dx, dscale, dvariance, vec_in);
    return pb::check_launch("gram_bwd");
}

}  // namespace

extern "C" {

int64_t pinot_b200_gram_workspace_bytes(int64_t n, int64_t n2, int64_t d, int same_operand, int backward, int need_dy) {
    if (n < 0 || n2 < 0 || d <= 0) return -1;
    if (!pb::gram_tc_applicable(n, n2, d, backward != 0)) return 0;
    return (int64_t)sizeof(float) * pb::gram_tc_workspace_floats(n, n2, d, same_operand != 0, backward != 0, need_dy != 0);
}

int pinot_b200_gram_fwd(int kind, const float* x, const float* y, const float* scale,
                        int64_t scale_len, const float* variance, float* out, int64_t ldo, int64_t n,
                        int64_t n2, int64_t d, float diag_add, const float* diag_add_log, int lower_only,
                        void* workspace, int64_t workspace_bytes, cudaStream_t stream) {
    PB_REQUIRE(kind == 0 || kind == 1, "gram_fwd: kind must be 0 (RBF) or 1 (Matern52)");
    PB_REQUIRE(n >= 0 && n2 >= 0 && d > 0 && d < (1 << 20), "gram_fwd: invalid sizes");
    PB_REQUIRE(scale_len == 1 || scale_len == d, "gram_fwd: scale must have 1 or d entries");
    if (n == 0 || n2 == 0) return pb::PB_OK;
    PB_REQUIRE(x && y && scale && variance && out && ldo >= n2, "gram_fwd: null pointer or ldo < n2");
    const int vec_in = pb::aligned16(x) && pb::aligned16(y) && (d % 4 == 0);
    const int vec_out = pb::aligned16(out) && (ldo % 4 == 0);
    // tensor-core route: needs the scratch of pinot_b200_gram_workspace_bytes and 16-byte aligned rows
    if (workspace && pb::gram_tc_applicable(n, n2, d, false) && vec_out && pb::aligned16(workspace) &&
        workspace_bytes >= pinot_b200_gram_workspace_bytes(n, n2, d, (x == y && n == n2) ? 1 : 0, 0, 0))
        return pb::gram_fwd_tc(kind, x, y, scale, scale_len == 1 ? 0 : 1, variance, out, ldo, n, n2, (int)d, diag_add,
                               diag_add_log, lower_only, static_cast<float*>(workspace), stream);
    dim3 grid((unsigned)pb::ceil_div(n2, TS), (unsigned)pb::ceil_div(n, TS));
    PB_REQUIRE(grid.y <= 65535, "gram_fwd: too many rows for one launch");
    gram_fwd_kernel<<<grid, 256, 0, stream>>>(kind, x, y, scale, scale_len == 1 ? 0 : 1, variance, out, ldo,
                                              n, n2, (int)d, diag_add, diag_add_log, vec_in, vec_out, lower_only ? 1 : 0);
    return pb::check_launch("gram_fwd");
}

int pinot_b200_gram_bwd(int kind, const float* x, const float* y, const float* scale,
                        int64_t scale_len, const float* variance, const float* dk, int64_t ldk,
                        int64_t n, int64_t n2, int64_t d, float* dx, float* dy, float* dscale,
                        float* dvariance, int symmetric, void* workspace, int64_t workspace_bytes,
                        cudaStream_t stream) {
    PB_REQUIRE(kind == 0 || kind == 1, "gram_bwd: kind must be 0 (RBF) or 1 (Matern52)");
    PB_REQUIRE(n >= 0 && n2 >= 0 && d > 0, "gram_bwd: invalid sizes");
    PB_REQUIRE(d <= 256, "gram_bwd: feature dimension %lld exceeds the supported maximum of 256", (long long)d);
    PB_REQUIRE(scale_len == 1 || scale_len == d, "gram_bwd: scale must have 1 or d entries");
    PB_REQUIRE(symmetric >= 0 && symmetric <= 2, "gram_bwd: symmetric must be 0, 1 or 2");
    PB_REQUIRE(symmetric != 1 || (n == n2 && x == y), "gram_bwd: symmetric mode needs x == y");
    const bool tc_route = workspace && n > 0 && n2 > 0 && x && y && pb::gram_tc_applicable(n, n2, d, true) && dx &&
                          pb::aligned16(workspace) && pb::aligned16(dk) && (ldk % 4 == 0) &&
                          workspace_bytes >= pinot_b200_gram_workspace_bytes(n, n2, d, (x == y && n == n2) ? 1 : 0, 1,
                                                                             (symmetric == 0 && (dy || dscale)) ? 1 : 0) &&
                          (symmetric != 0 || dy);
    if (dx && !tc_route) PB_CUDA(cudaMemsetAsync(dx, 0, sizeof(float) * n * d, stream));
    if (dy && !tc_route) PB_CUDA(cudaMemsetAsync(dy, 0, sizeof(float) * n2 * d, stream));
    if (dscale) PB_CUDA(cudaMemsetAsync(dscale, 0, sizeof(float) * scale_len, stream));
    if (dvariance) PB_CUDA(cudaMemsetAsync(dvariance, 0, sizeof(float), stream));
    if (n == 0 || n2 == 0) return pb::PB_OK;
    PB_REQUIRE(x && y && scale && variance && dk && ldk >= n2, "gram_bwd: null pointer or ldk < n2");
    const int vec_in = pb::aligned16(x) && pb::aligned16(y) && (d % 4 == 0);
    const int ss = scale_len == 1 ? 0 : 1;
    if (tc_route)
        return pb::gram_bwd_tc(kind, x, y, scale, ss, variance, dk, ldk, n, n2, (int)d, dx, dy, dscale, dvariance, symmetric,
                               static_cast<float*>(workspace), stream);
    auto run = [&](const float* a, const float* b, int trans, int64_t na, int64_t nb, float* da,
                   float* dsc, float* dvar) -> int {
        if (d <= 32) return launch_bwd<32>(kind, a, b, scale, ss, variance, dk, ldk, trans, na, nb, (int)d, da, dsc, dvar, vec_in, stream);
        if (d <= 64) return launch_bwd<64>(kind, a, b, scale, ss, variance, dk, ldk, trans, na, nb, (int)d, da, dsc, dvar, vec_in, stream);
        if (d <= 128) return launch_bwd<128>(kind, a, b, scale, ss, variance, dk, ldk, trans, na, nb, (int)d, da, dsc, dvar, vec_in, stream);
        return launch_bwd<256>(kind, a, b, scale, ss, variance, dk, ldk, trans, na, nb, (int)d, da, dsc, dvar, vec_in, stream);
    };
    // row-side pass: dx, and the row half of dscale; dvariance is complete after this pass
    int rc = run(x, y, 0, n, n2, dx, dscale, dvariance);
    if (rc) return rc;
    // column-side pass (roles swapped, dK read transposed): dy and the column half of dscale.
    // In symmetric mode the caller promises x == y and dK == dK^T, so the column pass would
    // reproduce the row pass: it is skipped and the Python layer doubles dx / dscale.  Mode 2 is the
    // same promise for a ROW BLOCK of such a problem (x = rows of y, dK = those rows of a symmetric
    // matrix): row pass only, the caller doubles and sums over row blocks.
    if (symmetric == 0 && (dy || dscale)) {
        rc = run(y, x, 1, n2, n, dy, dscale, nullptr);
        if (rc) return rc;
    }
    return pb::PB_OK;
}

}  // extern "C"
