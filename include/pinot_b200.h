/*
 * pinot_b200 -- C-ABI of the B200 (sm_100a) kernels behind pinot's
 * graph-convolution -> Gaussian-process hot path.
 *
 * Conventions:
 *   - every pointer is a DEVICE pointer unless stated otherwise; matrices are row-major fp32 with an
 *     explicit leading dimension (in elements); CSR indices are int32;
 *   - every entry point is asynchronous on `stream`, returns 0 on success and a non-zero code
 *     (1 bad argument, 2 CUDA error, 3 documented limit exceeded) otherwise; the message is
 *     available from pinot_b200_last_error() (thread-local);
 *   - the library never allocates or frees device memory and never takes ownership: outputs and
 *     scratch are provided by the caller (see the *_workspace_bytes queries).
 *
 * The reference (choderalab/pinot) is pure Python; the "reference interface each entry point
 * replaces" is therefore the torch / DGL call made at the cited pinot file:line.
 */
#ifndef PINOT_B200_H_
#define PINOT_B200_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#ifndef __DRIVER_TYPES_H__
typedef struct CUstream_st* cudaStream_t;
#endif

#if defined(__GNUC__)
#define PINOT_B200_API __attribute__((visibility("default")))
#else
#define PINOT_B200_API
#endif

/* Last error message of the calling thread ("" if none). */
PINOT_B200_API const char* pinot_b200_last_error(void);
/* ABI version of this header (bumped on any signature change). */
PINOT_B200_API int pinot_b200_abi_version(void);
/* Number of CUDA kernels this library has launched so far in the calling process. */
PINOT_B200_API int64_t pinot_b200_launch_count(void);

/* ---------------------------------------------------------------------------------------------
 * Kernel (1): message passing.
 * out[i,:] = dst_scale[i] * sum_{e in [indptr[i],indptr[i+1])} src_scale[indices[e]] * x[indices[e],:]
 * (either scale pointer may be NULL = 1).  CSR is sorted by destination node.  Source and destination
 * nodes are ONE node set: x and out both have n_dst rows and every entry of indices lies in [0, n_dst)
 * (molecular graphs; a bipartite / block graph with more source rows than n_dst is not supported -- the
 * kernel chooses 32-bit row offsets from n_dst * feat).
 * Replaces dgl.nn.pytorch.conv.GraphConv(norm="both") / SAGEConv("mean")'s
 * update_all(copy_src, sum|mean) reached from pinot/representation/dgl_legacy.py:40-50
 * (GraphConv: src_scale = out_deg^-1/2, dst_scale = in_deg^-1/2, degrees clamped to >= 1;
 *  SAGEConv mean: src_scale = NULL, dst_scale = 1/max(in_deg,1)).
 * --------------------------------------------------------------------------------------------- */
PINOT_B200_API int pinot_b200_mp_csr_aggregate_fwd(const float* x, float* out, const int32_t* indptr,
                                    const int32_t* indices, const float* src_scale,
                                    const float* dst_scale, int64_t n_dst, int64_t feat,
                                    cudaStream_t stream);
/* Gradient of the above w.r.t. x: dx[j,:] = src_scale[j] * sum_{i: j->i} dst_scale[i] * dout[i,:].
 * indptr_t / indices_t is the CSR of the TRANSPOSED graph (sorted by source node).  Replaces the
 * autograd of DGL's SpMM (same call sites). */
PINOT_B200_API int pinot_b200_mp_csr_aggregate_bwd(const float* dout, float* dx, const int32_t* indptr_t,
                                    const int32_t* indices_t, const float* src_scale,
                                    const float* dst_scale, int64_t n_src, int64_t feat,
                                    cudaStream_t stream);

/* Graph index build from a batched edge list (device int64 src / dst in DGL's batched numbering,
 * pinot/graph.py:238-239 + dgl.batch via pinot/data/utils.py:260-276): CSR by destination
 * (indptr / indices) and by source (indptr_t / indices_t) with the original edge order kept inside
 * every row (= a stable sort by the row key), in_norm / out_norm = clamp(deg, 1)^-1/2 (GraphConv
 * norm="both"), inv_in_deg = 1 / clamp(in_deg, 1) (SAGEConv "mean"), in_deg (int64), seg_offsets =
 * exclusive scan of batch_num_nodes (n_graphs + 1) and node_seg = graph of every node (dgl.sum_nodes).
 * Replaces the torch sort / bincount / cumsum sequence a DGL graph object performs lazily; one call,
 * no host synchronisation.  Edges with an endpoint outside [0, n_nodes) are skipped.
 * Limits: n_nodes, n_graphs <= 67 108 864.  workspace >= pinot_b200_csr_build_workspace_bytes(...). */
PINOT_B200_API int64_t pinot_b200_csr_build_workspace_bytes(int64_t n_nodes, int64_t n_edges);
PINOT_B200_API int pinot_b200_csr_build(const int64_t* src, const int64_t* dst, int64_t n_edges, int64_t n_nodes,
                         const int64_t* batch_num_nodes, int64_t n_graphs, int32_t* indptr, int32_t* indices,
                         int32_t* indptr_t, int32_t* indices_t, float* in_norm, float* out_norm,
                         float* inv_in_deg, int64_t* in_deg, int32_t* seg_offsets, int32_t* node_seg,
                         void* workspace, int64_t workspace_bytes, cudaStream_t stream);
/* Atom features from their compact form: out[a, :] = one_hot(z[a], n_onehot) ++ float(fp[a, 0 .. n_fp-1]),
 * (n_atoms x (n_onehot + n_fp)) fp32 with leading dimension ld; fp holds fp16 (fp_bytes = 2) or fp32
 * (fp_bytes = 4) values.  The reference builds exactly this per atom on the host (one-hot of the atomic number,
 * pinot/graph.py:163-168 / :217-222, followed by the 17 integers and flags of fp_oe / fp_rdkit, :50-69 /
 * :108-128, concatenated at :173 / :227) and ships the dense (n, 117) fp32 matrix to the device; shipping
 * 1 + fp_bytes * n_fp bytes per atom instead of 4 * (n_onehot + n_fp) cuts that transfer 13x (fp16) or 6.8x
 * (fp32).  The host side (pinot_b200/data/packed.py) only uses a compact form that reproduces the dense matrix
 * bit for bit.  Limits: n_onehot <= 256 (z is uint8), n_fp < 65536. */
PINOT_B200_API int pinot_b200_expand_atom_features(const uint8_t* z, const void* fp, int fp_bytes, int64_t n_atoms,
                                    int64_t n_onehot, int64_t n_fp, float* out, int64_t ld, cudaStream_t stream);
/* Row gather out[i,:] = x[idx[i],:] (idx: int64, device).  The data movement of re-batching a
 * selection of molecules, i.e. of dgl.batch([dgl.unbatch(g)[i] for i in idxs]) in
 * pinot/active/experiment.py:83-104 (_slice_fn_graph). */
PINOT_B200_API int pinot_b200_gather_rows(const float* x, const int64_t* idx, float* out, int64_t n_out,
                                          int64_t feat, cudaStream_t stream);

/* Readout: out[g,:] = sum (or mean) of x[n,:] over the nodes n in [offsets[g], offsets[g+1]).
 * Replaces dgl.sum_nodes(g, "h") at pinot/representation/sequential.py:78,92-103. */
PINOT_B200_API int pinot_b200_segment_readout_fwd(const float* x, float* out, const int32_t* offsets,
                                   int64_t n_seg, int64_t feat, int mean, cudaStream_t stream);
/* dx[n,:] = dout[node_seg[n],:] (divided by the segment size for mean). */
PINOT_B200_API int pinot_b200_segment_readout_bwd(const float* dout, float* dx, const int32_t* offsets,
                                   const int32_t* node_seg, int64_t n_nodes, int64_t feat, int mean,
                                   cudaStream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Kernel (2): node update  y = act(x W + b)  (act: 0 none, 1 tanh).
 * w_is_out_in = 1: W is a torch.nn.Linear weight (out, in)  -- Sequential.f_in,
 *                  pinot/representation/sequential.py:41-43,84, and SAGEConv fc_self / fc_neigh;
 * w_is_out_in = 0: W is a DGL GraphConv weight (in, out)    -- pinot/representation/dgl_legacy.py:40.
 * accumulate = 1 adds the product to the existing contents of y before bias/activation.
 * --------------------------------------------------------------------------------------------- */
PINOT_B200_API int pinot_b200_node_update_fwd(const float* x, const float* w, const float* bias, float* y,
                               int64_t n_rows, int64_t in_feats, int64_t out_feats, int w_is_out_in,
                               int act, int accumulate, cudaStream_t stream);
/* Backward: dz = dy * act'(y); dx = dz W^T; dw = x^T dz (in W's layout); db = colsum(dz).
 * dz_scratch (n_rows x out_feats) is required when act != 0; dx / dw / db may be NULL. */
PINOT_B200_API int pinot_b200_node_update_bwd(const float* dy, const float* y, const float* x, const float* w,
                               float* dz_scratch, float* dx, float* dw, float* db, int64_t n_rows,
                               int64_t in_feats, int64_t out_feats, int w_is_out_in, int act,
                               cudaStream_t stream);

/* General row-major fp32 GEMM  C = alpha op(A) op(B) + beta C  (transX != 0: X is stored transposed).
 * Replaces torch.matmul / `@` in pinot/regressors/gaussian_process_regressor.py:212,218,392-414. */
PINOT_B200_API int pinot_b200_sgemm(int transA, int transB, int64_t M, int64_t N, int64_t K, float alpha,
                     const float* A, int64_t lda, const float* B, int64_t ldb, float beta, float* C,
                     int64_t ldc, cudaStream_t stream);

/* C = alpha A B^T + beta C for LARGE products with A (M x K) and B (N x K) both row-major: the same
 * 3xTF32 arithmetic as pinot_b200_sgemm(0, 1, ...), but the operands are split ONCE into hi / lo arrays
 * in `workspace` and streamed into the tensor cores by TMA (cp.async.bulk.tensor) from a producer warp.
 * Same reference call sites as pinot_b200_sgemm (the `@` products of
 * pinot/regressors/gaussian_process_regressor.py:212,218,392-414).  All pointers 16-byte aligned,
 * K and the leading dimensions multiples of 4; workspace_bytes >= ..._workspace_bytes(M, N, K). */
PINOT_B200_API int64_t pinot_b200_sgemm_nt_tma_workspace_bytes(int64_t M, int64_t N, int64_t K);
PINOT_B200_API int pinot_b200_sgemm_nt_tma(int64_t M, int64_t N, int64_t K, float alpha, const float* A, int64_t lda,
                            const float* B, int64_t ldb, float beta, float* C, int64_t ldc, void* workspace,
                            int64_t workspace_bytes, cudaStream_t stream);

/* The two halves of pinot_b200_sgemm_nt_tma for callers that reuse one operand in several products
 * (the panel of a blocked / sharded Cholesky: split once, then one update per column block).
 * tf32_split: hi = tf32(v), lo = tf32(v - hi), v = x * f(colscale[col]) with f = 1 (scale_mode 0),
 * identity (1) or sqrt (2); hi / lo are rows x ceil4(K) arrays (leading dimension ceil4(K)).
 * sgemm_nt_presplit: C (M x N) = alpha A B^T + beta C from such arrays; lower != 0 computes only
 * elements with column <= row; tile_counter is one int of device scratch (set by the call).
 * No direct counterpart in the reference (the `@` products of gaussian_process_regressor.py:212-218). */
PINOT_B200_API int pinot_b200_tf32_split(const float* x, int64_t ldx, int64_t rows, int64_t K, const float* colscale,
                          int scale_mode, float* hi, float* lo, cudaStream_t stream);
PINOT_B200_API int pinot_b200_sgemm_nt_presplit(int64_t M, int64_t N, int64_t K, float alpha, const float* a_hi,
                                 const float* a_lo, const float* b_hi, const float* b_lo, float beta, float* C,
                                 int64_t ldc, int lower, int* tile_counter, cudaStream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Kernel (3): Gram matrix.  kind 0 = RBF (pinot/regressors/kernels/rbf.py:31-55: exp(variance) *
 * exp(-0.5 * ||(x - x') * exp(scale)||_2)), kind 1 = Matern-5/2 (pinot/regressors/kernels/matern.py:30-63).
 * scale has scale_len = d (ARD) or 1 entries; variance is a device scalar.
 * out[i,j] = K(x_i, y_j) + (i == j) * (diag_add + (diag_add_log ? exp(*diag_add_log) : 0)):
 * the fused `_perturb` (gaussian_process_regressor.py:29-46, rectangular matrices included) and
 * noise term exp(log_sigma) * I (:103-105).
 * lower_only != 0: whole tiles (64 x 64 or 128 x 128) strictly above the main diagonal are skipped (left
 * untouched) -- for a symmetric matrix that goes straight into the Cholesky sweep, which reads the lower
 * triangle only.
 * Two routes, same values to within the 1e-4 parity bar:
 *   - workspace == NULL (or too small, or fewer than ~2000^2 pairs forward / ~1200^2 backward, where the
 *     CUDA-core kernel is the faster one, or d outside 8..256): CUDA cores, distances accumulated from explicit
 *     differences exactly like the reference; pinot_b200_gram_workspace_bytes returns 0 for these cases;
 *   - workspace of pinot_b200_gram_workspace_bytes(...) bytes, 16-byte aligned: the dot products run as a
 *     TMA-fed tcgen05 GEMM (3xTF32) over rows that are centred, scaled and split once; the epilogue turns
 *     |x|^2 + |y|^2 - 2 x.y into the kernel value and recomputes in difference form every pair whose squared
 *     distance is too small against the norms for the GEMM form to hold (csrc/gram_tc.cu).
 * --------------------------------------------------------------------------------------------- */
PINOT_B200_API int64_t pinot_b200_gram_workspace_bytes(int64_t n, int64_t n2, int64_t d, int same_operand,
                                                       int backward, int need_dy);
PINOT_B200_API int pinot_b200_gram_fwd(int kind, const float* x, const float* y, const float* scale,
                        int64_t scale_len, const float* variance, float* out, int64_t ldo, int64_t n,
                        int64_t n2, int64_t d, float diag_add, const float* diag_add_log, int lower_only,
                        void* workspace, int64_t workspace_bytes, cudaStream_t stream);
/* Backward of the Gram matrix given dk = dLoss/dK (n x n2).  Outputs (any may be NULL, all are
 * overwritten): dx (n x d), dy (n2 x d), dscale (scale_len), dvariance (1).  symmetric = 1 promises
 * x == y and dk == dk^T: only the row-side pass runs and the caller doubles dx and dscale.
 * symmetric = 2: x is a ROW BLOCK of y and dk the same rows of a symmetric matrix (row-sharded GP):
 * row-side pass only; doubled dx is the full gradient of those rows, doubled dscale sums to the full
 * gradient over all row blocks.
 * Replaces autograd through torch.norm / exp at kernels/rbf.py:31-53 (sub-gradient 0 at d = 0).
 * Limit: d <= 256.  With a workspace (pinot_b200_gram_workspace_bytes(n, n2, d, x == y, 1, need_dy), need_dy =
 * symmetric == 0 && (dy || dscale)) the tensor-core route runs: the coefficient matrix c_ij = dk_ij k'(d_ij) / d_ij
 * comes out of the same GEMM + epilogue as the forward pass and dx = rowsum(c) * xs - c ys is a second GEMM;
 * near pairs (where that difference of products cancels) are added in difference form. */
PINOT_B200_API int pinot_b200_gram_bwd(int kind, const float* x, const float* y, const float* scale,
                        int64_t scale_len, const float* variance, const float* dk, int64_t ldk,
                        int64_t n, int64_t n2, int64_t d, float* dx, float* dy, float* dscale,
                        float* dvariance, int symmetric, void* workspace, int64_t workspace_bytes,
                        cudaStream_t stream);

/* ---------------------------------------------------------------------------------------------
 * Kernel (4): blocked Cholesky sweep over the stacked matrix e = [A ; X], (n + r) x n, ld >= n:
 *   A (n x n; lower triangle read)  -> L, A = L L^T (strict upper triangle left untouched)
 *   X (r x n)                       -> X L^{-T}
 * r = n_ext_dense (+ n if ext_identity: the caller stores I in those rows and gets L^{-T} back;
 * they must FOLLOW the dense rows).  diag_out (n, optional) receives diag(L).  *info (device int)
 * is 0 on success or 1 + the first non-positive pivot (LAPACK convention).  workspace (16-byte
 * aligned, workspace_bytes long) must hold at least pinot_b200_potrf_workspace_bytes(n) bytes; when
 * it holds pinot_b200_potrf_tma_scratch_bytes(n, n_ext_dense) MORE, the faster schedule for that size
 * runs (same arithmetic, results equal up to the order of fp32 additions):
 *   n < 6144:  with PINOT_B200_SWEEP=persistent in the environment the whole sweep is ONE persistent
 *              cooperative kernel (csrc/chol_sweep.cu; the tail holds its progress flags) instead of one
 *              launch per 64-column panel -- opt-in, it is not the faster of the two yet;
 *   n >= 6144: the rank-512 trailing updates run on the TMA-fed tensor-core kernel from hi / lo panel
 *              copies kept in the tail.
 * The single-launch sweep is a cooperative launch: it occupies up to two CTAs per SM of the whole GPU and
 * its CTAs wait on one another, so it needs the device to itself for its duration (a concurrent kernel
 * on another stream delays its start rather than breaking it).
 * Replaces torch.cholesky + torch.triangular_solve at gaussian_process_regressor.py:108-119,215,373-390.
 * --------------------------------------------------------------------------------------------- */
PINOT_B200_API int64_t pinot_b200_potrf_workspace_bytes(int64_t n);
PINOT_B200_API int64_t pinot_b200_potrf_tma_scratch_bytes(int64_t n, int64_t n_ext_dense);
PINOT_B200_API int pinot_b200_potrf_ext(float* e, int64_t ld, int64_t n, int64_t n_ext_dense, int ext_identity,
                         float* diag_out, void* workspace, int64_t workspace_bytes, int* info,
                         cudaStream_t stream);
/* The factorisation of pinot_b200_potrf_ext (same arguments, same outputs up to rounding) for callers
 * that cannot hand over scratch memory.
 * (No counterpart in the reference: pinot lets torch allocate whatever cuSOLVER asks for.)
 *
 * Who needs it: code that receives device buffers from a host framework but has no way to obtain
 * more -- e.g. a plug-in running inside a runtime whose memory pool owns the whole GPU -- and loops
 * that must not call cudaMalloc because it synchronises the device.  The deferred diagonal blocks
 * are kept in an array compiled into the library (2 MiB of device memory per GPU that loads it,
 * held until the process exits), so no workspace argument exists and nothing is allocated or
 * freed, here or in any routine it calls.  Inputs and outputs (e, diag_out, info) are still the
 * caller's.
 *
 * It is NOT the schedule of pinot_b200_potrf_ext: it runs the plain right-looking sweep, two launches
 * per 64-column panel (one kernel factors the diagonal block and solves the panel on CUDA cores, one
 * tensor-core GEMM applies the trailing update), without the look-ahead factor, the tensor-core panel
 * solve or the two-level blocking that need flags and inverse blocks in a workspace.  It is therefore
 * slower than pinot_b200_potrf_ext (DESIGN.md section 4 has the measured pair).
 *
 * Limits and rules:
 *   - n <= PINOT_B200_POTRF_NOALLOC_MAX_N; larger n returns code 3 with a message naming the limit;
 *     n + n_ext_dense (+ n with ext_identity) must stay below 2^30 rows;
 *   - the GPU that holds `e` must be the CURRENT device of the calling thread (the built-in array is
 *     per device; a mismatch returns code 1 instead of touching another GPU's memory);
 *   - the first call on each device looks up the array's device address (once, cached); make that
 *     first call before starting any stream capture on that device.  The cached address does not
 *     survive cudaDeviceReset(): do not reset a device between calls;
 *   - every call on a device uses the same array: two calls that may overlap on the GPU (different
 *     streams, or different host threads) race with each other -- serialise them.  Back-to-back
 *     calls on one stream are fine; nothing needs clearing in between because each call overwrites
 *     every scratch entry before reading it. */
#define PINOT_B200_POTRF_NOALLOC_MAX_N 8192
PINOT_B200_API int pinot_b200_potrf_ext_noalloc(float* e, int64_t ld, int64_t n, int64_t n_ext_dense,
                                                int ext_identity, float* diag_out, int* info,
                                                cudaStream_t stream);
/* X (r x n) <- X L^{-T} for an already factored lower-triangular L (torch.triangular_solve(
 * input=X^T, A=L, upper=False) transposed; gaussian_process_regressor.py:112-114,215,376-390). */
PINOT_B200_API int pinot_b200_trsm_rlt(const float* l, int64_t ldl, int64_t n, float* x, int64_t ldx, int64_t r,
                        cudaStream_t stream);
/* The same solve with caller scratch (pinot_b200_trsm_rlt_workspace_bytes(n, r) bytes, 16-byte aligned; 0 means the
 * sizes are too small for it to pay and NULL / 0 may be passed): the updates of everything right of a 512-column
 * block run as ONE product from pre-split operands on the TMA-fed tensor-core GEMM whose accumulators are flushed
 * into registers every 128 K-columns -- the accuracy of 128-wide updates (the posterior's 1e-4 bar needs that: the
 * tensor core truncates when it accumulates) with one pass over X per 512 columns.  Same values as
 * pinot_b200_trsm_rlt to within fp32 rounding; needs ldl, ldx multiples of 4 and 16-byte aligned l, x (otherwise
 * it falls back to the plain schedule). */
PINOT_B200_API int64_t pinot_b200_trsm_rlt_workspace_bytes(int64_t n, int64_t r);
PINOT_B200_API int pinot_b200_trsm_rlt_ws(const float* l, int64_t ldl, int64_t n, float* x, int64_t ldx, int64_t r,
                           void* workspace, int64_t workspace_bytes, cudaStream_t stream);
/* Zero the strict upper triangle of an n x n matrix (torch.cholesky's output convention). */
PINOT_B200_API int pinot_b200_zero_upper(float* a, int64_t ld, int64_t n, cudaStream_t stream);
/* y[i] = sum_k m[i,k] v[k] (upper != 0: only k >= i), one warp per row.  alpha = L^{-T} z
 * (gaussian_process_regressor.py:117-119) and small matrix-vector products. */
PINOT_B200_API int pinot_b200_gemv_rows(const float* m, int64_t ld, int64_t rows, int64_t cols, int upper,
                         const float* v, float* y, cudaStream_t stream);
/* out3[0] = 0.5 z.z + trace(L) + 0.5 n log(2 pi)   (pinot's objective, gaussian_process_regressor.py:160-164)
 * out3[1] = 0.5 z.z + sum_i log L_ii + 0.5 n log(2 pi)   (exact negative log marginal likelihood)
 * out3[2] = z.z                          with z = L^{-1} y, diag_l = diag(L). */
PINOT_B200_API int pinot_b200_gp_nll(const float* diag_l, const float* z, int64_t n, float* out3,
                      cudaStream_t stream);
/* Closed-form gradient of the objective w.r.t. K (full symmetric n x n, scaled by *upstream if
 * given): 0.5 L^{-T} diag(L) L^{-1} - 0.5 alpha alpha^T   (exact_logdet = 0, trace(L) objective)
 *         0.5 K^{-1}                 - 0.5 alpha alpha^T   (exact_logdet = 1).
 * linv_t = L^{-T} (upper triangular, from the identity extension rows of the sweep).
 * workspace is optional (NULL / 0): given pinot_b200_gp_nll_grad_workspace_bytes(n) bytes (0 for
 * small n), the product runs on the TMA-fed tensor-core kernel from a pre-split copy of
 * L^{-T} diag(L)^{1/2}; without it the register-staged kernel is used.
 * Replaces autograd's cholesky_backward / triangular_solve_backward for :123-166. */
PINOT_B200_API int64_t pinot_b200_gp_nll_grad_workspace_bytes(int64_t n);
PINOT_B200_API int pinot_b200_gp_nll_grad(const float* linv_t, int64_t ld, int64_t n, const float* diag_l,
                           const float* alpha, const float* upstream, int exact_logdet, float* dk,
                           int64_t lddk, void* workspace, int64_t workspace_bytes, cudaStream_t stream);
/* Predictive marginals from vt = K_te,tr L^{-T} (n_te x n) and z = L^{-1} y:
 * mean[i] = vt[i,:].z ; var[i] = exp(*variance) + eps - |vt[i,:]|^2  (var may be NULL).
 * Equals the mean and the diagonal of the covariance at gaussian_process_regressor.py:212-221. */
PINOT_B200_API int pinot_b200_gp_predict_mean_var(const float* vt, int64_t ld, int64_t n_te, int64_t n,
                                   const float* z, const float* variance, float eps, float* mean,
                                   float* var, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* PINOT_B200_H_ */
