// Kernel (1): CSR edge-message gather + segment-sum + degree normalisation, and the
// per-graph sum/mean readout.  Replaces what DGL's GraphConv / SAGEConv("mean") /
// dgl.sum_nodes dispatch to (update_all(copy_src, sum|mean) -> SpMM; segment reduce) on the
// path pinot/representation/dgl_legacy.py:40-50 and pinot/representation/sequential.py:78-103.
//
// HBM-bound: one pass over X (read) and R (write).  Each thread owns one 16-byte column slice
// of one destination row, so a warp reads/writes whole 512-byte rows (H = 128) fully coalesced;
// neighbour rows belong to the same molecule and are served from L1/L2.
#include "common.cuh"
#include "../../include/pinot_b200.h"
#include <stdlib.h>

namespace {

// out[i, :] = dst_scale[i] * sum_{e in [indptr[i], indptr[i+1])} src_scale[idx[e]] * x[idx[e], :]
//
// Each thread owns one 16-byte (VEC = 4) column slice of ITEMS destination rows that are
// `row_stride` apart, and walks their edge lists in lock-step: the ITEMS row-pointer loads, then
// the ITEMS index loads, then the ITEMS neighbour-row gathers of one edge slot are all issued
// before any of them is consumed.  The three dependent load levels (pointer -> index -> row) are
// what limits a one-row-per-thread version to about half of the HBM rate; ITEMS independent chains
// per thread keep enough bytes in flight.

// FV = float4 slices per row when known at compile time (32 for H = 128, 16 for H = 64, 8 for H = 32),
// 0 = run-time `feat_v`.  Offsets are 32-bit (in units of one slice): the host falls back to the wide
// kernel when n_rows * feat_v does not fit.
template <int VEC, int FV, int MP_ITEMS, int MINB>
__global__ void __launch_bounds__(256, MINB)
mp_aggregate_kernel(const float* __restrict__ x, float* __restrict__ out,
                    const int32_t* __restrict__ indptr, const int32_t* __restrict__ indices,
                    const float* __restrict__ src_scale, const float* __restrict__ dst_scale,
                    int n_dst, int feat_v_rt) {
    const int feat_v = FV > 0 ? FV : feat_v_rt;
    const int rows_per_block = 256 / feat_v;             // feat_v <= 256 on this path
    const int lrow = threadIdx.x / feat_v;
    const int c = threadIdx.x - lrow * feat_v;
    if (lrow >= rows_per_block) return;
    const int row0 = (int)blockIdx.x * MP_ITEMS * rows_per_block + lrow;

    int beg[MP_ITEMS], deg[MP_ITEMS];
    float acc[MP_ITEMS][VEC];
#pragma unroll
    for (int i = 0; i < MP_ITEMS; ++i) {
        const int row = row0 + i * rows_per_block;
        beg[i] = 0;
        deg[i] = 0;
        if (row < n_dst) {
            const int2 pq = make_int2(__ldg(indptr + row), __ldg(indptr + row + 1));
            beg[i] = pq.x;
            deg[i] = pq.y - pq.x;
        }
#pragma unroll
        for (int v = 0; v < VEC; ++v) acc[i][v] = 0.f;
    }
    int maxdeg = 0;
#pragma unroll
    for (int i = 0; i < MP_ITEMS; ++i) maxdeg = max(maxdeg, deg[i]);

    for (int e = 0; e < maxdeg; ++e) {
        int j[MP_ITEMS];
        float s[MP_ITEMS];
#pragma unroll
        for (int i = 0; i < MP_ITEMS; ++i) j[i] = (e < deg[i]) ? __ldg(indices + beg[i] + e) : -1;
        if (src_scale) {
#pragma unroll
            for (int i = 0; i < MP_ITEMS; ++i) s[i] = (j[i] >= 0) ? __ldg(src_scale + j[i]) : 0.f;
        } else {
#pragma unroll
            for (int i = 0; i < MP_ITEMS; ++i) s[i] = (j[i] >= 0) ? 1.f : 0.f;
        }
        if (VEC == 4) {
            const float4* x4 = reinterpret_cast<const float4*>(x);
            float4 v[MP_ITEMS];
#pragma unroll
            for (int i = 0; i < MP_ITEMS; ++i) v[i] = __ldg(x4 + (unsigned)(max(j[i], 0) * feat_v + c));
#pragma unroll
            for (int i = 0; i < MP_ITEMS; ++i) {
                acc[i][0] = fmaf(s[i], v[i].x, acc[i][0]);
                acc[i][1 % VEC] = fmaf(s[i], v[i].y, acc[i][1 % VEC]);
                acc[i][2 % VEC] = fmaf(s[i], v[i].z, acc[i][2 % VEC]);
                acc[i][3 % VEC] = fmaf(s[i], v[i].w, acc[i][3 % VEC]);
            }
        } else {
            float v[MP_ITEMS];
#pragma unroll
            for (int i = 0; i < MP_ITEMS; ++i) v[i] = __ldg(x + (unsigned)(max(j[i], 0) * feat_v + c));
#pragma unroll
            for (int i = 0; i < MP_ITEMS; ++i) acc[i][0] = fmaf(s[i], v[i], acc[i][0]);
        }
    }
#pragma unroll
    for (int i = 0; i < MP_ITEMS; ++i) {
        const int row = row0 + i * rows_per_block;
        if (row >= n_dst) continue;
        const float d = dst_scale ? __ldg(dst_scale + row) : 1.f;
        if (VEC == 4) {
            reinterpret_cast<float4*>(out)[(unsigned)(row * feat_v + c)] =
                make_float4(acc[i][0] * d, acc[i][1 % VEC] * d, acc[i][2 % VEC] * d, acc[i][3 % VEC] * d);
        } else {
            out[(unsigned)(row * feat_v + c)] = acc[i][0] * d;
        }
    }
}

// Wide rows (more than 256 column slices per row): one thread per (row, slice), plain loop.
template <int VEC>
__global__ void __launch_bounds__(256)
mp_aggregate_wide_kernel(const float* __restrict__ x, float* __restrict__ out,
                         const int32_t* __restrict__ indptr, const int32_t* __restrict__ indices,
                         const float* __restrict__ src_scale, const float* __restrict__ dst_scale,
                         int64_t n_dst, int feat_v) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t row = t / feat_v;
    if (row >= n_dst) return;
    const int c = (int)(t - row * feat_v);
    const int beg = __ldg(indptr + row), end = __ldg(indptr + row + 1);
    float acc[VEC];
#pragma unroll
    for (int v = 0; v < VEC; ++v) acc[v] = 0.f;
    for (int e = beg; e < end; ++e) {
        const int j = __ldg(indices + e);
        const float s = src_scale ? __ldg(src_scale + j) : 1.f;
        if (VEC == 4) {
            const float4 v = pb::ldg4(x + ((int64_t)j * feat_v + c) * 4);
            acc[0] = fmaf(s, v.x, acc[0]);
            acc[1 % VEC] = fmaf(s, v.y, acc[1 % VEC]);
            acc[2 % VEC] = fmaf(s, v.z, acc[2 % VEC]);
            acc[3 % VEC] = fmaf(s, v.w, acc[3 % VEC]);
        } else {
            acc[0] = fmaf(s, __ldg(x + (int64_t)j * feat_v + c), acc[0]);
        }
    }
    const float d = dst_scale ? __ldg(dst_scale + row) : 1.f;
    if (VEC == 4) {
        *reinterpret_cast<float4*>(out + (row * feat_v + c) * 4) =
            make_float4(acc[0] * d, acc[1 % VEC] * d, acc[2 % VEC] * d, acc[3 % VEC] * d);
    } else {
        out[row * feat_v + c] = acc[0] * d;
    }
}

// out[g, :] = (mean ? 1/max(n_g,1) : 1) * sum_{n in [off[g], off[g+1])} x[n, :]
template <int VEC>
__global__ void __launch_bounds__(256)
segment_readout_kernel(const float* __restrict__ x, float* __restrict__ out,
                       const int32_t* __restrict__ offsets, int64_t n_seg, int feat_v, int mean) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t g = t / feat_v;
    if (g >= n_seg) return;
    const int c = (int)(t - g * feat_v);
    const int beg = __ldg(offsets + g), end = __ldg(offsets + g + 1);
    float acc[4] = {0.f, 0.f, 0.f, 0.f};
    int n = beg;
    for (; n + 4 <= end; n += 4) {
        if (VEC == 4) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = pb::ldg4(x + ((int64_t)(n + u) * feat_v + c) * 4);
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                acc[0] += v[u].x; acc[1] += v[u].y; acc[2] += v[u].z; acc[3] += v[u].w;
            }
        } else {
            float v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = __ldg(x + (int64_t)(n + u) * feat_v + c);
#pragma unroll
            for (int u = 0; u < 4; ++u) acc[0] += v[u];
        }
    }
    for (; n < end; ++n) {
        if (VEC == 4) {
            const float4 v = pb::ldg4(x + ((int64_t)n * feat_v + c) * 4);
            acc[0] += v.x; acc[1] += v.y; acc[2] += v.z; acc[3] += v.w;
        } else {
            acc[0] += __ldg(x + (int64_t)n * feat_v + c);
        }
    }
    const float sc = mean ? 1.f / (float)max(end - beg, 1) : 1.f;
    if (VEC == 4) {
        *reinterpret_cast<float4*>(out + (g * feat_v + c) * 4) =
            make_float4(acc[0] * sc, acc[1] * sc, acc[2] * sc, acc[3] * sc);
    } else {
        out[g * feat_v + c] = acc[0] * sc;
    }
}

// dx[n, :] = (mean ? 1/max(n_g,1) : 1) * dout[node_seg[n], :]
template <int VEC>
__global__ void __launch_bounds__(256)
segment_readout_bwd_kernel(const float* __restrict__ dout, float* __restrict__ dx,
                           const int32_t* __restrict__ offsets, const int32_t* __restrict__ node_seg,
                           int64_t n_nodes, int feat_v, int mean) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t n = t / feat_v;
    if (n >= n_nodes) return;
    const int c = (int)(t - n * feat_v);
    const int g = __ldg(node_seg + n);
    float sc = 1.f;
    if (mean) sc = 1.f / (float)max(__ldg(offsets + g + 1) - __ldg(offsets + g), 1);
    if (VEC == 4) {
        float4 v = pb::ldg4(dout + ((int64_t)g * feat_v + c) * 4);
        v.x *= sc; v.y *= sc; v.z *= sc; v.w *= sc;
        *reinterpret_cast<float4*>(dx + (n * feat_v + c) * 4) = v;
    } else {
        dx[n * feat_v + c] = sc * __ldg(dout + (int64_t)g * feat_v + c);
    }
}

// out[i, :] = x[idx[i], :]  (row gather: the feature / embedding side of re-batching a selection of
// molecules out of a device-resident pool)
template <int VEC>
__global__ void __launch_bounds__(256)
gather_rows_kernel(const float* __restrict__ x, const int64_t* __restrict__ idx, float* __restrict__ out,
                   int64_t n_out, int feat_v) {
    const int64_t t = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t i = t / feat_v;
    if (i >= n_out) return;
    const int c = (int)(t - i * feat_v);
    const int64_t j = __ldg(idx + i);
    if (VEC == 4) {
        *reinterpret_cast<float4*>(out + (i * feat_v + c) * 4) = pb::ldg4(x + (j * feat_v + c) * 4);
    } else {
        out[i * feat_v + c] = __ldg(x + j * feat_v + c);
    }
}

inline bool can_vec4(int64_t feat, const void* a, const void* b) {
    return (feat % 4 == 0) && pb::aligned16(a) && pb::aligned16(b);
}

int launch_aggregate(const float* x, float* out, const int32_t* indptr, const int32_t* indices,
                     const float* src_scale, const float* dst_scale, int64_t n_dst, int64_t feat,
                     cudaStream_t stream, const char* who) {
    PB_REQUIRE(n_dst >= 0 && feat > 0, "%s: n_dst=%lld feat=%lld invalid", who, (long long)n_dst, (long long)feat);
    PB_REQUIRE(feat <= (1 << 20), "%s: feat too large", who);
    if (n_dst == 0) return pb::PB_OK;
    PB_REQUIRE(x && out && indptr, "%s: null pointer", who);
    PB_REQUIRE(x != out, "%s: in-place aggregation is not supported", who);
    const bool vec = can_vec4(feat, x, out);
    const int64_t fv = vec ? feat / 4 : feat;
    // the row-batched kernel indexes with 32-bit slice offsets; padded rows (j = -1 -> row 0) read row 0
    const bool fits32 = fv <= 256 && n_dst * fv < (1ll << 31);
    if (fits32) {
        static const int variant = [] { const char* v = getenv("PINOT_B200_MP_VARIANT"); return v ? atoi(v) : 0; }();
        const int rows_per_block = 256 / (int)fv;
#define PB_LAUNCH_MP(V, F, I, B)                                                                              \
        mp_aggregate_kernel<V, F, I, B><<<(unsigned)pb::ceil_div(n_dst, (int64_t)rows_per_block * I), 256, 0, stream>>>( \
            x, out, indptr, indices, src_scale, dst_scale, (int)n_dst, (int)fv)
#define PB_LAUNCH_MP_V(I, B)                                   \
        do {                                                   \
            if (vec && fv == 32) PB_LAUNCH_MP(4, 32, I, B);    \
            else if (vec && fv == 16) PB_LAUNCH_MP(4, 16, I, B); \
            else if (vec && fv == 8) PB_LAUNCH_MP(4, 8, I, B); \
            else if (vec) PB_LAUNCH_MP(4, 0, I, B);            \
            else PB_LAUNCH_MP(1, 0, I, B);                     \
        } while (0)
        if (variant == 1) PB_LAUNCH_MP_V(2, 8);
        else if (variant == 2) PB_LAUNCH_MP_V(2, 6);
        else if (variant == 3) PB_LAUNCH_MP_V(1, 8);
        else PB_LAUNCH_MP_V(4, 5);
#undef PB_LAUNCH_MP_V
#undef PB_LAUNCH_MP
    } else if (vec) {
        mp_aggregate_wide_kernel<4><<<(unsigned)pb::ceil_div(n_dst * fv, 256), 256, 0, stream>>>(
            x, out, indptr, indices, src_scale, dst_scale, n_dst, (int)fv);
    } else {
        mp_aggregate_wide_kernel<1><<<(unsigned)pb::ceil_div(n_dst * fv, 256), 256, 0, stream>>>(
            x, out, indptr, indices, src_scale, dst_scale, n_dst, (int)fv);
    }
    return pb::check_launch(who);
}

}  // namespace

extern "C" {

int pinot_b200_mp_csr_aggregate_fwd(const float* x, float* out, const int32_t* indptr,
                                    const int32_t* indices, const float* src_scale,
                                    const float* dst_scale, int64_t n_dst, int64_t feat,
                                    cudaStream_t stream) {
    return launch_aggregate(x, out, indptr, indices, src_scale, dst_scale, n_dst, feat, stream,
                            "mp_csr_aggregate_fwd");
}

int pinot_b200_mp_csr_aggregate_bwd(const float* dout, float* dx, const int32_t* indptr_t,
                                    const int32_t* indices_t, const float* src_scale,
                                    const float* dst_scale, int64_t n_src, int64_t feat,
                                    cudaStream_t stream) {
    // d x[j] = src_scale[j] * sum_{i : j->i} dst_scale[i] * dout[i]: the same gather on the
    // transposed CSR with the two scale vectors trading places.
    return launch_aggregate(dout, dx, indptr_t, indices_t, dst_scale, src_scale, n_src, feat, stream,
                            "mp_csr_aggregate_bwd");
}

int pinot_b200_gather_rows(const float* x, const int64_t* idx, float* out, int64_t n_out, int64_t feat,
                           cudaStream_t stream) {
    PB_REQUIRE(n_out >= 0 && feat > 0, "gather_rows: invalid sizes");
    if (n_out == 0) return pb::PB_OK;
    PB_REQUIRE(x && idx && out, "gather_rows: null pointer");
    if (can_vec4(feat, x, out)) {
        const int fv = (int)(feat / 4);
        gather_rows_kernel<4><<<(unsigned)pb::ceil_div(n_out * fv, 256), 256, 0, stream>>>(x, idx, out, n_out, fv);
    } else {
        gather_rows_kernel<1><<<(unsigned)pb::ceil_div(n_out * feat, 256), 256, 0, stream>>>(x, idx, out, n_out,
                                                                                         (int)feat);
    }
    return pb::check_launch("gather_rows");
}

int pinot_b200_segment_readout_fwd(const float* x, float* out, const int32_t* offsets,
                                   int64_t n_seg, int64_t feat, int mean, cudaStream_t stream) {
    PB_REQUIRE(n_seg >= 0 && feat > 0, "segment_readout_fwd: invalid sizes");
    if (n_seg == 0) return pb::PB_OK;
    PB_REQUIRE(x && out && offsets, "segment_readout_fwd: null pointer");
    if (can_vec4(feat, x, out)) {
        const int fv = (int)(feat / 4);
        segment_readout_kernel<4><<<(unsigned)pb::ceil_div(n_seg * fv, 256), 256, 0, stream>>>(
            x, out, offsets, n_seg, fv, mean);
    } else {
        segment_readout_kernel<1><<<(unsigned)pb::ceil_div(n_seg * feat, 256), 256, 0, stream>>>(
            x, out, offsets, n_seg, (int)feat, mean);
    }
    return pb::check_launch("segment_readout_fwd");
}

int pinot_b200_segment_readout_bwd(const float* dout, float* dx, const int32_t* offsets,
                                   const int32_t* node_seg, int64_t n_nodes, int64_t feat, int mean,
                                   cudaStream_t stream) {
    PB_REQUIRE(n_nodes >= 0 && feat > 0, "segment_readout_bwd: invalid sizes");
    if (n_nodes == 0) return pb::PB_OK;
    PB_REQUIRE(dout && dx && offsets && node_seg, "segment_readout_bwd: null pointer");
    if (can_vec4(feat, dout, dx)) {
        const int fv = (int)(feat / 4);
        segment_readout_bwd_kernel<4><<<(unsigned)pb::ceil_div(n_nodes * fv, 256), 256, 0, stream>>>(
            dout, dx, offsets, node_seg, n_nodes, fv, mean);
    } else {
        segment_readout_bwd_kernel<1><<<(unsigned)pb::ceil_div(n_nodes * feat, 256), 256, 0, stream>>>(
            dout, dx, offsets, node_seg, n_nodes, (int)feat, mean);
    }
    return pb::check_launch("segment_readout_bwd");
}

}  // extern "C"
