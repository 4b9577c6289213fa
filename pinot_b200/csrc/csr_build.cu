// Graph index build: everything the message-passing / readout kernels need from a batched edge list,
// in one C-ABI call (a dozen small launches, no host synchronisation):
//   CSR by destination and by source with DGL's edge order kept inside every row (what a STABLE sort by
//   dst / src gives: pinot/graph.py:238-239 edge order + dgl.batch concatenation, pinot/data/utils.py:260-276),
//   deg^-1/2 (clamped >= 1) for GraphConv(norm="both"), 1/deg for SAGEConv("mean"), per-graph node offsets
//   and the node -> graph map for dgl.sum_nodes.
// Counting sort: histogram -> exclusive scan -> scatter with atomic cursors (unordered inside a row) ->
// per-row insertion sort by edge id (rows of molecular graphs hold <= 6 entries).
#include "common.cuh"
#include <cuda_fp16.h>
#include "../../include/pinot_b200.h"

namespace {

constexpr int SCAN_THREADS = 512;
constexpr int SCAN_ITEMS = 8;
constexpr int SCAN_CHUNK = SCAN_THREADS * SCAN_ITEMS;   // 4096 elements per block
constexpr int MAX_PARTIALS = 16384;                     // one block scans the block sums: n <= 67 M

__global__ void __launch_bounds__(256)
degree_histogram_kernel(const int64_t* __restrict__ src, const int64_t* __restrict__ dst, int64_t n_edges, int64_t n_nodes,
                        int* __restrict__ cnt_in, int* __restrict__ cnt_out) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_edges) return;
    const int64_t s = __ldg(src + e), d = __ldg(dst + e);
    if ((uint64_t)s >= (uint64_t)n_nodes || (uint64_t)d >= (uint64_t)n_nodes) return;   // BatchedGraph(validate=True) reports these
    atomicAdd(cnt_in + d, 1);
    atomicAdd(cnt_out + s, 1);
}

template <typename T>
__device__ __forceinline__ int block_sum(const T* __restrict__ v, int64_t n, int64_t base, int* smem) {
    int acc = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        const int64_t k = base + (int64_t)i * SCAN_THREADS + threadIdx.x;
        if (k < n) acc += (int)v[k];
    }
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if ((threadIdx.x & 31) == 0) smem[threadIdx.x >> 5] = acc;
    __syncthreads();
    int tot = 0;
    if (threadIdx.x < 32) {
        tot = threadIdx.x < SCAN_THREADS / 32 ? smem[threadIdx.x] : 0;
        for (int o = 16; o > 0; o >>= 1) tot += __shfl_xor_sync(0xffffffffu, tot, o);
    }
    return tot;   // valid in warp 0
}

template <typename T>
__global__ void __launch_bounds__(SCAN_THREADS)
scan_partials_kernel(const T* __restrict__ v, int64_t n, int* __restrict__ partials) {
    __shared__ int smem[SCAN_THREADS / 32];
    const int tot = block_sum(v, n, (int64_t)blockIdx.x * SCAN_CHUNK, smem);
    if (threadIdx.x == 0) partials[blockIdx.x] = tot;
}

// exclusive scan of the block sums, in place, by one block
__global__ void __launch_bounds__(1024)
scan_block_sums_kernel(int* __restrict__ partials, int nb) {
    __shared__ int warp_tot[32];
    constexpr int PER = MAX_PARTIALS / 1024;   // 16 consecutive entries per thread
    int local[PER];
    int sum = 0;
    const int base = threadIdx.x * PER;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        local[i] = (base + i < nb) ? partials[base + i] : 0;
        sum += local[i];
    }
    int incl = sum;
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if ((threadIdx.x & 31) >= o) incl += t;
    }
    if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = incl;
    __syncthreads();
    if (threadIdx.x < 32) {
        int w = warp_tot[threadIdx.x];
        int wi = w;
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, wi, o);
            if (threadIdx.x >= o) wi += t;
        }
        warp_tot[threadIdx.x] = wi - w;   // exclusive prefix of the warp totals
    }
    __syncthreads();
    int run = warp_tot[threadIdx.x >> 5] + incl - sum;
#pragma unroll
    for (int i = 0; i < PER; ++i) {
        if (base + i < nb) partials[base + i] = run;
        run += local[i];
    }
}

// out[0] = 0, out[k + 1] = v[0] + ... + v[k]
template <typename T>
__global__ void __launch_bounds__(SCAN_THREADS)
scan_write_kernel(const T* __restrict__ v, int64_t n, const int* __restrict__ partials, int* __restrict__ out) {
    __shared__ int warp_tot[SCAN_THREADS / 32];
    const int64_t base = (int64_t)blockIdx.x * SCAN_CHUNK + (int64_t)threadIdx.x * SCAN_ITEMS;
    int local[SCAN_ITEMS];
    int sum = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        local[i] = (base + i < n) ? (int)v[base + i] : 0;
        sum += local[i];
    }
    int incl = sum;
    for (int o = 1; o < 32; o <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, o);
        if ((threadIdx.x & 31) >= o) incl += t;
    }
    if ((threadIdx.x & 31) == 31) warp_tot[threadIdx.x >> 5] = incl;
    __syncthreads();
    if (threadIdx.x < 32) {
        const int w = threadIdx.x < SCAN_THREADS / 32 ? warp_tot[threadIdx.x] : 0;
        int wi = w;
        for (int o = 1; o < 32; o <<= 1) {
            const int t = __shfl_up_sync(0xffffffffu, wi, o);
            if (threadIdx.x >= o) wi += t;
        }
        if (threadIdx.x < SCAN_THREADS / 32) warp_tot[threadIdx.x] = wi - w;
    }
    __syncthreads();
    int run = partials[blockIdx.x] + warp_tot[threadIdx.x >> 5] + incl - sum;
    if (blockIdx.x == 0 && threadIdx.x == 0) out[0] = 0;
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
        run += local[i];
        if (base + i < n) out[base + i + 1] = run;
    }
}

template <typename T>
int exclusive_scan(const T* v, int64_t n, int* out, int* partials, cudaStream_t stream) {
    if (n == 0) {
        PB_CUDA(cudaMemsetAsync(out, 0, sizeof(int), stream));
        return pb::PB_OK;
    }
    const int nb = (int)pb::ceil_div(n, SCAN_CHUNK);
    scan_partials_kernel<T><<<nb, SCAN_THREADS, 0, stream>>>(v, n, partials);
    int rc = pb::check_launch("csr_build/scan_partials");
    if (rc) return rc;
    scan_block_sums_kernel<<<1, 1024, 0, stream>>>(partials, nb);
    rc = pb::check_launch("csr_build/scan_block_sums");
    if (rc) return rc;
    scan_write_kernel<T><<<nb, SCAN_THREADS, 0, stream>>>(v, n, partials, out);
    return pb::check_launch("csr_build/scan_write");
}

__global__ void __launch_bounds__(256)
scatter_edge_ids_kernel(const int64_t* __restrict__ src, const int64_t* __restrict__ dst, int64_t n_edges, int64_t n_nodes,
                        const int* __restrict__ indptr, const int* __restrict__ indptr_t, int* __restrict__ cur_in,
                        int* __restrict__ cur_out, int* __restrict__ eid_in, int* __restrict__ eid_out) {
    const int64_t e = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= n_edges) return;
    const int64_t s = __ldg(src + e), d = __ldg(dst + e);
    if ((uint64_t)s >= (uint64_t)n_nodes || (uint64_t)d >= (uint64_t)n_nodes) return;
    eid_in[indptr[d] + atomicAdd(cur_in + d, 1)] = (int)e;
    eid_out[indptr_t[s] + atomicAdd(cur_out + s, 1)] = (int)e;
}

// One thread per row: order the row's edge ids ascending (= the original edge order, what a stable sort
// by the row key keeps) and emit the other endpoint of every edge.
__global__ void __launch_bounds__(256)
order_rows_kernel(const int* __restrict__ indptr, int* __restrict__ eid, const int64_t* __restrict__ other,
                  int* __restrict__ indices, int64_t n_nodes) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    const int beg = indptr[i], end = indptr[i + 1];
    for (int a = beg + 1; a < end; ++a) {
        const int key = eid[a];
        int b = a - 1;
        while (b >= beg && eid[b] > key) {
            eid[b + 1] = eid[b];
            --b;
        }
        eid[b + 1] = key;
    }
    for (int a = beg; a < end; ++a) indices[a] = (int)__ldg(other + eid[a]);
}

__global__ void __launch_bounds__(256)
node_vectors_kernel(const int* __restrict__ indptr, const int* __restrict__ indptr_t, const int* __restrict__ seg_offsets,
                    int n_graphs, int64_t n_nodes, float* __restrict__ in_norm, float* __restrict__ out_norm,
                    float* __restrict__ inv_in_deg, int64_t* __restrict__ in_deg, int* __restrict__ node_seg) {
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_nodes) return;
    const int din = indptr[i + 1] - indptr[i], dout = indptr_t[i + 1] - indptr_t[i];
    const float fin = fmaxf((float)din, 1.f), fout = fmaxf((float)dout, 1.f);
    in_norm[i] = rsqrtf(fin);
    out_norm[i] = rsqrtf(fout);
    inv_in_deg[i] = 1.0f / fin;
    in_deg[i] = din;
    // graph of node i: last g with seg_offsets[g] <= i
    int lo = 0, hi = n_graphs;          // invariant: seg_offsets[lo] <= i < seg_offsets[hi]
    while (hi - lo > 1) {
        const int mid = (lo + hi) >> 1;
        if (seg_offsets[mid] <= (int)i) lo = mid; else hi = mid;
    }
    node_seg[i] = lo;
}

// Atom features from their compact host form (the reference builds them per atom: a one-hot of the atomic number,
// pinot/graph.py:163-168, followed by 17 small integers / flags, :50-69 and :173): one output element per thread,
// coalesced fp32 stores; the 1 + 2 * n_fp bytes per atom read here are re-read from L1.
__device__ __forceinline__ float fp_value(const __half* p) { return __half2float(__ldg(p)); }
__device__ __forceinline__ float fp_value(const float* p) { return __ldg(p); }

template <typename T>
__global__ void __launch_bounds__(256)
expand_atom_features_kernel(const uint8_t* __restrict__ z, const T* __restrict__ fp, int64_t n_atoms, int n_onehot, int n_fp,
                            float* __restrict__ out, int64_t ld) {
    const int width = n_onehot + n_fp;
    const int64_t total = n_atoms * width;
    for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
        const int64_t a = i / width;
        const int c = (int)(i - a * width);
        float v;
        if (c < n_onehot) v = (__ldg(z + a) == c) ? 1.f : 0.f;
        else v = fp_value(fp + a * n_fp + (c - n_onehot));
        out[a * ld + c] = v;
    }
}

}  // namespace

extern "C" {

int pinot_b200_expand_atom_features(const uint8_t* z, const void* fp, int fp_bytes, int64_t n_atoms, int64_t n_onehot,
                                    int64_t n_fp, float* out, int64_t ld, cudaStream_t stream) {
    PB_REQUIRE(n_atoms >= 0 && n_onehot >= 0 && n_onehot <= 256 && n_fp >= 0 && n_fp < (1 << 16) && ld >= n_onehot + n_fp,
               "expand_atom_features: invalid sizes");
    PB_REQUIRE(fp_bytes == 2 || fp_bytes == 4, "expand_atom_features: fp_bytes must be 2 (fp16) or 4 (fp32)");
    if (n_atoms == 0 || n_onehot + n_fp == 0) return pb::PB_OK;
    PB_REQUIRE(out && (n_onehot == 0 || z) && (n_fp == 0 || fp), "expand_atom_features: null pointer");
    const int64_t total = n_atoms * (n_onehot + n_fp);
    const int64_t blocks = pb::ceil_div(total, 256);
    const unsigned grid = (unsigned)(blocks < 148 * 16 ? blocks : 148 * 16);
    if (fp_bytes == 2)
        expand_atom_features_kernel<__half><<<grid, 256, 0, stream>>>(z, static_cast<const __half*>(fp), n_atoms, (int)n_onehot,
                                                                      (int)n_fp, out, ld);
    else
        expand_atom_features_kernel<float><<<grid, 256, 0, stream>>>(z, static_cast<const float*>(fp), n_atoms, (int)n_onehot,
                                                                     (int)n_fp, out, ld);
    return pb::check_launch("expand_atom_features");
}


int64_t pinot_b200_csr_build_workspace_bytes(int64_t n_nodes, int64_t n_edges) {
    if (n_nodes < 0 || n_edges < 0) return -1;
    return (int64_t)sizeof(int) * (2 * n_nodes + 2 * n_edges + MAX_PARTIALS) + 256;
}

int pinot_b200_csr_build(const int64_t* src, const int64_t* dst, int64_t n_edges, int64_t n_nodes,
                         const int64_t* batch_num_nodes, int64_t n_graphs, int32_t* indptr, int32_t* indices,
                         int32_t* indptr_t, int32_t* indices_t, float* in_norm, float* out_norm, float* inv_in_deg,
                         int64_t* in_deg, int32_t* seg_offsets, int32_t* node_seg, void* workspace,
                         int64_t workspace_bytes, cudaStream_t stream) {
    PB_REQUIRE(n_edges >= 0 && n_nodes >= 0 && n_graphs >= 0, "csr_build: negative size");
    PB_REQUIRE(n_nodes < (1ll << 31) && n_edges < (1ll << 31), "csr_build: graphs are limited to 2^31 nodes / edges (int32 CSR)");
    PB_REQUIRE(n_nodes <= (int64_t)SCAN_CHUNK * MAX_PARTIALS && n_graphs <= (int64_t)SCAN_CHUNK * MAX_PARTIALS,
               "csr_build: more than %lld nodes / graphs is not supported", (long long)SCAN_CHUNK * MAX_PARTIALS);
    PB_REQUIRE(indptr && indptr_t && seg_offsets && workspace, "csr_build: null pointer");
    PB_REQUIRE(workspace_bytes >= pinot_b200_csr_build_workspace_bytes(n_nodes, n_edges), "csr_build: workspace too small");
    PB_REQUIRE(n_edges == 0 || (src && dst && indices && indices_t), "csr_build: null edge arrays");
    PB_REQUIRE(n_nodes == 0 || (in_norm && out_norm && inv_in_deg && in_deg && node_seg), "csr_build: null node arrays");
    PB_REQUIRE(n_graphs == 0 || batch_num_nodes, "csr_build: null batch_num_nodes");
    int* cnt_in = static_cast<int*>(workspace);
    int* cnt_out = cnt_in + n_nodes;
    int* eid_in = cnt_out + n_nodes;
    int* eid_out = eid_in + n_edges;
    int* partials = eid_out + n_edges;
    int rc;
    if (n_nodes > 0) PB_CUDA(cudaMemsetAsync(cnt_in, 0, sizeof(int) * 2 * n_nodes, stream));
    if (n_edges > 0) {
        degree_histogram_kernel<<<(unsigned)pb::ceil_div(n_edges, 256), 256, 0, stream>>>(src, dst, n_edges, n_nodes, cnt_in, cnt_out);
        if ((rc = pb::check_launch("csr_build/histogram"))) return rc;
    }
    if ((rc = exclusive_scan<int>(cnt_in, n_nodes, indptr, partials, stream))) return rc;
    if ((rc = exclusive_scan<int>(cnt_out, n_nodes, indptr_t, partials, stream))) return rc;
    if ((rc = exclusive_scan<int64_t>(batch_num_nodes, n_graphs, seg_offsets, partials, stream))) return rc;
    if (n_edges > 0) {
        PB_CUDA(cudaMemsetAsync(cnt_in, 0, sizeof(int) * 2 * n_nodes, stream));   // the counters become cursors
        scatter_edge_ids_kernel<<<(unsigned)pb::ceil_div(n_edges, 256), 256, 0, stream>>>(
            src, dst, n_edges, n_nodes, indptr, indptr_t, cnt_in, cnt_out, eid_in, eid_out);
        if ((rc = pb::check_launch("csr_build/scatter"))) return rc;
        const unsigned gb = (unsigned)pb::ceil_div(n_nodes, 256);
        order_rows_kernel<<<gb, 256, 0, stream>>>(indptr, eid_in, src, indices, n_nodes);
        if ((rc = pb::check_launch("csr_build/order_rows"))) return rc;
        order_rows_kernel<<<gb, 256, 0, stream>>>(indptr_t, eid_out, dst, indices_t, n_nodes);
        if ((rc = pb::check_launch("csr_build/order_rows_t"))) return rc;
    }
    if (n_nodes > 0) {
        node_vectors_kernel<<<(unsigned)pb::ceil_div(n_nodes, 256), 256, 0, stream>>>(
            indptr, indptr_t, seg_offsets, (int)n_graphs, n_nodes, in_norm, out_norm, inv_in_deg, in_deg, node_seg);
        if ((rc = pb::check_launch("csr_build/node_vectors"))) return rc;
    }
    return pb::PB_OK;
}

}  // extern "C"
