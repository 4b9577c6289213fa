// Interface of the single-launch (persistent, cooperative) Cholesky sweep, chol_sweep.cu.
#pragma once
#include "common.cuh"

namespace pb {

constexpr int64_t SWEEP_PERSISTENT_MAX_N = 6144;   // above this the rank-512 TMA-fed two-level sweep of chol.cu is faster
constexpr int SWEEP_MAX_ROW_TILES = 1024;          // 128-row tiles whose progress counters a CTA caches in shared memory

struct SweepParams {
    float* E; int64_t ld;
    int n, nd, ident, rows;        // matrix order, dense extension rows, identity rows present, total rows
    int NP, CT, RT;                // 64-column panels, 128-column tiles, 128-row tiles
    float* wdiag; float* winv;     // per panel: L11 and M = L11^-T (64 x 64 each)
    int* nfact;                    // number of factored panels (monotone)
    int* progress;                 // per tile row: panels solved so far (monotone)
    int* info; float* diag_out;
    int fused;                     // 1: regular diagonal tiles are processed on chip (diag_tile_fused)
};

bool sweep_persistent_enabled();
bool sweep_persistent_applicable(int64_t n, int64_t rows_total);
int64_t sweep_persistent_flag_bytes(int64_t n, int64_t n_ext_dense);
int launch_sweep_persistent(float* e, int64_t ld, int64_t n, int64_t n_ext_dense, int ext_identity, float* diag_out,
                            float* wdiag, float* winv, int* flags, int* info, cudaStream_t stream);

}  // namespace pb
