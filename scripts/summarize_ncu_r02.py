"""Turns the round-2 `ncu --set full` reports (scripts/gpu_run_ncu_r02.sh) into profiles/r02/ncu_full_gram.txt,
profiles/r02/ncu_full_sweep.txt and profiles/r02/ncu_traffic.json.
Usage: ncu -i gpurun_out/r02_gram.ncu-rep --page raw --csv > /tmp/r02_gram.csv   (same for r02_sweep)
       python scripts/summarize_ncu_r02.py /tmp/r02_gram.csv /tmp/r02_sweep.csv"""
import csv
import json
import sys

KEYS = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'lts__t_bytes.sum',
        'l1tex__m_xbar2l1tex_read_bytes_mem_global_op_tma_ld.sum',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'dram__throughput.avg.pct_of_peak_sustained_elapsed',
        'lts__t_sector_hit_rate.pct', 'launch__registers_per_thread', 'launch__occupancy_limit_shared_mem',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'smsp__inst_executed.sum', 'smsp__inst_executed_op_tma_ld.sum', 'sm__cycles_elapsed.avg',
        'smsp__average_warp_latency_issue_stalled_no_instruction.ratio', 'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio']
MULT = {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1}
TIME = {'us': 1.0, 'ms': 1e3, 'ns': 1e-3, 's': 1e6, 'usecond': 1.0, 'msecond': 1e3, 'nsecond': 1e-3, 'second': 1e6}


def load(path):
    rows = list(csv.reader(open(path)))
    hdr, units = rows[0], rows[1]
    idx = {h: i for i, h in enumerate(hdr)}
    return rows[2:], units, idx


def us(r, units, idx):
    return float(r[idx['gpu__time_duration.sum']]) * TIME[units[idx['gpu__time_duration.sum']]]


def dram(r, units, idx):
    return sum(float(r[idx[k]]) * MULT[units[idx[k]]] for k in ('dram__bytes_read.sum', 'dram__bytes_write.sum'))


def block(r, units, idx, what):
    out = ["== %s" % r[idx['Kernel Name']][:110], "   what: %s" % what, "   grid %s block %s" % (r[idx['Grid Size']], r[idx['Block Size']])]
    for k in KEYS:
        if k in idx:
            out.append("   %-82s %s %s" % (k, r[idx[k]], units[idx[k]]))
    out.append("   DRAM traffic of this launch (read + write): %.1f MB" % (dram(r, units, idx) / 1e6))
    out.append("")
    return out


traffic = {}
# ---------------- Gram ----------------
rows, units, idx = load(sys.argv[1])
what = {
    '<3, 0>': "norm pass: diagonal tiles of x x^T only (33 at N = 4200, 128 at N = 16384)",
    '<1, 0>': "forward, RBF: 3xTF32 GEMM + distance-to-kernel epilogue (launches 1-2: full K(x,x) by lower tiles + mirrored "
              "stores, 3-4: lower tiles only, both N = 4200 on the bench's embeddings with 15 381 near pairs; last: N = 16384 full, random rows)",
    '<2, 0>': "backward coefficient pass, RBF, N = 4200, symmetric mode (reads dK, writes the coefficient matrix C, row sums, near pairs)",
}
lines = ["# ncu --set full --clock-control none --import-source on -k regex:gram_tma_kernel, python scripts/profile_r02_kernels.py gram",
         "# every captured launch, in launch order; values as ncu reports them (units alongside); cold cache (ncu flushes L2 between launches)", ""]
fwd, bwd = [], []
for r in rows:
    name = r[idx['Kernel Name']]
    key = '<3, 0>' if 'gram_tma_kernel<3' in name else '<1, 0>' if 'gram_tma_kernel<1' in name else '<2, 0>'
    lines += block(r, units, idx, what[key])
    if key == '<1, 0>':
        fwd.append(r)
    if key == '<2, 0>':
        bwd.append(r)
open('profiles/r02/ncu_full_gram.txt', 'w').write("\n".join(lines))
lower = fwd[2:4]      # lower tiles only at N = 4200: what the training step launches
traffic["gram_fwd"] = {"dram_bytes_per_launch": sum(dram(r, units, idx) for r in lower) / len(lower),
                       "duration_us": sum(us(r, units, idx) for r in lower) / len(lower),
                       "what": "gram_tma_kernel<1,0>, lower tiles, N = 4200 (the GEMM + epilogue launch of pinot_b200_gram_fwd)"}
traffic["gram_fwd_n16384"] = {"dram_bytes_per_launch": dram(fwd[-1], units, idx), "duration_us": us(fwd[-1], units, idx),
                              "what": "gram_tma_kernel<1,0>, full K(x,x), N = 16384: 1.07 GB written"}
traffic["gram_bwd"] = {"dram_bytes_per_launch": sum(dram(r, units, idx) for r in bwd) / len(bwd),
                       "duration_us": sum(us(r, units, idx) for r in bwd) / len(bwd),
                       "what": "gram_tma_kernel<2,0>, N = 4200 (the coefficient pass of pinot_b200_gram_bwd; the C Ys GEMM and the finish kernel follow it)"}
# ---------------- sweep ----------------
rows, units, idx = load(sys.argv[2])
lines = ["# ncu --set full --clock-control none --import-source on -k regex:tgemm_kernel|chol_diag_kernel --launch-skip 67 -c 8,",
         "# python scripts/profile_r02_kernels.py sweep: the first launches of one sweep at N = 4200 with y row and identity rows",
         "# (chol_diag_kernel, the first panel solve, then steps 1-6 of the 65 look-ahead updates); cold cache", ""]
steps = []
for r in rows:
    name = r[idx['Kernel Name']]
    if 'chol_diag_kernel' in name:
        w = "stand-alone factor of the first 64 x 64 diagonal block (+ its transposed inverse)"
    elif 'tgemm_kernel<0, 1, 0, 1' in name:
        w = "one sweep step: rank-64 update of the trailing matrix with the look-ahead factor of the next diagonal block and the tensor-core panel solve fused (tile column 0)"
        steps.append(r)
    else:
        w = "panel solve of the first block column as a GEMM with L11^-T"
    lines += block(r, units, idx, w)
open('profiles/r02/ncu_full_sweep.txt', 'w').write("\n".join(lines))
per = sum(dram(r, units, idx) for r in steps) / len(steps)
traffic["tgemm_kernel"] = {"dram_bytes_per_launch": per, "duration_us": sum(us(r, units, idx) for r in steps) / len(steps),
                           "launches_per_call": 65,
                           "what": "tgemm_kernel<0,1,0,1,128>, steps 1-6 of the sweep at N = 4200; COLD cache: ncu flushes L2 between "
                                   "launches, so every step re-reads its trailing matrix from DRAM (39 MB); inside the training "
                                   "step the 71 MB stacked matrix stays in the 126 MB L2 across the 65 launches"}
# kernels that did not change since round 1 keep their round-1 capture
old = json.load(open('profiles/r01/ncu_traffic.json'))["kernels"]
for k in ("tgemm_tma_kernel", "split_tf32", "mp_aggregate", "node_gemm_c2", "node_gemm_pool"):
    traffic[k] = dict(old[k], what="round-1 capture (profiles/r01/ncu_full_top_kernels_final.txt); kernel unchanged")
json.dump({"source": "profiles/r02/ncu_full_gram.txt, profiles/r02/ncu_full_sweep.txt (ncu --set full, scripts/profile_r02_kernels.py)",
           "kernels": traffic}, open('profiles/r02/ncu_traffic.json', 'w'), indent=1)
for k, v in traffic.items():
    print(k, v["dram_bytes_per_launch"], v["duration_us"])
