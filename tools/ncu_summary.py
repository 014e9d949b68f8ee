#!/usr/bin/env python
"""Summarise ncu outputs: `launches` = per-kernel share of a launch list csv; `raw` = key metrics of a .ncu-rep."""
import collections
import csv
import subprocess
import sys

KEYS = ['gpu__time_duration.sum', 'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
        'smsp__thread_inst_executed_per_inst_executed.ratio', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct', 'lts__t_bytes.sum',
        'smsp__sass_thread_inst_executed_op_ffma_pred_on.sum', 'smsp__sass_thread_inst_executed_op_fmul_pred_on.sum',
        'smsp__sass_thread_inst_executed_op_fadd_pred_on.sum', 'smsp__sass_thread_inst_executed_op_dfma_pred_on.sum',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_dispatch_stall_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_imc_miss_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_sleeping_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_membar_per_issue_active.ratio']


def launches(path, last=None):
    rows = list(csv.reader(open(path)))
    h = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
    H = rows[h]
    ki, vi = H.index('Kernel Name'), H.index('Metric Value')
    seq = [(r[ki].split('(')[0][:64], float(r[vi].replace(',', ''))) for r in rows[h + 2:] if len(r) > vi]
    if last:
        seq = seq[-last:]
    agg = collections.OrderedDict()
    for n, v in seq:
        a = agg.setdefault(n, [0, 0.0])
        a[0] += 1
        a[1] += v
    tot = sum(a[1] for a in agg.values())
    print(f"{len(seq)} launches, {tot / 1e6:.3f} ms total (cold-cache, serialised: compare shares)")
    for n, a in sorted(agg.items(), key=lambda kv: -kv[1][1]):
        print(f"{n:66s} n={a[0]:3d} avg={a[1] / a[0] / 1e3:9.1f} us share={a[1] / tot * 100:5.1f}%")


def raw(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    H, U = rows[0], rows[1]
    ki = H.index('Kernel Name')
    for r in rows[2:]:
        print('==', r[ki].split('(')[0])
        for k in KEYS:
            if k in H:
                i = H.index(k)
                print(f"   {k:84s} {r[i]:>18s} {U[i]}")


def source(path, kernel, top=25):
    """hottest CUDA source lines of one kernel (needs -lineinfo and --import-source on)"""
    out = subprocess.run(['ncu', '-i', path, '--page', 'source', '--print-source', 'cuda,sass', '--csv',
                          '--kernel-name', 'regex:' + kernel], capture_output=True, text=True).stdout
    cur, H, acc = None, None, {}
    for r in csv.reader(out.splitlines()):
        if len(r) == 2 and r[0] == 'File Path':
            cur = r[1].split('/')[-1]
        elif len(r) > 8 and r[0] == 'Line No':
            H = r
        elif H and len(r) > 8 and r[2] == '-':
            try:
                smp, ins = float(r[H.index('# Samples')]), float(r[H.index('Instructions Executed')])
                thr = float(r[H.index('Thread Instructions Executed')])
            except ValueError:
                continue
            a = acc.setdefault((cur, r[0], r[1].strip()[:80]), [0.0, 0.0, 0.0])
            a[0] += smp
            a[1] += ins
            a[2] += thr
    ts, ti = sum(a[0] for a in acc.values()) or 1, sum(a[1] for a in acc.values()) or 1
    print(f"{kernel}: {ts:.0f} samples, {ti:.3g} warp instructions (summed over captured launches)")
    for k, a in sorted(acc.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{a[0] / ts * 100:5.1f}% smp {a[1] / ti * 100:5.1f}% inst {a[2] / max(a[1], 1):5.1f} thr/inst  {k[0]}:{k[1]}  {k[2]}")


if __name__ == '__main__':
    if sys.argv[1] == 'source':
        source(sys.argv[2], sys.argv[3], int(sys.argv[4]) if len(sys.argv) > 4 else 25)
    elif sys.argv[1] == 'launches':
        launches(sys.argv[2], int(sys.argv[3]) if len(sys.argv) > 3 else None)
    else:
        raw(sys.argv[2])
