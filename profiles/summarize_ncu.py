#!/usr/bin/env python
"""Summarise ncu captures into the text files committed under profiles/.

    python profiles/summarize_ncu.py launches gpurun_out/launches_v5.csv          > profiles/r01_launches_v5.txt
    python profiles/summarize_ncu.py full gpurun_out/prof_pds_f32_v5.ncu-rep      > profiles/r01_ncu_pds_f32_v5.txt

`launches`: per-kernel totals of the `--metrics gpu__time_duration.sum` launch list (share of the step per kernel).
`full`: the metrics of a `--set full` capture that the roofline discussion in DESIGN.md uses.
"""
import csv
import io
import subprocess
import sys
from collections import defaultdict

KEYS = [
    'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
    'dram__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
    'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__throughput.avg.pct_of_peak_sustained_active',
    'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct',
    'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'launch__grid_size',
    'launch__block_size', 'launch__shared_mem_per_block_dynamic', 'launch__occupancy_limit_shared_mem',
    'launch__occupancy_limit_registers', 'launch__waves_per_multiprocessor',
    'smsp__issue_active.avg.pct_of_peak_sustained_active', 'smsp__inst_executed.sum',
    'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active',
    'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
    'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed',
    'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
    'smsp__cycles_active.avg', 'sm__cycles_elapsed.max',
]


def launches(path):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    hdr = rows[0]
    ik, iv = hdr.index('Kernel Name'), hdr.index('Metric Value')
    tot, cnt = defaultdict(float), defaultdict(int)
    for r in rows[1:]:
        name = r[ik].split('(')[0][-70:]
        tot[name] += float(r[iv])
        cnt[name] += 1
    total = sum(tot.values())
    print(f'# {path}: {len(rows) - 1} launches, {total / 1e3:.1f} us total (cold-cache, serialised: shares, not absolutes)')
    print(f'{"kernel":70s} {"launches":>8s} {"total us":>10s} {"avg us":>9s} {"share":>7s}')
    for name in sorted(tot, key=tot.get, reverse=True):
        print(f'{name:70s} {cnt[name]:8d} {tot[name] / 1e3:10.1f} {tot[name] / cnt[name] / 1e3:9.2f} '
              f'{100 * tot[name] / total:6.1f}%')


def full(path):
    out = subprocess.run(['ncu', '-i', path, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print(f'# {path}\nkernel: {r[hdr.index("Kernel Name")]}')
        for k in KEYS:
            if k in hdr:
                print(f'{k:85s} {r[hdr.index(k)]:>16s} {units[hdr.index(k)]}')
        stalls = [(float(r[i]), h) for i, h in enumerate(hdr)
                  if h.startswith('smsp__average_warps_issue_stalled') and h.endswith('_per_issue_active.ratio')
                  or h.startswith('smsp__average_warp_latency_issue_stalled') and h.endswith('.ratio')]
        for v, h in sorted(stalls, reverse=True)[:8]:
            print(f'{h:85s} {v:16.3f}')
        print()


if __name__ == '__main__':
    {'launches': launches, 'full': full}[sys.argv[1]](sys.argv[2])
