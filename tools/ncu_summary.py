#!/usr/bin/env python
"""Summarise an `ncu -i <report> --page raw --csv` export of one timed step into a markdown table
and the JSON bench.py reads for `roofline.traffic`.

    ncu -i gpurun_out/step.ncu-rep --page raw --csv > gpurun_out/step.csv
    python tools/ncu_summary.py gpurun_out/step.csv --key r18-bf16-b16 \
        --md profiles/r02_ncu_launches.md --json profiles/r02_ncu_launches.json --source "..."
"""
import argparse
import csv
import json
import os
import re


def num(v):
    try:
        return float(str(v).replace(',', ''))
    except ValueError:
        return None


def find(cols, *needles):
    for c in cols:
        if all(n in c for n in needles):
            return c
    return None


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument('csv')
    ap.add_argument('--key', required=True)
    ap.add_argument('--md', required=True)
    ap.add_argument('--json', required=True)
    ap.add_argument('--source', default='')
    args = ap.parse_args()
    rows = list(csv.reader(open(args.csv, newline='')))
    # skip ncu's "==PROF==" preamble: the header is the first row that has a "Kernel Name" column
    hi = next(i for i, r in enumerate(rows) if 'Kernel Name' in r)
    header, units, data = rows[hi], rows[hi + 1], rows[hi + 2:]
    col = {name: i for i, name in enumerate(header)}
    cols = list(col)

    def get(row, name, scale_from_unit=True):
        if name is None:
            return None
        v = num(row[col[name]])
        if v is None:
            return None
        u = units[col[name]].lower()
        if scale_from_unit:
            mult = {'kbyte': 1e3, 'mbyte': 1e6, 'gbyte': 1e9, 'byte': 1.0, 'ns': 1e-3, 'us': 1.0,
                    'usecond': 1.0, 'nsecond': 1e-3, 'msecond': 1e3, 'ms': 1e3}.get(u)
            if mult is not None:
                v *= mult
        return v

    c_time = find(cols, 'gpu__time_duration.sum')
    c_rd = find(cols, 'dram__bytes_read.sum') if 'dram__bytes_read.sum' in col else find(cols, 'dram__bytes_read.sum')
    c_wr = find(cols, 'dram__bytes_write.sum')
    c_dram = find(cols, 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed')
    c_tensor = find(cols, 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active') or \
        find(cols, 'sm__pipe_tensor', 'pct')
    c_regs = find(cols, 'launch__registers_per_thread')
    c_grid = find(cols, 'launch__grid_size')
    c_l2hit = find(cols, 'lts__t_sector_hit_rate.pct')
    out, total_us, conv_us, conv_bytes, conv_n = [], 0.0, 0.0, 0.0, 0
    for r in data:
        if len(r) < len(header):
            continue
        name = re.sub(r'\(.*$', '', r[col['Kernel Name']])
        name = re.sub(r'^void |rnb::|<unnamed>::|\(anonymous namespace\)::', '', name).replace('unnamed>::', '')
        us = get(r, c_time) or 0.0
        rd, wr = get(r, c_rd) or 0.0, get(r, c_wr) or 0.0
        is_conv = name.startswith('conv_')
        total_us += us
        if is_conv:
            conv_us += us
            conv_bytes += rd + wr
            conv_n += 1
        out.append({'kernel': name, 'grid': get(r, c_grid, False), 'us': us, 'dram_read_mb': rd / 1e6,
                    'dram_write_mb': wr / 1e6, 'dram_pct': get(r, c_dram, False),
                    'tensor_pipe_pct': get(r, c_tensor, False), 'l2_hit_pct': get(r, c_l2hit, False),
                    'regs': get(r, c_regs, False)})
    with open(args.md, 'w') as f:
        f.write('# ncu launch list of one timed step (%s)\n\n%s\n\nDurations under ncu are cold-cache '
                'and serialised: shares, not absolutes.\n\n' % (args.key, args.source))
        f.write('| # | kernel | grid | us | DRAM rd MB | DRAM wr MB | DRAM % | tensor pipe % | L2 hit % | regs |\n')
        f.write('|---|---|---|---|---|---|---|---|---|---|\n')

        def fmt(v, p='%.1f'):
            return '-' if v is None else p % v
        for i, k in enumerate(out):
            f.write('| %d | `%s` | %s | %.1f | %.1f | %.1f | %s | %s | %s | %s |\n' % (
                i, k['kernel'], fmt(k['grid'], '%d'), k['us'], k['dram_read_mb'], k['dram_write_mb'],
                fmt(k['dram_pct']), fmt(k['tensor_pipe_pct']), fmt(k['l2_hit_pct']), fmt(k['regs'], '%d')))
        f.write('\nSum: %.1f us over %d launches; convolution kernels %d launches, %.1f us = %.3f of the '
                'step; DRAM traffic of the convolution launches %.1f MB = %.1f MB per launch on average.\n'
                % (total_us, len(out), conv_n, conv_us, conv_us / total_us if total_us else 0,
                   conv_bytes / 1e6, conv_bytes / 1e6 / max(conv_n, 1)))
    cap = {}
    if os.path.exists(args.json):
        cap = json.load(open(args.json))
    cap[args.key] = {'dram_bytes_per_conv_launch': conv_bytes / max(conv_n, 1),
                     'dram_bytes_per_step': sum((k['dram_read_mb'] + k['dram_write_mb']) * 1e6 for k in out),
                     'conv_launches': conv_n, 'conv_share_of_step': conv_us / total_us if total_us else None,
                     'source': args.source or os.path.basename(args.md), 'launches': out}
    json.dump(cap, open(args.json, 'w'), indent=1)
    print('wrote %s and %s (%d launches)' % (args.md, args.json, len(out)))


if __name__ == '__main__':
    main()
