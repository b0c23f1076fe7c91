#!/usr/bin/env python
"""One markdown table row per kernel launch of an .ncu-rep (raw page).  usage: python scripts/ncu_table.py file.ncu-rep"""
import csv, io, re, subprocess, sys
out = subprocess.run(['ncu', '-i', sys.argv[1], '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr = rows[0]
ix = {h: i for i, h in enumerate(hdr)}
cols = [("us", 'gpu__time_duration.sum', 1.0), ("grid", 'launch__grid_size', 1), ("blk", 'launch__block_size', 1), ("regs", 'launch__registers_per_thread', 1),
        ("warps act %", 'sm__warps_active.avg.pct_of_peak_sustained_active', 1), ("issue act %", 'smsp__issue_active.avg.pct_of_peak_sustained_active', 1),
        ("SM thr %", 'sm__throughput.avg.pct_of_peak_sustained_elapsed', 1), ("L1TEX thr %", 'l1tex__throughput.avg.pct_of_peak_sustained_active', 1),
        ("L2 thr %", 'lts__throughput.avg.pct_of_peak_sustained_elapsed', 1), ("DRAM thr %", 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 1),
        ("tensor pipe %", 'sm__pipe_tensor_subpipe_hmma_cycles_active.avg.pct_of_peak_sustained_active', 1),
        ("tensor inst", 'sm__inst_executed_pipe_tensor.sum', 1),
        ("dram rd MB", 'dram__bytes_read.sum', 1), ("dram wr MB", 'dram__bytes_write.sum', 1), ("warp inst", 'smsp__inst_executed.sum', 1), ("lanes/inst", 'smsp__thread_inst_executed_per_inst_executed.ratio', 1)]
units = rows[1]
print("| kernel | " + " | ".join(c[0] for c in cols) + " |")
print("|---|" + "---:|" * len(cols))
for r in rows[2:]:
    name = re.sub(r'\(.*', '', r[ix['Kernel Name']]).replace('ngp::', '').replace('void ', '')
    vals = []
    for label, key, _ in cols:
        if key not in ix:
            vals.append("n/a"); continue
        v, u = r[ix[key]], units[ix[key]]
        try:
            f = float(v.replace(',', ''))
        except ValueError:
            vals.append(v); continue
        if key == 'gpu__time_duration.sum':
            f = f / 1000 if u in ('ns', 'nsecond') else f * 1000 if u in ('ms', 'msecond') else f
        if 'bytes' in key:
            f = f / 1e6 if u == 'byte' else f * 1e-3 if u == 'Kbyte' else f if u == 'Mbyte' else f * 1e3 if u == 'Gbyte' else f
        vals.append(f"{f:.1f}" if abs(f) < 1e5 and f != int(f) else f"{int(f):,}".replace(',', ' '))
    print(f"| `{name}` | " + " | ".join(vals) + " |")
