"""Summarise an .ncu-rep (run where ncu is installed): key metrics of the last captured launch and
the SASS instructions holding >= 1.5 % of the warp-stall samples.

    python profiles/ncu_summary.py gpurun_out/prof.ncu-rep
"""
import csv
import subprocess
import sys

rep = sys.argv[1]
KEYS = ('gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem',
        'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active')
raw = subprocess.run(['ncu', '-i', rep, '--page', 'raw', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h, units, r = rows[0], rows[1], rows[-1]
print('kernel:', r[h.index('Kernel Name')][:90])
for k in KEYS:
    if k in h:
        i = h.index(k)
        print(f'{k:70s} {r[i]} {units[i]}')
src = subprocess.run(['ncu', '-i', rep, '--page', 'source', '--csv'], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
his = [i for i, x in enumerate(rows) if x and x[0] == 'Address']
if his:
    hi = his[-1]
    h = rows[hi]
    si, so = h.index('# Samples'), h.index('Source')
    data = []
    for x in rows[hi + 1:]:
        if len(x) > si:
            try:
                data.append((int(float(x[si])), x[so].strip()))
            except ValueError:
                pass
    tot = sum(d[0] for d in data) or 1
    print('stall samples:', tot, 'instructions:', len(data))
    for k, (n, s) in enumerate(data):
        if n > tot * 0.015:
            print(f'{k:5d} {n:6d} {100 * n / tot:5.1f}%  {s}')
