#!/usr/bin/env python
"""Per-kernel table of the LAST ApplyDeferred frame of an ncu launch list (gpu__time_duration + dram bytes)."""
import csv
import sys
from collections import defaultdict

rows = list(csv.reader(l for l in open(sys.argv[1]) if not l.startswith('==')))
hdr, rows = rows[0], rows[1:]
ki, mi, vi, ii = hdr.index('Kernel Name'), hdr.index('Metric Name'), hdr.index('Metric Value'), hdr.index('ID')
d = defaultdict(dict)
for r in rows:
    d[int(r[ii])][r[mi]] = float(r[vi].replace(',', ''))
    d[int(r[ii])]['name'] = r[ki]
ids = sorted(d)
first_name, last_name = sys.argv[2], sys.argv[3]
last = [i for i in ids if last_name in d[i]['name']][-1]
first = [i for i in ids if first_name in d[i]['name'] and i < last][-1]
tot = 0
print("| kernel | us | DRAM read MB | DRAM write MB |\n|---|---:|---:|---:|")
for i in range(first, last + 1):
    x = d[i]
    t = x['gpu__time_duration.sum'] / 1e3
    tot += t
    print(f"| `{x['name'][:70]}` | {t:.1f} | {x.get('dram__bytes_read.sum', 0) / 1e6:.1f} | {x.get('dram__bytes_write.sum', 0) / 1e6:.1f} |")
print(f"| total | {tot:.1f} | | |")
