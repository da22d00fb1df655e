#!/usr/bin/env python
"""Kernel durations of the last contacts step in an ncu launch list: last_step.py gpurun_out/<tag>_launches.csv"""
import csv, sys
rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')) if len(r) > 5]
h = rows[0]; k = h.index('Kernel Name'); v = h.index('Metric Value')
seq = [(r[k].split('(')[0].replace('arpd::', ''), int(r[v])) for r in rows[1:]]
for j in range(len(seq) - 1, -1, -1):
    if seq[j][0] == 'k_cell_assign' and seq[j][1] > 8000: break
step = seq[j:j + 9]
print(step, sum(t for _, t in step))
