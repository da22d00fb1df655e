#!/usr/bin/env python
"""Join an ncu SASS source page (csv) with nvdisasm -g line info: instruction and stall-sample share per source line.
usage: ncu_lines.py <src.csv> <dis.txt> <kernel mangled name> [top]"""
import csv, re, sys, collections
src, dis, kname = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40
# address -> (file, line) from nvdisasm
addr2line = {}
cur = None; inside = False
for ln in open(dis):
    if ln.startswith('//---') and '.text.' in ln:
        inside = kname in ln
    if not inside: continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m: cur = (m.group(1).split('/')[-1], int(m.group(2))); continue
    m = re.match(r'\s*/\*([0-9a-f]{4,})\*/\s+(\S.*?);', ln)
    if m: addr2line[int(m.group(1), 16)] = (cur, m.group(2))
rows = list(csv.reader(open(src)))
h = next(i for i, r in enumerate(rows) if 'Instructions Executed' in r)
hdr = rows[h]; ia = hdr.index('Address'); ie = hdr.index('Instructions Executed'); isamp = hdr.index('# Samples')
ith = hdr.index('Thread Instructions Executed')
base = None
agg = collections.defaultdict(lambda: [0, 0, 0])
tot = [0, 0, 0]
for r in rows[h + 1:]:
    try:
        a = int(r[ia], 16) if r[ia].startswith('0x') else int(r[ia])
        n = int(r[ie].replace(',', '')); s = int(r[isamp].replace(',', '')); t = int(r[ith].replace(',', ''))
    except Exception:
        continue
    if base is None: base = a
    key = addr2line.get(a - base, (None, ''))[0]
    agg[key][0] += n; agg[key][1] += s; agg[key][2] += t
    tot[0] += n; tot[1] += s; tot[2] += t
print(f"warp instructions {tot[0]:,}  thread instr {tot[2]:,}  avg active {tot[2]/max(tot[0],1):.1f}  samples {tot[1]:,}")
srcs = {}
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    f, l = k if k else ('?', 0)
    text = ''
    try:
        if f not in srcs:
            import glob
            p = glob.glob(f'/root/repo/arpeggia_b200/csrc/cuda/{f}')
            srcs[f] = open(p[0]).read().split('\n') if p else []
        text = srcs[f][l - 1].strip()[:90] if srcs[f] else ''
    except Exception:
        pass
    print(f"{100*v[0]/tot[0]:5.1f}% inst {100*v[1]/max(tot[1],1):5.1f}% smp  act {v[2]/max(v[0],1):4.1f}  {f}:{l:<4} {text}")

# ---- by region (pair_kernel.cuh line ranges given as name:lo-hi on argv[5:]) ----
if len(sys.argv) > 5:
    regs = []
    for spec in sys.argv[5:]:
        name, rng = spec.split(':'); lo, hi = map(int, rng.split('-')); regs.append((name, lo, hi))
    out = collections.defaultdict(lambda: [0, 0, 0])
    for k, v in agg.items():
        f, l = k if k else ('?', 0)
        name = f
        if f == 'pair_kernel.cuh':
            name = next((n for n, lo, hi in regs if lo <= l <= hi), 'pair_kernel:other')
        for j in range(3): out[name][j] += v[j]
    print('--- regions')
    for k, v in sorted(out.items(), key=lambda kv: -kv[1][0]):
        print(f"{100*v[0]/tot[0]:5.1f}% inst {100*v[1]/max(tot[1],1):5.1f}% smp  act {v[2]/max(v[0],1):4.1f}  {v[0]/1e6:7.2f}M  {k}")
