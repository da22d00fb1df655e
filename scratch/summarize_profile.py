#!/usr/bin/env python
"""profiles/ from one round's captures: summarize_profile.py <tag>   (reads gpurun_out/<tag>_full.ncu-rep, <tag>_launches.csv)"""
import csv, json, subprocess, sys, os, shutil
tag = sys.argv[1]
rep = f"gpurun_out/{tag}_full.ncu-rep"
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h, u = rows[0], rows[1]
col = h.index
want = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'smsp__inst_executed.sum', 'sm__inst_executed.avg.per_cycle_elapsed',
        'smsp__issue_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active', 'sm__warps_active.avg.pct_of_peak_sustained_active',
        'launch__registers_per_thread', 'launch__grid_size', 'launch__block_size', 'l1tex__t_sector_hit_rate.pct', 'lts__t_sector_hit_rate.pct',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__thread_inst_executed_per_inst_executed.ratio',
        # SURVEY 8(d)'s list: DRAM throughput, shared-memory wavefronts, FP64 pipe, L1/TEX throughput
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'l1tex__data_pipe_lsu_wavefronts_mem_shared.sum',
        'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active', 'smsp__inst_executed_pipe_fp64.sum', 'l1tex__throughput.avg.pct_of_peak_sustained_elapsed']
want = [w for w in want if w in h]
def tobytes(v, unit): return float(v.replace(',', '')) * {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}[unit]
out, kern, seen = [], {}, {}
for r in rows[2:]:
    name = r[col('Kernel Name')].split('(')[0]
    seen[name] = seen.get(name, 0) + 1
    out.append(name + ('' if seen[name] == 1 else ' (ring grid)'))
    for w in want: out.append(f'    {w:75s} {r[col(w)]:>16s} {u[col(w)]}')
    if seen[name] == 1:
        kern[name] = {'dram_read_bytes': tobytes(r[col('dram__bytes_read.sum')], u[col('dram__bytes_read.sum')]),
                      'dram_write_bytes': tobytes(r[col('dram__bytes_write.sum')], u[col('dram__bytes_write.sum')]),
                      'duration_us_under_ncu': float(r[col('gpu__time_duration.sum')])}
open(f'profiles/{tag}_kernels_summary.txt', 'w').write(
    'ncu --set full --clock-control none --import-source on -k regex:^k_ --launch-skip 48 --launch-count 9 python bench.py --no-cpu --no-counting --steps 2 --warmup 3\n'
    '(config 3, 1 022 500 atoms; per-launch values; caches flushed between replays, so durations are cold-cache)\n\n' + '\n'.join(out) + '\n')
pair = {k: v for k, v in kern.items() if k in ('k_screen_atoms', 'k_screen', 'k_filter', 'k_exact')}
other = {k: v for k, v in kern.items() if k not in pair}
prev = json.load(open('profiles/pair_stage_traffic.json')) if os.path.exists('profiles/pair_stage_traffic.json') else {}
cm = {k: v for k, v in {**prev.get('kernels', {}), **prev.get('counting_mode_kernels', {})}.items() if k in ('k_screen', 'k_filter', 'k_exact') and k not in pair or k == 'k_exact' and 'k_screen_atoms' in pair}
json.dump({'counting_mode_kernels': cm, 'counting_mode_source': prev.get('counting_mode_source', prev.get('source', '')), 'source': f'profiles/{tag}_kernels_summary.txt (ncu --set full, one launch of each kernel of a config-3 step)', 'kernels': pair,
           'other_kernels_of_the_step': other}, open('profiles/pair_stage_traffic.json', 'w'), indent=1)
shutil.copy(f'gpurun_out/{tag}_launches.csv', f'profiles/{tag}_launches.csv')
os.makedirs('/tmp/prof', exist_ok=True)
subprocess.run('cd /tmp/prof && rm -f *.cubin && cuobjdump -xelf all /root/repo/arpeggia_b200/_build/libarpeggia_b200.so >/dev/null && nvdisasm -g -c api.sm_100a.cubin > dis.txt 2>/dev/null', shell=True)
for k in ('k_screen_atoms', 'k_screen', 'k_filter', 'k_exact', 'k_rings', 'k_scatter_atoms'):
    if k not in kern: continue
    src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "-k", f"regex:^{k}$"], capture_output=True, text=True).stdout
    open(f'/tmp/prof/src_{k}.csv', 'w').write(src)
    t = subprocess.run(["python", "scratch/ncu_lines.py", f"/tmp/prof/src_{k}.csv", "/tmp/prof/dis.txt", k, "25"], capture_output=True, text=True).stdout
    open(f'profiles/{tag}_{k}_by_line.txt', 'w').write('\n'.join(l[:200] for l in t.splitlines()) + '\n')
print(sum(v['dram_read_bytes'] + v['dram_write_bytes'] for v in pair.values()) / 1e6, 'MB pair-stage DRAM traffic')
for k, v in kern.items(): print(k, v)
