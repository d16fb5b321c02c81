"""Turn gpurun_out ncu outputs into the tracked summaries under profiles/.
usage: python scripts/summarize_ncu.py <round-tag> <launches.csv> <prof.ncu-rep> [kernel-regex]"""
import csv, collections, json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag, launches, rep = sys.argv[1], sys.argv[2], sys.argv[3]
out = os.path.join(ROOT, 'profiles')
rows = [r for r in csv.reader(open(launches)) if len(r) > 5]
hdr = rows[0]; ki, vi, ui = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
for r in rows[1:]:
    v = float(r[vi].replace(',', '')); u = r[ui]
    ms = v / 1e6 if u.startswith('ns') else v / 1e3 if u.startswith('us') else v if u.startswith('ms') else v * 1e3
    a = agg[r[ki].split('(')[0]]; a[0] += 1; a[1] += ms; a[2] = max(a[2], ms)
tot = sum(a[1] for a in agg.values())
with open(os.path.join(out, tag + '_launches.md'), 'w') as f:
    f.write('# %s: ncu launch list of the exact-GP bench command (n=40000, d=8; the first 4000 launches = warm-up + timed + e2e evaluations)\n\n' % tag)
    f.write('`ncu --metrics gpu__time_duration.sum --clock-control none` on `python bench.py --steps 1 --warmup 1 --no-cpu --no-kissgp` (-c 4000); '
            'per-launch times are cold-cache and serialised: compare SHARES.\n\n| kernel | launches | total ms | share | max ms |\n|---|---:|---:|---:|---:|\n')
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        f.write('| %s | %d | %.2f | %.1f%% | %.3f |\n' % (k, a[0], a[1], 100 * a[1] / tot, a[2]))
    f.write('| **total** | %d | %.2f | | |\n' % (len(rows) - 1, tot))
raw = subprocess.check_output(['ncu', '-i', rep, '--page', 'raw', '--csv']).decode()
rr = list(csv.reader(raw.splitlines())); h = rr[0]; units = rr[1]
want = ['Kernel Name', 'Grid Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct', 'launch__registers_per_thread',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio']
idx = [h.index(w) for w in want if w in h]
with open(os.path.join(out, tag + '_dgemm_ncu_full.md'), 'w') as f:
    f.write('# %s: `ncu --set full` of dgemm_nt_kernel launches (same bench command)\n\n| metric | unit | ' % tag + ' | '.join('launch %d' % i for i in range(len(rr) - 2)) + ' |\n|---|---|' + '---|' * (len(rr) - 2) + '\n')
    for i in idx:
        f.write('| %s | %s | %s |\n' % (h[i], units[i], ' | '.join(r[i] for r in rr[2:])))
def col(name):
    i = h.index(name); mult = {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1}[units[i]]
    return [float(r[i].replace(',', '')) * mult for r in rr[2:]]
tr = [a + b for a, b in zip(col('dram__bytes_read.sum'), col('dram__bytes_write.sum'))]
json.dump({'dram_bytes_per_launch': sum(tr) / len(tr), 'launches_captured': len(tr), 'per_launch': tr,
           'note': 'mean over the captured launches (the first big SYRK launches of potrf plus the small panel updates between them)'},
          open(os.path.join(out, tag + '_dgemm_traffic.json'), 'w'), indent=1)
print(open(os.path.join(out, tag + '_launches.md')).read())
