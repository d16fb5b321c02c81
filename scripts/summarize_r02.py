"""Turn the round-2 gpurun outputs (launch list csv, ncu raw csv exports) into the tracked summaries under profiles/.
usage: python scripts/summarize_r02.py <gpurun_out/r02 dir>"""
import csv, collections, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = sys.argv[1]; out = os.path.join(ROOT, 'profiles'); tag = 'r02'
rows = [r for r in csv.reader(open(os.path.join(src, 'launches_exact.csv'))) if len(r) > 5]
hdr = rows[0]; ki, vi, ui = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
for r in rows[1:]:
    v = float(r[vi].replace(',', '')); u = r[ui]
    ms = v / 1e6 if u.startswith('ns') else v / 1e3 if u.startswith('us') else v if u.startswith('ms') else v * 1e3
    nm = r[ki].split('(')[0].replace('void ', '').replace('kgp::', '')
    a = agg[nm]; a[0] += 1; a[1] += ms; a[2] = max(a[2], ms)
tot = sum(a[1] for a in agg.values())
with open(os.path.join(out, tag + '_launches.md'), 'w') as f:
    f.write('# %s: ncu launch list of two exact-GP evaluations + predict (n=40000, d=8, covSEard)\n\n' % tag)
    f.write('`ncu --metrics gpu__time_duration.sum --clock-control none -c 4000` on `python scripts/time_exact.py 40000` (the first 4000 launches: '
            'two evaluations of nlZ + dnlZ + dnlZ/dX and the start of a 1000-point prediction); per-launch times are cold-cache and serialised: '
            'compare SHARES.\n\n| kernel | launches | total ms | share | max ms |\n|---|---:|---:|---:|---:|\n')
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        f.write('| %s | %d | %.2f | %.1f%% | %.3f |\n' % (k, a[0], a[1], 100 * a[1] / tot, a[2]))
    f.write('| **total** | %d | %.2f | | |\n' % (len(rows) - 1, tot))
rr = list(csv.reader(open(os.path.join(src, 'dgemm_full_raw.csv')))); h = rr[0]; units = rr[1]
want = ['Kernel Name', 'Grid Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_tensor_subpipe_dmma.avg.pct_of_peak_sustained_active',
        'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed', 'lts__t_sector_hit_rate.pct', 'launch__registers_per_thread',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'smsp__inst_executed.sum',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio']
idx = [h.index(w) for w in want if w in h]
with open(os.path.join(out, tag + '_dgemm_ncu_full.md'), 'w') as f:
    f.write('# %s: `ncu --set full --clock-control none` of dgemm_nt_tma_kernel launches (`python scripts/time_exact.py 40000`, -k regex:dgemm_nt_tma -s 3 -c 6)\n\n'
            'Launch 1 is the first big trailing SYRK of the Cholesky (one CTA per tile: the look-ahead form); the others are K = 128 panel updates.\n\n'
            '| metric | unit | ' % tag + ' | '.join('launch %d' % i for i in range(len(rr) - 2)) + ' |\n|---|---|' + '---|' * (len(rr) - 2) + '\n')
    for i in idx:
        f.write('| %s | %s | %s |\n' % (h[i], units[i], ' | '.join(r[i].replace('kgp::', '') for r in rr[2:])))
def col(name):
    i = h.index(name); mult = {'Gbyte': 1e9, 'Mbyte': 1e6, 'Kbyte': 1e3, 'byte': 1}[units[i]]
    return [float(r[i].replace(',', '')) * mult for r in rr[2:]]
tr = [a + b for a, b in zip(col('dram__bytes_read.sum'), col('dram__bytes_write.sum'))]
big = max(range(len(tr)), key=lambda i: tr[i])
gi = h.index('Grid Size'); tiles = int(rr[2 + big][gi].strip('() ').split(',')[0])
alg = tiles * 128 * 128 * 8 * 2 + 40064 * 512 * 8          # C tiles read + written once, the 512-wide panel read once
json.dump({'dram_bytes_per_launch': tr[big], 'algorithmic_bytes_per_launch': alg, 'ratio': tr[big] / alg, 'tiles': tiles,
           'launch': 'first trailing SYRK of the blocked Cholesky at n = 40 000 (K = 512, C read-modify-write of the trailing lower triangle)',
           'all_captured_launches_bytes': tr,
           'note': 'per-launch figure of the big trailing SYRK (not a mean over unlike launches): dram__bytes_read + dram__bytes_write from one ncu --set full '
                   'capture; algorithmic = every C tile read and written once + the panel read once'},
          open(os.path.join(out, tag + '_dgemm_traffic.json'), 'w'), indent=1)
print(open(os.path.join(out, tag + '_launches.md')).read()); print(open(os.path.join(out, tag + '_dgemm_traffic.json')).read())
