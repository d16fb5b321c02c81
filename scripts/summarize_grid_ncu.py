"""KISS-GP evidence for profiles/: launch list of one n=1M evaluation run and the ncu --set full metrics of the
persistent CG kernel.  usage: python scripts/summarize_grid_ncu.py <tag> <launches.csv> <raw.csv>"""
import csv, collections, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag, launches, raw = sys.argv[1:4]
rows = [r for r in csv.reader(open(launches)) if len(r) > 5]
hdr = rows[0]; ki, vi, ui = hdr.index('Kernel Name'), hdr.index('Metric Value'), hdr.index('Metric Unit')
agg = collections.defaultdict(lambda: [0, 0.0, 0.0])
for r in rows[1:]:
    v = float(r[vi].replace(',', '')); u = r[ui]
    ms = v / 1e6 if u.startswith('ns') else v / 1e3 if u.startswith('us') else v if u.startswith('ms') else v * 1e3
    a = agg[r[ki].split('(')[0]]; a[0] += 1; a[1] += ms; a[2] = max(a[2], ms)
tot = sum(a[1] for a in agg.values())
with open(os.path.join(ROOT, 'profiles', tag + '_kissgp_launches.md'), 'w') as f:
    f.write('# %s: ncu launch list of the KISS-GP run (n=1 000 000, 70x70 grid, covSM Q=6 x2)\n\n' % tag)
    f.write('`MAXIT=30000 ncu --metrics gpu__time_duration.sum --clock-control none python scripts/time_grid.py 1000000` '
            '(grid creation + sort once, then 3 evaluations = 3 launches of the persistent CG kernel, each a whole solve of ~5400 iterations).\n\n'
            '| kernel | launches | total ms | share | max ms |\n|---|---:|---:|---:|---:|\n')
    for k, a in sorted(agg.items(), key=lambda x: -x[1][1]):
        f.write('| %s | %d | %.2f | %.1f%% | %.3f |\n' % (k, a[0], a[1], 100 * a[1] / tot, a[2]))
    f.write('| **total** | %d | %.2f | | |\n' % (len(rows) - 1, tot))
rr = list(csv.reader(open(raw))); h, units, vals = rr[0], rr[1], rr[2]
want = ['Kernel Name', 'Grid Size', 'Block Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'launch__registers_per_thread', 'smsp__inst_executed.sum', 'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active',
        'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'lts__t_sector_hit_rate.pct', 'lts__throughput.avg.pct_of_peak_sustained_elapsed',
        'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_wait_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio',
        'smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio', 'smsp__average_warps_issue_stalled_membar_per_issue_active.ratio']
with open(os.path.join(ROOT, 'profiles', tag + '_cg_persistent2_ncu_full.md'), 'w') as f:
    f.write('# %s: `ncu --set full --clock-control none` of cg_persistent2_kernel (ONE launch = one whole CG solve at n=1M, ~5400 iterations)\n\n'
            '`scripts/ncu_grid.sh 1000000`.  DRAM bytes are for the WHOLE solve: divide by the iteration count for the per-iteration traffic '
            '(algorithmic bytes per iteration: 120.2 MB).\n\n| metric | unit | value |\n|---|---|---|\n' % tag)
    for w in want:
        if w in h:
            i = h.index(w); f.write('| %s | %s | %s |\n' % (w, units[i], vals[i]))
print(open(os.path.join(ROOT, 'profiles', tag + '_kissgp_launches.md')).read())
print(open(os.path.join(ROOT, 'profiles', tag + '_cg_persistent2_ncu_full.md')).read())
