"""Summarise gpurun_out/cycles.csv (tools/ncu_cycles.sh): cycles, instructions, issue activity per launch."""
import csv, sys
rows = [r for r in csv.reader(open(sys.argv[1] if len(sys.argv) > 1 else 'gpurun_out/cycles.csv')) if len(r) > 10]
h = rows[0]; idx = {k: i for i, k in enumerate(h)}
d = {}
for r in rows[1:]:
    d.setdefault(int(r[idx['ID']]), {'name': r[idx['Kernel Name']]})[r[idx['Metric Name']]] = float(r[idx['Metric Value']].replace(',', ''))
tot = 0.0
for k in sorted(d):
    v = d[k]
    tot += v['sm__cycles_elapsed.avg']
    print("%2d %-28s %8.0f kcyc %7.1f Minst  issue %4.1f%%  icc %5.1f%%  %6.1f us" % (
        k, v['name'][22:50], v['sm__cycles_elapsed.avg'] / 1e3, v['smsp__inst_executed.sum'] / 1e6,
        v['smsp__issue_active.avg.pct_of_peak_sustained_active'], v['sm__icc_request_hit_rate.pct'], v['gpu__time_duration.sum'] / 1e3))
print("total %.0f kcycles" % (tot / 1e3))
