# per-launch cycles / instructions of one bf16 step (ncu serialises and times each launch cold-cache)
# usage: bash tools/ncu_cycles.sh [tag]
ncu --metrics sm__cycles_elapsed.avg,smsp__inst_executed.sum,gpu__time_duration.sum,sm__icc_request_hit_rate.pct,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -k "regex:tc_(conv|du_pair)" --launch-count 40 --csv --log-file gpurun_out/cycles$1.csv python bench.py --steps 1 --warmup 0 --no-cpu > gpurun_out/ncu_cycles.log 2>&1
