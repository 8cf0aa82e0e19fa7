# One GPU box: the whole -m gpu suite, smoke(), the default bench (both arms), the per-launch cycle list, the DRAM
# traffic of one step, the other configurations.  Everything lands in gpurun_out/final_*; copy what is to be
# judged into profiles/.   usage (through gpurun): bash tools/final_validation.sh
set -x
timeout -k 5 900 python -m pytest tests -m gpu -x -q --timeout 120 > gpurun_out/final_pytest.log 2>&1; echo "pytest rc=$?"
tail -n 2 gpurun_out/final_pytest.log
timeout -k 5 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/final_smoke.log 2>&1; echo "smoke rc=$?"
timeout -k 5 300 python bench.py > gpurun_out/final_bench.json 2> gpurun_out/final_bench.err; echo "bench rc=$?"
timeout -k 5 600 python bench.py --impl reference > gpurun_out/final_bench_reference.json 2> gpurun_out/final_bench_reference.err; echo "reference rc=$?"
timeout -k 5 300 bash tools/ncu_cycles.sh _final; python tools/cycles.py gpurun_out/cycles_final.csv > gpurun_out/final_cycles.txt
timeout -k 5 300 ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k "regex:tc_(conv|du_pair)" --launch-count 40 --csv --log-file gpurun_out/final_traffic.csv python bench.py --steps 1 --warmup 0 --no-cpu > gpurun_out/final_traffic.log 2>&1
python tools/traffic.py gpurun_out/final_traffic.csv gpurun_out/final_tc_conv_traffic.json
timeout -k 5 900 python bench_configs.py > gpurun_out/final_configs.jsonl 2> gpurun_out/final_configs.err; echo "configs rc=$?"
cut -c1-300 gpurun_out/final_bench.json; cut -c1-300 gpurun_out/final_bench_reference.json; tail -n 1 gpurun_out/final_cycles.txt
