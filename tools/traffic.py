"""Sums dram__bytes_read / dram__bytes_write over the tc_conv launches of one step (ncu csv with those two
metrics + gpu__time_duration) -> profiles/rNN_tc_conv_traffic.json.
usage: ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum --clock-control none -k regex:tc_conv
       --launch-count 46 --csv --log-file gpurun_out/traffic.csv python bench.py --steps 1 --warmup 0 --no-cpu
       python tools/traffic.py gpurun_out/traffic.csv profiles/r02_tc_conv_traffic.json"""
import csv
import json
import sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
h = rows[0]
idx = {k: i for i, k in enumerate(h)}
unit = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1.0, "msecond": 1.0,
        "usecond": 1e-3, "nsecond": 1e-6}
tot = {"dram__bytes_read.sum": 0.0, "dram__bytes_write.sum": 0.0, "gpu__time_duration.sum": 0.0}
ids = set()
for r in rows[1:]:
    m = r[idx["Metric Name"]]
    if m in tot:
        tot[m] += float(r[idx["Metric Value"]].replace(",", "")) * unit[r[idx["Metric Unit"]]]
        ids.add(r[idx["ID"]])
out = {"source": "%s (ncu, one forward step = %d tc_conv_kernel launches, 512 streams x 2^18)" % (sys.argv[1], len(ids)),
       "launches": len(ids), "dram_bytes_read": tot["dram__bytes_read.sum"], "dram_bytes_write": tot["dram__bytes_write.sum"],
       "dram_bytes_per_step": tot["dram__bytes_read.sum"] + tot["dram__bytes_write.sum"],
       "ncu_time_ms": tot["gpu__time_duration.sum"]}
json.dump(out, open(sys.argv[2], "w"), indent=1)
print(json.dumps(out))
