"""Per-layer-class roofline table of DESIGN.md section 3.1 from profiles/rNN_bf16_cycles.txt (tools/ncu_cycles.sh +
tools/cycles.py: the 40 tc launches of one C2 step, 512 streams x 2^18 samples) and MEASURED_PEAKS.json.
usage: python tools/roofline_table.py [profiles/r02_bf16_cycles.txt]"""
import json
import os
import re
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
us = {}
for l in open(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "profiles", "r02_bf16_cycles.txt")):
    m = re.match(r"\s*(\d+)\s.*?([\d.]+) us\s*$", l)
    if m:
        us[int(m.group(1))] = float(m.group(2))
pk = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
PK, HB, B = pk["bf16_tflops_sustained"] * 1e12 / 2, pk["hbm_gbs"], 512


def mac(rows, cin, cout, k):
    return B * rows * cin * cout * k


def gb(ch, rows):
    return B * rows * ch * 2 / 1e9


CLS = [
    ("PQMF analysis", [0], mac(16384, 1, 16, 513), gb(1, 262144) + gb(16, 16384)),
    ("first conv 16->64 k7", [1], mac(16384, 16, 64, 7), gb(16, 16384) + gb(64, 16384)),
    ("DilatedUnit C=64 (6)", [2, 3, 4, 35, 36, 37], 6 * mac(16384, 64, 64, 4), 12 * gb(64, 16384)),
    ("64->128 k8 s4", [5], mac(4096, 64, 128, 8), gb(64, 16384) + gb(128, 4096)),
    ("DilatedUnit C=128 (6)", [6, 7, 8, 31, 32, 33], 6 * mac(4096, 128, 128, 4), 12 * gb(128, 4096)),
    ("128->256 k8 s4", [9], mac(1024, 128, 256, 8), gb(128, 4096) + gb(256, 1024)),
    ("DilatedUnit C=256, pair kernel (6)", [10, 11, 12, 27, 28, 29], 6 * mac(1024, 256, 256, 4), 12 * gb(256, 1024)),
    ("256->512 k4 s2", [13], mac(512, 256, 512, 4), gb(256, 1024) + gb(512, 512)),
    ("conv3 C=512 (4)", [14, 16, 22, 24], 4 * mac(512, 512, 512, 3), 8 * gb(512, 512)),
    ("conv1 + res C=512 (4)", [15, 17, 23, 25], 4 * mac(512, 512, 512, 1), 12 * gb(512, 512)),
    ("512->1024 k4 s2", [18], mac(256, 512, 1024, 4), gb(512, 512) + gb(1024, 256)),
    ("1024->64 k3", [19], mac(256, 1024, 64, 3), gb(1024, 256) + gb(64, 256)),
    ("320->1024 k3 (speaker channels folded: K = 64)", [20], mac(256, 320, 1024, 3), gb(64, 256) + gb(1024, 256)),
    ("1024->512 x2 transposed", [21], mac(512, 1024, 512, 2), gb(1024, 256) + gb(512, 512)),
    ("512->256 x2 transposed", [26], mac(1024, 512, 256, 2), gb(512, 512) + gb(256, 1024)),
    ("256->128 x4 transposed", [30], mac(4096, 256, 128, 2), gb(256, 1024) + gb(128, 4096)),
    ("128->64 x4 transposed", [34], mac(16384, 128, 64, 2), gb(128, 4096) + gb(64, 16384)),
    ("last conv 64->32 k7 + head", [38], mac(16384, 64, 32, 7), gb(64, 16384) + gb(16, 16384)),
    ("PQMF synthesis", [39], mac(16384, 16, 16, 33), gb(16, 16384) + gb(1, 262144)),
]
print("| layer class (launches) | GFLOP | GB | bound | floor µs | ncu µs | frac |\n|---|---|---|---|---|---|---|")
tm = tf = 0.0
for name, ids, m, g in CLS:
    t = sum(us[i] for i in ids)
    ft, fh = m / PK * 1e6, g / HB * 1e6
    fl = max(ft, fh)
    tm += t
    tf += fl
    print("| %s | %.0f | %.2f | %s | %.0f | %.0f | %.2f |" % (name, 2 * m / 1e9, g, "tensor" if ft > fh else "HBM", fl, t, fl / t))
print("| **step (40 launches)** | 10 684 | 37.6 | | **%.0f** | **%.0f** | **%.2f** |" % (tf, tm, tf / tm))
