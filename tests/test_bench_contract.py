"""bench.py's contract with the driver, checked on the CPU through the reference arm (the only arm that runs
without a GPU): exactly ONE line on stdout, that line is the JSON record with the keys the driver reads, and
whatever else the run prints goes to stderr."""
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_reference_arm_prints_one_json_line():
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--steps", "1",
                        "--warmup", "0", "--streams", "2", "--samples", "16384"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stderr[-2000:]
    lines = [l for l in r.stdout.split("\n") if l.strip()]
    assert len(lines) == 1, r.stdout[:2000]
    rec = json.loads(lines[0])
    assert rec["impl"] == "reference" and rec["higher_is_better"] is True
    assert rec["metric"] == "audio-sec/sec v2 encode+decode" and rec["unit"] == "audio-sec/sec"
    assert rec["value"] > 0 and rec["steps"] == 1 and rec["n_gpus"] == 1
    assert rec["cpu_baseline"]["kind"] == "port" and rec["cpu_baseline"]["value"] == rec["value"]
    assert rec["e2e"] == {"value": rec["value"], "unit": "audio-sec/sec", "h2d_bytes_per_step": 0,
                          "d2h_bytes_per_step": 0}
    assert rec["gpu_launches"] == 0


def test_reference_arm_other_ranks_exit_quietly():
    """Under torchrun only rank 0 runs the CPU baseline; the other ranks print nothing and exit 0."""
    env = dict(os.environ, RANK="1", LOCAL_RANK="1", WORLD_SIZE="2")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                        "--steps", "1", "--warmup", "0", "--streams", "2", "--samples", "16384"],
                       capture_output=True, text=True, timeout=600, cwd=ROOT, env=env)
    assert r.returncode == 0, r.stderr[-2000:]
    assert r.stdout.strip() == ""
