#!/usr/bin/env python
"""BASELINE.json configs[4], the throughput half: micro-batch sweep 1..512 of the ResNet-50 classify step
(device-resident value + end to end) on the GPUs bench.py is launched on.  The latency half (open-loop
Poisson arrivals) is tools/openloop_bench.py; configs[3] is tools/landcover_bench.py.
Writes one JSON object.   python tools/extra_benches.py [out.json] [--gpus N] [--batches 1,8,64,512]"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "aiforearth-api-platform_b200")]


def batch_sweep():
    rows = []
    gpus = int(sys.argv[sys.argv.index("--gpus") + 1]) if "--gpus" in sys.argv else 1
    launcher = [sys.executable] if gpus == 1 else [sys.executable, "-m", "torch.distributed.run", "--nnodes=1",
                                                   f"--nproc-per-node={gpus}", "--master-addr", "127.0.0.1", "--master-port", "29511"]
    batches = (1, 2, 4, 8, 16, 32, 64, 128, 256, 512)
    if "--batches" in sys.argv:
        batches = tuple(int(v) for v in sys.argv[sys.argv.index("--batches") + 1].split(","))
    for b in batches:
        r = subprocess.run(launcher + [os.path.join(ROOT, "bench.py"), "--gpus", str(gpus), "--batch", str(b), "--steps", "60", "--warmup", "6",
                                       "--no-cpu-baseline", "--sustained-seconds", "0", "--async-seconds", "0"],
                           capture_output=True, text=True, timeout=900)
        if r.returncode != 0:
            rows.append({"batch": b, "error": r.stderr[-300:]})
            continue
        d = json.loads(r.stdout.strip().splitlines()[-1])
        rows.append({"batch": b, "gpus": gpus, "images_per_s": d["value"], "ms_per_step": d["ms_per_step"], "e2e_images_per_s": d["e2e"]["value"],
                     "latency_ms": d["latency_ms"], "conv_tflops_whole_step": d["roofline"]["achieved_over_whole_step"]})
    return rows


if __name__ == "__main__":
    res = {"batch_sweep": batch_sweep()}
    s = json.dumps(res, indent=1)
    if len(sys.argv) > 1 and not sys.argv[1].startswith("--"):
        open(sys.argv[1], "w").write(s)
    print(s)
