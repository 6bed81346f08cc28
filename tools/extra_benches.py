#!/usr/bin/env python
"""The two BASELINE.json configurations bench.py does not print:
  configs[4]  micro-batch sweep of the ResNet-50 classify step (device-resident value + end to end)
  configs[3]  land-cover: 4-band 2048x2048 tiles cut into 256x256 chips (strided submit, no copy of
              the tile), ResNet-50 with a 4-channel stem and 8 classes, tiles/s and chips/s end to end
Writes one JSON object.   python tools/extra_benches.py [out.json]"""
import json
import os
import subprocess
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "aiforearth-api-platform_b200")]


def batch_sweep():
    rows = []
    for b in (1, 4, 16, 32, 64, 128, 256):
        r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--batch", str(b), "--steps", "60", "--warmup", "6",
                            "--no-cpu-baseline"], capture_output=True, text=True, timeout=600)
        if r.returncode != 0:
            rows.append({"batch": b, "error": r.stderr[-300:]})
            continue
        d = json.loads(r.stdout.strip().splitlines()[-1])
        rows.append({"batch": b, "images_per_s": d["value"], "ms_per_step": d["ms_per_step"], "e2e_images_per_s": d["e2e"]["value"],
                     "latency_ms": d["latency_ms"], "conv_tflops_whole_step": d["roofline"]["achieved_over_whole_step"]})
    return rows


def landcover():
    import numpy as np
    import torch  # noqa: F401
    from ai4e_b200.engine import Engine
    from ai4e_b200.landcover import classify_tile
    from ai4e_b200.models import build_resnet_program, synthetic_state_dict
    prog = build_resnet_program("resnet50", synthetic_state_dict("resnet50", seed=3, in_c=4, num_classes=8), 256, 256)
    rng = np.random.default_rng(7)
    tiles = [rng.integers(0, 256, (2048, 2048, 4), dtype=np.uint8) for _ in range(3)]
    out = {}
    with Engine(prog, device=0, max_batch=64, max_wait_us=2000) as eng:
        classify_tile(eng, tiles[0])  # warm-up
        t0 = time.perf_counter()
        n = 0
        for rep in range(4):
            for t in tiles:
                classify_tile(eng, t)
                n += 1
        dt = time.perf_counter() - t0
        out = {"tile": "2048x2048x4 uint8 -> 64 chips of 256x256", "tiles": n, "seconds": dt, "tiles_per_s": n / dt,
               "chips_per_s": 64 * n / dt, "model": "ResNet-50, 4-band stem, 8 classes, 256x256 input",
               "path": "landcover.classify_tile: strided per-chip submit -> micro-batcher -> wait (one tile at a time)"}
    return out


if __name__ == "__main__":
    res = {"batch_sweep": batch_sweep(), "landcover": landcover()}
    s = json.dumps(res, indent=1)
    if len(sys.argv) > 1:
        open(sys.argv[1], "w").write(s)
    print(s)
