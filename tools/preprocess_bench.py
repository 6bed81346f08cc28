#!/usr/bin/env python
"""K1 stand-alone (ai4e_preprocess through the C ABI): achieved HBM GB/s at B = 64 and B = 512
(SURVEY 7.2 asks for both, honestly), for the NCHW ABI layout and the engine's NHWC4 layout.

    python tools/preprocess_bench.py [--out file.json]
    ncu --metrics gpu__time_duration.sum --clock-control none --csv python tools/preprocess_bench.py --once

Timing by CUDA events over back-to-back launches (B = 512: 77 MB in, 154 / 205 MB out, larger than
L2); `--once` launches each case once more after a warm-up for an ncu pass (gpu__time_duration is the
figure to quote: a 10-40 us kernel timed by events carries its launch gap).
Algorithmic bytes per image: 150,528 read + 301,056 (NCHW bf16) or 401,408 (NHWC4 bf16) written.
"""
import argparse
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "aiforearth-api-platform_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402

from ai4e_b200 import ops  # noqa: E402

MEAN, STD = (0.485, 0.456, 0.406), (0.229, 0.224, 0.225)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--once", action="store_true")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    peak = 6542.1
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        peak = json.load(open(p))["hbm_gbs"]
    rows = []
    for b in (64, 512):
        x = torch.from_numpy(np.random.default_rng(b).integers(0, 256, (b, 224, 224, 3), dtype=np.uint8)).cuda()
        for layout in ("nchw", "nhwc"):
            byts = b * (224 * 224 * 3 + 224 * 224 * (3 if layout == "nchw" else 4) * 2)
            ops.preprocess(x, MEAN, STD, layout=layout)
            torch.cuda.synchronize()
            if a.once:
                ops.preprocess(x, MEAN, STD, layout=layout)
                torch.cuda.synchronize()
                continue
            reps = 50
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(reps):
                ops.preprocess(x, MEAN, STD, layout=layout)  # the caching allocator hands back the same output block
            e1.record()
            torch.cuda.synchronize()
            us = e0.elapsed_time(e1) * 1e3 / reps
            rows.append({"batch": b, "layout": layout, "bytes": byts, "us_events_back_to_back": us, "gbs": byts / us / 1e3,
                         "frac_of_hbm_peak": byts / us / 1e3 / peak, "hbm_peak_gbs": peak})
            print(json.dumps(rows[-1]), flush=True)
    if a.out and rows:
        json.dump(rows, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
