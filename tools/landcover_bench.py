#!/usr/bin/env python
"""BASELINE.json configs[3] at its stated size: a 4-band 4096x4096 uint8 tile (SURVEY 8(d):
`rng(4321).integers(0, 256, (4096, 4096, 4))`) -> 256 chips of 256x256x4 -> ResNet-50 with a
4-channel stem per chip.  The tile crosses PCIe once per replica band and K1 reads the chips in place
(K6); with N GPUs one Pool deals contiguous ranges of 256/N chips (32 each on 8).

    python tools/landcover_bench.py --gpus 1 [--seconds 5] [--tile 4096] [--out file.jsonl]

Prints one JSON line per (source memory, tiles in flight) combination: tiles/s, chips/s and the
fraction of the conv-bound chip rate (the classify rate of bench.py scaled by 8.174 / 10.78 GFLOP).
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "aiforearth-api-platform_b200")]

import numpy as np  # noqa: E402


def run(pool, tiles, seconds, in_flight, chips_per_tile, summary=False):
    """summary: tile -> chips -> classify -> aggregate (K7), one ticket and one small record per tile"""
    if summary:
        submit = lambda t: pool.submit_tile_summary(t, 256)
        wait = lambda h: pool.wait_tile_summary(*h)
    else:
        submit = lambda t: pool.submit_tile(t, 256)
        wait = pool.wait_many
    pending, n = [], 0
    for t in tiles[:2]:  # warm-up (tile buffers are allocated on first use)
        wait(submit(t))
    t0 = time.perf_counter()
    i = 0
    while time.perf_counter() - t0 < seconds:
        pending.append(submit(tiles[i % len(tiles)]))
        i += 1
        while len(pending) >= in_flight:
            wait(pending.pop(0))
            n += 1
    while pending:
        wait(pending.pop(0))
        n += 1
    dt = time.perf_counter() - t0
    return {"tiles": n, "seconds": dt, "tiles_per_s": n / dt, "chips_per_s": chips_per_tile * n / dt}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--seconds", type=float, default=5.0)
    ap.add_argument("--tile", type=int, default=4096)
    ap.add_argument("--classify-rate", type=float, default=None, help="images/s of bench.py on one GPU, for the conv-bound fraction")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    import torch
    from ai4e_b200.engine import Pool
    from ai4e_b200.models import build_resnet_program, synthetic_state_dict
    prog = build_resnet_program("resnet50", synthetic_state_dict("resnet50", seed=3, in_c=4, num_classes=8), 256, 256)
    rng = np.random.default_rng(4321)
    tiles = [rng.integers(0, 256, (a.tile, a.tile, 4), dtype=np.uint8) for _ in range(3)]
    pinned = [torch.from_numpy(t).pin_memory().numpy() for t in tiles]
    chips = (a.tile // 256) ** 2
    rows = []
    with Pool(prog, devices=list(range(a.gpus)), max_batch=64, max_wait_us=2000) as pool:
        # parity in passing: the pool's per-chip classes equal a plain batch forward of the copied chips
        eng0 = pool.engines()[0]
        t_idx, _ = pool.wait_many(pool.submit_tile(tiles[0], 256))
        copied = np.ascontiguousarray(tiles[0].reshape(a.tile // 256, 256, a.tile // 256, 256, 4).transpose(0, 2, 1, 3, 4)).reshape(chips, 256, 256, 4)
        want = np.concatenate([eng0.forward(copied[i:i + 64])[0] for i in range(0, chips, 64)])
        same = bool(np.array_equal(t_idx, want))
        for mem, src in (("pageable", tiles), ("page-locked", pinned)):
            for k in (1, 3):
                r = run(pool, src, a.seconds, k, chips)
                r.update({"config": f"{a.tile}x{a.tile}x4 uint8 tile -> {chips} chips of 256x256x4 -> ResNet-50 (4-band stem, 8 classes)",
                          "gpus": a.gpus, "source_memory": mem, "tiles_in_flight": k, "chips_per_gpu_per_tile": chips // a.gpus,
                          "tile_path_equals_copied_chips": same,
                          "path": "Pool.submit_tile: one H2D of each replica's band, chips read in place by K1 (ai4e_preprocess_tile)"})
                if a.classify_rate:
                    bound = a.classify_rate * 8.174 / 10.78 * a.gpus
                    r["conv_bound_chips_per_s"] = bound
                    r["frac_of_conv_bound"] = r["chips_per_s"] / bound
                rows.append(r)
                print(json.dumps(r), flush=True)
        # K7: the same tiles with the aggregate computed on the GPUs; parity in passing against the per-chip results
        t_idx, t_prob = pool.wait_many(pool.submit_tile(tiles[0], 256))
        got = pool.wait_tile_summary(*pool.submit_tile_summary(tiles[0], 256))
        q = np.rint(t_prob.astype(np.float64) * 4294967296.0).astype(np.uint64)
        mass = np.zeros(prog.num_classes, np.uint64)
        np.add.at(mass, t_idx.ravel(), q.ravel())
        same_sum = bool(np.array_equal(got["labels"].ravel(), t_idx[:, 0]) and np.array_equal(got["mass_q32"], mass)
                        and np.array_equal(got["hist"], np.bincount(t_idx[:, 0], minlength=prog.num_classes)))
        r = run(pool, pinned, a.seconds, 3, chips, summary=True)
        r.update({"config": f"{a.tile}x{a.tile}x4 uint8 tile -> {chips} chips -> ResNet-50 -> tile summary (label map, histogram, mass)",
                  "gpus": a.gpus, "source_memory": "page-locked", "tiles_in_flight": 3, "summary_equals_aggregate_of_chip_results": same_sum,
                  "path": "Pool.submit_tile_summary: K6 + forward + K7 aggregate on each replica, one record per tile"})
        rows.append(r)
        print(json.dumps(r), flush=True)
    if a.out:
        with open(a.out, "a") as fh:
            for r in rows:
                fh.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()
