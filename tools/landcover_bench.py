#!/usr/bin/env python
"""configs[3]: 4-band 2048x2048 tiles -> 64 chips of 256x256 -> ResNet-50 (4-channel stem, 8 classes)."""
import sys, time, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "aiforearth-api-platform_b200")]
import numpy as np
from ai4e_b200.engine import Engine
from ai4e_b200.landcover import classify_tile
from ai4e_b200.models import build_resnet_program, synthetic_state_dict
prog = build_resnet_program("resnet50", synthetic_state_dict("resnet50", seed=3, in_c=4, num_classes=8), 256, 256)
rng = np.random.default_rng(7)
tiles = [rng.integers(0, 256, (2048, 2048, 4), dtype=np.uint8) for _ in range(3)]
with Engine(prog, device=0, max_batch=64, max_wait_us=2000) as eng:
    classify_tile(eng, tiles[0])
    t0 = time.perf_counter(); n = 0
    for rep in range(8):
        for t in tiles:
            classify_tile(eng, t); n += 1
    dt = time.perf_counter() - t0
    print("one tile at a time: %.0f tiles/s, %.0f chips/s" % (n / dt, 64 * n / dt))
    t0 = time.perf_counter(); n = 0
    for rep in range(8):
        tk = [eng.submit_tile(t) for t in tiles]
        eng.flush()
        for k in tk:
            eng.wait_many(k); n += 1
    dt = time.perf_counter() - t0
    print("three tiles in flight: %.0f tiles/s, %.0f chips/s" % (n / dt, 64 * n / dt))
