#!/usr/bin/env python
"""K0 throughput: (1) host entropy decode alone, per core and with T threads (the GIL is released in
the C call); (2) GPU reconstruction kernels for a batch of 64 records; (3) end to end: T threads
calling Engine.classify_jpeg on 224x224 4:2:0 JPEGs; with Pillow's decode timed beside it.
    python tools/jpeg_bench.py [out.json]"""
import io
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "aiforearth-api-platform_b200")]
import numpy as np  # noqa: E402
import torch  # noqa: E402
from PIL import Image  # noqa: E402

from ai4e_b200 import ops  # noqa: E402
from ai4e_b200.engine import Engine  # noqa: E402
from ai4e_b200.models import build_resnet_program, synthetic_state_dict  # noqa: E402


def make_jpegs(n, hw=224, quality=90):
    rng = np.random.default_rng(0)
    out = []
    for i in range(n):
        yy, xx = np.mgrid[0:hw, 0:hw]
        img = np.stack([(xx * (i % 5 + 1) + yy) % 256, (yy * 2 + 3 * i) % 256, (xx + yy * (i % 3 + 1)) % 256], -1)
        img = np.clip(img + rng.integers(-20, 21, img.shape), 0, 255).astype(np.uint8)
        buf = io.BytesIO()
        Image.fromarray(img).save(buf, format="JPEG", quality=quality, subsampling=2)
        out.append(buf.getvalue())
    return out


def threaded(fn, items, threads, reps):
    def work(k):
        for _ in range(reps):
            for it in items[k::threads]:
                fn(it)
    ts = [threading.Thread(target=work, args=(k,)) for k in range(threads)]
    t0 = time.perf_counter()
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    return reps * len(items) / (time.perf_counter() - t0)


def main():
    jp = make_jpegs(256)
    res = {"image": "224x224 RGB, quality 90, 4:2:0", "mean_bytes": float(np.mean([len(j) for j in jp])), "host_cores": os.cpu_count()}
    res["entropy_decode_img_s"] = {str(t): threaded(ops.jpeg_entropy_decode, jp, t, 4) for t in (1, 4, 16)}
    res["pillow_decode_img_s"] = {str(t): threaded(lambda d: np.asarray(Image.open(io.BytesIO(d)).convert("RGB")), jp, t, 4)
                                  for t in (1, 16)}
    info, _ = ops.jpeg_entropy_decode(jp[0])
    recs = torch.from_numpy(np.stack([ops.jpeg_entropy_decode(d)[1] for d in jp[:64]])).cuda()
    for _ in range(3):
        ops.jpeg_reconstruct(recs, info)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        ops.jpeg_reconstruct(recs, info)
    e1.record()
    torch.cuda.synchronize()
    res["gpu_reconstruct_us_per_batch64"] = e0.elapsed_time(e1) / 20 * 1e3
    prog = build_resnet_program("resnet50", synthetic_state_dict("resnet50", seed=0), 224, 224)
    with Engine(prog, device=0, max_batch=64, max_wait_us=2000) as eng:
        threaded(eng.classify_jpeg, jp, 64, 1)  # warm-up
        res["classify_jpeg_img_s"] = {str(t): threaded(eng.classify_jpeg, jp, t, 8) for t in (64, 256)}
        raw = [np.asarray(Image.open(io.BytesIO(d)).convert("RGB")) for d in jp]
        res["classify_raw_pixels_img_s"] = {str(t): threaded(eng.classify, raw, t, 8) for t in (64, 256)}
    s = json.dumps(res, indent=1)
    if len(sys.argv) > 1:
        open(sys.argv[1], "w").write(s)
    print(s)


if __name__ == "__main__":
    main()
