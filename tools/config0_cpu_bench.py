#!/usr/bin/env python
"""BASELINE.json configs[0]: ResNet-18 through the ai4e SYNC API, batch 1, synthetic 224x224 RGB, on the
CPU - the reference's own CPU-runnable case (last paragraph).

    python tools/config0_cpu_bench.py --shell reference   # the UNMODIFIED reference APIService under stubs
    python tools/config0_cpu_bench.py --shell dropin      # this repo's ai4e_service, same decorators
    python tools/config0_cpu_bench.py --shell none        # bare model call, no host shell

One request = POST of a raw 224x224x3 image to a sync endpoint whose task function is what a PyTorch
CPU container runs inside the callback (reference call site ai4e_service.py:203): preprocess ->
torchvision resnet18 fp32 eager -> softmax -> top-5 -> JSON.  Timed with all host cores and with ONE
thread (torch.set_num_threads); 10 warm-up requests, then >= --seconds of requests; the median request
time is reported next to the mean rate.  Inputs: SURVEY 8(d) row 1 (rng(1234), 256 distinct images,
cycled).  `--shell reference` needs /root/reference (this container; it does not exist on the GPU
box): the reference file is imported from where it lies through the stub loader the host-shell
trace uses (tests/golden/make_host_shell_trace.py).
"""
import argparse
import io
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "aiforearth-api-platform_b200")]

import numpy as np  # noqa: E402


class _Log:
    tracer = None

    def log_exception(self, *a, **k):
        pass


def build(shell):
    import torch  # noqa: F401
    from oracle import resnet_oracle as ro
    model = ro.build_model("resnet18")

    def classify_image(img):
        idx, prob, _ = ro.classify_fp32(model, img[None])
        return json.dumps({"top5": idx[0].tolist(), "prob": [round(float(p), 6) for p in prob[0]]})

    if shell == "none":
        return None, classify_image
    if shell == "reference":
        sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
        import make_host_shell_trace as mk
        mod, Flask = mk.load_reference_service()
        APIService = mod.APIService
    else:
        from ai4e_b200.compat.miniflask import Flask
        from ai4e_service import APIService
    app = Flask("config0")
    svc = APIService(app, _Log())

    def read_image(request):
        return {"image": np.frombuffer(request.get_data(), np.uint8).reshape(224, 224, 3)}

    @svc.api_sync_func(api_path="/v1/org/classify", methods=["POST"], request_processing_function=read_image,
                       maximum_concurrent_requests=16, content_types=["application/octet-stream"],
                       content_max_length=1 << 20, trace_name="post:classify")
    def classify(*args, **kwargs):
        return classify_image(kwargs["image"])

    return app, classify_image


def run(app, classify_image, images, seconds, warmup=10):
    env0 = {"REQUEST_METHOD": "POST", "SCRIPT_NAME": "", "PATH_INFO": "/v1/org/classify", "QUERY_STRING": "",
            "SERVER_NAME": "container", "SERVER_PORT": "80", "SERVER_PROTOCOL": "HTTP/1.1", "wsgi.version": (1, 0),
            "wsgi.url_scheme": "http", "wsgi.errors": sys.stderr, "wsgi.multithread": False, "wsgi.multiprocess": False,
            "wsgi.run_once": False, "CONTENT_TYPE": "application/octet-stream", "HTTP_HOST": "container"}
    bodies = [im.tobytes() for im in images]
    status = []

    def one(i):
        if app is None:
            return classify_image(images[i % len(images)])
        env = dict(env0)
        body = bodies[i % len(bodies)]
        env["wsgi.input"] = io.BytesIO(body)
        env["CONTENT_LENGTH"] = str(len(body))
        return b"".join(app(env, lambda s, h, e=None: status.append(s)))

    for i in range(warmup):
        first = one(i)
    times = []
    t0 = time.perf_counter()
    i = warmup
    while time.perf_counter() - t0 < seconds or len(times) < 20:
        t = time.perf_counter()
        one(i)
        times.append(time.perf_counter() - t)
        i += 1
    dt = time.perf_counter() - t0
    assert all(s.startswith("200") for s in status), set(status)
    return {"requests": len(times), "seconds": dt, "images_per_s": len(times) / dt,
            "median_ms": float(np.median(times) * 1e3), "p99_ms": float(np.sort(times)[int((len(times) - 1) * 0.99)] * 1e3),
            "first_response": (first.decode() if isinstance(first, bytes) else first)[:80]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--shell", default="dropin", choices=["reference", "dropin", "none"])
    ap.add_argument("--seconds", type=float, default=15.0)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    import torch
    images = np.random.default_rng(1234).integers(0, 256, (256, 224, 224, 3), dtype=np.uint8)
    app, fn = build(a.shell)
    cores = os.cpu_count() or 1
    res = {"config": "ResNet-18 via ai4e sync API, batch 1, synthetic 224x224 RGB, CPU (torchvision fp32 eager, preprocessing + top-5 included)",
           "shell": {"reference": "unmodified /root/reference/APIs/1.0/base-py/ai4e_service.py under sys.modules stubs",
                     "dropin": "aiforearth-api-platform_b200/ai4e_service.py", "none": "no host shell (bare task function)"}[a.shell],
           "host_cores": cores, "torch": torch.__version__}
    for label, threads in (("all_cores", cores), ("one_thread", 1)):
        torch.set_num_threads(threads)
        r = run(app, fn, images, a.seconds)
        r["threads"] = threads
        res[label] = r
    print(json.dumps(res), flush=True)
    if a.out:
        with open(a.out, "a") as fh:
            fh.write(json.dumps(res) + "\n")


if __name__ == "__main__":
    main()
