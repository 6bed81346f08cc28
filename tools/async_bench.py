#!/usr/bin/env python
"""BASELINE.json configs[2]: the async task path through the drop-in `APIService` with micro-batching,
one dispatcher (`Pool`) over N GPUs in ONE process, in-process task store, no network.

    python tools/async_bench.py --gpus 8 --seconds 8 --clients 64 --mode later

Every task goes the way the platform sends it (reference call stack): the task is created
in the store (APIM's request policy), the body is POSTed to the endpoint with a `taskId` header
(BackendQueueProcessor.cs:48-52) -> before_request -> AddTask -> worker thread -> task function ->
UpdateTaskStatus('running') -> engine -> CompleteTask.  Task latency = CompleteTask time - POST time.

Modes (what the task function does with the image):
  blocked     engine.classify(image): one worker thread blocked per task
  later       return engine.classify_later(image, then=CompleteTask): no blocked thread, one reaper
  inline      the same, registered with inline=True: the (non-blocking) task function runs on the request
              thread, so a task needs no thread hand-off on its way in
  shell_only  no engine at all (the task completes at once): the host shell's own ceiling in Python
  native      no shell: feeder threads hand batches to Pool.submit_many and one thread drains
              Pool.poll - what the C side of the dispatcher sustains behind the Python shell
`--images-per-task K` lets one task body carry K images (batch-style APIs): the per-task cost of the
Python shell is then paid once per K images.

The clients call the WSGI application directly with a prebuilt environ (what a WSGI server hands
over after parsing a request); `werkzeug.test.Client` costs more per request than the shell does.
"""
from __future__ import annotations

import argparse
import io
import json
import os
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "aiforearth-api-platform_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

HW, IN_C = 224, 3
PATH = "/v1/org/classify"


def _percentiles(lat_s):
    if not len(lat_s):
        return None, None
    a = np.sort(np.asarray(lat_s)) * 1e3
    return float(a[int((len(a) - 1) * 0.50)]), float(a[int((len(a) - 1) * 0.99)])


class _NullLog:
    tracer = None

    def log_exception(self, *a, **k):
        pass


def build_service(engine, mode, store, per_task=1):
    """the API project: an unmodified-signature async endpoint over the engine."""
    os.environ.setdefault("AI4E_B200_MAX_WORKERS", "2048")
    from ai4e_b200.compat.miniflask import Flask
    from ai4e_service import APIService
    from task_management.api_task import TaskManager
    app = Flask("async_bench")
    svc = APIService(app, _NullLog())
    tm = TaskManager(store=store)
    svc.api_task_manager = tm

    def read_image(request):
        a = np.frombuffer(request.get_data(), np.uint8)
        return {"image": a.reshape(HW, HW, IN_C)} if per_task == 1 else {"images": a.reshape(per_task, HW, HW, IN_C)}

    @svc.api_async_func(api_path=PATH, methods=["POST"], request_processing_function=read_image,
                        maximum_concurrent_requests=1 << 20, content_types=["application/octet-stream"],
                        content_max_length=1 << 26, trace_name="post:classify", inline=(mode == "inline"))
    def classify(*args, **kwargs):
        tid = kwargs["taskId"]
        tm.UpdateTaskStatus(tid, "running")
        if mode == "shell_only":
            tm.CompleteTask(tid, "completed - []")
        elif per_task > 1:  # a task that carries several images (camera-trap style batches): one C call, one completion
            from ai4e_b200.engine import Pending
            return Pending.gather(engine.classify_many_later(kwargs["images"]),
                                  then=lambda idx, prob: tm.CompleteTask(tid, "completed - " + json.dumps(idx[:, 0].tolist())))
        elif mode == "blocked":
            idx, prob = engine.classify(kwargs["image"])
            tm.CompleteTask(tid, "completed - " + json.dumps(idx.tolist()))
        else:
            return engine.classify_later(
                kwargs["image"], then=lambda idx, prob: tm.CompleteTask(tid, "completed - " + json.dumps(idx.tolist())))

    return app, svc, tm


def run_shell(engine, mode, seconds, clients, inflight, images, per_task=1):
    from task_management.local_task_store import LocalTaskStore
    store = LocalTaskStore()
    app, svc, tm = build_service(engine, mode, store, per_task)
    t_post, lat, lock = {}, [], threading.Lock()
    slots = threading.Semaphore(inflight)
    counted = [False]
    real_complete = tm.distributed_api_task.CompleteTask

    def complete(task_id, status):  # the moment the task is done, as the platform would see it
        r = real_complete(task_id, status)
        now = time.perf_counter()
        t0 = t_post.pop(task_id, None)
        if t0 is not None and counted[0]:
            with lock:
                lat.append(now - t0)
        slots.release()
        return r

    tm.distributed_api_task.CompleteTask = complete
    base_env = {"REQUEST_METHOD": "POST", "SCRIPT_NAME": "", "PATH_INFO": PATH, "QUERY_STRING": "",
                "SERVER_NAME": "container", "SERVER_PORT": "80", "SERVER_PROTOCOL": "HTTP/1.1",
                "wsgi.version": (1, 0), "wsgi.url_scheme": "http", "wsgi.errors": sys.stderr,
                "wsgi.multithread": True, "wsgi.multiprocess": False, "wsgi.run_once": False,
                "CONTENT_TYPE": "application/octet-stream", "HTTP_HOST": "container"}
    if per_task == 1:
        bodies = [im.tobytes() for im in images]
    else:
        bodies = [images[(np.arange(per_task) + k * per_task) % len(images)].tobytes() for k in range(max(1, 64 // per_task))]
    stop = [False]
    posted = [0] * clients
    bad = []

    def client(ci):
        def start_response(status, headers, exc_info=None):
            if not status.startswith("200"):
                bad.append(status)
        k = ci
        while not stop[0]:
            if not slots.acquire(timeout=0.05):
                continue
            task = store.create_task("http://container" + PATH)
            tid = task["TaskId"]
            body = bodies[k % len(bodies)]
            k += clients
            env = dict(base_env)
            env["wsgi.input"] = io.BytesIO(body)
            env["CONTENT_LENGTH"] = str(len(body))
            env["HTTP_TASKID"] = tid
            t_post[tid] = time.perf_counter()
            for _ in app(env, start_response):
                pass
            posted[ci] += 1

    threads = [threading.Thread(target=client, args=(i,), daemon=True) for i in range(clients)]
    for t in threads:
        t.start()
    time.sleep(min(2.0, seconds * 0.25))  # warm-up: worker threads, autotuned engine, caches
    counted[0] = True
    with lock:
        lat.clear()
    t0 = time.perf_counter()
    time.sleep(seconds)
    with lock:
        n_done = len(lat)
        done_lat = list(lat)
    dt = time.perf_counter() - t0
    stop[0] = True
    for t in threads:
        t.join(timeout=2)
    deadline = time.time() + 10
    while t_post and time.time() < deadline:
        time.sleep(0.01)
    p50, p99 = _percentiles(done_lat)
    return {"tasks_per_s": n_done / dt, "images_per_s": n_done * per_task / dt, "images_per_task": per_task,
            "tasks": n_done, "seconds": dt, "p50_ms": p50, "p99_ms": p99,
            "latency_of": "POST with taskId header -> CompleteTask written to the store", "rejected": len(bad),
            "store_upserts": store.upserts, "store_gets": store.gets}


def run_native(engine, seconds, feeders, batch, images):
    """no shell: `feeders` threads submit_many(batch) with notify, one thread drains poll()."""
    done, lat, lock = [0], [], threading.Lock()
    t_sub = {}
    stop = [False]
    window = threading.Semaphore(feeders * 4)
    stack = np.ascontiguousarray(images[:batch]) if len(images) >= batch else np.ascontiguousarray(
        images[np.arange(batch) % len(images)])

    def feeder(fi):
        seq = fi << 40
        while not stop[0]:
            if not window.acquire(timeout=0.05):
                continue
            tags = np.arange(seq, seq + batch, dtype=np.uint64)
            seq += batch
            t_sub[int(tags[-1])] = time.perf_counter()
            engine.submit_many(stack, tags=tags, notify=True)

    def reaper():
        while not stop[0] or t_sub:
            tk, tg, st, idx, prob = engine.poll(4096, 0.05)
            if not len(tk):
                if stop[0]:
                    break
                continue
            now = time.perf_counter()
            with lock:
                done[0] += len(tk)
                for t in tg.tolist():
                    t0 = t_sub.pop(t, None)
                    if t0 is not None:
                        lat.append(now - t0)
                        window.release()

    th = [threading.Thread(target=feeder, args=(i,), daemon=True) for i in range(feeders)]
    rp = threading.Thread(target=reaper, daemon=True)
    rp.start()
    for t in th:
        t.start()
    time.sleep(min(2.0, seconds * 0.25))
    with lock:
        d0 = done[0]
        lat.clear()
    t0 = time.perf_counter()
    time.sleep(seconds)
    with lock:
        n = done[0] - d0
        l = list(lat)
    dt = time.perf_counter() - t0
    stop[0] = True
    for t in th:
        t.join(timeout=2)
    rp.join(timeout=5)
    p50, p99 = _percentiles(l)
    return {"tasks_per_s": n / dt, "images_per_s": n / dt, "images_per_task": 1, "tasks": n, "seconds": dt,
            "p50_ms": p50, "p99_ms": p99,
            "latency_of": f"submit_many({batch}) -> last image of the call announced by poll"}


def measure(gpus: int, mode: str, seconds: float, clients: int, inflight: int | None = None, max_batch: int = 64,
            max_wait_us: int = 2000, devices=None, arch: str = "resnet50", per_task: int = 1) -> dict:
    from ai4e_b200.engine import Pool
    from ai4e_b200.models import build_resnet_program, synthetic_state_dict
    rng = np.random.default_rng(1234)
    images = rng.integers(0, 256, (256, HW, HW, IN_C), dtype=np.uint8)  # 256 distinct images, cycled
    out = {"mode": mode, "gpus": gpus, "clients": clients, "max_batch": max_batch, "max_wait_us": max_wait_us,
           "process": "one Python process, one Pool (one replica per GPU), in-process task store"}
    if mode == "shell_only":
        inflight = inflight or 4096
        out.update(run_shell(None, mode, seconds, clients, inflight, images))
        out["inflight_limit"] = inflight
        return out
    prog = build_resnet_program(arch, synthetic_state_dict(arch, seed=0), HW, HW)
    devs = list(devices) if devices else list(range(gpus))
    with Pool(prog, devices=devs, max_batch=max_batch, max_wait_us=max_wait_us) as pool:
        if mode == "native":
            out.update(run_native(pool, seconds, clients, max_batch, images))
        else:
            inflight = inflight or max(2 * max_batch * gpus // per_task, 2 * gpus)
            out.update(run_shell(pool, mode, seconds, clients, inflight, images, per_task))
            out["inflight_limit"] = inflight
        st = pool.stats()
        out["batch_hist"] = np.sum([s["batch_hist"] for s in st], axis=0).tolist()
        out["images_per_replica"] = [s["images"] for s in st]
        out["engine_lat_ms_p50_p99"] = [[round(s["lat_ms_p50"], 3), round(s["lat_ms_p99"], 3)] for s in st]
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--seconds", type=float, default=6.0)
    ap.add_argument("--clients", type=int, default=64)
    ap.add_argument("--inflight", type=int, default=None)
    ap.add_argument("--mode", default="later", choices=["blocked", "later", "inline", "shell_only", "native"])
    ap.add_argument("--max-batch", type=int, default=64)
    ap.add_argument("--max-wait-us", type=int, default=2000)
    ap.add_argument("--images-per-task", type=int, default=1,
                    help="images carried by one task body (shell modes); 1 = BASELINE's one image per task")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    r = measure(a.gpus, a.mode, a.seconds, a.clients, a.inflight, a.max_batch, a.max_wait_us, per_task=a.images_per_task)
    r["cores"] = os.cpu_count()
    print(json.dumps(r), flush=True)
    if a.out:
        with open(a.out, "a") as fh:
            fh.write(json.dumps(r) + "\n")


if __name__ == "__main__":
    main()
