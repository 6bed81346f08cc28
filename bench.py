#!/usr/bin/env python
"""Benchmark of the hot path: images/sec end-to-end (preprocess + classify + top-5).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--batch B] [--impl b200|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Workload (BASELINE.json configs[1]): ResNet-50 bf16 classification, micro-batch 64, synthetic
224x224 RGB, one replica per B200.  A "step" is one micro-batch of 64 images through
preprocess -> 53 convolutions -> head.  Weights are the CPU oracle's seeded synthetic ResNet-50
(oracle.resnet_oracle.build_model: the same network the parity tests and goldens use); images are
SURVEY 8(d)'s uniform-noise generator, except batch 0 of the rotation, which holds the 64 golden
images of tests/golden/classify_resnet50_b64.npz: after each timed loop its outputs are compared
with the fp32 oracle's top-5 (`parity_checked`), so a kernel that went wrong at this batch size
cannot post a number.

Printed keys (one JSON line, rank 0):
  value         images/s with the uint8 batches already resident in HBM (ai4e_engine_forward_device),
                two micro-batches in flight on two streams (the engine's lanes), CUDA events
                bracketing both streams, max over ranks
  e2e           the same metric through the public API with HOST buffers: page-locked uint8
                batches -> Engine.submit_batch / wait (H2D + kernels + D2H of top-5 inside the
                timed region, three batches in flight)
  roofline      dominant kernel = the tcgen05 implicit-GEMM convolution (all 53 convolutions):
                algorithmic conv FLOPs of one step / summed device time of its launches (per-layer
                CUDA events, ai4e_engine_profile: each launch timed ALONE, so the overlap the
                engine gets from programmatic dependent launch and from the second lane is NOT
                credited), against the measured BURST bf16 peak in MEASURED_PEAKS.json (the figure
                for a kernel timed alone); `frac_over_whole_step` divides the same FLOPs by the whole
                pipelined step and by the SUSTAINED peak (a kernel inside a long step); `traffic` =
                DRAM bytes of those launches from the ncu capture recorded in profiles/
  sustained     the `value` loop and the `e2e` loop again, each for >= --sustained-seconds (5 s), with
                the SM clock, power and throttle reasons sampled over that window: the warmed-up
                state a service runs in, next to the cold-burst figures of the K timed steps
  async_path    (N = 1) BASELINE.json configs[2] in small: tasks through the drop-in APIService
                async endpoint, in-process task store, p50/p99 AddTask -> CompleteTask
                (tools/async_bench.py; `--gpus N` there runs one Pool over N GPUs in one process)
  open_loop     (N = 1) p50 / p99 TASK latency under Poisson arrivals at 50 % and 80 % of async_path.native
                (BASELINE.json configs[4]; native load generator, tools/openloop_bench.py --native runs it on N GPUs)
  cpu_baseline  the CPU oracle (torchvision fp32, preprocessing included) on the host cores,
                bounded sample, rank 0 at N=1 only
`--impl reference` times that CPU path alone (the reference's own arithmetic lives in a
user-supplied PyTorch image; the oracle restates it).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
PKG = os.path.join(ROOT, "aiforearth-api-platform_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

ARCH, HW, IN_C = "resnet50", 224, 3
POOL_BATCHES = 16  # distinct input batches rotated through: 16 x 64 x 150,528 B = 154 MB > 126 MB of L2


def measured_peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        d = json.load(open(path))
        return {"hbm_gbs": d["hbm_gbs"], "bf16_tflops": d["bf16_tflops"],
                "bf16_tflops_sustained": d.get("bf16_tflops_sustained", d["bf16_tflops"]), "source": "measured"}
    return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0, "bf16_tflops_sustained": 1400.0, "source": "fallback"}


class ClockSampler:
    """nvidia-smi clock/throttle samples DURING the timed region (B200_PROFILING.md recipe)."""
    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap,power.draw")

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-i", str(index), "-lms", "100"], stdout=subprocess.PIPE, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), line.strip()))

    def window(self, t0: float, t1: float) -> dict:
        """clocks over [t0, t1] without stopping the sampler"""
        time.sleep(0.15)
        return self._summarise(t0, t1)

    def stop(self, t0: float, t1: float) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        return self._summarise(t0, t1)

    def _summarise(self, t0: float, t1: float) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm, mx, reasons, power = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, line in self.rows:
            f = [x.strip() for x in line.split(",")]
            if len(f) < 6 or not (t0 - 0.05 <= ts <= t1 + 0.15):
                continue
            try:
                sm.append(float(f[0])); mx.append(float(f[1]))
            except ValueError:
                continue
            for n, v in zip(names, f[2:6]):
                if v.lower().startswith("active"):
                    reasons.add(n)
            try:
                power.append(float(f[6]))
            except (ValueError, IndexError):
                pass
        if not sm:  # region shorter than one sample: fall back to everything seen
            for ts, line in self.rows:
                f = [x.strip() for x in line.split(",")]
                try:
                    sm.append(float(f[0])); mx.append(float(f[1]))
                except (ValueError, IndexError):
                    pass
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm), "power_w_max": max(power) if power else None}


def workload_config(batch: int, replicas: int) -> dict:
    """the `config` object both arms print (the reference arm runs the same workload on the CPU)."""
    return {"workload": f"ResNet-50 bf16 classification, micro-batch {batch}, synthetic 224x224 RGB, one replica per GPU",
            "batch": batch, "replicas": replicas, "in_flight": os.environ.get("AI4E_B200_LANES", "2") + " micro-batches per replica, one stream each (engine lanes)", "parallelism": f"{replicas} independent replica(s), no collective",
            "l2": f"{POOL_BATCHES} distinct input batches rotated ({POOL_BATCHES * batch * HW * HW * IN_C / 1e6:.0f} MB > 126 MB L2); "
                  f"activations of one step ({batch * 43.6:.0f} MB) exceed L2 as well"}


def noise_batches(n_batches: int, batch: int, seed: int = 1234) -> np.ndarray:
    rng = np.random.default_rng(seed)
    return rng.integers(0, 256, (n_batches, batch, HW, HW, IN_C), dtype=np.uint8)


def cpu_oracle_rate(batch: int, budget_s: float, max_batches: int, threads: int):
    """images/s of the CPU oracle (preprocess + torchvision fp32 forward + softmax/top-5)."""
    import torch
    from oracle import resnet_oracle as ro
    torch.set_num_threads(threads)
    model = ro.build_model(ARCH)
    imgs = noise_batches(2, batch, seed=4321)
    ro.classify_fp32(model, imgs[0][: min(batch, 8)])  # warm-up
    n, t0 = 0, time.perf_counter()
    while n < max_batches and (time.perf_counter() - t0 < budget_s or n == 0):
        ro.classify_fp32(model, imgs[n % 2])
        n += 1
    dt = time.perf_counter() - t0
    return batch * n / dt, n, dt


def run_reference(args, rank: int):
    if rank != 0:
        return
    threads = os.cpu_count() or 1
    import torch
    from oracle import resnet_oracle as ro
    torch.set_num_threads(threads)
    model = ro.build_model(ARCH)
    imgs = noise_batches(2, args.batch, seed=4321)
    for i in range(args.warmup):
        ro.classify_fp32(model, imgs[i % 2])
    t0 = time.perf_counter()
    for i in range(args.steps):
        ro.classify_fp32(model, imgs[i % 2])
    dt = time.perf_counter() - t0
    v = args.batch * args.steps / dt
    sample = f"{args.steps} steps x {args.batch} images, torchvision {ARCH} fp32 eager on CPU, preprocessing + top-5 included"
    print(json.dumps({
        "impl": "reference", "metric": "images/sec end-to-end (preprocess+classify)", "value": v, "unit": "images/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.batch, max(args.gpus, 1)),
        "cpu_baseline": {"value": v, "unit": "images/s", "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


def golden_batch(batch: int):
    """the golden images of tests/golden/classify_resnet50_b64.npz, cycled to `batch`, with the fp32
    oracle's top-5 for them (the checker of the timed loops; never timed itself)."""
    from oracle import resnet_oracle as ro
    g = np.load(os.path.join(ROOT, "tests", "golden", "classify_resnet50_b64.npz"))
    imgs = ro.synthetic_images([int(i) for i in g["indices"]], HW, HW, IN_C, seed=int(g["image_seed"]))
    pick = np.arange(batch) % len(imgs)
    return imgs[pick], g["top5_idx"][pick], g["top5_prob"][pick]


def parity_of(idx, prob, want_idx, want_prob, what: str) -> dict:
    same = bool(np.array_equal(idx, want_idx))
    close = bool(np.allclose(prob, want_prob, atol=2e-4, rtol=3e-2))
    if not (same and close):
        raise SystemExit(f"bench.py: {what}: outputs of the timed loop differ from the fp32 oracle "
                         f"(top-5 identical: {same}, probabilities close: {close}) - the number would be meaningless")
    return {"top5_identical": same, "prob_close": close}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=60)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--sustained-seconds", type=float, default=5.0, help="0 skips the sustained block")
    ap.add_argument("--async-seconds", type=float, default=4.0, help="0 skips the async_path block (N = 1 only)")
    ap.add_argument("--dump-layers", default=None, help="write per-layer device times (JSON) to this path")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))

    if args.impl == "reference":
        run_reference(args, rank)
        return

    import torch
    from ai4e_b200.engine import Engine
    from ai4e_b200.models import build_resnet_program
    from oracle import resnet_oracle as ro  # the checker's model: the engine runs the network the goldens were minted on
    from tools.replica_group import ReplicaGroup

    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the engine has no CPU fallback")
    torch.cuda.set_device(local_rank)
    group = ReplicaGroup("nccl")  # measurement plumbing only: barrier + max of the per-rank times
    barrier, max_over_ranks = group.barrier, group.max_over_ranks

    B, K, W = args.batch, args.steps, args.warmup
    prog = build_resnet_program(ARCH, ro.build_model(ARCH).state_dict(), HW, HW)
    eng = Engine(prog, device=local_rank, max_batch=B, max_wait_us=2000)
    peaks = measured_peaks()

    pool_np = noise_batches(POOL_BATCHES, B, seed=1234 + rank)
    gold_imgs, gold_idx, gold_prob = golden_batch(B)
    pool_np[0] = gold_imgs  # batch 0 of the rotation is checked against the fp32 oracle after each loop
    host = torch.from_numpy(pool_np).pin_memory()
    dev = host.cuda(non_blocking=True)
    # The engine keeps two sets of activation slots ("lanes"): two micro-batches are in flight, each
    # on its own stream, so that the second fills the SMs the first leaves idle in its small layers.
    N_INFLIGHT = max(1, min(4, int(os.environ.get("AI4E_B200_LANES", "2"))))  # same switch the engine reads
    d_idx = [torch.empty((B, 5), dtype=torch.int32, device="cuda") for _ in range(N_INFLIGHT)]
    d_prob = [torch.empty((B, 5), dtype=torch.float32, device="cuda") for _ in range(N_INFLIGHT)]
    g_idx = torch.zeros((B, 5), dtype=torch.int32, device="cuda")    # outputs of the golden batch, kept apart
    g_prob = torch.zeros((B, 5), dtype=torch.float32, device="cuda")
    streams = [torch.cuda.Stream() for _ in range(N_INFLIGHT)]
    torch.cuda.synchronize()

    # ------------------------------------------------------------------ device-resident value
    def step_dev(i):
        j, b = i % N_INFLIGHT, i % POOL_BATCHES
        if b == 0:
            eng.forward_device(dev[0], B, HW, HW, g_idx, g_prob, None, streams[j])
        else:
            eng.forward_device(dev[b], B, HW, HW, d_idx[j], d_prob[j], None, streams[j])

    def timed_dev(n_steps):
        """n_steps forwards over the lanes' streams, bracketed by events on stream 0 -> device ms"""
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        ev0.record(streams[0])          # the timed region starts on stream 0 ...
        for st in streams[1:]:
            st.wait_event(ev0)          # ... and no other stream starts before it
        for i in range(n_steps):
            step_dev(i)
        for st in streams[1:]:          # ... and ends when every stream has finished
            e_end = torch.cuda.Event()
            e_end.record(st)
            streams[0].wait_event(e_end)
        ev1.record(streams[0])
        return ev0, ev1

    for i in range(W):
        step_dev(i + 1)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    time.sleep(0.25)
    barrier()
    t_wall0 = time.time()
    ev0, ev1 = timed_dev(K)
    barrier()
    t_wall1 = time.time()
    ms_dev = max_over_ranks(ev0.elapsed_time(ev1))
    value = world * B * K / (ms_dev / 1e3)
    parity = {"value_loop": parity_of(g_idx.cpu().numpy(), g_prob.cpu().numpy(), gold_idx, gold_prob, "value loop")}

    # ------------------------------------------------------------------ end to end, host buffers
    host_np = host.numpy()
    E2E_INFLIGHT = int(os.environ.get("AI4E_BENCH_E2E_INFLIGHT", "3"))
    in_flight, gold_seen = [], []

    def drain(n_keep):
        while len(in_flight) > n_keep:
            b, tk = in_flight.pop(0)
            if b == 0 and not gold_seen:  # the golden batch: read every ticket once (outside the timed loops' cost model: 64 waits)
                gold_seen.append(eng.wait_many(np.arange(tk[0], tk[-1] + 1, dtype=np.uint64)))
            else:
                eng.wait(tk[-1])  # the batch completes as a whole; its last ticket is enough

    def e2e_steps(n_steps, first=0):
        for i in range(first, first + n_steps):
            b = i % POOL_BATCHES
            in_flight.append((b, eng.submit_batch(host_np[b])))
            drain(E2E_INFLIGHT - 1)
        drain(0)

    e2e_steps(W, first=1)
    gold_seen.clear()
    barrier()
    t0 = time.perf_counter()
    e2e_steps(K)
    torch.cuda.synchronize()
    dt_e2e = max_over_ranks(time.perf_counter() - t0)
    t_wall2 = time.time()
    e2e = world * B * K / dt_e2e
    parity["e2e_loop"] = parity_of(gold_seen[0][0], gold_seen[0][1], gold_idx, gold_prob, "e2e loop")
    clocks = sampler.window(t_wall0, t_wall2) if sampler else None
    stats = eng.stats()

    # ------------------------------------------------------------------ sustained: the same two loops, >= S seconds each
    sustained = None
    if args.sustained_seconds > 0:
        S = args.sustained_seconds
        chunk = max(K, 200)
        barrier()
        tw0 = time.time()
        n_dev, ms_sum = 0, 0.0
        while True:                      # every rank runs the same number of chunks: the decision is rank 0's clock via max_over_ranks
            a, b_ = timed_dev(chunk)
            torch.cuda.synchronize()
            ms_sum += a.elapsed_time(b_)
            n_dev += chunk
            if max_over_ranks(time.time() - tw0) >= S:
                break
        ms_sum = max_over_ranks(ms_sum)
        tw1 = time.time()
        barrier()
        n_e2e, t0 = 0, time.perf_counter()
        while True:
            e2e_steps(chunk)
            n_e2e += chunk
            if max_over_ranks(time.perf_counter() - t0) >= S:
                break
        torch.cuda.synchronize()
        dt_s = max_over_ranks(time.perf_counter() - t0)
        tw2 = time.time()
        sustained = {"seconds_each": S, "value": world * B * n_dev / (ms_sum / 1e3), "value_steps": n_dev,
                     "value_ms_per_step": ms_sum / n_dev, "e2e": world * B * n_e2e / dt_s, "e2e_steps": n_e2e,
                     "unit": "images/s",
                     "clocks_value": sampler.window(tw0, tw1) if sampler else None,
                     "clocks_e2e": sampler.window(tw1, tw2) if sampler else None}
        parity["sustained_value_loop"] = parity_of(g_idx.cpu().numpy(), g_prob.cpu().numpy(), gold_idx, gold_prob,
                                                   "sustained value loop")
    if sampler:
        sampler.stop(t_wall0, time.time())

    # ------------------------------------------------------------------ per-kernel roofline
    ms = eng.profile_layers(dev[0], B, HW, HW, reps=5)
    from ai4e_b200 import _lib as L
    conv_flops, igemm_ms, igemm_launches = 0, 0.0, 0
    shapes = {0: (HW, HW)}
    for li, l in enumerate(prog.layers):
        h, w = shapes[l.in_buf]
        if l.op == L.OP_HEAD:
            continue
        p_, q_ = (h + 2 * l.pad - l.kh) // l.stride + 1, (w + 2 * l.pad - l.kw) // l.stride + 1
        shapes[l.out_buf] = (p_, q_)
        if l.op == L.OP_CONV:  # every convolution runs on conv_igemm_kernel (the stem in its row-buffer mode)
            cin = IN_C if li == 0 else l.cin  # the stem's 4th input channel is padding, not work
            conv_flops += 2 * B * p_ * q_ * l.cout * cin * l.kh * l.kw
            igemm_ms += float(ms[li + 1])
            igemm_launches += 1
    if args.dump_layers and rank == 0:
        rows, shp = [], {0: (HW, HW)}
        for li, l in enumerate(prog.layers):
            h, w = shp[l.in_buf]
            if l.op == L.OP_HEAD:
                rows.append({"name": l.name, "ms": float(ms[li + 1])})
                continue
            p_, q_ = (h + 2 * l.pad - l.kh) // l.stride + 1, (w + 2 * l.pad - l.kw) // l.stride + 1
            shp[l.out_buf] = (p_, q_)
            fl = 2 * B * p_ * q_ * l.cout * l.cin * l.kh * l.kw if l.op == L.OP_CONV else 0
            byt = 2 * B * (h * w * l.cin + p_ * q_ * l.cout * (2 if l.res_buf >= 0 else 1)) + 2 * l.cout * l.cin * l.kh * l.kw
            t = float(ms[li + 1])
            rows.append({"name": l.name, "in": [h, w, l.cin], "out": [p_, q_, l.cout], "k": l.kh, "s": l.stride,
                         "ms": t, "tflops": fl / t / 1e9 if t > 0 else 0, "gbs": byt / t / 1e6 if t > 0 else 0,
                         "M": B * p_ * q_})
        json.dump({"batch": B, "preprocess_ms": float(ms[0]), "layers": rows}, open(args.dump_layers, "w"), indent=1)
    traffic, traffic_src = None, None  # DRAM bytes of the conv launches of one step, from the committed ncu capture
    for name in ("r02_traffic.json", "r01_traffic.json"):
        tpath = os.path.join(ROOT, "profiles", name)
        if os.path.exists(tpath) and B == 64 and ARCH == "resnet50":
            traffic, traffic_src = json.load(open(tpath)).get("dram_bytes_per_step"), "profiles/" + name
            break
    achieved = conv_flops / (igemm_ms / 1e3) / 1e12
    peak_burst, peak_sust = peaks["bf16_tflops"], peaks["bf16_tflops_sustained"]
    whole = conv_flops / (ms_dev / K / 1e3) / 1e12
    # K1 stand-alone (ai4e_preprocess, the engine's NHWC4 layout).  In the engine path a raw batch of the
    # model's input size never launches it: its table look-up runs inside the stem kernel (layer_ms.preprocess = 0).
    from ai4e_b200 import ops
    pre_bytes = B * (HW * HW * IN_C + HW * HW * 4 * 2)  # u8 in + bf16 NHWC4 out
    ops.preprocess(dev[1], ro.IMAGENET_MEAN, ro.IMAGENET_STD, layout="nhwc", out_channels=4)
    pe0, pe1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    pe0.record()
    for i in range(20):
        ops.preprocess(dev[1 + i % (POOL_BATCHES - 1)], ro.IMAGENET_MEAN, ro.IMAGENET_STD, layout="nhwc", out_channels=4)
    pe1.record()
    torch.cuda.synchronize()
    pre_ms = pe0.elapsed_time(pe1) / 20
    pre_gbs = pre_bytes / (pre_ms / 1e3) / 1e9

    out = {
        "metric": "images/sec end-to-end (preprocess+classify)", "value": value, "unit": "images/s",
        "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms_dev / K, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
        "value_is": "device-resident: uint8 batches already in HBM, results left in HBM; `e2e` is the end-to-end figure",
        "config": workload_config(B, world),
        "e2e": {"value": e2e, "unit": "images/s", "h2d_bytes_per_step": B * HW * HW * IN_C,
                "d2h_bytes_per_step": B * (5 * 4 + 5 * 4 + prog.num_classes * 4),
                "api": f"Engine.submit_batch/wait, page-locked host batches, {E2E_INFLIGHT} batches in flight"},
        "gpu_launches": stats["launches_per_forward"] * K,
        "parity_checked": True, "parity": parity,
        "roofline": {"bound": "tensor", "kernel": f"tcgen05 convolutions: stem_pool_kernel + conv_igemm_kernel + bottleneck_tail_kernel ({igemm_launches} layers in "
                               f"{stats['launches_per_forward'] - 4} launches per step; the other 4: preprocess, avg-pool, FC, softmax/top-5)", "achieved": achieved,
                     "peak": peak_burst, "unit": "TFLOP/s", "frac": achieved / peak_burst, "traffic": traffic,
                     "traffic_source": traffic_src,
                     "peak_source": f"{peaks['source']} bf16 burst (launches timed alone, one by one)", "ms_per_step_in_kernel": igemm_ms,
                     "flops_per_step": conv_flops,
                     # the same FLOPs over the whole pipelined step (every kernel, two micro-batches in flight)
                     "achieved_over_whole_step": whole,
                     "frac_over_whole_step": whole / peak_sust, "frac_over_whole_step_vs_burst": whole / peak_burst,
                     "whole_step_peak": peak_sust, "whole_step_peak_source": f"{peaks['source']} bf16 sustained"},
        "roofline_preprocess": {"bound": "hbm", "kernel": "preprocess_lut_kernel, stand-alone (ai4e_preprocess), 20 launches back to back by events",
                                "achieved": pre_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s", "frac": pre_gbs / peaks["hbm_gbs"],
                                "bytes_per_launch": pre_bytes, "ms": pre_ms,
                                "in_engine_path": "not launched for raw batches of the input size: K1's table look-up runs inside the stem + max-pool kernel"
                                                  if float(ms[0]) < 1e-3 else "launched",
                                "at_batch_512_by_ncu": "profiles/r02_b_preprocess_ncu.csv: 53.1 us for 282.6 MB = 5.32 TB/s (NHWC4), 54.8 us for 231.2 MB = 4.22 TB/s (NCHW)"},
        "layer_ms": {"preprocess": float(ms[0]), "stem": float(ms[1]), "maxpool": float(ms[2]),
                     "igemm_convs": igemm_ms, "head": float(ms[-1]), "sum": float(ms.sum())},
        "latency_ms": {"p50": stats["lat_ms_p50"], "p99": stats["lat_ms_p99"], "of": "submit_batch -> results on host (a batch, not a task; task latency is async_path)"},
        "clocks": clocks,
    }
    if sustained:
        sustained["conv_tflops_over_whole_step"] = conv_flops / (sustained["value_ms_per_step"] / 1e3) / 1e12
        sustained["frac_of_sustained_peak"] = sustained["conv_tflops_over_whole_step"] / peak_sust
        out["sustained"] = sustained
    eng.close()

    if rank == 0 and world == 1 and args.async_seconds > 0:
        # BASELINE.json configs[2] on this GPU: tasks through the drop-in APIService async endpoint
        try:
            from tools import async_bench
            out["async_path"] = {m: async_bench.measure(1, m, args.async_seconds, 8 if m != "native" else 4, devices=[local_rank])
                                 for m in ("later", "native")}
            out["async_path"]["later_16_images_per_task"] = async_bench.measure(1, "later", args.async_seconds, 8,
                                                                                devices=[local_rank], per_task=16)
            out["async_path"]["shell_only"] = async_bench.measure(1, "shell_only", min(2.0, args.async_seconds), 8)
        except Exception as err:  # noqa: BLE001 - the headline numbers above stand on their own
            out["async_path"] = {"error": repr(err)}
        # the latency half of the metric (p50 / p99 TASK latency, BASELINE.json configs[4]): Poisson arrivals, one image per
        # arrival, at 50 % and 80 % of the closed-loop rate just measured, generated by native threads through the pool's
        # own submit_many / poll (ai4e_pool_openloop); latency = scheduled arrival -> result announced
        try:
            from ai4e_b200.engine import Pool
            peak = float(out["async_path"]["native"]["images_per_s"])
            imgs = np.ascontiguousarray(pool_np[:4].reshape(-1, HW, HW, IN_C))  # 256 distinct images, cycled
            with Pool(prog, devices=[local_rank], max_batch=B, max_wait_us=2000) as pool:
                pool.openloop(imgs, 0.2 * peak, 0.5, feeders=4, seed=1)  # warm-up
                out["open_loop"] = {"peak_per_s": peak, "of": "one Pool on this GPU, max_wait_us 2000, 4 native feeder threads",
                                    **{f"load_{int(f * 100)}": pool.openloop(imgs, f * peak, min(3.0, args.async_seconds), feeders=4, seed=int(f * 100))
                                       for f in (0.5, 0.8)}}
        except Exception as err:  # noqa: BLE001
            out["open_loop"] = {"error": repr(err)}

    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        threads = os.cpu_count() or 1
        v, n, dt = cpu_oracle_rate(B, budget_s=12.0, max_batches=40, threads=threads)
        out["cpu_baseline"] = {"value": v, "unit": "images/s", "cores": threads, "kind": "port",
                               "sample": f"{n} batches of {B} images in {dt:.1f} s: oracle (torchvision {ARCH} fp32 eager, "
                                         f"preprocessing + softmax/top-5 included)"}
    if rank == 0:
        print(json.dumps(out), flush=True)
    group.close()


if __name__ == "__main__":
    main()
