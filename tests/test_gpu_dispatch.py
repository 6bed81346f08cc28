"""GPU tests of the dispatch side of the hot path, through the C ABI: many images per call with
announced completions (`submit_many` / `poll`), deferred tasks (`classify_later`), bounded waits,
the device-side chip gather of the land-cover case (K6 inside K1), the pool's bulk entry points,
the fault drill, and the whole network at the micro-batch `bench.py` times (64).

Reference call sites: the per-task thread + per-image model call of
/root/reference/APIs/1.0/base-py/ai4e_service.py:180-183,203 and the failure mapping at :185-195.
"""
import json
import os
import threading
import time

import numpy as np
import pytest

from oracle import preprocess_oracle as pp
from oracle import resnet_oracle as ro
from tests import util

pytestmark = pytest.mark.gpu


def _engine(arch, in_ch=3, hw=224, max_batch=32, max_wait_us=2000):
    from ai4e_b200.engine import Engine
    return Engine(util.program_for(arch, in_ch, hw), device=0, max_batch=max_batch, max_wait_us=max_wait_us)


def _gpu_count():
    import torch
    return torch.cuda.device_count()


# ------------------------------------------------------------------ submit_many / poll / later
def test_submit_many_announces_every_ticket_with_its_tag():
    g, imgs = util.load_golden("resnet18")
    n = 20
    order = [i % 8 for i in range(n)]
    with _engine("resnet18", max_batch=8, max_wait_us=500) as e:
        tags = np.arange(1000, 1000 + n, dtype=np.uint64)
        tickets = e.submit_many(np.ascontiguousarray(imgs[order]), tags=tags, notify=True)
        assert tickets.shape == (n,) and len(set(tickets.tolist())) == n
        seen = {}
        deadline = time.time() + 30
        while len(seen) < n and time.time() < deadline:
            tk, tg, st, idx, prob = e.poll(7, 0.5)  # a small capacity: several polls are needed
            for j in range(len(tk)):
                assert st[j] == 0
                seen[int(tg[j])] = (int(tk[j]), idx[j].copy())
        assert sorted(seen) == tags.tolist()
        for j, tag in enumerate(tags.tolist()):
            assert seen[tag][0] == int(tickets[j])
            assert np.array_equal(seen[tag][1], g["top5_idx"][order[j]])
        # nothing is announced twice, and tickets stay waitable
        assert len(e.poll(8, 0.05)[0]) == 0
        assert np.array_equal(e.wait(int(tickets[3]))[0], g["top5_idx"][order[3]])
        # a list of separate arrays and strided views take the same path
        big = np.zeros((230, 240, 3), np.uint8)
        big[3:227, 5:229] = imgs[2]
        t2 = e.submit_many([imgs[1], big[3:227, 5:229]])
        i2, _ = e.wait_many(t2)
        assert np.array_equal(i2, g["top5_idx"][[1, 2]])
        # without notify nothing reaches the queue
        assert len(e.poll(8, 0.05)[0]) == 0


def test_flipped_view_is_copied_not_read_backwards():
    g, imgs = util.load_golden("resnet18")
    with _engine("resnet18", max_batch=4, max_wait_us=300) as e:
        flipped = imgs[0][::-1]            # negative row stride
        want = e.forward(np.ascontiguousarray(flipped)[None])[0][0]
        assert np.array_equal(e.classify(flipped)[0], want)


def test_classify_later_resolves_on_the_reaper_thread():
    g, imgs = util.load_golden("resnet18")
    with _engine("resnet18", max_batch=8, max_wait_us=500) as e:
        got, threads = {}, set()

        def then(k):
            def fn(idx, prob):
                got[k] = idx
                threads.add(threading.current_thread().name)
            return fn

        ps = [e.classify_later(imgs[k % 8], then=then(k)) for k in range(12)]
        for p in ps:
            p.result(timeout=30)
        assert threads == {"ai4e-reaper"}
        for k in range(12):
            assert np.array_equal(got[k], g["top5_idx"][k % 8])
        many = e.classify_many_later(imgs[:5])
        for k, p in enumerate(many):
            assert np.array_equal(p.result(timeout=30)[0], g["top5_idx"][k])
        # an exception in `then` is carried by the Pending
        bad = e.classify_later(imgs[0], then=lambda idx, prob: 1 / 0)
        with pytest.raises(ZeroDivisionError):
            bad.result(timeout=30)


def test_wait_times_out_and_can_be_retried():
    from ai4e_b200 import Ai4eError, _lib as L
    g, imgs = util.load_golden("resnet18")
    with _engine("resnet18", max_batch=16, max_wait_us=3_000_000) as e:  # the batch stays open for 3 s
        t = e.submit(imgs[0])
        t0 = time.perf_counter()
        with pytest.raises(Ai4eError) as ei:
            e.wait(t, timeout_s=0.05)
        assert ei.value.code == L.ERR_TIMEOUT and time.perf_counter() - t0 < 1.0
        e.flush()
        assert np.array_equal(e.wait(t, timeout_s=5.0)[0], g["top5_idx"][0])


# ------------------------------------------------------------------ K6: chips read in place
@pytest.mark.parametrize("tile_hw,chip,c,first,n", [
    ((1024, 1024), 256, 4, 0, 16),     # streaming kernel, strided rows (the configs[3] geometry in small)
    ((1024, 1280), 256, 4, 7, 9),      # a chip range that starts and ends mid-row
    ((512, 768), 256, 3, 0, 6),        # 3 bands: 768-byte rows
    ((300, 500), 100, 4, 2, 11),       # ragged: generic kernel, remainder ignored
])
def test_preprocess_tile_is_bit_exact(tile_hw, chip, c, first, n):
    """ai4e_preprocess_tile == oracle chips_from_tile + preprocess, bit for bit, NCHW and NHWC4."""
    import torch
    from ai4e_b200 import ops
    rng = np.random.default_rng(4321)
    tile = rng.integers(0, 256, (*tile_hw, c), dtype=np.uint8)
    mean, std = util.mean_std(c)
    chips = pp.chips_from_tile(tile, chip)[first:first + n]
    want = pp.bf16_bits(pp.preprocess_nchw(chips, mean, std))
    d_tile = torch.from_numpy(tile).cuda()
    got = ops.preprocess_tile(d_tile, chip, mean, std, first_chip=first, n_chips=n)
    assert np.array_equal(got.view(torch.int16).cpu().numpy().view(np.uint16), want)
    nhwc = ops.preprocess_tile(d_tile, chip, mean, std, first_chip=first, n_chips=n, layout="nhwc", out_channels=4)
    got4 = nhwc.view(torch.int16).cpu().numpy().view(np.uint16)
    assert np.array_equal(got4[..., :c].transpose(0, 3, 1, 2), want)
    if c == 3:
        assert (got4[..., 3] == 0).all()
    # a tile that is itself a view with a wider row pitch
    wide = torch.zeros((tile_hw[0], tile_hw[1] + 16, c), dtype=torch.uint8, device="cuda")
    wide[:, :tile_hw[1]] = d_tile
    got_w = ops.preprocess_tile(wide[:, :tile_hw[1]], chip, mean, std, first_chip=first, n_chips=n)
    assert torch.equal(got_w, got)


def test_preprocess_tile_resize_matches_oracle():
    import torch
    from ai4e_b200 import ops
    rng = np.random.default_rng(99)
    tile = rng.integers(0, 256, (160, 240, 3), dtype=np.uint8)
    chips = pp.chips_from_tile(tile, 80)
    want = pp.bf16_bits(pp.preprocess_resize_crop_nchw(chips, 64, 64, 56, 56, ro.IMAGENET_MEAN, ro.IMAGENET_STD))
    got = ops.preprocess_tile(torch.from_numpy(tile).cuda(), 80, ro.IMAGENET_MEAN, ro.IMAGENET_STD,
                              resize=(64, 64), crop=(56, 56))
    assert np.array_equal(got.view(torch.int16).cpu().numpy().view(np.uint16), want)


def test_preprocess_tile_rejects_bad_ranges():
    import torch
    from ai4e_b200 import ops, Ai4eError
    t = torch.zeros((512, 512, 4), dtype=torch.uint8, device="cuda")
    with pytest.raises(Ai4eError):
        ops.preprocess_tile(t, 256, ro.MEAN4, ro.STD4, first_chip=3, n_chips=2)
    with pytest.raises(Ai4eError):
        ops.preprocess_tile(t, 1024, ro.MEAN4, ro.STD4)


def _golden_tile(imgs, gh, gw):
    tile = np.zeros((gh * 256, gw * 256, 4), np.uint8)
    for k in range(gh * gw):
        r, c = divmod(k, gw)
        tile[r * 256:(r + 1) * 256, c * 256:(c + 1) * 256] = imgs[k % len(imgs)]
    return tile


def test_tile_submit_pageable_pinned_and_pool():
    """tile -> device once -> chips in place -> per-chip top-5: pageable and page-locked tiles, an
    engine and a two-replica pool (contiguous chip ranges), all equal to the golden chips."""
    import torch
    from ai4e_b200.engine import Pool
    from ai4e_b200.landcover import classify_tile
    g, imgs = util.load_golden("resnet50_landcover4")
    tile = _golden_tile(imgs, 3, 4)          # 12 chips, max_batch 5: three batches, the last ragged
    want = g["top5_idx"][[k % len(imgs) for k in range(12)]]
    with _engine("resnet50", in_ch=4, hw=256, max_batch=5, max_wait_us=500) as e:
        for rep in range(3):                 # more tiles than tile buffers: buffers are recycled
            cls, idx, prob = classify_tile(e, tile)
            assert np.array_equal(idx.reshape(12, 5), want)
        pinned = torch.from_numpy(tile).pin_memory().numpy()
        tk = e.submit_tile(pinned, 256)
        idx2, prob2 = e.wait_many(tk)
        assert np.array_equal(idx2, want) and np.array_equal(prob2, prob.reshape(12, 5))
        assert e.stats()["images"] == 48
    with Pool(util.program_for("resnet50", 4, 256), devices=[0, 0], max_batch=4, max_wait_us=500) as p:
        tk = p.submit_tile(tile, 256)
        assert sorted(set((tk >> np.uint64(56)).tolist())) == [0, 1]  # both replicas got a contiguous half
        assert ((tk[:6] >> np.uint64(56)) == (tk[0] >> np.uint64(56))).all()
        idx3, _ = p.wait_many(tk)
        assert np.array_equal(idx3, want)


def test_native_open_loop_generator_completes_every_arrival():
    """ai4e_pool_openloop (measurement aid): every scheduled arrival is submitted, announced once and timed."""
    from ai4e_b200.engine import Pool
    g, imgs = util.load_golden("resnet18")
    with Pool(util.program_for("resnet18"), devices=[0, 0], max_batch=8, max_wait_us=500) as p:
        r = p.openloop(np.ascontiguousarray(imgs[:16]), rate_per_s=3000.0, seconds=0.4, feeders=3, seed=5)
        assert r["tasks"] == 1200 and r["completed"] == 1200 and r["failed"] == 0
        assert 0.0 < r["p50_ms"] <= r["p99_ms"] <= r["max_ms"] < 2000.0
        assert 2000.0 < r["achieved_per_s"] < 4000.0
        assert sum(st["images"] for st in p.stats()) == 1200
        # the pool is usable afterwards and its queue is empty
        assert np.array_equal(p.classify(imgs[0])[0], g["top5_idx"][0])
        assert len(p.poll(8, 0.0)[0]) == 0


# ------------------------------------------------------------------ pool: bulk entry points, fault drill
def test_pool_bulk_entry_points():
    import torch
    from ai4e_b200.engine import Pool
    g, imgs = util.load_golden("resnet18")
    with Pool(util.program_for("resnet18"), devices=[0, 0], max_batch=8, max_wait_us=400) as p:
        order = [i % 8 for i in range(40)]
        tags = np.arange(40, dtype=np.uint64) + 7
        tk = p.submit_many(np.ascontiguousarray(imgs[order]), tags=tags, notify=True)
        seen = {}
        deadline = time.time() + 30
        while len(seen) < 40 and time.time() < deadline:
            t, tg, st, idx, _ = p.poll(64, 0.5)       # ONE queue for both replicas
            for j in range(len(t)):
                assert st[j] == 0
                seen[int(tg[j])] = (int(t[j]), idx[j].copy())
        assert sorted(seen) == tags.tolist()
        for j in range(40):
            assert seen[int(tags[j])][0] == int(tk[j])
            assert np.array_equal(seen[int(tags[j])][1], g["top5_idx"][order[j]])
        per = [s["images"] for s in p.stats()]
        assert sum(per) == 40 and all(v > 0 for v in per), per
        # a caller-made batch goes to one replica as a whole; page-locked memory is read in place
        batch = torch.from_numpy(np.ascontiguousarray(imgs[:8])).pin_memory().numpy()
        tb = p.submit_batch(batch)
        assert len(set((tb >> np.uint64(56)).tolist())) == 1
        ib, _ = p.wait_many(tb)
        assert np.array_equal(ib, g["top5_idx"][:8])
        # deferred tasks work on a pool as on an engine
        ps = p.classify_many_later(imgs[:6])
        for k, pend in enumerate(ps):
            assert np.array_equal(pend.result(timeout=30)[0], g["top5_idx"][k])


def test_fault_fails_only_that_batch_and_the_pool_routes_around_the_replica():
    """SURVEY 5's failure model: a CUDA error on a replica fails that batch's tasks only, marks the
    replica unhealthy, the dispatcher routes around it, and the host shell turns the error into
    FailTask(taskId, 'Task failed - try again') (reference ai4e_service.py:185-195)."""
    from ai4e_b200 import Ai4eError, _lib as L
    from ai4e_b200.engine import Pool
    from ai4e_service import APIService
    from tests.host_shell_scenario import RecordingTaskManager
    try:
        from flask import Flask
    except ImportError:
        from ai4e_b200.compat.miniflask import Flask
    g, imgs = util.load_golden("resnet18")
    with Pool(util.program_for("resnet18"), devices=[0, 0], max_batch=4, max_wait_us=300) as p:
        engines = p.engines()
        ok_before = [p.submit(imgs[i]) for i in range(4)]
        for i, t in enumerate(ok_before):
            assert np.array_equal(p.wait(t)[0], g["top5_idx"][i])
        engines[0].inject_fault(1)
        # route four images to replica 0 directly: its next batch fails as a whole
        doomed = engines[0].submit_many(np.ascontiguousarray(imgs[:4]), tags=np.arange(4, dtype=np.uint64), notify=True)
        for t in doomed.tolist():
            with pytest.raises(Ai4eError) as ei:
                engines[0].wait(t)
            assert ei.value.code == L.ERR_CUDA
        _, tg, st, _, _ = p.poll(8, 1.0)
        assert sorted(tg.tolist()) == [0, 1, 2, 3] and (st == L.ERR_CUDA).all()
        assert [s["healthy"] for s in p.stats()] == [False, True]
        # results that finished before the fault are still readable
        assert np.array_equal(p.wait(ok_before[0])[0], g["top5_idx"][0])
        with pytest.raises(Ai4eError):
            engines[0].submit(imgs[0])
        # the pool keeps serving from the healthy replica
        after = [p.submit(imgs[i % 8]) for i in range(12)]
        assert {t >> 56 for t in after} == {1}
        for i, t in enumerate(after):
            assert np.array_equal(p.wait(t)[0], g["top5_idx"][i % 8])

    # the same fault seen from the host shell: blocked classify and deferred classify both fail the task
    app = Flask("f")
    svc = APIService(app, type("L", (), {"tracer": None, "log_exception": lambda *a, **k: None})())
    tm = RecordingTaskManager()
    svc.api_task_manager = tm
    with _engine("resnet18", max_batch=2, max_wait_us=200) as e:
        @svc.api_async_func(api_path="/c", methods=["POST"], maximum_concurrent_requests=16,
                            request_processing_function=lambda request: {"image": imgs[0]})
        def c_(*args, **kwargs):
            if kwargs["taskId"] == "later":
                return e.classify_later(kwargs["image"], then=lambda idx, prob: tm.CompleteTask("later", "completed"))
            e.classify(kwargs["image"])
            tm.CompleteTask(kwargs["taskId"], "completed")

        e.inject_fault(1)
        cl = app.test_client()
        assert cl.post("/c", headers={"taskId": "blocked"}).get_data(as_text=True) == "TaskId: blocked"
        assert cl.post("/c", headers={"taskId": "later"}).get_data(as_text=True) == "TaskId: later"
        deadline = time.time() + 20
        while time.time() < deadline:
            fails = sorted(c[1] for c in tm.calls if c[0] == "FailTask")
            if fails == ["blocked", "later"] and svc.func_request_counts["/c"] == 0:
                break
            time.sleep(0.01)
        assert sorted(c[1] for c in tm.calls if c[0] == "FailTask") == ["blocked", "later"]
        assert all(c[2] == "Task failed - try again" for c in tm.calls if c[0] == "FailTask")
        assert not [c for c in tm.calls if c[0] == "CompleteTask"]
        assert svc.func_request_counts["/c"] == 0


@pytest.mark.skipif("_gpu_count() < 2")
def test_pool_over_two_gpus():
    """one dispatcher, one replica per GPU (north_star item 4), in ONE process."""
    from ai4e_b200.engine import Pool
    g, imgs = util.load_golden("resnet18")
    n_dev = min(_gpu_count(), 8)
    with Pool(util.program_for("resnet18"), devices=list(range(n_dev)), max_batch=8, max_wait_us=400) as p:
        order = [i % 8 for i in range(16 * n_dev)]
        tk = p.submit_many(np.ascontiguousarray(imgs[order]))
        idx, _ = p.wait_many(tk)
        assert np.array_equal(idx, g["top5_idx"][order])
        per = [s["images"] for s in p.stats()]
        assert all(v > 0 for v in per), per
    g4, chips = util.load_golden("resnet50_landcover4")
    tile = _golden_tile(chips, 2, 4)
    with Pool(util.program_for("resnet50", 4, 256), devices=list(range(n_dev)), max_batch=4, max_wait_us=400) as p:
        tk = p.submit_tile(tile, 256)
        idx, _ = p.wait_many(tk)
        assert np.array_equal(idx, g4["top5_idx"][[k % len(chips) for k in range(8)]])
        assert len(set((tk >> np.uint64(56)).tolist())) == min(n_dev, 8)


# ------------------------------------------------------------------ the benchmarked configuration
def _check_b64(g, idx, logits, what):
    emul, ref = g["logits_bf16_emul"], g["logits_fp32"]
    ok = util.within_north_star(logits, emul)
    assert ok.all(), f"{what}: {(~ok).sum()} logits outside 1e-2/1e-3 of the bf16 oracle; max {np.abs(logits - emul).max():.4f}"
    assert np.array_equal(idx, g["top5_idx"]), f"{what}: top-5 differs from the fp32 reference"
    d = np.abs(logits - ref)
    assert d.max() <= 2.5e-2, (what, d.max())


def test_resnet50_microbatch_64_as_benchmarked():
    """BASELINE.json configs[1] exactly as bench.py runs it: ResNet-50, max_batch 64, tilings autotuned
    for 64, two lanes, 132-SM grids - through forward_host, forward_device on two streams at once, and
    submit_batch from page-locked memory.  Every logit against the bf16 oracle at north_star's
    tolerance (1e-2 abs / 1e-3 rel), top-5 identical to the fp32 reference, on all 64 images."""
    import torch
    g, imgs = util.load_golden("resnet50_b64")
    assert len(imgs) == 64
    with _engine("resnet50", max_batch=64) as e:
        idx, prob, logits = e.forward(imgs, want_logits=True)
        _check_b64(g, idx, logits, "forward_host")
        assert np.allclose(prob, g["top5_prob"], atol=2e-4, rtol=3e-2)
        # two micro-batches in flight on the two lanes (what `value` times): the batch and its reverse
        a = torch.from_numpy(np.ascontiguousarray(imgs)).cuda()
        b = torch.from_numpy(np.ascontiguousarray(imgs[::-1])).cuda()
        outs = []
        streams = [torch.cuda.Stream(), torch.cuda.Stream()]
        for rep in range(3):
            for x, st in ((a, streams[0]), (b, streams[1])):
                o = (torch.empty((64, 5), dtype=torch.int32, device="cuda"), torch.empty((64, 5), dtype=torch.float32, device="cuda"),
                     torch.zeros((64, 1000), dtype=torch.float32, device="cuda"))
                e.forward_device(x, 64, 224, 224, *o, stream=st)
                outs.append((x is b, o))
        torch.cuda.synchronize()
        for rev, o in outs:
            i_, l_ = o[0].cpu().numpy(), o[2].cpu().numpy()
            _check_b64(g, i_[::-1] if rev else i_, l_[::-1] if rev else l_, "forward_device, two lanes")
        # the e2e path of the bench: page-locked batches, three in flight
        host = torch.from_numpy(np.ascontiguousarray(imgs)).pin_memory().numpy()
        tickets = [e.submit_batch(host) for _ in range(4)]
        for tk in tickets:
            res = [e.wait(t, want_logits=True) for t in tk]
            _check_b64(g, np.stack([r[0] for r in res]), np.stack([r[2] for r in res]), "submit_batch")
        assert e.stats()["healthy"]


def test_engine_with_and_without_stem_pool_fusion_agree_bit_for_bit(monkeypatch):
    """the fused stem + max-pool kernel inside the engine - behind K1 (the default) or, with
    AI4E_B200_U8_STEM=1, with the uint8 batch as its own input for the 3-band model - gives exactly the
    logits of the three separate kernels (AI4E_B200_NO_STEM_POOL=1: K1, stem, max-pool)."""
    from ai4e_b200.engine import Engine
    for name, arch, in_ch, hw in (("resnet50_b64", "resnet50", 3, 224), ("resnet50_landcover4", "resnet50", 4, 256)):
        g, imgs = util.load_golden(name)
        imgs = imgs[:16]
        prog = util.program_for(arch, in_ch, hw)
        monkeypatch.setenv("AI4E_B200_U8_STEM", "1")
        with Engine(prog, device=0, max_batch=16) as e:
            fused = e.forward(imgs, want_logits=True)
            launches_fused = e.stats()["launches_per_forward"]
        monkeypatch.delenv("AI4E_B200_U8_STEM")
        with Engine(prog, device=0, max_batch=16) as e:
            no_u8 = e.forward(imgs, want_logits=True)
            launches_no_u8 = e.stats()["launches_per_forward"]
        monkeypatch.setenv("AI4E_B200_NO_STEM_POOL", "1")
        with Engine(prog, device=0, max_batch=16) as e:
            plain = e.forward(imgs, want_logits=True)
            launches_plain = e.stats()["launches_per_forward"]
        monkeypatch.delenv("AI4E_B200_NO_STEM_POOL")
        assert launches_no_u8 == launches_plain - 1
        assert launches_fused == launches_plain - (2 if in_ch == 3 else 1)   # K1 disappears too for 3-band pixels
        for other in (no_u8, plain):
            assert np.array_equal(fused[2], other[2]) and np.array_equal(fused[0], other[0])
        assert np.array_equal(fused[0], g["top5_idx"][:16])


def test_gain1_model_report():
    """SURVEY 8(d)'s literal recipe leaves the residual branches at gain 1: a far worse conditioned
    network than the golden model.  This test REPORTS how the engine agrees with the
    fp32 and bf16-emulated CPU forwards on it (top-1 / top-5 agreement, max |dlogit|, the margins) into
    gpurun_out/gain1_report.json; it asserts only that the engine runs and is finite."""
    from ai4e_b200.engine import Engine
    from ai4e_b200.models import build_resnet_program
    m = ro.build_model("resnet50", residual_gain=1.0)
    imgs = ro.synthetic_images(32, seed=555)
    x = pp.preprocess_nchw(imgs, ro.IMAGENET_MEAN, ro.IMAGENET_STD)
    ref = ro.forward_fp32(m, x)
    emul = ro.forward_bf16_emul(m, x)
    with Engine(build_resnet_program("resnet50", m.state_dict(), 224, 224), device=0, max_batch=32) as e:
        idx, prob, logits = e.forward(imgs, want_logits=True)
    assert np.isfinite(logits).all()
    r_idx, _ = ro.softmax_top5(ref)
    e_idx, _ = ro.softmax_top5(emul)
    srt = -np.sort(-ref, axis=1)
    scale = float(np.abs(ref).max())
    rep = {
        "model": "resnet50, residual_gain=1.0 (SURVEY 8(d) literal recipe), 32 structured images, seed 555",
        "logit_scale_max_abs": scale,
        "engine_vs_fp32": {"top1_agree": float((idx[:, 0] == r_idx[:, 0]).mean()), "top5_identical": float((idx == r_idx).all(1).mean()),
                           "max_abs_dlogit": float(np.abs(logits - ref).max()), "max_rel_to_scale": float(np.abs(logits - ref).max() / scale)},
        "engine_vs_bf16_emul": {"top1_agree": float((idx[:, 0] == e_idx[:, 0]).mean()), "top5_identical": float((idx == e_idx).all(1).mean()),
                                "max_abs_dlogit": float(np.abs(logits - emul).max()),
                                "within_1e-2_1e-3": float(util.within_north_star(logits, emul).mean())},
        "bf16_emul_vs_fp32": {"top1_agree": float((e_idx[:, 0] == r_idx[:, 0]).mean()), "top5_identical": float((e_idx == r_idx).all(1).mean()),
                              "max_abs_dlogit": float(np.abs(emul - ref).max())},
        "fp32_margins": {"top1_minus_top2_min": float((srt[:, 0] - srt[:, 1]).min()), "top5_minus_top6_min": float((srt[:, 4] - srt[:, 5]).min()),
                         "top1_minus_top2_median": float(np.median(srt[:, 0] - srt[:, 1]))},
    }
    out = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "gpurun_out")
    os.makedirs(out, exist_ok=True)
    with open(os.path.join(out, "gain1_report.json"), "w") as fh:
        json.dump(rep, fh, indent=1)
    print("gain-1 report:", json.dumps(rep))
