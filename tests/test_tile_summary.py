"""K7, the on-device aggregate of the land-cover pipeline (tile -> chips -> classify ->
aggregate without a host hop between the stages; the reference chains stages as tasks through AddPipelineTask,
/root/reference/APIs/1.0/Common/task_management/distributed_api_task.py:67-100).

CPU: the oracle on hand-checked cases.  GPU: the kernel against the oracle bit for bit (every output is an integer
or a copied float), the engine's summary against the oracle applied to the per-chip results of `submit_tile`, the
same summary from one replica and from two (integer sums: replica count must not matter), and the failure path."""
import numpy as np
import pytest

from oracle import tile_summary_oracle as tso
from tests import util


def test_oracle_on_a_hand_checked_case():
    idx = np.array([[2, 0, 1, 3, 4], [2, 1, 0, 3, 4], [0, 2, 1, 3, 4]], np.int32)
    prob = np.array([[0.5, 0.25, 0.125, 0.0625, 0.0625], [1.0, 0, 0, 0, 0], [0.75, 0.25, 0, 0, 0]], np.float32)
    labels, lp, hist, mass = tso.tile_summary(idx, prob, 6)
    assert labels.tolist() == [2, 2, 0] and lp.tolist() == [0.5, 1.0, 0.75]
    assert hist.tolist() == [1, 0, 2, 0, 0, 0]
    one = 1 << 32
    assert mass.tolist() == [one // 4 + (3 * one) // 4, one // 8, one // 2 + one + one // 4, one // 16, one // 16, 0]
    assert mass.dtype == np.uint64 and hist.dtype == np.uint32


def test_oracle_ignores_bad_entries():
    idx = np.array([[7, -1, 1, 0, 2]], np.int32)         # classes outside [0, 3) add nothing; NaN adds nothing
    prob = np.array([[0.5, 0.25, np.nan, 2.0, -1.0]], np.float32)
    labels, lp, hist, mass = tso.tile_summary(idx, prob, 3)
    assert labels.tolist() == [7] and hist.tolist() == [0, 0, 0]
    assert mass.tolist() == [1 << 32, 0, 0]              # 2.0 clamps to 1, NaN and -1 to 0


def _random_top5(rng, n, classes):
    idx = np.stack([rng.permutation(classes)[:5] for _ in range(n)]).astype(np.int32)
    raw = np.sort(rng.random((n, 5)).astype(np.float32), axis=1)[:, ::-1]
    prob = (raw / raw.sum(1, keepdims=True) * rng.random((n, 1)).astype(np.float32)).astype(np.float32)
    return idx, np.ascontiguousarray(prob)


@pytest.mark.gpu
@pytest.mark.parametrize("n,classes", [(1, 5), (64, 10), (256, 1000), (1000, 17), (4096, 1000)])
def test_kernel_matches_oracle_bit_for_bit(n, classes):
    import torch
    from ai4e_b200 import ops
    rng = np.random.default_rng(n + classes)
    idx, prob = _random_top5(rng, n, classes)
    if n >= 64:
        prob[3, 1] = np.nan
        prob[5, 0] = 1.0
        idx[7, 4] = classes + 3                                  # out of range: ignored
    want = tso.tile_summary(idx, prob, classes)
    # in one launch, and in three ragged parts folded into the same summary (batches of one tile)
    got = ops.tile_summary(torch.from_numpy(idx).cuda(), torch.from_numpy(prob).cuda(), classes)
    cuts = [0, n // 3, n // 3 + n // 5, n]
    acc = None
    for a, b in zip(cuts[:-1], cuts[1:]):
        if b == a:
            continue
        if acc is None:
            acc = (torch.zeros(n, dtype=torch.int32, device="cuda"), torch.zeros(n, dtype=torch.float32, device="cuda"),
                   torch.zeros(classes, dtype=torch.int32, device="cuda"), torch.zeros(classes, dtype=torch.int64, device="cuda"))
        ops.tile_summary(torch.from_numpy(idx[a:b].copy()).cuda(), torch.from_numpy(prob[a:b].copy()).cuda(), classes, first=a, out=acc)
    torch.cuda.synchronize()
    for res in (got, acc):
        labels, lp, hist, mass = (t.cpu().numpy() for t in res)
        assert np.array_equal(labels, want[0])
        assert np.array_equal(lp.view(np.uint32), want[1].view(np.uint32))
        assert np.array_equal(hist.view(np.uint32), want[2])
        assert np.array_equal(mass.view(np.uint64), want[3])


def _golden_tile(imgs, gh, gw):
    tile = np.zeros((gh * 256, gw * 256, 4), np.uint8)
    for k in range(gh * gw):
        r, c = divmod(k, gw)
        tile[r * 256:(r + 1) * 256, c * 256:(c + 1) * 256] = imgs[k % len(imgs)]
    return tile


@pytest.mark.gpu
def test_engine_tile_summary_equals_oracle_of_the_per_chip_results():
    import torch
    from ai4e_b200.engine import Engine, Pool
    from ai4e_b200.landcover import summarize_tile
    g, imgs = util.load_golden("resnet50_landcover4")
    tile = _golden_tile(imgs, 3, 4)                       # 12 chips; max_batch 5: three batches, the last ragged
    prog = util.program_for("resnet50", 4, 256)
    K = prog.num_classes
    with Engine(prog, device=0, max_batch=5, max_wait_us=500) as e:
        idx, prob = e.wait_many(e.submit_tile(tile, 256))
        assert np.array_equal(idx, g["top5_idx"][[k % len(imgs) for k in range(12)]])
        want = tso.tile_summary(idx, prob, K)
        for rep in range(4):                              # more tiles than tile buffers: the summary buffers are recycled
            s = summarize_tile(e, tile)
            assert s["labels"].shape == (3, 4) and np.array_equal(s["labels"].ravel(), want[0])
            assert np.array_equal(s["label_prob"].ravel().view(np.uint32), want[1].view(np.uint32))
            assert np.array_equal(s["hist"], want[2]) and np.array_equal(s["mass_q32"], want[3])
            assert int(s["hist"].sum()) == 12 and abs(s["mass"].sum() - prob.astype(np.float64).sum()) < 1e-6
        # several tiles in flight, page-locked source, waited for out of order
        pinned = torch.from_numpy(tile).pin_memory().numpy()
        flipped = np.ascontiguousarray(tile[:, ::-1])     # another tile: the chip order (and so the label map) differs
        t1, grid = e.submit_tile_summary(pinned, 256)
        t2, _ = e.submit_tile_summary(flipped, 256)
        t3, _ = e.submit_tile_summary(pinned, 256)
        s3, s2, s1 = e.wait_tile_summary(t3, grid), e.wait_tile_summary(t2, grid), e.wait_tile_summary(t1, grid)
        for s in (s1, s3):
            assert np.array_equal(s["labels"].ravel(), want[0]) and np.array_equal(s["mass_q32"], want[3])
        idx_f, prob_f = e.wait_many(e.submit_tile(flipped, 256))
        want_f = tso.tile_summary(idx_f, prob_f, K)
        assert np.array_equal(s2["labels"].ravel(), want_f[0]) and np.array_equal(s2["hist"], want_f[2])
        assert np.array_equal(s2["mass_q32"], want_f[3])
        single = s1
    # two replicas share the tile (6 chips each): integer sums, so the summary is the same to the bit
    with Pool(prog, devices=[0, 0], max_batch=4, max_wait_us=500) as p:
        for rep in range(3):
            s = summarize_tile(p, tile)
            for k in ("labels", "hist", "mass_q32"):
                assert np.array_equal(s[k], single[k]), k
            assert np.array_equal(s["label_prob"].view(np.uint32), single["label_prob"].view(np.uint32))
        assert sum(st["images"] for st in p.stats()) == 36 and all(st["images"] == 18 for st in p.stats())


@pytest.mark.gpu
def test_tile_summary_failure_and_bad_tickets():
    from ai4e_b200 import Ai4eError, _lib as L
    from ai4e_b200.engine import Engine
    g, imgs = util.load_golden("resnet50_landcover4")
    tile = _golden_tile(imgs, 2, 3)
    with Engine(util.program_for("resnet50", 4, 256), device=0, max_batch=4, max_wait_us=300) as e:
        with pytest.raises(Ai4eError):
            e.wait_tile_summary(0, (2, 3))                # nothing submitted yet
        ok, grid = e.submit_tile_summary(tile, 256)
        assert int(e.wait_tile_summary(ok, grid)["hist"].sum()) == 6
        with pytest.raises(Ai4eError):
            e.submit_tile_summary(tile[:, :, :3].copy(), 256)   # band count the model does not take
        with pytest.raises(Ai4eError):
            e.submit_tile_summary(tile[:100], 256)              # smaller than one chip
        e.inject_fault(1)                                 # the first of the tile's two batches fails like a CUDA fault
        bad, _ = e.submit_tile_summary(tile, 256)
        with pytest.raises(Ai4eError) as ei:
            e.wait_tile_summary(bad, grid)
        assert ei.value.code == L.ERR_CUDA
        assert not e.stats()["healthy"]
