"""N > 1: one process per GPU, replicas only (DESIGN.md "multi-GPU").  The cross-rank plumbing of
bench.py is exercised here with world_size 2 over gloo on the CPU."""
import json
import os
import socket
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "aiforearth-api-platform_b200")


def _free_port() -> int:
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _spawn(script: str, world: int, extra_env=None, timeout=240):
    port = _free_port()
    procs = []
    for rank in range(world):
        env = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world),
                   MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), PYTHONPATH=os.pathsep.join([ROOT, PKG]))
        env.update(extra_env or {})
        procs.append(subprocess.Popen([sys.executable, "-c", script], env=env, stdout=subprocess.PIPE,
                                      stderr=subprocess.PIPE, text=True))
    outs = []
    for p in procs:
        try:
            o, e = p.communicate(timeout=timeout)
        except subprocess.TimeoutExpired:
            for q in procs:
                q.kill()
            raise
        assert p.returncode == 0, e[-2000:]
        outs.append(o)
    return outs


WORKER = r"""
import json, os, time
from tools.replica_group import ReplicaGroup, shard_range
g = ReplicaGroup("gloo")
assert g.world == 2 and g.backend == "gloo"
# rank 1 is late to the barrier; nobody may pass before it arrives
if g.rank == 1:
    time.sleep(0.5)
t0 = time.time(); g.barrier(); waited = time.time() - t0
# per-rank "device times": the job's time is the slowest rank's
ms = [10.0, 14.0][g.rank]
worst = g.max_over_ranks(ms)
rate = g.aggregate_rate(64 * 5, ms)          # 5 steps of 64 images on each rank
mine = list(shard_range(11, g.world, g.rank))
g.barrier()
print(json.dumps({"rank": g.rank, "waited": waited, "worst": worst, "rate": rate, "mine": mine}))
g.close()
"""


def test_replica_group_world2_gloo():
    outs = [json.loads(o.strip().splitlines()[-1]) for o in _spawn(WORKER, 2)]
    by_rank = {o["rank"]: o for o in outs}
    assert by_rank[0]["waited"] > 0.3            # rank 0 really waited for rank 1
    for o in outs:
        assert o["worst"] == 14.0
        assert o["rate"] == pytest.approx(2 * 64 * 5 / 14e-3)
    assert by_rank[0]["mine"] + by_rank[1]["mine"] == list(range(11))   # request sharding covers everything once
    assert len(by_rank[0]["mine"]) - len(by_rank[1]["mine"]) in (0, 1)


def test_replica_group_world1_needs_no_process_group():
    from tools.replica_group import ReplicaGroup, shard_range, merge_shards
    env = {k: os.environ.pop(k, None) for k in ("RANK", "LOCAL_RANK", "WORLD_SIZE")}
    try:
        g = ReplicaGroup()
        g.barrier()
        assert g.world == 1 and g.max_over_ranks(3.5) == 3.5 and g.aggregate_rate(640, 10.0) == 64000.0
        g.close()
    finally:
        for k, v in env.items():
            if v is not None:
                os.environ[k] = v
    parts = [list(shard_range(10, 4, r)) for r in range(4)]
    assert [len(p) for p in parts] == [3, 3, 2, 2] and merge_shards(parts) == list(range(10))
    assert list(shard_range(2, 4, 3)) == []
    with pytest.raises(ValueError):
        shard_range(4, 2, 2)


def test_bench_reference_arm_prints_once_under_two_ranks():
    """`bench.py --impl reference` under torchrun: rank 0 alone runs and prints, the others exit 0."""
    port = _free_port()
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2",
                                       "--steps", "1", "--warmup", "0", "--batch", "2"], env=env,
                                      stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = []
    for p in procs:
        o, e = p.communicate(timeout=600)
        assert p.returncode == 0, e[-2000:]
        outs.append(o.strip())
    assert outs[1] == ""
    line = json.loads(outs[0].splitlines()[-1])
    assert line["impl"] == "reference" and line["n_gpus"] == 2 and line["value"] > 0
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["cpu_baseline"]["kind"] == "port"
