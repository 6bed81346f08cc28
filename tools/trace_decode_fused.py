#!/usr/bin/env python
"""Decode a pipeline trace of bottleneck_tail_kernel (AI4E_B200_PIPE_TRACE=<file> with a -DAI4E_PIPE_TRACE build;
slot layout in csrc/conv_fused.cu).  Prints, relative to the kernel's entry: per ring unit when its slot was free,
how long the MMA warp waited for it and when its MMAs were issued; per pixel tile the waits of the MMA warp and the
epilogue's milestones; and a summary of where the MMA warp's time went.

    python tools/trace_decode_fused.py trace.txt [nk nu]     (nk K blocks + nu second-GEMM units per tile: 18 4)"""
import sys

import numpy as np


def main():
    v = np.zeros(2048, np.int64)
    for line in open(sys.argv[1]):
        i, t = line.split()
        v[int(i)] = int(t)
    t0 = v[1600]
    rel = lambda x: int(x - t0) if x else -1
    print("phases: setup done %d, previous kernel finished (A producer) %d, all warps done %d" % (rel(v[1601]), rel(v[1602]), rel(v[1603])))
    n = int((v[768:1024] > 0).sum())
    free, wait0, seen, issued = v[0:256], v[256:512], v[512:768], v[768:1024]
    print(f"{n} ring units traced")
    print("unit  slot_free  mma_wait_from  data_seen  issued   waited  gap_to_prev_issue")
    for g in range(n):
        gap = int(issued[g] - issued[g - 1]) if g else 0
        print(f"{g:4d} {rel(free[g]):9d} {rel(wait0[g]):9d} {rel(seen[g]):9d} {rel(issued[g]):9d} {int(seen[g] - wait0[g]):7d} {gap:7d}")
    if n > 8:
        waited = (seen[:n] - wait0[:n]).astype(np.int64)
        span = int(issued[n - 1] - wait0[0])
        print(f"MMA warp over {n} units: {span} cycles, {int(waited.sum())} waiting for data ({waited.sum() / span:.0%}), "
              f"median gap between issues {int(np.median(np.diff(issued[:n])))}")
    print("tile  acc1_empty_wait  mid_full_wait  acc2_grants(p0..p3)            | epi: acc1_full mid_empty step2_done acc2_full(p0..p3) last_store")
    for i in range(32):
        b = 1024 + 16 * i
        if not v[b + 8] and not v[b + 1]:
            break
        w1 = int(v[b + 1] - v[b + 0]) if v[b + 1] else -1
        w2 = int(v[b + 3] - v[b + 2]) if v[b + 3] else -1
        print(f"{i:4d} {w1:8d} (at {rel(v[b + 1])}) {w2:8d} (at {rel(v[b + 3])})  {[rel(x) for x in v[b + 4:b + 8]]} | "
              f"{rel(v[b + 8])} {rel(v[b + 9])} {rel(v[b + 10])} {[rel(x) for x in v[b + 11:b + 15]]} {rel(v[b + 15])}")


if __name__ == "__main__":
    main()
