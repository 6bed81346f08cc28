#!/usr/bin/env python
"""Time one bottleneck tail (3x3 + 1x1 + residual, ai4e_bottleneck_tail) with CUDA events; used under ncu.

    python tools/fused_microbench.py B H W Cin Cmid Cout STRIDE [reps]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "aiforearth-api-platform_b200")]
import torch  # noqa: E402

from ai4e_b200 import ops  # noqa: E402


def main():
    b, h, w, cin, cmid, cout, stride = [int(v) for v in sys.argv[1:8]]
    reps = int(sys.argv[8]) if len(sys.argv) > 8 else 10
    g = torch.Generator().manual_seed(0)
    x = (torch.randn((b, h, w, cin), generator=g) * 0.5).to(torch.bfloat16).cuda()
    w2 = (torch.randn((cmid, 3, 3, cin), generator=g) * 0.05).to(torch.bfloat16).cuda()
    b2 = torch.randn(cmid, generator=g).cuda()
    w3 = (torch.randn((cout, cmid), generator=g) * 0.05).to(torch.bfloat16).cuda()
    b3 = torch.randn(cout, generator=g).cuda()
    p, q = (h + 2 - 3) // stride + 1, (w + 2 - 3) // stride + 1
    r = (torch.randn((b, p, q, cout), generator=g) * 0.5).to(torch.bfloat16).cuda()
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    for _ in range(2):
        ops.bottleneck_tail(x, w2, b2, stride, 1, w3, b3, r)
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.bottleneck_tail(x, w2, b2, stride, 1, w3, b3, r)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    ms = ts[len(ts) // 2]
    flops = 2 * b * p * q * (cmid * cin * 9 + cout * cmid)
    print(f"tail {sys.argv[1:8]} median {ms * 1e3:.1f} us  {flops / ms / 1e9:.1f} TFLOP/s (cold L2)")


if __name__ == "__main__":
    main()
