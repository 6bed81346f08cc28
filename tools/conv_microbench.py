#!/usr/bin/env python
"""Time one convolution shape through the C ABI (ai4e_conv_layer) with CUDA events.

    python tools/conv_microbench.py B H W Cin Cout K STRIDE PAD RELU RES [reps]

Used under `ncu --set full -k regex:conv_igemm` for the per-kernel profiles in profiles/."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "aiforearth-api-platform_b200")]
import torch  # noqa: E402

from ai4e_b200 import ops, _lib as L  # noqa: E402


def main():
    b, h, w, cin, cout, k, stride, pad, relu, res = [int(v) for v in sys.argv[1:11]]
    reps = int(sys.argv[11]) if len(sys.argv) > 11 else 20
    g = torch.Generator().manual_seed(0)
    x = (torch.randn((b, h, w, cin), generator=g) * 0.5).to(torch.bfloat16).cuda()
    wt = (torch.randn((cout, k, k, cin), generator=g) * 0.05).to(torch.bfloat16).cuda()
    bias = torch.randn(cout, generator=g).cuda()
    p, q = (h + 2 * pad - k) // stride + 1, (w + 2 * pad - k) // stride + 1
    r = (torch.randn((b, p, q, cout), generator=g) * 0.5).to(torch.bfloat16).cuda() if res else None
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    for _ in range(3):
        ops.conv_layer(x, wt, bias, stride, pad, r, bool(relu), algo=L.CONV_IGEMM)
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        flush.zero_()  # evict L2 between timed launches
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        ops.conv_layer(x, wt, bias, stride, pad, r, bool(relu), algo=L.CONV_IGEMM)
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    ms = ts[len(ts) // 2]
    flops = 2 * b * p * q * cout * cin * k * k
    byts = 2 * (b * h * w * cin + b * p * q * cout * (2 if res else 1) + cout * cin * k * k)
    print(f"shape {sys.argv[1:11]} median {ms * 1e3:.1f} us  {flops / ms / 1e9:.1f} TFLOP/s  {byts / ms / 1e6:.0f} GB/s (cold L2)")


if __name__ == "__main__":
    main()
