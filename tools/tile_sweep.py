#!/usr/bin/env python
"""Time every tiling of the implicit-GEMM kernel on the distinct ResNet-50 convolution shapes
(batch 64) through the C ABI, cold L2.  python tools/tile_sweep.py [batch]"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "aiforearth-api-platform_b200")]
import torch  # noqa: E402

from ai4e_b200 import ops, _lib as L  # noqa: E402

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
# (name, h, w, cin, cout, k, stride, pad, relu, res, count in ResNet-50)
SHAPES = [
    ("l1.ds", 56, 56, 64, 256, 1, 1, 0, 0, 0, 1), ("l1.c1a", 56, 56, 64, 64, 1, 1, 0, 1, 0, 1),
    ("l1.c2", 56, 56, 64, 64, 3, 1, 1, 1, 0, 3), ("l1.c3", 56, 56, 64, 256, 1, 1, 0, 1, 1, 3),
    ("l1.c1", 56, 56, 256, 64, 1, 1, 0, 1, 0, 2), ("l2.ds", 56, 56, 256, 512, 1, 2, 0, 0, 0, 1),
    ("l2.0.c1", 56, 56, 256, 128, 1, 1, 0, 1, 0, 1), ("l2.0.c2", 56, 56, 128, 128, 3, 2, 1, 1, 0, 1),
    ("l2.c3", 28, 28, 128, 512, 1, 1, 0, 1, 1, 4), ("l2.c1", 28, 28, 512, 128, 1, 1, 0, 1, 0, 3),
    ("l2.c2", 28, 28, 128, 128, 3, 1, 1, 1, 0, 3), ("l3.ds", 28, 28, 512, 1024, 1, 2, 0, 0, 0, 1),
    ("l3.0.c1", 28, 28, 512, 256, 1, 1, 0, 1, 0, 1), ("l3.0.c2", 28, 28, 256, 256, 3, 2, 1, 1, 0, 1),
    ("l3.c3", 14, 14, 256, 1024, 1, 1, 0, 1, 1, 6), ("l3.c1", 14, 14, 1024, 256, 1, 1, 0, 1, 0, 5),
    ("l3.c2", 14, 14, 256, 256, 3, 1, 1, 1, 0, 5), ("l4.ds", 14, 14, 1024, 2048, 1, 2, 0, 0, 0, 1),
    ("l4.0.c1", 14, 14, 1024, 512, 1, 1, 0, 1, 0, 1), ("l4.0.c2", 14, 14, 512, 512, 3, 2, 1, 1, 0, 1),
    ("l4.c3", 7, 7, 512, 2048, 1, 1, 0, 1, 1, 3), ("l4.c1", 7, 7, 2048, 512, 1, 1, 0, 1, 0, 2),
    ("l4.c2", 7, 7, 512, 512, 3, 1, 1, 1, 0, 2),
]
TILINGS = ["64", "128", "256", "128p", "256p"]


def main():
    g = torch.Generator().manual_seed(0)
    flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")
    total_auto = total_best = 0.0
    print(f"{'shape':9s} " + " ".join(f"{t:>7s}" for t in ["auto"] + TILINGS) + "   best")
    for name, h, w, cin, cout, k, stride, pad, relu, res, cnt in SHAPES:
        x = (torch.randn((B, h, w, cin), generator=g) * 0.5).to(torch.bfloat16).cuda()
        wt = (torch.randn((cout, k, k, cin), generator=g) * 0.05).to(torch.bfloat16).cuda()
        bias = torch.randn(cout, generator=g).cuda()
        p, q = (h + 2 * pad - k) // stride + 1, (w + 2 * pad - k) // stride + 1
        r = (torch.randn((B, p, q, cout), generator=g) * 0.5).to(torch.bfloat16).cuda() if res else None
        row = {}
        for t in ["auto"] + TILINGS:
            if t == "auto":
                os.environ.pop("AI4E_B200_TILE", None)
            else:
                bn = int(t.rstrip("p"))
                if cout % bn or (res and bn > 128 and not t.endswith("p") and False):
                    row[t] = None
                    continue
                os.environ["AI4E_B200_TILE"] = t
            try:
                for _ in range(2):
                    ops.conv_layer(x, wt, bias, stride, pad, r, bool(relu), algo=L.CONV_IGEMM)
                torch.cuda.synchronize()
                ts = []
                for _ in range(9):
                    flush.zero_()
                    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                    e0.record()
                    ops.conv_layer(x, wt, bias, stride, pad, r, bool(relu), algo=L.CONV_IGEMM)
                    e1.record()
                    torch.cuda.synchronize()
                    ts.append(e0.elapsed_time(e1) * 1e3)
                ts.sort()
                row[t] = ts[len(ts) // 2]
            except Exception as e:  # tiling not available for this shape
                row[t] = None
        os.environ.pop("AI4E_B200_TILE", None)
        best = min((v, t) for t, v in row.items() if v is not None and t != "auto")
        total_auto += cnt * row["auto"]
        total_best += cnt * best[0]
        print(f"{name:9s} " + " ".join(f"{row[t]:7.1f}" if row[t] is not None else "      -" for t in ["auto"] + TILINGS)
              + f"   {best[1]} x{cnt}")
    print(f"sum over ResNet-50 (x count): auto {total_auto:.1f} us, best-of {total_best:.1f} us")


if __name__ == "__main__":
    main()
