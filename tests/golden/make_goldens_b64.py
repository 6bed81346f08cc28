"""Mint tests/golden/classify_resnet50_b64.npz: 64 ResNet-50 images, the micro-batch bench.py times
(BASELINE.json configs[1]), with the same recipe and selection rule as make_goldens.py.

Run from the repo root:  python tests/golden/make_goldens_b64.py
"""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import make_goldens as mg  # noqa: E402

if __name__ == "__main__":
    torch.set_num_threads(os.cpu_count() or 1)
    mg.mint_classify("resnet50", 32768, n_keep=64, tag="resnet50_b64")
