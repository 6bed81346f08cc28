"""Mint the golden fixtures under tests/golden/ from the CPU oracle.

Run from the repo root:  python tests/golden/make_goldens.py

The reference holds no golden vectors for this path (parity unpinned), so
these are minted from `oracle/` (torchvision fp32 CPU + the numpy preprocessing
restatement).  Weights and images are regenerated from seeds at test time; only indices,
logits, top-5 and small preprocessing vectors are stored.

Image selection: out of `N_CAND` structured synthetic images, keep those whose fp32 top-6
logits are separated by more than `MIN_GAP` between adjacent ranks, so that "identical
top-5" is a statement about the engine and not about coin-flip ties (a bf16-storage
engine carries ~1e-2 absolute logit error on this model: see `fmt_err_max` in each file).
"""
from __future__ import annotations

import os
import sys
import time

import numpy as np
import torch
import torchvision

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import preprocess_oracle as pp  # noqa: E402
from oracle import resnet_oracle as ro  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
MIN_GAP = 0.022
N_KEEP = 32


def mint_classify(arch: str, n_cand: int, hw: int = 224, in_ch: int = 3, n_keep: int = N_KEEP,
                  tag: str | None = None, seed_images: int = 1234):
    t0 = time.time()
    model = ro.build_model(arch, in_ch=in_ch, calib_hw=hw)
    mean = ro.MEAN4 if in_ch == 4 else ro.IMAGENET_MEAN
    std = ro.STD4 if in_ch == 4 else ro.IMAGENET_STD
    keep_idx, keep_logits = [], []
    bs = 32
    for start in range(0, n_cand, bs):
        idx = list(range(start, min(start + bs, n_cand)))
        imgs = ro.synthetic_images(idx, hw, hw, in_ch, seed=seed_images)
        y = ro.forward_fp32(model, pp.preprocess_nchw(imgs, mean, std))
        srt = -np.sort(-y, axis=1)[:, :6]
        gap = (srt[:, :-1] - srt[:, 1:]).min(axis=1)
        for j, i in enumerate(idx):
            if gap[j] > MIN_GAP:
                keep_idx.append(i)
                keep_logits.append(y[j])
        if len(keep_idx) >= n_keep:
            break
    keep_idx = keep_idx[:n_keep]
    logits = np.stack(keep_logits[:n_keep]).astype(np.float32)
    imgs = ro.synthetic_images(keep_idx, hw, hw, in_ch, seed=seed_images)
    x = pp.preprocess_nchw(imgs, mean, std)
    emul, stages = ro.forward_bf16_emul(model, x, return_stages=True)
    top_idx, top_prob = ro.softmax_top5(logits)
    e_idx, _ = ro.softmax_top5(emul)
    d = np.abs(emul - logits)
    name = tag or arch
    path = os.path.join(OUT, f"classify_{name}.npz")
    np.savez_compressed(
        path,
        arch=arch, hw=hw, in_ch=in_ch, image_seed=seed_images, indices=np.array(keep_idx, np.int64),
        logits_fp32=logits, logits_bf16_emul=emul.astype(np.float32),
        top5_idx=top_idx, top5_prob=top_prob,
        fmt_err_max=np.float32(d.max()), fmt_err_mean=np.float32(d.mean()),
        # cheap per-stage fingerprints of the bf16-emulated forward (sum and abs-sum, fp64)
        stage_names=np.array(list(stages.keys())),
        stage_sum=np.array([float(v.astype(np.float64).sum()) for v in stages.values()]),
        stage_abssum=np.array([float(np.abs(v.astype(np.float64)).sum()) for v in stages.values()]),
        torch_version=torch.__version__, torchvision_version=torchvision.__version__,
        min_gap=np.float32(MIN_GAP), residual_gain=np.float32(ro.RESIDUAL_GAIN),
    )
    print(f"{name}: kept {len(keep_idx)} of {start + bs} candidates; fmt err max {d.max():.4f} "
          f"mean {d.mean():.5f}; emul top5 == fp32 top5 on {(e_idx == top_idx).all(1).sum()}/{len(keep_idx)}; "
          f"{time.time() - t0:.1f}s -> {path}")


def mint_preprocess():
    rng = np.random.default_rng(7)
    cases = {}
    # plain normalise, ragged sizes (not multiples of 16 pixels), 3 and 4 bands
    for name, (n, h, w, c) in {"a": (2, 37, 53, 3), "b": (1, 224, 224, 3), "c": (3, 16, 24, 4),
                               "d": (1, 1, 1, 3)}.items():
        img = rng.integers(0, 256, (n, h, w, c), dtype=np.uint8)
        mean = ro.MEAN4 if c == 4 else ro.IMAGENET_MEAN
        std = ro.STD4 if c == 4 else ro.IMAGENET_STD
        out = pp.bf16_bits(pp.preprocess_nchw(img, mean, std))
        cases[f"plain_{name}_in"] = img
        cases[f"plain_{name}_out_nchw_bf16"] = out
    # all 256 byte values through every channel: the complete truth table of the LUT
    ramp = np.tile(np.arange(256, dtype=np.uint8)[None, None, :, None], (1, 1, 1, 4))
    cases["ramp_in"] = ramp
    cases["ramp_out_nchw_bf16"] = pp.bf16_bits(pp.preprocess_nchw(ramp, ro.MEAN4, ro.STD4))
    # resize + centre crop
    img = rng.integers(0, 256, (2, 40, 56, 3), dtype=np.uint8)
    cases["resize_in"] = img
    cases["resize_params"] = np.array([32, 44, 28, 28], np.int32)  # resize_h, resize_w, crop_h, crop_w
    cases["resize_out_nchw_bf16"] = pp.bf16_bits(
        pp.preprocess_resize_crop_nchw(img, 32, 44, 28, 28, ro.IMAGENET_MEAN, ro.IMAGENET_STD))
    img = ro.synthetic_images(1, 256, 256, 3, seed=5)
    cases["resize256_in_seed"] = np.array([5], np.int64)
    cases["resize256_out_nchw_bf16"] = pp.bf16_bits(
        pp.preprocess_resize_crop_nchw(img, 256, 256, 224, 224, ro.IMAGENET_MEAN, ro.IMAGENET_STD))
    path = os.path.join(OUT, "preprocess.npz")
    np.savez_compressed(path, **cases)
    print("preprocess ->", path)


if __name__ == "__main__":
    torch.set_num_threads(os.cpu_count() or 1)
    mint_preprocess()
    mint_classify("resnet18", 8192)
    mint_classify("resnet50", 8192)
    # land-cover config: 256x256x4 chips, ResNet-50 with a 4-channel stem
    mint_classify("resnet50", 4096, hw=256, in_ch=4, n_keep=8, tag="resnet50_landcover4", seed_images=4321)
