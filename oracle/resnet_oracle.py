"""CPU oracle for the classify hot path (TEST INFRASTRUCTURE — never shipped, never timed
as the product).

Only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference`
legs may import this module.  The product path (aiforearth-api-platform_b200/) must never
import anything under `oracle/`.

What it restates
----------------
The reference platform contains no arithmetic of its own: the model call is the opaque
user callback `r = func(*args, **kwargs)` at
`/root/reference/APIs/1.0/base-py/ai4e_service.py:203`, running whatever deep-learning
framework the user's base image ships (`Containers/Dockerfile:1`, unpinned
`mcr.microsoft.com/aiforearth/base-py:latest`).  BASELINE.json nominates "the reference
container's own PyTorch CPU path", so the oracle is:

  * third-party dependency: torchvision's `resnet18` / `resnet50` (ResNet v1.5: stride 2
    on the 3x3 conv of a down-sampling bottleneck), fp32, eager, `eval()`, `no_grad()`;
    versions in this image: torch 2.11.0+cu128, torchvision 0.26.0+cu128 (the reference
    pins none);
  * preprocessing: `oracle/preprocess_oracle.py` (ToTensor -> Normalize restated in numpy).

PARITY UNPINNED BY THE REFERENCE: the reference holds no tests, golden vectors or
known-answer fixtures for this path.  The goldens under
`tests/golden/` are minted by `tests/golden/make_goldens.py` from this oracle.

Two oracles are provided:

  `forward_fp32`      the nominated reference: torchvision fp32 forward, BN unfolded.
  `forward_bf16_emul` a restatement of what a bf16-storage engine computes: BN folded in
                      fp32 then weights rounded to bf16, every layer output rounded to
                      bf16, accumulation in fp32.  It bounds the error that is due to the
                      bf16 storage format alone, so kernel bugs can be told apart from
                      format error.

Synthetic weights (no pretrained checkpoints exist offline)
-----------------------------------------------------------
`torch.manual_seed(seed)` + `resnetNN(weights=None)`; BN affine parameters are drawn as
SURVEY.md 8(d) says (weight ~ U(0.5,1.5), bias ~ N(0,0.1), generator seed 1), the last BN
of every residual block then scaled by RESIDUAL_GAIN = 0.1.  BN running statistics are
*calibrated*: one pass of 16 structured synthetic images sets them to the
statistics of each layer's input, the way a trained network's statistics match its data.
SURVEY's literal recipe (running_mean ~ N(0,0.1), running_var ~ U(0.5,1.5)) was probed and
rejected: it yields input-independent logits of magnitude ~1e3 whose top-5 is the same for
every image, which tests nothing.  See DESIGN.md "Synthetic model".
"""
from __future__ import annotations

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F
import torchvision

from . import preprocess_oracle as pp

IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)
# 4-band land-cover config: 4th entry (0.5, 0.25)
MEAN4 = IMAGENET_MEAN + (0.5,)
STD4 = IMAGENET_STD + (0.25,)


def synthetic_image(index: int, h: int = 224, w: int = 224, c: int = 3, seed: int = 1234,
                    structured: bool = True) -> np.ndarray:
    """One uint8 HWC synthetic image, a pure function of (seed, index, h, w, c).

    structured=False is SURVEY 8(d)'s uniform-noise generator: fine for throughput, useless
    for parity because every noise image looks alike to a CNN.  structured=True adds
    per-image low-frequency content so that logits depend on the image."""
    rng = np.random.default_rng([seed, index])
    if not structured:
        return rng.integers(0, 256, (h, w, c), dtype=np.uint8)
    yy = np.linspace(0.0, 1.0, h, dtype=np.float32)[:, None]
    xx = np.linspace(0.0, 1.0, w, dtype=np.float32)[None, :]
    img = np.zeros((h, w, c), dtype=np.float32)
    for ch in range(c):
        acc = np.zeros((h, w), dtype=np.float32)
        for _ in range(4):
            fy, fx = rng.uniform(0.5, 6.0, 2)
            ph = rng.uniform(0, 2 * np.pi)
            amp = rng.uniform(0.2, 1.0)
            acc += (amp * np.cos(2 * np.pi * (fy * yy + fx * xx) + ph)).astype(np.float32)
        base = rng.uniform(0.2, 0.8)
        img[:, :, ch] = base + 0.18 * acc
    img += rng.normal(0.0, 0.04, (h, w, c)).astype(np.float32)
    return np.clip(np.rint(img * 255.0), 0, 255).astype(np.uint8)


def synthetic_images(n_or_indices, h: int = 224, w: int = 224, c: int = 3, seed: int = 1234,
                     structured: bool = True) -> np.ndarray:
    """[N,H,W,C] uint8; `n_or_indices` is a count (indices 0..n-1) or an index list."""
    idx = range(n_or_indices) if isinstance(n_or_indices, (int, np.integer)) else n_or_indices
    return np.stack([synthetic_image(int(i), h, w, c, seed, structured) for i in idx])


def _widen_stem(model: nn.Module, in_ch: int, seed: int) -> None:
    if in_ch == 3:
        return
    old = model.conv1
    g = torch.Generator().manual_seed(seed + 77)
    conv = nn.Conv2d(in_ch, old.out_channels, kernel_size=7, stride=2, padding=3, bias=False)
    w = torch.empty_like(conv.weight)
    std = (2.0 / (old.out_channels * 49)) ** 0.5  # kaiming_normal_, fan_out, relu
    w.normal_(0.0, std, generator=g)
    conv.weight.data.copy_(w)
    model.conv1 = conv


RESIDUAL_GAIN = 0.1


def build_model(arch: str = "resnet50", seed: int = 0, bn_seed: int = 1, in_ch: int = 3,
                calib_hw: int = 224, residual_gain: float = RESIDUAL_GAIN) -> nn.Module:
    """Seeded synthetic torchvision ResNet in eval mode (see module docstring)."""
    torch.manual_seed(seed)
    model = getattr(torchvision.models, arch)(weights=None)
    _widen_stem(model, in_ch, seed)
    g = torch.Generator().manual_seed(bn_seed)
    bns = [m for m in model.modules() if isinstance(m, nn.BatchNorm2d)]
    for m in bns:
        m.weight.data.copy_(torch.rand(m.weight.shape, generator=g) + 0.5)
        m.bias.data.copy_(torch.randn(m.bias.shape, generator=g) * 0.1)
    # trained residual nets have small residual-branch gain (torchvision's zero_init_residual
    # sets the last BN weight of a block to 0); a gain of 1 gives a chaotic random net in
    # which a 1e-7 relative change of summation order moves logits by 5e-2 (probe)
    for layer in (model.layer1, model.layer2, model.layer3, model.layer4):
        for blk in layer:
            last = blk.bn3 if hasattr(blk, "bn3") else blk.bn2
            last.weight.data.mul_(residual_gain)
    # calibrate running statistics on structured synthetic data
    mean = MEAN4[:in_ch] if in_ch == 4 else IMAGENET_MEAN
    std = STD4[:in_ch] if in_ch == 4 else IMAGENET_STD
    imgs = synthetic_images(16, calib_hw, calib_hw, in_ch, seed=999)
    x = torch.from_numpy(pp.preprocess_nchw(imgs, mean, std))
    for m in bns:
        m.momentum = 1.0
    model.train()
    with torch.no_grad():
        model(x)
    model.eval()
    for m in bns:
        m.momentum = 0.1
    for p in model.parameters():
        p.requires_grad_(False)
    return model


def forward_fp32(model: nn.Module, x_nchw: np.ndarray) -> np.ndarray:
    """The nominated reference: fp32 torchvision forward.  x is fp32 NCHW (normalised)."""
    with torch.no_grad():
        return model(torch.from_numpy(np.ascontiguousarray(x_nchw))).numpy()


def softmax_top5(logits: np.ndarray, k: int = 5):
    """softmax then top-k; ties resolve to the lower class index (stable)."""
    z = logits.astype(np.float64)
    z = z - z.max(axis=1, keepdims=True)
    p = np.exp(z)
    p /= p.sum(axis=1, keepdims=True)
    order = np.argsort(-logits, axis=1, kind="stable")[:, :k]
    return order.astype(np.int32), np.take_along_axis(p, order, axis=1).astype(np.float32)


# ----------------------------------------------------------------------------------------
# bf16-storage emulation
# ----------------------------------------------------------------------------------------

def bf16_round(t: torch.Tensor) -> torch.Tensor:
    return t.to(torch.bfloat16).to(torch.float32)


def fold_bn(conv: nn.Conv2d, bn: nn.BatchNorm2d):
    """(w_folded fp32 OIHW, bias fp32).  fold in fp32; caller rounds w to bf16."""
    scale = bn.weight / torch.sqrt(bn.running_var + bn.eps)
    w = conv.weight * scale[:, None, None, None]
    b = bn.bias - bn.running_mean * scale
    if conv.bias is not None:
        b = b + conv.bias * scale
    return w, b


def _conv_bf16(x, conv, bn, relu, residual=None):
    w, b = fold_bn(conv, bn)
    y = F.conv2d(x, bf16_round(w), None, conv.stride, conv.padding)
    y = y + b[None, :, None, None]
    if residual is not None:
        y = y + residual
    if relu:
        y = F.relu(y)
    return bf16_round(y)


def forward_bf16_emul(model: nn.Module, x_nchw: np.ndarray, return_stages: bool = False):
    """What an engine that stores activations/weights in bf16 and accumulates in fp32
    computes, layer by layer (input rounded to bf16 first).  Returns fp32 logits."""
    stages = {}
    with torch.no_grad():
        x = bf16_round(torch.from_numpy(np.ascontiguousarray(x_nchw)))
        stages["input"] = x
        x = _conv_bf16(x, model.conv1, model.bn1, relu=True)
        stages["stem"] = x
        x = F.max_pool2d(x, 3, 2, 1)
        stages["pool"] = x
        for li, layer in enumerate([model.layer1, model.layer2, model.layer3, model.layer4]):
            for blk in layer:
                identity = x
                if blk.downsample is not None:
                    identity = _conv_bf16(x, blk.downsample[0], blk.downsample[1], relu=False)
                if hasattr(blk, "conv3"):
                    t = _conv_bf16(x, blk.conv1, blk.bn1, relu=True)
                    t = _conv_bf16(t, blk.conv2, blk.bn2, relu=True)
                    x = _conv_bf16(t, blk.conv3, blk.bn3, relu=True, residual=identity)
                else:
                    t = _conv_bf16(x, blk.conv1, blk.bn1, relu=True)
                    x = _conv_bf16(t, blk.conv2, blk.bn2, relu=True, residual=identity)
            stages[f"layer{li + 1}"] = x
        pooled = x.mean(dim=(2, 3))  # fp32 average of bf16 values
        stages["pooled"] = pooled
        logits = pooled @ bf16_round(model.fc.weight).t() + model.fc.bias
        stages["logits"] = logits
    if return_stages:
        return logits.numpy(), {k: v.numpy() for k, v in stages.items()}
    return logits.numpy()


def classify_fp32(model: nn.Module, images_u8_hwc: np.ndarray, mean=None, std=None):
    """End-to-end reference path: preprocess -> fp32 forward -> softmax -> top-5."""
    c = images_u8_hwc.shape[-1]
    mean = mean if mean is not None else (MEAN4 if c == 4 else IMAGENET_MEAN)
    std = std if std is not None else (STD4 if c == 4 else IMAGENET_STD)
    x = pp.preprocess_nchw(images_u8_hwc, mean, std)
    logits = forward_fp32(model, x)
    idx, prob = softmax_top5(logits)
    return idx, prob, logits
