"""Turns a torchvision-style ResNet `state_dict` into the engine's layer program + weight blob.

Weight preparation happens once, at container start-up, on the host (it is not the hot path):
BatchNorm is folded into the preceding convolution in fp32, the folded weights are laid out
[Cout][kh][kw][Cin] and rounded to bf16, biases stay fp32.  torch is used only as the host
tensor container.  The arithmetic this reproduces is torchvision's `resnet.py` (BasicBlock /
Bottleneck, v1.5: stride on the 3x3 convolution), which is what a PyTorch CPU container runs
inside the callback the reference calls at APIs/1.0/base-py/ai4e_service.py:203.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field

import numpy as np
import torch

from . import _lib as L

ARCH_BLOCKS = {
    "resnet18": ("basic", [2, 2, 2, 2]),
    "resnet34": ("basic", [3, 4, 6, 3]),
    "resnet50": ("bottleneck", [3, 4, 6, 3]),
    "resnet101": ("bottleneck", [3, 4, 23, 3]),
    "resnet152": ("bottleneck", [3, 8, 36, 3]),
}
BN_EPS = 1e-5
IMAGENET_MEAN = (0.485, 0.456, 0.406)
IMAGENET_STD = (0.229, 0.224, 0.225)


@dataclass
class Layer:
    op: int
    in_buf: int
    out_buf: int
    res_buf: int = -1
    cin: int = 0
    cout: int = 0
    kh: int = 1
    kw: int = 1
    stride: int = 1
    pad: int = 0
    relu: int = 0
    w_off: int = 0
    b_off: int = 0
    name: str = ""


@dataclass
class Program:
    arch: str
    in_h: int
    in_w: int
    in_c: int
    in_c_pad: int
    num_classes: int
    mean: tuple
    std: tuple
    layers: list = field(default_factory=list)
    n_bufs: int = 0
    blob: bytes = b""
    stage_bufs: dict = field(default_factory=dict)  # stage name -> (slot, h, w, c) for parity tests
    keep_stages: bool = False  # every layer's output stays in memory (the engine then fuses nothing across layers' outputs)

    def to_ctypes(self):
        arr = (L.LayerDesc * len(self.layers))()
        for i, l in enumerate(self.layers):
            arr[i] = L.LayerDesc(l.op, l.in_buf, l.out_buf, l.res_buf, l.cin, l.cout, l.kh, l.kw, l.stride, l.pad,
                                 l.relu, 0, l.w_off, l.b_off)
        md = L.ModelDesc()
        md.n_layers = len(self.layers)
        md.n_bufs = self.n_bufs
        md.in_h, md.in_w, md.in_c, md.in_c_pad = self.in_h, self.in_w, self.in_c, self.in_c_pad
        md.num_classes = self.num_classes
        md.reserved = 1 if self.keep_stages else 0
        for i in range(4):
            md.mean[i] = float(self.mean[i]) if i < len(self.mean) else 0.0
            md.std[i] = float(self.std[i]) if i < len(self.std) else 1.0
        md.layers = C.cast(arr, C.POINTER(L.LayerDesc))
        return md, arr  # keep `arr` alive as long as `md` is used

    def flops_per_image(self) -> int:
        """algorithmic FLOPs (2 per MAC) of the convolutions + FC at this input size."""
        shapes = {0: (self.in_h, self.in_w)}
        total = 0
        for l in self.layers:
            h, w = shapes[l.in_buf]
            if l.op == L.OP_HEAD:
                total += 2 * l.cin * l.cout
                continue
            p = (h + 2 * l.pad - l.kh) // l.stride + 1
            q = (w + 2 * l.pad - l.kw) // l.stride + 1
            shapes[l.out_buf] = (p, q)
            if l.op == L.OP_CONV:
                cin = self.in_c if l.in_buf == 0 else l.cin  # the stem's padded channel does no work
                total += 2 * p * q * l.cout * cin * l.kh * l.kw
        return total


class _Blob:
    def __init__(self):
        self.parts = []
        self.size = 0

    def add(self, arr: np.ndarray) -> int:
        pad = (-self.size) % 256
        if pad:
            self.parts.append(b"\0" * pad)
            self.size += pad
        off = self.size
        b = np.ascontiguousarray(arr).tobytes()
        self.parts.append(b)
        self.size += len(b)
        return off

    def bytes(self) -> bytes:
        return b"".join(self.parts) + b"\0" * 256


def _bf16_bits(t: torch.Tensor) -> np.ndarray:
    return t.contiguous().to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)


def fold_conv_bn(sd: dict, conv: str, bn: str):
    """-> (weights [Cout][kh][kw][Cin] fp32 folded, bias [Cout] fp32)."""
    w = sd[conv + ".weight"].detach().to(torch.float32)
    gamma = sd[bn + ".weight"].detach().to(torch.float32)
    beta = sd[bn + ".bias"].detach().to(torch.float32)
    mu = sd[bn + ".running_mean"].detach().to(torch.float32)
    var = sd[bn + ".running_var"].detach().to(torch.float32)
    scale = gamma / torch.sqrt(var + BN_EPS)
    wf = w * scale[:, None, None, None]
    b = beta - mu * scale
    if conv + ".bias" in sd:
        b = b + sd[conv + ".bias"].detach().to(torch.float32) * scale
    return wf.permute(0, 2, 3, 1).contiguous(), b.contiguous()


def build_resnet_program(arch: str, state_dict: dict, in_h: int = 224, in_w: int = 224, mean=None,
                         std=None, keep_stages: bool = False) -> Program:
    """Layer program + weight blob for a torchvision-layout ResNet state_dict.

    keep_stages=True gives every stage output (stem, pool, layer1..4) a slot of its own that
    later layers never overwrite, so parity tests can read them back after a forward."""
    if arch not in ARCH_BLOCKS:
        raise ValueError(f"unknown arch {arch!r}; supported: {sorted(ARCH_BLOCKS)}")
    kind, counts = ARCH_BLOCKS[arch]
    sd = state_dict
    in_c = int(sd["conv1.weight"].shape[1])
    if in_c not in (3, 4):
        raise ValueError(f"stem takes {in_c} channels; 3 or 4 supported")
    if mean is None:
        mean = IMAGENET_MEAN if in_c == 3 else IMAGENET_MEAN + (0.5,)
    if std is None:
        std = IMAGENET_STD if in_c == 3 else IMAGENET_STD + (0.25,)
    num_classes = int(sd["fc.weight"].shape[0])
    prog = Program(arch, in_h, in_w, in_c, 4, num_classes, tuple(mean), tuple(std), keep_stages=keep_stages)
    blob = _Blob()
    slots_in_use = {0}
    pinned = set()

    def release(slot: int) -> None:
        if slot not in pinned:
            slots_in_use.discard(slot)

    def stage(name: str, slot: int, hh: int, ww: int, cc: int) -> None:
        prog.stage_bufs[name] = (slot, hh, ww, cc)
        if keep_stages:
            pinned.add(slot)

    def take() -> int:
        s = 1
        while s in slots_in_use:
            s += 1
        slots_in_use.add(s)
        prog.n_bufs = max(prog.n_bufs, s + 1)
        return s

    def conv(name, bn, src, cin_pad=None, stride=1, pad=0, relu=True, res=-1) -> int:
        wf, b = fold_conv_bn(sd, name, bn)
        if cin_pad and wf.shape[3] < cin_pad:  # zero weights for the stem's padding channel
            wf = torch.cat([wf, torch.zeros(*wf.shape[:3], cin_pad - wf.shape[3])], dim=3)
        cout, kh, kw, cin = wf.shape
        dst = take()
        prog.layers.append(Layer(L.OP_CONV, src, dst, res, cin, cout, kh, kw, stride, pad, int(relu),
                                 blob.add(_bf16_bits(wf)), blob.add(b.numpy().astype(np.float32)), name))
        return dst

    h, w = in_h, in_w
    x = conv("conv1", "bn1", 0, cin_pad=4, stride=2, pad=3)
    h, w = (h + 6 - 7) // 2 + 1, (w + 6 - 7) // 2 + 1
    stage("stem", x, h, w, 64)
    pooled = take()
    prog.layers.append(Layer(L.OP_MAXPOOL, x, pooled, -1, 64, 64, 3, 3, 2, 1, 0, name="maxpool"))
    release(x)
    x = pooled
    h, w = (h + 2 - 3) // 2 + 1, (w + 2 - 3) // 2 + 1
    stage("pool", x, h, w, 64)
    c = 64
    for li, n_blocks in enumerate(counts):
        for bi in range(n_blocks):
            pre = f"layer{li + 1}.{bi}"
            stride = 2 if (li > 0 and bi == 0) else 1
            identity = x
            ds = -1
            if pre + ".downsample.0.weight" in sd:
                ds = conv(pre + ".downsample.0", pre + ".downsample.1", x, stride=stride, pad=0, relu=False)
                identity = ds
            if kind == "bottleneck":
                t1 = conv(pre + ".conv1", pre + ".bn1", x)
                t2 = conv(pre + ".conv2", pre + ".bn2", t1, stride=stride, pad=1)
                # conv2 and conv3 may run as one kernel that reads t1 while it writes y (K4b): y must
                # not land in t1's slot, so t1 is released only after y has its own
                y = conv(pre + ".conv3", pre + ".bn3", t2, res=identity)
                release(t1)
                release(t2)
            else:
                t1 = conv(pre + ".conv1", pre + ".bn1", x, stride=stride, pad=1)
                y = conv(pre + ".conv2", pre + ".bn2", t1, pad=1, res=identity)
                release(t1)
            release(x)
            if ds >= 0:
                release(ds)
            x = y
            c = prog.layers[-1].cout
            h, w = (h + 2 - 3) // stride + 1, (w + 2 - 3) // stride + 1
        stage(f"layer{li + 1}", x, h, w, c)
    fc_w = sd["fc.weight"].detach().to(torch.float32)
    fc_b = sd["fc.bias"].detach().to(torch.float32)
    prog.layers.append(Layer(L.OP_HEAD, x, 0, -1, c, num_classes, 1, 1, 1, 0, 0, blob.add(_bf16_bits(fc_w)),
                             blob.add(fc_b.numpy().astype(np.float32)), "fc"))
    prog.blob = blob.bytes()
    return prog


def synthetic_state_dict(arch: str, seed: int = 0, in_c: int = 3, num_classes: int = 1000,
                         residual_gain: float = 0.1) -> dict:
    """Seeded random weights with torchvision's parameter names and shapes (no checkpoints can be
    downloaded offline).  Enough for throughput runs and smoke tests: kernel time does not depend
    on the values.  Kaiming-normal convolutions, BN weight ~ U(0.5,1.5) (x residual_gain on the
    last BN of a block), bias ~ N(0,0.1), running_mean ~ N(0,0.1), running_var ~ U(0.5,1.5)."""
    kind, counts = ARCH_BLOCKS[arch]
    g = torch.Generator().manual_seed(seed)
    sd = {}

    def conv(name, cout, cin, k):
        std = (2.0 / (cout * k * k)) ** 0.5
        sd[name + ".weight"] = torch.randn((cout, cin, k, k), generator=g) * std

    def bn(name, c, gain=1.0):
        sd[name + ".weight"] = (torch.rand(c, generator=g) + 0.5) * gain
        sd[name + ".bias"] = torch.randn(c, generator=g) * 0.1
        sd[name + ".running_mean"] = torch.randn(c, generator=g) * 0.1
        sd[name + ".running_var"] = torch.rand(c, generator=g) + 0.5

    conv("conv1", 64, in_c, 7)
    bn("bn1", 64)
    exp = 4 if kind == "bottleneck" else 1
    cin = 64
    for li, n_blocks in enumerate(counts):
        width = 64 * 2 ** li
        for bi in range(n_blocks):
            pre = f"layer{li + 1}.{bi}"
            stride = 2 if (li > 0 and bi == 0) else 1
            if kind == "bottleneck":
                conv(pre + ".conv1", width, cin, 1); bn(pre + ".bn1", width)
                conv(pre + ".conv2", width, width, 3); bn(pre + ".bn2", width)
                conv(pre + ".conv3", width * 4, width, 1); bn(pre + ".bn3", width * 4, residual_gain)
            else:
                conv(pre + ".conv1", width, cin, 3); bn(pre + ".bn1", width)
                conv(pre + ".conv2", width, width, 3); bn(pre + ".bn2", width, residual_gain)
            if stride != 1 or cin != width * exp:
                conv(pre + ".downsample.0", width * exp, cin, 1); bn(pre + ".downsample.1", width * exp)
            cin = width * exp
    bound = 1.0 / cin ** 0.5
    sd["fc.weight"] = (torch.rand((num_classes, cin), generator=g) * 2 - 1) * bound
    sd["fc.bias"] = (torch.rand(num_classes, generator=g) * 2 - 1) * bound
    return sd
