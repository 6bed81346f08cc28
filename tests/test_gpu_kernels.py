"""GPU parity tests of the stateless C-ABI kernels against the CPU oracle.

Integer/byte work (K1 preprocessing) is bit-exact; floating-point kernels carry the
tolerance written beside each assert."""
import numpy as np
import pytest
import torch

from oracle import preprocess_oracle as pp
from oracle import resnet_oracle as ro

pytestmark = pytest.mark.gpu


def _bits(t: torch.Tensor) -> np.ndarray:
    return t.contiguous().view(torch.int16).cpu().numpy().view(np.uint16)


def _dev(a: np.ndarray, dtype=None):
    t = torch.from_numpy(np.ascontiguousarray(a)).cuda()
    return t if dtype is None else t.to(dtype)


# ------------------------------------------------------------------ K1 preprocess
@pytest.mark.parametrize("case", ["a", "b", "c", "d"])
def test_preprocess_golden_nchw(golden_dir, case):
    from ai4e_b200 import ops
    g = np.load(f"{golden_dir}/preprocess.npz")
    img = g[f"plain_{case}_in"]
    c = img.shape[-1]
    mean, std = (ro.MEAN4, ro.STD4) if c == 4 else (ro.IMAGENET_MEAN, ro.IMAGENET_STD)
    out = ops.preprocess(_dev(img), mean, std, layout="nchw")
    torch.cuda.synchronize()
    assert np.array_equal(_bits(out), g[f"plain_{case}_out_nchw_bf16"])  # bit-exact


def test_preprocess_truth_table(golden_dir):
    """all 256 byte values x 4 channels: the complete truth table of the LUT."""
    from ai4e_b200 import ops
    g = np.load(f"{golden_dir}/preprocess.npz")
    out = ops.preprocess(_dev(g["ramp_in"]), ro.MEAN4, ro.STD4, layout="nchw")
    assert np.array_equal(_bits(out), g["ramp_out_nchw_bf16"])


@pytest.mark.parametrize("shape", [(3, 224, 224, 3), (2, 256, 256, 4), (5, 48, 80, 3), (1, 7, 9, 3)])
@pytest.mark.parametrize("layout", ["nchw", "nhwc"])
def test_preprocess_vs_oracle(shape, layout):
    from ai4e_b200 import ops
    rng = np.random.default_rng(11)
    img = rng.integers(0, 256, shape, dtype=np.uint8)
    c = shape[-1]
    mean, std = (ro.MEAN4, ro.STD4) if c == 4 else (ro.IMAGENET_MEAN, ro.IMAGENET_STD)
    out = ops.preprocess(_dev(img), mean, std, layout=layout, out_channels=4)
    if layout == "nchw":
        want = pp.bf16_bits(pp.preprocess_nchw(img, mean, std))
    else:
        x = pp.preprocess_nhwc(img, mean, std)
        if c == 3:
            x = np.concatenate([x, np.zeros(shape[:3] + (1,), np.float32)], axis=-1)
        want = pp.bf16_bits(x)
    assert np.array_equal(_bits(out), want)  # bit-exact


def test_preprocess_resize_crop_golden(golden_dir):
    from ai4e_b200 import ops
    g = np.load(f"{golden_dir}/preprocess.npz")
    rh, rw, ch, cw = [int(v) for v in g["resize_params"]]
    out = ops.preprocess(_dev(g["resize_in"]), ro.IMAGENET_MEAN, ro.IMAGENET_STD, layout="nchw",
                         resize=(rh, rw), crop=(ch, cw))
    assert np.array_equal(_bits(out), g["resize_out_nchw_bf16"])  # bit-exact (same fp32 op order)
    img = ro.synthetic_images(1, 256, 256, 3, seed=int(g["resize256_in_seed"][0]))
    out = ops.preprocess(_dev(img), ro.IMAGENET_MEAN, ro.IMAGENET_STD, layout="nchw", resize=(256, 256),
                         crop=(224, 224))
    assert np.array_equal(_bits(out), g["resize256_out_nchw_bf16"])


def test_preprocess_strided_rows():
    """a chip cut out of a larger tile: row stride > w*c (land-cover config)."""
    from ai4e_b200 import ops
    rng = np.random.default_rng(3)
    tile = rng.integers(0, 256, (1, 64, 96, 4), dtype=np.uint8)
    d = _dev(tile)
    chip = d[:, 16:48, 32:64, :]
    out = ops.preprocess(chip, ro.MEAN4, ro.STD4, layout="nchw")
    want = pp.bf16_bits(pp.preprocess_nchw(np.ascontiguousarray(tile[:, 16:48, 32:64, :]), ro.MEAN4, ro.STD4))
    assert np.array_equal(_bits(out), want)


def test_preprocess_rejects_bad_args():
    from ai4e_b200 import ops, Ai4eError
    img = torch.zeros((1, 8, 8, 2), dtype=torch.uint8, device="cuda")
    with pytest.raises(Ai4eError):
        ops.preprocess(img, (0.5, 0.5), (0.2, 0.2))
    img = torch.zeros((1, 8, 8, 3), dtype=torch.uint8, device="cuda")
    with pytest.raises(Ai4eError):
        ops.preprocess(img, ro.IMAGENET_MEAN, (0.2, 0.0, 0.2))


# ------------------------------------------------------------------ convolutions
def _conv_case(b, h, w, cin, cout, k, stride, pad, relu, residual, seed=0):
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn((b, h, w, cin), generator=g) * 0.7).to(torch.bfloat16)
    wt = (torch.randn((cout, k, k, cin), generator=g) * (2.0 / (k * k * cin)) ** 0.5).to(torch.bfloat16)
    bias = torch.randn((cout,), generator=g) * 0.1
    p = (h + 2 * pad - k) // stride + 1
    q = (w + 2 * pad - k) // stride + 1
    res = (torch.randn((b, p, q, cout), generator=g) * 0.5).to(torch.bfloat16) if residual else None
    y = torch.nn.functional.conv2d(x.float().permute(0, 3, 1, 2), wt.float().permute(0, 3, 1, 2), bias,
                                   stride, pad).permute(0, 2, 3, 1)
    if res is not None:
        y = y + res.float()
    if relu:
        y = torch.relu(y)
    return x, wt, bias, res, y.contiguous()


def _check_conv(out: torch.Tensor, want: torch.Tensor):
    got = out.float().cpu()
    # fp32 accumulation, one bf16 rounding of the result: |err| <= 2^-8 |want| + accumulation noise
    tol = 2.0 ** -7 * want.abs() + 2e-3
    bad = (got - want).abs() > tol
    assert not bad.any(), f"{int(bad.sum())} of {bad.numel()} outside tolerance; max err {(got - want).abs().max():.4f}"


DIRECT_CASES = [
    (2, 32, 32, 4, 64, 7, 2, 3, True, False),    # stem shape, small image
    (1, 224, 224, 4, 64, 7, 2, 3, True, False),  # stem at full size
    (2, 9, 11, 16, 8, 3, 1, 1, False, True),
]


@pytest.mark.parametrize("case", DIRECT_CASES)
def test_conv_direct_vs_oracle(case):
    from ai4e_b200 import ops, _lib as L
    b, h, w, cin, cout, k, stride, pad, relu, residual = case
    x, wt, bias, res, want = _conv_case(*case)
    out = ops.conv_layer(x.cuda(), wt.cuda(), bias.cuda(), stride, pad, res.cuda() if res is not None else None,
                         relu, algo=L.CONV_DIRECT)
    torch.cuda.synchronize()
    _check_conv(out, want)


STEM_CASES = [
    (2, 224, 224, 4, 64, 7, 2, 3, True, False),   # the stem at full size (M = 25088, 196 tiles)
    (3, 32, 32, 4, 64, 7, 2, 3, True, False),     # small: most taps touch the zero padding
    (1, 256, 256, 4, 64, 7, 2, 3, False, False),  # land-cover chip size, no ReLU
    (2, 30, 50, 4, 64, 7, 2, 3, True, False),     # ragged: M = 2*15*25 = 750, rows wrap mid-tile
    (1, 8, 8, 4, 64, 7, 2, 3, True, False),       # tiny: M = 16
    (2, 40, 64, 4, 64, 7, 2, 3, True, False),     # W % 16 == 0: row-buffer mode, 32 outputs per row
    (1, 18, 16, 4, 64, 7, 2, 3, False, False),    # row mode, 8 outputs per row, rows mostly padding
    (1, 64, 240, 4, 64, 7, 2, 3, True, False),    # row mode, 120 outputs per row
]


@pytest.mark.parametrize("case", STEM_CASES, ids=lambda c: "x".join(str(v) for v in c[:3]))
def test_conv_stem_tensor_core_vs_oracle(case):
    """7x7/2 stem on tcgen05 through super-pixel im2col loads (AUTO picks it for Cin = 4)."""
    from ai4e_b200 import ops, _lib as L
    b, h, w, cin, cout, k, stride, pad, relu, residual = case
    x, wt, bias, res, want = _conv_case(*case, seed=9)
    out = ops.conv_layer(x.cuda(), wt.cuda(), bias.cuda(), stride, pad, None, relu, algo=L.CONV_AUTO)
    torch.cuda.synchronize()
    _check_conv(out, want)
    direct = ops.conv_layer(x.cuda(), wt.cuda(), bias.cuda(), stride, pad, None, relu, algo=L.CONV_DIRECT)
    assert (out.float() - direct.float()).abs().max() <= 2.0 ** -6 * want.abs().max()


STEM_POOL_CASES = [
    (2, 224, 224, True),    # the network's stem: 112 x 112 -> 56 x 56, 7 runs of 8 pooled rows per image
    (3, 256, 256, True),    # land-cover chips: 128 output columns fill all four TMEM lane quadrants
    (1, 224, 224, False),   # no ReLU: negative values, -inf padding must not leak
    (5, 32, 64, True),      # small: pooled 8 x 16, one run per image, most taps in the padding
    (2, 40, 48, True),      # pooled 10 x 12: the second run of an image is ragged (2 rows) and starts with a seed tile
    (150, 64, 32, True),    # more runs than SMs: CTAs take several runs, accumulator stages wrap many times
]


@pytest.mark.parametrize("case", STEM_POOL_CASES, ids=lambda c: "x".join(str(v) for v in c))
def test_stem_pool_fused_is_bit_identical_to_stem_then_maxpool(case):
    """K2 + K3 in one kernel == ai4e_conv_layer (tensor-core stem) followed by ai4e_maxpool, bit for bit,
    and within one bf16 rounding of conv2d + relu + max_pool2d in fp32."""
    from ai4e_b200 import ops, _lib as L
    b, h, w, relu = case
    g = torch.Generator().manual_seed(h * 1000 + w + b)
    x = (torch.randn((b, h, w, 4), generator=g) * 0.8).to(torch.bfloat16)
    x[..., 3] = 0
    wt = (torch.randn((64, 7, 7, 4), generator=g) * 0.08).to(torch.bfloat16)
    bias = torch.randn(64, generator=g) * 0.3
    stem = ops.conv_layer(x.cuda(), wt.cuda(), bias.cuda(), 2, 3, None, relu, algo=L.CONV_AUTO)
    want = ops.maxpool(stem)
    got = ops.stem_pool(x.cuda(), wt.cuda(), bias.cuda(), relu)
    torch.cuda.synchronize()
    assert got.shape == want.shape == (b, h // 4, w // 4, 64)
    assert torch.equal(got, want), f"{(got != want).sum().item()} of {got.numel()} values differ"
    y = torch.nn.functional.conv2d(x.float().permute(0, 3, 1, 2), wt.float().permute(0, 3, 1, 2), bias, stride=2, padding=3)
    if relu:
        y = torch.relu(y)
    ref = torch.nn.functional.max_pool2d(y, 3, 2, 1).permute(0, 2, 3, 1)
    err = (got.float().cpu() - ref).abs()
    assert (err <= 2.0 ** -7 * ref.abs() + 2e-3).all(), err.max()


@pytest.mark.parametrize("b,h,w,relu", [(3, 224, 224, True), (2, 256, 256, True), (70, 32, 64, False), (1, 12, 16, True)])
def test_preprocess_stem_pool_is_bit_identical_to_the_three_kernels(b, h, w, relu):
    """K1 + K2 + K3 in one kernel (uint8 in, pooled bf16 out) == ai4e_preprocess (NHWC4) -> ai4e_stem_pool."""
    from ai4e_b200 import ops
    rng = np.random.default_rng(b * 7 + h)
    imgs = torch.from_numpy(rng.integers(0, 256, (b, h, w, 3), dtype=np.uint8)).cuda()
    imgs[0, :3] = 0          # a zero byte is lut[0], not the zero of the padding
    imgs[-1, -2:] = 255
    g = torch.Generator().manual_seed(w)
    wt = (torch.randn((64, 7, 7, 4), generator=g) * 0.08).to(torch.bfloat16).cuda()
    bias = (torch.randn(64, generator=g) * 0.3).cuda()
    x = ops.preprocess(imgs, ro.IMAGENET_MEAN, ro.IMAGENET_STD, layout="nhwc", out_channels=4)
    want = ops.stem_pool(x, wt, bias, relu)
    got = ops.preprocess_stem_pool(imgs, ro.IMAGENET_MEAN, ro.IMAGENET_STD, wt, bias, relu)
    torch.cuda.synchronize()
    assert torch.equal(got, want), f"{(got != want).sum().item()} of {got.numel()} values differ"


def test_stem_pool_rejects_unsupported_sizes():
    from ai4e_b200 import ops, Ai4eError, _lib as L
    wt = torch.zeros((64, 7, 7, 4), dtype=torch.bfloat16, device="cuda")
    bias = torch.zeros(64, device="cuda")
    for h, w in ((30, 50), (224, 520), (18, 16)):   # width not a multiple of 16 / too wide / height not a multiple of 4
        with pytest.raises(Ai4eError) as ei:
            ops.stem_pool(torch.zeros((1, h, w, 4), dtype=torch.bfloat16, device="cuda"), wt, bias)
        assert ei.value.code == L.ERR_UNSUPPORTED


# (b, h, w, cin, cout, k, stride, pad, relu, residual): every conv geometry of ResNet-18/50
# after the stem, at small batch, plus ragged tails (M not a multiple of 128)
IGEMM_CASES = [
    (2, 56, 56, 64, 64, 1, 1, 0, True, False),
    (2, 56, 56, 64, 64, 3, 1, 1, True, False),
    (2, 56, 56, 64, 256, 1, 1, 0, True, True),
    (2, 56, 56, 256, 64, 1, 1, 0, True, False),
    (2, 56, 56, 256, 128, 1, 1, 0, True, False),
    (2, 56, 56, 128, 128, 3, 2, 1, True, False),
    (2, 56, 56, 256, 512, 1, 2, 0, False, False),
    (2, 28, 28, 128, 512, 1, 1, 0, True, True),
    (2, 28, 28, 512, 128, 1, 1, 0, True, False),
    (2, 28, 28, 128, 128, 3, 1, 1, True, False),
    (2, 28, 28, 256, 256, 3, 2, 1, True, False),
    (2, 28, 28, 512, 1024, 1, 2, 0, False, False),
    (2, 14, 14, 256, 1024, 1, 1, 0, True, True),
    (2, 14, 14, 1024, 256, 1, 1, 0, True, False),
    (2, 14, 14, 256, 256, 3, 1, 1, True, False),
    (2, 14, 14, 512, 512, 3, 2, 1, True, False),
    (2, 14, 14, 1024, 2048, 1, 2, 0, False, False),
    (3, 7, 7, 512, 2048, 1, 1, 0, True, True),
    (3, 7, 7, 2048, 512, 1, 1, 0, True, False),
    (3, 7, 7, 512, 512, 3, 1, 1, True, False),
    (1, 7, 7, 512, 512, 3, 1, 1, True, True),     # M = 49: a single, mostly empty tile
    (1, 5, 9, 64, 64, 3, 1, 1, False, False),     # tiny, ragged
    (5, 13, 17, 128, 192, 3, 2, 1, True, True),   # odd sizes, Cout = 3 x 64
    (1, 64, 64, 64, 128, 3, 1, 1, True, False),   # M = 4096: exact multiple of 128
    (2, 8, 8, 64, 64, 1, 1, 0, False, False),     # M = 128: exactly one tile
    # big enough that the wide tiles are chosen (>= 148 tiles at BLOCK_N = 256 / 128) and every
    # pipeline stage and mbarrier phase wraps several times per CTA
    (8, 56, 56, 64, 256, 1, 1, 0, True, True),    # BLOCK_N = 256, 3 stages, K = 64
    (7, 56, 56, 64, 256, 3, 1, 1, True, False),   # BLOCK_N = 256, 9 K blocks over 3 stages
    (8, 56, 56, 256, 128, 1, 1, 0, True, False),  # BLOCK_N = 128, 5 stages, 4 K blocks
    (6, 56, 56, 128, 256, 3, 1, 1, True, True),   # BLOCK_N = 128, 18 K blocks
]


@pytest.mark.parametrize("case", IGEMM_CASES, ids=lambda c: "x".join(str(v) for v in c[:8]))
def test_conv_igemm_vs_oracle(case):
    from ai4e_b200 import ops, _lib as L
    b, h, w, cin, cout, k, stride, pad, relu, residual = case
    x, wt, bias, res, want = _conv_case(*case, seed=sum(case[:8]))
    out = ops.conv_layer(x.cuda(), wt.cuda(), bias.cuda(), stride, pad, res.cuda() if res is not None else None,
                         relu, algo=L.CONV_IGEMM)
    torch.cuda.synchronize()
    _check_conv(out, want)


def test_conv_igemm_many_tiles_multi_wave():
    """more tiles than SMs, so every CTA loops over several tiles and both TMEM stages cycle."""
    from ai4e_b200 import ops, _lib as L
    case = (16, 56, 56, 64, 64, 3, 1, 1, True, True)  # M = 50176 -> 392 tiles
    x, wt, bias, res, want = _conv_case(*case, seed=5)
    out = ops.conv_layer(x.cuda(), wt.cuda(), bias.cuda(), 1, 1, res.cuda(), True, algo=L.CONV_IGEMM)
    torch.cuda.synchronize()
    _check_conv(out, want)


# CTA-pair kernel (tcgen05 cta_group::2, 256-row tiles): forced through the developer override,
# whatever the automatic choice would be.  (tiling, case)
PAIR_CASES = [
    ("128p", (2, 56, 56, 64, 256, 1, 1, 0, True, True)),     # 24.5 pair tiles: the last one is half empty
    ("256p", (2, 56, 56, 64, 256, 1, 1, 0, True, False)),
    ("128p", (8, 56, 56, 256, 128, 1, 1, 0, True, False)),   # more pair tiles than SM pairs: phases wrap
    ("128p", (2, 14, 14, 256, 256, 3, 1, 1, True, False)),
    ("256p", (2, 14, 14, 256, 256, 3, 1, 1, True, True)),
    ("128p", (3, 7, 7, 512, 512, 3, 1, 1, True, False)),     # M = 147: one pair tile, second CTA mostly padding
    ("128p", (1, 7, 7, 512, 2048, 1, 1, 0, True, True)),     # M = 49: the second CTA of every pair has no rows
    ("256p", (5, 13, 17, 128, 256, 3, 2, 1, False, True)),   # ragged geometry, stride 2
    ("256p", (6, 56, 56, 128, 256, 3, 1, 1, True, True)),    # 18 K blocks, 74 pair tiles
    ("128p", (64, 14, 14, 1024, 256, 1, 1, 0, True, False)), # ResNet-50 layer3 conv1 at full batch
]


@pytest.mark.parametrize("tiling,case", PAIR_CASES, ids=lambda c: c if isinstance(c, str) else "x".join(str(v) for v in c[:8]))
def test_conv_igemm_cta_pair_vs_oracle(tiling, case, monkeypatch):
    from ai4e_b200 import ops, _lib as L
    b, h, w, cin, cout, k, stride, pad, relu, residual = case
    x, wt, bias, res, want = _conv_case(*case, seed=sum(case[:8]) + 1)
    monkeypatch.setenv("AI4E_B200_TILE", tiling)
    out = ops.conv_layer(x.cuda(), wt.cuda(), bias.cuda(), stride, pad, res.cuda() if res is not None else None,
                         relu, algo=L.CONV_IGEMM)
    torch.cuda.synchronize()
    _check_conv(out, want)


# K4b: conv (3x3) + relu + conv 1x1 + residual + relu in one kernel.  (b, h, w, cin, cmid, k, stride, pad, cout)
TAIL_CASES = [
    (2, 56, 56, 64, 64, 3, 1, 1, 256),      # ResNet-50 layer1 conv2+conv3, 49 tiles
    (2, 28, 28, 128, 128, 3, 1, 1, 512),    # layer2
    (2, 56, 56, 128, 128, 3, 2, 1, 512),    # layer2.0 (stride 2)
    (1, 7, 9, 64, 64, 3, 1, 1, 128),        # one ragged tile, a single 128-channel pass
    (3, 13, 17, 192, 128, 3, 2, 1, 384),    # ragged geometry, three passes
    (2, 14, 14, 64, 64, 1, 1, 0, 256),      # 1x1 first convolution: one K block per tile
    (16, 56, 56, 64, 64, 3, 1, 1, 256),     # 392 tiles: every CTA loops, all barrier phases wrap
    (24, 28, 28, 128, 128, 3, 1, 1, 512),   # 147 tiles of the Cmid = 128 variant
]


@pytest.mark.parametrize("case", TAIL_CASES, ids=lambda c: "x".join(str(v) for v in c))
def test_bottleneck_tail_vs_oracle_and_unfused(case):
    """the fused kernel against (a) the fp32 oracle with y rounded to bf16 between the layers and
    (b) the two layers run separately through ai4e_conv_layer: (b) must match bit for bit."""
    from ai4e_b200 import ops, _lib as L
    b, h, w, cin, cmid, k, stride, pad, cout = case
    g = torch.Generator().manual_seed(sum(case))
    x = (torch.randn((b, h, w, cin), generator=g) * 0.7).to(torch.bfloat16)
    w2 = (torch.randn((cmid, k, k, cin), generator=g) * (2.0 / (k * k * cin)) ** 0.5).to(torch.bfloat16)
    b2 = torch.randn((cmid,), generator=g) * 0.1
    w3 = (torch.randn((cout, cmid), generator=g) * (2.0 / cmid) ** 0.5).to(torch.bfloat16)
    b3 = torch.randn((cout,), generator=g) * 0.1
    p, q = (h + 2 * pad - k) // stride + 1, (w + 2 * pad - k) // stride + 1
    res = (torch.randn((b, p, q, cout), generator=g) * 0.5).to(torch.bfloat16)
    y = torch.relu(torch.nn.functional.conv2d(x.float().permute(0, 3, 1, 2), w2.float().permute(0, 3, 1, 2), b2, stride, pad))
    y = y.to(torch.bfloat16).float()  # the intermediate activation is bf16 in both implementations
    want = torch.relu(torch.nn.functional.conv2d(y, w3.float()[:, :, None, None], b3).permute(0, 2, 3, 1) + res.float())
    xd, w2d, b2d, w3d, b3d, resd = (t.cuda() for t in (x, w2, b2, w3, b3, res))
    out = ops.bottleneck_tail(xd, w2d, b2d, stride, pad, w3d, b3d, resd)
    torch.cuda.synchronize()
    y_dev = ops.conv_layer(xd, w2d, b2d, stride, pad, None, True, algo=L.CONV_IGEMM)
    z_dev = ops.conv_layer(y_dev, w3d.reshape(cout, 1, 1, cmid).contiguous(), b3d, 1, 0, resd, True, algo=L.CONV_IGEMM)
    torch.cuda.synchronize()
    assert torch.equal(out, z_dev), f"{int((out != z_dev).sum())} elements differ from the unfused path"
    # a bf16 flip of y (rounding of a slightly different fp32 sum) moves z by |w3| * ulp(y): looser than _check_conv
    got = out.float().cpu()
    tol = 2.0 ** -6 * want.abs() + 2e-2
    bad = (got - want.contiguous()).abs() > tol
    assert not bad.any(), f"{int(bad.sum())} of {bad.numel()} outside tolerance; max err {(got - want).abs().max():.4f}"


def test_conv_igemm_rejects_unsupported():
    from ai4e_b200 import ops, _lib as L, Ai4eError
    x = torch.zeros((1, 8, 8, 48), dtype=torch.bfloat16, device="cuda")
    wt = torch.zeros((64, 1, 1, 48), dtype=torch.bfloat16, device="cuda")
    with pytest.raises(Ai4eError):
        ops.conv_layer(x, wt, torch.zeros(64, device="cuda"), 1, 0, algo=L.CONV_IGEMM)


# ------------------------------------------------------------------ pooling + head
@pytest.mark.parametrize("shape", [(2, 112, 112, 64), (3, 9, 7, 8), (1, 128, 128, 64)])
def test_maxpool_vs_oracle(shape):
    from ai4e_b200 import ops
    g = torch.Generator().manual_seed(1)
    x = torch.randn(shape, generator=g).to(torch.bfloat16)
    want = torch.nn.functional.max_pool2d(x.float().permute(0, 3, 1, 2), 3, 2, 1).permute(0, 2, 3, 1)
    out = ops.maxpool(x.cuda())
    assert torch.equal(out.float().cpu(), want.contiguous())  # exact: max of bf16 values


@pytest.mark.parametrize("b,hw,c,classes", [(1, 7, 2048, 1000), (9, 7, 512, 1000), (64, 8, 2048, 1000), (3, 2, 64, 10)])
def test_head_vs_oracle(b, hw, c, classes):
    from ai4e_b200 import ops
    g = torch.Generator().manual_seed(2)
    x = torch.relu(torch.randn((b, hw, hw, c), generator=g)).to(torch.bfloat16)
    w = (torch.randn((classes, c), generator=g) * 0.02).to(torch.bfloat16)
    bias = torch.randn((classes,), generator=g) * 0.1
    logits, idx, prob = ops.head(x.cuda(), w.cuda(), bias.cuda())
    pooled = x.float().mean(dim=(1, 2))
    want = pooled @ w.float().t() + bias
    got = logits.cpu()
    # fp32 throughout; only the summation order differs
    assert torch.allclose(got, want, atol=2e-4, rtol=1e-4), (got - want).abs().max()
    want_idx, want_prob = ro.softmax_top5(got.numpy(), k=min(5, classes))
    assert np.array_equal(idx.cpu().numpy(), want_idx)
    assert np.allclose(prob.cpu().numpy(), want_prob, atol=1e-6, rtol=1e-4)


def test_head_topk_tie_break():
    """equal logits resolve to the lower class index, like a stable sort."""
    from ai4e_b200 import ops
    x = torch.zeros((2, 1, 1, 64), dtype=torch.bfloat16)
    w = torch.zeros((40, 64), dtype=torch.bfloat16)
    bias = torch.zeros(40)
    bias[[3, 17, 30]] = 1.0
    _, idx, prob = ops.head(x.cuda(), w.cuda(), bias.cuda())
    assert idx.cpu().tolist() == [[3, 17, 30, 0, 1]] * 2
