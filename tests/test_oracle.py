"""CPU tests that pin the oracle: against the committed goldens and against torch itself.

The reference holds no golden vectors for this path, so the pins are (a) the
fixtures minted by tests/golden/make_goldens.py and (b) torch / torchvision, the third-party
dependency BASELINE.json nominates as "the reference container's own PyTorch CPU path"."""
import numpy as np
import pytest
import torch
import torch.nn.functional as F

from oracle import preprocess_oracle as pp
from oracle import resnet_oracle as ro
from tests import util


def test_bf16_bits_match_torch():
    rng = np.random.default_rng(0)
    x = np.concatenate([rng.standard_normal(20000).astype(np.float32) * 3,
                        np.array([0.0, -0.0, 1.0, 1.00390625, 1.001953125, 3.4e38, -3.4e38, 1e-40, np.inf, -np.inf],
                                 np.float32)])
    want = torch.from_numpy(x).to(torch.bfloat16).view(torch.int16).numpy().view(np.uint16)
    assert np.array_equal(pp.bf16_bits(x), want)
    assert np.array_equal(pp.bf16_to_f32(want), torch.from_numpy(x).to(torch.bfloat16).float().numpy())


def test_preprocess_matches_torch_to_tensor_normalize():
    rng = np.random.default_rng(1)
    img = rng.integers(0, 256, (3, 31, 45, 3), dtype=np.uint8)
    got = pp.preprocess_nchw(img, ro.IMAGENET_MEAN, ro.IMAGENET_STD)
    t = torch.from_numpy(img).permute(0, 3, 1, 2).float().div(255)
    mean = torch.tensor(ro.IMAGENET_MEAN).view(1, 3, 1, 1)
    std = torch.tensor(ro.IMAGENET_STD).view(1, 3, 1, 1)
    want = ((t - mean) / std).numpy()
    assert np.array_equal(got, want)  # same three fp32 operations


@pytest.mark.parametrize("src,dst", [((40, 56), (32, 44)), ((256, 256), (224, 224)), ((17, 23), (40, 31))])
def test_resize_matches_torch_interpolate(src, dst):
    rng = np.random.default_rng(2)
    img = rng.integers(0, 256, (2, src[0], src[1], 3), dtype=np.uint8)
    got = pp.resize_bilinear(img, dst[0], dst[1])
    want = F.interpolate(torch.from_numpy(img).permute(0, 3, 1, 2).float(), size=dst, mode="bilinear",
                         align_corners=False, antialias=False).permute(0, 2, 3, 1).numpy()
    # same formula; torch rounds the source coordinate slightly differently (lambda differs by
    # ~1e-6), which times a pixel step of up to 255 is a few 1e-4 on the 0..255 scale
    assert np.abs(got - want).max() < 5e-4


def test_center_crop_offsets_match_torchvision():
    from torchvision.transforms import functional as TF
    for h, w, ch, cw in [(256, 256, 224, 224), (33, 47, 28, 30), (32, 44, 28, 28), (10, 10, 10, 10), (9, 12, 4, 7)]:
        t = torch.arange(h * w).view(1, h, w)
        c = TF.center_crop(t, [ch, cw])
        top, left = pp.center_crop_offsets(h, w, ch, cw)
        assert int(c[0, 0, 0]) == top * w + left


def test_preprocess_goldens(golden_dir):
    g = np.load(f"{golden_dir}/preprocess.npz")
    for case in "abcd":
        img = g[f"plain_{case}_in"]
        mean, std = util.mean_std(img.shape[-1])
        assert np.array_equal(pp.bf16_bits(pp.preprocess_nchw(img, mean, std)), g[f"plain_{case}_out_nchw_bf16"])
    rh, rw, ch, cw = [int(v) for v in g["resize_params"]]
    out = pp.preprocess_resize_crop_nchw(g["resize_in"], rh, rw, ch, cw, ro.IMAGENET_MEAN, ro.IMAGENET_STD)
    assert np.array_equal(pp.bf16_bits(out), g["resize_out_nchw_bf16"])


def test_chips_from_tile():
    tile = np.arange(8 * 12 * 2, dtype=np.uint8).reshape(8, 12, 2)
    chips = pp.chips_from_tile(tile, 4)
    assert chips.shape == (6, 4, 4, 2)
    assert np.array_equal(chips[4], tile[4:8, 4:8])


def test_synthetic_images_are_pure_functions_of_seed_and_index():
    a = ro.synthetic_images([3, 9], seed=1234)
    b = ro.synthetic_images(10, seed=1234)
    assert np.array_equal(a[0], b[3]) and np.array_equal(a[1], b[9])
    n = ro.synthetic_images(2, structured=False)
    assert n.dtype == np.uint8 and n.shape == (2, 224, 224, 3)


@pytest.mark.parametrize("name", ["resnet18", "resnet50", "resnet50_landcover4"])
def test_oracle_reproduces_goldens(name):
    """weights regenerated from seeds + fp32 forward reproduce the committed logits and top-5;
    the bf16-storage emulation stays within the recorded format error."""
    g, imgs = util.load_golden(name)
    n = min(8, len(imgs))
    in_ch, hw = int(g["in_ch"]), int(g["hw"])
    m = util.oracle_model(str(g["arch"]), in_ch, hw)
    mean, std = util.mean_std(in_ch)
    idx, prob, logits = ro.classify_fp32(m, imgs[:n], mean, std)
    assert np.abs(logits - g["logits_fp32"][:n]).max() < 2e-4
    assert np.array_equal(idx, g["top5_idx"][:n])
    assert np.allclose(prob, g["top5_prob"][:n], rtol=1e-3, atol=1e-7)
    emul = ro.forward_bf16_emul(m, pp.preprocess_nchw(imgs[:n], mean, std))
    assert np.abs(emul - g["logits_bf16_emul"][:n]).max() < 5e-3
    assert np.abs(emul - logits).max() <= float(g["fmt_err_max"]) * 1.25 + 1e-3
    srt = -np.sort(-logits, axis=1)[:, :6]
    assert (srt[:, :-1] - srt[:, 1:]).min() > float(g["min_gap"]) * 0.98  # the margin the images were chosen for


def test_softmax_top5_is_stable_on_ties():
    z = np.zeros((1, 10), np.float32)
    z[0, [7, 2, 5]] = 1.0
    idx, prob = ro.softmax_top5(z)
    assert idx.tolist() == [[2, 5, 7, 0, 1]]
    assert np.isclose(prob.sum() + (10 - 5) * prob[0, 4], 1.0)
