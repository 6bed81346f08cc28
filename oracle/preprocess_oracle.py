"""CPU oracle for the preprocessing step (TEST INFRASTRUCTURE — see resnet_oracle.py header;
only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import this).

The reference has no preprocessing code (the whole model function is the opaque callback at
`/root/reference/APIs/1.0/base-py/ai4e_service.py:203`).  This restates the standard
torchvision pipeline a PyTorch CPU container would run inside that callback:

    [Resize(bilinear) -> CenterCrop] -> ToTensor (uint8 HWC -> float /255, CHW) -> Normalize(mean, std)

in numpy fp32 with a fixed operation order so a GPU kernel can be bit-exact against it.
PARITY UNPINNED BY THE REFERENCE (no golden vectors exist upstream); `tests/test_oracle.py`
pins this restatement against torch (`F.interpolate`, tensor arithmetic) instead.
"""
from __future__ import annotations

import numpy as np

F32 = np.float32


def bf16_bits(x: np.ndarray) -> np.ndarray:
    """fp32 -> bf16 bit pattern (uint16), round-to-nearest-even, NaN preserved."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    u = x.view(np.uint32)
    bias = np.uint32(0x7FFF) + ((u >> np.uint32(16)) & np.uint32(1))
    r = ((u + bias) >> np.uint32(16)).astype(np.uint16)
    nan = np.isnan(x)
    if nan.any():
        r = np.where(nan, np.uint16(0x7FC0), r)
    return r


def bf16_to_f32(bits: np.ndarray) -> np.ndarray:
    return (np.ascontiguousarray(bits, dtype=np.uint16).astype(np.uint32) << np.uint32(16)).view(np.float32)


def bf16_round(x: np.ndarray) -> np.ndarray:
    return bf16_to_f32(bf16_bits(x))


def normalize_hwc(x_f32_hwc: np.ndarray, mean, std) -> np.ndarray:
    """(x/255 - mean)/std, three separately rounded fp32 operations, in that order."""
    m = np.asarray(mean, dtype=F32)
    s = np.asarray(std, dtype=F32)
    x = x_f32_hwc.astype(F32) / F32(255.0)
    x = x - m
    x = x / s
    return x.astype(F32)


def preprocess_nhwc(images_u8: np.ndarray, mean, std) -> np.ndarray:
    """[N,H,W,C] uint8 -> [N,H,W,C] fp32 normalised."""
    assert images_u8.dtype == np.uint8 and images_u8.ndim == 4
    return normalize_hwc(images_u8.astype(F32), mean, std)


def preprocess_nchw(images_u8: np.ndarray, mean, std) -> np.ndarray:
    """[N,H,W,C] uint8 -> [N,C,H,W] fp32 normalised (ToTensor + Normalize)."""
    return np.ascontiguousarray(preprocess_nhwc(images_u8, mean, std).transpose(0, 3, 1, 2))


def _src_index(out_size: int, in_size: int):
    """torch upsample_bilinear2d, align_corners=False: (i0, i1, lambda0, lambda1) in fp32."""
    scale = F32(in_size) / F32(out_size)
    dst = np.arange(out_size, dtype=F32)
    src = (dst + F32(0.5)) * scale - F32(0.5)
    src = np.maximum(src, F32(0.0)).astype(F32)
    i0 = np.minimum(np.floor(src).astype(np.int64), in_size - 1)
    i1 = np.minimum(i0 + 1, in_size - 1)
    l1 = (src - i0.astype(F32)).astype(F32)
    l0 = (F32(1.0) - l1).astype(F32)
    return i0, i1, l0, l1


def resize_bilinear(images_u8: np.ndarray, out_h: int, out_w: int) -> np.ndarray:
    """[N,H,W,C] uint8 -> [N,out_h,out_w,C] fp32 (0..255 scale), no antialias, no rounding.

    out = l0h*(l0w*p00 + l1w*p01) + l1h*(l0w*p10 + l1w*p11), every product and sum rounded
    to fp32 separately (no fused multiply-add)."""
    n, h, w, c = images_u8.shape
    y0, y1, ly0, ly1 = _src_index(out_h, h)
    x0, x1, lx0, lx1 = _src_index(out_w, w)
    img = images_u8.astype(F32)
    p00 = img[:, y0][:, :, x0]
    p01 = img[:, y0][:, :, x1]
    p10 = img[:, y1][:, :, x0]
    p11 = img[:, y1][:, :, x1]
    lx0 = lx0[None, None, :, None]
    lx1 = lx1[None, None, :, None]
    ly0 = ly0[None, :, None, None]
    ly1 = ly1[None, :, None, None]
    top = (lx0 * p00).astype(F32) + (lx1 * p01).astype(F32)
    bot = (lx0 * p10).astype(F32) + (lx1 * p11).astype(F32)
    return ((ly0 * top.astype(F32)).astype(F32) + (ly1 * bot.astype(F32)).astype(F32)).astype(F32)


def center_crop_offsets(h: int, w: int, crop_h: int, crop_w: int):
    """torchvision CenterCrop offsets: int(round((h - crop_h) / 2.0))."""
    top = int(round((h - crop_h) / 2.0))
    left = int(round((w - crop_w) / 2.0))
    return top, left


def preprocess_resize_crop_nhwc(images_u8: np.ndarray, resize_h: int, resize_w: int,
                                crop_h: int, crop_w: int, mean, std) -> np.ndarray:
    """Resize(bilinear) -> CenterCrop -> /255 -> normalise.  [N,crop_h,crop_w,C] fp32."""
    r = resize_bilinear(images_u8, resize_h, resize_w)
    top, left = center_crop_offsets(resize_h, resize_w, crop_h, crop_w)
    r = r[:, top:top + crop_h, left:left + crop_w, :]
    return normalize_hwc(r, mean, std)


def preprocess_resize_crop_nchw(images_u8, resize_h, resize_w, crop_h, crop_w, mean, std):
    x = preprocess_resize_crop_nhwc(images_u8, resize_h, resize_w, crop_h, crop_w, mean, std)
    return np.ascontiguousarray(x.transpose(0, 3, 1, 2))


def chips_from_tile(tile_u8: np.ndarray, chip: int = 256) -> np.ndarray:
    """Land-cover config: [H,W,C] tile -> [(H/chip)*(W/chip), chip, chip, C] non-overlapping
    chips, row-major over the chip grid."""
    h, w, c = tile_u8.shape
    gh, gw = h // chip, w // chip
    t = tile_u8[:gh * chip, :gw * chip].reshape(gh, chip, gw, chip, c)
    return np.ascontiguousarray(t.transpose(0, 2, 1, 3, 4)).reshape(gh * gw, chip, chip, c)
