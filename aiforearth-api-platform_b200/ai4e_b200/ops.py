"""Thin Python wrappers over the stateless C-ABI kernels.  torch tensors are used purely as
device-memory containers (`data_ptr()`), never for compute."""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib as L


def _stream_ptr(stream=None):
    s = stream if stream is not None else torch.cuda.current_stream()
    return C.c_void_p(s.cuda_stream)


def _ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else C.c_void_p(0)


def _farr(vals, n):
    return (C.c_float * n)(*[float(v) for v in vals][:n])


def preprocess(src_u8: torch.Tensor, mean, std, layout: str = "nchw", out_channels: int | None = None,
               resize: tuple[int, int] | None = None, crop: tuple[int, int] | None = None,
               stream=None) -> torch.Tensor:
    """uint8 [B,H,W,C] cuda tensor -> bf16 NCHW [B,C,h,w] or NHWC [B,h,w,out_channels]."""
    lib = L.load()
    assert src_u8.is_cuda and src_u8.dtype == torch.uint8 and src_u8.dim() == 4
    b, h, w, c = src_u8.shape
    assert src_u8.stride(3) == 1 and src_u8.stride(2) == c, "pixels must be dense HWC"
    rh, rw = resize if resize else (0, 0)
    full_h, full_w = (rh, rw) if resize else (h, w)
    oh, ow = crop if crop else (full_h, full_w)
    if layout == "nchw":
        oc = c
        out = torch.empty((b, c, oh, ow), dtype=torch.bfloat16, device=src_u8.device)
        lay = L.LAYOUT_NCHW
    else:
        oc = out_channels or 4
        out = torch.empty((b, oh, ow, oc), dtype=torch.bfloat16, device=src_u8.device)
        lay = L.LAYOUT_NHWC
    L.check(lib.ai4e_preprocess(_ptr(src_u8), b, h, w, c, src_u8.stride(0), src_u8.stride(1), rh, rw, oh, ow,
                                _farr(mean, c), _farr(std, c), lay, oc, _ptr(out), _stream_ptr(stream)))
    return out


def preprocess_tile(tile_u8: torch.Tensor, chip: int, mean, std, first_chip: int = 0, n_chips: int | None = None,
                    layout: str = "nchw", out_channels: int | None = None, resize: tuple[int, int] | None = None,
                    crop: tuple[int, int] | None = None, stream=None) -> torch.Tensor:
    """K6 fused into K1: uint8 [H,W,C] cuda tile -> the preprocessed chips [first_chip, first_chip + n_chips)
    of its chip x chip grid (row-major), read in place - no chip is copied."""
    lib = L.load()
    assert tile_u8.is_cuda and tile_u8.dtype == torch.uint8 and tile_u8.dim() == 3
    th, tw, c = tile_u8.shape
    assert tile_u8.stride(2) == 1 and tile_u8.stride(1) == c, "pixels must be dense HWC"
    total = (th // chip) * (tw // chip)
    n = total - first_chip if n_chips is None else n_chips
    rh, rw = resize if resize else (0, 0)
    full_h, full_w = (rh, rw) if resize else (chip, chip)
    oh, ow = crop if crop else (full_h, full_w)
    if layout == "nchw":
        oc, lay = c, L.LAYOUT_NCHW
        out = torch.empty((n, c, oh, ow), dtype=torch.bfloat16, device=tile_u8.device)
    else:
        oc, lay = out_channels or 4, L.LAYOUT_NHWC
        out = torch.empty((n, oh, ow, oc), dtype=torch.bfloat16, device=tile_u8.device)
    L.check(lib.ai4e_preprocess_tile(_ptr(tile_u8), th, tw, c, tile_u8.stride(0), chip, first_chip, n, rh, rw, oh, ow,
                                     _farr(mean, c), _farr(std, c), lay, oc, _ptr(out), _stream_ptr(stream)))
    return out


def conv_layer(x_nhwc: torch.Tensor, weight_krsc: torch.Tensor, bias: torch.Tensor, stride: int, pad: int,
               residual: torch.Tensor | None = None, relu: bool = True, algo: int = L.CONV_AUTO,
               stream=None) -> torch.Tensor:
    """x [B,H,W,Cin] bf16, weight [Cout,kh,kw,Cin] bf16, bias [Cout] fp32 -> [B,P,Q,Cout] bf16."""
    lib = L.load()
    assert x_nhwc.is_contiguous() and weight_krsc.is_contiguous() and bias.is_contiguous()
    assert x_nhwc.dtype == torch.bfloat16 and weight_krsc.dtype == torch.bfloat16 and bias.dtype == torch.float32
    b, h, w, cin = x_nhwc.shape
    cout, kh, kw, cin2 = weight_krsc.shape
    assert cin == cin2
    p = (h + 2 * pad - kh) // stride + 1
    q = (w + 2 * pad - kw) // stride + 1
    out = torch.empty((b, p, q, cout), dtype=torch.bfloat16, device=x_nhwc.device)
    if residual is not None:
        assert residual.is_contiguous() and residual.shape == out.shape and residual.dtype == torch.bfloat16
    L.check(lib.ai4e_conv_layer(_ptr(x_nhwc), b, h, w, cin, _ptr(weight_krsc), _ptr(bias), cout, kh, kw, stride,
                                pad, _ptr(residual), int(relu), _ptr(out), algo, _stream_ptr(stream)))
    return out


def png_decode(data: bytes):
    """host only: PNG bytes -> uint8 [H, W, 3] RGB, what Pillow's `Image.open(...).convert("RGB")` gives."""
    import numpy as np
    lib = L.load()
    buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
    h, w = C.c_int32(), C.c_int32()
    L.check(lib.ai4e_png_decode(buf, len(data), None, 0, C.byref(h), C.byref(w)))
    out = np.empty((h.value, w.value, 3), np.uint8)
    L.check(lib.ai4e_png_decode(buf, len(data), out.ctypes.data_as(C.c_void_p), out.nbytes, C.byref(h), C.byref(w)))
    return out


def image_probe(data: bytes):
    """host only, headers only: ("jpeg" | "png", height, width) when the engine can decode the payload;
    raises Ai4eError(ERR_UNSUPPORTED) otherwise.  See `require_decodable` for the 415 answer."""
    lib = L.load()
    buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
    kind, h, w = C.c_int32(), C.c_int32(), C.c_int32()
    L.check(lib.ai4e_image_probe(buf, len(data), C.byref(kind), C.byref(h), C.byref(w)))
    return {L.IMAGE_JPEG: "jpeg", L.IMAGE_PNG: "png"}[kind.value], h.value, w.value


def jpeg_entropy_decode(data: bytes):
    """host only: JPEG bytes (Huffman-coded sequential or progressive) -> (JpegInfo, int16 record [3 x 64 quantisation values, then
    the quantised coefficients per component, natural order]).  Raises Ai4eError for unsupported
    streams (arithmetic coding, CMYK, ...)."""
    import ctypes as C
    import numpy as np
    lib = L.load()
    info = L.JpegInfo()
    buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
    L.check(lib.ai4e_jpeg_entropy_decode(buf, len(data), None, 0, C.byref(info)))
    rec = np.zeros(int(info.record_int16), np.int16)
    L.check(lib.ai4e_jpeg_entropy_decode(buf, len(data), rec.ctypes.data_as(C.c_void_p), rec.size, C.byref(info)))
    return info, rec


def jpeg_decode(data: bytes, stream=None) -> torch.Tensor:
    """JPEG bytes -> uint8 [H, W, 3] RGB on the GPU, bit-identical to Pillow / libjpeg-turbo
    (entropy decode on this thread, IDCT + upsampling + colour conversion on the device)."""
    import ctypes as C
    lib = L.load()
    info = L.JpegInfo()
    buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
    L.check(lib.ai4e_jpeg_entropy_decode(buf, len(data), None, 0, C.byref(info)))
    out = torch.empty((info.height, info.width, 3), dtype=torch.uint8, device="cuda")
    h, w = C.c_int32(), C.c_int32()
    L.check(lib.ai4e_jpeg_decode(buf, len(data), _ptr(out), out.numel(), C.byref(h), C.byref(w), _stream_ptr(stream)))
    return out


def jpeg_reconstruct(records: torch.Tensor, info, stream=None) -> torch.Tensor:
    """[batch, record_int16] int16 records of ONE geometry on the GPU -> uint8 [batch, H, W, 3]."""
    import ctypes as C
    lib = L.load()
    assert records.is_cuda and records.dtype == torch.int16 and records.is_contiguous()
    b = records.shape[0]
    assert records.shape[1] == info.record_int16
    scratch = torch.empty((b, int(info.plane_bytes)), dtype=torch.uint8, device=records.device)
    out = torch.empty((b, info.height, info.width, 3), dtype=torch.uint8, device=records.device)
    L.check(lib.ai4e_jpeg_reconstruct(_ptr(records), b, C.byref(info), _ptr(scratch), _ptr(out), _stream_ptr(stream)))
    return out


def bottleneck_tail(x_nhwc: torch.Tensor, w2_krsc: torch.Tensor, b2: torch.Tensor, stride: int, pad: int,
                    w3_kc: torch.Tensor, b3: torch.Tensor, residual: torch.Tensor, stream=None) -> torch.Tensor:
    """relu(conv1x1(relu(conv(x; w2) + b2); w3) + b3 + residual) in one kernel (K4b).
    x [B,H,W,Cin] bf16, w2 [Cmid,kh,kw,Cin] bf16, w3 [Cout,Cmid] bf16, residual [B,P,Q,Cout] bf16."""
    lib = L.load()
    for t in (x_nhwc, w2_krsc, w3_kc, residual):
        assert t.is_contiguous() and t.dtype == torch.bfloat16
    b, h, w, cin = x_nhwc.shape
    cmid, kh, kw, _ = w2_krsc.shape
    cout = w3_kc.shape[0]
    p = (h + 2 * pad - kh) // stride + 1
    q = (w + 2 * pad - kw) // stride + 1
    assert tuple(residual.shape) == (b, p, q, cout)
    out = torch.empty((b, p, q, cout), dtype=torch.bfloat16, device=x_nhwc.device)
    L.check(lib.ai4e_bottleneck_tail(_ptr(x_nhwc), b, h, w, cin, _ptr(w2_krsc), _ptr(b2), cmid, kh, kw, stride, pad,
                                     _ptr(w3_kc), _ptr(b3), cout, _ptr(residual), _ptr(out), _stream_ptr(stream)))
    return out


def stem_pool(x_nhwc4: torch.Tensor, weight_krsc: torch.Tensor, bias: torch.Tensor, relu: bool = True, stream=None) -> torch.Tensor:
    """7x7/2 stem + bias (+ ReLU) + 3x3/2 max-pool in one kernel: x [B,H,W,4] bf16, weight [64,7,7,4] bf16,
    bias [64] fp32 -> [B,H/4,W/4,64] bf16; bit-identical to conv_layer followed by maxpool."""
    lib = L.load()
    assert x_nhwc4.is_contiguous() and x_nhwc4.dtype == torch.bfloat16 and x_nhwc4.shape[3] == 4
    assert weight_krsc.is_contiguous() and tuple(weight_krsc.shape) == (64, 7, 7, 4) and weight_krsc.dtype == torch.bfloat16
    b, h, w, _ = x_nhwc4.shape
    out = torch.empty((b, h // 4, w // 4, 64), dtype=torch.bfloat16, device=x_nhwc4.device)
    L.check(lib.ai4e_stem_pool(_ptr(x_nhwc4), b, h, w, _ptr(weight_krsc), _ptr(bias), int(relu), _ptr(out),
                               _stream_ptr(stream)))
    return out


def preprocess_stem_pool(images_u8: torch.Tensor, mean, std, weight_krsc: torch.Tensor, bias: torch.Tensor,
                         relu: bool = True, stream=None) -> torch.Tensor:
    """uint8 [B,H,W,3] -> normalise + 7x7/2 stem + bias (+ ReLU) + 3x3/2 max-pool in ONE kernel ->
    [B,H/4,W/4,64] bf16; bit-identical to preprocess(layout="nhwc", out_channels=4) followed by stem_pool."""
    lib = L.load()
    assert images_u8.is_cuda and images_u8.is_contiguous() and images_u8.dtype == torch.uint8 and images_u8.shape[3] == 3
    b, h, w, _ = images_u8.shape
    out = torch.empty((b, h // 4, w // 4, 64), dtype=torch.bfloat16, device=images_u8.device)
    L.check(lib.ai4e_preprocess_stem_pool(_ptr(images_u8), b, h, w, _farr(mean, 3), _farr(std, 3), _ptr(weight_krsc),
                                          _ptr(bias), int(relu), _ptr(out), _stream_ptr(stream)))
    return out


def maxpool(x_nhwc: torch.Tensor, k: int = 3, stride: int = 2, pad: int = 1, stream=None) -> torch.Tensor:
    lib = L.load()
    assert x_nhwc.is_contiguous() and x_nhwc.dtype == torch.bfloat16
    b, h, w, c = x_nhwc.shape
    p = (h + 2 * pad - k) // stride + 1
    q = (w + 2 * pad - k) // stride + 1
    out = torch.empty((b, p, q, c), dtype=torch.bfloat16, device=x_nhwc.device)
    L.check(lib.ai4e_maxpool(_ptr(x_nhwc), b, h, w, c, k, stride, pad, _ptr(out), _stream_ptr(stream)))
    return out


def head(x_nhwc: torch.Tensor, fc_w: torch.Tensor, fc_b: torch.Tensor, topk: int = L.TOPK, stream=None):
    """x [B,H,W,C] bf16, fc_w [classes,C] bf16, fc_b fp32 -> (logits fp32, top_idx i32, top_prob fp32)."""
    lib = L.load()
    assert x_nhwc.is_contiguous() and fc_w.is_contiguous() and fc_b.is_contiguous()
    b, h, w, c = x_nhwc.shape
    classes = fc_w.shape[0]
    logits = torch.empty((b, classes), dtype=torch.float32, device=x_nhwc.device)
    idx = torch.empty((b, max(topk, 1)), dtype=torch.int32, device=x_nhwc.device)
    prob = torch.empty((b, max(topk, 1)), dtype=torch.float32, device=x_nhwc.device)
    L.check(lib.ai4e_head(_ptr(x_nhwc), b, h * w, c, _ptr(fc_w), _ptr(fc_b), classes, _ptr(logits), _ptr(idx),
                          _ptr(prob), topk, _stream_ptr(stream)))
    return logits, idx[:, :topk], prob[:, :topk]


def head_noalloc(x_nhwc: torch.Tensor, fc_w: torch.Tensor, fc_b: torch.Tensor, topk: int = L.TOPK, stream=None):
    """`head` for callers that must not allocate device memory: the pooled-feature scratch is a
    fixed array inside the library (batch*C <= L.HEAD_NOALLOC_MAX_FLOATS), shared by every call on
    the device - do not overlap calls.  The output tensors here are made by the caller (torch)."""
    lib = L.load()
    assert x_nhwc.is_contiguous() and fc_w.is_contiguous() and fc_b.is_contiguous()
    b, h, w, c = x_nhwc.shape
    classes = fc_w.shape[0]
    logits = torch.empty((b, classes), dtype=torch.float32, device=x_nhwc.device)
    idx = torch.empty((b, max(topk, 1)), dtype=torch.int32, device=x_nhwc.device)
    prob = torch.empty((b, max(topk, 1)), dtype=torch.float32, device=x_nhwc.device)
    L.check(lib.ai4e_head_noalloc(_ptr(x_nhwc), b, h * w, c, _ptr(fc_w), _ptr(fc_b), classes, _ptr(logits),
                                  _ptr(idx), _ptr(prob), topk, _stream_ptr(stream)))
    return logits, idx[:, :topk], prob[:, :topk]


def tile_summary(top_idx5: torch.Tensor, top_prob5: torch.Tensor, num_classes: int, first: int = 0, out=None, stream=None):
    """K7 stand-alone: folds the top-5 of `count` chips into (labels int32 [n], label_prob float32 [n], hist
    int32-as-uint32 [classes], mass_q32 int64-as-uint64 [classes]) on the GPU.  `out` = a previous result to add
    to (chips [first, first + count) of a larger summary); otherwise fresh zeroed tensors sized first + count."""
    lib = L.load()
    assert top_idx5.is_cuda and top_idx5.dtype == torch.int32 and top_idx5.is_contiguous() and top_idx5.shape[1] == L.TOPK
    assert top_prob5.is_cuda and top_prob5.dtype == torch.float32 and top_prob5.is_contiguous() and top_prob5.shape == top_idx5.shape
    count = top_idx5.shape[0]
    if out is None:
        dev = top_idx5.device
        out = (torch.zeros(first + count, dtype=torch.int32, device=dev), torch.zeros(first + count, dtype=torch.float32, device=dev),
               torch.zeros(num_classes, dtype=torch.int32, device=dev), torch.zeros(num_classes, dtype=torch.int64, device=dev))
    labels, lprob, hist, mass = out
    assert labels.numel() >= first + count and hist.numel() == num_classes and mass.numel() == num_classes
    L.check(lib.ai4e_tile_summary(_ptr(top_idx5), _ptr(top_prob5), count, first, num_classes, _ptr(labels), _ptr(lprob),
                                  _ptr(hist), _ptr(mass), _stream_ptr(stream)))
    return out
