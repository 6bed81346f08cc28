"""Python face of the engine: `Engine` (one replica on one GPU) and `Pool` (one replica per
GPU behind one dispatcher).  A user `func` decorated with `APIService.api_sync_func` /
`api_async_func` calls `engine.classify(image)`; that single call stands where the per-image
PyTorch forward stood inside the reference's callback (APIs/1.0/base-py/ai4e_service.py:203).

ctypes releases the GIL for the duration of every C call, so request threads overlap in
`submit`/`wait`.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

from . import _lib as L
from .models import Program


def _u8_hwc(image) -> np.ndarray:
    a = np.asarray(image)
    if a.dtype != np.uint8 or a.ndim != 3:
        raise ValueError(f"expected a uint8 HWC image, got dtype {a.dtype} shape {a.shape}")
    if a.strides[2] != 1 or a.strides[1] != a.shape[2]:
        a = np.ascontiguousarray(a)
    return a


class _Base:
    _h = None

    def _results(self, n=1):
        return (np.empty((n, L.TOPK), np.int32), np.empty((n, L.TOPK), np.float32))


class Engine(_Base):
    """One model replica on one GPU with its micro-batching queue."""

    def __init__(self, program: Program, device: int = 0, max_batch: int | None = None,
                 max_wait_us: int | None = None):
        self._lib = L.load()
        self.program = program
        self.max_batch = int(max_batch if max_batch is not None else os.environ.get("AI4E_B200_MAX_BATCH", 64))
        self.max_wait_us = int(max_wait_us if max_wait_us is not None else os.environ.get("AI4E_B200_MAX_WAIT_US", 2000))
        md, keep = program.to_ctypes()
        h = C.c_void_p()
        L.check(self._lib.ai4e_engine_create(device, C.byref(md), program.blob, len(program.blob), self.max_batch,
                                             self.max_wait_us, C.byref(h)))
        self._h = h
        self._owned = True
        self.device = device
        self.num_classes = program.num_classes

    @classmethod
    def _borrow(cls, handle, program, device, max_batch):
        self = cls.__new__(cls)
        self._lib = L.load()
        self._h, self._owned = handle, False
        self.program, self.device, self.max_batch = program, device, max_batch
        self.num_classes = program.num_classes
        return self

    # ---- synchronous batch (host memory in, host memory out)
    def forward(self, images_u8: np.ndarray, want_logits: bool = False):
        """[B,H,W,C] uint8 host array -> (top5_idx [B,5] i32, top5_prob [B,5] f32[, logits [B,K] f32])."""
        a = np.ascontiguousarray(images_u8)
        if a.dtype != np.uint8 or a.ndim != 4:
            raise ValueError("expected a uint8 [B,H,W,C] array")
        b, h, w, c = a.shape
        if c != self.program.in_c:
            raise ValueError(f"model takes {self.program.in_c} channels, got {c}")
        idx, prob = self._results(b)
        logits = np.empty((b, self.num_classes), np.float32) if want_logits else None
        L.check(self._lib.ai4e_engine_forward_host(self._h, a.ctypes.data_as(C.c_void_p), b, h, w,
                                                   idx.ctypes.data_as(C.c_void_p), prob.ctypes.data_as(C.c_void_p),
                                                   logits.ctypes.data_as(C.c_void_p) if want_logits else None))
        return (idx, prob, logits) if want_logits else (idx, prob)

    def forward_device(self, d_images, batch, h, w, d_idx, d_prob, d_logits=None, stream=None):
        """everything resident on the device (torch tensors as containers); returns once enqueued."""
        sp = C.c_void_p(stream.cuda_stream) if stream is not None else C.c_void_p(0)
        L.check(self._lib.ai4e_engine_forward_device(
            self._h, C.c_void_p(d_images.data_ptr()), batch, h, w, C.c_void_p(d_idx.data_ptr()),
            C.c_void_p(d_prob.data_ptr()), C.c_void_p(d_logits.data_ptr()) if d_logits is not None else None, sp))

    # ---- micro-batched path
    def submit(self, image, tag: int = 0) -> int:
        a = _u8_hwc(image)
        t = C.c_uint64()
        L.check(self._lib.ai4e_engine_submit(self._h, a.ctypes.data_as(C.c_void_p), a.shape[0], a.shape[1], a.shape[2],
                                             a.strides[0], tag, C.byref(t)))
        return t.value

    def submit_tile(self, tile_u8_hwc: np.ndarray, chip: int = 256) -> np.ndarray:
        """every chip x chip chip of a [H, W, C] uint8 tile (row-major over the chip grid), submitted
        in one C call with the tile's row stride -> uint64 tickets [gh * gw]."""
        a = np.asarray(tile_u8_hwc)
        if a.dtype != np.uint8 or a.ndim != 3 or a.strides[2] != 1 or a.strides[1] != a.shape[2]:
            raise ValueError("expected a uint8 [H, W, C] tile with dense pixels")
        gh, gw = a.shape[0] // chip, a.shape[1] // chip
        tickets = np.empty(max(gh * gw, 1), np.uint64)
        n = C.c_int32()
        L.check(self._lib.ai4e_engine_submit_tile(self._h, a.ctypes.data_as(C.c_void_p), a.shape[0], a.shape[1], a.shape[2],
                                                  a.strides[0], chip, tickets.ctypes.data_as(C.c_void_p), tickets.size,
                                                  C.byref(n)))
        return tickets[:n.value]

    def wait_many(self, tickets: np.ndarray):
        """-> (idx [n, 5] int32, prob [n, 5] float32) for the tickets, in order."""
        t = np.ascontiguousarray(tickets, np.uint64)
        idx = np.empty((t.size, L.TOPK), np.int32)
        prob = np.empty((t.size, L.TOPK), np.float32)
        L.check(self._lib.ai4e_engine_wait_many(self._h, t.ctypes.data_as(C.c_void_p), t.size,
                                                idx.ctypes.data_as(C.c_void_p), prob.ctypes.data_as(C.c_void_p)))
        return idx, prob

    def submit_jpeg(self, data: bytes, tag: int = 0) -> int:
        """one baseline JPEG (bytes) -> ticket.  The Huffman decode runs on this thread with the GIL
        released; the GPU reconstructs the pixels (bit-identical to Pillow) in front of the model."""
        buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
        t = C.c_uint64()
        L.check(self._lib.ai4e_engine_submit_jpeg(self._h, buf, len(data), tag, C.byref(t)))
        return t.value

    def classify_jpeg(self, data: bytes, want_logits: bool = False):
        """blocking: submit_jpeg + wait."""
        return self.wait(self.submit_jpeg(data), want_logits)

    def submit_batch(self, images_u8: np.ndarray) -> range:
        """dense [B,H,W,C] uint8 host array -> tickets (a range).  Page-locked arrays are read in
        place by the H2D copy: keep them unchanged until the last ticket has been waited for."""
        a = images_u8
        if a.dtype != np.uint8 or a.ndim != 4 or not a.flags.c_contiguous:
            raise ValueError("expected a C-contiguous uint8 [B,H,W,C] array")
        t = C.c_uint64()
        L.check(self._lib.ai4e_engine_submit_batch(self._h, a.ctypes.data_as(C.c_void_p), a.shape[0], a.shape[1],
                                                   a.shape[2], a.shape[3], C.byref(t)))
        return range(t.value, t.value + a.shape[0])

    def profile_layers(self, d_images, batch: int, h: int, w: int, reps: int = 5) -> np.ndarray:
        """mean device ms of [preprocess, layer 0, layer 1, ...] on a device-resident batch."""
        n = len(self.program.layers) + 1
        out = np.zeros(n, np.float32)
        L.check(self._lib.ai4e_engine_profile(self._h, C.c_void_p(d_images.data_ptr()), batch, h, w, reps,
                                              out.ctypes.data_as(C.POINTER(C.c_float)), n))
        return out

    def wait(self, ticket: int, want_logits: bool = False):
        idx, prob = self._results(1)
        logits = np.empty((self.num_classes,), np.float32) if want_logits else None
        L.check(self._lib.ai4e_engine_wait(self._h, ticket, idx.ctypes.data_as(C.c_void_p),
                                           prob.ctypes.data_as(C.c_void_p),
                                           logits.ctypes.data_as(C.c_void_p) if want_logits else None))
        return (idx[0], prob[0], logits) if want_logits else (idx[0], prob[0])

    def classify(self, image, want_logits: bool = False):
        """submit + wait for one image: the call a user `func` makes."""
        return self.wait(self.submit(image), want_logits)

    def flush(self):
        L.check(self._lib.ai4e_engine_flush(self._h))

    def stats(self) -> dict:
        s = L.EngineStats()
        L.check(self._lib.ai4e_engine_stats_get(self._h, C.byref(s)))
        return {"images": s.images, "batches": s.batches, "batch_hist": list(s.batch_hist),
                "queue_depth": s.queue_depth, "gpu_ms_total": s.gpu_ms_total, "lat_ms_p50": s.lat_ms_p50,
                "lat_ms_p99": s.lat_ms_p99, "healthy": bool(s.healthy),
                "launches_per_forward": s.launches_per_forward}

    def read_stage(self, name: str, batch: int) -> np.ndarray:
        """bf16 bits [batch,h,w,c] (uint16) of a stage output of the last forward (parity tests;
        build the program with keep_stages=True)."""
        slot, h, w, c = self.program.stage_bufs[name]
        out = np.empty((batch, h, w, c), np.uint16)
        L.check(self._lib.ai4e_engine_read_activation(self._h, slot, out.ctypes.data_as(C.c_void_p), out.nbytes))
        return out

    def close(self):
        if self._h is not None and self._owned:
            self._lib.ai4e_engine_destroy(self._h)
        self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass


class Pool(_Base):
    """N independent replicas (one per GPU) fed from one dispatcher; no collective anywhere."""

    def __init__(self, program: Program, devices=None, max_batch: int | None = None, max_wait_us: int | None = None):
        self._lib = L.load()
        if devices is None:
            env = os.environ.get("AI4E_B200_DEVICES")
            if env:
                devices = [int(v) for v in env.split(",") if v.strip()]
            else:
                n = C.c_int()
                L.check(self._lib.ai4e_device_count(C.byref(n)))
                devices = list(range(n.value))
        self.devices = list(devices)
        self.program = program
        self.num_classes = program.num_classes
        self.max_batch = int(max_batch if max_batch is not None else os.environ.get("AI4E_B200_MAX_BATCH", 64))
        self.max_wait_us = int(max_wait_us if max_wait_us is not None else os.environ.get("AI4E_B200_MAX_WAIT_US", 2000))
        md, keep = program.to_ctypes()
        arr = (C.c_int * len(self.devices))(*self.devices)
        h = C.c_void_p()
        L.check(self._lib.ai4e_pool_create(arr, len(self.devices), C.byref(md), program.blob, len(program.blob),
                                           self.max_batch, self.max_wait_us, C.byref(h)))
        self._h = h

    def submit(self, image, tag: int = 0) -> int:
        a = _u8_hwc(image)
        t = C.c_uint64()
        L.check(self._lib.ai4e_pool_submit(self._h, a.ctypes.data_as(C.c_void_p), a.shape[0], a.shape[1], a.shape[2],
                                           a.strides[0], tag, C.byref(t)))
        return t.value

    def submit_jpeg(self, data: bytes, tag: int = 0) -> int:
        """one baseline JPEG -> ticket, routed to the least-loaded replica (see Engine.submit_jpeg)."""
        buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
        t = C.c_uint64()
        L.check(self._lib.ai4e_pool_submit_jpeg(self._h, buf, len(data), tag, C.byref(t)))
        return t.value

    def classify_jpeg(self, data: bytes, want_logits: bool = False):
        return self.wait(self.submit_jpeg(data), want_logits)

    def wait(self, ticket: int, want_logits: bool = False):
        idx, prob = self._results(1)
        logits = np.empty((self.num_classes,), np.float32) if want_logits else None
        L.check(self._lib.ai4e_pool_wait(self._h, ticket, idx.ctypes.data_as(C.c_void_p),
                                         prob.ctypes.data_as(C.c_void_p),
                                         logits.ctypes.data_as(C.c_void_p) if want_logits else None))
        return (idx[0], prob[0], logits) if want_logits else (idx[0], prob[0])

    def classify(self, image, want_logits: bool = False):
        return self.wait(self.submit(image), want_logits)

    def flush(self):
        L.check(self._lib.ai4e_pool_flush(self._h))

    def engines(self):
        out = []
        for i in range(self._lib.ai4e_pool_size(self._h)):
            h = C.c_void_p()
            L.check(self._lib.ai4e_pool_engine(self._h, i, C.byref(h)))
            out.append(Engine._borrow(h, self.program, self.devices[i], self.max_batch))
        return out

    def stats(self):
        return [e.stats() for e in self.engines()]

    def close(self):
        if self._h is not None:
            self._lib.ai4e_pool_destroy(self._h)
        self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def __del__(self):  # pragma: no cover
        try:
            self.close()
        except Exception:
            pass
