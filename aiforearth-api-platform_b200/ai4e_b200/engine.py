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
import threading

import numpy as np

from . import _lib as L
from .models import Program

_REAPER_GUARD = threading.Lock()


def _u8_hwc(image) -> np.ndarray:
    a = np.asarray(image)
    if a.dtype != np.uint8 or a.ndim != 3:
        raise ValueError(f"expected a uint8 HWC image, got dtype {a.dtype} shape {a.shape}")
    # dense pixels and rows that go FORWARD in memory by at least a row; anything else (a flipped or
    # transposed view) is copied - the C side reads rows at +row_stride
    if a.strides[2] != 1 or a.strides[1] != a.shape[2] or a.strides[0] < a.shape[1] * a.shape[2]:
        a = np.ascontiguousarray(a)
    return a


def _vp(a: np.ndarray):
    return C.c_void_p(a.ctypes.data)


class Pending:
    """Result of `classify_later`: resolved by the engine's reaper thread, never by a thread blocked
    per task.  `APIService` understands it: a task function that returns a Pending keeps its request
    counted as in flight until the result (and the `then` callback) is through, and an error in
    either marks the task failed exactly like an exception raised by the function itself."""

    __slots__ = ("ticket", "idx", "prob", "error", "_then", "_done", "_event", "_watchers")
    _guard = threading.Lock()  # one lock for all Pendings: held for a few instructions at a time

    def __init__(self, then=None):
        self.ticket = None
        self.idx = self.prob = self.error = None
        self._then = then
        self._done = False
        self._event = None      # made only if somebody blocks in result()
        self._watchers = None

    def done(self) -> bool:
        return self._done

    def result(self, timeout=None):
        if not self._done:
            with Pending._guard:
                if not self._done and self._event is None:
                    self._event = threading.Event()
                ev = self._event
            if not self._done and not ev.wait(timeout):
                raise TimeoutError("classify_later: no result yet")
        if self.error is not None:
            raise self.error
        return self.idx, self.prob

    def add_done_callback(self, fn):
        """fn(pending) once the result and `then` are through (at once if they already are)."""
        with Pending._guard:
            if not self._done:
                if self._watchers is None:
                    self._watchers = [fn]
                else:
                    self._watchers.append(fn)
                return
        fn(self)

    @classmethod
    def gather(cls, pendings, then=None):
        """one Pending for a task made of several images: resolves when all of `pendings` have, with
        idx / prob stacked in order ([n, 5]); `then(idx, prob)` runs once; the first error wins."""
        pendings = list(pendings)
        whole = cls(then)
        left = [len(pendings)]
        lock = threading.Lock()

        def one_done(_p):
            with lock:
                left[0] -= 1
                last = left[0] == 0
            if last:
                err = next((q.error for q in pendings if q.error is not None), None)
                if err is not None:
                    whole._resolve(None, None, err)
                else:
                    whole._resolve(np.stack([q.idx for q in pendings]), np.stack([q.prob for q in pendings]), None)

        if not pendings:
            whole._resolve(np.empty((0, L.TOPK), np.int32), np.empty((0, L.TOPK), np.float32), None)
        for q in pendings:
            q.add_done_callback(one_done)
        return whole

    def _resolve(self, idx, prob, error):
        self.idx, self.prob, self.error = idx, prob, error
        if error is None and self._then is not None:
            try:
                self._then(idx, prob)
            except BaseException as e:  # noqa: BLE001 - reported through the Pending, like a task function's exception
                self.error = e
        with Pending._guard:
            self._done = True
            watchers, self._watchers = self._watchers, None
            ev = self._event
        if ev is not None:
            ev.set()
        if watchers:
            for fn in watchers:
                try:
                    fn(self)
                except BaseException:  # noqa: BLE001 - a watcher must not kill the reaper
                    pass


def require_decodable(data: bytes):
    """For a `request_processing_function`: returns ("jpeg" | "png", height, width) when the engine can
    decode the payload and otherwise ends the request with 415 Unsupported Media Type - while the
    request is still open, instead of accepting a task that can only fail later (the reference's
    content gate, ai4e_service.py:127-129, sees the MIME type only)."""
    from . import ops
    try:
        return ops.image_probe(data)
    except L.Ai4eError as e:
        from werkzeug.exceptions import UnsupportedMediaType
        raise UnsupportedMediaType(description=str(e)) from e


class _Base:
    """what Engine and Pool share: both C handles answer the same calls under two prefixes."""
    _h = None
    _prefix = "ai4e_engine_"

    def _fn(self, name):
        return getattr(self._lib, self._prefix + name)

    def _results(self, n=1):
        return (np.empty((n, L.TOPK), np.int32), np.empty((n, L.TOPK), np.float32))

    # ---- many images per call, completions through one queue
    def submit_many(self, images, tags=None, notify: bool = False) -> np.ndarray:
        """`images`: a uint8 [N,H,W,C] array or a sequence of equal-sized uint8 HWC arrays -> uint64
        tickets [N], in ONE C call.  `tags` (uint64 [N]) come back from `poll`; with `notify` every
        ticket is announced there when its batch has finished."""
        if isinstance(images, tuple):
            # (stack, index): images stack[index[0]], stack[index[1]], ... - pointers by vector arithmetic,
            # no per-image Python work (what a load generator or a frame ring uses)
            a, pick = images
            if a.dtype != np.uint8 or a.ndim != 4 or not a.flags.c_contiguous:
                raise ValueError("expected a C-contiguous uint8 [N,H,W,C] stack")
            pick = np.asarray(pick, np.int64)
            if pick.size and (pick.min() < 0 or pick.max() >= a.shape[0]):
                raise IndexError("submit_many: index outside the stack")
            n, (h, w, c) = pick.size, a.shape[1:]
            if n == 0:
                return np.empty(0, np.uint64)
            ptrs = (np.uint64(a.ctypes.data) + pick.astype(np.uint64) * np.uint64(a.strides[0]))
            row_stride, keep = a.strides[1], a
        elif isinstance(images, np.ndarray) and images.ndim == 4:
            a = images
            if a.dtype != np.uint8:
                raise ValueError("expected uint8 images")
            if a.strides[3] != 1 or a.strides[2] != a.shape[3] or a.strides[1] < a.shape[2] * a.shape[3] or a.strides[0] <= 0:
                a = np.ascontiguousarray(a)
            n, h, w, c = a.shape
            ptrs = a.ctypes.data + np.arange(n, dtype=np.uint64) * np.uint64(a.strides[0])
            row_stride, keep = a.strides[1], a
        else:
            keep = [_u8_hwc(im) for im in images]
            n = len(keep)
            if n == 0:
                return np.empty(0, np.uint64)
            h, w, c = keep[0].shape
            row_stride = keep[0].strides[0]
            for im in keep:
                if im.shape != (h, w, c):
                    raise ValueError("submit_many: all images of one call must have the same size")
                if im.strides[0] != row_stride:
                    keep = [np.ascontiguousarray(x) for x in keep]
                    row_stride = w * c
                    break
            ptrs = np.fromiter((im.__array_interface__["data"][0] for im in keep), np.uint64, n)
        tickets = np.empty(n, np.uint64)
        t = None if tags is None else np.ascontiguousarray(tags, np.uint64)
        if t is not None and t.size != n:
            raise ValueError("submit_many: one tag per image")
        L.check(self._fn("submit_many")(self._h, _vp(ptrs), n, h, w, c, row_stride, _vp(t) if t is not None else None,
                                        int(bool(notify)), _vp(tickets)))
        del keep
        return tickets

    def poll(self, capacity: int = 256, timeout_s: float | None = 0.1):
        """announced completions, oldest first -> (tickets u64 [n], tags u64 [n], status i32 [n],
        idx i32 [n,5], prob f32 [n,5]); n may be 0 after `timeout_s` (None = wait for ever)."""
        tk, tg = np.empty(capacity, np.uint64), np.empty(capacity, np.uint64)
        st = np.empty(capacity, np.int32)
        idx, prob = self._results(capacity)
        n = C.c_int()
        us = -1 if timeout_s is None else int(timeout_s * 1e6)
        L.check(self._fn("poll")(self._h, capacity, us, _vp(tk), _vp(tg), _vp(st), _vp(idx), _vp(prob), C.byref(n)))
        n = n.value
        return tk[:n], tg[:n], st[:n], idx[:n], prob[:n]

    # ---- deferred classification: no thread blocked per task
    def classify_later(self, image, then=None) -> Pending:
        """submit one image and return at once; `then(idx, prob)` runs on the reaper thread when the
        micro-batch has finished.  A task function returns the Pending to `APIService`."""
        p = Pending(then)
        if getattr(self, "_reaper", None) is None:
            self._ensure_reaper()
        a = _u8_hwc(image)
        with self._pend_lock:
            self._pend_seq += 1
            tag = self._pend_seq
            self._pending[tag] = p
        # one image: the same C entry point as submit_many, without its array plumbing
        ptr = C.c_void_p(a.__array_interface__["data"][0])
        tag_c, ticket = C.c_uint64(tag), C.c_uint64()
        rc = self._fn_submit_many(self._h, C.byref(ptr), 1, a.shape[0], a.shape[1], a.shape[2], a.strides[0],
                                  C.byref(tag_c), 1, C.byref(ticket))
        if rc != L.OK:
            with self._pend_lock:
                self._pending.pop(tag, None)
            L.check(rc)
        p.ticket = ticket.value
        return p

    def classify_many_later(self, images, then=None):
        """`classify_later` for N images in one C call; `then(i, idx, prob)` per image."""
        imgs = images if isinstance(images, np.ndarray) else list(images)
        n = len(imgs)
        self._ensure_reaper()
        pend = [Pending((lambda idx, prob, i=i: then(i, idx, prob)) if then else None) for i in range(n)]
        with self._pend_lock:
            tags = np.arange(self._pend_seq + 1, self._pend_seq + 1 + n, dtype=np.uint64)
            self._pend_seq += n
            for t, p in zip(tags.tolist(), pend):
                self._pending[t] = p
        try:
            tickets = self.submit_many(imgs, tags=tags, notify=True)
        except BaseException:
            with self._pend_lock:
                for t in tags.tolist():
                    self._pending.pop(t, None)
            raise
        for p, t in zip(pend, tickets.tolist()):
            p.ticket = t
        return pend

    def _ensure_reaper(self):
        if getattr(self, "_reaper", None) is not None:
            return
        with _REAPER_GUARD:
            if getattr(self, "_reaper", None) is not None:
                return
            self._pend_lock = threading.Lock()
            self._pending, self._pend_seq = {}, 0
            self._fn_submit_many = self._fn("submit_many")
            self._reaper_stop = False
            self._reaper = threading.Thread(target=self._reap, daemon=True, name="ai4e-reaper")
            self._reaper.start()

    def _reap(self):
        cap = 1024
        tg, st = np.empty(cap, np.uint64), np.empty(cap, np.int32)
        idx, prob = self._results(cap)
        n_c = C.c_int()
        poll, args = self._fn("poll"), (cap, 50_000, None, _vp(tg), _vp(st), _vp(idx), _vp(prob), C.byref(n_c))
        while not self._reaper_stop:
            if poll(self._h, *args) != L.OK:  # the handle went away under us: stop
                return
            n = n_c.value
            if not n:
                continue
            tags, status = tg[:n].tolist(), st[:n].tolist()
            ii, pp = idx[:n].copy(), prob[:n].copy()  # rows handed to the callbacks outlive the next poll
            with self._pend_lock:
                ps = [self._pending.pop(t, None) for t in tags]
            for i, p in enumerate(ps):
                if p is None:
                    continue
                if status[i] == L.OK:
                    p._resolve(ii[i], pp[i], None)
                else:
                    p._resolve(None, None, L.Ai4eError(status[i], "the micro-batch of this task failed on the GPU"))

    def _stop_reaper(self):
        r = getattr(self, "_reaper", None)
        if r is not None:
            self._reaper_stop = True
            r.join(timeout=1.0)
            self._reaper = None
            with self._pend_lock:
                left, self._pending = list(self._pending.values()), {}
            for p in left:
                p._resolve(None, None, L.Ai4eError(L.ERR_STATE, "engine closed before the task finished"))

    def wait_many(self, tickets: np.ndarray):
        """-> (idx [n, 5] int32, prob [n, 5] float32) for the tickets, in order."""
        t = np.ascontiguousarray(tickets, np.uint64)
        idx, prob = self._results(t.size)
        L.check(self._fn("wait_many")(self._h, _vp(t), t.size, _vp(idx), _vp(prob)))
        return idx, prob

    def wait(self, ticket: int, want_logits: bool = False, timeout_s: float | None = None):
        """blocks until the ticket's micro-batch has finished.  `timeout_s` None = the library's
        default limit (AI4E_B200_WAIT_TIMEOUT_MS, 60 s); raises Ai4eError(ERR_TIMEOUT) when it passes."""
        idx, prob = self._results(1)
        logits = np.empty((self.num_classes,), np.float32) if want_logits else None
        lp = _vp(logits) if want_logits else None
        if timeout_s is None:
            L.check(self._fn("wait")(self._h, ticket, _vp(idx), _vp(prob), lp))
        else:
            L.check(self._fn("wait_for")(self._h, ticket, int(timeout_s * 1e6), _vp(idx), _vp(prob), lp))
        return (idx[0], prob[0], logits) if want_logits else (idx[0], prob[0])

    def classify(self, image, want_logits: bool = False):
        """submit + wait for one image: the call a user `func` makes."""
        return self.wait(self.submit(image), want_logits)

    def submit_encoded(self, data: bytes, tag: int = 0) -> int:
        """one encoded image (baseline / progressive JPEG or PNG, sniffed from its first bytes) -> ticket.
        The sequential part of the decode (Huffman / inflate + un-filter) runs on this thread with
        the GIL released.  Unsupported payloads raise Ai4eError(ERR_UNSUPPORTED) and queue nothing."""
        buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
        t = C.c_uint64()
        L.check(self._fn("submit_image")(self._h, buf, len(data), tag, C.byref(t)))
        return t.value

    def classify_encoded(self, data: bytes, want_logits: bool = False):
        """blocking: submit_encoded + wait."""
        return self.wait(self.submit_encoded(data), want_logits)

    def submit_tile(self, tile_u8_hwc: np.ndarray, chip: int = 256) -> np.ndarray:
        """every chip x chip chip of a [H, W, C] uint8 tile (row-major over the chip grid) in one C
        call -> uint64 tickets [gh * gw].  The tile goes to the GPU once and the preprocessing kernel
        reads the chips in place; a Pool deals contiguous chip ranges to its replicas.  A page-locked
        tile must stay unchanged until its tickets are done."""
        a = np.asarray(tile_u8_hwc)
        if a.dtype != np.uint8 or a.ndim != 3 or a.strides[2] != 1 or a.strides[1] != a.shape[2] or a.strides[0] < a.shape[1] * a.shape[2]:
            raise ValueError("expected a uint8 [H, W, C] tile with dense pixels and forward rows")
        gh, gw = a.shape[0] // chip, a.shape[1] // chip
        tickets = np.empty(max(gh * gw, 1), np.uint64)
        n = C.c_int32()
        L.check(self._fn("submit_tile")(self._h, _vp(a), a.shape[0], a.shape[1], a.shape[2], a.strides[0], chip,
                                        _vp(tickets), tickets.size, C.byref(n)))
        return tickets[:n.value]

    def submit_tile_summary(self, tile_u8_hwc: np.ndarray, chip: int = 256):
        """tile -> chips -> classify -> aggregate with no host hop between the stages: like `submit_tile`, but
        the chips' results are folded into one summary ON THE GPU and one ticket stands for the whole tile.
        -> (tile_ticket, (gh, gw)); results with `wait_tile_summary`."""
        a = np.asarray(tile_u8_hwc)
        if a.dtype != np.uint8 or a.ndim != 3 or a.strides[2] != 1 or a.strides[1] != a.shape[2] or a.strides[0] < a.shape[1] * a.shape[2]:
            raise ValueError("expected a uint8 [H, W, C] tile with dense pixels and forward rows")
        t, n = C.c_uint64(), C.c_int32()
        L.check(self._fn("submit_tile_summary")(self._h, _vp(a), a.shape[0], a.shape[1], a.shape[2], a.strides[0], chip,
                                                C.byref(t), C.byref(n)))
        gh, gw = a.shape[0] // chip, a.shape[1] // chip
        assert n.value == gh * gw
        return t.value, (gh, gw)

    def wait_tile_summary(self, tile_ticket: int, grid: tuple[int, int]) -> dict:
        """-> {"labels": int32 [gh, gw] top-1 class per chip, "label_prob": float32 [gh, gw], "hist": uint32
        [classes] chips per top-1 class, "mass_q32": uint64 [classes] top-5 probability mass per class in 32.32
        fixed point (exact integer sums), "mass": the same as float64}."""
        gh, gw = grid
        n = gh * gw
        labels, prob = np.empty(n, np.int32), np.empty(n, np.float32)
        hist, mass = np.empty(self.num_classes, np.uint32), np.empty(self.num_classes, np.uint64)
        got = C.c_int32()
        L.check(self._fn("wait_tile_summary")(self._h, C.c_uint64(int(tile_ticket)), _vp(labels), _vp(prob), n, _vp(hist),
                                              _vp(mass), C.byref(got)))
        assert got.value == n
        return {"labels": labels.reshape(gh, gw), "label_prob": prob.reshape(gh, gw), "hist": hist, "mass_q32": mass,
                "mass": mass.astype(np.float64) / 4294967296.0}

    def submit_batch(self, images_u8: np.ndarray):
        """dense [B,H,W,C] uint8 host array -> tickets (consecutive; a range for an Engine, an array
        for a Pool).  Page-locked arrays are read in place by the H2D copy: keep them unchanged until
        the last ticket has been waited for."""
        a = images_u8
        if a.dtype != np.uint8 or a.ndim != 4 or not a.flags.c_contiguous:
            raise ValueError("expected a C-contiguous uint8 [B,H,W,C] array")
        t = C.c_uint64()
        L.check(self._fn("submit_batch")(self._h, _vp(a), a.shape[0], a.shape[1], a.shape[2], a.shape[3], C.byref(t)))
        if self._prefix == "ai4e_engine_":
            return range(t.value, t.value + a.shape[0])
        return np.uint64(t.value) + np.arange(a.shape[0], dtype=np.uint64)


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

    def submit_jpeg(self, data: bytes, tag: int = 0) -> int:
        """one JPEG (bytes; baseline or progressive) -> ticket.  The Huffman decode runs on this thread with the GIL
        released; the GPU reconstructs the pixels (bit-identical to Pillow) in front of the model."""
        buf = (C.c_uint8 * len(data)).from_buffer_copy(data)
        t = C.c_uint64()
        L.check(self._lib.ai4e_engine_submit_jpeg(self._h, buf, len(data), tag, C.byref(t)))
        return t.value

    def classify_jpeg(self, data: bytes, want_logits: bool = False):
        """blocking: submit_jpeg + wait."""
        return self.wait(self.submit_jpeg(data), want_logits)

    def profile_layers(self, d_images, batch: int, h: int, w: int, reps: int = 5) -> np.ndarray:
        """mean device ms of [preprocess, layer 0, layer 1, ...] on a device-resident batch."""
        n = len(self.program.layers) + 1
        out = np.zeros(n, np.float32)
        L.check(self._lib.ai4e_engine_profile(self._h, C.c_void_p(d_images.data_ptr()), batch, h, w, reps,
                                              out.ctypes.data_as(C.POINTER(C.c_float)), n))
        return out

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

    def inject_fault(self, n_batches: int = 1):
        """fault drill: the next `n_batches` micro-batches fail like a sticky CUDA error (tests)."""
        L.check(self._lib.ai4e_engine_debug_inject_fault(self._h, n_batches))

    def close(self):
        self._stop_reaper()
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
    _prefix = "ai4e_pool_"

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

    def openloop(self, images_u8: np.ndarray, rate_per_s: float, seconds: float, feeders: int = 4, seed: int = 0) -> dict:
        """Measurement aid (BASELINE.json configs[4]): Poisson arrivals at `rate_per_s` for `seconds`, generated by
        `feeders` native threads through the pool's own submit_many / poll (no interpreter on the path; nobody else
        may poll meanwhile).  images_u8: dense [N, H, W, C], cycled.  -> offered / completed / failed counts, achieved
        rate, latency percentiles (scheduled arrival -> result announced, ms) and how late the feeders ran."""
        a = np.ascontiguousarray(images_u8)
        assert a.dtype == np.uint8 and a.ndim == 4
        cap = int(rate_per_s * seconds) + 1
        lat = np.empty(cap, np.float32)
        offered, completed, failed = C.c_int64(), C.c_int64(), C.c_int64()
        span, late = C.c_double(), C.c_double()
        L.check(self._lib.ai4e_pool_openloop(self._h, _vp(a), a.shape[0], a.shape[1], a.shape[2], a.shape[3], float(rate_per_s),
                                             float(seconds), int(feeders), int(seed), _vp(lat), cap, C.byref(offered),
                                             C.byref(completed), C.byref(failed), C.byref(span), C.byref(late)))
        ok = np.sort(lat[:offered.value][lat[:offered.value] >= 0])
        q = lambda f: float(ok[int((len(ok) - 1) * f)]) if len(ok) else None  # noqa: E731
        return {"offered_per_s": float(rate_per_s), "tasks": offered.value, "completed": completed.value, "failed": failed.value,
                "achieved_per_s": completed.value / span.value if span.value > 0 else 0.0, "p50_ms": q(0.50), "p90_ms": q(0.90),
                "p99_ms": q(0.99), "max_ms": q(1.0), "feeder_max_lateness_ms": late.value, "feeders": int(feeders)}

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
        self._stop_reaper()
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
