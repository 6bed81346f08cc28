"""A small PNG ENCODER for the tests: every colour type / bit depth of the PNG specification, plain or Adam7-interlaced,
with a chosen (or random) filter type per line, so that the decoder in csrc/png.cu sees what Pillow never writes
(interlaced files, 16-bit RGB, sub-byte depths, all five filters on every kind of line).  Test infrastructure; the
expected pixels always come from Pillow reading the same bytes."""
import struct
import zlib

import numpy as np

CHANNELS = {0: 1, 2: 3, 3: 1, 4: 2, 6: 4}
ADAM7 = ((0, 0, 8, 8), (4, 0, 8, 8), (0, 4, 4, 8), (2, 0, 4, 4), (0, 2, 2, 4), (1, 0, 2, 2), (0, 1, 1, 2))  # x0, y0, dx, dy


def _chunk(kind: bytes, body: bytes) -> bytes:
    return struct.pack(">I", len(body)) + kind + body + struct.pack(">I", zlib.crc32(kind + body) & 0xFFFFFFFF)


def _pack_rows(samples: np.ndarray, depth: int) -> list[bytes]:
    """samples [h, w, channels] of ints below 2**depth -> the unfiltered bytes of each line"""
    h, w, c = samples.shape
    if depth == 8:
        return [samples[r].astype(np.uint8).tobytes() for r in range(h)]
    if depth == 16:
        return [samples[r].astype(">u2").tobytes() for r in range(h)]
    per = 8 // depth
    rows = []
    for r in range(h):
        v = samples[r, :, 0].astype(np.uint8)
        pad = (-w) % per
        v = np.concatenate([v, np.zeros(pad, np.uint8)]).reshape(-1, per)
        out = np.zeros(len(v), np.uint8)
        for k in range(per):
            out |= v[:, k] << (8 - depth * (k + 1))
        rows.append(out.tobytes())
    return rows


def _paeth(a, b, c):
    p = a + b - c
    pa, pb, pc = abs(p - a), abs(p - b), abs(p - c)
    return a if pa <= pb and pa <= pc else (b if pb <= pc else c)


def _filter(rows: list[bytes], fbpp: int, choose) -> bytes:
    out = bytearray()
    prev = bytes(len(rows[0])) if rows else b""
    for r, row in enumerate(rows):
        ft = choose(r)
        line = bytearray(len(row))
        for i, x in enumerate(row):
            a = row[i - fbpp] if i >= fbpp else 0
            b = prev[i]
            c = prev[i - fbpp] if i >= fbpp else 0
            pred = (0, a, b, (a + b) >> 1, _paeth(a, b, c))[ft]
            line[i] = (x - pred) & 0xFF
        out.append(ft)
        out += line
        prev = row
    return bytes(out)


def encode(samples: np.ndarray, depth: int, ctype: int, interlace: bool = False, palette: bytes | None = None,
           filters="random", seed: int = 0, extra_chunks=(), idat_split: int = 0) -> bytes:
    """samples: [h, w, channels(ctype)] integer array with values below 2**depth"""
    h, w, c = samples.shape
    assert c == CHANNELS[ctype]
    rng = np.random.default_rng(seed)
    choose = (lambda r: int(rng.integers(0, 5))) if filters == "random" else (lambda r: int(filters))
    fbpp = max(1, depth * c // 8)
    raw = b""
    if interlace:
        for x0, y0, dx, dy in ADAM7:
            sub = samples[y0::dy, x0::dx]
            if sub.shape[0] and sub.shape[1]:
                raw += _filter(_pack_rows(sub, depth), fbpp, choose)
    else:
        raw = _filter(_pack_rows(samples, depth), fbpp, choose)
    comp = zlib.compress(raw, 6)
    out = b"\x89PNG\r\n\x1a\n" + _chunk(b"IHDR", struct.pack(">IIBBBBB", w, h, depth, ctype, 0, 0, 1 if interlace else 0))
    if palette is not None:
        out += _chunk(b"PLTE", palette)
    for kind, body in extra_chunks:
        out += _chunk(kind, body)
    if idat_split:  # the IDAT stream may be cut anywhere
        for i in range(0, len(comp), idat_split):
            out += _chunk(b"IDAT", comp[i:i + idat_split])
    else:
        out += _chunk(b"IDAT", comp)
    return out + _chunk(b"IEND", b"")


VALID = [(0, d) for d in (1, 2, 4, 8, 16)] + [(2, 8), (2, 16)] + [(3, d) for d in (1, 2, 4, 8)] + [(4, 8), (4, 16), (6, 8), (6, 16)]


def random_image(rng, h, w, depth, ctype):
    """-> (samples, palette or None); smooth along x so that the filters matter"""
    c = CHANNELS[ctype]
    hi = 1 << depth
    base = np.cumsum(rng.integers(-3, 4, (h, w, c)) * max(1, hi // 64), axis=1) + rng.integers(0, hi, (h, 1, c))
    samples = np.mod(base, hi).astype(np.int64)
    if ctype == 3:
        n = int(rng.integers(1, hi + 1))          # a palette shorter than the index range: indices past it read zeros
        return samples, rng.integers(0, 256, n * 3, dtype=np.uint8).tobytes()
    if depth == 16 and ctype == 0:
        samples[::2] = samples[::2] % 300         # values around the 255 clip of Pillow's I;16 -> RGB
    return samples, None
