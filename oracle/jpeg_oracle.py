"""CPU oracle for K0, baseline JPEG decode (encoded images on the path).

TEST INFRASTRUCTURE ONLY - the product never imports this module.

The reference gates requests on image MIME types (APIs/1.0/base-py/ai4e_service.py:72,127-129)
and leaves decoding to the user callback it calls at ai4e_service.py:203; in a PyTorch container
that is Pillow, i.e. libjpeg-turbo.  This file restates libjpeg's baseline decoder in numpy -
Huffman decoding (ITU T.81 F.2.2), the accurate integer IDCT (jidctint.c, "islow", the library's
default), "fancy" triangle-filter chroma upsampling (jdsample.c h2v1 / h2v2) and the fixed-point
YCbCr -> RGB conversion (jdcolor.c) - so that every intermediate (coefficients, sample planes)
can be checked, not only the final pixels.

Pinned: tests/test_jpeg.py compares `decode()` bit for bit with Pillow 12.2 / libjpeg-turbo on
the committed fixtures in tests/golden/jpeg/ (made by tests/golden/make_jpeg_goldens.py).
Supported: 8-bit baseline / extended sequential (SOF0, SOF1) and progressive (SOF2) Huffman JPEG, single- or
multi-scan, 1 or 3 components, chroma subsampling 4:4:4, 4:2:2 (h2v1), 4:2:0 (h2v2), restart intervals.
Progressive scans follow jdphuff.c (DC / AC, first pass / refinement, end-of-band runs).  Arithmetic coding,
CMYK and 12-bit are rejected (ValueError) - as the CUDA path rejects them.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List

import numpy as np

ZIGZAG = np.array([
    0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14, 21,
    28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61,
    54, 47, 55, 62, 63], dtype=np.int32)  # zigzag position -> natural (row-major) index


@dataclass
class Component:
    cid: int
    h: int
    v: int
    tq: int
    td: int = 0
    ta: int = 0
    blocks_w: int = 0   # padded to whole MCUs
    blocks_h: int = 0
    width: int = 0      # real (downsampled) size in samples
    height: int = 0
    coefs: np.ndarray = field(default=None, repr=False)  # [blocks_h, blocks_w, 64] int16, natural order, quantised


@dataclass
class Frame:
    width: int
    height: int
    comps: List[Component]
    qt: Dict[int, np.ndarray]          # table id -> [64] uint16, natural order
    restart_interval: int
    hmax: int
    vmax: int
    mcus_x: int
    mcus_y: int
    progressive: bool = False
    scans: int = 0


class _Huff:
    """ITU T.81 Annex C / F.2.2.3 decoding tables."""

    def __init__(self, counts, symbols):
        self.symbols = list(symbols)
        self.mincode = [0] * 17
        self.maxcode = [-1] * 18
        self.valptr = [0] * 17
        code, k = 0, 0
        for length in range(1, 17):
            self.valptr[length] = k
            self.mincode[length] = code
            code += counts[length - 1]
            k += counts[length - 1]
            self.maxcode[length] = code - 1 if counts[length - 1] else -1
            code <<= 1
        self.maxcode[17] = 0xFFFFF


class _Bits:
    """bit reader over one entropy-coded segment (byte stuffing already removed)."""

    def __init__(self, data: bytes):
        self.d, self.pos, self.acc, self.n = data, 0, 0, 0

    def bit(self) -> int:
        if self.n == 0:
            self.acc = self.d[self.pos] if self.pos < len(self.d) else 0  # libjpeg feeds zeros past the end
            self.pos += 1
            self.n = 8
        self.n -= 1
        return (self.acc >> self.n) & 1

    def bits(self, k: int) -> int:
        v = 0
        for _ in range(k):
            v = (v << 1) | self.bit()
        return v


def _decode_symbol(br: _Bits, h: _Huff) -> int:
    code, length = br.bit(), 1
    while code > h.maxcode[length]:
        code = (code << 1) | br.bit()
        length += 1
        if length > 16:
            raise ValueError("bad Huffman code")
    return h.symbols[h.valptr[length] + code - h.mincode[length]]


def _extend(v: int, s: int) -> int:
    return v if v >= (1 << (s - 1)) else v - (1 << s) + 1


def parse(data: bytes) -> Frame:
    """markers + entropy decoding -> quantised coefficients per component."""
    if data[:2] != b"\xff\xd8":
        raise ValueError("not a JPEG (no SOI)")
    pos = 2
    qt: Dict[int, np.ndarray] = {}
    huff: Dict[tuple, _Huff] = {}
    frame = None
    ri = 0
    while pos < len(data):
        if data[pos] != 0xFF:
            raise ValueError(f"marker expected at {pos}")
        while data[pos] == 0xFF:
            pos += 1
        m = data[pos]
        pos += 1
        if m == 0xD9:
            break
        if m in (0x01,) or 0xD0 <= m <= 0xD7:
            continue
        seg_len = (data[pos] << 8) | data[pos + 1]
        seg = data[pos + 2:pos + seg_len]
        pos += seg_len
        if m == 0xDB:
            i = 0
            while i < len(seg):
                pq, tq = seg[i] >> 4, seg[i] & 15
                i += 1
                t = np.zeros(64, np.uint16)
                for k in range(64):
                    if pq:
                        t[ZIGZAG[k]] = (seg[i] << 8) | seg[i + 1]
                        i += 2
                    else:
                        t[ZIGZAG[k]] = seg[i]
                        i += 1
                qt[tq] = t
        elif m == 0xC4:
            i = 0
            while i < len(seg):
                tc, th = seg[i] >> 4, seg[i] & 15
                counts = list(seg[i + 1:i + 17])
                n = sum(counts)
                huff[(tc, th)] = _Huff(counts, seg[i + 17:i + 17 + n])
                i += 17 + n
        elif m in (0xC0, 0xC1, 0xC2):
            if seg[0] != 8:
                raise ValueError("only 8-bit samples are supported")
            h, w, nf = (seg[1] << 8) | seg[2], (seg[3] << 8) | seg[4], seg[5]
            if nf not in (1, 3):
                raise ValueError(f"{nf} components are not supported")
            comps = [Component(seg[6 + 3 * i], seg[7 + 3 * i] >> 4, seg[7 + 3 * i] & 15, seg[8 + 3 * i]) for i in range(nf)]
            hmax, vmax = max(c.h for c in comps), max(c.v for c in comps)
            if nf == 1:  # a single component is never subsampled against itself
                comps[0].h = comps[0].v = hmax = vmax = 1
            mcus_x, mcus_y = -(-w // (8 * hmax)), -(-h // (8 * vmax))
            for c in comps:
                c.blocks_w, c.blocks_h = mcus_x * c.h, mcus_y * c.v
                c.width, c.height = -(-w * c.h // hmax), -(-h * c.v // vmax)
                c.coefs = np.zeros((c.blocks_h, c.blocks_w, 64), np.int16)
            frame = Frame(w, h, comps, qt, 0, hmax, vmax, mcus_x, mcus_y, progressive=(m == 0xC2))
        elif 0xC3 <= m <= 0xCF and m not in (0xC4, 0xC8, 0xCC):
            raise ValueError("only Huffman-coded sequential or progressive JPEG is supported (SOF0, SOF1, SOF2)")
        elif m == 0xDD:
            ri = (seg[0] << 8) | seg[1]
        elif m == 0xDA:
            if frame is None:
                raise ValueError("SOS before SOF")
            ns = seg[0]
            scan_comps = []
            for i in range(ns):
                c = next(c for c in frame.comps if c.cid == seg[1 + 2 * i])
                c.td, c.ta = seg[2 + 2 * i] >> 4, seg[2 + 2 * i] & 15
                scan_comps.append(c)
            ss, se, ah, al = seg[1 + 2 * ns], seg[2 + 2 * ns], seg[3 + 2 * ns] >> 4, seg[3 + 2 * ns] & 15
            if not frame.progressive and (ss, se, ah, al) != (0, 63, 0, 0):
                raise ValueError("sequential scan with a spectral selection")
            frame.restart_interval = ri
            frame.qt = dict(qt)
            frame.scans += 1
            pos = _decode_scan(data, pos, frame, huff, scan_comps, ss, se, ah, al)
    if frame is None:
        raise ValueError("no frame header")
    return frame


def _decode_scan(data: bytes, pos: int, f: Frame, huff, comps, ss: int, se: int, ah: int, al: int) -> int:
    """one scan (sequential: all 64 coefficients of the listed components; progressive: one of the four
    kinds of jdphuff.c) into the components' coefficient arrays; returns the position of the next marker."""
    # split the entropy-coded data at restart markers, removing the stuffed zero bytes
    segs, cur = [], bytearray()
    while pos < len(data):
        b = data[pos]
        if b != 0xFF:
            cur.append(b)
            pos += 1
            continue
        nb = data[pos + 1]
        if nb == 0x00:
            cur.append(0xFF)
            pos += 2
        elif 0xD0 <= nb <= 0xD7:
            segs.append(bytes(cur))
            cur = bytearray()
            pos += 2
        elif nb == 0xFF:
            pos += 1
        else:
            break  # a real marker (EOI, DHT, SOS ...): the scan ends, `pos` stays on it
    segs.append(bytes(cur))
    if len(comps) > 1:  # interleaved: MCUs of the frame
        total = f.mcus_x * f.mcus_y

        def blocks(mcu):
            my, mx = divmod(mcu, f.mcus_x)
            for c in comps:
                for by in range(c.v):
                    for bx in range(c.h):
                        yield c, my * c.v + by, mx * c.h + bx
    else:  # one component: its own blocks in raster order, without the padding to whole MCUs
        c0 = comps[0]
        bw, bh = -(-c0.width // 8), -(-c0.height // 8)
        total = bw * bh

        def blocks(mcu):
            yield (c0, *divmod(mcu, bw))
    if f.progressive and ss > 0 and len(comps) != 1:
        raise ValueError("progressive AC scans carry one component")
    per_seg = f.restart_interval or total
    p1, m1 = 1 << al, -(1 << al)
    mcu = 0
    for seg in segs:
        br = _Bits(seg)
        pred = {c.cid: 0 for c in comps}
        eobrun = 0
        for _ in range(min(per_seg, total - mcu)):
            for c, by, bx in blocks(mcu):
                blk = c.coefs[by, bx]
                if not f.progressive:
                    dc, ac = huff[(0, c.td)], huff[(1, c.ta)]
                    s = _decode_symbol(br, dc)
                    pred[c.cid] += _extend(br.bits(s), s) if s else 0
                    blk[0] = pred[c.cid]
                    k = 1
                    while k < 64:
                        rs = _decode_symbol(br, ac)
                        r, s = rs >> 4, rs & 15
                        if s == 0:
                            if r != 15:
                                break
                            k += 16
                            continue
                        k += r
                        if k > 63:
                            raise ValueError("AC coefficient index past 63")
                        blk[ZIGZAG[k]] = _extend(br.bits(s), s)
                        k += 1
                elif ss == 0:
                    if ah == 0:  # DC, first pass (decode_mcu_DC_first)
                        s = _decode_symbol(br, huff[(0, c.td)])
                        pred[c.cid] += _extend(br.bits(s), s) if s else 0
                        blk[0] = pred[c.cid] * p1
                    elif br.bit():  # DC refinement: one more bit
                        blk[0] |= p1
                elif ah == 0:  # AC, first pass (decode_mcu_AC_first)
                    if eobrun > 0:
                        eobrun -= 1
                        continue
                    ac = huff[(1, c.ta)]
                    k = ss
                    while k <= se:
                        rs = _decode_symbol(br, ac)
                        r, s = rs >> 4, rs & 15
                        if s:
                            k += r
                            if k > 63:
                                raise ValueError("AC coefficient index past 63")
                            blk[ZIGZAG[k]] = _extend(br.bits(s), s) * p1
                        elif r == 15:
                            k += 15
                        else:
                            eobrun = (1 << r) + (br.bits(r) if r else 0) - 1
                            break
                        k += 1
                else:  # AC refinement (decode_mcu_AC_refine)
                    ac = huff[(1, c.ta)]
                    k = ss
                    if eobrun == 0:
                        while k <= se:
                            rs = _decode_symbol(br, ac)
                            r, s = rs >> 4, rs & 15
                            if s:
                                s = p1 if br.bit() else m1
                            elif r != 15:
                                eobrun = (1 << r) + (br.bits(r) if r else 0)
                                break
                            while k <= se:  # over nonzero coefficients (one correction bit each) and r zero ones
                                z = ZIGZAG[k]
                                if blk[z] != 0:
                                    if br.bit() and (int(blk[z]) & p1) == 0:
                                        blk[z] += p1 if blk[z] >= 0 else m1
                                else:
                                    r -= 1
                                    if r < 0:
                                        break
                                k += 1
                            if s:
                                if k > 63:
                                    raise ValueError("AC coefficient index past 63")
                                blk[ZIGZAG[k]] = s
                            k += 1
                    if eobrun > 0:  # the rest of the band: correction bits only
                        while k <= se:
                            z = ZIGZAG[k]
                            if blk[z] != 0 and br.bit() and (int(blk[z]) & p1) == 0:
                                blk[z] += p1 if blk[z] >= 0 else m1
                            k += 1
                        eobrun -= 1
            mcu += 1
        if mcu >= total:
            break
    return pos


# ------------------------------------------------------------------------------------------
# reconstruction: jidctint.c (islow), jdsample.c (fancy upsampling), jdcolor.c
# ------------------------------------------------------------------------------------------
CONST_BITS, PASS1_BITS = 13, 2
FIX_0_298631336, FIX_0_390180644, FIX_0_541196100 = 2446, 3196, 4433
FIX_0_765366865, FIX_0_899976223, FIX_1_175875602 = 6270, 7373, 9633
FIX_1_501321110, FIX_1_847759065, FIX_1_961570560 = 12299, 15137, 16069
FIX_2_053119869, FIX_2_562915447, FIX_3_072711026 = 16819, 20995, 25172


def _descale(x, n):
    return (x + (1 << (n - 1))) >> n


def _idct_1d(i0, i1, i2, i3, i4, i5, i6, i7, shift):
    """one pass of jpeg_idct_islow over eight int32 arrays; returns the eight outputs, descaled"""
    z1 = (i2 + i6) * FIX_0_541196100
    tmp2 = z1 + i6 * (-FIX_1_847759065)
    tmp3 = z1 + i2 * FIX_0_765366865
    tmp0 = (i0 + i4) << CONST_BITS
    tmp1 = (i0 - i4) << CONST_BITS
    tmp10, tmp13, tmp11, tmp12 = tmp0 + tmp3, tmp0 - tmp3, tmp1 + tmp2, tmp1 - tmp2
    t0, t1, t2, t3 = i7, i5, i3, i1
    z1, z2, z3, z4 = t0 + t3, t1 + t2, t0 + t2, t1 + t3
    z5 = (z3 + z4) * FIX_1_175875602
    t0 = t0 * FIX_0_298631336
    t1 = t1 * FIX_2_053119869
    t2 = t2 * FIX_3_072711026
    t3 = t3 * FIX_1_501321110
    z1 = z1 * (-FIX_0_899976223)
    z2 = z2 * (-FIX_2_562915447)
    z3 = z3 * (-FIX_1_961570560) + z5
    z4 = z4 * (-FIX_0_390180644) + z5
    t0 = t0 + z1 + z3
    t1 = t1 + z2 + z4
    t2 = t2 + z2 + z3
    t3 = t3 + z1 + z4
    return [_descale(v, shift) for v in (tmp10 + t3, tmp11 + t2, tmp12 + t1, tmp13 + t0,
                                         tmp13 - t0, tmp12 - t1, tmp11 - t2, tmp10 - t3)]


def idct_islow(coefs: np.ndarray, q: np.ndarray) -> np.ndarray:
    """[..., 64] quantised int16 coefficients (natural order) x [64] table -> [..., 8, 8] uint8 samples.
    The library's zero-AC shortcuts give the same numbers as the full computation, so none here."""
    x = (coefs.astype(np.int32) * q.astype(np.int32)).reshape(coefs.shape[:-1] + (8, 8))
    # pass 1: columns (input index = row), results scaled up by 2^PASS1_BITS
    cols = _idct_1d(*[x[..., r, :] for r in range(8)], CONST_BITS - PASS1_BITS)
    ws = np.stack(cols, axis=-2)  # [..., 8 rows, 8 cols]
    # pass 2: rows
    rows = _idct_1d(*[ws[..., :, c] for c in range(8)], CONST_BITS + PASS1_BITS + 3)
    out = np.stack(rows, axis=-1)
    # range_limit[(x) & RANGE_MASK]: a 1024-entry table centred on 128 (jdmaster.c prepare_range_limit_table)
    idx = out & 1023
    lim = np.where(idx < 128, idx + 128, np.where(idx < 512, 255, np.where(idx < 896, 0, idx - 896)))
    return lim.astype(np.uint8)


def component_plane(c: Component, q: np.ndarray) -> np.ndarray:
    s = idct_islow(c.coefs, q)  # [bh, bw, 8, 8]
    plane = s.transpose(0, 2, 1, 3).reshape(c.blocks_h * 8, c.blocks_w * 8)
    return plane[:c.height, :c.width]


def upsample_box(p: np.ndarray, hr: int, vr: int) -> np.ndarray:
    """jdsample.c int_upsample / h2v1_upsample / h2v2_upsample: every sample replicated hr x vr times"""
    return np.repeat(np.repeat(p, vr, axis=0), hr, axis=1)


def upsample_h1v2(p: np.ndarray) -> np.ndarray:
    """jdsample.c h1v2_fancy_upsample (libjpeg-turbo): 3/4 nearer row + 1/4 the next nearest, bias 1 (upper output row) /
    2 (lower); the row above the first / below the last is that row itself"""
    a = p.astype(np.int32)
    above = np.concatenate([a[:1], a[:-1]], axis=0)
    below = np.concatenate([a[1:], a[-1:]], axis=0)
    out = np.empty((2 * a.shape[0], a.shape[1]), np.int32)
    out[0::2] = (3 * a + above + 1) >> 2
    out[1::2] = (3 * a + below + 2) >> 2
    return out.astype(np.uint8)


def upsample_h2v1(p: np.ndarray) -> np.ndarray:
    """jdsample.c h2v1_fancy_upsample: 3/4 nearer + 1/4 further, rounding alternates 1 / 2"""
    a = p.astype(np.int32)
    h, w = a.shape
    out = np.empty((h, 2 * w), np.int32)
    left = np.concatenate([a[:, :1], a[:, :-1]], axis=1)
    right = np.concatenate([a[:, 1:], a[:, -1:]], axis=1)
    out[:, 0::2] = (3 * a + left + 1) >> 2
    out[:, 1::2] = (3 * a + right + 2) >> 2
    out[:, 0] = a[:, 0]
    out[:, -1] = a[:, -1]
    return out.astype(np.uint8)


def upsample_h2v2(p: np.ndarray) -> np.ndarray:
    """jdsample.c h2v2_fancy_upsample: 9/16, 3/16, 3/16, 1/16; the row above / below the image is the
    edge row itself (jdmainct.c context rows)."""
    a = p.astype(np.int32)
    h, w = a.shape
    above = np.concatenate([a[:1], a[:-1]], axis=0)
    below = np.concatenate([a[1:], a[-1:]], axis=0)
    out = np.empty((2 * h, 2 * w), np.int32)
    for v, other in ((0, above), (1, below)):
        colsum = 3 * a + other  # "thiscolsum"
        last = np.concatenate([colsum[:, :1], colsum[:, :-1]], axis=1)
        nxt = np.concatenate([colsum[:, 1:], colsum[:, -1:]], axis=1)
        even = (3 * colsum + last + 8) >> 4
        odd = (3 * colsum + nxt + 7) >> 4
        even[:, 0] = (colsum[:, 0] * 4 + 8) >> 4
        odd[:, -1] = (colsum[:, -1] * 4 + 7) >> 4
        out[v::2, 0::2] = even
        out[v::2, 1::2] = odd
    return out.astype(np.uint8)


def ycc_to_rgb(y: np.ndarray, cb: np.ndarray, cr: np.ndarray) -> np.ndarray:
    """jdcolor.c build_ycc_rgb_table / ycc_rgb_convert (SCALEBITS = 16)"""
    one_half = 1 << 15
    fix = lambda v: int(v * 65536 + 0.5)
    x = np.arange(256, dtype=np.int64) - 128
    cr_r = (fix(1.40200) * x + one_half) >> 16
    cb_b = (fix(1.77200) * x + one_half) >> 16
    cr_g = -fix(0.71414) * x
    cb_g = -fix(0.34414) * x + one_half
    yy = y.astype(np.int64)
    r = yy + cr_r[cr]
    g = yy + ((cb_g[cb] + cr_g[cr]) >> 16)
    b = yy + cb_b[cb]
    return np.clip(np.stack([r, g, b], axis=-1), 0, 255).astype(np.uint8)


def reconstruct(f: Frame) -> np.ndarray:
    planes = []
    for c in f.comps:
        p = component_plane(c, f.qt[c.tq])
        hr, vr = f.hmax // c.h, f.vmax // c.v
        if f.hmax % c.h or f.vmax % c.v or hr not in (1, 2, 4) or vr not in (1, 2):
            raise ValueError(f"sampling ratio {f.hmax}/{c.h} x {f.vmax}/{c.v} is not supported")
        # the choice of jdsample.c jinit_upsampler: the triangle filters where libjpeg has them - and, for the two
        # horizontal ones, only when the component is more than two samples wide - replication otherwise
        if (hr, vr) == (1, 1):
            pass
        elif (hr, vr) == (2, 1) and c.width > 2:
            p = upsample_h2v1(p)
        elif (hr, vr) == (1, 2):
            p = upsample_h1v2(p)
        elif (hr, vr) == (2, 2) and c.width > 2:
            p = upsample_h2v2(p)
        else:
            p = upsample_box(p, hr, vr)
        planes.append(p[:f.height, :f.width])
    if len(planes) == 1:
        return np.repeat(planes[0][:, :, None], 3, axis=2)
    return ycc_to_rgb(*planes)


def decode(data: bytes) -> np.ndarray:
    """JPEG bytes -> uint8 [H, W, 3] RGB, as PIL.Image.open(...).convert('RGB')"""
    return reconstruct(parse(data))
