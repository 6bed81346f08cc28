"""A progressive-JPEG ENCODER for the tests: re-codes the quantised coefficients of an existing JPEG (as the oracle
parsed them) into a progressive file with ANY legal scan script - the scripts other encoders write (mozjpeg's, or
no successive approximation at all, or DC scans of single components), which Pillow never produces - optionally with a
restart interval.  Entropy coding follows ITU T.81 annex G.1.2 (the procedures libjpeg's jcphuff.c implements):
DC first / refinement, AC first with end-of-band runs, AC refinement with buffered correction bits.  Every scan uses one
fixed pair of Huffman tables (sixteen 5-bit DC codes; all 256 AC symbols in 8 and 9 bits), legal if wasteful.  Test infrastructure: the
expected pixels always come from Pillow decoding the same bytes."""
import struct

import numpy as np

from oracle import jpeg_oracle as jo

def _codes(counts, symbols):
    table, code, k = {}, 0, 0
    for length in range(1, 17):
        for _ in range(counts[length - 1]):
            table[symbols[k]] = (code, length)
            code += 1
            k += 1
        code <<= 1
    return table


_AC_COUNTS = [0] * 7 + [128, 128] + [0] * 7         # 128 codes of 8 bits, 128 of 9 (Kraft sum 0.75, no all-ones code)
_AC_SYMBOLS = list(range(256))
_DC_COUNTS = [0] * 4 + [16] + [0] * 11              # the sixteen DC categories, 5 bits each (libjpeg refuses DC symbols above 15)
_DC_SYMBOLS = list(range(16))
_AC_CODE, _DC_CODE = _codes(_AC_COUNTS, _AC_SYMBOLS), _codes(_DC_COUNTS, _DC_SYMBOLS)


class _Out:
    def __init__(self):
        self.buf = bytearray()
        self.acc = 0
        self.n = 0

    def put(self, value: int, bits: int):
        if bits == 0:
            return
        self.acc = (self.acc << bits) | (value & ((1 << bits) - 1))
        self.n += bits
        while self.n >= 8:
            b = (self.acc >> (self.n - 8)) & 0xFF
            self.buf.append(b)
            if b == 0xFF:
                self.buf.append(0)
            self.n -= 8
        self.acc &= (1 << self.n) - 1

    def symbol(self, s: int, dc: bool = False):
        self.put(*(_DC_CODE if dc else _AC_CODE)[s])

    def flush(self):
        if self.n:
            self.put((1 << (8 - self.n)) - 1, 8 - self.n)  # pad with ones


def _nbits(v: int) -> int:
    return int(v).bit_length()


class _ScanCoder:
    """one scan's entropy coder (jcphuff.c's state: EOBRUN and the buffered correction bits BE)"""

    def __init__(self, out: _Out, ss, se, ah, al):
        self.o, self.ss, self.se, self.ah, self.al = out, ss, se, ah, al
        self.eobrun = 0
        self.be = []
        self.pred = {}

    def emit_eobrun(self):
        if self.eobrun > 0:
            nb = _nbits(self.eobrun) - 1
            self.o.symbol(nb << 4)
            self.o.put(self.eobrun, nb)
            self.eobrun = 0
        for b in self.be:
            self.o.put(b, 1)
        self.be = []

    def restart(self):
        self.emit_eobrun()
        self.pred = {}

    def dc_first(self, c, blk):
        v = int(blk[0]) >> self.al                      # arithmetic shift, as the point transform of the standard
        d = v - self.pred.get(c, 0)
        self.pred[c] = v
        nb = _nbits(abs(d))
        self.o.symbol(nb, dc=True)
        self.o.put(d if d >= 0 else d - 1, nb)

    def dc_refine(self, blk):
        self.o.put((int(blk[0]) >> self.al) & 1, 1)

    def ac_first(self, blk):
        r = 0
        for k in range(self.ss, self.se + 1):
            v = int(blk[jo.ZIGZAG[k]])
            t = (-v if v < 0 else v) >> self.al
            if t == 0:
                r += 1
                continue
            if self.eobrun:
                self.emit_eobrun()
            while r > 15:
                self.o.symbol(0xF0)
                r -= 16
            nb = _nbits(t)
            self.o.symbol((r << 4) | nb)
            self.o.put(t if v >= 0 else ~t, nb)
            r = 0
        if r > 0:
            self.eobrun += 1
            if self.eobrun == 0x7FFF:
                self.emit_eobrun()

    def ac_refine(self, blk):
        absv = {}
        eob = 0
        for k in range(self.ss, self.se + 1):
            v = int(blk[jo.ZIGZAG[k]])
            t = (-v if v < 0 else v) >> self.al
            absv[k] = t
            if t == 1:
                eob = k                                  # the last newly non-zero coefficient
        r = 0
        br = []
        for k in range(self.ss, self.se + 1):
            t = absv[k]
            if t == 0:
                r += 1
                continue
            while r > 15 and k <= eob:
                self.emit_eobrun()
                self.o.symbol(0xF0)
                r -= 16
                for b in br:
                    self.o.put(b, 1)
                br = []
            if t > 1:                                    # already non-zero: one correction bit
                br.append(t & 1)
                continue
            self.emit_eobrun()
            self.o.symbol((r << 4) | 1)
            self.o.put(0 if int(blk[jo.ZIGZAG[k]]) < 0 else 1, 1)
            for b in br:
                self.o.put(b, 1)
            br = []
            r = 0
        if r > 0 or br:
            self.eobrun += 1
            self.be += br
            if self.eobrun == 0x7FFF or len(self.be) > 900:
                self.emit_eobrun()


def _segment(marker: int, body: bytes) -> bytes:
    return bytes([0xFF, marker]) + struct.pack(">H", len(body) + 2) + body


# (component indices, Ss, Se, Ah, Al) per scan
SCRIPT_PLAIN = [((0, 1, 2), 0, 0, 0, 0), ((0,), 1, 63, 0, 0), ((1,), 1, 63, 0, 0), ((2,), 1, 63, 0, 0)]          # spectral selection only
SCRIPT_MOZ = [((0, 1, 2), 0, 0, 0, 1), ((0,), 1, 8, 0, 0), ((1,), 1, 8, 0, 0), ((2,), 1, 8, 0, 0), ((0,), 9, 63, 0, 2),
              ((0,), 9, 63, 2, 1), ((0, 1, 2), 0, 0, 1, 0), ((1,), 9, 63, 0, 1), ((2,), 9, 63, 0, 1), ((0,), 9, 63, 1, 0),
              ((1,), 9, 63, 1, 0), ((2,), 9, 63, 1, 0)]                                                             # bands split at 8, mozjpeg style
SCRIPT_DEEP = [((0,), 0, 0, 0, 3), ((1, 2), 0, 0, 0, 3), ((0,), 1, 63, 0, 3), ((1,), 1, 63, 0, 3), ((2,), 1, 63, 0, 3)] + [
    s for al in (2, 1, 0) for s in (((0, 1, 2), 0, 0, al + 1, al), ((0,), 1, 63, al + 1, al), ((1,), 1, 5, al + 1, al),
                                    ((1,), 6, 63, al + 1, al), ((2,), 1, 63, al + 1, al))]                        # three refinement rounds, DC by component
SCRIPT_GRAY = [((0,), 0, 0, 0, 2), ((0,), 1, 20, 0, 1), ((0,), 21, 63, 0, 1), ((0,), 0, 0, 2, 1), ((0,), 0, 0, 1, 0),
               ((0,), 1, 63, 1, 0)]


SEQUENTIAL = "sequential, one scan per component"   # a `script` value: SOF0 with non-interleaved scans (jpegtran -scans style)
BASELINE = "sequential, one interleaved scan"       # a `script` value: the ordinary baseline layout


def synthetic_frame(width: int, height: int, sampling, seed: int = 0, quality_scale: int = 4):
    """a Frame of random (sparse, low-frequency) quantised coefficients with the given (h, v) sampling factors per component:
    the files no encoder at hand writes (4:4:0, 4:1:1, ...); what they look like does not matter, Pillow decodes them too"""
    rng = np.random.default_rng(seed)
    comps = [jo.Component(ci + 1, h, v, 0 if ci == 0 else 1) for ci, (h, v) in enumerate(sampling)]
    hmax, vmax = max(c.h for c in comps), max(c.v for c in comps)
    mcus_x, mcus_y = -(-width // (8 * hmax)), -(-height // (8 * vmax))
    for c in comps:
        c.blocks_w, c.blocks_h = mcus_x * c.h, mcus_y * c.v
        c.width, c.height = -(-width * c.h // hmax), -(-height * c.v // vmax)
        coefs = np.zeros((c.blocks_h, c.blocks_w, 64), np.int16)
        coefs[:, :, 0] = np.cumsum(rng.integers(-6, 7, (c.blocks_h, c.blocks_w)), axis=1) + rng.integers(-20, 20)
        for k in range(1, 12):                           # a few low frequencies, most of them zero
            mask = rng.random((c.blocks_h, c.blocks_w)) < 0.35
            coefs[:, :, jo.ZIGZAG[k]] = np.where(mask, rng.integers(-9, 10, (c.blocks_h, c.blocks_w)), 0)
        c.coefs = coefs
    qt = {0: np.clip(np.arange(64, dtype=np.uint16) // 4 * quality_scale + 4, 1, 255).astype(np.uint16),
          1: np.clip(np.arange(64, dtype=np.uint16) // 3 * quality_scale + 6, 1, 255).astype(np.uint16)}
    return jo.Frame(width, height, comps, qt, 0, hmax, vmax, mcus_x, mcus_y)



def transcode(jpeg: bytes, script, restart_interval: int = 0) -> bytes:
    """the same image (same quantised coefficients, tables, sampling) as a progressive file with this scan script, or
    (script = SEQUENTIAL) as a baseline file whose components come in scans of their own"""
    return encode_frame(jo.parse(jpeg), script, restart_interval)


def encode_frame(f, script, restart_interval: int = 0) -> bytes:
    """a parsed (or synthetic) Frame as JPEG bytes: progressive with `script`, or SEQUENTIAL / BASELINE"""
    sequential = script in (SEQUENTIAL, BASELINE)
    if script == SEQUENTIAL:
        script = [((ci,), 0, 63, 0, 0) for ci in range(len(f.comps))]
    elif script == BASELINE:
        script = [(tuple(range(len(f.comps))), 0, 63, 0, 0)]
    out = bytearray(b"\xff\xd8")
    for tq in sorted({c.tq for c in f.comps}):
        q = f.qt[tq]
        wide = bool((q > 255).any())
        body = bytes([(16 if wide else 0) | tq])
        for k in range(64):
            v = int(q[jo.ZIGZAG[k]])
            body += struct.pack(">H", v) if wide else bytes([v])
        out += _segment(0xDB, body)
    body = bytes([8]) + struct.pack(">HH", f.height, f.width) + bytes([len(f.comps)])
    for c in f.comps:
        body += bytes([c.cid, (c.h << 4) | c.v, c.tq])
    out += _segment(0xC0 if sequential else 0xC2, body)
    out += _segment(0xC4, bytes([0x00]) + bytes(_DC_COUNTS) + bytes(_DC_SYMBOLS))
    out += _segment(0xC4, bytes([0x10]) + bytes(_AC_COUNTS) + bytes(_AC_SYMBOLS))
    if restart_interval:
        out += _segment(0xDD, struct.pack(">H", restart_interval))
    for comps, ss, se, ah, al in script:
        hdr = bytes([len(comps)])
        for ci in comps:
            hdr += bytes([f.comps[ci].cid, 0x00])
        out += _segment(0xDA, hdr + bytes([ss, se, (ah << 4) | al]))
        o = _Out()
        coder = _ScanCoder(o, ss, se, ah, al)

        def block(ci, by, bx):
            blk = f.comps[ci].coefs[by, bx]
            if sequential:                               # DC difference, then the block's AC run, closed by its own EOB
                coder.dc_first(ci, blk)
                coder.ss, coder.se = 1, 63
                coder.ac_first(blk)
                coder.emit_eobrun()
            elif ss == 0:
                coder.dc_first(ci, blk) if ah == 0 else coder.dc_refine(blk)
            else:
                coder.ac_first(blk) if ah == 0 else coder.ac_refine(blk)

        units = []
        if len(comps) > 1:                               # interleaved: MCU order over the padded block grid
            for my in range(f.mcus_y):
                for mx in range(f.mcus_x):
                    units.append([(ci, my * f.comps[ci].v + y, mx * f.comps[ci].h + x) for ci in comps
                                  for y in range(f.comps[ci].v) for x in range(f.comps[ci].h)])
        else:                                            # one component: its own blocks, without the MCU padding
            c = f.comps[comps[0]]
            bw, bh = (c.width + 7) // 8, (c.height + 7) // 8
            units = [[(comps[0], y, x)] for y in range(bh) for x in range(bw)]
        rst = 0
        for i, unit in enumerate(units):
            if restart_interval and i and i % restart_interval == 0:
                coder.restart()
                o.flush()
                o.buf += bytes([0xFF, 0xD0 + (rst & 7)])
                rst += 1
            for ci, by, bx in unit:
                block(ci, by, bx)
        coder.emit_eobrun()
        o.flush()
        out += o.buf
    return bytes(out) + b"\xff\xd9"
