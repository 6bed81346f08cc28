"""K0: baseline and progressive JPEG decode.  CPU part: the numpy oracle is pinned, bit for
bit, to Pillow / libjpeg-turbo's output on the committed fixtures (and to Pillow itself when it is
importable), and the library's host-side entropy decoder must produce the oracle's coefficients."""
import glob
import io
import os

import numpy as np
import pytest

from oracle import jpeg_oracle as jo

HERE = os.path.dirname(os.path.abspath(__file__))
FIXTURES = sorted(glob.glob(os.path.join(HERE, "golden", "jpeg", "*.jpg")))
BASELINE0 = os.path.join(HERE, "golden", "jpeg", "q90_420_64x64.jpg")


def _name(p):
    return os.path.basename(p)[:-4]


@pytest.mark.parametrize("path", FIXTURES, ids=_name)
def test_oracle_matches_pillow_fixture(path):
    want = np.load(path[:-4] + ".npy")
    got = jo.decode(open(path, "rb").read())
    assert got.shape == want.shape and np.array_equal(got, want)


def test_fixtures_cover_the_supported_formats():
    assert len(FIXTURES) >= 9
    kinds = set()
    for p in FIXTURES:
        f = jo.parse(open(p, "rb").read())
        kinds.add((len(f.comps), f.hmax, f.vmax, f.restart_interval > 0))
    assert {(3, 2, 2, False), (3, 2, 2, True), (3, 2, 1, False), (3, 1, 1, False), (1, 1, 1, False)} <= kinds
    prog = {(len(f.comps), f.hmax, f.vmax, f.restart_interval > 0)
            for f in (jo.parse(open(p, "rb").read()) for p in FIXTURES) if f.progressive}
    assert {(3, 2, 2, False), (3, 2, 2, True), (3, 2, 1, False), (3, 1, 1, False), (1, 1, 1, False)} <= prog


def test_oracle_matches_live_pillow_when_available():
    PIL = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(5)
    for (h, w), kw in (((37, 53), dict(quality=70, subsampling=2)), ((16, 24), dict(quality=99, subsampling=1)),
                       ((9, 9), dict(quality=50, subsampling=0)), ((8, 130), dict(quality=88, subsampling=2, restart_marker_blocks=1))):
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        img = (img // 4 + np.linspace(0, 190, w, dtype=np.int32)[None, :, None]).astype(np.uint8)
        buf = io.BytesIO()
        PIL.fromarray(img).save(buf, format="JPEG", **kw)
        want = np.asarray(PIL.open(io.BytesIO(buf.getvalue())).convert("RGB"))
        assert np.array_equal(jo.decode(buf.getvalue()), want), (h, w, kw)


def test_oracle_matches_live_pillow_progressive():
    """SOF2: every scan kind of T.81 annex G as libjpeg's default script writes them (DC first + refinement,
    AC first with end-of-band runs, AC refinement), with and without restart markers / optimised tables."""
    PIL = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(6)
    for (h, w), kw in (((37, 53), dict(quality=70, subsampling=2)), ((16, 24), dict(quality=99, subsampling=1)),
                       ((9, 9), dict(quality=50, subsampling=0)), ((24, 130), dict(quality=88, subsampling=2, restart_marker_blocks=1)),
                       ((40, 40), dict(quality=20, subsampling=2, optimize=True)), ((31, 17), dict(quality=90))):
        img = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        img = (img // 4 + np.linspace(0, 190, w, dtype=np.int32)[None, :, None]).astype(np.uint8)
        pil = PIL.fromarray(img)
        if "subsampling" not in kw:
            pil = pil.convert("L")
        buf = io.BytesIO()
        pil.save(buf, format="JPEG", progressive=True, **kw)
        f = jo.parse(buf.getvalue())
        assert f.progressive and f.scans > 1
        want = np.asarray(PIL.open(io.BytesIO(buf.getvalue())).convert("RGB"))
        assert np.array_equal(jo.decode(buf.getvalue()), want), (h, w, kw)


@pytest.mark.parametrize("src,script,restart", [
    ("q75_420_ragged_61x83", "SCRIPT_MOZ", 0), ("q75_420_ragged_61x83", "SCRIPT_DEEP", 3), ("q85_422_50x70", "SCRIPT_DEEP", 0),
    ("q85_422_50x70", "SCRIPT_PLAIN", 5), ("q95_444_33x47", "SCRIPT_MOZ", 3), ("q60_420_rst2_72x104", "SCRIPT_PLAIN", 0),
    ("q92_gray_45x37", "SCRIPT_GRAY", 0), ("q92_gray_45x37", "SCRIPT_GRAY", 2),
    ("q75_420_ragged_61x83", "SEQUENTIAL", 0), ("q85_422_50x70", "SEQUENTIAL", 4)])
def test_progressive_scan_scripts_of_other_encoders(src, script, restart):
    """Pillow only writes libjpeg's default progression.  tests/jpeg_writer.py re-codes a baseline fixture's coefficients
    with other legal scripts - spectral selection only, a mozjpeg-style script (bands split at 8, DC with one refinement),
    three refinement rounds with DC scans of one and of two components - with and without restart markers: same
    coefficients, so the pixels must be those of the original for Pillow (when importable), for the oracle, and the
    library's host decoder must produce the oracle's coefficients."""
    from ai4e_b200 import ops
    from tests import jpeg_writer as jw
    base = os.path.join(HERE, "golden", "jpeg", src)
    data = jw.transcode(open(base + ".jpg", "rb").read(), getattr(jw, script), restart)
    want = np.load(base + ".npy")
    try:
        from PIL import Image
        assert np.array_equal(np.asarray(Image.open(io.BytesIO(data)).convert("RGB")), want)   # the writer itself
    except ImportError:
        pass
    f = jo.parse(data)
    if script == "SEQUENTIAL":       # a baseline file whose components come in scans of their own (jpegtran -scans style)
        assert not f.progressive and f.scans == len(f.comps)
    else:
        assert f.progressive and f.scans == len(getattr(jw, script))
    assert f.restart_interval == restart
    assert np.array_equal(jo.decode(data), want)
    info, rec = ops.jpeg_entropy_decode(data)
    for c, comp in enumerate(f.comps):
        n = comp.blocks_w * comp.blocks_h * 64
        assert np.array_equal(rec[info.coef_off[c]:info.coef_off[c] + n].reshape(comp.blocks_h, comp.blocks_w, 64), comp.coefs)


def test_other_chroma_samplings_and_narrow_components_match_live_pillow():
    """4:4:0, 4:1:1, 4:1:0, differently sampled chroma components, baseline / progressive / one scan per component, and
    sizes whose chroma is one or two samples wide (libjpeg replicates there instead of filtering): synthetic coefficients
    from tests/jpeg_writer.py, expected pixels from Pillow; the library's host decoder must give the oracle's coefficients"""
    Image = pytest.importorskip("PIL.Image")
    from ai4e_b200 import ops
    from tests import jpeg_writer as jw
    samplings = {"440": [(1, 2), (1, 1), (1, 1)], "411": [(4, 1), (1, 1), (1, 1)], "410": [(4, 2), (1, 1), (1, 1)],
                 "mixed": [(2, 2), (2, 1), (1, 1)], "mixed2": [(2, 2), (1, 2), (1, 1)], "420": [(2, 2), (1, 1), (1, 1)],
                 "422": [(2, 1), (1, 1), (1, 1)]}
    n = 0
    for name, samp in samplings.items():
        for h, w in ((1, 1), (3, 3), (5, 4), (17, 33), (24, 61)):
            script = (jw.BASELINE, jw.SCRIPT_MOZ, jw.SEQUENTIAL)[n % 3]
            data = jw.encode_frame(jw.synthetic_frame(w, h, samp, seed=n), script, restart_interval=(2 if n % 2 else 0))
            n += 1
            want = np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))
            assert np.array_equal(jo.decode(data), want), (name, h, w)
            f = jo.parse(data)
            info, rec = ops.jpeg_entropy_decode(data)
            for c, comp in enumerate(f.comps):
                m = comp.blocks_w * comp.blocks_h * 64
                assert np.array_equal(rec[info.coef_off[c]:info.coef_off[c] + m].reshape(comp.blocks_h, comp.blocks_w, 64), comp.coefs)
    # tiny images as Pillow itself writes them: every size from 1 x 1 to 6 x 6 in the three samplings it offers
    rng = np.random.default_rng(2)
    for sub in (0, 1, 2):
        for h in range(1, 7):
            for w in range(1, 7):
                buf = io.BytesIO()
                Image.fromarray(rng.integers(0, 256, (h, w, 3), dtype=np.uint8)).save(buf, format="JPEG", quality=90, subsampling=sub)
                want = np.asarray(Image.open(io.BytesIO(buf.getvalue())).convert("RGB"))
                assert np.array_equal(jo.decode(buf.getvalue()), want), (sub, h, w)


def test_sixteen_bit_quantisation_tables():
    """DQT with Pq = 1 (entries above 255), baseline and progressive: oracle == Pillow, the library's record carries the table"""
    Image = pytest.importorskip("PIL.Image")
    from ai4e_b200 import ops
    from tests import jpeg_writer as jw
    f = jw.synthetic_frame(33, 17, [(2, 2), (1, 1), (1, 1)], seed=5)
    f.qt[0] = f.qt[0].copy()
    f.qt[0][40:] = np.arange(300, 324, dtype=np.uint16)
    for script in (jw.BASELINE, jw.SCRIPT_MOZ):
        data = jw.encode_frame(f, script)
        assert data[data.index(b"\xff\xdb") + 4] >> 4 == 1
        want = np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))
        assert np.array_equal(jo.decode(data), want)
        info, rec = ops.jpeg_entropy_decode(data)
        g = jo.parse(data)
        for c, comp in enumerate(g.comps):
            assert np.array_equal(rec[c * 64:(c + 1) * 64].astype(np.uint16), g.qt[comp.tq])


def _cmyk_jpeg():
    PIL = pytest.importorskip("PIL.Image")
    buf = io.BytesIO()
    PIL.fromarray(np.zeros((16, 16, 4), np.uint8), mode="CMYK").save(buf, format="JPEG")
    return buf.getvalue()


def _as_arithmetic(data):
    """the same headers with the frame marker rewritten to SOF9 (arithmetic coding, sequential)"""
    i = data.index(b"\xff\xc0")
    return data[:i] + b"\xff\xc9" + data[i + 2:]


def test_oracle_rejects_what_it_does_not_support():
    with pytest.raises(ValueError):
        jo.decode(_cmyk_jpeg())
    with pytest.raises(ValueError):
        jo.decode(_as_arithmetic(open(BASELINE0, "rb").read()))


# ------------------------------------------------------------------ library, host side (no GPU needed)
@pytest.mark.parametrize("path", FIXTURES, ids=_name)
def test_host_entropy_decoder_matches_oracle_coefficients(path):
    """ai4e_jpeg_entropy_decode runs on the host: its record must hold exactly the oracle's
    quantisation tables and quantised coefficients."""
    from ai4e_b200 import ops
    data = open(path, "rb").read()
    info, rec = ops.jpeg_entropy_decode(data)
    f = jo.parse(data)
    assert (info.width, info.height, info.components, info.hmax, info.vmax, info.restart_interval) == \
           (f.width, f.height, len(f.comps), f.hmax, f.vmax, f.restart_interval)
    assert info.record_int16 % 64 == 0 and rec.size == info.record_int16
    for c, comp in enumerate(f.comps):
        assert (info.blocks_w[c], info.blocks_h[c], info.comp_w[c], info.comp_h[c]) == \
               (comp.blocks_w, comp.blocks_h, comp.width, comp.height)
        n = comp.blocks_w * comp.blocks_h * 64
        got = rec[info.coef_off[c]:info.coef_off[c] + n].reshape(comp.blocks_h, comp.blocks_w, 64)
        assert np.array_equal(got, comp.coefs)
        assert np.array_equal(rec[c * 64:(c + 1) * 64].astype(np.uint16), f.qt[comp.tq])


def test_host_decoder_rejects_what_it_does_not_support():
    from ai4e_b200 import ops, Ai4eError
    with pytest.raises(Ai4eError, match="4 components"):
        ops.jpeg_entropy_decode(_cmyk_jpeg())
    with pytest.raises(Ai4eError, match="arithmetic"):
        ops.jpeg_entropy_decode(_as_arithmetic(open(BASELINE0, "rb").read()))
    with pytest.raises(Ai4eError, match="SOI"):
        ops.jpeg_entropy_decode(b"\x89PNG\r\n\x1a\n" + b"\0" * 32)
    good = open(BASELINE0, "rb").read()
    with pytest.raises(Ai4eError):
        ops.jpeg_entropy_decode(good[:200])  # truncated inside the tables / scan


def test_host_decoder_survives_mutated_streams():
    """Random byte edits, bit flips, cuts and deletions over every fixture (baseline and progressive): each
    stream is either decoded or refused with an error - never a crash, never a write outside the record
    (the record is allocated exactly and followed by a canary)."""
    import ctypes as C
    from ai4e_b200 import _lib as L
    lib = L.load()
    rng = np.random.default_rng(77)
    small = [open(p, "rb").read() for p in FIXTURES if os.path.getsize(p) < 4000]
    decoded = refused = 0
    for it in range(3000):
        d = bytearray(small[it % len(small)])
        kind = it % 4
        if kind == 0:
            for _ in range(int(rng.integers(1, 6))):
                d[int(rng.integers(2, len(d)))] = int(rng.integers(256))
        elif kind == 1:
            d = d[:int(rng.integers(2, len(d)))]
        elif kind == 2:
            d[int(rng.integers(2, len(d)))] ^= 1 << int(rng.integers(8))
        else:
            i = int(rng.integers(2, len(d)))
            del d[i:i + int(rng.integers(1, 40))]
        buf = (C.c_uint8 * len(d)).from_buffer_copy(bytes(d))
        info = L.JpegInfo()
        if lib.ai4e_jpeg_entropy_decode(buf, len(d), None, 0, C.byref(info)) != 0:
            refused += 1
            continue
        n = int(info.record_int16)
        if n > (1 << 24):       # a mutated frame header can ask for a huge record: not this test's business
            continue
        rec = np.full(n + 64, 0x5A5A, np.int16)
        rc = lib.ai4e_jpeg_entropy_decode(buf, len(d), rec.ctypes.data_as(C.c_void_p), n, C.byref(info))
        assert np.all(rec[n:] == 0x5A5A)
        decoded += rc == 0
        refused += rc != 0
    assert decoded > 300 and refused > 300


def _dht_stream(counts, nsyms=None, tc_th=0x00):
    """SOI + one DHT segment with the given 16 code-length counts (+ EOI)."""
    counts = list(counts) + [0] * (16 - len(counts))
    n = sum(counts) if nsyms is None else nsyms
    seg = bytes([tc_th]) + bytes(counts) + bytes(range(256))[:n]
    return b"\xff\xd8" + b"\xff\xc4" + (len(seg) + 2).to_bytes(2, "big") + seg + b"\xff\xd9"


@pytest.mark.parametrize("length", range(1, 17))
def test_host_decoder_refuses_oversubscribed_huffman_tables(length):
    """A DHT whose counts claim more codes of one length than that length has must be refused
    BEFORE the look-up table is filled: such a table used to write past it (a crafted request could
    crash the serving process)."""
    from ai4e_b200 import ops, Ai4eError
    counts = [0] * 16
    counts[length - 1] = min(255, (1 << length) + 1)
    if counts[length - 1] <= (1 << length):  # 255 codes fit lengths >= 8: over-subscribe with the shorter ones
        counts[0] = 2
        counts[1] = 1
    with pytest.raises(Ai4eError, match="Huffman"):
        ops.jpeg_entropy_decode(_dht_stream(counts))


def test_host_decoder_survives_malformed_segments():
    from ai4e_b200 import ops, Ai4eError
    with pytest.raises(Ai4eError):                       # counts[0] = 255: the reproducer of the overflow
        ops.jpeg_entropy_decode(_dht_stream([255]))
    with pytest.raises(Ai4eError):                       # 256 symbols declared, segment too short for them
        ops.jpeg_entropy_decode(_dht_stream([0, 0, 0, 0, 0, 0, 0, 255, 1], nsyms=10))
    with pytest.raises(Ai4eError):                       # a full, valid 256-symbol table alone: no frame header
        ops.jpeg_entropy_decode(_dht_stream([0, 0, 0, 0, 0, 0, 0, 255, 2][:9]))
    good = open(BASELINE0, "rb").read()
    sos = good.index(b"\xff\xda")
    with pytest.raises(Ai4eError):                       # SOS as the last four bytes: header length 2, no payload
        ops.jpeg_entropy_decode(good[:sos] + b"\xff\xda\x00\x02")
    # a scan that names the first component twice leaves the others without tables
    hdr = bytearray(good[sos:sos + 14])
    assert hdr[4] == 3
    hdr[7] = hdr[5]
    with pytest.raises(Ai4eError, match="twice"):
        ops.jpeg_entropy_decode(good[:sos] + bytes(hdr) + good[sos + 14:])
    for cut in range(2, min(len(good), 700), 7):         # truncation anywhere in the headers never crashes
        try:
            ops.jpeg_entropy_decode(good[:cut])
        except Ai4eError:
            pass


# ------------------------------------------------------------------ GPU: bit-exact reconstruction
@pytest.mark.gpu
@pytest.mark.parametrize("path", FIXTURES, ids=_name)
def test_gpu_decode_is_bit_exact_with_pillow(path):
    from ai4e_b200 import ops
    want = np.load(path[:-4] + ".npy")
    got = ops.jpeg_decode(open(path, "rb").read()).cpu().numpy()
    assert got.shape == want.shape
    assert np.array_equal(got, want), f"{int((got != want).sum())} bytes differ, max {np.abs(got.astype(int) - want).max()}"


@pytest.mark.gpu
def test_gpu_reconstruct_batch_of_one_geometry():
    """several images of one geometry (different content and quality -> different quantisation
    tables) reconstructed by one launch pair, each bit-exact with Pillow."""
    import torch
    from ai4e_b200 import ops
    PIL = pytest.importorskip("PIL.Image")
    rng = np.random.default_rng(3)
    recs, wants, info0 = [], [], None
    for i, q in enumerate((35, 60, 75, 90, 97)):
        img = rng.integers(0, 256, (120, 200, 3), dtype=np.uint8)
        img = (img // 3 + np.linspace(0, 160, 200, dtype=np.int32)[None, :, None] + 5 * i).astype(np.uint8)
        buf = io.BytesIO()
        PIL.fromarray(img).save(buf, format="JPEG", quality=q, subsampling=2)
        wants.append(np.asarray(PIL.open(io.BytesIO(buf.getvalue())).convert("RGB")))
        info, rec = ops.jpeg_entropy_decode(buf.getvalue())
        info0 = info0 or info
        assert info.record_int16 == info0.record_int16
        recs.append(rec)
    out = ops.jpeg_reconstruct(torch.from_numpy(np.stack(recs)).cuda(), info0).cpu().numpy()
    for got, want in zip(out, wants):
        assert np.array_equal(got, want)


@pytest.mark.gpu
def test_engine_submit_jpeg_equals_submit_of_decoded_pixels():
    """the engine classifies JPEG bytes exactly as it classifies Pillow's decode of them: mixed
    sizes and samplings (each closes the open micro-batch), the resize path for non-224 images,
    many threads submitting at once, and a bad stream that must fail without queueing anything."""
    import threading
    from tests import util
    from ai4e_b200 import Ai4eError
    from ai4e_b200.engine import Engine
    PIL = pytest.importorskip("PIL.Image")
    g, imgs = util.load_golden("resnet50")
    cases = []
    for i, img in enumerate(imgs[:6]):  # 224 x 224, 4:2:0, differing quality
        buf = io.BytesIO()
        PIL.fromarray(img).save(buf, format="JPEG", quality=80 + 3 * i, subsampling=2)
        cases.append(buf.getvalue())
    for i, (hw, sub) in enumerate((((224, 224), 1), ((200, 256), 2), ((200, 256), 2), ((224, 224), 0))):  # other sampling / size
        buf = io.BytesIO()
        PIL.fromarray(imgs[i]).resize((hw[1], hw[0])).save(buf, format="JPEG", quality=88, subsampling=sub)
        cases.append(buf.getvalue())
    for i, (hw, sub) in enumerate((((224, 224), 2), ((224, 224), 2), ((180, 240), 1))):  # progressive files among the baseline ones
        buf = io.BytesIO()
        PIL.fromarray(imgs[6 + i]).resize((hw[1], hw[0])).save(buf, format="JPEG", quality=85, subsampling=sub, progressive=True)
        cases.append(buf.getvalue())
    decoded = [np.asarray(PIL.open(io.BytesIO(d)).convert("RGB")) for d in cases]
    with Engine(util.program_for("resnet50", 3, 224, False), device=0, max_batch=8, max_wait_us=3000) as e:
        want = [e.classify(d, want_logits=True) for d in decoded]
        got = [None] * len(cases)

        def work(k):
            got[k] = e.classify_jpeg(cases[k], want_logits=True)

        threads = [threading.Thread(target=work, args=(k,)) for k in range(len(cases))]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        for k, ((wi, wp, wl), (gi, gp, gl)) in enumerate(zip(want, got)):
            assert np.array_equal(gi, wi), k
            assert np.array_equal(gl, wl), (k, np.abs(gl - wl).max())  # same pixels -> same bits
        before = e.stats()["images"]
        with pytest.raises(Ai4eError, match="4 components"):
            e.submit_jpeg(_cmyk_jpeg())
        e.flush()
        assert e.stats()["images"] == before and e.stats()["healthy"]


@pytest.mark.gpu
def test_pool_submit_jpeg_routes_like_submit():
    from tests import util
    from ai4e_b200.engine import Pool
    PIL = pytest.importorskip("PIL.Image")
    g, imgs = util.load_golden("resnet18")
    jpegs = []
    for img in imgs[:8]:
        buf = io.BytesIO()
        PIL.fromarray(img).save(buf, format="JPEG", quality=93, subsampling=2)
        jpegs.append(buf.getvalue())
    decoded = [np.asarray(PIL.open(io.BytesIO(d)).convert("RGB")) for d in jpegs]
    with Pool(util.program_for("resnet18"), devices=[0, 0], max_batch=4, max_wait_us=300) as p:
        want = [p.classify(d, want_logits=True) for d in decoded]
        tickets = [p.submit_jpeg(d) for d in jpegs * 2]
        got = [p.wait(t, want_logits=True) for t in tickets]
        for k, (gi, gp, gl) in enumerate(got):
            assert np.array_equal(gi, want[k % 8][0]) and np.array_equal(gl, want[k % 8][2])
        assert all(s["images"] > 0 for s in p.stats())
