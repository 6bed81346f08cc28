"""PNG payloads: the host-side decoder of libai4e_b200 (ai4e_png_decode, no GPU
needed) against fixtures minted by Pillow (tests/golden/make_png_goldens.py), against live Pillow
when it is importable, and through the engine on the GPU (a PNG and its decoded pixels classify
identically).  The admission probe (`ai4e_image_probe`, 415 helper) is tested here too."""
import glob
import io
import os

import numpy as np
import pytest

from tests import util

PNG_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "png")
FIXTURES = sorted(p for p in glob.glob(os.path.join(PNG_DIR, "*.png")) if "unsupported" not in p)


def _name(p):
    return os.path.basename(p)[:-4]


def test_fixtures_cover_the_supported_flavours():
    names = " ".join(_name(p) for p in FIXTURES)
    for flavour in ("rgb_", "rgba_", "gray_", "graya_", "palette_", "adam7_", "rgb16_", "gray16_", "gray1_", "gray2_", "palette4_", "palette1_"):
        assert flavour in names


@pytest.mark.parametrize("path", FIXTURES, ids=_name)
def test_host_png_decoder_matches_pillow_fixture(path):
    from ai4e_b200 import ops
    got = ops.png_decode(open(path, "rb").read())
    want = np.load(path[:-4] + ".npy")
    assert got.shape == want.shape and np.array_equal(got, want)


def test_every_png_image_type_plain_and_interlaced_matches_live_pillow():
    """all fifteen colour type / bit depth combinations of the specification, plain and Adam7, sizes that leave some of
    the seven passes empty, a random filter type on every line, IDAT cut into pieces, palettes shorter than the index
    range, 16-bit grey around Pillow's clip at 255 - written by tests/png_writer.py, expected pixels from Pillow"""
    Image = pytest.importorskip("PIL.Image")
    from ai4e_b200 import ops
    from tests import png_writer as pw
    rng = np.random.default_rng(3)
    n = 0
    for ctype, depth in pw.VALID:
        for interlace in (False, True):
            for h, w in ((1, 1), (2, 3), (7, 5), (8, 8), (5, 1), (1, 9), (19, 33)):
                samples, pal = pw.random_image(rng, h, w, depth, ctype)
                data = pw.encode(samples, depth, ctype, interlace, pal, seed=n, idat_split=(97 if n % 3 == 0 else 0))
                want = np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))
                got = ops.png_decode(data)
                assert got.shape == want.shape and np.array_equal(got, want), (ctype, depth, interlace, h, w)
                n += 1
    assert n == 15 * 2 * 7


def test_host_png_decoder_survives_mutated_streams():
    """byte edits, cuts and deletions behind the IHDR chunk of every fixture: decoded or refused, never a crash and never
    a write outside the output (a canary follows it)"""
    import ctypes as C
    from ai4e_b200 import _lib as L
    lib = L.load()
    rng = np.random.default_rng(5)
    small = [open(p, "rb").read() for p in FIXTURES if os.path.getsize(p) < 6000]
    decoded = refused = 0
    for it in range(1500):
        d = bytearray(small[it % len(small)])
        kind = it % 3
        lo = 33  # keep the header: a mutated IHDR only asks for another (possibly huge) size
        if kind == 0:
            for _ in range(int(rng.integers(1, 5))):
                d[int(rng.integers(lo, len(d)))] = int(rng.integers(256))
        elif kind == 1:
            d = d[:int(rng.integers(lo, len(d)))]
        else:
            i = int(rng.integers(lo, len(d)))
            del d[i:i + int(rng.integers(1, 30))]
        buf = (C.c_uint8 * len(d)).from_buffer_copy(bytes(d))
        h, w = C.c_int32(), C.c_int32()
        if lib.ai4e_png_decode(buf, len(d), None, 0, C.byref(h), C.byref(w)) != 0:
            refused += 1
            continue
        n = h.value * w.value * 3
        out = np.full(n + 64, 0x5A, np.uint8)
        rc = lib.ai4e_png_decode(buf, len(d), out.ctypes.data_as(C.c_void_p), n, C.byref(h), C.byref(w))
        assert np.all(out[n:] == 0x5A)
        decoded += rc == 0
        refused += rc != 0
    assert decoded > 30 and refused > 300


def test_host_png_decoder_matches_live_pillow():
    Image = pytest.importorskip("PIL.Image")
    from ai4e_b200 import ops
    rng = np.random.default_rng(11)
    for shape, mode in (((61, 83, 3), "RGB"), ((17, 5, 4), "RGBA"), ((9, 130), "L")):
        smooth = np.cumsum(rng.integers(-3, 4, shape), axis=1).astype(np.uint8)  # filters 1-4 get picked
        for level in (1, 9):
            buf = io.BytesIO()
            Image.fromarray(smooth, mode).save(buf, format="PNG", compress_level=level)
            want = np.asarray(Image.open(io.BytesIO(buf.getvalue())).convert("RGB"))
            assert np.array_equal(ops.png_decode(buf.getvalue()), want)


def test_probe_and_refusals():
    from ai4e_b200 import ops, Ai4eError, _lib as L
    from ai4e_b200.engine import require_decodable
    from werkzeug.exceptions import UnsupportedMediaType
    good = open(os.path.join(PNG_DIR, "rgb_33x47.png"), "rb").read()
    assert ops.image_probe(good) == ("png", 33, 47)
    jpg = open(os.path.join(util.GOLDEN, "jpeg", "q90_420_64x64.jpg"), "rb").read()
    assert ops.image_probe(jpg) == ("jpeg", 64, 64)
    assert require_decodable(jpg) == ("jpeg", 64, 64)
    prog = open(os.path.join(util.GOLDEN, "jpeg", "prog_q75_422_50x70.jpg"), "rb").read()
    assert ops.image_probe(prog) == ("jpeg", 50, 70)    # progressive files are admitted (and decoded: test_jpeg.py)
    bad16 = open(os.path.join(PNG_DIR, "unsupported_rgb4_8x8.png"), "rb").read()   # 4-bit RGB: not a PNG image type
    for payload in (bad16, b"GIF89a" + b"\0" * 64, b""):
        with pytest.raises(Ai4eError) as ei:
            ops.image_probe(payload) if payload else ops.image_probe(b"\0")
        assert ei.value.code == L.ERR_UNSUPPORTED
        with pytest.raises(UnsupportedMediaType):       # what a request_processing_function lets propagate: 415
            require_decodable(payload or b"\0")
    with pytest.raises(Ai4eError):
        ops.png_decode(bad16)
    # corrupt streams are errors, never crashes
    for cut in range(8, len(good), 97):
        with pytest.raises(Ai4eError):
            ops.png_decode(good[:cut])
    flipped = bytearray(good)
    flipped[len(good) // 2] ^= 0x55
    try:
        ops.png_decode(bytes(flipped))                   # either zlib notices or the pixels differ; no crash
    except Ai4eError:
        pass


def test_shell_answers_415_for_an_undecodable_payload():
    """the probe inside request_processing_function ends the request with 415 before a task is accepted."""
    from ai4e_service import APIService
    from ai4e_b200.engine import require_decodable
    from tests.host_shell_scenario import RecordingTaskManager
    try:
        from flask import Flask
    except ImportError:
        from ai4e_b200.compat.miniflask import Flask
    app = Flask("p")
    svc = APIService(app, type("L", (), {"tracer": None, "log_exception": lambda *a, **k: None})())
    tm = RecordingTaskManager()
    svc.api_task_manager = tm
    ran = []

    def read(request):
        data = request.get_data()
        require_decodable(data)
        return {"data": data}

    @svc.api_async_func(api_path="/c", methods=["POST"], request_processing_function=read, maximum_concurrent_requests=4)
    def c_(*args, **kwargs):
        ran.append(kwargs["taskId"])

    cl = app.test_client()
    r = cl.post("/c", data=b"GIF89a" + b"\0" * 32, headers={"taskId": "t1"})
    assert r.status_code == 415 and not tm.calls and not ran
    ok = open(os.path.join(PNG_DIR, "rgb_1x1.png"), "rb").read()
    assert cl.post("/c", data=ok, headers={"taskId": "t2"}).get_data(as_text=True) == "TaskId: t2"


@pytest.mark.gpu
def test_engine_classifies_png_like_its_pixels():
    from ai4e_b200 import Ai4eError, _lib as L
    from ai4e_b200.engine import Engine
    data = open(os.path.join(PNG_DIR, "rgb_224x224.png"), "rb").read()
    pixels = np.load(os.path.join(PNG_DIR, "rgb_224x224.npy"))
    small = open(os.path.join(PNG_DIR, "rgb_33x47.png"), "rb").read()      # resized to the model input by K1
    small_px = np.load(os.path.join(PNG_DIR, "rgb_33x47.npy"))
    jpg = open(os.path.join(util.GOLDEN, "jpeg", "q90_420_224x224.jpg"), "rb").read()
    with Engine(util.program_for("resnet18"), device=0, max_batch=4, max_wait_us=300) as e:
        i0, p0, l0 = e.classify(pixels, want_logits=True)
        i1, p1, l1 = e.classify_encoded(data, want_logits=True)
        assert np.array_equal(i0, i1) and np.array_equal(l0, l1)
        i2, _, l2 = e.classify(small_px, want_logits=True)
        i3, _, l3 = e.classify_encoded(small, want_logits=True)
        assert np.array_equal(i2, i3) and np.array_equal(l2, l3)
        ij, _, lj = e.classify_encoded(jpg, want_logits=True)              # the same entry point takes JPEG
        ik, _, lk = e.classify_jpeg(jpg, want_logits=True)
        assert np.array_equal(ij, ik) and np.array_equal(lj, lk)
        with pytest.raises(Ai4eError) as ei:
            e.submit_encoded(b"GIF89a" + b"\0" * 64)
        assert ei.value.code == L.ERR_UNSUPPORTED
        assert e.stats()["queue_depth"] == 0
