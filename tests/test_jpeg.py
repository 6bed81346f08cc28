"""K0: baseline JPEG decode.  CPU part: the numpy oracle is pinned, bit for
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


def test_oracle_rejects_progressive():
    PIL = pytest.importorskip("PIL.Image")
    buf = io.BytesIO()
    PIL.fromarray(np.zeros((16, 16, 3), np.uint8)).save(buf, format="JPEG", progressive=True)
    with pytest.raises(ValueError):
        jo.decode(buf.getvalue())
