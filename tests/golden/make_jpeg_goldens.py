#!/usr/bin/env python
"""Mint the JPEG fixtures of tests/golden/jpeg/ with Pillow (libjpeg-turbo), the decoder a PyTorch
API container uses: small synthetic images saved with different chroma subsampling, quality,
restart intervals, (optimised) Huffman tables and progressive scan scripts, plus Pillow's decode of each as .npy.
    python tests/golden/make_jpeg_goldens.py
"""
import io
import os
import sys

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
from oracle import resnet_oracle as ro  # noqa: E402

OUT = os.path.join(HERE, "jpeg")
CASES = [  # name, (h, w), save kwargs
    ("q90_420_64x64", (64, 64), dict(quality=90, subsampling=2)),
    ("q75_420_ragged_61x83", (61, 83), dict(quality=75, subsampling=2)),
    ("q85_422_50x70", (50, 70), dict(quality=85, subsampling=1)),
    ("q95_444_33x47", (33, 47), dict(quality=95, subsampling=0)),
    ("q60_420_rst2_72x104", (72, 104), dict(quality=60, subsampling=2, restart_marker_blocks=2)),
    ("q80_420_opt_rstrow_96x80", (96, 80), dict(quality=80, subsampling=2, optimize=True, restart_marker_rows=1)),
    ("q30_420_40x40", (40, 40), dict(quality=30, subsampling=2)),
    ("q92_gray_45x37", (45, 37), dict(quality=92)),
    ("q90_420_224x224", (224, 224), dict(quality=90, subsampling=2)),
    # progressive (SOF2): spectral selection + successive approximation as libjpeg's default script writes them
    ("prog_q85_420_ragged_61x83", (61, 83), dict(quality=85, subsampling=2, progressive=True)),
    ("prog_q75_422_50x70", (50, 70), dict(quality=75, subsampling=1, progressive=True)),
    ("prog_q95_444_33x47", (33, 47), dict(quality=95, subsampling=0, progressive=True)),
    ("prog_q60_420_rst2_72x104", (72, 104), dict(quality=60, subsampling=2, progressive=True, restart_marker_blocks=2)),
    ("prog_q92_gray_45x37", (45, 37), dict(quality=92, progressive=True)),
    ("prog_q90_420_224x224", (224, 224), dict(quality=90, subsampling=2, progressive=True)),
]


def main():
    os.makedirs(OUT, exist_ok=True)
    rng = np.random.default_rng(11)
    for i, (name, (h, w), kw) in enumerate(CASES):
        img = ro.synthetic_images(1, h, w, 3, seed=100 + i)[0]
        img = np.clip(img.astype(np.int32) + rng.integers(-12, 13, img.shape), 0, 255).astype(np.uint8)  # some texture
        pil = Image.fromarray(img)
        if "gray" in name:
            pil = pil.convert("L")
        buf = io.BytesIO()
        pil.save(buf, format="JPEG", **kw)
        data = buf.getvalue()
        open(os.path.join(OUT, name + ".jpg"), "wb").write(data)
        dec = np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))
        np.save(os.path.join(OUT, name + ".npy"), dec)
        print(name, len(data), "bytes", dec.shape)


def transcoded():
    """progressive files with the scan scripts OTHER encoders write (tests/jpeg_writer.py re-codes the coefficients of a
    baseline fixture): Pillow must decode them to the pixels of the original, and so must this repo"""
    from PIL import Image as _I
    from tests import jpeg_writer as jw
    for name, src, script, ri in (("progmoz_420_ragged_61x83", "q75_420_ragged_61x83", jw.SCRIPT_MOZ, 0),
                                  ("progdeep_422_rst3_50x70", "q85_422_50x70", jw.SCRIPT_DEEP, 3),
                                  ("progplain_444_33x47", "q95_444_33x47", jw.SCRIPT_PLAIN, 0),
                                  ("proggray_rst3_45x37", "q92_gray_45x37", jw.SCRIPT_GRAY, 3)):
        data = jw.transcode(open(os.path.join(OUT, src + ".jpg"), "rb").read(), script, ri)
        open(os.path.join(OUT, name + ".jpg"), "wb").write(data)
        dec = np.asarray(_I.open(io.BytesIO(data)).convert("RGB"))
        assert np.array_equal(dec, np.load(os.path.join(OUT, src + ".npy")))  # same coefficients, same pixels
        np.save(os.path.join(OUT, name + ".npy"), dec)
        print(name, len(data), "bytes", dec.shape)


def other_samplings():
    """chroma samplings Pillow does not write (4:4:0 - what a losslessly rotated 4:2:2 photo becomes -, 4:1:1, 4:1:0, the
    two chroma components sampled differently) and components so narrow that libjpeg replicates instead of filtering:
    synthetic coefficients (tests/jpeg_writer.py), expected pixels from Pillow"""
    from PIL import Image as _I
    from tests import jpeg_writer as jw
    for name, samp, (h, w), script, ri in (
            ("samp440_40x61", [(1, 2), (1, 1), (1, 1)], (40, 61), jw.BASELINE, 0),
            ("samp440_prog_17x33", [(1, 2), (1, 1), (1, 1)], (17, 33), jw.SCRIPT_MOZ, 2),
            ("samp411_40x61", [(4, 1), (1, 1), (1, 1)], (40, 61), jw.BASELINE, 3),
            ("samp410_33x70", [(4, 2), (1, 1), (1, 1)], (33, 70), jw.BASELINE, 0),
            ("sampmixed_24x40", [(2, 2), (2, 1), (1, 1)], (24, 40), jw.SEQUENTIAL, 0),
            ("samp420_narrow_9x3", [(2, 2), (1, 1), (1, 1)], (9, 3), jw.BASELINE, 0),
            ("samp422_narrow_5x4", [(2, 1), (1, 1), (1, 1)], (5, 4), jw.BASELINE, 0)):
        data = jw.encode_frame(jw.synthetic_frame(w, h, samp, seed=len(name)), script, ri)
        open(os.path.join(OUT, name + ".jpg"), "wb").write(data)
        dec = np.asarray(_I.open(io.BytesIO(data)).convert("RGB"))
        np.save(os.path.join(OUT, name + ".npy"), dec)
        print(name, len(data), "bytes", dec.shape)


if __name__ == "__main__":
    main()
    transcoded()
    other_samplings()
