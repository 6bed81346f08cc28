"""Mint tests/golden/png/*: small PNG files written by Pillow (and, for what Pillow does not write, by tests/png_writer.py), each next to
the RGB pixels `Image.open(path).convert("RGB")` returns for it (Pillow 12.2 here).

    python tests/golden/make_png_goldens.py
"""
import io
import os

import numpy as np
from PIL import Image

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "png")


def structured(h, w, c, seed):
    rng = np.random.default_rng(seed)
    y, x = np.mgrid[0:h, 0:w]
    base = np.stack([(x * 3 + y * 5 + 40 * k) % 256 for k in range(c)], -1)
    return ((base + rng.integers(0, 40, (h, w, c))) % 256).astype(np.uint8)  # smooth + noise: every filter type gets chosen


def main():
    os.makedirs(OUT, exist_ok=True)
    cases = {
        "rgb_33x47": Image.fromarray(structured(33, 47, 3, 1), "RGB"),
        "rgba_20x31": Image.fromarray(structured(20, 31, 4, 2), "RGBA"),
        "gray_45x37": Image.fromarray(structured(45, 37, 1, 3)[..., 0], "L"),
        "graya_18x22": Image.fromarray(structured(18, 22, 2, 4), "LA"),
        "palette_40x40": Image.fromarray(structured(40, 40, 3, 5), "RGB").quantize(colors=37),
        "rgb_224x224": Image.fromarray(structured(224, 224, 3, 6), "RGB"),
        "rgb_1x1": Image.fromarray(structured(1, 1, 3, 7), "RGB"),
    }
    for name, im in cases.items():
        buf = io.BytesIO()
        im.save(buf, format="PNG", optimize=(name != "rgb_224x224"))
        data = buf.getvalue()
        open(os.path.join(OUT, name + ".png"), "wb").write(data)
        np.save(os.path.join(OUT, name + ".npy"), np.asarray(Image.open(io.BytesIO(data)).convert("RGB")))
        print(name, len(data), "bytes")
    # what Pillow does not write: Adam7-interlaced files, 16-bit RGB(A), sub-byte depths, every filter on every kind of
    # line - encoded by tests/png_writer.py, expected pixels from Pillow reading those bytes
    import sys
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
    from tests import png_writer as pw
    rng = np.random.default_rng(21)
    for name, (ctype, depth, interlace, h, w) in {
            "adam7_rgb_33x47": (2, 8, True, 33, 47), "adam7_rgba16_9x17": (6, 16, True, 9, 17), "rgb16_20x31": (2, 16, False, 20, 31),
            "gray16_clip_12x40": (0, 16, False, 12, 40), "adam7_graya16_7x5": (4, 16, True, 7, 5), "gray1_19x43": (0, 1, False, 19, 43),
            "adam7_gray2_10x23": (0, 2, True, 10, 23), "palette4_short_25x25": (3, 4, False, 25, 25), "adam7_palette1_16x9": (3, 1, True, 16, 9),
            "adam7_palette8_5x1": (3, 8, True, 5, 1)}.items():
        samples, pal = pw.random_image(rng, h, w, depth, ctype)
        data = pw.encode(samples, depth, ctype, interlace, pal, seed=len(name), idat_split=(61 if "adam7" in name else 0))
        open(os.path.join(OUT, name + ".png"), "wb").write(data)
        np.save(os.path.join(OUT, name + ".npy"), np.asarray(Image.open(io.BytesIO(data)).convert("RGB")))
        print(name, len(data), "bytes")
    # refused: not one of the image types of the specification (4-bit RGB)
    samples, _ = pw.random_image(rng, 8, 8, 4, 0)
    data = bytearray(pw.encode(samples, 4, 0))
    data[25] = 2  # colour type byte of IHDR (the CRC is not checked by anybody here)
    open(os.path.join(OUT, "unsupported_rgb4_8x8.png"), "wb").write(bytes(data))


if __name__ == "__main__":
    main()
