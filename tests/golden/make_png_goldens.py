"""Mint tests/golden/png/*: small PNG files written by Pillow in every supported flavour, each next to
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
    # refused flavours: 16-bit and interlaced (written by hand-free Pillow options where it can)
    im16 = Image.fromarray((structured(8, 8, 1, 8)[..., 0].astype(np.uint16) * 257), "I;16")
    buf = io.BytesIO()
    im16.save(buf, format="PNG")
    open(os.path.join(OUT, "unsupported_gray16_8x8.png"), "wb").write(buf.getvalue())


if __name__ == "__main__":
    main()
