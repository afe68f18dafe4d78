#!/usr/bin/env python
"""Regenerates tests/golden/jpeg/cases.npz: small JPEG files written by Pillow (libjpeg-turbo) over the encoder options
the decoder has to cope with, each with the RGB image libjpeg-turbo decodes from it.  These fixtures pin
oracle/jpeg_oracle.c (and through it the CUDA decoder) to libjpeg's integer slow IDCT, fancy upsampling and colour
tables -- the defaults DecodeJpeg(channels=3) runs with (EX/common/GraphBuilder.java:41-47).  The reference itself holds
no JPEG fixture.  Needs Pillow only; the .npz is committed and travels to the GPU box."""
import io
import os

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))


def picture(rng, h, w, gray=False, noise=24.0):
    """smooth colour ramps + blobs + noise: every AC band populated, Huffman codes of every length class"""
    y, x = np.mgrid[0:h, 0:w].astype(np.float64)
    ch = []
    for c in range(1 if gray else 3):
        f = rng.uniform(0.02, 0.3, 4)
        p = rng.uniform(0, 6.28, 4)
        v = 128 + 60 * np.sin(f[0] * x + p[0]) * np.cos(f[1] * y + p[1]) + 50 * np.sin(f[2] * (x + y) + p[2])
        v += rng.normal(0, noise, (h, w))
        ch.append(v)
    a = np.clip(np.stack(ch, -1), 0, 255).astype(np.uint8)
    return a[:, :, 0] if gray else a


def main():
    rng = np.random.default_rng(20261107)
    cases = []

    def add(name, arr, **kw):
        b = io.BytesIO()
        Image.fromarray(arr).save(b, "JPEG", **kw)
        data = b.getvalue()
        rgb = np.array(Image.open(io.BytesIO(data)).convert("RGB"))
        cases.append((name, np.frombuffer(data, np.uint8), rgb))

    for ss, tag in ((0, "444"), (1, "422"), (2, "420")):
        for (h, w) in ((48, 64), (17, 33), (33, 17), (8, 8), (5, 3), (1, 1), (16, 5), (40, 41)):
            add(f"{tag}_{h}x{w}_q75", picture(rng, h, w), quality=75, subsampling=ss)
        add(f"{tag}_q30", picture(rng, 56, 72), quality=30, subsampling=ss)
        add(f"{tag}_q98", picture(rng, 56, 72, noise=60), quality=98, subsampling=ss)
        add(f"{tag}_q100_sat", (rng.integers(0, 2, (40, 48, 3)) * 255).astype(np.uint8), quality=100, subsampling=ss)
        add(f"{tag}_opt", picture(rng, 64, 80), quality=85, subsampling=ss, optimize=True)
        add(f"{tag}_rst_blocks", picture(rng, 64, 80), quality=80, subsampling=ss, restart_marker_blocks=3)
        add(f"{tag}_rst_rows", picture(rng, 70, 90), quality=80, subsampling=ss, restart_marker_rows=1)
        add(f"{tag}_flat", np.full((24, 40, 3), 77, np.uint8), quality=75, subsampling=ss)
    for (h, w) in ((48, 64), (17, 33), (1, 1), (9, 70)):
        add(f"gray_{h}x{w}", picture(rng, h, w, gray=True), quality=75)
    add("gray_rst", picture(rng, 40, 56, gray=True), quality=60, restart_marker_blocks=5)
    add("refuse_progressive", picture(rng, 32, 32), quality=75, progressive=True)

    out = {}
    for name, data, rgb in cases:
        out[name + ".jpg"] = data
        out[name + ".rgb"] = rgb
    path = os.path.join(HERE, "jpeg", "cases.npz")
    np.savez_compressed(path, **out)
    print(len(cases), "cases,", os.path.getsize(path), "bytes")


if __name__ == "__main__":
    main()
