#!/usr/bin/env python
"""Writes tests/golden/reference_fixtures.npz and reference_fixture_report.json from the reference's OWN test
images (decoded here, where /root/reference exists; the GPU box has no reference tree, so the decoded pixels
travel as a small fixture file).

    python tests/golden/make_reference_fixtures.py [/root/reference]

Fixtures (reference test that uses each):
  alpha_original / alpha_half   tests/regression/image/resize-with-alpha-original{,-half-size}.png
                                tests/regression.rs:154-174 (resizing_with_alpha), 188-214 (resizing_with_catmul)
  tp1n3p08                      tests/images/png/transparency/tp1n3p08.png  (palette + tRNS -> Rgba8, as png's
                                decoder expands it) -- src/imageops/sample.rs:1861-1897 (resize_transparent_image)
  tinycross                     examples/scaleup/tinycross.png -- examples/scaleup/main.rs
  fractal                       examples/fractal.png (800x800 Rgb8) -- a real image for the blur tests
  rgb16                         tests/images/tiff/testsuite/rgb-3c-16b.tiff (157x151 Rgb16, uncompressed strips, read by
                                the few lines below) -- src/imageops/mod.rs:535-555
                                (fast_blur_approximates_gaussian_blur_well)
Report: the committed outputs of examples/scaledown/main.rs (scaledown-test-{near,tri,cmr,gauss,lcz2}.png, written
by 2019 code from a JPEG decoded by 2019's decoder) against the oracle's imageops::resize of PIL's decode of
test.jpg -- INFORMATIONAL (two JPEG decoders differ by a level or two per sample before any resampling); the
12 MP decoded JPEG is too large to commit, so these numbers are computed here, once, by this script.
"""
import json
import os
import struct
import sys

import numpy as np
from PIL import Image

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"


def png(path, mode=None):
    im = Image.open(os.path.join(REF, path))
    if mode:
        im = im.convert(mode)
    return np.ascontiguousarray(np.asarray(im))


def tiff_rgb16(path):
    """Baseline TIFF, uncompressed, chunky RGB, 16 bits per sample."""
    d = open(os.path.join(REF, path), "rb").read()
    bo = "<" if d[:2] == b"II" else ">"
    off = struct.unpack(bo + "I", d[4:8])[0]
    n = struct.unpack(bo + "H", d[off:off + 2])[0]
    tags = {}
    for i in range(n):
        tag, typ, cnt = struct.unpack(bo + "HHI", d[off + 2 + 12 * i:off + 10 + 12 * i])
        raw = d[off + 10 + 12 * i:off + 14 + 12 * i]
        size = {1: 1, 2: 1, 3: 2, 4: 4}[typ] * cnt
        if size > 4:
            p = struct.unpack(bo + "I", raw)[0]
            raw = d[p:p + size]
        fmt = {1: "B", 2: "c", 3: "H", 4: "I"}[typ]
        tags[tag] = struct.unpack(bo + fmt * cnt, raw[:size])
    w, h = tags[256][0], tags[257][0]
    assert tags[259][0] == 1 and tags[262][0] == 2 and tags[277][0] == 3 and tags[258] == (16, 16, 16) and tags[284][0] == 1
    data = b"".join(d[o:o + c] for o, c in zip(tags[273], tags[279]))
    a = np.frombuffer(data, dtype=(bo + "u2")).astype(np.uint16).reshape(h, w, 3)
    return np.ascontiguousarray(a)


def main():
    import oracle
    fx = {
        "alpha_original": png("tests/regression/image/resize-with-alpha-original.png"),
        "alpha_half": png("tests/regression/image/resize-with-alpha-original-half-size.png"),
        "tp1n3p08": png("tests/images/png/transparency/tp1n3p08.png", "RGBA"),
        "tinycross": png("examples/scaleup/tinycross.png"),
        "fractal": png("examples/fractal.png"),
        "rgb16": tiff_rgb16("tests/images/tiff/testsuite/rgb-3c-16b.tiff"),
    }
    for k, v in fx.items():
        print(k, v.shape, v.dtype)
    np.savez_compressed(os.path.join(HERE, "reference_fixtures.npz"), **fx)

    # scaledown (informational): oracle imageops::resize of PIL's decode against the committed outputs
    report = {"note": "informational: JPEG decoders differ before resampling; 2019 outputs; acc_init = 0 (the intended filter; "
                      "the 2019 code had no accumulator offset)"}
    src = png("examples/scaledown/test.jpg")
    for name, filt in (("near", "nearest"), ("tri", "triangle"), ("cmr", "catmullrom"), ("gauss", "gaussian"), ("lcz2", "lanczos3")):
        want = png(f"examples/scaledown/scaledown-test-{name}.png")
        h, w = want.shape[:2]
        entry = {"golden_shape": [int(h), int(w)]}
        for acc in (0.0, 1.0):
            got = oracle.resize(src, w, h, filt, acc_init=acc, threads=8)
            sd = got.astype(int) - want.astype(int)
            d = np.abs(sd)
            entry[f"acc_init_{acc:g}"] = {"max_abs": int(d.max()), "mean_abs": round(float(d.mean()), 4),
                                          "mean_signed": round(float(sd.mean()), 4),
                                          "within_2": round(float((d <= 2).mean()), 4), "exact": round(float((d == 0).mean()), 4)}
        report[name] = entry
        print(name, entry)
    json.dump(report, open(os.path.join(HERE, "reference_fixture_report.json"), "w"), indent=1)


if __name__ == "__main__":
    main()
