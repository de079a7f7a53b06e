"""Golden-vector cases: seeded inputs, the oracle call and the CUDA call (bit-exact configuration) of each.

The reference is Rust and cannot run here, so these vectors freeze the ORACLE's output (pinned in turn
by the reference's own known-answer tests, tests/test_oracle_kat.py).  `make_golden.py` writes
`digests.json` from the oracle; `tests/test_golden.py` checks that the oracle still reproduces it
(compiler, flags or libm drift) and -- on the GPU, without the oracle in the loop -- that the CUDA
path produces the same bytes."""
import hashlib

import numpy as np


def lcg_image(seed, h, w, c, dtype):
    """The reference's test LCG (src/imageops/fast_blur.rs:519-522), literally, one value per subpixel."""
    n = h * w * c
    state = seed & 0xFFFFFFFFFFFFFFFF
    out = np.empty(n, np.uint64)
    for i in range(n):
        state = (state * 6364136223846793005 + 1) & 0xFFFFFFFFFFFFFFFF
        out[i] = state >> 32
    dtype = np.dtype(dtype)
    if dtype == np.uint8:
        a = (out & 0xFF).astype(np.uint8)
    elif dtype == np.uint16:
        a = (out & 0xFFFF).astype(np.uint16)
    else:
        a = (out.astype(np.float64) / 2.0 ** 32).astype(np.float32)
    return a.reshape(h, w, c)


def digest(a) -> str:
    a = np.ascontiguousarray(a)
    return hashlib.sha256(str(a.shape).encode() + str(a.dtype).encode() + a.tobytes()).hexdigest()


def cases():
    """yields (name, input, oracle_fn, gpu_fn or None); gpu_fn runs the bit-exact configuration"""
    import oracle as o
    from image_b200 import FilterType, GaussianBlurParameters, imageops
    from image_b200 import _lib as L
    S = L.FP_STRICT
    U8, U16, F32 = np.uint8, np.uint16, np.float32
    filt = [("nearest", FilterType.Nearest), ("triangle", FilterType.Triangle), ("catmullrom", FilterType.CatmullRom),
            ("gaussian", FilterType.Gaussian), ("lanczos3", FilterType.Lanczos3)]
    shapes = [("u8c4_down", (48, 64, 4, U8), (31, 23)), ("u16c3_up", (20, 24, 3, U16), (50, 41)), ("f32c1_mixed", (30, 18, 1, F32), (9, 44)),
              ("u8c3_down", (45, 61, 3, U8), (20, 19))]
    for sname, (h, w, c, dt), (nw, nh) in shapes:
        img = lcg_image(11 + h + c, h, w, c, dt)
        for fname, f in filt:
            for acc in (1.0, 0.0):
                yield (f"resize/{sname}/{fname}/acc{acc:g}", img,
                       lambda a, fname=fname, nw=nw, nh=nh, acc=acc: o.resize(a, nw, nh, fname, acc_init=acc),
                       lambda a, f=f, nw=nw, nh=nh, acc=acc: imageops.resize(a, nw, nh, f, acc_init=acc, flags=S))
    for (h, w, c, dt) in ((40, 52, 3, U8), (33, 29, 4, U16), (25, 31, 2, F32), (37, 40, 1, U8)):
        img = lcg_image(101 + c, h, w, c, dt)
        tag = f"{np.dtype(dt).name}c{c}"
        for pname, pg, po in (("radius2.5", GaussianBlurParameters.new_from_radius(2.5), o.gauss_params_from_radius(2.5)),
                              ("sigma3", GaussianBlurParameters.new_from_sigma(3.0), o.gauss_params_from_sigma(3.0)),
                              ("aniso5x9", GaussianBlurParameters.new_anisotropic_kernel_size(5.0, 9.0), o.gauss_params_from_anisotropic(5.0, 9.0)),
                              ("tiny", GaussianBlurParameters.new_from_sigma(0.1), o.gauss_params_from_sigma(0.1))):
            yield (f"blur_generic/{tag}/{pname}", img, lambda a, po=po: o.gaussian_blur(a, po, "generic"),
                   lambda a, pg=pg: imageops.blur_advanced(a, pg, flags=S))
            yield (f"blur_dyn/{tag}/{pname}", img, lambda a, po=po: o.gaussian_blur(a, po, "dyn"),
                   lambda a, pg=pg: imageops.blur_dyn(a, pg, flags=S))
        for sigma in (0.7, 2.0, 10.0):
            # f32 / u16 box sums accumulate in f32 in a streaming order the GPU reproduces for u8 only
            yield (f"fast_blur/{tag}/sigma{sigma:g}", img, lambda a, sigma=sigma: o.fast_blur(a, sigma),
                   (lambda a, sigma=sigma: imageops.fast_blur(a, sigma)) if dt == U8 else None)
        yield (f"filter3x3/{tag}/sharpen", img, lambda a: o.filter3x3(a, [0, -1, 0, -1, 5, -1, 0, -1, 0]),
               lambda a: imageops.filter3x3(a, [0, -1, 0, -1, 5, -1, 0, -1, 0]))
        yield (f"unsharpen/{tag}/s1.5t3", img, lambda a: o.unsharpen(a, 1.5, 3), lambda a: imageops.unsharpen(a, 1.5, 3, flags=S))
        for (nw, nh) in ((13, 9), (80, 70), (20, 60)):
            yield (f"thumbnail/{tag}/{nw}x{nh}", img, lambda a, nw=nw, nh=nh: o.thumbnail(a, nw, nh),
                   (lambda a, nw=nw, nh=nh: imageops.thumbnail(a, nw, nh)) if dt != F32 else None)
        yield (f"invert/{tag}", img, o.invert, lambda a: imageops.invert(a.copy()))
        yield (f"brighten/{tag}/-37", img, lambda a: o.brighten(a, -37), lambda a: imageops.brighten(a, -37))
        yield (f"contrast/{tag}/22.5", img, lambda a: o.contrast(a, 22.5), lambda a: imageops.contrast(a, 22.5))
        yield (f"huerotate/{tag}/100", img, lambda a: o.huerotate(a, 100), lambda a: imageops.huerotate(a, 100))
        yield (f"grayscale/{tag}", img, o.grayscale, imageops.grayscale)
        yield (f"grayscale_alpha/{tag}", img, lambda a: o.grayscale(a, alpha=True), imageops.grayscale_alpha)
    for sc, sdt in ((4, U16), (3, U8), (1, F32), (2, U8), (4, F32)):
        src = lcg_image(7 + sc, 16, 33, sc, sdt)
        for dc, ddt in ((3, U8), (4, F32), (1, U16), (2, U8), (4, U16)):
            for mode, mname in ((o.CAST, "cast"), (o.CONVERT, "convert")):
                yield (f"convert/{np.dtype(sdt).name}c{sc}_to_{np.dtype(ddt).name}c{dc}/{mname}", src,
                       lambda a, sc=sc, dc=dc, ddt=ddt, mode=mode: o.convert(a.reshape(-1, sc), sc, ddt, dc, mode),
                       lambda a, dc=dc, ddt=ddt, mode=mode: imageops.convert_pixels(a, dc, ddt, mode=mode).reshape(-1, dc))
    for c, dt in ((3, U8), (4, U8), (4, U16), (3, F32)):
        buf = lcg_image(55 + c, 1, 101, c, dt).reshape(-1)
        yield (f"swizzle/{np.dtype(dt).name}c{c}", buf, lambda a, c=c: o.swizzle_02(a, c),
               lambda a, c=c: imageops.swizzle_02(a.copy(), c))
