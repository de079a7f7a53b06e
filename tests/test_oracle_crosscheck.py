"""A second, independent restatement of the resampler (numpy float32, written from SURVEY.md
Appendix A.1 / A.2 and src/imageops/sample.rs:146-198) against the C++ oracle, bit for bit.
Only the libm-free filters (Nearest, Triangle, CatmullRom) are used, so the comparison does not
depend on which sin / exp implementation numpy links."""
import math

import numpy as np
import pytest

import oracle as o

F = np.float32


def _triangle(x):
    a = F(abs(x))
    return F(1.0) - a if a < F(1.0) else F(0.0)


def _catmullrom(x):                      # bc_cubic_spline(x, b = 0, c = 0.5), sample.rs:148-165
    a, b, c = F(abs(x)), F(0.0), F(0.5)
    a2, a3 = a * a, a * (a * a)          # powi(2), powi(3) of compiler-rt: square-and-multiply
    if a < F(1.0):
        k = (F(12.0) - F(9.0) * b - F(6.0) * c) * a3 + (F(-18.0) + F(12.0) * b + F(6.0) * c) * a2 + (F(6.0) - F(2.0) * b)
    elif a < F(2.0):
        k = (-b - F(6.0) * c) * a3 + (F(6.0) * b + F(30.0) * c) * a2 + (F(-12.0) * b - F(48.0) * c) * a + (F(8.0) * b + F(24.0) * c)
    else:
        k = F(0.0)
    return k / F(6.0)


KERNELS = {"nearest": (lambda x: F(1.0), F(0.0)), "triangle": (_triangle, F(1.0)), "catmullrom": (_catmullrom, F(2.0))}


def weights(name, src_len, dst_len, axis):
    """Appendix A.1: axis 0 = vertical_sample's rule (f32 centre), 1 = horizontal_sample's (f64 centre)"""
    kern, support = KERNELS[name]
    ratio = F(src_len) / F(dst_len)
    sratio = F(1.0) if ratio < F(1.0) else ratio
    ss = support * sratio
    out = []
    for o_ in range(dst_len):
        if axis == 0:
            c = (F(o_) + F(0.5)) * ratio
            left = min(max(int(math.floor(c - ss)), 0), src_len - 1)
            right = min(max(int(math.ceil(c + ss)), left + 1), src_len)
            c = c - F(0.5)
        else:
            c64 = (np.float64(o_) + 0.5) * np.float64(ratio)
            lf = c64 - np.float64(ss)
            left = 0 if not lf > 0 else min(int(lf), 2 ** 32 - 1)     # `as u32`: truncating, saturating
            left = min(max(left, 0), src_len - 1)
            right = min(max(int(math.ceil(c64 + np.float64(ss))), left + 1), src_len)
            c = F(c64) - F(0.5)
        w = [kern((F(i) - c) / sratio) for i in range(left, right)]
        s = F(0.0)
        for v in w:
            s = s + v
        out.append((left, [v / s for v in w]))
    return out


def resize(img, nw, nh, name, acc_init):
    """Appendix A.2: vertical pass into f32 (accumulator starts at acc_init), then horizontal, then the store rule"""
    h, w, c = img.shape
    wv, wh = weights(name, h, nh, 0), weights(name, w, nw, 1)
    tmp = np.empty((nh, w, c), F)
    for oy, (left, ws) in enumerate(wv):
        acc = np.full((w, c), F(acc_init), F)
        for i, wt in enumerate(ws):
            acc = acc + img[left + i].astype(F) * wt           # numpy keeps every product and sum in f32
        tmp[oy] = acc
    out = np.empty((nh, nw, c), img.dtype)
    for ox, (left, ws) in enumerate(wh):
        t = np.zeros((nh, c), F)
        for i, wt in enumerate(ws):
            t = t + tmp[:, left + i] * wt
        if img.dtype == np.float32:
            out[:, ox] = np.minimum(np.maximum(t, F(0.0)), F(1.0))
        else:
            mx = np.iinfo(img.dtype).max
            v = t + F(0.5)
            out[:, ox] = np.where(v >= mx + 1, mx, np.where(v > 0, np.trunc(np.minimum(v, F(mx + 1))), 0)).astype(img.dtype)
    return out


@pytest.mark.parametrize("name", ["nearest", "triangle", "catmullrom"])
@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32])
def test_oracle_resize_matches_independent_numpy(name, dtype):
    rng = np.random.default_rng(17)
    for (h, w, c, nw, nh) in ((13, 17, 3, 7, 5), (9, 11, 4, 23, 19), (16, 5, 1, 5, 40), (6, 30, 2, 11, 6)):
        if dtype == np.float32:
            img = rng.random((h, w, c), dtype=np.float32)
        else:
            img = rng.integers(0, np.iinfo(dtype).max + 1, (h, w, c), dtype=dtype)
        for acc in (1.0, 0.0):
            want = resize(img, nw, nh, name, acc)
            got = o.resize(img, nw, nh, name, acc_init=acc)
            assert np.array_equal(got, want), (name, dtype, (h, w, c, nw, nh), acc, np.abs(got.astype(np.float64) - want).max())


@pytest.mark.parametrize("name", ["nearest", "triangle", "catmullrom"])
def test_oracle_weight_tables_match_independent_numpy(name):
    for axis in (0, 1):
        for (src, dst) in ((17, 7), (7, 17), (100, 33), (5, 5), (1, 9), (64, 1)):
            left, cnt, w = o.resize_weights(name, src, dst, axis)
            mine = weights(name, src, dst, axis)
            for i, (l, ws) in enumerate(mine):
                assert (left[i], cnt[i]) == (l, len(ws)), (name, axis, src, dst, i)
                assert np.array_equal(w[i, :len(ws)], np.array(ws, F)), (name, axis, src, dst, i)


# ---- the separable convolution engine (src/imageops/filter_1d.rs) ----------
# The 1-D kernels come from the oracle (they need expf / powf); what is rebuilt here is everything
# after them: Q0.15 transform, zero-end stripping, replicate padding, the paired summation order of
# the row and column passes, the intermediate of the pixel type, and the store rules.
def _strip(k):
    k = list(k)
    while len(k) > 2 and k[0] == 0 and k[-1] == 0:
        k = k[1:-1]
    return k


def _pass(a, k, axis, q15, maxv):
    """a: (h, w, c) array of f32 or uint64 values; symmetric pass along `axis` with clamp-to-edge"""
    n = a.shape[axis]
    half = len(k) // 2
    idx = np.arange(n)
    take = lambda off: np.take(a, np.clip(idx + off, 0, n - 1), axis=axis)
    if q15:
        acc = take(0) * np.uint64(k[half])
        for i in range(half):
            acc = acc + (take(-(half - i)) + take(half - i)) * np.uint64(k[i])
        return np.minimum((acc + np.uint64(16384)) >> np.uint64(15), np.uint64(maxv))
    acc = take(0) * F(k[half])
    for i in range(half):
        acc = acc + (take(-(half - i)) + take(half - i)) * F(k[i])   # pair first, then multiply, then add: all f32
    return acc


def blur(img, kx, ky, path):
    q15 = path == "dyn" and img.dtype != np.float32
    if q15:
        maxv = np.iinfo(img.dtype).max
        tr = lambda k: _strip([int(min(max(F(v) * F(32768.0), F(0.0)), F(32767.0))) for v in k])
        mid = _pass(img.astype(np.uint64), tr(kx), 1, True, maxv)
        return _pass(mid, tr(ky), 0, True, maxv).astype(img.dtype)
    out = _pass(_pass(img.astype(F), _strip(kx), 1, False, 0), _strip(ky), 0, False, 0)
    if img.dtype == np.float32:
        return out
    maxv = np.iinfo(img.dtype).max
    v = out + F(0.5)
    return np.where(v >= maxv + 1, maxv, np.where(v > 0, np.trunc(np.minimum(v, F(maxv + 1))), 0)).astype(img.dtype)


@pytest.mark.parametrize("dtype", [np.uint8, np.uint16, np.float32])
@pytest.mark.parametrize("path", ["generic", "dyn"])
def test_oracle_gaussian_engine_matches_independent_numpy(dtype, path):
    rng = np.random.default_rng(23)
    for (h, w, c) in ((9, 14, 3), (20, 7, 1), (12, 12, 4)):
        img = rng.random((h, w, c), dtype=np.float32) if dtype == np.float32 else \
            rng.integers(0, np.iinfo(dtype).max + 1, (h, w, c), dtype=dtype)
        for p in (o.gauss_params_from_sigma(1.0), o.gauss_params_from_radius(3.0), o.gauss_params_from_sigma(2.6),
                  o.gauss_params_from_anisotropic(9.0, 3.0), o.gauss_params_from_sigma(10.0)):
            kx = o.gaussian_kernel_1d(p.x_kernel_size, p.x_sigma)
            ky = o.gaussian_kernel_1d(p.y_kernel_size, p.y_sigma)
            want = blur(img, kx, ky, path)
            got = o.gaussian_blur(img, p, path)
            assert np.array_equal(got, want), (dtype, path, (h, w, c), p)
