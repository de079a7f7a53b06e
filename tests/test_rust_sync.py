"""rust/image-b200 (authored, not compilable in this image -- no Rust toolchain) against the reference's own
source text: every hot-path method of `DynamicImage` exists in the crate's `DynamicImageB200` trait with the
reference's receiver, argument names, argument types and return type; `GaussianBlurParameters` has the
reference's constructors; the free functions mirror `imageops`; `DeviceBatch` has the owning-type surface.
Reference signatures are the literal lines of /root/reference at the cited positions (restated here because the
GPU box has no reference tree)."""
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
RS = os.path.join(ROOT, "rust", "image-b200", "src")

# src/images/dynimage.rs:320-568, 1116-1337 (receiver, args, return type)
DYNIMAGE = {
    "to_rgb8": ("&self", "", "RgbImage"), "to_rgb16": ("&self", "", "Rgb16Image"), "to_rgb32f": ("&self", "", "Rgb32FImage"),
    "to_rgba8": ("&self", "", "RgbaImage"), "to_rgba16": ("&self", "", "Rgba16Image"), "to_rgba32f": ("&self", "", "Rgba32FImage"),
    "to_luma8": ("&self", "", "GrayImage"), "to_luma16": ("&self", "", "Gray16Image"), "to_luma32f": ("&self", "", "Gray32FImage"),
    "to_luma_alpha8": ("&self", "", "GrayAlphaImage"), "to_luma_alpha16": ("&self", "", "GrayAlpha16Image"),
    "to_luma_alpha32f": ("&self", "", "GrayAlpha32FImage"),
    "into_rgb8": ("self", "", "RgbImage"), "into_rgb16": ("self", "", "Rgb16Image"), "into_rgb32f": ("self", "", "Rgb32FImage"),
    "into_rgba8": ("self", "", "RgbaImage"), "into_rgba16": ("self", "", "Rgba16Image"), "into_rgba32f": ("self", "", "Rgba32FImage"),
    "into_luma8": ("self", "", "GrayImage"), "into_luma16": ("self", "", "Gray16Image"), "into_luma32f": ("self", "", "Gray32FImage"),
    "into_luma_alpha8": ("self", "", "GrayAlphaImage"), "into_luma_alpha16": ("self", "", "GrayAlpha16Image"),
    "into_luma_alpha32f": ("self", "", "GrayAlpha32FImage"),
    "resize": ("&mut self", "nwidth: u32, nheight: u32, filter: FilterType", ""),
    "resize_exact": ("&mut self", "nwidth: u32, nheight: u32, filter: FilterType", ""),
    "resize_to_fill": ("&mut self", "nwidth: u32, nheight: u32, filter: FilterType", ""),
    "thumbnail": ("&self", "nwidth: u32, nheight: u32", "DynamicImage"),
    "thumbnail_exact": ("&self", "nwidth: u32, nheight: u32", "DynamicImage"),
    "blur": ("&self", "radius: f32", "DynamicImage"),
    "blur_advanced": ("&self", "parameters: GaussianBlurParameters", "DynamicImage"),
    "fast_blur": ("&self", "sigma: f32", "DynamicImage"),
    "unsharpen": ("&self", "sigma: f32, threshold: i32", "DynamicImage"),
    "filter3x3": ("&self", "kernel: &[f32; 9]", "DynamicImage"),
    "adjust_contrast": ("&self", "c: f32", "DynamicImage"),
    "brighten": ("&self", "value: i32", "DynamicImage"),
    "huerotate": ("&self", "value: i32", "DynamicImage"),
    "invert": ("&mut self", "", ""),
}
# the reference's image type aliases (src/images/buffer.rs) in the spelled-out form the crate may use
ALIASES = {"RgbImage": ["image::RgbImage"], "RgbaImage": ["image::RgbaImage"], "GrayImage": ["image::GrayImage"],
           "GrayAlphaImage": ["image::GrayAlphaImage"], "Rgb32FImage": ["image::Rgb32FImage"], "Rgba32FImage": ["image::Rgba32FImage"],
           "Rgb16Image": ["ImageBuffer<Rgb<u16>, Vec<u16>>"], "Rgba16Image": ["ImageBuffer<Rgba<u16>, Vec<u16>>"],
           "Gray16Image": ["ImageBuffer<Luma<u16>, Vec<u16>>"], "GrayAlpha16Image": ["ImageBuffer<LumaA<u16>, Vec<u16>>"],
           "Gray32FImage": ["ImageBuffer<Luma<f32>, Vec<f32>>"], "GrayAlpha32FImage": ["ImageBuffer<LumaA<f32>, Vec<f32>>"],
           "DynamicImage": ["DynamicImage"], "": [""]}


def _trait_sigs():
    txt = open(os.path.join(RS, "dynimage.rs")).read()
    body = txt[txt.index("pub trait DynamicImageB200"):txt.index("impl DynamicImageB200 for DynamicImage")]
    out = {}
    for m in re.finditer(r"fn b200_(\w+)\((&mut self|&self|self)(?:,\s*([^)]*))?\)(?:\s*->\s*([^;]+))?;", body):
        out[m.group(1)] = (m.group(2), (m.group(3) or "").strip(), (m.group(4) or "").strip())
    return out


def test_dynamic_image_trait_mirrors_the_reference():
    got = _trait_sigs()
    assert set(got) == set(DYNIMAGE), set(got) ^ set(DYNIMAGE)
    for name, (recv, args, ret) in DYNIMAGE.items():
        g = got[name]
        assert g[0] == recv, (name, g[0], recv)
        assert g[1] == args, (name, g[1], args)
        assert g[2] in ALIASES[ret], (name, g[2], ret)


def test_reference_signatures_are_the_reference_s(tmp_path):
    """Where the reference tree is present, the table above is checked against its source text."""
    ref = "/root/reference/src/images/dynimage.rs"
    if not os.path.exists(ref):
        import pytest
        pytest.skip("no reference tree here")
    txt = open(ref).read()
    for name, (recv, args, ret) in DYNIMAGE.items():
        m = re.search(r"pub fn %s\(\s*(&mut self|&self|self)(?:,\s*([^)]*?))?,?\s*\)(?:\s*->\s*(\w+))?\s*\{" % name, txt)
        assert m, name
        ref_args = re.sub(r"\s+", " ", (m.group(2) or "").strip()).replace("imageops::FilterType", "FilterType")
        assert (m.group(1), ref_args, m.group(3) or "") == (recv, args, ret), (name, m.groups())


def test_every_trait_method_is_implemented_and_gaussian_parameters_and_batch_exist():
    dyn = open(os.path.join(RS, "dynimage.rs")).read()
    impl = dyn[dyn.index("impl DynamicImageB200 for DynamicImage"):]
    for name in DYNIMAGE:
        assert re.search(r"fn b200_%s\(" % name, impl), name
    lib = open(os.path.join(RS, "lib.rs")).read()
    for ctor in ("new_from_radius(radius: f32)", "new_from_kernel_size(kernel_size: f32)", "new_from_sigma(sigma: f32)",
                 "new_anisotropic_kernel_size(x_axis_kernel_size: f32, y_axis_kernel_size: f32)", "SMOOTHING_3", "SMOOTHING_5", "SMOOTHING_7"):
        assert ctor in lib, ctor
    for fn in ("pub fn resize<I: GenericImageView>(image: &I, nwidth: u32, nheight: u32, filter: FilterType)",
               "pub fn blur<I: GenericImageView>(image: &I, radius: f32)",
               "pub fn blur_advanced<I: GenericImageView>(image: &I, p: GaussianBlurParameters)",
               "pub fn fast_blur<P: Pixel>(image: &ImageBuffer<P, Vec<P::Subpixel>>, sigma: f32)",
               "pub fn thumbnail<I: GenericImageView>(image: &I, new_width: u32, new_height: u32)",
               "pub fn from_raw_bgr<", "pub fn into_raw_bgr<", "pub fn from_raw_bgra<", "pub fn into_raw_bgra<"):
        assert fn in lib, fn
    dev = open(os.path.join(RS, "device.rs")).read()
    for fn in ("pub struct DeviceBatch<P: Pixel>", "pub fn from_host(", "pub fn to_host(", "pub fn resize(", "pub fn blur(", "pub fn blur_advanced(",
               "pub fn fast_blur(", "pub fn thumbnail(", "pub fn convert<To: Pixel>(", "pub fn swizzle_02(", "impl<P: Pixel> Drop for DeviceBatch<P>"):
        assert fn in dev, fn
    # every FFI symbol the wrappers call is declared
    ffi = open(os.path.join(RS, "ffi.rs")).read()
    declared = set(re.findall(r"fn (b200img_[a-z0-9_]+)", ffi))
    for src in (lib, dyn, dev):
        for sym in re.findall(r"ffi::(b200img_[a-z0-9_]+)\(", src):
            assert sym in declared, sym
