//! `image-b200`: the `image` crate's data-parallel pixel hot path on an NVIDIA B200.
//!
//! Same signatures as the reference over `ImageBuffer<P, Vec<S>>` and `DynamicImage`
//! (`imageops::resize`, `blur`, `blur_advanced`, `fast_blur`, `ImageBuffer::convert`,
//! `DynamicImage::to_*`, `from_raw_bgra` / `into_raw_bgra`), forwarding to `libb200img`'s C ABI.
//! There is no CPU fallback: without a usable GPU every call panics with the library's message.
//!
//! NOTE: written in an environment without `rustc`/`cargo`; it mirrors `image_b200/imageops.py`
//! (which is exercised by the test-suite) line for line but has not itself been compiled.
pub mod ffi;

use std::ffi::{c_int, c_void, CStr};

use image::imageops::FilterType;
use image::{GenericImageView, ImageBuffer, Pixel, Primitive};

/// Subpixel types the kernels support.
pub trait B200Subpixel: Primitive + 'static {
    const DTYPE: c_int;
}
impl B200Subpixel for u8 { const DTYPE: c_int = ffi::U8; }
impl B200Subpixel for u16 { const DTYPE: c_int = ffi::U16; }
impl B200Subpixel for f32 { const DTYPE: c_int = ffi::F32; }

/// The reference's literal vertical-accumulator start (`Rgba([1.0; 4])`, sample.rs:547).
pub const ACC_INIT_REFERENCE: f32 = 1.0;

fn check(rc: c_int) {
    if rc == 0 { return; }
    let msg = unsafe { CStr::from_ptr(ffi::b200img_last_error()) }.to_string_lossy().into_owned();
    // -1: the reference panics on this input; -2: ImageError::Limits(DimensionError); -3/-4: no memory / no GPU
    panic!("b200img (status {rc}): {msg}");
}

fn filter_code(f: FilterType) -> c_int {
    match f {
        FilterType::Nearest => 0, FilterType::Triangle => 1, FilterType::CatmullRom => 2,
        FilterType::Gaussian => 3, FilterType::Lanczos3 => 4,
    }
}

/// Packs any view into a contiguous row-major buffer (the reference's `to_pixel_view` fast path
/// applies when the view already is one).
fn packed<I: GenericImageView>(image: &I) -> Vec<<I::Pixel as Pixel>::Subpixel> {
    let (w, h) = image.dimensions();
    let c = <I::Pixel as Pixel>::CHANNEL_COUNT as usize;
    let mut v = Vec::with_capacity(w as usize * h as usize * c);
    for y in 0..h { for x in 0..w { v.extend_from_slice(image.get_pixel(x, y).channels()); } }
    v
}

pub mod imageops {
    use super::*;
    pub use crate::ffi::b200img_gauss_params as GaussianBlurParameters;

    /// `image::imageops::resize` (src/imageops/sample.rs:970).
    pub fn resize<I: GenericImageView>(image: &I, nwidth: u32, nheight: u32, filter: FilterType)
        -> ImageBuffer<I::Pixel, Vec<<I::Pixel as Pixel>::Subpixel>>
    where <I::Pixel as Pixel>::Subpixel: B200Subpixel
    {
        resize_with(image, nwidth, nheight, filter, ACC_INIT_REFERENCE, ffi::FP_FAST)
    }

    /// `resize` with the accumulator start and arithmetic flags exposed.
    pub fn resize_with<I: GenericImageView>(image: &I, nwidth: u32, nheight: u32, filter: FilterType, acc_init: f32, flags: c_int)
        -> ImageBuffer<I::Pixel, Vec<<I::Pixel as Pixel>::Subpixel>>
    where <I::Pixel as Pixel>::Subpixel: B200Subpixel
    {
        let src = packed(image);
        let mut out = ImageBuffer::new(nwidth, nheight);
        check(unsafe { ffi::b200img_resize(src.as_ptr() as *const c_void, image.width(), image.height(),
            <I::Pixel as Pixel>::CHANNEL_COUNT as c_int, <<I::Pixel as Pixel>::Subpixel as B200Subpixel>::DTYPE,
            out.as_mut_ptr() as *mut c_void, nwidth, nheight, filter_code(filter), acc_init, flags) });
        out
    }

    pub fn params_from_radius(radius: f32) -> GaussianBlurParameters {
        let mut p = GaussianBlurParameters { x_kernel_size: 0, x_sigma: 0.0, y_kernel_size: 0, y_sigma: 0.0 };
        check(unsafe { ffi::b200img_gauss_params_from_radius(radius, &mut p) });
        p
    }
    pub fn params_from_sigma(sigma: f32) -> GaussianBlurParameters {
        let mut p = GaussianBlurParameters { x_kernel_size: 0, x_sigma: 0.0, y_kernel_size: 0, y_sigma: 0.0 };
        check(unsafe { ffi::b200img_gauss_params_from_sigma(sigma, &mut p) });
        p
    }

    fn gauss<I: GenericImageView>(image: &I, p: GaussianBlurParameters, path: c_int)
        -> ImageBuffer<I::Pixel, Vec<<I::Pixel as Pixel>::Subpixel>>
    where <I::Pixel as Pixel>::Subpixel: B200Subpixel
    {
        let src = packed(image);
        let mut out = ImageBuffer::new(image.width(), image.height());
        check(unsafe { ffi::b200img_gaussian_blur(src.as_ptr() as *const c_void, image.width(), image.height(),
            <I::Pixel as Pixel>::CHANNEL_COUNT as c_int, <<I::Pixel as Pixel>::Subpixel as B200Subpixel>::DTYPE,
            out.as_mut_ptr() as *mut c_void, &p, path, ffi::FP_FAST) });
        out
    }

    /// `image::imageops::blur` -- `radius`, not sigma, at this revision (sample.rs:1034).
    pub fn blur<I: GenericImageView>(image: &I, radius: f32) -> ImageBuffer<I::Pixel, Vec<<I::Pixel as Pixel>::Subpixel>>
    where <I::Pixel as Pixel>::Subpixel: B200Subpixel { gauss(image, params_from_radius(radius), ffi::BLUR_GENERIC) }

    /// `image::imageops::blur_advanced` (sample.rs:1050).
    pub fn blur_advanced<I: GenericImageView>(image: &I, p: GaussianBlurParameters) -> ImageBuffer<I::Pixel, Vec<<I::Pixel as Pixel>::Subpixel>>
    where <I::Pixel as Pixel>::Subpixel: B200Subpixel { gauss(image, p, ffi::BLUR_GENERIC) }

    /// The arithmetic of `DynamicImage::blur_advanced` (Q0.15 for u8 / u16; sample.rs:1297).
    pub fn blur_dyn<I: GenericImageView>(image: &I, p: GaussianBlurParameters) -> ImageBuffer<I::Pixel, Vec<<I::Pixel as Pixel>::Subpixel>>
    where <I::Pixel as Pixel>::Subpixel: B200Subpixel { gauss(image, p, ffi::BLUR_DYN) }

    /// `image::imageops::fast_blur` (src/imageops/fast_blur.rs:111).
    pub fn fast_blur<P: Pixel>(image: &ImageBuffer<P, Vec<P::Subpixel>>, sigma: f32) -> ImageBuffer<P, Vec<P::Subpixel>>
    where P::Subpixel: B200Subpixel
    {
        let (w, h) = image.dimensions();
        if w == 0 || h == 0 { return image.clone(); }
        let mut out = ImageBuffer::new(w, h);
        check(unsafe { ffi::b200img_fast_blur(image.as_ptr() as *const c_void, w, h, P::CHANNEL_COUNT as c_int,
            <P::Subpixel as B200Subpixel>::DTYPE, out.as_mut_ptr() as *mut c_void, sigma) });
        out
    }

    /// `image::imageops::thumbnail` (src/imageops/sample.rs:595).
    pub fn thumbnail<I: GenericImageView>(image: &I, new_width: u32, new_height: u32)
        -> ImageBuffer<I::Pixel, Vec<<I::Pixel as Pixel>::Subpixel>>
    where <I::Pixel as Pixel>::Subpixel: B200Subpixel
    {
        let src = packed(image);
        let mut out = ImageBuffer::new(new_width, new_height);
        check(unsafe { ffi::b200img_thumbnail(src.as_ptr() as *const c_void, image.width(), image.height(),
            <I::Pixel as Pixel>::CHANNEL_COUNT as c_int, <<I::Pixel as Pixel>::Subpixel as B200Subpixel>::DTYPE,
            out.as_mut_ptr() as *mut c_void, new_width, new_height, 0) });
        out
    }

    /// `image::imageops::filter3x3` (src/imageops/sample.rs:897).
    pub fn filter3x3<I: GenericImageView>(image: &I, kernel: &[f32; 9])
        -> ImageBuffer<I::Pixel, Vec<<I::Pixel as Pixel>::Subpixel>>
    where <I::Pixel as Pixel>::Subpixel: B200Subpixel
    {
        let src = packed(image);
        let mut out = ImageBuffer::new(image.width(), image.height());
        check(unsafe { ffi::b200img_filter3x3(src.as_ptr() as *const c_void, image.width(), image.height(),
            <I::Pixel as Pixel>::CHANNEL_COUNT as c_int, <<I::Pixel as Pixel>::Subpixel as B200Subpixel>::DTYPE,
            out.as_mut_ptr() as *mut c_void, kernel.as_ptr()) });
        out
    }

    /// `image::imageops::unsharpen` (src/imageops/sample.rs:1617).
    pub fn unsharpen<I: GenericImageView>(image: &I, sigma: f32, threshold: i32)
        -> ImageBuffer<I::Pixel, Vec<<I::Pixel as Pixel>::Subpixel>>
    where <I::Pixel as Pixel>::Subpixel: B200Subpixel
    {
        let src = packed(image);
        let mut out = ImageBuffer::new(image.width(), image.height());
        check(unsafe { ffi::b200img_unsharpen(src.as_ptr() as *const c_void, image.width(), image.height(),
            <I::Pixel as Pixel>::CHANNEL_COUNT as c_int, <<I::Pixel as Pixel>::Subpixel as B200Subpixel>::DTYPE,
            out.as_mut_ptr() as *mut c_void, sigma, threshold as c_int, ffi::FP_FAST) });
        out
    }

    // ---- colorops (src/imageops/colorops.rs); `f` is one of the elementwise entry points ----
    fn colorop<I: GenericImageView>(image: &I, f: impl FnOnce(*const c_void, u32, u32, c_int, c_int, *mut c_void) -> c_int)
        -> ImageBuffer<I::Pixel, Vec<<I::Pixel as Pixel>::Subpixel>>
    where <I::Pixel as Pixel>::Subpixel: B200Subpixel
    {
        let src = packed(image);
        let mut out = ImageBuffer::new(image.width(), image.height());
        check(f(src.as_ptr() as *const c_void, image.width(), image.height(), <I::Pixel as Pixel>::CHANNEL_COUNT as c_int,
                <<I::Pixel as Pixel>::Subpixel as B200Subpixel>::DTYPE, out.as_mut_ptr() as *mut c_void));
        out
    }

    /// `image::imageops::brighten` (colorops.rs:161).
    pub fn brighten<I: GenericImageView>(image: &I, value: i32) -> ImageBuffer<I::Pixel, Vec<<I::Pixel as Pixel>::Subpixel>>
    where <I::Pixel as Pixel>::Subpixel: B200Subpixel
    { colorop(image, |s, w, h, c, d, o| unsafe { ffi::b200img_brighten(s, w, h, c, d, o, value) }) }

    /// `image::imageops::contrast` (colorops.rs:92).
    pub fn contrast<I: GenericImageView>(image: &I, contrast: f32) -> ImageBuffer<I::Pixel, Vec<<I::Pixel as Pixel>::Subpixel>>
    where <I::Pixel as Pixel>::Subpixel: B200Subpixel
    { colorop(image, |s, w, h, c, d, o| unsafe { ffi::b200img_contrast(s, w, h, c, d, o, contrast) }) }

    /// `image::imageops::huerotate` (colorops.rs:232).
    pub fn huerotate<I: GenericImageView>(image: &I, value: i32) -> ImageBuffer<I::Pixel, Vec<<I::Pixel as Pixel>::Subpixel>>
    where <I::Pixel as Pixel>::Subpixel: B200Subpixel
    { colorop(image, |s, w, h, c, d, o| unsafe { ffi::b200img_huerotate(s, w, h, c, d, o, value) }) }

    /// `image::imageops::invert` (colorops.rs:70): in place.
    pub fn invert<P: Pixel>(image: &mut ImageBuffer<P, Vec<P::Subpixel>>)
    where P::Subpixel: B200Subpixel
    {
        let (w, h) = image.dimensions();
        let p = image.as_mut_ptr() as *mut c_void;
        check(unsafe { ffi::b200img_invert(p as *const c_void, w, h, P::CHANNEL_COUNT as c_int, <P::Subpixel as B200Subpixel>::DTYPE, p) });
    }

    /// `has_constant_alpha` of `resize_impl` (src/imageops/resize.rs:155).
    pub fn has_constant_alpha<P: Pixel>(image: &ImageBuffer<P, Vec<P::Subpixel>>) -> bool
    where P::Subpixel: B200Subpixel
    {
        let (w, h) = image.dimensions();
        let mut r: c_int = 1;
        check(unsafe { ffi::b200img_has_constant_alpha(image.as_ptr() as *const c_void, w, h, P::CHANNEL_COUNT as c_int,
            <P::Subpixel as B200Subpixel>::DTYPE, &mut r) });
        r != 0
    }
}

/// `ImageBuffer::convert::<To>()` (src/images/buffer.rs:1643): FromColor arithmetic.
pub fn convert<From: Pixel, To: Pixel>(image: &ImageBuffer<From, Vec<From::Subpixel>>) -> ImageBuffer<To, Vec<To::Subpixel>>
where From::Subpixel: B200Subpixel, To::Subpixel: B200Subpixel
{
    convert_mode(image, ffi::CONVERT, None)
}

/// `DynamicImage::to_*` (src/images/dynimage.rs:320): cast_pixels arithmetic; `coefs` = the
/// image's `derived_luminance()` (src/metadata/cicp.rs:1240), `None` = sRGB fallback.
pub fn cast<From: Pixel, To: Pixel>(image: &ImageBuffer<From, Vec<From::Subpixel>>, coefs: Option<[f32; 3]>) -> ImageBuffer<To, Vec<To::Subpixel>>
where From::Subpixel: B200Subpixel, To::Subpixel: B200Subpixel
{
    convert_mode(image, ffi::CAST, coefs)
}

fn convert_mode<From: Pixel, To: Pixel>(image: &ImageBuffer<From, Vec<From::Subpixel>>, mode: c_int, coefs: Option<[f32; 3]>)
    -> ImageBuffer<To, Vec<To::Subpixel>>
where From::Subpixel: B200Subpixel, To::Subpixel: B200Subpixel
{
    let (w, h) = image.dimensions();
    let mut out = ImageBuffer::new(w, h);
    let cp = coefs.as_ref().map_or(std::ptr::null(), |c| c.as_ptr());
    check(unsafe { ffi::b200img_convert(image.as_ptr() as *const c_void, <From::Subpixel as B200Subpixel>::DTYPE,
        From::CHANNEL_COUNT as c_int, out.as_mut_ptr() as *mut c_void, <To::Subpixel as B200Subpixel>::DTYPE,
        To::CHANNEL_COUNT as c_int, w as usize * h as usize, mode, cp) });
    out
}

/// `ImageBuffer::from_raw_bgra` / `into_raw_bgra` core (src/images/buffer.rs:1564-1582): in-place swap of
/// subpixels 0 and 2 of every full chunk of `channels` (3 or 4).
pub fn swizzle_02<S: B200Subpixel>(buf: &mut [S], channels: usize) {
    check(unsafe { ffi::b200img_swizzle_02(buf.as_mut_ptr() as *mut c_void, S::DTYPE, channels as c_int, buf.len()) });
}

/// Device-resident batch variant: `n` equally shaped images already in device memory.
/// # Safety
/// `d_src` / `d_dst` must be valid device pointers on the current CUDA device; `stream` a valid `cudaStream_t` or null.
pub unsafe fn resize_batch_dev(d_src: *const c_void, w: u32, h: u32, channels: u32, dtype: c_int, d_dst: *mut c_void,
                               nw: u32, nh: u32, n: u32, filter: FilterType, acc_init: f32, flags: c_int, stream: ffi::Stream) {
    check(ffi::b200img_resize_batch_dev(d_src, 0, w, h, channels as c_int, dtype, d_dst, 0, nw, nh, n,
                                        filter_code(filter), acc_init, flags, stream));
}
