//! `image-b200`: the `image` crate's data-parallel pixel hot path on an NVIDIA B200.
//!
//! Same signatures as the reference over `ImageBuffer<P, Vec<S>>` (`imageops::resize`, `blur`, `blur_advanced`,
//! `fast_blur`, `thumbnail`, `ImageBuffer::convert`, `from_raw_bgr(a)` / `into_raw_bgr(a)`) in this file and over
//! `DynamicImage` (`to_*` / `into_*`, `resize*`, `blur*`, `fast_blur`, `thumbnail*`) in `dynimage.rs`, plus the
//! device-resident batch type `DeviceBatch<P>` in `device.rs`, all forwarding to `libb200img`'s C ABI.
//! There is no CPU fallback: without a usable GPU every call panics with the library's message.
//!
//! NOTE: written in an environment without `rustc`/`cargo`; it mirrors `image_b200/imageops.py`
//! (which is exercised by the test-suite) line for line but has not itself been compiled.
pub mod device;
pub mod dynimage;
pub mod ffi;

pub use device::DeviceBatch;
pub use dynimage::DynamicImageB200;

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

pub(crate) fn check(rc: c_int) {
    if rc == 0 { return; }
    let msg = unsafe { CStr::from_ptr(ffi::b200img_last_error()) }.to_string_lossy().into_owned();
    // -1: the reference panics on this input; -2: ImageError::Limits(DimensionError); -3/-4: no memory / no GPU
    panic!("b200img (status {rc}): {msg}");
}

pub(crate) fn filter_code(f: FilterType) -> c_int {
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

    /// `image::imageops::GaussianBlurParameters` (src/imageops/sample.rs:1081-1295): same constructors, same asserts
    /// (the library validates and the wrapper panics, as the reference does).
    #[derive(Clone, Copy, Debug, PartialEq)]
    pub struct GaussianBlurParameters(ffi::b200img_gauss_params);
    impl GaussianBlurParameters {
        pub const SMOOTHING_3: GaussianBlurParameters = GaussianBlurParameters(ffi::b200img_gauss_params { x_kernel_size: 3, x_sigma: 0.8, y_kernel_size: 3, y_sigma: 0.8 });
        pub const SMOOTHING_5: GaussianBlurParameters = GaussianBlurParameters(ffi::b200img_gauss_params { x_kernel_size: 5, x_sigma: 1.1, y_kernel_size: 5, y_sigma: 1.1 });
        pub const SMOOTHING_7: GaussianBlurParameters = GaussianBlurParameters(ffi::b200img_gauss_params { x_kernel_size: 7, x_sigma: 1.4, y_kernel_size: 7, y_sigma: 1.4 });
        fn make(f: impl FnOnce(*mut ffi::b200img_gauss_params) -> c_int) -> GaussianBlurParameters {
            let mut p = ffi::b200img_gauss_params { x_kernel_size: 0, x_sigma: 0.0, y_kernel_size: 0, y_sigma: 0.0 };
            check(f(&mut p));
            GaussianBlurParameters(p)
        }
        pub fn new_from_radius(radius: f32) -> GaussianBlurParameters { Self::make(|p| unsafe { ffi::b200img_gauss_params_from_radius(radius, p) }) }
        pub fn new_from_kernel_size(kernel_size: f32) -> GaussianBlurParameters { Self::make(|p| unsafe { ffi::b200img_gauss_params_from_kernel_size(kernel_size, p) }) }
        pub fn new_anisotropic_kernel_size(x_axis_kernel_size: f32, y_axis_kernel_size: f32) -> GaussianBlurParameters {
            Self::make(|p| unsafe { ffi::b200img_gauss_params_from_anisotropic(x_axis_kernel_size, y_axis_kernel_size, p) })
        }
        pub fn new_from_sigma(sigma: f32) -> GaussianBlurParameters { Self::make(|p| unsafe { ffi::b200img_gauss_params_from_sigma(sigma, p) }) }
        pub(crate) fn raw(&self) -> ffi::b200img_gauss_params { self.0 }
    }

    /// `resize_impl`'s control flow around the resampler (src/imageops/resize.rs:27-34, 88-90, 105-218): images with
    /// a varying alpha channel are premultiplied before and unpremultiplied after any filter but Nearest.
    pub fn resize_alpha_aware<P: Pixel>(image: &ImageBuffer<P, Vec<P::Subpixel>>, nwidth: u32, nheight: u32, filter: FilterType)
        -> ImageBuffer<P, Vec<P::Subpixel>>
    where P::Subpixel: B200Subpixel
    {
        let has_alpha = P::CHANNEL_COUNT == 2 || P::CHANNEL_COUNT == 4;
        if !has_alpha || matches!(filter, FilterType::Nearest) || has_constant_alpha(image) {
            return resize(image, nwidth, nheight, filter);
        }
        let (w, h) = image.dimensions();
        let mut pre = image.clone();
        check(unsafe { ffi::b200img_premultiply_alpha(pre.as_mut_ptr() as *mut c_void, w, h, P::CHANNEL_COUNT as c_int, <P::Subpixel as B200Subpixel>::DTYPE) });
        let mut out = resize(&pre, nwidth, nheight, filter);
        check(unsafe { ffi::b200img_unpremultiply_alpha(out.as_mut_ptr() as *mut c_void, nwidth, nheight, P::CHANNEL_COUNT as c_int, <P::Subpixel as B200Subpixel>::DTYPE) });
        out
    }

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


    fn gauss<I: GenericImageView>(image: &I, p: GaussianBlurParameters, path: c_int)
        -> ImageBuffer<I::Pixel, Vec<<I::Pixel as Pixel>::Subpixel>>
    where <I::Pixel as Pixel>::Subpixel: B200Subpixel
    {
        let src = packed(image);
        let mut out = ImageBuffer::new(image.width(), image.height());
        check(unsafe { ffi::b200img_gaussian_blur(src.as_ptr() as *const c_void, image.width(), image.height(),
            <I::Pixel as Pixel>::CHANNEL_COUNT as c_int, <<I::Pixel as Pixel>::Subpixel as B200Subpixel>::DTYPE,
            out.as_mut_ptr() as *mut c_void, &p.raw(), path, ffi::FP_FAST) });
        out
    }

    /// `image::imageops::blur` -- `radius`, not sigma, at this revision (sample.rs:1034).
    pub fn blur<I: GenericImageView>(image: &I, radius: f32) -> ImageBuffer<I::Pixel, Vec<<I::Pixel as Pixel>::Subpixel>>
    where <I::Pixel as Pixel>::Subpixel: B200Subpixel { gauss(image, GaussianBlurParameters::new_from_radius(radius), ffi::BLUR_GENERIC) }

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

/// `ImageBuffer::<Rgb<S>, _>::from_raw_bgr` (src/images/buffer.rs:1551): a BGR buffer becomes an RGB image.
pub fn from_raw_bgr<S: B200Subpixel>(width: u32, height: u32, mut container: Vec<S>) -> Option<ImageBuffer<image::Rgb<S>, Vec<S>>>
where image::Rgb<S>: Pixel<Subpixel = S>
{
    if (container.len() as u64) < width as u64 * height as u64 * 3 { return None; }
    swizzle_02(&mut container, 3);
    ImageBuffer::from_raw(width, height, container)
}
/// `into_raw_bgr` (buffer.rs:1558).
pub fn into_raw_bgr<S: B200Subpixel>(image: ImageBuffer<image::Rgb<S>, Vec<S>>) -> Vec<S>
where image::Rgb<S>: Pixel<Subpixel = S>
{
    let mut raw = image.into_raw();
    swizzle_02(&mut raw, 3);
    raw
}
/// `ImageBuffer::<Rgba<S>, _>::from_raw_bgra` (buffer.rs:1571).
pub fn from_raw_bgra<S: B200Subpixel>(width: u32, height: u32, mut container: Vec<S>) -> Option<ImageBuffer<image::Rgba<S>, Vec<S>>>
where image::Rgba<S>: Pixel<Subpixel = S>
{
    if (container.len() as u64) < width as u64 * height as u64 * 4 { return None; }
    swizzle_02(&mut container, 4);
    ImageBuffer::from_raw(width, height, container)
}
/// `into_raw_bgra` (buffer.rs:1578).
pub fn into_raw_bgra<S: B200Subpixel>(image: ImageBuffer<image::Rgba<S>, Vec<S>>) -> Vec<S>
where image::Rgba<S>: Pixel<Subpixel = S>
{
    let mut raw = image.into_raw();
    swizzle_02(&mut raw, 4);
    raw
}

/// Device-resident batch on raw pointers (see `DeviceBatch` for the owning type): `n` equally shaped images already in device memory.
/// # Safety
/// `d_src` / `d_dst` must be valid device pointers on the current CUDA device; `stream` a valid `cudaStream_t` or null.
pub unsafe fn resize_batch_dev(d_src: *const c_void, w: u32, h: u32, channels: u32, dtype: c_int, d_dst: *mut c_void,
                               nw: u32, nh: u32, n: u32, filter: FilterType, acc_init: f32, flags: c_int, stream: ffi::Stream) {
    check(ffi::b200img_resize_batch_dev(d_src, 0, w, h, channels as c_int, dtype, d_dst, 0, nw, nh, n,
                                        filter_code(filter), acc_init, flags, stream));
}
