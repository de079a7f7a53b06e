//! `DynamicImage` on the B200: the methods of `image::DynamicImage` that lie on the pixel hot path, with the
//! reference's names, receivers and argument lists (src/images/dynimage.rs:320-568, 1127-1261), dispatching over
//! its variants the way `dynamic_map!` does (dynimage.rs:111-146).  `DynamicImage` is a foreign type, so the
//! methods come as an extension trait: `use image_b200::DynamicImageB200;` and call them as
//! `img.b200_resize_exact(..)` -- or bring the un-prefixed aliases into a module of your own with the
//! `b200_dynimage_aliases!` pattern shown in INTEGRATION.md when switching a code base over wholesale.
//!
//! Variants whose subpixels the kernels take (u8 / u16 / f32: Luma, LumaA, Rgb, Rgba of each) run on the
//! GPU; any other variant a future `image` adds is converted through `to_rgba32f` first, as the reference's
//! own fallbacks do.
//!
//! NOTE: authored without a Rust toolchain (see lib.rs); tests/test_rust_sync.py checks every signature here
//! against the reference's source text.
use image::imageops::FilterType;
use image::{DynamicImage, GenericImageView, ImageBuffer, Luma, LumaA, Pixel, Rgb, Rgba};

use crate::imageops::{self, GaussianBlurParameters};
use crate::{cast, B200Subpixel};

/// `dynamic_map!(img, ref p => expr)`: the same expression on whichever buffer the variant holds, re-wrapped.
macro_rules! b200_dynamic_map {
    ($img:expr, $p:ident => $action:expr) => {
        match $img {
            DynamicImage::ImageLuma8($p) => DynamicImage::ImageLuma8($action),
            DynamicImage::ImageLumaA8($p) => DynamicImage::ImageLumaA8($action),
            DynamicImage::ImageRgb8($p) => DynamicImage::ImageRgb8($action),
            DynamicImage::ImageRgba8($p) => DynamicImage::ImageRgba8($action),
            DynamicImage::ImageLuma16($p) => DynamicImage::ImageLuma16($action),
            DynamicImage::ImageLumaA16($p) => DynamicImage::ImageLumaA16($action),
            DynamicImage::ImageRgb16($p) => DynamicImage::ImageRgb16($action),
            DynamicImage::ImageRgba16($p) => DynamicImage::ImageRgba16($action),
            DynamicImage::ImageRgb32F($p) => DynamicImage::ImageRgb32F($action),
            DynamicImage::ImageRgba32F($p) => DynamicImage::ImageRgba32F($action),
            other => {
                let $p = &other.to_rgba32f();
                DynamicImage::ImageRgba32F($action)
            }
        }
    };
}

/// `DynamicImage::to_*` (dynimage.rs:298-316 `to::<T>()` -> `CicpRgb::cast_pixels`, src/metadata/cicp.rs:706-1033).
fn to_buffer<To: Pixel>(img: &DynamicImage) -> ImageBuffer<To, Vec<To::Subpixel>>
where To::Subpixel: B200Subpixel
{
    // the luma coefficients of the image's own primaries (cicp.rs:1240-1277); None = the sRGB fallback of buffer.rs:1743
    let coefs: Option<[f32; 3]> = None;
    match img {
        DynamicImage::ImageLuma8(p) => cast::<Luma<u8>, To>(p, coefs),
        DynamicImage::ImageLumaA8(p) => cast::<LumaA<u8>, To>(p, coefs),
        DynamicImage::ImageRgb8(p) => cast::<Rgb<u8>, To>(p, coefs),
        DynamicImage::ImageRgba8(p) => cast::<Rgba<u8>, To>(p, coefs),
        DynamicImage::ImageLuma16(p) => cast::<Luma<u16>, To>(p, coefs),
        DynamicImage::ImageLumaA16(p) => cast::<LumaA<u16>, To>(p, coefs),
        DynamicImage::ImageRgb16(p) => cast::<Rgb<u16>, To>(p, coefs),
        DynamicImage::ImageRgba16(p) => cast::<Rgba<u16>, To>(p, coefs),
        DynamicImage::ImageRgb32F(p) => cast::<Rgb<f32>, To>(p, coefs),
        DynamicImage::ImageRgba32F(p) => cast::<Rgba<f32>, To>(p, coefs),
        other => cast::<Rgba<f32>, To>(&other.to_rgba32f(), coefs),
    }
}

/// The hot-path methods of `image::DynamicImage`, B200-backed.  Every `b200_x` has the signature of the reference's `x`.
pub trait DynamicImageB200 {
    // ---- conversions: dynimage.rs:320-424 (to_*) and 431-568 (into_*) ----
    fn b200_to_rgb8(&self) -> image::RgbImage;
    fn b200_to_rgb16(&self) -> ImageBuffer<Rgb<u16>, Vec<u16>>;
    fn b200_to_rgb32f(&self) -> image::Rgb32FImage;
    fn b200_to_rgba8(&self) -> image::RgbaImage;
    fn b200_to_rgba16(&self) -> ImageBuffer<Rgba<u16>, Vec<u16>>;
    fn b200_to_rgba32f(&self) -> image::Rgba32FImage;
    fn b200_to_luma8(&self) -> image::GrayImage;
    fn b200_to_luma16(&self) -> ImageBuffer<Luma<u16>, Vec<u16>>;
    fn b200_to_luma32f(&self) -> ImageBuffer<Luma<f32>, Vec<f32>>;
    fn b200_to_luma_alpha8(&self) -> image::GrayAlphaImage;
    fn b200_to_luma_alpha16(&self) -> ImageBuffer<LumaA<u16>, Vec<u16>>;
    fn b200_to_luma_alpha32f(&self) -> ImageBuffer<LumaA<f32>, Vec<f32>>;
    fn b200_into_rgb8(self) -> image::RgbImage;
    fn b200_into_rgb16(self) -> ImageBuffer<Rgb<u16>, Vec<u16>>;
    fn b200_into_rgb32f(self) -> image::Rgb32FImage;
    fn b200_into_rgba8(self) -> image::RgbaImage;
    fn b200_into_rgba16(self) -> ImageBuffer<Rgba<u16>, Vec<u16>>;
    fn b200_into_rgba32f(self) -> image::Rgba32FImage;
    fn b200_into_luma8(self) -> image::GrayImage;
    fn b200_into_luma16(self) -> ImageBuffer<Luma<u16>, Vec<u16>>;
    fn b200_into_luma32f(self) -> ImageBuffer<Luma<f32>, Vec<f32>>;
    fn b200_into_luma_alpha8(self) -> image::GrayAlphaImage;
    fn b200_into_luma_alpha16(self) -> ImageBuffer<LumaA<u16>, Vec<u16>>;
    fn b200_into_luma_alpha32f(self) -> ImageBuffer<LumaA<f32>, Vec<f32>>;
    // ---- resampling: dynimage.rs:1127, 1143, 1188 (in place, like the reference at this revision) ----
    fn b200_resize(&mut self, nwidth: u32, nheight: u32, filter: FilterType);
    fn b200_resize_exact(&mut self, nwidth: u32, nheight: u32, filter: FilterType);
    fn b200_resize_to_fill(&mut self, nwidth: u32, nheight: u32, filter: FilterType);
    // ---- thumbnails: dynimage.rs:1159, 1175 ----
    fn b200_thumbnail(&self, nwidth: u32, nheight: u32) -> DynamicImage;
    fn b200_thumbnail_exact(&self, nwidth: u32, nheight: u32) -> DynamicImage;
    // ---- blurs: dynimage.rs:1226, 1243, 1259 ----
    fn b200_blur(&self, radius: f32) -> DynamicImage;
    fn b200_blur_advanced(&self, parameters: GaussianBlurParameters) -> DynamicImage;
    fn b200_fast_blur(&self, sigma: f32) -> DynamicImage;
    // ---- the callers next to the path: dynimage.rs:1277, 1291, 1302, 1313, 1337, 1116 ----
    fn b200_unsharpen(&self, sigma: f32, threshold: i32) -> DynamicImage;
    fn b200_filter3x3(&self, kernel: &[f32; 9]) -> DynamicImage;
    fn b200_adjust_contrast(&self, c: f32) -> DynamicImage;
    fn b200_brighten(&self, value: i32) -> DynamicImage;
    fn b200_huerotate(&self, value: i32) -> DynamicImage;
    fn b200_invert(&mut self);
}

/// `resize_dimensions` (src/math/utils.rs:56-84), computed by the library with the reference's arithmetic.
fn resize_dimensions(width: u32, height: u32, nwidth: u32, nheight: u32, fill: bool) -> (u32, u32) {
    let (mut ow, mut oh) = (0u32, 0u32);
    crate::check(unsafe { crate::ffi::b200img_resize_dimensions(width, height, nwidth, nheight, fill as std::ffi::c_int, &mut ow, &mut oh) });
    (ow, oh)
}

impl DynamicImageB200 for DynamicImage {
    fn b200_to_rgb8(&self) -> image::RgbImage { to_buffer::<Rgb<u8>>(self) }
    fn b200_to_rgb16(&self) -> ImageBuffer<Rgb<u16>, Vec<u16>> { to_buffer::<Rgb<u16>>(self) }
    fn b200_to_rgb32f(&self) -> image::Rgb32FImage { to_buffer::<Rgb<f32>>(self) }
    fn b200_to_rgba8(&self) -> image::RgbaImage { to_buffer::<Rgba<u8>>(self) }
    fn b200_to_rgba16(&self) -> ImageBuffer<Rgba<u16>, Vec<u16>> { to_buffer::<Rgba<u16>>(self) }
    fn b200_to_rgba32f(&self) -> image::Rgba32FImage { to_buffer::<Rgba<f32>>(self) }
    fn b200_to_luma8(&self) -> image::GrayImage { to_buffer::<Luma<u8>>(self) }
    fn b200_to_luma16(&self) -> ImageBuffer<Luma<u16>, Vec<u16>> { to_buffer::<Luma<u16>>(self) }
    fn b200_to_luma32f(&self) -> ImageBuffer<Luma<f32>, Vec<f32>> { to_buffer::<Luma<f32>>(self) }
    fn b200_to_luma_alpha8(&self) -> image::GrayAlphaImage { to_buffer::<LumaA<u8>>(self) }
    fn b200_to_luma_alpha16(&self) -> ImageBuffer<LumaA<u16>, Vec<u16>> { to_buffer::<LumaA<u16>>(self) }
    fn b200_to_luma_alpha32f(&self) -> ImageBuffer<LumaA<f32>, Vec<f32>> { to_buffer::<LumaA<f32>>(self) }

    // into_*: the variant that already is the target moves out untouched (dynimage.rs:431-440), others convert
    fn b200_into_rgb8(self) -> image::RgbImage { match self { DynamicImage::ImageRgb8(x) => x, x => x.b200_to_rgb8() } }
    fn b200_into_rgb16(self) -> ImageBuffer<Rgb<u16>, Vec<u16>> { match self { DynamicImage::ImageRgb16(x) => x, x => x.b200_to_rgb16() } }
    fn b200_into_rgb32f(self) -> image::Rgb32FImage { match self { DynamicImage::ImageRgb32F(x) => x, x => x.b200_to_rgb32f() } }
    fn b200_into_rgba8(self) -> image::RgbaImage { match self { DynamicImage::ImageRgba8(x) => x, x => x.b200_to_rgba8() } }
    fn b200_into_rgba16(self) -> ImageBuffer<Rgba<u16>, Vec<u16>> { match self { DynamicImage::ImageRgba16(x) => x, x => x.b200_to_rgba16() } }
    fn b200_into_rgba32f(self) -> image::Rgba32FImage { match self { DynamicImage::ImageRgba32F(x) => x, x => x.b200_to_rgba32f() } }
    fn b200_into_luma8(self) -> image::GrayImage { match self { DynamicImage::ImageLuma8(x) => x, x => x.b200_to_luma8() } }
    fn b200_into_luma16(self) -> ImageBuffer<Luma<u16>, Vec<u16>> { match self { DynamicImage::ImageLuma16(x) => x, x => x.b200_to_luma16() } }
    fn b200_into_luma32f(self) -> ImageBuffer<Luma<f32>, Vec<f32>> { self.b200_to_luma32f() }
    fn b200_into_luma_alpha8(self) -> image::GrayAlphaImage { match self { DynamicImage::ImageLumaA8(x) => x, x => x.b200_to_luma_alpha8() } }
    fn b200_into_luma_alpha16(self) -> ImageBuffer<LumaA<u16>, Vec<u16>> { match self { DynamicImage::ImageLumaA16(x) => x, x => x.b200_to_luma_alpha16() } }
    fn b200_into_luma_alpha32f(self) -> ImageBuffer<LumaA<f32>, Vec<f32>> { self.b200_to_luma_alpha32f() }

    fn b200_resize(&mut self, nwidth: u32, nheight: u32, filter: FilterType) {
        if (nwidth, nheight) == self.dimensions() {
            return;
        }
        let (width2, height2) = resize_dimensions(self.width(), self.height(), nwidth, nheight, false);
        self.b200_resize_exact(width2, height2, filter)
    }

    /// The reference forwards to `pic-scale-safe` here (src/imageops/resize.rs:36-87, unvendored); this runs the
    /// in-tree resampler `imageops::resize` instead, inside the same control flow (resize.rs:27-34, 88-90): nothing
    /// to do for equal dimensions; premultiply / unpremultiply around a non-Nearest filter when alpha varies.
    fn b200_resize_exact(&mut self, nwidth: u32, nheight: u32, filter: FilterType) {
        if (nwidth, nheight) == self.dimensions() {
            return;
        }
        let resized = b200_dynamic_map!(&*self, p => imageops::resize_alpha_aware(p, nwidth, nheight, filter));
        *self = resized;
    }

    fn b200_resize_to_fill(&mut self, nwidth: u32, nheight: u32, filter: FilterType) {
        let (width2, height2) = resize_dimensions(self.width(), self.height(), nwidth, nheight, true);
        self.b200_resize_exact(width2, height2, filter);
        let (iwidth, iheight) = self.dimensions();
        let ratio = u64::from(iwidth) * u64::from(nheight);
        let nratio = u64::from(nwidth) * u64::from(iheight);
        *self = if nratio > ratio {
            self.crop_imm(0, (iheight - nheight) / 2, nwidth, nheight)
        } else {
            self.crop_imm((iwidth - nwidth) / 2, 0, nwidth, nheight)
        };
    }

    fn b200_thumbnail(&self, nwidth: u32, nheight: u32) -> DynamicImage {
        let (width2, height2) = resize_dimensions(self.width(), self.height(), nwidth, nheight, false);
        self.b200_thumbnail_exact(width2, height2)
    }

    fn b200_thumbnail_exact(&self, nwidth: u32, nheight: u32) -> DynamicImage {
        b200_dynamic_map!(self, p => imageops::thumbnail(p, nwidth, nheight))
    }

    fn b200_blur(&self, radius: f32) -> DynamicImage {
        self.b200_blur_advanced(GaussianBlurParameters::new_from_radius(radius))
    }

    /// `gaussian_blur_dyn_image` (src/imageops/sample.rs:1297-1490): Q0.15 for the u8 / u16 variants, f32 engine for f32.
    fn b200_blur_advanced(&self, parameters: GaussianBlurParameters) -> DynamicImage {
        b200_dynamic_map!(self, p => imageops::blur_dyn(p, parameters))
    }

    fn b200_fast_blur(&self, sigma: f32) -> DynamicImage {
        b200_dynamic_map!(self, p => imageops::fast_blur(p, sigma))
    }

    fn b200_unsharpen(&self, sigma: f32, threshold: i32) -> DynamicImage {
        b200_dynamic_map!(self, p => imageops::unsharpen(p, sigma, threshold))
    }

    fn b200_filter3x3(&self, kernel: &[f32; 9]) -> DynamicImage {
        b200_dynamic_map!(self, p => imageops::filter3x3(p, kernel))
    }

    fn b200_adjust_contrast(&self, c: f32) -> DynamicImage {
        b200_dynamic_map!(self, p => imageops::contrast(p, c))
    }

    fn b200_brighten(&self, value: i32) -> DynamicImage {
        b200_dynamic_map!(self, p => imageops::brighten(p, value))
    }

    fn b200_huerotate(&self, value: i32) -> DynamicImage {
        b200_dynamic_map!(self, p => imageops::huerotate(p, value))
    }

    fn b200_invert(&mut self) {
        match self {
            DynamicImage::ImageLuma8(p) => imageops::invert(p),
            DynamicImage::ImageLumaA8(p) => imageops::invert(p),
            DynamicImage::ImageRgb8(p) => imageops::invert(p),
            DynamicImage::ImageRgba8(p) => imageops::invert(p),
            DynamicImage::ImageLuma16(p) => imageops::invert(p),
            DynamicImage::ImageLumaA16(p) => imageops::invert(p),
            DynamicImage::ImageRgb16(p) => imageops::invert(p),
            DynamicImage::ImageRgba16(p) => imageops::invert(p),
            DynamicImage::ImageRgb32F(p) => imageops::invert(p),
            DynamicImage::ImageRgba32F(p) => imageops::invert(p),
            other => image::imageops::invert(other),
        }
    }
}
