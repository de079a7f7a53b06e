//! Raw bindings: one declaration per symbol of `include/b200img.h`.
#![allow(non_camel_case_types)]
use std::ffi::{c_char, c_float, c_int, c_void};

pub const U8: c_int = 0;
pub const U16: c_int = 1;
pub const F32: c_int = 2;
pub const FP_FAST: c_int = 0;
pub const FP_STRICT: c_int = 1;
pub const FP_FUSED_MAC: c_int = 2;
/// Kernel-choice flags (tests and callers that want one rounding behaviour across shapes)
pub const RESIZE_UNFUSED: c_int = 256;
pub const RESIZE_TILED: c_int = 512;
pub const THUMBNAIL_PER_PIXEL: c_int = 1024;
pub const GAUSS_TWO_KERNELS: c_int = 2048;
pub const GAUSS_FUSED: c_int = 4096;
pub const RESIZE_GENERAL: c_int = 8192;
pub const BLUR_GENERIC: c_int = 0;
pub const BLUR_DYN: c_int = 1;
pub const CAST: c_int = 0;
pub const CONVERT: c_int = 1;

#[repr(C)]
#[derive(Clone, Copy, Debug, PartialEq)]
pub struct b200img_gauss_params {
    pub x_kernel_size: u32,
    pub x_sigma: f32,
    pub y_kernel_size: u32,
    pub y_sigma: f32,
}

pub type Stream = *mut c_void; // cudaStream_t

/// One device's share of a row-band split (`b200img_band`).
#[repr(C)]
#[derive(Clone, Copy, Debug)]
pub struct b200img_band {
    pub device: c_int,
    pub d_src: *mut c_void,
    pub src_row0: u32,
    pub src_rows: u32,
    pub own_row0: u32,
    pub own_rows: u32,
    pub d_dst: *mut c_void,
    pub dst_row0: u32,
    pub dst_rows: u32,
}

extern "C" {
    pub fn b200img_last_error() -> *const c_char;
    pub fn b200img_version() -> c_int;
    pub fn b200img_device_count(count: *mut c_int) -> c_int;
    pub fn b200img_pinned_alloc(bytes: usize, out: *mut *mut c_void) -> c_int;
    pub fn b200img_pinned_free(p: *mut c_void) -> c_int;
    pub fn b200img_release_caches() -> c_int;

    pub fn b200img_filter_support(filter: c_int) -> c_float;
    pub fn b200img_filter_kernel(filter: c_int, x: c_float) -> c_float;
    pub fn b200img_resize_max_taps(filter: c_int, src_len: u32, dst_len: u32) -> c_int;
    pub fn b200img_build_weights(filter: c_int, src_len: u32, dst_len: u32, axis: c_int, left: *mut i32,
                                 count: *mut i32, weights: *mut c_float, stride: c_int) -> c_int;
    pub fn b200img_resize_dimensions(w: u32, h: u32, nw: u32, nh: u32, fill: c_int, ow: *mut u32, oh: *mut u32) -> c_int;
    pub fn b200img_gauss_params_from_radius(radius: c_float, out: *mut b200img_gauss_params) -> c_int;
    pub fn b200img_gauss_params_from_kernel_size(ks: c_float, out: *mut b200img_gauss_params) -> c_int;
    pub fn b200img_gauss_params_from_anisotropic(xks: c_float, yks: c_float, out: *mut b200img_gauss_params) -> c_int;
    pub fn b200img_gauss_params_from_sigma(sigma: c_float, out: *mut b200img_gauss_params) -> c_int;
    pub fn b200img_gaussian_kernel_1d(width: u32, sigma: c_float, out: *mut c_float) -> c_int;
    pub fn b200img_boxes_for_gauss(sigma: c_float, out_radii: *mut u32) -> c_int;

    pub fn b200img_resize(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void,
                          nw: u32, nh: u32, filter: c_int, acc_init: c_float, flags: c_int) -> c_int;
    pub fn b200img_resize_dev(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void,
                              nw: u32, nh: u32, filter: c_int, acc_init: c_float, flags: c_int, stream: Stream) -> c_int;
    pub fn b200img_resize_batch_dev(src: *const c_void, src_stride: usize, w: u32, h: u32, channels: c_int, dtype: c_int,
                                    dst: *mut c_void, dst_stride: usize, nw: u32, nh: u32, n_images: u32, filter: c_int,
                                    acc_init: c_float, flags: c_int, stream: Stream) -> c_int;
    pub fn b200img_resize_batch(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void,
                                nw: u32, nh: u32, n_images: u32, filter: c_int, acc_init: c_float, flags: c_int) -> c_int;
    pub fn b200img_resize_band_src_rows(h: u32, nh: u32, filter: c_int, dst_row0: u32, dst_rows: u32,
                                        src_row0: *mut u32, src_rows: *mut u32) -> c_int;
    pub fn b200img_resize_band_dev(src_band: *const c_void, src_row0: u32, src_rows: u32, w: u32, h: u32, channels: c_int,
                                   dtype: c_int, dst_band: *mut c_void, dst_row0: u32, dst_rows: u32, nw: u32, nh: u32,
                                   filter: c_int, acc_init: c_float, flags: c_int, stream: Stream) -> c_int;

    pub fn b200img_gaussian_blur(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void,
                                 params: *const b200img_gauss_params, path: c_int, flags: c_int) -> c_int;
    pub fn b200img_gaussian_blur_dev(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void,
                                     params: *const b200img_gauss_params, path: c_int, flags: c_int, stream: Stream) -> c_int;
    pub fn b200img_gaussian_blur_batch(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void,
                                       n_images: u32, params: *const b200img_gauss_params, path: c_int, flags: c_int) -> c_int;
    pub fn b200img_gaussian_blur_batch_dev(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int,
                                           dst: *mut c_void, n_images: u32, params: *const b200img_gauss_params,
                                           path: c_int, flags: c_int, stream: Stream) -> c_int;

    pub fn b200img_fast_blur(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void,
                             sigma: c_float) -> c_int;
    pub fn b200img_fast_blur_dev(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void,
                                 sigma: c_float, stream: Stream) -> c_int;
    pub fn b200img_fast_blur_batch(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void,
                                   n_images: u32, sigma: c_float) -> c_int;
    pub fn b200img_fast_blur_batch_dev(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void,
                                       n_images: u32, sigma: c_float, stream: Stream) -> c_int;
    pub fn b200img_fast_blur_noalloc_dev(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int,
                                         dst: *mut c_void, sigma: c_float, stream: Stream) -> c_int;

    pub fn b200img_plan_resize_bands(h: u32, nh: u32, filter: c_int, n_bands: c_int, bands: *mut b200img_band) -> c_int;
    pub fn b200img_plan_gaussian_blur_bands(h: u32, params: *const b200img_gauss_params, n_bands: c_int,
                                            bands: *mut b200img_band) -> c_int;
    pub fn b200img_plan_fast_blur_bands(h: u32, sigma: c_float, n_bands: c_int, bands: *mut b200img_band) -> c_int;
    pub fn b200img_exchange_ghost_rows(bands: *const b200img_band, n_bands: c_int, h: u32, row_bytes: usize,
                                       streams: *const Stream) -> c_int;
    pub fn b200img_resize_bands(bands: *const b200img_band, n_bands: c_int, w: u32, h: u32, channels: c_int, dtype: c_int,
                                nw: u32, nh: u32, filter: c_int, acc_init: c_float, flags: c_int,
                                streams: *const Stream) -> c_int;
    pub fn b200img_gaussian_blur_bands(bands: *const b200img_band, n_bands: c_int, w: u32, h: u32, channels: c_int,
                                       dtype: c_int, params: *const b200img_gauss_params, path: c_int, flags: c_int,
                                       streams: *const Stream) -> c_int;
    pub fn b200img_fast_blur_bands(bands: *const b200img_band, n_bands: c_int, w: u32, h: u32, channels: c_int, dtype: c_int,
                                   sigma: c_float, streams: *const Stream) -> c_int;
    pub fn b200img_gaussian_blur_band_src_rows(h: u32, params: *const b200img_gauss_params, dst_row0: u32, dst_rows: u32,
                                               src_row0: *mut u32, src_rows: *mut u32) -> c_int;
    pub fn b200img_gaussian_blur_band_dev(src_band: *const c_void, src_row0: u32, src_rows: u32, w: u32, h: u32,
                                          channels: c_int, dtype: c_int, dst_band: *mut c_void, dst_row0: u32, dst_rows: u32,
                                          params: *const b200img_gauss_params, path: c_int, flags: c_int,
                                          stream: Stream) -> c_int;
    pub fn b200img_fast_blur_band_src_rows(h: u32, sigma: c_float, dst_row0: u32, dst_rows: u32, src_row0: *mut u32,
                                           src_rows: *mut u32) -> c_int;
    pub fn b200img_fast_blur_band_dev(src_band: *const c_void, src_row0: u32, src_rows: u32, w: u32, h: u32, channels: c_int,
                                      dtype: c_int, dst_band: *mut c_void, dst_row0: u32, dst_rows: u32, sigma: c_float,
                                      stream: Stream) -> c_int;

    pub fn b200img_thumbnail(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void,
                             nw: u32, nh: u32, flags: c_int) -> c_int;
    pub fn b200img_thumbnail_dev(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void,
                                 nw: u32, nh: u32, flags: c_int, stream: Stream) -> c_int;
    pub fn b200img_thumbnail_batch_dev(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int,
                                       dst: *mut c_void, nw: u32, nh: u32, n_images: u32, flags: c_int,
                                       stream: Stream) -> c_int;
    pub fn b200img_filter3x3(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void,
                             kernel: *const c_float) -> c_int;
    pub fn b200img_filter3x3_dev(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void,
                                 kernel: *const c_float, stream: Stream) -> c_int;
    pub fn b200img_unsharpen(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void,
                             sigma: c_float, threshold: c_int, flags: c_int) -> c_int;
    pub fn b200img_unsharpen_dev(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void,
                                 sigma: c_float, threshold: c_int, flags: c_int, stream: Stream) -> c_int;

    pub fn b200img_has_constant_alpha(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, result: *mut c_int) -> c_int;
    pub fn b200img_has_constant_alpha_dev(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, result: *mut c_int, stream: Stream) -> c_int;
    pub fn b200img_premultiply_alpha(buf: *mut c_void, w: u32, h: u32, channels: c_int, dtype: c_int) -> c_int;
    pub fn b200img_premultiply_alpha_dev(buf: *mut c_void, w: u32, h: u32, channels: c_int, dtype: c_int, stream: Stream) -> c_int;
    pub fn b200img_unpremultiply_alpha(buf: *mut c_void, w: u32, h: u32, channels: c_int, dtype: c_int) -> c_int;
    pub fn b200img_unpremultiply_alpha_dev(buf: *mut c_void, w: u32, h: u32, channels: c_int, dtype: c_int, stream: Stream) -> c_int;
    pub fn b200img_convert_resize(src: *const c_void, src_dtype: c_int, src_channels: c_int, w: u32, h: u32,
                                  dst: *mut c_void, dst_dtype: c_int, dst_channels: c_int, nw: u32, nh: u32,
                                  filter: c_int, acc_init: c_float, flags: c_int) -> c_int;
    pub fn b200img_invert(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void) -> c_int;
    pub fn b200img_invert_dev(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void, stream: Stream) -> c_int;
    pub fn b200img_brighten(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void, value: i32) -> c_int;
    pub fn b200img_brighten_dev(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void, value: i32, stream: Stream) -> c_int;
    pub fn b200img_contrast(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void, contrast: c_float) -> c_int;
    pub fn b200img_contrast_dev(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void, contrast: c_float, stream: Stream) -> c_int;
    pub fn b200img_huerotate(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void, degrees: i32) -> c_int;
    pub fn b200img_huerotate_dev(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void, degrees: i32, stream: Stream) -> c_int;
    pub fn b200img_grayscale(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void, keep_alpha: c_int) -> c_int;
    pub fn b200img_grayscale_dev(src: *const c_void, w: u32, h: u32, channels: c_int, dtype: c_int, dst: *mut c_void, keep_alpha: c_int, stream: Stream) -> c_int;

    pub fn b200img_convert(src: *const c_void, src_dtype: c_int, src_channels: c_int, dst: *mut c_void, dst_dtype: c_int,
                           dst_channels: c_int, npixels: usize, mode: c_int, coefs: *const c_float) -> c_int;
    pub fn b200img_convert_dev(src: *const c_void, src_dtype: c_int, src_channels: c_int, dst: *mut c_void,
                               dst_dtype: c_int, dst_channels: c_int, npixels: usize, mode: c_int, coefs: *const c_float,
                               stream: Stream) -> c_int;
    pub fn b200img_swizzle_02(buf: *mut c_void, dtype: c_int, channels: c_int, nsubpixels: usize) -> c_int;
    pub fn b200img_swizzle_02_dev(buf: *mut c_void, dtype: c_int, channels: c_int, nsubpixels: usize, stream: Stream) -> c_int;
}
