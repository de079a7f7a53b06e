//! The device-resident batch variant: `DeviceBatch<P>` owns `n` equally shaped images in the memory of one
//! B200 and runs the hot-path operators on them without leaving the device (one launch per operator for the
//! whole batch; nothing crosses PCIe between `from_host` and `to_host`).
//!
//! Device memory comes from the CUDA runtime the application already links (`cudaMalloc` / `cudaFree` /
//! `cudaMemcpy`, declared below); every operator goes through `libb200img`'s `*_dev` entry points on `stream`
//! (NULL = the legacy default stream).  `Drop` frees the buffer.
use std::ffi::{c_int, c_void};
use std::marker::PhantomData;

use image::imageops::FilterType;
use image::{ImageBuffer, Pixel};

use crate::imageops::GaussianBlurParameters;
use crate::{check, ffi, filter_code, B200Subpixel, ACC_INIT_REFERENCE};

extern "C" {
    fn cudaMalloc(ptr: *mut *mut c_void, bytes: usize) -> c_int;
    fn cudaFree(ptr: *mut c_void) -> c_int;
    fn cudaMemcpy(dst: *mut c_void, src: *const c_void, bytes: usize, kind: c_int) -> c_int;
    fn cudaStreamSynchronize(stream: *mut c_void) -> c_int;
}
const H2D: c_int = 1;
const D2H: c_int = 2;

fn cuda(rc: c_int, what: &str) {
    if rc != 0 {
        panic!("CUDA runtime error {rc} in {what}");
    }
}

/// `n` images of `width x height` pixels of type `P`, tightly packed, resident on the current CUDA device.
pub struct DeviceBatch<P: Pixel>
where P::Subpixel: B200Subpixel
{
    ptr: *mut c_void,
    n: u32,
    width: u32,
    height: u32,
    stream: ffi::Stream,
    _p: PhantomData<P>,
}

impl<P: Pixel> DeviceBatch<P>
where P::Subpixel: B200Subpixel
{
    fn image_bytes(width: u32, height: u32) -> usize {
        width as usize * height as usize * P::CHANNEL_COUNT as usize * std::mem::size_of::<P::Subpixel>()
    }

    /// Uninitialised batch (the destination of an operator).
    pub fn alloc(n: u32, width: u32, height: u32, stream: ffi::Stream) -> Self {
        let mut ptr = std::ptr::null_mut();
        cuda(unsafe { cudaMalloc(&mut ptr, (Self::image_bytes(width, height) * n as usize).max(1)) }, "cudaMalloc");
        DeviceBatch { ptr, n, width, height, stream, _p: PhantomData }
    }

    /// Uploads equally shaped host images (panics, like the reference's asserts, if the shapes differ).
    pub fn from_host(images: &[ImageBuffer<P, Vec<P::Subpixel>>], stream: ffi::Stream) -> Self {
        assert!(!images.is_empty(), "DeviceBatch::from_host: no images");
        let (width, height) = images[0].dimensions();
        let batch = Self::alloc(images.len() as u32, width, height, stream);
        let bytes = Self::image_bytes(width, height);
        for (i, img) in images.iter().enumerate() {
            assert_eq!(img.dimensions(), (width, height), "DeviceBatch::from_host: images differ in shape");
            cuda(unsafe { cudaMemcpy((batch.ptr as *mut u8).add(i * bytes) as *mut c_void, img.as_ptr() as *const c_void, bytes, H2D) }, "cudaMemcpy (upload)");
        }
        batch
    }

    /// Downloads the batch (waits for the work queued on its stream first).
    pub fn to_host(&self) -> Vec<ImageBuffer<P, Vec<P::Subpixel>>> {
        cuda(unsafe { cudaStreamSynchronize(self.stream) }, "cudaStreamSynchronize");
        let bytes = Self::image_bytes(self.width, self.height);
        (0..self.n as usize).map(|i| {
            let mut img = ImageBuffer::new(self.width, self.height);
            cuda(unsafe { cudaMemcpy(img.as_mut_ptr() as *mut c_void, (self.ptr as *const u8).add(i * bytes) as *const c_void, bytes, D2H) }, "cudaMemcpy (download)");
            img
        }).collect()
    }

    pub fn len(&self) -> u32 { self.n }
    pub fn is_empty(&self) -> bool { self.n == 0 }
    pub fn dimensions(&self) -> (u32, u32) { (self.width, self.height) }
    /// The raw device pointer (for interop with other CUDA code on the same stream).
    pub fn as_device_ptr(&self) -> *mut c_void { self.ptr }

    fn dtype() -> c_int { <P::Subpixel as B200Subpixel>::DTYPE }
    fn channels() -> c_int { P::CHANNEL_COUNT as c_int }

    /// `imageops::resize` on every image of the batch (one launch).
    pub fn resize(&self, nwidth: u32, nheight: u32, filter: FilterType) -> DeviceBatch<P> {
        let out = Self::alloc(self.n, nwidth, nheight, self.stream);
        check(unsafe { ffi::b200img_resize_batch_dev(self.ptr, 0, self.width, self.height, Self::channels(), Self::dtype(), out.ptr, 0,
            nwidth, nheight, self.n, filter_code(filter), ACC_INIT_REFERENCE, ffi::FP_FAST, self.stream) });
        out
    }

    /// `imageops::blur_advanced` (f32 engine) on every image.
    pub fn blur_advanced(&self, parameters: GaussianBlurParameters) -> DeviceBatch<P> {
        let out = Self::alloc(self.n, self.width, self.height, self.stream);
        check(unsafe { ffi::b200img_gaussian_blur_batch_dev(self.ptr, self.width, self.height, Self::channels(), Self::dtype(), out.ptr,
            self.n, &parameters.raw(), ffi::BLUR_GENERIC, ffi::FP_FAST, self.stream) });
        out
    }

    /// `imageops::blur` (radius) on every image.
    pub fn blur(&self, radius: f32) -> DeviceBatch<P> { self.blur_advanced(GaussianBlurParameters::new_from_radius(radius)) }

    /// `imageops::fast_blur` on every image.
    pub fn fast_blur(&self, sigma: f32) -> DeviceBatch<P> {
        let out = Self::alloc(self.n, self.width, self.height, self.stream);
        check(unsafe { ffi::b200img_fast_blur_batch_dev(self.ptr, self.width, self.height, Self::channels(), Self::dtype(), out.ptr,
            self.n, sigma, self.stream) });
        out
    }

    /// `imageops::thumbnail` on every image.
    pub fn thumbnail(&self, nwidth: u32, nheight: u32) -> DeviceBatch<P> {
        let out = Self::alloc(self.n, nwidth, nheight, self.stream);
        check(unsafe { ffi::b200img_thumbnail_batch_dev(self.ptr, self.width, self.height, Self::channels(), Self::dtype(), out.ptr,
            nwidth, nheight, self.n, 0, self.stream) });
        out
    }

    /// `ImageBuffer::convert::<To>()` on every image (`cast = true`: the arithmetic of `DynamicImage::to_*`).
    pub fn convert<To: Pixel>(&self, cast: bool) -> DeviceBatch<To>
    where To::Subpixel: B200Subpixel
    {
        let out = DeviceBatch::<To>::alloc(self.n, self.width, self.height, self.stream);
        check(unsafe { ffi::b200img_convert_dev(self.ptr, Self::dtype(), Self::channels(), out.ptr, <To::Subpixel as B200Subpixel>::DTYPE,
            To::CHANNEL_COUNT as c_int, self.n as usize * self.width as usize * self.height as usize,
            if cast { ffi::CAST } else { ffi::CONVERT }, std::ptr::null(), self.stream) });
        out
    }

    /// BGR(A) <-> RGB(A) in place (`from_raw_bgra` / `into_raw_bgra`, src/images/buffer.rs:1564-1582).
    pub fn swizzle_02(&mut self) {
        let nsub = self.n as usize * self.width as usize * self.height as usize * P::CHANNEL_COUNT as usize;
        check(unsafe { ffi::b200img_swizzle_02_dev(self.ptr, Self::dtype(), Self::channels(), nsub, self.stream) });
    }
}

impl<P: Pixel> Drop for DeviceBatch<P>
where P::Subpixel: B200Subpixel
{
    fn drop(&mut self) {
        unsafe {
            cudaStreamSynchronize(self.stream);  // queued kernels may still use the buffer
            cudaFree(self.ptr);
        }
    }
}

// A batch is tied to the CUDA context of the thread that made it; moving it between threads is fine
// (the pointer is process-wide), sharing it without synchronisation is not.
unsafe impl<P: Pixel> Send for DeviceBatch<P> where P::Subpixel: B200Subpixel {}
