// host_tables.h -- the small per-call tables the reference computes on the CPU
// (filter weights, Gaussian kernels, box radii).  The product computes them on
// the host too, with the reference's f32/f64 arithmetic, and ships them to the
// device; only the O(pixels) work runs on the GPU.
#pragma once

#include <stdint.h>
#include <vector>

#include "../../include/b200img.h"

namespace b200 {

float filter_support(int filter);            // < 0: unknown filter
float filter_kernel(int filter, float x);
int resize_max_taps(int filter, uint32_t src_len, uint32_t dst_len);

// Taps of every output sample along one axis.
struct AxisTable {
    int stride = 0;     // floats per output sample in `w` (>= max_count)
    int max_count = 0;  // largest tap count actually used
    std::vector<int32_t> left, count;
    std::vector<float> w;  // dst_len x stride, zero padded
    // Shape of the table, worked out once when it is built (the fixed-window kernels ask on every call):
    // fixed_R > 0: a downscale by the integer fixed_R whose every window is [fixed_R o + fixed_L0, + max_count) cut to the
    // image, the whole windows all carrying bitwise the weights of the middle one, and those symmetric;
    // up2: a 2x upscale whose every window is [(o >> 1) - S, + max_count) cut to the image (max_count = 2 S + 1 = 3, 5 or 7), the
    // whole windows of each parity carrying bitwise the weights of the middle window of that parity.
    int fixed_R = 0, fixed_L0 = 0;
    bool up2 = false;
};
// axis 0: vertical_sample's rule; axis 1: horizontal_sample's rule.
void build_axis_table(int filter, uint32_t src_len, uint32_t dst_len, int axis, AxisTable &t);

int gauss_params_from_radius(float radius, b200img_gauss_params *out);
int gauss_params_from_kernel_size(float ks, b200img_gauss_params *out);
int gauss_params_from_anisotropic(float xks, float yks, b200img_gauss_params *out);
int gauss_params_from_sigma(float sigma, b200img_gauss_params *out);
std::vector<float> gaussian_kernel_1d(uint32_t width, float sigma);
// KernelTransformer for u8/u16 + prepare_symmetric_kernel
std::vector<uint32_t> q15_scanned_kernel(const std::vector<float> &k);
// prepare_symmetric_kernel on the f32 kernel
std::vector<float> f32_scanned_kernel(const std::vector<float> &k);
void boxes_for_gauss(float sigma, uint32_t out[3]);
void resize_dimensions(uint32_t w, uint32_t h, uint32_t nw, uint32_t nh, bool fill, uint32_t *ow, uint32_t *oh);

}  // namespace b200
