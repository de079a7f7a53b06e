// Parity tests written the way the reference's own tests read (src/imageops/sample.rs:1812-1919,
// src/imageops/mod.rs:473-555, src/metadata/cicp.rs:1328-1390), against the C++ mirror.
#include <cstdio>
#include <cstdlib>

#include "image_b200.hpp"

using namespace image_b200;
#define EXPECT(c) do { if (!(c)) { std::printf("FAILED %s:%d: %s\n", __FILE__, __LINE__, #c); std::exit(1); } } while (0)

int main() {
    {   // test_issue_186
        ImageBuffer<uint8_t, 3> img(100, 100);
        auto out = imageops::resize(img, 50, 50, FilterType::Lanczos3);
        EXPECT(out.width == 50 && out.height == 50);
        for (auto v : out.data) EXPECT(v == 1);  // literal reference: accumulator starts at 1.0
        out = imageops::resize(img, 50, 50, FilterType::Lanczos3, 0.0f);
        for (auto v : out.data) EXPECT(v == 0);
    }
    {   // bug_1600
        ImageBuffer<uint8_t, 4> img(629, 627);
        for (auto &v : img.data) v = 255;
        auto out = imageops::resize(img, 22, 22, FilterType::Lanczos3);
        bool any = false;
        for (auto v : out.data) any |= v != 0;
        EXPECT(any);
    }
    {   // test_resize_same_size
        ImageBuffer<uint8_t, 3> img(17, 13);
        for (size_t i = 0; i < img.data.size(); ++i) img.data[i] = (uint8_t)(i * 7);
        auto out = imageops::resize(img, 17, 13, FilterType::CatmullRom);
        EXPECT(out.data == img.data);
    }
    {   // transform_pixels_srgb (cicp.rs:1344-1355)
        ImageBuffer<uint8_t, 4> px(1, 1);
        px.data = {255, 0, 0, 255};
        EXPECT((px.cast<uint8_t, 3>().data == std::vector<uint8_t>{255, 0, 0}));
        EXPECT((px.cast<uint8_t, 1>().data == std::vector<uint8_t>{54}));
        EXPECT((px.cast<uint8_t, 2>().data == std::vector<uint8_t>{54, 255}));
    }
    {   // tests/conversions.rs: u8 max -> u16 max; bgra swizzle
        ImageBuffer<uint8_t, 4> img(2, 2);
        for (auto &v : img.data) v = 255;
        for (auto v : img.convert<uint16_t, 4>().data) EXPECT(v == 65535);
        ImageBuffer<uint8_t, 4> b(1, 1);
        b.data = {1, 2, 3, 4};
        b.swizzle_02();
        EXPECT((b.data == std::vector<uint8_t>{3, 2, 1, 4}));
    }
    {   // test_fast_blur_* / constant image stays constant; blur(img, 0.0) does not panic
        ImageBuffer<uint8_t, 3> img(60, 40);
        for (auto &v : img.data) v = 143;
        for (auto v : imageops::fast_blur(img, 7.0f).data) EXPECT(v == 143);
        for (auto v : imageops::blur(img, 3.0f).data) EXPECT(v == 143);
        for (auto v : imageops::blur_dyn(img, GaussianBlurParameters::new_from_sigma(5.0f)).data) EXPECT(v == 143);
        (void)imageops::blur(img, 0.0f);
        EXPECT(imageops::fast_blur(ImageBuffer<uint8_t, 3>(0, 0), 3.0f).data.empty());
    }
    {   // the widened operators: thumbnail_huge_block (sample.rs:1922), constant-image invariants
        ImageBuffer<uint16_t, 1> huge(1024, 1024);
        for (auto &v : huge.data) v = 65535;
        EXPECT(imageops::thumbnail(huge, 1, 1).data[0] == 65535);
        ImageBuffer<uint8_t, 3> img(60, 40);
        for (auto &v : img.data) v = 90;
        for (auto v : imageops::thumbnail(img, 17, 9).data) EXPECT(v == 90);
        for (auto v : imageops::thumbnail(img, 100, 70).data) EXPECT(v == 90);
        for (auto v : imageops::unsharpen(img, 2.0f, 0).data) EXPECT(v == 90);
        const float box[9] = {1, 1, 1, 1, 1, 1, 1, 1, 1};
        auto f = imageops::filter3x3(img, box);
        EXPECT(f.data[(1 * 60 + 1) * 3] == 90 && f.data[0] == 0);   // interior averaged, border zero
        bool threw = false;   // the reference divides by a zero count here
        try { (void)imageops::thumbnail(ImageBuffer<uint8_t, 1>(100, 1000), 1, 1); } catch (const Panic &) { threw = true; }
        EXPECT(threw);
    }
    {   // panics of the reference surface as exceptions
        bool threw = false;
        try { GaussianBlurParameters::new_from_sigma(-1.0f); } catch (const Panic &) { threw = true; }
        EXPECT(threw);
    }
    std::printf("cpp mirror ok\n");
    return 0;
}
