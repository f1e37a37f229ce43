// TEST INFRASTRUCTURE ONLY (oracle/).  extern "C" entry points over the reference's own CPU
// implementation, compiled UNMODIFIED and IN PLACE from /root/reference/include/cpp/*.hpp against
// oracle/cvshim (see oracle/Makefile).  Output: oracle/_ref/libvip_ref_cpu.so (git-ignored).
// Used by tests/ and by bench.py's cpu_baseline / --impl reference legs only.
#include <cstdint>
#include <cstring>

#include <opencv2/core.hpp>

#include "cpp/adaptive_bilateral_filter.hpp"
#include "cpp/bilateral_filter.hpp"
#include "cpp/bilateral_texture_filter.hpp"
#include "cpp/border_replicated_integral_image.hpp"
#include "cpp/gradient.hpp"

namespace {
inline cv::Mat3b wrap3b(const std::uint8_t* p, int w, int h) {
    return cv::Mat3b(h, w, reinterpret_cast<cv::Vec3b*>(const_cast<std::uint8_t*>(p)));
}
inline cv::Mat3f wrap3f(const float* p, int w, int h) {
    return cv::Mat3f(h, w, reinterpret_cast<cv::Vec3f*>(const_cast<float*>(p)));
}
inline cv::Mat1f wrap1f(const float* p, int w, int h) { return cv::Mat1f(h, w, const_cast<float*>(p)); }
}  // namespace

extern "C" {

void ref_set_num_threads(int n) { cv::setNumThreads(n); }
int ref_get_num_threads() { return cv::getNumThreads(); }

// include/cpp/bilateral_filter.hpp:41-124
void ref_bilateral_filter(const std::uint8_t* src, std::uint8_t* dst, int w, int h, int ksize, float sigma_space,
                          float sigma_color) {
    const cv::Mat3b s = wrap3b(src, w, h);
    cv::Mat3b d = wrap3b(dst, w, h);
    bilateral_filter(s, d, ksize, sigma_space, sigma_color);
}

// include/cpp/bilateral_filter.hpp:126-207
void ref_joint_bilateral_filter(const std::uint8_t* src, const std::uint8_t* guide, std::uint8_t* dst, int w, int h,
                                int ksize, float sigma_space, float sigma_color) {
    const cv::Mat3b s = wrap3b(src, w, h);
    const cv::Mat3b g = wrap3b(guide, w, h);
    cv::Mat3b d = wrap3b(dst, w, h);
    joint_bilateral_filter(s, g, d, ksize, sigma_space, sigma_color);
}

// include/cpp/adaptive_bilateral_filter.hpp:13-104
void ref_adaptive_bilateral_filter(const std::uint8_t* src, std::uint8_t* dst, int w, int h, int ksize,
                                   float sigma_space, float sigma_color) {
    const cv::Mat3b s = wrap3b(src, w, h);
    cv::Mat3b d = wrap3b(dst, w, h);
    adaptive_bilateral_filter(s, d, ksize, sigma_space, sigma_color);
}

// include/cpp/gradient.hpp:89-108 ; elem: 0 = u8, 1 = f32 ; channels 1 or 3
int ref_gradient(const void* src, float* dst, int w, int h, int channels, int elem) {
    int type;
    if (elem == 0 && channels == 1) type = CV_8UC1;
    else if (elem == 0 && channels == 3) type = CV_8UC3;
    else if (elem == 1 && channels == 1) type = CV_32FC1;
    else if (elem == 1 && channels == 3) type = CV_32FC3;
    else return -1;
    const cv::Mat s(h, w, type, const_cast<void*>(src));
    cv::Mat d(h, w, CV_32FC1, dst);
    gradient(s, d);
    return 0;
}

// include/cpp/bilateral_texture_filter.hpp:17-83
void ref_compute_blur_and_rtv(const std::uint8_t* image, const float* magnitude, float* blurred, float* rtv, int w,
                              int h, int ksize) {
    const cv::Mat3b img = wrap3b(image, w, h);
    const cv::Mat1f mag = wrap1f(magnitude, w, h);
    cv::Mat3f bl = wrap3f(blurred, w, h);
    cv::Mat1f rt = wrap1f(rtv, w, h);
    internal::compute_blur_and_rtv(img, mag, bl, rt, ksize);
}

// include/cpp/bilateral_texture_filter.hpp:85-145
void ref_compute_guide(const float* blurred, const float* rtv, std::uint8_t* guide, int w, int h, int ksize) {
    const cv::Mat3f bl = wrap3f(blurred, w, h);
    const cv::Mat1f rt = wrap1f(rtv, w, h);
    cv::Mat3b g = wrap3b(guide, w, h);
    internal::compute_guide(bl, rt, g, ksize);
}

// include/cpp/bilateral_texture_filter.hpp:151-178 (JBF step routed by the ximgproc shim)
void ref_bilateral_texture_filter(const std::uint8_t* src, std::uint8_t* dst, int w, int h, int ksize, int nitr) {
    const cv::Mat3b s = wrap3b(src, w, h);
    cv::Mat3b d = wrap3b(dst, w, h);
    bilateral_texture_filter(s, d, ksize, nitr);
}

// include/cpp/border_replicated_integral_image.hpp:6-85, u8 x 3 channels.
// Evaluates get(x0,y0,x1,y1) for n rectangles; rects = n x 4 ints; out = n x 3 ints.
void ref_integral_u8c3_get(const std::uint8_t* src, int w, int h, int radius, const int* rects, int n, int* out) {
    const cv::Mat3b s = wrap3b(src, w, h);
    const BorderReplicatedIntegralImage<std::uint8_t, 3> ii(s, radius);
    for (int i = 0; i < n; i++) {
        const auto v = ii.get(rects[4 * i + 0], rects[4 * i + 1], rects[4 * i + 2], rects[4 * i + 3]);
        out[3 * i + 0] = v[0];
        out[3 * i + 1] = v[1];
        out[3 * i + 2] = v[2];
    }
}

// Same class, u8 x 1 channel, dumping the whole (h+2r+1) x (w+2r+1) table.
void ref_integral_u8c3_table(const std::uint8_t* src, int w, int h, int radius, int* table) {
    // The class keeps its buffer private; rebuild the table through get(): cell (Y,X) of the padded
    // table equals get(-r, -r, X-r-1, Y-r-1) for X,Y >= 1 and 0 on the first row / column.
    const cv::Mat3b s = wrap3b(src, w, h);
    const BorderReplicatedIntegralImage<std::uint8_t, 3> ii(s, radius);
    const int tw = w + 2 * radius + 1, th = h + 2 * radius + 1;
    std::memset(table, 0, sizeof(int) * 3 * static_cast<std::size_t>(tw) * th);
    for (int Y = 1; Y < th; Y++) {
        for (int X = 1; X < tw; X++) {
            const auto v = ii.get(-radius, -radius, X - radius - 1, Y - radius - 1);
            int* t = table + 3 * (static_cast<std::size_t>(Y) * tw + X);
            t[0] = v[0]; t[1] = v[1]; t[2] = v[2];
        }
    }
}

}  // extern "C"
