// CudaBilateralFilter::Impl with the member names of the reference's src/bilateral_filter_impl.cuh:9-35, so that code
// written against it -- the reference's own tests subclass the public class and call impl_->bilateral_filter(...) /
// impl_->joint_bilateral_filter(...) after including "bilateral_filter_impl.cuh" (test/bilateral_filter.cu:6-33) --
// compiles against this library by putting include/cuda/impl on the include path instead of the reference's src/.
// Header-only on top of the C ABI; the calls launch and return (no synchronisation), like the reference's Impl.
#ifndef BILATERAL_FILTER_IMPL_CUH
#define BILATERAL_FILTER_IMPL_CUH

#include <cstdint>

#include "../../vip_b200.hpp"
#include "vip_impl_common.hpp"

class CudaBilateralFilter::Impl {
public:
    Impl(const int width, const int height, const int ksize = 9, const float sigma_space = 10.f,
         const float sigma_color = 30.f) {
        VIP_CHECK(vip_bf_create(&handle, width, height, ksize, sigma_space, sigma_color));
    }
    ~Impl() { vip_bf_destroy(handle); }
    Impl(const Impl&) = delete;
    Impl& operator=(const Impl&) = delete;

    void bilateral_filter(const std::uint8_t* const d_src, std::uint8_t* const d_dst) const {
        VIP_CHECK(vip_bf_run(handle, d_src, d_dst, nullptr));
    }
    void joint_bilateral_filter(const std::uint8_t* const d_src, const std::uint8_t* const d_guide,
                                std::uint8_t* const d_dst) const {
        VIP_CHECK(vip_jbf_run(handle, d_src, d_guide, d_dst, nullptr));
    }
    // additions: explicit stream
    void bilateral_filter(const std::uint8_t* d_src, std::uint8_t* d_dst, void* stream) const {
        VIP_CHECK(vip_bf_run(handle, d_src, d_dst, stream));
    }
    void joint_bilateral_filter(const std::uint8_t* d_src, const std::uint8_t* d_guide, std::uint8_t* d_dst,
                                void* stream) const {
        VIP_CHECK(vip_jbf_run(handle, d_src, d_guide, d_dst, stream));
    }

    vip_bf* handle = nullptr;
};

#endif  // BILATERAL_FILTER_IMPL_CUH
