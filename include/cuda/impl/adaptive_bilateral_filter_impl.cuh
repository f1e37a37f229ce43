// CudaAdaptiveBilateralFilter::Impl with the member names of the reference's src/adaptive_bilateral_filter_impl.cuh:9-30
// (its tests call impl_->execute(...), test/adaptive_bilateral_filter.cu:121-137).  Header-only on top of the C ABI.
#ifndef ADAPTIVE_BILATERAL_FILTER_IMPL_CUH
#define ADAPTIVE_BILATERAL_FILTER_IMPL_CUH

#include <cstdint>

#include "../../vip_b200.hpp"
#include "vip_impl_common.hpp"

class CudaAdaptiveBilateralFilter::Impl {
public:
    Impl(const int width, const int height, const int ksize = 9, const float sigma_space = 10.f,
         const float sigma_color = 30.f) {
        VIP_CHECK(vip_abf_create(&handle, width, height, ksize, sigma_space, sigma_color));
    }
    ~Impl() { vip_abf_destroy(handle); }
    Impl(const Impl&) = delete;
    Impl& operator=(const Impl&) = delete;

    void execute(const std::uint8_t* const d_src, std::uint8_t* const d_dst) const {
        VIP_CHECK(vip_abf_run(handle, d_src, d_dst, nullptr));
    }
    void execute(const std::uint8_t* d_src, std::uint8_t* d_dst, void* stream) const {
        VIP_CHECK(vip_abf_run(handle, d_src, d_dst, stream));
    }

    vip_abf* handle = nullptr;
};

#endif  // ADAPTIVE_BILATERAL_FILTER_IMPL_CUH
