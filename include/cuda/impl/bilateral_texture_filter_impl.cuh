// CudaBilateralTextureFilter::Impl with the member names of the reference's src/bilateral_texture_filter_impl.cuh:10-46:
// execute() and the two stage entry points its tests reach through impl_ (test/bilateral_texture_filter.cu:115-136) --
// compute_blur_and_rtv / compute_guide on thrust::device_vector (declared when compiled by nvcc) and on raw device
// pointers.  Header-only on top of the C ABI; the calls launch and return, like the reference's Impl.
#ifndef BILATERAL_TEXTURE_FILTER_IMPL_CUH
#define BILATERAL_TEXTURE_FILTER_IMPL_CUH

#include <cstdint>

#include "../../vip_b200.hpp"
#include "vip_impl_common.hpp"

class CudaBilateralTextureFilter::Impl {
public:
    Impl(const int width, const int height, const int ksize = 9, const int nitr = 3) {
        VIP_CHECK(vip_btf_create(&handle, width, height, ksize, nitr));
    }
    ~Impl() { vip_btf_destroy(handle); }
    Impl(const Impl&) = delete;
    Impl& operator=(const Impl&) = delete;

    void execute(const std::uint8_t* const d_src, std::uint8_t* const d_dst) { VIP_CHECK(vip_btf_run(handle, d_src, d_dst, nullptr)); }
    void execute(const std::uint8_t* d_src, std::uint8_t* d_dst, void* stream) { VIP_CHECK(vip_btf_run(handle, d_src, d_dst, stream)); }

    void compute_blur_and_rtv(const std::uint8_t* d_image, const float* d_magnitude, float* d_blurred, float* d_rtv,
                              void* stream = nullptr) {
        VIP_CHECK(vip_btf_blur_rtv(handle, d_image, d_magnitude, d_blurred, d_rtv, stream));
    }
    void compute_guide(const float* d_blurred, const float* d_rtv, std::uint8_t* d_guide, void* stream = nullptr) {
        VIP_CHECK(vip_btf_guide(handle, d_blurred, d_rtv, d_guide, stream));
    }
#ifdef VIP_HAVE_THRUST
    void compute_blur_and_rtv(const thrust::device_vector<std::uint8_t>& d_image,
                              const thrust::device_vector<float>& d_magnitude, thrust::device_vector<float>& d_blurred,
                              thrust::device_vector<float>& d_rtv) {
        compute_blur_and_rtv(thrust::raw_pointer_cast(d_image.data()), thrust::raw_pointer_cast(d_magnitude.data()),
                             thrust::raw_pointer_cast(d_blurred.data()), thrust::raw_pointer_cast(d_rtv.data()));
    }
    void compute_guide(const thrust::device_vector<float>& d_blurred, const thrust::device_vector<float>& d_rtv,
                       thrust::device_vector<std::uint8_t>& d_guide) {
        compute_guide(thrust::raw_pointer_cast(d_blurred.data()), thrust::raw_pointer_cast(d_rtv.data()),
                      thrust::raw_pointer_cast(d_guide.data()));
    }
#endif

    vip_btf* handle = nullptr;
};

#endif  // BILATERAL_TEXTURE_FILTER_IMPL_CUH
