// Compile-only check of the drop-in `Impl` surface: the pattern of the reference's own tests --
// include "<name>_impl.cuh", subclass the public class, reach the stages through the protected impl_ member with
// thrust::device_vector arguments (reference: test/bilateral_filter.cu:9-33, test/adaptive_bilateral_filter.cu:121-137,
// test/bilateral_texture_filter.cu:115-136) -- written against THIS library's headers: include/ and include/cuda/impl
// take the place of the reference's include/ and src/ on the include path.  Built by tests/test_abi.py with nvcc.
#include <cstdint>

#include <thrust/device_vector.h>

#include "adaptive_bilateral_filter_impl.cuh"
#include "bilateral_filter_impl.cuh"
#include "bilateral_texture_filter_impl.cuh"
#include "cuda/device_image.hpp"
#include "cuda/gradient.hpp"

class BfUnderTest : public CudaBilateralFilter {
public:
    BfUnderTest(const int width, const int height, const int ksize = 9, const float sigma_space = 10.f,
                const float sigma_color = 30.f)
        : CudaBilateralFilter(width, height, ksize, sigma_space, sigma_color) {}
    void bilateral_filter(const std::uint8_t* const d_src, std::uint8_t* const d_dst) { impl_->bilateral_filter(d_src, d_dst); }
    void joint_bilateral_filter(const std::uint8_t* const d_src, const std::uint8_t* const d_guide, std::uint8_t* const d_dst) {
        impl_->joint_bilateral_filter(d_src, d_guide, d_dst);
    }
};

class AbfUnderTest : public CudaAdaptiveBilateralFilter {
public:
    AbfUnderTest(const int width, const int height) : CudaAdaptiveBilateralFilter(width, height) {}
    void execute(const std::uint8_t* const d_src, std::uint8_t* const d_dst) { impl_->execute(d_src, d_dst); }
};

class BtfUnderTest : public CudaBilateralTextureFilter {
public:
    BtfUnderTest(const int width, const int height, const int ksize = 9, const int nitr = 3)
        : CudaBilateralTextureFilter(width, height, ksize, nitr) {}
    void compute_blur_and_rtv(const thrust::device_vector<std::uint8_t>& d_image, const thrust::device_vector<float>& d_magnitude,
                              thrust::device_vector<float>& d_blurred, thrust::device_vector<float>& d_rtv) {
        impl_->compute_blur_and_rtv(d_image, d_magnitude, d_blurred, d_rtv);
    }
    void compute_guide(const thrust::device_vector<float>& d_blurred, const thrust::device_vector<float>& d_rtv,
                       thrust::device_vector<std::uint8_t>& d_guide) {
        impl_->compute_guide(d_blurred, d_rtv, d_guide);
    }
};

extern "C" int vip_impl_surface_pattern(int width, int height) {
    const int n = width * height;
    thrust::device_vector<std::uint8_t> d_image(3 * n), d_out(3 * n), d_guide(3 * n);
    thrust::device_vector<float> d_magnitude(n), d_blurred(3 * n), d_rtv(n);
    BfUnderTest bf(width, height);
    bf.bilateral_filter(thrust::raw_pointer_cast(d_image.data()), thrust::raw_pointer_cast(d_out.data()));
    bf.joint_bilateral_filter(thrust::raw_pointer_cast(d_image.data()), thrust::raw_pointer_cast(d_guide.data()),
                              thrust::raw_pointer_cast(d_out.data()));
    AbfUnderTest abf(width, height);
    abf.execute(thrust::raw_pointer_cast(d_image.data()), thrust::raw_pointer_cast(d_out.data()));
    BtfUnderTest btf(width, height);
    cuda_gradient(thrust::raw_pointer_cast(d_image.data()), thrust::raw_pointer_cast(d_magnitude.data()), width, height, 3);
    btf.compute_blur_and_rtv(d_image, d_magnitude, d_blurred, d_rtv);
    btf.compute_guide(d_blurred, d_rtv, d_guide);
    return cudaDeviceSynchronize() == cudaSuccess ? 0 : 1;
}
