// TEST INFRASTRUCTURE ONLY (oracle/).  extern "C" entry points over the reference's own CUDA
// implementation: /root/reference/src/*.cu are compiled UNMODIFIED and IN PLACE for sm_100a together
// with this wrapper (see oracle/Makefile) into oracle/_ref/libvip_ref_cuda.so (git-ignored, travels to
// the GPU box).  It is (i) the bit-exactness oracle for the new kernels on a B200 and (ii) "the kernel
// to beat" in bench.py's optional --with-ref-cuda comparison.  Never linked into the product library.
//
// All pointers are DEVICE pointers, images are packed interleaved (pitch = width*channels).
#include <cstdint>
#include <cuda_runtime.h>
#include <thrust/copy.h>
#include <thrust/device_ptr.h>
#include <thrust/device_vector.h>

#include "cuda/adaptive_bilateral_filter.hpp"
#include "cuda/bilateral_filter.hpp"
#include "cuda/bilateral_texture_filter.hpp"
#include <type_traits>
#include "cuda/gradient.hpp"

#include "bilateral_texture_filter_impl.cuh"

namespace {
// Same access route as /root/reference/test/bilateral_texture_filter.cu:115-136 (protected impl_).
class BtfStages : public CudaBilateralTextureFilter {
public:
    BtfStages(int w, int h, int ksize, int nitr) : CudaBilateralTextureFilter(w, h, ksize, nitr) {}
    void blur_rtv(const thrust::device_vector<std::uint8_t>& img, const thrust::device_vector<float>& mag,
                  thrust::device_vector<float>& blurred, thrust::device_vector<float>& rtv) {
        impl_->compute_blur_and_rtv(img, mag, blurred, rtv);
    }
    void guide(const thrust::device_vector<float>& blurred, const thrust::device_vector<float>& rtv,
               thrust::device_vector<std::uint8_t>& g) {
        impl_->compute_guide(blurred, rtv, g);
    }
};
template <typename T>
thrust::device_vector<T> to_vec(const T* d, std::size_t n) {
    thrust::device_vector<T> v(n);
    cudaMemcpy(v.data().get(), d, n * sizeof(T), cudaMemcpyDeviceToDevice);
    return v;
}
template <typename T>
void from_vec(const thrust::device_vector<T>& v, T* d) {
    cudaMemcpy(d, v.data().get(), v.size() * sizeof(T), cudaMemcpyDeviceToDevice);
}
}  // namespace

extern "C" {

void* refcu_bf_create(int w, int h, int ksize, float sigma_space, float sigma_color) {
    return new CudaBilateralFilter(w, h, ksize, sigma_space, sigma_color);
}
void refcu_bf_destroy(void* p) { delete static_cast<CudaBilateralFilter*>(p); }
void refcu_bf_run(void* p, const std::uint8_t* d_src, std::uint8_t* d_dst) {
    static_cast<CudaBilateralFilter*>(p)->bilateral_filter(d_src, d_dst);
}
void refcu_jbf_run(void* p, const std::uint8_t* d_src, const std::uint8_t* d_guide, std::uint8_t* d_dst) {
    static_cast<CudaBilateralFilter*>(p)->joint_bilateral_filter(d_src, d_guide, d_dst);
}

void* refcu_abf_create(int w, int h, int ksize, float sigma_space, float sigma_color) {
    return new CudaAdaptiveBilateralFilter(w, h, ksize, sigma_space, sigma_color);
}
void refcu_abf_destroy(void* p) { delete static_cast<CudaAdaptiveBilateralFilter*>(p); }
void refcu_abf_run(void* p, const std::uint8_t* d_src, std::uint8_t* d_dst) {
    static_cast<CudaAdaptiveBilateralFilter*>(p)->execute(d_src, d_dst);
}

void* refcu_btf_create(int w, int h, int ksize, int nitr) { return new BtfStages(w, h, ksize, nitr); }
void refcu_btf_destroy(void* p) { delete static_cast<BtfStages*>(p); }
void refcu_btf_run(void* p, const std::uint8_t* d_src, std::uint8_t* d_dst) {
    static_cast<BtfStages*>(p)->execute(d_src, d_dst);
}
void refcu_btf_blur_rtv(void* p, const std::uint8_t* d_img, const float* d_mag, float* d_blurred, float* d_rtv, int w,
                        int h) {
    const std::size_t n = static_cast<std::size_t>(w) * h;
    auto img = to_vec(d_img, n * 3);
    auto mag = to_vec(d_mag, n);
    thrust::device_vector<float> bl(n * 3), rt(n);
    static_cast<BtfStages*>(p)->blur_rtv(img, mag, bl, rt);
    cudaDeviceSynchronize();
    from_vec(bl, d_blurred);
    from_vec(rt, d_rtv);
}
void refcu_btf_guide(void* p, const float* d_blurred, const float* d_rtv, std::uint8_t* d_guide, int w, int h) {
    const std::size_t n = static_cast<std::size_t>(w) * h;
    auto bl = to_vec(d_blurred, n * 3);
    auto rt = to_vec(d_rtv, n);
    thrust::device_vector<std::uint8_t> g(n * 3);
    static_cast<BtfStages*>(p)->guide(bl, rt, g);
    cudaDeviceSynchronize();
    from_vec(g, d_guide);
}

void refcu_gradient_u8(const std::uint8_t* d_src, float* d_dst, int w, int h, int ch) {
    cuda_gradient<std::uint8_t>(d_src, d_dst, w, h, ch);
    cudaDeviceSynchronize();
}
void refcu_gradient_f32(const float* d_src, float* d_dst, int w, int h, int ch) {
    cuda_gradient<float>(d_src, d_dst, w, h, ch);
    cudaDeviceSynchronize();
}

}  // extern "C"
