// C++ drop-in classes (include/vip_b200.hpp) implemented on the C ABI (include/vip_c_api.h).
// Plain C++: no CUDA headers.  Error convention of the reference (src/host_utilities.hpp:7-13): report on
// stderr with file and line, never throw, carry on.
#include <cstdio>
#include <utility>

#include "../../include/vip_b200.hpp"
// the Impl classes are public headers (include/cuda/impl/*.cuh) with the reference's member names, so that code written
// against the reference's src/*_impl.cuh -- its own tests -- compiles against this library too
#include "../../include/cuda/impl/adaptive_bilateral_filter_impl.cuh"
#include "../../include/cuda/impl/bilateral_filter_impl.cuh"
#include "../../include/cuda/impl/bilateral_texture_filter_impl.cuh"

// ---------------------------------------------------------------- DeviceImage
template <typename T>
class DeviceImage<T>::Impl {
public:
    vip_image* img = nullptr;
    std::size_t elems = 0;
};

template <typename T>
DeviceImage<T>::DeviceImage(const int width, const int height, const int channels) : impl_(new Impl()) {
    VIP_CHECK(vip_image_create(&impl_->img, width, height, channels, static_cast<int>(sizeof(T))));
    if (impl_->img) impl_->elems = vip_image_bytes(impl_->img) / sizeof(T);
}
template <typename T>
DeviceImage<T>::~DeviceImage() {
    if (impl_) {
        vip_image_destroy(impl_->img);
        delete impl_;
    }
}
template <typename T>
DeviceImage<T>::DeviceImage(DeviceImage&& other) noexcept : impl_(other.impl_) {
    other.impl_ = nullptr;
}
template <typename T>
void DeviceImage<T>::upload(const T* const data) {
    VIP_CHECK(vip_image_upload(impl_->img, data));
}
template <typename T>
void DeviceImage<T>::download(T* const data) {
    VIP_CHECK(vip_image_download(impl_->img, data));
}
template <typename T>
void DeviceImage<T>::upload_async(const T* const data, void* stream) {
    VIP_CHECK(vip_image_upload_async(impl_->img, data, stream));
}
template <typename T>
void DeviceImage<T>::download_async(T* const data, void* stream) {
    VIP_CHECK(vip_image_download_async(impl_->img, data, stream));
}
template <typename T>
T* DeviceImage<T>::get() {
    return static_cast<T*>(vip_image_data(impl_->img));
}
template <typename T>
std::size_t DeviceImage<T>::size() const {
    return impl_->elems;
}
template class VIP_API DeviceImage<std::uint8_t>;
template class VIP_API DeviceImage<float>;

// ---------------------------------------------------------------- CudaBilateralFilter

CudaBilateralFilter::CudaBilateralFilter(const int width, const int height, const int ksize, const float sigma_space,
                                         const float sigma_color)
    : impl_(std::make_unique<Impl>(width, height, ksize, sigma_space, sigma_color)) {}
CudaBilateralFilter::~CudaBilateralFilter() = default;
void CudaBilateralFilter::bilateral_filter(const std::uint8_t* const d_src, std::uint8_t* const d_dst) const {
    impl_->bilateral_filter(d_src, d_dst);
    VIP_CHECK(vip_device_synchronize());  // src/bilateral_filter_impl.cu:299
}
void CudaBilateralFilter::joint_bilateral_filter(const std::uint8_t* const d_src, const std::uint8_t* const d_guide,
                                                 std::uint8_t* const d_dst) const {
    impl_->joint_bilateral_filter(d_src, d_guide, d_dst);
    VIP_CHECK(vip_device_synchronize());  // :309
}
void CudaBilateralFilter::bilateral_filter_async(const std::uint8_t* d_src, std::uint8_t* d_dst, void* stream) const {
    impl_->bilateral_filter(d_src, d_dst, stream);
}
void CudaBilateralFilter::joint_bilateral_filter_async(const std::uint8_t* d_src, const std::uint8_t* d_guide,
                                                       std::uint8_t* d_dst, void* stream) const {
    impl_->joint_bilateral_filter(d_src, d_guide, d_dst, stream);
}
void CudaBilateralFilter::bilateral_filter_strip(const std::uint8_t* d_src, std::uint8_t* d_dst, const vip_strip& strip,
                                                 void* stream) const {
    VIP_CHECK(vip_bf_run_strip(impl_->handle, d_src, d_dst, &strip, stream));
}
int CudaBilateralFilter::halo_rows() const { return vip_bf_halo_rows(impl_->handle); }

// ---------------------------------------------------------------- CudaAdaptiveBilateralFilter

CudaAdaptiveBilateralFilter::CudaAdaptiveBilateralFilter(const int width, const int height, const int ksize,
                                                         const float sigma_space, const float sigma_color)
    : impl_(std::make_unique<Impl>(width, height, ksize, sigma_space, sigma_color)) {}
CudaAdaptiveBilateralFilter::~CudaAdaptiveBilateralFilter() = default;
void CudaAdaptiveBilateralFilter::execute(const std::uint8_t* const d_src, std::uint8_t* const d_dst) const {
    impl_->execute(d_src, d_dst);
    VIP_CHECK(vip_device_synchronize());  // src/adaptive_bilateral_filter_impl.cu:190
}
void CudaAdaptiveBilateralFilter::execute_async(const std::uint8_t* d_src, std::uint8_t* d_dst, void* stream) const {
    impl_->execute(d_src, d_dst, stream);
}
void CudaAdaptiveBilateralFilter::execute_strip(const std::uint8_t* d_src, std::uint8_t* d_dst, const vip_strip& strip,
                                                void* stream) const {
    VIP_CHECK(vip_abf_run_strip(impl_->handle, d_src, d_dst, &strip, stream));
}
int CudaAdaptiveBilateralFilter::halo_rows() const { return vip_abf_halo_rows(impl_->handle); }

// ---------------------------------------------------------------- CudaBilateralTextureFilter

CudaBilateralTextureFilter::CudaBilateralTextureFilter(const int width, const int height, const int ksize,
                                                       const int nitr)
    : impl_(std::make_unique<Impl>(width, height, ksize, nitr)) {}
CudaBilateralTextureFilter::~CudaBilateralTextureFilter() = default;
void CudaBilateralTextureFilter::execute(const std::uint8_t* const d_src, std::uint8_t* const d_dst) {
    impl_->execute(d_src, d_dst);
    VIP_CHECK(vip_device_synchronize());  // src/bilateral_texture_filter_impl.cu:274
}
void CudaBilateralTextureFilter::execute_async(const std::uint8_t* d_src, std::uint8_t* d_dst, void* stream) {
    impl_->execute(d_src, d_dst, stream);
}
int CudaBilateralTextureFilter::halo_rows_per_iteration() const { return vip_btf_halo_rows_per_iteration(impl_->handle); }

// ---------------------------------------------------------------- gradient (asynchronous, like the reference)
template <>
VIP_API void cuda_gradient_impl<std::uint8_t>(const std::uint8_t* const d_src, float* const d_dst, const int width,
                                      const int height, const int src_ch) {
    VIP_CHECK(vip_gradient_u8(d_src, d_dst, width, height, src_ch, nullptr));
}
template <>
VIP_API void cuda_gradient_impl<float>(const float* const d_src, float* const d_dst, const int width, const int height,
                               const int src_ch) {
    VIP_CHECK(vip_gradient_f32(d_src, d_dst, width, height, src_ch, nullptr));
}

// ---------------------------------------------------------------- CudaIntegralImage
CudaIntegralImage::CudaIntegralImage(const int width, const int height, const int radius)
    : width_(width), height_(height), radius_(radius), table_(nullptr) {
    VIP_CHECK(vip_image_create(&table_, width + 2 * radius + 1, height + 2 * radius + 1, 3, 4));
}
CudaIntegralImage::~CudaIntegralImage() { vip_image_destroy(table_); }
void CudaIntegralImage::build(const std::uint8_t* d_src, void* stream) {
    VIP_CHECK(vip_integral_u8c3(d_src, static_cast<std::uint32_t*>(vip_image_data(table_)), width_, height_, radius_, stream));
}
void CudaIntegralImage::get(const std::int32_t* d_rects, int n, std::uint32_t* d_out, void* stream) const {
    VIP_CHECK(vip_integral_get_u8c3(table(), width_, height_, radius_, d_rects, n, d_out, stream));
}
void CudaIntegralImage::box_mean(float* d_mean, void* stream) const {
    VIP_CHECK(vip_integral_box_mean_u8c3(table(), d_mean, width_, height_, radius_, stream));
}
const std::uint32_t* CudaIntegralImage::table() const { return static_cast<const std::uint32_t*>(vip_image_data(table_)); }

// ---------------------------------------------------------------- MultiGpuFilter
MultiGpuFilter::MultiGpuFilter(const std::vector<int>& device_ids, Kind kind, int width, int height, int ksize,
                               float sigma_space, float sigma_color, int nitr, BtfHalo btf_halo)
    : handle_(nullptr) {
    VIP_CHECK(vip_mg_create_ex(&handle_, device_ids.data(), static_cast<int>(device_ids.size()), static_cast<int>(kind),
                               width, height, ksize, sigma_space, sigma_color, nitr, static_cast<int>(btf_halo)));
}
MultiGpuFilter::~MultiGpuFilter() { vip_mg_destroy(handle_); }
void MultiGpuFilter::execute(const std::uint8_t* h_src, std::uint8_t* h_dst) {
    VIP_CHECK(vip_mg_scatter_host(handle_, h_src));
    VIP_CHECK(vip_mg_run(handle_));
    VIP_CHECK(vip_mg_gather_host(handle_, h_dst));
}
void MultiGpuFilter::execute_batch(const std::uint8_t* h_src, std::uint8_t* h_dst, int n_frames) {
    VIP_CHECK(vip_mg_run_batch_host(handle_, h_src, h_dst, n_frames));
}
void MultiGpuFilter::execute_batch_set(int filter_set, const std::uint8_t* h_src, std::uint8_t* const h_dst[3], int n_frames) {
    VIP_CHECK(vip_mg_run_batch_set_host(handle_, filter_set, h_src, h_dst, n_frames));
}
bool MultiGpuFilter::peer_access() const { return vip_mg_peer_access(handle_) == 1; }
float MultiGpuFilter::last_run_ms() const {
    float ms = 0.f;
    VIP_CHECK(vip_mg_last_run_ms(handle_, &ms));
    return ms;
}
