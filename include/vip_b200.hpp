// C++ host API of the B200-native edge-preserving-filter library.
//
// Drop-in for the reference's include/cuda/ headers (yuyuyu-bot/various_image_processings): the class and function
// names, constructor / method signatures, default arguments, the `protected: class Impl; std::unique_ptr<Impl>
// impl_;` layout the reference's tests rely on (test/bilateral_filter.cu:9-33), the synchronous behaviour (every
// filter call returns after cudaDeviceSynchronize(), cuda_gradient returns right after the launch) and the error
// convention (CUDA errors are printed to stderr, never thrown: src/host_utilities.hpp:7-13) are kept:
//
//   CudaBilateralFilter            include/cuda/bilateral_filter.hpp:7-29
//   CudaAdaptiveBilateralFilter    include/cuda/adaptive_bilateral_filter.hpp:7-24
//   CudaBilateralTextureFilter     include/cuda/bilateral_texture_filter.hpp:7-17
//   DeviceImage<T>                 include/cuda/device_image.hpp:4-16
//   cuda_gradient<T>               include/cuda/gradient.hpp:4-23
//
// The old include paths ("cuda/bilateral_filter.hpp", ...) still work: they forward to this file.  Everything is
// implemented in plain C++ on top of the C ABI in vip_c_api.h; user code needs neither nvcc nor CUDA headers.
//
// Additions (not in the reference): stream-aware `*_async` methods, row-strip methods for the multi-GPU driver,
// CudaIntegralImage (GPU twin of the CPU-only BorderReplicatedIntegralImage) and MultiGpuFilter.
#ifndef VIP_B200_HPP
#define VIP_B200_HPP

#include <cstddef>
#include <cstdint>
#include <memory>
#include <type_traits>
#include <vector>

#include "vip_c_api.h"

// ------------------------------------------------------------------------------------------------ DeviceImage
template <typename ElemType>
class VIP_API DeviceImage {
    static_assert(std::is_same<ElemType, std::uint8_t>::value || std::is_same<ElemType, float>::value,
                  "DeviceImage is provided for uint8_t and float (src/device_image.cu:51-52)");

public:
    DeviceImage(const int width, const int height, const int channels = 1);
    ~DeviceImage();
    DeviceImage(const DeviceImage&) = delete;             // the reference's raw owning pointer would double-free
    DeviceImage& operator=(const DeviceImage&) = delete;  // on copy; copying is disabled here
    DeviceImage(DeviceImage&& other) noexcept;

    void upload(const ElemType* const data);    // blocking, like thrust::copy (src/device_image.cu:10-12)
    void download(ElemType* const data);        // blocking (src/device_image.cu:14-16)
    ElemType* get();

    // additions
    void upload_async(const ElemType* const pinned_data, void* stream);
    void download_async(ElemType* const pinned_data, void* stream);
    std::size_t size() const;  // elements

private:
    class Impl;
    Impl* impl_;
};

// ------------------------------------------------------------------------------------------------ bilateral
class VIP_API CudaBilateralFilter {
public:
    CudaBilateralFilter(const int width, const int height, const int ksize = 9, const float sigma_space = 10.f,
                        const float sigma_color = 30.f);
    ~CudaBilateralFilter();

    void bilateral_filter(const std::uint8_t* const d_src, std::uint8_t* const d_dst) const;
    void joint_bilateral_filter(const std::uint8_t* const d_src, const std::uint8_t* const d_guide,
                                std::uint8_t* const d_dst) const;

    // additions: no implicit synchronisation; `stream` is a cudaStream_t
    void bilateral_filter_async(const std::uint8_t* d_src, std::uint8_t* d_dst, void* stream) const;
    void joint_bilateral_filter_async(const std::uint8_t* d_src, const std::uint8_t* d_guide, std::uint8_t* d_dst,
                                      void* stream) const;
    void bilateral_filter_strip(const std::uint8_t* d_src, std::uint8_t* d_dst, const vip_strip& strip,
                                void* stream) const;
    int halo_rows() const;

protected:
    class Impl;
    std::unique_ptr<Impl> impl_;
};

// ------------------------------------------------------------------------------------------------ adaptive bilateral
class VIP_API CudaAdaptiveBilateralFilter {
public:
    CudaAdaptiveBilateralFilter(const int width, const int height, const int ksize = 9, const float sigma_space = 10.f,
                                const float sigma_color = 30.f);
    ~CudaAdaptiveBilateralFilter();

    void execute(const std::uint8_t* const d_src, std::uint8_t* const d_dst) const;

    void execute_async(const std::uint8_t* d_src, std::uint8_t* d_dst, void* stream) const;
    void execute_strip(const std::uint8_t* d_src, std::uint8_t* d_dst, const vip_strip& strip, void* stream) const;
    int halo_rows() const;

protected:
    class Impl;
    std::unique_ptr<Impl> impl_;
};

// ------------------------------------------------------------------------------------------------ bilateral texture
class VIP_API CudaBilateralTextureFilter {
public:
    CudaBilateralTextureFilter(const int width, const int height, const int ksize = 9, const int nitr = 3);
    ~CudaBilateralTextureFilter();

    void execute(const std::uint8_t* const d_src, std::uint8_t* const d_dst);

    void execute_async(const std::uint8_t* d_src, std::uint8_t* d_dst, void* stream);
    int halo_rows_per_iteration() const;

protected:
    class Impl;
    std::unique_ptr<Impl> impl_;
};

// ------------------------------------------------------------------------------------------------ gradient
template <typename SrcType>
VIP_API void cuda_gradient_impl(const SrcType* const d_src, float* const d_dst, const int width, const int height,
                        const int src_ch);

template <typename SrcType>
void cuda_gradient(const SrcType* const d_src, float* const d_dst, const int width, const int height,
                   const int src_ch = 1) {
    static_assert(std::is_same_v<SrcType, std::uint8_t> || std::is_same_v<SrcType, float>);
    cuda_gradient_impl(d_src, d_dst, width, height, src_ch);
}

// ------------------------------------------------------------------------------------------------ additions
// GPU twin of BorderReplicatedIntegralImage<uint8_t, 3> (include/cpp/border_replicated_integral_image.hpp:6-85).
class VIP_API CudaIntegralImage {
public:
    CudaIntegralImage(const int width, const int height, const int radius);
    ~CudaIntegralImage();
    CudaIntegralImage(const CudaIntegralImage&) = delete;
    CudaIntegralImage& operator=(const CudaIntegralImage&) = delete;

    void build(const std::uint8_t* d_src, void* stream = nullptr);
    // get(x0, y0, x1, y1) for n rectangles (device arrays: n x 4 int32 in, n x 3 uint32 out)
    void get(const std::int32_t* d_rects, int n, std::uint32_t* d_out, void* stream = nullptr) const;
    void box_mean(float* d_mean, void* stream = nullptr) const;
    const std::uint32_t* table() const;
    int table_width() const { return width_ + 2 * radius_ + 1; }
    int table_height() const { return height_ + 2 * radius_ + 1; }

private:
    int width_, height_, radius_;
    vip_image* table_;
};

// BorderReplicatedIntegralImage<SrcType, Channels> on the GPU for every instantiation the reference tests
// (test/border_replicated_integral_image.cpp:63-95): uint8_t / uint16_t sources -> uint32 table (modulo 2^32),
// float -> float table.  Header-only on top of vip_integral_build / vip_integral_get.
template <typename SrcType, int Channels>
class CudaBorderReplicatedIntegralImage {
    static_assert(std::is_same<SrcType, std::uint8_t>::value || std::is_same<SrcType, std::uint16_t>::value ||
                      std::is_same<SrcType, float>::value, "uint8_t, uint16_t or float");
    static_assert(Channels >= 1 && Channels <= 4, "1..4 channels");

public:
    using AccType = typename std::conditional<std::is_same<SrcType, float>::value, float, std::uint32_t>::type;
    static constexpr int kSrcType = std::is_same<SrcType, float>::value ? VIP_SRC_F32
                                    : (std::is_same<SrcType, std::uint16_t>::value ? VIP_SRC_U16 : VIP_SRC_U8);

    CudaBorderReplicatedIntegralImage(const int width, const int height, const int radius)
        : width_(width), height_(height), radius_(radius), table_(nullptr) {
        vip_image_create(&table_, width + 2 * radius + 1, height + 2 * radius + 1, Channels, 4);
    }
    ~CudaBorderReplicatedIntegralImage() { vip_image_destroy(table_); }
    CudaBorderReplicatedIntegralImage(const CudaBorderReplicatedIntegralImage&) = delete;
    CudaBorderReplicatedIntegralImage& operator=(const CudaBorderReplicatedIntegralImage&) = delete;

    int build(const SrcType* d_src, void* stream = nullptr) {
        return vip_integral_build(d_src, vip_image_data(table_), width_, height_, Channels, radius_, kSrcType, stream);
    }
    // n rectangles (x0, y0, x1, y1) on the device -> n x Channels inclusive box sums (get(), :34-47)
    int get(const std::int32_t* d_rects, int n, AccType* d_out, void* stream = nullptr) const {
        return vip_integral_get(vip_image_data(table_), width_, height_, Channels, radius_, kSrcType, d_rects, n, d_out,
                                stream);
    }
    const AccType* table() const { return static_cast<const AccType*>(vip_image_data(table_)); }

private:
    int width_, height_, radius_;
    vip_image* table_;
};

// Row strips with halo exchange / frame batches over the GPUs of one box (vip_mg_* in vip_c_api.h).
class VIP_API MultiGpuFilter {
public:
    enum Kind { Bilateral = VIP_MG_BF, AdaptiveBilateral = VIP_MG_ABF, BilateralTexture = VIP_MG_BTF };
    // BilateralTexture strips: deep halos (nitr x 17 rows from the scatter, overlap recomputed, no communication) or a
    // peer-to-peer halo exchange after every iteration; Automatic = deep while the redundant rows are <= 35 % of a strip
    enum BtfHalo { Automatic = 0, Exchange = VIP_MG_FLAG_BTF_EXCHANGE, DeepHalo = VIP_MG_FLAG_BTF_DEEP_HALO };
    MultiGpuFilter(const std::vector<int>& device_ids, Kind kind, int width, int height, int ksize = 9,
                   float sigma_space = 10.f, float sigma_color = 30.f, int nitr = 3, BtfHalo btf_halo = Automatic);
    ~MultiGpuFilter();
    MultiGpuFilter(const MultiGpuFilter&) = delete;
    MultiGpuFilter& operator=(const MultiGpuFilter&) = delete;

    bool ok() const { return handle_ != nullptr; }
    void execute(const std::uint8_t* h_src, std::uint8_t* h_dst);  // scatter, run, gather
    void execute_batch(const std::uint8_t* h_src, std::uint8_t* h_dst, int n_frames);
    // every frame through a set of filters (VIP_MG_SET_BF | VIP_MG_SET_ABF | VIP_MG_SET_BTF), one upload per frame;
    // h_dst[j] receives filter j of the set (entries of filters not in the set may be null)
    void execute_batch_set(int filter_set, const std::uint8_t* h_src, std::uint8_t* const h_dst[3], int n_frames);
    bool peer_access() const;   // halo rows move GPU<->GPU with peer access enabled (false: staged through the host)
    float last_run_ms() const;

private:
    vip_mg* handle_;
};

#endif  // VIP_B200_HPP
