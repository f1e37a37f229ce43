// Shared by the headless sample programs: file -> device BGR image and back.
// Replaces cv::imread(filename, cv::IMREAD_COLOR) + DeviceImage::upload (sample/bilateral_filter/main.cpp:19,35) and the
// cv::imshow windows (:40-42) -- no display exists on the GPU box, results are written as image files instead.
#ifndef VIP_APPS_COMMON_HPP
#define VIP_APPS_COMMON_HPP

#include <cstdint>
#include <iostream>
#include <memory>
#include <string>
#include <vector>

#include "cuda/device_image.hpp"
#include "image_io.hpp"
#include "vip_c_api.h"

namespace vipapp {

// Decoded file (any of gray / gray+alpha / RGB / RGBA) -> packed 8-bit BGR on the device, converted there.
inline std::unique_ptr<DeviceImage<std::uint8_t>> upload_as_bgr(vipio::Image& img) {
    if (img.channels == 2) {  // gray + alpha: drop alpha on the host (rare)
        std::vector<std::uint8_t> g(static_cast<size_t>(img.width) * img.height);
        for (size_t i = 0; i < g.size(); ++i) g[i] = img.data[2 * i];
        img.data.swap(g);
        img.channels = 1;
    }
    DeviceImage<std::uint8_t> raw(img.width, img.height, img.channels);
    raw.upload(img.data.data());
    std::unique_ptr<DeviceImage<std::uint8_t>> bgr(new DeviceImage<std::uint8_t>(img.width, img.height, 3));
    const int st = vip_convert_to_bgr_u8(raw.get(), img.channels, bgr->get(), img.width, img.height, /*swap_rb=*/1, nullptr);
    if (st != VIP_OK) std::cerr << "vip_convert_to_bgr_u8: " << vip_status_string(st) << std::endl;
    vip_device_synchronize();
    return bgr;
}

// Device BGR -> file (RGB order on disk).
inline bool save_bgr(const std::string& path, DeviceImage<std::uint8_t>& d_bgr, int width, int height) {
    DeviceImage<std::uint8_t> rgb(width, height, 3);
    vip_convert_to_bgr_u8(d_bgr.get(), 3, rgb.get(), width, height, 1, nullptr);
    std::vector<std::uint8_t> host(static_cast<size_t>(width) * height * 3);
    rgb.download(host.data());
    const bool ok = vipio::imwrite(path, host.data(), width, height, 3);
    if (!ok) std::cerr << "Failed to write " << path << std::endl;
    return ok;
}

inline std::string stem(const std::string& path) {
    const size_t slash = path.find_last_of('/');
    const std::string base = slash == std::string::npos ? path : path.substr(slash + 1);
    const size_t dot = base.find_last_of('.');
    return dot == std::string::npos ? base : base.substr(0, dot);
}

}  // namespace vipapp
#endif
