// Headless twin of sample/gradient/main.cpp: same argv (filename), the same DeviceImage / cuda_gradient calls (:40-47)
// and the same display scaling (convert_to_u8, :9-18: img * 255 / max -> 8 bit), done on the device; the result goes
// to <stem>_gradient.png (or the path given with -o) instead of cv::imshow windows.
#include <algorithm>
#include <cstdlib>
#include <iostream>
#include <string>

#include "app_common.hpp"
#include "cuda/gradient.hpp"

int main(int argc, char** argv) {
    std::string out_path;
    std::vector<char*> pos;
    for (int i = 1; i < argc; ++i) {
        if (std::string(argv[i]) == "-o" && i + 1 < argc) out_path = argv[++i];
        else pos.push_back(argv[i]);
    }
    if (pos.size() != 1) {
        std::cerr << "[Usage] ./gradient filename [-o output]" << std::endl;
        std::exit(EXIT_FAILURE);
    }
    const auto filename = std::string(pos[0]);
    vipio::Image input_image = vipio::imread(filename);
    if (input_image.empty()) {
        std::cerr << "Failed to load " << filename << std::endl;
        return 0;
    }
    const int cols = input_image.width, rows = input_image.height;

    auto d_src = vipapp::upload_as_bgr(input_image);
    DeviceImage<float> d_dst(cols, rows);
    cuda_gradient(d_src->get(), d_dst.get(), cols, rows, 3);
    vip_device_synchronize();  // cuda_gradient returns right after the launch, like the reference's

    std::vector<float> dst_cuda(static_cast<size_t>(cols) * rows);
    d_dst.download(dst_cuda.data());
    const float max = *std::max_element(dst_cuda.begin(), dst_cuda.end());
    DeviceImage<std::uint8_t> d_u8(cols, rows, 1);
    std::vector<std::uint8_t> u8(dst_cuda.size(), 0);
    if (max > 0.f) {
        vip_scale_to_u8_f32(d_dst.get(), d_u8.get(), dst_cuda.size(), max, nullptr);
        d_u8.download(u8.data());
    }
    if (out_path.empty()) out_path = vipapp::stem(filename) + "_gradient.png";
    if (vipio::imwrite(out_path, u8.data(), cols, rows, 1)) std::cout << "wrote " << out_path << " (max magnitude " << max << ")" << std::endl;
    return 0;
}
