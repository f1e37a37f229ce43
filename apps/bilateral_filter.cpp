// Headless twin of sample/bilateral_filter/main.cpp: same argv (filename [ksize] [sigma_space] [sigma_color]) and the
// same DeviceImage / CudaBilateralFilter calls (:30-37); the result goes to <stem>_bilateral_filter.png (or the path
// given with -o) instead of cv::imshow windows.  The OpenCV comparison image of the sample (:24-27) is not produced.
#include <cstdlib>
#include <iostream>
#include <string>

#include "app_common.hpp"
#include "cuda/bilateral_filter.hpp"

int main(int argc, char** argv) {
    std::string out_path;
    std::vector<char*> pos;
    for (int i = 1; i < argc; ++i) {
        if (std::string(argv[i]) == "-o" && i + 1 < argc) out_path = argv[++i];
        else pos.push_back(argv[i]);
    }
    if (pos.empty()) {
        std::cerr << "[Usage] ./bilateral_filter filename [ksize] [sigma_space] [sigma_color] [-o output]" << std::endl;
        std::exit(EXIT_FAILURE);
    }
    const auto filename    = std::string(pos[0]);
    const auto ksize       = pos.size() >= 2 ? std::stoi(pos[1]) : 9;
    const auto sigma_space = pos.size() >= 3 ? std::stof(pos[2]) : 10.f;
    const auto sigma_color = pos.size() >= 4 ? std::stof(pos[3]) : 30.f;

    vipio::Image input_image = vipio::imread(filename);
    if (input_image.empty()) {
        std::cerr << "Failed to load " << filename << std::endl;
        return 0;
    }
    const int cols = input_image.width, rows = input_image.height;

    auto d_input_image = vipapp::upload_as_bgr(input_image);
    DeviceImage<std::uint8_t> d_dst(cols, rows, 3);
    CudaBilateralFilter filter(cols, rows, ksize, sigma_space, sigma_color);
    filter.bilateral_filter(d_input_image->get(), d_dst.get());

    if (out_path.empty()) out_path = vipapp::stem(filename) + "_bilateral_filter.png";
    if (vipapp::save_bgr(out_path, d_dst, cols, rows)) std::cout << "wrote " << out_path << std::endl;
    return 0;
}
