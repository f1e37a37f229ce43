// Headless twin of sample/bilateral_texture_filter/main.cpp: same argv (filename [ksize] [nitr]) and the same
// DeviceImage / CudaBilateralTextureFilter calls (:30-37); the result goes to <stem>_bilateral_texture_filter.png (or
// the path given with -o) instead of cv::imshow windows.
#include <cstdlib>
#include <iostream>
#include <string>

#include "app_common.hpp"
#include "cuda/bilateral_texture_filter.hpp"

int main(int argc, char** argv) {
    std::string out_path;
    std::vector<char*> pos;
    for (int i = 1; i < argc; ++i) {
        if (std::string(argv[i]) == "-o" && i + 1 < argc) out_path = argv[++i];
        else pos.push_back(argv[i]);
    }
    if (pos.empty()) {
        std::cerr << "[Usage] ./bilateral_texture_filter filename [ksize] [nitr] [-o output]" << std::endl;
        std::exit(EXIT_FAILURE);
    }
    const auto filename = std::string(pos[0]);
    const auto ksize    = pos.size() >= 2 ? std::stoi(pos[1]) : 9;
    const auto nitr     = pos.size() >= 3 ? std::stoi(pos[2]) : 3;  // the reference parses this with std::stof (:17)

    vipio::Image input_image = vipio::imread(filename);
    if (input_image.empty()) {
        std::cerr << "Failed to load " << filename << std::endl;
        return 0;
    }
    const int cols = input_image.width, rows = input_image.height;

    auto d_input_image = vipapp::upload_as_bgr(input_image);
    DeviceImage<std::uint8_t> d_dst(cols, rows, 3);
    CudaBilateralTextureFilter filter(cols, rows, ksize, nitr);
    filter.execute(d_input_image->get(), d_dst.get());

    if (out_path.empty()) out_path = vipapp::stem(filename) + "_bilateral_texture_filter.png";
    if (vipapp::save_bgr(out_path, d_dst, cols, rows)) std::cout << "wrote " << out_path << std::endl;
    return 0;
}
