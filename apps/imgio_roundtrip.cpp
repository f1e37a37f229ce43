// Decoder/encoder self-check without a GPU: read any supported file, write it back (alpha dropped) as PNG or PNM.
#include <iostream>

#include "image_io.hpp"

int main(int argc, char** argv) {
    if (argc != 3) {
        std::cerr << "[Usage] ./imgio_roundtrip input output" << std::endl;
        return 2;
    }
    vipio::Image img = vipio::imread(argv[1]);
    if (img.empty()) {
        std::cerr << "Failed to load " << argv[1] << std::endl;
        return 1;
    }
    const int oc = img.channels <= 2 ? 1 : 3;
    std::vector<std::uint8_t> out(static_cast<size_t>(img.width) * img.height * oc);
    for (size_t p = 0; p < static_cast<size_t>(img.width) * img.height; ++p)
        for (int c = 0; c < oc; ++c) out[p * oc + c] = img.data[p * img.channels + c];
    std::cout << img.width << " " << img.height << " " << img.channels << std::endl;
    return vipio::imwrite(argv[2], out.data(), img.width, img.height, oc) ? 0 : 1;
}
