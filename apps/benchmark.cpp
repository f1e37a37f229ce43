// Benchmark harness with the shape of sample/benchmark/main.cpp: the same `Parameters` struct and config keys
// (:39-89, sample/benchmark/config.toml), the same MEASURE protocol (1 warm-up + execute_times timed runs, wall clock,
// mean in msec, :20-33), the same "name [cuda] : x [msec]" lines (:35-37) for gradient / bilateral / adaptive bilateral /
// bilateral texture filter on the same default input (100x100, uniform [100,120), :210-213).
//
// Differences, all deliberate:
//  * only the [cuda] lines are produced -- the C++ CPU path belongs to the reference (bench.py times it as the baseline);
//  * the config path is argv[1] (the reference documents argv[1] but reads argv[2]); a tiny parser for the
//    config.toml subset replaces toml11;
//  * `gradient [cuda]` is synchronised before the clock stops (the reference times the launch only, :120-121);
//  * --width/--height/--image choose a realistic input (a 100x100 image cannot load a B200), --device-time adds a
//    CUDA-event timing through the async API, --json prints one JSON object with everything.
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <iostream>
#include <random>
#include <sstream>
#include <string>
#include <vector>

#include "app_common.hpp"
#include "cuda/adaptive_bilateral_filter.hpp"
#include "cuda/bilateral_filter.hpp"
#include "cuda/bilateral_texture_filter.hpp"
#include "cuda/gradient.hpp"

#define MEASURE(num_itr, fn, duration)                                                                   \
    do {                                                                                                 \
        std::int64_t sum_duration = 0ll;                                                                 \
        for (int i = 0; i <= num_itr; i++) {                                                             \
            const auto start = std::chrono::steady_clock::now();                                         \
            fn;                                                                                          \
            const auto end = std::chrono::steady_clock::now();                                           \
            if (i != 0) sum_duration += std::chrono::duration_cast<std::chrono::nanoseconds>(end - start).count(); \
        }                                                                                                \
        duration = sum_duration / num_itr / 1e6f;                                                        \
    } while (false)

static void print_duration(const std::string& name, const float duration) {
    std::printf("%-40s : %10.6f [msec]\n", name.c_str(), duration);
}

struct Parameters {
    int execute_times = 50;
    struct {
        int ksize = 9;
    } BilateralFilter;
    struct {
        int ksize = 9;
    } AdaptiveBilateralFilter;
    struct {
        int ksize = 9;
        int nitr = 3;
    } BilateralTextureFilter;
};

// `key = integer` lines under optional `[Section]` headers; '#' starts a comment.  Unknown keys are ignored.
static Parameters parse_config(const std::string& filename) {
    Parameters params;
    std::ifstream in(filename);
    if (!in) {
        std::cerr << "cannot open config " << filename << ", using defaults" << std::endl;
        return params;
    }
    std::string line, section;
    while (std::getline(in, line)) {
        const size_t hash = line.find('#');
        if (hash != std::string::npos) line.erase(hash);
        const size_t a = line.find_first_not_of(" \t\r");
        if (a == std::string::npos) continue;
        if (line[a] == '[') {
            const size_t b = line.find(']', a);
            section = line.substr(a + 1, b == std::string::npos ? std::string::npos : b - a - 1);
            continue;
        }
        const size_t eq = line.find('=');
        if (eq == std::string::npos) continue;
        std::string key = line.substr(a, eq - a);
        key.erase(key.find_last_not_of(" \t") + 1);
        const int value = std::atoi(line.c_str() + eq + 1);
        if (section.empty() && key == "execute_times") params.execute_times = value;
        else if (section == "BilateralFilter" && key == "ksize") params.BilateralFilter.ksize = value;
        else if (section == "AdaptiveBilateralFilter" && key == "ksize") params.AdaptiveBilateralFilter.ksize = value;
        else if (section == "BilateralTextureFilter" && key == "ksize") params.BilateralTextureFilter.ksize = value;
        else if (section == "BilateralTextureFilter" && key == "nitr") params.BilateralTextureFilter.nitr = value;
    }
    return params;
}

struct Result {
    std::string name;
    float wall_ms;
    float device_ms;  // < 0: not measured
};

int main(int argc, char** argv) {
    std::string config, image_path;
    int width = 100, height = 100;
    bool json = false, device_time = false;
    for (int i = 1; i < argc; ++i) {
        const std::string a = argv[i];
        if (a == "--width" && i + 1 < argc) width = std::atoi(argv[++i]);
        else if (a == "--height" && i + 1 < argc) height = std::atoi(argv[++i]);
        else if (a == "--image" && i + 1 < argc) image_path = argv[++i];
        else if (a == "--json") json = true;
        else if (a == "--device-time") device_time = true;
        else if (a == "--help") {
            std::cerr << "[Usage] ./benchmark [config_path] [--width W --height H | --image file] [--device-time] [--json]" << std::endl;
            return 0;
        } else config = a;
    }
    const Parameters params = config.empty() ? Parameters() : parse_config(config);

    // input: cv::randu(input_image, 100, 120) in the reference (:213); here mt19937(42) uniform in [100, 120)
    vipio::Image input;
    if (!image_path.empty()) {
        input = vipio::imread(image_path);
        if (input.empty()) {
            std::cerr << "Failed to load " << image_path << std::endl;
            return 1;
        }
        width = input.width;
        height = input.height;
    } else {
        input.width = width;
        input.height = height;
        input.channels = 3;
        input.data.resize(static_cast<size_t>(width) * height * 3);
        std::mt19937 rng(42);
        for (auto& v : input.data) v = static_cast<std::uint8_t>(100 + rng() % 20);
    }

    std::cout << "Parameters" << std::endl;
    std::cout << "\twidth         : " << width << std::endl;
    std::cout << "\theight        : " << height << std::endl;
    std::cout << "\texecute times : " << params.execute_times << std::endl << std::endl;
    std::cout << "\t[bilateral filter]" << std::endl;
    std::cout << "\t\tksize : " << params.BilateralFilter.ksize << std::endl;
    std::cout << "\t[adaptive bilateral filter]" << std::endl;
    std::cout << "\t\tksize : " << params.AdaptiveBilateralFilter.ksize << std::endl;
    std::cout << "\t[bilateral texture filter]" << std::endl;
    std::cout << "\t\tksize : " << params.BilateralTextureFilter.ksize << std::endl;
    std::cout << "\t\tnitr  : " << params.BilateralTextureFilter.nitr << std::endl << std::endl;

    auto d_input_image = vipapp::upload_as_bgr(input);  // convert_to_3ch (:91-103), on the device
    DeviceImage<std::uint8_t> d_dst(width, height, 3);
    DeviceImage<float> d_grad(width, height);
    const int n = params.execute_times;
    std::vector<Result> results;
    float duration = 0.f;

    // device-side timing of `times` back-to-back asynchronous calls: wall clock around enqueue + one final sync
    auto device_ms = [&](auto&& enqueue) -> float {
        if (!device_time) return -1.f;
        enqueue();
        vip_device_synchronize();
        const auto t0 = std::chrono::steady_clock::now();
        for (int i = 0; i < n; ++i) enqueue();
        vip_device_synchronize();
        const auto t1 = std::chrono::steady_clock::now();
        return std::chrono::duration_cast<std::chrono::nanoseconds>(t1 - t0).count() / 1e6f / n;
    };

    MEASURE(n, (cuda_gradient(d_input_image->get(), d_grad.get(), width, height, 3), vip_device_synchronize()), duration);
    print_duration("gradient [cuda]", duration);
    results.push_back({"gradient", duration, device_ms([&] { cuda_gradient(d_input_image->get(), d_grad.get(), width, height, 3); })});

    {
        CudaBilateralFilter filter(width, height, params.BilateralFilter.ksize);
        MEASURE(n, filter.bilateral_filter(d_input_image->get(), d_dst.get()), duration);
        print_duration("bilateral filter [cuda]", duration);
        results.push_back({"bilateral filter", duration,
                           device_ms([&] { filter.bilateral_filter_async(d_input_image->get(), d_dst.get(), nullptr); })});
    }
    {
        CudaAdaptiveBilateralFilter filter(width, height, params.AdaptiveBilateralFilter.ksize);
        MEASURE(n, filter.execute(d_input_image->get(), d_dst.get()), duration);
        print_duration("adaptive bilateral filter [cuda]", duration);
        results.push_back({"adaptive bilateral filter", duration,
                           device_ms([&] { filter.execute_async(d_input_image->get(), d_dst.get(), nullptr); })});
    }
    {
        CudaBilateralTextureFilter filter(width, height, params.BilateralTextureFilter.ksize, params.BilateralTextureFilter.nitr);
        MEASURE(n, filter.execute(d_input_image->get(), d_dst.get()), duration);
        print_duration("bilateral texture filter [cuda]", duration);
        results.push_back({"bilateral texture filter", duration,
                           device_ms([&] { filter.execute_async(d_input_image->get(), d_dst.get(), nullptr); })});
    }

    // the reference's harness ends with a CPU-only SuperpixelSLIC section (sample/benchmark/main.cpp:188-201, 240).  SLIC
    // (and Wexler inpainting) are plain compute that COULD be CUDA kernels; they are outside the edge-preserving-filter hot
    // path this library covers and are not built -- said here instead of silently dropping the section
    std::cout << "superpixel SLIC [cuda] : not implemented (outside the hot-path scope of this library; no CUDA twin in the reference either)"
              << std::endl;

    if (device_time)
        for (const auto& r : results) print_duration(r.name + " [cuda, async x" + std::to_string(n) + "]", r.device_ms);
    if (json) {
        std::ostringstream os;
        os << "{\"width\": " << width << ", \"height\": " << height << ", \"execute_times\": " << n << ", \"results\": [";
        for (size_t i = 0; i < results.size(); ++i) {
            const double mpix = static_cast<double>(width) * height / 1e6;
            os << (i ? ", " : "") << "{\"name\": \"" << results[i].name << "\", \"wall_ms\": " << results[i].wall_ms
               << ", \"mpixel_per_s\": " << mpix / (results[i].wall_ms * 1e-3);
            if (results[i].device_ms >= 0) os << ", \"async_ms\": " << results[i].device_ms;
            os << "}";
        }
        os << "], \"launches\": " << vip_launch_count() << "}";
        std::cout << os.str() << std::endl;
    }
    return 0;
}
