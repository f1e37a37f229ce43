// TEST INFRASTRUCTURE ONLY (oracle/): a minimal stand-in for the subset of the OpenCV core API
// that the reference's CPU headers (/root/reference/include/cpp/*.hpp) use, so that those headers
// compile UNMODIFIED, in place, without OpenCV (which is absent from this image).
//
// This is not OpenCV code; it is an independent, header-only re-statement of the handful of
// types the reference touches: Mat / Mat_<T> / Vec<T,N> / Range / ParallelLoopBody /
// parallel_for_ / Scalar / Size and the CV_8UC* / CV_32FC* / CV_32SC* type codes.
// Never linked into the product library.
#ifndef VIP_ORACLE_CVSHIM_CORE_HPP
#define VIP_ORACLE_CVSHIM_CORE_HPP

#include <algorithm>
#include <array>
#include <cstdint>
#include <cstring>
#include <functional>
#include <iostream>
#include <limits>
#include <memory>
#include <thread>
#include <tuple>
#include <type_traits>
#include <vector>

#define CV_OVERRIDE override

#define CV_CN_SHIFT 3
#define CV_8U 0
#define CV_16U 2
#define CV_32S 4
#define CV_32F 5
#define CV_MAKETYPE(depth, cn) ((depth) + (((cn)-1) << CV_CN_SHIFT))
#define CV_8UC1 CV_MAKETYPE(CV_8U, 1)
#define CV_8UC3 CV_MAKETYPE(CV_8U, 3)
#define CV_8UC(n) CV_MAKETYPE(CV_8U, (n))
#define CV_16UC(n) CV_MAKETYPE(CV_16U, (n))
#define CV_32SC(n) CV_MAKETYPE(CV_32S, (n))
#define CV_32FC1 CV_MAKETYPE(CV_32F, 1)
#define CV_32FC3 CV_MAKETYPE(CV_32F, 3)
#define CV_32FC(n) CV_MAKETYPE(CV_32F, (n))

namespace cv {

inline int shim_depth_size(int type) {
    switch (type & ((1 << CV_CN_SHIFT) - 1)) {
        case CV_8U: return 1;
        case CV_16U: return 2;
        case CV_32S: return 4;
        case CV_32F: return 4;
        default: return 1;
    }
}
inline int shim_channels(int type) { return (type >> CV_CN_SHIFT) + 1; }

template <typename T> struct shim_depth;
template <> struct shim_depth<std::uint8_t> { static constexpr int value = CV_8U; };
template <> struct shim_depth<std::uint16_t> { static constexpr int value = CV_16U; };
template <> struct shim_depth<int> { static constexpr int value = CV_32S; };
template <> struct shim_depth<float> { static constexpr int value = CV_32F; };

template <typename T, int N>
struct Vec {
    T val[N];
    Vec() { for (int i = 0; i < N; i++) val[i] = T(0); }
    Vec(T a, T b, T c) { static_assert(N == 3); val[0] = a; val[1] = b; val[2] = c; }
    T& operator[](int i) { return val[i]; }
    const T& operator[](int i) const { return val[i]; }
    Vec operator+(const Vec& o) const { Vec r; for (int i = 0; i < N; i++) r.val[i] = val[i] + o.val[i]; return r; }
    Vec operator-(const Vec& o) const { Vec r; for (int i = 0; i < N; i++) r.val[i] = val[i] - o.val[i]; return r; }
};
using Vec3b = Vec<std::uint8_t, 3>;
using Vec3f = Vec<float, 3>;

template <typename T> struct shim_type { static constexpr int value = CV_MAKETYPE(shim_depth<T>::value, 1); };
template <typename T, int N> struct shim_type<Vec<T, N>> { static constexpr int value = CV_MAKETYPE(shim_depth<T>::value, N); };

struct Size {
    int width = 0, height = 0;
    Size() = default;
    Size(int w, int h) : width(w), height(h) {}
    bool operator==(const Size& o) const { return width == o.width && height == o.height; }
};

struct Scalar {
    double v[4];
    static Scalar all(double x) { Scalar s; s.v[0] = s.v[1] = s.v[2] = s.v[3] = x; return s; }
};

struct Range {
    int start = 0, end = 0;
    Range() = default;
    Range(int s, int e) : start(s), end(e) {}
};

class ParallelLoopBody {
public:
    virtual ~ParallelLoopBody() = default;
    virtual void operator()(const Range& range) const = 0;
};

inline int& shim_num_threads() {
    static int n = static_cast<int>(std::max(1u, std::thread::hardware_concurrency()));
    return n;
}
inline void setNumThreads(int n) {
    shim_num_threads() = n > 0 ? n : static_cast<int>(std::max(1u, std::thread::hardware_concurrency()));
}
inline int getNumThreads() { return shim_num_threads(); }

// Contiguous stripes, one std::thread per stripe (the calling thread takes the first).
inline void parallel_for_(const Range& range, const ParallelLoopBody& body, double = -1.) {
    const int total = range.end - range.start;
    if (total <= 0) return;
    const int nthreads = std::max(1, std::min(shim_num_threads(), total));
    if (nthreads == 1) { body(range); return; }
    // finer stripes than threads for load balance, handed out round-robin
    const int nstripes = std::min(total, nthreads * 8);
    std::vector<std::thread> workers;
    workers.reserve(nthreads - 1);
    auto run = [&](int tid) {
        for (int s = tid; s < nstripes; s += nthreads) {
            const int b = range.start + static_cast<int>(static_cast<long long>(total) * s / nstripes);
            const int e = range.start + static_cast<int>(static_cast<long long>(total) * (s + 1) / nstripes);
            if (e > b) body(Range(b, e));
        }
    };
    for (int t = 1; t < nthreads; t++) workers.emplace_back(run, t);
    run(0);
    for (auto& w : workers) w.join();
}

class Mat {
public:
    int rows = 0, cols = 0, dims = 0;
    std::uint8_t* data = nullptr;

    Mat() = default;
    Mat(int r, int c, int type) { create(r, c, type); }
    Mat(Size s, int type) { create(s.height, s.width, type); }
    Mat(int r, int c, int type, void* ext) : rows(r), cols(c), dims(2), data(static_cast<std::uint8_t*>(ext)), type_(type) {}

    int type() const { return type_; }
    int channels() const { return shim_channels(type_); }
    std::size_t elemSize() const { return static_cast<std::size_t>(shim_depth_size(type_)) * shim_channels(type_); }
    Size size() const { return Size(cols, rows); }
    bool empty() const { return data == nullptr || rows == 0 || cols == 0; }
    bool isContinuous() const { return true; }

    void create(int r, int c, int type) {
        if (data != nullptr && r == rows && c == cols && type == type_) return;
        rows = r; cols = c; dims = 2; type_ = type;
        store_ = std::make_shared<std::vector<std::uint8_t>>(static_cast<std::size_t>(r) * c * elemSize());
        data = store_->data();
    }
    void create(Size s, int type) { create(s.height, s.width, type); }
    // n-d form as used by the reference's integral image: only the first `ndims` sizes count.
    void create(int ndims, const int* sizes, int type) {
        create(sizes[0], ndims >= 2 ? sizes[1] : 1, type);
    }

    Mat& setTo(const Scalar& s) {
        if (s.v[0] == 0.0) { std::memset(data, 0, static_cast<std::size_t>(rows) * cols * elemSize()); return *this; }
        const int cn = channels();
        for (std::size_t i = 0; i < static_cast<std::size_t>(rows) * cols; i++) {
            for (int ch = 0; ch < cn; ch++) {
                switch (type_ & ((1 << CV_CN_SHIFT) - 1)) {
                    case CV_8U: data[i * cn + ch] = static_cast<std::uint8_t>(s.v[ch & 3]); break;
                    case CV_16U: reinterpret_cast<std::uint16_t*>(data)[i * cn + ch] = static_cast<std::uint16_t>(s.v[ch & 3]); break;
                    case CV_32S: reinterpret_cast<int*>(data)[i * cn + ch] = static_cast<int>(s.v[ch & 3]); break;
                    case CV_32F: reinterpret_cast<float*>(data)[i * cn + ch] = static_cast<float>(s.v[ch & 3]); break;
                }
            }
        }
        return *this;
    }

    void copyTo(Mat& dst) const {
        dst.create(rows, cols, type_);
        if (dst.data != data) std::memcpy(dst.data, data, static_cast<std::size_t>(rows) * cols * elemSize());
    }

    template <typename T> T& at(int y, int x) { return reinterpret_cast<T*>(data)[static_cast<std::size_t>(y) * cols + x]; }
    template <typename T> const T& at(int y, int x) const { return reinterpret_cast<const T*>(data)[static_cast<std::size_t>(y) * cols + x]; }
    template <typename T = std::uint8_t> T* ptr(int y = 0) { return reinterpret_cast<T*>(data + static_cast<std::size_t>(y) * cols * elemSize()); }
    template <typename T = std::uint8_t> const T* ptr(int y = 0) const { return reinterpret_cast<const T*>(data + static_cast<std::size_t>(y) * cols * elemSize()); }

protected:
    int type_ = 0;
    std::shared_ptr<std::vector<std::uint8_t>> store_;
};

template <typename T>
class Mat_ : public Mat {
public:
    Mat_() { type_ = shim_type<T>::value; }
    Mat_(int r, int c) { Mat::create(r, c, shim_type<T>::value); }
    explicit Mat_(Size s) { Mat::create(s.height, s.width, shim_type<T>::value); }
    Mat_(int r, int c, T* ext) : Mat(r, c, shim_type<T>::value, ext) {}
    Mat_(const Mat& m) : Mat(m) {}

    using Mat::create;
    void create(Size s) { Mat::create(s.height, s.width, shim_type<T>::value); }
    void create(int r, int c) { Mat::create(r, c, shim_type<T>::value); }

    T& operator()(int y, int x) { return reinterpret_cast<T*>(data)[static_cast<std::size_t>(y) * cols + x]; }
    const T& operator()(int y, int x) const { return reinterpret_cast<const T*>(data)[static_cast<std::size_t>(y) * cols + x]; }
};

using Mat1b = Mat_<std::uint8_t>;
using Mat3b = Mat_<Vec3b>;
using Mat1f = Mat_<float>;
using Mat3f = Mat_<Vec3f>;

enum BorderTypes { BORDER_CONSTANT = 0, BORDER_REPLICATE = 1, BORDER_DEFAULT = 4 };

}  // namespace cv

#endif  // VIP_ORACLE_CVSHIM_CORE_HPP
