// TEST INFRASTRUCTURE ONLY (oracle/). Stand-in for opencv_contrib's ximgproc edge_filter.hpp.
//
// The reference's CPU bilateral texture filter calls cv::ximgproc::jointBilateralFilter
// (/root/reference/include/cpp/bilateral_texture_filter.hpp:162), a third-party function that is
// neither vendored in the reference nor installed here (opencv_contrib, version unpinned by
// /root/reference/test/CMakeLists.txt:5).  The reference's CUDA path does NOT use it: it runs
// the reference's own joint bilateral kernel with (ksize 2k-1, sigma_space k-1, sigma_color sqrt 3)
// (/root/reference/src/bilateral_texture_filter_impl.cu:188,210).  This shim therefore routes the
// call to the reference's own CPU joint_bilateral_filter
// (/root/reference/include/cpp/bilateral_filter.hpp:126-207) with the same three parameters, which
// is exactly what the reference's CUDA implementation computes.
#ifndef VIP_ORACLE_CVSHIM_XIMGPROC_EDGE_FILTER_HPP
#define VIP_ORACLE_CVSHIM_XIMGPROC_EDGE_FILTER_HPP

#include <opencv2/core.hpp>

#include "cpp/bilateral_filter.hpp"

namespace cv {
namespace ximgproc {

// OpenCV argument order: (joint, src, dst, d, sigmaColor, sigmaSpace, borderType)
inline void jointBilateralFilter(const cv::Mat3b& joint, const cv::Mat3b& src, cv::Mat3b& dst, int d,
                                 double sigma_color, double sigma_space, int /*border_type*/ = BORDER_DEFAULT) {
    ::joint_bilateral_filter(src, joint, dst, d, static_cast<float>(sigma_space), static_cast<float>(sigma_color));
}

}  // namespace ximgproc
}  // namespace cv

#endif  // VIP_ORACLE_CVSHIM_XIMGPROC_EDGE_FILTER_HPP
