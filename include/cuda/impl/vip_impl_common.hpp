// Shared by the Impl headers of this directory: the reference's error convention (src/host_utilities.hpp:7-13 --
// report on stderr with file and line, never throw, carry on) on top of the C ABI's status codes.
#ifndef VIP_IMPL_COMMON_HPP
#define VIP_IMPL_COMMON_HPP

#include <cstdio>

#include "../../vip_c_api.h"

namespace vip_detail {
inline void report(int status, const char* what, const char* file, int line) {
    if (status != VIP_OK)
        std::fprintf(stderr, "vip_b200: %s failed: %s (%d) at %s:%d\n", what, vip_status_string(status), status, file, line);
}
}  // namespace vip_detail
#define VIP_CHECK(expr) ::vip_detail::report((expr), #expr, __FILE__, __LINE__)

// thrust::device_vector overloads (the reference's BTF Impl takes them, src/bilateral_texture_filter_impl.cuh:19-28) are
// only declared for translation units compiled by nvcc, where <thrust/device_vector.h> is usable
#if defined(__CUDACC__) && defined(__has_include)
#if __has_include(<thrust/device_vector.h>)
#include <thrust/device_vector.h>
#define VIP_HAVE_THRUST 1
#endif
#endif

#endif  // VIP_IMPL_COMMON_HPP
