// Internal declarations shared by the kernel translation units and the C-ABI layer (capi.cu).
// Nothing here is visible outside libvip_b200.so.
#ifndef VIP_INTERNAL_H
#define VIP_INTERNAL_H

#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>

#include <cstdlib>

namespace vip {

enum StencilMode { kBF = 0, kJBF = 1, kABF = 2 };

// Programmatic dependent launch (sm_90+).  Kernels launched through launch_pdl() may be SCHEDULED while the previous
// kernel of the stream drains (launch latency and constant-table prologues overlap its tail); each of them executes
// pdl_wait() before it touches any image buffer, which blocks until the previous grid has completed and flushed, and
// pdl_release() to let its own successor be scheduled early.  Both are no-ops in a normally launched grid.
// VIP_PDL=0 in the environment launches everything the classic way (A/B measurements).
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_release() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

inline bool pdl_enabled() {
    static const bool on = [] {
        const char* v = std::getenv("VIP_PDL");
        return !(v && v[0] == '0');
    }();
    return on;
}

template <typename... KArgs, typename... Args>
cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled() ? 1 : 0;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
    return e != cudaSuccess ? e : cudaGetLastError();
}
#endif

// Arguments of the bilateral-family stencil kernels (stencil.cu).  Row numbers are rows of the FULL image;
// the buffers hold the rows a strip describes (include/vip_c_api.h, struct vip_strip).
struct StencilArgs {
    const std::uint8_t* src;    // values (BF/ABF: also the range source)
    const std::uint8_t* guide;  // JBF: range source; otherwise unused
    std::uint8_t* dst;
    int width, height;          // full image
    int src_y0, src_rows;       // rows held by src / guide
    int dst_y0;                 // full-image row of dst's first row
    int out_y0, out_rows;       // rows to produce
    int radius;
    const float* lut;           // device: colour table (768 or 1536 floats)
    int lut_n;                  // entries replicated into shared memory
    int lut_kind;               // LutKind of stencil.cu: replication of the table and whether the index is clamped to lut_n-1
    int lut_rows;               // 1; kLutPremul: rows of the premultiplied table (distinct spatial weights), lut = [lut_rows][lut_n]
    const int* hw;              // device: 2r+1 half widths, hw[ky+r] = floor(sqrt(r^2-ky^2))
    const float* ks_full;       // device: ksize*ksize spatial kernel (the generic large-radius kernel reads it)
    const float* w_rows;        // device: ksize rows of stencil_w_pitch(radius) floats: row ky = its 2*hw+1 non-zero weights from index 0, zero padded
    // Peer halos (BF / ABF, *_strip_peer entry points): rows above src_y0 / from src_y0 + src_rows on are read straight
    // from the neighbouring ranks' buffers (peer / IPC-mapped device pointers, NVLink loads) -- nullptr: clamp into src.
    const std::uint8_t* above;  // holds rows [above_y0, above_y0 + above_rows)
    const std::uint8_t* below;  // holds rows [below_y0, below_y0 + below_rows)
    int above_y0, above_rows, below_y0, below_rows;
};

cudaError_t launch_stencil(StencilMode mode, const StencilArgs& a, int sm_count, cudaStream_t stream);
// Largest radius the tiled kernels cover; beyond it the generic one-thread-per-pixel kernel runs.
int stencil_max_tiled_radius();
// Launch configuration of the tiled kernels for a table of `lut_n` entries of kind `lut_kind` (host logic only).
void stencil_describe(StencilMode mode, int radius, int lut_n, int lut_rows, int lut_kind, int* copies, int* clamped,
                      int* groups, int* smem_bytes);
// Floats per row of StencilArgs::w_rows.
int stencil_w_pitch(int radius);
// Decides how many colour-table entries are replicated in shared memory and whether the index is clamped.
// n_ks: distinct non-zero values of the spatial kernel (rows a premultiplied table would need).
void stencil_pick_lut(StencilMode mode, int radius, int table_len, int last_nz, int n_ks, int* lut_n, int* lut_kind);
constexpr int kStencilLutPremul = 3;   // LutKind kLutPremul

// gradient.cu
cudaError_t launch_gradient_u8(const std::uint8_t* src, float* dst, int width, int height, int channels, int src_y0,
                               int src_rows, int dst_y0, int out_y0, int out_rows, cudaStream_t stream);
cudaError_t launch_gradient_f32(const float* src, float* dst, int width, int height, int channels,
                                cudaStream_t stream);

// btf.cu
cudaError_t launch_blur_rtv(const std::uint8_t* image, const float* magnitude, float* blurred, float* rtv, int width,
                            int height, int ksize, int src_y0, int src_rows, int dst_y0, int out_y0, int out_rows,
                            cudaStream_t stream);
cudaError_t launch_guide(const float* blurred, const float* rtv, std::uint8_t* guide, int width, int height, int ksize,
                         int src_y0, int src_rows, int dst_y0, int out_y0, int out_rows, cudaStream_t stream);

// integral.cu
cudaError_t launch_integral_u8c3(const std::uint8_t* src, std::uint32_t* table, int width, int height, int radius,
                                 cudaStream_t stream);
// src_type: 0 = u8, 1 = u16 (uint32 table), 2 = f32 (float table); channels 1..4
cudaError_t launch_integral_any(const void* src, void* table, int width, int height, int channels, int radius,
                                int src_type, cudaStream_t stream);
cudaError_t launch_integral_get_any(const void* table, int width, int height, int channels, int radius, int is_float,
                                    const std::int32_t* rects, int n, void* out, cudaStream_t stream);
cudaError_t launch_integral_get(const std::uint32_t* table, int width, int height, int radius,
                                const std::int32_t* rects, int n, std::uint32_t* out, cudaStream_t stream);
cudaError_t launch_integral_box_mean(const std::uint32_t* table, float* mean, int width, int height, int radius,
                                     cudaStream_t stream);

// convert.cu
cudaError_t launch_convert_to_bgr(const std::uint8_t* src, int src_channels, std::uint8_t* dst, int width, int height,
                                  int swap_rb, cudaStream_t stream);
cudaError_t launch_scale_to_u8(const float* src, std::uint8_t* dst, size_t n, float max_value, cudaStream_t stream);

void count_launch(int n = 1);

}  // namespace vip

#endif
