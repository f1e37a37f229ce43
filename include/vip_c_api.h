/*
 * vip_c_api.h -- C ABI of the B200-native edge-preserving-filter library (libvip_b200.so).
 *
 * This is the drop-in boundary.  The reference (yuyuyu-bot/various_image_processings) has no
 * FFI layer of its own: its boundary is the set of C++ classes in include/cuda/ (CudaBilateralFilter,
 * CudaAdaptiveBilateralFilter, CudaBilateralTextureFilter, DeviceImage<T>, cuda_gradient<T>).  Those classes
 * are re-provided, with the same names and signatures, by include/cuda/ (hpp) in this repository and are
 * implemented in plain C++ on top of the entry points below; every entry point cites the reference interface
 * it replaces.  The library is linked against the CUDA runtime only: no torch types, no C++ types, no
 * exceptions cross this boundary.
 *
 * Conventions
 *   - every function returns 0 on success, a positive cudaError_t value when the CUDA runtime failed, or a
 *     negative VIP_E_* code for argument errors.  Nothing is printed, nothing throws.  (The reference prints
 *     cudaGetLastError() to stderr and carries on -- src/host_utilities.hpp:7-13; the C++ wrappers in
 *     include/cuda keep that behaviour on top of these codes.)
 *   - images are tightly packed, interleaved 8-bit BGR (3 bytes / pixel, pitch = 3*width), like the
 *     reference (src/bilateral_filter_impl.cu:20); borders replicate (clamp), (:50-51).
 *   - `d_` pointers are device pointers on the CURRENT device; `stream` is a cudaStream_t passed as void*
 *     (NULL = legacy default stream).  All run calls are asynchronous with respect to the host; the C++
 *     wrappers add the cudaDeviceSynchronize() the reference performs after every public call.
 *   - in-place operation (d_src == d_dst) is not supported by the bilateral / adaptive bilateral filters (same as the
 *     reference: neighbours are read while outputs are written); the bilateral texture filter accepts it, like the
 *     reference, whose execute() copies the source into its own buffers first.
 *   - ksize must be odd and >= 1 (the reference silently misbehaves for even ksize, SURVEY.md 9.4;
 *     VIP_E_KSIZE is returned instead).
 *
 * Row strips (the multi-GPU driver): every filter has a *_strip variant.  The filter object is
 * created with the FULL image size; a strip call describes which rows the device buffers hold and which rows
 * are to be produced (struct vip_strip).  Coordinates are clamped against the full image only, so strip
 * results are bit-identical to the corresponding rows of a whole-image call.
 */
#ifndef VIP_C_API_H
#define VIP_C_API_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define VIP_API __attribute__((visibility("default")))

/* ---- status codes ------------------------------------------------------------------------------------ */
#define VIP_OK 0
#define VIP_E_NULL (-1)      /* NULL handle / pointer                                   */
#define VIP_E_SIZE (-2)      /* width/height/channels/rows out of range                 */
#define VIP_E_KSIZE (-3)     /* ksize even or < 1                                       */
#define VIP_E_SIGMA (-4)     /* sigma <= 0 or not finite                                */
#define VIP_E_STRIP (-5)     /* strip rows outside the buffers / missing halo rows      */
#define VIP_E_ALLOC (-6)     /* host allocation failed                                  */
#define VIP_E_UNSUPPORTED (-7)
#define VIP_E_DEVICE (-8)    /* multi-GPU: bad device list / no peer access             */

/* How the spatial / range lookup tables are built:
 *   VIP_LUT_CUDA : float coefficients + expf   -- src/bilateral_filter_impl.cu:217-238 (default; bit parity
 *                  with the reference's CUDA classes)
 *   VIP_LUT_CPU  : double coefficients + exp, rounded to float -- include/cpp/bilateral_filter.hpp:12-37 */
#define VIP_LUT_CUDA 0
#define VIP_LUT_CPU 1

/* Which rows a strip call reads and writes.  All row numbers are rows of the FULL image.
 *   the source (and guide) buffers hold rows [src_y0, src_y0 + src_rows)
 *   the destination buffer's first row is row dst_y0 of the full image
 *   rows [out_y0, out_y0 + out_rows) are produced
 * The source rows must cover [out_y0 - halo, out_y0 + out_rows + halo) intersected with the image, where halo
 * is vip_*_halo_rows() of the filter. */
typedef struct vip_strip {
    int src_y0;
    int src_rows;
    int dst_y0;
    int out_y0;
    int out_rows;
} vip_strip;

/* Where the halo rows of a strip come from when the strip buffer holds only the rank's OWN rows (one process per
 * GPU): device pointers into the buffers of the ranks above / below -- peer pointers of the same process or pointers
 * opened with vip_ipc_open() -- and the image rows those buffers hold.  NULL = no neighbour on that side (the image
 * border: rows clamp).  The *_strip_peer entry points stage these rows straight from the neighbour's memory inside the
 * filter kernel (NVLink loads): no halo copy, no halo rows in the local buffer. */
typedef struct vip_halo {
    const uint8_t* d_above;
    int above_y0, above_rows;
    const uint8_t* d_below;
    int below_y0, below_rows;
} vip_halo;

/* ---- library ------------------------------------------------------------------------------------------ */
VIP_API const char* vip_version(void);
/* Human-readable text for a status returned by any function of this header. */
VIP_API const char* vip_status_string(int status);
/* Number of kernels this library has launched in the calling process (all threads), for benchmark
 * accounting ("gpu_launches"). */
VIP_API unsigned long long vip_launch_count(void);
/* How a bilateral-family filter with these parameters will run on the device -- pure host logic, no CUDA call, so
 * it also answers on a machine without a GPU (the tests sweep it to prove that every valid parameter set has a launch
 * configuration).  mode: 0 bilateral, 1 joint bilateral, 2 adaptive bilateral. */
typedef struct vip_stencil_plan {
    int tiled;        /* 1: persistent tiled kernel (radius <= 15); 0: one-thread-per-pixel kernel                 */
    int lut_entries;  /* range-table entries kept in shared memory (the all-zero tail may be cut)                  */
    int lut_copies;   /* copies of every entry: 32 (bank-conflict free for any content) or 8                       */
    int lut_clamped;  /* 1: the index is clamped to lut_entries - 1                                                */
    int warp_groups;  /* 256-thread warp groups per CTA (one CTA per SM)                                           */
    int smem_bytes;   /* dynamic shared memory per CTA                                                             */
    int lut_rows;     /* 1, or the rows of a PREMULTIPLIED table: one row RN(ks * lut[i]) per distinct spatial weight, */
                      /* which takes the multiply out of every tap (short tables only: sigma_color of a few levels)   */
} vip_stencil_plan;
VIP_API int vip_stencil_plan_query(int mode, int ksize, float sigma_space, float sigma_color, vip_stencil_plan* out);

/* ---- device image container ---------------------------------------------------------------------------
 * replaces DeviceImage<ElemType> (include/cuda/device_image.hpp:4-16, src/device_image.cu:5-52).
 * elem_size is 1 (uint8_t) or 4 (float).  upload/download copy exactly width*height*channels elements;
 * the plain versions block like the reference's thrust::copy, the _async versions enqueue on `stream`
 * (use pinned host memory for real overlap: vip_host_alloc). */
typedef struct vip_image vip_image;
VIP_API int vip_image_create(vip_image** out, int width, int height, int channels, int elem_size);
VIP_API int vip_image_destroy(vip_image* img);
VIP_API int vip_image_upload(vip_image* img, const void* host_data);
VIP_API int vip_image_download(const vip_image* img, void* host_data);
VIP_API int vip_image_upload_async(vip_image* img, const void* host_data, void* stream);
VIP_API int vip_image_download_async(const vip_image* img, void* host_data, void* stream);
/* Whole rows [row0, row0 + rows) of the image <-> a packed host buffer of rows * width * channels elements (the own rows
 * of a strip buffer that also has room for halo rows). */
VIP_API int vip_image_upload_rows_async(vip_image* img, const void* host_rows, int row0, int rows, void* stream);
VIP_API int vip_image_download_rows_async(const vip_image* img, void* host_rows, int row0, int rows, void* stream);
VIP_API void* vip_image_data(const vip_image* img);
VIP_API size_t vip_image_bytes(const vip_image* img);
/* pinned host memory for the async copies */
VIP_API int vip_host_alloc(void** out, size_t bytes);
VIP_API int vip_host_free(void* p);
VIP_API int vip_stream_synchronize(void* stream);
VIP_API int vip_device_synchronize(void);

/* ---- bilateral / joint bilateral filter ---------------------------------------------------------------
 * replaces CudaBilateralFilter (include/cuda/bilateral_filter.hpp:7-29):
 *   ctor -> CudaBilateralFilter::Impl::Impl, LUT precompute   src/bilateral_filter_impl.cu:204-239
 *   vip_bf_run  -> Impl::bilateral_filter + kernel            src/bilateral_filter_impl.cu:7-96, 241-258
 *   vip_jbf_run -> Impl::joint_bilateral_filter + kernel      src/bilateral_filter_impl.cu:98-202, 260-280 */
typedef struct vip_bf vip_bf;
VIP_API int vip_bf_create(vip_bf** out, int width, int height, int ksize, float sigma_space, float sigma_color);
VIP_API int vip_bf_create_ex(vip_bf** out, int width, int height, int ksize, float sigma_space, float sigma_color,
                             int lut_mode);
VIP_API int vip_bf_destroy(vip_bf* f);
VIP_API int vip_bf_run(const vip_bf* f, const uint8_t* d_src, uint8_t* d_dst, void* stream);
VIP_API int vip_jbf_run(const vip_bf* f, const uint8_t* d_src, const uint8_t* d_guide, uint8_t* d_dst, void* stream);
VIP_API int vip_bf_run_strip(const vip_bf* f, const uint8_t* d_src, uint8_t* d_dst, const vip_strip* strip,
                             void* stream);
VIP_API int vip_jbf_run_strip(const vip_bf* f, const uint8_t* d_src, const uint8_t* d_guide, uint8_t* d_dst,
                              const vip_strip* strip, void* stream);
/* Strip whose halo rows are read from the neighbouring ranks' device memory (see struct vip_halo).  src rows + halo
 * rows together must cover [out_y0 - radius, out_y0 + out_rows + radius) clipped to the image. */
VIP_API int vip_bf_run_strip_peer(const vip_bf* f, const uint8_t* d_src, const vip_halo* halo, uint8_t* d_dst,
                                  const vip_strip* strip, void* stream);
VIP_API int vip_bf_halo_rows(const vip_bf* f); /* = ksize/2 */
/* Copies of the tables the object holds (for tests): kernel_space is ksize*ksize floats, color_table 768. */
VIP_API int vip_bf_get_luts(const vip_bf* f, float* kernel_space, float* color_table);

/* ---- adaptive bilateral filter ------------------------------------------------------------------------
 * replaces CudaAdaptiveBilateralFilter (include/cuda/adaptive_bilateral_filter.hpp:7-24):
 *   ctor -> Impl::Impl (1536-entry colour table)              src/adaptive_bilateral_filter_impl.cu:117-152
 *   vip_abf_run -> Impl::execute + kernel                     src/adaptive_bilateral_filter_impl.cu:7-115, 154-171 */
typedef struct vip_abf vip_abf;
VIP_API int vip_abf_create(vip_abf** out, int width, int height, int ksize, float sigma_space, float sigma_color);
VIP_API int vip_abf_create_ex(vip_abf** out, int width, int height, int ksize, float sigma_space, float sigma_color,
                              int lut_mode);
VIP_API int vip_abf_destroy(vip_abf* f);
VIP_API int vip_abf_run(const vip_abf* f, const uint8_t* d_src, uint8_t* d_dst, void* stream);
VIP_API int vip_abf_run_strip(const vip_abf* f, const uint8_t* d_src, uint8_t* d_dst, const vip_strip* strip,
                              void* stream);
VIP_API int vip_abf_run_strip_peer(const vip_abf* f, const uint8_t* d_src, const vip_halo* halo, uint8_t* d_dst,
                                   const vip_strip* strip, void* stream);
VIP_API int vip_abf_halo_rows(const vip_abf* f); /* = ksize/2 */

/* ---- gradient magnitude -------------------------------------------------------------------------------
 * replaces cuda_gradient<T> / cuda_gradient_impl<T> (include/cuda/gradient.hpp:4-23, src/gradient_impl.cu:7-112).
 * channels is 1 or 3 in the reference's tests; any channels >= 1 is accepted.  dst is float [height][width].
 * Like the reference, these return right after the launch. */
VIP_API int vip_gradient_u8(const uint8_t* d_src, float* d_dst, int width, int height, int channels, void* stream);
VIP_API int vip_gradient_f32(const float* d_src, float* d_dst, int width, int height, int channels, void* stream);
VIP_API int vip_gradient_u8_strip(const uint8_t* d_src, float* d_dst, int width, int height, int channels,
                                  const vip_strip* strip, void* stream);

/* ---- host-I/O edge: layout conversion ------------------------------------------------------------------
 * vip_convert_to_bgr_u8: device twin of convert_to_3ch (sample/benchmark/main.cpp:91-103): src_channels 1 (GRAY2BGR:
 * the channel is replicated), 3 (copy) or 4 (BGRA2BGR: alpha dropped) -> packed 3-channel.  swap_rb != 0 also swaps
 * channels 0 and 2 (a PNG / PPM decoder delivers RGB(A), the filters' convention -- cv::imread -- is BGR).
 * vip_scale_to_u8_f32: dst = saturate_u8(round_half_even(src * 255 / max_value)), the convert_to_u8 of
 * sample/gradient/main.cpp:9-18 used to write a gradient magnitude as an 8-bit image. */
VIP_API int vip_convert_to_bgr_u8(const uint8_t* d_src, int src_channels, uint8_t* d_dst, int width, int height,
                                  int swap_rb, void* stream);
VIP_API int vip_scale_to_u8_f32(const float* d_src, uint8_t* d_dst, size_t n, float max_value, void* stream);

/* ---- border-replicated integral image -----------------------------------------------------------------
 * GPU twin of BorderReplicatedIntegralImage<uint8_t,3> (include/cpp/border_replicated_integral_image.hpp:6-85;
 * the reference has this on the CPU only).  The table is (height+2r+1) x (width+2r+1) x 3 uint32, row 0 and
 * column 0 zero, cell (Y,X) = sum over the clamp-padded image rows < Y, columns < X.  Arithmetic is modulo
 * 2^32, so box sums stay exact for any image size.
 * vip_integral_get_u8c3 evaluates get(x0,y0,x1,y1) (inclusive box sums, :34-47) for n rectangles given as
 * n x 4 int32 (x0,y0,x1,y1) on the device; out is n x 3 uint32. */
VIP_API int vip_integral_u8c3(const uint8_t* d_src, uint32_t* d_table, int width, int height, int radius,
                              void* stream);
VIP_API int vip_integral_get_u8c3(const uint32_t* d_table, int width, int height, int radius, const int32_t* d_rects,
                                  int n, uint32_t* d_out, void* stream);
VIP_API size_t vip_integral_table_elems(int width, int height, int radius);
/* k x k box mean through the integral image: out[y][x][c] = float(get(x-r,y-r,x+r,y+r)[c]) / (k*k), the
 * quantity include/cpp/adaptive_bilateral_filter.hpp:53-57 reads from the table. */
VIP_API int vip_integral_box_mean_u8c3(const uint32_t* d_table, float* d_mean, int width, int height, int radius,
                                       void* stream);

/* vip_integral_u8c3_noalloc -- no counterpart in the reference; it exists for hosts that must not allocate.
 *
 * vip_integral_u8c3 above borrows its aggregate arrays (about a third of the table's size) from a CUDA memory pool on
 * every call.  Some callers cannot afford that: a plug-in inside an application that has already given the whole GPU to
 * its own allocator (the pool request then fails or steals memory the host did not budget for), a real-time or
 * safety-reviewed loop where no driver allocation may happen after start-up, a process under an outer memory
 * accountant.  This variant computes the identical table (bit for bit) and allocates NOTHING, neither device nor
 * pinned memory, directly or through a library: the aggregates live in a fixed array of
 * VIP_INTEGRAL_NOALLOC_SCRATCH_BYTES that is part of libvip_b200.so's device image.  You still pass the source
 * image and the table (vip_integral_table_elems() uint32 elements), allocated however your host allows.
 *
 * Cost: that array occupies device memory on every GPU that loads the library's kernels, until the process exits,
 * even if this function is never called.
 * Limit: with tw = width+2r+1, th = height+2r+1 the call needs
 *     (ceil(th/8) + ceil(th/32)) * tw * 3  +  2 * ceil(tw/256) * 3 * th      32-bit words;
 * more than VIP_INTEGRAL_NOALLOC_SCRATCH_BYTES / 4 returns VIP_E_SIZE before anything is launched (3840x2160 fits
 * up to r = 15 with room to spare; 7680x4320 does not fit -- use vip_integral_u8c3 there).
 * First call on a device: looks up the array's address on the current device (cudaGetSymbolAddress, once per
 * device, cached).  Make that first call outside any stream capture, with the device you will use current.
 * Sharing: every call on a device uses the same array.  Calls are safe back to back on one stream (nothing needs
 * clearing: each call writes every word before it reads it), but two calls that can overlap on the GPU -- different
 * streams without an event between them, or two host threads -- race and produce garbage tables.  Serialise them. */
#define VIP_INTEGRAL_NOALLOC_SCRATCH_BYTES (20u << 20)
VIP_API int vip_integral_u8c3_noalloc(const uint8_t* d_src, uint32_t* d_table, int width, int height, int radius,
                                      void* stream);

/* Any source type / channel count of the reference template (BorderReplicatedIntegralImage<SrcType, Channels>,
 * instantiated by its tests for uint8_t / uint16_t / float x 1..3 channels,
 * test/border_replicated_integral_image.cpp:63-95).  src_type VIP_SRC_U8 / VIP_SRC_U16: uint32 table (modulo 2^32);
 * VIP_SRC_F32: float table.  channels 1..4.  The table has (height+2r+1) x (width+2r+1) x channels elements.
 * vip_integral_get: n rectangles (x0,y0,x1,y1) -> n x channels sums (uint32 or float like the table). */
#define VIP_SRC_U8 0
#define VIP_SRC_U16 1
#define VIP_SRC_F32 2
VIP_API int vip_integral_build(const void* d_src, void* d_table, int width, int height, int channels, int radius,
                               int src_type, void* stream);
VIP_API int vip_integral_get(const void* d_table, int width, int height, int channels, int radius, int src_type,
                             const int32_t* d_rects, int n, void* d_out, void* stream);

/* ---- bilateral texture filter -------------------------------------------------------------------------
 * replaces CudaBilateralTextureFilter (include/cuda/bilateral_texture_filter.hpp:7-17):
 *   ctor/execute -> Impl::Impl / Impl::execute                src/bilateral_texture_filter_impl.cu:179-214
 *   vip_btf_blur_rtv -> Impl::compute_blur_and_rtv + kernel   src/bilateral_texture_filter_impl.cu:10-104, 216-237
 *   vip_btf_guide    -> Impl::compute_guide + kernel          src/bilateral_texture_filter_impl.cu:106-177, 239-256
 * The joint bilateral step uses (ksize 2k-1, sigma_space k-1, sigma_color sqrt(3)) like the reference (:188);
 * ksize 1 would make sigma_space 0 (the reference divides by zero building its table): VIP_E_KSIZE, ksize >= 3.
 * The object owns its scratch buffers; a run is not re-entrant on the same object (same as the reference). */
typedef struct vip_btf vip_btf;
VIP_API int vip_btf_create(vip_btf** out, int width, int height, int ksize, int nitr);
VIP_API int vip_btf_destroy(vip_btf* f);
VIP_API int vip_btf_run(vip_btf* f, const uint8_t* d_src, uint8_t* d_dst, void* stream);
/* Replay the 3 x nitr launches of vip_btf_run as ONE captured CUDA graph per (d_src, d_dst) pair (up to 32 pairs are
 * kept).  Off by default; only on non-default streams.  Results are identical either way. */
VIP_API int vip_btf_set_graph(vip_btf* f, int on);
VIP_API int vip_btf_blur_rtv(const vip_btf* f, const uint8_t* d_image, const float* d_magnitude, float* d_blurred,
                             float* d_rtv, void* stream);
VIP_API int vip_btf_guide(const vip_btf* f, const float* d_blurred, const float* d_rtv, uint8_t* d_guide,
                          void* stream);
/* Rows of halo one iteration needs on each side: (k-1) + k/2 + k/2 + 1. */
VIP_API int vip_btf_halo_rows_per_iteration(const vip_btf* f);
/* One iteration on a strip: the buffers hold rows [src_y0, src_y0+src_rows); rows [out_y0, out_y0+out_rows)
 * of the next iterate are written.  The object must have been created with the FULL image size. */
VIP_API int vip_btf_iterate_strip(vip_btf* f, const uint8_t* d_src_n, uint8_t* d_dst_n, const vip_strip* strip,
                                  void* stream);
VIP_API size_t vip_btf_scratch_bytes(const vip_btf* f);

/* ---- multi-GPU driver (new layer) --------------------------------------------------------
 * One host process drives n devices of one box.  Large images are cut into contiguous row strips, one per
 * device, each with a halo of vip_*_halo_rows() rows; halos move by peer-to-peer copies (NVLink) between
 * neighbouring devices; frames of a batch are dealt round-robin to the devices without communication.
 * kind: VIP_MG_BF / VIP_MG_ABF / VIP_MG_BTF.  For BF/ABF `nitr` is ignored. */
#define VIP_MG_BF 0
#define VIP_MG_ABF 1
#define VIP_MG_BTF 2
typedef struct vip_mg vip_mg;
VIP_API int vip_mg_create(vip_mg** out, const int* device_ids, int n_devices, int kind, int width, int height,
                          int ksize, float sigma_space, float sigma_color, int nitr);
/* Same with flags.  BTF strips get the 17 (k = 9) rows of the previous iterate every iteration needs either from
 * buffers that carry nitr x 17 halo rows from the scatter, recomputing the shrinking overlap redundantly (no
 * communication, VIP_MG_FLAG_BTF_DEEP_HALO), or by a peer-to-peer halo exchange after every iteration
 * (VIP_MG_FLAG_BTF_EXCHANGE).  0 = automatic: deep halo while the redundant rows are <= 35 % of a strip.  The
 * results are bit-identical either way. */
#define VIP_MG_FLAG_BTF_EXCHANGE 1
#define VIP_MG_FLAG_BTF_DEEP_HALO 2
/* Fail with VIP_E_DEVICE when two neighbouring slots sit on different devices without peer access (by default
 * cudaMemcpyPeerAsync then stages the halo rows through the host; vip_mg_peer_access() reports which one a handle uses) */
#define VIP_MG_FLAG_REQUIRE_P2P 4
VIP_API int vip_mg_create_ex(vip_mg** out, const int* device_ids, int n_devices, int kind, int width, int height,
                             int ksize, float sigma_space, float sigma_color, int nitr, int flags);
VIP_API int vip_mg_destroy(vip_mg* m);
/* Row range [*y0, *y0 + *rows) owned by device slot `i`. */
VIP_API int vip_mg_strip_rows(const vip_mg* m, int i, int* y0, int* rows);
/* The strip plan vip_mg_create_ex would make, as pure host logic (no CUDA call, no devices needed): what slot `i`
 * of `n_devices` owns, what its buffers hold, and for BTF which halo mode was chosen. */
typedef struct vip_mg_plan {
    int y0, rows;          /* rows owned (and produced) by the slot                                               */
    int buf_y0, buf_rows;  /* rows held by its strip buffers: owned rows + halo rows on each side, clipped         */
    int halo;              /* halo rows ONE pass needs on each side: radius, or (k-1) + 2 (k/2) + 1 per BTF iteration */
    int btf_deep_halo;     /* BTF: 1 = the buffers carry nitr x halo rows and the overlap is recomputed, 0 = exchange */
} vip_mg_plan;
VIP_API int vip_mg_plan_query(int n_devices, int slot, int kind, int height, int ksize, int nitr, int flags,
                              vip_mg_plan* out);
/* Scatter a full host image (pinned or pageable) into the per-device strip buffers (own rows + halos). */
VIP_API int vip_mg_scatter_host(vip_mg* m, const uint8_t* h_src);
/* Scatter from a full image resident on device slot 0 using peer copies. */
VIP_API int vip_mg_scatter_device(vip_mg* m, const uint8_t* d_src_on_dev0);
/* Run the filter on every strip (all devices concurrently).  For BTF the iterations exchange
 * halo rows between neighbouring devices with peer copies.  Blocks until all devices are done. */
VIP_API int vip_mg_run(vip_mg* m);
/* Same, but returns after enqueueing; vip_mg_wait() blocks. */
VIP_API int vip_mg_run_async(vip_mg* m);
VIP_API int vip_mg_wait(vip_mg* m);
VIP_API int vip_mg_gather_host(vip_mg* m, uint8_t* h_dst);
/* Device time of the last run, max over devices, from CUDA events (ms). */
VIP_API int vip_mg_last_run_ms(const vip_mg* m, float* ms);
/* Independent frames: frame f goes to device slot f % n.  h_src/h_dst hold n_frames packed images.  Uploads, kernels
 * and downloads of different frames overlap on two streams per device -- when the host memory is PINNED (vip_host_alloc
 * or cudaHostRegister); with pageable memory the copies serialise.  One enqueueing host thread per device. */
VIP_API int vip_mg_run_batch_host(vip_mg* m, const uint8_t* h_src, uint8_t* h_dst, int n_frames);
/* The same through a SET of filters (BASELINE configs[4]: every frame through all three): each frame is uploaded once,
 * filter j of the set writes h_dst[j] (entries of filters not in the set may be NULL).  The filters share the handle's
 * ksize / sigmas / nitr; whatever its own kind, the missing filter objects are created on first use. */
#define VIP_MG_SET_BF 1
#define VIP_MG_SET_ABF 2
#define VIP_MG_SET_BTF 4
VIP_API int vip_mg_run_batch_set_host(vip_mg* m, int filter_set, const uint8_t* h_src, uint8_t* const* h_dst /* [3] */,
                                      int n_frames);
/* 1: every pair of neighbouring slots on different devices has peer access enabled (halo rows move GPU<->GPU over
 * NVLink); 0: at least one pair stages through the host. */
VIP_API int vip_mg_peer_access(const vip_mg* m);

/* ---- one process per GPU: cross-process device memory and neighbour ordering (csrc/link.cu) -------------------
 * Under torchrun every rank owns one GPU and one row strip.  vip_ipc_export() turns a device allocation (the base
 * pointer of a vip_image, or of any cudaMalloc'ed buffer) into a 64-byte handle that another PROCESS of the same box
 * opens with vip_ipc_open(); the pointer it gets is an ordinary device pointer there (peer-to-peer over NVLink when the
 * two ranks sit on different GPUs).  VIP_E_DEVICE: the two devices have no peer path -- pick another transport, nothing
 * falls back silently.  vip_copy_rows() is the explicit halo pull (one asynchronous copy of whole rows). */
#define VIP_IPC_HANDLE_BYTES 64
VIP_API int vip_ipc_export(const void* d_ptr, unsigned char* handle /* [VIP_IPC_HANDLE_BYTES] */);
VIP_API int vip_ipc_open(const unsigned char* handle, void** d_ptr);
VIP_API int vip_ipc_close(void* d_ptr);
VIP_API int vip_peer_can_access(int device, int peer_device); /* 1 / 0 */
VIP_API int vip_copy_rows(uint8_t* d_dst, int dst_y0, const uint8_t* d_src, int src_y0, int y0, int rows, int width,
                          void* stream);
/* Ordering between a rank and its two strip neighbours without the host: step numbers stored into the neighbour's flag
 * block by a one-thread kernel, waits polling the rank's own block -- all enqueued on streams.
 *   ready(slot, step)     "my own rows of `step` are in buffer `slot`"          (after the upload / the producing kernel)
 *   consumed(slot, step)  "I have finished reading your rows of `step`"         (after the filter kernel)
 * A frame pipeline with n_slots buffers in flight runs, per step s on slot s % n_slots:
 *   wait_consumed(slot, s - n_slots) ; upload own rows ; signal_ready(slot, s) ; wait_ready(slot, s) ;
 *   filter (halo rows from the neighbours) ; signal_consumed(slot, s) ; download.
 * Steps count from 1.  A wait that sees no progress for the timeout (default 2000 ms of device time) gives up and
 * latches an error that vip_link_status() returns as VIP_E_DEVICE. */
typedef struct vip_link vip_link;
VIP_API int vip_link_create(vip_link** out, int n_slots);
VIP_API int vip_link_destroy(vip_link* l);
VIP_API int vip_link_export(const vip_link* l, unsigned char* handle /* [VIP_IPC_HANDLE_BYTES] */);
/* handles of the ranks above / below (NULL at the image border), as exported by THEIR vip_link_export */
VIP_API int vip_link_connect(vip_link* l, const unsigned char* above_handle, const unsigned char* below_handle);
/* the same inside one process (several ranks of a test, or devices driven by one process) */
VIP_API int vip_link_connect_local(vip_link* l, vip_link* above, vip_link* below);
VIP_API int vip_link_set_timeout_ms(vip_link* l, int timeout_ms);
VIP_API int vip_link_signal_ready(vip_link* l, int slot, unsigned step, void* stream);
VIP_API int vip_link_wait_ready(vip_link* l, int slot, unsigned step, void* stream);
VIP_API int vip_link_signal_consumed(vip_link* l, int slot, unsigned step, void* stream);
VIP_API int vip_link_wait_consumed(vip_link* l, int slot, unsigned step, void* stream);
/* vip_link_status: VIP_OK, or VIP_E_DEVICE once any wait of this link has timed out (synchronises the device).
 * A timed-out wait does NOT stop the stream: the filter and the download enqueued behind it still run, on halo rows
 * that are stale or missing.  So a frame produced through a link is only good if vip_link_status() is VIP_OK AFTER the
 * frame's work has completed -- check it before the frame is consumed (per frame, or per batch of frames that is
 * dropped as a whole).  The error stays latched until vip_link_reset().
 * Same-device links (vip_link_connect_local on one GPU, or two processes sharing a GPU): the waits are spin kernels and
 * need the signalling kernel of the other stream / process to run concurrently, which CUDA does not promise; that set-up
 * is for functional tests, and the time-out is what bounds it.
 * vip_link_reset: clears the latch (after synchronising the device); step counters are kept. */
VIP_API int vip_link_status(vip_link* l);
VIP_API int vip_link_reset(vip_link* l);

#ifdef __cplusplus
}
#endif
#endif /* VIP_C_API_H */
