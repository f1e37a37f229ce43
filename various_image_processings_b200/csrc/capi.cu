// C-ABI layer of libvip_b200.so (include/vip_c_api.h): handles, host-side LUT construction, launch
// orchestration.  Host logic only -- the kernels live in stencil.cu / gradient.cu / btf.cu / integral.cu.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cmath>
#include <cstdlib>
#include <cstring>
#include <new>
#include <vector>

#include "../../include/vip_c_api.h"
#include "vip_internal.h"

namespace vip {
static std::atomic<unsigned long long> g_launches{0};
void count_launch(int n) { g_launches.fetch_add(static_cast<unsigned long long>(n), std::memory_order_relaxed); }
}  // namespace vip

namespace {

inline int cu(cudaError_t e) { return static_cast<int>(e); }

struct DeviceBuf {
    void* p = nullptr;
    size_t bytes = 0;
    cudaError_t alloc(size_t n) {
        release();
        if (n == 0) return cudaSuccess;
        cudaError_t e = cudaMalloc(&p, n);
        if (e == cudaSuccess) bytes = n;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        bytes = 0;
    }
    ~DeviceBuf() { release(); }
    DeviceBuf() = default;
    DeviceBuf(const DeviceBuf&) = delete;
    DeviceBuf& operator=(const DeviceBuf&) = delete;
};

// Tables shared by BF / JBF / ABF objects.
struct StencilTables {
    int width = 0, height = 0, ksize = 0, radius = 0;
    float sigma_space = 0.f, sigma_color = 0.f;
    int table_len = 0;  // 768 (BF/JBF) or 1536 (ABF)
    int last_nz = 0;    // last non-zero entry of the colour table
    int device = 0, sm_count = 0;
    std::vector<float> h_ks, h_lut;
    DeviceBuf d_lut, d_hw, d_ks, d_wrows;
    // premultiplied table (stencil.cu kLutPremul): one row of pm_n entries RN(ks * lut[i]) per distinct spatial weight,
    // and the weight rows with every tap's row offset (bytes in the 32x replicated shared-memory table) as int bits
    int n_ks = 0, pm_n = 0;
    DeviceBuf d_lut_pm, d_wrows_pm;
};

int validate(int width, int height, int ksize, float ss, float sc) {
    if (width <= 0 || height <= 0 || static_cast<long long>(width) * height > (1ll << 31) / 3) return VIP_E_SIZE;
    if (ksize < 1 || (ksize & 1) == 0) return VIP_E_KSIZE;
    if (!(ss > 0.f) || !(sc > 0.f) || !std::isfinite(ss) || !std::isfinite(sc)) return VIP_E_SIGMA;
    return VIP_OK;
}

// Spatial kernel (circular mask) and colour table.
//   VIP_LUT_CUDA: src/bilateral_filter_impl.cu:217-238 (float coefficients, std::exp(float) == expf)
//   VIP_LUT_CPU : include/cpp/bilateral_filter.hpp:12-37 (double coefficients, exp(double) rounded to float)
void build_luts(int ksize, float ss, float sc, int table_len, int mode, std::vector<float>& ks, std::vector<float>& lut) {
    const int r = ksize / 2;
    ks.assign(static_cast<size_t>(ksize) * ksize, 0.f);
    lut.assign(table_len, 0.f);
    if (mode == VIP_LUT_CPU) {
        const double cc = -1. / (2 * sc * sc);
        const double cs = -1. / (2 * ss * ss);
        for (int ky = -r; ky <= r; ky++)
            for (int kx = -r; kx <= r; kx++) {
                const int r2 = kx * kx + ky * ky;
                if (r2 > r * r) continue;
                ks[(ky + r) * ksize + (kx + r)] = static_cast<float>(std::exp(r2 * cs));
            }
        for (int i = 0; i < table_len; i++) lut[i] = static_cast<float>(std::exp((i * i) * cc));
    } else {
        const float cc = -1.f / (2 * sc * sc);
        const float cs = -1.f / (2 * ss * ss);
        for (int ky = -r; ky <= r; ky++)
            for (int kx = -r; kx <= r; kx++) {
                const int r2 = kx * kx + ky * ky;
                if (r2 > r * r) continue;
                ks[(ky + r) * ksize + (kx + r)] = std::exp(r2 * cs);
            }
        for (int i = 0; i < table_len; i++) lut[i] = std::exp((i * i) * cc);
    }
}

int init_tables(StencilTables& T, int width, int height, int ksize, float ss, float sc, int table_len, int lut_mode) {
    const int v = validate(width, height, ksize, ss, sc);
    if (v != VIP_OK) return v;
    if (lut_mode != VIP_LUT_CUDA && lut_mode != VIP_LUT_CPU) return VIP_E_UNSUPPORTED;
    T.width = width;
    T.height = height;
    T.ksize = ksize;
    T.radius = ksize / 2;
    T.sigma_space = ss;
    T.sigma_color = sc;
    T.table_len = table_len;
    cudaError_t e = cudaGetDevice(&T.device);
    if (e != cudaSuccess) return cu(e);
    e = cudaDeviceGetAttribute(&T.sm_count, cudaDevAttrMultiProcessorCount, T.device);
    if (e != cudaSuccess) return cu(e);
    build_luts(ksize, ss, sc, table_len, lut_mode, T.h_ks, T.h_lut);

    const int r = T.radius;
    std::vector<int> hw(ksize);
    for (int ky = -r; ky <= r; ky++) {
        int h = 0;
        while ((h + 1) * (h + 1) + ky * ky <= r * r) h++;
        hw[ky + r] = h;
    }
    // Trailing zeros of the colour table: every index > last_nz has weight exactly 0, so clamping the index
    // to last_nz+1 is bit-neutral and lets the 32x replicated table shrink (vip::stencil_pick_lut decides).
    T.last_nz = table_len - 1;
    while (T.last_nz > 0 && T.h_lut[T.last_nz] == 0.f) T.last_nz--;
    if ((e = T.d_lut.alloc(table_len * sizeof(float))) != cudaSuccess) return cu(e);
    if ((e = T.d_hw.alloc(hw.size() * sizeof(int))) != cudaSuccess) return cu(e);
    if ((e = T.d_ks.alloc(T.h_ks.size() * sizeof(float))) != cudaSuccess) return cu(e);
    if ((e = cudaMemcpy(T.d_lut.p, T.h_lut.data(), table_len * sizeof(float), cudaMemcpyHostToDevice)) != cudaSuccess)
        return cu(e);
    if ((e = cudaMemcpy(T.d_hw.p, hw.data(), hw.size() * sizeof(int), cudaMemcpyHostToDevice)) != cudaSuccess)
        return cu(e);
    if ((e = cudaMemcpy(T.d_ks.p, T.h_ks.data(), T.h_ks.size() * sizeof(float), cudaMemcpyHostToDevice)) !=
        cudaSuccess)
        return cu(e);
    // the rows of the spatial kernel as the tiled kernels keep them in shared memory (one coalesced load per launch)
    const int wp = vip::stencil_w_pitch(r);
    std::vector<float> w_rows(static_cast<size_t>(ksize) * wp, 0.f);
    for (int row = 0; row < ksize; row++)
        for (int j = 0; j <= 2 * hw[row]; j++) w_rows[static_cast<size_t>(row) * wp + j] = T.h_ks[row * ksize + (r - hw[row] + j)];
    if ((e = T.d_wrows.alloc(w_rows.size() * sizeof(float))) != cudaSuccess) return cu(e);
    if ((e = cudaMemcpy(T.d_wrows.p, w_rows.data(), w_rows.size() * sizeof(float), cudaMemcpyHostToDevice)) != cudaSuccess)
        return cu(e);
    // distinct non-zero spatial weights; when the colour table is short enough for one premultiplied row per weight
    // (vip::stencil_pick_lut decides, per mode), build those rows: RN(ks * lut[i]) is the product the tap's FMUL rounds to
    std::vector<float> dk;
    for (float v : T.h_ks)
        if (v != 0.f && std::find(dk.begin(), dk.end(), v) == dk.end()) dk.push_back(v);
    T.n_ks = static_cast<int>(dk.size());
    bool want_pm = false;
    if (table_len == 768) {
        for (int mode = 0; mode < 2; mode++) {
            int n = 0, kind = 0;
            vip::stencil_pick_lut(static_cast<vip::StencilMode>(mode), r, table_len, T.last_nz, T.n_ks, &n, &kind);
            if (kind == vip::kStencilLutPremul) {
                // BF and JBF share ONE premultiplied table: both modes must have picked the same row length
                if (want_pm && T.pm_n != n) return VIP_E_UNSUPPORTED;
                want_pm = true;
                T.pm_n = n;
            }
        }
    }
    if (want_pm) {
        std::vector<float> pm(static_cast<size_t>(T.n_ks) * T.pm_n);
        for (int d = 0; d < T.n_ks; d++)
            for (int i = 0; i < T.pm_n; i++) {
                volatile float prod = dk[d] * T.h_lut[i];   // one IEEE single multiply, no contraction / excess precision
                pm[static_cast<size_t>(d) * T.pm_n + i] = prod;
            }
        std::vector<std::uint32_t> offs(w_rows.size(), 0u);
        for (size_t i = 0; i < w_rows.size(); i++) {
            if (w_rows[i] == 0.f) continue;
            const int d = static_cast<int>(std::find(dk.begin(), dk.end(), w_rows[i]) - dk.begin());
            offs[i] = static_cast<std::uint32_t>(d) * T.pm_n * 128u;
        }
        if ((e = T.d_lut_pm.alloc(pm.size() * sizeof(float))) != cudaSuccess) return cu(e);
        if ((e = cudaMemcpy(T.d_lut_pm.p, pm.data(), pm.size() * sizeof(float), cudaMemcpyHostToDevice)) != cudaSuccess) return cu(e);
        if ((e = T.d_wrows_pm.alloc(offs.size() * sizeof(std::uint32_t))) != cudaSuccess) return cu(e);
        if ((e = cudaMemcpy(T.d_wrows_pm.p, offs.data(), offs.size() * sizeof(std::uint32_t), cudaMemcpyHostToDevice)) != cudaSuccess)
            return cu(e);
    }
    return VIP_OK;
}

// Rows [out_y0-halo, out_y0+out_rows+halo) clipped to the image must lie inside the source buffer -- or, with peer
// halos, inside the union of the neighbour-above rows, the source buffer and the neighbour-below rows.
int check_strip(const vip_strip& s, int height, int halo, const vip_halo* peer = nullptr) {
    if (s.out_rows < 0 || s.src_rows <= 0) return VIP_E_STRIP;
    if (s.out_rows == 0) return VIP_OK;
    if (s.out_y0 < 0 || s.out_y0 + s.out_rows > height) return VIP_E_STRIP;
    if (s.dst_y0 > s.out_y0 || s.dst_y0 < 0 || s.src_y0 < 0) return VIP_E_STRIP;
    const int need0 = std::max(0, s.out_y0 - halo), need1 = std::min(height, s.out_y0 + s.out_rows + halo);
    int have0 = s.src_y0, have1 = s.src_y0 + s.src_rows;
    if (peer && peer->d_above) {
        if (peer->above_rows <= 0 || peer->above_y0 < 0 || peer->above_y0 + peer->above_rows < have0) return VIP_E_STRIP;
        have0 = std::min(have0, peer->above_y0);
    }
    if (peer && peer->d_below) {
        if (peer->below_rows <= 0 || peer->below_y0 > have1) return VIP_E_STRIP;
        have1 = std::max(have1, peer->below_y0 + peer->below_rows);
    }
    if (have0 > need0 || have1 < need1) return VIP_E_STRIP;
    // the centre pixels (and the ABF box of the outputs) are read from the local buffer: outputs must be own rows
    if (peer && (s.out_y0 < s.src_y0 || s.out_y0 + s.out_rows > s.src_y0 + s.src_rows)) return VIP_E_STRIP;
    return VIP_OK;
}

int run_stencil(const StencilTables& T, vip::StencilMode mode, const uint8_t* d_src, const uint8_t* d_guide,
                uint8_t* d_dst, const vip_strip& s, void* stream, const vip_halo* peer = nullptr) {
    if (!d_src || !d_dst || (mode == vip::kJBF && !d_guide)) return VIP_E_NULL;
    if (peer && mode == vip::kJBF) return VIP_E_UNSUPPORTED;
    const int c = check_strip(s, T.height, T.radius, peer);
    if (c != VIP_OK) return c;
    vip::StencilArgs a;
    a.above = peer ? peer->d_above : nullptr;
    a.below = peer ? peer->d_below : nullptr;
    a.above_y0 = peer ? peer->above_y0 : 0;
    a.above_rows = peer ? peer->above_rows : 0;
    a.below_y0 = peer ? peer->below_y0 : 0;
    a.below_rows = peer ? peer->below_rows : 0;
    a.src = d_src;
    a.guide = d_guide;
    a.dst = d_dst;
    a.width = T.width;
    a.height = T.height;
    a.src_y0 = s.src_y0;
    a.src_rows = s.src_rows;
    a.dst_y0 = s.dst_y0;
    a.out_y0 = s.out_y0;
    a.out_rows = s.out_rows;
    a.radius = T.radius;
    a.lut = static_cast<const float*>(T.d_lut.p);
    vip::stencil_pick_lut(mode, T.radius, T.table_len, T.last_nz, T.n_ks, &a.lut_n, &a.lut_kind);
    a.lut_rows = 1;
    a.hw = static_cast<const int*>(T.d_hw.p);
    a.ks_full = static_cast<const float*>(T.d_ks.p);
    a.w_rows = static_cast<const float*>(T.d_wrows.p);
    if (a.lut_kind == vip::kStencilLutPremul) {
        a.lut = static_cast<const float*>(T.d_lut_pm.p);
        a.w_rows = static_cast<const float*>(T.d_wrows_pm.p);
        a.lut_rows = T.n_ks;
    }
    return cu(vip::launch_stencil(mode, a, T.sm_count, static_cast<cudaStream_t>(stream)));
}

inline vip_strip whole(int height) { return vip_strip{0, height, 0, 0, height}; }

}  // namespace

struct vip_bf {
    StencilTables T;
};
struct vip_abf {
    StencilTables T;
};
struct vip_image {
    int width, height, channels, elem_size;
    DeviceBuf buf;
};
struct BtfGraph {
    const uint8_t* src;
    uint8_t* dst;
    cudaGraphExec_t exec;
};
struct vip_btf {
    int width, height, ksize, nitr, scratch_rows;
    vip_bf* jbf;
    DeviceBuf ping, pong, blurred, rtv, guide;
    // CUDA-graph replay of the 3 x nitr launches of a run: one captured graph per (src, dst) pair,
    // programmatic-dependent-launch edges included.  Off by default (see DESIGN.md 3.2 for the measured A/B);
    // vip_btf_set_graph() or VIP_BTF_GRAPH=1 switches it on.
    bool use_graph = false;
    std::vector<BtfGraph> graphs;
};

extern "C" {

const char* vip_version(void) { return "vip_b200 0.1 (sm_100a)"; }

const char* vip_status_string(int status) {
    if (status > 0) return cudaGetErrorString(static_cast<cudaError_t>(status));
    switch (status) {
        case VIP_OK: return "ok";
        case VIP_E_NULL: return "null handle or pointer";
        case VIP_E_SIZE: return "image size out of range";
        case VIP_E_KSIZE: return "ksize must be odd and >= 1";
        case VIP_E_SIGMA: return "sigma must be finite and > 0";
        case VIP_E_STRIP: return "strip rows outside the buffers or missing halo rows";
        case VIP_E_ALLOC: return "host allocation failed";
        case VIP_E_UNSUPPORTED: return "unsupported argument";
        case VIP_E_DEVICE: return "bad device list or no peer access";
        default: return "unknown status";
    }
}

unsigned long long vip_launch_count(void) { return vip::g_launches.load(std::memory_order_relaxed); }

int vip_stencil_plan_query(int mode, int ksize, float sigma_space, float sigma_color, vip_stencil_plan* out) {
    if (!out) return VIP_E_NULL;
    if (mode < 0 || mode > 2) return VIP_E_UNSUPPORTED;
    const int v = validate(1, 1, ksize, sigma_space, sigma_color);
    if (v != VIP_OK) return v;
    const int table_len = mode == 2 ? 1536 : 768;
    std::vector<float> ks, lut;
    build_luts(ksize, sigma_space, sigma_color, table_len, VIP_LUT_CUDA, ks, lut);
    int last_nz = table_len - 1;
    while (last_nz > 0 && lut[last_nz] == 0.f) last_nz--;
    const int radius = ksize / 2;
    int lut_n = table_len, kind = 0;
    std::vector<float> dk;
    for (float v : ks)
        if (v != 0.f && std::find(dk.begin(), dk.end(), v) == dk.end()) dk.push_back(v);
    vip::stencil_pick_lut(static_cast<vip::StencilMode>(mode), radius, table_len, last_nz, static_cast<int>(dk.size()), &lut_n,
                          &kind);
    std::memset(out, 0, sizeof(*out));
    out->tiled = radius <= vip::stencil_max_tiled_radius() ? 1 : 0;
    out->lut_entries = lut_n;
    out->lut_rows = kind == vip::kStencilLutPremul ? static_cast<int>(dk.size()) : 1;
    if (out->tiled)
        vip::stencil_describe(static_cast<vip::StencilMode>(mode), radius, lut_n, out->lut_rows, kind, &out->lut_copies,
                              &out->lut_clamped, &out->warp_groups, &out->smem_bytes);
    return VIP_OK;
}

/* ---------------------------------------------------------------- images */
int vip_image_create(vip_image** out, int width, int height, int channels, int elem_size) {
    if (!out) return VIP_E_NULL;
    *out = nullptr;
    if (width <= 0 || height <= 0 || channels <= 0) return VIP_E_SIZE;
    if (elem_size != 1 && elem_size != 4) return VIP_E_UNSUPPORTED;
    vip_image* img = new (std::nothrow) vip_image();
    if (!img) return VIP_E_ALLOC;
    img->width = width;
    img->height = height;
    img->channels = channels;
    img->elem_size = elem_size;
    const size_t bytes = static_cast<size_t>(width) * height * channels * elem_size;
    cudaError_t e = img->buf.alloc(bytes);
    // thrust::device_vector value-initialises (src/device_image.cu:8); keep that observable behaviour
    if (e == cudaSuccess) e = cudaMemset(img->buf.p, 0, bytes);
    if (e != cudaSuccess) {
        delete img;
        return cu(e);
    }
    *out = img;
    return VIP_OK;
}
int vip_image_destroy(vip_image* img) {
    delete img;
    return VIP_OK;
}
int vip_image_upload(vip_image* img, const void* h) {
    if (!img || !h) return VIP_E_NULL;
    return cu(cudaMemcpy(img->buf.p, h, img->buf.bytes, cudaMemcpyHostToDevice));
}
int vip_image_download(const vip_image* img, void* h) {
    if (!img || !h) return VIP_E_NULL;
    return cu(cudaMemcpy(h, img->buf.p, img->buf.bytes, cudaMemcpyDeviceToHost));
}
int vip_image_upload_async(vip_image* img, const void* h, void* stream) {
    if (!img || !h) return VIP_E_NULL;
    return cu(cudaMemcpyAsync(img->buf.p, h, img->buf.bytes, cudaMemcpyHostToDevice, static_cast<cudaStream_t>(stream)));
}
int vip_image_download_async(const vip_image* img, void* h, void* stream) {
    if (!img || !h) return VIP_E_NULL;
    return cu(cudaMemcpyAsync(h, img->buf.p, img->buf.bytes, cudaMemcpyDeviceToHost, static_cast<cudaStream_t>(stream)));
}
int vip_image_upload_rows_async(vip_image* img, const void* h, int row0, int rows, void* stream) {
    if (!img || !h) return VIP_E_NULL;
    if (row0 < 0 || rows < 0 || row0 + rows > img->height) return VIP_E_SIZE;
    const size_t rb = static_cast<size_t>(img->width) * img->channels * img->elem_size;
    return cu(cudaMemcpyAsync(static_cast<char*>(img->buf.p) + rb * row0, h, rb * rows, cudaMemcpyHostToDevice,
                              static_cast<cudaStream_t>(stream)));
}
int vip_image_download_rows_async(const vip_image* img, void* h, int row0, int rows, void* stream) {
    if (!img || !h) return VIP_E_NULL;
    if (row0 < 0 || rows < 0 || row0 + rows > img->height) return VIP_E_SIZE;
    const size_t rb = static_cast<size_t>(img->width) * img->channels * img->elem_size;
    return cu(cudaMemcpyAsync(h, static_cast<const char*>(img->buf.p) + rb * row0, rb * rows, cudaMemcpyDeviceToHost,
                              static_cast<cudaStream_t>(stream)));
}
void* vip_image_data(const vip_image* img) { return img ? img->buf.p : nullptr; }
size_t vip_image_bytes(const vip_image* img) { return img ? img->buf.bytes : 0; }
int vip_host_alloc(void** out, size_t bytes) {
    if (!out) return VIP_E_NULL;
    return cu(cudaMallocHost(out, bytes));
}
int vip_host_free(void* p) { return cu(cudaFreeHost(p)); }
int vip_stream_synchronize(void* stream) { return cu(cudaStreamSynchronize(static_cast<cudaStream_t>(stream))); }
int vip_device_synchronize(void) { return cu(cudaDeviceSynchronize()); }

/* ---------------------------------------------------------------- BF / JBF */
int vip_bf_create_ex(vip_bf** out, int width, int height, int ksize, float ss, float sc, int lut_mode) {
    if (!out) return VIP_E_NULL;
    *out = nullptr;
    vip_bf* f = new (std::nothrow) vip_bf();
    if (!f) return VIP_E_ALLOC;
    const int rc = init_tables(f->T, width, height, ksize, ss, sc, 768, lut_mode);
    if (rc != VIP_OK) {
        delete f;
        return rc;
    }
    *out = f;
    return VIP_OK;
}
int vip_bf_create(vip_bf** out, int width, int height, int ksize, float ss, float sc) {
    return vip_bf_create_ex(out, width, height, ksize, ss, sc, VIP_LUT_CUDA);
}
int vip_bf_destroy(vip_bf* f) {
    delete f;
    return VIP_OK;
}
int vip_bf_run(const vip_bf* f, const uint8_t* d_src, uint8_t* d_dst, void* stream) {
    if (!f) return VIP_E_NULL;
    return run_stencil(f->T, vip::kBF, d_src, nullptr, d_dst, whole(f->T.height), stream);
}
int vip_jbf_run(const vip_bf* f, const uint8_t* d_src, const uint8_t* d_guide, uint8_t* d_dst, void* stream) {
    if (!f) return VIP_E_NULL;
    return run_stencil(f->T, vip::kJBF, d_src, d_guide, d_dst, whole(f->T.height), stream);
}
int vip_bf_run_strip(const vip_bf* f, const uint8_t* d_src, uint8_t* d_dst, const vip_strip* strip, void* stream) {
    if (!f || !strip) return VIP_E_NULL;
    return run_stencil(f->T, vip::kBF, d_src, nullptr, d_dst, *strip, stream);
}
int vip_jbf_run_strip(const vip_bf* f, const uint8_t* d_src, const uint8_t* d_guide, uint8_t* d_dst,
                      const vip_strip* strip, void* stream) {
    if (!f || !strip) return VIP_E_NULL;
    return run_stencil(f->T, vip::kJBF, d_src, d_guide, d_dst, *strip, stream);
}
int vip_bf_run_strip_peer(const vip_bf* f, const uint8_t* d_src, const vip_halo* halo, uint8_t* d_dst,
                          const vip_strip* strip, void* stream) {
    if (!f || !strip || !halo) return VIP_E_NULL;
    return run_stencil(f->T, vip::kBF, d_src, nullptr, d_dst, *strip, stream, halo);
}
int vip_bf_halo_rows(const vip_bf* f) { return f ? f->T.radius : VIP_E_NULL; }
int vip_bf_get_luts(const vip_bf* f, float* kernel_space, float* color_table) {
    if (!f) return VIP_E_NULL;
    if (kernel_space) std::memcpy(kernel_space, f->T.h_ks.data(), f->T.h_ks.size() * sizeof(float));
    if (color_table) std::memcpy(color_table, f->T.h_lut.data(), f->T.h_lut.size() * sizeof(float));
    return VIP_OK;
}

/* ---------------------------------------------------------------- ABF */
int vip_abf_create_ex(vip_abf** out, int width, int height, int ksize, float ss, float sc, int lut_mode) {
    if (!out) return VIP_E_NULL;
    *out = nullptr;
    vip_abf* f = new (std::nothrow) vip_abf();
    if (!f) return VIP_E_ALLOC;
    const int rc = init_tables(f->T, width, height, ksize, ss, sc, 1536, lut_mode);
    if (rc != VIP_OK) {
        delete f;
        return rc;
    }
    *out = f;
    return VIP_OK;
}
int vip_abf_create(vip_abf** out, int width, int height, int ksize, float ss, float sc) {
    return vip_abf_create_ex(out, width, height, ksize, ss, sc, VIP_LUT_CUDA);
}
int vip_abf_destroy(vip_abf* f) {
    delete f;
    return VIP_OK;
}
int vip_abf_run(const vip_abf* f, const uint8_t* d_src, uint8_t* d_dst, void* stream) {
    if (!f) return VIP_E_NULL;
    return run_stencil(f->T, vip::kABF, d_src, nullptr, d_dst, whole(f->T.height), stream);
}
int vip_abf_run_strip(const vip_abf* f, const uint8_t* d_src, uint8_t* d_dst, const vip_strip* strip, void* stream) {
    if (!f || !strip) return VIP_E_NULL;
    return run_stencil(f->T, vip::kABF, d_src, nullptr, d_dst, *strip, stream);
}
int vip_abf_run_strip_peer(const vip_abf* f, const uint8_t* d_src, const vip_halo* halo, uint8_t* d_dst,
                           const vip_strip* strip, void* stream) {
    if (!f || !strip || !halo) return VIP_E_NULL;
    return run_stencil(f->T, vip::kABF, d_src, nullptr, d_dst, *strip, stream, halo);
}
int vip_abf_halo_rows(const vip_abf* f) { return f ? f->T.radius : VIP_E_NULL; }

/* ---------------------------------------------------------------- gradient */
int vip_gradient_u8(const uint8_t* d_src, float* d_dst, int width, int height, int channels, void* stream) {
    if (!d_src || !d_dst) return VIP_E_NULL;
    if (width <= 0 || height <= 0 || channels <= 0) return VIP_E_SIZE;
    return cu(vip::launch_gradient_u8(d_src, d_dst, width, height, channels, 0, height, 0, 0, height,
                                      static_cast<cudaStream_t>(stream)));
}
int vip_gradient_f32(const float* d_src, float* d_dst, int width, int height, int channels, void* stream) {
    if (!d_src || !d_dst) return VIP_E_NULL;
    if (width <= 0 || height <= 0 || channels <= 0) return VIP_E_SIZE;
    return cu(vip::launch_gradient_f32(d_src, d_dst, width, height, channels, static_cast<cudaStream_t>(stream)));
}
int vip_gradient_u8_strip(const uint8_t* d_src, float* d_dst, int width, int height, int channels,
                          const vip_strip* s, void* stream) {
    if (!d_src || !d_dst || !s) return VIP_E_NULL;
    if (width <= 0 || height <= 0 || channels <= 0) return VIP_E_SIZE;
    const int c = check_strip(*s, height, 1);
    if (c != VIP_OK) return c;
    return cu(vip::launch_gradient_u8(d_src, d_dst, width, height, channels, s->src_y0, s->src_rows, s->dst_y0,
                                      s->out_y0, s->out_rows, static_cast<cudaStream_t>(stream)));
}

/* ---------------------------------------------------------------- host-I/O edge: layout conversion */
int vip_convert_to_bgr_u8(const uint8_t* d_src, int src_channels, uint8_t* d_dst, int width, int height, int swap_rb,
                          void* stream) {
    if (!d_src || !d_dst) return VIP_E_NULL;
    if (width <= 0 || height <= 0) return VIP_E_SIZE;
    if (src_channels != 1 && src_channels != 3 && src_channels != 4) return VIP_E_UNSUPPORTED;
    return cu(vip::launch_convert_to_bgr(d_src, src_channels, d_dst, width, height, swap_rb,
                                         static_cast<cudaStream_t>(stream)));
}
int vip_scale_to_u8_f32(const float* d_src, uint8_t* d_dst, size_t n, float max_value, void* stream) {
    if (!d_src || !d_dst) return VIP_E_NULL;
    if (!(max_value > 0.f)) return VIP_E_SIGMA;
    return cu(vip::launch_scale_to_u8(d_src, d_dst, n, max_value, static_cast<cudaStream_t>(stream)));
}

/* ---------------------------------------------------------------- integral image */
size_t vip_integral_table_elems(int width, int height, int radius) {
    if (width <= 0 || height <= 0 || radius < 0) return 0;
    return static_cast<size_t>(width + 2 * radius + 1) * (height + 2 * radius + 1) * 3;
}
int vip_integral_u8c3(const uint8_t* d_src, uint32_t* d_table, int width, int height, int radius, void* stream) {
    if (!d_src || !d_table) return VIP_E_NULL;
    if (width <= 0 || height <= 0 || radius < 0) return VIP_E_SIZE;
    return cu(vip::launch_integral_u8c3(d_src, d_table, width, height, radius, static_cast<cudaStream_t>(stream)));
}
int vip_integral_get_u8c3(const uint32_t* d_table, int width, int height, int radius, const int32_t* d_rects, int n,
                          uint32_t* d_out, void* stream) {
    if (!d_table || !d_rects || !d_out) return VIP_E_NULL;
    if (width <= 0 || height <= 0 || radius < 0 || n < 0) return VIP_E_SIZE;
    return cu(vip::launch_integral_get(d_table, width, height, radius, d_rects, n, d_out,
                                       static_cast<cudaStream_t>(stream)));
}
int vip_integral_box_mean_u8c3(const uint32_t* d_table, float* d_mean, int width, int height, int radius,
                               void* stream) {
    if (!d_table || !d_mean) return VIP_E_NULL;
    if (width <= 0 || height <= 0 || radius < 0) return VIP_E_SIZE;
    return cu(vip::launch_integral_box_mean(d_table, d_mean, width, height, radius, static_cast<cudaStream_t>(stream)));
}

int vip_integral_build(const void* d_src, void* d_table, int width, int height, int channels, int radius, int src_type,
                       void* stream) {
    if (!d_src || !d_table) return VIP_E_NULL;
    if (width <= 0 || height <= 0 || radius < 0 || channels < 1 || channels > 4) return VIP_E_SIZE;
    if (src_type != VIP_SRC_U8 && src_type != VIP_SRC_U16 && src_type != VIP_SRC_F32) return VIP_E_UNSUPPORTED;
    return cu(vip::launch_integral_any(d_src, d_table, width, height, channels, radius, src_type,
                                       static_cast<cudaStream_t>(stream)));
}
int vip_integral_get(const void* d_table, int width, int height, int channels, int radius, int src_type,
                     const int32_t* d_rects, int n, void* d_out, void* stream) {
    if (!d_table || !d_rects || !d_out) return VIP_E_NULL;
    if (width <= 0 || height <= 0 || radius < 0 || n < 0 || channels < 1 || channels > 4) return VIP_E_SIZE;
    if (src_type != VIP_SRC_U8 && src_type != VIP_SRC_U16 && src_type != VIP_SRC_F32) return VIP_E_UNSUPPORTED;
    return cu(vip::launch_integral_get_any(d_table, width, height, channels, radius, src_type == VIP_SRC_F32, d_rects, n,
                                           d_out, static_cast<cudaStream_t>(stream)));
}

/* ---------------------------------------------------------------- BTF */
static int btf_create_rows(vip_btf** out, int width, int height, int ksize, int nitr, int scratch_rows) {
    if (!out) return VIP_E_NULL;
    *out = nullptr;
    if (width <= 0 || height <= 0 || static_cast<long long>(width) * height > (1ll << 31) / 12) return VIP_E_SIZE;
    if (ksize < 1 || (ksize & 1) == 0) return VIP_E_KSIZE;
    if (nitr < 0) return VIP_E_UNSUPPORTED;
    vip_btf* f = new (std::nothrow) vip_btf();
    if (!f) return VIP_E_ALLOC;
    f->width = width;
    f->height = height;
    f->ksize = ksize;
    f->nitr = nitr;
    f->scratch_rows = scratch_rows;
    f->jbf = nullptr;
    {
        const char* v = std::getenv("VIP_BTF_GRAPH");
        f->use_graph = v && v[0] == '1';
    }
    // src/bilateral_texture_filter_impl.cu:188 and .cuh:31: (ksize 2k-1, sigma_space k-1, sigma_color sqrt(3))
    int rc = VIP_OK;
    if (ksize > 1) {
        rc = vip_bf_create(&f->jbf, width, height, 2 * ksize - 1, static_cast<float>(ksize - 1), 1.73205080757f);
    } else {
        rc = VIP_E_KSIZE;  // k = 1 gives sigma_space 0: the reference divides by zero building its table
    }
    const size_t n = static_cast<size_t>(width) * scratch_rows;
    cudaError_t e = cudaSuccess;
    if (rc == VIP_OK) {
        if ((e = f->ping.alloc(n * 3)) == cudaSuccess && (e = f->pong.alloc(n * 3)) == cudaSuccess &&
            (e = f->guide.alloc(n * 3)) == cudaSuccess && (e = f->blurred.alloc(n * 12)) == cudaSuccess &&
            (e = f->rtv.alloc(n * 4)) == cudaSuccess) {
        }
        rc = cu(e);
    }
    if (rc != VIP_OK) {
        if (f->jbf) vip_bf_destroy(f->jbf);
        delete f;
        return rc;
    }
    *out = f;
    return VIP_OK;
}

int vip_btf_create(vip_btf** out, int width, int height, int ksize, int nitr) {
    return btf_create_rows(out, width, height, ksize, nitr, height);
}
/* internal (multi-GPU driver): scratch sized for a strip */
int vip_btf_create_rows_internal(vip_btf** out, int width, int height, int ksize, int nitr, int scratch_rows) {
    return btf_create_rows(out, width, height, ksize, nitr, scratch_rows);
}
int vip_btf_destroy(vip_btf* f) {
    if (f)
        for (BtfGraph& g : f->graphs) cudaGraphExecDestroy(g.exec);
    if (f && f->jbf) vip_bf_destroy(f->jbf);
    delete f;
    return VIP_OK;
}
size_t vip_btf_scratch_bytes(const vip_btf* f) {
    if (!f) return 0;
    return f->ping.bytes + f->pong.bytes + f->guide.bytes + f->blurred.bytes + f->rtv.bytes;
}
int vip_btf_halo_rows_per_iteration(const vip_btf* f) {
    if (!f) return VIP_E_NULL;
    return (f->ksize - 1) + 2 * (f->ksize / 2) + 1;
}
int vip_btf_blur_rtv(const vip_btf* f, const uint8_t* d_image, const float* d_magnitude, float* d_blurred,
                     float* d_rtv, void* stream) {
    if (!f || !d_image || !d_magnitude || !d_blurred || !d_rtv) return VIP_E_NULL;
    return cu(vip::launch_blur_rtv(d_image, d_magnitude, d_blurred, d_rtv, f->width, f->height, f->ksize, 0, f->height,
                                   0, 0, f->height, static_cast<cudaStream_t>(stream)));
}
int vip_btf_guide(const vip_btf* f, const float* d_blurred, const float* d_rtv, uint8_t* d_guide, void* stream) {
    if (!f || !d_blurred || !d_rtv || !d_guide) return VIP_E_NULL;
    return cu(vip::launch_guide(d_blurred, d_rtv, d_guide, f->width, f->height, f->ksize, 0, f->height, 0, 0,
                                f->height, static_cast<cudaStream_t>(stream)));
}

// One iteration (src/bilateral_texture_filter_impl.cu:205-211) on a strip.  Stage row ranges, from the
// output backwards; all scratch buffers are indexed relative to the source strip's first row.
int vip_btf_iterate_strip(vip_btf* f, const uint8_t* d_src_n, uint8_t* d_dst_n, const vip_strip* s, void* stream) {
    if (!f || !d_src_n || !d_dst_n || !s) return VIP_E_NULL;
    const int H = f->height, W = f->width, k = f->ksize, r = k / 2, R = k - 1;
    const int c = check_strip(*s, H, vip_btf_halo_rows_per_iteration(f));
    if (c != VIP_OK) return c;
    if (s->src_rows > f->scratch_rows) return VIP_E_STRIP;
    if (s->out_rows == 0) return VIP_OK;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int g0 = std::max(0, s->out_y0 - R), g1 = std::min(H, s->out_y0 + s->out_rows + R);  // guide rows
    const int b0 = std::max(0, g0 - r), b1 = std::min(H, g1 + r);                              // blurred / rtv rows
    const int base = s->src_y0;
    float* bl = static_cast<float*>(f->blurred.p);
    float* rt = static_cast<float*>(f->rtv.p);
    uint8_t* gd = static_cast<uint8_t*>(f->guide.p);
    cudaError_t e;
    // the gradient stage (:207) is fused into blur / mRTV: magnitude == nullptr computes it from the image tile, so the
    // magnitude image (rows [b0 - r, b1 + r)) is never materialised
    if ((e = vip::launch_blur_rtv(d_src_n, nullptr, bl, rt, W, H, k, base, s->src_rows, base, b0, b1 - b0, st)) !=
        cudaSuccess)
        return cu(e);
    if ((e = vip::launch_guide(bl, rt, gd, W, H, k, base, s->src_rows, base, g0, g1 - g0, st)) != cudaSuccess)
        return cu(e);
    return vip_jbf_run_strip(f->jbf, d_src_n, gd, d_dst_n, s, stream);
}

static int btf_enqueue(vip_btf* f, const uint8_t* d_src, uint8_t* d_dst, void* stream);

int vip_btf_set_graph(vip_btf* f, int on) {
    if (!f) return VIP_E_NULL;
    f->use_graph = on != 0;
    return VIP_OK;
}

int vip_btf_run(vip_btf* f, const uint8_t* d_src, uint8_t* d_dst, void* stream) {
    if (!f || !d_src || !d_dst) return VIP_E_NULL;
    if (f->scratch_rows < f->height) return VIP_E_STRIP;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    // graph replay: not on the legacy default stream (it cannot be captured) and not for the trivial nitr = 0 copy
    if (!f->use_graph || f->nitr == 0 || st == nullptr || st == cudaStreamLegacy) return btf_enqueue(f, d_src, d_dst, stream);
    for (const BtfGraph& g : f->graphs)
        if (g.src == d_src && g.dst == d_dst) {
            vip::count_launch(3 * f->nitr);
            return cu(cudaGraphLaunch(g.exec, st));
        }
    cudaGraph_t graph = nullptr;
    cudaError_t e = cudaStreamBeginCapture(st, cudaStreamCaptureModeRelaxed);
    if (e != cudaSuccess) return cu(e);
    const int rc = btf_enqueue(f, d_src, d_dst, stream);
    e = cudaStreamEndCapture(st, &graph);
    if (rc != VIP_OK || e != cudaSuccess) {
        if (graph) cudaGraphDestroy(graph);
        cudaGetLastError();
        return rc != VIP_OK ? rc : cu(e);
    }
    cudaGraphExec_t exec = nullptr;
    e = cudaGraphInstantiate(&exec, graph, 0);
    cudaGraphDestroy(graph);
    if (e != cudaSuccess) return cu(e);
    if (f->graphs.size() >= 32) {   // bounded cache: forget the oldest pair
        cudaGraphExecDestroy(f->graphs.front().exec);
        f->graphs.erase(f->graphs.begin());
    }
    f->graphs.push_back(BtfGraph{d_src, d_dst, exec});
    return cu(cudaGraphLaunch(exec, st));
}

static int btf_enqueue(vip_btf* f, const uint8_t* d_src, uint8_t* d_dst, void* stream) {
    const size_t bytes = static_cast<size_t>(f->width) * f->height * 3;
    if (f->nitr == 0)  // the reference copies src -> dst_n -> dst (:203, :213)
        return cu(cudaMemcpyAsync(d_dst, d_src, bytes, cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(stream)));
    const vip_strip s = whole(f->height);
    const uint8_t* in = d_src;
    for (int it = 0; it < f->nitr; it++) {
        // ping-pong instead of the reference's nitr+2 device-to-device copies; the last iterate lands in d_dst --
        // unless it would be written while the same pass still reads it (in-place call with one iteration: the
        // reference works in place because it copies src into its own buffers first, :203-213), then it goes through
        // the scratch image and one copy
        uint8_t* outp = (it == f->nitr - 1) ? d_dst : static_cast<uint8_t*>((it & 1) ? f->pong.p : f->ping.p);
        const bool alias = (outp == in);
        if (alias) outp = static_cast<uint8_t*>(f->ping.p);
        const int rc = vip_btf_iterate_strip(f, in, outp, &s, stream);
        if (rc != VIP_OK) return rc;
        if (alias) {
            const cudaError_t e = cudaMemcpyAsync(d_dst, outp, bytes, cudaMemcpyDeviceToDevice, static_cast<cudaStream_t>(stream));
            if (e != cudaSuccess) return cu(e);
        }
        in = outp;
    }
    return VIP_OK;
}

}  // extern "C"
