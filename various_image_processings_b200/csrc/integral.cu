// Border-replicated integral image on the GPU (sm_100a).  GPU twin of the reference's CPU-only
// BorderReplicatedIntegralImage<SrcType, Channels> (include/cpp/border_replicated_integral_image.hpp:6-85): a
// (H+2r+1) x (W+2r+1) x C table over the clamp-padded image, first row / column zero; get(x0,y0,x1,y1) is the
// inclusive box sum from four corners (:34-47).  Integer sources (u8, u16) accumulate in uint32 modulo 2^32:
// differences of corners are exact for any image size (the reference's int32 table overflows -- UB -- beyond ~4K,
// SURVEY.md 8a/a11).  Float sources accumulate in fp32 (the reference's own test allows 1 % there,
// test/border_replicated_integral_image.cpp:36-39).
//
// The stage is HBM-bound: 12 B written + C bytes read per padded cell is the floor.  Every (8-row, 256-column) tile of the
// table is produced independently from three small sets of aggregates -- four launches, all but the last tiny:
//  K1 band sums   32-row bands x 256-column chunks, one thread per padded column: vertical sums of the clamped source,
//                 written at every 8th row (the running sum inside the band) plus the band total; the same pass reduces
//                 every row of the chunk to its ROW SUM (one warp reduction per row and channel);
//  K2 band scan   exclusive prefix of the band totals over the bands (tiny: th/32 entries per column);
//  K2b row carry  for every chunk k and row Y: the table value just left of the chunk, T(Y, 256 k - 1) = the sum over all
//                 rows <= Y of the row sums of the chunks left of k -- one block scan per (chunk, channel) over th rows;
//  K3 scan+store  one CTA per tile.  Phase A: a thread per column turns the 8 source rows into vertical prefix sums
//                 (start value from K1/K2) in shared memory; phase B: a WARP per row, a lane per 8 adjacent columns:
//                 lane-local prefix, one __shfl_up_sync scan of the lane totals, plus the row's carry from K2b;
//                 phase C: fully coalesced 128-byte stores.  (Round 1 let one CTA walk along a whole 8-row band with a
//                 running carry: 272 CTAs at 4K, 23 % occupancy, 49 us of the 77; tiles are 16x as many CTAs.)
// Traffic: the source is read twice (the second time from L2 for anything up to 8K), the table written once; the
// aggregate arrays are 1/8 of the table and stay in L2.
#include <cuda_runtime.h>

#include <cstdint>
#include <mutex>
#include <type_traits>

#include "../../include/vip_c_api.h"
#include "vip_internal.h"

namespace vip {
namespace {

__device__ __forceinline__ int clampi(int v, int lo, int hi) { return v < lo ? lo : (v > hi ? hi : v); }

constexpr int kScanThreads = 256;
constexpr int kChunk = 256;     // padded columns per chunk
constexpr int kSubRows = 8;     // rows of a K3 band (= warps per CTA)
constexpr int kBandRows = 32;   // rows of a K1/K2 band
constexpr int kSubPerBand = kBandRows / kSubRows;

template <typename S>
struct AccOf {
    using type = std::uint32_t;
};
template <>
struct AccOf<float> {
    using type = float;
};

// value of padded cell (Y, X): 0 on the first row / column, else the clamped source pixel
template <typename S, int C>
__device__ __forceinline__ const S* cell_ptr(const S* __restrict__ src, const int W, const int H, const int r,
                                             const int Y, const int X) {
    const int sy = clampi(Y - 1 - r, 0, H - 1), sx = clampi(X - 1 - r, 0, W - 1);
    return src + (static_cast<size_t>(sy) * W + sx) * C;
}

template <typename A>
__device__ __forceinline__ A warp_sum(A v) {
    if constexpr (std::is_same<A, std::uint32_t>::value) {
        return __reduce_add_sync(0xffffffffu, v);   // REDUX: one instruction
    } else {
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) v += __shfl_xor_sync(0xffffffffu, v, d);
        return v;
    }
}

// K1: sub[(Y/8)][X][c] = sum of padded rows [32*band, Y) of column X (Y a multiple of 8); tot[band][X][c] = band total;
// rowsum[(chunk*C + c)*th + Y] = sum of the cells of padded row Y inside this CTA's 256 columns
template <typename S, int C>
__global__ void __launch_bounds__(kScanThreads) integral_bands_kernel(const S* __restrict__ src,
                                                                     typename AccOf<S>::type* __restrict__ sub,
                                                                     typename AccOf<S>::type* __restrict__ tot,
                                                                     typename AccOf<S>::type* __restrict__ rowsum,
                                                                     const int W, const int H, const int r) {
    using A = typename AccOf<S>::type;
    __shared__ A part[kScanThreads / 32][kBandRows][C];   // per-warp row sums, added in a fixed order (deterministic for float)
    const int tw = W + 2 * r + 1, th = H + 2 * r + 1;
    const int X = blockIdx.x * kScanThreads + threadIdx.x;
    const bool in = X < tw;
    const int band = blockIdx.y;
    const int Y0 = band * kBandRows;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane < kBandRows) {
#pragma unroll
        for (int c = 0; c < C; ++c) part[warp][lane][c] = A(0);   // rows past the table's end stay zero
    }
    __syncwarp();
    A acc[C];
#pragma unroll
    for (int c = 0; c < C; ++c) acc[c] = A(0);
#pragma unroll
    for (int q = 0; q < kSubPerBand; ++q) {
        const int Ys = Y0 + q * kSubRows;
        if (Ys >= th) break;   // uniform over the CTA
        if (in) {
            A* o = sub + (static_cast<size_t>(Ys / kSubRows) * tw + X) * C;
#pragma unroll
            for (int c = 0; c < C; ++c) o[c] = acc[c];
        }
        S pix[kSubRows][C];
#pragma unroll
        for (int j = 0; j < kSubRows; ++j) {
            const int Y = Ys + j;
            const bool ok = in && Y >= 1 && Y < th && X >= 1;
            const S* p = cell_ptr<S, C>(src, W, H, r, ok ? Y : 1, ok ? X : 1);
#pragma unroll
            for (int c = 0; c < C; ++c) pix[j][c] = ok ? __ldg(p + c) : S(0);
        }
#pragma unroll
        for (int j = 0; j < kSubRows; ++j) {
#pragma unroll
            for (int c = 0; c < C; ++c) {
                const A v = static_cast<A>(pix[j][c]);
                acc[c] += v;
                const A rs = warp_sum<A>(v);
                if (lane == 0) part[warp][q * kSubRows + j][c] = rs;
            }
        }
    }
    if (in) {
        A* t = tot + (static_cast<size_t>(band) * tw + X) * C;
#pragma unroll
        for (int c = 0; c < C; ++c) t[c] = acc[c];
    }
    __syncthreads();
    if (threadIdx.x < kBandRows * C) {
        const int row = threadIdx.x / C, c = threadIdx.x - row * C;
        const int Y = Y0 + row;
        A sum = A(0);
#pragma unroll
        for (int w = 0; w < kScanThreads / 32; ++w) sum += part[w][row][c];
        if (Y < th) rowsum[(static_cast<size_t>(blockIdx.x) * C + c) * th + Y] = sum;
    }
}

// K2b: carry[(k*C + c)*th + Y] = sum over rows Y' <= Y of the row sums of the chunks k' < k: the table value just left of
// chunk k in row Y.  One CTA per (chunk k >= 1, channel): an inclusive block scan over the th rows.
constexpr int kCarryThreads = 1024;
template <typename A>
__global__ void __launch_bounds__(kCarryThreads) integral_row_carry_kernel(const A* __restrict__ rowsum, A* __restrict__ carry,
                                                                           const int th, const int C) {
    __shared__ A warp_tot[kCarryThreads / 32];
    __shared__ A run_s;
    const int k = blockIdx.x + 1, c = blockIdx.y;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    if (tid == 0) run_s = A(0);
    __syncthreads();
    for (int base = 0; base < th; base += kCarryThreads) {
        const int Y = base + tid;
        A v = A(0);
        if (Y < th) {
#pragma unroll 8
            for (int kk = 0; kk < k; ++kk) v += __ldg(rowsum + (static_cast<size_t>(kk) * C + c) * th + Y);   // independent loads
        }
        A incl = v;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const A up = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += up;
        }
        if (lane == 31) warp_tot[warp] = incl;
        __syncthreads();
        A before = run_s;
        for (int q = 0; q < warp; ++q) before += warp_tot[q];
        if (Y < th) carry[(static_cast<size_t>(k) * C + c) * th + Y] = incl + before;
        __syncthreads();
        if (tid == kCarryThreads - 1) run_s = incl + before;
        __syncthreads();
    }
}

// K2: in-place exclusive scan of tot over the bands.  A CTA owns 32 (column, channel) entries; its 8 warps split the
// bands into 8 contiguous segments (all loads of a segment are independent, so the kernel costs about two memory
// latencies instead of nbands of them), exchange the segment totals through shared memory and write the prefixes.
constexpr int kScanSeg = 8;
template <typename A>
__global__ void __launch_bounds__(kScanThreads) integral_band_scan_kernel(A* __restrict__ tot, const int pitch,
                                                                         const int nbands) {
    __shared__ A seg_tot[kScanSeg][32];
    const int lane = threadIdx.x & 31, seg = threadIdx.x >> 5;
    const int i = blockIdx.x * 32 + lane;
    const int per = (nbands + kScanSeg - 1) / kScanSeg;
    const int b0 = min(seg * per, nbands), b1 = min(b0 + per, nbands);
    constexpr int kKeep = 24;  // bands of a segment kept in registers (nbands <= 192 <=> image height <= ~6100 rows)
    A v[kKeep];
    A sum = A(0);
    if (i < pitch) {
#pragma unroll
        for (int j = 0; j < kKeep; ++j) v[j] = (b0 + j < b1) ? tot[static_cast<size_t>(b0 + j) * pitch + i] : A(0);
#pragma unroll
        for (int j = 0; j < kKeep; ++j) sum += v[j];
        for (int b = b0 + kKeep; b < b1; ++b) sum += tot[static_cast<size_t>(b) * pitch + i];
    }
    seg_tot[seg][lane] = sum;
    __syncthreads();
    if (i >= pitch) return;
    A run = A(0);
    for (int q = 0; q < seg; ++q) run += seg_tot[q][lane];
#pragma unroll
    for (int j = 0; j < kKeep; ++j) {
        if (b0 + j < b1) {
            tot[static_cast<size_t>(b0 + j) * pitch + i] = run;
            run += v[j];
        }
    }
    for (int b = b0 + kKeep; b < b1; ++b) {
        const A x = tot[static_cast<size_t>(b) * pitch + i];
        tot[static_cast<size_t>(b) * pitch + i] = run;
        run += x;
    }
}

// K3: scan + store of one (8-row, 256-column) tile (see the header).  Sm holds the tile as [row][pad(column*C + channel)],
// one spare word after every 8*C so that the 32 lanes of phase B start in 32 different banks.
template <typename S, int C>
__global__ void __launch_bounds__(kScanThreads) integral_scan_kernel(const S* __restrict__ src,
                                                                    const typename AccOf<S>::type* __restrict__ sub,
                                                                    const typename AccOf<S>::type* __restrict__ tot,
                                                                    const typename AccOf<S>::type* __restrict__ carry_in,
                                                                    typename AccOf<S>::type* __restrict__ table,
                                                                    const int W, const int H, const int r) {
    using A = typename AccOf<S>::type;
    constexpr int kLaneWords = 8 * C;               // words one lane owns in phase B
    constexpr int kRowWords = kChunk * C + kChunk / 8;
    __shared__ A Sm[kSubRows][kRowWords];
    const int tw = W + 2 * r + 1, th = H + 2 * r + 1;
    const int chunk = blockIdx.x, X0 = chunk * kChunk;
    const int b8 = blockIdx.y, Y0 = b8 * kSubRows;
    const int band = Y0 / kBandRows;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    // phase A: vertical prefix of this thread's column, starting from the sum of everything above the tile
    {
        const int X = X0 + tid;
        const bool in = X >= 1 && X < tw;
        A acc[C];
#pragma unroll
        for (int c = 0; c < C; ++c) acc[c] = A(0);
        if (in) {
            const A* t = tot + (static_cast<size_t>(band) * tw + X) * C;
            const A* u = sub + (static_cast<size_t>(b8) * tw + X) * C;
#pragma unroll
            for (int c = 0; c < C; ++c) acc[c] = __ldg(t + c) + __ldg(u + c);
        }
        const S* const colp = src + static_cast<size_t>(clampi(X - 1 - r, 0, W - 1)) * C;
        S pix[kSubRows][C];
#pragma unroll
        for (int j = 0; j < kSubRows; ++j) {
            const int Y = Y0 + j;
            const bool ok = in && Y >= 1 && Y < th;
            const S* p = colp + static_cast<size_t>(clampi(Y - 1 - r, 0, H - 1)) * W * C;
#pragma unroll
            for (int c = 0; c < C; ++c) pix[j][c] = ok ? __ldg(p + c) : S(0);
        }
        const int w0 = tid * C + tid / 8;
#pragma unroll
        for (int j = 0; j < kSubRows; ++j) {
#pragma unroll
            for (int c = 0; c < C; ++c) {
                acc[c] += static_cast<A>(pix[j][c]);
                Sm[j][w0 + c] = acc[c];
            }
        }
    }
    __syncthreads();
    // phase B: warp = row, lane = 8 adjacent columns; the row starts from the table value just left of the chunk
    {
        const int Y = Y0 + warp;
        A* row = &Sm[warp][lane * (kLaneWords + 1)];
        A v[8][C], run[C];
#pragma unroll
        for (int c = 0; c < C; ++c) run[c] = A(0);
#pragma unroll
        for (int m = 0; m < 8; ++m) {
#pragma unroll
            for (int c = 0; c < C; ++c) {
                run[c] += row[m * C + c];
                v[m][c] = run[c];
            }
        }
#pragma unroll
        for (int c = 0; c < C; ++c) {
            A incl = run[c];
#pragma unroll
            for (int d = 1; d < 32; d <<= 1) {
                const A up = __shfl_up_sync(0xffffffffu, incl, d);
                if (lane >= d) incl += up;
            }
            A excl = __shfl_up_sync(0xffffffffu, incl, 1);
            if (lane == 0) excl = A(0);
            if (chunk > 0 && Y < th) excl += __ldg(carry_in + (static_cast<size_t>(chunk) * C + c) * th + Y);
#pragma unroll
            for (int m = 0; m < 8; ++m) row[m * C + c] = v[m][c] + excl;
        }
    }
    __syncthreads();
    // phase C: coalesced stores
    {
        const int nwords = min(kChunk, tw - X0) * C;
        const int nrows = min(kSubRows, th - Y0);
        A* const tbase = table + (static_cast<size_t>(Y0) * tw + X0) * C;
#pragma unroll
        for (int q = 0; q < C; ++q) {
            const int i = tid + q * kScanThreads;
            if (i < nwords) {
                const int si = i + i / kLaneWords;
#pragma unroll
                for (int j = 0; j < kSubRows; ++j)
                    if (j < nrows) tbase[static_cast<size_t>(j) * tw * C + i] = Sm[j][si];
            }
        }
    }
}

// Stream-ordered scratch from a library-private memory pool that keeps its memory between calls (the default pool
// gives freed blocks back to the driver at the next synchronisation; re-acquiring ~16 MB per call cost more than the
// three kernels together).
cudaError_t scratch_alloc(void** p, size_t bytes, cudaStream_t stream) {
    static std::mutex mu;
    static cudaMemPool_t pools[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaMallocAsync(p, bytes, stream);
    {
        std::lock_guard<std::mutex> lock(mu);
        if (!pools[dev]) {
            cudaMemPoolProps props = {};
            props.allocType = cudaMemAllocationTypePinned;
            props.handleTypes = cudaMemHandleTypeNone;
            props.location.type = cudaMemLocationTypeDevice;
            props.location.id = dev;
            e = cudaMemPoolCreate(&pools[dev], &props);
            if (e != cudaSuccess) return e;
            unsigned long long keep = ~0ull;
            cudaMemPoolSetAttribute(pools[dev], cudaMemPoolAttrReleaseThreshold, &keep);
        }
    }
    return cudaMallocFromPoolAsync(p, bytes, pools[dev], stream);
}

template <typename S, int C>
cudaError_t launch_integral_t(const S* src, typename AccOf<S>::type* table, int width, int height, int radius,
                              cudaStream_t stream) {
    using A = typename AccOf<S>::type;
    const int tw = width + 2 * radius + 1, th = height + 2 * radius + 1;
    const int nsub = (th + kSubRows - 1) / kSubRows, nbands = (th + kBandRows - 1) / kBandRows;
    const int nchunks = (tw + kChunk - 1) / kChunk;
    const size_t row_elems = static_cast<size_t>(tw) * C;
    const size_t rs_elems = static_cast<size_t>(nchunks) * C * th;   // row sums and row carries: [chunk][channel][row]
    A* aux = nullptr;
    cudaError_t e = scratch_alloc(reinterpret_cast<void**>(&aux), ((nsub + nbands) * row_elems + 2 * rs_elems) * sizeof(A), stream);
    if (e != cudaSuccess) return e;
    A* sub = aux;
    A* tot = aux + nsub * row_elems;
    A* rowsum = tot + nbands * row_elems;
    A* carry = rowsum + rs_elems;
    const dim3 g1(nchunks, nbands);
    integral_bands_kernel<S, C><<<g1, kScanThreads, 0, stream>>>(src, sub, tot, rowsum, width, height, radius);
    integral_band_scan_kernel<A><<<(static_cast<int>(row_elems) + 31) / 32, kScanThreads, 0, stream>>>(
        tot, static_cast<int>(row_elems), nbands);
    if (nchunks > 1)
        integral_row_carry_kernel<A><<<dim3(nchunks - 1, C), kCarryThreads, 0, stream>>>(rowsum, carry, th, C);
    integral_scan_kernel<S, C><<<dim3(nchunks, nsub), kScanThreads, 0, stream>>>(src, sub, tot, carry, table, width, height,
                                                                                  radius);
    count_launch(nchunks > 1 ? 4 : 3);
    e = cudaGetLastError();
    cudaFreeAsync(aux, stream);
    return e;
}

template <typename S>
cudaError_t launch_integral_c(const S* src, typename AccOf<S>::type* table, int width, int height, int channels,
                              int radius, cudaStream_t stream) {
    switch (channels) {
        case 1: return launch_integral_t<S, 1>(src, table, width, height, radius, stream);
        case 2: return launch_integral_t<S, 2>(src, table, width, height, radius, stream);
        case 3: return launch_integral_t<S, 3>(src, table, width, height, radius, stream);
        case 4: return launch_integral_t<S, 4>(src, table, width, height, radius, stream);
        default: return cudaErrorInvalidValue;
    }
}

__device__ __forceinline__ void box3(const std::uint32_t* __restrict__ table, const int tw, const int r, const int x0,
                                     const int y0, const int x1, const int y1, std::uint32_t out[3]) {
    const std::uint32_t* a = table + (static_cast<size_t>(y1 + r + 1) * tw + (x1 + r + 1)) * 3;
    const std::uint32_t* b = table + (static_cast<size_t>(y1 + r + 1) * tw + (x0 + r)) * 3;
    const std::uint32_t* c = table + (static_cast<size_t>(y0 + r) * tw + (x1 + r + 1)) * 3;
    const std::uint32_t* d = table + (static_cast<size_t>(y0 + r) * tw + (x0 + r)) * 3;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) out[ch] = __ldg(a + ch) - __ldg(b + ch) - __ldg(c + ch) + __ldg(d + ch);
}

__global__ void integral_get_kernel(const std::uint32_t* __restrict__ table, const int W, const int r,
                                    const std::int32_t* __restrict__ rects, const int n,
                                    std::uint32_t* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    std::uint32_t v[3];
    box3(table, W + 2 * r + 1, r, rects[4 * i], rects[4 * i + 1], rects[4 * i + 2], rects[4 * i + 3], v);
    out[3 * i] = v[0];
    out[3 * i + 1] = v[1];
    out[3 * i + 2] = v[2];
}

__global__ void integral_box_mean_kernel(const std::uint32_t* __restrict__ table, float* __restrict__ mean,
                                         const int W, const int H, const int r) {
    const int x = blockIdx.x * blockDim.x + threadIdx.x;
    const int y = blockIdx.y * blockDim.y + threadIdx.y;
    if (x >= W || y >= H) return;
    std::uint32_t v[3];
    box3(table, W + 2 * r + 1, r, x - r, y - r, x + r, y + r, v);
    const float kk = static_cast<float>((2 * r + 1) * (2 * r + 1));
    float* m = mean + (static_cast<size_t>(y) * W + x) * 3;
    m[0] = __fdiv_rn(static_cast<float>(v[0]), kk);
    m[1] = __fdiv_rn(static_cast<float>(v[1]), kk);
    m[2] = __fdiv_rn(static_cast<float>(v[2]), kk);
}

}  // namespace

cudaError_t launch_integral_u8c3(const std::uint8_t* src, std::uint32_t* table, int width, int height, int radius,
                                 cudaStream_t stream) {
    return launch_integral_t<std::uint8_t, 3>(src, table, width, height, radius, stream);
}

cudaError_t launch_integral_any(const void* src, void* table, int width, int height, int channels, int radius,
                                int src_type, cudaStream_t stream) {
    switch (src_type) {
        case 0: return launch_integral_c(static_cast<const std::uint8_t*>(src), static_cast<std::uint32_t*>(table), width,
                                         height, channels, radius, stream);
        case 1: return launch_integral_c(static_cast<const std::uint16_t*>(src), static_cast<std::uint32_t*>(table), width,
                                         height, channels, radius, stream);
        case 2: return launch_integral_c(static_cast<const float*>(src), static_cast<float*>(table), width, height,
                                         channels, radius, stream);
        default: return cudaErrorInvalidValue;
    }
}

// get() for any channel count / accumulator type: out is n x channels
template <typename A>
__global__ void integral_get_any_kernel(const A* __restrict__ table, const int W, const int C, const int r,
                                        const std::int32_t* __restrict__ rects, const int n, A* __restrict__ out) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n * C) return;
    const int q = i / C, ch = i - q * C;
    const int tw = W + 2 * r + 1;
    const int x0 = rects[4 * q], y0 = rects[4 * q + 1], x1 = rects[4 * q + 2], y1 = rects[4 * q + 3];
    auto at = [&](int Y, int X) { return __ldg(table + (static_cast<size_t>(Y) * tw + X) * C + ch); };
    // include/cpp/border_replicated_integral_image.hpp:34-47, same left-to-right order (matters for float)
    out[i] = at(y1 + r + 1, x1 + r + 1) - at(y1 + r + 1, x0 + r) - at(y0 + r, x1 + r + 1) + at(y0 + r, x0 + r);
}

cudaError_t launch_integral_get_any(const void* table, int width, int height, int channels, int radius, int is_float,
                                    const std::int32_t* rects, int n, void* out, cudaStream_t stream) {
    (void)height;
    if (n <= 0) return cudaSuccess;
    const int total = n * channels;
    if (is_float)
        integral_get_any_kernel<float><<<(total + 255) / 256, 256, 0, stream>>>(
            static_cast<const float*>(table), width, channels, radius, rects, n, static_cast<float*>(out));
    else
        integral_get_any_kernel<std::uint32_t><<<(total + 255) / 256, 256, 0, stream>>>(
            static_cast<const std::uint32_t*>(table), width, channels, radius, rects, n, static_cast<std::uint32_t*>(out));
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_integral_get(const std::uint32_t* table, int width, int height, int radius,
                                const std::int32_t* rects, int n, std::uint32_t* out, cudaStream_t stream) {
    (void)height;
    if (n <= 0) return cudaSuccess;
    integral_get_kernel<<<(n + 255) / 256, 256, 0, stream>>>(table, width, radius, rects, n, out);
    count_launch();
    return cudaGetLastError();
}

cudaError_t launch_integral_box_mean(const std::uint32_t* table, float* mean, int width, int height, int radius,
                                     cudaStream_t stream) {
    const dim3 block(32, 8);
    const dim3 grid((width + 31) / 32, (height + 7) / 8);
    integral_box_mean_kernel<<<grid, block, 0, stream>>>(table, mean, width, height, radius);
    count_launch();
    return cudaGetLastError();
}

}  // namespace vip

// ---------------------------------------------------------------------------------------------------------------
// vip_integral_u8c3_noalloc: the integral image for callers that are not allowed to allocate device memory.
//
// vip_integral_u8c3 takes its aggregate arrays (band sums, band totals, row sums, row carries -- about 1/3 of the
// table) from a stream-ordered memory pool.  That is a device allocation: it can fail when the process has handed all
// of the GPU's memory to its own allocator, it may synchronise, and some hosts simply forbid it -- a plug-in running
// inside an application that pre-reserves the whole device (a video pipeline with a fixed memory budget, an
// inference server), a real-time loop that must not call the driver's allocator after start-up, or a process whose
// memory is accounted by an outer framework that would not see ours.  This variant runs the same four kernels with the
// aggregates in a fixed array that is part of the library image itself.
//
// The array costs kNoAllocWords * 4 bytes (20 MiB) on EVERY GPU on which the library's kernels get loaded, from the
// first use until the process exits, whether or not this entry point is ever called.  Images whose aggregates do not
// fit are refused (VIP_E_SIZE); a 3840x2160 frame fits up to radius 15, a 7680x4320 frame does not.
constexpr size_t kNoAllocWords = size_t(5) << 20;
static_assert(kNoAllocWords * sizeof(std::uint32_t) == VIP_INTEGRAL_NOALLOC_SCRATCH_BYTES, "header and array disagree");
static __device__ __align__(256) std::uint32_t g_noalloc_scratch[kNoAllocWords];

namespace {
// In host code `g_noalloc_scratch` is only a handle for the runtime, not somewhere a kernel could write; the real
// device address differs per GPU and is looked up once per device with that device current.
cudaError_t noalloc_scratch_ptr(std::uint32_t** out) {
    static std::mutex mu;
    static std::uint32_t* ptrs[64] = {};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        void* p = nullptr;
        e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess) return e;
        ptrs[dev] = static_cast<std::uint32_t*>(p);
    }
    *out = ptrs[dev];
    return cudaSuccess;
}
}  // namespace

extern "C" VIP_API int vip_integral_u8c3_noalloc(const uint8_t* d_src, uint32_t* d_table, int width, int height,
                                                  int radius, void* stream_) {
    using namespace vip;
    using A = std::uint32_t;
    constexpr int C = 3;
    if (!d_src || !d_table) return VIP_E_NULL;
    if (width <= 0 || height <= 0 || radius < 0) return VIP_E_SIZE;
    // 64-bit throughout: a caller probing the limit with absurd sizes must get VIP_E_SIZE, not a wrapped product
    const long long tw = static_cast<long long>(width) + 2ll * radius + 1, th = static_cast<long long>(height) + 2ll * radius + 1;
    const long long nsub = (th + kSubRows - 1) / kSubRows, nbands = (th + kBandRows - 1) / kBandRows;
    const long long nchunks = (tw + kChunk - 1) / kChunk;
    const long long row_elems = tw * C, rs_elems = nchunks * C * th;
    if (tw > (1 << 24) || th > (1 << 24) || nsub > 65535) return VIP_E_SIZE;   // nsub is a grid y-dimension
    const long long need = (nsub + nbands) * row_elems + 2 * rs_elems;
    if (need > static_cast<long long>(kNoAllocWords)) return VIP_E_SIZE;
    A* aux = nullptr;
    const cudaError_t e = noalloc_scratch_ptr(&aux);   // first call on a device: one cudaGetSymbolAddress
    if (e != cudaSuccess) return static_cast<int>(e);
    cudaStream_t stream = static_cast<cudaStream_t>(stream_);
    // Same carve-up and launch sequence as launch_integral_t.  Nothing has to be cleared between calls: every word
    // a kernel reads (sub, tot, rowsum, carry) is written earlier in the same call by K1 / K2 / K2b for exactly the
    // (row, column, chunk) ranges K3 reads, so whatever the previous call left behind is never looked at.
    A* sub = aux;
    A* tot = aux + nsub * row_elems;
    A* rowsum = tot + nbands * row_elems;
    A* carry = rowsum + rs_elems;
    const dim3 g1(static_cast<unsigned>(nchunks), static_cast<unsigned>(nbands));
    integral_bands_kernel<std::uint8_t, C><<<g1, kScanThreads, 0, stream>>>(d_src, sub, tot, rowsum, width, height, radius);
    integral_band_scan_kernel<A><<<(static_cast<int>(row_elems) + 31) / 32, kScanThreads, 0, stream>>>(
        tot, static_cast<int>(row_elems), static_cast<int>(nbands));
    if (nchunks > 1)
        integral_row_carry_kernel<A><<<dim3(static_cast<unsigned>(nchunks - 1), C), kCarryThreads, 0, stream>>>(
            rowsum, carry, static_cast<int>(th), C);
    integral_scan_kernel<std::uint8_t, C><<<dim3(static_cast<unsigned>(nchunks), static_cast<unsigned>(nsub)), kScanThreads, 0,
                                            stream>>>(d_src, sub, tot, carry, d_table, width, height, radius);
    count_launch(nchunks > 1 ? 4 : 3);
    return static_cast<int>(cudaGetLastError());
}
