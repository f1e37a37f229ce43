// Cross-process strip neighbours (new layer).  With one PROCESS per GPU (torchrun) a rank holds only its
// own rows of the image on its device; the radius-deep halo rows live in the device memory of the ranks above and below.
// This file is what lets the rank reach them over NVLink without going through the host:
//
//   vip_ipc_*    export / open a device allocation across processes (CUDA IPC).  An opened pointer is a plain device
//                pointer in the importing process: cudaMemcpyAsync from it is a peer-to-peer copy (NVLink through the
//                NVSwitch), and a kernel may load from it directly (the *_strip_peer entry points of the filters stage
//                their halo rows straight from the neighbours' memory -- no halo copy at all).
//   vip_link_*   ordering between neighbouring ranks WITHOUT the host: every rank owns a small block of 32-bit flags in
//                device memory, exported to its two neighbours.  "My rows of step s are uploaded" / "I have finished
//                reading your rows of step s" are single-thread kernels that store a step number into the NEIGHBOUR's
//                block (st.release.sys over NVLink); the matching waits poll the rank's OWN block (ld.acquire.sys, local
//                memory).  All of it is enqueued on streams: the host never blocks inside a frame pipeline.
//                A wait gives up after `timeout_ms` of device time and latches an error flag that vip_link_status()
//                reports, so a lost neighbour cannot hang the GPU.
//
// The same objects connect two "virtual ranks" inside one process by raw pointers (vip_link_connect_local), which is how
// the single-GPU tests exercise the protocol.
#include <cuda_runtime.h>

#include <cstdint>
#include <cstring>
#include <new>

#include "../../include/vip_c_api.h"
#include "vip_internal.h"

namespace {

inline int cu(cudaError_t e) { return static_cast<int>(e); }

// flag block: kFlagsPerSlot words per slot
enum { kReadyFromAbove = 0, kReadyFromBelow = 1, kConsumedByAbove = 2, kConsumedByBelow = 3, kFlagsPerSlot = 4 };

__device__ __forceinline__ void store_release_sys(std::uint32_t* p, const std::uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ std::uint32_t load_acquire_sys(const std::uint32_t* p) {
    std::uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}

// One thread: everything earlier in the stream (uploads, kernels) is complete and flushed when it runs.
__global__ void link_signal_kernel(std::uint32_t* a, std::uint32_t* b, const std::uint32_t value) {
    __threadfence_system();
    if (a) store_release_sys(a, value);
    if (b) store_release_sys(b, value);
}

// One thread polls up to two local flags until both are >= value (step numbers only grow; wrap-around safe compare).
__global__ void link_wait_kernel(const std::uint32_t* a, const std::uint32_t* b, const std::uint32_t value,
                                 const long long timeout_ns, std::uint32_t* err) {
    long long t0;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t0));
    for (;;) {
        const bool oka = !a || static_cast<std::int32_t>(load_acquire_sys(a) - value) >= 0;
        const bool okb = !b || static_cast<std::int32_t>(load_acquire_sys(b) - value) >= 0;
        if (oka && okb) return;
        long long t;
        asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
        if (t - t0 > timeout_ns) {
            atomicExch(err, value ? value : 1u);
            return;
        }
        __nanosleep(200);
    }
}

}  // namespace

struct vip_link {
    int n_slots = 0;
    int device = 0;
    std::uint32_t* flags = nullptr;   // own block: n_slots x kFlagsPerSlot, then one error word
    std::uint32_t* above = nullptr;   // neighbour blocks (IPC-mapped or local), nullptr at the image border
    std::uint32_t* below = nullptr;
    bool above_ipc = false, below_ipc = false;
    long long timeout_ns = 2000ll * 1000 * 1000;
};

extern "C" {

/* ---------------------------------------------------------------- CUDA IPC */
int vip_ipc_export(const void* d_ptr, unsigned char* handle) {
    if (!d_ptr || !handle) return VIP_E_NULL;
    static_assert(sizeof(cudaIpcMemHandle_t) == VIP_IPC_HANDLE_BYTES, "handle size");
    cudaIpcMemHandle_t h;
    const cudaError_t e = cudaIpcGetMemHandle(&h, const_cast<void*>(d_ptr));
    if (e != cudaSuccess) return cu(e);
    std::memcpy(handle, &h, sizeof(h));
    return VIP_OK;
}

int vip_ipc_open(const unsigned char* handle, void** d_ptr) {
    if (!handle || !d_ptr) return VIP_E_NULL;
    *d_ptr = nullptr;
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof(h));
    const cudaError_t e = cudaIpcOpenMemHandle(d_ptr, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
        cudaGetLastError();
        // no peer path between the two devices (or not the same box): the caller must pick another transport
        return (e == cudaErrorPeerAccessUnsupported || e == cudaErrorInvalidDevice) ? VIP_E_DEVICE : cu(e);
    }
    return VIP_OK;
}

int vip_ipc_close(void* d_ptr) {
    if (!d_ptr) return VIP_OK;
    return cu(cudaIpcCloseMemHandle(d_ptr));
}

int vip_peer_can_access(int device, int peer_device) {
    int can = 0;
    if (device == peer_device) return 1;
    if (cudaDeviceCanAccessPeer(&can, device, peer_device) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return can;
}

/* Rows [y0, y0+rows) of a packed width x 3 image from a (peer / IPC-mapped / local) buffer whose first row is src_y0
 * into a buffer whose first row is dst_y0: the explicit halo pull (one cudaMemcpyAsync, P2P when the pointers live on
 * different devices). */
int vip_copy_rows(uint8_t* d_dst, int dst_y0, const uint8_t* d_src, int src_y0, int y0, int rows, int width,
                  void* stream) {
    if (!d_dst || !d_src) return VIP_E_NULL;
    if (rows <= 0) return VIP_OK;
    if (width <= 0 || y0 < dst_y0 || y0 < src_y0) return VIP_E_STRIP;
    const size_t rb = static_cast<size_t>(width) * 3;
    return cu(cudaMemcpyAsync(d_dst + rb * (y0 - dst_y0), d_src + rb * (y0 - src_y0), rb * rows, cudaMemcpyDefault,
                              static_cast<cudaStream_t>(stream)));
}

/* ---------------------------------------------------------------- links */
int vip_link_create(vip_link** out, int n_slots) {
    if (!out) return VIP_E_NULL;
    *out = nullptr;
    if (n_slots < 1 || n_slots > 4096) return VIP_E_SIZE;
    vip_link* l = new (std::nothrow) vip_link();
    if (!l) return VIP_E_ALLOC;
    l->n_slots = n_slots;
    cudaError_t e = cudaGetDevice(&l->device);
    const size_t bytes = (static_cast<size_t>(n_slots) * kFlagsPerSlot + 1) * sizeof(std::uint32_t);
    if (e == cudaSuccess) e = cudaMalloc(reinterpret_cast<void**>(&l->flags), bytes);
    if (e == cudaSuccess) e = cudaMemset(l->flags, 0, bytes);
    if (e == cudaSuccess) e = cudaDeviceSynchronize();
    if (e != cudaSuccess) {
        if (l->flags) cudaFree(l->flags);
        delete l;
        return cu(e);
    }
    *out = l;
    return VIP_OK;
}

int vip_link_destroy(vip_link* l) {
    if (!l) return VIP_OK;
    if (l->above && l->above_ipc) cudaIpcCloseMemHandle(l->above);
    if (l->below && l->below_ipc) cudaIpcCloseMemHandle(l->below);
    if (l->flags) cudaFree(l->flags);
    delete l;
    return VIP_OK;
}

int vip_link_export(const vip_link* l, unsigned char* handle) {
    if (!l) return VIP_E_NULL;
    return vip_ipc_export(l->flags, handle);
}

int vip_link_connect(vip_link* l, const unsigned char* above_handle, const unsigned char* below_handle) {
    if (!l) return VIP_E_NULL;
    if (above_handle) {
        void* p = nullptr;
        const int rc = vip_ipc_open(above_handle, &p);
        if (rc != VIP_OK) return rc;
        l->above = static_cast<std::uint32_t*>(p);
        l->above_ipc = true;
    }
    if (below_handle) {
        void* p = nullptr;
        const int rc = vip_ipc_open(below_handle, &p);
        if (rc != VIP_OK) return rc;
        l->below = static_cast<std::uint32_t*>(p);
        l->below_ipc = true;
    }
    return VIP_OK;
}

int vip_link_connect_local(vip_link* l, vip_link* above, vip_link* below) {
    if (!l) return VIP_E_NULL;
    l->above = above ? above->flags : nullptr;
    l->below = below ? below->flags : nullptr;
    l->above_ipc = l->below_ipc = false;
    return VIP_OK;
}

int vip_link_set_timeout_ms(vip_link* l, int timeout_ms) {
    if (!l) return VIP_E_NULL;
    if (timeout_ms < 1) return VIP_E_SIZE;
    l->timeout_ns = static_cast<long long>(timeout_ms) * 1000 * 1000;
    return VIP_OK;
}

static int link_signal(vip_link* l, int slot, int which_in_above, int which_in_below, unsigned value, void* stream) {
    if (!l) return VIP_E_NULL;
    if (slot < 0 || slot >= l->n_slots) return VIP_E_SIZE;
    if (!l->above && !l->below) return VIP_OK;
    std::uint32_t* a = l->above ? l->above + slot * kFlagsPerSlot + which_in_above : nullptr;
    std::uint32_t* b = l->below ? l->below + slot * kFlagsPerSlot + which_in_below : nullptr;
    link_signal_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(a, b, value);
    vip::count_launch();
    return cu(cudaGetLastError());
}

static int link_wait(vip_link* l, int slot, int which_above, int which_below, unsigned value, void* stream) {
    if (!l) return VIP_E_NULL;
    if (slot < 0 || slot >= l->n_slots) return VIP_E_SIZE;
    if (!l->above && !l->below) return VIP_OK;
    const std::uint32_t* a = l->above ? l->flags + slot * kFlagsPerSlot + which_above : nullptr;
    const std::uint32_t* b = l->below ? l->flags + slot * kFlagsPerSlot + which_below : nullptr;
    link_wait_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(a, b, value, l->timeout_ns,
                                                                       l->flags + l->n_slots * kFlagsPerSlot);
    vip::count_launch();
    return cu(cudaGetLastError());
}

/* "my rows of step `value` are in place": seen by the rank above as ready-from-below, by the rank below as
 * ready-from-above */
int vip_link_signal_ready(vip_link* l, int slot, unsigned value, void* stream) {
    return link_signal(l, slot, kReadyFromBelow, kReadyFromAbove, value, stream);
}
int vip_link_wait_ready(vip_link* l, int slot, unsigned value, void* stream) {
    return link_wait(l, slot, kReadyFromAbove, kReadyFromBelow, value, stream);
}
/* "I have finished reading your rows of step `value`" */
int vip_link_signal_consumed(vip_link* l, int slot, unsigned value, void* stream) {
    return link_signal(l, slot, kConsumedByBelow, kConsumedByAbove, value, stream);
}
int vip_link_wait_consumed(vip_link* l, int slot, unsigned value, void* stream) {
    return link_wait(l, slot, kConsumedByAbove, kConsumedByBelow, value, stream);
}

/* 0 while no wait has timed out; otherwise VIP_E_DEVICE (the step number that timed out is lost on purpose: the
 * pipeline is dead either way).  Synchronises the device. */
int vip_link_status(vip_link* l) {
    if (!l) return VIP_E_NULL;
    std::uint32_t err = 0;
    const cudaError_t e = cudaMemcpy(&err, l->flags + l->n_slots * kFlagsPerSlot, sizeof(err), cudaMemcpyDeviceToHost);
    if (e != cudaSuccess) return cu(e);
    return err ? VIP_E_DEVICE : VIP_OK;
}

/* Clears the latched time-out so that the link can be used again after the cause has been dealt with (neighbour
 * restarted, steps re-agreed).  Synchronises the device first: no wait of this rank may still be spinning.  The step
 * counters are NOT touched -- they are monotonic and shared with the neighbours. */
int vip_link_reset(vip_link* l) {
    if (!l) return VIP_E_NULL;
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) return cu(e);
    const std::uint32_t zero = 0;
    e = cudaMemcpy(l->flags + l->n_slots * kFlagsPerSlot, &zero, sizeof(zero), cudaMemcpyHostToDevice);
    return cu(e);
}

}  // extern "C"
