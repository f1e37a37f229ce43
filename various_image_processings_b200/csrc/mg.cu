// Multi-GPU driver (new layer; the reference is single-GPU).  One host process drives the devices
// of one box:
//   * a large image is cut into contiguous row strips, one per device slot; every strip buffer carries `halo`
//     rows of its neighbours on each side; kernels clamp against the FULL image, so the strip results are
//     bit-identical to a single-GPU run;
//   * BF / ABF need their halo once (it comes with the scatter).  BTF needs 17 rows (k = 9) of the previous
//     iterate per iteration and has two ways to get them:
//       deep halo  the strip buffers carry nitr x 17 rows per side from the scatter; iteration `it` recomputes the
//                  shrinking overlap (nitr-1-it) x 17 rows per side redundantly -- no communication and no
//                  cross-device ordering at all (12.6 % extra rows at 4K on 8 devices);
//       exchange   17 halo rows of the new iterate are refreshed after every iteration: the receiving device's
//                  stream waits for the neighbour's "iteration done" event and pulls the rows with
//                  cudaMemcpyPeerAsync (NVLink P2P when peer access is enabled) -- events only, no host
//                  synchronisation inside the iteration loop;
//     vip_mg_create picks deep halo while the redundant rows stay under 35 % of a strip (VIP_MG_FLAG_* overrides);
//   * every device slot has its own host worker thread that enqueues the slot's work (kernels, event records, peer
//     copies): with one enqueueing thread the 8 x 3 x (3 launches + events + 2 peer copies) calls of a BTF run cost
//     more host time than the GPUs need for the work.  Workers only order themselves through one atomic counter per
//     slot ("my iteration-done event has been recorded"), which the neighbours poll before cudaStreamWaitEvent;
//   * frame batches are dealt round-robin to the devices, two in-flight slots per device so that the upload of
//     frame i+1 and the download of frame i-1 overlap the kernels of frame i; no inter-device traffic.
// The same device id may appear several times in the list ("virtual" devices): the strip logic is then
// exercised on a single GPU.
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <climits>
#include <condition_variable>
#include <memory>
#include <mutex>
#include <new>
#include <thread>
#include <vector>

#include "../../include/vip_c_api.h"
#include "vip_internal.h"

extern "C" int vip_btf_create_rows_internal(vip_btf** out, int width, int height, int ksize, int nitr,
                                            int scratch_rows);

namespace {

inline int cu(cudaError_t e) { return static_cast<int>(e); }

// std::atomic<bool> that a std::vector element may be copied with (slots are only resized before any thread runs)
struct SharedFlag {
    std::atomic<bool> v{false};
    SharedFlag() = default;
    SharedFlag(const SharedFlag& o) : v(o.v.load(std::memory_order_relaxed)) {}
    SharedFlag& operator=(const SharedFlag& o) {
        v.store(o.v.load(std::memory_order_relaxed), std::memory_order_relaxed);
        return *this;
    }
};

struct Slot {
    int device = 0;
    int y0 = 0, rows = 0;      // owned rows
    int sy0 = 0, srows = 0;    // rows held by the strip buffers (owned + halos, clipped)
    cudaStream_t stream = nullptr;
    cudaEvent_t ev_start = nullptr, ev_stop = nullptr;
    std::vector<cudaEvent_t> ev_iter;  // BTF: iterate `it` is complete on this slot
    cudaEvent_t ev_pulled = nullptr;   // BTF exchange: this slot's last halo pull of a run is complete
    SharedFlag pulled_valid;           // ev_pulled has been recorded by the previous run (written by this slot's worker, read by the neighbours')
    int peer_up = -1, peer_dn = -1;    // transport to the neighbouring slots: 1 P2P enabled, 0 staged through the host, -1 same device / none
    uint8_t* src = nullptr;            // rows [sy0, sy0+srows)
    uint8_t* buf[2] = {nullptr, nullptr};  // BF/ABF: buf[0] = dst (owned rows only); BTF: ping/pong over [sy0, ..)
    vip_bf* bf = nullptr;
    vip_abf* abf = nullptr;
    vip_btf* btf = nullptr;
    // batch mode (whole frames): two in-flight lanes
    cudaStream_t bstream[2] = {nullptr, nullptr};
    uint8_t* bsrc[2] = {nullptr, nullptr};
    uint8_t* bdst[2][3] = {{nullptr, nullptr, nullptr}, {nullptr, nullptr, nullptr}};   // per lane: BF / ABF / BTF output
    vip_btf* bbtf[2] = {nullptr, nullptr};
};

}  // namespace

struct Worker {
    std::thread th;
    std::mutex mu;
    std::condition_variable cv;
    int cmd_gen = 0, done_gen = 0;  // a run is requested by cmd_gen++, finished when done_gen == cmd_gen
    bool quit = false;
    int rc = VIP_OK;
};

struct vip_mg {
    int n = 0, kind = 0, width = 0, height = 0, ksize = 0, nitr = 0, halo = 0;
    float ss = 0.f, sc = 0.f;
    int result_buf = 0;  // BTF: which of buf[0/1] holds the final iterate (-1: src)
    bool btf_deep = false;  // BTF: deep-halo mode (no exchange), see the header comment
    std::vector<Slot> slots;
    std::vector<std::unique_ptr<Worker>> workers;          // one enqueueing thread per slot
    std::unique_ptr<std::atomic<int>[]> iter_recorded;     // BTF: iterations whose "done" event slot i has recorded
    int unreported_rc = VIP_OK;  // first enqueue error of a run that was followed by another run before anybody asked
};

namespace {

struct DeviceGuard {
    int prev = 0;
    DeviceGuard() { cudaGetDevice(&prev); }
    ~DeviceGuard() { cudaSetDevice(prev); }
};

int free_slot(vip_mg* m, Slot& s) {
    cudaSetDevice(s.device);
    if (s.bf) vip_bf_destroy(s.bf);
    if (s.abf) vip_abf_destroy(s.abf);
    if (s.btf) vip_btf_destroy(s.btf);
    for (int i = 0; i < 2; i++) {
        if (s.bbtf[i]) vip_btf_destroy(s.bbtf[i]);
        if (s.bsrc[i]) cudaFree(s.bsrc[i]);
        for (int j = 0; j < 3; j++)
            if (s.bdst[i][j]) cudaFree(s.bdst[i][j]);
        if (s.bstream[i]) cudaStreamDestroy(s.bstream[i]);
        if (s.buf[i]) cudaFree(s.buf[i]);
    }
    if (s.src) cudaFree(s.src);
    for (cudaEvent_t e : s.ev_iter) cudaEventDestroy(e);
    if (s.ev_pulled) cudaEventDestroy(s.ev_pulled);
    if (s.ev_start) cudaEventDestroy(s.ev_start);
    if (s.ev_stop) cudaEventDestroy(s.ev_stop);
    if (s.stream) cudaStreamDestroy(s.stream);
    (void)m;
    return VIP_OK;
}

int run_one(vip_mg* m, Slot& s, const uint8_t* src, uint8_t* dst, const vip_strip& st) {
    switch (m->kind) {
        case VIP_MG_BF: return vip_bf_run_strip(s.bf, src, dst, &st, s.stream);
        case VIP_MG_ABF: return vip_abf_run_strip(s.abf, src, dst, &st, s.stream);
        default: return VIP_E_UNSUPPORTED;
    }
}

// The strip plan (host logic only; vip_mg_plan_query exposes it so that it can be checked without devices).
int plan_slot(int n_devices, int slot, int kind, int height, int ksize, int nitr, int flags, vip_mg_plan* out) {
    if (n_devices < 1 || n_devices > 64 || slot < 0 || slot >= n_devices) return VIP_E_DEVICE;
    if (kind != VIP_MG_BF && kind != VIP_MG_ABF && kind != VIP_MG_BTF) return VIP_E_UNSUPPORTED;
    if (height <= 0) return VIP_E_SIZE;
    if (ksize < 1 || (ksize & 1) == 0) return VIP_E_KSIZE;
    if (kind == VIP_MG_BTF && nitr < 0) return VIP_E_UNSUPPORTED;
    if ((flags & VIP_MG_FLAG_BTF_EXCHANGE) && (flags & VIP_MG_FLAG_BTF_DEEP_HALO)) return VIP_E_UNSUPPORTED;
    if (flags & ~(VIP_MG_FLAG_BTF_EXCHANGE | VIP_MG_FLAG_BTF_DEEP_HALO | VIP_MG_FLAG_REQUIRE_P2P)) return VIP_E_UNSUPPORTED;
    const int halo = (kind == VIP_MG_BTF) ? (ksize - 1) + 2 * (ksize / 2) + 1 : ksize / 2;
    bool deep = false;
    if (kind == VIP_MG_BTF && n_devices > 1 && nitr > 1) {
        // redundant rows of the deep-halo mode, averaged over the iterations: (nitr-1) x halo per strip
        const long long extra = static_cast<long long>(nitr - 1) * halo, strip = height / n_devices;
        deep = (flags & VIP_MG_FLAG_BTF_DEEP_HALO) || (!(flags & VIP_MG_FLAG_BTF_EXCHANGE) && extra * 100 <= strip * 35);
    }
    // BTF halo exchange: every strip must be able to serve its neighbour's halo from its own rows (BF / ABF and the
    // deep-halo mode take their halos from the scattered source, any strip height works)
    if (kind == VIP_MG_BTF && n_devices > 1 && nitr > 1 && !deep && height / n_devices < halo) return VIP_E_SIZE;
    const int buf_halo = deep ? nitr * halo : halo;  // rows the strip buffers carry on each side
    out->y0 = static_cast<int>(static_cast<long long>(height) * slot / n_devices);
    out->rows = static_cast<int>(static_cast<long long>(height) * (slot + 1) / n_devices) - out->y0;
    out->buf_y0 = std::max(0, out->y0 - buf_halo);
    out->buf_rows = std::min(height, out->y0 + out->rows + buf_halo) - out->buf_y0;
    out->halo = halo;
    out->btf_deep_halo = deep ? 1 : 0;
    return VIP_OK;
}

// Everything slot i contributes to one run, enqueued on its stream by its worker thread.
int enqueue_slot(vip_mg* m, int i) {
    Slot& s = m->slots[i];
    const size_t row_bytes = static_cast<size_t>(m->width) * 3;
    cudaError_t e;
    if ((e = cudaSetDevice(s.device)) != cudaSuccess) return cu(e);
    if ((e = cudaEventRecord(s.ev_start, s.stream)) != cudaSuccess) return cu(e);
    if (m->kind != VIP_MG_BTF) {
        if (s.rows > 0) {  // more slots than image rows: this one owns nothing
            const vip_strip st{s.sy0, s.srows, s.y0, s.y0, s.rows};
            const int rc = run_one(m, s, s.src, s.buf[0], st);
            if (rc != VIP_OK) return rc;
        }
    } else {
        auto wait_recorded = [&](int j, int want) {  // the neighbour's event must be RECORDED before it can be waited on
            while (m->iter_recorded[j].load(std::memory_order_acquire) < want) std::this_thread::yield();
        };
        // Exchange mode, run N+1 after run N without a gather in between: the neighbours' last halo pulls of run N read
        // this slot's ping / pong buffers; they must be complete before iteration 0 overwrites buf[0].  Their events were
        // recorded while run N was being enqueued, which vip_mg_run_async has waited for.
        if (!m->btf_deep) {
            for (int j = i - 1; j <= i + 1; j += 2) {
                if (j < 0 || j >= m->n || !m->slots[j].pulled_valid.v.load(std::memory_order_acquire)) continue;
                if ((e = cudaStreamWaitEvent(s.stream, m->slots[j].ev_pulled, 0)) != cudaSuccess) return cu(e);
            }
        }
        bool pulled = false;
        for (int it = 0; it < m->nitr; it++) {
            const int ob = it & 1;
            const uint8_t* in = (it == 0) ? s.src : s.buf[ob ^ 1];
            // deep halo: this iterate must also cover what the remaining iterations will read around the owned rows
            const int margin = m->btf_deep ? (m->nitr - 1 - it) * m->halo : 0;
            const int oy0 = std::max(0, s.y0 - margin), oy1 = std::min(m->height, s.y0 + s.rows + margin);
            const vip_strip st{s.sy0, s.srows, s.sy0, oy0, oy1 - oy0};
            const int rc = vip_btf_iterate_strip(s.btf, in, s.buf[ob], &st, s.stream);
            if (rc != VIP_OK) return rc;
            if (m->btf_deep) continue;
            if ((e = cudaEventRecord(s.ev_iter[it], s.stream)) != cudaSuccess) return cu(e);
            m->iter_recorded[i].store(it + 1, std::memory_order_release);
            if (it + 1 < m->nitr) {
                // halo refresh of the new iterate: pull the rows [sy0, y0) from the slot above and
                // [y0+rows, sy0+srows) from the slot below, on the RECEIVER's stream
                if (i > 0 && s.y0 > s.sy0) {
                    Slot& up = m->slots[i - 1];
                    wait_recorded(i - 1, it + 1);
                    if ((e = cudaStreamWaitEvent(s.stream, up.ev_iter[it], 0)) != cudaSuccess) return cu(e);
                    e = cudaMemcpyPeerAsync(s.buf[ob], s.device, up.buf[ob] + row_bytes * (s.sy0 - up.sy0), up.device,
                                            row_bytes * (s.y0 - s.sy0), s.stream);
                    if (e != cudaSuccess) return cu(e);
                    pulled = true;
                }
                const int below0 = s.y0 + s.rows, below1 = s.sy0 + s.srows;
                if (i + 1 < m->n && below1 > below0) {
                    Slot& dn = m->slots[i + 1];
                    wait_recorded(i + 1, it + 1);
                    if ((e = cudaStreamWaitEvent(s.stream, dn.ev_iter[it], 0)) != cudaSuccess) return cu(e);
                    e = cudaMemcpyPeerAsync(s.buf[ob] + row_bytes * (below0 - s.sy0), s.device,
                                            dn.buf[ob] + row_bytes * (below0 - dn.sy0), dn.device,
                                            row_bytes * (below1 - below0), s.stream);
                    if (e != cudaSuccess) return cu(e);
                    pulled = true;
                }
            }
        }
        if (pulled) {
            if ((e = cudaEventRecord(s.ev_pulled, s.stream)) != cudaSuccess) return cu(e);
        }
        s.pulled_valid.v.store(pulled, std::memory_order_release);
    }
    if ((e = cudaEventRecord(s.ev_stop, s.stream)) != cudaSuccess) return cu(e);
    return VIP_OK;
}

void worker_main(vip_mg* m, Worker* wp, int i) {
    Worker& w = *wp;
    for (;;) {
        int gen;
        {
            std::unique_lock<std::mutex> lock(w.mu);
            w.cv.wait(lock, [&] { return w.quit || w.cmd_gen != w.done_gen; });
            if (w.quit) return;
            gen = w.cmd_gen;
        }
        const int rc = enqueue_slot(m, i);
        if (rc != VIP_OK) m->iter_recorded[i].store(INT_MAX, std::memory_order_release);  // never block the neighbours
        {
            std::lock_guard<std::mutex> lock(w.mu);
            w.rc = rc;
            w.done_gen = gen;
        }
        w.cv.notify_all();
    }
}

// Blocks until every worker has finished ENQUEUEING the current run; returns the first error.
int wait_workers(vip_mg* m) {
    int rc = VIP_OK;
    for (auto& wp : m->workers) {
        std::unique_lock<std::mutex> lock(wp->mu);
        wp->cv.wait(lock, [&] { return wp->cmd_gen == wp->done_gen; });
        if (rc == VIP_OK) rc = wp->rc;
    }
    return rc;
}

// wait_workers + hand out, exactly once, the first error since the last report: a run whose enqueueing failed and that
// was followed by another vip_mg_run_async (no wait / gather in between) is reported by the next wait / gather.
int take_run_status(vip_mg* m) {
    int rc = VIP_OK;
    for (auto& wp : m->workers) {
        std::unique_lock<std::mutex> lock(wp->mu);
        wp->cv.wait(lock, [&] { return wp->cmd_gen == wp->done_gen; });
        if (rc == VIP_OK) rc = wp->rc;
        wp->rc = VIP_OK;
    }
    if (m->unreported_rc != VIP_OK) rc = m->unreported_rc;   // the older error first
    m->unreported_rc = VIP_OK;
    return rc;
}

void stop_workers(vip_mg* m) {
    for (auto& wp : m->workers) {
        {
            std::lock_guard<std::mutex> lock(wp->mu);
            wp->quit = true;
        }
        wp->cv.notify_all();
        if (wp->th.joinable()) wp->th.join();
    }
    m->workers.clear();
}

}  // namespace

extern "C" {

int vip_mg_create(vip_mg** out, const int* device_ids, int n_devices, int kind, int width, int height, int ksize,
                  float sigma_space, float sigma_color, int nitr) {
    return vip_mg_create_ex(out, device_ids, n_devices, kind, width, height, ksize, sigma_space, sigma_color, nitr, 0);
}

int vip_mg_create_ex(vip_mg** out, const int* device_ids, int n_devices, int kind, int width, int height, int ksize,
                     float sigma_space, float sigma_color, int nitr, int flags) {
    if (!out || !device_ids) return VIP_E_NULL;
    *out = nullptr;
    if (n_devices < 1 || n_devices > 64) return VIP_E_DEVICE;
    if (kind != VIP_MG_BF && kind != VIP_MG_ABF && kind != VIP_MG_BTF) return VIP_E_UNSUPPORTED;
    if (width <= 0 || height <= 0) return VIP_E_SIZE;
    if (ksize < 1 || (ksize & 1) == 0) return VIP_E_KSIZE;
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess) return cu(e);
    for (int i = 0; i < n_devices; i++)
        if (device_ids[i] < 0 || device_ids[i] >= ndev) return VIP_E_DEVICE;
    vip_mg_plan plan0;
    const int prc = plan_slot(n_devices, 0, kind, height, ksize, nitr, flags, &plan0);
    if (prc != VIP_OK) return prc;
    const int halo = plan0.halo;
    const bool deep = plan0.btf_deep_halo != 0;

    DeviceGuard guard;
    vip_mg* m = new (std::nothrow) vip_mg();
    if (!m) return VIP_E_ALLOC;
    m->n = n_devices;
    m->kind = kind;
    m->width = width;
    m->height = height;
    m->ksize = ksize;
    m->nitr = nitr;
    m->halo = halo;
    m->btf_deep = deep;
    m->ss = sigma_space;
    m->sc = sigma_color;
    m->slots.resize(n_devices);
    int rc = VIP_OK;
    const size_t row_bytes = static_cast<size_t>(width) * 3;
    for (int i = 0; i < n_devices && rc == VIP_OK; i++) {
        Slot& s = m->slots[i];
        s.device = device_ids[i];
        vip_mg_plan pl;
        plan_slot(n_devices, i, kind, height, ksize, nitr, flags, &pl);
        s.y0 = pl.y0;
        s.rows = pl.rows;
        s.sy0 = pl.buf_y0;
        s.srows = pl.buf_rows;
        if ((e = cudaSetDevice(s.device)) != cudaSuccess) { rc = cu(e); break; }
        // P2P with the neighbouring slots (an error only means "already enabled" or "not supported":
        // cudaMemcpyPeerAsync then falls back to staging through the host)
        for (int j = i - 1; j <= i + 1; j += 2) {
            if (j < 0 || j >= n_devices || device_ids[j] == s.device) continue;
            int can = 0, ok = 0;
            if (cudaDeviceCanAccessPeer(&can, s.device, device_ids[j]) == cudaSuccess && can) {
                const cudaError_t pe = cudaDeviceEnablePeerAccess(device_ids[j], 0);
                ok = (pe == cudaSuccess || pe == cudaErrorPeerAccessAlreadyEnabled) ? 1 : 0;
            }
            cudaGetLastError();
            (j < i ? s.peer_up : s.peer_dn) = ok;
            // without peer access cudaMemcpyPeerAsync stages through the host: allowed unless the caller asked for P2P
            if (!ok && (flags & VIP_MG_FLAG_REQUIRE_P2P)) { rc = VIP_E_DEVICE; break; }
        }
        if (rc != VIP_OK) break;
        if ((e = cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking)) != cudaSuccess) { rc = cu(e); break; }
        if ((e = cudaEventCreate(&s.ev_start)) != cudaSuccess) { rc = cu(e); break; }
        if ((e = cudaEventCreate(&s.ev_stop)) != cudaSuccess) { rc = cu(e); break; }
        if ((e = cudaMalloc(reinterpret_cast<void**>(&s.src), row_bytes * s.srows)) != cudaSuccess) { rc = cu(e); break; }
        if (kind == VIP_MG_BTF) {
            for (int b = 0; b < 2; b++)
                if ((e = cudaMalloc(reinterpret_cast<void**>(&s.buf[b]), row_bytes * s.srows)) != cudaSuccess) break;
            if (e != cudaSuccess) { rc = cu(e); break; }
            s.ev_iter.resize(std::max(nitr, 1));
            for (auto& ev : s.ev_iter)
                if ((e = cudaEventCreateWithFlags(&ev, cudaEventDisableTiming)) != cudaSuccess) break;
            if (e != cudaSuccess) { rc = cu(e); break; }
            if ((e = cudaEventCreateWithFlags(&s.ev_pulled, cudaEventDisableTiming)) != cudaSuccess) { rc = cu(e); break; }
            rc = vip_btf_create_rows_internal(&s.btf, width, height, ksize, nitr, s.srows);
        } else {
            if ((e = cudaMalloc(reinterpret_cast<void**>(&s.buf[0]), row_bytes * s.rows)) != cudaSuccess) { rc = cu(e); break; }
            if (kind == VIP_MG_BF) rc = vip_bf_create(&s.bf, width, height, ksize, sigma_space, sigma_color);
            else rc = vip_abf_create(&s.abf, width, height, ksize, sigma_space, sigma_color);
        }
    }
    if (rc != VIP_OK) {
        vip_mg_destroy(m);
        return rc;
    }
    m->iter_recorded.reset(new (std::nothrow) std::atomic<int>[n_devices]);
    if (!m->iter_recorded) {
        vip_mg_destroy(m);
        return VIP_E_ALLOC;
    }
    for (int i = 0; i < n_devices; i++) m->iter_recorded[i].store(0);
    // all Worker objects exist before the first thread starts: a running worker reads m->workers[i]
    for (int i = 0; i < n_devices; i++) m->workers.emplace_back(new Worker());
    for (int i = 0; i < n_devices; i++) m->workers[i]->th = std::thread(worker_main, m, m->workers[i].get(), i);
    *out = m;
    return VIP_OK;
}

int vip_mg_plan_query(int n_devices, int slot, int kind, int height, int ksize, int nitr, int flags, vip_mg_plan* out) {
    if (!out) return VIP_E_NULL;
    return plan_slot(n_devices, slot, kind, height, ksize, nitr, flags, out);
}

int vip_mg_destroy(vip_mg* m) {
    if (!m) return VIP_OK;
    DeviceGuard guard;
    wait_workers(m);
    stop_workers(m);
    for (Slot& s : m->slots) free_slot(m, s);
    delete m;
    return VIP_OK;
}

int vip_mg_strip_rows(const vip_mg* m, int i, int* y0, int* rows) {
    if (!m) return VIP_E_NULL;
    if (i < 0 || i >= m->n) return VIP_E_DEVICE;
    if (y0) *y0 = m->slots[i].y0;
    if (rows) *rows = m->slots[i].rows;
    return VIP_OK;
}

int vip_mg_scatter_host(vip_mg* m, const uint8_t* h_src) {
    if (!m || !h_src) return VIP_E_NULL;
    wait_workers(m);
    DeviceGuard guard;
    const size_t row_bytes = static_cast<size_t>(m->width) * 3;
    for (Slot& s : m->slots) {
        cudaError_t e = cudaSetDevice(s.device);
        if (e == cudaSuccess)
            e = cudaMemcpyAsync(s.src, h_src + row_bytes * s.sy0, row_bytes * s.srows, cudaMemcpyHostToDevice, s.stream);
        if (e != cudaSuccess) return cu(e);
    }
    return VIP_OK;
}

int vip_mg_scatter_device(vip_mg* m, const uint8_t* d_src_on_dev0) {
    if (!m || !d_src_on_dev0) return VIP_E_NULL;
    wait_workers(m);
    DeviceGuard guard;
    const size_t row_bytes = static_cast<size_t>(m->width) * 3;
    const int dev0 = m->slots[0].device;
    for (Slot& s : m->slots) {
        cudaError_t e = cudaSetDevice(s.device);
        if (e == cudaSuccess)
            e = cudaMemcpyPeerAsync(s.src, s.device, d_src_on_dev0 + row_bytes * s.sy0, dev0, row_bytes * s.srows, s.stream);
        if (e != cudaSuccess) return cu(e);
    }
    return VIP_OK;
}

int vip_mg_run_async(vip_mg* m) {
    if (!m) return VIP_E_NULL;
    m->unreported_rc = take_run_status(m);  // a previous run still being enqueued; its error, if nobody has asked yet, stays pending
    for (int i = 0; i < m->n; i++) m->iter_recorded[i].store(0, std::memory_order_release);
    if (m->kind == VIP_MG_BTF) m->result_buf = m->nitr > 0 ? ((m->nitr - 1) & 1) : -1;
    for (auto& wp : m->workers) {
        {
            std::lock_guard<std::mutex> lock(wp->mu);
            wp->rc = VIP_OK;
            wp->cmd_gen++;
        }
        wp->cv.notify_all();
    }
    return VIP_OK;
}

int vip_mg_wait(vip_mg* m) {
    if (!m) return VIP_E_NULL;
    const int rc = take_run_status(m);
    if (rc != VIP_OK) return rc;
    DeviceGuard guard;
    for (Slot& s : m->slots) {
        cudaError_t e = cudaSetDevice(s.device);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s.stream);
        if (e != cudaSuccess) return cu(e);
    }
    return VIP_OK;
}

int vip_mg_run(vip_mg* m) {
    const int rc = vip_mg_run_async(m);
    if (rc != VIP_OK) return rc;
    return vip_mg_wait(m);
}

int vip_mg_gather_host(vip_mg* m, uint8_t* h_dst) {
    if (!m || !h_dst) return VIP_E_NULL;
    const int wrc = take_run_status(m);
    if (wrc != VIP_OK) return wrc;
    DeviceGuard guard;
    const size_t row_bytes = static_cast<size_t>(m->width) * 3;
    for (Slot& s : m->slots) {
        cudaError_t e = cudaSetDevice(s.device);
        if (e != cudaSuccess) return cu(e);
        const uint8_t* from;
        if (m->kind != VIP_MG_BTF) from = s.buf[0];
        else if (m->result_buf < 0) from = s.src + row_bytes * (s.y0 - s.sy0);
        else from = s.buf[m->result_buf] + row_bytes * (s.y0 - s.sy0);
        if (s.rows == 0) continue;
        e = cudaMemcpyAsync(h_dst + row_bytes * s.y0, from, row_bytes * s.rows, cudaMemcpyDeviceToHost, s.stream);
        if (e != cudaSuccess) return cu(e);
    }
    return vip_mg_wait(m);
}

int vip_mg_last_run_ms(const vip_mg* m, float* ms) {
    if (!m || !ms) return VIP_E_NULL;
    wait_workers(const_cast<vip_mg*>(m));
    DeviceGuard guard;
    float worst = 0.f;
    for (const Slot& s : m->slots) {
        cudaError_t e = cudaSetDevice(s.device);
        float t = 0.f;
        if (e == cudaSuccess) e = cudaEventSynchronize(s.ev_stop);
        if (e == cudaSuccess) e = cudaEventElapsedTime(&t, s.ev_start, s.ev_stop);
        if (e != cudaSuccess) return cu(e);
        worst = std::max(worst, t);
    }
    *ms = worst;
    return VIP_OK;
}

// Lazy, idempotent set-up of the frame-batch lanes of one slot for the filters in `set` (a failure leaves what exists
// in place; the next call continues where this one stopped).
static int batch_setup_slot(vip_mg* m, Slot& s, int set) {
    const size_t frame_bytes = static_cast<size_t>(m->width) * m->height * 3;
    cudaError_t e;
    if ((e = cudaSetDevice(s.device)) != cudaSuccess) return cu(e);
    for (int l = 0; l < 2; l++) {
        if (!s.bstream[l] && (e = cudaStreamCreateWithFlags(&s.bstream[l], cudaStreamNonBlocking)) != cudaSuccess) return cu(e);
        if (!s.bsrc[l] && (e = cudaMalloc(reinterpret_cast<void**>(&s.bsrc[l]), frame_bytes)) != cudaSuccess) return cu(e);
        for (int j = 0; j < 3; j++)
            if ((set >> j & 1) && !s.bdst[l][j] &&
                (e = cudaMalloc(reinterpret_cast<void**>(&s.bdst[l][j]), frame_bytes)) != cudaSuccess)
                return cu(e);
        if ((set & VIP_MG_SET_BTF) && !s.bbtf[l]) {
            const int rc = vip_btf_create(&s.bbtf[l], m->width, m->height, m->ksize, m->nitr);
            if (rc != VIP_OK) return rc;
        }
    }
    if ((set & VIP_MG_SET_BF) && !s.bf) {
        const int rc = vip_bf_create(&s.bf, m->width, m->height, m->ksize, m->ss, m->sc);
        if (rc != VIP_OK) return rc;
    }
    if ((set & VIP_MG_SET_ABF) && !s.abf) {
        const int rc = vip_abf_create(&s.abf, m->width, m->height, m->ksize, m->ss, m->sc);
        if (rc != VIP_OK) return rc;
    }
    return VIP_OK;
}

// Frames i, i + n, i + 2n, ... on slot i: upload ONCE, every filter of the set, one download per filter; two lanes so
// that the copies of one frame overlap the kernels of the other.  Runs on its own host thread (one per device).
static int batch_run_slot(vip_mg* m, int i, int set, const uint8_t* h_src, uint8_t* const* h_dst, int n_frames) {
    Slot& s = m->slots[i];
    const size_t frame_bytes = static_cast<size_t>(m->width) * m->height * 3;
    cudaError_t e;
    int rc = batch_setup_slot(m, s, set);
    if (rc != VIP_OK) return rc;
    // On an error the loop stops, but copies into the caller's h_dst that are already enqueued must have landed before
    // this returns (the caller may free or reuse that memory): both lanes are synchronised on every path.
    int lane = 0;
    for (int f = i; f < n_frames && rc == VIP_OK; f += m->n, lane ^= 1) {
        cudaStream_t st = s.bstream[lane];
        if ((e = cudaMemcpyAsync(s.bsrc[lane], h_src + frame_bytes * f, frame_bytes, cudaMemcpyHostToDevice, st)) != cudaSuccess) {
            rc = cu(e);
            break;
        }
        if (rc == VIP_OK && (set & VIP_MG_SET_BF)) rc = vip_bf_run(s.bf, s.bsrc[lane], s.bdst[lane][0], st);
        if (rc == VIP_OK && (set & VIP_MG_SET_ABF)) rc = vip_abf_run(s.abf, s.bsrc[lane], s.bdst[lane][1], st);
        if (rc == VIP_OK && (set & VIP_MG_SET_BTF)) rc = vip_btf_run(s.bbtf[lane], s.bsrc[lane], s.bdst[lane][2], st);
        for (int j = 0; j < 3 && rc == VIP_OK; j++)
            if (set >> j & 1)
                if ((e = cudaMemcpyAsync(h_dst[j] + frame_bytes * f, s.bdst[lane][j], frame_bytes, cudaMemcpyDeviceToHost, st)) !=
                    cudaSuccess)
                    rc = cu(e);
    }
    for (int l = 0; l < 2; l++)
        if (s.bstream[l] && (e = cudaStreamSynchronize(s.bstream[l])) != cudaSuccess && rc == VIP_OK) rc = cu(e);
    return rc;
}

int vip_mg_run_batch_set_host(vip_mg* m, int filter_set, const uint8_t* h_src, uint8_t* const* h_dst, int n_frames) {
    if (!m || !h_src || !h_dst) return VIP_E_NULL;
    if (n_frames < 0) return VIP_E_SIZE;
    if (filter_set <= 0 || filter_set > 7) return VIP_E_UNSUPPORTED;
    for (int j = 0; j < 3; j++)
        if ((filter_set >> j & 1) && !h_dst[j]) return VIP_E_NULL;
    if ((filter_set & VIP_MG_SET_BTF) && (m->ksize < 3 || m->nitr < 0)) return VIP_E_KSIZE;
    wait_workers(m);
    DeviceGuard guard;
    std::vector<int> rcs(m->n, VIP_OK);
    std::vector<std::thread> th;
    th.reserve(m->n);
    for (int i = 0; i < m->n; i++)
        th.emplace_back([=, &rcs] { rcs[i] = batch_run_slot(m, i, filter_set, h_src, h_dst, n_frames); });
    for (auto& t : th) t.join();
    for (int rc : rcs)
        if (rc != VIP_OK) return rc;
    return VIP_OK;
}

int vip_mg_run_batch_host(vip_mg* m, const uint8_t* h_src, uint8_t* h_dst, int n_frames) {
    if (!m) return VIP_E_NULL;
    uint8_t* dst[3] = {nullptr, nullptr, nullptr};
    dst[m->kind] = h_dst;   // VIP_MG_BF / ABF / BTF = 0 / 1 / 2 = the bit positions of VIP_MG_SET_*
    return vip_mg_run_batch_set_host(m, 1 << m->kind, h_src, dst, n_frames);
}

int vip_mg_peer_access(const vip_mg* m) {
    if (!m) return VIP_E_NULL;
    int all = 1;
    for (const Slot& s : m->slots)
        if (s.peer_up == 0 || s.peer_dn == 0) all = 0;
    return all;
}

}  // extern "C"
