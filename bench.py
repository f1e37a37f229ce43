#!/usr/bin/env python
"""Benchmark of the edge-preserving-filter hot path (BASELINE.json).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]

Default workload (N = 1): BASELINE.json configs[1] -- adaptive bilateral filter, radius 7 (ksize 15), on a synthetic
3840x2160 RGB image, one B200.  A "step" is one pass of the filter over one frame.  At N > 1 (torchrun, one process
per GPU) every rank filters its own frames (frames are independent: no data-path collective, weak scaling) and the
time is the max over ranks.  Other workloads (--workload): bf_8k_r15 (configs[3]; at N > 1 the image is cut into row
strips with radius-deep halos, strong scaling), btf_4k (configs[2]), bf_lenna (configs[0] geometry), bf_1080p /
abf_1080p / btf_1080p (configs[4] frames).

One JSON line on stdout.  `value` = whole-job Mpixel/s with inputs resident in HBM; `e2e` = the same through the
public API with pinned HOST buffers, H2D + kernel + D2H inside the timed region, four frames in flight; `roofline` =
the dominant kernel against the measured HBM peak (MEASURED_PEAKS.json) -- the stencil kernels are issue-bound, not
HBM-bound, so `issue` gives the achieved taps/s and the estimated issue-slot utilisation next to it; `cpu_baseline`
= the reference's own C++ CPU path (oracle/_ref/libvip_ref_cpu.so, compiled unmodified from the reference) on a
bounded sample, on this box's host cores.

--impl reference times the reference's CPU implementation of the same workload (bounded sample per step).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {
    # name: (filter, width, height, ksize, nitr, description)
    "abf_4k_r7": ("abf", 3840, 2160, 15, 0, "adaptive bilateral filter radius=7 on synthetic 3840x2160 RGB (BASELINE configs[1])"),
    "btf_4k": ("btf", 3840, 2160, 9, 3, "bilateral texture filter k=9, 3 iterations on synthetic 3840x2160 RGB (configs[2])"),
    "bf_8k_r15": ("bf", 7680, 4320, 31, 0, "bilateral filter radius=15 on synthetic 7680x4320 RGB, row strips + halos (configs[3])"),
    "bf_lenna": ("bf", 512, 512, 11, 0, "bilateral filter radius=5 on sample_image/lenna.png 512x512 BGR (configs[0]; committed fixture)"),
    "bf_4k_r5": ("bf", 3840, 2160, 11, 0, "bilateral filter radius=5 on synthetic 3840x2160 RGB"),
    "bf_1080p": ("bf", 1920, 1080, 9, 0, "bilateral filter k=9 on 1080p frames (configs[4])"),
    "abf_1080p": ("abf", 1920, 1080, 9, 0, "adaptive bilateral filter k=9 on 1080p frames (configs[4])"),
    "btf_1080p": ("btf", 1920, 1080, 9, 3, "bilateral texture filter k=9 nitr=3 on 1080p frames (configs[4])"),
    "stream_1080p": ("all", 1920, 1080, 9, 3, "1080p RGB frame stream through all three filters (BF k=9, ABF k=9, BTF k=9 x3), frames dealt to the GPUs (configs[4]; --steps = frames per GPU)"),
}
L2_BYTES = 126 * 1024 * 1024

# non-zero taps of the circular mask and the per-tap instruction estimates of DESIGN.md
def mask_taps(r):
    return sum(1 for y in range(-r, r + 1) for x in range(-r, r + 1) if x * x + y * y <= r * r)


def synth_frame(w, h, seed):
    """S-rand: i.i.d. uniform bytes (content-independent, like the reference's test fixtures)."""
    return np.random.RandomState(seed).randint(0, 256, size=(h, w, 3), dtype=np.uint8)


def lenna_frame():
    """configs[0]: sample_image/lenna.png decoded to BGR (committed fixture, tests/golden/make_golden_lenna.py)."""
    return np.ascontiguousarray(np.load(os.path.join(ROOT, "tests", "golden", "lenna_vectors.npz"))["lenna_bgr"])


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.index)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def mark(self):
        return len(self.rows)

    def stop(self, first=0):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = self.rows[first:] or self.rows
        sm, mx, reasons = [], [], set()
        for r in rows:
            try:
                sm.append(float(r[1]))
                mx.append(float(r[2]))
            except (ValueError, IndexError):
                continue
            for name, col in (("hw_slowdown", 5), ("hw_thermal_slowdown", 6), ("sw_thermal_slowdown", 7), ("sw_power_cap", 8)):
                if len(r) > col and r[col].lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": sorted(reasons), "samples": len(sm)}


def cpu_reference(filt, w, h, ksize, nitr, sample_rows, threads, repeats=1):
    """The reference's own CPU path on a full-width crop of `sample_rows` rows; returns (Mpixel/s, kind, cores, what)."""
    import _oracle as O
    img = np.ascontiguousarray(lenna_frame()[:sample_rows]) if (w, h, filt) == (512, 512, "bf") else synth_frame(w, sample_rows, 42)
    dst = np.empty_like(img)
    if O.have_ref_cpu():
        R = O.ref_cpu()
        kind = "reference"
        cores = threads or (os.cpu_count() or 1)
        R.ref_set_num_threads(cores)
        r_bf = lambda: R.ref_bilateral_filter(img.ctypes.data, dst.ctypes.data, w, sample_rows, ksize, 10.0, 30.0)
        r_abf = lambda: R.ref_adaptive_bilateral_filter(img.ctypes.data, dst.ctypes.data, w, sample_rows, ksize, 10.0, 30.0)
        r_btf = lambda: R.ref_bilateral_texture_filter(img.ctypes.data, dst.ctypes.data, w, sample_rows, ksize, nitr)
        run = {"bf": r_bf, "abf": r_abf, "btf": r_btf, "all": lambda: (r_bf(), r_abf(), r_btf())}[filt]
    else:
        kind = "port"
        L = O.oracle()
        L.vipo_set_num_threads(threads or 0)
        cores = L.vipo_get_num_threads()
        r_bf = lambda: L.vipo_bilateral_filter(img.ctypes.data, dst.ctypes.data, w, sample_rows, ksize, 10.0, 30.0, 1)
        r_abf = lambda: L.vipo_adaptive_bilateral_filter(img.ctypes.data, dst.ctypes.data, w, sample_rows, ksize, 10.0, 30.0, 1)
        r_btf = lambda: L.vipo_bilateral_texture_filter(img.ctypes.data, dst.ctypes.data, w, sample_rows, ksize, nitr, 1)
        run = {"bf": r_bf, "abf": r_abf, "btf": r_btf, "all": lambda: (r_bf(), r_abf(), r_btf())}[filt]
    best = float("inf")
    for _ in range(repeats):
        t0 = time.perf_counter()
        run()
        best = min(best, time.perf_counter() - t0)
    mpix = w * sample_rows / best / 1e6
    what = f"{w}x{sample_rows} full-width crop of the workload frame (seed 42), best of {repeats}, {best * 1e3:.1f} ms"
    return mpix, kind, cores, what, best


def sample_rows_for(filt, w, h, ksize):
    taps = ksize * ksize * {"bf": 1, "abf": 2, "btf": 4, "all": 7}[filt]
    # ~1e9 tap evaluations per sample: about a second on a many-core host, bounded on a small one
    rows = int(1e9 / (taps * w))
    return max(4 * ksize, min(h, rows - rows % 8))


def run_reference_arm(args, wl):
    filt, w, h, ksize, nitr, desc = wl
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    rows = sample_rows_for(filt, w, h, ksize)
    # calibrate once so that warmup + steps stay within ~2 minutes whatever the host is
    _, _, _, _, t_cal = cpu_reference(filt, w, h, ksize, nitr, rows, args.cpu_threads)
    budget_s = 120.0 / max(1, args.warmup + args.steps)
    if t_cal > budget_s:
        rows = max(4 * ksize, int(rows * budget_s / t_cal) // 4 * 4)
    times = []
    cores = kind = what = None
    for i in range(args.warmup + args.steps):
        mp, kind, cores, what, t = cpu_reference(filt, w, h, ksize, nitr, rows, args.cpu_threads)
        if i >= args.warmup:
            times.append(t)
    ms = 1e3 * float(np.mean(times))
    value = w * rows / (ms * 1e-3) / 1e6
    out = {
        "impl": "reference", "metric": f"{filt}_throughput", "value": value, "unit": "Mpixel/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "description": desc, "width": w, "height": h, "ksize": ksize, "nitr": nitr,
                   "sample": f"each step filters a {w}x{rows} full-width crop (bounded sample of the workload frame)"},
        "cpu_baseline": {"value": value, "unit": "Mpixel/s", "cores": cores, "kind": kind, "sample": what},
        "e2e": {"value": value, "unit": "Mpixel/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default="abf_4k_r7", choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-threads", type=int, default=0, help="threads of the CPU baseline (0 = all host cores)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-probe", action="store_true", help="skip the issue-rate probe kernels (for ncu launch lists)")
    ap.add_argument("--content", default="rand", choices=["rand", "smooth"], help="synthetic content")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    wl = WORKLOADS[args.workload]
    if args.impl == "reference":
        run_reference_arm(args, wl)
        return

    import torch
    import torch.distributed as dist
    import various_image_processings_b200 as vip

    filt, w, h, ksize, nitr, desc = wl
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: the product has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    radius = ksize // 2

    # ---- work decomposition
    strong = (filt == "bf" and args.workload == "bf_8k_r15" and world > 1)
    if strong:
        plan = vip.plan_strips(h, world, vip.halo_rows("bf", ksize))[rank]
        y0, y1, s0, s1 = plan
        strip = vip.to_strip(plan)
    else:
        y0, y1, s0, s1 = 0, h, 0, h
        strip = None
    out_rows = y1 - y0
    frame_bytes = w * h * 3
    # inputs resident in HBM; rotate over enough distinct frames that in+out exceed the L2
    n_pool = max(2, int(np.ceil(1.2 * L2_BYTES / (2.0 * w * (s1 - s0) * 3))))
    n_pool = min(n_pool, 16)
    if args.workload == "bf_lenna":
        frames = [lenna_frame()] * n_pool
    elif args.content == "rand":
        frames = [synth_frame(w, h, 42 + 1000 * rank + i) for i in range(n_pool)]
    else:
        import _oracle as O
        frames = [O.smooth_image(w, h, seed=42 + 1000 * rank + i) for i in range(n_pool)]
    d_src = [torch.from_numpy(np.ascontiguousarray(f[s0:s1])).cuda() for f in frames]
    d_dst = [torch.empty((out_rows, w, 3), dtype=torch.uint8, device="cuda") for _ in range(n_pool)]

    if filt == "bf":
        F = vip.CudaBilateralFilter(w, h, ksize)
        if strong:
            step = lambda i: F.bilateral_filter_strip(d_src[i % n_pool], d_dst[i % n_pool], strip)
        else:
            step = lambda i: F.bilateral_filter_async(d_src[i % n_pool], d_dst[i % n_pool])
    elif filt == "abf":
        F = vip.CudaAdaptiveBilateralFilter(w, h, ksize)
        step = lambda i: F.execute_async(d_src[i % n_pool], d_dst[i % n_pool])
    elif filt == "btf":
        F = vip.CudaBilateralTextureFilter(w, h, ksize, nitr)
        step = lambda i: F.execute_async(d_src[i % n_pool], d_dst[i % n_pool])
    else:
        F = (vip.CudaBilateralFilter(w, h, ksize), vip.CudaAdaptiveBilateralFilter(w, h, ksize),
             vip.CudaBilateralTextureFilter(w, h, ksize, nitr))
        d_dst2 = [torch.empty_like(d) for d in d_dst]
        d_dst3 = [torch.empty_like(d) for d in d_dst]

        def step(i):
            j = i % n_pool
            F[0].bilateral_filter_async(d_src[j], d_dst[j])
            F[1].execute_async(d_src[j], d_dst2[j])
            F[2].execute_async(d_src[j], d_dst3[j])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    sampler = ClockSampler(local_rank)
    sampler.start()
    for i in range(args.warmup):
        step(i)
    barrier()
    mark = sampler.mark()
    n0 = vip.launch_count()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record()
    for i in range(args.steps):
        step(i)
    ev1.record()
    barrier()
    launches = vip.launch_count() - n0
    ms_total = ev0.elapsed_time(ev1)
    clocks = sampler.stop(mark)
    if world > 1:
        t = torch.tensor([ms_total], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_total = float(t.item())
    ms_step = ms_total / args.steps
    pixels_step = (w * h) if strong else (w * h * world)   # strong: one image over all ranks; weak: one frame per rank
    value = pixels_step / (ms_step * 1e-3) / 1e6

    # ---- end to end through the public API with pinned host buffers, two frames in flight
    e2e_steps = max(8, min(args.steps, 40))
    slots = []
    NS = 4   # frames in flight: upload, filter and download of different frames overlap on NS streams
    for s in range(NS):
        src_img, dst_img = vip.DeviceImage(w, s1 - s0, 3), vip.DeviceImage(w, out_rows, 3)
        hin, hout = vip.PinnedBuffer(w * (s1 - s0) * 3), vip.PinnedBuffer(w * out_rows * 3)
        hin.array[:] = frames[s % n_pool][s0:s1].reshape(-1)
        slots.append((src_img, dst_img, hin, hout, torch.cuda.Stream()))
    if filt == "btf":
        F2 = [F] + [vip.CudaBilateralTextureFilter(w, h, ksize, nitr) for _ in range(NS - 1)]   # scratch is per object
    n_out = 3 if filt == "all" else 1                                 # result images downloaded per step
    if filt == "all":
        F2 = [F[2]] + [vip.CudaBilateralTextureFilter(w, h, ksize, nitr) for _ in range(NS - 1)]
        extra = [[(vip.DeviceImage(w, out_rows, 3), vip.PinnedBuffer(w * out_rows * 3)) for _ in range(2)] for _ in range(NS)]

    def e2e_step(i):
        src_img, dst_img, hin, hout, st = slots[i % NS]
        src_img.upload(hin.array, stream=st)
        if filt == "bf":
            if strong:
                F.bilateral_filter_strip(src_img, dst_img, strip, stream=st)
            else:
                F.bilateral_filter_async(src_img, dst_img, stream=st)
        elif filt == "abf":
            F.execute_async(src_img, dst_img, stream=st)
        elif filt == "btf":
            F2[i % NS].execute_async(src_img, dst_img, stream=st)
        else:
            (d2, h2), (d3, h3) = extra[i % NS]
            F[0].bilateral_filter_async(src_img, dst_img, stream=st)
            F[1].execute_async(src_img, d2, stream=st)
            F2[i % NS].execute_async(src_img, d3, stream=st)
            d2.download(h2.array, stream=st)
            d3.download(h3.array, stream=st)
        dst_img.download(hout.array, stream=st)

    for i in range(NS):
        e2e_step(i)
    barrier()
    t0 = time.perf_counter()
    for i in range(e2e_steps):
        e2e_step(i)
    for sl in slots:
        sl[4].synchronize()
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_value = pixels_step * e2e_steps / e2e_s / 1e6

    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return

    # ---- roofline of the dominant kernel
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)"
    local_pixels = w * out_rows
    kernels_per_step = launches / max(args.steps, 1)
    taps = mask_taps(radius if filt not in ("btf", "all") else ksize - 1)
    if filt in ("btf", "all"):
        alg_bytes = local_pixels * 9.0                    # dominant kernel = joint bilateral: src + guide in, dst out
        # several kernels per step: the dominant one (BTF's joint bilateral: ksize 2k-1, sigma_space k-1, sigma_color
        # sqrt 3, src/bilateral_texture_filter_impl.cu:188) is timed alone, live, on the same resident frames (the
        # filtered frame of the timed region serves as the guide), CUDA events on the launching stream
        J = vip.CudaBilateralFilter(w, h, 2 * ksize - 1, float(ksize - 1), 3.0 ** 0.5)
        d_tmp = torch.empty_like(d_dst[0])
        jstep = lambda i: J.joint_bilateral_filter_async(d_src[i % n_pool], d_dst[i % n_pool], d_tmp)
        for i in range(3):
            jstep(i)
        torch.cuda.synchronize()
        jn = max(10, min(args.steps, 50))
        j0, j1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        j0.record()
        for i in range(jn):
            jstep(i)
        j1.record()
        torch.cuda.synchronize()
        kernel_ms = j0.elapsed_time(j1) / jn
        instr_per_tap = 8.5                               # JBF 17x17: smsp__inst_executed x 32 / taps (ncu, profiles/r01d)
        packed_per_tap = 2.0
        taps_per_px = taps * nitr + (2 * mask_taps(radius) if filt == "all" else 0)
    else:
        alg_bytes = local_pixels * 6.0                    # 3 B in + 3 B out per pixel
        kernel_ms = ms_step
        instr_per_tap = 18.1 if filt == "abf" else 7.8   # whole-kernel SASS instructions per tap measured by ncu (profiles/r01d; tile staging, box sums and epilogue included)
        packed_per_tap = 4.2 if filt == "abf" else 2.0   # fp32x2 instructions among them (SASS of the tap loop): ABF 6 FADD2 + 4 FFMA2 per tap PAIR less the scalar edge taps; BF/JBF 1 FFMA2 {s0,s1} + 2 per pair {s2},{sk}
        taps_per_px = taps
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get(args.workload)
    except (OSError, ValueError):
        pass
    roof = {"bound": "hbm", "unit": "GB/s", "peak": hbm_peak, "peak_source": peak_src, "traffic": traffic,
            "kernel": {"bf": "stencil_kernel<BF>", "abf": "stencil_kernel<ABF>", "btf": "stencil_kernel<JBF>", "all": "stencil_kernel<JBF>"}[filt],
            "note": "the bilateral stencils are instruction-issue bound (see 'issue'); HBM traffic is 6-9 B/pixel"}
    roof["achieved"] = alg_bytes / (kernel_ms * 1e-3) / 1e9
    roof["frac"] = roof["achieved"] / hbm_peak
    roof["kernel_ms"] = kernel_ms
    sm_mhz = clocks.get("sm_mhz") or float(peaks.get("sm_max_mhz", 1965.0))
    sm_count = torch.cuda.get_device_properties(local_rank).multi_processor_count
    taps_s = local_pixels * taps / (kernel_ms * 1e-3)    # of the dominant kernel, from its own live timing
    issue_peak = sm_count * 4 * 32 * sm_mhz * 1e6        # thread-instructions / s at the sampled clock
    issue = {"kernel": roof["kernel"], "kernel_ms": kernel_ms, "taps_per_pixel": taps, "step_taps_per_pixel": taps_per_px,
             "gtaps_per_s": taps_s / 1e9, "est_instr_per_tap": instr_per_tap,
             "est_issue_util": taps_s * instr_per_tap / issue_peak, "sm_mhz_used": sm_mhz,
             "peak_thread_instr_per_s": issue_peak,
             "note": "taps/s x SASS instructions per tap (ncu smsp__inst_executed of the same kernel, profiles/) / (SMs x 4 x 32 x clock); packed fp32x2 instructions count once here but hold the dispatch port twice, see DESIGN.md 3"}
    # measured issue roofs (SURVEY.md 8d asks for an FFMA micro-kernel on the box): tools/probe/issue_probe.cu, a few ms each
    try:
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import peaks as _peaks
        _pl = _peaks.load()
        if _pl is not None and rank == 0 and not args.no_probe:
            ffma, _ = _peaks.measure(_pl, 0)
            ffma2, _ = _peaks.measure(_pl, 1)
            issue["measured_ffma_thread_instr_per_s"] = ffma
            issue["measured_ffma2_packed_instr_per_s"] = ffma2
            # a packed fp32x2 instruction holds the dispatch port for two slots (measured: the FFMA2 rate is half the FFMA
            # rate), so the slots a tap costs are instr + packed; their rate against the measured FFMA rate is the fraction of
            # the stencil's own roofline -- the number to read instead of roofline.frac for these kernels
            issue["packed_instr_per_tap"] = packed_per_tap
            issue["dispatch_slots_per_tap"] = instr_per_tap + packed_per_tap
            issue["frac_of_measured_issue_roof"] = taps_s * (instr_per_tap + packed_per_tap) / ffma
            # the same inside the roofline object, for a reader who looks only there: the bound that binds this kernel
            roof["relevant"] = {"bound": "issue", "unit": "thread-instr/s", "achieved": taps_s * (instr_per_tap + packed_per_tap),
                                "peak": ffma, "peak_source": "measured live (tools/probe/issue_probe.cu, FFMA)",
                                "frac": issue["frac_of_measured_issue_roof"]}
    except Exception as exc:                                 # the probe is optional evidence, never a reason to fail the bench
        issue["probe_error"] = str(exc)

    cpu = None
    if not args.no_cpu_baseline:
        rows = sample_rows_for(filt, w, h, ksize)
        mp, kind, cores, what, _ = cpu_reference(filt, w, h, ksize, nitr, rows, args.cpu_threads, repeats=2)
        cpu = {"value": mp, "unit": "Mpixel/s", "cores": cores, "kind": kind, "sample": what}

    out = {
        "metric": f"{filt}_throughput", "value": value, "unit": "Mpixel/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong" if strong else "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": args.workload, "description": desc, "width": w, "height": h, "ksize": ksize, "nitr": nitr,
                   "content": args.content, "parallelism": ("row strips + halos" if strong else f"{world} independent frame stream(s)"),
                   "l2": f"rotating over {n_pool} distinct resident frames ({2 * n_pool * w * (s1 - s0) * 3 >> 20} MiB in+out > 126 MiB L2)",
                   "e2e": "pinned host buffers, upload + filter + download per step, four frames in flight on four streams"},
        "e2e": {"value": e2e_value, "unit": "Mpixel/s", "h2d_bytes_per_step": w * (s1 - s0) * 3 * world,
                "d2h_bytes_per_step": w * out_rows * 3 * world * n_out, "steps": e2e_steps},
        "gpu_launches": int(launches), "kernels_per_step": kernels_per_step,
        "roofline": roof, "issue": issue, "cpu_baseline": cpu, "clocks": clocks,
    }
    print(json.dumps(out), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
