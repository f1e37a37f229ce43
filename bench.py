#!/usr/bin/env python
"""Benchmark of the edge-preserving-filter hot path (BASELINE.json).

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--workload NAME]
                  [--content rand|narrow|nat|smooth] [--halo fused|copy] [--no-per-workload]

Default workload: BASELINE.json configs[3] -- bilateral filter, radius 15 (ksize 31), on a synthetic 7680x4320 RGB image:
the largest single-GPU configuration the metric is quoted on and the one the 8-GPU target is stated for.  A "step" is one
pass of the filter over one frame.  N = 1: the whole frame on one B200.  N > 1 (torchrun, one process per GPU): the frame
is cut into N row strips (STRONG scaling); a rank keeps only its OWN rows on its device and gets the radius-deep halo rows
from the neighbouring ranks' device memory over NVLink inside the timed region -- either staged by the filter kernel
itself straight from the neighbour's buffer (CUDA IPC pointers, --halo fused, the default for BF / ABF) or pulled by
peer-to-peer row copies (--halo copy; always for BTF, whose strips carry nitr x 17 halo rows and recompute the
overlap).  Outside the timed region the strips are gathered and compared with the single-GPU result, the reference's own
CUDA kernel and oracle spot rows ("strip_parity").

One JSON line on stdout.  `value` = whole-job Mpixel/s with inputs resident in HBM; `e2e` = the same through the public
API from pinned HOST buffers (every step uploads the rank's own rows, filters, downloads its rows; neighbours order
themselves with device-side flags -- vip_link_*); `roofline` = the dominant kernel's ALGORITHMIC work (taps/s x the
minimal per-tap instruction mix of SURVEY.md 8d) against the FP32 issue rate measured live on this GPU, with the HBM
figures next to it; `ref_cuda` = the reference's own kernel recompiled for sm_100a on the same GPU (the kernel to beat);
`cpu_baseline` = the reference's own C++ CPU path on a bounded sample; `per_workload` = the same figures for the other
configured workloads (ABF 4K r=7, BTF 4K k=9 x3, the 1080p all-three-filters frame stream).

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
sys.path.insert(0, os.path.join(ROOT, "tools"))

WORKLOADS = {
    # name: (filter, width, height, ksize, nitr, description)
    "bf_8k_r15": ("bf", 7680, 4320, 31, 0, "bilateral filter radius=15 on synthetic 7680x4320 RGB, row strips + halos at N > 1 (BASELINE configs[3])"),
    "abf_4k_r7": ("abf", 3840, 2160, 15, 0, "adaptive bilateral filter radius=7 on synthetic 3840x2160 RGB (configs[1])"),
    "btf_4k": ("btf", 3840, 2160, 9, 3, "bilateral texture filter k=9, 3 iterations on synthetic 3840x2160 RGB (configs[2])"),
    "bf_lenna": ("bf", 512, 512, 11, 0, "bilateral filter radius=5 on sample_image/lenna.png 512x512 BGR (configs[0]; committed fixture)"),
    "bf_4k_r5": ("bf", 3840, 2160, 11, 0, "bilateral filter radius=5 on synthetic 3840x2160 RGB"),
    "bf_1080p": ("bf", 1920, 1080, 9, 0, "bilateral filter k=9 on 1080p frames (configs[4])"),
    "abf_1080p": ("abf", 1920, 1080, 9, 0, "adaptive bilateral filter k=9 on 1080p frames (configs[4])"),
    "btf_1080p": ("btf", 1920, 1080, 9, 3, "bilateral texture filter k=9 nitr=3 on 1080p frames (configs[4])"),
    "stream_1080p": ("all", 1920, 1080, 9, 3, "1080p RGB frame stream through all three filters (BF k=9, ABF k=9, BTF k=9 x3), frames dealt to the GPUs (configs[4]; a step = one frame per GPU)"),
}
DEFAULT_WORKLOAD = "bf_8k_r15"
PER_WORKLOAD = ["abf_4k_r7", "btf_4k", "stream_1080p"]
L2_BYTES = 126 * 1024 * 1024
NS = 3   # frames in flight of the end-to-end legs


def mask_taps(r):
    return sum(1 for y in range(-r, r + 1) for x in range(-r, r + 1) if x * x + y * y <= r * r)


# minimal thread-instructions per tap (SURVEY.md 8d "Roofline formulas"): LUT LDS, SAD, FMUL, 3 FFMA, FADD + the pixel LDS
# and 3 I2F amortised over the P outputs a thread produces
def i_min(kind, P):
    return {"bf": 7 + 4.0 / P, "jbf": 7 + 5.0 / P, "abf": 15 + 4.0 / P}[kind]


OUTPUTS_PER_THREAD = {"bf": 8, "jbf": 8, "abf": 4}   # csrc/stencil.cu outputs_per_thread()
FLOP_PER_TAP = {"bf": 8, "jbf": 8, "abf": 16}


# ---------------------------------------------------------------- synthetic content
def lenna_frame():
    """configs[0]: sample_image/lenna.png decoded to BGR (committed fixture, tests/golden/make_golden_lenna.py)."""
    return np.ascontiguousarray(np.load(os.path.join(ROOT, "tests", "golden", "lenna_vectors.npz"))["lenna_bgr"])


def synth_frame(w, h, seed, content="rand"):
    """S-rand: i.i.d. uniform bytes (the headline: content-independent, like the reference's test fixtures);
    S-narrow: uniform [100, 120) like the reference benchmark (sample/benchmark/main.cpp:210-213); S-nat: lenna, mirrored
    tiling; smooth: the tests' synthetic natural-ish image."""
    if content == "rand":
        return np.random.RandomState(seed).randint(0, 256, size=(h, w, 3), dtype=np.uint8)
    if content == "narrow":
        return np.random.RandomState(seed).randint(100, 120, size=(h, w, 3), dtype=np.uint8)
    if content == "nat":
        t = lenna_frame()
        t = np.roll(t, seed % 97, axis=1)
        t = np.concatenate([np.concatenate([t, t[:, ::-1]], 1), np.concatenate([t[::-1], t[::-1, ::-1]], 1)], 0)
        return np.ascontiguousarray(np.tile(t, (h // t.shape[0] + 1, w // t.shape[1] + 1, 1))[:h, :w])
    import _oracle as O
    t = O.smooth_image(min(w, 512), min(h, 432), seed=seed)
    return np.ascontiguousarray(np.tile(t, (h // t.shape[0] + 1, w // t.shape[1] + 1, 1))[:h, :w])


class ClockSampler:
    """nvidia-smi clocks / throttle reasons DURING the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.rows, self.proc = index, [], None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "50",
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

    def stop(self, first=0, last=None):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.12)
        self.proc.terminate()
        rows = self.rows[first:last] or self.rows[first:] or self.rows
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


# ---------------------------------------------------------------- the reference's CPU path (baseline / reference arm)
def cpu_reference(filt, w, h, ksize, nitr, sample_rows, threads, repeats=1, content="rand"):
    """The reference's own CPU path on a full-width crop of `sample_rows` rows; returns (Mpixel/s, kind, cores, what, s)."""
    import _oracle as O
    img = np.ascontiguousarray(lenna_frame()[:sample_rows]) if (w, h, filt) == (512, 512, "bf") else synth_frame(w, sample_rows, 42, content)
    dst = np.empty_like(img)
    if O.have_ref_cpu():
        R = O.ref_cpu()
        kind = "reference"
        cores = threads or (os.cpu_count() or 1)
        R.ref_set_num_threads(cores)
        r_bf = lambda: R.ref_bilateral_filter(img.ctypes.data, dst.ctypes.data, w, sample_rows, ksize, 10.0, 30.0)
        r_abf = lambda: R.ref_adaptive_bilateral_filter(img.ctypes.data, dst.ctypes.data, w, sample_rows, ksize, 10.0, 30.0)
        r_btf = lambda: R.ref_bilateral_texture_filter(img.ctypes.data, dst.ctypes.data, w, sample_rows, ksize, nitr)
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


def sample_rows_for(filt, w, h, ksize, budget_taps=1e9, min_rows=0):
    """rows of the full-width crop the CPU legs filter; min_rows = 256 for the headline (a crop shorter than
    that is mostly border rows of a radius-15 window), never more than the frame has"""
    taps = ksize * ksize * {"bf": 1, "abf": 2, "btf": 4, "all": 7}[filt]
    rows = int(budget_taps / (taps * w))
    return min(h, max(4 * ksize, min_rows, rows - rows % 8))


def run_reference_arm(args, wl):
    filt, w, h, ksize, nitr, desc = wl
    if int(os.environ.get("RANK", "0")) != 0:
        return
    rows = sample_rows_for(filt, w, h, ksize, min_rows=256)
    # calibrate once so that warmup + steps stay within ~2 minutes whatever the host is
    _, _, _, _, t_cal = cpu_reference(filt, w, h, ksize, nitr, rows, args.cpu_threads, content=args.content)
    budget_s = 120.0 / max(1, args.warmup + args.steps)
    if t_cal > budget_s:
        rows = max(4 * ksize, int(rows * budget_s / t_cal) // 4 * 4)
    times = []
    cores = kind = what = None
    for i in range(args.warmup + args.steps):
        mp, kind, cores, what, t = cpu_reference(filt, w, h, ksize, nitr, rows, args.cpu_threads, content=args.content)
        if i >= args.warmup:
            times.append(t)
    ms = 1e3 * float(np.mean(times))
    value = w * rows / (ms * 1e-3) / 1e6
    out = {
        "impl": "reference", "metric": f"{filt}_throughput", "value": value, "unit": "Mpixel/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True,
        "scaling": "strong" if filt != "all" else "weak",      # same label as our arm's line for this workload (one frame's work is fixed)
        "vs_baseline": None, "dtype": "f32", "data": "lenna.png" if args.workload == "bf_lenna" else "synthetic",
        "config": {"workload": args.workload, "description": desc, "width": w, "height": h, "ksize": ksize, "nitr": nitr,
                   "content": args.content,
                   "sample": f"each step filters a {w}x{rows} full-width crop (bounded sample of the workload frame); Mpixel/s normalises"},
        "cpu_baseline": {"value": value, "unit": "Mpixel/s", "cores": cores, "kind": kind, "sample": what},
        "e2e": {"value": value, "unit": "Mpixel/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(out), flush=True)


# ---------------------------------------------------------------- host placement
def bind_to_gpu_numa_node(index):
    """Pin this process (and therefore the pages cudaHostAlloc first-touches) to the NUMA node / CPUs the GPU hangs off.
    Returns what was done, for the JSON line."""
    info = {"numa_node": None, "cpus": None}
    try:
        bus = subprocess.run(["nvidia-smi", "--query-gpu=pci.bus_id", "--format=csv,noheader", "-i", str(index)],
                             capture_output=True, text=True, timeout=10).stdout.strip().lower()
        if bus.startswith("00000000:"):
            bus = bus[4:]
        base = f"/sys/bus/pci/devices/{bus}"
        node = int(open(f"{base}/numa_node").read().strip())
        cpus = open(f"{base}/local_cpulist").read().strip()
        info["numa_node"], info["cpus"] = node, cpus
        info["affinity_before"] = sorted(os.sched_getaffinity(0))
        if cpus:
            ids = set()
            for part in cpus.split(","):
                a, _, b = part.partition("-")
                ids.update(range(int(a), int(b or a) + 1))
            ids &= os.sched_getaffinity(0)
            if ids:
                os.sched_setaffinity(0, ids)
                info["bound"] = True
    except Exception as exc:  # placement is an optimisation, never a reason to fail
        info["error"] = str(exc)[:80]
    return info


# ---------------------------------------------------------------- jobs
class Ctx:
    def __init__(self, args):
        import torch
        import torch.distributed as dist
        self.torch, self.dist, self.args = torch, dist, args
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available():
            raise SystemExit("bench.py needs a CUDA device: the product has no CPU path (use --impl reference for the CPU baseline)")
        # fewer GPUs than ranks (the 1-GPU test box running the 2-rank strip test): ranks share devices and the plumbing
        # falls back to gloo (NCCL refuses two ranks on one device); the data path -- CUDA IPC + flags -- is the same
        self.shared_gpu = torch.cuda.device_count() < min(self.world, int(os.environ.get("LOCAL_WORLD_SIZE", self.world)))
        self.device_index = self.local_rank % torch.cuda.device_count()
        torch.cuda.set_device(self.device_index)
        self.placement = bind_to_gpu_numa_node(self.device_index)
        if self.world > 1:
            if self.shared_gpu:
                dist.init_process_group("gloo")
            else:
                dist.init_process_group("nccl", device_id=torch.device("cuda", self.device_index))
        self.sm_count = torch.cuda.get_device_properties(self.device_index).multi_processor_count

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def max_over_ranks(self, x):
        if self.world == 1:
            return float(x)
        t = self.torch.tensor([float(x)], device="cpu" if self.shared_gpu else "cuda", dtype=self.torch.float64)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def all_gather_object(self, obj):
        if self.world == 1:
            return [obj]
        out = [None] * self.world
        self.dist.all_gather_object(out, obj)
        return out


def timed_steps(ctx, step, steps, warmup, sampler=None):
    """W untimed + K timed steps, bracketed by barrier + synchronize, CUDA events on the launching stream, max over ranks.
    Returns (ms per step, launches, (clock mark before, after))."""
    import various_image_processings_b200 as vip
    torch = ctx.torch
    for i in range(warmup):
        step(i)
    ctx.barrier()
    m0 = sampler.mark() if sampler else 0
    n0 = vip.launch_count()
    evs = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]   # one event per step boundary
    evs[0].record()
    for i in range(steps):
        step(warmup + i)
        evs[i + 1].record()
    ctx.barrier()
    m1 = sampler.mark() if sampler else 0
    launches = vip.launch_count() - n0
    per = [evs[i].elapsed_time(evs[i + 1]) for i in range(steps)]
    stats = {"ms_median": ctx.max_over_ranks(float(np.median(per))), "ms_best": ctx.max_over_ranks(float(np.min(per)))}
    return ctx.max_over_ranks(evs[0].elapsed_time(evs[steps])) / steps, launches, (m0, m1), stats


class LinkTimeout(RuntimeError):
    """a device-side neighbour wait gave up (vip_link_status), seen by every rank at the same point"""


class Job:
    """One workload on this rank: buffers, the device-resident step, the end-to-end step, parity of the split."""

    def __init__(self, ctx, name, halo_mode):
        import various_image_processings_b200 as vip
        self.vip, self.ctx, self.name = vip, ctx, name
        self.filt, self.w, self.h, self.ksize, self.nitr, self.desc = WORKLOADS[name]
        torch = ctx.torch
        world, rank = ctx.world, ctx.rank
        content = ctx.args.content
        w, h, ksize, nitr, filt = self.w, self.h, self.ksize, self.nitr, self.filt
        self.rb = w * 3
        # ---- decomposition: strips of one frame (strong) when the filter and the frame allow it, else a frame per rank
        self.halo = 0 if filt == "all" else vip.halo_rows(filt, ksize, nitr, exchange_between_iterations=False)
        self.strong = world > 1 and filt != "all" and h // world >= max(self.halo, 32)
        if self.strong:
            plans = vip.plan_strips(h, world, self.halo)
            self.plans = plans
            self.y0, self.y1, self.s0, self.s1 = plans[rank]
        else:
            self.plans = None
            self.y0, self.y1, self.s0, self.s1 = 0, h, 0, h
        self.own = self.y1 - self.y0
        self.halo_mode = "none" if not self.strong else ("copy" if filt == "btf" else halo_mode)
        self.has_above = self.strong and rank > 0
        self.has_below = self.strong and rank < world - 1
        # ---- frames: rotate over enough distinct resident frames that in + out exceed the L2
        n_pool = max(2, int(np.ceil(1.2 * L2_BYTES / (2.0 * self.rb * self.own))))
        self.n_pool = n_pool = min(n_pool, 12)
        seed0 = 42 if self.strong else 42 + 1000 * rank      # strips: every rank cuts the SAME frames
        self.frame0 = None
        self.host_own = []
        for i in range(n_pool):
            f = lenna_frame() if name == "bf_lenna" else synth_frame(w, h, seed0 + i, content)
            if i == 0:
                self.frame0 = f
            self.host_own.append(np.ascontiguousarray(f[self.y0:self.y1]))
        # ---- filters
        if filt == "bf":
            self.F = vip.CudaBilateralFilter(w, h, ksize)
        elif filt == "abf":
            self.F = vip.CudaAdaptiveBilateralFilter(w, h, ksize)
        elif filt == "btf":
            self.F = vip.CudaBilateralTextureFilter(w, h, ksize, nitr)
        else:
            self.F = (vip.CudaBilateralFilter(w, h, ksize), vip.CudaAdaptiveBilateralFilter(w, h, ksize),
                      vip.CudaBilateralTextureFilter(w, h, ksize, nitr))
        self.n_out = 3 if filt == "all" else 1
        # ---- device buffers: source strips hold [s0, s1) (own rows uploaded, halo rows only ever written by the halo
        #      pull), destinations the own rows.  DeviceImage = one cudaMalloc each, so the base pointer can be exported.
        srows = self.s1 - self.s0
        self.src = [vip.DeviceImage(w, srows, 3) for _ in range(n_pool)]
        self.dst = [[vip.DeviceImage(w, self.own, 3) for _ in range(self.n_out)] for _ in range(n_pool)]
        for j in range(n_pool):
            self.src[j].upload_rows(self.host_own[j], self.y0 - self.s0, self.own, stream=0)
        if filt == "btf" and self.strong:
            self.pp = [vip.DeviceImage(w, srows, 3), vip.DeviceImage(w, srows, 3)]
        # ---- end-to-end slots
        self.slots = []
        for s in range(NS):
            sl = {"src": vip.DeviceImage(w, srows, 3), "dst": [vip.DeviceImage(w, self.own, 3) for _ in range(self.n_out)],
                  "hin": vip.PinnedBuffer(self.rb * self.own), "hout": [vip.PinnedBuffer(self.rb * self.own) for _ in range(self.n_out)],
                  "stream": torch.cuda.Stream()}
            sl["hin"].array[:] = self.host_own[s % n_pool].reshape(-1)
            if filt in ("btf", "all"):   # BTF scratch is per object: one object per frame in flight
                sl["btf"] = vip.CudaBilateralTextureFilter(w, h, ksize, nitr)
                if self.strong:
                    sl["pp"] = [vip.DeviceImage(w, srows, 3), vip.DeviceImage(w, srows, 3)]
            self.slots.append(sl)
        # ---- neighbours: IPC handles of every source buffer + one flag block per rank
        self.link = None
        self.transport = "none"
        self.e2e_i = 0                      # running step counter of the end-to-end legs (same on every rank)
        self.prev_step = [0] * NS           # last step number that used each slot
        self.nbr = {}
        if self.strong:
            self.link = vip.StripLink(NS)
            mine = {"src": [vip.ipc_export(b) for b in self.src], "slot": [vip.ipc_export(sl["src"]) for sl in self.slots],
                    "link": self.link.export(), "plan": tuple(plans[rank]), "device": ctx.device_index}
            every = ctx.all_gather_object(mine)
            ok = os.environ.get("VIP_BENCH_TRANSPORT", "ipc") != "nccl"
            if ok:
                try:
                    for side, r in (("above", rank - 1), ("below", rank + 1)):
                        if 0 <= r < world:
                            o = every[r]
                            self.nbr[side] = {"src": [vip.ipc_open(hd) for hd in o["src"]],
                                              "slot": [vip.ipc_open(hd) for hd in o["slot"]], "plan": o["plan"]}
                    self.link.connect(every[rank - 1]["link"] if rank > 0 else None,
                                      every[rank + 1]["link"] if rank < world - 1 else None)
                except vip.VipError as exc:      # no peer path / IPC not permitted here: said in the line, never silent
                    ok = False
                    self.ipc_error = str(exc)
            # every rank takes the same transport
            if not all(ctx.all_gather_object(ok)):
                self._close_ipc()
                self.link = None
                self.halo_mode = "nccl"
                for side, r in (("above", rank - 1), ("below", rank + 1)):
                    if 0 <= r < world:
                        self.nbr[side] = {"rank": r, "plan": every[r]["plan"]}
                self.transport = ("NCCL send/recv of the halo rows (torch.distributed), because CUDA IPC between the ranks is not available here: "
                                  + getattr(self, "ipc_error", os.environ.get("VIP_BENCH_TRANSPORT", "")))
            else:
                self.transport = "cuda-ipc peer pointers over NVLink (%s)" % (
                    "halo rows staged by the filter kernel straight from the neighbours' buffers" if self.halo_mode == "fused"
                    else "cudaMemcpyAsync peer copies of the halo rows")
        ctx.barrier()

    def _close_ipc(self):
        for side in self.nbr.values():
            for p in side.get("src", []) + side.get("slot", []):
                try:
                    self.vip.ipc_close(p)
                except self.vip.VipError:
                    pass
        self.nbr = {}

    def _view(self, img, row0, rows):
        """torch uint8 view of rows [row0, row0 + rows) of a DeviceImage (zero copy), for the NCCL fallback"""
        class V:
            pass
        v = V()
        v.__cuda_array_interface__ = {"shape": (rows * self.rb,), "typestr": "|u1", "data": (img.row_ptr(row0), False), "version": 2}
        return self.ctx.torch.as_tensor(v, device="cuda")

    def _exchange_halo_nccl(self, img, stream):
        """fallback transport: the rows the neighbours need are sent, the halo rows received, through NCCL"""
        torch, dist = self.ctx.torch, self.ctx.dist
        ops = []
        with torch.cuda.stream(stream) if stream is not None else torch.cuda.stream(torch.cuda.current_stream()):
            if "above" in self.nbr:
                r = self.nbr["above"]["rank"]
                ops.append(dist.P2POp(dist.isend, self._view(img, self.y0 - self.s0, self.halo), r))
                ops.append(dist.P2POp(dist.irecv, self._view(img, 0, self.y0 - self.s0), r))
            if "below" in self.nbr:
                r = self.nbr["below"]["rank"]
                ops.append(dist.P2POp(dist.isend, self._view(img, self.y1 - self.s0 - self.halo, self.halo), r))
                ops.append(dist.P2POp(dist.irecv, self._view(img, self.y1 - self.s0, self.s1 - self.y1), r))
            for w in dist.batch_isend_irecv(ops):
                w.wait()

    # strip descriptors ---------------------------------------------------------------------------------------------
    def _peer_halo(self, which, j):
        """struct vip_halo over the neighbours' OWN rows of pool frame / slot j."""
        vip = self.vip

        def side(name):
            if name not in self.nbr:
                return None
            ny0, ny1, ns0, _ = self.nbr[name]["plan"]
            return (self.nbr[name][which][j] + (ny0 - ns0) * self.rb, ny0, ny1 - ny0)
        return vip.make_halo(above=side("above"), below=side("below"))

    def _pull_halo(self, which, j, dst_img, stream):
        """explicit halo pull: rows [s0, y0) from the rank above and [y1, s1) from the rank below, peer-to-peer"""
        vip = self.vip
        if "above" in self.nbr and self.y0 > self.s0:
            ny0, ny1, ns0, _ = self.nbr["above"]["plan"]
            vip.copy_rows(dst_img, self.s0, self.nbr["above"][which][j], ns0, self.s0, self.y0 - self.s0, self.w, stream=stream)
        if "below" in self.nbr and self.s1 > self.y1:
            ny0, ny1, ns0, _ = self.nbr["below"]["plan"]
            vip.copy_rows(dst_img, self.s0, self.nbr["below"][which][j], ns0, self.y1, self.s1 - self.y1, self.w, stream=stream)

    def _filter(self, which, j, src_img, dst_imgs, stream, btf=None, pp=None):
        """the filter pass of one step on this rank's rows; returns the (image, first row) holding the result"""
        vip, filt = self.vip, self.filt
        if not self.strong:
            if filt == "bf":
                self.F.bilateral_filter_async(src_img, dst_imgs[0], stream=stream)
            elif filt == "abf":
                self.F.execute_async(src_img, dst_imgs[0], stream=stream)
            elif filt == "btf":
                (btf or self.F).execute_async(src_img, dst_imgs[0], stream=stream)
            else:
                self.F[0].bilateral_filter_async(src_img, dst_imgs[0], stream=stream)
                self.F[1].execute_async(src_img, dst_imgs[1], stream=stream)
                (btf or self.F[2]).execute_async(src_img, dst_imgs[2], stream=stream)
            return
        if self.halo_mode == "fused":
            own_ptr = src_img.row_ptr(self.y0 - self.s0)
            st = vip.Strip(self.y0, self.own, self.y0, self.y0, self.own)
            halo = self._peer_halo(which, j)
            if filt == "bf":
                self.F.bilateral_filter_strip_peer(own_ptr, halo, dst_imgs[0], st, stream=stream)
            else:
                self.F.execute_strip_peer(own_ptr, halo, dst_imgs[0], st, stream=stream)
            return
        if self.halo_mode == "nccl":
            self._exchange_halo_nccl(src_img, stream)
        else:
            self._pull_halo(which, j, src_img, stream)
        if filt in ("bf", "abf"):
            st = vip.Strip(self.s0, self.s1 - self.s0, self.y0, self.y0, self.own)
            if filt == "bf":
                self.F.bilateral_filter_strip(src_img, dst_imgs[0], st, stream=stream)
            else:
                self.F.execute_strip(src_img, dst_imgs[0], st, stream=stream)
            return
        # BTF strip, deep halo: iteration `it` also produces the rows the remaining iterations will read around the own rows
        B = btf or self.F
        per_it = self.halo // max(1, self.nitr)
        cur = src_img
        for it in range(self.nitr):
            margin = (self.nitr - 1 - it) * per_it
            oy0, oy1 = max(0, self.y0 - margin), min(self.h, self.y1 + margin)
            last = it == self.nitr - 1
            out = dst_imgs[0] if last else pp[it & 1]
            st = vip.Strip(self.s0, self.s1 - self.s0, self.y0 if last else self.s0, oy0, oy1 - oy0)
            B.iterate_strip(cur, out, st, stream=stream)
            cur = out

    # the two legs ---------------------------------------------------------------------------------------------------
    def step(self, i):
        j = i % self.n_pool
        self._filter("src", j, self.src[j], self.dst[j], None, pp=getattr(self, "pp", None))

    def ordered_step(self, i):
        """device-resident step of the strip split WITH the neighbour protocol: this rank's rows of frame i arrive fresh
        (a device-to-device copy from the resident pool into the slot buffer the neighbours read), the neighbours are told,
        their rows are awaited, the filter runs, the neighbours are released -- all on the launching stream.  Step
        numbers continue the end-to-end counter (the flags are monotonic per slot)."""
        n = self.e2e_i + i
        s = n % NS
        sl = self.slots[s]
        step_no = n + 1
        if self.link is not None and self.prev_step[s]:
            self.link.wait_consumed(s, self.prev_step[s], stream=None)
        self.prev_step[s] = step_no
        self.vip.copy_rows(sl["src"], self.s0, self.src[i % self.n_pool], self.s0, self.y0, self.own, self.w, stream=None)
        if self.link is not None:
            self.link.signal_ready(s, step_no, stream=None)
            self.link.wait_ready(s, step_no, stream=None)
        self._filter("slot", s, sl["src"], sl["dst"], None, btf=sl.get("btf"), pp=sl.get("pp"))
        if self.link is not None:
            self.link.signal_consumed(s, step_no, stream=None)

    def run_ordered(self, steps, warmup, sampler=None):
        """timed_steps over ordered_step; advances the shared step counter"""
        res = timed_steps(self.ctx, self.ordered_step, steps, warmup, sampler)
        self.ctx.torch.cuda.synchronize()
        self.e2e_i += steps + warmup
        self.check_link("device-resident ordered leg")
        return res

    def e2e_step(self, i):
        """step i (from 0): upload own rows -> (neighbours ready) -> filter -> (tell neighbours) -> download"""
        s = i % NS
        sl = self.slots[s]
        st = sl["stream"]
        step_no = i + 1
        if self.link is not None and self.prev_step[s]:
            self.link.wait_consumed(s, self.prev_step[s], stream=st)   # neighbours are done with the rows this upload overwrites
        self.prev_step[s] = step_no
        sl["src"].upload_rows(sl["hin"].array.reshape(self.own, self.w, 3), self.y0 - self.s0, self.own, stream=st)
        if self.link is not None:
            self.link.signal_ready(s, step_no, stream=st)
            self.link.wait_ready(s, step_no, stream=st)
        self._filter("slot", s, sl["src"], sl["dst"], st, btf=sl.get("btf"), pp=sl.get("pp"))
        if self.link is not None:
            self.link.signal_consumed(s, step_no, stream=st)
        for d, hb in zip(sl["dst"], sl["hout"]):
            d.download_rows(hb.array.reshape(self.own, self.w, 3), 0, self.own, stream=st)

    def run_e2e(self, steps):
        """returns seconds per step (max over ranks) of the pipelined end-to-end leg"""
        ctx = self.ctx
        base = self.e2e_i
        for i in range(NS):
            self.e2e_step(base + i)
        for sl in self.slots:
            sl["stream"].synchronize()
        ctx.barrier()
        t0 = time.perf_counter()
        for i in range(NS, NS + steps):
            self.e2e_step(base + i)
        for sl in self.slots:
            sl["stream"].synchronize()
        dt = time.perf_counter() - t0
        self.e2e_i = base + NS + steps
        self.check_link("end-to-end leg")
        return ctx.max_over_ranks(dt) / steps

    def check_link(self, where):
        """A neighbour wait that timed out (vip_link_status) is an error of the RUN on every rank: said loudly, and the
        same on all ranks so that no rank is left waiting in a collective."""
        if self.link is None:
            return
        ok = True
        try:
            self.link.status()
        except self.vip.VipError:
            ok = False
        if not all(self.ctx.all_gather_object(ok)):
            raise LinkTimeout(where)

    def fallback_to_nccl(self, why):
        """switch this job's halo transport to NCCL send / recv (all ranks together), keeping the plans"""
        plans = {side: d["plan"] for side, d in self.nbr.items()}
        self._close_ipc()
        self.link = None
        self.halo_mode = "nccl"
        for side, r in (("above", self.ctx.rank - 1), ("below", self.ctx.rank + 1)):
            if side in plans:
                self.nbr[side] = {"rank": r, "plan": plans[side]}
        self.transport = "NCCL send/recv of the halo rows (torch.distributed): " + why
        self.prev_step = [0] * NS
        self.ctx.barrier()

    def run_copy_ceiling(self, steps):
        """the same host<->device traffic with no kernels: what the box's host side allows at this N"""
        ctx = self.ctx

        def one(i):
            sl = self.slots[i % NS]
            st = sl["stream"]
            sl["src"].upload_rows(sl["hin"].array.reshape(self.own, self.w, 3), self.y0 - self.s0, self.own, stream=st)
            for d, hb in zip(sl["dst"], sl["hout"]):
                d.download_rows(hb.array.reshape(self.own, self.w, 3), 0, self.own, stream=st)
        for i in range(NS):
            one(i)
        for sl in self.slots:
            sl["stream"].synchronize()
        ctx.barrier()
        t0 = time.perf_counter()
        for i in range(steps):
            one(i)
        for sl in self.slots:
            sl["stream"].synchronize()
        return ctx.max_over_ranks(time.perf_counter() - t0) / steps

    def close(self):
        """unmap the neighbours' buffers before anybody frees them"""
        self.ctx.barrier()
        self._close_ipc()
        self.link = None
        self.ctx.barrier()

    # parity of the split --------------------------------------------------------------------------------------------
    def check_parity(self):
        """Outside the timed region: frame 0 through the strip path (device-resident AND end-to-end), gathered on rank 0
        and compared with (a) the whole frame filtered on one GPU, (b) the reference's own CUDA kernel on that GPU,
        (c) oracle rows at the strip boundaries.  Returns a dict for the JSON line (rank 0) or None."""
        import _oracle as O
        ctx, vip, torch = self.ctx, self.vip, self.ctx.torch
        w, h, filt = self.w, self.h, self.filt
        if filt == "all":
            return None
        self.step(0)                          # pool frame 0 -> dst[0]
        torch.cuda.synchronize()
        mine = self.dst[0][0].download()
        self.ctx.barrier()
        # the e2e path's result for slot 0 (holds pool frame 0 as input)
        self.e2e_i = (self.e2e_i + NS - 1) // NS * NS      # next step that lands on slot 0
        self.e2e_step(self.e2e_i)
        self.e2e_i += 1
        for sl in self.slots:
            sl["stream"].synchronize()
        self.check_link("parity check")
        mine_e2e = self.slots[0]["hout"][0].array.reshape(self.own, w, 3).copy()
        if ctx.world > 1:
            parts = [None] * ctx.world if ctx.rank == 0 else None
            ctx.dist.gather_object((mine, mine_e2e), parts, dst=0)
        else:
            parts = [(mine, mine_e2e)]
        if ctx.rank != 0:
            return None
        got = np.concatenate([p[0] for p in parts], 0)
        got_e2e = np.concatenate([p[1] for p in parts], 0)
        res = {"n_strips": ctx.world if self.strong else 1, "frame": "pool frame 0"}
        d_src = torch.from_numpy(self.frame0).cuda()
        d_one = torch.empty_like(d_src)
        if self.strong:
            if filt == "bf":
                self.F.bilateral_filter_async(d_src, d_one)
            elif filt == "abf":
                self.F.execute_async(d_src, d_one)
            else:
                self.F.execute_async(d_src, d_one)
            torch.cuda.synchronize()
            one = d_one.cpu().numpy()
            res["vs_single_gpu"] = "exact" if np.array_equal(got, one) else "DIFF(%d px)" % int((got != one).any(-1).sum())
        res["e2e_vs_resident"] = "exact" if np.array_equal(got, got_e2e) else "DIFF(%d px)" % int((got != got_e2e).any(-1).sum())
        # reference CUDA kernel on the same GPU: the parity oracle O2 and the kernel to beat
        ref_cuda = None
        if O.have_ref_cuda():
            R = O.ref_cuda()
            d_ref = torch.empty_like(d_src)
            if filt == "bf":
                p = R.refcu_bf_create(w, h, self.ksize, 10.0, 30.0)
                run = lambda: R.refcu_bf_run(p, d_src.data_ptr(), d_ref.data_ptr())
            elif filt == "abf":
                p = R.refcu_abf_create(w, h, self.ksize, 10.0, 30.0)
                run = lambda: R.refcu_abf_run(p, d_src.data_ptr(), d_ref.data_ptr())
            else:
                p = R.refcu_btf_create(w, h, self.ksize, self.nitr)
                run = lambda: R.refcu_btf_run(p, d_src.data_ptr(), d_ref.data_ptr())
            run()
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(3):
                run()               # the reference's public calls end with cudaDeviceSynchronize (src/*_impl.cu)
            e1.record()
            torch.cuda.synchronize()
            ref_ms = e0.elapsed_time(e1) / 3
            ref = d_ref.cpu().numpy()
            same = np.array_equal(got, ref)
            res["vs_ref_cuda"] = "exact" if same else "DIFF(%d px, max %d)" % (int((got != ref).any(-1).sum()), int(np.abs(got.astype(int) - ref.astype(int)).max()))
            ref_cuda = {"ms": ref_ms, "exact": bool(same), "what": "the reference's own kernel(s) (src/*_impl.cu compiled unmodified for sm_100a, oracle/_ref/libvip_ref_cuda.so), whole frame on this GPU, outside the timed region"}
        # oracle rows at the strip boundaries (bounded CPU work)
        if filt in ("bf", "abf"):
            ys = sorted({0, h - 1} | ({pl.y0 for pl in self.plans[1:]} | {pl.y0 - 1 for pl in self.plans[1:]} if self.strong else {h // 2}))
            ys = ys[:6]
            bad = 0
            for y in ys:
                want = O.o_rows("abf" if filt == "abf" else "bf", self.frame0, y, 1, self.ksize)
                bad += int((want[0] != got[y]).any(-1).sum())
            res["vs_oracle_rows"] = {"rows": ys, "mismatching_pixels": bad}
        ok = all(v == "exact" for k, v in res.items() if k.startswith("vs_") and isinstance(v, str)) and \
            res["e2e_vs_resident"] == "exact" and res.get("vs_oracle_rows", {}).get("mismatching_pixels", 0) == 0
        res["verdict"] = "exact" if ok else "MISMATCH"
        return res, ref_cuda


def _events_ms(torch, fn, n, warm=3):
    for i in range(warm):
        fn(i)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        fn(warm + i)
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def time_btf_stages(vip, torch, job, srcs, jbf_ms):
    """The kernels of the hot path that are not the dominant one, each timed alone with CUDA events on frames of the BTF
    workload: one whole texture-filter iteration (3 launches), the guide kernel, the gradient and the integral image.
    The fused gradient + blur/mRTV kernel has no entry point of its own: its time is iteration - guide - joint bilateral
    (derived, and said so in the line)."""
    w, h, k = job.w, job.h, job.ksize
    n = 30
    B = vip.CudaBilateralTextureFilter(w, h, k, 1)
    outs = [torch.empty_like(srcs[0]) for _ in range(2)]
    st = vip.Strip(0, h, 0, 0, h)
    it_ms = _events_ms(torch, lambda i: B.iterate_strip(srcs[i % len(srcs)], outs[i & 1], st), n)
    blurred = torch.rand((h, w, 3), dtype=torch.float32, device="cuda") * 255.0
    rtv = torch.rand((h, w), dtype=torch.float32, device="cuda")
    guide = torch.empty((h, w, 3), dtype=torch.uint8, device="cuda")
    g_ms = _events_ms(torch, lambda i: B.compute_guide(blurred, rtv, guide), n)
    mags = [torch.empty((h, w), dtype=torch.float32, device="cuda") for _ in range(4)]
    grad_ms = _events_ms(torch, lambda i: vip.cuda_gradient(srcs[i % len(srcs)], mags[i % 4], w, h, 3), n)
    r = 7
    ii = vip.CudaIntegralImage(w, h, r)
    int_ms = _events_ms(torch, lambda i: ii.build(srcs[i % len(srcs)]), n)
    return {"iteration_ms": it_ms, "guide_ms": g_ms, "jbf_ms": jbf_ms, "gradient_ms": grad_ms, "integral_ms": int_ms, "integral_radius": r}


def other_rooflines(t, w, h, k, issue_peak, issue_src, hbm_peak):
    """roofline objects of the non-dominant kernels (round-2 review, item 5).  HBM side: algorithmic bytes / time against
    the measured copy bandwidth.  Issue side (the BTF stages are instruction-bound): pixels/s x I_min
    against the same measured FFMA issue rate as the stencils, with I_min(K) = the instructions the ARITHMETIC DEFINITION
    needs per output pixel when nothing is wasted:
      blur/mRTV  K^2/2 (the raster-order magnitude sum cannot be reordered: one FADD per tap, two outputs per packed FADD2)
                 + K^2/8 (their shared-memory operand, one load per 8 outputs) + 2 x 6 x (K-1) (three box sums, min, max of
                 b+g+r and max magnitude, separable: K-1 ops per quantity and pass) + 2K (record loads of the two passes)
                 + 17 (gradient magnitude of the tile incl. halo) + 30 (three exact divides, the fp64 mRTV divide, stores)
      guide      2 x 3 x (K-1) (lexicographic (value, index) min: compare + two selects per step, two passes) + 2K (loads)
                 + 16 (alpha: expf, reciprocal) + 6 (gather of Bq, Bp) + 18 (blend, round, clamp of 3 channels) + 2 (store)"""
    px = float(w * h)
    K = float(k)
    imin_blur = K * K / 2 + K * K / 8 + 12 * (K - 1) + 2 * K + 17 + 30
    imin_guide = 6 * (K - 1) + 2 * K + 16 + 6 + 18 + 2
    blur_ms = t["iteration_ms"] - t["guide_ms"] - t["jbf_ms"]
    r = t["integral_radius"]
    cells = float((w + 2 * r + 1) * (h + 2 * r + 1))

    def hbm(bytes_, ms):
        a = bytes_ / (ms * 1e-3) / 1e9
        return {"achieved": a, "peak": hbm_peak, "unit": "GB/s", "frac": a / hbm_peak, "algorithmic_bytes_per_launch": bytes_}

    def issue(imin, ms):
        a = px * imin / (ms * 1e-3)
        return {"achieved": a, "peak": issue_peak, "unit": "thread-instr/s", "frac": a / issue_peak if issue_peak else None,
                "i_min_per_pixel": imin, "peak_source": issue_src}
    out = {
        "btf_iteration_ms": t["iteration_ms"],
        "blur_rtv_kernel": {"bound": "issue", "kernel_ms": blur_ms, "timing": "derived: one whole iteration - guide kernel - joint bilateral kernel (each timed alone with CUDA events)",
                            **issue(imin_blur, blur_ms), "hbm": hbm(px * (3 + 12 + 4), blur_ms)},
        "guide_kernel": {"bound": "issue", "kernel_ms": t["guide_ms"], "timing": "alone, CUDA events, 30 launches",
                         **issue(imin_guide, t["guide_ms"]), "hbm": hbm(px * (12 + 4 + 3), t["guide_ms"])},
        "gradient_u8_kernel": {"bound": "hbm", "kernel_ms": t["gradient_ms"], "timing": "alone, CUDA events, 30 launches, rotating frames",
                               **hbm(px * 7, t["gradient_ms"])},
        "integral_kernels": {"bound": "hbm", "kernel_ms": t["integral_ms"], "radius": r, "launches": 4,
                             "timing": "alone (4 launches + pool allocation), CUDA events, 30 builds, rotating frames",
                             **hbm(cells * 15, t["integral_ms"])},
    }
    return out


def measure_job(ctx, name, steps, warmup, halo_mode, sampler=None, with_cpu=True, cpu_budget=1e9):
    """All figures of one workload; rank 0 gets the dict, the others None."""
    import various_image_processings_b200 as vip
    torch = ctx.torch
    job = Job(ctx, name, halo_mode)
    filt, w, h, ksize, nitr = job.filt, job.w, job.h, job.ksize, job.nitr
    ms_step, launches, marks, step_stats = timed_steps(ctx, job.step, steps, warmup, sampler)
    pixels_step = w * h if job.strong else w * h * ctx.world
    # Strip split (N > 1): job.step reads halos from neighbours' pool frames that never change -- N independent strip
    # kernels, no rank waits for another.  The HEADLINE of a strong-scaled frame is the ordered leg: own rows rewritten
    # every step, ready / consumed flags between neighbours (the unordered figure overstates a frame pipeline).
    ordered = None
    if job.strong:
        try:
            o_ms, o_launches, _, o_stats = job.run_ordered(steps, warmup)
        except LinkTimeout as exc:
            job.fallback_to_nccl(f"device-side neighbour flags timed out during the {exc}")
            o_ms, o_launches, _, o_stats = job.run_ordered(steps, warmup)
        ordered = {"ms_per_step": o_ms, "median": o_stats["ms_median"], "best": o_stats["ms_best"], "launches": int(o_launches)}
    unordered_ms = ms_step
    head_ms = ordered["ms_per_step"] if ordered else ms_step
    value = pixels_step / (head_ms * 1e-3) / 1e6
    e2e_steps = max(8, min(steps, 40))
    e2e_note = None
    try:
        e2e_s = job.run_e2e(e2e_steps)
    except LinkTimeout as exc:
        # never a silent fallback: the line says that the end-to-end leg ran over NCCL and why
        e2e_note = f"device-side neighbour flags timed out during the {exc}; the end-to-end leg was re-run with NCCL send/recv halos"
        job.fallback_to_nccl(e2e_note)
        e2e_s = job.run_e2e(e2e_steps)
    e2e_value = pixels_step / e2e_s / 1e6
    mark_end = sampler.mark() if sampler else 0
    ceil_s = job.run_copy_ceiling(e2e_steps)
    try:
        par = job.check_parity()
    except LinkTimeout as exc:
        job.fallback_to_nccl(f"device-side neighbour flags timed out during the {exc}")
        par = job.check_parity()
    job.close()
    if ctx.rank != 0:
        return None, job
    parity, ref_cuda = par if par else (None, None)

    # ---- the dominant kernel, timed alone where a step is several kernels (BTF: the joint bilateral, 2k-1 window,
    #      sigma_space k-1, sigma_color sqrt 3 -- src/bilateral_texture_filter_impl.cu:188), CUDA events, same frames
    local_rows = job.own
    stage_times = None
    if filt in ("btf", "all"):
        kind = "jbf"
        radius = ksize - 1
        J = vip.CudaBilateralFilter(w, h, 2 * ksize - 1, float(ksize - 1), 3.0 ** 0.5)
        srcs = [torch.from_numpy(job.host_own[j]).cuda() for j in range(job.n_pool)] if not job.strong else None
        if job.strong:
            # strip of the same height as this rank's rows, cut from a resident frame (kernel time of a strip)
            full = torch.from_numpy(job.frame0).cuda()
            g = torch.flip(full, dims=[0]).contiguous()
            d_tmp = torch.empty((job.own, w, 3), dtype=torch.uint8, device="cuda")
            stp = vip.Strip(0, h, job.y0, job.y0, job.own)
            jstep = lambda i: J.joint_bilateral_filter_strip(full, g, d_tmp, stp)
        else:
            guides = [torch.flip(s, dims=[0]).contiguous() for s in srcs]
            d_tmp = torch.empty_like(srcs[0])
            jstep = lambda i: J.joint_bilateral_filter_async(srcs[i % job.n_pool], guides[i % job.n_pool], d_tmp)
        for i in range(3):
            jstep(i)
        torch.cuda.synchronize()
        jn = max(10, min(steps, 50))
        j0, j1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        j0.record()
        for i in range(jn):
            jstep(i)
        j1.record()
        torch.cuda.synchronize()
        kernel_ms = j0.elapsed_time(j1) / jn
        launches_of_kernel_per_step = nitr
        alg_bytes = w * local_rows * 9.0
        if filt == "btf" and not job.strong and nitr > 0:
            stage_times = time_btf_stages(vip, torch, job, srcs, kernel_ms)
    else:
        kind = filt
        radius = ksize // 2
        kernel_ms = ms_step
        launches_of_kernel_per_step = 1
        alg_bytes = w * local_rows * 6.0
    taps = mask_taps(radius)
    P = OUTPUTS_PER_THREAD[kind]
    imin = i_min(kind, P)
    taps_s = w * local_rows * taps / (kernel_ms * 1e-3)        # this rank's kernel
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except (OSError, ValueError):
        pass
    hbm_peak = float(peaks.get("hbm_gbs", 6650.0))
    ncu = {}
    try:
        ncu = json.load(open(os.path.join(ROOT, "profiles", "ncu_constants.json"))).get("stencil_" + kind, {})
    except (OSError, ValueError):
        pass
    roof = {"bound": "issue", "unit": "thread-instr/s",
            "kernel": {"bf": "stencil_kernel<BF>", "abf": "stencil_kernel<ABF>", "jbf": "stencil_kernel<JBF>"}[kind],
            "kernel_ms": kernel_ms, "taps_per_pixel": taps, "taps_per_s": taps_s, "i_min_per_tap": imin, "outputs_per_thread": P,
            "achieved": taps_s * imin, "peak": None, "frac": None,
            "definition": "achieved = taps/s x I_min (SURVEY.md 8d: BF 7+4/P, JBF 7+5/P, ABF 15+4/P thread-instructions per tap); peak = FFMA issue rate measured live on this GPU (tools/probe/issue_probe.cu); MEASURED_PEAKS.json has no FP32 / issue figure",
            "traffic": ncu.get("dram_bytes_per_launch"),
            "hbm": {"achieved": alg_bytes / (kernel_ms * 1e-3) / 1e9, "peak": hbm_peak, "unit": "GB/s",
                    "peak_source": "measured (MEASURED_PEAKS.json)" if "hbm_gbs" in peaks else "fallback (B200_PROFILING.md)",
                    "algorithmic_bytes_per_launch": alg_bytes}}
    roof["hbm"]["frac"] = roof["hbm"]["achieved"] / hbm_peak
    roof["fp32"] = {"achieved_tflops": taps_s * FLOP_PER_TAP[kind] / 1e12, "flop_per_tap": FLOP_PER_TAP[kind]}
    issue = {"util_ncu": ncu or None}
    try:
        import peaks as _peaks
        _pl = _peaks.load()
        if _pl is not None and not ctx.args.no_probe:
            ffma, _ = _peaks.measure(_pl, 0)
            roof["peak"] = ffma
            roof["peak_source"] = "measured live: FFMA thread-instructions/s (tools/probe/issue_probe.cu)"
            roof["frac"] = roof["achieved"] / ffma
            roof["fp32"]["peak_tflops"] = 2 * ffma / 1e12
            roof["fp32"]["frac"] = roof["fp32"]["achieved_tflops"] / roof["fp32"]["peak_tflops"]
    except Exception as exc:                                 # the probe is evidence, never a reason to fail the bench
        issue["probe_error"] = str(exc)
    if roof["peak"] is None:                                 # nominal issue rate at the max clock, said so
        nominal = ctx.sm_count * 128 * float(peaks.get("sm_max_mhz", 1965.0)) * 1e6
        roof["peak"], roof["peak_source"] = nominal, "nominal: SMs x 128 thread-instr/clk x max SM clock (probe not run)"
        roof["frac"] = roof["achieved"] / nominal

    other = other_rooflines(stage_times, w, h, ksize, roof["peak"], roof.get("peak_source"), hbm_peak) if stage_times else None

    out = {
        "metric": f"{filt}_throughput", "value": value, "unit": "Mpixel/s", "ms_per_step": head_ms, "steps": steps, "warmup": warmup,
        # one frame's work is fixed and cut into N strips (strong) wherever the filter and the frame allow it -- at N = 1 the
        # cut is the whole frame; only the frame stream / tiny frames give every GPU its own frames (weak)
        "scaling": "strong" if (job.strong or (ctx.world == 1 and filt != "all")) else "weak",
        "config": {"workload": name, "description": job.desc, "width": w, "height": h, "ksize": ksize, "nitr": nitr,
                   "content": ctx.args.content,
                   "parallelism": (f"{ctx.world} row strips of one frame, one process per {'rank (ranks SHARE a GPU: functional test, not a measurement)' if ctx.shared_gpu else 'GPU'}; halo {job.halo} rows per side; {job.transport}"
                                   if job.strong else f"{ctx.world} independent frame stream(s)"),
                   "l2": f"rotating over {job.n_pool} distinct resident frames ({2 * job.n_pool * job.rb * job.own >> 20} MiB in+out per GPU > 126 MiB L2)",
                   "e2e": f"pinned host buffers; per step every rank uploads its own rows, filters, downloads its rows; {NS} frames in flight on {NS} streams; neighbours ordered by device-side flags (vip_link_*)"},
        "e2e": {"value": e2e_value, "unit": "Mpixel/s", "h2d_bytes_per_step": job.rb * (h if job.strong else h * ctx.world),
                "d2h_bytes_per_step": job.rb * (h if job.strong else h * ctx.world) * job.n_out, "steps": e2e_steps,
                "ms_per_step": e2e_s * 1e3,
                "copy_ceiling": {"value": pixels_step / ceil_s / 1e6, "unit": "Mpixel/s", "ms_per_step": ceil_s * 1e3,
                                 "gbs_each_way": job.rb * (h if job.strong else h * ctx.world) / ceil_s / 1e9,
                                 "what": "the same pinned uploads + downloads on the same streams with no kernels, all ranks at once"},
                "frac_of_copy_ceiling": ceil_s / e2e_s, "note": e2e_note},
        "gpu_launches": int(launches), "kernels_per_step": launches / max(steps, 1),
        "step_ms": {"mean": ms_step, "median": step_stats["ms_median"], "best": step_stats["ms_best"],
                    "what": "per-step CUDA events on the launching stream, max over ranks (SURVEY.md 8d asks for median and best)"},
        "dominant_kernel_launches_per_step": launches_of_kernel_per_step,
        "roofline": roof, "issue": issue, "_marks": (marks[0], marks[1], mark_end),
    }
    if ordered is not None:
        out["resident_legs"] = {
            "ordered": {"value": value, "ms_per_step": ordered["ms_per_step"], "ms_median": ordered["median"], "ms_best": ordered["best"],
                        "what": "HEADLINE (value): every step each rank's own rows are rewritten device-to-device in the buffer its neighbours read, then signal_ready / wait_ready -> strip filter -> signal_consumed (vip_link_* flags over NVLink, or NCCL halos after a fallback); ranks are coupled every frame"},
            "unordered": {"value": pixels_step / (unordered_ms * 1e-3) / 1e6, "ms_per_step": unordered_ms,
                          "what": "N independent strip kernels whose halo rows come from neighbours' STATIC pool frames: no neighbour ordering, no per-step coupling -- kernel cost of a strip only (this was `value` in rounds 1-2); roofline.kernel_ms is this leg and, in halo mode 'copy', includes the halo copy"}}
        out["gpu_launches"] = ordered["launches"]
        out["kernels_per_step"] = ordered["launches"] / max(steps, 1)
    if other:
        out["other_kernels"] = other
    if parity is not None:
        out["strip_parity"] = parity["verdict"]
        out["parity"] = parity
    if ref_cuda is not None:
        ref_cuda["speedup_whole_job"] = ref_cuda["ms"] / head_ms
        out["ref_cuda"] = ref_cuda
    if with_cpu:
        # the CPU baseline gets ALL host cores back (the process was pinned to the GPU's NUMA-local CPUs for its own
        # pinned buffers; on a two-socket host that would halve the reference's threads)
        before = ctx.placement.get("affinity_before")
        if before:
            try:
                os.sched_setaffinity(0, before)
            except OSError:
                pass
        headline = cpu_budget >= 1e9
        rows = sample_rows_for(filt, w, h, ksize, cpu_budget, min_rows=256 if headline else 0)
        mp, ckind, cores, what, t_all = cpu_reference(filt, w, h, ksize, nitr, rows, ctx.args.cpu_threads, repeats=2, content=ctx.args.content)
        out["cpu_baseline"] = {"value": mp, "unit": "Mpixel/s", "cores": cores, "kind": ckind, "sample": what}
        if headline and cores > 1:
            # the ONE-core row: same code, one thread, on a crop sized from the all-cores time so that it
            # takes about 8 s whatever the host is
            rows1 = int(rows * 8.0 / max(t_all * cores, 1e-3)) // 8 * 8
            rows1 = min(rows, max(2 * ksize, rows1))
            mp1, _, _, what1, _ = cpu_reference(filt, w, h, ksize, nitr, rows1, 1, repeats=1, content=ctx.args.content)
            out["cpu_baseline"]["one_core"] = {"value": mp1, "unit": "Mpixel/s", "cores": 1, "sample": what1,
                                               "all_cores_speedup": mp / mp1}
            out["cpu_baseline"]["not_measured"] = "a -march=native build of the reference: oracle/_ref is compiled where /root/reference exists (not on the GPU box), for a generic x86-64 target"
    return out, job


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--workload", default=DEFAULT_WORKLOAD, choices=sorted(WORKLOADS))
    ap.add_argument("--cpu-threads", type=int, default=0, help="threads of the CPU baseline (0 = all host cores)")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-probe", action="store_true", help="skip the issue-rate probe kernels (for ncu launch lists)")
    ap.add_argument("--no-per-workload", action="store_true", help="only the main workload (for ncu / quick runs)")
    ap.add_argument("--content", default="rand", choices=["rand", "narrow", "nat", "smooth"], help="synthetic content (SURVEY.md 8d)")
    ap.add_argument("--halo", default="fused", choices=["fused", "copy"],
                    help="N > 1: halo rows staged by the kernel from the neighbours' memory, or pulled by peer copies")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference_arm(args, WORKLOADS[args.workload])
        return

    ctx = Ctx(args)
    sampler = ClockSampler(ctx.device_index)
    sampler.start()
    time.sleep(0.2)
    main_out, job = measure_job(ctx, args.workload, args.steps, args.warmup, args.halo, sampler=sampler,
                                with_cpu=not args.no_cpu_baseline and ctx.world == 1)
    if main_out is not None:
        m0, m1, m2 = main_out.pop("_marks")
        # the timed region alone can be shorter than two sampling periods (20 steps x 0.9 ms at N = 8): then the window is
        # widened to the device-timed region + the end-to-end leg that follows it, and the line says so
        if m1 - m0 >= 3:
            clocks = sampler.stop(m0, m1 + 1)
            clocks["window"] = "device-timed region"
        else:
            clocks = sampler.stop(max(0, m0 - 1), m2 + 1)
            clocks["window"] = "device-timed region + end-to-end leg (the timed region alone spans < 3 samples of 50 ms)"
    else:
        clocks = sampler.stop()
    del job
    per = {}
    if not args.no_per_workload and args.workload == DEFAULT_WORKLOAD:
        for name in PER_WORKLOAD:
            try:
                o, j = measure_job(ctx, name, 20, 3, args.halo, with_cpu=not args.no_cpu_baseline and ctx.world == 1, cpu_budget=4e8)
                del j
                if o is not None:
                    for k in ("steps", "warmup", "gpu_launches", "kernels_per_step", "issue", "_marks"):
                        o.pop(k, None)
                    per[name] = o
            except Exception as exc:       # a side workload never takes the headline down; it is reported instead
                per[name] = {"error": f"{type(exc).__name__}: {exc}"[:300]}
                ctx.barrier()
    if ctx.rank == 0:
        out = {"metric": main_out.pop("metric"), "value": main_out.pop("value"), "unit": main_out.pop("unit"), "n_gpus": ctx.world,
               "steps": main_out.pop("steps"), "warmup": main_out.pop("warmup"), "ms_per_step": main_out.pop("ms_per_step"),
               "higher_is_better": True, "scaling": main_out.pop("scaling"), "vs_baseline": None, "dtype": "f32",
               "data": "lenna.png" if args.workload == "bf_lenna" else "synthetic"}
        out.update(main_out)
        out["clocks"] = clocks
        out["host_placement"] = {k: v for k, v in ctx.placement.items() if k != "affinity_before"}
        if per:
            out["per_workload"] = per
        print(json.dumps(out), flush=True)
    if ctx.world > 1:
        ctx.dist.destroy_process_group()


if __name__ == "__main__":
    main()
