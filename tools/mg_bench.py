#!/usr/bin/env python
"""Single-process multi-GPU driver (vip_mg_*, csrc/mg.cu) on the GPUs of one box:
  * configs[3]: bilateral filter radius 15 on a 7680x4320 image cut into row strips with radius-deep halos;
  * configs[2] as strips: BTF k=9 x3 on 3840x2160 with halo exchange between iterations (peer copies);
  * configs[4]: a batch of 1080p frames dealt round-robin to the devices, per filter (pinned host buffers, H2D + D2H inside).
Prints one JSON line per measurement.   python tools/mg_bench.py [--gpus N] [--frames F]
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import various_image_processings_b200 as vip  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=torch.cuda.device_count())
    ap.add_argument("--frames", type=int, default=256)
    ap.add_argument("--reps", type=int, default=5)
    args = ap.parse_args()
    devs = list(range(args.gpus))
    rs = np.random.RandomState(42)

    ap_only = os.environ.get("MG_ONLY", "")
    for name, kind, w, h, k, nitr, mode in (("bf_8k_r15_strips", "bf", 7680, 4320, 31, 0, None), ("abf_4k_r7_strips", "abf", 3840, 2160, 15, 0, None),
                                            ("btf_4k_strips_deep_halo", "btf", 3840, 2160, 9, 3, "deep"),
                                            ("btf_4k_strips_exchange", "btf", 3840, 2160, 9, 3, "exchange"),
                                            ("btf_8k_strips_deep_halo", "btf", 7680, 4320, 9, 3, "deep"),
                                            ("btf_8k_strips_exchange", "btf", 7680, 4320, 9, 3, "exchange")):
        if ap_only and not name.startswith(ap_only):
            continue
        img = rs.randint(0, 256, size=(h, w, 3)).astype(np.uint8)
        hin, hout = vip.PinnedBuffer(img.size), vip.PinnedBuffer(img.size)
        hin.array[:] = img.reshape(-1)
        mg = vip.MultiGpuFilter(devs, kind, w, h, ksize=k, nitr=nitr, btf_halo=mode)
        src, dst = hin.array.reshape(h, w, 3), hout.array.reshape(h, w, 3)
        mg.scatter_host(src)
        mg.run()
        run_ms, e2e_ms = [], []
        for _ in range(args.reps):
            mg.run()
            run_ms.append(mg.last_run_ms())
        for _ in range(args.reps):
            t0 = time.perf_counter()
            mg.scatter_host(src)
            mg.run()
            mg.gather_host(dst)
            e2e_ms.append((time.perf_counter() - t0) * 1e3)
        # spot check: first / last rows of every strip boundary against a single-GPU run of the same rows is done by the tests
        print(json.dumps({"workload": name, "n_gpus": args.gpus, "device_ms": float(np.median(run_ms)),
                          "mpixel_per_s": w * h / np.median(run_ms) / 1e3, "e2e_ms_scatter_run_gather": float(np.median(e2e_ms)),
                          "e2e_mpixel_per_s": w * h / np.median(e2e_ms) / 1e3}), flush=True)
        del mg

    if ap_only:
        return
    w, h, nf = 1920, 1080, args.frames
    frames = vip.PinnedBuffer(nf * w * h * 3)
    out = vip.PinnedBuffer(nf * w * h * 3)
    frames.array[:] = rs.randint(0, 256, size=frames.array.size, dtype=np.uint8)
    total = 0.0
    for kind in ("bf", "abf", "btf"):
        mg = vip.MultiGpuFilter(devs, kind, w, h, ksize=9, nitr=3)
        mg.execute_batch(frames.array, out.array, min(nf, 4 * args.gpus))   # warm-up
        t0 = time.perf_counter()
        mg.execute_batch(frames.array, out.array, nf)
        dt = time.perf_counter() - t0
        total += dt
        print(json.dumps({"workload": f"stream_1080p_{kind}", "n_gpus": args.gpus, "frames": nf, "seconds": dt,
                          "frames_per_s": nf / dt, "mpixel_per_s": nf * w * h / dt / 1e6,
                          "note": "pinned host buffers, H2D + kernels + D2H, two streams per device"}), flush=True)
        del mg
    print(json.dumps({"workload": "stream_1080p_all_three", "n_gpus": args.gpus, "frames": nf, "seconds": total,
                      "frames_per_s": nf / total, "mpixel_per_s": nf * w * h / total / 1e6,
                      "note": "three passes over the host frames, one filter kind per pass (round-1 path)"}), flush=True)
    # configs[4] as one pass: every frame uploaded ONCE and sent through all three filters (vip_mg_run_batch_set_host)
    out2, out3 = vip.PinnedBuffer(nf * w * h * 3), vip.PinnedBuffer(nf * w * h * 3)   # kept alive: .array is only a view
    outs = {"bf": out.array, "abf": out2.array, "btf": out3.array}
    mg = vip.MultiGpuFilter(devs, "bf", w, h, ksize=9, nitr=3)
    mg.execute_batch_set(frames.array, outs, min(nf, 4 * args.gpus))   # warm-up
    t0 = time.perf_counter()
    mg.execute_batch_set(frames.array, outs, nf)
    dt = time.perf_counter() - t0
    print(json.dumps({"workload": "stream_1080p_filter_set", "n_gpus": args.gpus, "frames": nf, "seconds": dt,
                      "frames_per_s": nf / dt, "mpixel_per_s": nf * w * h / dt / 1e6, "peer_access": mg.peer_access(),
                      "note": "one upload per frame, BF + ABF + BTF, three downloads, one host thread per device"}), flush=True)


if __name__ == "__main__":
    main()
