#!/usr/bin/env python
"""Developer loop on the GPU box: time each kernel of the hot path at BASELINE sizes and check it bit-for-bit against
the reference's own CUDA kernels (oracle/_ref/libvip_ref_cuda.so -- checker only) on the same random input.
  python tools/kbench.py [bf:3840x2160:11 abf:3840x2160:15:10:100 abf:3840x2160:15 jbf:3840x2160:17 btf:3840x2160:9 ...]
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oracle as O  # noqa: E402
import various_image_processings_b200 as vip  # noqa: E402

DEFAULT = ["bf:3840x2160:11", "abf:3840x2160:15", "jbf:3840x2160:17", "bf:7680x4320:31", "btf:3840x2160:9", "bf:1920x1080:9", "abf:1920x1080:9", "grad:3840x2160:3", "integral:3840x2160:7"]


def taps(r):
    return sum(1 for y in range(-r, r + 1) for x in range(-r, r + 1) if x * x + y * y <= r * r)


def timed(fn, n=10, warm=2):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


def main():
    R = O.ref_cuda() if O.have_ref_cuda() else None
    content = os.environ.get("KB_CONTENT", "rand")
    for spec in (sys.argv[1:] or DEFAULT):
        kind, size, k, *sig = spec.split(":")          # optional sigma_space, sigma_color after the window size
        ss, sc = (float(sig[0]), float(sig[1])) if len(sig) == 2 else (10.0, 30.0)
        w, h = map(int, size.split("x"))
        k = int(k)
        rs = np.random.RandomState(42)
        if content == "rand":                                     # S-rand, the headline content
            img = rs.randint(0, 256, size=(h, w, 3)).astype(np.uint8)
        elif content == "narrow":                                 # S-narrow: the reference benchmark's uniform [100, 120)
            img = rs.randint(100, 120, size=(h, w, 3)).astype(np.uint8)
        elif content == "lenna":                                  # S-nat: the reference's sample image, mirrored tiling
            t = np.load(os.path.join(ROOT, "tests", "golden", "lenna_vectors.npz"))["lenna_bgr"]
            t = np.concatenate([np.concatenate([t, t[:, ::-1]], 1), np.concatenate([t[::-1], t[::-1, ::-1]], 1)], 0)
            img = np.ascontiguousarray(np.tile(t, (h // t.shape[0] + 1, w // t.shape[1] + 1, 1))[:h, :w])
        else:
            t = O.smooth_image(256, 216, seed=8)
            img = np.ascontiguousarray(np.tile(t, (h // 216 + 1, w // 256 + 1, 1))[:h, :w])
        guide = np.ascontiguousarray(img[::-1, ::-1])
        d_src, d_guide = torch.from_numpy(img).cuda(), torch.from_numpy(guide).cuda()
        d_dst, d_ref = torch.empty_like(d_src), torch.empty_like(d_src)
        ref_ms = float("nan")
        if kind in ("grad", "integral"):
            # streaming stages: report achieved GB/s of algorithmic bytes (k = channels for grad, radius for integral)
            if kind == "grad":
                g = torch.empty((h, w), dtype=torch.float32, device="cuda")
                run = lambda: vip.cuda_gradient(d_src, g, w, h, 3)
                nbytes = w * h * 7
                same = "n/a"
                if R:
                    gr = torch.empty_like(g)
                    R.refcu_gradient_u8(d_src.data_ptr(), gr.data_ptr(), w, h, 3)
                    run()
                    torch.cuda.synchronize()
                    same = "EXACT" if torch.equal(g, gr) else "DIFF"
            else:
                ii = vip.CudaIntegralImage(w, h, k)
                run = lambda: ii.build(d_src)
                nbytes = (w + 2 * k + 1) * (h + 2 * k + 1) * 15
                run()
                torch.cuda.synchronize()
                tab = ii.download_table().reshape(h + 2 * k + 1, w + 2 * k + 1, 3)
                pad = np.pad(img[:64].astype(np.uint64), ((k, 0), (k, k), (0, 0)), mode="edge")
                want = np.cumsum(np.cumsum(pad, 0), 1).astype(np.uint32)
                same = "EXACT(top rows)" if np.array_equal(tab[1:1 + pad.shape[0], 1:], want) else "DIFF"
            ms = timed(run)
            print(f"{spec:22s} {ms:8.3f} ms  {w * h / ms / 1e3:9.0f} Mpix/s  {nbytes / ms / 1e6:8.1f} GB/s algorithmic  {same}", flush=True)
            continue
        if kind in ("bf", "jbf"):
            f = vip.CudaBilateralFilter(w, h, k, ss, sc)
            run = (lambda: f.bilateral_filter_async(d_src, d_dst)) if kind == "bf" else (lambda: f.joint_bilateral_filter_async(d_src, d_guide, d_dst))
            t_per_px = taps(k // 2)
            if R:
                p = R.refcu_bf_create(w, h, k, ss, sc)
                rr = (lambda: R.refcu_bf_run(p, d_src.data_ptr(), d_ref.data_ptr())) if kind == "bf" else (lambda: R.refcu_jbf_run(p, d_src.data_ptr(), d_guide.data_ptr(), d_ref.data_ptr()))
        elif kind == "abf":
            f = vip.CudaAdaptiveBilateralFilter(w, h, k, ss, sc)
            run = lambda: f.execute_async(d_src, d_dst)
            t_per_px = taps(k // 2)
            if R:
                p = R.refcu_abf_create(w, h, k, ss, sc)
                rr = lambda: R.refcu_abf_run(p, d_src.data_ptr(), d_ref.data_ptr())
        else:
            f = vip.CudaBilateralTextureFilter(w, h, k, 3)
            run = lambda: f.execute_async(d_src, d_dst)
            t_per_px = 3 * taps(k - 1)
            if R:
                p = R.refcu_btf_create(w, h, k, 3)
                rr = lambda: R.refcu_btf_run(p, d_src.data_ptr(), d_ref.data_ptr())
        ms = timed(run)
        same = "n/a"
        if R:
            ref_ms = timed(rr, n=2, warm=1)
            same = "EXACT" if torch.equal(d_dst, d_ref) else "DIFF(%d px)" % int((d_dst != d_ref).any(-1).sum().item())
        alg_bytes = w * h * (9 if kind == "jbf" else 6)     # u8 BGR in (+ guide) and out: the HBM floor
        print(f"{spec:22s} {ms:8.3f} ms  {w * h / ms / 1e3:9.0f} Mpix/s  {w * h * t_per_px / ms / 1e6:8.1f} Gtap/s  {alg_bytes / ms / 1e6:7.1f} GB/s alg   ref-cuda {ref_ms:9.3f} ms  x{ref_ms / ms:5.1f}  {same}", flush=True)


if __name__ == "__main__":
    if os.environ.get("KB_STREAM", "0") == "1":      # a non-default stream (CUDA-graph replay needs one)
        with torch.cuda.stream(torch.cuda.Stream()):
            main()
    else:
        main()
