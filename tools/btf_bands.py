#!/usr/bin/env python
"""Does running the three BTF stages band by band keep the intermediates (blurred 12 B/px + rtv 4 B/px + guide 3 B/px) in
the 126 MB L2 and pay off?  (Round-2 review, item N1: "the cheap variant ... was not tried".)
Per iteration the frame is cut into `nb` row bands; vip_btf_iterate_strip runs gradient+blur/mRTV -> guide -> joint
bilateral for one band before the next band starts, so a band's intermediates are written and re-read while they are
still L2-resident (4K: a 540-row band holds 39 MB of them).  The price: every band recomputes the stage halos
(4 + 8 rows of blur/mRTV, 8 rows of guide on each side) and the iteration is 3 x nb launches instead of 3.
Output is checked against the whole-frame call (bit for bit).
  python tools/btf_bands.py [WxH] [k] [nitr]
"""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import various_image_processings_b200 as vip  # noqa: E402


def timed(fn, n=20, warm=3):
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
    w, h = map(int, (sys.argv[1] if len(sys.argv) > 1 else "3840x2160").split("x"))
    k = int(sys.argv[2]) if len(sys.argv) > 2 else 9
    nitr = int(sys.argv[3]) if len(sys.argv) > 3 else 3
    # several frames so that a new step does not find its input in L2 (in + out of 6 frames at 4K = 300 MB)
    frames = [torch.from_numpy(np.random.RandomState(40 + i).randint(0, 256, size=(h, w, 3)).astype(np.uint8)).cuda() for i in range(6)]
    dst = torch.empty_like(frames[0])
    want = torch.empty_like(frames[0])
    f = vip.CudaBilateralTextureFilter(w, h, k, nitr)
    pp = [torch.empty_like(frames[0]), torch.empty_like(frames[0])]
    state = {"i": 0}

    def whole():
        state["i"] += 1
        f.execute_async(frames[state["i"] % len(frames)], dst)

    def banded(nb):
        edges = [h * b // nb for b in range(nb + 1)]

        def run():
            state["i"] += 1
            cur = frames[state["i"] % len(frames)]
            for it in range(nitr):
                out = dst if it == nitr - 1 else pp[it & 1]
                for b in range(nb):
                    f.iterate_strip(cur, out, vip.Strip(0, h, 0, edges[b], edges[b + 1] - edges[b]))
                cur = out
        return run

    f.execute_async(frames[0], want)
    torch.cuda.synchronize()
    ms0 = timed(whole)
    print(f"BTF {w}x{h} k={k} nitr={nitr}: whole frame (3 launches / iteration)  {ms0:7.3f} ms", flush=True)
    for nb in (1, 2, 4, 8, 16, 32):
        run = banded(nb)
        state["i"] = -1
        run()
        torch.cuda.synchronize()
        same = "EXACT" if torch.equal(dst, want) else "DIFF"
        ms = timed(run)
        band_mb = (h // nb) * w * 19 / 2**20
        print(f"  {nb:3d} bands of {h // nb:5d} rows ({band_mb:6.1f} MiB of intermediates per band)  {ms:7.3f} ms  x{ms0 / ms:5.3f} vs whole  {same}", flush=True)


if __name__ == "__main__":
    main()
