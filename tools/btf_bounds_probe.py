#!/usr/bin/env python
"""Prints the measured differences behind the two bounded (not exact) BTF assertions of tests/test_gpu_parity.py, so the
bounds can be set to what is measured: 64x48 smooth image vs the reference CPU golden, lenna crop vs the glibc-expf oracle."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import _oracle as O  # noqa: E402
import various_image_processings_b200 as vip  # noqa: E402

G = np.load(os.path.join(ROOT, "tests", "golden", "reference_vectors.npz"))
LENNA = np.load(os.path.join(ROOT, "tests", "golden", "lenna_vectors.npz"))


def btf(img, k, n):
    h, w, _ = img.shape
    d = torch.empty((h, w, 3), dtype=torch.uint8, device="cuda")
    vip.CudaBilateralTextureFilter(w, h, k, n).execute(torch.from_numpy(np.ascontiguousarray(img)).cuda(), d)
    return d.cpu().numpy()


img = O.smooth_image(64, 48, seed=5)
d = np.abs(btf(img, 9, 3).astype(int) - G["btf_k9_n3_smooth_cpu"].astype(int))
print("64x48 smooth vs CPU golden: share >0 %.6f  share >1 %.6f  max %d  (n = %d)" % ((d > 0).mean(), (d > 1).mean(), d.max(), d.size))
crop = np.ascontiguousarray(LENNA["lenna_bgr"][128:384, 128:384])
got, want = btf(crop, 9, 3), O.o_btf(crop, 9, 3)
dd = np.abs(got.astype(int) - want.astype(int))
print("lenna crop vs glibc-expf oracle: share != %.6f  max %d  count %d of %d" % ((got != want).mean(), dd.max(), int((got != want).sum()), got.size))
