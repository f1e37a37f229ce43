"""BASELINE.json configs[0]: bilateral filter radius=5 (ksize 11, sigma_space 10, sigma_color 30) on
sample_image/lenna.png (512x512, RGBA on disk -> BGR with cv2.IMREAD_COLOR, like sample/bilateral_filter/main.cpp:19)
through the reference's own C++ CPU path (include/cpp/bilateral_filter.hpp:41-124, compiled unmodified into
oracle/_ref/libvip_ref_cpu.so).  Also the adaptive bilateral filter r=7 and the BTF (k=9, 3 iterations, own-JBF
substitution) on the centre 256x256 crop.  /root/reference does not exist on the GPU box, so the
decoded input and the outputs are committed:  python tests/golden/make_golden_lenna.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
import _oracle as O  # noqa: E402


def main():
    import cv2
    img = cv2.imread("/root/reference/sample_image/lenna.png", cv2.IMREAD_COLOR)
    assert img is not None and img.shape == (512, 512, 3) and img.dtype == np.uint8
    img = np.ascontiguousarray(img)
    R = O.ref_cpu()
    R.ref_set_num_threads(os.cpu_count() or 1)
    out = {"lenna_bgr": img}
    d = np.empty_like(img)
    R.ref_bilateral_filter(img.ctypes.data, d.ctypes.data, 512, 512, 11, 10.0, 30.0)
    out["bf_r5_cpu"] = d.copy()
    crop = np.ascontiguousarray(img[128:384, 128:384])
    dc = np.empty_like(crop)
    R.ref_adaptive_bilateral_filter(crop.ctypes.data, dc.ctypes.data, 256, 256, 15, 10.0, 30.0)
    out["abf_r7_crop_cpu"] = dc.copy()
    R.ref_bilateral_texture_filter(crop.ctypes.data, dc.ctypes.data, 256, 256, 9, 3)
    out["btf_k9_n3_crop_cpu"] = dc.copy()
    np.savez_compressed(os.path.join(HERE, "lenna_vectors.npz"), **out)
    for k in sorted(out):
        print(f"{k:22s} {out[k].shape} fnv1a64={O.fnv1a64(np.ascontiguousarray(out[k]).tobytes())}")


if __name__ == "__main__":
    main()
