"""The headless sample programs and the benchmark harness (apps/).

CPU: the PNG / PNM codec the programs use instead of cv::imread is checked against cv2 on files cv2 wrote.
GPU: each program is run like the reference's samples (same argv) on sample-sized inputs and its output file is compared
bit for bit with the Python mirror of the same class; the benchmark prints the reference's line format.
"""
import json
import os
import re
import subprocess

import numpy as np
import pytest

cv2 = pytest.importorskip("cv2")

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BIN = os.path.join(ROOT, "apps", "bin")


@pytest.fixture(scope="module")
def apps():
    subprocess.check_call(["make", "-s", "-C", os.path.join(ROOT, "apps")])
    return BIN


def lenna():
    return np.load(os.path.join(ROOT, "tests", "golden", "lenna_vectors.npz"))["lenna_bgr"]


@pytest.mark.parametrize("kind", ["rgba", "rgb", "gray", "ppm", "pgm"])
@pytest.mark.parametrize("out_ext", [".png", ".ppm"])
def test_image_codec_against_cv2(apps, tmp_path, kind, out_ext):
    rs = np.random.RandomState(5)
    h, w = 37, 53
    bgr = rs.randint(0, 256, size=(h, w, 3)).astype(np.uint8)
    bgr[5:20, 7:30] = 200                                       # flat area: exercises the PNG filters
    src = str(tmp_path / ("in" + {"ppm": ".ppm", "pgm": ".pgm"}.get(kind, ".png")))
    if kind == "rgba":
        cv2.imwrite(src, np.dstack([bgr, rs.randint(0, 256, size=(h, w)).astype(np.uint8)]))
    elif kind in ("rgb", "ppm"):
        cv2.imwrite(src, bgr)
    else:
        cv2.imwrite(src, bgr[..., 0])
    gray = kind in ("gray", "pgm")
    dst = str(tmp_path / ("out" + (".pgm" if gray and out_ext == ".ppm" else out_ext)))
    out = subprocess.check_output([os.path.join(apps, "imgio_roundtrip"), src, dst], text=True).split()
    assert out[:2] == [str(w), str(h)]
    back = cv2.imread(dst, cv2.IMREAD_UNCHANGED)
    assert np.array_equal(back, bgr[..., 0] if gray else bgr)


def test_lenna_png_decodes_like_cv2(apps, tmp_path):
    """cv2 writes the committed lenna fixture as PNG (with the alpha channel the reference's file has); our decoder + the
    RGBA->BGR device conversion must give back what cv::imread(IMREAD_COLOR) gives."""
    img = lenna()
    src = str(tmp_path / "lenna_rgba.png")
    cv2.imwrite(src, np.dstack([img, np.full(img.shape[:2], 255, np.uint8)]))
    dst = str(tmp_path / "lenna.ppm")
    subprocess.check_call([os.path.join(apps, "imgio_roundtrip"), src, dst], stdout=subprocess.DEVNULL)
    assert np.array_equal(cv2.imread(dst, cv2.IMREAD_COLOR), img)


# ------------------------------------------------------------------------------------------------ GPU
def _gpu_vip():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import various_image_processings_b200 as vip
    return torch, vip


@pytest.mark.gpu
@pytest.mark.parametrize("prog,args", [("bilateral_filter", ["11", "10", "30"]), ("adaptive_bilateral_filter", ["15"]),
                                       ("bilateral_texture_filter", ["9", "2"]), ("gradient", [])])
def test_sample_programs_match_the_oracle(apps, tmp_path, prog, args):
    """The headless sample programs (decode -> layout conversion -> filter -> encode) against the ORACLE on the same
    pixels, not against this library called a second way: a decode / channel-order slip shared by apps/image_io.hpp and
    the Python side would otherwise go unseen.  BF / ABF / gradient: the C restatement in CUDA arithmetic, exact.
    BTF: the reference's own CUDA kernels (oracle/_ref, required on a GPU box), exact; the restatement (glibc expf
    instead of the device's) bounds it to 1 level on a handful of samples."""
    torch, vip = _gpu_vip()
    import _oracle as O
    img = np.ascontiguousarray(lenna()[100:356, 60:380])          # 320 x 256
    h, w, _ = img.shape
    src = str(tmp_path / "in.png")
    cv2.imwrite(src, np.dstack([img, np.full((h, w), 255, np.uint8)]))   # RGBA on disk, like sample_image/lenna.png
    dst = str(tmp_path / "out.png")
    subprocess.check_call([os.path.join(apps, prog), src, *args, "-o", dst], stdout=subprocess.DEVNULL)
    if prog == "gradient":
        g = O.o_gradient(img)
        want = np.clip(np.rint(g * np.float32(255.0) / np.float32(g.max())), 0, 255).astype(np.uint8)
        assert np.array_equal(cv2.imread(dst, cv2.IMREAD_UNCHANGED), want)
        return
    got = cv2.imread(dst, cv2.IMREAD_COLOR)
    if prog == "bilateral_filter":
        assert np.array_equal(got, O.o_bf(img, 11, 10.0, 30.0))
    elif prog == "adaptive_bilateral_filter":
        assert np.array_equal(got, O.o_abf(img, 15))
    else:
        assert O.have_ref_cuda(), "oracle/_ref/libvip_ref_cuda.so must travel to the GPU box (make -C oracle ref)"
        R = O.ref_cuda()
        d_src = torch.from_numpy(img).cuda()
        d_ref = torch.empty_like(d_src)
        p = R.refcu_btf_create(w, h, 9, 2)
        R.refcu_btf_run(p, d_src.data_ptr(), d_ref.data_ptr())
        R.refcu_btf_destroy(p)
        torch.cuda.synchronize()
        assert np.array_equal(got, d_ref.cpu().numpy())
        soft = O.o_btf(img, 9, 2)
        diff = np.abs(got.astype(int) - soft.astype(int))
        assert diff.max() <= 1 and (diff != 0).mean() < 2e-4


@pytest.mark.gpu
def test_benchmark_harness_output(apps, tmp_path):
    _gpu_vip()
    cfg = tmp_path / "config.toml"
    cfg.write_text("execute_times = 3\n[BilateralFilter]\nksize = 11\n[AdaptiveBilateralFilter]\nksize = 7\n"
                   "[BilateralTextureFilter]\nksize = 5\nnitr  = 2\n[SuperpixelSLIC]\nsuperpixel_size = 10\n")
    out = subprocess.check_output([os.path.join(apps, "benchmark"), str(cfg), "--width", "640", "--height", "360",
                                   "--device-time", "--json"], text=True)
    assert "execute times : 3" in out and "ksize : 11" in out and "nitr  : 2" in out
    for name in ("gradient", "bilateral filter", "adaptive bilateral filter", "bilateral texture filter"):
        assert re.search(r"^%s \[cuda\]\s+:\s+\d+\.\d{6} \[msec\]$" % re.escape(name), out, re.M), name   # sample/benchmark/main.cpp:35-37
    js = json.loads(out.strip().splitlines()[-1])
    assert js["width"] == 640 and len(js["results"]) == 4 and js["launches"] > 0
