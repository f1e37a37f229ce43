"""CPU tests: pin the oracle (oracle/vip_oracle.c) to the reference.

  1. against the committed golden vectors, which were produced by the reference's OWN CPU headers compiled
     unmodified (tests/golden/make_golden.py) -- bit-for-bit in ARITH_CPU mode;
  2. against the reference's known-answer test (cv2.bilateralFilter on the mt19937(42) fixture,
     /root/reference/test/bilateral_filter.cu:35-63, tolerance 1);
  3. live against oracle/_ref/libvip_ref_cpu.so on other sizes, when that build is present;
  4. ARITH_CUDA mode (the arithmetic of the reference's CUDA kernels) stays within 1 LSB of the CPU mode.
"""
import os

import numpy as np
import pytest

import _oracle as O

G = np.load(os.path.join(os.path.dirname(__file__), "golden", "reference_vectors.npz"))
W = H = 50


def test_fixture_generator_matches_reference_stream():
    # std::mt19937(42): first outputs 1608637542, 3421126067 -> % 255
    assert O.mt_u8(6).tolist() == [132, 167, 181, 199, 126, 125]
    assert np.array_equal(O.mt_u8(W * H * 3).reshape(H, W, 3), G["src_mt42"])


@pytest.mark.parametrize("k,key", [(9, "bf_k9"), (11, "bf_k11")])
def test_bf_cpu_mode_bit_exact(k, key):
    assert np.array_equal(O.o_bf(G["src_mt42"], k, 10.0, 30.0, O.ARITH_CPU), G[key])


def test_bf_known_answer_cv2():
    got = O.o_bf(G["src_mt42"], 9, 10.0, 30.0, O.ARITH_CPU)
    assert np.abs(got.astype(int) - G["cv2_bilateral_k9"].astype(int)).max() <= 1
    got = O.o_bf(G["src_mt42"], 9, 10.0, 30.0, O.ARITH_CUDA)
    assert np.abs(got.astype(int) - G["cv2_bilateral_k9"].astype(int)).max() <= 1


def test_jbf_cpu_mode_bit_exact():
    src = G["src_mt42"]
    assert np.array_equal(O.o_jbf(src, src, 9, 10.0, 30.0, O.ARITH_CPU), G["jbf_k9_self"])
    assert np.array_equal(O.o_jbf(src, G["guide_seed7"], 9, 10.0, 30.0, O.ARITH_CPU), G["jbf_k9_guide7"])


@pytest.mark.parametrize("k,key", [(9, "abf_k9"), (15, "abf_k15")])
def test_abf_cpu_mode_bit_exact(k, key):
    assert np.array_equal(O.o_abf(G["src_mt42"], k, 10.0, 30.0, O.ARITH_CPU), G[key])


def test_smooth_image_cases():
    s = G["smooth_64x48"]
    assert np.array_equal(O.o_bf(s, 11, 10.0, 30.0, O.ARITH_CPU), G["bf_k11_smooth"])
    assert np.array_equal(O.o_abf(s, 15, 10.0, 30.0, O.ARITH_CPU), G["abf_k15_smooth"])
    assert np.array_equal(O.o_btf(s, 9, 3, O.ARITH_CPU), G["btf_k9_n3_smooth_cpu"])


@pytest.mark.parametrize("ch", [1, 3])
def test_gradient(ch):
    s8 = O.mt_u8(W * H * ch).reshape(H, W, ch)
    assert np.array_equal(O.o_gradient(s8 if ch > 1 else s8[..., 0]), G[f"grad_u8_c{ch}"])
    sf = O.mt_f32(W * H * ch).reshape(H, W, ch)
    assert np.array_equal(O.o_gradient(sf if ch > 1 else sf[..., 0], O.ARITH_CPU), G[f"grad_f32_c{ch}"])
    # CUDA arithmetic (fma) differs by at most a few ulp: the reference's test tolerance is 4 ulp
    a = O.o_gradient(sf if ch > 1 else sf[..., 0], O.ARITH_CUDA)
    np.testing.assert_allclose(a, G[f"grad_f32_c{ch}"], rtol=4 * 1.2e-7)


def test_btf_stages_cpu_mode_bit_exact():
    mag = O.mt_f32(W * H).reshape(H, W)
    bl, rt = O.o_blur_rtv(G["src_mt42"], mag, 9, O.ARITH_CPU)
    assert np.array_equal(bl, G["btf_blurred_cpu"])
    assert np.array_equal(rt, G["btf_rtv_cpu"])
    blurred_in = O.mt_f32(W * H * 3).reshape(H, W, 3)
    rtv_in = O.mt_f32(W * H, 1.0).reshape(H, W)
    assert np.array_equal(O.o_guide(blurred_in, rtv_in, 9, O.ARITH_CPU), G["btf_guide_cpu"])
    # CUDA arithmetic: rtv within the reference's 4-ulp test tolerance
    bl2, rt2 = O.o_blur_rtv(G["src_mt42"], mag, 9, O.ARITH_CUDA)
    assert np.array_equal(bl2, bl)
    np.testing.assert_allclose(rt2, rt, rtol=4 * 1.2e-7)


@pytest.mark.parametrize("r", [1, 3, 5])
def test_integral_all_rectangles(r):
    src = G["ii_src_5x5"]
    table = O.o_integral(src, r)
    rects, sums = G[f"ii_rects_r{r}"], G[f"ii_sums_r{r}"]
    out = np.empty(3, np.uint32)
    for (x0, y0, x1, y1), want in zip(rects[::7], sums[::7]):
        O.oracle().vipo_integral_get(table.ctypes.data, 5, r, int(x0), int(y0), int(x1), int(y1), out.ctypes.data)
        assert out.astype(np.int64).tolist() == want.tolist()
    # brute force over the clamp-padded image as an independent check of the golden sums
    pad = np.pad(src.astype(np.int64), ((r, r), (r, r), (0, 0)), mode="edge")
    for (x0, y0, x1, y1), want in zip(rects[::97], sums[::97]):
        assert pad[y0 + r:y1 + r + 1, x0 + r:x1 + r + 1].sum(axis=(0, 1)).tolist() == want.tolist()


def test_cuda_mode_within_one_lsb_of_cpu_mode():
    src = G["src_mt42"]
    for fn, k in ((O.o_bf, 9), (O.o_abf, 9), (O.o_abf, 15)):
        a = fn(src, k, 10.0, 30.0, O.ARITH_CUDA).astype(int)
        b = fn(src, k, 10.0, 30.0, O.ARITH_CPU).astype(int)
        assert np.abs(a - b).max() <= 1


def test_rows_variant_matches_full():
    img = O.smooth_image(70, 60, seed=11)
    full = O.o_bf(img, 7)
    part = O.o_rows("bf", img, 13, 20, 7)
    assert np.array_equal(full[13:33], part)
    full = O.o_abf(img, 7)
    part = O.o_rows("abf", img, 40, 20, 7)
    assert np.array_equal(full[40:60], part)


@pytest.mark.skipif(not O.have_ref_cpu(), reason="oracle/_ref/libvip_ref_cpu.so not built (needs /root/reference)")
@pytest.mark.parametrize("w,h,k", [(37, 23, 5), (64, 31, 13), (20, 20, 21)])
def test_live_against_reference_cpu_build(w, h, k):
    R = O.ref_cpu()
    R.ref_set_num_threads(2)
    img = O.rand_image(w, h, seed=w * h + k)
    d = np.empty_like(img)
    R.ref_bilateral_filter(img.ctypes.data, d.ctypes.data, w, h, k, 6.0, 25.0)
    assert np.array_equal(d, O.o_bf(img, k, 6.0, 25.0, O.ARITH_CPU))
    R.ref_adaptive_bilateral_filter(img.ctypes.data, d.ctypes.data, w, h, k, 6.0, 25.0)
    assert np.array_equal(d, O.o_abf(img, k, 6.0, 25.0, O.ARITH_CPU))
    if k <= 13:
        R.ref_bilateral_texture_filter(img.ctypes.data, d.ctypes.data, w, h, 5, 2)
        assert np.array_equal(d, O.o_btf(img, 5, 2, O.ARITH_CPU))


# ------------------------------------------------------------------ BASELINE configs[0]: lenna.png through the reference CPU path
LENNA = np.load(os.path.join(os.path.dirname(__file__), "golden", "lenna_vectors.npz"))


def test_lenna_bf_r5_cpu_mode_bit_exact():
    """configs[0]: the restatement reproduces the reference's C++ CPU path on lenna.png bit for bit."""
    img = LENNA["lenna_bgr"]
    assert img.shape == (512, 512, 3)
    assert np.array_equal(O.o_bf(img, 11, 10.0, 30.0, O.ARITH_CPU), LENNA["bf_r5_cpu"])


def test_lenna_crop_abf_btf_cpu_mode_bit_exact():
    crop = np.ascontiguousarray(LENNA["lenna_bgr"][128:384, 128:384])
    assert np.array_equal(O.o_abf(crop, 15, 10.0, 30.0, O.ARITH_CPU), LENNA["abf_r7_crop_cpu"])
    assert np.array_equal(O.o_btf(crop, 9, 3, O.ARITH_CPU), LENNA["btf_k9_n3_crop_cpu"])
