"""Property-based GPU parity (hypothesis): random sizes, window sizes, sigmas, content classes and strip splits, every case
bit-exact against the oracle in CUDA arithmetic (oracle/vip_oracle.c) through the C ABI.  Sizes stay small enough for the
scalar oracle; the fixed-size / full-size cases live in test_gpu_parity.py.  derandomize=True: every run draws the same
examples (a new random failure at round end would be noise, not signal; widen max_examples when hunting)."""
import os

import numpy as np
import pytest

import _oracle as O

torch = pytest.importorskip("torch")
hyp = pytest.importorskip("hypothesis")
from hypothesis import HealthCheck, given, settings, strategies as st  # noqa: E402

pytestmark = pytest.mark.gpu
HUNT = int(os.environ.get("VIP_HYP_HUNT", "0"))          # e.g. VIP_HYP_HUNT=10: ten times the examples, fresh random seeds
SET = dict(max_examples=60 * max(HUNT, 1), deadline=None, derandomize=not HUNT, suppress_health_check=[HealthCheck.function_scoped_fixture, HealthCheck.too_slow])


@pytest.fixture(scope="module")
def vip():
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    import various_image_processings_b200 as v
    torch.cuda.set_device(0)
    return v


def make_image(w, h, kind, seed):
    rs = np.random.RandomState(seed)
    if kind == "rand":
        return rs.randint(0, 256, size=(h, w, 3)).astype(np.uint8)
    if kind == "narrow":                                   # the reference benchmark's content (sample/benchmark/main.cpp:213)
        return rs.randint(100, 120, size=(h, w, 3)).astype(np.uint8)
    if kind == "const":
        return np.full((h, w, 3), rs.randint(0, 256), np.uint8)
    if kind == "extreme":                                  # only 0 and 255: largest colour distances, ABF index up to 1530
        return (rs.randint(0, 2, size=(h, w, 3)) * 255).astype(np.uint8)
    return O.smooth_image(w, h, seed=seed)


def dev(a):
    return torch.from_numpy(np.ascontiguousarray(a)).cuda()


sizes = st.tuples(st.integers(1, 150), st.integers(1, 90))
kinds = st.sampled_from(["rand", "narrow", "const", "extreme", "smooth"])
odd_k = st.integers(0, 20).map(lambda r: 2 * r + 1)        # radius <= 15: tiled kernels, beyond: the generic kernel
sigma = st.floats(0.5, 200.0, allow_nan=False, allow_infinity=False, width=32)


@settings(**SET)
@given(wh=sizes, k=odd_k, ss=sigma, sc=sigma, kind=kinds, seed=st.integers(0, 10**6))
def test_bf_jbf_random_cases(vip, wh, k, ss, sc, kind, seed):
    w, h = wh
    img, guide = make_image(w, h, kind, seed), make_image(w, h, "smooth", seed + 1)
    d = torch.empty((h, w, 3), dtype=torch.uint8, device="cuda")
    f = vip.CudaBilateralFilter(w, h, k, ss, sc)
    f.bilateral_filter(dev(img), d)
    assert np.array_equal(d.cpu().numpy(), O.o_bf(img, k, ss, sc))
    f.joint_bilateral_filter(dev(img), dev(guide), d)
    assert np.array_equal(d.cpu().numpy(), O.o_jbf(img, guide, k, ss, sc))


@settings(**SET)
@given(wh=sizes, k=odd_k, ss=sigma, sc=sigma, kind=kinds, seed=st.integers(0, 10**6))
def test_abf_random_cases(vip, wh, k, ss, sc, kind, seed):
    w, h = wh
    img = make_image(w, h, kind, seed)
    d = torch.empty((h, w, 3), dtype=torch.uint8, device="cuda")
    vip.CudaAdaptiveBilateralFilter(w, h, k, ss, sc).execute(dev(img), d)
    assert np.array_equal(d.cpu().numpy(), O.o_abf(img, k, ss, sc))


@settings(max_examples=12 * max(HUNT, 1), deadline=None, derandomize=not HUNT, suppress_health_check=SET["suppress_health_check"])
@given(wh=st.tuples(st.integers(1, 120), st.integers(1, 90)), k=st.sampled_from([3, 5, 9, 13]), nitr=st.integers(0, 3),
       kind=st.sampled_from(["rand", "narrow", "smooth"]), seed=st.integers(0, 10**6))
def test_btf_stages_random_cases(vip, wh, k, nitr, kind, seed):
    """blur / mRTV and guide stages exact (integer + ordered-float arithmetic); the end-to-end filter equals the oracle
    except where device expf differs from glibc expf by one level of the guide: bounded."""
    w, h = wh
    img = make_image(w, h, kind, seed)
    f = vip.CudaBilateralTextureFilter(w, h, k, nitr)
    mag = O.o_gradient(img)
    bl = torch.empty((h, w, 3), dtype=torch.float32, device="cuda")
    rt = torch.empty((h, w), dtype=torch.float32, device="cuda")
    f.compute_blur_and_rtv(dev(img), dev(mag), bl, rt)
    torch.cuda.synchronize()
    obl, ort = O.o_blur_rtv(img, mag, k)
    assert np.array_equal(bl.cpu().numpy(), obl) and np.array_equal(rt.cpu().numpy(), ort)
    g = torch.empty((h, w, 3), dtype=torch.uint8, device="cuda")
    f.compute_guide(bl, rt, g)
    torch.cuda.synchronize()
    og = O.o_guide(obl, ort, k)
    dg = np.abs(g.cpu().numpy().astype(int) - og.astype(int))
    assert dg.max() <= 1 and (dg != 0).sum() <= max(2, 1e-3 * dg.size)   # argmin exact; alpha blend within one expf ulp
    d = torch.empty((h, w, 3), dtype=torch.uint8, device="cuda")
    f.execute(dev(img), d)
    want = O.o_btf(img, k, nitr)
    got = d.cpu().numpy()
    assert (got != want).sum() <= max(6, 5e-3 * got.size)


@settings(max_examples=15 * max(HUNT, 1), deadline=None, derandomize=not HUNT, suppress_health_check=SET["suppress_health_check"])
@given(w=st.integers(8, 140), h=st.integers(30, 120), k=st.sampled_from([3, 9, 15]), cut=st.floats(0.05, 0.95),
       kind=st.sampled_from(["bf", "abf"]), seed=st.integers(0, 10**6))
def test_two_strips_equal_whole_image(vip, w, h, k, cut, kind, seed):
    """Any split into two row strips with radius-deep halos reproduces the whole-image result bit for bit."""
    img = make_image(w, h, "rand", seed)
    r = k // 2
    y_cut = min(max(int(h * cut), 1), h - 1)
    F = vip.CudaBilateralFilter(w, h, k) if kind == "bf" else vip.CudaAdaptiveBilateralFilter(w, h, k)
    whole = torch.empty((h, w, 3), dtype=torch.uint8, device="cuda")
    (F.bilateral_filter if kind == "bf" else F.execute)(dev(img), whole)
    out = np.empty_like(img)
    for y0, y1 in ((0, y_cut), (y_cut, h)):
        s0, s1 = max(0, y0 - r), min(h, y1 + r)
        d = torch.empty((y1 - y0, w, 3), dtype=torch.uint8, device="cuda")
        strip = vip.Strip(s0, s1 - s0, y0, y0, y1 - y0)
        (F.bilateral_filter_strip if kind == "bf" else F.execute_strip)(dev(img[s0:s1]), d, strip)
        torch.cuda.synchronize()
        out[y0:y1] = d.cpu().numpy()
    assert np.array_equal(out, whole.cpu().numpy())


# ------------------------------------------------------------------ streaming stages (numpy restatements are exact)
@settings(**SET)
@given(wh=st.tuples(st.integers(1, 700), st.integers(1, 60)), ch=st.integers(1, 4), f32=st.booleans(), seed=st.integers(0, 10**6))
def test_gradient_random_cases(vip, wh, ch, f32, seed):
    w, h = wh
    rs = np.random.RandomState(seed)
    src = rs.randint(0, 256, size=(h, w, ch)).astype(np.uint8)
    if f32:
        src = (src.astype(np.float32) * np.float32(1.37)).astype(np.float32)
    d = torch.empty((h, w), dtype=torch.float32, device="cuda")
    vip.cuda_gradient(dev(src), d, w, h, ch)
    torch.cuda.synchronize()
    assert np.array_equal(d.cpu().numpy(), O.o_gradient(src if ch > 1 else src[..., 0]))


@settings(**SET)
@given(wh=st.tuples(st.integers(1, 900), st.integers(1, 300)), r=st.integers(0, 20), ch=st.integers(1, 4),
       dtype=st.sampled_from([np.uint8, np.uint16]), seed=st.integers(0, 10**6))
def test_integral_table_random_cases(vip, wh, r, ch, dtype, seed):
    """Border-replicated integral table of any size / radius / channel count, exact modulo 2^32 (the integer tables of
    the reference are int, include/cpp/border_replicated_integral_image.hpp)."""
    w, h = wh
    src = np.random.RandomState(seed).randint(0, np.iinfo(dtype).max + 1, size=(h, w, ch)).astype(dtype)
    ii = vip.CudaBorderReplicatedIntegralImage(w, h, r, ch, dtype)
    ii.build(dev(src))
    torch.cuda.synchronize()
    got = ii.download_table().reshape(h + 2 * r + 1, w + 2 * r + 1, ch)
    pad = np.pad(src.astype(np.uint64), ((r, r), (r, r), (0, 0)), mode="edge")
    want = np.pad(np.cumsum(np.cumsum(pad, 0), 1), ((1, 0), (1, 0), (0, 0)))
    assert np.array_equal(got, (want & 0xFFFFFFFF).astype(np.uint32))


@settings(**SET)
@given(wh=st.tuples(st.integers(1, 500), st.integers(1, 40)), ch=st.sampled_from([1, 3, 4]), swap=st.booleans(),
       so=st.integers(0, 3), do=st.integers(0, 3), seed=st.integers(0, 10**6))
def test_convert_random_cases(vip, wh, ch, swap, so, do, seed):
    """convert_to_3ch on the device (sample/benchmark/main.cpp:91-103) at any size and any byte alignment of either buffer."""
    w, h = wh
    src = np.random.RandomState(seed).randint(0, 256, size=(h, w, ch)).astype(np.uint8)
    sbuf = torch.zeros((h * w * ch + so,), dtype=torch.uint8, device="cuda")
    sbuf[so:] = dev(src).reshape(-1)
    dbuf = torch.full((h * w * 3 + do + 8,), 77, dtype=torch.uint8, device="cuda")
    vip.convert_to_bgr(sbuf[so:], ch, dbuf[do:], w, h, swap)
    torch.cuda.synchronize()
    want = np.repeat(src, 3, axis=2) if ch == 1 else (src[..., 2::-1] if swap else src[..., :3])
    out = dbuf.cpu().numpy()
    assert np.array_equal(out[do:do + h * w * 3].reshape(h, w, 3), want)
    assert np.all(out[:do] == 77) and np.all(out[do + h * w * 3:] == 77)          # nothing written outside the image


# ------------------------------------------------------------------ multi-GPU driver on virtual device slots
@settings(max_examples=20 * max(HUNT, 1), deadline=None, derandomize=not HUNT, suppress_health_check=SET["suppress_health_check"])
@given(n=st.integers(1, 5), h=st.integers(90, 260), w=st.integers(8, 130), kind=st.sampled_from(["bf", "abf", "btf"]),
       k=st.sampled_from([3, 5, 9]), nitr=st.integers(0, 3), mode=st.sampled_from(["auto", "deep", "exchange"]),
       seed=st.integers(0, 10**6))
def test_multigpu_strips_random_cases(vip, n, h, w, kind, k, nitr, mode, seed):
    """Any number of device slots / image height / window: the strip planner, both BTF halo modes and the gather give the
    single-device result bit for bit.  Slots are dealt to the devices that exist (one GPU: all on cuda:0)."""
    img = make_image(w, h, "smooth" if seed & 1 else "rand", seed)
    ndev = torch.cuda.device_count()
    mg = vip.MultiGpuFilter([i % ndev for i in range(n)], kind, w, h, ksize=k, nitr=nitr, btf_halo=mode)
    got = mg.execute(img)
    d = torch.empty((h, w, 3), dtype=torch.uint8, device="cuda")
    if kind == "bf":
        vip.CudaBilateralFilter(w, h, k).bilateral_filter(dev(img), d)
    elif kind == "abf":
        vip.CudaAdaptiveBilateralFilter(w, h, k).execute(dev(img), d)
    else:
        vip.CudaBilateralTextureFilter(w, h, k, nitr).execute(dev(img), d)
    assert np.array_equal(got, d.cpu().numpy())


@settings(max_examples=20 * max(HUNT, 1), deadline=None, derandomize=not HUNT, suppress_health_check=SET["suppress_health_check"])
@given(w=st.integers(4, 120), h=st.integers(20, 150), k=st.sampled_from([3, 5, 9]), a=st.floats(0, 1), b=st.floats(0, 1),
       slack=st.tuples(st.integers(0, 9), st.integers(0, 9)), kind=st.sampled_from(["jbf", "btf", "grad"]), seed=st.integers(0, 10**6))
def test_strip_calls_random_row_ranges(vip, w, h, k, a, b, slack, kind, seed):
    """Any output row range [y0, y1) from buffers that hold exactly the rows it needs (+ random slack): joint bilateral,
    one BTF iteration and the gradient write the same rows a whole-image call produces (the contract of struct vip_strip)."""
    img, guide = make_image(w, h, "rand", seed), make_image(w, h, "smooth", seed + 1)
    y0, y1 = sorted((int(a * h), int(b * h)))
    y1 = min(h, max(y1, y0 + 1))
    f_btf = vip.CudaBilateralTextureFilter(w, h, k, 1)
    halo = {"jbf": k // 2, "btf": f_btf.halo_rows_per_iteration(), "grad": 1}[kind]
    s0, s1 = max(0, y0 - halo - slack[0]), min(h, y1 + halo + slack[1])
    strip = vip.Strip(s0, s1 - s0, y0, y0, y1 - y0)
    if kind == "grad":
        whole = torch.empty((h, w), dtype=torch.float32, device="cuda")
        vip.cuda_gradient(dev(img), whole, w, h, 3)
        part = torch.full((y1 - y0, w), -1.0, dtype=torch.float32, device="cuda")
        vip.cuda_gradient_strip(dev(img[s0:s1]), part, w, h, 3, strip)
    else:
        whole = torch.empty((h, w, 3), dtype=torch.uint8, device="cuda")
        part = torch.empty((y1 - y0, w, 3), dtype=torch.uint8, device="cuda")
        if kind == "jbf":
            F = vip.CudaBilateralFilter(w, h, k)
            F.joint_bilateral_filter(dev(img), dev(guide), whole)
            F.joint_bilateral_filter_strip(dev(img[s0:s1]), dev(guide[s0:s1]), part, strip)
        else:
            f_btf.execute(dev(img), whole)                       # nitr = 1: one iteration
            vip.CudaBilateralTextureFilter(w, h, k, 1).iterate_strip(dev(img[s0:s1]), part, strip)
    torch.cuda.synchronize()
    assert torch.equal(part, whole[y0:y1])


@settings(max_examples=30 * max(HUNT, 1), deadline=None, derandomize=not HUNT, suppress_health_check=SET["suppress_health_check"])
@given(w=st.integers(4, 160), n=st.integers(2, 5), rows=st.integers(1, 60), k=st.sampled_from([3, 5, 9, 15, 21, 31, 35]),
       kind=st.sampled_from(["bf", "abf"]), content=kinds, seed=st.integers(0, 10**6))
def test_peer_halo_strips_random_cases(vip, w, n, rows, k, kind, content, seed):
    """Strips whose halo rows are staged by the kernel from the NEIGHBOURS' buffers (struct vip_halo): every rank holds only
    its own rows.  Any width, strip height >= radius (a halo never spans more than one neighbour), tiled and generic
    kernels (k = 35), every content class: bit-exact against the oracle."""
    r = k // 2
    rows = max(rows, r)                      # a neighbour must be able to serve a whole halo
    h = rows * n + (seed % 3)
    img = make_image(w, h, content, seed)
    want = O.o_abf(img, k) if kind == "abf" else O.o_bf(img, k)
    f = vip.CudaAdaptiveBilateralFilter(w, h, k) if kind == "abf" else vip.CudaBilateralFilter(w, h, k)
    plans = vip.plan_strips(h, n, r)
    own = [dev(img[p.y0:p.y1]) for p in plans]
    for i, p in enumerate(plans):
        above = (own[i - 1], plans[i - 1].y0, plans[i - 1].y1 - plans[i - 1].y0) if i > 0 else None
        below = (own[i + 1], plans[i + 1].y0, plans[i + 1].y1 - plans[i + 1].y0) if i + 1 < n else None
        d = torch.zeros((p.y1 - p.y0, w, 3), dtype=torch.uint8, device="cuda")
        st_ = vip.Strip(p.y0, p.y1 - p.y0, p.y0, p.y0, p.y1 - p.y0)
        (f.execute_strip_peer if kind == "abf" else f.bilateral_filter_strip_peer)(own[i], vip.make_halo(above, below), d, st_)
        torch.cuda.synchronize()
        assert np.array_equal(d.cpu().numpy(), want[p.y0:p.y1]), (i, p)
