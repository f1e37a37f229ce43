"""CPU tests of the drop-in boundary: the C-ABI library loads, exports every symbol include/vip_c_api.h declares,
the ctypes table covers the header, and argument validation answers without a GPU (no compute calls here)."""
import ctypes as C
import os
import re
import subprocess

import pytest

import various_image_processings_b200 as vip
from various_image_processings_b200 import capi

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "vip_c_api.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"VIP_API\s+[\w\s\*]+?\b(vip_\w+)\s*\(", text)))


def test_header_symbols_are_exported_and_bound():
    names = declared_symbols()
    assert len(names) >= 50
    L = capi.lib()
    for n in names:
        assert hasattr(L, n), f"{n} declared in vip_c_api.h but not exported"
        assert n in capi.SIGNATURES, f"{n} has no ctypes signature in capi.py"
    for n in capi.SIGNATURES:
        assert n in names, f"{n} bound in capi.py but not declared in the header"


def test_dynamic_symbol_table_has_no_stray_c_symbols():
    out = subprocess.check_output(["nm", "-D", "--defined-only", capi.LIB_PATH], text=True)
    c_syms = [l.split()[-1] for l in out.splitlines() if " T " in l and not l.split()[-1].startswith("_Z")]
    assert all(s.startswith("vip_") for s in c_syms), c_syms


def test_version_and_status_strings():
    L = capi.lib()
    assert b"sm_100a" in L.vip_version()
    assert capi.status_string(0) == "ok"
    assert "odd" in capi.status_string(-3)
    assert capi.status_string(2)  # a cudaError_t value maps to the runtime's text


def test_argument_validation_precedes_cuda():
    L = capi.lib()
    h = C.c_void_p()
    assert L.vip_bf_create(C.byref(h), 64, 64, 8, 10.0, 30.0) == -3      # even ksize (reference: UB)
    assert L.vip_bf_create(C.byref(h), 0, 64, 9, 10.0, 30.0) == -2
    assert L.vip_bf_create(C.byref(h), 64, 64, 9, 0.0, 30.0) == -4
    assert L.vip_abf_create(C.byref(h), 64, 64, 9, 10.0, float("nan")) == -4
    assert L.vip_btf_create(C.byref(h), 64, 64, 4, 3) == -3
    assert L.vip_bf_run(None, None, None, None) == -1
    assert L.vip_gradient_u8(None, None, 4, 4, 3, None) == -1
    assert L.vip_integral_table_elems(5, 5, 1) == 8 * 8 * 3
    assert L.vip_image_create(C.byref(h), 4, 4, 3, 2) == -7
    assert h.value is None


def test_python_mirror_has_reference_interface():
    # names and defaults of include/cuda/*.hpp
    import inspect
    sig = inspect.signature(vip.CudaBilateralFilter.__init__)
    assert [p for p in sig.parameters][1:6] == ["width", "height", "ksize", "sigma_space", "sigma_color"]
    assert sig.parameters["ksize"].default == 9 and sig.parameters["sigma_space"].default == 10.0
    assert sig.parameters["sigma_color"].default == 30.0
    sig = inspect.signature(vip.CudaBilateralTextureFilter.__init__)
    assert sig.parameters["ksize"].default == 9 and sig.parameters["nitr"].default == 3
    for cls, meth in ((vip.CudaBilateralFilter, "bilateral_filter"), (vip.CudaBilateralFilter, "joint_bilateral_filter"),
                      (vip.CudaAdaptiveBilateralFilter, "execute"), (vip.CudaBilateralTextureFilter, "execute"),
                      (vip.DeviceImage, "upload"), (vip.DeviceImage, "download"), (vip.DeviceImage, "get")):
        assert callable(getattr(cls, meth))


def test_no_fallback_on_host_tensors():
    import torch
    with pytest.raises(ValueError):
        vip.filters._ptr(torch.zeros(4, dtype=torch.uint8))


def test_cpp_dropin_headers_compile_without_cuda(tmp_path):
    """A user translation unit written against the reference's include paths compiles with plain g++."""
    src = tmp_path / "user.cpp"
    src.write_text('#include <cstdint>\n#include "cuda/bilateral_filter.hpp"\n#include "cuda/adaptive_bilateral_filter.hpp"\n'
                   '#include "cuda/bilateral_texture_filter.hpp"\n#include "cuda/device_image.hpp"\n#include "cuda/gradient.hpp"\n'
                   "int user(int w, int h, const std::uint8_t* host, std::uint8_t* out) {\n"
                   "  DeviceImage<std::uint8_t> a(w, h, 3), b(w, h, 3); a.upload(host);\n"
                   "  CudaBilateralFilter f(w, h, 11); f.bilateral_filter(a.get(), b.get());\n"
                   "  CudaAdaptiveBilateralFilter g(w, h); g.execute(a.get(), b.get());\n"
                   "  CudaBilateralTextureFilter t(w, h, 9, 3); t.execute(a.get(), b.get());\n"
                   "  DeviceImage<float> m(w, h); cuda_gradient(a.get(), m.get(), w, h, 3);\n"
                   "  CudaBorderReplicatedIntegralImage<std::uint8_t, 3> ii(w, h, 7); ii.build(a.get());\n"
                   "  CudaBorderReplicatedIntegralImage<float, 1> fi(w, h, 2); fi.build(m.get());\n"
                   "  b.download(out); return 0; }\n")
    so = tmp_path / "libuser.so"
    subprocess.check_call(["g++", "-std=c++17", "-shared", "-fPIC", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(so),
                           "-L", os.path.dirname(capi.LIB_PATH), "-lvip_b200"])
    assert so.exists()


def test_impl_surface_of_the_reference_tests_compiles(tmp_path):
    """The reference's tests include "<name>_impl.cuh", subclass the public classes and call impl_->... with
    thrust::device_vector arguments (test/bilateral_texture_filter.cu:115-136).  The same pattern builds against this
    library with include/cuda/impl on the include path in place of the reference's src/ (nvcc cross-compiles here)."""
    import shutil
    if shutil.which("nvcc") is None:
        pytest.skip("nvcc not on PATH")
    so = tmp_path / "libpattern.so"
    subprocess.check_call(["nvcc", "-std=c++17", "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-Xcompiler", "-fPIC",
                           "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "include", "cuda", "impl"),
                           os.path.join(ROOT, "tests", "impl_surface_pattern.cu"), "-o", str(so),
                           "-L", os.path.dirname(capi.LIB_PATH), "-lvip_b200"])
    assert so.exists()


def test_division_by_constant_shortcut_is_exact(tmp_path):
    """csrc/btf.cu replaces __fdiv_rn(x, K*K) and __fdiv_rn(x, 3.f) by a 3-instruction sequence; exhaustive proof."""
    exe = tmp_path / "div_const_check"
    subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", os.path.join(ROOT, "tests", "div_const_check.c"), "-o", str(exe), "-lm"])
    out = subprocess.check_output([str(exe)], text=True)
    assert "mismatches 0" in out


def test_every_valid_filter_has_a_launch_configuration():
    """vip_stencil_plan_query is pure host logic: sweep window sizes 1..41 x sigma_color 0.5..1000 x the three filters and
    require a configuration that fits the 227 KB of shared memory of an SM.  (ABF with sigma_color >~ 50 once had none:
    1536 non-zero table entries x 32 copies = 192 KB; it now runs with 8 copies.)"""
    import ctypes as C
    import various_image_processings_b200 as vip
    L = vip.capi.lib()
    seen = set()
    for mode in (0, 1, 2):
        for k in range(1, 43, 2):
            for sc in (0.5, 1.7320508, 5.0, 30.0, 50.0, 60.0, 106.0, 200.0, 400.0, 1000.0):
                plan = vip.capi.StencilPlan()
                assert L.vip_stencil_plan_query(mode, k, 10.0, sc, C.byref(plan)) == vip.capi.VIP_OK
                if k // 2 > 15:
                    assert plan.tiled == 0
                    continue
                assert plan.tiled == 1 and plan.warp_groups >= 1, (mode, k, sc)
                assert 0 < plan.smem_bytes <= 227 * 1024, (mode, k, sc, plan.smem_bytes)
                assert plan.lut_copies in (8, 32) and 2 <= plan.lut_entries <= (1536 if mode == 2 else 768)
                seen.add((plan.lut_copies, plan.lut_clamped))
    assert seen == {(32, 0), (32, 1), (8, 1)}           # all three table layouts occur
    # the configured workloads keep three warp groups (ABF: four, at 64 registers per thread) and a conflict-free table
    for mode, k, ss, sc in ((0, 11, 10.0, 30.0), (2, 15, 10.0, 30.0), (1, 17, 8.0, 1.7320508), (0, 31, 10.0, 30.0)):
        plan = vip.capi.StencilPlan()
        assert L.vip_stencil_plan_query(mode, k, ss, sc, C.byref(plan)) == 0
        assert (plan.warp_groups, plan.lut_copies) == (4 if mode == 2 else 3, 32), (mode, k, plan.warp_groups, plan.lut_copies)
    bad = vip.capi.StencilPlan()
    assert L.vip_stencil_plan_query(0, 8, 10.0, 30.0, C.byref(bad)) == vip.capi.VIP_E_KSIZE
    assert L.vip_stencil_plan_query(3, 9, 10.0, 30.0, C.byref(bad)) == vip.capi.VIP_E_UNSUPPORTED
    assert L.vip_stencil_plan_query(0, 9, 10.0, 30.0, None) == vip.capi.VIP_E_NULL


def test_multigpu_strip_plan_invariants():
    """vip_mg_plan_query (host logic of csrc/mg.cu): the strips tile the image; every pass finds its halo rows in the
    buffers; with deep halos iteration `it` can produce owned rows +- (nitr-1-it) x halo from what iteration it-1 produced;
    with the exchange every neighbour owns enough rows to serve a halo."""
    import ctypes as C
    import random
    import various_image_processings_b200 as vip
    L, cap = vip.capi.lib(), vip.capi
    rnd = random.Random(7)
    modes_seen = set()
    for _ in range(3000):
        n, h = rnd.randint(1, 8), rnd.randint(1, 5000)
        kind = rnd.choice([cap.VIP_MG_BF, cap.VIP_MG_ABF, cap.VIP_MG_BTF])
        k, nitr = rnd.choice([1, 3, 5, 9, 15, 31]), rnd.randint(0, 4)
        flags = rnd.choice([0, cap.VIP_MG_FLAG_BTF_EXCHANGE, cap.VIP_MG_FLAG_BTF_DEEP_HALO])
        plans, rc = [], 0
        for i in range(n):
            p = cap.MgPlan()
            rc = L.vip_mg_plan_query(n, i, kind, h, k, nitr, flags, C.byref(p))
            if rc != 0:
                break
            plans.append(p)
        halo = (k - 1) + 2 * (k // 2) + 1 if kind == cap.VIP_MG_BTF else k // 2
        if rc != 0:
            # the only refusal for valid arguments: BTF halo exchange with strips too thin to serve a neighbour's halo
            assert rc == cap.VIP_E_SIZE and kind == cap.VIP_MG_BTF and n > 1 and nitr > 1 and h // n < halo
            assert flags != cap.VIP_MG_FLAG_BTF_DEEP_HALO
            continue
        assert plans[0].y0 == 0 and plans[-1].y0 + plans[-1].rows == h
        deep = plans[0].btf_deep_halo
        modes_seen.add((kind == cap.VIP_MG_BTF, deep))
        for i, p in enumerate(plans):
            assert p.halo == halo and p.btf_deep_halo == deep and p.rows >= 0
            if i:
                assert p.y0 == plans[i - 1].y0 + plans[i - 1].rows                  # contiguous, no overlap
            b0, b1 = p.buf_y0, p.buf_y0 + p.buf_rows
            assert 0 <= b0 <= p.y0 and p.y0 + p.rows <= b1 <= h
            if kind != cap.VIP_MG_BTF or not deep:
                assert b0 == max(0, p.y0 - halo) and b1 == min(h, p.y0 + p.rows + halo)
                if n > 1 and kind == cap.VIP_MG_BTF and nitr > 1:
                    assert p.rows >= halo                                            # can serve a neighbour's halo
            else:
                assert flags != cap.VIP_MG_FLAG_BTF_EXCHANGE and nitr > 1 and n > 1
                have = (b0, b1)                                                      # rows valid in the input of iteration 0
                for it in range(nitr):
                    m = (nitr - 1 - it) * halo
                    o0, o1 = max(0, p.y0 - m), min(h, p.y0 + p.rows + m)             # what csrc/mg.cu asks iteration `it` for
                    assert have[0] <= max(0, o0 - halo) and min(h, o1 + halo) <= have[1]
                    have = (o0, o1)
                assert have == (p.y0, p.y0 + p.rows)
    assert modes_seen == {(False, 0), (True, 0), (True, 1)}
