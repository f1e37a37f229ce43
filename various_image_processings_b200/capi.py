"""ctypes binding of libvip_b200.so (the C ABI declared in include/vip_c_api.h).

The library is the product: there is NO Python / CPU fallback.  Importing this module without a built
library raises; calling a compute entry point without a CUDA device returns a CUDA error status which is
turned into VipError.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# VIP_B200_LIB: developer override to A/B a differently built library (tools/kbench.py); never a fallback
LIB_PATH = os.environ.get("VIP_B200_LIB") or os.path.join(_HERE, "lib", "libvip_b200.so")

VIP_OK = 0
(VIP_E_NULL, VIP_E_SIZE, VIP_E_KSIZE, VIP_E_SIGMA, VIP_E_STRIP, VIP_E_ALLOC, VIP_E_UNSUPPORTED,
 VIP_E_DEVICE) = range(-1, -9, -1)
VIP_LUT_CUDA, VIP_LUT_CPU = 0, 1
VIP_MG_BF, VIP_MG_ABF, VIP_MG_BTF = 0, 1, 2
VIP_MG_FLAG_BTF_EXCHANGE, VIP_MG_FLAG_BTF_DEEP_HALO, VIP_MG_FLAG_REQUIRE_P2P = 1, 2, 4
VIP_MG_SET_BF, VIP_MG_SET_ABF, VIP_MG_SET_BTF = 1, 2, 4
VIP_SRC_U8, VIP_SRC_U16, VIP_SRC_F32 = 0, 1, 2
VIP_IPC_HANDLE_BYTES = 64


class VipError(RuntimeError):
    def __init__(self, status, what):
        self.status = status
        super().__init__(f"{what}: {status_string(status)} (status {status})")


class Strip(C.Structure):
    """struct vip_strip: rows held by the buffers and rows to produce (all in full-image row numbers)."""
    _fields_ = [("src_y0", C.c_int), ("src_rows", C.c_int), ("dst_y0", C.c_int), ("out_y0", C.c_int),
                ("out_rows", C.c_int)]


class Halo(C.Structure):
    """struct vip_halo: the neighbouring ranks' buffers a *_strip_peer call stages its halo rows from."""
    _fields_ = [("d_above", C.c_void_p), ("above_y0", C.c_int), ("above_rows", C.c_int),
                ("d_below", C.c_void_p), ("below_y0", C.c_int), ("below_rows", C.c_int)]


class StencilPlan(C.Structure):
    """struct vip_stencil_plan: how a bilateral-family filter will run (host logic, answers without a GPU)."""
    _fields_ = [("tiled", C.c_int), ("lut_entries", C.c_int), ("lut_copies", C.c_int), ("lut_clamped", C.c_int),
                ("warp_groups", C.c_int), ("smem_bytes", C.c_int), ("lut_rows", C.c_int)]


class MgPlan(C.Structure):
    """struct vip_mg_plan: rows a device slot owns / holds and the BTF halo mode (host logic, answers without a GPU)."""
    _fields_ = [("y0", C.c_int), ("rows", C.c_int), ("buf_y0", C.c_int), ("buf_rows", C.c_int), ("halo", C.c_int),
                ("btf_deep_halo", C.c_int)]


_vp, _i, _f, _sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
_pp = C.POINTER(C.c_void_p)

# name -> (restype, argtypes); this table is also what tests/test_abi.py checks against include/vip_c_api.h
SIGNATURES = {
    "vip_version": (C.c_char_p, []),
    "vip_status_string": (C.c_char_p, [_i]),
    "vip_launch_count": (C.c_ulonglong, []),
    "vip_stencil_plan_query": (_i, [_i, _i, _f, _f, C.POINTER(StencilPlan)]),
    "vip_image_create": (_i, [_pp, _i, _i, _i, _i]),
    "vip_image_destroy": (_i, [_vp]),
    "vip_image_upload": (_i, [_vp, _vp]),
    "vip_image_download": (_i, [_vp, _vp]),
    "vip_image_upload_async": (_i, [_vp, _vp, _vp]),
    "vip_image_download_async": (_i, [_vp, _vp, _vp]),
    "vip_image_upload_rows_async": (_i, [_vp, _vp, _i, _i, _vp]),
    "vip_image_download_rows_async": (_i, [_vp, _vp, _i, _i, _vp]),
    "vip_image_data": (_vp, [_vp]),
    "vip_image_bytes": (_sz, [_vp]),
    "vip_host_alloc": (_i, [_pp, _sz]),
    "vip_host_free": (_i, [_vp]),
    "vip_stream_synchronize": (_i, [_vp]),
    "vip_device_synchronize": (_i, []),
    "vip_bf_create": (_i, [_pp, _i, _i, _i, _f, _f]),
    "vip_bf_create_ex": (_i, [_pp, _i, _i, _i, _f, _f, _i]),
    "vip_bf_destroy": (_i, [_vp]),
    "vip_bf_run": (_i, [_vp, _vp, _vp, _vp]),
    "vip_jbf_run": (_i, [_vp, _vp, _vp, _vp, _vp]),
    "vip_bf_run_strip": (_i, [_vp, _vp, _vp, C.POINTER(Strip), _vp]),
    "vip_jbf_run_strip": (_i, [_vp, _vp, _vp, _vp, C.POINTER(Strip), _vp]),
    "vip_bf_run_strip_peer": (_i, [_vp, _vp, C.POINTER(Halo), _vp, C.POINTER(Strip), _vp]),
    "vip_bf_halo_rows": (_i, [_vp]),
    "vip_bf_get_luts": (_i, [_vp, _vp, _vp]),
    "vip_abf_create": (_i, [_pp, _i, _i, _i, _f, _f]),
    "vip_abf_create_ex": (_i, [_pp, _i, _i, _i, _f, _f, _i]),
    "vip_abf_destroy": (_i, [_vp]),
    "vip_abf_run": (_i, [_vp, _vp, _vp, _vp]),
    "vip_abf_run_strip": (_i, [_vp, _vp, _vp, C.POINTER(Strip), _vp]),
    "vip_abf_run_strip_peer": (_i, [_vp, _vp, C.POINTER(Halo), _vp, C.POINTER(Strip), _vp]),
    "vip_abf_halo_rows": (_i, [_vp]),
    "vip_gradient_u8": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "vip_gradient_f32": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "vip_gradient_u8_strip": (_i, [_vp, _vp, _i, _i, _i, C.POINTER(Strip), _vp]),
    "vip_convert_to_bgr_u8": (_i, [_vp, _i, _vp, _i, _i, _i, _vp]),
    "vip_scale_to_u8_f32": (_i, [_vp, _vp, _sz, _f, _vp]),
    "vip_integral_u8c3": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "vip_integral_u8c3_noalloc": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "vip_integral_get_u8c3": (_i, [_vp, _i, _i, _i, _vp, _i, _vp, _vp]),
    "vip_integral_table_elems": (_sz, [_i, _i, _i]),
    "vip_integral_build": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp]),
    "vip_integral_get": (_i, [_vp, _i, _i, _i, _i, _i, _vp, _i, _vp, _vp]),
    "vip_integral_box_mean_u8c3": (_i, [_vp, _vp, _i, _i, _i, _vp]),
    "vip_btf_create": (_i, [_pp, _i, _i, _i, _i]),
    "vip_btf_destroy": (_i, [_vp]),
    "vip_btf_run": (_i, [_vp, _vp, _vp, _vp]),
    "vip_btf_set_graph": (_i, [_vp, _i]),
    "vip_btf_blur_rtv": (_i, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "vip_btf_guide": (_i, [_vp, _vp, _vp, _vp, _vp]),
    "vip_btf_halo_rows_per_iteration": (_i, [_vp]),
    "vip_btf_iterate_strip": (_i, [_vp, _vp, _vp, C.POINTER(Strip), _vp]),
    "vip_btf_scratch_bytes": (_sz, [_vp]),
    "vip_mg_create": (_i, [_pp, C.POINTER(C.c_int), _i, _i, _i, _i, _i, _f, _f, _i]),
    "vip_mg_create_ex": (_i, [_pp, C.POINTER(C.c_int), _i, _i, _i, _i, _i, _f, _f, _i, _i]),
    "vip_mg_plan_query": (_i, [_i, _i, _i, _i, _i, _i, _i, C.POINTER(MgPlan)]),
    "vip_mg_destroy": (_i, [_vp]),
    "vip_mg_strip_rows": (_i, [_vp, _i, C.POINTER(C.c_int), C.POINTER(C.c_int)]),
    "vip_mg_scatter_host": (_i, [_vp, _vp]),
    "vip_mg_scatter_device": (_i, [_vp, _vp]),
    "vip_mg_run": (_i, [_vp]),
    "vip_mg_run_async": (_i, [_vp]),
    "vip_mg_wait": (_i, [_vp]),
    "vip_mg_gather_host": (_i, [_vp, _vp]),
    "vip_mg_last_run_ms": (_i, [_vp, C.POINTER(C.c_float)]),
    "vip_mg_run_batch_host": (_i, [_vp, _vp, _vp, _i]),
    "vip_mg_run_batch_set_host": (_i, [_vp, _i, _vp, C.POINTER(C.c_void_p), _i]),
    "vip_mg_peer_access": (_i, [_vp]),
    "vip_ipc_export": (_i, [_vp, _vp]),
    "vip_ipc_open": (_i, [_vp, _pp]),
    "vip_ipc_close": (_i, [_vp]),
    "vip_peer_can_access": (_i, [_i, _i]),
    "vip_copy_rows": (_i, [_vp, _i, _vp, _i, _i, _i, _i, _vp]),
    "vip_link_create": (_i, [_pp, _i]),
    "vip_link_destroy": (_i, [_vp]),
    "vip_link_export": (_i, [_vp, _vp]),
    "vip_link_connect": (_i, [_vp, _vp, _vp]),
    "vip_link_connect_local": (_i, [_vp, _vp, _vp]),
    "vip_link_set_timeout_ms": (_i, [_vp, _i]),
    "vip_link_signal_ready": (_i, [_vp, _i, C.c_uint, _vp]),
    "vip_link_wait_ready": (_i, [_vp, _i, C.c_uint, _vp]),
    "vip_link_signal_consumed": (_i, [_vp, _i, C.c_uint, _vp]),
    "vip_link_wait_consumed": (_i, [_vp, _i, C.c_uint, _vp]),
    "vip_link_status": (_i, [_vp]),
    "vip_link_reset": (_i, [_vp]),
}

_lib = None


def lib():
    """The loaded library (loads on first use; raises if it has not been built)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise ImportError(
                f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                "or `make -C various_image_processings_b200/csrc`. There is no CPU fallback.")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def status_string(status):
    return lib().vip_status_string(int(status)).decode()


def check(status, what="vip call"):
    if status != VIP_OK:
        raise VipError(status, what)


def release(fn_name, handle):
    """Destructor helper: frees a handle; silent during interpreter shutdown (module globals already torn down)."""
    L = _lib
    if L is not None and handle:
        getattr(L, fn_name)(handle)


def launch_count():
    return int(lib().vip_launch_count())
