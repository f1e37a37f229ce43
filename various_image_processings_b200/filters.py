"""Host-side mirror of the reference's include/cuda classes on top of the C ABI (libvip_b200.so).

Same names, argument order, defaults and synchronisation behaviour as the reference
(include/cuda/{bilateral_filter,adaptive_bilateral_filter,bilateral_texture_filter,device_image,gradient}.hpp),
so the parity tests read like the reference's own tests (test/*.cu).  Device memory is addressed by raw pointers:
anything exposing `data_ptr()` (a torch CUDA tensor), a DeviceImage, or an int is accepted.  torch is only used by
callers for allocation / streams -- nothing here computes with it and there is no CPU fallback.
"""
import ctypes as C

import numpy as np

from . import capi
from .capi import Halo, Strip, check, lib


def _ptr(x):
    if x is None:
        return None
    if isinstance(x, DeviceImage):
        return x.get()
    if hasattr(x, "data_ptr"):
        if hasattr(x, "is_cuda") and not x.is_cuda:
            raise ValueError("expected a CUDA tensor (device pointer); host tensors are not accepted")
        if hasattr(x, "is_contiguous") and not x.is_contiguous():
            raise ValueError("expected a contiguous (tightly packed) tensor")
        return x.data_ptr()
    return int(x)


def _stream(stream):
    """cudaStream_t as int.  None -> torch's current stream if torch is imported and has CUDA, else the default."""
    if stream is not None:
        return int(getattr(stream, "cuda_stream", stream))
    import sys
    torch = sys.modules.get("torch")
    if torch is not None and torch.cuda.is_available():
        return int(torch.cuda.current_stream().cuda_stream)
    return 0


def device_synchronize():
    check(lib().vip_device_synchronize(), "vip_device_synchronize")


class DeviceImage:
    """DeviceImage<ElemType> (include/cuda/device_image.hpp:4-16): owning device buffer of w*h*c elements."""

    def __init__(self, width, height, channels=1, dtype=np.uint8):
        self.dtype = np.dtype(dtype)
        if self.dtype not in (np.dtype(np.uint8), np.dtype(np.float32)):
            raise TypeError("DeviceImage supports uint8 and float32 (src/device_image.cu:51-52)")
        self.width, self.height, self.channels = width, height, channels
        self._h = C.c_void_p()
        check(lib().vip_image_create(C.byref(self._h), width, height, channels, self.dtype.itemsize), "vip_image_create")

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            if capi is not None:
                capi.release('vip_image_destroy', h)

    @property
    def shape(self):
        return (self.height, self.width, self.channels) if self.channels > 1 else (self.height, self.width)

    def size(self):
        return self.width * self.height * self.channels

    def upload(self, data, stream=None):
        a = np.ascontiguousarray(data, dtype=self.dtype)
        if a.size != self.size():
            raise ValueError("upload: element count mismatch")
        if stream is None:
            check(lib().vip_image_upload(self._h, a.ctypes.data), "vip_image_upload")
        else:
            check(lib().vip_image_upload_async(self._h, a.ctypes.data, _stream(stream)), "vip_image_upload_async")

    def download(self, out=None, stream=None):
        if out is None:
            out = np.empty(self.shape, self.dtype)
        if out.size != self.size() or out.dtype != self.dtype or not out.flags.c_contiguous:
            raise ValueError("download: bad output array")
        if stream is None:
            check(lib().vip_image_download(self._h, out.ctypes.data), "vip_image_download")
        else:
            check(lib().vip_image_download_async(self._h, out.ctypes.data, _stream(stream)), "vip_image_download_async")
        return out

    def upload_rows(self, host_rows, row0, rows, stream=None):
        """Rows [row0, row0+rows) of the buffer from a packed host array (pinned for real overlap); asynchronous."""
        a = host_rows if isinstance(host_rows, np.ndarray) else np.asarray(host_rows)
        if a.dtype != self.dtype or not a.flags.c_contiguous or a.size != rows * self.width * self.channels:
            raise ValueError("upload_rows: bad host array")
        check(lib().vip_image_upload_rows_async(self._h, a.ctypes.data, row0, rows, _stream(stream)),
              "vip_image_upload_rows_async")

    def download_rows(self, out, row0, rows, stream=None):
        if out.dtype != self.dtype or not out.flags.c_contiguous or out.size != rows * self.width * self.channels:
            raise ValueError("download_rows: bad output array")
        check(lib().vip_image_download_rows_async(self._h, out.ctypes.data, row0, rows, _stream(stream)),
              "vip_image_download_rows_async")
        return out

    def row_ptr(self, row):
        """Device address of row `row` of the buffer."""
        return self.get() + row * self.width * self.channels * self.dtype.itemsize

    def get(self):
        return lib().vip_image_data(self._h)


class PinnedBuffer:
    """Page-locked host memory (vip_host_alloc) exposed as a numpy array; for the asynchronous copies."""

    def __init__(self, nbytes):
        self._p = C.c_void_p()
        check(lib().vip_host_alloc(C.byref(self._p), nbytes), "vip_host_alloc")
        self.nbytes = nbytes
        self.array = np.ctypeslib.as_array((C.c_uint8 * nbytes).from_address(self._p.value))

    def __del__(self):
        p, self._p = getattr(self, "_p", None), None
        if p:
            self.array = None
            if capi is not None:
                capi.release('vip_host_free', p)


class CudaBilateralFilter:
    """CudaBilateralFilter (include/cuda/bilateral_filter.hpp:7-29)."""

    def __init__(self, width, height, ksize=9, sigma_space=10.0, sigma_color=30.0, lut_mode=capi.VIP_LUT_CUDA):
        self.width, self.height, self.ksize = width, height, ksize
        self._h = C.c_void_p()
        check(lib().vip_bf_create_ex(C.byref(self._h), width, height, ksize, sigma_space, sigma_color, lut_mode),
              "vip_bf_create")

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            if capi is not None:
                capi.release('vip_bf_destroy', h)

    def bilateral_filter(self, d_src, d_dst):
        """Blocking, like the reference (src/bilateral_filter_impl.cu:294-300)."""
        self.bilateral_filter_async(d_src, d_dst)
        device_synchronize()

    def joint_bilateral_filter(self, d_src, d_guide, d_dst):
        self.joint_bilateral_filter_async(d_src, d_guide, d_dst)
        device_synchronize()

    def bilateral_filter_async(self, d_src, d_dst, stream=None):
        check(lib().vip_bf_run(self._h, _ptr(d_src), _ptr(d_dst), _stream(stream)), "vip_bf_run")

    def joint_bilateral_filter_async(self, d_src, d_guide, d_dst, stream=None):
        check(lib().vip_jbf_run(self._h, _ptr(d_src), _ptr(d_guide), _ptr(d_dst), _stream(stream)), "vip_jbf_run")

    def bilateral_filter_strip(self, d_src, d_dst, strip, stream=None):
        check(lib().vip_bf_run_strip(self._h, _ptr(d_src), _ptr(d_dst), C.byref(strip), _stream(stream)),
              "vip_bf_run_strip")

    def bilateral_filter_strip_peer(self, d_src, halo, d_dst, strip, stream=None):
        """Strip whose halo rows are staged from the neighbouring ranks' device memory (struct vip_halo)."""
        check(lib().vip_bf_run_strip_peer(self._h, _ptr(d_src), C.byref(halo), _ptr(d_dst), C.byref(strip), _stream(stream)),
              "vip_bf_run_strip_peer")

    def joint_bilateral_filter_strip(self, d_src, d_guide, d_dst, strip, stream=None):
        check(lib().vip_jbf_run_strip(self._h, _ptr(d_src), _ptr(d_guide), _ptr(d_dst), C.byref(strip), _stream(stream)),
              "vip_jbf_run_strip")

    def halo_rows(self):
        return lib().vip_bf_halo_rows(self._h)

    def luts(self):
        ks = np.empty((self.ksize, self.ksize), np.float32)
        ct = np.empty(768, np.float32)
        check(lib().vip_bf_get_luts(self._h, ks.ctypes.data, ct.ctypes.data), "vip_bf_get_luts")
        return ks, ct


class CudaAdaptiveBilateralFilter:
    """CudaAdaptiveBilateralFilter (include/cuda/adaptive_bilateral_filter.hpp:7-24)."""

    def __init__(self, width, height, ksize=9, sigma_space=10.0, sigma_color=30.0, lut_mode=capi.VIP_LUT_CUDA):
        self.width, self.height, self.ksize = width, height, ksize
        self._h = C.c_void_p()
        check(lib().vip_abf_create_ex(C.byref(self._h), width, height, ksize, sigma_space, sigma_color, lut_mode),
              "vip_abf_create")

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            if capi is not None:
                capi.release('vip_abf_destroy', h)

    def execute(self, d_src, d_dst):
        self.execute_async(d_src, d_dst)
        device_synchronize()

    def execute_async(self, d_src, d_dst, stream=None):
        check(lib().vip_abf_run(self._h, _ptr(d_src), _ptr(d_dst), _stream(stream)), "vip_abf_run")

    def execute_strip(self, d_src, d_dst, strip, stream=None):
        check(lib().vip_abf_run_strip(self._h, _ptr(d_src), _ptr(d_dst), C.byref(strip), _stream(stream)),
              "vip_abf_run_strip")

    def execute_strip_peer(self, d_src, halo, d_dst, strip, stream=None):
        check(lib().vip_abf_run_strip_peer(self._h, _ptr(d_src), C.byref(halo), _ptr(d_dst), C.byref(strip), _stream(stream)),
              "vip_abf_run_strip_peer")

    def halo_rows(self):
        return lib().vip_abf_halo_rows(self._h)


class CudaBilateralTextureFilter:
    """CudaBilateralTextureFilter (include/cuda/bilateral_texture_filter.hpp:7-17) plus the two stage entry
    points the reference's tests reach through impl_ (src/bilateral_texture_filter_impl.cuh:19-28)."""

    def __init__(self, width, height, ksize=9, nitr=3):
        self.width, self.height, self.ksize, self.nitr = width, height, ksize, nitr
        self._h = C.c_void_p()
        check(lib().vip_btf_create(C.byref(self._h), width, height, ksize, nitr), "vip_btf_create")

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            if capi is not None:
                capi.release('vip_btf_destroy', h)

    def execute(self, d_src, d_dst):
        self.execute_async(d_src, d_dst)
        device_synchronize()

    def execute_async(self, d_src, d_dst, stream=None):
        check(lib().vip_btf_run(self._h, _ptr(d_src), _ptr(d_dst), _stream(stream)), "vip_btf_run")

    def set_graph(self, on=True):
        """Replay a run's 3 x nitr launches as one captured CUDA graph per (src, dst) pair (non-default streams only)."""
        check(lib().vip_btf_set_graph(self._h, int(bool(on))), "vip_btf_set_graph")

    def compute_blur_and_rtv(self, d_image, d_magnitude, d_blurred, d_rtv, stream=None):
        check(lib().vip_btf_blur_rtv(self._h, _ptr(d_image), _ptr(d_magnitude), _ptr(d_blurred), _ptr(d_rtv),
                                     _stream(stream)), "vip_btf_blur_rtv")

    def compute_guide(self, d_blurred, d_rtv, d_guide, stream=None):
        check(lib().vip_btf_guide(self._h, _ptr(d_blurred), _ptr(d_rtv), _ptr(d_guide), _stream(stream)),
              "vip_btf_guide")

    def iterate_strip(self, d_src_n, d_dst_n, strip, stream=None):
        check(lib().vip_btf_iterate_strip(self._h, _ptr(d_src_n), _ptr(d_dst_n), C.byref(strip), _stream(stream)),
              "vip_btf_iterate_strip")

    def halo_rows_per_iteration(self):
        return lib().vip_btf_halo_rows_per_iteration(self._h)

    def scratch_bytes(self):
        return lib().vip_btf_scratch_bytes(self._h)


def cuda_gradient(d_src, d_dst, width, height, src_ch=1, dtype=None, stream=None):
    """cuda_gradient<T> (include/cuda/gradient.hpp:13-23): asynchronous like the reference.  The element type is
    taken from the tensor / DeviceImage dtype (uint8 or float32), or from `dtype` for raw pointers."""
    if dtype is None:
        dtype = getattr(d_src, "dtype", np.uint8)
    name = str(dtype)
    if "uint8" in name:
        check(lib().vip_gradient_u8(_ptr(d_src), _ptr(d_dst), width, height, src_ch, _stream(stream)), "vip_gradient_u8")
    elif "float32" in name:
        check(lib().vip_gradient_f32(_ptr(d_src), _ptr(d_dst), width, height, src_ch, _stream(stream)), "vip_gradient_f32")
    else:
        raise TypeError("cuda_gradient: SrcType must be uint8 or float32 (include/cuda/gradient.hpp:21)")


def cuda_gradient_strip(d_src, d_dst, width, height, src_ch, strip, stream=None):
    check(lib().vip_gradient_u8_strip(_ptr(d_src), _ptr(d_dst), width, height, src_ch, C.byref(strip), _stream(stream)),
          "vip_gradient_u8_strip")


def convert_to_bgr(d_src, src_channels, d_dst, width, height, swap_rb=False, stream=None):
    """Device twin of convert_to_3ch (sample/benchmark/main.cpp:91-103): gray / 3-channel / 4-channel -> packed BGR."""
    check(lib().vip_convert_to_bgr_u8(_ptr(d_src), src_channels, _ptr(d_dst), width, height, int(bool(swap_rb)),
                                      _stream(stream)), "vip_convert_to_bgr_u8")


def scale_to_u8(d_src, d_dst, n, max_value, stream=None):
    """convert_to_u8 of sample/gradient/main.cpp:9-18: round(src * 255 / max) saturated to uint8."""
    check(lib().vip_scale_to_u8_f32(_ptr(d_src), _ptr(d_dst), n, float(max_value), _stream(stream)), "vip_scale_to_u8_f32")


class CudaIntegralImage:
    """GPU twin of BorderReplicatedIntegralImage<uint8_t,3> (include/cpp/border_replicated_integral_image.hpp)."""

    def __init__(self, width, height, radius):
        self.width, self.height, self.radius = width, height, radius
        self.table = DeviceImage(width + 2 * radius + 1, height + 2 * radius + 1, 3, np.float32)  # 4-byte cells

    def build(self, d_src, stream=None):
        check(lib().vip_integral_u8c3(_ptr(d_src), self.table.get(), self.width, self.height, self.radius,
                                      _stream(stream)), "vip_integral_u8c3")

    def get(self, d_rects, n, d_out, stream=None):
        check(lib().vip_integral_get_u8c3(self.table.get(), self.width, self.height, self.radius, _ptr(d_rects), n,
                                          _ptr(d_out), _stream(stream)), "vip_integral_get_u8c3")

    def box_mean(self, d_mean, stream=None):
        check(lib().vip_integral_box_mean_u8c3(self.table.get(), _ptr(d_mean), self.width, self.height, self.radius,
                                               _stream(stream)), "vip_integral_box_mean_u8c3")

    def download_table(self):
        raw = self.table.download()
        return raw.view(np.uint32)


class CudaBorderReplicatedIntegralImage:
    """BorderReplicatedIntegralImage<SrcType, Channels> on the GPU for the source types / channel counts the reference
    template is instantiated with (uint8, uint16, float32; 1..4 channels;
    test/border_replicated_integral_image.cpp:63-95).  Integer tables are uint32 (modulo 2^32), float tables float32."""
    SRC = {np.dtype(np.uint8): capi.VIP_SRC_U8, np.dtype(np.uint16): capi.VIP_SRC_U16, np.dtype(np.float32): capi.VIP_SRC_F32}

    def __init__(self, width, height, radius, channels=3, dtype=np.uint8):
        self.dtype = np.dtype(dtype)
        if self.dtype not in self.SRC:
            raise TypeError("integral image source must be uint8, uint16 or float32")
        self.width, self.height, self.radius, self.channels = width, height, radius, channels
        self.src_type = self.SRC[self.dtype]
        self.table = DeviceImage(width + 2 * radius + 1, height + 2 * radius + 1, channels, np.float32)  # 4-byte cells

    @property
    def table_dtype(self):
        return np.dtype(np.float32) if self.src_type == capi.VIP_SRC_F32 else np.dtype(np.uint32)

    def build(self, d_src, stream=None):
        check(lib().vip_integral_build(_ptr(d_src), self.table.get(), self.width, self.height, self.channels, self.radius,
                                       self.src_type, _stream(stream)), "vip_integral_build")

    def get(self, d_rects, n, d_out, stream=None):
        check(lib().vip_integral_get(self.table.get(), self.width, self.height, self.channels, self.radius, self.src_type,
                                     _ptr(d_rects), n, _ptr(d_out), _stream(stream)), "vip_integral_get")

    def download_table(self):
        return self.table.download().view(self.table_dtype)


class MultiGpuFilter:
    """Row strips + halo exchange / frame batches across the GPUs of one box, driven by ONE host process
    (vip_mg_* in include/vip_c_api.h).  kind: 'bf' | 'abf' | 'btf'."""
    KINDS = {"bf": capi.VIP_MG_BF, "abf": capi.VIP_MG_ABF, "btf": capi.VIP_MG_BTF}

    BTF_HALO = {None: 0, "auto": 0, "exchange": capi.VIP_MG_FLAG_BTF_EXCHANGE, "deep": capi.VIP_MG_FLAG_BTF_DEEP_HALO}

    def __init__(self, device_ids, kind, width, height, ksize=9, sigma_space=10.0, sigma_color=30.0, nitr=3,
                 btf_halo=None, require_p2p=False):
        """btf_halo: 'deep' (nitr x 17 halo rows, redundant recompute, no communication), 'exchange' (17 rows refreshed
        peer-to-peer after every iteration) or None / 'auto' (vip_mg_create_ex decides)."""
        self.width, self.height, self.n = width, height, len(device_ids)
        ids = (C.c_int * len(device_ids))(*device_ids)
        self._h = C.c_void_p()
        check(lib().vip_mg_create_ex(C.byref(self._h), ids, len(device_ids), self.KINDS[kind], width, height, ksize,
                                     sigma_space, sigma_color, nitr,
                                     self.BTF_HALO[btf_halo] | (capi.VIP_MG_FLAG_REQUIRE_P2P if require_p2p else 0)), "vip_mg_create_ex")

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            if capi is not None:
                capi.release('vip_mg_destroy', h)

    def strip_rows(self, i):
        y0, rows = C.c_int(), C.c_int()
        check(lib().vip_mg_strip_rows(self._h, i, C.byref(y0), C.byref(rows)), "vip_mg_strip_rows")
        return y0.value, rows.value

    def scatter_host(self, h_src):
        check(lib().vip_mg_scatter_host(self._h, h_src.ctypes.data), "vip_mg_scatter_host")

    def run(self):
        check(lib().vip_mg_run(self._h), "vip_mg_run")

    def run_async(self):
        check(lib().vip_mg_run_async(self._h), "vip_mg_run_async")

    def wait(self):
        check(lib().vip_mg_wait(self._h), "vip_mg_wait")

    def gather_host(self, h_dst):
        check(lib().vip_mg_gather_host(self._h, h_dst.ctypes.data), "vip_mg_gather_host")

    def last_run_ms(self):
        ms = C.c_float()
        check(lib().vip_mg_last_run_ms(self._h, C.byref(ms)), "vip_mg_last_run_ms")
        return ms.value

    def execute(self, h_src, h_dst=None):
        h_src = np.ascontiguousarray(h_src, np.uint8)
        if h_dst is None:
            h_dst = np.empty_like(h_src)
        self.scatter_host(h_src)
        self.run()
        self.gather_host(h_dst)
        return h_dst

    def peer_access(self):
        """True: halo rows move GPU<->GPU with peer access enabled; False: some neighbour pair stages through the host."""
        return bool(lib().vip_mg_peer_access(self._h))

    def execute_batch_set(self, h_src, h_dsts, n_frames, filters=("bf", "abf", "btf")):
        """Every frame through a SET of filters (configs[4]): one upload per frame, h_dsts[name] receives filter `name`."""
        bits = {"bf": capi.VIP_MG_SET_BF, "abf": capi.VIP_MG_SET_ABF, "btf": capi.VIP_MG_SET_BTF}
        mask, arr = 0, (C.c_void_p * 3)()
        for j, name in enumerate(("bf", "abf", "btf")):
            if name in filters:
                mask |= bits[name]
                arr[j] = h_dsts[name].ctypes.data
        check(lib().vip_mg_run_batch_set_host(self._h, mask, h_src.ctypes.data, arr, n_frames), "vip_mg_run_batch_set_host")
        return h_dsts

    def execute_batch(self, h_src, h_dst, n_frames):
        check(lib().vip_mg_run_batch_host(self._h, h_src.ctypes.data, h_dst.ctypes.data, n_frames),
              "vip_mg_run_batch_host")
        return h_dst


# ---------------------------------------------------------------- one process per GPU: IPC memory + neighbour links
def ipc_export(d_ptr):
    """64-byte CUDA IPC handle (bytes) of a device allocation's BASE pointer (a DeviceImage, or a raw cudaMalloc pointer)."""
    h = (C.c_ubyte * capi.VIP_IPC_HANDLE_BYTES)()
    check(lib().vip_ipc_export(_ptr(d_ptr), h), "vip_ipc_export")
    return bytes(h)


def ipc_open(handle):
    """Device pointer (int) of another process's allocation; VipError(VIP_E_DEVICE) when there is no peer path."""
    p = C.c_void_p()
    buf = (C.c_ubyte * capi.VIP_IPC_HANDLE_BYTES).from_buffer_copy(handle)
    check(lib().vip_ipc_open(buf, C.byref(p)), "vip_ipc_open")
    return p.value


def ipc_close(ptr):
    check(lib().vip_ipc_close(ptr), "vip_ipc_close")


def copy_rows(d_dst, dst_y0, d_src, src_y0, y0, rows, width, stream=None):
    """Rows [y0, y0+rows) of a packed BGR image from one strip buffer into another (the explicit halo pull; P2P over NVLink
    when the two pointers live on different devices)."""
    check(lib().vip_copy_rows(_ptr(d_dst), dst_y0, _ptr(d_src), src_y0, y0, rows, width, _stream(stream)), "vip_copy_rows")


def make_halo(above=None, below=None):
    """struct vip_halo from (ptr, y0, rows) tuples (None at the image border)."""
    h = Halo()
    if above is not None:
        h.d_above, h.above_y0, h.above_rows = _ptr(above[0]), above[1], above[2]
    if below is not None:
        h.d_below, h.below_y0, h.below_rows = _ptr(below[0]), below[1], below[2]
    return h


class StripLink:
    """vip_link: stream-ordered "ready" / "consumed" step counters between a rank and its two strip neighbours, stored
    into the neighbour's flag block over NVLink -- no host synchronisation inside a frame pipeline."""

    def __init__(self, n_slots):
        self._h = C.c_void_p()
        check(lib().vip_link_create(C.byref(self._h), n_slots), "vip_link_create")

    def __del__(self):
        h, self._h = getattr(self, "_h", None), None
        if h:
            if capi is not None:
                capi.release('vip_link_destroy', h)

    def export(self):
        h = (C.c_ubyte * capi.VIP_IPC_HANDLE_BYTES)()
        check(lib().vip_link_export(self._h, h), "vip_link_export")
        return bytes(h)

    def connect(self, above_handle, below_handle):
        mk = lambda b: (C.c_ubyte * capi.VIP_IPC_HANDLE_BYTES).from_buffer_copy(b) if b is not None else None
        check(lib().vip_link_connect(self._h, mk(above_handle), mk(below_handle)), "vip_link_connect")

    def connect_local(self, above, below):
        check(lib().vip_link_connect_local(self._h, above._h if above else None, below._h if below else None),
              "vip_link_connect_local")

    def set_timeout_ms(self, ms):
        check(lib().vip_link_set_timeout_ms(self._h, ms), "vip_link_set_timeout_ms")

    def signal_ready(self, slot, step, stream=None):
        check(lib().vip_link_signal_ready(self._h, slot, step, _stream(stream)), "vip_link_signal_ready")

    def wait_ready(self, slot, step, stream=None):
        check(lib().vip_link_wait_ready(self._h, slot, step, _stream(stream)), "vip_link_wait_ready")

    def signal_consumed(self, slot, step, stream=None):
        check(lib().vip_link_signal_consumed(self._h, slot, step, _stream(stream)), "vip_link_signal_consumed")

    def wait_consumed(self, slot, step, stream=None):
        check(lib().vip_link_wait_consumed(self._h, slot, step, _stream(stream)), "vip_link_wait_consumed")

    def status(self):
        """Raises VipError once any wait of this link has timed out; check it before a frame is consumed."""
        check(lib().vip_link_status(self._h), "vip_link_status")

    def reset(self):
        """Clears the latched time-out (the step counters stay)."""
        check(lib().vip_link_reset(self._h), "vip_link_reset")
