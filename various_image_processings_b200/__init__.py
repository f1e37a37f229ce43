"""B200-native edge-preserving filters: bilateral, joint/adaptive bilateral, bilateral texture filter.

Drop-in for the CUDA classes of yuyuyu-bot/various_image_processings (include/cuda/*.hpp): the compute lives in
lib/libvip_b200.so (hand-written sm_100a kernels behind the C ABI of include/vip_c_api.h); this package is the
Python mirror of the reference's host interface.  There is no CPU fallback.
"""
from . import capi
from .capi import Halo, Strip, VipError, launch_count
from .strips import StripPlan, halo_rows, plan_strips, to_strip
from .filters import (CudaAdaptiveBilateralFilter, CudaBilateralFilter, CudaBilateralTextureFilter, CudaBorderReplicatedIntegralImage,
                      CudaIntegralImage,
                      DeviceImage, MultiGpuFilter, PinnedBuffer, convert_to_bgr, cuda_gradient, cuda_gradient_strip, scale_to_u8,
                      device_synchronize, StripLink, copy_rows, ipc_close, ipc_export, ipc_open, make_halo)

__all__ = [
    "capi", "Halo", "StripLink", "copy_rows", "ipc_close", "ipc_export", "ipc_open", "make_halo", "Strip", "VipError", "launch_count", "CudaAdaptiveBilateralFilter", "CudaBilateralFilter",
    "CudaBilateralTextureFilter", "CudaBorderReplicatedIntegralImage", "CudaIntegralImage", "DeviceImage", "MultiGpuFilter", "PinnedBuffer",
    "convert_to_bgr", "scale_to_u8", "cuda_gradient", "cuda_gradient_strip", "device_synchronize", "StripPlan", "halo_rows", "plan_strips", "to_strip",
]
