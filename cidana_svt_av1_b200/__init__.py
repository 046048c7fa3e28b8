"""cidana_svt_av1_b200 -- B200-native residual-coding path of SVT-AV1 (forward 2-D transform,
quantize/dequantize/eob, inverse transform + reconstruction) behind the encoder's own entry
points.  The compute lives in libsvtgpu.so (hand-written CUDA for sm_100a, C ABI in
include/svt_gpu.h); this package is the thin host-side mirror of the reference interface.

There is no CPU fallback: importing works anywhere, but every compute call needs the CUDA
library and a B200 and raises loudly otherwise.
"""
from .api import (SvtGpu, SvtGpuError, SvtGpuMulti, av1_estimate_transform, av1_inv_transform_recon,  # noqa: F401
                  av1_inv_transform_recon8bit, av1_quantize_inv_quantize, lib, load_library)
from . import workloads  # noqa: F401

__all__ = ["SvtGpu", "SvtGpuError", "SvtGpuMulti", "lib", "load_library", "workloads", "av1_estimate_transform",
           "av1_quantize_inv_quantize", "av1_inv_transform_recon", "av1_inv_transform_recon8bit"]
