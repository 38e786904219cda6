"""ctypes binding of tools/_build/libsynthgen.so (tools/synthgen.cu): seeded synthetic rectified stereo frames, the
same bytes from the device kernels and from the host loop.  Bench / test input only."""
import ctypes as C
from pathlib import Path

import numpy as np

_LIB = None


def lib():
    global _LIB
    if _LIB is None:
        so = Path(__file__).resolve().parent / "_build" / "libsynthgen.so"
        if not so.exists():
            raise RuntimeError(f"{so} missing: run `make -C tools` (or __graft_entry__.build())")
        L = C.CDLL(str(so))
        for f in (L.synthgen_device, L.synthgen_host):
            f.restype = C.c_int
        L.synthgen_device.argtypes = [C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p,
                                      C.c_size_t, C.c_void_p]
        L.synthgen_host.argtypes = [C.c_uint64, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_size_t, C.c_void_p,
                                    C.c_size_t, C.c_int]
        _LIB = L
    return _LIB


def host_frames(seed0, n, w=752, h=480, threads=0):
    """frames seed0 .. seed0+n-1 as (left, right) uint8 [n, h, w]"""
    import os
    left = np.empty((n, h, w), np.uint8); right = np.empty((n, h, w), np.uint8)
    rc = lib().synthgen_host(seed0, n, w, h, left.ctypes.data, w * h, right.ctypes.data, w * h,
                             threads or (os.cpu_count() or 1))
    if rc:
        raise RuntimeError(f"synthgen_host: {rc}")
    return left, right


def device_frames_into(seed0, n, w, h, dev_left, stride_l, dev_right, stride_r, stream=0):
    """generate into device memory (raw pointers, image strides in bytes), asynchronously on `stream`"""
    rc = lib().synthgen_device(seed0, n, w, h, dev_left, stride_l, dev_right, stride_r, stream)
    if rc:
        raise RuntimeError(f"synthgen_device: cudaError {rc}")


def device_batch_input(seed0, n, w, h, dev_in, stream=0):
    """generate n stereo frames in gforb_batch's input layout: image 2f = left, 2f+1 = right, tight w*h images"""
    device_frames_into(seed0, n, w, h, dev_in, 2 * w * h, dev_in + w * h, 2 * w * h, stream)
