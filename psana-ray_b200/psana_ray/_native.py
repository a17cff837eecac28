"""ctypes binding of libpsana_ray_cuda.so (declarations: include/psana_ray_cuda.h).

There is no CPU fallback: if the library is missing or will not load, importing callers get a
``NativeLibraryError`` that says how to build it.
"""
from __future__ import annotations

import ctypes as C
import os

PSR_OK = 0
PSR_WAIT_INPUT, PSR_WAIT_ALL = 0, 1
PSR_DET_EPIX10KA, PSR_DET_JUNGFRAU = 0, 1
PSR_OUT_CALIB, PSR_OUT_IMAGE = 0, 1
PSR_CM_ROWS, PSR_CM_COLS, PSR_CM_BANKS = 1, 2, 4
PSR_IPC_HANDLE_BYTES = 64
PSR_NOALLOC_SCRATCH_BYTES = 32 << 20

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "lib", "libpsana_ray_cuda.so")


class NativeLibraryError(RuntimeError):
    pass


class PsrError(RuntimeError):
    def __init__(self, code, message):
        super().__init__(f"libpsana_ray_cuda error {code}: {message}")
        self.code = code


_vp, _i, _f, _sz, _u32 = C.c_void_p, C.c_int, C.c_float, C.c_size_t, C.c_uint32
_pvp = C.POINTER(C.c_void_p)
_pi = C.POINTER(C.c_int)

# name -> (restype, argtypes); mirrors include/psana_ray_cuda.h one to one
SIGNATURES = {
    "psr_version": (_i, []),
    "psr_last_error": (C.c_char_p, []),
    "psr_device_count": (_i, [_pi]),
    "psr_create": (_i, [_pvp, _i, _i, _i, _i, _i]),
    "psr_destroy": (None, [_vp]),
    "psr_set_constants": (_i, [_vp, _vp, _vp, _vp, _vp, _vp, _i]),
    "psr_set_common_mode": (_i, [_vp, _i, _f, _i, _u32]),
    "psr_set_geometry": (_i, [_vp, _vp, _vp]),
    "psr_image_shape": (_i, [_vp, _pi, _pi]),
    "psr_set_keep_mask": (_i, [_vp, _vp, _sz, _i]),
    "psr_reserve": (_i, [_vp, _i]),
    "psr_set_option": (_i, [_vp, C.c_char_p, _i]),
    "psr_calib_batch": (_i, [_vp, _vp, _i, _vp, _i, _vp]),
    "psr_calib_image_noalloc": (_i, [_vp, _vp, _i, _vp, _vp]),
    "psr_decode_gain_modes": (_i, [_vp, _vp, _i, _vp, _vp]),
    "psr_calib_batch_host": (_i, [_vp, _vp, _i, _vp, _vp, _i]),
    "psr_calib_batch_host_begin": (_i, [_vp, _vp, _i, _vp, _vp, _i, _vp]),
    "psr_calib_batch_host_wait": (_i, [_vp, _i]),
    "psr_set_host_chunk": (_i, [_vp, _i, _i]),
    "psr_host_alloc": (_i, [_pvp, _sz]),
    "psr_host_alloc_wc": (_i, [_pvp, _sz]),
    "psr_host_free": (_i, [_vp]),
    "psr_host_register": (_i, [_vp, _sz]),
    "psr_host_unregister": (_i, [_vp]),
    "psr_device_alloc": (_i, [_pvp, _sz, _i]),
    "psr_device_free": (_i, [_vp, _i]),
    "psr_memcpy_h2d": (_i, [_vp, _vp, _sz, _vp]),
    "psr_memcpy_d2h": (_i, [_vp, _vp, _sz, _vp]),
    "psr_stream_synchronize": (_i, [_vp]),
    "psr_ipc_export_mem": (_i, [_vp, _vp]),
    "psr_ipc_open_mem": (_i, [_vp, _i, _pvp]),
    "psr_ipc_close_mem": (_i, [_vp, _i]),
    "psr_ipc_event_create": (_i, [_pvp, _vp, _i]),
    "psr_ipc_event_open": (_i, [_vp, _i, _pvp]),
    "psr_event_record": (_i, [_vp, _vp]),
    "psr_event_wait": (_i, [_vp, _vp]),
    "psr_event_synchronize": (_i, [_vp]),
    "psr_event_destroy": (_i, [_vp]),
}

_lib = None


def load(path: str | None = None):
    """Load the shared library (once) and set every prototype.  Raises NativeLibraryError if absent."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or os.environ.get("PSANA_RAY_CUDA_LIB") or LIB_PATH
    if not os.path.exists(p):
        raise NativeLibraryError(
            f"{p} not found. Build it with `python psana-ray_b200/build.py` (needs nvcc; sm_100a). "
            "This package has no CPU fallback.")
    try:
        lib = C.CDLL(p)        # RTLD_LOCAL: several builds of the library can coexist in one process
    except OSError as e:
        raise NativeLibraryError(f"cannot load {p}: {e}") from e
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:
            raise NativeLibraryError(f"{p} does not export {name}") from e
        fn.restype = res
        fn.argtypes = args
    if path is None:
        _lib = lib
    return lib


def check(rc: int):
    if rc != PSR_OK:
        msg = load().psr_last_error()
        raise PsrError(rc, msg.decode("utf-8", "replace") if msg else "")
    return rc
