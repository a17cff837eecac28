"""GpuCalibrator -- host-side handle on the CUDA calibration path (one per rank / GPU).

Stands where the reference has ``PsanaWrapperSmd`` + ``psana.Detector`` doing ``det.calib(evt)`` /
``det.image(evt)`` one event at a time on the CPU (ref: psana_ray/producer.py:88,150-159): here a
whole batch of raw uint16 frames is calibrated by the sm_100a kernels behind the C ABI in
``include/psana_ray_cuda.h``.  PyTorch only provides device memory and streams.
"""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _native as N
from .synthetic import DetectorSpec, detector_spec


def _np_ptr(a: np.ndarray | None):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def _as_np(a, dtype, shape=None, name="array"):
    if a is None:
        return None
    a = np.ascontiguousarray(a, dtype=dtype)
    if shape is not None and tuple(a.shape) != tuple(shape):
        raise ValueError(f"{name} has shape {a.shape}, expected {tuple(shape)}")
    return a


class GpuCalibrator:
    def __init__(self, detector, device: int = 0):
        self.spec: DetectorSpec = detector if isinstance(detector, DetectorSpec) else detector_spec(detector)
        self.device = int(device)
        self._lib = N.load()
        ctx = C.c_void_p()
        kind = N.PSR_DET_EPIX10KA if self.spec.kind == "epix10ka" else N.PSR_DET_JUNGFRAU
        N.check(self._lib.psr_create(C.byref(ctx), kind, self.spec.nseg, self.spec.rows, self.spec.cols, self.device))
        self._ctx = ctx
        self.image_shape = None
        self.cm_mode = 0
        self._reserved = 0

    # -- per-run set-up ---------------------------------------------------------------------------
    def set_constants(self, peds, gain=None, offs=None, cfg=None, mask=None):
        s = self.spec
        nm = s.nmodes
        peds = _as_np(peds, np.float32, (nm,) + s.shape, "peds")
        gain = _as_np(gain, np.float32, (nm,) + s.shape, "gain")
        offs = _as_np(offs, np.float32, (nm,) + s.shape, "offs")
        cfg = _as_np(cfg, np.uint8, s.shape, "cfg")
        mask = _as_np(mask, np.uint8, s.shape, "mask")
        N.check(self._lib.psr_set_constants(self._ctx, _np_ptr(peds), _np_ptr(gain), _np_ptr(offs),
                                            _np_ptr(cfg), _np_ptr(mask), nm))

    def set_common_mode(self, cmpars=None, eligible_modes_mask: int = 0):
        """cmpars = (alg, mode, cormax[, npix_min=10]) as psana's det.calib(evt, cmpars=...); None = off."""
        if cmpars is None:
            mode, cormax, npix_min = 0, 0.0, 10
        else:
            mode, cormax = int(cmpars[1]), float(cmpars[2])
            npix_min = int(cmpars[3]) if len(cmpars) > 3 else 10
        N.check(self._lib.psr_set_common_mode(self._ctx, mode, cormax, npix_min, eligible_modes_mask))
        self.cm_mode = mode

    def set_geometry(self, iX, iY):
        iX = _as_np(iX, np.int32, self.spec.shape, "iX")
        iY = _as_np(iY, np.int32, self.spec.shape, "iY")
        N.check(self._lib.psr_set_geometry(self._ctx, _np_ptr(iX), _np_ptr(iY)))
        h, w = C.c_int(), C.c_int()
        N.check(self._lib.psr_image_shape(self._ctx, C.byref(h), C.byref(w)))
        self.image_shape = (h.value, w.value)
        return self.image_shape

    def set_keep_mask(self, keep, image: bool = False):
        """Logical AND of the producer's bad-pixel and manual masks (ref: producer.py:92-95); None clears."""
        kind = N.PSR_OUT_IMAGE if image else N.PSR_OUT_CALIB
        if keep is None:
            N.check(self._lib.psr_set_keep_mask(self._ctx, None, 0, kind))
            return
        k = np.ascontiguousarray(np.asarray(keep) != 0, dtype=np.uint8)
        N.check(self._lib.psr_set_keep_mask(self._ctx, _np_ptr(k), k.size, kind))

    def reserve(self, max_frames: int):
        """Sizes the scratch that image output with a row / column common-mode word goes through.  ``calib()`` grows it
        on demand (an allocation on that call); reserve the largest batch up front to keep the hot call allocation-free."""
        N.check(self._lib.psr_reserve(self._ctx, int(max_frames)))
        self._reserved = max(self._reserved, int(max_frames))

    def set_option(self, name: str, value: int):
        N.check(self._lib.psr_set_option(self._ctx, name.encode(), int(value)))

    def set_host_chunk(self, chunk_frames: int, depth: int = 3):
        N.check(self._lib.psr_set_host_chunk(self._ctx, int(chunk_frames), int(depth)))

    # -- the hot call -----------------------------------------------------------------------------
    def out_shape(self, nframes: int, image: bool):
        if image:
            if self.image_shape is None:
                raise RuntimeError("image mode needs set_geometry()")
            return (nframes,) + self.image_shape
        return (nframes,) + self.spec.shape

    def calib(self, raw, out=None, image: bool = False, stream=None):
        """raw: CUDA tensor (B, nseg, rows, cols) uint16/int16 -> CUDA float32 tensor (calib or image)."""
        import torch
        if not raw.is_cuda or raw.device.index != self.device:
            raise ValueError(f"raw must live on cuda:{self.device}")
        if raw.dtype not in (torch.uint16, torch.int16) or not raw.is_contiguous():
            raise ValueError("raw must be a contiguous uint16/int16 tensor")
        if tuple(raw.shape[1:]) != self.spec.shape:
            raise ValueError(f"raw has frame shape {tuple(raw.shape[1:])}, expected {self.spec.shape}")
        nframes = raw.shape[0]
        shape = self.out_shape(nframes, image)
        if out is None:
            out = torch.empty(shape, dtype=torch.float32, device=raw.device)
        elif tuple(out.shape) != shape or out.dtype != torch.float32 or not out.is_contiguous() or out.device != raw.device:
            raise ValueError(f"out must be a contiguous float32 CUDA tensor of shape {shape}")
        st = stream if stream is not None else torch.cuda.current_stream(raw.device).cuda_stream
        if image and self.cm_mode not in (0, 4) and nframes > self._reserved:
            self.reserve(nframes)                   # the C ABI refuses to allocate inside the hot call
        N.check(self._lib.psr_calib_batch(self._ctx, raw.data_ptr(), nframes, out.data_ptr(),
                                          N.PSR_OUT_IMAGE if image else N.PSR_OUT_CALIB, st))
        return out

    def calib_image_noalloc(self, raw, out, stream=None):
        """Image-mode calibration that allocates nothing on the device (psr_calib_image_noalloc).

        For consumers that share a GPU whose memory is already spoken for, or loops that cannot afford
        the device-wide stall of an allocation.  ``out`` must be supplied by the caller; at most
        ``PSR_NOALLOC_SCRATCH_BYTES // (4 * npix)`` frames per call (3 for epix10k2M, 2 for Jungfrau 4M).
        Calls on one device share one scratch array: keep them on one stream.
        """
        import torch
        if raw.dtype not in (torch.uint16, torch.int16) or not raw.is_contiguous() or not raw.is_cuda or raw.device.index != self.device:
            raise ValueError(f"raw must be a contiguous uint16/int16 tensor on cuda:{self.device}")
        if tuple(raw.shape[1:]) != self.spec.shape:
            raise ValueError(f"raw has frame shape {tuple(raw.shape[1:])}, expected {self.spec.shape}")
        nframes = raw.shape[0]
        shape = self.out_shape(nframes, True)
        if (tuple(out.shape) != shape or out.dtype != torch.float32 or not out.is_contiguous()
                or not out.is_cuda or out.device != raw.device):
            raise ValueError(f"out must be a contiguous float32 tensor of shape {shape} on cuda:{self.device}")
        st = stream if stream is not None else torch.cuda.current_stream(raw.device).cuda_stream
        N.check(self._lib.psr_calib_image_noalloc(self._ctx, raw.data_ptr(), nframes, out.data_ptr(), st))
        return out

    def calib_ptr(self, d_raw: int, nframes: int, d_out: int, image: bool = False, stream: int = 0):
        """Raw-pointer form of calib() for callers that manage their own device memory."""
        N.check(self._lib.psr_calib_batch(self._ctx, d_raw, nframes, d_out,
                                          N.PSR_OUT_IMAGE if image else N.PSR_OUT_CALIB, stream))

    def decode(self, raw, stream=None):
        import torch
        nframes = raw.shape[0]
        out = torch.empty(tuple(raw.shape), dtype=torch.uint8, device=raw.device)
        st = stream if stream is not None else torch.cuda.current_stream(raw.device).cuda_stream
        N.check(self._lib.psr_decode_gain_modes(self._ctx, raw.data_ptr(), nframes, out.data_ptr(), st))
        return out

    def launches_per_batch(self, nframes: int, image: bool) -> int:
        """Kernel launches one ``calib()`` call of this configuration makes (for launch accounting)."""
        if not image:
            return 1
        if self.cm_mode in (0, 4):
            return 2                    # scatter / fused common-mode kernel + the gap-fill kernel beside it
        return 2                        # common mode into scratch + gather

    def calib_host(self, h_raw, h_out=None, d_out=None, image: bool = False, begin_only: bool = False, done_event=None):
        """Host-fed form: ``h_raw`` is a NumPy array or CPU tensor (pinned for full speed), ``h_out`` likewise,
        ``d_out`` a CUDA tensor on this calibrator's device.  Shapes, dtypes and contiguity are checked here: the
        C ABI below takes bare pointers."""
        def check_host(x, name, dtypes_np, dtypes_t, shape_tail):
            if isinstance(x, np.ndarray):
                ok = x.dtype in dtypes_np and x.flags["C_CONTIGUOUS"]
            else:
                ok = (not x.is_cuda) and x.dtype in dtypes_t and x.is_contiguous()
            if not ok:
                raise ValueError(f"{name} must be a C-contiguous host array of dtype {' / '.join(str(d) for d in dtypes_np)}")
            if tuple(x.shape[1:]) != tuple(shape_tail):
                raise ValueError(f"{name} has frame shape {tuple(x.shape[1:])}, expected {tuple(shape_tail)}")

        def ptr_n(x):
            if x is None:
                return None, 0
            if isinstance(x, np.ndarray):
                return x.ctypes.data, x.shape[0]
            return x.data_ptr(), x.shape[0]
        import torch
        if h_out is None and d_out is None:
            raise ValueError("need h_out and/or d_out")
        check_host(h_raw, "h_raw", (np.dtype(np.uint16), np.dtype(np.int16)), (torch.uint16, torch.int16), self.spec.shape)
        nframes = h_raw.shape[0]
        oshape = self.out_shape(nframes, image)
        if h_out is not None:
            check_host(h_out, "h_out", (np.dtype(np.float32),), (torch.float32,), oshape[1:])
            if h_out.shape[0] != nframes:
                raise ValueError(f"h_out holds {h_out.shape[0]} frames, h_raw {nframes}")
        if d_out is not None:
            if (not d_out.is_cuda or d_out.device.index != self.device or d_out.dtype != torch.float32
                    or not d_out.is_contiguous() or tuple(d_out.shape) != oshape):
                raise ValueError(f"d_out must be a contiguous float32 tensor of shape {oshape} on cuda:{self.device}")
        pr, n = ptr_n(h_raw)
        po, _ = ptr_n(h_out)
        pd = None if d_out is None else d_out.data_ptr()
        kind = N.PSR_OUT_IMAGE if image else N.PSR_OUT_CALIB
        if begin_only:
            N.check(self._lib.psr_calib_batch_host_begin(self._ctx, pr, n, po, pd, kind, done_event))
        else:
            N.check(self._lib.psr_calib_batch_host(self._ctx, pr, n, po, pd, kind))

    def calib_host_begin(self, h_raw, h_out=None, d_out=None, image: bool = False, done_event=None):
        """Asynchronous ``calib_host``: queues the batch on the library's streams and returns.  ``done_event`` (a raw
        ``cudaEvent_t`` value, e.g. a slab's inter-process event) is recorded where the batch is final.  The host buffer
        must stay untouched until ``calib_host_wait(input_only=True)``."""
        self.calib_host(h_raw, h_out=h_out, d_out=d_out, image=image, begin_only=True, done_event=done_event)

    def calib_host_wait(self, input_only: bool = False):
        N.check(self._lib.psr_calib_batch_host_wait(self._ctx, N.PSR_WAIT_INPUT if input_only else N.PSR_WAIT_ALL))

    def close(self):
        if getattr(self, "_ctx", None):
            self._lib.psr_destroy(self._ctx)
            self._ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()
