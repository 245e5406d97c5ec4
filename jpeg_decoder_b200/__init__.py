"""jpeg_decoder_b200 -- B200-native scan-to-pixels path for baseline JPEG.

The product is ``libb200jpeg.so`` (hand-written sm_100a CUDA kernels behind the C ABI of
``include/b200jpeg.h``).  This package is the thin Python harness over that ABI used by the tests
and ``bench.py``; host-side C++ users bind the header directly.

There is no CPU fallback: if the library is missing the import of :func:`load` fails loudly, and
every decode call fails when no sm_100-class GPU is present.
"""
import ctypes as C
import os

import numpy as np

from . import abi
from .abi import ImageDesc  # noqa: F401

_HERE = os.path.dirname(os.path.abspath(__file__))
# B200JPEG_LIB: another build of the SAME sources (kernel tuning experiments, scripts/gpu_variants.sh)
LIB_SO = os.environ.get("B200JPEG_LIB") or os.path.join(_HERE, "libb200jpeg.so")
_lib = None


class B200JpegError(RuntimeError):
    def __init__(self, code, what=""):
        super().__init__("b200jpeg error %d %s" % (code, what))
        self.code = code


def load():
    """dlopen the in-tree CUDA library; never falls back to anything else."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_SO):
            raise ImportError("%s is missing: build it with `python -m jpeg_decoder_b200.build` "
                              "(there is no CPU fallback)" % LIB_SO)
        _lib = abi.bind(C.CDLL(LIB_SO))
        if _lib.b200jpeg_abi_version() != abi.ABI_VERSION:
            raise ImportError("libb200jpeg.so ABI version mismatch")
    return _lib


def parse_header(data, flags=abi.PARSE_REFERENCE):
    """Host-side header parse (mirror of Image::readJPEG up to SOS, jpeg.cpp:826-867)."""
    buf = np.frombuffer(data, dtype=np.uint8)
    d = ImageDesc()
    rc = load().b200jpeg_parse_header(buf.ctypes.data, len(buf), flags, C.byref(d))
    if rc:
        raise B200JpegError(rc, "parse_header")
    return d


def parse_headers(bufs, flags=abi.PARSE_REFERENCE, threads=0, descs=None):
    """b200jpeg_parse_headers: n headers on several host threads -> ((ImageDesc*n), [rc])."""
    n = len(bufs)
    if descs is None:
        descs = (ImageDesc * n)()
    files = (C.c_void_p * n)(*[b.ctypes.data for b in bufs])
    sizes = (C.c_size_t * n)(*[b.size for b in bufs])
    rcs = (C.c_int32 * n)()
    load().b200jpeg_parse_headers(files, sizes, n, flags, threads, descs, rcs)
    return descs, list(rcs)


class MultiDecoder:
    """One context + one host thread per entry of `devices` (b200jpeg_multi_*): a batch is cut into contiguous image
    ranges, one per device; nothing moves between devices."""

    def __init__(self, devices):
        self.lib = load()
        devs = (C.c_int * len(devices))(*devices)
        h = C.c_void_p()
        rc = self.lib.b200jpeg_multi_create(devs, len(devices), C.byref(h))
        if rc:
            raise B200JpegError(rc, "b200jpeg_multi_create(%r)" % (list(devices),))
        self.h = h
        self.devices = list(devices)

    def close(self):
        if self.h:
            self.lib.b200jpeg_multi_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def decode_batch(self, descs, bufs, rgbs):
        n = len(bufs)
        files = (C.c_void_p * n)(*[b.ctypes.data for b in bufs])
        sizes = (C.c_size_t * n)(*[b.size for b in bufs])
        prgb = (C.c_void_p * n)(*[r.ctypes.data for r in rgbs])
        st = (C.c_int32 * n)()
        rc = self.lib.b200jpeg_multi_decode_batch(self.h, descs, n, files, sizes, prgb, st)
        if rc:
            raise B200JpegError(rc, "multi_decode_batch")
        return list(st)


class Decoder:
    """One context per GPU (b200jpeg_create)."""

    def __init__(self, device=0):
        self.lib = load()
        h = C.c_void_p()
        rc = self.lib.b200jpeg_create(device, C.byref(h))
        if rc:
            raise B200JpegError(rc, "b200jpeg_create(device=%d): no usable sm_100 GPU" % device)
        self.h = h
        self._keep = None

    def close(self):
        if self.h:
            self.lib.b200jpeg_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc, what):
        if rc:
            raise B200JpegError(rc, "%s: %s" % (what, self.lib.b200jpeg_last_error(self.h).decode()))

    # ---- staged API -------------------------------------------------------------------------
    def prepare(self, descs, bufs):
        """descs: (ImageDesc*n) array; bufs: list of uint8 numpy arrays (whole files)."""
        n = len(bufs)
        files = (C.c_void_p * n)(*[b.ctypes.data for b in bufs])
        sizes = (C.c_size_t * n)(*[b.size for b in bufs])
        self._keep = (descs, bufs, files, sizes)
        self._check(self.lib.b200jpeg_batch_prepare(self.h, descs, n, files, sizes), "prepare")
        return n

    def run(self):
        self._check(self.lib.b200jpeg_batch_run(self.h), "run")

    def sync(self):
        self._check(self.lib.b200jpeg_sync(self.h), "sync")

    def fetch(self, rgbs=None, coefs=None):
        n = len(self._keep[1])
        prgb = (C.c_void_p * n)(*[r.ctypes.data for r in rgbs]) if rgbs is not None else None
        pco = (C.c_void_p * n)(*[c.ctypes.data for c in coefs]) if coefs is not None else None
        st = (C.c_int32 * n)()
        self._check(self.lib.b200jpeg_batch_fetch(self.h, prgb, pco, st), "fetch")
        return list(st)

    def stats(self):
        s = abi.BatchStats()
        self.lib.b200jpeg_batch_get_stats(self.h, C.byref(s))
        return {k: getattr(s, k) for k, _ in s._fields_}

    def set_entropy_path(self, mode=abi.ENTROPY_AUTO):
        """abi.ENTROPY_AUTO / ENTROPY_SEGMENT / ENTROPY_SYNC (b200jpeg_set_entropy_path)."""
        self._check(self.lib.b200jpeg_set_entropy_path(self.h, int(mode)), "set_entropy_path")

    def set_timing(self, on=True):
        self.lib.b200jpeg_set_timing(self.h, int(on))

    def last_timing(self):
        ms = (C.c_float * abi.T_COUNT)()
        self.lib.b200jpeg_last_timing(self.h, ms, abi.T_COUNT)
        return dict(zip(("unstuff", "sync", "write", "pixel", "total"), list(ms)))

    def stream(self):
        return self.lib.b200jpeg_stream(self.h)

    def decode_batch(self, descs, bufs, rgbs):
        """b200jpeg_decode_batch: whole files in host memory in, RGB in host memory out, one call
        (chunked and double-buffered inside the library).  Returns the per-image status list."""
        n = len(bufs)
        files = (C.c_void_p * n)(*[b.ctypes.data for b in bufs])
        sizes = (C.c_size_t * n)(*[b.size for b in bufs])
        prgb = (C.c_void_p * n)(*[r.ctypes.data for r in rgbs])
        st = (C.c_int32 * n)()
        self._check(self.lib.b200jpeg_decode_batch(self.h, descs, n, files, sizes, prgb, None, st), "decode_batch")
        return list(st)

    def export_coefficients_noalloc(self, i, out):
        """b200jpeg_export_coefficients_noalloc: coefficients of image i of the current batch through the library's
        static device scratch (no allocation).  `out`: int16 numpy array [positions, 3, 64]."""
        rc = self.lib.b200jpeg_export_coefficients_noalloc(self.h, i, out.ctypes.data, out.nbytes)
        self._check(rc, "export_coefficients_noalloc")
        return out

    def decode_progressive(self, datas, want_coef=False):
        """Progressive (SOF2) files: b200jpeg_parse_progressive + b200jpeg_decode_progressive.
        Returns (status list, rgb arrays [H,W,3], coef arrays or None) like decode()."""
        n = len(datas)
        descs = (abi.ProgressiveDesc * n)()
        bufs = [np.frombuffer(d, dtype=np.uint8) for d in datas]
        for i, b in enumerate(bufs):
            rc = self.lib.b200jpeg_parse_progressive(b.ctypes.data, b.size, C.byref(descs[i]))
            if rc:
                raise B200JpegError(rc, "parse_progressive(image %d)" % i)
        fr = [descs[i].frame for i in range(n)]
        rgbs = [np.empty((f.image_height, f.image_width, 3), np.uint8) for f in fr]
        coefs = [np.empty((f.mcuWidthReal * f.mcuHeightReal, 3, 64), np.int16) for f in fr] if want_coef else None
        files = (C.c_void_p * n)(*[b.ctypes.data for b in bufs])
        sizes = (C.c_size_t * n)(*[b.size for b in bufs])
        prgb = (C.c_void_p * n)(*[r.ctypes.data for r in rgbs])
        pco = (C.c_void_p * n)(*[c.ctypes.data for c in coefs]) if want_coef else None
        st = (C.c_int32 * n)()
        self._check(self.lib.b200jpeg_decode_progressive(self.h, descs, n, files, sizes, prgb, pco, st), "decode_progressive")
        return list(st), rgbs, coefs

    def copy_floor(self, descs, bufs, rgbs):
        """b200jpeg_copy_floor: the copies of decode_batch without its kernels; returns milliseconds."""
        n = len(bufs)
        files = (C.c_void_p * n)(*[b.ctypes.data for b in bufs])
        sizes = (C.c_size_t * n)(*[b.size for b in bufs])
        prgb = (C.c_void_p * n)(*[r.ctypes.data for r in rgbs])
        ms = C.c_float(0)
        self._check(self.lib.b200jpeg_copy_floor(self.h, descs, n, files, sizes, prgb, C.byref(ms)), "copy_floor")
        return float(ms.value)

    # ---- one-call API -----------------------------------------------------------------------
    def decode(self, datas, flags=abi.PARSE_REFERENCE, want_coef=False):
        """datas: list of bytes.  Returns (status list, rgb arrays [H,W,3], coef arrays or None);
        coef is int16 [mcuWidthReal*mcuHeightReal, 3, 64] in the reference's MCU-array layout."""
        n = len(datas)
        descs = (ImageDesc * n)()
        bufs = [np.frombuffer(d, dtype=np.uint8) for d in datas]
        for i, b in enumerate(bufs):
            rc = self.lib.b200jpeg_parse_header(b.ctypes.data, b.size, flags, C.byref(descs[i]))
            if rc:
                raise B200JpegError(rc, "parse_header(image %d)" % i)
        rgbs = [np.empty((descs[i].image_height, descs[i].image_width, 3), np.uint8) for i in range(n)]
        coefs = ([np.empty((descs[i].mcuWidthReal * descs[i].mcuHeightReal, 3, 64), np.int16) for i in range(n)]
                 if want_coef else None)
        self.prepare(descs, bufs)
        self.run()
        st = self.fetch(rgbs, coefs)
        return st, rgbs, coefs
