"""ctypes front end of harness/encoder.c: deterministic synthetic pictures -> baseline JPEG bytes.

Used by tests and bench.py to make the inputs BASELINE.json's configs name (C2..C5).  Nothing on
the decode path calls it.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
HARNESS_SO = os.path.join(_HERE, "libb200jpeg_harness.so")
_SRC = os.path.join(_HERE, "encoder.c")
_lib = None


def build_harness(force=False):
    if force or not os.path.exists(HARNESS_SO) or os.path.getmtime(HARNESS_SO) < os.path.getmtime(_SRC):
        # fixed flags, no -march=native / FMA contraction: the streams must be the same on every box
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-o", HARNESS_SO, _SRC, "-lm"])
    return HARNESS_SO


def _load():
    global _lib
    if _lib is None:
        build_harness()
        lib = C.CDLL(HARNESS_SO)
        lib.b200enc_synth_jpeg.restype = C.c_long
        lib.b200enc_synth_jpeg.argtypes = [C.c_uint64] + [C.c_int] * 7 + [C.c_void_p, C.c_long]
        lib.b200enc_encode_rgb.restype = C.c_long
        lib.b200enc_encode_rgb.argtypes = [C.c_void_p] + [C.c_int] * 7 + [C.c_void_p, C.c_long]
        lib.b200enc_synth_rgb.restype = None
        lib.b200enc_synth_rgb.argtypes = [C.c_uint64, C.c_int, C.c_int, C.c_void_p]
        _lib = lib
    return _lib


def synth_rgb(seed, w, h):
    out = np.empty((h, w, 3), np.uint8)
    _load().b200enc_synth_rgb(seed, w, h, out.ctypes.data)
    return out


def encode_rgb(rgb, quality=90, hs=2, vs=2, gray=False, restart=0):
    """restart: 0 none, >0 DRI value in MCUs, -1 one MCU row."""
    rgb = np.ascontiguousarray(rgb, np.uint8)
    h, w = rgb.shape[:2]
    buf = np.empty(w * h * 8 + 65536, np.uint8)      # white noise at quality 100 takes more bytes than the pixels it encodes
    n = _load().b200enc_encode_rgb(rgb.ctypes.data, w, h, quality, hs, vs, int(gray), restart, buf.ctypes.data, len(buf))
    if n < 0:
        raise RuntimeError("encoder output buffer too small")
    return buf[:n].tobytes()


def synth_jpeg(seed, w, h, quality=90, hs=2, vs=2, gray=False, restart=0, out=None):
    """One synthetic JPEG.  With `out` (a uint8 numpy buffer) the bytes are written there and the
    length is returned (no copy, GIL released during the call)."""
    buf = out if out is not None else np.empty(w * h * 3 + 65536, np.uint8)
    n = _load().b200enc_synth_jpeg(seed, w, h, quality, hs, vs, int(gray), restart, buf.ctypes.data, len(buf))
    if n < 0:
        raise RuntimeError("encoder output buffer too small")
    return n if out is not None else buf[:n].tobytes()
