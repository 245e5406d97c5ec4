"""TEST INFRASTRUCTURE: Python front end of the emulation build (tests/emul/emul_lib.cpp)."""
import ctypes as C
import os
import subprocess

import numpy as np

from jpeg_decoder_b200 import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
_ROOT = os.path.dirname(os.path.dirname(_HERE))
SO = os.path.join(_HERE, "_build", "libb200jpeg_emul.so")
SO_STRESS = os.path.join(_HERE, "_build", "libb200jpeg_emul_stress.so")
_lib = None
_libs = {}


def build(force=False):
    csrc = os.path.join(_ROOT, "jpeg_decoder_b200", "csrc")
    deps = [os.path.join(csrc, f) for f in os.listdir(csrc)] + [os.path.join(_HERE, "cuda_emul.h"),
                                                                os.path.join(_HERE, "emul_lib.cpp")]
    if force or not os.path.exists(SO) or any(os.path.getmtime(d) > os.path.getmtime(SO) for d in deps):
        subprocess.check_call(["sh", os.path.join(_HERE, "build.sh")])
    return SO


def _load(stress=False):
    if stress not in _libs:
        build()
        lib = C.CDLL(SO_STRESS if stress else SO)
        abi.bind_parse(lib)
        lib.emu_decode_batch.restype = C.c_int
        _libs[stress] = lib
    return _libs[stress]


def decode(datas, flags=0, want_coef=True, stress=False, path=0):
    """Same contract as jpeg_decoder_b200.Decoder.decode, run on the emulation.
    stress=True: the build with 512-bit subs and no run-in (exercises the repair/fix paths)."""
    L = _load(stress)
    n = len(datas)
    descs = (abi.ImageDesc * n)()
    bufs = [np.frombuffer(d, dtype=np.uint8).copy() for d in datas]
    for i, b in enumerate(bufs):
        rc = L.b200jpeg_parse_header(b.ctypes.data, len(b), flags, C.byref(descs[i]))
        if rc:
            raise RuntimeError("parse_header rc=%d" % rc)
    files = (C.c_void_p * n)(*[b.ctypes.data for b in bufs])
    sizes = (C.c_size_t * n)(*[len(b) for b in bufs])
    rgbs = [np.zeros((descs[i].image_height, descs[i].image_width, 3), np.uint8) for i in range(n)]
    coefs = [np.zeros((descs[i].mcuWidthReal * descs[i].mcuHeightReal, 3, 64), np.int16) for i in range(n)]
    prgb = (C.c_void_p * n)(*[r.ctypes.data for r in rgbs])
    pco = (C.c_void_p * n)(*[c.ctypes.data for c in coefs]) if want_coef else None
    st = (C.c_int32 * n)()
    info = (C.c_int32 * 3)()
    info[2] = path
    rc = L.emu_decode_batch(descs, n, files, sizes, prgb, pco, st, info)
    if rc:
        raise RuntimeError("emu_decode_batch rc=%d" % rc)
    return list(st), rgbs, (coefs if want_coef else None), {"launches": info[0], "sync_rounds": info[1]}


def decode_progressive(datas, want_coef=True):
    """Progressive (SOF2) files through the emulation: same contract as Decoder.decode_progressive."""
    L = _load()
    L.emu_decode_progressive.restype = C.c_int
    n = len(datas)
    descs = (abi.ProgressiveDesc * n)()
    bufs = [np.frombuffer(d, dtype=np.uint8).copy() for d in datas]
    for i, b in enumerate(bufs):
        rc = L.b200jpeg_parse_progressive(b.ctypes.data, len(b), C.byref(descs[i]))
        if rc:
            raise RuntimeError("parse_progressive rc=%d" % rc)
    fr = [descs[i].frame for i in range(n)]
    files = (C.c_void_p * n)(*[b.ctypes.data for b in bufs])
    sizes = (C.c_size_t * n)(*[len(b) for b in bufs])
    rgbs = [np.zeros((f.image_height, f.image_width, 3), np.uint8) for f in fr]
    coefs = [np.zeros((f.mcuWidthReal * f.mcuHeightReal, 3, 64), np.int16) for f in fr]
    prgb = (C.c_void_p * n)(*[r.ctypes.data for r in rgbs])
    pco = (C.c_void_p * n)(*[c.ctypes.data for c in coefs]) if want_coef else None
    st = (C.c_int32 * n)()
    rc = L.emu_decode_progressive(descs, n, files, sizes, prgb, pco, st)
    if rc:
        raise RuntimeError("emu_decode_progressive rc=%d" % rc)
    return list(st), rgbs, (coefs if want_coef else None)
