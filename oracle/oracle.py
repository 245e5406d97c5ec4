"""oracle/oracle.py -- TEST INFRASTRUCTURE: Python access to the parity oracle.

Two checkers live behind this module:

* ``port_decode`` -- the plain-C restatement in ``oracle/jpeg_oracle.c`` (``_ref/liboracle_port.so``),
  which travels to the GPU box and is what the ``-m gpu`` tests, ``smoke()`` and the
  ``cpu_baseline`` leg of ``bench.py`` call;
* ``ref_dump`` / ``ref_bench`` -- the UNTOUCHED reference compiled headless
  (``_ref/ref_decode``, built from /root/reference by ``oracle/Makefile``), used to pin the port.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py`` (cpu_baseline / --impl reference)
may import this module.  Nothing under ``jpeg_decoder_b200/`` does.
"""
import ctypes
import json
import os
import subprocess
import tempfile

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_DIR = os.path.join(_HERE, "_ref")
PORT_SO = os.path.join(REF_DIR, "liboracle_port.so")
REF_BIN = os.path.join(REF_DIR, "ref_decode")

F_SPEC = 1  # ORACLE_F_SPEC


class OracleInfo(ctypes.Structure):
    _fields_ = [(n, ctypes.c_int) for n in (
        "width", "height", "bw", "bh", "bwr", "bhr", "hs", "vs", "ncomp", "restart_interval",
        "scan_offset", "scan_unstuffed")] + [(n, ctypes.c_double) for n in (
            "ms_parse_unstuff", "ms_huffman", "ms_dequant", "ms_idct", "ms_colour")]

    def as_dict(self):
        return {n: getattr(self, n) for n, _ in self._fields_}


def build(force=False):
    """Compile the port (and, where /root/reference exists, the reference binary)."""
    if force or not os.path.exists(PORT_SO) or (
            os.path.getmtime(PORT_SO) < os.path.getmtime(os.path.join(_HERE, "jpeg_oracle.c"))):
        subprocess.check_call(["make", "-s", "-C", _HERE, "port"])
    if os.path.exists("/root/reference/jpeg.cpp") and (force or not os.path.exists(REF_BIN) or (
            os.path.getmtime(REF_BIN) < os.path.getmtime(os.path.join(_HERE, "ref_wrapper_post.h")))):
        subprocess.check_call(["make", "-s", "-C", _HERE, "ref"])


_lib = None


def _port():
    global _lib
    if _lib is None:
        if not os.path.exists(PORT_SO):
            build()
        lib = ctypes.CDLL(PORT_SO)
        lib.oracle_decode.restype = ctypes.c_int
        lib.oracle_decode.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_int,
                                      ctypes.POINTER(OracleInfo),
                                      ctypes.POINTER(ctypes.POINTER(ctypes.c_int32)),
                                      ctypes.POINTER(ctypes.POINTER(ctypes.c_uint8))]
        lib.oracle_decode_progressive.restype = ctypes.c_int
        lib.oracle_decode_progressive.argtypes = lib.oracle_decode.argtypes
        lib.oracle_free.argtypes = [ctypes.c_void_p]
        lib.oracle_idct_block.argtypes = [ctypes.c_void_p]
        lib.oracle_dequant_block.argtypes = [ctypes.c_void_p, ctypes.c_void_p]
        lib.oracle_idct_constants.argtypes = [ctypes.c_void_p]
        lib.oracle_unstuff.restype = ctypes.c_long
        lib.oracle_unstuff.argtypes = [ctypes.c_void_p, ctypes.c_size_t, ctypes.c_void_p,
                                       ctypes.POINTER(ctypes.c_size_t)]
        _lib = lib
    return _lib


def have_ref():
    return os.path.exists(REF_BIN)


def port_decode_progressive(data, want_coef=True, want_rgb=True):
    """Progressive (SOF2) files through the Annex-G restatement (port only: the reference rejects them)."""
    return port_decode(data, 0, want_coef, want_rgb, _fn="oracle_decode_progressive")


def port_decode(data, flags=0, want_coef=True, want_rgb=True, _fn="oracle_decode"):
    """Decode ``data`` (bytes) with the C restatement.

    Returns ``(rc, info_dict, coef, rgb)``; ``coef`` is int32 ``[bwr*bhr, 3, 64]`` in the reference's
    MCU-array layout (natural order), ``rgb`` is uint8 ``[H, W, 3]``; both ``None`` on error.
    """
    lib = _port()
    buf = (ctypes.c_uint8 * len(data)).from_buffer_copy(data)
    info = OracleInfo()
    pc = ctypes.POINTER(ctypes.c_int32)()
    pr = ctypes.POINTER(ctypes.c_uint8)()
    rc = getattr(lib, _fn)(buf, len(data), flags, ctypes.byref(info),
                           ctypes.byref(pc) if want_coef else None,
                           ctypes.byref(pr) if want_rgb else None)
    coef = rgb = None
    if rc == 0:
        npos = info.bwr * info.bhr
        if want_coef:
            coef = np.ctypeslib.as_array(pc, shape=(npos, 3, 64)).copy()
        if want_rgb:
            rgb = np.ctypeslib.as_array(pr, shape=(info.height, info.width, 3)).copy()
    if pc:
        lib.oracle_free(pc)
    if pr:
        lib.oracle_free(pr)
    return rc, info.as_dict(), coef, rgb


def port_idct_block(block):
    b = np.ascontiguousarray(block, dtype=np.int32).copy()
    _port().oracle_idct_block(b.ctypes.data)
    return b


def port_idct_constants():
    out = np.zeros(14, dtype=np.uint32)
    _port().oracle_idct_constants(out.ctypes.data)
    return out


def port_unstuff(scan):
    src = np.frombuffer(scan, dtype=np.uint8)
    out = np.zeros(len(src) + 1, dtype=np.uint8)
    used = ctypes.c_size_t(0)
    n = _port().oracle_unstuff(src.ctypes.data, len(src), out.ctypes.data, ctypes.byref(used))
    return n, (out[:n].copy() if n >= 0 else None), used.value


def ref_dump(path_or_bytes):
    """Run the untouched reference; returns ``(info_dict, coef[npos,3,64] int32, rgb[H,W,3] uint8)``.

    Raises ``RuntimeError`` when the reference exits non-zero (it calls ``exit(1)`` on every
    rejected input, e.g. jpeg.cpp:69-73)."""
    with tempfile.TemporaryDirectory() as td:
        if isinstance(path_or_bytes, (bytes, bytearray, memoryview)):
            path = os.path.join(td, "in.jpg")
            with open(path, "wb") as f:
                f.write(path_or_bytes)
        else:
            path = path_or_bytes
        cpath, rpath = os.path.join(td, "o.coef"), os.path.join(td, "o.rgb")
        p = subprocess.run([REF_BIN, "dump", path, cpath, rpath], capture_output=True, text=True,
                           timeout=600)
        if p.returncode != 0:
            raise RuntimeError("reference rejected input (rc=%d): %s" % (p.returncode, p.stderr[-200:]))
        info = json.loads(p.stdout.strip().splitlines()[-1])
        coef = np.fromfile(cpath, dtype=np.int32).reshape(info["bwr"] * info["bhr"], 3, 64)
        rgb = np.fromfile(rpath, dtype=np.uint8).reshape(info["h"], info["w"], 3)
    return info, coef, rgb


def ref_bench(paths, reps=1, timeout=3600):
    """Time ctor + readJPEG() of the untouched reference over ``paths``; one process."""
    p = subprocess.run([REF_BIN, "bench", str(reps)] + list(paths), capture_output=True, text=True,
                       timeout=timeout)
    if p.returncode != 0:
        raise RuntimeError("reference bench failed rc=%d: %s" % (p.returncode, p.stderr[-300:]))
    return json.loads(p.stdout.strip().splitlines()[-1])
