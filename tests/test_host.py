"""CPU tier: the host surface -- header parser vs the oracle, ABI exports, rejection codes,
batch planner arithmetic.  No GPU compute calls."""
import ctypes as C
import re
import os

import numpy as np
import pytest

from tests.conftest import ROOT, fixture_bytes
from jpeg_decoder_b200 import abi

FIXTURES = ["img.jpg", "img2.jpg", "img3.jpg", "img4.jpg", "img5.jpg", "img6.jpg"]


@pytest.fixture(scope="module")
def lib(lib_so):
    import jpeg_decoder_b200 as j
    return j.load()


def test_library_exports_every_declared_symbol(lib):
    hdr = open(os.path.join(ROOT, "include", "b200jpeg.h")).read()
    declared = set(re.findall(r"\b(b200jpeg_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(abi.EXPORTS)
    for s in declared:
        assert hasattr(lib, s), s
    assert lib.b200jpeg_abi_version() == abi.ABI_VERSION


def test_struct_sizes_match_header(lib_so):
    # compile a tiny C probe against the header and compare sizeof with the ctypes mirror
    import subprocess, tempfile
    src = '#include "b200jpeg.h"\n#include <stdio.h>\nint main(){printf("%zu %zu %zu %zu %zu %zu %zu\\n",sizeof(b200jpeg_huffman_table),sizeof(b200jpeg_quant_table),sizeof(b200jpeg_component),sizeof(b200jpeg_image_desc),sizeof(b200jpeg_batch_stats),sizeof(b200jpeg_scan),sizeof(b200jpeg_progressive_desc));return 0;}'
    with tempfile.TemporaryDirectory() as td:
        open(os.path.join(td, "p.c"), "w").write(src)
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", os.path.join(td, "p"), os.path.join(td, "p.c")])
        got = [int(x) for x in subprocess.check_output([os.path.join(td, "p")]).split()]
    want = [C.sizeof(abi.HuffmanTable), C.sizeof(abi.QuantTable), C.sizeof(abi.Component), C.sizeof(abi.ImageDesc),
            C.sizeof(abi.BatchStats), C.sizeof(abi.Scan), C.sizeof(abi.ProgressiveDesc)]
    assert got == want


@pytest.mark.parametrize("name", FIXTURES)
def test_parse_header_matches_oracle(lib, oracle_mod, name):
    import jpeg_decoder_b200 as j
    data = fixture_bytes(name)
    d = j.parse_header(data)
    rc, info, _, _ = oracle_mod.port_decode(data, want_coef=False, want_rgb=False)
    assert rc == 0
    assert (d.image_width, d.image_height, d.mcuWidth, d.mcuHeight, d.mcuWidthReal, d.mcuHeightReal) == (
        info["width"], info["height"], info["bw"], info["bh"], info["bwr"], info["bhr"])
    assert (d.horizontalSamplingFactor, d.verticalSamplingFactor, d.restartInterval) == (
        info["hs"], info["vs"], info["restart_interval"])
    assert d.scan_offset == info["scan_offset"] and d.scan_offset + d.scan_length == len(data)
    # canonical codes as HuffmanTable::generateCodes (huffman.hpp:17-29): strictly increasing per length
    t = d.huffmanACTable[d.color_components[0].huffmanACTableID]
    assert t.set == 1
    for l in range(16):
        codes = [t.codes[k] for k in range(t.offset[l], t.offset[l + 1])]
        assert codes == list(range(codes[0], codes[0] + len(codes))) if codes else True
        assert all(c < (1 << (l + 1)) for c in codes)


def test_parse_header_rejections(lib, harness):
    import jpeg_decoder_b200 as j
    def code(data, flags=0):
        try:
            j.parse_header(data, flags)
            return 0
        except j.B200JpegError as e:
            return e.code
    assert code(fixture_bytes("prograssive-sample-2.jpg")) == abi.E_PROGRESSIVE   # jpeg.cpp:69-73
    assert code(b"\x00\x01\x02\x03") == abi.E_NOT_JPEG                             # jpeg.cpp:800-805
    gray = harness.synth_jpeg(5, 64, 48, 80, 1, 1, True, 0)
    assert code(gray) == abi.E_COMPONENTS                                          # jpeg.cpp:83-87
    assert code(gray, abi.PARSE_SPEC) == 0
    good = bytearray(harness.synth_jpeg(6, 64, 48, 80, 2, 2, False, 0))
    sof = good.find(b"\xff\xc0")
    bad = bytearray(good); bad[sof + 11] = 0x31                                    # luma 3x1
    assert code(bytes(bad)) == abi.E_SAMPLING                                      # jpeg.cpp:110-117
    bad = bytearray(good); bad[sof + 14] = 0x21                                    # chroma 2x1
    assert code(bytes(bad)) == abi.E_SAMPLING                                      # jpeg.cpp:129-136
    sos = good.find(b"\xff\xda")
    bad = bytearray(good); bad[sos + 12] = 62                                      # Se != 63
    assert code(bytes(bad)) == abi.E_SOS                                           # jpeg.cpp:255-259
    assert code(bytes(good[:sos])) in (abi.E_NO_SOS, abi.E_TRUNCATED)
    dht = good.find(b"\xff\xc4")
    bad = bytearray(good); bad[dht + 4] = 0x05                                     # table id 5
    assert code(bytes(bad)) == abi.E_TABLE_ID                                      # jpeg.cpp:156-160
    com = bytes(good[:2]) + b"\xff\xfe\x00\x04hi" + bytes(good[2:])
    assert code(com) == abi.E_NO_SOS        # COM is INVALID to the reference (file.hpp:252-253)
    assert code(com, abi.PARSE_SPEC) == 0


def test_decoder_fails_loudly_without_gpu(lib):
    import torch
    import jpeg_decoder_b200 as j
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(j.B200JpegError) as e:
        j.Decoder(0)
    assert e.value.code == abi.E_CUDA     # no CPU fallback exists


def test_spec_parser_component_ids_and_wide_dqt(lib, harness):
    """under B200JPEG_PARSE_SPEC component identifiers may be any byte (tables bound by the scan's
    selectors, the frame's first component is luma) and a 16-bit DQT keeps its full values; REFERENCE mode keeps the
    reference's rules (identifier <= 3, jpeg.cpp:95-104; low byte, jpeg.cpp:216)."""
    import jpeg_decoder_b200 as j
    from tests import jpeg_craft as jc
    d = harness.synth_jpeg(61, 200, 120, 85, 2, 1, False, 0)
    ids = jc.with_component_ids(d, (7, 9, 200))
    with pytest.raises(j.B200JpegError) as e:
        j.parse_header(ids)
    assert e.value.code == abi.E_COMPONENT_ID
    a, b = j.parse_header(d, abi.PARSE_SPEC), j.parse_header(ids, abi.PARSE_SPEC)
    assert (b.horizontalSamplingFactor, b.verticalSamplingFactor) == (2, 1)
    assert bytes(a)[:C.sizeof(abi.ImageDesc) - 16] == bytes(b)[:C.sizeof(abi.ImageDesc) - 16]   # same state at the seam
    sos = [s for s in jc.segments(ids) if s[0] == 0xDA][0]
    bad = bytearray(ids); bad[sos[1] + 3] = 8                                    # a selector the frame does not hold
    with pytest.raises(j.B200JpegError) as e:
        j.parse_header(bytes(bad), abi.PARSE_SPEC)
    assert e.value.code == abi.E_SOS
    w = jc.with_wide_dqt(d, lambda t, i, v: min(65535, v * 40) if i > 20 else v)
    lo, hi = j.parse_header(w), j.parse_header(w, abi.PARSE_SPEC)
    tq = lo.color_components[0].quantizationTableID
    t_lo, t_hi = list(lo.quantization_tables[tq].table), list(hi.quantization_tables[tq].table)
    assert max(t_hi) > 255 and t_lo == [v & 255 for v in t_hi]


def test_parse_headers_batch_equals_single(lib, harness):
    import jpeg_decoder_b200 as j
    datas = [harness.synth_jpeg(900 + i, 64 + 8 * (i % 7), 48, 80, 2, 2, False, i % 3) for i in range(150)]
    datas[17] = b"\x00\x01garbage"
    datas[99] = fixture_bytes("prograssive-sample-2.jpg")
    bufs = [np.frombuffer(d, dtype=np.uint8) for d in datas]
    for threads in (0, 1, 5):
        descs, rcs = j.parse_headers(bufs, threads=threads)
        assert rcs[17] == abi.E_NOT_JPEG and rcs[99] == abi.E_PROGRESSIVE
        for i, d in enumerate(datas):
            if i in (17, 99):
                continue
            assert rcs[i] == 0 and bytes(descs[i]) == bytes(j.parse_header(d)), i


def test_plan_rejects_malformed_hand_built_tables(lib, harness):
    """Descriptors may be hand-built through the C ABI: a Huffman table whose cumulative counts are not monotone, or a
    table id outside 0..3, must be refused before anything reaches the device (no GPU needed to see the refusal order:
    without a device the call fails with E_CUDA at create, so this checks the planner through the emulation build)."""
    from tests.emul import emul_harness
    emul_harness.build()
    L = emul_harness._load()
    d = np.frombuffer(harness.synth_jpeg(62, 64, 48, 80, 2, 2, False, 0), dtype=np.uint8).copy()
    for kind in ("offsets", "table_id", "quant"):
        desc = (abi.ImageDesc * 1)()
        assert L.b200jpeg_parse_header(d.ctypes.data, d.size, 0, C.byref(desc[0])) == 0
        if kind == "offsets":
            desc[0].huffmanACTable[0].offset[5] = 200
        elif kind == "table_id":
            desc[0].color_components[1].huffmanDCTTableID = 9
        else:
            desc[0].quantization_tables[0].table[3] = 70000
        files = (C.c_void_p * 1)(d.ctypes.data)
        sizes = (C.c_size_t * 1)(d.size)
        rgb = np.zeros((48, 64, 3), np.uint8)
        prgb = (C.c_void_p * 1)(rgb.ctypes.data)
        st = (C.c_int32 * 1)()
        info = (C.c_int32 * 3)()
        rc = L.emu_decode_batch(desc, 1, files, sizes, prgb, None, st, info)
        assert rc == {"offsets": abi.E_DHT, "table_id": abi.E_TABLE_ID, "quant": abi.E_ARG}[kind], (kind, rc)


def test_parse_progressive(lib, golden_prog, harness):
    """b200jpeg_parse_progressive: frame + scan list + table pool of SOF2 files; the baseline parser keeps the reference's
    refusal (jpeg.cpp:69-73)."""
    import jpeg_decoder_b200 as j
    from tests.conftest import GOLDEN_DIR
    for name, g in golden_prog.items():
        data = open(os.path.join(GOLDEN_DIR, "prog", name + ".jpg"), "rb").read()
        with pytest.raises(j.B200JpegError) as e:
            j.parse_header(data, abi.PARSE_SPEC)
        assert e.value.code == abi.E_PROGRESSIVE
        buf = np.frombuffer(data, np.uint8)
        d = abi.ProgressiveDesc()
        assert lib.b200jpeg_parse_progressive(buf.ctypes.data, buf.size, C.byref(d)) == 0
        assert (d.frame.image_width, d.frame.image_height, d.nscans) == (g["w"], g["h"], g["scans"])
        assert d.ntables >= 2 and d.frame.scan_offset == d.scans[0].offset
        end = 0
        for k in range(d.nscans):
            s = d.scans[k]
            assert s.offset >= end and s.length > 0 and data[s.offset - 2 * s.ns - 7] == 0xDA            # right behind its SOS header
            assert data[s.offset + s.length] == 0xFF and data[s.offset + s.length + 1] not in (0x00, 0xFF) + tuple(range(0xD0, 0xD8))
            assert (s.ss == 0) == (s.se == 0) and (s.ns == 1 or s.ss == 0)
            end = s.offset + s.length
        if "rst" in name:
            assert d.scans[0].restart_interval > 0
    base = np.frombuffer(harness.synth_jpeg(6, 64, 48, 80, 2, 2, False, 0), np.uint8)
    assert lib.b200jpeg_parse_progressive(base.ctypes.data, base.size, C.byref(abi.ProgressiveDesc())) == abi.E_ARG
    junk = np.frombuffer(b"\x00\x01\x02", np.uint8)
    assert lib.b200jpeg_parse_progressive(junk.ctypes.data, junk.size, C.byref(abi.ProgressiveDesc())) == abi.E_NOT_JPEG
