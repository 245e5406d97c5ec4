"""CPU tier: pin the oracle.  The C restatement (oracle/jpeg_oracle.c) must reproduce the golden
digests that tests/golden/make_golden.py took from the UNTOUCHED reference binary, and, where that
binary is present (this container), match it byte for byte when both are run now."""
import hashlib

import numpy as np
import pytest

from tests.conftest import fixture_bytes

FIXTURES = ["img.jpg", "img2.jpg", "img3.jpg", "img4.jpg", "img5.jpg", "img6.jpg"]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


@pytest.mark.parametrize("name", FIXTURES)
def test_port_matches_reference_golden_on_fixtures(oracle_mod, golden, name):
    rc, info, coef, rgb = oracle_mod.port_decode(fixture_bytes(name))
    g = golden["fixtures"][name]
    assert rc == 0
    assert (info["width"], info["height"], info["bwr"], info["bhr"], info["hs"], info["vs"]) == (
        g["w"], g["h"], g["bwr"], g["bhr"], g["hs"], g["vs"])
    assert info["restart_interval"] == g["ri"] and info["scan_unstuffed"] == g["scan_unstuffed"]
    assert sha(coef) == g["coef_sha256"]      # MCU array after decodeHuffman(), bit-exact
    assert sha(rgb) == g["rgb_sha256"]        # RGB gather, bit-exact


def test_port_matches_reference_golden_on_synthetic(oracle_mod, golden, harness):
    for name, g in golden["synthetic"].items():
        data = harness.synth_jpeg(**g["args"])
        assert hashlib.sha256(data).hexdigest() == g["jpeg_sha256"], "bundled encoder drifted: " + name
        rc, info, coef, rgb = oracle_mod.port_decode(data)
        assert rc == 0 and sha(coef) == g["coef_sha256"] and sha(rgb) == g["rgb_sha256"], name


def test_port_matches_live_reference_binary(oracle_mod, harness):
    if not oracle_mod.have_ref():
        pytest.skip("oracle/_ref/ref_decode not built (reference sources absent)")
    datas = [fixture_bytes("img2.jpg"), harness.synth_jpeg(77, 160, 96, 80, 2, 1, False, 5),
             harness.synth_jpeg(78, 97, 55, 92, 1, 1, False, 0)]
    for d in datas:
        info, coef, rgb = oracle_mod.ref_dump(d)
        rc, pi, pc, pr = oracle_mod.port_decode(d)
        assert rc == 0 and np.array_equal(coef, pc) and np.array_equal(rgb, pr)
        assert info["sizeof_mcu"] == 768


def test_port_matches_live_reference_on_transcoded_huffman_tables(oracle_mod, harness):
    # the same coefficients under AC tables with 140 twelve-bit codes (tests/jpeg_craft.py): the reference's bit-by-bit
    # getNextSymbol (jpeg.cpp:300-320) and the port agree with each other and with the untranscoded file
    if not oracle_mod.have_ref():
        pytest.skip("oracle/_ref/ref_decode not built (reference sources absent)")
    from tests import jpeg_craft as jc
    for args in [(71, 96, 64, 98, 2, 2, False, 0), (72, 88, 56, 95, 1, 1, False, 5), (74, 160, 120, 99, 2, 1, False, 0)]:
        d = harness.synth_jpeg(*args)
        t = jc.with_ac_code_lengths(d, jc.OVERFLOW_AC_BITS)
        assert t != d
        info, coef, rgb = oracle_mod.ref_dump(t)
        rc, _, pc, pr = oracle_mod.port_decode(t)
        rc0, _, pc0, pr0 = oracle_mod.port_decode(d)
        assert rc == 0 and rc0 == 0
        assert np.array_equal(coef, pc) and np.array_equal(rgb, pr)
        assert np.array_equal(pc, pc0) and np.array_equal(pr, pr0)


def test_idct_constants_bit_patterns(oracle_mod):
    # m0,m1,m2,m3,m4,m5,s0..s7 as the reference's double->float initialisers give them
    want = [0x3fec835e, 0x3fb504f3, 0x3f8a8bd4, 0x3fb504f3, 0x40273d74, 0x3f43ef15,
            0x3eb504f3, 0x3efb14be, 0x3eec835e, 0x3ed4db31, 0x3eb504f3, 0x3e8e39da, 0x3e43ef15, 0x3dc7c5c2]
    assert [int(x) for x in oracle_mod.port_idct_constants()] == want


def test_unstuff_rules(oracle_mod):
    # file.hpp:59-104: FF00->FF, RSTn dropped, FF FF fill collapses, FF D9 ends
    scan = bytes([0x12, 0xff, 0x00, 0x34, 0xff, 0xd3, 0x56, 0xff, 0xff, 0x00, 0x78, 0xff, 0xd9, 0x99])
    n, out, used = oracle_mod.port_unstuff(scan)
    assert n == 6 and bytes(out) == bytes([0x12, 0xff, 0x34, 0x56, 0xff, 0x78]) and used == 13
    n, _, _ = oracle_mod.port_unstuff(bytes([0x01, 0xff, 0xc0, 0xff, 0xd9]))
    assert n == -11   # ORACLE_E_SCAN_MARKER, file.hpp:91-95
    n, _, _ = oracle_mod.port_unstuff(bytes([0x01, 0x02, 0x03]))
    assert n == -12   # no EOI


def test_reference_rejections(oracle_mod, harness):
    assert oracle_mod.port_decode(fixture_bytes("prograssive-sample-2.jpg"))[0] == -2   # jpeg.cpp:69-73
    gray = harness.synth_jpeg(5, 64, 48, 80, 1, 1, True, 0)
    assert oracle_mod.port_decode(gray)[0] == -3                                      # jpeg.cpp:83-87
    assert oracle_mod.port_decode(gray, flags=oracle_mod.F_SPEC)[0] == 0
    assert oracle_mod.port_decode(b"\x89PNG....")[0] == -1                             # jpeg.cpp:800-805
    if oracle_mod.have_ref():
        with pytest.raises(RuntimeError):
            oracle_mod.ref_dump(gray)


def test_gray_equals_444_with_zero_chroma(oracle_mod, harness):
    # grey cross-check -- a grey picture coded as 3 components has the same Y
    rgbsrc = harness.synth_rgb(9, 80, 56)
    y = (0.299 * rgbsrc[..., 0] + 0.587 * rgbsrc[..., 1] + 0.114 * rgbsrc[..., 2] + 0.5).astype(np.uint8)
    g3 = np.repeat(y[..., None], 3, axis=2)
    d1 = harness.encode_rgb(g3, 88, 1, 1, True, 0)
    d3 = harness.encode_rgb(g3, 88, 1, 1, False, 0)
    rc1, _, c1, r1 = oracle_mod.port_decode(d1, flags=oracle_mod.F_SPEC)
    rc3, _, c3, r3 = oracle_mod.port_decode(d3)
    assert rc1 == 0 and rc3 == 0
    assert np.array_equal(c1[:, 0], c3[:, 0])
    assert np.abs(r1.astype(int) - r3.astype(int)).max() <= 1


def test_port_matches_live_reference_on_quirk_inputs(oracle_mod, harness):
    """Two reference quirks no encoder exercises, pinned against the reference binary: a 16-bit DQT keeps only the low
    byte of every entry (jpeg.cpp:216), and the padding bits before an RSTn marker are never looked at (segments end by
    block count + align(), jpeg.cpp:419-425) -- 0-bits decode like 1-bits."""
    if not oracle_mod.have_ref():
        pytest.skip("oracle/_ref/ref_decode not built (reference sources absent)")
    from tests import jpeg_craft as jc
    d = harness.synth_jpeg(61, 200, 120, 85, 1, 1, False, 3)
    rc, _, c0, r0 = oracle_mod.port_decode(d)
    assert rc == 0

    def same(x):
        rc, _, c, r = oracle_mod.port_decode(x)
        return rc == 0 and np.array_equal(c, c0) and np.array_equal(r, r0)
    z, cleared = jc.with_zero_padding(d, same)
    assert cleared > 100
    _, c, r = oracle_mod.ref_dump(z)
    assert np.array_equal(c, c0) and np.array_equal(r, r0)
    w = jc.with_wide_dqt(d, lambda t, i, v: min(65535, v * 40) if i > 20 else v)
    _, c, r = oracle_mod.ref_dump(w)
    rc, _, pc, pr = oracle_mod.port_decode(w)
    assert rc == 0 and np.array_equal(c, pc) and np.array_equal(r, pr)
    rc, _, sc, sr = oracle_mod.port_decode(w, flags=oracle_mod.F_SPEC)       # T.81: the whole 16-bit value
    assert rc == 0 and np.array_equal(sc, pc) and not np.array_equal(sr, pr)


def test_progressive_port_is_pinned_through_baseline_twins(oracle_mod, golden_prog):
    """SOF2 (T.81 Annex G) restatement.  The reference refuses progressive files (jpeg.cpp:69-73), so the witness is
    indirect: Pillow wrote each fixture twice from the same pixels, progressive and baseline -- the same quantised
    coefficients -- and the digests in golden_prog.json come from the REFERENCE BINARY on the baseline twin."""
    import os
    from tests.conftest import GOLDEN_DIR
    for name, g in golden_prog.items():
        prog = open(os.path.join(GOLDEN_DIR, "prog", name + ".jpg"), "rb").read()
        twin = open(os.path.join(GOLDEN_DIR, "prog", name + "_twin.jpg"), "rb").read()
        assert hashlib.sha256(prog).hexdigest() == g["prog_sha256"] and hashlib.sha256(twin).hexdigest() == g["twin_sha256"]
        assert oracle_mod.port_decode(prog, flags=1)[0] == -2                      # the baseline entry point still refuses it
        rc, info, coef, rgb = oracle_mod.port_decode_progressive(prog)
        assert rc == 0 and info["scan_unstuffed"] == g["scans"]
        assert sha(coef) == g["coef_sha256"] and sha(rgb) == g["rgb_sha256"], name
        rc, _, tc, tr = oracle_mod.port_decode(twin, flags=1)
        assert rc == 0 and np.array_equal(tc, coef) and np.array_equal(tr, rgb), name
        if oracle_mod.have_ref() and g["witness"].startswith("reference"):
            _, rcf, rrgb = oracle_mod.ref_dump(twin)                                # the live reference binary, here
            assert sha(rcf) == g["coef_sha256"] and sha(rrgb) == g["rgb_sha256"], name


def test_progressive_port_decodes_the_references_own_fixtures(oracle_mod):
    # images/prograssive-sample*.jpg: the files the reference ships and refuses.  No pixel oracle exists for them; this
    # checks geometry, scan count and that the result is a picture (low-frequency energy in every component)
    for name, (w, h, scans) in {"prograssive-sample-2.jpg": (427, 640, 8), "prograssive-sample.jpg": (2832, 4256, 10)}.items():
        rc, info, coef, rgb = oracle_mod.port_decode_progressive(fixture_bytes(name))
        assert rc == 0 and (info["width"], info["height"], info["scan_unstuffed"]) == (w, h, scans)
        assert rgb.std() > 20 and np.abs(coef[:, 0, 0]).max() > 100
