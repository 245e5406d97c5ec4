"""GPU tier (-m gpu): parity of the CUDA path with the oracle, through the C ABI.

Bars (BASELINE.json north_star): quantised coefficients BIT-EXACT with the reference's MCU decode;
RGB within +-1 LSB per channel of the reference's float IDCT + colour (the kernels replicate the
operation order, so the expected number of differing samples is 0 -- asserted where noted)."""
import hashlib

import numpy as np
import pytest

from tests.conftest import fixture_bytes

pytestmark = pytest.mark.gpu
FIXTURES = ["img.jpg", "img2.jpg", "img3.jpg", "img4.jpg", "img5.jpg", "img6.jpg"]
RGB_TOL = 1   # LSB per channel, north_star


@pytest.fixture(scope="module", params=["sync", "segment"])
def dec(lib_so, request):
    """Every parity test runs through BOTH entropy decoders: the self-synchronising path
    (kernels_entropy.cuh) and the one-thread-per-restart-segment path (kernels_segdec.cuh)."""
    import jpeg_decoder_b200 as jd
    from jpeg_decoder_b200 import abi
    d = jd.Decoder(0)
    d.set_entropy_path(abi.ENTROPY_SYNC if request.param == "sync" else abi.ENTROPY_SEGMENT)
    yield d
    d.close()


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def compare(dec, oracle_mod, datas, flags=0, exact_rgb=True):
    st, rgbs, coefs = dec.decode(datas, flags=flags, want_coef=True)
    worst, ndiff = 0, 0
    for i, d in enumerate(datas):
        rc, _, oc, orgb = oracle_mod.port_decode(d, flags=flags)
        assert rc == 0 and st[i] == 0, (i, rc, st[i])
        assert np.array_equal(coefs[i].astype(np.int32), oc), "coefficients differ (image %d)" % i
        err = np.abs(rgbs[i].astype(np.int16) - orgb.astype(np.int16))
        worst = max(worst, int(err.max())); ndiff += int((err > 0).sum())
    assert worst <= RGB_TOL
    if exact_rgb:
        assert ndiff == 0, "RGB within tolerance but %d samples differ (max %d)" % (ndiff, worst)
    return worst, ndiff


@pytest.mark.parametrize("name", FIXTURES)
def test_reference_fixtures(dec, oracle_mod, golden, name):
    data = fixture_bytes(name)
    st, rgbs, coefs = dec.decode([data], want_coef=True)
    assert st == [0]
    g = golden["fixtures"][name]
    # digests taken from the UNTOUCHED reference binary (tests/golden/make_golden.py)
    assert sha(coefs[0].astype(np.int32)) == g["coef_sha256"]
    assert sha(rgbs[0]) == g["rgb_sha256"]


def test_golden_synthetic(dec, golden, harness):
    for name, g in golden["synthetic"].items():
        data = harness.synth_jpeg(**g["args"])
        st, rgbs, coefs = dec.decode([data], want_coef=True)
        assert st == [0], name
        assert sha(coefs[0].astype(np.int32)) == g["coef_sha256"], name
        assert sha(rgbs[0]) == g["rgb_sha256"], name


def test_mixed_batch_all_sampling_formats(dec, oracle_mod, harness):
    cases = [(64, 64, 90, 2, 2, 0, 0), (200, 120, 90, 2, 2, 0, -1), (200, 120, 90, 2, 1, 0, 7), (200, 120, 75, 1, 2, 0, 3),
             (123, 77, 95, 1, 1, 0, 0), (8, 8, 50, 2, 2, 0, 0), (1, 1, 90, 1, 1, 0, 0), (17, 9, 90, 2, 2, 0, 1),
             (300, 200, 50, 1, 1, 1, 0), (300, 200, 90, 1, 1, 1, 5), (679, 451, 30, 2, 2, 0, 0), (1000, 30, 99, 2, 1, 0, 0)]
    datas = [harness.synth_jpeg(100 + i, *c) for i, c in enumerate(cases)]
    compare(dec, oracle_mod, datas, flags=1)


def test_file_cut_right_behind_a_restart_marker_is_flagged(dec, harness):
    # an EMPTY restart segment (file cut behind an RSTn, or two RSTn in a row): flagged by both entropy paths, black
    # pixels, neighbours untouched
    from tests import jpeg_craft as jc
    good = harness.synth_jpeg(81, 200, 120, 85, 2, 2, False, 1)
    marks = jc.restart_markers(good)
    cut = good[:marks[-1] + 2] + b"\xff\xd9"
    mid = good[:marks[len(marks) // 2] + 2] + good[marks[len(marks) // 2 + 1]:]
    st, rgbs, _ = dec.decode([good, cut, mid, good], flags=1)
    assert st[0] == 0 and st[3] == 0 and np.array_equal(rgbs[0], rgbs[3])
    assert st[1] != 0 and st[2] != 0, st
    assert not rgbs[1].any() and not rgbs[2].any()


def test_streams_the_soak_found_are_flagged(dec, oracle_mod):
    # tests/golden/soak_*.jpg: what scripts/dev/soak.py found (see test_emul_kernels.py); flagged, black, by both paths
    import os
    from tests.conftest import GOLDEN_DIR
    datas = [open(os.path.join(GOLDEN_DIR, n), "rb").read() for n in ("soak_cut_behind_rst.jpg", "soak_leftover_behind_last_block.jpg")]
    st, rgbs, _ = dec.decode(datas, flags=1)
    for i, d in enumerate(datas):
        assert oracle_mod.port_decode(d, flags=1)[0] != 0
        assert st[i] != 0 and not rgbs[i].any()


def test_extreme_aspect_ratios_and_tiny_restart_intervals(dec, oracle_mod, harness):
    # one MCU wide or one MCU high, a restart marker after every MCU (thousands of segments of a few bytes), and a
    # restart interval that does not divide the MCU count
    cases = [(8, 3001, 90, 1, 1, 0, 1), (3001, 8, 90, 1, 1, 0, 1), (16, 2049, 85, 2, 2, 0, 1), (4100, 16, 85, 2, 2, 0, 0),
             (2047, 9, 92, 2, 1, 0, 1), (9, 1999, 92, 1, 2, 0, 7), (5, 4000, 80, 1, 1, 1, 3), (4000, 3, 80, 1, 1, 1, 0)]
    datas = [harness.synth_jpeg(900 + i, *c) for i, c in enumerate(cases)]
    compare(dec, oracle_mod, datas, flags=1)


def test_extreme_content(dec, oracle_mod, harness):
    """Pictures at the ends of the symbol statistics, many restart segments each:
    flat ones (every block is DC + EOB: a lane finishes four to five blocks per FIFO round -- the most the segment
    decoder's ring of block images ever holds, its room check has to wait), white noise at q98 (long codes and large
    magnitudes in nearly every block: the second table level in every round's ninth step), and both alternating MCU by MCU."""
    rng = np.random.default_rng(11)
    w, h = 640, 352
    flat = np.full((h, w, 3), 128, np.uint8)
    black = np.zeros((h, w, 3), np.uint8)
    noise = rng.integers(0, 256, size=(h, w, 3), dtype=np.uint8)
    mixed = flat.copy()
    for y in range(0, h, 16):
        for x in range(0, w, 16):
            if ((x + y) // 16) % 2:
                mixed[y:y + 16, x:x + 16] = noise[y:y + 16, x:x + 16]
    datas = [harness.encode_rgb(flat, 90, 2, 2, False, -1), harness.encode_rgb(black, 50, 2, 2, False, 1),
             harness.encode_rgb(noise, 98, 2, 2, False, -1), harness.encode_rgb(noise, 98, 1, 1, False, 3),
             harness.encode_rgb(mixed, 95, 2, 2, False, -1), harness.encode_rgb(mixed, 90, 2, 1, False, 2),
             harness.encode_rgb(flat, 90, 1, 1, True, 5), harness.encode_rgb(noise, 98, 1, 1, True, -1),
             # q100: quantisers of 1, 99 % of all coefficients non-zero -- a block takes eight rounds of a lane, the ring
             # of rounds between the writers and the flusher is what bounds them then
             harness.encode_rgb(noise, 100, 1, 1, False, -1), harness.encode_rgb(noise, 100, 1, 1, False, 0),
             harness.encode_rgb(noise, 100, 2, 2, False, 7)]
    compare(dec, oracle_mod, datas, flags=1)


def test_fixtures_in_one_batch_and_repeatable(dec, oracle_mod):
    datas = [fixture_bytes(n) for n in FIXTURES]
    compare(dec, oracle_mod, datas)
    a = dec.decode(datas)[1]
    b = dec.decode(datas)[1]
    assert all(np.array_equal(x, y) for x, y in zip(a, b))     # idempotent / deterministic


def test_sampling_coverage_c4_reduced(dec, oracle_mod, harness):
    # BASELINE config 4 (sampling-factor coverage) at 1024^2 so the CPU oracle finishes in seconds
    datas, flags = [], 1
    for q in (50, 90):
        datas += [harness.synth_jpeg(200 + q, 1024, 1024, q, 1, 1, True, 0), harness.synth_jpeg(300 + q, 1024, 1024, q, 2, 1, False, 0),
                  harness.synth_jpeg(400 + q, 1024, 1024, q, 1, 1, False, 0), harness.synth_jpeg(500 + q, 1024, 1024, q, 2, 1, False, 16)]
    compare(dec, oracle_mod, datas, flags=flags)


def test_corrupt_streams_flagged(dec, harness):
    from jpeg_decoder_b200 import abi
    good = bytearray(harness.synth_jpeg(41, 160, 96, 85, 2, 2, False, -1))
    body = good.find(b"\xff\xda") + 14
    a = bytearray(good); a[-2:] = b"\x00\x00"
    b = bytearray(good); b[body + 40:body + 42] = b"\xff\xc4"
    c = bytearray(good); k = c.find(b"\xff\xd1", body); c[k:k + 2] = b"\x12\x34"
    e = bytearray(good); e[body + 100:body + 140] = bytes(40)           # garbage bits: decode errors, no crash
    st, rgbs, _ = dec.decode([bytes(good), bytes(a), bytes(b), bytes(c), bytes(e), bytes(good)])
    assert st[0] == 0 and st[5] == 0
    assert st[1] & abi.ST_NO_EOI and st[2] & abi.ST_BAD_MARKER and st[3] & abi.ST_BAD_RST_COUNT
    assert np.array_equal(rgbs[0], rgbs[5])                              # neighbours in the batch unaffected


def test_truncated_and_random_garbage_do_not_crash(dec, harness):
    rng = np.random.default_rng(3)
    good = harness.synth_jpeg(43, 256, 256, 80, 2, 2, False, 0)
    body = good.find(b"\xff\xda") + 14
    datas = [good]
    for k in range(6):
        m = bytearray(good)
        pos = rng.integers(body, len(good) - 2, size=20)
        for p in pos:
            m[p] = int(rng.integers(0, 255))           # avoids 0xFF: stays a syntactically valid scan
        datas.append(bytes(m))
    # a header that promises 1024 blocks-rows of data over a scan that stops after 1 KB: the decoders must count the
    # blocks off without reading outside the scan buffer, and say so
    big = harness.synth_jpeg(44, 1024, 1024, 90, 2, 2, False, 0)
    cut = big[:big.find(b"\xff\xda") + 14 + 1000] + b"\xff\xd9"
    datas.append(cut)
    st, rgbs, _ = dec.decode(datas)
    assert st[0] == 0 and len(rgbs) == 8
    assert st[7] != 0


def test_decode_batch_is_chunked_and_matches_single_batch(dec, harness):
    # b200jpeg_decode_batch cuts batches of more than 8 images into chunks that rotate over four
    # pipelines (H2D / kernels / D2H overlap): results must equal the staged single-batch API
    import ctypes as C
    from jpeg_decoder_b200 import abi
    datas = [harness.synth_jpeg(700 + i, 96 + 8 * (i % 5), 64 + 16 * (i % 3), 85, 2, 2, False, -1 if i % 2 else 0) for i in range(21)]
    st, want, _ = dec.decode(datas)
    assert st == [0] * len(datas)
    n = len(datas)
    bufs = [np.frombuffer(d, dtype=np.uint8) for d in datas]
    descs = (abi.ImageDesc * n)()
    for i, b in enumerate(bufs):
        assert dec.lib.b200jpeg_parse_header(b.ctypes.data, b.size, 0, C.byref(descs[i])) == 0
    rgbs = [np.zeros_like(w) for w in want]
    st2 = dec.decode_batch(descs, bufs, rgbs)
    assert st2 == [0] * n
    for a, b in zip(want, rgbs):
        assert np.array_equal(a, b)
    s = dec.stats()
    assert s["images"] == n and s["rgb_bytes"] == sum(w.size for w in want) and s["d2h_bytes"] == s["rgb_bytes"]


def test_zero_padded_restart_segments(dec, oracle_mod, harness):
    # a stream padded with 0-bits before RSTn is non-conforming, but the reference decodes it (block count +
    # align(), jpeg.cpp:419-425; pinned in tests/test_oracle.py): both decoders must take it, status 0, same result
    from tests import jpeg_craft as jc
    d = harness.synth_jpeg(61, 200, 120, 85, 1, 1, False, 3)
    rc, _, c0, r0 = oracle_mod.port_decode(d)

    def same(x):
        rc, _, c, r = oracle_mod.port_decode(x)
        return rc == 0 and np.array_equal(c, c0) and np.array_equal(r, r0)
    z, cleared = jc.with_zero_padding(d, same)
    assert cleared > 100
    st, rgbs, coefs = dec.decode([z, d], want_coef=True)
    assert st == [0, 0]
    assert np.array_equal(coefs[0].astype(np.int32), c0) and np.array_equal(rgbs[0], r0) and np.array_equal(rgbs[1], r0)


def test_tables_with_more_long_codes_than_the_second_level_holds(dec, oracle_mod, harness):
    # 140 twelve-bit AC codes over 35 ten-bit prefixes: beyond the lookup tables' second level, resolved from the DHT data
    from tests import jpeg_craft as jc
    datas = [jc.with_ac_code_lengths(harness.synth_jpeg(71, 96, 64, 98, 2, 2, False, 0), jc.OVERFLOW_AC_BITS),
             jc.with_ac_code_lengths(harness.synth_jpeg(72, 88, 56, 95, 1, 1, False, 5), jc.OVERFLOW_AC_BITS),
             jc.with_ac_code_lengths(harness.synth_jpeg(74, 160, 120, 99, 2, 1, False, 0), jc.OVERFLOW_AC_BITS),
             harness.synth_jpeg(73, 64, 64, 90, 2, 2, False, 4)]
    compare(dec, oracle_mod, datas)


def test_spec_superset_component_ids_and_16bit_dqt(dec, oracle_mod, harness):
    # SURVEY.md 8(f) rank 1, "port only" (the reference rejects / truncates these): any component identifiers, 16-bit DQT
    from tests import jpeg_craft as jc
    d = harness.synth_jpeg(63, 120, 88, 85, 2, 2, False, 4)
    ids = jc.with_component_ids(d, (7, 9, 200))
    wide = jc.with_wide_dqt(d, lambda t, i, v: min(65535, v * 40) if i > 20 else v)
    compare(dec, oracle_mod, [ids, wide, d], flags=1)
    compare(dec, oracle_mod, [jc.with_wide_dqt(harness.synth_jpeg(64, 96, 40, 80, 1, 1, False, 0))])   # REFERENCE mode: low byte


def test_no_pixels_of_an_earlier_batch_come_back(dec, harness):
    # the device arenas are reused from batch to batch; an image the decoder gives up on must not
    # return what an earlier batch -- possibly another caller's picture -- left in its slot
    rng = np.random.default_rng(5)
    noise = rng.integers(1, 256, size=(256, 256, 3), dtype=np.uint8)
    first = [harness.encode_rgb(noise, 95, 2, 2, False, 0)] * 3
    st, rgbs, _ = dec.decode(first)
    assert st == [0, 0, 0] and rgbs[0].any()
    good = harness.synth_jpeg(45, 256, 256, 90, 2, 2, False, 0)
    body = good.find(b"\xff\xda") + 14
    no_eoi = good[:-2] + b"\x00\x00"
    cut = good[:body + 700] + b"\xff\xd9"                        # the header promises 1024 MCUs, the scan holds a few
    bad_marker = bytearray(good); bad_marker[body + 300:body + 302] = b"\xff\xc4"
    st, rgbs, coefs = dec.decode([no_eoi, cut, bytes(bad_marker)], want_coef=True)
    assert all(s != 0 for s in st)
    for r, c in zip(rgbs, coefs):
        assert not r.any(), "pixels of the earlier batch came back"
        assert not c.any(), "coefficients of the earlier batch came back"


def test_multi_device_entry_point(lib_so, oracle_mod, harness):
    # b200jpeg_multi_*: one context + one host thread per entry; on a one-GPU box the two entries are two contexts on
    # device 0 (the threading and the range arithmetic are what is under test); with more devices, the first two
    import ctypes as C
    import torch
    import jpeg_decoder_b200 as jd
    from jpeg_decoder_b200 import abi
    devs = [0, 1] if torch.cuda.device_count() >= 2 else [0, 0]
    datas = [harness.synth_jpeg(800 + i, 160 + 16 * (i % 4), 96, 85, 2, 2, False, -1 if i % 2 else 0) for i in range(37)]
    bufs = [np.frombuffer(d, dtype=np.uint8) for d in datas]
    descs, rcs = jd.parse_headers(bufs)
    assert not any(rcs)
    rgbs = [np.zeros((descs[i].image_height, descs[i].image_width, 3), np.uint8) for i in range(len(datas))]
    m = jd.MultiDecoder(devs)
    st = m.decode_batch(descs, bufs, rgbs)
    m.close()
    assert st == [0] * len(datas)
    for i in (0, 18, 19, 36):                                      # both sides of the range boundary
        rc, _, _, orgb = oracle_mod.port_decode(datas[i])
        assert rc == 0 and np.array_equal(rgbs[i], orgb), i
    # the handle-free form
    lib = jd.load()
    n = len(datas)
    files = (C.c_void_p * n)(*[b.ctypes.data for b in bufs])
    sizes = (C.c_size_t * n)(*[b.size for b in bufs])
    again = [np.zeros_like(r) for r in rgbs]
    prgb = (C.c_void_p * n)(*[r.ctypes.data for r in again])
    st2 = (C.c_int32 * n)()
    dv = (C.c_int * len(devs))(*devs)
    assert lib.b200jpeg_decode_batch_multi(dv, len(devs), descs, n, files, sizes, prgb, st2) == 0
    assert list(st2) == [0] * n and all(np.array_equal(a, b) for a, b in zip(again, rgbs))


def test_progressive_files(lib_so, oracle_mod, golden_prog):
    """SOF2 files, which the reference refuses (jpeg.cpp:69-73).  Coefficients and pixels against the
    digests the REFERENCE BINARY produced on each fixture's baseline twin; the reference's own two progressive fixtures
    (427x640 and 2832x4256) against the Annex-G port ("port only"); a batch mixing sizes, formats, grey and restarts."""
    import os
    import jpeg_decoder_b200 as jd
    from tests.conftest import GOLDEN_DIR
    dec = jd.Decoder(0)
    try:
        names = sorted(golden_prog)
        datas = [open(os.path.join(GOLDEN_DIR, "prog", n + ".jpg"), "rb").read() for n in names]
        own = [fixture_bytes("prograssive-sample-2.jpg"), fixture_bytes("prograssive-sample.jpg")]
        for rep in range(2):                                         # twice: the arenas are reused
            st, rgbs, coefs = dec.decode_progressive(datas + own, want_coef=True)
            assert st == [0] * (len(datas) + 2)
            for n, r, c in zip(names, rgbs, coefs):
                assert sha(c.astype(np.int32)) == golden_prog[n]["coef_sha256"], n
                assert sha(r) == golden_prog[n]["rgb_sha256"], n
        for d, r, c in zip(own, rgbs[len(datas):], coefs[len(datas):]):
            rc, _, oc, orgb = oracle_mod.port_decode_progressive(d)
            assert rc == 0 and np.array_equal(c.astype(np.int32), oc) and np.array_equal(r, orgb)
        # a baseline batch right after, in the same context
        st, rgbs, _ = dec.decode([fixture_bytes("img2.jpg")])
        rc, _, _, orgb = oracle_mod.port_decode(fixture_bytes("img2.jpg"))
        assert st == [0] and np.array_equal(rgbs[0], orgb)
        # corrupt progressive scan: status bits, no crash
        bad = bytearray(datas[0]); k = bad.rfind(b"\xff\xda"); bad[k + 20:k + 60] = bytes(40)
        st, _, _ = dec.decode_progressive([bytes(bad), datas[1]])
        assert st[1] == 0
        # what scripts/dev/soak_progressive.py found: around a destroyed RSTn the result is the reference's reading of the file
        # (BitStream::align, then on from the same byte) or a status; a scan that ends before its blocks do is flagged and black
        from tests import jpeg_craft as jc
        zeroed, cut = jc.damaged_progressive(open(os.path.join(GOLDEN_DIR, "prog", "p420_320x240_q90_rst.jpg"), "rb").read())
        st, rgbs, coefs = dec.decode_progressive([zeroed, cut, datas[1]], want_coef=True)
        rc, _, oc, orgb = oracle_mod.port_decode_progressive(zeroed)
        assert st[0] != 0 or (rc == 0 and np.array_equal(coefs[0].astype(np.int32), oc) and np.array_equal(rgbs[0], orgb))   # flagged, or the reference's reading
        assert oracle_mod.port_decode_progressive(cut)[0] != 0 and st[1] != 0 and not rgbs[1].any()
        assert st[2] == 0
    finally:
        dec.close()
