"""CPU tier: the product's kernel SOURCES (unstuff, entropy sync/fix/scan/write, fused pixel) and
launch sequence, compiled against the SIMT emulation in tests/emul, checked against the oracle.
This checks logic only; the same comparisons run on the real GPU in test_gpu_parity.py."""
import numpy as np
import pytest

from tests.conftest import fixture_bytes


def check(emul, oracle_mod, datas, flags=0, stress=False, path=2):
    """path: 2 = self-synchronising entropy decoder, 1 = segment-parallel decoder (b200jpeg.h)."""
    st, rgbs, coefs, info = emul.decode(datas, flags=flags, stress=stress, path=path)
    for i, d in enumerate(datas):
        rc, _, oc, orgb = oracle_mod.port_decode(d, flags=flags)
        assert rc == 0 and st[i] == 0, (i, rc, st[i])
        assert np.array_equal(coefs[i].astype(np.int32), oc), "coefficients differ, image %d" % i   # bit-exact
        assert np.array_equal(rgbs[i], orgb), "RGB differs, image %d" % i                              # 0 differing samples
    return info


def test_fixture_no_restart(emul, oracle_mod):
    check(emul, oracle_mod, [fixture_bytes("img2.jpg")])          # 4:2:0, optimised tables, self-sync path


def test_fixture_restart_444(emul, oracle_mod):
    check(emul, oracle_mod, [fixture_bytes("img4.jpg")])          # 4:4:4, DRI=100, 49 RSTn


def test_mixed_batch_all_formats(emul, oracle_mod, harness):
    cases = [(64, 64, 90, 2, 2, 0, 0), (200, 120, 90, 2, 2, 0, -1), (200, 120, 90, 2, 1, 0, 7),
             (200, 120, 75, 1, 2, 0, 3), (123, 77, 95, 1, 1, 0, 0), (8, 8, 50, 2, 2, 0, 0), (1, 1, 90, 1, 1, 0, 0),
             (17, 9, 90, 2, 2, 0, 1), (300, 200, 50, 1, 1, 1, 0), (300, 200, 90, 1, 1, 1, 5)]
    datas = [harness.synth_jpeg(100 + i, *c) for i, c in enumerate(cases)]
    check(emul, oracle_mod, datas, flags=1)


def test_segment_path_all_formats_and_fixtures(emul, oracle_mod, harness):
    # one thread per restart segment, straight from the stuffed bytes (kernels_segdec.cuh): RSTn every MCU row,
    # every 3/5/7 MCUs, none at all; optimised tables with long codes (fixtures); grey; odd sizes
    cases = [(64, 64, 90, 2, 2, 0, 0), (200, 120, 90, 2, 2, 0, -1), (200, 120, 90, 2, 1, 0, 7),
             (200, 120, 75, 1, 2, 0, 3), (123, 77, 95, 1, 1, 0, 0), (8, 8, 50, 2, 2, 0, 0), (1, 1, 90, 1, 1, 0, 0),
             (17, 9, 90, 2, 2, 0, 1), (300, 200, 50, 1, 1, 1, 0), (300, 200, 90, 1, 1, 1, 5)]
    datas = [harness.synth_jpeg(100 + i, *c) for i, c in enumerate(cases)]
    datas += [fixture_bytes("img2.jpg"), fixture_bytes("img4.jpg")]
    info = check(emul, oracle_mod, datas, flags=1, path=1)
    assert info["sync_rounds"] == 0


def test_segment_path_extreme_content(emul, oracle_mod, harness):
    # flat pictures (DC + EOB blocks: four to five blocks per lane and FIFO round, the fullest the ring of block images
    # gets) and white noise at q98 (long codes: the rounds' ninth step) through the segment decoder
    rng = np.random.default_rng(5)
    flat = np.full((48, 96, 3), 200, np.uint8)
    noise = rng.integers(0, 256, size=(48, 96, 3), dtype=np.uint8)
    datas = [harness.encode_rgb(flat, 90, 2, 2, False, -1), harness.encode_rgb(noise, 98, 2, 2, False, -1),
             harness.encode_rgb(flat, 50, 1, 1, True, 2), harness.encode_rgb(noise, 98, 1, 1, False, 3)]
    check(emul, oracle_mod, datas, flags=1, path=1)


def test_segment_path_flags_corrupt_streams(emul, harness):
    from jpeg_decoder_b200 import abi
    good = bytearray(harness.synth_jpeg(41, 160, 96, 85, 2, 2, False, -1))
    body = good.find(b"\xff\xda") + 14
    a = bytearray(good); a[-2:] = b"\x00\x00"
    b = bytearray(good); b[body + 40:body + 42] = b"\xff\xc4"
    c = bytearray(good); k = c.find(b"\xff\xd1", body); c[k:k + 2] = b"\x12\x34"
    e = bytearray(good); e[body + 100:body + 140] = bytes(40)
    st, rgbs, _, _ = emul.decode([bytes(good), bytes(a), bytes(b), bytes(c), bytes(e), bytes(good)], path=1)
    assert st[0] == 0 and st[5] == 0
    assert st[1] & abi.ST_NO_EOI and st[2] & abi.ST_BAD_MARKER and st[3] & abi.ST_BAD_RST_COUNT
    assert np.array_equal(rgbs[0], rgbs[5])


def test_multi_cta_scan_without_restart(emul, oracle_mod, harness):
    # > 16 KiB of scan: several entropy CTAs whose start states are only known after the fix launches
    d = harness.synth_jpeg(31, 640, 480, 92, 2, 2, False, 0)
    assert len(d) > 5 * 16384
    info = check(emul, oracle_mod, [d])
    assert info["sync_rounds"] == 2


def test_corrupt_streams_are_flagged_not_fatal(emul, harness):
    from jpeg_decoder_b200 import abi
    good = bytearray(harness.synth_jpeg(41, 160, 96, 85, 2, 2, False, -1))
    sos = good.find(b"\xff\xda")
    body = sos + 14
    a = bytearray(good); a[-2:] = b"\x00\x00"                 # EOI removed
    b = bytearray(good); b[body + 40:body + 42] = b"\xff\xc4"  # illegal marker inside the scan
    c = bytearray(good)
    k = c.find(b"\xff\xd1", body); c[k:k + 2] = b"\x12\x34"    # one RSTn destroyed
    st, _, _, _ = emul.decode([bytes(good), bytes(a), bytes(b), bytes(c)])
    assert st[0] == 0
    assert st[1] & abi.ST_NO_EOI
    assert st[2] & abi.ST_BAD_MARKER
    assert st[3] & abi.ST_BAD_RST_COUNT


def test_repair_and_fix_paths_under_stress(emul, oracle_mod, harness):
    # 512-bit subs, no run-in: most subs start out of sync, so the in-CTA Jacobi repair, the
    # cross-CTA fix launches and the host-side "keep repairing" loop all have to work
    datas = [harness.synth_jpeg(51, 320, 240, 90, 2, 2, False, 0), harness.synth_jpeg(52, 200, 160, 85, 2, 1, False, 5),
             fixture_bytes("img2.jpg")]
    info = check(emul, oracle_mod, datas, stress=True)
    assert info["sync_rounds"] >= 2


def test_zero_padded_restart_segments_decode_on_both_paths(emul, oracle_mod, harness):
    # a non-conforming stream the reference decodes (it counts blocks and never looks at the padding, jpeg.cpp:419-425):
    # both entropy decoders must take it, with the same status and the same result
    from tests import jpeg_craft as jc
    d = harness.synth_jpeg(61, 200, 120, 85, 1, 1, False, 3)
    rc, _, c0, r0 = oracle_mod.port_decode(d)

    def same(x):
        rc, _, c, r = oracle_mod.port_decode(x)
        return rc == 0 and np.array_equal(c, c0) and np.array_equal(r, r0)
    z, cleared = jc.with_zero_padding(d, same)
    assert cleared > 100
    for path in (1, 2):
        st, rgbs, coefs, _ = emul.decode([z, d], path=path)
        assert st == [0, 0]
        assert np.array_equal(coefs[0].astype(np.int32), c0) and np.array_equal(rgbs[0], r0)
        assert np.array_equal(rgbs[1], r0)


def test_file_cut_right_behind_a_restart_marker_is_flagged_on_both_paths(emul, harness):
    # the last restart segment holds no bit at all (found by scripts/dev/soak.py): the self-synchronising path has no
    # sub-sequence there, so the planner itself has to say that the segment's blocks are missing
    from jpeg_decoder_b200 import abi
    from tests import jpeg_craft as jc
    good = harness.synth_jpeg(81, 200, 120, 85, 2, 2, False, 1)
    marks = jc.restart_markers(good)
    cut = good[:marks[-1] + 2] + b"\xff\xd9"                   # ... RSTn FF D9
    mid = good[:marks[len(marks) // 2] + 2] + good[marks[len(marks) // 2 + 1]:]   # a segment's bytes removed: RSTn RSTn+1 in a row
    for path in (1, 2):
        st, rgbs, _, _ = emul.decode([good, cut, mid, good], flags=1, path=path)
        assert st[0] == 0 and st[3] == 0 and np.array_equal(rgbs[0], rgbs[3])
        assert st[1] != 0 and st[2] != 0, (path, st)
        assert not rgbs[1].any() or (st[1] & (abi.ST_BLOCK_COUNT | abi.ST_OVERRUN | abi.ST_BAD_RST_COUNT))


def test_streams_the_soak_found_are_flagged_on_both_paths(emul, oracle_mod):
    # scripts/dev/soak.py (random corruption, both entropy paths against the oracle) found two kinds of stream that the
    # reference's logic rejects and the self-synchronising path let through: a file cut right behind an RSTn, and a
    # restart segment with more than padding left over behind its last block.  The files are kept as fixtures.
    import os
    from tests.conftest import GOLDEN_DIR
    for name in ("soak_cut_behind_rst.jpg", "soak_leftover_behind_last_block.jpg"):
        d = open(os.path.join(GOLDEN_DIR, name), "rb").read()
        assert oracle_mod.port_decode(d, flags=1)[0] != 0
        for path in (1, 2):
            st, _, _, _ = emul.decode([d], flags=1, path=path)
            assert st[0] != 0, (name, path)


def test_extreme_aspect_ratios_and_tiny_restart_intervals(emul, oracle_mod, harness):
    # one MCU wide / one MCU high, a restart marker after every MCU; both entropy paths
    cases = [(8, 600, 90, 1, 1, 0, 1), (600, 8, 90, 1, 1, 0, 1), (16, 401, 85, 2, 2, 0, 1), (520, 16, 85, 2, 2, 0, 0),
             (9, 333, 92, 1, 2, 0, 7), (5, 300, 80, 1, 1, 1, 3)]
    datas = [harness.synth_jpeg(900 + i, *c) for i, c in enumerate(cases)]
    check(emul, oracle_mod, datas, flags=1, path=1)
    check(emul, oracle_mod, datas, flags=1, path=2)


def test_tables_with_more_long_codes_than_the_second_level_holds(emul, oracle_mod, harness):
    # AC tables whose 140 twelve-bit codes cover 35 ten-bit prefixes (kernels_huff.cuh provides a second level for 8):
    # the codes beyond them are resolved from the DHT data (huff_slow) -- in the ninth step of the segment decoder's
    # rounds and in the self-synchronising walk.  The reference decodes these files bit by bit (test_oracle.py).
    from tests import jpeg_craft as jc
    datas = [jc.with_ac_code_lengths(harness.synth_jpeg(71, 96, 64, 98, 2, 2, False, 0), jc.OVERFLOW_AC_BITS),
             jc.with_ac_code_lengths(harness.synth_jpeg(72, 88, 56, 95, 1, 1, False, 5), jc.OVERFLOW_AC_BITS),
             harness.synth_jpeg(73, 64, 64, 90, 2, 2, False, 4)]          # a CTA of its own: standard tables, no overflow
    check(emul, oracle_mod, datas, path=1)
    check(emul, oracle_mod, datas, path=2)


def test_spec_superset_component_ids_and_16bit_dqt(emul, oracle_mod, harness):
    from tests import jpeg_craft as jc
    d = harness.synth_jpeg(63, 120, 88, 85, 2, 2, False, 4)
    ids = jc.with_component_ids(d, (7, 9, 200))
    wide = jc.with_wide_dqt(d, lambda t, i, v: min(65535, v * 40) if i > 20 else v)
    check(emul, oracle_mod, [ids, wide, d], flags=1)
    check(emul, oracle_mod, [ids, wide, d], flags=1, path=1)
    check(emul, oracle_mod, [jc.with_wide_dqt(harness.synth_jpeg(64, 96, 40, 80, 1, 1, False, 0))])   # REFERENCE mode: low byte


def test_sub_sizes_of_the_self_synchronising_path(emul, oracle_mod, harness):
    # several CTAs of subs, units of more than one sub, restart segments shorter and longer than a sub
    datas = [harness.synth_jpeg(71, 512, 384, 92, 2, 2, False, 0), harness.synth_jpeg(72, 400, 300, 90, 2, 1, False, 11),
             harness.synth_jpeg(73, 256, 256, 95, 1, 1, False, 1), fixture_bytes("img4.jpg")]
    check(emul, oracle_mod, datas, flags=1)


def test_progressive_scans_against_reference_digests(emul, oracle_mod, golden_prog):
    # kernels_prog.cuh (T.81 Annex G) through the emulation: the digests are the reference binary's, taken on each file's
    # baseline twin (tests/golden/make_prog_golden.py); plus one of the reference's own progressive fixtures against the port
    import hashlib
    import os
    from tests.conftest import GOLDEN_DIR
    names = sorted(golden_prog)
    datas = [open(os.path.join(GOLDEN_DIR, "prog", n + ".jpg"), "rb").read() for n in names] + [fixture_bytes("prograssive-sample-2.jpg")]
    st, rgbs, coefs = emul.decode_progressive(datas)
    assert st == [0] * len(datas)
    sha = lambda a: hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()
    for n, r, c in zip(names, rgbs, coefs):
        assert sha(c.astype(np.int32)) == golden_prog[n]["coef_sha256"] and sha(r) == golden_prog[n]["rgb_sha256"], n
    rc, _, oc, orgb = oracle_mod.port_decode_progressive(datas[-1])
    assert rc == 0 and np.array_equal(coefs[-1].astype(np.int32), oc) and np.array_equal(rgbs[-1], orgb)
    # damaged files: around a destroyed RSTn the result is the reference's reading of the file or a status; a scan that ends early is flagged
    from tests import jpeg_craft as jc
    zeroed, cut = jc.damaged_progressive(open(os.path.join(GOLDEN_DIR, "prog", "p420_320x240_q90_rst.jpg"), "rb").read())
    st, rgbs, coefs = emul.decode_progressive([zeroed, cut])
    rc, _, oc, orgb = oracle_mod.port_decode_progressive(zeroed)
    assert st[0] != 0 or (rc == 0 and np.array_equal(coefs[0].astype(np.int32), oc) and np.array_equal(rgbs[0], orgb))   # flagged, or the reference's reading
    assert oracle_mod.port_decode_progressive(cut)[0] != 0 and st[1] != 0



def test_mini_soak_of_both_paths_against_the_oracle(emul, oracle_mod, harness):
    # a small fixed-seed edition of scripts/dev/soak.py (which runs thousands of batches on the GPU): random pictures and
    # corrupted copies of them through both entropy paths -- a clean one is bit-exact, a corrupted one is flagged or
    # decoded to what the reference's logic makes of it, and whatever the oracle rejects is flagged
    rng = np.random.default_rng(5)
    datas, clean = [], []
    for i in range(10):
        w, h = int(rng.integers(8, 160)), int(rng.integers(8, 120))
        hs, vs = [(1, 1), (2, 1), (1, 2), (2, 2)][int(rng.integers(0, 4))]
        d = harness.synth_jpeg(600 + i, w, h, int(rng.integers(30, 99)), hs, vs, False, int([0, -1, 1, 3][int(rng.integers(0, 4))]))
        datas.append(d); clean.append(True)
        m = bytearray(d)
        body = d.find(b"\xff\xda") + 14
        at = body + int(rng.integers(0, max(1, len(d) - 2 - body - 8)))
        ln = int(rng.integers(1, 60))
        kind = i % 5
        if kind == 0:
            m[at:at + ln] = bytes(ln)
        elif kind == 1:
            m = m[:at] + b"\xff\xd9"
        elif kind == 2:
            m[at:at + ln] = (b"\xff\x00" * ln)[:ln]
        elif kind == 3:
            m[at:at + ln] = b"\xff" * ln
        else:
            for p in rng.integers(body, len(d) - 2, size=6):
                m[p] = int(rng.integers(0, 255))
        datas.append(bytes(m)); clean.append(False)
    for path in (1, 2):
        st, rgbs, coefs, _ = emul.decode(datas, flags=1, path=path)
        for i, d in enumerate(datas):
            rc, _, oc, orgb = oracle_mod.port_decode(d, flags=1)
            if clean[i]:
                assert rc == 0 and st[i] == 0 and np.array_equal(coefs[i].astype(np.int32), oc) and np.array_equal(rgbs[i], orgb), (path, i)
            elif st[i] == 0:
                assert rc == 0 and np.array_equal(coefs[i].astype(np.int32), oc), (path, i, rc)
