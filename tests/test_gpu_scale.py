"""GPU tier (-m gpu): BASELINE.json's configurations at their NAMED sizes.  Digests taken from the reference binary
itself at those sizes (tests/golden/golden.json "named_size", made by make_golden.py --named), full CPU-oracle
comparison on samples, size-independent properties on the rest (every status clean, duplicates decode identically,
checksum of checksums stable across repeated runs)."""
import hashlib
from concurrent.futures import ThreadPoolExecutor

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def dec(lib_so):
    import jpeg_decoder_b200 as jd
    d = jd.Decoder(0)
    yield d
    d.close()


def check_sample(dec_out, datas, oracle_mod, idx, flags=0):
    st, rgbs, coefs = dec_out
    for i in idx:
        rc, _, oc, orgb = oracle_mod.port_decode(datas[i], flags=flags)
        assert rc == 0 and st[i] == 0
        assert np.array_equal(coefs[i].astype(np.int32), oc), "coefficients differ (image %d)" % i
        err = np.abs(rgbs[i].astype(np.int16) - orgb.astype(np.int16))
        assert err.max() <= 1 and (err > 0).sum() == 0, (i, int(err.max()), int((err > 0).sum()))


def test_c2_batch_1080p_restart_rows(dec, oracle_mod, harness):
    # config 2 at 64 images (same per-image structure as the 256-image bench batch)
    n = 64
    datas = [harness.synth_jpeg(1000 + i, 1920, 1080, 90, 2, 2, False, -1) for i in range(n - 2)]
    datas += [datas[0], datas[5]]                                   # duplicates
    out = dec.decode(datas, want_coef=True)
    assert all(s == 0 for s in out[0])
    check_sample(out, datas, oracle_mod, [0, 17, 61])
    assert np.array_equal(out[1][62], out[1][0]) and np.array_equal(out[1][63], out[1][5])
    digest = hashlib.sha256(b"".join(hashlib.sha256(r.tobytes()).digest() for r in out[1])).hexdigest()
    again = dec.decode(datas)
    assert digest == hashlib.sha256(b"".join(hashlib.sha256(r.tobytes()).digest() for r in again[1])).hexdigest()


def test_c3_single_large_444_no_restart(dec, oracle_mod, harness):
    # config 3 (self-synchronising path over one long scan): 4096^2 here keeps the CPU oracle ~10 s;
    # the 8192^2 case is exercised in test_c3_full_size_properties
    d = harness.synth_jpeg(3000, 4096, 4096, 95, 1, 1, False, 0)
    out = dec.decode([d], want_coef=True)
    check_sample(out, [d], oracle_mod, [0])
    assert dec.stats()["sync_rounds"] >= 2


def test_c3_full_size_properties(dec, harness):
    d = harness.synth_jpeg(3001, 8192, 8192, 95, 1, 1, False, 0)
    st, rgbs, _ = dec.decode([d])
    assert st == [0]
    st2, rgbs2, _ = dec.decode([d, d])
    assert st2 == [0, 0] and np.array_equal(rgbs2[0], rgbs[0]) and np.array_equal(rgbs2[1], rgbs[0])
    # the same picture encoded WITH restart markers (one per MCU row) carries the same coefficients in a
    # different bit stream (DC prediction restarts, byte alignment, RSTn): the segment-parallel decoder on it
    # and the self-synchronising decoder on the marker-free stream must produce identical pixels
    from jpeg_decoder_b200 import abi
    dr = harness.synth_jpeg(3001, 8192, 8192, 95, 1, 1, False, -1)
    dec.set_entropy_path(abi.ENTROPY_SEGMENT)
    st3, rgbs3, _ = dec.decode([dr])
    dec.set_entropy_path(abi.ENTROPY_AUTO)
    assert st3 == [0] and np.array_equal(rgbs3[0], rgbs[0])
    # the decoded picture must resemble the source it was encoded from (q95: high PSNR)
    src = harness.synth_rgb(3001, 8192, 8192)
    mse = np.mean((rgbs[0][::8, ::8].astype(np.float32) - src[::8, ::8].astype(np.float32)) ** 2)
    assert 10 * np.log10(255.0 ** 2 / mse) > 35.0


def test_c5_many_small_images(dec, oracle_mod, harness):
    n = 1024
    datas = [harness.synth_jpeg(5000 + i, 512, 512, 85, 2, 2, False, 0) for i in range(n)]
    out = dec.decode(datas, want_coef=True)
    assert all(s == 0 for s in out[0])
    check_sample(out, datas, oracle_mod, [0, 1, 511, 1023])


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def synth_many(harness, seeds, *args):
    harness.build_harness()
    with ThreadPoolExecutor(max_workers=32) as ex:
        return list(ex.map(lambda s: harness.synth_jpeg(s, *args), seeds))


@pytest.mark.parametrize("name", ["c3_444_8192_q95", "c4_422_4096_q50", "c4_422_4096_q90", "c4_444_4096_q50", "c4_444_4096_q90"])
def test_named_size_single_images_against_reference_digests(dec, golden, harness, name):
    # C3 at 8192^2 (32-bit bit positions past 2^27, one scan of 28 MB through the self-synchronising path) and C4's
    # 4:2:2 / 4:4:4 at 4096^2, q50 and q90: coefficients and pixels against what the reference binary produced
    g = golden["named_size"][name]
    data = harness.synth_jpeg(**g["args"])
    assert hashlib.sha256(data).hexdigest() == g["jpeg_sha256"]
    st, rgbs, coefs = dec.decode([data], want_coef=True)
    assert st == [0]
    assert sha(rgbs[0]) == g["rgb_sha256"]
    assert sha(coefs[0].astype(np.int32)) == g["coef_sha256"]


@pytest.mark.parametrize("q", [50, 90])
def test_c4_gray_4096_against_port(dec, oracle_mod, harness, q):
    # grey is rejected by the reference (jpeg.cpp:83-87): port only, pinned through the zero-chroma cross-check of test_oracle.py
    d = harness.synth_jpeg(4300 + q, 4096, 4096, q, 1, 1, True, 0)
    st, rgbs, coefs = dec.decode([d], flags=1, want_coef=True)
    rc, _, oc, orgb = oracle_mod.port_decode(d, flags=1)
    assert st == [0] and rc == 0
    assert np.array_equal(coefs[0].astype(np.int32), oc) and np.array_equal(rgbs[0], orgb)


def test_c2_named_size_256_images(dec, oracle_mod, golden, harness):
    n = 256
    datas = synth_many(harness, range(1000, 1000 + n), 1920, 1080, 90, 2, 2, False, -1)
    st, rgbs, _ = dec.decode(datas)
    assert st == [0] * n
    g = golden["named_size"]["c2_420_1080p_rstrow_255"]
    assert hashlib.sha256(datas[255]).hexdigest() == g["jpeg_sha256"] and sha(rgbs[255]) == g["rgb_sha256"]
    for i in (0, 100, 254):
        rc, _, _, orgb = oracle_mod.port_decode(datas[i])
        assert rc == 0 and np.array_equal(rgbs[i], orgb), i
    s2, c255 = dec.decode([datas[255]], want_coef=True)[0::2]
    assert s2 == [0] and sha(c255[0].astype(np.int32)) == g["coef_sha256"]


def test_c5_named_size_16384_images(dec, oracle_mod, golden, harness):
    # the whole sharding-sweep batch on one GPU: 12.9 GB of coefficients and of RGB in the arenas, 400 k sub slots
    n = 16384
    datas = synth_many(harness, range(5000, 5000 + n), 512, 512, 85, 2, 2, False, 0)
    st, rgbs, _ = dec.decode(datas)
    assert st == [0] * n
    g = golden["named_size"]["c5_420_512_q85_16383"]
    assert hashlib.sha256(datas[n - 1]).hexdigest() == g["jpeg_sha256"] and sha(rgbs[n - 1]) == g["rgb_sha256"]
    for i in (0, 1, 2047, 2048, 8191, 8192, 12345, 16382):
        rc, _, _, orgb = oracle_mod.port_decode(datas[i])
        assert rc == 0 and np.array_equal(rgbs[i], orgb), i
    digest = hashlib.sha256(b"".join(hashlib.sha256(r.tobytes()).digest() for r in rgbs[::64])).hexdigest()
    del rgbs
    st2, again, _ = dec.decode(datas)
    assert st2 == [0] * n
    assert digest == hashlib.sha256(b"".join(hashlib.sha256(r.tobytes()).digest() for r in again[::64])).hexdigest()
