"""GPU tier (-m gpu): BASELINE.json's full-size configurations.  Full CPU-oracle comparison on a
sample, size-independent properties on the rest (every status clean, duplicates decode identically,
checksum of checksums stable across repeated runs)."""
import hashlib

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
