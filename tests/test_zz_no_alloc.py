"""GPU tier, collected LAST on purpose (a failure here must not hide another test).

b200jpeg_export_coefficients_noalloc -- the entry point added for callers that
cannot call cudaMalloc: the coefficient export kernel (k_export_coef) with its staging area in a file-scope static
device array of B200JPEG_NOALLOC_MAX_BYTES instead of a buffer allocated on demand.  Checked through the C ABI against
the existing entry point (b200jpeg_batch_fetch with coef_out), bit for bit -- the tolerance every coefficient test of
this repository uses -- on two calls in a row, with a larger image exported in between so that a call that did not
clear the shared array would show the leftovers."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def test_noalloc_export_equals_batch_fetch_twice_in_a_row(lib_so, harness):
    import jpeg_decoder_b200 as jd
    from jpeg_decoder_b200 import abi
    dec = jd.Decoder(0)
    try:
        # 1080p 4:2:0 (12.5 MB of MCU array: close to the 16 MiB limit), a small 4:4:4 without restart markers, a 4:2:2
        # with odd geometry and restarts, and one picture too large for the static array
        datas = [harness.synth_jpeg(2001, 1920, 1080, 90, 2, 2, False, -1), harness.synth_jpeg(2002, 123, 77, 95, 1, 1, False, 0),
                 harness.synth_jpeg(2003, 333, 211, 80, 2, 1, False, 7), harness.synth_jpeg(2004, 2560, 2048, 75, 2, 2, False, 0)]
        st, _, want = dec.decode(datas, want_coef=True)              # existing entry point: b200jpeg_batch_fetch(..., coef_out, ...)
        assert st == [0, 0, 0, 0]
        assert want[0].nbytes <= abi.NOALLOC_MAX_BYTES < want[3].nbytes
        for order in ([0, 1, 0, 2, 1], [2, 1, 0]):                   # big, small, big again, ...: every call sees the last one's bytes
            for i in order:
                for call in range(2):                                # two calls in a row
                    got = np.full_like(want[i], 0x5A5A)
                    dec.export_coefficients_noalloc(i, got)
                    assert np.array_equal(got, want[i]), "image %d, call %d: differs from b200jpeg_batch_fetch" % (i, call)
        # beyond the stated maximum: a clear error, nothing written
        got = np.full_like(want[3], 0x5A5A)
        rc = dec.lib.b200jpeg_export_coefficients_noalloc(dec.h, 3, got.ctypes.data, got.nbytes)
        assert rc == abi.E_TOO_LARGE
        assert b"B200JPEG_NOALLOC_MAX_BYTES" in dec.lib.b200jpeg_last_error(dec.h)
        assert (got == 0x5A5A).all()
        # argument checks: image index, output size
        small = np.zeros_like(want[1])
        assert dec.lib.b200jpeg_export_coefficients_noalloc(dec.h, 9, small.ctypes.data, small.nbytes) == abi.E_ARG
        assert dec.lib.b200jpeg_export_coefficients_noalloc(dec.h, 0, small.ctypes.data, small.nbytes) == abi.E_ARG
        # a second batch in the same context, and a corrupt image in it: zeros, as the existing entry point gives
        good = harness.synth_jpeg(2005, 200, 120, 85, 2, 2, False, 0)
        st, _, want2 = dec.decode([good[:-2] + b"\x00\x00", good], want_coef=True)
        assert st[0] != 0 and st[1] == 0 and not want2[0].any()
        for i in (1, 0, 1):
            got = np.full_like(want2[i], 0x5A5A)
            dec.export_coefficients_noalloc(i, got)
            assert np.array_equal(got, want2[i]), i
    finally:
        dec.close()


def test_noalloc_needs_a_batch(lib_so):
    import jpeg_decoder_b200 as jd
    from jpeg_decoder_b200 import abi
    dec = jd.Decoder(0)
    try:
        out = np.zeros((4, 3, 64), np.int16)
        assert dec.lib.b200jpeg_export_coefficients_noalloc(dec.h, 0, out.ctypes.data, out.nbytes) == abi.E_STATE
        assert dec.lib.b200jpeg_export_coefficients_noalloc(None, 0, out.ctypes.data, out.nbytes) == abi.E_ARG
    finally:
        dec.close()
