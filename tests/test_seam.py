"""The drop-in proof: the REFERENCE's own marker reader and read_* functions parse the file, INTEGRATION.md's fill_desc()
(tests/seam/fill_desc.inc) turns the reference's Image members into a b200jpeg_image_desc, b200jpeg_decode_batch() decodes,
and the reference's own decodeHuffman / dequantize / inverseDCT / YCbCrToRGB in the same process are the judge
(tests/seam/ref_seam_post.h).  The binaries are built where /root/reference exists (tests/seam/build.sh) and travel
prebuilt to the GPU box."""
import json
import os
import subprocess

import pytest

from tests.conftest import GOLDEN_DIR, ROOT

SEAM = os.path.join(ROOT, "tests", "seam", "_build", "ref_seam")
SEAM_EMU = SEAM + "_emu"
FIXTURES = ["img.jpg", "img2.jpg", "img3.jpg", "img4.jpg", "img5.jpg", "img6.jpg"]


def build(emul=None):
    if os.path.exists("/root/reference/jpeg.cpp"):
        subprocess.check_call(["sh", os.path.join(ROOT, "tests", "seam", "build.sh")], stdout=subprocess.DEVNULL)


def run(exe, paths):
    p = subprocess.run([exe] + paths, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=900)
    lines = [json.loads(l) for l in p.stdout.splitlines() if l.startswith("{")]
    return p.returncode, lines, p.stderr


def test_seam_logic_against_the_emulation(lib_so, emul, harness, tmp_path):
    # CPU tier: fill_desc() on the reference's members + the kernels' emulation; checks the stub, not the GPU
    build()
    if not os.path.exists(SEAM_EMU):
        pytest.skip("tests/seam/_build/ref_seam_emu not built (reference sources absent)")
    dri = tmp_path / "dri.jpg"
    dri.write_bytes(harness.synth_jpeg(91, 200, 120, 88, 2, 1, False, 7))
    rc, lines, err = run(SEAM_EMU, [os.path.join(GOLDEN_DIR, "img2.jpg"), os.path.join(GOLDEN_DIR, "img4.jpg"), str(dri)])
    assert rc == 0 and len(lines) == 3, (lines, err)
    for l in lines:
        assert l["ok"] and l["coef_mismatch"] == 0 and l["rgb_differing"] == 0, l


@pytest.mark.gpu
def test_reference_parser_feeds_the_library(lib_so, harness, tmp_path):
    if not os.path.exists(SEAM):
        pytest.skip("tests/seam/_build/ref_seam not built (reference sources absent when the tree was prepared)")
    extra = []
    for k, (args, name) in enumerate([((92, 1920, 1080, 90, 2, 2, False, -1), "c2.jpg"), ((93, 640, 480, 95, 1, 1, False, 0), "s444.jpg"),
                                      ((94, 333, 211, 75, 2, 1, False, 7), "s422_dri7.jpg")]):
        p = tmp_path / name
        p.write_bytes(harness.synth_jpeg(*args))
        extra.append(str(p))
    rc, lines, err = run(SEAM, [os.path.join(GOLDEN_DIR, n) for n in FIXTURES] + extra)
    assert rc == 0 and len(lines) == len(FIXTURES) + len(extra), (lines, err)
    for l in lines:
        assert l["ok"] and l["coef_mismatch"] == 0, l           # coefficients bit-exact with the reference's MCU decode
        assert l["rgb_maxerr"] <= 1, l                          # north_star: +-1 LSB
        assert l["rgb_differing"] == 0, l                       # measured: none differ
