"""include/b200jpeg.hpp + tools/b200jpeg_cli.cpp: the reference's C++ surface (class Image, jpeg.cpp:19-914,
and main(), jpeg.cpp:916-929) over the C ABI.  CPU tier: it builds, parses headers like the reference and
FAILS LOUDLY without a GPU.  GPU tier: the files it writes equal the oracle's pixels."""
import os
import subprocess

import numpy as np
import pytest

from tests.conftest import GOLDEN_DIR, ROOT

PROBE = r"""
#include <iostream>
#include "b200jpeg.hpp"
int main(int argc, char **argv) {
  try {
    b200jpeg::Image img(argv[1]);
    if (argc > 2 && std::string(argv[2]) == "markers") {
      for (b200jpeg::Marker &m : img.markers()) {
        std::cout << "M " << m.type << " " << (int)m.marker << " " << m.length << " " << (m.read(1) ? (int)m.read(0)[-1] : -1) << " " << m.scan << std::endl;
      }
      return 0;
    }
    img.readHeaders();
    std::cout << img.image_width << " " << img.image_height << " " << img.mcuWidthReal << " " << img.mcuHeightReal << " "
              << img.restartInterval << " " << img.color_components[0].horizontal_sampling_factor
              << img.color_components[0].vertical_sampling_factor << " " << img.quantization_tables[0][0] << " "
              << img.quantization_tables[0][8] << " " << (int)img.huffmanDCTTable[0].offset[16] << " "
              << img.huffmanACTable[0].codes[0] << std::endl;
    if (argc > 2) { img.readJPEG(); std::cout << "decoded" << std::endl; }
    return 0;
  } catch (const b200jpeg::Error &e) { std::cerr << "ERR " << e.code << " " << e.what() << std::endl; return 1; }
}
"""


@pytest.fixture(scope="module")
def cli(lib_so):
    from jpeg_decoder_b200 import build
    return build.build_cli()


@pytest.fixture(scope="module")
def probe(lib_so, tmp_path_factory):
    d = tmp_path_factory.mktemp("probe")
    src = d / "probe.cpp"
    src.write_text(PROBE)
    exe = str(d / "probe")
    pkg = os.path.join(ROOT, "jpeg_decoder_b200")
    subprocess.check_call(["g++", "-std=c++17", "-O1", "-Wall", "-Werror", "-I" + os.path.join(ROOT, "include"), "-o", exe, str(src),
                           "-L" + pkg, "-lb200jpeg", "-Wl,-rpath," + pkg])
    return exe


def run(cmd):
    p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True, timeout=300)
    return p.returncode, p.stdout, p.stderr


def test_image_headers_match_python_parser(probe, lib_so):
    import jpeg_decoder_b200 as jd
    for name in ["img.jpg", "img2.jpg", "img5.jpg"]:
        path = os.path.join(GOLDEN_DIR, name)
        rc, out, err = run([probe, path])
        assert rc == 0, err
        f = out.split()
        d = jd.parse_header(open(path, "rb").read())
        assert [int(f[0]), int(f[1]), int(f[2]), int(f[3]), int(f[4])] == [
            d.image_width, d.image_height, d.mcuWidthReal, d.mcuHeightReal, d.restartInterval]
        c0 = d.color_components[0]
        assert f[5] == "%d%d" % (c0.horizontal_sampling_factor, c0.vertical_sampling_factor)
        q = d.quantization_tables[c0.quantizationTableID]
        assert [int(f[6]), int(f[7])] == [q.table[0], q.table[8]]      # natural order, types.hpp:88-95
        assert int(f[8]) == d.huffmanDCTTable[0].offset[16]


def test_reference_error_behaviour(probe, cli, tmp_path):
    # not a JPEG (jpeg.cpp:801-805), progressive (jpeg.cpp:843-846): refused before any GPU work
    bad = tmp_path / "bad.jpg"
    bad.write_bytes(b"\x89PNG....")
    rc, out, err = run([probe, str(bad)])
    assert rc == 1 and "Invalid JPEG" in err
    rc, out, err = run([probe, os.path.join(GOLDEN_DIR, "prograssive-sample-2.jpg")])
    assert rc == 1 and "ERR" in err
    rc, out, err = run([cli])                    # usage error, jpeg.cpp:918-922
    assert rc == 1 and "Usage" in err


@pytest.mark.skipif(os.path.exists("/dev/nvidia0"), reason="only meaningful without a GPU")
def test_no_gpu_is_a_loud_failure(probe):
    rc, out, err = run([probe, os.path.join(GOLDEN_DIR, "img2.jpg"), "decode"])
    assert rc == 1 and "no usable sm_100 GPU" in err and "decoded" not in out


@pytest.mark.gpu
def test_cli_writes_the_oracles_pixels(cli, oracle_mod, tmp_path):
    names = ["img.jpg", "img2.jpg", "img4.jpg"]
    rc, out, err = run([cli] + [os.path.join(GOLDEN_DIR, n) for n in names] + ["--out-dir", str(tmp_path)])
    assert rc == 0, err
    for n in names:
        _, _, _, orgb = oracle_mod.port_decode(open(os.path.join(GOLDEN_DIR, n), "rb").read())
        raw = open(tmp_path / (n[:-4] + ".ppm"), "rb").read()
        head = b"P6\n%d %d\n255\n" % (orgb.shape[1], orgb.shape[0])
        assert raw.startswith(head)
        got = np.frombuffer(raw[len(head):], np.uint8).reshape(orgb.shape)
        assert np.array_equal(got, orgb), n
    # BMP: bottom-up B,G,R rows padded to 4 bytes (what jpeg.cpp:448-509 means to write)
    rc, out, err = run([cli, os.path.join(GOLDEN_DIR, "img2.jpg"), "--out-dir", str(tmp_path), "--bmp"])
    assert rc == 0, err
    _, _, _, orgb = oracle_mod.port_decode(open(os.path.join(GOLDEN_DIR, "img2.jpg"), "rb").read())
    h, w = orgb.shape[:2]
    raw = open(tmp_path / "img2.bmp", "rb").read()
    pitch = (w * 3 + 3) // 4 * 4
    assert raw[:2] == b"BM" and len(raw) == 26 + pitch * h
    rows = np.frombuffer(raw[26:], np.uint8).reshape(h, pitch)[:, :w * 3].reshape(h, w, 3)
    assert np.array_equal(rows[::-1, :, ::-1], orgb)


@pytest.mark.gpu
def test_image_readjpeg_decodes(probe):
    rc, out, err = run([probe, os.path.join(GOLDEN_DIR, "img3.jpg"), "decode"])
    assert rc == 0 and "decoded" in out, err


def test_marker_reader_classifies_like_the_reference(probe, harness, tmp_path):
    # Marker / MarkerType (include/file.hpp:167-258): APPn -> META (-2), COM -> INVALID (-1), SOF0/SOF2 -> SOF (0xc0),
    # `length` counts its own two bytes, the payload starts behind it; SOS reports where the scan starts
    from tests import jpeg_craft as jc
    for name in ["img.jpg", "img4.jpg"]:
        path = os.path.join(GOLDEN_DIR, name)
        data = open(path, "rb").read()
        rc, out, err = run([probe, path, "markers"])
        assert rc == 0, err
        got = [[int(x) for x in l.split()[1:]] for l in out.splitlines() if l.startswith("M ")]
        segs = jc.segments(data)
        kinds = {0xdb: 0xdb, 0xc0: 0xc0, 0xc2: 0xc0, 0xc4: 0xc4, 0xda: 0xda, 0xdd: 0xdd, 0xd9: 0xd9}
        assert len(got) == len(segs)
        for g, (m, off, n) in zip(got, segs):
            t = -2 if 0xe0 <= m < 0xf0 else kinds.get(m, -1)
            assert g[0] == t and g[1] == m and g[2] == n + 2 and g[3] == data[off]
            assert g[4] == (off + n if m == 0xda else 0)
    com = tmp_path / "com.jpg"
    good = harness.synth_jpeg(6, 64, 48, 80, 2, 2, False, 0)
    com.write_bytes(good[:2] + b"\xff\xfe\x00\x04hi" + good[2:])
    rc, out, err = run([probe, str(com), "markers"])
    assert rc == 0 and [l.split()[1] for l in out.splitlines() if l.startswith("M ")][-1] == "-1"   # COM ends parsing, file.hpp:252-253


@pytest.mark.gpu
def test_cli_many_files_through_the_pinned_feed(cli, oracle_mod, harness, tmp_path):
    # b200jpeg::decode_files: files read into one pinned arena by host threads, headers parsed on host threads, one
    # b200jpeg_decode_batch_multi call; --gpus takes as many devices as the box has (1 here)
    import torch
    n = 40
    datas = [harness.synth_jpeg(1500 + i, 320 + 16 * (i % 3), 200, 88, 2, 2, False, -1 if i % 2 else 0) for i in range(n)]
    paths = []
    for i, d in enumerate(datas):
        p = tmp_path / ("f%03d.jpg" % i)
        p.write_bytes(d)
        paths.append(str(p))
    (tmp_path / "broken.jpg").write_bytes(b"\xff\xd8\xff\xd9")
    gpus = min(2, torch.cuda.device_count())
    rc, out, err = run([cli] + paths + [str(tmp_path / "broken.jpg"), "--out-dir", str(tmp_path), "--gpus", str(gpus)])
    assert rc == 1 and "broken.jpg: not decodable" in err and "%d of %d files decoded" % (n, n + 1) in out, (out, err)
    for i in (0, 19, 20, 39):
        _, _, _, orgb = oracle_mod.port_decode(datas[i])
        raw = open(tmp_path / ("f%03d.ppm" % i), "rb").read()
        head = b"P6\n%d %d\n255\n" % (orgb.shape[1], orgb.shape[0])
        assert raw.startswith(head) and np.array_equal(np.frombuffer(raw[len(head):], np.uint8).reshape(orgb.shape), orgb), i


@pytest.mark.gpu
def test_cli_spec_mode_lifts_the_progressive_refusal(cli, oracle_mod, tmp_path):
    # `b200jpeg file.jpg` refuses a progressive file like the reference (jpeg.cpp:69-73); `--spec` decodes it
    src = os.path.join(GOLDEN_DIR, "prograssive-sample-2.jpg")
    rc, out, err = run([cli, src, "--out-dir", str(tmp_path)])
    assert rc == 1 and "not decodable, b200jpeg error -11" in err
    rc, out, err = run([cli, src, os.path.join(GOLDEN_DIR, "img2.jpg"), "--out-dir", str(tmp_path), "--spec"])
    assert rc == 0, err
    _, _, _, orgb = oracle_mod.port_decode_progressive(open(src, "rb").read())
    raw = open(tmp_path / "prograssive-sample-2.ppm", "rb").read()
    head = b"P6\n%d %d\n255\n" % (orgb.shape[1], orgb.shape[0])
    assert raw.startswith(head) and np.array_equal(np.frombuffer(raw[len(head):], np.uint8).reshape(orgb.shape), orgb)
