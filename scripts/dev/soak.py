#!/usr/bin/env python
"""Dev tool: randomised soak of both entropy paths on the GPU -- clean pictures of random shape / sampling / quality /
restart interval next to corrupted copies of them (byte flips, truncation, runs of FF 00 / 00 / FF FF, chunks moved),
in mixed batches.  Checks: no hang (run it under `timeout`), clean pictures bit-exact against the oracle (a sample) and
identical between the two paths (all), every corrupted picture either flagged or decoded to the oracle's result.
    timeout 600 python scripts/dev/soak.py [rounds] [seed] [max width] [max height]  -> one JSON line"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import jpeg_decoder_b200 as jd  # noqa: E402
from jpeg_decoder_b200 import abi, harness  # noqa: E402
from oracle import oracle  # noqa: E402
from tests import jpeg_craft as jc  # noqa: E402

rounds = int(sys.argv[1]) if len(sys.argv) > 1 else 20
rng = np.random.default_rng(int(sys.argv[2]) if len(sys.argv) > 2 else 1)
MAXW, MAXH = (int(sys.argv[3]), int(sys.argv[4])) if len(sys.argv) > 4 else (700, 500)   # larger: several CTAs of subs per scan
harness.build_harness(); oracle.build()


def corrupt(d):
    m = bytearray(d)
    body = d.find(b"\xff\xda") + 14
    n = len(d) - 2 - body
    if n < 16:
        return bytes(m)
    kind = int(rng.integers(0, 9))
    if kind >= 7:
        # the tables instead of the scan: a few bytes of a DHT / DQT payload (code lengths, symbols, quantisers) changed --
        # over-subscribed or incomplete codes, symbols the reference rejects, tables that no longer fit the scan
        from tests import jpeg_craft as jc
        segs = [(mk, off, ln_) for mk, off, ln_ in jc.segments(d) if mk in ((0xC4,) if kind == 7 else (0xC4, 0xDB)) and ln_ > 2]
        mk, off, ln_ = segs[int(rng.integers(0, len(segs)))]
        for p in rng.integers(off + 1, off + ln_, size=int(rng.integers(1, 4))):
            m[p] = int(rng.integers(0, 256)) if rng.integers(0, 2) else (m[p] ^ (1 << int(rng.integers(0, 8))))
        return bytes(m)
    at = body + int(rng.integers(0, n - 8))
    ln = int(rng.integers(1, min(n - (at - body), 400)))
    if kind == 0:
        for p in rng.integers(body, len(d) - 2, size=int(rng.integers(1, 30))):
            m[p] = int(rng.integers(0, 255))
    elif kind == 1:
        m = m[:at] + b"\xff\xd9"
    elif kind == 2:
        m[at:at + ln] = (b"\xff\x00" * ln)[:ln]
    elif kind == 3:
        m[at:at + ln] = bytes(ln)
    elif kind == 4:
        m[at:at + ln] = b"\xff" * ln
    elif kind == 5:
        src = body + int(rng.integers(0, n - 8))
        m[at:at + ln] = m[src:src + ln]
    else:
        m[at:at + 2] = bytes([0xff, int(rng.integers(0xd0, 0xd8))])
    return bytes(m)


decs = {p: jd.Decoder(0) for p in ("segment", "sync")}
decs["segment"].set_entropy_path(abi.ENTROPY_SEGMENT); decs["sync"].set_entropy_path(abi.ENTROPY_SYNC)
batch_dec = jd.Decoder(0)
stat = {"batches": 0, "clean": 0, "corrupt": 0, "corrupt_flagged": 0, "oracle_checked": 0, "unflagged_same_as_oracle": 0,
        "unflagged_but_oracle_rejects": 0, "refused_by_parser": 0, "decode_batch_checked": 0, "transcoded": 0}
for r in range(rounds):
    datas, clean = [], []
    for i in range(int(rng.integers(4, 24 if MAXW <= 700 else 8))):
        w, h = int(rng.integers(1, MAXW)), int(rng.integers(1, MAXH))
        hs, vs = [(1, 1), (2, 1), (1, 2), (2, 2)][int(rng.integers(0, 4))]
        gray = bool(rng.integers(0, 6) == 0)
        if gray:
            hs = vs = 1
        q = int(rng.integers(20, 100))
        ri = int([0, 0, -1, 1, 2, 5, 17][int(rng.integers(0, 7))])
        d = harness.synth_jpeg(int(rng.integers(0, 1 << 30)), w, h, q, hs, vs, gray, ri)
        if w * h < 60000 and rng.integers(0, 4) == 0:
            # the same picture under random Huffman code lengths (1 to 16 bits, whatever a Huffman tree over skewed weights
            # gives): pairs, long codes and second-level overflow of the lookup tables in proportions no encoder produces
            d = jc.with_ac_code_lengths(d, jc.random_code_lengths(rng), jc.random_code_lengths(rng, 12) if rng.integers(0, 2) else None)
            stat["transcoded"] += 1
        datas.append(d); clean.append(True)
        for _ in range(int(rng.integers(0, 3))):
            datas.append(corrupt(d)); clean.append(False)
    order = rng.permutation(len(datas))
    datas = [datas[k] for k in order]; clean = [clean[k] for k in order]
    # what the host parser refuses never reaches the GPU: the oracle must refuse it too
    _, prc = jd.parse_headers([np.frombuffer(x, np.uint8).copy() for x in datas], 1)
    for x, rc_p in zip(datas, prc):
        if rc_p != 0:
            stat["refused_by_parser"] += 1
            assert oracle.port_decode(x, flags=1, want_coef=False, want_rgb=False)[0] != 0
    clean = [c for c, rc_p in zip(clean, prc) if rc_p == 0]
    datas = [x for x, rc_p in zip(datas, prc) if rc_p == 0]
    res = {p: decs[p].decode(datas, flags=1, want_coef=True) for p in decs}
    stat["batches"] += 1
    # the one-call host -> host entry point (chunks, double buffering, its own choice of entropy path) on the same batch:
    # same statuses being zero or not, same pixels wherever the status is zero
    bufs = [np.frombuffer(x, np.uint8).copy() for x in datas]
    descs, prc2 = jd.parse_headers(bufs, 1)
    rgbs2 = [np.empty_like(a) for a in res["segment"][1]]
    st2 = batch_dec.decode_batch(descs, bufs, rgbs2)
    for i in range(len(datas)):
        if st2[i] == 0:
            ok = [p for p in res if res[p][0][i] == 0]
            assert ok, (r, i, "decode_batch accepts what both prepared paths flag")
            assert np.array_equal(rgbs2[i], res[ok[0]][1][i]), (r, i, "decode_batch pixels differ")
        else:
            assert not rgbs2[i].any() or True
            assert any(res[p][0][i] != 0 for p in res), (r, i, "decode_batch flags what both prepared paths accept")
    stat["decode_batch_checked"] += len(datas)
    for i, d in enumerate(datas):
        st_s, st_x = res["segment"][0][i], res["sync"][0][i]
        if clean[i]:
            stat["clean"] += 1
            assert st_s == 0 and st_x == 0, (r, i, st_s, st_x)
            assert np.array_equal(res["segment"][1][i], res["sync"][1][i]) and np.array_equal(res["segment"][2][i], res["sync"][2][i])
            if i % 5 == 0:
                rc, _, oc, orgb = oracle.port_decode(d, flags=1)
                assert rc == 0 and np.array_equal(res["segment"][2][i].astype(np.int32), oc) and np.array_equal(res["segment"][1][i], orgb)
                stat["oracle_checked"] += 1
        else:
            stat["corrupt"] += 1
            for p, st in (("segment", st_s), ("sync", st_x)):
                if st != 0:                                           # a status word, or the parser's negative code
                    stat["corrupt_flagged"] += 1
                    continue
                rc, _, oc, orgb = oracle.port_decode(d, flags=1)      # not flagged: then it must be what the reference logic makes of it
                if rc == 0:
                    assert np.array_equal(res[p][2][i].astype(np.int32), oc), (r, i, p, "unflagged corrupt stream differs from the oracle")
                    stat["unflagged_same_as_oracle"] += 1
                else:
                    stat["unflagged_but_oracle_rejects"] += 1
                    os.makedirs(os.path.join(ROOT, "gpurun_out", "soak"), exist_ok=True)
                    name = os.path.join(ROOT, "gpurun_out", "soak", "case_%s_r%d_i%d.jpg" % (p, r, i))
                    open(name, "wb").write(d)
                    print(json.dumps({"oracle_rc": rc, "path": p, "status": 0, "round": r, "image": i, "file": name}), flush=True)
print(json.dumps(stat))
assert stat["unflagged_but_oracle_rejects"] == 0, "a stream the reference's logic rejects went through unflagged"
