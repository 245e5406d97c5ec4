"""TEST INFRASTRUCTURE: byte-level surgery on baseline JPEG files, for the parser / decoder edge cases no encoder
emits: arbitrary component identifiers (T.81 B.2.2), 16-bit quantisation tables (Pq = 1, B.2.4.1), restart
segments padded with 0-bits instead of 1-bits (F.1.2.3)."""
import struct

import numpy as np


def segments(data):
    """[(marker, payload offset, payload length)] of the header segments up to and including SOS."""
    out, pos = [], 2
    while pos < len(data):
        assert data[pos] == 0xFF, pos
        m = data[pos + 1]
        pos += 2
        if m == 0xFF:
            pos -= 1
            continue
        n = struct.unpack(">H", data[pos:pos + 2])[0]
        out.append((m, pos + 2, n - 2))
        pos += n
        if m == 0xDA:
            break
    return out


def with_component_ids(data, ids):
    """The same picture with the frame's component identifiers (and the scan's selectors) renamed."""
    d = bytearray(data)
    for m, off, n in segments(data):
        if m == 0xC0:
            for i in range(d[off + 5]):
                d[off + 6 + 3 * i] = ids[i]
        if m == 0xDA:
            for i in range(d[off]):
                d[off + 1 + 2 * i] = ids[i]
    return bytes(d)


def with_wide_dqt(data, scale=None):
    """Every DQT rewritten with 16-bit entries (Pq = 1).  scale(table id, zig-zag index, value) -> new value lets a
    test push entries above 255."""
    out = bytearray(data[:2])
    pos = 2
    for m, off, n in segments(data):
        seg = data[off:off + n]
        if m == 0xDB:
            new = bytearray()
            p = 0
            while p < len(seg):
                pq, tq = seg[p] >> 4, seg[p] & 15
                assert pq == 0
                new.append(0x10 | tq)
                for i in range(64):
                    v = seg[p + 1 + i]
                    if scale:
                        v = scale(tq, i, v)
                    new += struct.pack(">H", v)
                p += 65
            seg = bytes(new)
        out += bytes([0xFF, m]) + struct.pack(">H", len(seg) + 2) + seg
        pos = off + n
    return bytes(out) + data[pos:]


def restart_markers(data):
    """Offsets of the FF Dn markers inside the scan."""
    segs = segments(data)
    m, off, n = segs[-1]
    pos, out = off + n, []
    while pos + 1 < len(data):
        if data[pos] == 0xFF and 0xD0 <= data[pos + 1] <= 0xD7:
            out.append(pos)
            pos += 2
        else:
            pos += 1
    return out


def with_zero_padding(data, same_decode):
    """Clears the padding bits in the last byte of every restart segment (as many low bits of the byte before each
    RSTn as can be cleared without changing what `same_decode(candidate)` -- a block-counting decoder such as the
    reference, jpeg.cpp:419-425 -- makes of the file).  Returns (new file, number of bits cleared)."""
    d = bytearray(data)
    cleared = 0
    for at in restart_markers(data):
        b = at - 1
        if d[b] == 0x00 and d[b - 1] == 0xFF:      # a stuffed FF: leave it
            continue
        best = 0
        for k in range(1, 8):
            if (d[b] & ((1 << k) - 1)) != (1 << k) - 1:
                break                              # padding is 1-bits: the run of low 1-bits bounds it
            trial = bytearray(d)
            trial[b] &= ~((1 << k) - 1) & 0xFF
            if same_decode(bytes(trial)):
                best = k
        if best:
            d[b] &= ~((1 << best) - 1) & 0xFF
            cleared += best
    return bytes(d), cleared


# ---------------------------------------------------------------------------------------------------------------
# Huffman transcoding: the same coefficients under other code lengths (what no encoder here emits: AC tables whose
# long codes spread over more 10-bit prefixes than a decoder's second table level provides for)
def _canon(bits, vals):
    """(code, length) per symbol, and {(length, code): symbol}: the canonical assignment of T.81 Annex C."""
    enc, dec, code, k = {}, {}, 0, 0
    for ln in range(1, 17):
        for _ in range(bits[ln - 1]):
            enc[vals[k]] = (code, ln)
            dec[(ln, code)] = vals[k]
            code += 1
            k += 1
        code <<= 1
    return enc, dec


class _Bits:
    def __init__(self, raw):
        self.raw, self.pos = raw, 0

    def get(self, n):
        v = 0
        for _ in range(n):
            byte = self.raw[self.pos >> 3] if (self.pos >> 3) < len(self.raw) else 0xFF
            v = (v << 1) | ((byte >> (7 - (self.pos & 7))) & 1)
            self.pos += 1
        return v

    def symbol(self, dec):
        code = 0
        for ln in range(1, 17):
            code = (code << 1) | self.get(1)
            if (ln, code) in dec:
                return dec[(ln, code)]
        raise ValueError("no such code")


class _Out:
    def __init__(self):
        self.acc, self.n, self.out = 0, 0, bytearray()

    def put(self, v, n):
        self.acc = (self.acc << n) | (v & ((1 << n) - 1))
        self.n += n
        while self.n >= 8:
            b = (self.acc >> (self.n - 8)) & 0xFF
            self.out.append(b)
            if b == 0xFF:
                self.out.append(0)
            self.n -= 8
        self.acc &= (1 << self.n) - 1

    def align(self):
        if self.n:
            self.put((1 << (8 - self.n)) - 1, 8 - self.n)


def with_ac_code_lengths(data, bits, dc_bits=None):
    """The same picture with every AC table's code LENGTHS replaced by `bits` (16 counts; sum = the table's symbol
    count, symbols keep their order) and the scan transcoded to the new codes; dc_bits: the same for every DC table.
    Baseline, one scan, RSTn allowed."""
    segs = segments(data)
    tabs, frame, scan, ri = {}, None, None, 0
    for m, off, n in segs:
        seg = data[off:off + n]
        if m == 0xC4:
            p = 0
            while p < len(seg):
                tcth, b = seg[p], list(seg[p + 1:p + 17])
                cnt = sum(b)
                tabs[tcth] = (b, list(seg[p + 17:p + 17 + cnt]))
                p += 17 + cnt
        elif m == 0xC0:
            h, w, nc = struct.unpack(">HH", seg[1:5]) + (seg[5],)
            frame = (w, h, [(seg[6 + 3 * i], seg[7 + 3 * i] >> 4, seg[7 + 3 * i] & 15) for i in range(nc)])
        elif m == 0xDA:
            scan = [(seg[1 + 2 * i], seg[2 + 2 * i] >> 4, seg[2 + 2 * i] & 15) for i in range(seg[0])]
        elif m == 0xDD:
            ri = struct.unpack(">H", seg[:2])[0]
    w, h, comps = frame
    hmax, vmax = max(c[1] for c in comps), max(c[2] for c in comps)
    nmcu = ((w + 8 * hmax - 1) // (8 * hmax)) * ((h + 8 * vmax - 1) // (8 * vmax))
    blocks = []                                     # (dc table, ac table) per block of an MCU
    for (cid, hs, vs), (sid, td, ta) in zip(comps, scan):
        assert cid == sid
        blocks += [(td, 0x10 | ta)] * (hs * vs if len(comps) > 1 else 1)
    new_tabs = {k: ((list(bits), v[1]) if k >> 4 else ((list(dc_bits), v[1]) if dc_bits else v)) for k, v in tabs.items()}
    for k, (b, v) in new_tabs.items():
        assert sum(b) == len(v), (k, sum(b), len(v))
    dec = {k: _canon(*v)[1] for k, v in tabs.items()}
    enc = {k: _canon(*v)[0] for k, v in new_tabs.items()}
    # the scan, cut at the restart markers
    sos_m, sos_off, sos_n = segs[-1]
    body = data[sos_off + sos_n:data.rindex(b"\xff\xd9")]
    pieces, cur, i = [], bytearray(), 0
    while i < len(body):
        if body[i] == 0xFF and i + 1 < len(body) and body[i + 1] == 0x00:
            cur.append(0xFF); i += 2
        elif body[i] == 0xFF and i + 1 < len(body) and 0xD0 <= body[i + 1] <= 0xD7:
            pieces.append(bytes(cur)); cur = bytearray(); i += 2
        else:
            cur.append(body[i]); i += 1
    pieces.append(bytes(cur))
    out, done = _Out(), 0
    for k, piece in enumerate(pieces):
        rd = _Bits(piece)
        for _ in range(min(ri, nmcu - done) if ri else nmcu):
            for td, ta in blocks:
                s = rd.symbol(dec[td])
                out.put(*enc[td][s]); out.put(rd.get(s), s)
                z = 1
                while z < 64:
                    rs = rd.symbol(dec[ta])
                    out.put(*enc[ta][rs])
                    if rs & 15:
                        out.put(rd.get(rs & 15), rs & 15)
                        z += (rs >> 4) + 1
                    elif rs == 0xF0:
                        z += 16
                    else:
                        break
            done += 1
        out.align()
        if k + 1 < len(pieces):
            out.out += bytes([0xFF, 0xD0 + (k & 7)])
    head, dht_done = bytearray(data[:2]), False
    for m, off, n in segs:
        seg = data[off:off + n]
        if m == 0xC4:
            if dht_done:
                continue
            seg, dht_done = b"".join(bytes([k]) + bytes(b) + bytes(v) for k, (b, v) in new_tabs.items()), True   # all tables in the first DHT
        head += bytes([0xFF, m]) + struct.pack(">H", len(seg) + 2) + seg
    return bytes(head) + bytes(out.out) + b"\xff\xd9"


# 22 codes of 2..8 bits, 140 codes of 12 bits: the long codes cover 35 ten-bit prefixes
OVERFLOW_AC_BITS = [0, 1, 1, 2, 2, 4, 4, 8, 0, 0, 0, 140, 0, 0, 0, 0]


def damaged_progressive(data):
    """Two damaged copies of a progressive file with restart markers (what scripts/dev/soak_progressive.py found):
    (zeros over a stretch of the LAST scan that holds one RSTn -- the reference's logic reads on over it and so must we,
     the file cut in the middle of the last scan -- the scan ends before its blocks do)."""
    sos = [i for i in range(len(data) - 1) if data[i] == 0xFF and data[i + 1] == 0xDA][-1]
    marks = [i for i in range(sos + 14, len(data) - 2) if data[i] == 0xFF and 0xD0 <= data[i + 1] <= 0xD7]
    m = marks[len(marks) // 2]
    zeroed = bytearray(data)
    zeroed[m - 60:m + 60] = bytes(120)
    cut = data[:m - 30] + b"\xff\xd9"
    return bytes(zeroed), cut


def random_code_lengths(rng, nsym=162):
    """16 counts of a random prefix code over nsym symbols: lengths of a Huffman tree over random (skewed) weights, one
    extra symbol reserving the all-ones code (T.81 Annex K.2), no code longer than 16 bits."""
    import heapq
    while True:
        lam = float(rng.uniform(0.01, 0.35)) * 162.0 / nsym    # geometric weights: from nearly flat to 1-bit-to-16-bit codes
        w = list(np.exp(-lam * np.arange(nsym + 1)) * (0.5 + rng.random(nsym + 1)))
        w[-1] = 1e-12                                  # the reserved code point: the lightest, so it ends up all ones
        heap = [(x, i, (i,)) for i, x in enumerate(w)]
        heapq.heapify(heap)
        depth = [0] * (nsym + 1)
        nxt = nsym + 1
        while len(heap) > 1:
            a = heapq.heappop(heap); b = heapq.heappop(heap)
            for i in a[2] + b[2]:
                depth[i] += 1
            heapq.heappush(heap, (a[0] + b[0], nxt, a[2] + b[2])); nxt += 1
        if max(depth) <= 16:
            bits = [0] * 16
            for d in sorted(depth[:-1]):
                bits[d - 1] += 1
            return bits
