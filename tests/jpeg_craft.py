"""TEST INFRASTRUCTURE: byte-level surgery on baseline JPEG files, for the parser / decoder edge cases no encoder
emits: arbitrary component identifiers (T.81 B.2.2), 16-bit quantisation tables (Pq = 1, B.2.4.1), restart
segments padded with 0-bits instead of 1-bits (F.1.2.3)."""
import struct


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
