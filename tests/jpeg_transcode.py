"""Test-side baseline JPEG re-encoder: takes the quantised coefficient planes of a decoded file
and writes them back as a baseline JPEG with a chosen SCAN STRUCTURE (non-interleaved scans, mixed
scans, per-scan restart intervals, DHT segments repeated in front of every SOS or not).  Pillow
and stb can only write one interleaved scan; BASELINE config 3 asks for non-interleaved ones
(the role SURVEY.md gives to "tools/jpegsynth").  Pure Python + numpy, small images only.

The coefficient values are known exactly (they are the input), so a decoder under test has a
known answer that does not depend on any other decoder.
"""
import numpy as np

ZIGZAG = np.array([0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48,
                   41, 34, 27, 20, 13, 6, 7, 14, 21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15,
                   23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53, 60, 61, 54, 47, 55, 62,
                   63])


def parse_tables(data: bytes):
    """DQT / DHT / SOF0 of an existing baseline file."""
    i = 2
    dqt, dht, sof = {}, {}, None
    while i < len(data):
        assert data[i] == 0xFF
        m = data[i + 1]
        if m == 0xD9:
            break
        length = (data[i + 2] << 8) | data[i + 3]
        seg = data[i + 4:i + 2 + length]
        if m == 0xDB:
            j = 0
            while j < len(seg):
                assert seg[j] >> 4 == 0
                dqt[seg[j] & 15] = bytes(seg[j + 1:j + 65])
                j += 65
        elif m == 0xC4:
            j = 0
            while j < len(seg):
                tc, th = seg[j] >> 4, seg[j] & 15
                counts = list(seg[j + 1:j + 17])
                n = sum(counts)
                dht[(tc, th)] = (counts, list(seg[j + 17:j + 17 + n]))
                j += 17 + n
        elif m == 0xC0:
            nc = seg[5]
            sof = dict(height=(seg[1] << 8) | seg[2], width=(seg[3] << 8) | seg[4],
                       comps=[dict(id=seg[6 + 3 * k], h=seg[7 + 3 * k] >> 4, v=seg[7 + 3 * k] & 15,
                                   tq=seg[8 + 3 * k]) for k in range(nc)])
        elif m == 0xDA:
            break
        i += 2 + length
    return dqt, dht, sof


def code_table(counts, vals):
    """symbol -> (code, length), canonical (T.81 Annex C)."""
    out, code, k = {}, 0, 0
    for length in range(1, 17):
        for _ in range(counts[length - 1]):
            out[vals[k]] = (code, length)
            code += 1
            k += 1
        code <<= 1
    return out


class BitWriter:
    def __init__(self):
        self.out = bytearray()
        self.acc = 0
        self.n = 0

    def put(self, value, length):
        self.acc = (self.acc << length) | (value & ((1 << length) - 1))
        self.n += length
        while self.n >= 8:
            b = (self.acc >> (self.n - 8)) & 0xFF
            self.out.append(b)
            if b == 0xFF:
                self.out.append(0)
            self.n -= 8
        self.acc &= (1 << self.n) - 1

    def flush(self):
        if self.n:
            self.put((1 << (8 - self.n)) - 1, 8 - self.n)

    def marker(self, m):
        self.flush()
        self.out += bytes([0xFF, m])


def _encode_block(bw, block_zz, pred, dc_codes, ac_codes):
    diff = int(block_zz[0]) - pred
    cat = abs(diff).bit_length()
    bw.put(*dc_codes[cat])
    if cat:
        bw.put(diff if diff >= 0 else diff + (1 << cat) - 1, cat)
    run = 0
    for k in range(1, 64):
        v = int(block_zz[k])
        if v == 0:
            run += 1
            continue
        while run > 15:
            bw.put(*ac_codes[0xF0])
            run -= 16
        size = abs(v).bit_length()
        bw.put(*ac_codes[(run << 4) | size])
        bw.put(v if v >= 0 else v + (1 << size) - 1, size)
        run = 0
    if run:
        bw.put(*ac_codes[0x00])
    return int(block_zz[0])


def _blocks(plane, by, bx):
    h, w = plane.shape
    if by * 8 >= h or bx * 8 >= w:
        return np.zeros(64, dtype=np.int32)  # padding block outside the stored plane
    return plane[by * 8:by * 8 + 8, bx * 8:bx * 8 + 8].reshape(64)[ZIGZAG]


def transcode(data: bytes, planes, scans, restart_interval=0, dht_every_scan=True,
              table_map=None):
    """data: source baseline JPEG (its DQT/DHT/SOF0 are reused); planes[c]: int16 quantised
    coefficient plane of frame component c (raster of natural-order 8x8 tiles, DC absolute), at
    least ceil(size/8) blocks in each direction; scans: list of lists of frame component indices,
    e.g. [[0], [1], [2]] or [[0], [1, 2]].  table_map[c] = (dc table id, ac table id); default
    0/0 for component 0 and 1/1 for the others (falls back to 0/0 if the source has one pair).
    restart_interval in MCUs of each scan (0 = none), or a list with one value per scan: a DRI
    segment is then written in front of every scan whose interval differs from the one in force
    (T.81 B.2.4.4)."""
    dqt, dht, sof = parse_tables(data)
    comps = sof["comps"]
    W, H = sof["width"], sof["height"]
    hmax = max(c["h"] for c in comps)
    vmax = max(c["v"] for c in comps)
    have_second = (0, 1) in dht and (1, 1) in dht
    if table_map is None:
        table_map = [(0, 0) if (c == 0 or not have_second) else (1, 1) for c in range(len(comps))]
    out = bytearray(b"\xff\xd8")
    for tq, tab in sorted(dqt.items()):
        out += b"\xff\xdb" + (67).to_bytes(2, "big") + bytes([tq]) + tab
    sof_body = bytes([8]) + H.to_bytes(2, "big") + W.to_bytes(2, "big") + bytes([len(comps)])
    for c in comps:
        sof_body += bytes([c["id"], (c["h"] << 4) | c["v"], c["tq"]])
    out += b"\xff\xc0" + (len(sof_body) + 2).to_bytes(2, "big") + sof_body
    per_scan = list(restart_interval) if isinstance(restart_interval, (list, tuple)) else [restart_interval] * len(scans)
    in_force = 0

    def dht_segment():
        body = bytearray()
        for (tc, th), (counts, vals) in sorted(dht.items()):
            body += bytes([(tc << 4) | th]) + bytes(counts) + bytes(vals)
        return b"\xff\xc4" + (len(body) + 2).to_bytes(2, "big") + bytes(body)

    codes = {k: code_table(*v) for k, v in dht.items()}
    for si, scan in enumerate(scans):
        if dht_every_scan or si == 0:
            out += dht_segment()
        restart_interval = per_scan[si]
        if restart_interval != in_force:
            out += b"\xff\xdd\x00\x04" + restart_interval.to_bytes(2, "big")
            in_force = restart_interval
        sos = bytes([len(scan)])
        for c in scan:
            sos += bytes([comps[c]["id"], (table_map[c][0] << 4) | table_map[c][1]])
        sos += bytes([0, 63, 0])
        out += b"\xff\xda" + (len(sos) + 2).to_bytes(2, "big") + sos
        bw = BitWriter()
        pred = {c: 0 for c in scan}
        if len(scan) == 1:  # T.81 A.2.3: one data unit per MCU, ceil(size/8) blocks per direction
            c = scan[0]
            cw = -(-W * comps[c]["h"] // hmax)
            ch = -(-H * comps[c]["v"] // vmax)
            mx, my = -(-cw // 8), -(-ch // 8)
            units = [[(c, by, bx)] for by in range(my) for bx in range(mx)]
        else:
            mx, my = -(-W // (8 * hmax)), -(-H // (8 * vmax))
            units = []
            for y in range(my):
                for x in range(mx):
                    mcu = []
                    for c in scan:
                        for iy in range(comps[c]["v"]):
                            for ix in range(comps[c]["h"]):
                                mcu.append((c, y * comps[c]["v"] + iy, x * comps[c]["h"] + ix))
                    units.append(mcu)
        rst = 0
        for n, mcu in enumerate(units):
            if restart_interval and n and n % restart_interval == 0:
                bw.marker(0xD0 + (rst & 7))
                rst += 1
                pred = {c: 0 for c in scan}
            for (c, by, bx) in mcu:
                pred[c] = _encode_block(bw, _blocks(planes[c], by, bx), pred[c],
                                        codes[(0, table_map[c][0])], codes[(1, table_map[c][1])])
        bw.flush()
        out += bw.out
    out += b"\xff\xd9"
    return bytes(out)
