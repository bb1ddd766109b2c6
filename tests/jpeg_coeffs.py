"""TEST INFRASTRUCTURE: a small baseline (SOF0, Huffman, 8-bit) JPEG entropy decoder that returns what libjpeg's
jpeg_read_coefficients() would: per component the quantised DCT coefficients of every block (natural order, whole
MCUs) and its quantisation table. Row N4's kernels start from there (the entropy decoder stays on the host).
Follows ITU-T T.81 (Annex F.2.2 decoding procedures, Figure A.6 zig-zag)."""
import numpy as np

ZIGZAG = np.array([0, 1, 8, 16, 9, 2, 3, 10, 17, 24, 32, 25, 18, 11, 4, 5, 12, 19, 26, 33, 40, 48, 41, 34, 27, 20, 13, 6, 7, 14,
                   21, 28, 35, 42, 49, 56, 57, 50, 43, 36, 29, 22, 15, 23, 30, 37, 44, 51, 58, 59, 52, 45, 38, 31, 39, 46, 53,
                   60, 61, 54, 47, 55, 62, 63])


class Component:
    def __init__(self, cid, h, v, tq):
        self.cid, self.h, self.v, self.tq = cid, h, v, tq
        self.td = self.ta = 0
        self.coeffs = None  # int16 [blocks_h, blocks_w, 64], natural order
        self.quant = None   # uint16 [64], natural order
        self.width = self.height = 0  # visible samples


def _huff_lut(bits, vals):
    """16-bit peek -> (code length, symbol); T.81 Annex C code assignment."""
    length = np.zeros(65536, np.uint8)
    symbol = np.zeros(65536, np.uint8)
    code, k = 0, 0
    for ln in range(1, 17):
        for _ in range(bits[ln - 1]):
            lo, hi = code << (16 - ln), (code + 1) << (16 - ln)
            length[lo:hi] = ln
            symbol[lo:hi] = vals[k]
            code += 1
            k += 1
        code <<= 1
    return length, symbol


def decode(data):
    """-> (width, height, [Component]) for a baseline JPEG byte string."""
    data = bytes(data)
    assert data[:2] == b"\xff\xd8", "not a JPEG"
    pos, qt, dc_t, ac_t, comps, frame, ri = 2, {}, {}, {}, [], None, 0
    while True:
        assert data[pos] == 0xFF
        m = data[pos + 1]
        pos += 2
        if m == 0xD8 or 0xD0 <= m <= 0xD7 or m == 0x01:
            continue
        ln = (data[pos] << 8) | data[pos + 1]
        seg = data[pos + 2: pos + ln]
        pos += ln
        if m == 0xDB:  # DQT
            i = 0
            while i < len(seg):
                pq, tq = seg[i] >> 4, seg[i] & 15
                i += 1
                if pq:
                    t = np.frombuffer(seg[i:i + 128], dtype=">u2").astype(np.uint16)
                    i += 128
                else:
                    t = np.frombuffer(seg[i:i + 64], dtype=np.uint8).astype(np.uint16)
                    i += 64
                nat = np.zeros(64, np.uint16)
                nat[ZIGZAG] = t
                qt[tq] = nat
        elif m == 0xC0 or m == 0xC1:  # SOF0 / SOF1 (sequential Huffman)
            assert seg[0] == 8, "8-bit samples only"
            frame = ((seg[1] << 8) | seg[2], (seg[3] << 8) | seg[4])  # height, width
            for c in range(seg[5]):
                cid, hv, tq = seg[6 + 3 * c: 9 + 3 * c]
                comps.append(Component(cid, hv >> 4, hv & 15, tq))
        elif m == 0xC2:
            raise ValueError("progressive JPEG: not handled by this test decoder")
        elif m == 0xC4:  # DHT
            i = 0
            while i < len(seg):
                tc, th = seg[i] >> 4, seg[i] & 15
                bits = list(seg[i + 1:i + 17])
                n = sum(bits)
                vals = list(seg[i + 17:i + 17 + n])
                i += 17 + n
                (ac_t if tc else dc_t)[th] = _huff_lut(bits, vals)
        elif m == 0xDD:
            ri = (seg[0] << 8) | seg[1]
        elif m == 0xDA:  # SOS
            ns = seg[0]
            for k in range(ns):
                cs, t = seg[1 + 2 * k], seg[2 + 2 * k]
                c = next(c for c in comps if c.cid == cs)
                c.td, c.ta = t >> 4, t & 15
            assert ns == len(comps), "non-interleaved scans are not handled by this test decoder"
            break
    height, width = frame
    hmax, vmax = max(c.h for c in comps), max(c.v for c in comps)
    mcus_x, mcus_y = -(-width // (8 * hmax)), -(-height // (8 * vmax))
    for c in comps:
        c.coeffs = np.zeros((mcus_y * c.v, mcus_x * c.h, 64), np.int16)
        c.quant = qt[c.tq]
        c.width, c.height = -(-width * c.h // hmax), -(-height * c.v // vmax)
    # ---- entropy-coded segment: remove byte stuffing, split at restart markers
    scan = bytearray()
    segments = []
    i = pos
    while True:
        b = data[i]
        if b != 0xFF:
            scan.append(b)
            i += 1
            continue
        nb = data[i + 1]
        if nb == 0x00:
            scan.append(0xFF)
            i += 2
        elif 0xD0 <= nb <= 0xD7:
            segments.append(bytes(scan))
            scan = bytearray()
            i += 2
        else:
            segments.append(bytes(scan))
            break
    mcu = 0
    for segb in segments:
        bits = np.unpackbits(np.frombuffer(segb + b"\xff\xff\xff\xff", dtype=np.uint8)).astype(np.uint32)
        n = bits.size - 16
        win = np.zeros(n, np.uint32)
        for k in range(16):
            win |= bits[k:k + n] << (15 - k)
        win = win.tolist()
        p = 0
        pred = [0] * len(comps)
        count = ri if ri else mcus_x * mcus_y
        luts = [(dc_t[c.td][0].tolist(), dc_t[c.td][1].tolist(), ac_t[c.ta][0].tolist(), ac_t[c.ta][1].tolist()) for c in comps]

        def receive(s, p):
            v = win[p] >> (16 - s)
            return v if v >= (1 << (s - 1)) else v - (1 << s) + 1

        for _ in range(count):
            if mcu >= mcus_x * mcus_y:
                break
            my, mx = divmod(mcu, mcus_x)
            for ci, c in enumerate(comps):
                dl, dsym, al, asym = luts[ci]
                for by in range(c.v):
                    for bx in range(c.h):
                        blk = [0] * 64
                        w = win[p]
                        s = dsym[w]
                        p += dl[w]
                        if s:
                            pred[ci] += receive(s, p)
                            p += s
                        blk[0] = pred[ci]
                        k = 1
                        while k < 64:
                            w = win[p]
                            rs = asym[w]
                            p += al[w]
                            r, s = rs >> 4, rs & 15
                            if s == 0:
                                if r != 15:
                                    break
                                k += 16
                                continue
                            k += r
                            blk[int(ZIGZAG[k])] = receive(s, p)
                            p += s
                            k += 1
                        c.coeffs[my * c.v + by, mx * c.h + bx] = blk
            mcu += 1
    return width, height, comps
