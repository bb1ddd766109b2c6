"""CPU study of the synchronisation chains of the device Huffman decoder (csrc/jpeg_huff.cu): how many bits a decoder that
starts at a subsequence border in a guessed state (coefficient index 0, block 0 of the MCU) needs before its state (bit
position, coefficient index, block-in-MCU) coincides with the true decoder's, on a 4:2:0 stream made here through Pillow.
The device decode time is set by the LONGEST such distance in the stream, not by the mean. Also reported: the distance when
the best of the six possible block-in-MCU guesses is taken at every border (what a decoder that tries all six phases and
keeps the one that agrees with the truth would see), and the same for a guess that starts half a subsequence early.
No GPU. python scripts/huff_chain_sim.py [quality] [width] [height]"""
import io
import json
import os
import sys

import numpy as np
from PIL import Image

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from libultrahdr_dev_b200 import synth  # noqa: E402

SUB = 1024  # bits per subsequence (kJpegSubseqBits)


def huff_lut(bits, vals):
    length, symbol = [0] * 65536, [0] * 65536
    code, k = 0, 0
    for ln in range(1, 17):
        for _ in range(bits[ln - 1]):
            lo, hi = code << (16 - ln), (code + 1) << (16 - ln)
            for i in range(lo, hi):
                length[i], symbol[i] = ln, vals[k]
            code += 1
            k += 1
        code <<= 1
    return length, symbol


def parse(data):
    pos, dc, ac, comps = 2, {}, {}, []
    while True:
        m = data[pos + 1]
        pos += 2
        ln = (data[pos] << 8) | data[pos + 1]
        seg = data[pos + 2: pos + ln]
        pos += ln
        if m == 0xC4:
            i = 0
            while i < len(seg):
                tc, th = seg[i] >> 4, seg[i] & 15
                bits = list(seg[i + 1: i + 17])
                n = sum(bits)
                (ac if tc else dc)[th] = huff_lut(bits, list(seg[i + 17: i + 17 + n]))
                i += 17 + n
        elif m == 0xC0:
            comps = [(seg[6 + 3 * i] , seg[7 + 3 * i] >> 4, seg[7 + 3 * i] & 15) for i in range(seg[5])]
        elif m == 0xDA:
            sel = {seg[1 + 2 * i]: (seg[2 + 2 * i] >> 4, seg[2 + 2 * i] & 15) for i in range(seg[0])}
            break
    scan = bytearray()
    i = pos
    while True:
        b = data[i]
        if b != 0xFF:
            scan.append(b)
            i += 1
        elif data[i + 1] == 0:
            scan.append(0xFF)
            i += 2
        else:
            break
    tables = []  # per block of the MCU: (dc length, dc symbol, ac length, ac symbol)
    for cid, h, v in comps:
        td, ta = sel[cid]
        tables += [(dc[td][0], dc[td][1], ac[ta][0], ac[ta][1])] * (h * v)
    return bytes(scan), tables


def main():
    quality = int(sys.argv[1]) if len(sys.argv) > 1 else 90
    W = int(sys.argv[2]) if len(sys.argv) > 2 else 1536
    H = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
    sdr, _ = synth.natural_pair(W, H, 0)
    y = sdr[: W * H].reshape(H, W)
    u = sdr[W * H: W * H * 5 // 4].reshape(H // 2, W // 2).repeat(2, 0).repeat(2, 1)
    v = sdr[W * H * 5 // 4:].reshape(H // 2, W // 2).repeat(2, 0).repeat(2, 1)
    rng = np.random.Generator(np.random.PCG64(0))
    y = np.clip(y.astype(np.int16) + rng.integers(-6, 7, y.shape), 0, 255).astype(np.uint8)  # the bench's sensor-like texture
    b = io.BytesIO()
    Image.fromarray(np.stack([y, u, v], -1), "YCbCr").save(b, "JPEG", quality=quality, subsampling=2)
    scan, tables = parse(b.getvalue())
    B = len(tables)
    bits = np.unpackbits(np.frombuffer(scan + b"\xff\xff\xff\xff", dtype=np.uint8)).astype(np.uint32)
    n = bits.size - 16
    win = np.zeros(n, np.uint32)
    for k in range(16):
        win |= bits[k:k + n] << (15 - k)
    win = win.tolist()
    total = len(scan) * 8

    def step(p, k, c):
        """one symbol from state (bit position, coefficient index, block in MCU)"""
        t = tables[c]
        w = win[p]
        if k == 0:
            s = t[1][w]
            ln = t[0][w]
            if ln == 0:
                ln = 16
            return p + ln + (s & 15), 1, c
        rs = t[3][w]
        ln = t[2][w]
        if ln == 0:
            ln = 16
        r, s = rs >> 4, rs & 15
        p += ln + s
        if s == 0 and r != 15:
            return p, 0, (c + 1) % B
        k += (16 if s == 0 else r + 1)
        if k >= 64:
            return p, 0, (c + 1) % B
        return p, k, c

    truth = {}
    p, k, c = 0, 0, 0
    while p < total:
        truth[p] = (k, c)
        p, k, c = step(p, k, c)

    def distance(p0, c0, limit=400 * SUB):
        p, k, c = p0, 0, c0
        while p < total and p - p0 < limit:
            if truth.get(p) == (k, c):
                return p - p0
            p, k, c = step(p, k, c)
        return min(total, p0 + limit) - p0

    n_sub = total // SUB
    starts = [j * SUB for j in range(1, n_sub - 1)]
    d0 = np.array([distance(p0, 0) for p0 in starts])
    d6 = np.array([min(distance(p0, c0) for c0 in range(B)) for p0 in starts])
    dh = np.array([min(distance(p0 - SUB // 2, c0) - SUB // 2 for c0 in range(B)) for p0 in starts]).clip(0)

    # a rule a decoder could apply: take the guess that equals the block-in-MCU of the block in progress at the border
    # (known from the left neighbour's exit state), or the block after it
    keys = sorted(truth)
    import bisect
    def true_state_at(p0):
        i = bisect.bisect_right(keys, p0) - 1
        return truth[keys[i]]
    d_same = np.array([distance(p0, true_state_at(p0)[1]) for p0 in starts])
    d_next = np.array([distance(p0, (true_state_at(p0)[1] + 1) % B) for p0 in starts])
    d_rule = np.minimum(d_same, d_next)

    best = [int(np.argmin([distance(p0, c0) for c0 in range(B)])) for p0 in starts]
    offs = np.bincount([(b - true_state_at(p0)[1]) % B for b, p0 in zip(best, starts)], minlength=B).tolist()

    def row(d):
        return {"mean_subseq": round(float(d.mean()) / SUB, 2), "p99_subseq": round(float(np.percentile(d, 99)) / SUB, 2),
                "max_subseq": round(float(d.max()) / SUB, 2)}

    # what the fixed-point rounds need: the longest stretch of borders no chain has reached in step
    def rounds(d):
        reach = np.array(starts) + d  # first bit from which the chain started at border j is right
        need = 0
        for idx, p0 in enumerate(starts):
            best = min((p0 - starts[j]) // SUB for j in range(max(0, idx - 200), idx + 1) if reach[j] <= p0 + SUB) if any(
                reach[j] <= p0 + SUB for j in range(max(0, idx - 200), idx + 1)) else 200
            need = max(need, best)
        return int(need)

    # ---- the decoder as built (jpeg_huff.cu: huff_hyp_kernel, huff_link_kernel, huff_select_kernel) and two variants
    def span(p, k, c, e):
        while p < e and p < total:
            p, k, c = step(p, k, c)
        return (p, k, c)

    def emulate(early=0, second_pass=False):
        borders = [j * SUB for j in range(n_sub + 1)]
        T = [None] * n_sub  # true exit states
        st = (0, 0, 0)
        for j in range(n_sub):
            st = span(st[0], st[1], st[2], borders[j + 1])
            T[j] = st
        X = [[span(max(0, borders[j] - early), 0, h, borders[j + 1]) if not (early and j == 0) else span(0, 0, h, borders[1])
              for h in range(B)] for j in range(n_sub)]
        X[0] = [span(0, 0, 0, borders[1])] * B
        exits, hyp, broken, repaired = [X[0][0]], 0, 0, 0
        j = 1
        while j < n_sub:
            prev = X[j - 1][hyp] if exits[-1] == X[j - 1][hyp] else exits[-1]
            y = span(X[j - 1][hyp][0], X[j - 1][hyp][1], X[j - 1][hyp][2], borders[j + 1])
            match = [h for h in range(B) if X[j][h] == y]
            if match:
                hyp = match[0]
            else:
                broken += 1
                same = [h for h in range(B) if X[j][h][2] == y[2]]
                hyp = same[0] if same else 0
                if second_pass and j + 1 < n_sub:  # the exact exit of j + 1 from the exact exit of j, linked there
                    y2 = span(y[0], y[1], y[2], borders[j + 2])
                    m2 = [h for h in range(B) if X[j + 1][h] == y2]
                    if m2:
                        exits += [y, y2]
                        hyp = m2[0]
                        repaired += 1
                        j += 2
                        continue
            exits.append(y)
            j += 1
        wrong = [exits[j] != T[j] for j in range(n_sub)]
        run = best = 0
        for w in wrong:
            run = run + 1 if w else 0
            best = max(best, run)
        return {"broken_links": broken, "second_pass_repairs": repaired, "wrong_exit_states": int(sum(wrong)), "longest_wrong_run": best}

    as_built = emulate()
    early_half = emulate(early=SUB // 2)
    two_step = emulate(second_pass=True)

    print(json.dumps({"quality": quality, "image": f"{W}x{H} 4:2:0", "scan_bits": total, "subsequences": n_sub, "blocks_per_mcu": B,
                      "best_guess_minus_block_in_progress_mod_B": offs,
                      "guess_block_0": dict(row(d0), rounds=rounds(d0)),
                      "best_of_all_block_guesses": dict(row(d6), rounds=rounds(d6)),
                      "guess_block_in_progress": dict(row(d_same), rounds=rounds(d_same)),
                      "guess_block_after_it": dict(row(d_next), rounds=rounds(d_next)),
                      "better_of_those_two": dict(row(d_rule), rounds=rounds(d_rule)),
                      "best_of_all_guesses_half_a_subsequence_early": dict(row(dh), rounds=rounds(dh)),
                      "decoder_as_built": as_built, "hypotheses_started_half_a_subsequence_early": early_half,
                      "with_a_second_link_pass_over_two_subsequences": two_step}))


if __name__ == "__main__":
    main()
