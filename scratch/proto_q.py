"""Numerical prototype of the q-domain apply kernel: out = P(q), q = Phi(c) * f^(1/2.4),
P(q) = 1023 * O(q^2.4); Phi = (c+0.055)/1.055 for c > 0.04045, (c/12.92)^(1/2.4) below."""
import sys

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "scratch")
from proto_apply import *  # noqa


def runq(kind, fmt, mP, w=512, h=288, max_boost=1000.0 / 203.0, seed=21, lo=-12):
    orc = CpuImpl.oracle()
    if kind == "natural":
        sdr_b, _ = synth.natural_pair(w, h, seed)
        mp_b = synth.natural_map(w // 4, h // 4, seed + 1)
    else:
        sdr_b, _ = synth.uniform_pair(w, h, seed + 2)
        mp_b = np.random.Generator(np.random.PCG64(seed + 3)).integers(0, 256, (h // 4, w // 4), dtype=np.uint8)
    sdr, mp = abi.host_yuv420(sdr_b, w, h, 0), abi.host_map(mp_b, w // 4, h // 4)
    md = abi.default_metadata(max_boost)
    st, want, _ = orc.apply_gainmap(sdr, mp, md, fmt, 3.4e38)
    want = want.view(np.uint32).reshape(h, w)
    Y = sdr_b[: w * h].reshape(h, w).astype(f32)
    U = sdr_b[w * h: w * h + w * h // 4].reshape(h // 2, w // 2)
    V = sdr_b[w * h + w * h // 4:].reshape(h // 2, w // 2)
    uf = ((U.astype(np.int32) - 128).astype(f32) * f32(1 / 255.0)).astype(f32)
    vf = ((V.astype(np.int32) - 128).astype(f32) * f32(1 / 255.0)).astype(f32)
    uf = np.repeat(np.repeat(uf, 2, 0), 2, 1)
    vf = np.repeat(np.repeat(vf, 2, 0), 2, 1)
    cr, cb = f32(1.402), f32(1.772)
    gcb, gcr = f32(f32(0.114) * f32(1.772) / f32(0.587)), f32(f32(0.299) * f32(1.402) / f32(0.587))
    inv255 = f32(1 / 255.0)
    chR = (cr * vf).astype(f32)
    chG = (-((gcb * uf).astype(f32) + (gcr * vf).astype(f32))).astype(f32)
    chB = (cb * uf).astype(f32)
    idw = np.stack([orc.idw_table(4, t) for t in range(4)]).reshape(4, 4, 4, 4)
    mw, mh = w // 4, h // 4
    ys, xs = np.mgrid[0:h, 0:w]
    xl, yl = xs // 4, ys // 4
    xu, yu = np.minimum(xl + 1, mw - 1), np.minimum(yl + 1, mh - 1)
    ox, oy = xs % 4, ys % 4
    tsel = np.where((xl == xu) & (yl == yu), 3, np.where(xl == xu, 1, np.where(yl == yu, 2, 0)))
    wts = idw[tsel, oy, ox]
    e255 = (np.arange(256, dtype=f32) / f32(255.0)).astype(f32)
    m = e255[mp_b]
    l2min, l2max = f32(np.log2(1.0)), f32(np.log2(np.float64(f32(max_boost))))
    db = f32(max_boost)
    k = np.float64(db) / np.float64(f32(max_boost))
    A = np.float64((np.float64(l2max) - np.float64(l2min)) * k)
    Bc = np.float64(np.float64(l2min) * k - np.log2(np.float64(db)))
    Aq, Bq = f32(A / 2.4), f32(Bc / 2.4)
    wq = (wts.astype(np.float64) * np.float64(Aq)).astype(f32)
    ex = fma3(m[yl, xl], wq[..., 0], Bq)
    ex = fma3(m[yu, xl], wq[..., 1], ex)
    ex = fma3(m[yl, xu], wq[..., 2], ex)
    ex = fma3(m[yu, xu], wq[..., 3], ex)
    fq = np.exp2(ex.astype(np.float64)).astype(f32)
    q0 = f32(2.0 ** lo)
    ofn = hlg_oetf if fmt == abi.OUTPUT_HDR_HLG else pq_oetf
    ptab = build_table(lambda x: 1023.0 * ofn(np.power(np.maximum(x - np.float64(q0), 0), 2.4)), lo, 1, mP)
    KP = f32(1.0 / (255.0 * 1.055))
    out = np.zeros((h, w), dtype=np.uint32)
    ndark = 0
    for sh, ch in ((0, chR), (10, chG), (20, chB)):
        off = ((ch.astype(np.float64) + 0.055) / 1.055).astype(f32)  # per-quad constant, computed in fp32 on device: fma(ch, K1, K0)
        off = fma3(ch, K1, K0)
        phi = np.clip(fma3(Y, KP, off), 0, 1)
        c = np.clip(fma3(Y, inv255, ch), 0, 1)
        dark = c <= f32(0.04045)
        ndark += dark.sum()
        x = (c * f32(1 / 12.92)).astype(f32)
        philow = np.exp2((np.log2(np.maximum(x.astype(np.float64), 1e-300)).astype(f32) * f32(1 / 2.4)).astype(np.float64)).astype(f32)
        philow = np.where(x == 0, f32(0), philow)
        phi = np.where(dark, philow, phi)
        q = fma3(phi, fq, q0)
        v = lookup(ptab, q, lo, mP)
        code = np.floor(np.maximum(v, 0)).astype(np.uint32) & 0x3FF
        out |= code << sh
    out |= 0xC0000000
    res = []
    for sh in (0, 10, 20):
        d = ((out >> sh) & 0x3FF).astype(np.int32) - ((want >> sh) & 0x3FF).astype(np.int32)
        res.append(((d == 0).mean(), np.abs(d).max()))
    return res, len(ptab[0]) * 8, ndark / (3 * w * h)


if __name__ == "__main__":
    for fmt in (abi.OUTPUT_HDR_HLG, abi.OUTPUT_HDR_PQ):
        for kind in ("natural", "uniform"):
            for mP in (6, 7, 8):
                r, sz, nd = runq(kind, fmt, mP)
                print("fmt", fmt, kind, "mP", mP, "bytes", sz, "dark", round(nd, 4), " ".join(f"{e:.5f}/{d}" for e, d in r))
