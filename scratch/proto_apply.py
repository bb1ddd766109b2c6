"""Numerical prototype of the table-driven apply kernel (numpy emulation of the fp32 FFMA chain),
compared with the oracle: exact-match fraction per table resolution."""
import sys

import numpy as np

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
from libultrahdr_dev_b200 import abi, synth  # noqa: E402
from oracle.pyoracle import CpuImpl  # noqa: E402

f32 = np.float32


def fma(a, b, c):
    return (a.astype(np.float64) * np.float64(b) + np.float64(c)).astype(f32) if not isinstance(b, np.ndarray) or True else None


def fma3(a, b, c):
    return (np.asarray(a, np.float64) * np.asarray(b, np.float64) + np.asarray(c, np.float64)).astype(f32)


K1 = f32(1.0 / 1.055)
K0 = f32(0.055 / 1.055)


def srgb_inv_of_t(t):  # exact math, t = (c+0.055)/1.055
    c = t * 1.055 - 0.055
    return np.where(c <= 0.04045, c / 12.92, np.power(np.maximum(t, 1e-30), 2.4))


def hlg_oetf(e):
    a, b, c = 0.17883277, 0.28466892, 0.55991073
    return np.where(e <= 1 / 12, np.sqrt(3 * np.maximum(e, 0)), a * np.log(np.maximum(12 * e - b, 1e-300)) + c)


def pq_oetf(e):
    m1, m2 = 2610 / 16384, 2523 / 4096 * 128
    c1, c2, c3 = 3424 / 4096, 2413 / 4096 * 32, 2392 / 4096 * 32
    p = np.power(np.maximum(e, 0), m1)
    return np.where(e <= 0, 0.0, np.power((c1 + c2 * p) / (1 + c3 * p), m2))


def build_table(fn, lo_exp, hi_exp, m):
    """log-spaced linear segments: index = (bits >> (23-m)) - (((lo_exp+127) << m)); returns a,b float32 arrays."""
    n = (hi_exp - lo_exp) << m
    i = np.arange(n + 1)
    bits0 = ((lo_exp + 127) << 23) + (i.astype(np.int64) << (23 - m))
    x0 = bits0.astype(np.uint32).view(f32).astype(np.float64)
    x1 = (bits0 + (1 << (23 - m))).astype(np.uint32).view(f32).astype(np.float64)
    h = x1 - x0
    n0, n1 = x0 + h * (0.5 - np.sqrt(2) / 4), x0 + h * (0.5 + np.sqrt(2) / 4)
    f0, f1 = fn(n0), fn(n1)
    b = (f1 - f0) / (n1 - n0)
    a = f0 - b * n0
    return a.astype(f32), b.astype(f32)


def lookup(tab, x, lo_exp, m):
    a, b = tab
    idx = (x.view(np.uint32) >> np.uint32(23 - m)).astype(np.int64) - ((lo_exp + 127) << m)
    idx = np.clip(idx, 0, len(a) - 1)
    return fma3(b[idx], x, a[idx])


def run(kind, fmt, mS, mO, w=512, h=288, max_boost=1000.0 / 203.0):
    orc = CpuImpl.oracle()
    if kind == "natural":
        sdr_b, _ = synth.natural_pair(w, h, 21)
        mp_b = synth.natural_map(w // 4, h // 4, 22)
    else:
        sdr_b, _ = synth.uniform_pair(w, h, 23)
        mp_b = np.random.Generator(np.random.PCG64(24)).integers(0, 256, (h // 4, w // 4), dtype=np.uint8)
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
    cR = np.clip(fma3(Y, inv255, (cr * vf).astype(f32)), 0, 1)
    cG = np.clip(fma3(Y, inv255, -((gcb * uf).astype(f32) + (gcr * vf).astype(f32)).astype(f32)), 0, 1)
    cB = np.clip(fma3(Y, inv255, (cb * uf).astype(f32)), 0, 1)
    # gain: IDW with fma chain
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
    g = (m[yl, xl] * wts[..., 0]).astype(f32)
    g = fma3(m[yu, xl], wts[..., 1], g)
    g = fma3(m[yl, xu], wts[..., 2], g)
    g = fma3(m[yu, xu], wts[..., 3], g)
    l2min, l2max = f32(np.log2(1.0)), f32(np.log2(np.float64(f32(max_boost))))
    db = f32(max_boost)
    k = np.float64(db) / np.float64(f32(max_boost))
    A = f32((np.float64(l2max) - np.float64(l2min)) * k)
    Bc = f32(np.float64(l2min) * k - np.log2(np.float64(db)))
    ex = fma3(g, A, Bc)
    f = np.exp2(ex.astype(np.float64)).astype(f32)  # MUFU.EX2 ~ 2 ulp; emulate exact

    S_LO = -5
    stab = build_table(srgb_inv_of_t, S_LO, 1, mS)
    O_LO, O_HI = -26, 1
    e0 = f32(2.0 ** O_LO)
    ofn = hlg_oetf if fmt == abi.OUTPUT_HDR_HLG else pq_oetf
    otab = build_table(lambda x: 1023.0 * ofn(x - np.float64(e0)), O_LO, O_HI, mO)
    out = np.zeros((h, w), dtype=np.uint32)
    for sh, c in ((0, cR), (10, cG), (20, cB)):
        t = fma3(c, K1, K0)
        s = lookup(stab, t, S_LO, mS)
        ep = fma3(s, f, e0)
        v = lookup(otab, ep, O_LO, mO)
        code = np.floor(np.maximum(v, 0)).astype(np.uint32) & 0x3FF
        out |= code << sh
    out |= 0xC0000000
    res = []
    for sh in (0, 10, 20):
        d = ((out >> sh) & 0x3FF).astype(np.int32) - ((want >> sh) & 0x3FF).astype(np.int32)
        res.append(((d == 0).mean(), np.abs(d).max()))
    return res, (len(stab[0]) * 8, len(otab[0]) * 8)


if __name__ == "__main__":
    for fmt in (abi.OUTPUT_HDR_HLG, abi.OUTPUT_HDR_PQ):
        for kind in ("natural", "uniform"):
            for mS, mO in ((7, 6), (8, 7), (9, 7), (9, 8), (10, 8)):
                r, sz = run(kind, fmt, mS, mO)
                print("fmt", fmt, kind, "mS", mS, "mO", mO, "bytes", sz, " ".join(f"{e:.5f}/{d}" for e, d in r))


def half_ref(x):
    b = x.view(np.uint32).astype(np.int64) + 0x1000
    e = (b & 0x7F800000) >> 23
    m = b & 0x007FFFFF
    r = (b & 0x80000000) >> 16
    r = np.where(e > 112, r | ((((e - 112) << 10) & 0x7C00) | (m >> 13)), r)
    sub = (e < 113) & (e > 101)
    r = np.where(sub, r | ((((0x007FF000 + m) >> np.clip(125 - e, 0, 40)) + 1) >> 1), r)
    r = np.where(e > 143, r | 0x7FFF, r)
    return r.astype(np.uint16)
