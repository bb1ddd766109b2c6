"""Generates tests/golden/golden_v1.npz from the UNMODIFIED reference hot path (oracle/_ref/
libuhdr_ref.so, built by oracle/Makefile from /root/reference). Run in the build container:

    python tests/golden/make_golden.py

Inputs are regenerated from seeds by libultrahdr_dev_b200.synth (their SHA-256 is stored so a
change of the generator is detected); only the reference's OUTPUTS are stored. Cases cover the
BASELINE.json configurations at reduced size plus strides, scale factors and boosts.
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from libultrahdr_dev_b200 import abi, synth  # noqa: E402

GENERATE_CASES = [
    # name, w, h, kind, tf, sdr_gamut, hdr_gamut, sdr_is_601, (sdr luma, sdr chroma, hdr luma, hdr chroma strides)
    ("cfg1_hlg_2100_709", 320, 176, "natural", abi.TF_HLG, 0, 2, False, None),
    ("cfg3_pq_2100_p3", 320, 176, "natural", abi.TF_PQ, 1, 2, False, None),
    ("cfg4_hlg_2100_2100", 256, 192, "natural", abi.TF_HLG, 2, 2, False, None),
    ("linear_709_709", 128, 64, "natural", abi.TF_LINEAR, 0, 0, False, None),
    ("uniform_hlg", 192, 128, "uniform", abi.TF_HLG, 0, 2, False, None),
    ("uniform_pq_601", 192, 128, "uniform", abi.TF_PQ, 0, 1, True, None),
    ("strided_hlg", 100, 60, "uniform", abi.TF_HLG, 0, 2, False, (118, 59, 118, 118)),
]
APPLY_CASES = [
    # name, w, h, kind, scale, max_boost, min_boost, display_boost
    ("cfg2_decode_default", 320, 176, "natural", 4, 1000.0 / 203.0, 1.0, 3.4028234663852886e38),
    ("cfg3_pq_boost", 320, 176, "natural", 4, 10000.0 / 203.0, 1.0, 3.4028234663852886e38),
    ("uniform_scale4", 192, 128, "uniform", 4, 1000.0 / 203.0, 1.0, 3.4028234663852886e38),
    ("scale2_boost2", 128, 64, "uniform", 2, 1000.0 / 203.0, 0.5, 2.0),
    ("scale1", 64, 32, "uniform", 1, 4.0, 1.0, 4.0),
    ("scale8_boost1", 128, 64, "natural", 8, 1000.0 / 203.0, 1.0, 1.0),
]
FORMATS = [abi.OUTPUT_HDR_LINEAR, abi.OUTPUT_HDR_PQ, abi.OUTPUT_HDR_HLG, abi.OUTPUT_HDR_LINEAR_RGB_10BIT]
TONEMAP_CASES = [("tm_packed", 256, 128, None, None, 256, 128), ("tm_strided", 130, 66, 136, 140, 144, 72)]
CONVERT_CASES = [("cv_%d_%d" % (s, d), 96, 64, s, d) for s in range(3) for d in range(3) if s != d]


def gen_inputs(case):
    name, w, h, kind, tf, sg, hg, is601, strides = case
    st = strides or (None, None, None, None)
    seed = int(hashlib.sha256(name.encode()).hexdigest()[:6], 16)
    if kind == "natural":
        return synth.natural_pair(w, h, seed, tf, sg, hg, *st)
    return synth.uniform_pair(w, h, seed, *st)


def apply_inputs(case):
    name, w, h, kind, scale, mx, mn, db = case
    seed = int(hashlib.sha256(name.encode()).hexdigest()[:6], 16)
    if kind == "natural":
        sdr = synth.natural_pair(w, h, seed)[0]
        mp = synth.natural_map(w // scale, h // scale, seed + 1)
    else:
        sdr = synth.uniform_pair(w, h, seed)[0]
        mp = np.random.Generator(np.random.PCG64(seed + 1)).integers(0, 256, (h // scale, w // scale), dtype=np.uint8)
    return sdr, mp


def tonemap_input(case):
    name, w, h, ls, cs, dls, dcs = case
    seed = int(hashlib.sha256(name.encode()).hexdigest()[:6], 16)
    return synth.uniform_pair(w, h, seed, hdr_luma_stride=ls, hdr_chroma_stride=cs, dirty_low_bits=True)[1]


def convert_input(case):
    name, w, h, s, d = case
    seed = int(hashlib.sha256(name.encode()).hexdigest()[:6], 16)
    return synth.uniform_pair(w, h, seed)[0]


def sha(*arrays):
    m = hashlib.sha256()
    for a in arrays:
        m.update(np.ascontiguousarray(a).tobytes())
    return m.hexdigest()


def run_all(impl):
    """-> dict name -> ndarray, computed with `impl` (a CpuImpl)."""
    out = {}
    for case in GENERATE_CASES:
        name, w, h, kind, tf, sg, hg, is601, strides = case
        st = strides or (None, None, None, None)
        sdr_b, hdr_b = gen_inputs(case)
        s, mp, md, _ = impl.generate_gainmap(abi.host_yuv420(sdr_b, w, h, sg, st[0], st[1]), abi.host_p010(hdr_b, w, h, hg, st[2], st[3]), tf, is601)
        assert s == 0
        out["gen/" + name] = mp.copy()
        out["gen/" + name + "/max_boost"] = np.float32(md.maxContentBoost)
        out["gen/" + name + "/in_sha"] = np.array(sha(sdr_b, hdr_b))
    for case in APPLY_CASES:
        name, w, h, kind, scale, mx, mn, db = case
        sdr_b, mp_b = apply_inputs(case)
        md = abi.default_metadata(mx, mn)
        for fmt in FORMATS:
            s, o, _ = impl.apply_gainmap(abi.host_yuv420(sdr_b, w, h, 0), abi.host_map(mp_b, w // scale, h // scale), md, fmt, db)
            assert s == 0
            out["apply/%s/fmt%d" % (name, fmt)] = o.copy()
        out["apply/" + name + "/in_sha"] = np.array(sha(sdr_b, mp_b))
    for case in TONEMAP_CASES:
        name, w, h, ls, cs, dls, dcs = case
        hdr_b = tonemap_input(case)
        buf = np.full(abi.yuv420_size(w, h, dls, dcs), 0x55, dtype=np.uint8)
        d = abi.host_yuv420(buf, w, h, -1, dls, dcs)
        assert impl.tonemap(abi.host_p010(hdr_b, w, h, 2, ls, cs), d) == 0
        out["tonemap/" + name] = buf
        out["tonemap/" + name + "/in_sha"] = np.array(sha(hdr_b))
    for case in CONVERT_CASES:
        name, w, h, s, d = case
        buf = convert_input(case)
        insha = sha(buf)
        assert impl.convert_yuv(abi.host_yuv420(buf, w, h, s), s, d) == 0
        out["convert/" + name] = buf
        out["convert/" + name + "/in_sha"] = np.array(insha)
    return out


if __name__ == "__main__":
    from oracle.pyoracle import CpuImpl

    ref = CpuImpl.reference()
    data = run_all(ref)
    path = os.path.join(HERE, "golden_v1.npz")
    np.savez_compressed(path, **data)
    print(path, os.path.getsize(path), "bytes,", len(data), "arrays")
