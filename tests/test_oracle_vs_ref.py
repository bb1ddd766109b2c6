"""Pins the oracle (oracle/uhdr_oracle.c, our plain-C restatement) against the UNMODIFIED
reference hot path compiled by oracle/Makefile into oracle/_ref/libuhdr_ref.so — bit for bit, per
L1 function and per whole image, in both numeric variants (as built = analytic; -DUSE_*_LUT=1).
Skips (not fails) where oracle/_ref is absent and cannot be built (no /root/reference)."""
import ctypes as C

import numpy as np
import pytest

from libultrahdr_dev_b200 import abi, synth

FORMATS = [abi.OUTPUT_HDR_LINEAR, abi.OUTPUT_HDR_PQ, abi.OUTPUT_HDR_HLG, abi.OUTPUT_HDR_LINEAR_RGB_10BIT]


def _grid(n, lo=0.0, hi=1.0, seed=7):
    rng = np.random.Generator(np.random.PCG64(seed))
    g = np.concatenate([np.linspace(lo, hi, n // 2, dtype=np.float32), rng.uniform(lo, hi, n // 2).astype(np.float32),
                        np.array([0.0, 1e-7, 1e-4, 0.0001, 0.00010001, 0.04045, 0.040451, 1 / 12, 0.5, 0.50000006, 1.0],
                                 dtype=np.float32)])
    return np.clip(g, lo, hi)


@pytest.mark.parametrize("fn", ["srgbInvOetf", "hlgOetf", "hlgInvOetf", "pqOetf", "pqInvOetf"])
def test_transfer_functions_bit_exact(orc, ref, fn):
    xs = _grid(6000)
    for suffix in ("", "LUT"):
        fo, fr = getattr(orc, "c_" + fn + suffix), getattr(ref, "c_" + fn + suffix)
        a = np.array([fo(float(x)) for x in xs], dtype=np.float32)
        b = np.array([fr(float(x)) for x in xs], dtype=np.float32)
        assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), fn + suffix


def test_hlg_oetf_above_one_bit_exact(orc, ref):
    """display_boost < maxContentBoost feeds the OETFs values above 1.0 (ultrahdr.cpp:451)."""
    for x in np.linspace(1.0, 50.0, 500, dtype=np.float32):
        for fn in ("hlgOetf", "pqOetf"):
            a, b = getattr(orc, "c_" + fn)(float(x)), getattr(ref, "c_" + fn)(float(x))
            assert np.float32(a).view(np.uint32) == np.float32(b).view(np.uint32)


def test_color_math_bit_exact(orc, ref):
    rng = np.random.Generator(np.random.PCG64(3))
    for _ in range(400):
        v = rng.uniform(-0.6, 1.2, 3).astype(np.float32)
        for g in range(3):
            assert np.array_equal(orc.vec3("yuvToRgb", g)(v), ref.vec3("yuvToRgb", g)(v))
            assert np.array_equal(orc.vec3("rgbToYuv", g)(v), ref.vec3("rgbToYuv", g)(v))
            assert orc.c_luminance(g, *map(float, v)) == ref.c_luminance(g, *map(float, v))
        for i in range(6):
            assert np.array_equal(orc.vec3("yuvToYuv", i)(v), ref.vec3("yuvToYuv", i)(v))
        for sg in range(3):
            for hg in range(3):
                o1, o2 = (C.c_float * 3)(), (C.c_float * 3)()
                i3 = (C.c_float * 3)(*map(float, v))
                assert orc.c_hdrConversion(sg, hg, i3, o1) == ref.c_hdrConversion(sg, hg, i3, o2) == 0
                assert list(o1) == list(o2)
    # selector returns nullptr for unspecified gamuts (gainmapmath_test.cpp:949-978)
    o = (C.c_float * 3)()
    for sg, hg in ((-1, 0), (0, -1), (-1, -1), (2, -1)):
        assert orc.c_hdrConversion(sg, hg, (C.c_float * 3)(0, 0, 0), o) == ref.c_hdrConversion(sg, hg, (C.c_float * 3)(0, 0, 0), o) == -1


def test_encode_apply_gain_bit_exact(orc, ref):
    rng = np.random.Generator(np.random.PCG64(5))
    bounds = [(0.25, 4.0), (0.5, 2.0), (0.125, 8.0), (1.0, 8.0), (0.5, 8.0), (1.0, 1000 / 203), (1.0, 10000 / 203)]
    for mn, mx in bounds:
        for _ in range(300):
            ys, yh = float(rng.uniform(0, 250)), float(rng.uniform(0, 10000))
            if rng.random() < 0.1:
                ys = 0.0
            assert orc.c_encodeGain(ys, yh, mn, mx) == ref.c_encodeGain(ys, yh, mn, mx)
            rgb = rng.uniform(0, 1, 3).astype(np.float32)
            g = float(np.float32(rng.integers(0, 256) / 255.0))
            for db in (None, mx, 1.0, 2.0):
                a, b = orc.apply_gain(rgb, g, mn, mx, db), ref.apply_gain(rgb, g, mn, mx, db)
                assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
            a, b = orc.apply_gain_lut(rgb, g, mn, mx, mx), ref.apply_gain_lut(rgb, g, mn, mx, mx)
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32))


def test_packers_bit_exact(orc, ref):
    rng = np.random.Generator(np.random.PCG64(9))
    vals = np.concatenate([rng.uniform(0, 1.2, 600), 10.0 ** rng.uniform(-9, 5, 600), -(10.0 ** rng.uniform(-9, 5, 200)),
                           [0.0, 1.0, -1.0, 0.1, 3.4028234e38, -3.4028234e38, 1.17549435e-38, 65504.0, 65520.0, 6.1e-5, 5.96e-8, 2.98e-8]]).astype(np.float32)
    for v in vals:
        assert orc.c_floatToHalf(float(v)) == ref.c_floatToHalf(float(v)), float(v)
    for i in range(0, len(vals) - 3, 3):
        r, g, b = (float(abs(x)) for x in vals[i:i + 3])
        assert orc.c_colorToRgbaF16(r, g, b) == ref.c_colorToRgbaF16(r, g, b)
        r, g, b = (min(x, 4.0e6) for x in (r, g, b))  # float->uint32 is UB beyond 2^32 in both
        assert orc.c_colorToRgba1010102(r, g, b) == ref.c_colorToRgba1010102(r, g, b)


@pytest.mark.parametrize("scale", [1, 2, 3, 4, 8])
def test_idw_tables_and_sample_map_bit_exact(orc, ref, scale):
    for t in range(4):
        assert np.array_equal(orc.idw_table(scale, t).view(np.uint32), ref.idw_table(scale, t).view(np.uint32))
    mw, mh = 7, 5
    mp_b = np.random.Generator(np.random.PCG64(scale)).integers(0, 256, (mh, mw), dtype=np.uint8)
    mp = abi.host_map(mp_b, mw, mh)
    for y in range(mh * scale):
        for x in range(mw * scale):
            a, b = orc.c_sampleMapIdw(C.byref(mp), scale, x, y), ref.c_sampleMapIdw(C.byref(mp), scale, x, y)
            assert np.float32(a).view(np.uint32) == np.float32(b).view(np.uint32)
            a, b = orc.c_sampleMapFloat(C.byref(mp), float(scale), x, y), ref.c_sampleMapFloat(C.byref(mp), float(scale), x, y)
            assert np.float32(a).view(np.uint32) == np.float32(b).view(np.uint32)


def test_pixel_accessors_bit_exact(orc, ref):
    w, h = 16, 12
    sdr_b, hdr_b = synth.uniform_pair(w, h, 77, 20, 11, 18, 22, dirty_low_bits=True)
    sdr = abi.host_yuv420(sdr_b, w, h, 0, 20, 11)
    hdr = abi.host_p010(hdr_b, w, h, 2, 18, 22)
    for y in range(h):
        for x in range(w):
            assert np.array_equal(orc.pixel("getYuv420Pixel", sdr, x, y), ref.pixel("getYuv420Pixel", sdr, x, y))
            assert np.array_equal(orc.pixel("getP010Pixel", hdr, x, y), ref.pixel("getP010Pixel", hdr, x, y))
    for s in (2, 4):
        for y in range(h // s):
            for x in range(w // s):
                assert np.array_equal(orc.pixel("sampleYuv420", sdr, s, x, y), ref.pixel("sampleYuv420", sdr, s, x, y))
                assert np.array_equal(orc.pixel("sampleP010", hdr, s, x, y), ref.pixel("sampleP010", hdr, s, x, y))


# ---------------------------------------------------------------------------------- whole images
def _impls(request, lut):
    from oracle.pyoracle import CpuImpl

    o = CpuImpl.oracle(lut=lut)
    r = request.getfixturevalue("ref_lut" if lut else "ref")
    return o, r


@pytest.mark.parametrize("lut", [False, True])
@pytest.mark.parametrize("tf", [abi.TF_HLG, abi.TF_PQ, abi.TF_LINEAR])
@pytest.mark.parametrize("sg,hg,is601", [(0, 2, False), (1, 2, False), (2, 2, False), (0, 0, False), (2, 0, False), (1, 0, True), (0, 1, True)])
def test_generate_whole_image_bit_exact(request, lut, tf, sg, hg, is601):
    o, r = _impls(request, lut)
    try:
        w, h = 96, 64
        for kind in ("natural", "uniform"):
            if kind == "natural":
                sdr_b, hdr_b = synth.natural_pair(w, h, 5 + tf, tf, sg, hg, 100, 52, 104, 108)
            else:
                sdr_b, hdr_b = synth.uniform_pair(w, h, 6 + tf, 100, 52, 104, 108)
            sdr = abi.host_yuv420(sdr_b, w, h, sg, 100, 52)
            hdr = abi.host_p010(hdr_b, w, h, hg, 104, 108)
            s1, m1, md1, d1 = o.generate_gainmap(sdr, hdr, tf, is601)
            s2, m2, md2, d2 = r.generate_gainmap(sdr, hdr, tf, is601)
            assert s1 == s2 == 0 and md1.as_dict() == md2.as_dict()
            assert (d1.width, d1.height, d1.luma_stride, d1.pixelFormat, d1.colorGamut) == (d2.width, d2.height, d2.luma_stride, d2.pixelFormat, d2.colorGamut)
            assert np.array_equal(m1, m2), f"{kind}: {(m1 != m2).sum()} map bytes differ"
    finally:
        o.lib.orc_set_lut_mode(0)


@pytest.mark.parametrize("lut", [False, True])
@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("scale,boost,min_boost", [(4, 3.4e38, 1.0), (2, 2.0, 0.5), (1, 3.4e38, 1.0), (8, 1.0, 1.0)])
def test_apply_whole_image_bit_exact(request, lut, fmt, scale, boost, min_boost):
    o, r = _impls(request, lut)
    try:
        w, h = 64, 48
        sdr_b, _ = synth.uniform_pair(w, h, 12, 70, 37)
        sdr = abi.host_yuv420(sdr_b, w, h, 1, 70, 37)
        mp_b = np.random.Generator(np.random.PCG64(13)).integers(0, 256, (h // scale, w // scale), dtype=np.uint8)
        mp = abi.host_map(mp_b, w // scale, h // scale)
        md = abi.default_metadata(1000.0 / 203.0, min_boost)
        s1, o1, d1 = o.apply_gainmap(sdr, mp, md, fmt, boost)
        s2, o2, d2 = r.apply_gainmap(sdr, mp, md, fmt, boost)
        assert s1 == s2 == 0
        assert (d1.width, d1.height, d1.colorGamut) == (d2.width, d2.height, d2.colorGamut) == (w, h, 1)
        assert np.array_equal(o1, o2), f"{(o1 != o2).sum()} output bytes differ"
    finally:
        o.lib.orc_set_lut_mode(0)


def test_tonemap_and_convert_yuv_bit_exact(orc, ref):
    w, h = 50, 34
    _, hdr_b = synth.uniform_pair(w, h, 14, hdr_luma_stride=56, hdr_chroma_stride=60, dirty_low_bits=True)
    src = abi.host_p010(hdr_b, w, h, 2, 56, 60)
    outs = []
    for impl in (orc, ref):
        buf = np.full(abi.yuv420_size(w, h, 64, 32), 0x55, dtype=np.uint8)
        d = abi.host_yuv420(buf, w, h, -1, 64, 32)
        assert impl.tonemap(src, d) == 0 and d.colorGamut == 2
        outs.append(buf)
    assert np.array_equal(outs[0], outs[1])
    base, _ = synth.uniform_pair(w, h, 15, 54, 29)
    for s in range(3):
        for d in range(3):
            a, b = base.copy(), base.copy()
            assert orc.convert_yuv(abi.host_yuv420(a, w, h, s, 54, 29), s, d) == ref.convert_yuv(abi.host_yuv420(b, w, h, s, 54, 29), s, d) == 0
            assert np.array_equal(a, b)
            assert (s == d) == np.array_equal(a, base)


def test_error_codes_match_reference(orc, ref):
    """Argument checks in the reference's order (ultrahdr.cpp:189-202, 245-247, 364-406, 518-523)."""
    w, h = 16, 8
    sdr_b, hdr_b = synth.uniform_pair(w, h, 1)
    mp_b = np.zeros((h // 4, w // 4), dtype=np.uint8)
    for impl in (orc, ref):
        sdr, hdr = abi.host_yuv420(sdr_b, w, h, 0), abi.host_p010(hdr_b, w, h, 2)
        assert impl.generate_gainmap(sdr, hdr, abi.TF_SRGB)[0] == abi.ERROR_INVALID_TRANS_FUNC
        assert impl.generate_gainmap(sdr, hdr, abi.TF_UNSPECIFIED)[0] == abi.ERROR_INVALID_TRANS_FUNC
        bad = abi.host_p010(hdr_b, w, h - 2, 2)
        assert impl.generate_gainmap(sdr, bad, abi.TF_HLG)[0] == abi.ERROR_RESOLUTION_MISMATCH
        bad = abi.host_yuv420(sdr_b, w, h, -1)
        assert impl.generate_gainmap(bad, hdr, abi.TF_HLG)[0] == abi.ERROR_INVALID_COLORGAMUT
        bad = abi.host_yuv420(sdr_b, w, h, 0)
        bad.data = None
        assert impl.generate_gainmap(bad, hdr, abi.TF_HLG)[0] == abi.ERROR_BAD_PTR
        mp = abi.host_map(mp_b, w // 4, h // 4)
        for field, val in (("gamma", 2.0), ("offsetSdr", 0.1), ("offsetHdr", 0.1), ("hdrCapacityMax", 3.0), ("hdrCapacityMin", 0.5)):
            md = abi.default_metadata(4.0)
            setattr(md, field, val)
            assert impl.apply_gainmap(sdr, mp, md, abi.OUTPUT_HDR_HLG, 4.0)[0] == abi.ERROR_BAD_METADATA
        md = abi.default_metadata(4.0)
        md.version = b"1.1"
        assert impl.apply_gainmap(sdr, mp, md, abi.OUTPUT_HDR_HLG, 4.0)[0] == abi.ERROR_BAD_METADATA
        md = abi.default_metadata(4.0)
        bad_map = abi.host_map(mp_b, 3, 2)
        assert impl.apply_gainmap(sdr, bad_map, md, abi.OUTPUT_HDR_HLG, 4.0)[0] == abi.ERROR_UNSUPPORTED_MAP_SCALE_FACTOR
        bad_map = abi.host_map(mp_b, 4, 1)  # factors 4 and 8
        assert impl.apply_gainmap(sdr, bad_map, md, abi.OUTPUT_HDR_HLG, 4.0)[0] == abi.ERROR_UNSUPPORTED_MAP_SCALE_FACTOR
        st, out, _ = impl.apply_gainmap(sdr, mp, md, abi.OUTPUT_SDR, 4.0)  # no case in the switch: writes nothing
        assert st == 0 and not out.any()
