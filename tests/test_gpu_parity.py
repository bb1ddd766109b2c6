"""GPU parity: the CUDA path (through the C ABI, host-pointer entry points) against the oracle on
the same seeded inputs. Integer stages bit-exact; float stages within +-1 LSB / 1e-3 relative."""
import ctypes as C

import numpy as np
import pytest

from libultrahdr_dev_b200 import abi, synth
from libultrahdr_dev_b200.api import take_map

import parity

pytestmark = pytest.mark.gpu

GAMUTS = [abi.COLORGAMUT_BT709, abi.COLORGAMUT_P3, abi.COLORGAMUT_BT2100]
FORMATS = [abi.OUTPUT_HDR_LINEAR, abi.OUTPUT_HDR_PQ, abi.OUTPUT_HDR_HLG, abi.OUTPUT_HDR_LINEAR_RGB_10BIT]


def gpu_generate(uh, sdr, hdr, tf, sdr_is_601=False):
    md, dest = abi.Metadata(), abi.Image()
    st = uh.generateGainMap(sdr, hdr, tf, md, dest, sdr_is_601)
    assert st == 0
    return take_map(dest), md, dest


def gpu_apply(uh, sdr, mp, md, fmt, boost):
    out = abi.alloc_output(fmt, sdr.width, sdr.height)
    dest = abi.Image()
    dest.data = out.ctypes.data
    st = uh.applyGainMap(sdr, mp, md, fmt, boost, dest)
    assert st == 0
    return out, dest


# ------------------------------------------------------------------------------ toneMap (bit exact)
@pytest.mark.parametrize("w,h,ls,cs,dls,dcs", [
    (1280, 720, None, None, None, None),
    (1920, 1080, None, None, 1920, 960),
    (130, 66, 130, 130, 144, 72),        # ALIGNM(w,16) destination stride, as jpegr.cpp:185
    (34, 18, 52, 56, 48, 24),            # padded source, padded destination
    (22, 10, 23, 26, 27, 15),            # odd strides -> scalar path
    (8, 8, None, None, None, None),
])
def test_tonemap_exact(uh, orc, w, h, ls, cs, dls, dcs):
    _, hdr = synth.uniform_pair(w, h, 3, hdr_luma_stride=ls, hdr_chroma_stride=cs, dirty_low_bits=True)
    src = abi.host_p010(hdr, w, h, abi.COLORGAMUT_BT2100, ls, cs)
    dls = dls or w
    dcs = dcs or dls // 2
    outs = []
    for impl in ("gpu", "orc"):
        buf = np.full(abi.yuv420_size(w, h, dls, dcs), 0x55, dtype=np.uint8)
        d = abi.host_yuv420(buf, w, h, abi.COLORGAMUT_UNSPECIFIED, dls, dcs)
        st = uh.toneMap(src, d) if impl == "gpu" else orc.tonemap(src, d)
        assert st == 0 and d.colorGamut == abi.COLORGAMUT_BT2100
        outs.append(buf)
    assert np.array_equal(outs[0], outs[1])


# ------------------------------------------------------------------------------ generateGainMap
@pytest.mark.parametrize("tf", [abi.TF_HLG, abi.TF_PQ, abi.TF_LINEAR])
@pytest.mark.parametrize("sg,hg", [(0, 2), (1, 2), (2, 2), (0, 0), (1, 0), (2, 1), (0, 1)])
@pytest.mark.parametrize("kind", ["natural", "uniform"])
def test_generate_parity(uh, orc, tf, sg, hg, kind):
    w, h = 640, 360
    if kind == "natural":
        sdr_b, hdr_b = synth.natural_pair(w, h, 11 + tf, tf, sg, hg)
    else:
        sdr_b, hdr_b = synth.uniform_pair(w, h, 17 + tf)
    sdr = abi.host_yuv420(sdr_b, w, h, sg)
    hdr = abi.host_p010(hdr_b, w, h, hg)
    st, want, md_o, _ = orc.generate_gainmap(sdr, hdr, tf)
    got, md_g, dest = gpu_generate(uh, sdr, hdr, tf)
    assert st == 0 and md_g.as_dict() == md_o.as_dict()
    assert (dest.width, dest.height, dest.luma_stride, dest.pixelFormat) == (w // 4, h // 4, w // 4, abi.PIX_FMT_MONOCHROME)
    exact = parity.compare_bytes(got, want, f"map tf={tf} {sg}<-{hg} {kind}", min_exact=0.999)
    print(f"generate tf={tf} sdr={sg} hdr={hg} {kind}: exact {exact:.5f}, map range {want.min()}..{want.max()}")


def test_generate_sdr_is_601(uh, orc):
    w, h = 320, 240
    sdr_b, hdr_b = synth.natural_pair(w, h, 5, abi.TF_HLG, 1, 2)
    sdr, hdr = abi.host_yuv420(sdr_b, w, h, 0), abi.host_p010(hdr_b, w, h, 2)
    _, want, _, _ = orc.generate_gainmap(sdr, hdr, abi.TF_HLG, sdr_is_601=True)
    got, _, _ = gpu_generate(uh, sdr, hdr, abi.TF_HLG, sdr_is_601=True)
    parity.compare_bytes(got, want, "map sdr_is_601", min_exact=0.999)
    _, other, _, _ = orc.generate_gainmap(sdr, hdr, abi.TF_HLG, sdr_is_601=False)
    assert not np.array_equal(want, other)  # the flag matters on this input


@pytest.mark.parametrize("w,h,sls,scs,hls,hcs", [
    (100, 60, 118, 59, 118, 118),     # width+18 strides as in jpegr_test.cpp:1489
    (64, 32, 67, 35, 65, 70),         # odd strides -> scalar path
    (36, 20, 40, 24, 44, 48),
    (18, 10, None, None, None, None), # dims not multiples of 4: map is floor(w/4) x floor(h/4)
    (4, 4, None, None, None, None),
])
def test_generate_strides(uh, orc, w, h, sls, scs, hls, hcs):
    """Strided inputs give byte-identical maps to packed ones (jpegr_test.cpp:1485-1582)."""
    sdr_p, hdr_p = synth.uniform_pair(w, h, 9)
    sdr_s, hdr_s = synth.uniform_pair(w, h, 9, sls, scs, hls, hcs)
    packed = gpu_generate(uh, abi.host_yuv420(sdr_p, w, h, 0), abi.host_p010(hdr_p, w, h, 2), abi.TF_HLG)[0]
    strided = gpu_generate(uh, abi.host_yuv420(sdr_s, w, h, 0, sls, scs), abi.host_p010(hdr_s, w, h, 2, hls, hcs), abi.TF_HLG)[0]
    assert np.array_equal(packed, strided)
    _, want, _, _ = orc.generate_gainmap(abi.host_yuv420(sdr_s, w, h, 0, sls, scs), abi.host_p010(hdr_s, w, h, 2, hls, hcs), abi.TF_HLG)
    parity.compare_bytes(strided, want, "map strided", min_exact=0.995)


# ------------------------------------------------------------------------------ applyGainMap
MODES = [abi.APPLY_AUTO, abi.APPLY_Q_PIPE, abi.APPLY_TWO_TABLE, abi.APPLY_GENERIC]


@pytest.fixture
def apply_mode(uh, request):
    """Selects the applyGainMap kernel family (uhdrb200_test_set_apply_mode) for one test."""
    assert uh.lib.uhdrb200_test_set_apply_mode(request.param) == 0
    yield request.param
    assert uh.lib.uhdrb200_test_set_apply_mode(abi.APPLY_AUTO) == 0


@pytest.mark.parametrize("apply_mode", MODES, indirect=True)
@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("kind", ["natural", "uniform", "dark"])
@pytest.mark.parametrize("max_boost", [1000.0 / 203.0, 10000.0 / 203.0])
def test_apply_parity(uh, orc, fmt, kind, max_boost, apply_mode):
    w, h = 512, 288
    if kind == "natural":
        sdr_b, _ = synth.natural_pair(w, h, 21)
        mp_b = synth.natural_map(w // 4, h // 4, 22)
    elif kind == "dark":  # luma 0..24 with mild chroma: every pixel near srgbInvOetf's linear toe (c <= 0.04045)
        rng = np.random.Generator(np.random.PCG64(25))
        sdr_b = np.concatenate([rng.integers(0, 25, w * h, dtype=np.uint8), rng.integers(120, 137, w * h // 2, dtype=np.uint8)])
        mp_b = synth.natural_map(w // 4, h // 4, 26)
    else:
        sdr_b, _ = synth.uniform_pair(w, h, 23)
        mp_b = np.random.Generator(np.random.PCG64(24)).integers(0, 256, (h // 4, w // 4), dtype=np.uint8)
    sdr, mp = abi.host_yuv420(sdr_b, w, h, 0), abi.host_map(mp_b, w // 4, h // 4)
    md = abi.default_metadata(max_boost)
    st, want, _ = orc.apply_gainmap(sdr, mp, md, fmt, 3.4e38)  # FLT_MAX default, ultrahdr_api.cpp:1289
    got, dest = gpu_apply(uh, sdr, mp, md, fmt, 3.4e38)
    assert st == 0 and (dest.width, dest.height, dest.colorGamut) == (w, h, 0)
    # floors: what DESIGN.md claims for the default (q-domain) kernels; the A/B and last-resort kernels a little lower
    floor = 0.995 if apply_mode in (abi.APPLY_AUTO, abi.APPLY_Q_PIPE) else 0.99
    exact = parity.compare_output(fmt, got, want, w * h, f"apply fmt={fmt} {kind} boost={max_boost:.2f}", min_exact=floor)
    print(f"apply mode={apply_mode} fmt={fmt} {kind} max_boost={max_boost:.2f}: exact {exact:.5f}")


@pytest.mark.parametrize("scale", [1, 2, 4, 8, 16])
@pytest.mark.parametrize("fmt", [abi.OUTPUT_HDR_HLG, abi.OUTPUT_HDR_LINEAR])
def test_apply_scale_factors(uh, orc, scale, fmt):
    w, h = 128 * 2, 64 * 2
    sdr_b, _ = synth.uniform_pair(w, h, 31)
    mp_b = np.random.Generator(np.random.PCG64(32)).integers(0, 256, (h // scale, w // scale), dtype=np.uint8)
    sdr, mp = abi.host_yuv420(sdr_b, w, h, 1), abi.host_map(mp_b, w // scale, h // scale)
    md = abi.default_metadata(4.926108)
    _, want, _ = orc.apply_gainmap(sdr, mp, md, fmt, 3.4e38)
    got, _ = gpu_apply(uh, sdr, mp, md, fmt, 3.4e38)
    parity.compare_output(fmt, got, want, w * h, f"apply scale={scale}", min_exact=0.995)


@pytest.mark.parametrize("apply_mode", [abi.APPLY_AUTO, abi.APPLY_Q_PIPE, abi.APPLY_TWO_TABLE], indirect=True)
@pytest.mark.parametrize("boost", [1.0, 2.0, 4.926108, 3.4e38])
@pytest.mark.parametrize("min_boost", [1.0, 0.5])
def test_apply_display_boost(uh, orc, boost, min_boost, apply_mode):
    """max_display_boost below maxContentBoost: values can exceed 1.0 and the 10-bit pack wraps
    through `& 0x3ff` (gainmapmath.cpp:722-727) — compare modulo that wrap."""
    w, h = 256, 128
    sdr_b, _ = synth.natural_pair(w, h, 41)
    mp_b = synth.natural_map(w // 4, h // 4, 42)
    sdr, mp = abi.host_yuv420(sdr_b, w, h, 0), abi.host_map(mp_b, w // 4, h // 4)
    md = abi.default_metadata(4.926108, min_boost)
    for fmt in (abi.OUTPUT_HDR_HLG, abi.OUTPUT_HDR_LINEAR):
        _, want, _ = orc.apply_gainmap(sdr, mp, md, fmt, boost)
        got, _ = gpu_apply(uh, sdr, mp, md, fmt, boost)
        if fmt == abi.OUTPUT_HDR_HLG:
            g, wv = parity.unpack1010102(got, w * h)[:, :3], parity.unpack1010102(want, w * h)[:, :3]
            d = np.abs(g - wv)
            d = np.minimum(d, 1024 - d)  # +-1 LSB across the 10-bit wrap
            assert d.max() <= 1 and (d == 0).mean() >= 0.995
        else:
            parity.compare_output(fmt, got, want, w * h, f"apply boost={boost}")


@pytest.mark.parametrize("w,h,ls,cs", [(100, 60, 118, 59), (66, 34, 67, 35), (18, 10, None, None), (36, 20, 40, 24)])
def test_apply_strides(uh, orc, w, h, ls, cs):
    scale = 2
    sdr_p, _ = synth.uniform_pair(w, h, 51)
    sdr_s, _ = synth.uniform_pair(w, h, 51, ls, cs)
    mp_b = np.random.Generator(np.random.PCG64(52)).integers(0, 256, (h // scale, w // scale), dtype=np.uint8)
    mp = abi.host_map(mp_b, w // scale, h // scale)
    md = abi.default_metadata(4.926108)
    for fmt in FORMATS:
        a, _ = gpu_apply(uh, abi.host_yuv420(sdr_p, w, h, 0), mp, md, fmt, 3.4e38)
        b, _ = gpu_apply(uh, abi.host_yuv420(sdr_s, w, h, 0, ls, cs), mp, md, fmt, 3.4e38)
        # byte-identical (jpegr_test.cpp:1485-1582): the host entry point re-strides rows the table-driven kernels cannot
        # read (odd strides) on the way up, so a strided image gets the same kernel as its packed twin
        assert np.array_equal(a, b)
        _, want, _ = orc.apply_gainmap(abi.host_yuv420(sdr_s, w, h, 0, ls, cs), mp, md, fmt, 3.4e38)
        parity.compare_output(fmt, b, want, w * h, f"apply strided fmt={fmt}")


@pytest.mark.parametrize("apply_mode", [abi.APPLY_AUTO, abi.APPLY_Q_PIPE, abi.APPLY_TWO_TABLE], indirect=True)
@pytest.mark.parametrize("w,h,ls,cs", [(128, 64, 132, 70), (36, 20, 40, 24), (260, 12, 264, 132), (4, 4, None, None)])
def test_apply_fast_path_strides(uh, orc, w, h, ls, cs, apply_mode):
    """Map scale 4 (the table-driven kernels): strided input == packed input, ragged tiles, borders."""
    sdr_p, _ = synth.uniform_pair(w, h, 53)
    sdr_s, _ = synth.uniform_pair(w, h, 53, ls, cs)
    mp_b = np.random.Generator(np.random.PCG64(54)).integers(0, 256, (h // 4, w // 4), dtype=np.uint8)
    mp = abi.host_map(mp_b, w // 4, h // 4)
    md = abi.default_metadata(4.926108)
    for fmt in FORMATS:
        a, _ = gpu_apply(uh, abi.host_yuv420(sdr_p, w, h, 0), mp, md, fmt, 3.4e38)
        b, _ = gpu_apply(uh, abi.host_yuv420(sdr_s, w, h, 0, ls, cs), mp, md, fmt, 3.4e38)
        # byte-identical (jpegr_test.cpp:1485-1582): the host entry point re-strides rows the table-driven kernels cannot
        # read (odd strides) on the way up, so a strided image gets the same kernel as its packed twin
        assert np.array_equal(a, b)
        _, want, _ = orc.apply_gainmap(abi.host_yuv420(sdr_s, w, h, 0, ls, cs), mp, md, fmt, 3.4e38)
        parity.compare_output(fmt, b, want, w * h, f"apply scale-4 strided fmt={fmt}")


# ------------------------------------------------------------------------------ convertYuv (N1)
@pytest.mark.parametrize("src,dst", [(0, 1), (0, 2), (1, 0), (1, 2), (2, 0), (2, 1)])
@pytest.mark.parametrize("w,h,ls,cs", [(256, 128, None, None), (70, 38, 75, 39), (64, 34, 68, 36), (132, 6, None, None)])
def test_convert_yuv(uh, orc, src, dst, w, h, ls, cs):
    base, _ = synth.uniform_pair(w, h, 61, ls, cs)
    a, b = base.copy(), base.copy()
    assert uh.convertYuv(abi.host_yuv420(a, w, h, src, ls, cs), src, dst) == 0
    assert orc.convert_yuv(abi.host_yuv420(b, w, h, src, ls, cs), src, dst) == 0
    # float arithmetic with the same roundings (-fmad=false): expected bit exact, bound is +-1
    exact = parity.compare_bytes(a, b, f"convertYuv {src}->{dst}", min_exact=0.9999)
    assert not np.array_equal(a, base)
    print(f"convertYuv {src}->{dst}: exact {exact:.6f}")
