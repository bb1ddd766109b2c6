"""GPU parity of the LUT numeric mode (UHDRB200_MODE_LUT, include/uhdr_b200.h): the CUDA path against the reference's
`-DUSE_*_LUT=1` build (oracle/_ref/libuhdr_ref_lut.so, what lib/src/jpegr.cpp:33-38 intends) and against the oracle's LUT
restatement (pinned to that build bit for bit, tests/test_oracle_vs_ref.py). In this mode the kernels follow the
reference's float arithmetic operation by operation and read tables built by the same host expressions, so the bar is
BYTE-IDENTICAL output, not +-1 LSB."""
import threading

import numpy as np
import pytest

from libultrahdr_dev_b200 import abi, refdata, synth
from test_gpu_parity import FORMATS, gpu_apply, gpu_generate

pytestmark = pytest.mark.gpu
FLT_MAX = 3.4028234663852886e38


@pytest.fixture
def lut(uh):
    """The calling thread's numeric mode set to LUT for one test."""
    assert uh.lib.uhdrb200_get_numeric_mode() == abi.MODE_ANALYTIC
    assert uh.lib.uhdrb200_set_numeric_mode(abi.MODE_LUT) == 0
    yield
    assert uh.lib.uhdrb200_set_numeric_mode(abi.MODE_ANALYTIC) == 0


@pytest.fixture(params=["oracle-lut", "reference-lut"])
def checker(request, orc):
    """Both checkers: the oracle switched to its LUT variant (a process-wide switch of liboracle.so, restored afterwards)
    and the reference's own LUT build when oracle/_ref travelled to this box."""
    from oracle.pyoracle import CpuImpl

    if request.param == "reference-lut":
        if not CpuImpl.have_reference(lut=True):
            pytest.skip("oracle/_ref/libuhdr_ref_lut.so not on this box")
        yield CpuImpl.reference(lut=True)
        return
    o = CpuImpl.oracle(lut=True)
    yield o
    o.lib.orc_set_lut_mode(0)


# ------------------------------------------------------------------------------ generateGainMap
@pytest.mark.parametrize("tf", [abi.TF_HLG, abi.TF_PQ, abi.TF_LINEAR])
@pytest.mark.parametrize("sg,hg", [(0, 2), (1, 2), (2, 2), (0, 0), (2, 1)])
@pytest.mark.parametrize("kind", ["natural", "uniform"])
def test_lut_generate_bit_exact(uh, lut, checker, tf, sg, hg, kind):
    w, h = 640, 360
    if kind == "natural":
        sdr_b, hdr_b = synth.natural_pair(w, h, 111 + tf, tf, sg, hg)
    else:
        sdr_b, hdr_b = synth.uniform_pair(w, h, 117 + tf)
    sdr, hdr = abi.host_yuv420(sdr_b, w, h, sg), abi.host_p010(hdr_b, w, h, hg)
    st, want, md_o, _ = checker.generate_gainmap(sdr, hdr, tf)
    got, md_g, _ = gpu_generate(uh, sdr, hdr, tf)
    assert st == 0 and md_g.as_dict() == md_o.as_dict()
    n_diff = int((got != want).sum())
    print(f"[{checker.name}] LUT generate tf={tf} {sg}<-{hg} {kind}: {n_diff} of {want.size} bytes differ")
    assert n_diff == 0


@pytest.mark.parametrize("w,h,sls,scs,hls,hcs", [
    (100, 60, 118, 59, 118, 118),
    (64, 32, 67, 35, 65, 70),          # odd strides -> scalar gather
    (18, 10, None, None, None, None),  # dims not multiples of 4
    (4, 4, None, None, None, None),
])
def test_lut_generate_strides(uh, lut, checker, w, h, sls, scs, hls, hcs):
    sdr_b, hdr_b = synth.uniform_pair(w, h, 109, sls, scs, hls, hcs)
    sdr, hdr = abi.host_yuv420(sdr_b, w, h, 0, sls, scs), abi.host_p010(hdr_b, w, h, 2, hls, hcs)
    for tf in (abi.TF_HLG, abi.TF_PQ):
        _, want, _, _ = checker.generate_gainmap(sdr, hdr, tf, sdr_is_601=True)
        got, _, _ = gpu_generate(uh, sdr, hdr, tf, sdr_is_601=True)
        assert np.array_equal(got, want)


# ------------------------------------------------------------------------------ applyGainMap
@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("kind", ["natural", "uniform", "dark"])
@pytest.mark.parametrize("max_boost", [1000.0 / 203.0, 10000.0 / 203.0])
def test_lut_apply_bit_exact(uh, lut, checker, fmt, kind, max_boost):
    w, h = 512, 288
    if kind == "natural":
        sdr_b, _ = synth.natural_pair(w, h, 121)
        mp_b = synth.natural_map(w // 4, h // 4, 122)
    elif kind == "dark":
        rng = np.random.Generator(np.random.PCG64(125))
        sdr_b = np.concatenate([rng.integers(0, 25, w * h, dtype=np.uint8), rng.integers(120, 137, w * h // 2, dtype=np.uint8)])
        mp_b = synth.natural_map(w // 4, h // 4, 126)
    else:
        sdr_b, _ = synth.uniform_pair(w, h, 123)
        mp_b = np.random.Generator(np.random.PCG64(124)).integers(0, 256, (h // 4, w // 4), dtype=np.uint8)
    sdr, mp = abi.host_yuv420(sdr_b, w, h, 0), abi.host_map(mp_b, w // 4, h // 4)
    md = abi.default_metadata(max_boost)
    st, want, _ = checker.apply_gainmap(sdr, mp, md, fmt, FLT_MAX)
    got, dest = gpu_apply(uh, sdr, mp, md, fmt, FLT_MAX)
    assert st == 0 and (dest.width, dest.height, dest.colorGamut) == (w, h, 0)
    n_diff = int((got != want).sum())
    print(f"[{checker.name}] LUT apply fmt={fmt} {kind} boost={max_boost:.2f}: {n_diff} of {want.size} bytes differ")
    assert n_diff == 0


@pytest.mark.parametrize("scale", [1, 2, 4, 8, 16, 32])
def test_lut_apply_scale_factors(uh, lut, checker, scale):
    w, h = 256, 128
    sdr_b, _ = synth.uniform_pair(w, h, 131)
    mp_b = np.random.Generator(np.random.PCG64(132)).integers(0, 256, (h // scale, w // scale), dtype=np.uint8)
    sdr, mp = abi.host_yuv420(sdr_b, w, h, 1), abi.host_map(mp_b, w // scale, h // scale)
    md = abi.default_metadata(4.926108)
    for fmt in (abi.OUTPUT_HDR_HLG, abi.OUTPUT_HDR_LINEAR):
        _, want, _ = checker.apply_gainmap(sdr, mp, md, fmt, FLT_MAX)
        got, _ = gpu_apply(uh, sdr, mp, md, fmt, FLT_MAX)
        assert np.array_equal(got, want), (scale, fmt)


@pytest.mark.parametrize("boost", [1.0, 2.0, 4.926108, FLT_MAX])
@pytest.mark.parametrize("min_boost", [1.0, 0.5])
def test_lut_apply_display_boost(uh, lut, checker, boost, min_boost):
    """max_display_boost below maxContentBoost: values above 1.0 index past the OETF tables (clipped to the last entry,
    gainmapmath.cpp:272) and the linear 10-bit pack wraps through `& 0x3ff` exactly like the reference."""
    w, h = 256, 128
    sdr_b, _ = synth.natural_pair(w, h, 141)
    mp_b = synth.natural_map(w // 4, h // 4, 142)
    sdr, mp = abi.host_yuv420(sdr_b, w, h, 0), abi.host_map(mp_b, w // 4, h // 4)
    md = abi.default_metadata(4.926108, min_boost)
    for fmt in FORMATS:
        _, want, _ = checker.apply_gainmap(sdr, mp, md, fmt, boost)
        got, _ = gpu_apply(uh, sdr, mp, md, fmt, boost)
        assert np.array_equal(got, want), (fmt, boost, min_boost)


@pytest.mark.parametrize("w,h,ls,cs", [(100, 60, 118, 59), (66, 34, 67, 35), (18, 10, None, None)])
def test_lut_apply_strides(uh, lut, checker, w, h, ls, cs):
    sdr_b, _ = synth.uniform_pair(w, h, 151, ls, cs)
    mp_b = np.random.Generator(np.random.PCG64(152)).integers(0, 256, (h // 2, w // 2), dtype=np.uint8)
    sdr, mp = abi.host_yuv420(sdr_b, w, h, 0, ls, cs), abi.host_map(mp_b, w // 2, h // 2)
    md = abi.default_metadata(4.926108)
    for fmt in FORMATS:
        _, want, _ = checker.apply_gainmap(sdr, mp, md, fmt, FLT_MAX)
        got, _ = gpu_apply(uh, sdr, mp, md, fmt, FLT_MAX)
        assert np.array_equal(got, want), fmt


# ------------------------------------------------------------------------------ the reference's own fixtures
@pytest.mark.skipif(not refdata.available(), reason="reference fixtures not staged (tests/_ref_data)")
@pytest.mark.parametrize("tf", [abi.TF_HLG, abi.TF_PQ])
def test_lut_fixture_chain(uh, lut, checker, tf):
    """raw_p010_image.p010 + raw_yuv420_image.yuv420 (jpegr_test.cpp:2257-2270): map, then all four output formats."""
    p010, w, h = refdata.load("p010")
    yuv, _, _ = refdata.load("yuv420")
    sdr, hdr = abi.host_yuv420(yuv, w, h, abi.COLORGAMUT_BT709), abi.host_p010(p010, w, h, abi.COLORGAMUT_BT2100)
    st, want_map, md, _ = checker.generate_gainmap(sdr, hdr, tf)
    got_map, md_g, _ = gpu_generate(uh, sdr, hdr, tf)
    assert st == 0 and md_g.as_dict() == md.as_dict()
    assert np.array_equal(got_map, want_map)
    mp = abi.host_map(want_map, w // 4, h // 4)
    for fmt in FORMATS:
        _, want, _ = checker.apply_gainmap(sdr, mp, md, fmt, FLT_MAX)
        got, _ = gpu_apply(uh, sdr, mp, md, fmt, FLT_MAX)
        n_diff = int((got != want).sum())
        print(f"[{checker.name}] LUT fixture apply tf={tf} fmt={fmt}: {n_diff} of {want.size} bytes differ")
        assert n_diff == 0


# ------------------------------------------------------------------------------ API-0 (fused toneMap + generate)
@pytest.mark.parametrize("tf", [abi.TF_HLG, abi.TF_PQ])
@pytest.mark.parametrize("to601", [False, True])
def test_lut_api0_fused(uh, lut, checker, tf, to601):
    """API-0 in LUT mode: SDR planes == toneMap (+ convertYuv), map == the LUT build's generateGainMap on them."""
    w, h = 256, 128
    _, hdr_b = synth.natural_pair(w, h, 161, tf, 2, 2)
    hdr = abi.host_p010(hdr_b, w, h, abi.COLORGAMUT_BT2100)
    sdr_g = np.zeros(abi.yuv420_size(w, h), dtype=np.uint8)
    sdr_d = abi.host_yuv420(sdr_g, w, h, -1)
    md, dest = abi.Metadata(), abi.Image()
    map_buf = np.zeros((h // 4) * (w // 4), dtype=np.uint8)
    assert uh.toneMapGenerateGainMapInto(hdr, tf, sdr_d, md, dest, map_buf, sdr_to_601=to601) == 0
    sdr_o = np.zeros_like(sdr_g)
    sdr_od = abi.host_yuv420(sdr_o, w, h, -1)
    assert checker.tonemap(hdr, sdr_od) == 0
    st, want, md_o, _ = checker.generate_gainmap(sdr_od, hdr, tf)
    assert st == 0 and md.as_dict() == md_o.as_dict()
    assert np.array_equal(map_buf.reshape(h // 4, w // 4), want.reshape(h // 4, w // 4))
    if to601:
        assert checker.convert_yuv(sdr_od, abi.COLORGAMUT_BT2100, abi.COLORGAMUT_P3) == 0
    assert np.array_equal(sdr_g, sdr_o)  # toneMap and convertYuv are mode-free and byte-identical


# ------------------------------------------------------------------------------ the mode itself
def test_lut_mode_is_a_mode(uh, orc):
    """The switch changes results (the two builds disagree by several LSB), is restored exactly, rejects
    unknown values and is a property of the calling thread."""
    w, h = 256, 128
    sdr_b, _ = synth.natural_pair(w, h, 171)
    mp_b = synth.natural_map(w // 4, h // 4, 172)
    sdr, mp = abi.host_yuv420(sdr_b, w, h, 0), abi.host_map(mp_b, w // 4, h // 4)
    md = abi.default_metadata(10000.0 / 203.0)
    a0, _ = gpu_apply(uh, sdr, mp, md, abi.OUTPUT_HDR_PQ, FLT_MAX)
    assert uh.lib.uhdrb200_set_numeric_mode(7) == abi.ERROR_UNSUPPORTED_FEATURE
    assert uh.lib.uhdrb200_get_numeric_mode() == abi.MODE_ANALYTIC
    assert uh.lib.uhdrb200_set_numeric_mode(abi.MODE_LUT) == 0
    try:
        seen = []
        t = threading.Thread(target=lambda: seen.append(uh.lib.uhdrb200_get_numeric_mode()))
        t.start()
        t.join()
        assert seen == [abi.MODE_ANALYTIC]  # another thread keeps the default
        a1, _ = gpu_apply(uh, sdr, mp, md, abi.OUTPUT_HDR_PQ, FLT_MAX)
    finally:
        assert uh.lib.uhdrb200_set_numeric_mode(abi.MODE_ANALYTIC) == 0
    a2, _ = gpu_apply(uh, sdr, mp, md, abi.OUTPUT_HDR_PQ, FLT_MAX)
    assert np.array_equal(a0, a2) and not np.array_equal(a0, a1)


def test_lut_mode_gain_table_cache(uh, lut, checker):
    """More distinct (metadata, display boost) combinations than the gain-table cache holds (capi.cu: kGainLutMax = 256):
    the cache is drained and rebuilt, results stay exact."""
    w, h = 64, 32
    sdr_b, _ = synth.natural_pair(w, h, 181)
    mp_b = synth.natural_map(w // 4, h // 4, 182)
    sdr, mp = abi.host_yuv420(sdr_b, w, h, 0), abi.host_map(mp_b, w // 4, h // 4)
    for i in range(300):
        mx = 2.0 + 0.01 * i
        md = abi.default_metadata(mx)
        got, _ = gpu_apply(uh, sdr, mp, md, abi.OUTPUT_HDR_HLG, FLT_MAX)
        if i % 37 == 0 or i == 299:
            _, want, _ = checker.apply_gainmap(sdr, mp, md, abi.OUTPUT_HDR_HLG, FLT_MAX)
            assert np.array_equal(got, want), i


def test_lut_mode_noalloc_rejected(uh, lut):
    """The no-allocation entry point serves the analytic mode only (the LUT tables live in allocated device memory)."""
    imgs, mds = (abi.Image * 1)(), (abi.Metadata * 1)()
    st = uh.lib.uhdrb200_generate_gainmap_batch_dev_noalloc(1, imgs, imgs, abi.TF_HLG, mds, imgs, 0, None)
    assert st == abi.ERROR_UNSUPPORTED_FEATURE
    assert b"MODE_ANALYTIC" in uh.lib.uhdrb200_last_error()
