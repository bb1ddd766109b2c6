"""GPU parity on the reference's OWN image content (SURVEY 8c "(i) the 720p fixtures"): the 1280x720 raw pair that
tests/jpegr_test.cpp:1450-1571 and ProfileGainMapFuncs (:2250-2304) load, and the only photograph in the tree
(minnie-320x240.yu12, tiled to 1280x960). The CUDA path is compared with the UNMODIFIED reference hot path compiled into
oracle/_ref (falls back to the oracle restatement, which is pinned bit-for-bit to it, when oracle/_ref is absent).
The fixtures are staged by __graft_entry__.build() into tests/_ref_data/ (never committed)."""
import numpy as np
import pytest

from libultrahdr_dev_b200 import abi, refdata, synth

import parity
from test_gpu_parity import FORMATS, gpu_apply, gpu_generate

pytestmark = [pytest.mark.gpu, pytest.mark.skipif(not refdata.available(), reason="reference fixtures not staged (tests/_ref_data)")]
FLT_MAX = 3.4028234663852886e38


@pytest.fixture(scope="module")
def checker(request):
    """oracle/_ref (the reference itself) when it travelled to this box, else the pinned restatement."""
    from oracle.pyoracle import CpuImpl

    if CpuImpl.have_reference():
        return CpuImpl.reference()
    return request.getfixturevalue("orc")


@pytest.fixture(scope="module")
def raw_pair():
    p010, w, h = refdata.load("p010")
    yuv, _, _ = refdata.load("yuv420")
    return yuv, p010, w, h


@pytest.mark.parametrize("tf", [abi.TF_HLG, abi.TF_PQ])
def test_fixture_generate(uh, checker, raw_pair, tf):
    """generateGainMap on raw_p010_image.p010 + raw_yuv420_image.yuv420 (BT.2100 HDR, BT.709 SDR as in jpegr_test.cpp:2257-2270)."""
    yuv, p010, w, h = raw_pair
    sdr = abi.host_yuv420(yuv, w, h, abi.COLORGAMUT_BT709)
    hdr = abi.host_p010(p010, w, h, abi.COLORGAMUT_BT2100)
    st, want, md_o, _ = checker.generate_gainmap(sdr, hdr, tf)
    got, md, dest = gpu_generate(uh, sdr, hdr, tf)
    assert st == 0 and md.as_dict() == md_o.as_dict() and (dest.width, dest.height) == (w // 4, h // 4)
    exact = parity.compare_bytes(got, want, f"fixture map tf={tf}", min_exact=0.999)
    print(f"[{checker.name}] fixture generate tf={tf}: exact {exact:.6f}, map range {want.min()}..{want.max()}, "
          f"saturated {(want == 255).mean():.3f}")


@pytest.fixture(params=[abi.APPLY_AUTO, abi.APPLY_Q_PIPE], ids=["q", "pipe"])
def q_family(uh, request):
    """Both q-domain kernel families: the default plain-load kernels and the TMA pipeline variant."""
    assert uh.lib.uhdrb200_test_set_apply_mode(request.param) == 0
    yield request.param
    assert uh.lib.uhdrb200_test_set_apply_mode(abi.APPLY_AUTO) == 0


@pytest.mark.parametrize("fmt", FORMATS)
@pytest.mark.parametrize("tf", [abi.TF_HLG, abi.TF_PQ])
def test_fixture_apply(uh, checker, raw_pair, fmt, tf, q_family):
    """applyGainMap on the fixture SDR image with the reference's own map of the pair, all four output formats."""
    yuv, p010, w, h = raw_pair
    sdr = abi.host_yuv420(yuv, w, h, abi.COLORGAMUT_BT709)
    hdr = abi.host_p010(p010, w, h, abi.COLORGAMUT_BT2100)
    st, mp_b, md, _ = checker.generate_gainmap(sdr, hdr, tf)
    assert st == 0
    mp = abi.host_map(mp_b, w // 4, h // 4)
    _, want, _ = checker.apply_gainmap(sdr, mp, md, fmt, FLT_MAX)
    got, _ = gpu_apply(uh, sdr, mp, md, fmt, FLT_MAX)
    exact = parity.compare_output(fmt, got, want, w * h, f"fixture apply fmt={fmt} tf={tf}", min_exact=0.995)
    print(f"[{checker.name}] fixture apply fmt={fmt} tf={tf}: exact {exact:.6f}")


@pytest.mark.parametrize("fmt", FORMATS)
def test_minnie_apply(uh, checker, fmt, q_family):
    """The photograph (mean |dx| of the luma 6.2 codes, 3.7 % of the pixels on srgbInvOetf's toe), tiled 4 x 4 to
    1280x960, with a smooth full-range map; HLG-sized and PQ-sized boosts."""
    buf, w0, h0 = refdata.load("minnie")
    sdr_b, w, h = refdata.tile_yuv420(buf, w0, h0, 4, 4)
    mp_b = synth.natural_map(w // 4, h // 4, 71)
    sdr, mp = abi.host_yuv420(sdr_b, w, h, abi.COLORGAMUT_BT709), abi.host_map(mp_b, w // 4, h // 4)
    for max_boost in (1000.0 / 203.0, 10000.0 / 203.0):
        md = abi.default_metadata(max_boost)
        _, want, _ = checker.apply_gainmap(sdr, mp, md, fmt, FLT_MAX)
        got, _ = gpu_apply(uh, sdr, mp, md, fmt, FLT_MAX)
        exact = parity.compare_output(fmt, got, want, w * h, f"minnie apply fmt={fmt} boost={max_boost:.2f}", min_exact=0.995)
        print(f"[{checker.name}] minnie apply fmt={fmt} boost={max_boost:.2f}: exact {exact:.6f}")


def test_minnie_generate_roundtrip(uh, checker):
    """The photograph as the SDR image against a synthetic HDR twin (its own planes scaled into P010): generate on
    real SDR texture, then apply with that map."""
    buf, w0, h0 = refdata.load("minnie")
    sdr_b, w, h = refdata.tile_yuv420(buf, w0, h0, 2, 2)
    y = sdr_b[: w * h].reshape(h, w).astype(np.uint16)
    u = sdr_b[w * h: w * h * 5 // 4].reshape(h // 2, w // 2).astype(np.uint16)
    v = sdr_b[w * h * 5 // 4:].reshape(h // 2, w // 2).astype(np.uint16)
    hdr_b = synth.pack_p010(np.minimum(64 + y * 3, 940), 64 + u * 7 // 2, 64 + v * 7 // 2)
    sdr, hdr = abi.host_yuv420(sdr_b, w, h, abi.COLORGAMUT_BT709), abi.host_p010(hdr_b, w, h, abi.COLORGAMUT_BT2100)
    st, want, md_o, _ = checker.generate_gainmap(sdr, hdr, abi.TF_HLG)
    got, md, _ = gpu_generate(uh, sdr, hdr, abi.TF_HLG)
    exact = parity.compare_bytes(got, want, "minnie map", min_exact=0.999)
    print(f"[{checker.name}] minnie generate: exact {exact:.6f}, map range {want.min()}..{want.max()}")
    mp = abi.host_map(want, w // 4, h // 4)
    _, want_o, _ = checker.apply_gainmap(sdr, mp, md_o, abi.OUTPUT_HDR_HLG, FLT_MAX)
    got_o, _ = gpu_apply(uh, sdr, mp, md_o, abi.OUTPUT_HDR_HLG, FLT_MAX)
    parity.compare_output(abi.OUTPUT_HDR_HLG, got_o, want_o, w * h, "minnie round trip", min_exact=0.995)
