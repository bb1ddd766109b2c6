"""Randomised (seeded) sweep of the CUDA path against the oracle, SURVEY 8c(iii): sizes (even, 8..~600), strides,
gamut pairs, transfer functions, sdr_is_601, the four output formats, max_display_boost in {1, 2, max, FLT_MAX}, map
scale factors in {1, 2, 4, 8}. Every case is small enough for the oracle to finish in milliseconds; tolerances are
the north_star ones (parity.py)."""
import numpy as np
import pytest

from libultrahdr_dev_b200 import abi, synth

import parity
from test_gpu_parity import FORMATS, gpu_apply, gpu_generate

pytestmark = pytest.mark.gpu
FLT_MAX = 3.4028234663852886e38


def _rng(seed):
    return np.random.Generator(np.random.PCG64(0xB200 + seed))


def _strides(rng, w, want_pad):
    if not want_pad:
        return None, None
    ls = w + int(rng.integers(0, 9))
    cs = max(w // 2, ls // 2) + int(rng.integers(0, 5))
    return ls, cs


@pytest.mark.parametrize("seed", range(24))
def test_sweep_generate(uh, orc, seed):
    rng = _rng(seed)
    w, h = 4 * int(rng.integers(2, 150)), 4 * int(rng.integers(2, 80))  # generate needs multiples of the map scale (4)
    tf = [abi.TF_HLG, abi.TF_PQ, abi.TF_LINEAR][int(rng.integers(0, 3))]
    sg, hg = int(rng.integers(0, 3)), int(rng.integers(0, 3))
    is601 = bool(rng.integers(0, 2))
    sls, scs = _strides(rng, w, rng.integers(0, 2))
    hls = None if rng.integers(0, 2) else w + 2 * int(rng.integers(0, 5))
    hcs = None if hls is None else hls + 2 * int(rng.integers(0, 3))
    kind = "natural" if rng.integers(0, 2) else "uniform"
    if kind == "natural" and sls is None and hls is None:
        sdr_b, hdr_b = synth.natural_pair(w, h, seed, tf, sg, hg)
    else:
        sdr_b, hdr_b = synth.uniform_pair(w, h, seed, sls, scs, hls, hcs)
    sdr = abi.host_yuv420(sdr_b, w, h, sg, sls, scs)
    hdr = abi.host_p010(hdr_b, w, h, hg, hls, hcs)
    st, want, md_o, _ = orc.generate_gainmap(sdr, hdr, tf, is601)
    got, md, dest = gpu_generate(uh, sdr, hdr, tf, is601)
    assert st == 0 and md.as_dict() == md_o.as_dict()
    assert (dest.width, dest.height) == (w // 4, h // 4)
    parity.compare_bytes(got, want, f"sweep generate seed={seed} {w}x{h} tf={tf} {sg}->{hg} 601={is601} {kind}", min_exact=0.999 if w * h >= 4096 else 0.99)


@pytest.mark.parametrize("seed", range(32))
def test_sweep_apply(uh, orc, seed):
    rng = _rng(1000 + seed)
    scale = [1, 2, 4, 4, 4, 8][int(rng.integers(0, 6))]
    w, h = 2 * scale * int(rng.integers(2, max(3, 300 // scale))), 2 * scale * int(rng.integers(2, max(3, 160 // scale)))
    w, h = max(w, 8), max(h, 8)
    fmt = FORMATS[int(rng.integers(0, 4))]
    max_boost = [1000.0 / 203.0, 10000.0 / 203.0, 2.5][int(rng.integers(0, 3))]
    min_boost = [1.0, 1.0, 0.5][int(rng.integers(0, 3))]
    disp = [1.0, 2.0, max_boost, FLT_MAX][int(rng.integers(0, 4))]
    ls, cs = _strides(rng, w, rng.integers(0, 2))
    kind = int(rng.integers(0, 3))
    if kind == 0 and ls is None:
        sdr_b, _ = synth.natural_pair(w, h, seed)
        mp_b = synth.natural_map(w // scale, h // scale, seed + 1)
    else:
        sdr_b, _ = synth.uniform_pair(w, h, seed, ls, cs)
        if kind == 2:  # dark: the linear toe of srgbInvOetf
            sdr_b = sdr_b.copy()
            n_luma = (ls or w) * h
            sdr_b[:n_luma] //= 10
            sdr_b[n_luma:] = 120 + sdr_b[n_luma:] % 17
        mp_b = rng.integers(0, 256, (h // scale, w // scale), dtype=np.uint8)
    sdr = abi.host_yuv420(sdr_b, w, h, int(rng.integers(0, 3)), ls, cs)
    mp = abi.host_map(mp_b, w // scale, h // scale)
    md = abi.default_metadata(max_boost, min_boost)
    st, want, _ = orc.apply_gainmap(sdr, mp, md, fmt, disp)
    got, dest = gpu_apply(uh, sdr, mp, md, fmt, disp)
    assert st == 0 and (dest.width, dest.height) == (w, h)
    what = f"sweep apply seed={seed} {w}x{h} scale={scale} fmt={fmt} boost={max_boost:.2f}/{min_boost} disp={disp:.3g} kind={kind}"
    if fmt in (abi.OUTPUT_HDR_HLG, abi.OUTPUT_HDR_PQ) and disp < max_boost:
        # values above 1.0 wrap through `& 0x3ff` (gainmapmath.cpp:722-727): +-1 LSB modulo 1024
        g, wv = parity.unpack1010102(got, w * h)[:, :3], parity.unpack1010102(want, w * h)[:, :3]
        d = np.abs(g - wv)
        d = np.minimum(d, 1024 - d)
        assert d.max() <= 1 and (d == 0).mean() >= 0.99, what
    elif fmt == abi.OUTPUT_HDR_LINEAR_RGB_10BIT and disp < max_boost:
        g, wv = parity.unpack_planar10(got, w * h), parity.unpack_planar10(want, w * h)
        d = np.abs(g - wv)
        d = np.minimum(d, 1024 - d)
        assert d.max() <= 1 and (d == 0).mean() >= 0.99, what
    else:
        parity.compare_output(fmt, got, want, w * h, what, min_exact=0.995 if w * h >= 4096 else 0.98)


# ------------------------------------------------------------------------------ the reference's size limits
# kMaxWidth = kMaxHeight = 8192 (lib/include/ultrahdr/jpegr.h:31-32, fuzzer/ultrahdr_enc_fuzzer.cpp:116): extreme aspect
# ratios and an almost-maximal image with padded strides, so that the 32-bit plane offsets of the kernels run near
# their limits. Uniform-random planes; the oracle (4 threads) needs a few seconds for the large case.
LIMITS = [(8192, 8, None, None, None, None), (8, 8192, None, None, None, None), (8190, 4094, 8200, 4104, 8198, 8204)]


@pytest.mark.parametrize("w,h,sls,scs,hls,hcs", LIMITS)
def test_limits_generate(uh, orc, w, h, sls, scs, hls, hcs):
    sdr_b, hdr_b = synth.uniform_pair(w, h, 77, sls, scs, hls, hcs)
    sdr = abi.host_yuv420(sdr_b, w, h, abi.COLORGAMUT_BT709, sls, scs)
    hdr = abi.host_p010(hdr_b, w, h, abi.COLORGAMUT_BT2100, hls, hcs)
    st, want, md_o, _ = orc.generate_gainmap(sdr, hdr, abi.TF_HLG)
    got, md, dest = gpu_generate(uh, sdr, hdr, abi.TF_HLG)
    assert st == 0 and md.as_dict() == md_o.as_dict() and (dest.width, dest.height) == (w // 4, h // 4)
    exact = parity.compare_bytes(got, want, f"limits generate {w}x{h}", min_exact=0.999)
    print(f"limits generate {w}x{h}: exact {exact:.6f}")


@pytest.mark.parametrize("w,h,sls,scs,scale", [(8192, 8, None, None, 4), (8, 8192, None, None, 4), (8190, 4094, 8200, 4104, 2),
                                                (8192, 4096, None, None, 4)])
@pytest.mark.parametrize("fmt", [abi.OUTPUT_HDR_HLG, abi.OUTPUT_HDR_LINEAR])
def test_limits_apply(uh, orc, w, h, sls, scs, scale, fmt):
    sdr_b, _ = synth.uniform_pair(w, h, 78, sls, scs)
    mp_b = _rng(79).integers(0, 256, (h // scale, w // scale), dtype=np.uint8)
    sdr = abi.host_yuv420(sdr_b, w, h, abi.COLORGAMUT_BT709, sls, scs)
    mp = abi.host_map(mp_b, w // scale, h // scale)
    md = abi.default_metadata(1000.0 / 203.0)
    st, want, _ = orc.apply_gainmap(sdr, mp, md, fmt, FLT_MAX)
    got, dest = gpu_apply(uh, sdr, mp, md, fmt, FLT_MAX)
    assert st == 0 and (dest.width, dest.height) == (w, h)
    exact = parity.compare_output(fmt, got, want, w * h, f"limits apply {w}x{h} scale={scale} fmt={fmt}", min_exact=0.995)
    print(f"limits apply {w}x{h} scale={scale} fmt={fmt}: exact {exact:.6f}")
