"""Committed golden vectors (tests/golden/golden_v1.npz, generated from the unmodified reference by
tests/golden/make_golden.py): the oracle must reproduce them bit for bit (CPU), the CUDA path must
match them within the north-star tolerance (GPU: +-1 LSB, F16 1e-3 relative; integer stages exact)."""
import os
import sys

import numpy as np
import pytest

from libultrahdr_dev_b200 import abi

import parity

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "golden"))
import make_golden as mg  # noqa: E402


@pytest.fixture(scope="module")
def golden():
    return np.load(os.path.join(HERE, "golden", "golden_v1.npz"))


def test_oracle_reproduces_golden_bit_for_bit(orc, golden):
    got = mg.run_all(orc)
    assert set(got) == set(golden.files)
    for k in golden.files:
        assert np.array_equal(np.asarray(got[k]), golden[k]), k


def test_reference_build_reproduces_golden(ref, golden):
    got = mg.run_all(ref)
    for k in golden.files:
        assert np.array_equal(np.asarray(got[k]), golden[k]), k


@pytest.mark.gpu
def test_cuda_generate_vs_golden(uh, golden):
    from libultrahdr_dev_b200.api import take_map

    for case in mg.GENERATE_CASES:
        name, w, h, kind, tf, sg, hg, is601, strides = case
        st = strides or (None, None, None, None)
        sdr_b, hdr_b = mg.gen_inputs(case)
        assert mg.sha(sdr_b, hdr_b) == str(golden["gen/" + name + "/in_sha"])
        md, dest = abi.Metadata(), abi.Image()
        assert uh.generateGainMap(abi.host_yuv420(sdr_b, w, h, sg, st[0], st[1]), abi.host_p010(hdr_b, w, h, hg, st[2], st[3]), tf, md, dest, is601) == 0
        got = take_map(dest)
        assert np.float32(md.maxContentBoost) == golden["gen/" + name + "/max_boost"]
        exact = parity.compare_bytes(got, golden["gen/" + name], "golden gen " + name, min_exact=0.995)
        print(f"golden generate {name}: exact {exact:.5f}")


@pytest.mark.gpu
def test_cuda_apply_vs_golden(uh, golden):
    for case in mg.APPLY_CASES:
        name, w, h, kind, scale, mx, mn, db = case
        sdr_b, mp_b = mg.apply_inputs(case)
        assert mg.sha(sdr_b, mp_b) == str(golden["apply/" + name + "/in_sha"])
        md = abi.default_metadata(mx, mn)
        for fmt in mg.FORMATS:
            out = abi.alloc_output(fmt, w, h)
            d = abi.Image()
            d.data = out.ctypes.data
            assert uh.applyGainMap(abi.host_yuv420(sdr_b, w, h, 0), abi.host_map(mp_b, w // scale, h // scale), md, fmt, db, d) == 0
            want = golden["apply/%s/fmt%d" % (name, fmt)]
            if fmt in (abi.OUTPUT_HDR_HLG, abi.OUTPUT_HDR_PQ) and db < mx:  # codes wrap through & 0x3ff
                g, wv = parity.unpack1010102(out, w * h)[:, :3], parity.unpack1010102(want, w * h)[:, :3]
                dd = np.abs(g - wv)
                assert np.minimum(dd, 1024 - dd).max() <= 1
                continue
            if fmt == abi.OUTPUT_HDR_LINEAR_RGB_10BIT and db < mx:
                g, wv = parity.unpack_planar10(out, w * h), parity.unpack_planar10(want, w * h)
                dd = np.abs(g - wv)
                assert np.minimum(dd, 1024 - dd).max() <= 1
                continue
            exact = parity.compare_output(fmt, out, want, w * h, f"golden apply {name} fmt{fmt}", min_exact=0.99)
            print(f"golden apply {name} fmt {fmt}: exact {exact:.5f}")


@pytest.mark.gpu
def test_cuda_tonemap_and_convert_vs_golden_exact(uh, golden):
    for case in mg.TONEMAP_CASES:
        name, w, h, ls, cs, dls, dcs = case
        hdr_b = mg.tonemap_input(case)
        buf = np.full(abi.yuv420_size(w, h, dls, dcs), 0x55, dtype=np.uint8)
        d = abi.host_yuv420(buf, w, h, -1, dls, dcs)
        assert uh.toneMap(abi.host_p010(hdr_b, w, h, 2, ls, cs), d) == 0
        assert np.array_equal(buf, golden["tonemap/" + name]), name
    for case in mg.CONVERT_CASES:
        name, w, h, s, d = case
        buf = mg.convert_input(case)
        assert uh.convertYuv(abi.host_yuv420(buf, w, h, s), s, d) == 0
        parity.compare_bytes(buf, golden["convert/" + name], "golden " + name, min_exact=0.9999)
