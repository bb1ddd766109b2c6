"""Device-side known-answer tests: the values of the reference's tests/gainmapmath_test.cpp
(:467-473 sRGB, :752-806 HLG / PQ, :980-1038 encodeGain, :1229-1262 packers) and dense grids against the oracle, applied
to the DEVICE functions of csrc/uhdr_math.cuh and to the shared-memory table evaluators of csrc/apply_q_common.cuh one
function at a time (uhdrb200_test_eval, include/uhdr_b200_test.h), so that a compensating pair of errors inside a kernel
cannot hide behind whole-image parity."""
import numpy as np
import pytest

from libultrahdr_dev_b200 import abi

pytestmark = pytest.mark.gpu
EPS = 1e-4  # ComparisonEpsilon, tests/gainmapmath_test.cpp:29
N_GRID = 1 << 20


@pytest.fixture(scope="module")
def ev(uh):
    import torch

    assert torch.cuda.is_available()

    def run(fn, x, pairs=False):
        x = np.ascontiguousarray(x, dtype=np.float32)
        n = x.size // (2 if pairs else 1)
        d_in = torch.from_numpy(x.ravel()).cuda()
        d_out = torch.empty(n, dtype=torch.float32, device="cuda")
        st = uh.lib.uhdrb200_test_eval(fn, d_in.data_ptr(), d_out.data_ptr(), n, torch.cuda.current_stream().cuda_stream)
        assert st == 0, uh.lib.uhdrb200_last_error()
        torch.cuda.synchronize()
        return d_out.cpu().numpy()

    return run


def orc_vec(orc, name, x):
    f = getattr(orc, "c_" + name)
    return np.array([f(float(v)) for v in x], dtype=np.float32)


def rel_err(got, want, floor):
    got, want = got.astype(np.float64), want.astype(np.float64)
    return float(np.max(np.abs(got - want) / np.maximum(np.abs(want), floor)))


# ---------------------------------------------------------------- the reference's own known answers
def test_kat_srgb(ev):  # gainmapmath_test.cpp:467-473
    got = ev(abi.FN_SRGB_INV_OETF, [0.0, 0.02, 0.04045, 0.5, 1.0])
    assert got[0] == 0.0
    for g, want in zip(got[1:4], (0.00154, 0.00313, 0.21404)):
        assert abs(g - want) < EPS
    assert got[4] == pytest.approx(1.0, rel=1e-6)


def test_kat_hlg(ev):  # :752-782
    x = [0.0, 0.04167, 0.08333, 0.5, 1.0]
    got = ev(abi.FN_HLG_OETF, x)
    assert got[0] == 0.0 and got[4] == pytest.approx(1.0, rel=1e-6)
    for g, want in zip(got[1:4], (0.35357, 0.5, 0.87164)):
        assert abs(g - want) < EPS
    back = ev(abi.FN_HLG_INV_OETF, got)
    assert np.max(np.abs(back - np.array(x, dtype=np.float32))) < EPS
    inv = ev(abi.FN_HLG_INV_OETF, [0.0, 0.25, 0.5, 0.75, 1.0])
    assert inv[0] == 0.0 and inv[4] == pytest.approx(1.0, rel=1e-6)
    for g, want in zip(inv[1:4], (0.02083, 0.08333, 0.26496)):
        assert abs(g - want) < EPS


def test_kat_pq(ev):  # :784-806, 941-947
    x = [0.0, 0.01, 0.5, 0.99, 1.0]
    got = ev(abi.FN_PQ_OETF, x)
    assert got[0] == 0.0 and got[4] == pytest.approx(1.0, rel=2e-6)
    for g, want in zip(got[1:4], (0.50808, 0.92655, 0.99895)):
        assert abs(g - want) < EPS
    back = ev(abi.FN_PQ_INV_OETF, got)
    assert np.max(np.abs(back - np.array(x, dtype=np.float32))) < EPS
    inv = ev(abi.FN_PQ_INV_OETF, [0.0, 0.01, 0.5, 0.99, 1.0])
    assert inv[0] == 0.0 and inv[4] == pytest.approx(1.0, rel=1e-5)
    for g, want in zip(inv[1:4], (2.31017e-7, 0.00922, 0.90903)):
        assert abs(g - want) < EPS


def test_kat_packers(ev):  # :1229-1262
    fmax = 3.4028234663852886e38
    vals = [0.1, 0.0, 1.0, -1.0, fmax, -fmax, 1.1754943508222875e-38, 0.2, 0.3]
    got = ev(abi.FN_FLOAT_TO_HALF, vals).astype(np.int64)
    assert [int(v) for v in got] == [0x2E66, 0x0, 0x3C00, 0xBC00, 0x7FFF, 0xFFFF, 0x0, 0x3266, 0x34CD]
    ten = ev(abi.FN_TO_10BIT, [0.0, 1.0, 0.1, 0.2, 0.3]).astype(np.int64)
    want = [0, 1023] + [int(np.float32(v) * np.float32(1023.0)) for v in (0.1, 0.2, 0.3)]
    assert [int(v) for v in ten] == want


def test_kat_encode_gain(ev, orc):  # :980-1038 restated for the metadata generateGainMap writes (min 1, max 1000/203)
    mx = float(np.float32(1000.0 / 203.0))
    cases = [(0.0, 0.0), (1.0, 0.0), (1.0, 1.0), (1.0, mx), (1.0, 10.0), (2.0, 1.0), (1.0, 2.0), (1.0, mx ** 0.5), (1.0, mx ** 0.25), (3.0, 9.0)]
    got = ev(abi.FN_ENCODE_GAIN, np.array(cases, dtype=np.float32), pairs=True).astype(np.int64)
    for (s, h), g in zip(cases, got):
        gain = float(np.float32(h) / np.float32(s)) if s > 0 else 1.0
        want = orc.c_encodeGain(float(s), float(h), 1.0, mx)
        assert abs(int(g) - int(want)) <= 1, (s, h, g, want)
        if gain <= 1.0 or gain >= mx:
            assert int(g) == int(want), (s, h, g, want)  # clamped gains: the host-computed exact codes


# ---------------------------------------------------------------- dense grids against the oracle (max error printed)
@pytest.mark.parametrize("fn,name,lo,hi,tol,floor", [
    (abi.FN_SRGB_INV_OETF, "srgbInvOetf", 0.0, 1.0, 3e-6, 1e-4),
    (abi.FN_HLG_OETF, "hlgOetf", 0.0, 1.0, 2e-6, 1e-3),
    (abi.FN_HLG_INV_OETF, "hlgInvOetf", 0.0, 1.0, 3e-6, 1e-4),
    (abi.FN_PQ_OETF, "pqOetf", 0.0, 1.0, 4e-6, 1e-2),
    (abi.FN_PQ_INV_OETF, "pqInvOetf", 0.0, 1.0, 3e-5, 1e-6),
])
def test_grid_transfer_functions(ev, orc, fn, name, lo, hi, tol, floor):
    x = np.linspace(lo, hi, N_GRID, dtype=np.float32)
    got = ev(fn, x)
    sel = np.random.Generator(np.random.PCG64(fn)).choice(N_GRID, 20000, replace=False)  # the scalar oracle on a sample
    want = orc_vec(orc, name, x[sel])
    e = rel_err(got[sel], want, floor)
    print(f"{name}: max relative error {e:.3e} on {sel.size} of {N_GRID} grid points (floor {floor})")
    assert e <= tol
    assert np.all(np.isfinite(got)) and np.all(np.diff(got.astype(np.float64)) >= -1e-6)  # monotone up to rounding


def test_grid_float_to_half_bit_exact(ev, orc):
    rng = np.random.Generator(np.random.PCG64(5))
    bits = rng.integers(0, 1 << 32, N_GRID, dtype=np.uint64).astype(np.uint32)
    bits = bits[((bits >> 23) & 0xFF) != 0xFF]  # finite values
    x = bits.view(np.float32)
    got = ev(abi.FN_FLOAT_TO_HALF, x).astype(np.int64)
    sel = np.concatenate([np.arange(0, x.size, 53), np.flatnonzero((np.abs(x) > 2.0 ** -26) & (np.abs(x) < 2.0 ** -13))[:20000]])
    want = np.array([orc.c_floatToHalf(float(v)) for v in x[sel]], dtype=np.int64)
    assert np.array_equal(got[sel], want)


def _q_grid():  # the kernels keep f < 3.9 (capi.cu), i.e. q < 3.9^(1/2.4) = 1.763
    return np.exp(np.linspace(np.log(2.0 ** -11), np.log(1.76), N_GRID)).astype(np.float32)


@pytest.mark.parametrize("fn,oetf", [(abi.FN_P_HLG, "hlgOetf"), (abi.FN_P_PQ, "pqOetf")])
def test_grid_p_table(ev, orc, fn, oetf):
    """P(q) = 1023 * OETF(q^2.4) (the apply kernels' single look-up) against the oracle's srgb-free chain in double."""
    q = _q_grid()
    got = ev(fn, q).astype(np.float64)
    sel = np.random.Generator(np.random.PCG64(fn)).choice(N_GRID, 20000, replace=False)
    lin = q[sel].astype(np.float64) ** 2.4
    want = 1023.0 * np.array([getattr(orc, "c_" + oetf)(float(v)) for v in lin.astype(np.float32)], dtype=np.float64)
    err = np.abs(got[sel] - want)
    print(f"P table ({oetf}): max abs error {err.max():.4f} codes, mean {err.mean():.5f}")
    assert err.max() < 0.02 and got.min() >= 0.0


def test_grid_l_table(ev):
    q = np.exp(np.linspace(np.log(2.0 ** -6), np.log(1.999), N_GRID)).astype(np.float32)
    got = ev(abi.FN_L_POW, q).astype(np.float64)
    want = q.astype(np.float64) ** 2.4
    e = float(np.max(np.abs(got - want) / want))
    print(f"L table: max relative error {e:.3e}")
    assert e < 2e-6


def test_grid_toe_fix(ev, orc):
    """Phi -> Phi' with Phi'^2.4 == srgbInvOetf(c), c = 1.055 Phi - 0.055: identity above the toe, the exact power on it."""
    phi = np.linspace(0.0, 0.2, 1 << 16, dtype=np.float32)
    got = ev(abi.FN_TOE_FIX, phi).astype(np.float64)
    c = np.clip(phi.astype(np.float64) * 1.055 - 0.055, 0.0, 1.0)
    want = np.where(c <= 0.04045, c / 12.92, ((c + 0.055) / 1.055) ** 2.4)
    err = np.abs(got ** 2.4 - want) / np.maximum(want, 1e-5)
    print(f"toe select: max relative error of Phi'^2.4 vs srgbInvOetf {err.max():.3e}")
    assert err.max() < 2e-5
    assert np.array_equal(got[phi > 0.0905], phi[phi > 0.0905].astype(np.float64))
