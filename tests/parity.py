"""Shared helpers of the parity tests: unpackers, tolerances, comparison reports.

Tolerances (BASELINE.json north_star): gain-map bytes and 8/10-bit packed outputs within +-1 LSB
(exact-match fraction reported and bounded below); RGBA F16 linear within 1e-3 relative.
"""
import ctypes as C

import numpy as np

from libultrahdr_dev_b200 import abi

F16_REL_TOL = 1e-3          # north_star: "F16 linear output within 1e-3 relative"
F16_ABS_FLOOR = 2.0 ** -14  # below the smallest normal half, relative error is meaningless:
F16_ABS_TOL = 2.0 ** -24    # there we allow one subnormal step (2^-24)
LSB_TOL = 1                 # north_star: "+-1 LSB"


def unpack1010102(buf, n):
    w = buf[: n * 4].view(np.uint32)
    return np.stack([w & 0x3FF, (w >> 10) & 0x3FF, (w >> 20) & 0x3FF, w >> 30], axis=-1).astype(np.int32)


def unpack_f16(buf, n):
    return buf[: n * 8].view(np.float16).reshape(n, 4).astype(np.float32)


def unpack_planar10(buf, n):
    return buf[: n * 6].view(np.uint16).reshape(3, n).astype(np.int32)


def compare_bytes(got, want, what, min_exact=0.0, tol=LSB_TOL):
    """+-tol LSB on integer arrays; returns the exact-match fraction."""
    got = np.asarray(got).astype(np.int32).ravel()
    want = np.asarray(want).astype(np.int32).ravel()
    assert got.shape == want.shape, f"{what}: shape {got.shape} vs {want.shape}"
    if got.size == 0:
        return 1.0
    d = np.abs(got - want)
    exact = float((d == 0).mean())
    worst = int(d.max())
    assert worst <= tol, f"{what}: max |diff| = {worst} LSB > {tol} at flat index {int(d.argmax())} (exact {exact:.5f})"
    assert exact >= min_exact, f"{what}: exact-match fraction {exact:.5f} < {min_exact}"
    return exact


def compare_output(fmt, got, want, n, what, min_exact=0.0):
    """Compare two packed apply outputs of n pixels. Returns exact-match fraction (of channels)."""
    if fmt in (abi.OUTPUT_HDR_HLG, abi.OUTPUT_HDR_PQ):
        g, w = unpack1010102(got, n), unpack1010102(want, n)
        assert (g[:, 3] == 3).all(), f"{what}: alpha bits"
        return compare_bytes(g[:, :3], w[:, :3], what, min_exact)
    if fmt == abi.OUTPUT_HDR_LINEAR_RGB_10BIT:
        return compare_bytes(unpack_planar10(got, n), unpack_planar10(want, n), what, min_exact)
    if fmt == abi.OUTPUT_HDR_LINEAR:
        g, w = unpack_f16(got, n), unpack_f16(want, n)
        assert (got[: n * 8].view(np.uint16).reshape(n, 4)[:, 3] == 0x3C00).all(), f"{what}: alpha half"
        big = np.abs(w) >= F16_ABS_FLOOR
        rel = np.abs(g - w)[big] / np.abs(w)[big]
        if rel.size:
            assert rel.max() <= F16_REL_TOL, f"{what}: F16 max rel err {rel.max():.3e} > {F16_REL_TOL}"
        small = np.abs(g - w)[~big]
        if small.size:
            assert small.max() <= F16_ABS_TOL * 1.0001, f"{what}: F16 subnormal abs err {small.max():.3e}"
        exact = float((got[: n * 8].view(np.uint16) == want[: n * 8].view(np.uint16)).mean())
        assert exact >= min_exact, f"{what}: F16 exact-match fraction {exact:.5f} < {min_exact}"
        return exact
    raise ValueError(fmt)


def device_image_like(img, dev_luma_ptr, dev_chroma_ptr):
    """Copy of a descriptor with its plane pointers swapped for device addresses."""
    d = abi.Image()
    C.memmove(C.byref(d), C.byref(img), C.sizeof(abi.Image))
    d.data = dev_luma_ptr
    d.chroma_data = dev_chroma_ptr
    return d
