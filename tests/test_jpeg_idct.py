"""Row N4, checker side (no GPU): the oracle's restatement of the IJG ISLOW inverse DCT against libjpeg-turbo's OWN output
for the reference's JPEG fixtures — committed CRCs (tests/golden/jpeg_planes_v1.json, made by make_jpeg_golden.py through
Pillow) and, where Pillow is importable, libjpeg-turbo directly — plus closed-form block known answers."""
import ctypes as C
import json
import os
import zlib

import numpy as np
import pytest

from libultrahdr_dev_b200 import refdata

import jpeg_coeffs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
need_fixtures = pytest.mark.skipif(not all(refdata.jpeg_path(n) for n in refdata.JPEGS), reason="JPEG fixtures not staged (tests/_ref_data)")


def oracle_planes(orc, comps):
    out = []
    for c in comps:
        plane = np.zeros((c.height, c.width), np.uint8)
        co, q = np.ascontiguousarray(c.coeffs), np.ascontiguousarray(c.quant)
        fn = orc.lib.orc_jpeg_idct_plane
        fn.restype = C.c_int
        fn.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_size_t, C.c_void_p, C.c_size_t, C.c_size_t, C.c_size_t]
        assert fn(co.ctypes.data, q.ctypes.data, co.shape[1], co.shape[0], plane.ctypes.data, c.width, c.height, c.width) == 0
        out.append(plane)
    return out


@need_fixtures
@pytest.mark.parametrize("name", refdata.JPEGS)
def test_oracle_idct_equals_libjpeg_turbo_output(orc, name):
    golden = json.load(open(os.path.join(ROOT, "tests", "golden", "jpeg_planes_v1.json")))["files"][name]
    w, h, comps = jpeg_coeffs.decode(open(refdata.jpeg_path(name), "rb").read())
    planes = oracle_planes(orc, comps)
    assert len(planes) == len(golden)
    for p, g in zip(planes, golden):
        assert list(p.shape) == g["shape"] and zlib.crc32(p.tobytes()) == g["crc32"]
    try:
        from golden.make_jpeg_golden import libjpeg_planes
    except ImportError:
        return
    for p, ref in zip(planes, libjpeg_planes(refdata.jpeg_path(name))):
        assert np.array_equal(p, ref)


def test_block_known_answers(orc):
    """DC-only blocks have a closed form: every sample = range_limit(((dc * q + 4) >> 3)); a zero block is mid-grey."""
    fn = orc.lib.orc_jpeg_idct_islow_block
    fn.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    out = np.zeros(64, np.uint8)
    q = np.full(64, 16, np.uint16)
    for dc in (0, 1, -1, 5, -7, 63, -64, 100, -100, 255, -256):  # |(dc * 16 + 4) >> 3| < 512: inside the range-limit table's linear part
        co = np.zeros(64, np.int16)
        co[0] = dc
        fn(co.ctypes.data, q.ctypes.data, out.ctypes.data)
        want = min(255, max(0, ((dc * 16 + 4) >> 3) + 128))
        assert (out == want).all(), (dc, out[:4], want)
    # one AC coefficient: a cosine along x, constant along y, antisymmetric about the block centre
    co = np.zeros(64, np.int16)
    co[1] = 40
    fn(co.ctypes.data, np.ones(64, np.uint16).ctypes.data, out.ctypes.data)
    b = out.reshape(8, 8).astype(int) - 128
    assert (b == b[0]).all() and np.array_equal(b[0], -b[0][::-1]) and b[0][0] > 0
