"""The C-ABI library: loads, exports every symbol include/uhdr_b200.h declares, validates arguments
in the reference's order BEFORE touching CUDA, and fails loudly (no CPU fallback) without a GPU."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from libultrahdr_dev_b200 import abi, api, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "uhdr_b200.h")


def declared_symbols(header=HEADER):
    text = open(header).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(uhdrb200_[a-z0-9_]+)\s*\(", text)))


def test_library_exports_every_declared_symbol():
    lib = api.load_library()
    syms = declared_symbols()
    assert len(syms) >= 20
    for s in syms:
        assert hasattr(lib, s), f"{s} declared in include/uhdr_b200.h but not exported"
    assert set(syms) == set(api.SYMBOLS), "python mirror and header disagree"
    test_syms = declared_symbols(os.path.join(ROOT, "include", "uhdr_b200_test.h"))
    assert set(test_syms) == set(api.TEST_SYMBOLS) and all(hasattr(lib, s) for s in test_syms)
    assert not set(test_syms) & set(syms), "test hooks must stay out of the drop-in header"
    out = subprocess.run(["nm", "-D", "--defined-only", api.LIB_PATH], capture_output=True, text=True).stdout
    exported = set(re.findall(r" T (uhdrb200_\w+)", out))
    assert set(syms) <= exported
    assert lib.uhdrb200_abi_version() == 1


def test_library_is_sm100a_only_and_does_not_link_the_oracle():
    out = subprocess.run(["cuobjdump", "-lelf", api.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs
    deps = subprocess.run(["ldd", api.LIB_PATH], capture_output=True, text=True).stdout
    assert "oracle" not in deps and "uhdr_ref" not in deps
    sym = subprocess.run(["nm", "-D", api.LIB_PATH], capture_output=True, text=True).stdout
    assert "orc_" not in sym and "ref_" not in sym


def test_struct_layouts_match_the_header():
    # ultrahdr_uncompressed_struct on LP64: ptr, size_t, size_t, int(+pad), ptr, size_t, size_t, int(+pad) = 64 bytes
    assert C.sizeof(abi.Image) == 64
    assert abi.Image.chroma_data.offset == 32 and abi.Image.luma_stride.offset == 40 and abi.Image.pixelFormat.offset == 56
    assert C.sizeof(abi.Metadata) == 36 and abi.Metadata.maxContentBoost.offset == 8


def _pair(w=32, h=16):
    sdr_b, hdr_b = synth.uniform_pair(w, h, 1)
    return sdr_b, hdr_b, abi.host_yuv420(sdr_b, w, h, 0), abi.host_p010(hdr_b, w, h, 2)


def test_argument_validation_precedes_cuda():
    """Same codes, same order as ultrahdr.cpp:189-202, 245-247, 364-406, 518-523 — none of these
    may need a device."""
    lib = api.load_library()
    uh = api.JpegR(lib)
    sdr_b, hdr_b, sdr, hdr = _pair()
    md, dest = abi.Metadata(), abi.Image()
    assert uh.generateGainMap(None, hdr, abi.TF_HLG, md, dest) == abi.ERROR_BAD_PTR
    assert uh.generateGainMap(sdr, None, abi.TF_HLG, md, dest) == abi.ERROR_BAD_PTR
    assert uh.generateGainMap(sdr, hdr, abi.TF_HLG, None, dest) == abi.ERROR_BAD_PTR
    assert uh.generateGainMap(sdr, hdr, abi.TF_HLG, md, None) == abi.ERROR_BAD_PTR
    bad = abi.host_yuv420(sdr_b, 32, 16, 0)
    bad.chroma_data = None
    assert uh.generateGainMap(bad, hdr, abi.TF_HLG, md, dest) == abi.ERROR_BAD_PTR
    assert uh.generateGainMap(sdr, abi.host_p010(hdr_b, 32, 14, 2), abi.TF_HLG, md, dest) == abi.ERROR_RESOLUTION_MISMATCH
    assert uh.generateGainMap(abi.host_yuv420(sdr_b, 32, 16, -1), hdr, abi.TF_HLG, md, dest) == abi.ERROR_INVALID_COLORGAMUT
    assert uh.generateGainMap(sdr, abi.host_p010(hdr_b, 32, 16, -1), abi.TF_HLG, md, dest) == abi.ERROR_INVALID_COLORGAMUT
    for tf in (abi.TF_SRGB, abi.TF_UNSPECIFIED, 7):
        assert uh.generateGainMap(sdr, hdr, tf, md, dest) == abi.ERROR_INVALID_TRANS_FUNC
    # resolution mismatch is reported before the gamut, the gamut before the transfer function
    assert uh.generateGainMap(abi.host_yuv420(sdr_b, 32, 16, -1), abi.host_p010(hdr_b, 32, 14, 2), abi.TF_SRGB, md, dest) == abi.ERROR_RESOLUTION_MISMATCH
    assert uh.generateGainMap(abi.host_yuv420(sdr_b, 32, 16, -1), hdr, abi.TF_SRGB, md, dest) == abi.ERROR_INVALID_COLORGAMUT

    mp_b = np.zeros((4, 8), dtype=np.uint8)
    mp = abi.host_map(mp_b, 8, 4)
    out = abi.alloc_output(abi.OUTPUT_HDR_HLG, 32, 16)
    d = abi.Image()
    d.data = out.ctypes.data
    ok = abi.default_metadata(4.0)
    assert uh.applyGainMap(None, mp, ok, abi.OUTPUT_HDR_HLG, 4.0, d) == abi.ERROR_BAD_PTR
    assert uh.applyGainMap(sdr, None, ok, abi.OUTPUT_HDR_HLG, 4.0, d) == abi.ERROR_BAD_PTR
    assert uh.applyGainMap(sdr, mp, None, abi.OUTPUT_HDR_HLG, 4.0, d) == abi.ERROR_BAD_PTR
    assert uh.applyGainMap(sdr, mp, ok, abi.OUTPUT_HDR_HLG, 4.0, None) == abi.ERROR_BAD_PTR
    for field, val in (("version", b"2.0"), ("gamma", 2.2), ("offsetSdr", 0.015625), ("offsetHdr", 0.015625),
                       ("hdrCapacityMin", 2.0), ("hdrCapacityMax", 2.0)):
        md2 = abi.default_metadata(4.0)
        setattr(md2, field, val)
        assert uh.applyGainMap(sdr, mp, md2, abi.OUTPUT_HDR_HLG, 4.0, d) == abi.ERROR_BAD_METADATA, field
    assert uh.applyGainMap(sdr, abi.host_map(mp_b, 5, 4), ok, abi.OUTPUT_HDR_HLG, 4.0, d) == abi.ERROR_UNSUPPORTED_MAP_SCALE_FACTOR
    assert uh.applyGainMap(sdr, abi.host_map(mp_b, 8, 2), ok, abi.OUTPUT_HDR_HLG, 4.0, d) == abi.ERROR_UNSUPPORTED_MAP_SCALE_FACTOR
    # unknown output format: dims are set, nothing is written, NO_ERROR (ultrahdr.cpp:411-413, 491-493) — no device needed
    d2 = abi.Image()
    d2.data = out.ctypes.data
    assert uh.applyGainMap(sdr, mp, ok, abi.OUTPUT_SDR, 4.0, d2) == abi.NO_ERROR
    assert (d2.width, d2.height, d2.colorGamut) == (32, 16, 0) and not out.any()

    tm = np.zeros(abi.yuv420_size(32, 16), dtype=np.uint8)
    assert uh.toneMap(None, abi.host_yuv420(tm, 32, 16, -1)) == abi.ERROR_BAD_PTR
    assert uh.toneMap(hdr, None) == abi.ERROR_BAD_PTR
    assert uh.toneMap(hdr, abi.host_yuv420(tm, 32, 14, -1)) == abi.ERROR_RESOLUTION_MISMATCH

    assert uh.convertYuv(None, 0, 1) == abi.ERROR_BAD_PTR
    assert uh.convertYuv(sdr, -1, 1) == abi.ERROR_INVALID_COLORGAMUT
    assert uh.convertYuv(sdr, 0, -1) == abi.ERROR_INVALID_COLORGAMUT
    keep = sdr_b.copy()
    assert uh.convertYuv(sdr, 1, 1) == abi.NO_ERROR and np.array_equal(keep, sdr_b)  # identity: untouched, no device
    assert lib.uhdrb200_test_set_apply_mode(7) != 0 and lib.uhdrb200_test_set_apply_mode(abi.APPLY_AUTO) == 0
    assert lib.uhdrb200_set_sync_mode(7) != 0 and lib.uhdrb200_set_sync_mode(abi.SYNC_BLOCKING) == 0
    assert lib.uhdrb200_set_sync_mode(abi.SYNC_SPIN) == 0


def test_no_gpu_means_loud_failure_not_a_cpu_fallback():
    lib = api.load_library()
    if lib.uhdrb200_device_count() > 0:
        pytest.skip("a CUDA device is present")
    uh = api.JpegR(lib)
    _, _, sdr, hdr = _pair()
    md, dest = abi.Metadata(), abi.Image()
    with pytest.raises(api.UltraHdrCudaError) as e:
        uh.generateGainMap(sdr, hdr, abi.TF_HLG, md, dest)
    assert "CUDA" in str(e.value) or "device" in str(e.value)
    # the raw C ABI: reference status code for resource failures + a reason string
    st = lib.uhdrb200_generate_gainmap_host(C.byref(sdr), C.byref(hdr), abi.TF_HLG, C.byref(md), C.byref(dest), 0)
    assert st == abi.ERROR_INSUFFICIENT_RESOURCE and lib.uhdrb200_last_error()
    assert dest.data is None  # nothing was produced


def test_missing_library_raises():
    with pytest.raises(FileNotFoundError):
        api.load_library("/nonexistent/libuhdr_b200.so")


def test_product_package_never_imports_the_oracle():
    pkg = os.path.join(ROOT, "libultrahdr_dev_b200")
    for dirpath, _, files in os.walk(pkg):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".cpp", ".h")):
                text = open(os.path.join(dirpath, f)).read()
                assert "pyoracle" not in text and "liboracle" not in text and "uhdr_oracle" not in text and "libuhdr_ref" not in text, f
