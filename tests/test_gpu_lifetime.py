"""Context lifetime of the C ABI: resources of a calling thread go away with the thread, a rebind to
another GPU tears down everything the old device owned, *_dev entry points follow the caller's current device, and the
API-0 host entry point stages enough of a P010 image whose luma_stride is 0."""
import ctypes as C
import threading

import numpy as np
import pytest

from libultrahdr_dev_b200 import abi, synth
from libultrahdr_dev_b200.api import take_map

import parity

pytestmark = pytest.mark.gpu
FLT_MAX = 3.4028234663852886e38


def _host_roundtrip(lib, w, h, seed):
    sdr_b, hdr_b = synth.uniform_pair(w, h, seed)
    s_img, h_img = abi.host_yuv420(sdr_b, w, h, 0), abi.host_p010(hdr_b, w, h, 2)
    mp, out = np.zeros(w * h // 16, np.uint8), np.zeros(w * h * 4, np.uint8)
    md, m_img, o_img = abi.Metadata(), abi.Image(), abi.Image()
    st = lib.uhdrb200_generate_gainmap_host_into(C.byref(s_img), C.byref(h_img), abi.TF_HLG, C.byref(md), C.byref(m_img), 0, mp.ctypes.data, mp.size)
    assert st == 0, lib.uhdrb200_last_error()
    o_img.data = out.ctypes.data
    st = lib.uhdrb200_apply_gainmap_host(C.byref(s_img), C.byref(m_img), C.byref(md), abi.OUTPUT_HDR_HLG, FLT_MAX, C.byref(o_img))
    assert st == 0, lib.uhdrb200_last_error()
    return mp, out


def test_thread_exit_releases_its_context(uh):
    """A thread-per-request caller must not leak: 24 short-lived threads each stage a 1080p pair (~25 MB of device staging
    and 16 MB of page-locked chunks per thread); free device memory afterwards is what it was after the first."""
    import torch

    lib = uh.lib
    results = []

    def worker(i):
        results.append(_host_roundtrip(lib, 1920, 1080, 700 + i % 2)[1][:64].copy())

    def run(n):
        for i in range(n):
            t = threading.Thread(target=worker, args=(i,))
            t.start()
            t.join()

    run(4)
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    run(24)
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    # a leaked context would cost ~25 MB of device staging per thread (600 MB here); the driver's own pools may grow a little
    assert free0 - free1 < (64 << 20), f"device memory shrank by {(free0 - free1) >> 20} MB over 24 thread lifetimes"
    assert all(np.array_equal(results[i], results[i % 2]) for i in range(len(results)))


def test_shutdown_then_reuse(uh):
    lib = uh.lib
    a = _host_roundtrip(lib, 256, 128, 3)
    lib.uhdrb200_shutdown()
    lib.uhdrb200_shutdown()  # idempotent
    b = _host_roundtrip(lib, 256, 128, 3)
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])


def test_api0_host_with_luma_stride_zero(uh, orc):
    """luma_stride == 0: toneMap takes it literally (every row reads row 0, ultrahdr.cpp:527) and generateGainMap reads it
    as `width` (gainmapmath.cpp:585); the host entry point must stage the whole image for the second."""
    w, h = 256, 64
    _, hdr_b = synth.uniform_pair(w, h, 41)
    outs = []
    for impl in ("gpu", "orc"):
        hdr = abi.host_p010(hdr_b, w, h, abi.COLORGAMUT_BT2100)
        hdr.luma_stride = 0
        sdr_buf = np.full(abi.yuv420_size(w, h), 0x55, dtype=np.uint8)
        sdr = abi.host_yuv420(sdr_buf, w, h, abi.COLORGAMUT_UNSPECIFIED)
        md, dest = abi.Metadata(), abi.Image()
        if impl == "gpu":
            mp = np.zeros(w * h // 16, np.uint8)
            assert uh.toneMapGenerateGainMapInto(hdr, abi.TF_HLG, sdr, md, dest, mp) == 0
        else:
            assert orc.tonemap(hdr, sdr) == 0
            st, mp, md, _ = orc.generate_gainmap(sdr, hdr, abi.TF_HLG)
            assert st == 0
        outs.append((sdr_buf, np.asarray(mp).ravel().copy()))
    assert np.array_equal(outs[0][0], outs[1][0])
    parity.compare_bytes(outs[0][1], outs[1][1], "API-0 map with luma_stride 0", min_exact=0.995)


def test_dev_entry_points_follow_the_callers_device(uh):
    """Two GPUs: a thread bound to device 0 passes device-1 buffers and a device-1 stream; the call must run there and
    leave the caller's current device alone. Then a rebind (init(1)) with pageable host calls and a batch call on both."""
    import torch

    lib = uh.lib
    if lib.uhdrb200_device_count() < 2:
        pytest.skip("needs two GPUs")
    w, h = 256, 128
    sdr_b, hdr_b = synth.natural_pair(w, h, 77)
    want = None
    assert lib.uhdrb200_init(0) == 0
    for dev in (0, 1, 0):
        torch.cuda.set_device(dev)
        ds = torch.from_numpy(sdr_b).cuda(dev)
        dh = torch.from_numpy(hdr_b.view(np.int16)).cuda(dev)
        n = 3
        d_map = torch.zeros((n, w * h // 16), dtype=torch.uint8, device=f"cuda:{dev}")
        sdrs, hdrs, maps, mds = (abi.Image * n)(), (abi.Image * n)(), (abi.Image * n)(), (abi.Metadata * n)()
        for i in range(n):
            s_img, h_img = abi.host_yuv420(sdr_b, w, h, 0), abi.host_p010(hdr_b, w, h, 2)
            s_img.data, s_img.chroma_data = ds.data_ptr(), ds.data_ptr() + w * h
            h_img.data, h_img.chroma_data = dh.data_ptr(), dh.data_ptr() + 2 * w * h
            sdrs[i], hdrs[i] = s_img, h_img
            maps[i].data = d_map[i].data_ptr()
        st = uh.generateGainMapBatch(n, sdrs, hdrs, abi.TF_HLG, mds, maps, False, torch.cuda.current_stream(dev).cuda_stream)
        assert st == 0 and torch.cuda.current_device() == dev
        torch.cuda.synchronize(dev)
        got = d_map.cpu().numpy()
        want = got if want is None else want
        assert np.array_equal(got, want)
    for dev in (1, 0):  # rebind the host-pointer context
        assert lib.uhdrb200_init(dev) == 0
        mp, _ = _host_roundtrip(lib, w, h, 5)
        assert mp.any()
    torch.cuda.set_device(0)
