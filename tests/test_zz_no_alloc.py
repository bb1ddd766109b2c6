"""uhdrb200_generate_gainmap_batch_dev_noalloc (no counterpart in the reference):
the batched generateGainMap with its per-image descriptor array in a file-scope __device__ array instead of
cudaMalloc'ed memory. Called through the C ABI and compared with the regular batched entry point on the same device
inputs, twice in a row (the shared array keeps the previous call's entries). Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest

from libultrahdr_dev_b200 import abi, synth

pytestmark = pytest.mark.gpu


def _device_batch(t, pairs, w, h):
    n = len(pairs)
    keep, sdrs, hdrs = [], (abi.Image * n)(), (abi.Image * n)()
    for i, (sdr_b, hdr_b) in enumerate(pairs):
        ds = t.from_numpy(sdr_b).cuda()
        dh = t.from_numpy(hdr_b.view(np.int16)).cuda()
        keep += [ds, dh]
        s_img, h_img = abi.host_yuv420(sdr_b, w, h, abi.COLORGAMUT_BT709), abi.host_p010(hdr_b, w, h, abi.COLORGAMUT_BT2100)
        s_img.data, s_img.chroma_data = ds.data_ptr(), ds.data_ptr() + w * h
        h_img.data, h_img.chroma_data = dh.data_ptr(), dh.data_ptr() + 2 * w * h
        sdrs[i], hdrs[i] = s_img, h_img
    return keep, sdrs, hdrs


def _run(t, lib, fn_name, n, sdrs, hdrs, tf, w, h):
    d_map = t.full((n, (w // 4) * (h // 4)), 0xEE, dtype=t.uint8, device="cuda")
    maps, mds = (abi.Image * n)(), (abi.Metadata * n)()
    for i in range(n):
        maps[i].data = d_map[i].data_ptr()
    st = getattr(lib, fn_name)(n, sdrs, hdrs, tf, mds, maps, 0, t.cuda.current_stream().cuda_stream)
    assert st == 0, lib.uhdrb200_last_error()
    t.cuda.synchronize()
    descs = [(m.width, m.height, m.luma_stride, m.pixelFormat, m.colorGamut) for m in maps]
    return d_map.cpu().numpy(), [m.as_dict() for m in mds], descs


def test_noalloc_generate_matches_regular_entry_point_twice(uh):
    import torch as t

    assert t.cuda.is_available()
    t.cuda.set_device(0)
    lib = uh.lib
    w, h = 256, 128
    pairs = [synth.natural_pair(w, h, 900 + i) for i in range(3)] + [synth.uniform_pair(w, h, 910)]
    keep, sdrs, hdrs = _device_batch(t, pairs, w, h)
    # call 1: four images; call 2: two images in another order with the other transfer function (stale entries 2, 3 of the
    # shared array must not matter); call 3: call 1 again
    calls = [(4, sdrs, hdrs, abi.TF_HLG)]
    keep2, sdrs2, hdrs2 = _device_batch(t, [pairs[3], pairs[1]], w, h)
    calls.append((2, sdrs2, hdrs2, abi.TF_PQ))
    calls.append((4, sdrs, hdrs, abi.TF_HLG))
    first = None
    for k, (n, s_arr, h_arr, tf) in enumerate(calls):
        got, md_g, desc_g = _run(t, lib, "uhdrb200_generate_gainmap_batch_dev_noalloc", n, s_arr, h_arr, tf, w, h)
        want, md_w, desc_w = _run(t, lib, "uhdrb200_generate_gainmap_batch_dev", n, s_arr, h_arr, tf, w, h)
        # same kernel, same inputs: byte-identical (the parity tests' +-1 LSB tolerance is not needed)
        assert np.array_equal(got, want), f"call {k}: maps differ"
        assert md_g == md_w and desc_g == desc_w
        assert not (got == 0xEE).all()
        if k == 0:
            first = got
    assert np.array_equal(first, got)  # call 3 reproduces call 1


def test_noalloc_rejects_oversize_and_bad_arguments(uh):
    lib = uh.lib
    n = abi.NOALLOC_MAX_IMAGES + 1
    imgs, mds = (abi.Image * n)(), (abi.Metadata * n)()
    st = lib.uhdrb200_generate_gainmap_batch_dev_noalloc(n, imgs, imgs, abi.TF_HLG, mds, imgs, 0, None)
    assert st == abi.ERROR_UNSUPPORTED_FEATURE
    assert b"UHDRB200_NOALLOC_MAX_IMAGES" in lib.uhdrb200_last_error()
    assert lib.uhdrb200_generate_gainmap_batch_dev_noalloc(0, None, None, abi.TF_HLG, None, None, 0, None) == abi.NO_ERROR
    assert lib.uhdrb200_generate_gainmap_batch_dev_noalloc(1, None, imgs, abi.TF_HLG, mds, imgs, 0, None) == abi.ERROR_BAD_PTR
    assert lib.uhdrb200_generate_gainmap_batch_dev_noalloc(1, imgs, imgs, abi.TF_HLG, mds, imgs, 0, None) == abi.ERROR_BAD_PTR  # null planes
