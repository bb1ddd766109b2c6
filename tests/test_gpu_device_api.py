"""GPU: the device-pointer (`*_dev`) and batched (`*_batch_dev`) entry points of the C ABI, and the
BASELINE.json configurations at FULL size against the oracle (1080p / 4K / 50 MP)."""
import ctypes as C

import numpy as np
import pytest

from libultrahdr_dev_b200 import abi, synth
from libultrahdr_dev_b200.api import take_map

import parity

pytestmark = pytest.mark.gpu
FLT_MAX = 3.4028234663852886e38


@pytest.fixture(scope="module")
def torch_cuda():
    import torch

    assert torch.cuda.is_available()
    torch.cuda.set_device(0)
    return torch


def dev_yuv420(t, buf, w, h, gamut, ls=None, cs=None):
    d = t.from_numpy(buf).cuda()
    img = abi.host_yuv420(buf, w, h, gamut, ls, cs)
    off = img.chroma_data - img.data
    img.data, img.chroma_data = d.data_ptr(), d.data_ptr() + off
    return d, img


def dev_p010(t, buf, w, h, gamut, ls=None, cs=None):
    d = t.from_numpy(buf.view(np.int16)).cuda()
    img = abi.host_p010(buf, w, h, gamut, ls, cs)
    off = img.chroma_data - img.data
    img.data, img.chroma_data = d.data_ptr(), d.data_ptr() + off
    return d, img


def cpu_impl(orc):
    """Full-size checks use the multi-threaded reference build when it is present (same results
    as the oracle bit for bit, tests/test_oracle_vs_ref.py), else the single-threaded oracle."""
    from oracle.pyoracle import CpuImpl

    return CpuImpl.reference() if CpuImpl.have_reference() else orc


def test_device_entry_points_match_host_entry_points(uh, torch_cuda):
    t = torch_cuda
    w, h = 384, 216  # 216 % 8 == 0
    sdr_b, hdr_b = synth.natural_pair(w, h, 7)
    hs, hh = abi.host_yuv420(sdr_b, w, h, 0), abi.host_p010(hdr_b, w, h, 2)
    md_h, dest_h = abi.Metadata(), abi.Image()
    assert uh.generateGainMap(hs, hh, abi.TF_HLG, md_h, dest_h) == 0
    map_h = take_map(dest_h)

    ds, s_img = dev_yuv420(t, sdr_b, w, h, 0)
    dh, h_img = dev_p010(t, hdr_b, w, h, 2)
    d_map = t.zeros(w * h // 16, dtype=t.uint8, device="cuda")
    md_d, m_img = abi.Metadata(), abi.Image()
    m_img.data = d_map.data_ptr()
    st = t.cuda.current_stream().cuda_stream
    assert uh.generateGainMapDevice(s_img, h_img, abi.TF_HLG, md_d, m_img, False, st) == 0
    t.cuda.synchronize()
    assert md_d.as_dict() == md_h.as_dict() and (m_img.width, m_img.height, m_img.luma_stride) == (w // 4, h // 4, w // 4)
    assert np.array_equal(d_map.cpu().numpy().reshape(h // 4, w // 4), map_h)

    for fmt in (abi.OUTPUT_HDR_LINEAR, abi.OUTPUT_HDR_PQ, abi.OUTPUT_HDR_HLG, abi.OUTPUT_HDR_LINEAR_RGB_10BIT):
        out_h = abi.alloc_output(fmt, w, h)
        d = abi.Image()
        d.data = out_h.ctypes.data
        assert uh.applyGainMap(hs, abi.host_map(map_h, w // 4, h // 4), md_h, fmt, FLT_MAX, d) == 0
        d_out = t.zeros(out_h.size, dtype=t.uint8, device="cuda")
        o_img = abi.Image()
        o_img.data = d_out.data_ptr()
        assert uh.applyGainMapDevice(s_img, m_img, md_d, fmt, FLT_MAX, o_img, st) == 0
        t.cuda.synchronize()
        assert (o_img.width, o_img.height, o_img.colorGamut) == (w, h, 0)
        assert np.array_equal(d_out.cpu().numpy(), out_h), fmt

    tm_h = np.zeros(abi.yuv420_size(w, h), dtype=np.uint8)
    assert uh.toneMap(hh, abi.host_yuv420(tm_h, w, h, -1)) == 0
    d_tm = t.zeros(tm_h.size, dtype=t.uint8, device="cuda")
    tm_img = abi.host_yuv420(tm_h, w, h, -1)
    tm_img.data, tm_img.chroma_data = d_tm.data_ptr(), d_tm.data_ptr() + w * h
    assert uh.toneMapDevice(h_img, tm_img, st) == 0
    t.cuda.synchronize()
    assert tm_img.colorGamut == 2 and np.array_equal(d_tm.cpu().numpy(), tm_h)
    del ds, dh


@pytest.mark.parametrize("n", [1, 5, 70])
def test_batched_launch_equals_per_image_calls(uh, orc, torch_cuda, n):
    """n independent images in ONE launch each for toneMap / generate / apply; image i must equal
    the single-image result, and image 0 the oracle."""
    t = torch_cuda
    w, h = 128, 96
    K = min(n, 5)
    pairs = [synth.natural_pair(w, h, 100 + i) if i % 2 == 0 else synth.uniform_pair(w, h, 100 + i) for i in range(K)]
    keep, sdrs, hdrs, maps, outs, tms = [], (abi.Image * n)(), (abi.Image * n)(), (abi.Image * n)(), (abi.Image * n)(), (abi.Image * n)()
    mds = (abi.Metadata * n)()
    d_map = t.zeros((n, w * h // 16), dtype=t.uint8, device="cuda")
    d_out = t.zeros((n, w * h * 4), dtype=t.uint8, device="cuda")
    d_tm = t.zeros((n, abi.yuv420_size(w, h)), dtype=t.uint8, device="cuda")
    for i in range(n):
        a, simg = dev_yuv420(t, pairs[i % K][0], w, h, 0)
        b, himg = dev_p010(t, pairs[i % K][1], w, h, 2)
        keep += [a, b]
        sdrs[i], hdrs[i] = simg, himg
        maps[i].data = d_map[i].data_ptr()
        outs[i].data = d_out[i].data_ptr()
        tms[i] = abi.host_yuv420(np.zeros(1, np.uint8), w, h, -1)
        tms[i].data, tms[i].chroma_data = d_tm[i].data_ptr(), d_tm[i].data_ptr() + w * h
    st = t.cuda.current_stream().cuda_stream
    n0 = uh.lib.uhdrb200_launch_count()
    assert uh.toneMapBatch(n, hdrs, tms, st) == 0
    assert uh.generateGainMapBatch(n, sdrs, hdrs, abi.TF_HLG, mds, maps, False, st) == 0
    assert uh.applyGainMapBatch(n, sdrs, maps, mds[0], abi.OUTPUT_HDR_HLG, FLT_MAX, outs, st) == 0
    t.cuda.synchronize()
    assert uh.lib.uhdrb200_launch_count() - n0 == 3  # one launch per stage, whatever n is
    got_map, got_out, got_tm = d_map.cpu().numpy(), d_out.cpu().numpy(), d_tm.cpu().numpy()
    for i in range(K):
        hs, hh = abi.host_yuv420(pairs[i][0], w, h, 0), abi.host_p010(pairs[i][1], w, h, 2)
        md, dest = abi.Metadata(), abi.Image()
        assert uh.generateGainMap(hs, hh, abi.TF_HLG, md, dest) == 0
        single = take_map(dest)
        for j in range(i, n, K):
            assert np.array_equal(got_map[j].reshape(h // 4, w // 4), single)
            assert mds[j].as_dict() == md.as_dict()
        o = abi.alloc_output(abi.OUTPUT_HDR_HLG, w, h)
        d = abi.Image()
        d.data = o.ctypes.data
        assert uh.applyGainMap(hs, abi.host_map(single, w // 4, h // 4), md, abi.OUTPUT_HDR_HLG, FLT_MAX, d) == 0
        for j in range(i, n, K):
            assert np.array_equal(got_out[j], o)
        tm = np.zeros(abi.yuv420_size(w, h), dtype=np.uint8)
        assert orc.tonemap(hh, abi.host_yuv420(tm, w, h, -1)) == 0
        for j in range(i, n, K):
            assert np.array_equal(got_tm[j], tm)
    s0, want, _, _ = orc.generate_gainmap(abi.host_yuv420(pairs[0][0], w, h, 0), abi.host_p010(pairs[0][1], w, h, 2), abi.TF_HLG)
    parity.compare_bytes(got_map[0], want, "batch map 0", min_exact=0.99)


def test_batch_rejects_mixed_layouts(uh, torch_cuda):
    t = torch_cuda
    a, s1 = dev_yuv420(t, synth.uniform_pair(64, 32, 1)[0], 64, 32, 0)
    b, s2 = dev_yuv420(t, synth.uniform_pair(32, 32, 1)[0], 32, 32, 0)
    c, h1 = dev_p010(t, synth.uniform_pair(64, 32, 1)[1], 64, 32, 2)
    d, h2 = dev_p010(t, synth.uniform_pair(32, 32, 1)[1], 32, 32, 2)
    sd, hd, mp = (abi.Image * 2)(s1, s2), (abi.Image * 2)(h1, h2), (abi.Image * 2)()
    dm = t.zeros(2 * 128, dtype=t.uint8, device="cuda")
    mp[0].data, mp[1].data = dm.data_ptr(), dm.data_ptr() + 128
    assert uh.generateGainMapBatch(2, sd, hd, abi.TF_HLG, (abi.Metadata * 2)(), mp, False, 0) == abi.ERROR_RESOLUTION_MISMATCH


# ------------------------------------------------------------------------- BASELINE.json configs, full size
def _full_generate(uh, cpu, w, h, tf, sg, hg, seed, kind="natural"):
    if kind == "natural":
        sdr_b, hdr_b = synth.natural_pair(w, h, seed, tf, sg, hg)
    else:
        sdr_b, hdr_b = synth.uniform_pair(w, h, seed)
    sdr, hdr = abi.host_yuv420(sdr_b, w, h, sg), abi.host_p010(hdr_b, w, h, hg)
    st, want, md_o, _ = cpu.generate_gainmap(sdr, hdr, tf)
    md, dest = abi.Metadata(), abi.Image()
    assert st == 0 and uh.generateGainMap(sdr, hdr, tf, md, dest) == 0
    got = take_map(dest)
    assert md.as_dict() == md_o.as_dict()
    return sdr_b, sdr, got, want, md


def test_config1_1080p_generate_hlg(uh, orc):
    """configs[0]: API-1 encode 1920x1080 P010 HLG BT.2100 + YUV420 BT.709 -> gain map."""
    _, _, got, want, md = _full_generate(uh, cpu_impl(orc), 1920, 1080, abi.TF_HLG, 0, 2, 1)
    exact = parity.compare_bytes(got, want, "cfg1 map", min_exact=0.998)
    assert abs(md.maxContentBoost - 1000.0 / 203.0) < 1e-6 and int(want.max()) - int(want.min()) > 150
    print(f"cfg1 1080p generate: exact {exact:.6f}")


def test_config2_4k_apply_hlg_1010102(uh, orc):
    """configs[1]: applyGainMap 3840x2160 -> RGBA1010102 HLG at the default max_display_boost."""
    cpu = cpu_impl(orc)
    w, h = 3840, 2160
    sdr_b, sdr, got_map, want_map, md = _full_generate(uh, cpu, w, h, abi.TF_HLG, 0, 2, 2)
    parity.compare_bytes(got_map, want_map, "cfg2 map", min_exact=0.998)
    mp = abi.host_map(want_map, w // 4, h // 4)  # the reference's map: isolates apply
    st, want, _ = cpu.apply_gainmap(sdr, mp, md, abi.OUTPUT_HDR_HLG, FLT_MAX)
    out = abi.alloc_output(abi.OUTPUT_HDR_HLG, w, h)
    d = abi.Image()
    d.data = out.ctypes.data
    assert st == 0 and uh.applyGainMap(sdr, mp, md, abi.OUTPUT_HDR_HLG, FLT_MAX, d) == 0
    exact = parity.compare_output(abi.OUTPUT_HDR_HLG, out, want, w * h, "cfg2 apply", min_exact=0.999)
    print(f"cfg2 4K apply HLG: exact {exact:.6f}")


def test_config3_4k_pq_p3_roundtrip_f16(uh, orc):
    """configs[2]: PQ 3840x2160 P010 BT.2100 + P3 SDR: generate + apply to RGBA F16 linear."""
    cpu = cpu_impl(orc)
    w, h = 3840, 2160
    sdr_b, sdr, got_map, want_map, md = _full_generate(uh, cpu, w, h, abi.TF_PQ, 1, 2, 3)
    exact_map = parity.compare_bytes(got_map, want_map, "cfg3 map", min_exact=0.998)
    assert abs(md.maxContentBoost - 10000.0 / 203.0) < 1e-5
    mp = abi.host_map(want_map, w // 4, h // 4)
    st, want, _ = cpu.apply_gainmap(sdr, mp, md, abi.OUTPUT_HDR_LINEAR, FLT_MAX)
    out = abi.alloc_output(abi.OUTPUT_HDR_LINEAR, w, h)
    d = abi.Image()
    d.data = out.ctypes.data
    assert st == 0 and uh.applyGainMap(sdr, mp, md, abi.OUTPUT_HDR_LINEAR, FLT_MAX, d) == 0
    exact = parity.compare_output(abi.OUTPUT_HDR_LINEAR, out, want, w * h, "cfg3 apply F16", min_exact=0.999)
    print(f"cfg3 4K PQ/P3: map exact {exact_map:.6f}, F16 exact {exact:.6f}")


def test_config4_50mp_api0_tonemap_generate(uh, orc):
    """configs[3]: API-0 8160x6144 P010: toneMap HDR->SDR + gain map generation (jpegr.cpp:167-206)."""
    cpu = cpu_impl(orc)
    w, h = 8160, 6144
    rng = np.random.Generator(np.random.PCG64(0xB200 + 4))
    # a smooth 10-bit ramp with noise keeps the 50 MP generation cheap on the host (no float pipeline)
    yy = (np.add.outer(np.arange(h, dtype=np.uint32) * 3, np.arange(w, dtype=np.uint32) * 2) // 48 % 877 + 64).astype(np.uint16)
    yy = np.clip(yy + rng.integers(-8, 9, (h, w), dtype=np.int16), 0, 1023).astype(np.uint16)
    u10 = rng.integers(400, 624, (h // 2, w // 2), dtype=np.uint16)
    v10 = rng.integers(400, 624, (h // 2, w // 2), dtype=np.uint16)
    hdr_b = synth.pack_p010(yy, u10, v10)
    hdr = abi.host_p010(hdr_b, w, h, 2)
    ls = (w + 15) // 16 * 16  # ALIGNM(width, 16), jpegr.cpp:185
    outs = []
    for tm in (uh.toneMap, cpu.tonemap):
        buf = np.zeros(abi.yuv420_size(w, h, ls, ls // 2), dtype=np.uint8)
        sdr = abi.host_yuv420(buf, w, h, -1, ls, ls // 2)
        assert tm(hdr, sdr) == 0 and sdr.colorGamut == 2
        outs.append((buf, sdr))
    assert np.array_equal(outs[0][0], outs[1][0])  # bit exact at 50 MP
    assert np.array_equal(outs[0][0][: w * h].reshape(h, w), (yy >> 2).astype(np.uint8))  # == top 8 of 10 bits
    sdr = outs[1][1]
    st, want, md_o, _ = cpu.generate_gainmap(sdr, hdr, abi.TF_HLG)
    md, dest = abi.Metadata(), abi.Image()
    assert st == 0 and uh.generateGainMap(sdr, hdr, abi.TF_HLG, md, dest) == 0
    got = take_map(dest)
    assert got.shape == (h // 4, w // 4)
    exact = parity.compare_bytes(got, want, "cfg4 map", min_exact=0.998)
    print(f"cfg4 50MP API-0: toneMap exact, map exact {exact:.6f}")


def test_config5_image_sharded_batch_checksum(uh, orc, torch_cuda):
    """configs[4] at reduced count: a batch of 4K pairs, generate + apply-HLG in two launches; a
    checksum of per-image checksums must equal that of per-image single calls (images are
    independent: sharding cannot change any byte)."""
    t = torch_cuda
    w, h, n, K = 3840, 2160, 12, 3
    pairs = [synth.natural_pair(w, h, 50 + i) for i in range(K)]
    keep, sdrs, hdrs, maps, outs = [], (abi.Image * n)(), (abi.Image * n)(), (abi.Image * n)(), (abi.Image * n)()
    mds = (abi.Metadata * n)()
    d_map = t.zeros((n, w * h // 16), dtype=t.uint8, device="cuda")
    d_out = t.zeros((n, w * h), dtype=t.int32, device="cuda")
    for i in range(n):
        a, simg = dev_yuv420(t, pairs[i % K][0], w, h, 0)
        b, himg = dev_p010(t, pairs[i % K][1], w, h, 2)
        keep += [a, b]
        sdrs[i], hdrs[i] = simg, himg
        maps[i].data, outs[i].data = d_map[i].data_ptr(), d_out[i].data_ptr()
    st = t.cuda.current_stream().cuda_stream
    assert uh.generateGainMapBatch(n, sdrs, hdrs, abi.TF_HLG, mds, maps, False, st) == 0
    assert uh.applyGainMapBatch(n, sdrs, maps, mds[0], abi.OUTPUT_HDR_HLG, FLT_MAX, outs, st) == 0
    t.cuda.synchronize()
    sums = d_out.to(t.int64).sum(dim=1).cpu().numpy()
    msums = d_map.to(t.int64).sum(dim=1).cpu().numpy()
    for i in range(K):
        assert (sums[i::K] == sums[i]).all() and (msums[i::K] == msums[i]).all()
        hs, hh = abi.host_yuv420(pairs[i][0], w, h, 0), abi.host_p010(pairs[i][1], w, h, 2)
        md, dest = abi.Metadata(), abi.Image()
        assert uh.generateGainMap(hs, hh, abi.TF_HLG, md, dest) == 0
        single = take_map(dest)
        assert int(single.astype(np.int64).sum()) == int(msums[i])
        o = abi.alloc_output(abi.OUTPUT_HDR_HLG, w, h)
        d = abi.Image()
        d.data = o.ctypes.data
        assert uh.applyGainMap(hs, abi.host_map(single, w // 4, h // 4), md, abi.OUTPUT_HDR_HLG, FLT_MAX, d) == 0
        assert int(o.view(np.int32).astype(np.int64).sum()) == int(sums[i])


def test_config5_job_64_images_16_seeds_sharded(uh, torch_cuda):
    """configs[4] as a FIXED job (strong scaling, bench.py --job-images): 64 4K pairs from 16 distinct seeds (each seed
    in 4 cyclic shifts: 64 byte-distinct images), image i -> rank i mod G (shard.py). For G = 1, 2, 4 every rank's
    batch launches run here one after the other; the merged per-image checksums must be identical for every G and
    equal to single-image calls: sharding cannot change a byte."""
    import concurrent.futures as cf
    import multiprocessing as mp

    from libultrahdr_dev_b200 import shard

    t = torch_cuda
    w, h, n, K = 3840, 2160, 64, 16
    with cf.ProcessPoolExecutor(8, mp_context=mp.get_context("spawn")) as ex:
        pairs = list(ex.map(synth.natural_pair, [w] * K, [h] * K, [500 + i for i in range(K)]))
    sdr_len, hdr_len = abi.yuv420_size(w, h), abi.p010_size(w, h)
    d_sdr = t.empty((n, sdr_len), dtype=t.uint8, device="cuda")
    d_hdr = t.empty((n, hdr_len), dtype=t.int16, device="cuda")
    d_map = t.zeros((n, w * h // 16), dtype=t.uint8, device="cuda")
    d_out = t.zeros((n, w * h), dtype=t.int32, device="cuda")
    for i in range(K):
        s = t.from_numpy(pairs[i][0]).cuda()
        hh = t.from_numpy(pairs[i][1].view(np.int16)).cuda()
        for j in range(i, n, K):
            dy, dx = 8 * (j // K), 16 * (j // K)
            for off, pw, ph, sc in ((0, w, h, 1), (w * h, w // 2, h // 2, 2), (w * h * 5 // 4, w // 2, h // 2, 2)):
                d_sdr[j, off:off + pw * ph].view(ph, pw).copy_(t.roll(s[off:off + pw * ph].view(ph, pw), (dy // sc, dx // sc), (0, 1)))
            d_hdr[j, :w * h].view(h, w).copy_(t.roll(hh[:w * h].view(h, w), (dy, dx), (0, 1)))
            d_hdr[j, w * h:].view(h // 2, w).copy_(t.roll(hh[w * h:].view(h // 2, w), (dy // 2, dx), (0, 1)))

    def img_arrays(idx):
        m = len(idx)
        sd, hd, mp_, ou = (abi.Image * m)(), (abi.Image * m)(), (abi.Image * m)(), (abi.Image * m)()
        for k, i in enumerate(idx):
            sb, hb = d_sdr[i].data_ptr(), d_hdr[i].data_ptr()
            sd[k].data, sd[k].chroma_data, sd[k].width, sd[k].height, sd[k].colorGamut = sb, sb + w * h, w, h, 0
            sd[k].luma_stride, sd[k].chroma_stride, sd[k].pixelFormat = w, w // 2, abi.PIX_FMT_YUV420
            hd[k].data, hd[k].chroma_data, hd[k].width, hd[k].height, hd[k].colorGamut = hb, hb + 2 * w * h, w, h, 2
            hd[k].luma_stride, hd[k].chroma_stride, hd[k].pixelFormat = w, w, abi.PIX_FMT_P010
            mp_[k].data, ou[k].data = d_map[i].data_ptr(), d_out[i].data_ptr()
        return sd, hd, mp_, ou

    st = t.cuda.current_stream().cuda_stream
    merged = {}
    for G in (1, 2, 4):
        d_map.zero_()
        d_out.zero_()
        per_rank = []
        for r in range(G):
            idx = shard.shard_round_robin(n, r, G).indices
            sd, hd, mp_, ou = img_arrays(idx)
            mds = (abi.Metadata * len(idx))()
            assert uh.generateGainMapBatch(len(idx), sd, hd, abi.TF_HLG, mds, mp_, False, st) == 0
            assert uh.applyGainMapBatch(len(idx), sd, mp_, mds[0], abi.OUTPUT_HDR_HLG, FLT_MAX, ou, st) == 0
            t.cuda.synchronize()
            osum = d_out[idx].to(t.int64).sum(dim=1).cpu().tolist()  # Python ints: no wrap-around in the combination
            msum = d_map[idx].to(t.int64).sum(dim=1).cpu().tolist()
            per_rank.append([(i, o * 1000003 + m) for i, o, m in zip(idx, osum, msum)])
        merged[G] = shard.merge_checksums(per_rank)
    assert merged[1] == merged[2] == merged[4]
    assert len(set(merged[1])) == n  # 64 byte-distinct images give 64 distinct checksums
    for i in (0, 5, 37, 63):  # single-image device calls agree with the batch launches
        sd, hd, mp_, ou = img_arrays([i])
        d_map[i].zero_()
        d_out[i].zero_()
        md1 = abi.Metadata()
        assert uh.generateGainMapDevice(sd[0], hd[0], abi.TF_HLG, md1, mp_[0], False, st) == 0
        assert uh.applyGainMapDevice(sd[0], mp_[0], md1, abi.OUTPUT_HDR_HLG, FLT_MAX, ou[0], st) == 0
        t.cuda.synchronize()
        c = int(d_out[i].to(t.int64).sum()) * 1000003 + int(d_map[i].to(t.int64).sum())
        assert c == merged[1][i]


# ------------------------------------------------------------------------- N2: fused API-0 (toneMap + generate)
@pytest.mark.parametrize("w,h,ls,cs,dls", [
    (256, 128, None, None, None),   # packed
    (136, 72, 144, 152, 144),       # ALIGNM(w,16) destination stride with zero-filled tails (jpegr.cpp:185), padded source
    (1280, 720, None, None, 1280),
    (130, 66, None, None, 144),     # not multiples of 4 -> two-kernel fallback, same results
    (8, 4, None, None, 16),
])
def test_api0_fused_equals_tonemap_then_generate(uh, orc, w, h, ls, cs, dls):
    _, hdr_b = synth.natural_pair(w, h, 9, abi.TF_HLG, 2, 2, hdr_luma_stride=ls, hdr_chroma_stride=cs)
    hdr = abi.host_p010(hdr_b, w, h, 2, ls, cs)
    dls = dls or w
    dcs = dls // 2
    # reference flow, two calls
    sdr_a = np.full(abi.yuv420_size(w, h, dls, dcs), 0x55, dtype=np.uint8)
    sa = abi.host_yuv420(sdr_a, w, h, -1, dls, dcs)
    assert uh.toneMap(hdr, sa) == 0
    md_a, dest = abi.Metadata(), abi.Image()
    assert uh.generateGainMap(sa, hdr, abi.TF_HLG, md_a, dest) == 0
    map_a = take_map(dest)
    # one call
    sdr_b = np.full(abi.yuv420_size(w, h, dls, dcs), 0x55, dtype=np.uint8)
    sb = abi.host_yuv420(sdr_b, w, h, -1, dls, dcs)
    map_b = np.zeros((h // 4) * (w // 4), dtype=np.uint8)
    md_b, mdesc = abi.Metadata(), abi.Image()
    n0 = uh.lib.uhdrb200_launch_count()
    assert uh.toneMapGenerateGainMapInto(hdr, abi.TF_HLG, sb, md_b, mdesc, map_b) == 0
    launches = uh.lib.uhdrb200_launch_count() - n0
    assert launches == (1 if (w % 4 == 0 and h % 4 == 0 and w >= 4) else 2)
    assert sb.colorGamut == 2 and (mdesc.width, mdesc.height, mdesc.pixelFormat) == (w // 4, h // 4, abi.PIX_FMT_MONOCHROME)
    assert np.array_equal(sdr_a, sdr_b), "fused SDR planes (incl. stride tails) differ from toneMap"
    assert np.array_equal(map_a.ravel(), map_b) and md_a.as_dict() == md_b.as_dict()
    # and against the oracle's two-step flow
    sdr_o = np.full(abi.yuv420_size(w, h, dls, dcs), 0x55, dtype=np.uint8)
    so = abi.host_yuv420(sdr_o, w, h, -1, dls, dcs)
    assert orc.tonemap(hdr, so) == 0 and np.array_equal(sdr_o, sdr_b)
    _, want, _, _ = orc.generate_gainmap(so, hdr, abi.TF_HLG)
    parity.compare_bytes(map_b, want, "api0 fused map", min_exact=0.99)


@pytest.mark.parametrize("gamut", [abi.COLORGAMUT_BT709, abi.COLORGAMUT_P3, abi.COLORGAMUT_BT2100])
@pytest.mark.parametrize("w,h,dls", [(256, 128, 256), (136, 72, 144), (130, 66, 144), (8, 4, 16)])
def test_api0_fused_with_convert_to_601(uh, orc, gamut, w, h, dls):
    """The whole pixel part of API-0 (jpegr.cpp:167-225) in one call: toneMap, generateGainMap on the
    un-converted image, convertYuv(gamut -> BT.601) of the SDR planes; byte-identical to the oracle's chain."""
    _, hdr_b = synth.natural_pair(w, h, 12, abi.TF_HLG, gamut, gamut)
    hdr = abi.host_p010(hdr_b, w, h, gamut)
    dcs = dls // 2
    sdr_o = np.full(abi.yuv420_size(w, h, dls, dcs), 0x55, dtype=np.uint8)
    so = abi.host_yuv420(sdr_o, w, h, -1, dls, dcs)
    assert orc.tonemap(hdr, so) == 0
    _, want_map, _, _ = orc.generate_gainmap(so, hdr, abi.TF_HLG)
    if gamut != abi.COLORGAMUT_P3:
        assert orc.convert_yuv(so, gamut, abi.COLORGAMUT_P3) == 0
    sdr_b = np.full(abi.yuv420_size(w, h, dls, dcs), 0x55, dtype=np.uint8)
    sb = abi.host_yuv420(sdr_b, w, h, -1, dls, dcs)
    map_b = np.zeros((h // 4) * (w // 4), dtype=np.uint8)
    md, mdesc = abi.Metadata(), abi.Image()
    n0 = uh.lib.uhdrb200_launch_count()
    assert uh.toneMapGenerateGainMapInto(hdr, abi.TF_HLG, sb, md, mdesc, map_b, sdr_to_601=True) == 0
    launches = uh.lib.uhdrb200_launch_count() - n0
    fused = w % 4 == 0 and h % 4 == 0
    assert launches == (1 if fused else (2 if gamut == abi.COLORGAMUT_P3 else 3))
    assert sb.colorGamut == gamut
    assert np.array_equal(sdr_b, sdr_o), "SDR planes after the fused convertYuv differ from the oracle chain"
    parity.compare_bytes(map_b, want_map, "api0 fused+convert map", min_exact=0.99)


def test_api0_fused_error_codes(uh):
    w, h = 32, 16
    _, hdr_b = synth.uniform_pair(w, h, 1)
    hdr = abi.host_p010(hdr_b, w, h, 2)
    sdr_b = np.zeros(abi.yuv420_size(w, h), dtype=np.uint8)
    mp = np.zeros(32, dtype=np.uint8)
    md, d = abi.Metadata(), abi.Image()
    assert uh.toneMapGenerateGainMapInto(None, abi.TF_HLG, abi.host_yuv420(sdr_b, w, h, -1), md, d, mp) == abi.ERROR_BAD_PTR
    assert uh.toneMapGenerateGainMapInto(hdr, abi.TF_HLG, abi.host_yuv420(sdr_b, w, h - 2, -1), md, d, mp) == abi.ERROR_RESOLUTION_MISMATCH
    assert uh.toneMapGenerateGainMapInto(abi.host_p010(hdr_b, w, h, -1), abi.TF_HLG, abi.host_yuv420(sdr_b, w, h, -1), md, d, mp) == abi.ERROR_INVALID_COLORGAMUT
    assert uh.toneMapGenerateGainMapInto(hdr, abi.TF_SRGB, abi.host_yuv420(sdr_b, w, h, -1), md, d, mp) == abi.ERROR_INVALID_TRANS_FUNC


def test_api0_fused_batch_device(uh, torch_cuda):
    t = torch_cuda
    w, h, n = 256, 128, 6
    hdr_bufs = [synth.natural_pair(w, h, 20 + i, abi.TF_PQ, 2, 2)[1] for i in range(n)]
    keep, hdrs, sdrs, maps = [], (abi.Image * n)(), (abi.Image * n)(), (abi.Image * n)()
    mds = (abi.Metadata * n)()
    d_sdr = t.zeros((n, abi.yuv420_size(w, h)), dtype=t.uint8, device="cuda")
    d_map = t.zeros((n, w * h // 16), dtype=t.uint8, device="cuda")
    for i in range(n):
        a, himg = dev_p010(t, hdr_bufs[i], w, h, 2)
        keep.append(a)
        hdrs[i] = himg
        sdrs[i] = abi.host_yuv420(np.zeros(1, np.uint8), w, h, -1)
        sdrs[i].data, sdrs[i].chroma_data = d_sdr[i].data_ptr(), d_sdr[i].data_ptr() + w * h
        maps[i].data = d_map[i].data_ptr()
    n0 = uh.lib.uhdrb200_launch_count()
    assert uh.toneMapGenerateGainMapBatch(n, hdrs, abi.TF_PQ, sdrs, mds, maps, t.cuda.current_stream().cuda_stream) == 0
    t.cuda.synchronize()
    assert uh.lib.uhdrb200_launch_count() - n0 == 1
    for i in range(n):
        hh = abi.host_p010(hdr_bufs[i], w, h, 2)
        sb = np.zeros(abi.yuv420_size(w, h), dtype=np.uint8)
        mb = np.zeros(w * h // 16, dtype=np.uint8)
        md, md_desc = abi.Metadata(), abi.Image()
        assert uh.toneMapGenerateGainMapInto(hh, abi.TF_PQ, abi.host_yuv420(sb, w, h, -1), md, md_desc, mb) == 0
        assert np.array_equal(d_sdr[i].cpu().numpy(), sb) and np.array_equal(d_map[i].cpu().numpy(), mb)
        assert sdrs[i].colorGamut == 2 and mds[i].as_dict() == md.as_dict()


# ---------------------------------------------------------------- pageable vs page-locked host planes
def test_pinned_and_pageable_host_planes_agree(uh):
    """The host entry points hand page-locked planes to the copy engine directly and stage pageable ones through
    per-thread 4 MB chunks (capi.cu copy_h2d / copy_d2h): same bytes out, multi-chunk sizes (1080p and 4K planes),
    including a size that is not a multiple of the chunk."""
    lib = uh.lib
    for w, h in ((1920, 1080), (3840, 2160)):
        sdr_b, hdr_b = synth.uniform_pair(w, h, 77)
        outs = []
        for kind in ("pageable", "pinned"):
            if kind == "pinned":
                ps, ph = lib.uhdrb200_host_alloc(sdr_b.nbytes), lib.uhdrb200_host_alloc(hdr_b.nbytes)
                po, pm = lib.uhdrb200_host_alloc(w * h * 4), lib.uhdrb200_host_alloc(w * h // 16)
                assert ps and ph and po and pm
                C.memmove(ps, sdr_b.ctypes.data, sdr_b.nbytes)
                C.memmove(ph, hdr_b.ctypes.data, hdr_b.nbytes)
            else:
                keep = [sdr_b.copy(), hdr_b.copy(), np.zeros(w * h * 4, np.uint8), np.zeros(w * h // 16, np.uint8)]
                ps, ph, po, pm = (k.ctypes.data for k in keep)
            s_img = abi.host_yuv420(sdr_b, w, h, abi.COLORGAMUT_BT709)
            h_img = abi.host_p010(hdr_b, w, h, abi.COLORGAMUT_BT2100)
            s_img.data, s_img.chroma_data = ps, ps + w * h
            h_img.data, h_img.chroma_data = ph, ph + 2 * w * h
            md, m_img, o_img = abi.Metadata(), abi.Image(), abi.Image()
            st = lib.uhdrb200_generate_gainmap_host_into(C.byref(s_img), C.byref(h_img), abi.TF_HLG, C.byref(md), C.byref(m_img), 0, pm, w * h // 16)
            assert st == 0, lib.uhdrb200_last_error()
            o_img.data = po
            st = lib.uhdrb200_apply_gainmap_host(C.byref(s_img), C.byref(m_img), C.byref(md), abi.OUTPUT_HDR_HLG, FLT_MAX, C.byref(o_img))
            assert st == 0, lib.uhdrb200_last_error()
            outs.append((np.ctypeslib.as_array(C.cast(pm, C.POINTER(C.c_uint8)), (w * h // 16,)).copy(),
                         np.ctypeslib.as_array(C.cast(po, C.POINTER(C.c_uint8)), (w * h * 4,)).copy()))
            if kind == "pinned":
                for q in (ps, ph, po, pm):
                    lib.uhdrb200_host_free(q)
        assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])
        assert outs[0][1].any()


# ---------------------------------------------------------------- concurrent callers (thread-local contexts)
def test_concurrent_host_callers_get_their_own_results(uh):
    """The reference API is re-entrant (no shared mutable state); here every host thread owns a stream, device
    staging buffers and page-locked chunks. Eight threads hammer generate + apply on different images and sizes
    (pageable planes, multi-chunk at the larger sizes); every result must equal the one computed alone."""
    import concurrent.futures as cf

    lib = uh.lib
    sizes = [(256, 128), (1920, 1080), (640, 360), (132, 68), (2048, 1024), (516, 36), (1280, 720), (64, 64)]
    jobs = []
    for i, (w, h) in enumerate(sizes):
        sdr_b, hdr_b = synth.uniform_pair(w, h, 300 + i)
        jobs.append((w, h, sdr_b, hdr_b))

    def run(job):
        w, h, sdr_b, hdr_b = job
        s_img = abi.host_yuv420(sdr_b, w, h, abi.COLORGAMUT_BT709)
        h_img = abi.host_p010(hdr_b, w, h, abi.COLORGAMUT_BT2100)
        mp = np.zeros(w * h // 16, np.uint8)
        out = np.zeros(w * h * 4, np.uint8)
        md, m_img, o_img = abi.Metadata(), abi.Image(), abi.Image()
        st = lib.uhdrb200_generate_gainmap_host_into(C.byref(s_img), C.byref(h_img), abi.TF_HLG, C.byref(md), C.byref(m_img), 0,
                                                     mp.ctypes.data, mp.size)
        assert st == 0, lib.uhdrb200_last_error()
        o_img.data = out.ctypes.data
        st = lib.uhdrb200_apply_gainmap_host(C.byref(s_img), C.byref(m_img), C.byref(md), abi.OUTPUT_HDR_HLG, FLT_MAX, C.byref(o_img))
        assert st == 0, lib.uhdrb200_last_error()
        return mp, out

    alone = [run(j) for j in jobs]
    with cf.ThreadPoolExecutor(8) as pool:
        for _ in range(3):
            order = list(range(len(jobs))) * 3
            got = list(pool.map(lambda k: run(jobs[k]), order))
            for k, (mp, out) in zip(order, got):
                assert np.array_equal(mp, alone[k][0]) and np.array_equal(out, alone[k][1]), f"job {k} differs under concurrency"
