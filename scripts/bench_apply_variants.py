"""A/B harness: time applyGainMap (batched, device-resident, 4K) for every variant library in build/variants/*.so on
natural / photo (minnie) / dark / uniform content, and check each against the oracle on a crop.
Build variants with scripts/mkvariants.sh name:"-DFLAG=..." ...   Usage: python scripts/bench_apply_variants.py [fmt] [batch]"""
import ctypes as C
import glob
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def worker(path, fmt, B):
    import numpy as np
    import torch
    from libultrahdr_dev_b200 import abi, api, refdata, synth
    from oracle.pyoracle import CpuImpl
    import parity

    lib = api.load_library(path)
    uh = api.JpegR(lib)
    W, H = 3840, 2160
    dev = torch.device("cuda", 0)
    res = {"lib": os.path.basename(path)}
    md = abi.default_metadata(1000.0 / 203.0)
    orc = CpuImpl.oracle()
    kinds = ["natural", "minnie", "dark", "uniform"]
    for kind in kinds:
        K = 4
        if kind == "natural":
            bufs = [synth.natural_pair(W, H, i)[0] for i in range(2)] * 2
            maps = [synth.natural_map(W // 4, H // 4, 10 + i) for i in range(K)]
        elif kind == "minnie":
            fm, w0, h0 = refdata.load("minnie")
            bufs = [refdata.tile_yuv420(fm, w0, h0, 12, 9, 8 * i, 4 * i)[0] for i in range(K)]
            maps = [synth.natural_map(W // 4, H // 4, 20 + i) for i in range(K)]
        elif kind == "dark":
            base = synth.natural_pair(W, H, 0)[0]
            b = base.copy()
            b[: W * H] //= 8
            b[W * H:] = (128 + (b[W * H:].astype(np.int16) - 128) // 8).astype(np.uint8)
            bufs = [b] * K
            maps = [synth.natural_map(W // 4, H // 4, 30 + i) for i in range(K)]
        else:
            bufs = [synth.uniform_pair(W, H, i)[0] for i in range(K)]
            maps = [np.random.Generator(np.random.PCG64(i)).integers(0, 256, (H // 4, W // 4), dtype=np.uint8) for i in range(K)]
        d_sdr = torch.empty((B, bufs[0].size), dtype=torch.uint8, device=dev)
        d_map = torch.empty((B, W * H // 16), dtype=torch.uint8, device=dev)
        bpp = abi.out_bytes_per_pixel(fmt)
        d_out = torch.zeros((B, W * H * bpp), dtype=torch.uint8, device=dev)
        for i in range(K):
            d_sdr[i::K] = torch.from_numpy(bufs[i]).to(dev)
            d_map[i::K] = torch.from_numpy(maps[i].ravel()).to(dev)
        Img = abi.Image * B
        sdrs, mps, outs = Img(), Img(), Img()
        for i in range(B):
            b = d_sdr[i].data_ptr()
            sdrs[i].data, sdrs[i].chroma_data = b, b + W * H
            sdrs[i].width, sdrs[i].height, sdrs[i].colorGamut = W, H, 0
            sdrs[i].luma_stride, sdrs[i].chroma_stride, sdrs[i].pixelFormat = W, W // 2, abi.PIX_FMT_YUV420
            mps[i].data, mps[i].width, mps[i].height, mps[i].luma_stride = d_map[i].data_ptr(), W // 4, H // 4, W // 4
            outs[i].data = d_out[i].data_ptr()
        st = torch.cuda.current_stream().cuda_stream
        for _ in range(3):
            assert uh.applyGainMapBatch(B, sdrs, mps, md, fmt, 3.4e38, outs, st) == 0
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n = 10
        e0.record()
        for _ in range(n):
            uh.applyGainMapBatch(B, sdrs, mps, md, fmt, 3.4e38, outs, st)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / n
        res[kind] = round(B * W * H / ms / 1e6, 1)  # Gpix/s
        # single-image call rate
        e0.record()
        for i in range(64):
            uh.applyGainMapDevice(sdrs[i % B], mps[i % B], md, fmt, 3.4e38, outs[i % B], st)
        e1.record()
        torch.cuda.synchronize()
        res[kind + "_1img_us"] = round(e0.elapsed_time(e1) / 64 * 1e3, 2)
        # exactness on a 256-row crop of image 0 via the oracle
        assert uh.applyGainMapBatch(B, sdrs, mps, md, fmt, 3.4e38, outs, st) == 0
        torch.cuda.synchronize()
        hh = 256
        crop = np.concatenate([bufs[0][: W * hh], bufs[0][W * H: W * H + (W // 2) * (hh // 2)],
                               bufs[0][W * H + W * H // 4: W * H + W * H // 4 + (W // 2) * (hh // 2)]])
        sd = abi.host_yuv420(crop, W, hh, 0)
        mp = abi.host_map(np.ascontiguousarray(maps[0][: hh // 4]), W // 4, hh // 4)
        _, want, _ = orc.apply_gainmap(sd, mp, md, fmt, 3.4e38)
        got = d_out[0].cpu().numpy()
        npx = W * (hh - 4)  # the last cell row of the crop sees a different lower neighbour
        if fmt in (2, 3):
            g, w_ = parity.unpack1010102(got, npx)[:, :3], parity.unpack1010102(want, npx)[:, :3]
        elif fmt == 1:
            g, w_ = got[: npx * 8].view(np.uint16).astype(np.int32), want[: npx * 8].view(np.uint16).astype(np.int32)
        else:
            g = np.stack([got.view(np.uint16)[c * W * H: c * W * H + npx] for c in range(3)]).astype(np.int32)
            w_ = np.stack([want.view(np.uint16)[c * W * hh: c * W * hh + npx] for c in range(3)]).astype(np.int32)
        d = np.abs(g - w_)
        res[kind + "_exact"] = round(float((d == 0).mean()), 5)
        res[kind + "_maxdiff"] = int(d.max())
    print(json.dumps(res))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--worker":
        worker(sys.argv[2], int(sys.argv[3]), int(sys.argv[4]))
    else:
        fmt = int(sys.argv[1]) if len(sys.argv) > 1 else 3
        B = int(sys.argv[2]) if len(sys.argv) > 2 else 48
        for p in sorted(glob.glob(os.path.join(ROOT, "build", "variants", "*.so"))):
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--worker", p, str(fmt), str(B)], capture_output=True, text=True, timeout=900)
            print(r.stdout.strip().splitlines()[-1] if r.stdout.strip() else ("FAILED " + p + r.stderr[-800:]), flush=True)
