"""Experiment harness: time applyGainMap (batch, device-resident) for each .so in scratch/variants/
and report exact-match vs the oracle. Usage: python scratch/bench_variants.py [fmt] [batch]"""
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
    from libultrahdr_dev_b200 import abi, api, synth
    from oracle.pyoracle import CpuImpl
    import parity

    lib = api.load_library(path)
    uh = api.JpegR(lib)
    W, H = 3840, 2160
    dev = torch.device("cuda", 0)
    res = {"lib": os.path.basename(path)}
    md = abi.default_metadata(1000.0 / 203.0)
    for kind in ("natural", "uniform"):
        K = 2
        if kind == "natural":
            bufs = [synth.natural_pair(W, H, i)[0] for i in range(K)]
            maps = [synth.natural_map(W // 4, H // 4, 10 + i) for i in range(K)]
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
        res[kind + "_gpix_s"] = round(B * W * H / ms / 1e6, 1)
        # exactness on a 512-row crop of image 0 via the oracle (rows 0..511 only depend on map rows 0..128)
        hh = 512
        crop = np.concatenate([bufs[0][: W * hh], bufs[0][W * H: W * H + (W // 2) * (hh // 2)],
                               bufs[0][W * H + W * H // 4: W * H + W * H // 4 + (W // 2) * (hh // 2)]])
        orc = CpuImpl.oracle()
        sd = abi.host_yuv420(crop, W, hh, 0)
        mp = abi.host_map(np.ascontiguousarray(maps[0][: hh // 4]), W // 4, hh // 4)
        _, want, _ = orc.apply_gainmap(sd, mp, md, fmt, 3.4e38)
        got = d_out[0].cpu().numpy()
        npx = W * (hh - 4)  # last cell row of the crop sees a different lower neighbour
        if fmt in (2, 3):
            g, w_ = parity.unpack1010102(got, npx)[:, :3], parity.unpack1010102(want, npx)[:, :3]
        elif fmt == 1:
            g, w_ = got[: npx * 8].view(np.uint16).astype(np.int32), want[: npx * 8].view(np.uint16).astype(np.int32)
        else:
            g = np.stack([got.view(np.uint16)[c * W * H: c * W * H + npx] for c in range(3)]).astype(np.int32)
            w_ = np.stack([want.view(np.uint16)[c * W * hh: c * W * hh + npx] for c in range(3)]).astype(np.int32)
        d = np.abs(g - w_)
        res[kind + "_exact"] = round(float((d == 0).mean()), 6)
        res[kind + "_maxdiff"] = int(d.max())
    print(json.dumps(res))


if __name__ == "__main__":
    if len(sys.argv) > 1 and sys.argv[1] == "--worker":
        worker(sys.argv[2], int(sys.argv[3]), int(sys.argv[4]))
    else:
        fmt = int(sys.argv[1]) if len(sys.argv) > 1 else 3
        B = int(sys.argv[2]) if len(sys.argv) > 2 else 32
        libs = sorted(glob.glob(os.path.join(ROOT, "scratch", "variants", "*.so")))
        for p in libs:
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--worker", p, str(fmt), str(B)], capture_output=True, text=True, timeout=600)
            print(r.stdout.strip().splitlines()[-1] if r.stdout.strip() else ("FAILED " + p + r.stderr[-800:]))
