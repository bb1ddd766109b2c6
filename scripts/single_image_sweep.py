"""Per-call latency of applyGainMap / generateGainMap on ONE 4K image (device pointers) for a few values of an environment
knob. Usage: python scripts/single_image_sweep.py ENVVAR v1 v2 ..."""
import json, os, subprocess, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def worker():
    import numpy as np, torch
    from libultrahdr_dev_b200 import abi, api, synth
    uh = api.JpegR(); W, H, B = 3840, 2160, 24
    dev = torch.device("cuda", 0)
    sdr_b, hdr_b = synth.natural_pair(W, H, 1)
    d_sdr = torch.from_numpy(sdr_b).to(dev).repeat(B, 1); d_hdr = torch.from_numpy(hdr_b.view(np.int16)).to(dev).repeat(B, 1)
    d_map = torch.zeros((B, W * H // 16), dtype=torch.uint8, device=dev); d_out = torch.zeros((B, W * H), dtype=torch.int32, device=dev)
    Img = abi.Image * B; sdrs, hdrs, maps, outs = Img(), Img(), Img(), Img()
    for i in range(B):
        b = d_sdr[i].data_ptr(); sdrs[i].data, sdrs[i].chroma_data = b, b + W * H
        sdrs[i].width, sdrs[i].height, sdrs[i].colorGamut, sdrs[i].luma_stride, sdrs[i].chroma_stride, sdrs[i].pixelFormat = W, H, 0, W, W // 2, abi.PIX_FMT_YUV420
        hb = d_hdr[i].data_ptr(); hdrs[i].data, hdrs[i].chroma_data = hb, hb + 2 * W * H
        hdrs[i].width, hdrs[i].height, hdrs[i].colorGamut, hdrs[i].luma_stride, hdrs[i].chroma_stride, hdrs[i].pixelFormat = W, H, 2, W, W, abi.PIX_FMT_P010
        maps[i].data, maps[i].width, maps[i].height, maps[i].luma_stride = d_map[i].data_ptr(), W // 4, H // 4, W // 4
        outs[i].data = d_out[i].data_ptr()
    st = torch.cuda.current_stream().cuda_stream; md = abi.Metadata()
    for i in range(B): assert uh.generateGainMapDevice(sdrs[i], hdrs[i], abi.TF_HLG, md, maps[i], False, st) == 0
    res = {}
    for name, fn in (("apply_us", lambda i: uh.applyGainMapDevice(sdrs[i], maps[i], md, abi.OUTPUT_HDR_HLG, 3.4e38, outs[i], st)),
                     ("generate_us", lambda i: uh.generateGainMapDevice(sdrs[i], hdrs[i], abi.TF_HLG, md, maps[i], False, st))):
        for i in range(8): fn(i)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        import time
        torch.cuda.synchronize(); e0.record(); t0 = time.perf_counter()
        for i in range(96): fn(i % B)
        t1 = time.perf_counter()  # host side only: what the calling thread spends per call (python + ctypes + C ABI + launch)
        e1.record(); torch.cuda.synchronize()
        res[name] = round(e0.elapsed_time(e1) / 96 * 1e3, 2)
        res[name.replace("_us", "_host_us")] = round((t1 - t0) / 96 * 1e6, 2)
    print(json.dumps(res))


if __name__ == "__main__":
    if sys.argv[1] == "--worker":
        worker()
    else:
        var = sys.argv[1]
        for v in sys.argv[2:]:
            env = dict(os.environ); env[var] = v
            r = subprocess.run([sys.executable, os.path.abspath(__file__), "--worker"], capture_output=True, text=True, env=env)
            print(var, v, r.stdout.strip().splitlines()[-1] if r.stdout.strip() else r.stderr[-400:], flush=True)
