"""Where do the CUDA path and the reference disagree on the 720p fixture (PQ map, HLG output)? Breakdown by cell kind."""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from libultrahdr_dev_b200 import abi, api, refdata
from oracle.pyoracle import CpuImpl
import parity
uh = api.JpegR(); ref = CpuImpl.reference() if CpuImpl.have_reference() else CpuImpl.oracle()
p010, w, h = refdata.load("p010"); yuv, _, _ = refdata.load("yuv420")
sdr = abi.host_yuv420(yuv, w, h, 0); hdr = abi.host_p010(p010, w, h, 2)
for tf, fmt in ((2, 3), (2, 2), (1, 3)):
    st, mp_b, md, _ = ref.generate_gainmap(sdr, hdr, tf)
    mp = abi.host_map(mp_b, w // 4, h // 4)
    _, want, _ = ref.apply_gainmap(sdr, mp, md, fmt, 3.4e38)
    out = abi.alloc_output(fmt, w, h); d = abi.Image(); d.data = out.ctypes.data
    assert uh.applyGainMap(sdr, mp, md, fmt, 3.4e38, d) == 0
    g = parity.unpack1010102(out, w * h)[:, :3].reshape(h, w, 3); wv = parity.unpack1010102(want, w * h)[:, :3].reshape(h, w, 3)
    bad = (g != wv)
    m = mp_b.astype(int); mpad = np.pad(m, ((0, 1), (0, 1)), mode="edge")
    sat = (mpad[:-1, :-1] == 255) & (mpad[1:, :-1] == 255) & (mpad[:-1, 1:] == 255) & (mpad[1:, 1:] == 255)
    satpx = np.repeat(np.repeat(sat, 4, 0), 4, 1)
    y = yuv[: w * h].reshape(h, w)
    print(f"tf={tf} fmt={fmt}: exact {1 - bad.mean():.5f}; in sat cells {1 - bad[satpx].mean():.5f} ({satpx.mean():.3f} of px); elsewhere {1 - bad[~satpx].mean():.5f}")
    print("   bad by channel", bad.reshape(-1, 3).mean(0), " want-got sign", np.sign(wv - g)[bad].mean())
    ys, cnt = np.unique(y[bad.any(2)], return_counts=True)
    print("   luma of bad px (top):", sorted(zip(cnt.tolist(), ys.tolist()), reverse=True)[:8])
    codes, cnt = np.unique(wv[bad], return_counts=True)
    print("   want codes of bad (top):", sorted(zip(cnt.tolist(), codes.tolist()), reverse=True)[:8])
