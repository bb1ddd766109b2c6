import sys
import numpy as np
sys.path.insert(0, '.'); sys.path.insert(0, 'tests')
from libultrahdr_dev_b200 import abi, synth
from libultrahdr_dev_b200.api import JpegR
from oracle.pyoracle import CpuImpl
import parity
uh, orc = JpegR(), CpuImpl.oracle()
w, h = 512, 288
for kind in ('natural', 'uniform'):
    if kind == 'natural':
        sdr_b, _ = synth.natural_pair(w, h, 21); mp_b = synth.natural_map(w // 4, h // 4, 22)
    else:
        sdr_b, _ = synth.uniform_pair(w, h, 23)
        mp_b = np.random.Generator(np.random.PCG64(24)).integers(0, 256, (h // 4, w // 4), dtype=np.uint8)
    sdr, mp = abi.host_yuv420(sdr_b, w, h, 0), abi.host_map(mp_b, w // 4, h // 4)
    md = abi.default_metadata(1000 / 203.0)
    for fmt in (1, 2, 3, 4):
        _, want, _ = orc.apply_gainmap(sdr, mp, md, fmt, 3.4e38)
        for mode in (1, 0):
            uh.lib.uhdrb200_set_apply_mode(mode)
            out = abi.alloc_output(fmt, w, h); d = abi.Image(); d.data = out.ctypes.data
            assert uh.applyGainMap(sdr, mp, md, fmt, 3.4e38, d) == 0
            if fmt == 1:
                g = out.view(np.uint16).reshape(h, w, 4).astype(np.int32); wv = want.view(np.uint16).reshape(h, w, 4).astype(np.int32)
                df = g - wv
                print(kind, 'fmt', fmt, 'mode', mode, 'exact', (df == 0).mean(), 'max|d|', np.abs(df).max(),
                      'per-ch', [(df[..., c] == 0).mean() for c in range(3)],
                      'interior exact', (df[:-4, :-4] == 0).mean())
                bad = np.argwhere(df[..., :3] != 0)
                if mode == 0 and len(bad):
                    vals = want.view(np.float16).reshape(h, w, 4)[bad[:, 0], bad[:, 1], bad[:, 2]].astype(np.float32)
                    print('   bad value quantiles', np.quantile(vals, [0, .1, .5, .9, 1]), 'n', len(bad))
                    allv = want.view(np.float16).reshape(h, w, 4)[..., :3].astype(np.float32)
                    for lo, hi in ((0, 6.1e-5), (6.1e-5, 1e-3), (1e-3, 1e-2), (1e-2, 1e-1), (1e-1, 2)):
                        m = (allv > lo) & (allv <= hi)
                        print('   range', lo, hi, 'n', m.sum(), 'exact', (df[..., :3][m] == 0).mean() if m.any() else None)
            else:
                n = w * h
                if fmt == 4:
                    g, wv = parity.unpack_planar10(out, n), parity.unpack_planar10(want, n)
                else:
                    g, wv = parity.unpack1010102(out, n)[:, :3], parity.unpack1010102(want, n)[:, :3]
                df = g - wv
                print(kind, 'fmt', fmt, 'mode', mode, 'exact', (df == 0).mean(), 'max|d|', np.abs(df).max())
