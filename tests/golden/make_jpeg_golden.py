"""Generates tests/golden/jpeg_planes_v1.json: CRC-32 of the raw Y / Cb / Cr planes libjpeg-turbo itself decodes
(through Pillow, dct_method ISLOW: luma = greyscale decode at full size; 4:2:0 chroma = the Cb / Cr channels of a 1/2-scale
decode, where libjpeg runs the full 8x8 IDCT on them and no upsampling) for the reference's baseline JPEG fixtures.
Run where /root/reference and Pillow exist:  python tests/golden/make_jpeg_golden.py"""
import json
import os
import sys
import zlib

import numpy as np
from PIL import Image, features

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from libultrahdr_dev_b200 import refdata  # noqa: E402


def libjpeg_planes(path):
    """-> list of raw component planes as libjpeg-turbo decodes them (1 or 3 arrays)."""
    im = Image.open(path)
    w, h = im.size
    n_comp = len(im.layer)
    lum = Image.open(path)
    lum.draft("L", (w, h))
    planes = [np.asarray(lum if lum.mode == "L" else lum.convert("L")).copy()]
    if n_comp == 3:
        half = Image.open(path)
        half.draft("YCbCr", (w // 2, h // 2))
        a = np.asarray(half)
        assert a.shape[:2] == (h // 2, w // 2)
        planes += [a[..., 1].copy(), a[..., 2].copy()]
    return planes


if __name__ == "__main__":
    out = {"generator": "Pillow %s, libjpeg-turbo %s" % (Image.__version__, features.version("jpg")), "files": {}}
    for name in refdata.JPEGS:
        planes = libjpeg_planes(refdata.jpeg_path(name))
        out["files"][name] = [{"shape": list(p.shape), "crc32": zlib.crc32(p.tobytes())} for p in planes]
    with open(os.path.join(ROOT, "tests", "golden", "jpeg_planes_v1.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))
