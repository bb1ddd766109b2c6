"""Generates tests/golden/jpeg_encode_v1.json: size and CRC-32 of the JPEG files libjpeg-turbo (through Pillow) writes for
deterministic 4:2:0 / grey planes under the reference's encoder settings (lib/src/jpegencoderhelper.cpp:113-130). The GPU
encoder must reproduce them byte for byte (tests/test_gpu_jpeg_encode.py), also on a box without Pillow.
Run where Pillow exists:  python tests/golden/make_jpeg_encode_golden.py"""
import json
import os
import sys
import zlib

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

CASES = [(64, 48, 90, False), (640, 368, 75, False), (640, 368, 100, False), (1280, 720, 95, False), (320, 240, 85, True),
         (960, 544, 50, True), (16, 16, 1, False), (1920, 1088, 92, False)]


def planes_of(w, h, quality, grey):
    from test_gpu_jpeg_encode import planes_420

    p = planes_420(w, h, w * 3 + quality)
    return p[:1] if grey else p


if __name__ == "__main__":
    from PIL import Image, features

    from test_gpu_jpeg_encode import pillow_encode

    out = {"generator": "Pillow %s, libjpeg-turbo %s" % (Image.__version__, features.version("jpg")), "cases": []}
    for w, h, q, grey in CASES:
        data = pillow_encode(planes_of(w, h, q, grey), q)
        out["cases"].append({"width": w, "height": h, "quality": q, "grey": grey, "bytes": len(data), "crc32": zlib.crc32(data)})
    with open(os.path.join(ROOT, "tests", "golden", "jpeg_encode_v1.json"), "w") as f:
        json.dump(out, f, indent=1)
    print(json.dumps(out, indent=1))
