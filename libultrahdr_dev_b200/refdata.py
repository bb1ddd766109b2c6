"""The reference's own image fixtures (tests/data of DichenZhang1/libultrahdr_dev), harness side only.

``/root/reference`` does not exist on the GPU box, so ``__graft_entry__.build()`` copies the three
raw fixtures the reference's tests load (``tests/jpegr_test.cpp:1450-1571, 2250-2304``) into the
git-ignored ``tests/_ref_data/`` (it travels with the snapshot like ``oracle/_ref``). Nothing here is
committed and nothing is read from ``/root/reference`` at test or bench time on the GPU box.

  raw_p010_image.p010      1280x720 P010 (HLG, BT.2100 in the reference's tests)
  raw_yuv420_image.yuv420  1280x720 I420 (BT.709), the SDR twin of the above
  minnie-320x240.yu12      320x240 I420, the only photograph in the tree
"""
import os
import shutil

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LOCAL = os.path.join(ROOT, "tests", "_ref_data")
REF = "/root/reference/tests/data"
FILES = {"p010": ("raw_p010_image.p010", 1280, 720), "yuv420": ("raw_yuv420_image.yuv420", 1280, 720),
         "minnie": ("minnie-320x240.yu12", 320, 240)}


# baseline JPEG fixtures of the reference's codec tests (row N4): 4:2:0 colour 1280x720 and 320x240, and a grey one
JPEGS = ["jpeg_image.jpg", "minnie-320x240-yuv.jpg", "minnie-320x240-y.jpg"]


def jpeg_path(name):
    for d in (LOCAL, REF):
        p = os.path.join(d, name)
        if os.path.exists(p):
            return p
    return None


def stage():
    """Copy the fixtures next to the tests (build time, only where the reference tree exists)."""
    if not os.path.isdir(REF):
        return False
    os.makedirs(LOCAL, exist_ok=True)
    for name in [f[0] for f in FILES.values()] + JPEGS:
        dst = os.path.join(LOCAL, name)
        if not os.path.exists(dst):
            shutil.copyfile(os.path.join(REF, name), dst)
    return True


def path_of(key):
    name = FILES[key][0]
    for d in (LOCAL, REF):
        p = os.path.join(d, name)
        if os.path.exists(p):
            return p
    return None


def available():
    return all(path_of(k) for k in FILES)


def load(key):
    """-> (flat array, width, height): u16 for 'p010', u8 otherwise. Raises FileNotFoundError when absent."""
    p = path_of(key)
    if p is None:
        raise FileNotFoundError(FILES[key][0] + ": run __graft_entry__.build() where /root/reference exists")
    _, w, h = FILES[key]
    a = np.fromfile(p, dtype=np.uint16 if key == "p010" else np.uint8)
    assert a.size == w * h * 3 // 2, (key, a.size)
    return a, w, h


def tile_yuv420(buf, w, h, nx, ny, dx=0, dy=0):
    """Tile a packed I420 buffer nx x ny times and roll it by (dx, dy) pixels (both even). -> (buffer, W, H)"""
    y = buf[: w * h].reshape(h, w)
    u = buf[w * h: w * h * 5 // 4].reshape(h // 2, w // 2)
    v = buf[w * h * 5 // 4:].reshape(h // 2, w // 2)
    planes = []
    for p, s in ((y, 1), (u, 2), (v, 2)):
        t = np.tile(p, (ny, nx))
        planes.append(np.roll(t, (dy // s, dx // s), axis=(0, 1)).ravel())
    return np.concatenate(planes), w * nx, h * ny


def tile_p010(buf, w, h, nx, ny, dx=0, dy=0):
    """Same for a packed P010 buffer (u16: Y plane, interleaved UV plane)."""
    y = buf[: w * h].reshape(h, w)
    uv = buf[w * h:].reshape(h // 2, w)  # pairs (u, v)
    ty = np.roll(np.tile(y, (ny, nx)), (dy, dx), axis=(0, 1))
    tuv = np.roll(np.tile(uv, (ny, nx)), (dy // 2, dx), axis=(0, 1))  # dx even keeps the (u, v) pairing
    return np.concatenate([ty.ravel(), tuv.ravel()]), w * nx, h * ny
