"""Cost of pinning caller memory in place (cudaHostRegister + cudaHostUnregister around one call) against the staged copy
it would replace: decides whether the *_host entry points should register pageable planes instead of copying them."""
import ctypes as C, json, time
import numpy as np, torch

rt = torch.cuda.cudart()
torch.cuda.init()
res = {}
for mb in (12, 33, 66):
    n = mb << 20
    a = np.empty(n, dtype=np.uint8); a[::4096] = 1  # touched pages, as a decoded image would be
    ts = []
    for _ in range(5):
        t0 = time.perf_counter()
        r = rt.cudaHostRegister(a.ctypes.data, n, 0)
        t1 = time.perf_counter()
        assert int(r) == 0, r
        rt.cudaHostUnregister(a.ctypes.data)
        t2 = time.perf_counter()
        ts.append((t1 - t0, t2 - t1))
    b = np.empty(n, dtype=np.uint8); b[::4096] = 1
    t0 = time.perf_counter()
    for _ in range(5): np.copyto(b, a)
    cp = (time.perf_counter() - t0) / 5
    res[mb] = {"register_ms": round(min(t[0] for t in ts) * 1e3, 3), "unregister_ms": round(min(t[1] for t in ts) * 1e3, 3), "memcpy_1thread_ms": round(cp * 1e3, 3)}
print(json.dumps(res))
