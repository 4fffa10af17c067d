"""How fast is cudaHostRegister / cudaHostUnregister on malloc'ed (touched) memory, against the staged copy path?
    python tools/hostregister_probe.py"""
import ctypes as C, json, time, glob
import numpy as np
rt = C.CDLL(sorted(glob.glob("/usr/local/cuda/lib64/libcudart.so*"))[0])
rt.cudaHostRegister.argtypes = [C.c_void_p, C.c_size_t, C.c_uint]
rt.cudaHostUnregister.argtypes = [C.c_void_p]
rt.cudaMalloc.argtypes = [C.POINTER(C.c_void_p), C.c_size_t]
rt.cudaMemcpyAsync.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_int, C.c_void_p]
rt.cudaFree(0)
for mb in (8, 32, 72, 256, 1024):
    n = mb << 20
    a = np.ones(n // 8)          # malloc'ed and touched
    d = C.c_void_p()
    assert rt.cudaMalloc(C.byref(d), n) == 0
    res = {"MB": mb}
    for flags, name in ((0, "default"), (8, "readonly")):   # cudaHostRegisterReadOnly = 8
        ts = []
        for rep in range(3):
            t0 = time.perf_counter(); st = rt.cudaHostRegister(a.ctypes.data, n, flags); t1 = time.perf_counter()
            if st != 0:
                res[name] = f"error {st}"; rt.cudaGetLastError(); break
            rt.cudaMemcpyAsync(d, a.ctypes.data, n, 1, None); rt.cudaDeviceSynchronize(); t2 = time.perf_counter()
            rt.cudaHostUnregister(a.ctypes.data); t3 = time.perf_counter()
            ts.append((t1 - t0, t2 - t1, t3 - t2))
        if ts:
            r, c, u = min(ts)[0], min(t[1] for t in ts), min(t[2] for t in ts)
            res[name] = {"register_ms": round(r * 1e3, 3), "h2d_ms": round(c * 1e3, 3), "unregister_ms": round(u * 1e3, 3),
                         "register_GBps": round(n / r / 1e9, 1), "h2d_GBps": round(n / c / 1e9, 1)}
    t0 = time.perf_counter(); rt.cudaMemcpyAsync(d, a.ctypes.data, n, 1, None); rt.cudaDeviceSynchronize()
    res["pageable_h2d_ms"] = round((time.perf_counter() - t0) * 1e3, 3)
    print(json.dumps(res), flush=True)
    rt.cudaFree(d)
