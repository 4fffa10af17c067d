"""A/B timing of two builds of libdgemm_b200.so on the same box (binds only the symbols every build has):
    python tools/ab_lib.py path/to/lib.so M N K [reps]"""
import ctypes as C, json, sys
lib = C.CDLL(sys.argv[1])
M, N, K = (int(x) for x in sys.argv[2:5])
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 5
vp, ll = C.c_void_p, C.c_longlong
lib.dgemm_b200_malloc.restype = vp; lib.dgemm_b200_malloc.argtypes = [C.c_size_t]
lib.dgemm_b200_fill_signed.argtypes = [vp, C.c_int, C.c_int, ll, C.c_ulonglong, C.c_ulonglong, C.c_ulonglong, vp]
lib.dgemm_b200_sync.argtypes = [vp]
lib.dgemm_b200_time_device.restype = C.c_double
lib.dgemm_b200_time_device.argtypes = [C.c_int, C.c_int, C.c_int, C.c_double, vp, ll, vp, ll, C.c_double, vp, ll, C.c_int,
                                       C.c_int, C.c_int, C.c_int, C.c_size_t]
dA, dB, dC = lib.dgemm_b200_malloc(8 * M * K), lib.dgemm_b200_malloc(8 * K * N), lib.dgemm_b200_malloc(8 * M * N)
lib.dgemm_b200_fill_signed(dA, M, K, M, 1, 0, 0, None); lib.dgemm_b200_fill_signed(dB, K, N, K, 2, 0, 0, None)
lib.dgemm_b200_fill_signed(dC, M, N, M, 3, 0, 0, None); lib.dgemm_b200_sync(None)
ms = lib.dgemm_b200_time_device(M, N, K, 1.0, dA, M, dB, K, 1.0, dC, M, 0, 0, 2, reps, 256 << 20)
print(json.dumps({"lib": sys.argv[1].split("/")[-1], "M": M, "N": N, "K": K, "ms": round(ms, 4), "tflops": round(2.0 * M * N * K / ms / 1e9, 3)}))
