import ctypes, torch, time
torch.cuda.init()
rt = torch.cuda.cudart()
H = 124_000_000
dev = torch.device("cuda:0")
din = torch.empty(H, dtype=torch.uint8, device=dev)
dout = torch.empty(72_000_000, dtype=torch.uint8, device=dev)
def host_alloc(n, flags):
    err, ptr = rt.cudaHostAlloc(n, flags) if hasattr(rt, "cudaHostAlloc") else (1, 0)
    return err, ptr
lib = ctypes.CDLL("libcudart.so.12")
def alloc(n, flags):
    p = ctypes.c_void_p()
    rc = lib.cudaHostAlloc(ctypes.byref(p), ctypes.c_size_t(n), ctypes.c_uint(flags))
    assert rc == 0, rc
    return p.value
def h2d(ptr, n, reps=5):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize(); t = time.perf_counter()
        rc = lib.cudaMemcpyAsync(ctypes.c_void_p(din.data_ptr()), ctypes.c_void_p(ptr), ctypes.c_size_t(n), 1, None)
        torch.cuda.synchronize(); best = min(best, time.perf_counter() - t)
    return best * 1e3
def d2h(ptr, n, reps=5):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize(); t = time.perf_counter()
        rc = lib.cudaMemcpyAsync(ctypes.c_void_p(ptr), ctypes.c_void_p(dout.data_ptr()), ctypes.c_size_t(n), 2, None)
        torch.cuda.synchronize(); best = min(best, time.perf_counter() - t)
    return best * 1e3
for name, flags in (("default", 0), ("portable", 1), ("write-combined", 4)):
    p = alloc(H, flags)
    ctypes.memset(p, 1, H)
    print(name, "h2d ms %.3f  (%.1f GB/s)" % (h2d(p, H), H / h2d(p, H) / 1e6), " d2h 72MB ms %.3f" % d2h(p, 72_000_000))
