import torch, time
torch.cuda.init()
dev = torch.device("cuda:0")
def timeit(fn, n=5):
    fn(); torch.cuda.synchronize()
    best = 1e9
    for _ in range(n):
        t = time.perf_counter(); fn(); torch.cuda.synchronize(); best = min(best, time.perf_counter() - t)
    return best * 1e3
H = 124_000_000; D = 68_000_000
hin = torch.empty(H, dtype=torch.uint8).pin_memory(); din = torch.empty(H, dtype=torch.uint8, device=dev)
hout = torch.empty(D, dtype=torch.uint8).pin_memory(); dout = torch.empty(D, dtype=torch.uint8, device=dev)
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
print("h2d 124MB one copy ms", timeit(lambda: din.copy_(hin, non_blocking=True)))
print("d2h 68MB one copy ms", timeit(lambda: hout.copy_(dout, non_blocking=True)))
def both():
    with torch.cuda.stream(s1): din.copy_(hin, non_blocking=True)
    with torch.cuda.stream(s2): hout.copy_(dout, non_blocking=True)
print("both concurrently ms", timeit(both))
def chunked(nc_in, nc_out):
    def f():
        with torch.cuda.stream(s1):
            for c in range(nc_in):
                a, b = H * c // nc_in, H * (c + 1) // nc_in
                din[a:b].copy_(hin[a:b], non_blocking=True)
        with torch.cuda.stream(s2):
            for c in range(nc_out):
                a, b = D * c // nc_out, D * (c + 1) // nc_out
                hout[a:b].copy_(dout[a:b], non_blocking=True)
    return f
for nci, nco in ((16, 16), (32, 272), (32, 68), (8, 8)):
    print(f"both, {nci} h2d pieces, {nco} d2h pieces ms", timeit(chunked(nci, nco)))
def d2h_only(nc):
    def f():
        with torch.cuda.stream(s2):
            for c in range(nc):
                a, b = D * c // nc, D * (c + 1) // nc
                hout[a:b].copy_(dout[a:b], non_blocking=True)
    return f
for nc in (1, 17, 68, 272):
    print(f"d2h only {nc} pieces ms", timeit(d2h_only(nc)))
