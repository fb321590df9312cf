import torch, time
dev = torch.device("cuda:0")
H = 124_000_000
hin = torch.empty(H, dtype=torch.uint8).pin_memory(); din = torch.empty(H, dtype=torch.uint8, device=dev)
widths = [20, 1, 4, 8, 8, 4, 4, 1, 8]
N = 1_000_000
hcols = [torch.empty(N * w, dtype=torch.uint8).pin_memory() for w in widths]
dcols = [torch.empty(N * w, dtype=torch.uint8, device=dev) for w in widths]
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(h2d_pieces, groups, delay_ms=0.0):
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True); e2 = torch.cuda.Event(enable_timing=True)
    e0.record()
    s1.wait_event(e0); s2.wait_event(e0)
    with torch.cuda.stream(s1):
        for c in range(h2d_pieces):
            a, b = H * c // h2d_pieces, H * (c + 1) // h2d_pieces
            din[a:b].copy_(hin[a:b], non_blocking=True)
        e1.record()
    with torch.cuda.stream(s2):
        if delay_ms: torch.cuda._sleep(int(delay_ms * 1.9e6))
        for g in range(groups):
            lo, hi = N * g // groups, N * (g + 1) // groups
            for w, hc, dc in zip(widths, hcols, dcols):
                hc[lo * w:hi * w].copy_(dc[lo * w:hi * w], non_blocking=True)
        e2.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1), e0.elapsed_time(e2)
for args in ((1, 1), (20, 1), (20, 4), (20, 10), (20, 16), (1, 16), (20, 4, 0.7)):
    run(*args)
    r = [run(*args) for _ in range(3)]
    print(args, "h2d done %.2f  d2h done %.2f" % min(r))
# one column only, big
def run1():
    torch.cuda.synchronize()
    e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True); e2 = torch.cuda.Event(enable_timing=True)
    e0.record(); s1.wait_event(e0); s2.wait_event(e0)
    with torch.cuda.stream(s1):
        din.copy_(hin, non_blocking=True); e1.record()
    with torch.cuda.stream(s2):
        hin2.copy_(dbig, non_blocking=True); e2.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1), e0.elapsed_time(e2)
hin2 = torch.empty(58_000_000, dtype=torch.uint8).pin_memory(); dbig = torch.empty(58_000_000, dtype=torch.uint8, device=dev)
run1(); print("single 58MB d2h + single h2d: h2d %.2f d2h %.2f" % min(run1() for _ in range(3)))
