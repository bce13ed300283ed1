import torch, time
torch.cuda.init()
for mb in (1.5, 3.67, 4.7, 17.3):
    n = int(mb * 1e6) // 8
    h = torch.empty(n, dtype=torch.float64).pin_memory()
    d = torch.empty(n, dtype=torch.float64, device="cuda")
    for name, fn in (("h2d", lambda: d.copy_(h, non_blocking=True)), ("d2h", lambda: h.copy_(d, non_blocking=True))):
        for _ in range(3): fn()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(20): fn()
        e.record(); torch.cuda.synchronize()
        ms = s.elapsed_time(e) / 20
        print("%5.2f MB %s: %.3f ms  %.1f GB/s" % (mb, name, ms, mb * 1e-3 / ms))
# concurrent h2d + d2h on two streams
n = int(17.3e6) // 8; m = int(3.67e6) // 8
h1 = torch.empty(n, dtype=torch.float64).pin_memory(); d1 = torch.empty(n, dtype=torch.float64, device="cuda")
h2 = torch.empty(m, dtype=torch.float64).pin_memory(); d2 = torch.empty(m, dtype=torch.float64, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
torch.cuda.synchronize(); t = time.perf_counter()
for _ in range(20):
    with torch.cuda.stream(s1): d1.copy_(h1, non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
torch.cuda.synchronize(); print("overlapped 17.3 h2d + 3.67 d2h: %.3f ms" % ((time.perf_counter() - t) / 20 * 1e3))
