"""H2D / D2H rate of pinned copies on this box (what bounds the end-to-end bench)."""
import torch
n = 466_616_000
h = torch.empty(n, dtype=torch.uint8).pin_memory()
d = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(177_544_000, dtype=torch.uint8).pin_memory()
d2 = torch.empty(177_544_000, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def t(fn, reps=5):
    best = 1e9
    for _ in range(reps):
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        fn()
        b.record()
        torch.cuda.synchronize()
        best = min(best, a.elapsed_time(b))
    return best
ms = t(lambda: d.copy_(h, non_blocking=True))
print("H2D 466 MB alone: %.2f ms = %.1f GB/s" % (ms, n / ms / 1e6))
ms = t(lambda: h2.copy_(d2, non_blocking=True))
print("D2H 177 MB alone: %.2f ms = %.1f GB/s" % (ms, 177_544_000 / ms / 1e6))
def both():
    cur = torch.cuda.current_stream()
    s1.wait_stream(cur); s2.wait_stream(cur)
    with torch.cuda.stream(s1):
        d.copy_(h, non_blocking=True)
    with torch.cuda.stream(s2):
        h2.copy_(d2, non_blocking=True)
    cur.wait_stream(s1); cur.wait_stream(s2)
ms = t(both)
print("H2D 466 MB + D2H 177 MB concurrently: %.2f ms" % ms)
def chunks():
    for i in range(0, n, n // 8):
        d[i:i + n // 8].copy_(h[i:i + n // 8], non_blocking=True)
ms = t(chunks)
print("H2D in 8 chunks: %.2f ms" % ms)
