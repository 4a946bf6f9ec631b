"""Pinned host<->device copy bandwidth of the box (the floor of the end-to-end number)."""
import torch, time
n = 68_000_000
h = torch.empty(n, dtype=torch.uint8).pin_memory(); d = torch.empty(n, dtype=torch.uint8, device="cuda")
h2 = torch.empty(16_000_000, dtype=torch.uint8).pin_memory(); d2 = torch.empty(16_000_000, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
for name, fn in (("D2H 68MB", lambda: h.copy_(d, non_blocking=True)), ("H2D 16MB", lambda: d2.copy_(h2, non_blocking=True))):
    for _ in range(3): fn()
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(20): fn()
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 20
    print(name, f"{dt*1e3:.3f} ms", f"{(n if 'D2H' in name else 16e6)/dt/1e9:.1f} GB/s")
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(20):
    with torch.cuda.stream(s1): h.copy_(d, non_blocking=True)
    with torch.cuda.stream(s2): d2.copy_(h2, non_blocking=True)
torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 20
print("both directions concurrently", f"{dt*1e3:.3f} ms per pair")
