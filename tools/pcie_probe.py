"""Developer tool: pinned-memory PCIe bandwidth of this box (H2D, D2H, both at once) for sizing the host-buffer pipeline."""
import torch, time
n = 64 << 20
h1 = torch.empty(n, dtype=torch.uint8).pin_memory(); h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
d1 = torch.empty(n, dtype=torch.uint8, device="cuda"); d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def t(fn, reps=10):
    fn(); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps
a = t(lambda: d1.copy_(h1, non_blocking=True)); b = t(lambda: h2.copy_(d2, non_blocking=True))
def both():
    with torch.cuda.stream(s1): d1.copy_(h1, non_blocking=True)
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
c = t(both)
print(f"H2D {n/a/1e9:.1f} GB/s  D2H {n/b/1e9:.1f} GB/s  both at once {n/c/1e9:.1f} GB/s each way")
for sz in (1 << 20, 4 << 20, 8 << 20):
    x = t(lambda: d1[:sz].copy_(h1[:sz], non_blocking=True), 50)
    print(f"  H2D {sz >> 20} MB: {sz/x/1e9:.1f} GB/s ({x*1e6:.0f} us)")
