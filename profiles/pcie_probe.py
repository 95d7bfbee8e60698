import torch, time
n = 1 << 30   # 1 GiB
h1 = torch.empty(n, dtype=torch.uint8).pin_memory(); h2 = torch.empty(n, dtype=torch.uint8).pin_memory()
d1 = torch.empty(n, dtype=torch.uint8, device="cuda"); d2 = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def t(fn, reps=5):
    fn(); torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize()
    return (time.perf_counter() - t0) / reps
def h2d():
    with torch.cuda.stream(s1): d1.copy_(h1, non_blocking=True)
def d2h():
    with torch.cuda.stream(s2): h2.copy_(d2, non_blocking=True)
def both(): h2d(); d2h()
a, b, c = t(h2d), t(d2h), t(both)
print(f"H2D {n/a/1e9:.1f} GB/s  D2H {n/b/1e9:.1f} GB/s  both: {2*n/c/1e9:.1f} GB/s total ({n/c/1e9:.1f} each)")
# chunked copies (64 MB pieces) as the pipeline issues them
ch = 64 << 20
def h2d_c():
    with torch.cuda.stream(s1):
        for o in range(0, n, ch): d1[o:o+ch].copy_(h1[o:o+ch], non_blocking=True)
def d2h_c():
    with torch.cuda.stream(s2):
        for o in range(0, n, ch): h2[o:o+ch].copy_(d2[o:o+ch], non_blocking=True)
def both_c(): h2d_c(); d2h_c()
a, b, c = t(h2d_c), t(d2h_c), t(both_c)
print(f"chunked 64MB: H2D {n/a/1e9:.1f}  D2H {n/b/1e9:.1f}  both total {2*n/c/1e9:.1f}")
