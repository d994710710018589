import torch, time
torch.cuda.set_device(0)
h_in = torch.empty(8_000_000_000, dtype=torch.uint8).pin_memory()
h_out = torch.empty(12_000_000_000, dtype=torch.uint8).pin_memory()
d_in = torch.empty(8_000_000_000, dtype=torch.uint8, device="cuda")
d_out = torch.empty(12_000_000_000, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(h2d, d2h):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    if h2d:
        with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
    if d2h:
        with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
    torch.cuda.synchronize(); return time.perf_counter() - t0
for _ in range(2):
    a, b, c = run(True, False), run(False, True), run(True, True)
    print(f"H2D 8 GB alone {a*1e3:.0f} ms ({8/a:.1f} GB/s); D2H 12 GB alone {b*1e3:.0f} ms ({12/b:.1f} GB/s); both at once {c*1e3:.0f} ms")
