"""PCIe floor for the e2e figure: pinned H2D, D2H and both at once (two streams), 2 GiB each."""
import torch, time
n = 2 ** 31
h_in = torch.empty(n, dtype=torch.uint8).pin_memory()
h_out = torch.empty(n, dtype=torch.uint8).pin_memory()
d_in = torch.empty(n, dtype=torch.uint8, device="cuda")
d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def run(h2d, d2h, chunk=None):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    if chunk is None:
        if h2d:
            with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
        if d2h:
            with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
    else:
        for o in range(0, n, chunk):
            if h2d:
                with torch.cuda.stream(s1): d_in[o:o+chunk].copy_(h_in[o:o+chunk], non_blocking=True)
            if d2h:
                with torch.cuda.stream(s2): h_out[o:o+chunk].copy_(d_out[o:o+chunk], non_blocking=True)
    torch.cuda.synchronize(); return (time.perf_counter() - t0) * 1e3
for name, a, b, c in (("H2D", 1, 0, None), ("D2H", 0, 1, None), ("both", 1, 1, None), ("both 48MiB chunks", 1, 1, 48 << 20), ("both 16MiB chunks", 1, 1, 16 << 20)):
    run(a, b, c); ms = min(run(a, b, c) for _ in range(3))
    print(f"{name}: {ms:.1f} ms for 2 GiB each -> {n/ms/1e6:.1f} GB/s per direction")
