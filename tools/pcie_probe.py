"""Host<->device copy bandwidth of the box (pinned memory), alone and both directions at once:
the ceiling of bench.py's e2e number."""
import torch, time
n = 369098752
h_in = torch.empty(n, dtype=torch.uint8, pin_memory=True); h_out = torch.empty(n, dtype=torch.uint8, pin_memory=True)
d_in = torch.empty(n, dtype=torch.uint8, device="cuda"); d_out = torch.empty(n, dtype=torch.uint8, device="cuda")
s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
def t(fn, reps=5):
    fn(); torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(reps): fn()
    torch.cuda.synchronize(); return (time.perf_counter() - t0) / reps
def h2d():
    with torch.cuda.stream(s1): d_in.copy_(h_in, non_blocking=True)
def d2h():
    with torch.cuda.stream(s2): h_out.copy_(d_out, non_blocking=True)
def both(): h2d(); d2h()
for name, fn in (("h2d", h2d), ("d2h", d2h), ("both", both)):
    dt = t(fn); print(f"{name}: {dt*1e3:.2f} ms  {n/dt/1e9:.1f} GB/s per direction")
# chunked both directions (37-block chunks of 13.3 MB)
ck = 37 * 22 * 16384
def chunked():
    for o in range(0, n, ck):
        with torch.cuda.stream(s1): d_in[o:o+ck].copy_(h_in[o:o+ck], non_blocking=True)
        with torch.cuda.stream(s2): h_out[o:o+ck].copy_(d_out[o:o+ck], non_blocking=True)
dt = t(chunked); print(f"chunked both: {dt*1e3:.2f} ms  {n/dt/1e9:.1f} GB/s per direction")
