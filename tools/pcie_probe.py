"""Host<->device copy bandwidth of this box for the frame size of the headline config (page-locked memory)."""
import torch, time
n = 3840 * 2160 * 3
d = torch.empty(n, dtype=torch.float64, device="cuda")
h = torch.empty(n, dtype=torch.float64).pin_memory()
for name, fn in (("d2h", lambda: h.copy_(d, non_blocking=True)), ("h2d", lambda: d.copy_(h, non_blocking=True))):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 10
    print(f"{name}: {n * 8 / 1e6:.1f} MB in {ms:.3f} ms = {n * 8 / ms / 1e6:.1f} GB/s")
# split into 6 chunks on one stream (what the render path does per batch)
chunks = list(zip(d.chunk(6), h.chunk(6)))
torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10):
    for a, b in chunks: b.copy_(a, non_blocking=True)
e1.record(); torch.cuda.synchronize()
print(f"d2h in 6 chunks: {e0.elapsed_time(e1) / 10:.3f} ms")
