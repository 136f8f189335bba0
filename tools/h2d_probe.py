"""Probe: pinned H2D bandwidth per 64 MiB slab of one large pinned allocation."""
import torch, time, sys
d = torch.empty(64 << 20, dtype=torch.uint8, device="cuda")
def bw(h, reps=3):
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    d.copy_(h, non_blocking=True); torch.cuda.synchronize()
    a.record()
    for _ in range(reps): d.copy_(h, non_blocking=True)
    b.record(); torch.cuda.synchronize()
    return h.numel() * reps / a.elapsed_time(b) / 1e6
gib = int(sys.argv[1]) if len(sys.argv) > 1 else 4
t0 = time.time()
big = torch.empty(gib << 30, dtype=torch.uint8, pin_memory=True)
print("alloc s", round(time.time() - t0, 2))
print([round(bw(big[i << 26:(i + 1) << 26])) for i in range(gib * 16)])
