"""GPU diagnostic: pinned host -> device copy bandwidth (what bounds the e2e arm of bench.py)."""
import torch
for mb in (16, 80, 320):
    h = torch.empty(mb << 20, dtype=torch.uint8).pin_memory()
    d = torch.empty(mb << 20, dtype=torch.uint8, device='cuda')
    for _ in range(3):
        d.copy_(h, non_blocking=True)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(10):
        d.copy_(h, non_blocking=True)
    e1.record()
    torch.cuda.synchronize()
    print('H2D %4d MiB: %.1f GB/s' % (mb, 10 * (mb << 20) / (e0.elapsed_time(e1) * 1e-3) / 1e9))
