"""Host-to-device copy rate of pinned feature blocks (the floor of every *_host entry point)."""
import torch
for mb in (53, 368):
    h = torch.empty(mb * 2**20, dtype=torch.uint8).pin_memory()
    d = torch.empty_like(h, device="cuda")
    d.copy_(h, non_blocking=True)
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(5):
        d.copy_(h, non_blocking=True)
    b.record()
    torch.cuda.synchronize()
    print(f"H2D {mb} MiB pinned: {mb * 2**20 * 5 / a.elapsed_time(b) / 1e6:.1f} GB/s")
