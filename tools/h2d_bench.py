"""Pinned host -> device copy rate of one 154 MB batch, as 1 / 2 / 4 chunks on as many streams (dev tool, GPU)."""
import torch

B = 256
host = torch.randn(B, 3, 224, 224).pin_memory()
dev = torch.empty_like(host, device="cuda")
for parts in (1, 2, 4, 8):
    streams = [torch.cuda.Stream() for _ in range(parts)]
    hs, ds = host.chunk(parts), dev.chunk(parts)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 20
    e0.record()
    for _ in range(reps):
        evs = []
        for s, h, d in zip(streams, hs, ds):
            s.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(s):
                d.copy_(h, non_blocking=True)
            evs.append(s)
        for s in evs:
            torch.cuda.current_stream().wait_stream(s)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / reps
    print(f"{parts} chunk(s): {ms:.3f} ms per batch, {host.numel() * 4 / ms / 1e6:.1f} GB/s")
