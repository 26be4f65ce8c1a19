"""Bring-up probe for a multi-GPU box: peer-access matrix and device-to-device copy bandwidth (GPU 0 -> every peer)."""
import time
import torch
n = torch.cuda.device_count()
print("devices", n)
print("peer access from 0:", [torch.cuda.can_device_access_peer(0, d) for d in range(1, n)])
src = torch.empty(256 << 20, dtype=torch.uint8, device="cuda:0")
for d in range(1, n):
    dst = torch.empty_like(src, device=f"cuda:{d}")
    for _ in range(2):
        dst.copy_(src)
    torch.cuda.synchronize(0); torch.cuda.synchronize(d)
    t = time.perf_counter()
    for _ in range(5):
        dst.copy_(src)
    torch.cuda.synchronize(0); torch.cuda.synchronize(d)
    dt = (time.perf_counter() - t) / 5
    print(f"0 -> {d}: {src.numel() / dt / 1e9:.1f} GB/s")
