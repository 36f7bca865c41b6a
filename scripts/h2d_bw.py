"""Host->device copy bandwidth of this box from pinned memory (what bounds the float32 host path: 8.29 MB per 1080p frame)."""
import time, torch
n = 1 << 30
h = torch.empty(n, dtype=torch.uint8, pin_memory=True); d = torch.empty(n, dtype=torch.uint8, device="cuda")
for _ in range(2): d.copy_(h, non_blocking=True)
torch.cuda.synchronize(); t0 = time.perf_counter()
for _ in range(5): d.copy_(h, non_blocking=True)
torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 5
print(f"H2D pinned 1 GiB: {n / dt / 1e9:.1f} GB/s  -> {n / dt / (1920 * 1080 * 4):.0f} float32 1080p frames/s, {n / dt / (1920 * 1080):.0f} 8-bit frames/s")
hp = torch.empty(n, dtype=torch.uint8)
t0 = time.perf_counter(); d.copy_(hp); torch.cuda.synchronize(); print(f"H2D pageable: {n / (time.perf_counter() - t0) / 1e9:.1f} GB/s")
