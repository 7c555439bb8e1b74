"""Raw peer-copy bandwidth between GPU 0 and GPU 1 (single process) + topology."""
import subprocess, torch
print(subprocess.run(["nvidia-smi", "topo", "-m"], capture_output=True, text=True).stdout)
n = torch.cuda.device_count()
print("devices", n, "p2p 0->1", torch.cuda.can_device_access_peer(0, 1) if n > 1 else None)
a = torch.empty(1 << 30, dtype=torch.uint8, device="cuda:0")
b = torch.empty(1 << 30, dtype=torch.uint8, device="cuda:1")
for name, dst, src in (("0->1", b, a), ("1->0", a, b)):
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    dst.copy_(src); torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.device(src.device):
        e0.record()
        for _ in range(10):
            dst.copy_(src, non_blocking=True)
        e1.record()
    torch.cuda.synchronize(0); torch.cuda.synchronize(1)
    print(name, "GB/s", 10 * (1 << 30) / 1e9 / (e0.elapsed_time(e1) * 1e-3))
# bidirectional
s0, s1 = torch.cuda.Stream(0), torch.cuda.Stream(1)
a2 = torch.empty(1 << 30, dtype=torch.uint8, device="cuda:0"); b2 = torch.empty(1 << 30, dtype=torch.uint8, device="cuda:1")
import time
torch.cuda.synchronize(0); torch.cuda.synchronize(1)
t0 = time.perf_counter()
for _ in range(10):
    with torch.cuda.stream(s0): b.copy_(a, non_blocking=True)
    with torch.cuda.stream(s1): a2.copy_(b2, non_blocking=True)
torch.cuda.synchronize(0); torch.cuda.synchronize(1)
dt = time.perf_counter() - t0
print("bidir per direction GB/s", 10 * (1 << 30) / 1e9 / dt)
