"""1 GPU: cost of the row pass in sharded (nparts=2) addressing vs the single-GPU plan on the same rows."""
import os, sys
import numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import blur_algorithms_b200 as bb
dev = torch.device("cuda", 0)
rows, cols, sigma = 16384, 16384, 50.0
img = torch.from_numpy(np.random.default_rng(0).integers(0, 256, (rows // 2, cols, 3), dtype=np.uint8)).to(dev)
def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n
p2 = bb.GaussPlan(rows, cols, sigma, 0, device=0, rank=0, nranks=2)
q2 = torch.zeros(2 * p2.info["block_floats"], dtype=torch.float32, device=dev)
print("sharded nranks=2 row pass (8192 rows):", round(t(lambda: p2.row_pass(img, q2)), 3), "ms")
p1 = bb.GaussPlan(rows // 2, cols, sigma, 0, device=0)
q1 = torch.zeros(p1.info["block_floats"], dtype=torch.float32, device=dev)
print("single plan 8192x16384 row pass       :", round(t(lambda: p1.row_pass(img, q1)), 3), "ms")
