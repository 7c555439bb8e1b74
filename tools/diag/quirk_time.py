"""ncu helper: one pffft-mode blur of 16384^2 sigma=50 (per-kernel durations of the quirk path)."""
import os, sys, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
import blur_algorithms_b200 as bb
d = torch.from_numpy(np.random.default_rng(0).integers(0, 256, (16384, 16384, 3), dtype=np.uint8)).cuda()
o = torch.empty_like(d)
p = bb.GaussPlan(16384, 16384, 50.0, bb.NYQUIST_PFFFT)
for _ in range(2):
    p.run(d, o)
torch.cuda.synchronize()
