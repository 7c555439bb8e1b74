import os, sys, numpy as np, torch
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import blur_algorithms_b200 as bb
dev = torch.device("cuda", 0)
w, h, batch = 1920, 1080, 256
d = torch.from_numpy(np.random.default_rng(0).integers(0, 256, (batch, h, w, 3), dtype=np.uint8)).to(dev)
s = torch.empty_like(d)
for _ in range(2): bb.fastboxblur_device(d, w, h, 3, 31, 2, batch, s)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5): bb.fastboxblur_device(d, w, h, 3, 31, 2, batch, s)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(os.environ.get("BLUR_B200_BOX_GROUP"), "ms", round(ms, 3), "Gpx/s", round(batch * w * h / ms / 1e6, 1))
