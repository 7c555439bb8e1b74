import os, sys, time, numpy as np, torch
sys.path.insert(0, "/root/repo")
import blur_algorithms_b200 as bb
rows, cols = 1080, 1920
img = np.random.default_rng(0).integers(0, 256, (rows, cols, 3), dtype=np.uint8)
h = [torch.from_numpy(img.copy()).pin_memory() for _ in range(2)]
o = [torch.empty_like(h[0]).pin_memory() for _ in range(2)]
for d in ("1", "0"):
    os.environ["BLUR_B200_DIRECT"] = d
    p = bb.GaussPlan(rows, cols, 5.0, 0)
    ins = [h[i % 2].numpy() for i in range(16)]; outs = [o[i % 2].numpy() for i in range(16)]
    p.run_host_batch(ins[:2], outs[:2])
    ts = []
    for rep in range(5):
        t0 = time.perf_counter(); p.run_host_batch(ins, outs); ts.append((time.perf_counter() - t0) / 16 * 1e3)
    print("direct", d, "ms/image", [round(t, 3) for t in ts])
    p.close()
