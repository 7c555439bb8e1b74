"""torchrun check + timing of the row-stripe (halo) sharded blur: bit-identity against the single-GPU plan, then ms per blur
(CUDA events, max over ranks) next to the contract configuration's default exchange.
    torchrun --nproc-per-node N tools/diag/halo_check.py [rows cols sigma]"""
import json, os, sys
import numpy as np
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import blur_algorithms_b200 as bb
from blur_algorithms_b200.dist import HaloShardedGaussBlur, ShardedGaussBlur

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
rank, world = dist.get_rank(), dist.get_world_size()
args = [a for a in sys.argv[1:]]
rows, cols, sigma = (int(args[0]), int(args[1]), float(args[2])) if len(args) >= 3 else (16384, 16384, 50.0)
img = np.random.default_rng(0).integers(0, 256, (rows, cols, 3), dtype=np.uint8)
hb = HaloShardedGaussBlur(rows, cols, sigma, device=dev)
stripe = torch.from_numpy(img[hb.row0:hb.row0 + hb.valid_rows].copy()).to(dev)
got = hb.run(stripe).clone()
full = torch.from_numpy(img).to(dev)
ref = hb.plan.run(full, torch.empty_like(full))
same = bool(torch.equal(got, ref[hb.row0:hb.row0 + hb.valid_rows]))
t = torch.tensor([int(same)], device=dev); dist.all_reduce(t, op=dist.ReduceOp.MIN)
del full, ref


def timed(fn, steps=20, warm=3):
    for _ in range(warm):
        fn()
    dist.barrier(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record(); dist.barrier(); torch.cuda.synchronize()
    ms = torch.tensor([e0.elapsed_time(e1) / steps], device=dev, dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return float(ms.item())


halo_ms = timed(lambda: hb.run(stripe))
sb = ShardedGaussBlur(rows, cols, sigma, bb.NYQUIST_EXACT, device=dev)
g = sb.geom
s2 = torch.from_numpy(img[g.row0:g.row0 + g.valid_rows].copy()).to(dev)
a2a_ms = timed(lambda: sb.run(s2))
if rank == 0:
    print(json.dumps({"check": "halo sharding", "n_gpus": world, "rows": rows, "cols": cols, "sigma": sigma, "identical": bool(t.item()),
                      "halo_ms_per_blur": halo_ms, "contract_exchange_ms_per_blur": a2a_ms, "contract_exchange": sb.exchange_mode,
                      "stripe_rows": hb.info["cap"], "halo_rows": hb.info["halo"],
                      "halo_bytes_out_per_gpu": 2 * hb.info["halo"] * hb.info["qpitch"] * 12}), flush=True)
dist.barrier(); dist.destroy_process_group()
sys.exit(0 if t.item() else 1)
