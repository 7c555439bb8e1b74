"""torchrun check: sharded blur over WORLD_SIZE GPUs is bit-identical to the single-GPU blur."""
import os, sys
import numpy as np
import torch
import torch.distributed as dist
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import blur_algorithms_b200 as bb
from blur_algorithms_b200.dist import ShardedGaussBlur

local = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local)
dev = torch.device("cuda", local)
dist.init_process_group("nccl", device_id=dev)
rank, world = dist.get_rank(), dist.get_world_size()
ok = True
cases = [(1080, 1920, 5.0), (2160, 3840, 25.0), (1001, 1531, 9.0), (4096, 4096, 50.0)]
if "--headline" in sys.argv:        # BASELINE config 3 (tests/test_gpu_parity.py::test_gauss_sharded_multi_gpu_bit_identity)
    cases = [(1001, 1531, 9.0), (16384, 16384, 50.0)]
for rows, cols, sigma in cases:
    img = np.random.default_rng(7).integers(0, 256, (rows, cols, 3), dtype=np.uint8)
    sb = ShardedGaussBlur(rows, cols, sigma, bb.NYQUIST_EXACT, device=dev)
    g = sb.geom
    stripe = torch.from_numpy(img[g.row0:g.row0 + g.valid_rows].copy()).to(dev)
    out_cols = sb.run(stripe).clone()
    full = torch.from_numpy(img).to(dev)
    plan = bb.GaussPlan(rows, cols, sigma, bb.NYQUIST_EXACT, device=local, direct=False)   # narrow kernels: FFT kernels like the sharded plans
    ref = plan.run(full, torch.empty_like(full))
    mine = ref[:, g.col0:g.col0 + g.valid_cols]
    same = bool(torch.equal(mine, out_cols)) if g.valid_cols > 0 else True
    # return exchange: back to the row stripes the input came in
    back = sb.to_row_stripe()
    same = same and bool(torch.equal(back, ref[g.row0:g.row0 + g.valid_rows]))
    # pffft mode (Nyquist corrections summed over the ranks) against the single-GPU pffft-mode plan
    if (rows, cols) != (16384, 16384) or "--headline" in sys.argv:
        sbq = ShardedGaussBlur(rows, cols, sigma, bb.NYQUIST_PFFFT, device=dev)
        planq = bb.GaussPlan(rows, cols, sigma, bb.NYQUIST_PFFFT, device=local, direct=False)
        refq = planq.run(full, torch.empty_like(full))
        gotq = sbq.run(stripe)
        same = same and (bool(torch.equal(refq[:, g.col0:g.col0 + g.valid_cols], gotq)) if g.valid_cols > 0 else True)
        planq.close(); del sbq, refq
    # streamed host path: three different images through run_host_stream == run() on each
    imgs = [np.random.default_rng(70 + i).integers(0, 256, (rows, cols, 3), dtype=np.uint8) for i in range(3)]
    h_ins = [torch.from_numpy(im[g.row0:g.row0 + g.valid_rows].copy()).pin_memory() for im in imgs]
    h_outs = [torch.empty((rows, 2 * g.cpp, 3), dtype=torch.uint8).pin_memory() for _ in imgs]
    sb.run_host_stream(h_ins, h_outs)
    for hi, ho in zip(h_ins, h_outs):
        want = sb.run(hi.to(dev)).clone()
        same = same and bool(torch.equal(want.cpu(), ho[:, :g.valid_cols]))
    t = torch.tensor([int(same)], device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MIN)
    if rank == 0:
        print(f"{rows}x{cols} sigma={sigma}: algo={g.algo} stripes={g.stripe_rows} cpp={g.cpp} identical={bool(t.item())} "
              f"exchange={sb.exchange_mode} trial_ms={sb.exchange_trial}", flush=True)
    ok = ok and bool(t.item())
dist.barrier()
dist.destroy_process_group()
sys.exit(0 if ok else 1)
