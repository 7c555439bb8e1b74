"""Small invocations of every kernel family, for compute-sanitizer (memcheck / racecheck) runs:

    compute-sanitizer --tool memcheck  python tools/diag/sanitize_cases.py
    compute-sanitizer --tool racecheck python tools/diag/sanitize_cases.py

Sizes are chosen so that every code path runs (edge blocks, ragged last blocks, odd widths, both kernel
families, pffft corrections, staged / peer / push exchange with all ranks emulated on one GPU, box kernels,
DFT_image) while the whole script stays within a minute under the sanitizer."""
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import blur_algorithms_b200 as bb  # noqa: E402

rng = np.random.default_rng(0)


def img(r, c):
    return rng.integers(0, 256, (r, c, 3), dtype=np.uint8)


for rows, cols, sigma, mode in [(270, 480, 5.0, 0), (270, 480, 5.0, 1), (301, 1203, 9.0, 0), (1100, 260, 3.0, 1), (400, 1300, 80.0, 0)]:
    d = torch.from_numpy(img(rows, cols)).cuda()
    p = bb.GaussPlan(rows, cols, sigma, mode)
    p.run(d)
    torch.cuda.synchronize()
    print("gauss", rows, cols, sigma, mode, p.info["ols_L_row"], p.info["ols_L_col"], int(d.sum()) > 0)
    p.close()
os.environ["BLUR_B200_COL_FUSE"] = "1"
d = torch.from_numpy(img(700, 1024)).cuda()
p = bb.GaussPlan(700, 1024, 6.0, 0)
p.run(d); p.run(d)
torch.cuda.synchronize()
p.close()
os.environ.pop("BLUR_B200_COL_FUSE")
print("fused interleave ok")
# sharded, all ranks in this process
for rows, cols, sigma, P, S, mode in [(260, 4096, 30.0, 4, 2, 0), (260, 4096, 30.0, 2, 4, 1)]:
    im = img(rows, cols)
    d_img = torch.from_numpy(im).cuda()
    for how in ("staged", "peer", "push"):
        plans = [bb.GaussPlan(rows, cols, sigma, mode, rank=r, nranks=P) for r in range(P)]
        bf = plans[0].info["block_floats"]
        recv = [torch.zeros(P * bf, dtype=torch.float32, device="cuda") for _ in range(P)]
        ptrs = [t.data_ptr() for t in recv]
        stripes = [d_img[pl.info["row0"]:pl.info["row0"] + pl.info["valid_rows"]].contiguous() for pl in plans]
        bufs = [pl.quirk_buffers() for pl in plans] if mode else None
        for r, pl in enumerate(plans):
            if mode:
                pl.quirk_local(stripes[r])
            if how == "staged" and pl.staged_supported(S):
                pl.row_pass_staged(stripes[r], ptrs, S)
            elif how == "push":
                pl.row_pass_push(stripes[r], ptrs, 2)
            else:
                pl.row_pass_peer(stripes[r], ptrs)
        torch.cuda.synchronize()
        if mode:
            tc, ti = sum(b[0].clone() for b in bufs), sum(b[1].clone() for b in bufs)
            for b in bufs:
                b[0].copy_(tc); b[1].copy_(ti)
        for r, pl in enumerate(plans):
            if mode:
                pl.quirk_finish()
            oc = torch.zeros((rows, 2 * pl.info["cpp"], 3), dtype=torch.uint8, device="cuda")
            if how == "staged" and pl.staged_supported(S):
                for k in range(S):
                    for q in plans:
                        q.stage_wait(k)
                    pl.col_pass_stage(recv[r], oc, k, S)
            else:
                pl.col_pass(recv[r], oc)
        torch.cuda.synchronize()
        for pl in plans:
            pl.close()
        print("sharded", how, mode, "ok")
b = img(270, 480)
bb.fastboxblur(b, 480, 270, 3, 31, 2)
b4 = rng.integers(0, 256, (100, 77, 1), dtype=np.uint8)
bb.fastboxblur(b4, 77, 100, 1, 9, 1)
print("box ok")
bb.dft_image(img(200, 300), 5.0)
bb.boxfft_blur(img(200, 300), 3.0)
pl = torch.empty((3, 1000), dtype=torch.float32, device="cuda")
bb.deinterleave_BGR(torch.from_numpy(rng.integers(0, 65536, 3000, dtype=np.uint16)).cuda(), pl, 1000)
torch.cuda.synchronize()
print("all cases ran")
