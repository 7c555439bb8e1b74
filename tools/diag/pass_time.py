"""Row pass / column pass / whole blur timings for kernel variants selected by environment knobs, in
one GPU visit.  Each variant is a fresh plan (the knobs are read at plan creation); outputs are
compared with the first variant's (max LSB difference, mismatching fraction).

    python tools/diag/pass_time.py [--rows R --cols C --sigma S] [--variants "OLS2=0;OLS2=1;..."] [--reps N]

A variant is a comma-separated list of NAME=VALUE pairs, NAME without the BLUR_B200_ prefix.
"""
import argparse
import json
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import blur_algorithms_b200 as bb  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=16384)
    ap.add_argument("--cols", type=int, default=16384)
    ap.add_argument("--sigma", type=float, default=50.0)
    ap.add_argument("--mode", type=int, default=0)
    ap.add_argument("--reps", type=int, default=10)
    ap.add_argument("--variants", default="OLS2=0;OLS2=1;OLS2=row;OLS2=col")
    a = ap.parse_args()
    rows, cols = a.rows, a.cols
    rng = np.random.default_rng(0)
    d_in = torch.from_numpy(rng.integers(0, 256, (rows, cols, 3), dtype=np.uint8)).cuda()
    first = None
    stream = torch.cuda.current_stream()
    for var in a.variants.split(";"):
        keys = []
        for kv in filter(None, var.split(",")):
            k, v = kv.split("=")
            os.environ["BLUR_B200_" + k] = v
            keys.append("BLUR_B200_" + k)
        plan = bb.GaussPlan(rows, cols, a.sigma, a.mode)
        d_out = torch.empty_like(d_in)
        q = torch.empty(plan.info["block_floats"], dtype=torch.float32, device="cuda")
        oc = torch.empty((rows, 2 * plan.info["cpp"], 3), dtype=torch.uint8, device="cuda")
        for _ in range(2):
            plan.run(d_in, d_out); plan.row_pass(d_in, q); plan.col_pass(q, oc)
        torch.cuda.synchronize()
        e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        e[0].record(stream)
        for _ in range(a.reps):
            plan.row_pass(d_in, q)
        e[1].record(stream)
        for _ in range(a.reps):
            plan.col_pass(q, oc)
        e[2].record(stream)
        for _ in range(a.reps):
            plan.run(d_in, d_out)
        e[3].record(stream)
        torch.cuda.synchronize()
        res = {"variant": var, "row_ms": e[0].elapsed_time(e[1]) / a.reps, "col_il_ms": e[1].elapsed_time(e[2]) / a.reps,
               "blur_ms": e[2].elapsed_time(e[3]) / a.reps,
               "L": (plan.info["ols_L_row"], plan.info["ols_L_col"]), "nblocks": (plan.info["ols_nblocks_row"], plan.info["ols_nblocks_col"])}
        out = d_out.cpu()
        if first is None:
            first = out
        else:
            d = (out.to(torch.int16) - first.to(torch.int16)).abs()
            res["max_lsb_vs_first"] = int(d.max())
            res["mismatch_frac_vs_first"] = float((d > 0).float().mean())
        print(json.dumps(res), flush=True)
        plan.close()
        del q, oc, d_out
        for k in keys:
            os.environ.pop(k, None)


if __name__ == "__main__":
    main()
