#!/usr/bin/env python
"""bench.py — megapixels/s of the RGB Gaussian blur (sigma = 50) hot path on B200.

  python bench.py --gpus N --steps K --warmup W             (N > 1: launched by torchrun)
  python bench.py --impl reference --gpus N --steps K --warmup W

Workload (BASELINE.json config 3): one 16384 x 16384 interleaved-BGR uint8 image, sigma = 50,
synthetic uniform noise (numpy default_rng(0)).  A "step" is one full blur of that image:
row pass -> (N > 1: all-to-all of the float32 intermediate) -> column pass -> interleave.
At N > 1 the single image is sharded by row stripes / column stripes over the ranks (strong
scaling), as the config prescribes.

`value`    : device-resident throughput (input already in HBM), CUDA events, max over ranks.
`e2e`      : same metric through the C-ABI host entry point with pinned HOST buffers — H2D of the
             image and D2H of the result inside the timed region.
`roofline` : dominant kernel (the slower of the two FFT passes): algorithmic bytes per launch
             (15 B/px: DESIGN.md) / its CUDA-event duration, against the measured HBM copy peak.
`cpu_baseline` (N = 1): the reference's own pffft_() pipeline (Source.cpp compiled unmodified,
             OpenMP, FFT engine = the repo's CPU stand-in because pffft is absent) on the host
             cores, on a bounded sample of the workload.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

ROWS = COLS = 16384
SIGMA = 50.0
ALG_BYTES_PER_PX_STEP = 30.0      # 3 (u8 in) + 12 (f32 out) + 12 (f32 in) + 3 (u8 out), SURVEY.md §8d
ALG_BYTES_PER_PX_PASS = 15.0      # each FFT pass: 3 + 12
METRIC = "megapixels/sec RGB Gaussian blur (sigma=50)"


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""

    def __init__(self, index: int):
        self.index = index
        self.rows = []
        self.proc = None

    def start(self):
        q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={q}",
                                          "--format=csv,noheader,nounits", "-lms", "100"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([c.strip() for c in line.split(",")])

    def stop(self) -> dict:
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0])); mx = float(r[1])
                for n, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except Exception:
                continue
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


def cpu_reference_run(steps: int, warmup: int, sample_rows: int | None = None):
    """Times the reference's own pffft_ pipeline (oracle/_ref; the oracle port if _ref is absent)
    on the host cores on a bounded sample: a sample_rows x sample_rows sub-image, sigma = 50."""
    import numpy as np
    from oracle_lib import Oracle, Reference
    cores = os.cpu_count() or 1
    n = sample_rows or (8192 if cores >= 32 else 4096)
    rng = np.random.default_rng(0)
    img = rng.integers(0, 256, (n, n, 3), dtype=np.uint8)
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(1)
    if Reference.available():
        ref = Reference()
        kind, threads = "reference", ref.num_threads()
        fn = lambda: ref.test(img, 3, SIGMA)           # noqa: E731  Test(flag=3) -> pffft_
        what = "Source.cpp pffft_() compiled unmodified (OpenMP) against API shims; FFT engine = repo CPU stand-in (pffft absent)"
    else:
        o = Oracle()
        kind, threads = "port", o.num_threads()
        fn = lambda: o.gauss_blur_1d(img, SIGMA, 1)    # noqa: E731
        what = "C restatement of pffft_() (oracle/blur_oracle.c, OpenMP)"
    os.dup2(devnull, 1)                                # the reference printf()s its own timing
    try:
        for _ in range(warmup):
            fn()
        t0 = time.perf_counter()
        for _ in range(steps):
            fn()
        dt = (time.perf_counter() - t0) / max(steps, 1)
    finally:
        os.dup2(saved, 1)
        os.close(devnull)
        os.close(saved)
    mps = n * n / 1e6 / dt
    return {"value": mps, "unit": "megapixels/sec", "cores": threads, "kind": kind,
            "sample": f"{n}x{n} RGB sub-image of the workload, sigma=50, {steps} timed steps after {warmup} warm-up; {what}"}, dt


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--rows", type=int, default=ROWS)
    ap.add_argument("--cols", type=int, default=COLS)
    ap.add_argument("--sigma", type=float, default=SIGMA)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rows, cols, sigma = args.rows, args.cols, args.sigma
    config = {"workload": f"gauss_blur_{rows}x{cols}_rgb_u8_sigma{sigma:g}", "rows": rows, "cols": cols, "sigma": sigma,
              "nyquist_mode": "exact", "l2_policy": "inputs larger than L2 (805 MB image, 3.2 GB float32 intermediate)",
              "sharding": "single GPU" if args.gpus == 1 else f"row stripes x{args.gpus} + all-to-all + column stripes"}

    if args.impl == "reference":
        if rank != 0:
            return
        base, dt = cpu_reference_run(max(1, args.steps), max(0, args.warmup))
        line = {"metric": METRIC, "value": base["value"], "unit": "megapixels/sec", "n_gpus": args.gpus, "steps": args.steps,
                "warmup": args.warmup, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": "strong",
                "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": config, "impl": "reference",
                "cpu_baseline": base,
                "e2e": {"value": base["value"], "unit": "megapixels/sec", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return

    import numpy as np
    import torch
    import torch.distributed as dist

    import blur_algorithms_b200 as bb
    from blur_algorithms_b200.dist import ShardedGaussBlur

    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    multi = world > 1
    if multi:
        dist.init_process_group("nccl", device_id=dev)
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"

    def barrier():
        if multi:
            dist.barrier()
        torch.cuda.synchronize()

    rng = np.random.default_rng(0)
    stream = torch.cuda.current_stream()
    K, W = max(1, args.steps), max(3, args.warmup)
    npx = rows * cols

    if not multi:
        h_img = torch.from_numpy(rng.integers(0, 256, (rows, cols, 3), dtype=np.uint8)).pin_memory()
        d_in = h_img.to(dev)
        d_out = torch.empty_like(d_in)
        plan = bb.GaussPlan(rows, cols, sigma, bb.NYQUIST_EXACT, device=local)
        step = lambda: plan.run(d_in, d_out)            # noqa: E731
    else:
        fused_env = os.environ.get("BLUR_B200_FUSED")        # "0": NCCL all_to_all exchange instead of the fused peer stores
        sb = ShardedGaussBlur(rows, cols, sigma, bb.NYQUIST_EXACT, device=dev,
                              fused=None if fused_env is None else fused_env != "0")
        g = sb.geom
        # every rank draws the same image stream and keeps its stripe
        full_rng = np.random.default_rng(0)
        stripe_np = None
        chunk = 1024
        parts = []
        for r0 in range(0, rows, chunk):
            blk = full_rng.integers(0, 256, (min(chunk, rows - r0), cols, 3), dtype=np.uint8)
            lo, hi = max(r0, g.row0), min(r0 + blk.shape[0], g.row0 + g.valid_rows)
            if hi > lo:
                parts.append(blk[lo - r0:hi - r0])
        stripe_np = np.concatenate(parts) if parts else np.zeros((0, cols, 3), np.uint8)
        h_img = torch.from_numpy(stripe_np).pin_memory()
        d_in = h_img.to(dev)
        plan = sb.plan
        if not sb.fused:
            config["exchange"] = "NCCL all_to_all of the float32 intermediate"
        elif sb.exchange_mode == "store":
            config["exchange"] = "fused: row-pass epilogue stores into the owners' symmetric-memory buffers over NVLink"
        else:
            config["exchange"] = (f"pipelined: row pass in {sb.nchunks} chunks, q_push_kernel copies each finished chunk into the "
                                  "owners' symmetric-memory buffers over NVLink (128-bit peer stores) under the next chunk's FFTs")
        step = lambda: sb.run(d_in)                      # noqa: E731

    # ---- device-resident timing
    for _ in range(W):
        step()
    barrier()
    sampler = ClockSampler(local)
    sampler.start()
    n0 = bb.kernel_launches()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    barrier()
    ev0.record(stream)
    for _ in range(K):
        step()
    ev1.record(stream)
    barrier()
    launches = bb.kernel_launches() - n0
    ms = ev0.elapsed_time(ev1)
    clocks = sampler.stop()
    if multi:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    ms_per_step = ms / K
    value = npx / 1e6 / (ms_per_step * 1e-3)

    # ---- per-kernel timing of the two FFT passes (same stream, CUDA events) for the roofline
    peak, peak_src = hbm_peak()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    reps = max(3, min(K, 10))
    if not multi:
        q = torch.empty(plan.info["block_floats"], dtype=torch.float32, device=dev)
        oc = torch.empty((rows, 2 * plan.info["cpp"], 3), dtype=torch.uint8, device=dev)
        plan.row_pass(d_in, q); plan.col_pass(q, oc)
        torch.cuda.synchronize()
        e[0].record(stream)
        for _ in range(reps):
            plan.row_pass(d_in, q)
        e[1].record(stream)
        for _ in range(reps):
            plan.col_pass(q, oc)     # column FFT kernel + interleave kernel
        e[2].record(stream)
        torch.cuda.synchronize()
        row_ms = e[0].elapsed_time(e[1]) / reps
        colil_ms = e[1].elapsed_time(e[2]) / reps
        pass_px = npx
        del q, oc
    else:
        g = sb.geom
        if not sb.fused:
            rp = lambda: sb.plan.row_pass(d_in, sb.q_send)                                   # noqa: E731
        elif sb.exchange_mode == "store":
            rp = lambda: sb.plan.row_pass_peer(d_in, sb.peer_ptrs)                           # noqa: E731
        else:
            rp = lambda: sb.plan.row_pass_push(d_in, sb.peer_ptrs, sb.nchunks)               # noqa: E731
        barrier()
        rp(); sb.plan.col_pass(sb.q_recv, sb.out_cols)
        barrier()
        e[0].record(stream)
        for _ in range(reps):
            rp()
        e[1].record(stream)
        for _ in range(reps):
            sb.plan.col_pass(sb.q_recv, sb.out_cols)
        e[2].record(stream)
        torch.cuda.synchronize()
        row_ms = e[0].elapsed_time(e[1]) / reps
        colil_ms = e[1].elapsed_time(e[2]) / reps
        pass_px = g.valid_rows * cols
    algo = plan.info["algo"]
    kn = ("ols_row_kernel", "ols_col_kernel (+interleave_oct_kernel)") if algo == 1 else ("fft_conv_kernel<RowIO>", "fft_conv_kernel<ColIO> (+interleave_u8_kernel)")
    dom_name, dom_ms = (kn[0], row_ms) if row_ms >= colil_ms else (kn[1], colil_ms)
    achieved = ALG_BYTES_PER_PX_PASS * pass_px / (dom_ms * 1e-3) / 1e9
    traffic = None
    tp = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp):
        try:
            traffic = json.load(open(tp)).get("dominant_kernel_dram_bytes_per_launch")
        except Exception:
            traffic = None
    roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": traffic, "kernel": dom_name, "kernel_ms": dom_ms, "row_pass_ms": row_ms,
                "col_pass_plus_interleave_ms": colil_ms, "peak_source": peak_src,
                "step_achieved_GBps": ALG_BYTES_PER_PX_STEP * npx / (ms_per_step * 1e-3) / 1e9 / max(world, 1),
                "step_frac": ALG_BYTES_PER_PX_STEP * npx / (ms_per_step * 1e-3) / 1e9 / max(world, 1) / peak}

    # ---- end to end through the host entry point (pinned host buffers, copies inside)
    e2e = None
    if not args.no_e2e:
        Ke = max(2, min(K, 5))
        e2e_extra = {}
        if not multi:
            # (1) one image per call: upload chunks -> row pass -> column blocks -> download blocks
            h_out = torch.empty_like(h_img).pin_memory()
            plan.run_host(h_img.numpy(), h_out.numpy())
            barrier()
            t0 = time.perf_counter()
            for _ in range(Ke):
                plan.run_host(h_img.numpy(), h_out.numpy())
            barrier()
            dt_single = (time.perf_counter() - t0) / Ke
            # (2) the headline: Ks images streamed through blur_b200_gauss_run_host_batch, every image
            # uploaded from and downloaded to pinned host memory inside the timed region; upload of
            # image i+1 / kernels of image i / download of image i-1 overlap (PCIe is full duplex)
            Ks = max(4, min(K, 16))
            h_in2 = torch.from_numpy(np.ascontiguousarray(h_img.numpy()[::-1])).pin_memory()      # a second, different image
            h_out2 = torch.empty_like(h_img).pin_memory()
            ins = [(h_img if i % 2 == 0 else h_in2).numpy() for i in range(Ks)]
            outs = [(h_out if i % 2 == 0 else h_out2).numpy() for i in range(Ks)]
            plan.run_host_batch(ins[:2], outs[:2])
            barrier()
            t0 = time.perf_counter()
            plan.run_host_batch(ins, outs)
            barrier()
            dt = (time.perf_counter() - t0) / Ks
            Ke = Ks
            h2d = d2h = rows * cols * 3
            e2e_extra = {"single_call_ms": dt_single * 1e3, "single_call_value": npx / 1e6 / dt_single,
                         "api": "blur_b200_gauss_run_host_batch (streamed: %d images, 2 distinct pinned inputs)" % Ks}
        else:
            g = sb.geom
            # every rank streams its row stripe in and its column stripe out through
            # ShardedGaussBlur.run_host_stream (upload of image i+1 under kernels / download of image i)
            Ks = max(4, min(K, 16))
            h_out = torch.empty((rows, 2 * g.cpp, 3), dtype=torch.uint8).pin_memory()
            h_out2 = torch.empty_like(h_out).pin_memory()
            h_in2 = torch.flip(h_img, dims=[0]).contiguous().pin_memory()
            ins = [h_img if i % 2 == 0 else h_in2 for i in range(Ks)]
            outs = [h_out if i % 2 == 0 else h_out2 for i in range(Ks)]
            sb.run_host_stream(ins[:2], outs[:2]); barrier()
            t0 = time.perf_counter()
            sb.run_host_stream(ins, outs)
            barrier()
            dt = (time.perf_counter() - t0) / Ks
            Ke = Ks
            e2e_extra = {"api": "ShardedGaussBlur.run_host_stream (%d images, per-rank stripes from pinned host memory)" % Ks}
            tt = torch.tensor([dt], device=dev)
            dist.all_reduce(tt, op=dist.ReduceOp.MAX)
            dt = float(tt.item())
            h2d, d2h = int(h_img.numel()), int(h_out.numel())
        e2e = {"value": npx / 1e6 / dt, "unit": "megapixels/sec", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
               "ms_per_step": dt * 1e3, "steps": Ke, "timing": "wall clock around synchronous C-ABI host calls"}
        e2e.update(e2e_extra)

    cpu = None
    if rank == 0 and not multi and not args.no_cpu_baseline:
        cpu, _ = cpu_reference_run(2, 1)

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": "megapixels/sec", "n_gpus": args.gpus, "steps": K, "warmup": W,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": config, "roofline": roofline, "clocks": clocks,
                "e2e": e2e, "gpu_launches": int(launches), "impl": "b200",
                "plan": {k: plan.info[k] for k in ("algo", "n_row", "n_col", "ols_L_row", "ols_L_col", "ols_V_row", "ols_V_col", "ols_nblocks_row", "ols_nblocks_col")}}
        if cpu is not None:
            line["cpu_baseline"] = cpu
        print(json.dumps(line))
    if multi:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
