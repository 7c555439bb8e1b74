#!/usr/bin/env python
"""bench.py — the reference's hot paths on B200, one JSON line per run.

  python bench.py --gpus N --steps K --warmup W [--workload W]      (N > 1: launched by torchrun)
  python bench.py --impl reference --gpus N --steps K --warmup W [--workload W]

Workloads (BASELINE.json configs):
  gauss_c3 (default)  16384 x 16384 RGB, sigma = 50 — the configuration the metric is quoted on.  A step is one
                      whole blur: row pass -> (N > 1: exchange of the float32 intermediate) -> column pass ->
                      interleave.  At N > 1 the ONE image is sharded by row / column stripes: strong scaling.
  gauss_c2, gauss_c1  3840 x 2160 sigma = 25 and 1920 x 1080 sigma = 5 (single GPU; L2 flushed between steps).
  box_c5              FastBoxBlur, 256 images 1920 x 1080 RGB, ksize 31, 2 passes; a step is the whole batch,
                      images dealt round-robin to the ranks (image i -> rank i mod N), no collective.

`value`    : device-resident throughput (inputs already in HBM), CUDA events on the launching stream, max over ranks.
`e2e`      : the same metric through the C-ABI host entry point with pinned HOST buffers: the H2D copy of every
             input and the D2H copy of every result are inside the timed region.
`roofline` : dominant kernel: algorithmic bytes per launch / its CUDA-event duration, against the measured HBM
             copy peak (MEASURED_PEAKS.json).  `traffic` = ncu dram bytes of that kernel (profiles/traffic.json),
             only where a capture of that launch shape exists.
`cpu_baseline` (rank 0, N = 1): the reference's own pipeline (Source.cpp compiled unmodified, OpenMP, all host
             cores; the FFT / box engines underneath are this repo's scalar C stand-ins because the reference's
             pffft / pocketfft / FastBoxBlur submodules are absent) on the same workload at full size.
`--impl reference`: that CPU pipeline as the timed arm, full-size steps, thread count forced to all host cores.
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

WORKLOADS = {
    # name: kind, rows, cols, sigma / ksize, batch
    "gauss_c3": dict(kind="gauss", rows=16384, cols=16384, sigma=50.0),
    "gauss_c2": dict(kind="gauss", rows=2160, cols=3840, sigma=25.0),
    "gauss_c1": dict(kind="gauss", rows=1080, cols=1920, sigma=5.0),
    "box_c5": dict(kind="box", rows=1080, cols=1920, ksize=31, passes=2, batch=256),
}
GAUSS_BYTES_PER_PX_STEP = 30.0    # 3 (u8 in) + 12 (f32 out) + 12 (f32 in) + 3 (u8 out), SURVEY.md §8d
GAUSS_BYTES_PER_PX_PASS = 15.0    # each FFT pass: 3 + 12
BOX_BYTES_PER_PX = 6.0            # uint8 in + out; the passes in between fit in L2
METRICS = {"gauss": "megapixels/sec RGB Gaussian blur (sigma=50)", "box": "megapixels/sec FastBoxBlur (ksize 31, 2 passes)"}


def hbm_peak():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except Exception:
        return os.cpu_count() or 1


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


# ------------------------------------------------------------------------------------------ CPU reference

def cpu_reference_run(wl: dict, steps: int, warmup: int, nyquist: str = "exact"):
    """Times the reference's own pipeline on the host cores: oracle/_ref (Source.cpp + Utils.hpp compiled
    unmodified) when it was built, else the C restatement.  FULL-SIZE steps of the workload (box: a bounded
    number of the batch's images); the OpenMP thread count is set here to every host core (torchrun exports
    OMP_NUM_THREADS=1).  Returns (cpu_baseline dict, seconds per step, megapixels per step)."""
    import numpy as np
    from oracle_lib import Oracle, Reference
    cores = host_cores()
    rng = np.random.default_rng(0)
    rows, cols = wl["rows"], wl["cols"]
    use_ref = Reference.available()
    impl = Reference() if use_ref else Oracle()
    impl.set_num_threads(cores)
    threads = impl.num_threads()
    assert threads == cores or cores == 1, (threads, cores)
    engine = "FFT engine = this repo's scalar C mixed-radix FFT (oracle/cpu_fft.c) standing in for the absent pffft / pocketfft"
    if wl["kind"] == "gauss":
        img = rng.integers(0, 256, (rows, cols, 3), dtype=np.uint8)
        # the same function as the GPU arm: exact multiply = pocketfft_1D (Test flag 5, Source.cpp:580-581),
        # pffft's Nyquist quirk = pffft_ (Test flag 3, :589-590)
        flag, fname, mode = (5, "pocketfft_1D", 0) if nyquist == "exact" else (3, "pffft_", 1)
        if use_ref:
            fn = lambda: impl.test(img, flag, wl["sigma"])              # noqa: E731
            what = f"Source.cpp Test(flag={flag}) -> {fname}() compiled unmodified (OpenMP) against API shims; " + engine
        else:
            fn = lambda: impl.gauss_blur_1d(img, wl["sigma"], mode)     # noqa: E731
            what = f"C restatement of {fname}() (oracle/blur_oracle.c, OpenMP); " + engine
        mpx = rows * cols / 1e6
        sample = f"{rows}x{cols} RGB, sigma={wl['sigma']:g}: the whole workload image per step"
    else:
        nimg = min(wl["batch"], 16)                                      # bounded sample of the batch
        imgs = [rng.integers(0, 256, (rows, cols, 3), dtype=np.uint8) for _ in range(nimg)]
        ns = float(np.sqrt(wl["ksize"] + 0.5))                           # Test(flag=4): ksize = int(nsmooth * nsmooth)
        assert int(ns * ns) == wl["ksize"]
        if use_ref:
            fn = lambda: [impl.test(im, 4, ns) for im in imgs]           # noqa: E731
            what = "Source.cpp Test(flag=4) -> fastboxblur() against the FastBoxBlur API shim (declared spec, SURVEY A.7; upstream absent: parity unpinned)"
        else:
            fn = lambda: [impl.fastboxblur(im, wl["ksize"], wl["passes"]) for im in imgs]   # noqa: E731
            what = "C restatement of the declared FastBoxBlur spec (oracle/blur_oracle.c)"
        mpx = nimg * rows * cols / 1e6
        sample = f"{nimg} of the batch's {wl['batch']} images ({rows}x{cols} RGB, ksize {wl['ksize']}, {wl['passes']} passes) per step"
    devnull = os.open(os.devnull, os.O_WRONLY)
    saved = os.dup(1)
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
    base = {"value": mpx / dt, "unit": "megapixels/sec", "cores": threads, "kind": "reference" if use_ref else "port",
            "sample": f"{sample}; {steps} timed steps after {warmup} warm-up; {what}"}
    return base, dt, mpx


# ------------------------------------------------------------------------------------------ helpers

def sample_exact_gauss(img, sigma, pts):
    """float64 direct separable convolution (reflect-101, the reference's float taps) at sample pixels."""
    import numpy as np
    import blur_algorithms_b200 as bb
    rows, cols = img.shape[:2]
    k = bb.gaussian_window(sigma, max(rows, cols))
    pad = (k - 1) // 2
    g = bb.getGaussian(sigma, k, 0).astype(np.float64)
    idx = np.arange(-pad, pad + 1)

    def refl(i, L):
        i = np.abs(i)
        return np.where(i >= L, 2 * (L - 1) - i, i)
    out = []
    for (r, c) in pts:
        patch = img[np.ix_(refl(r + idx, rows), refl(c + idx, cols))].astype(np.float64)
        out.append(np.einsum("i,j,ijc->c", g, g, patch))
    return np.array(out)


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="gauss_c3", choices=sorted(WORKLOADS))
    ap.add_argument("--rows", type=int, default=0)
    ap.add_argument("--cols", type=int, default=0)
    ap.add_argument("--sigma", type=float, default=0.0)
    ap.add_argument("--nyquist", default="exact", choices=["exact", "pffft"],
                    help="gauss workloads: exact multiply (pocketfft_1D, the parity default) or pffft_'s Nyquist quirk")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-verify", action="store_true")
    ap.add_argument("--no-halo", action="store_true", help="N > 1 gauss: skip the row-stripe (halo) sharding line printed beside the contract one")
    ap.add_argument("--sharding", default="contract", choices=["contract", "halo"],
                    help="N > 1 gauss: which sharding `value` reports; contract = row stripes -> all-to-all -> column stripes "
                         "(BASELINE.json), halo = row stripes in, row stripes out, 2*halo rows exchanged (the other one is printed beside it)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    wl = dict(WORKLOADS[args.workload])
    if args.rows: wl["rows"] = args.rows
    if args.cols: wl["cols"] = args.cols
    if args.sigma and wl["kind"] == "gauss": wl["sigma"] = args.sigma
    rows, cols = wl["rows"], wl["cols"]
    if wl["kind"] == "gauss":
        config = {"workload": f"gauss_blur_{rows}x{cols}_rgb_u8_sigma{wl['sigma']:g}", "rows": rows, "cols": cols,
                  "sigma": wl["sigma"], "nyquist_mode": args.nyquist,
                  "sharding": "single GPU" if args.gpus == 1 else
                              (f"one image: row stripes x{args.gpus} in and out, halo rows exchanged" if args.sharding == "halo" else
                               f"one image: row stripes x{args.gpus} -> exchange -> column stripes")}
    else:
        config = {"workload": f"fastboxblur_{wl['batch']}x{rows}x{cols}_rgb_u8_ksize{wl['ksize']}_passes{wl['passes']}",
                  "rows": rows, "cols": cols, "ksize": wl["ksize"], "passes": wl["passes"], "batch": wl["batch"],
                  "sharding": "single GPU" if args.gpus == 1 else f"image i -> rank i mod {args.gpus}, no collective",
                  "parity": "bit-exact against the declared spec (SURVEY A.7); upstream FastBoxBlur absent: parity unpinned"}
    # `config` names the workload and is the same dict in both arms (the driver compares them); what only one arm
    # can say (exchange variant, CPU engine, threads) goes into the line's "run" object
    if wl["kind"] == "gauss":
        small = rows * cols * 15 < 200e6                 # working set near the 126 MB L2: flush between steps
        config["l2_policy"] = ("L2 flushed (256 MB write) before every timed step" if small else
                               "inputs larger than L2 (%d MB image, %.1f GB float32 intermediate)" % (rows * cols * 3 // 1000000, rows * cols * 12 / 1e9))
    else:
        per_rank = -(-wl["batch"] // max(args.gpus, 1)) * rows * cols * 3
        config["l2_policy"] = "per-rank batch (%d MB) larger than L2" % (per_rank >> 20)
    run_info = {}
    metric = METRICS[wl["kind"]] if wl["kind"] == "box" or wl.get("sigma") == 50.0 else f"megapixels/sec RGB Gaussian blur (sigma={wl['sigma']:g})"
    scaling = "strong"          # gauss: one image cut over the ranks; box: one fixed batch dealt over the ranks

    if args.impl == "reference":
        if rank != 0:
            return
        K, W = max(1, args.steps), max(0, args.warmup)
        base, dt, mpx = cpu_reference_run(wl, K, W, args.nyquist)
        run_info["engine"] = "scalar C stand-in for pffft / pocketfft / FastBoxBlur (submodules absent)"
        run_info["sample"] = base["sample"].split(";")[0]
        run_info["threads"] = base["cores"]
        line = {"metric": metric, "value": base["value"], "unit": "megapixels/sec", "n_gpus": args.gpus, "steps": K,
                "warmup": W, "ms_per_step": dt * 1e3, "higher_is_better": True, "scaling": scaling,
                "vs_baseline": None, "dtype": "f32" if wl["kind"] == "gauss" else "u8", "data": "synthetic", "config": config,
                "impl": "reference", "run": run_info, "cpu_baseline": base,
                "e2e": {"value": base["value"], "unit": "megapixels/sec", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                "gpu_launches": 0}
        print(json.dumps(line))
        return

    import numpy as np
    import torch
    import torch.distributed as dist

    import blur_algorithms_b200 as bb

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

    def allmax(x: float) -> float:
        if not multi:
            return x
        t = torch.tensor([x], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def allmin_int(x: int) -> int:
        if not multi:
            return x
        t = torch.tensor([x], device=dev, dtype=torch.int64)
        dist.all_reduce(t, op=dist.ReduceOp.MIN)
        return int(t.item())

    stream = torch.cuda.current_stream()
    K, W = max(1, args.steps), max(3, args.warmup)
    peak, peak_src = hbm_peak()
    flush = None
    line_extra = {}

    # ================================================================================== Gaussian
    if wl["kind"] == "gauss":
        from blur_algorithms_b200.dist import ShardedGaussBlur
        sigma = wl["sigma"]
        nyq = bb.NYQUIST_PFFFT if args.nyquist == "pffft" else bb.NYQUIST_EXACT
        lsb_tol = 1 if args.nyquist == "exact" else 16    # pffft_'s Nyquist term moves uniform noise by several LSB off the exact convolution
        npx = rows * cols
        small = config["l2_policy"].startswith("L2 flushed")
        if small:
            flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
        full_img = None
        if not multi:
            full_img = np.random.default_rng(0).integers(0, 256, (rows, cols, 3), dtype=np.uint8)
            h_img = torch.from_numpy(full_img).pin_memory()
            d_in = h_img.to(dev)
            d_out = torch.empty_like(d_in)
            plan = bb.GaussPlan(rows, cols, sigma, nyq, device=local)
            step = lambda: plan.run(d_in, d_out)            # noqa: E731
        else:
            fused_env = os.environ.get("BLUR_B200_FUSED")        # "0": NCCL all_to_all exchange instead of peer memory
            sb = ShardedGaussBlur(rows, cols, sigma, nyq, device=dev,
                                  fused=None if fused_env is None else fused_env != "0")
            g = sb.geom
            full_img = np.random.default_rng(0).integers(0, 256, (rows, cols, 3), dtype=np.uint8)   # every rank: same image
            h_img = torch.from_numpy(np.ascontiguousarray(full_img[g.row0:g.row0 + g.valid_rows])).pin_memory()
            d_in = h_img.to(dev)
            plan = sb.plan
            step = lambda: sb.run(d_in)                      # noqa: E731
            # row-stripe (halo) sharding beside it: input and output row-sharded, only 2*halo rows of the float32
            # intermediate cross NVLink (dist.HaloShardedGaussBlur); exact mode, two-factor kernels
            hb = d_in_h = None
            if args.nyquist == "exact" and not args.no_halo:
                from blur_algorithms_b200.dist import HaloShardedGaussBlur
                try:
                    hb = HaloShardedGaussBlur(rows, cols, sigma, device=dev)
                    d_in_h = torch.from_numpy(np.ascontiguousarray(full_img[hb.row0:hb.row0 + hb.valid_rows])).to(dev)
                except ValueError as ex:
                    hb = None
                    line_extra["halo"] = {"unavailable": str(ex)}
            halo_ok = None

        # ---- verification, outside every timed region
        verified = None
        if not args.no_verify:
            if not multi:
                step(); torch.cuda.synchronize()
                out = d_out.cpu().numpy()
                rng = np.random.default_rng(1)
                pts = [(0, 0), (0, cols - 1), (rows - 1, 0), (rows - 1, cols - 1), (rows // 2, cols // 2)] + \
                      [(int(rng.integers(rows)), int(rng.integers(cols))) for _ in range(27)]
                exact = sample_exact_gauss(full_img, sigma, pts)
                got = np.array([out[r, c] for r, c in pts]).astype(np.float64)
                dmax = float(np.abs(got - np.floor(exact + 0.5)).max())
                verified = {"ok": bool(dmax <= lsb_tol), "check": "32 pixels against a float64 direct convolution (+-%d LSB)" % lsb_tol,
                            "max_abs_diff_lsb": dmax}
                del out
            else:
                # this rank's column stripe of the sharded result == the single-GPU plan's, bit for bit
                p1 = bb.GaussPlan(rows, cols, sigma, nyq, device=local, direct=False)   # FFT kernels like the sharded plans
                d_full = torch.from_numpy(full_img).to(dev)
                p1.run(d_full)
                mine = sb.run(d_in)
                torch.cuda.synchronize()
                same = bool(torch.equal(mine, d_full[:, g.col0:g.col0 + g.valid_cols])) if g.valid_cols > 0 else True
                rng = np.random.default_rng(1)
                pts = [(int(rng.integers(rows)), g.col0 + int(rng.integers(max(g.valid_cols, 1)))) for _ in range(8)] if g.valid_cols > 0 else []
                dmax = 0.0
                if pts:
                    exact = sample_exact_gauss(full_img, sigma, pts)
                    got = np.array([mine[r, c - g.col0].cpu().numpy() for r, c in pts]).astype(np.float64)
                    dmax = float(np.abs(got - np.floor(exact + 0.5)).max())
                ok = allmin_int(int(same and dmax <= lsb_tol))
                verified = {"ok": bool(ok), "check": "every rank: its column stripe bit-identical to the single-GPU plan's output, "
                            "plus 8 pixels per rank against a float64 direct convolution (+-1 LSB)", "max_abs_diff_lsb": allmax(dmax)}
                if hb is not None:
                    got = hb.run(d_in_h)
                    torch.cuda.synchronize()
                    halo_ok = bool(allmin_int(int(torch.equal(got, d_full[hb.row0:hb.row0 + hb.valid_rows]))))
                p1.close()
                del d_full
        del full_img

        # ---- device-resident timing
        for _ in range(W):
            step()
        if multi:
            # the exchange variant is chosen by a timed trial on the first blur (dist.ShardedGaussBlur.tune_exchange)
            run_info["exchange"] = sb.describe_exchange()
            run_info["exchange_mode"] = sb.exchange_mode if sb.fused else "nccl"
            run_info["exchange_trial_ms"] = sb.exchange_trial
        barrier()
        sampler = ClockSampler(local)
        sampler.start()
        n0 = bb.kernel_launches()
        if flush is None:
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            ev0.record(stream)
            for _ in range(K):
                step()
            ev1.record(stream)
            barrier()
            ms = ev0.elapsed_time(ev1)
        else:
            ms = 0.0
            for _ in range(K):
                flush.fill_(1)
                ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                barrier()
                ev0.record(stream)
                step()
                ev1.record(stream)
                barrier()
                ms += ev0.elapsed_time(ev1)
        launches = bb.kernel_launches() - n0
        clocks = sampler.stop()
        ms_per_step = allmax(ms) / K
        value = npx / 1e6 / (ms_per_step * 1e-3)
        if multi and hb is not None:
            for _ in range(W):
                hb.run(d_in_h)
            barrier()
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            ev0.record(stream)
            for _ in range(K):
                hb.run(d_in_h)
            ev1.record(stream)
            barrier()
            hms = allmax(ev0.elapsed_time(ev1)) / K
            line_extra["halo"] = {
                "ms_per_step": hms, "value": npx / 1e6 / (hms * 1e-3), "unit": "megapixels/sec",
                "sharding": "row stripes of whole column-pass blocks in AND out; every rank pushes 2*halo rows of the float32 "
                            "intermediate to its two neighbours (copy engines over NVLink), no all-to-all",
                "halo_rows": hb.info["halo"], "exchanged_bytes_per_gpu": 2 * hb.info["halo"] * cols * 12,
                "verified": None if halo_ok is None else {"ok": halo_ok, "check": "every rank: its row stripe bit-identical to the single-GPU plan's output"}}
            line_extra["contract"] = {"ms_per_step": ms_per_step, "value": value, "unit": "megapixels/sec"}
            if args.sharding == "halo":
                ms_per_step, value = hms, line_extra["halo"]["value"]

        # ---- per-pass timing for the roofline (same stream, CUDA events)
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        reps = max(3, min(K, 10))

        def time_passes(rp, cp):
            barrier()
            rp(); cp()
            barrier()
            if flush is None:
                e[0].record(stream)
                for _ in range(reps):
                    rp()
                e[1].record(stream)
                for _ in range(reps):
                    cp()
                e[2].record(stream)
                torch.cuda.synchronize()
                return e[0].elapsed_time(e[1]) / reps, e[1].elapsed_time(e[2]) / reps
            tr = tc = 0.0
            for _ in range(reps):
                flush.fill_(1)
                e[0].record(stream); rp(); e[1].record(stream)
                torch.cuda.synchronize()
                tr += e[0].elapsed_time(e[1])
                # the column pass reads what the row pass has just written: no flush in between (as in a whole step)
                e[1].record(stream); cp(); e[2].record(stream)
                torch.cuda.synchronize()
                tc += e[1].elapsed_time(e[2])
            return tr / reps, tc / reps

        direct = (not multi) and bool(plan.info.get("direct"))
        if direct:
            # narrow kernel: the whole blur is ONE fused direct-convolution kernel (csrc/direct_conv.cu); there are no passes to time
            row_ms = colil_ms = None
            pass_px = npx
        elif not multi:
            q = torch.empty(plan.info["block_floats"], dtype=torch.float32, device=dev)
            oc = torch.empty((rows, 2 * plan.info["cpp"], 3), dtype=torch.uint8, device=dev)
            row_ms, colil_ms = time_passes(lambda: plan.row_pass(d_in, q), lambda: plan.col_pass(q, oc))
            pass_px = npx
            del q, oc
        else:
            row_ms, colil_ms = time_passes(sb.row_pass_only, sb.col_pass_only)
            pass_px = sb.geom.valid_rows * cols
        algo = plan.info["algo"]
        two = plan.info["ols_L_row"] == 1024 and algo == 1
        kn = (("ols2_row_kernel" if two else "ols_row_kernel"),
              ("ols2_col_kernel" if plan.info["ols_L_col"] == 1024 else "ols_col_kernel") + " (+interleave_oct_kernel)") if algo == 1 else \
             ("fft_conv_kernel<RowIO>", "fft_conv_kernel<ColIO> (+interleave_u8_kernel)")
        if direct:
            dom_name, dom_ms = "direct_gauss_kernel (row + column convolution fused, float32 intermediate in shared memory)", ms_per_step
            achieved = GAUSS_BYTES_PER_PX_STEP * pass_px / (dom_ms * 1e-3) / 1e9      # the contract's 30 B/px for the whole blur
        else:
            dom_name, dom_ms = (kn[0], row_ms) if row_ms >= colil_ms else (kn[1], colil_ms)
            achieved = GAUSS_BYTES_PER_PX_PASS * pass_px / (dom_ms * 1e-3) / 1e9
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp) and not multi:
            try:
                tj = json.load(open(tp))
                traffic = tj.get(args.workload, {}).get(dom_name.split(" ")[0])
            except Exception:
                traffic = None
        step_gbps = GAUSS_BYTES_PER_PX_STEP * npx / (ms_per_step * 1e-3) / 1e9 / max(world, 1)
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                    "traffic": traffic, "kernel": dom_name, "kernel_ms": dom_ms, "row_pass_ms": row_ms,
                    "col_pass_plus_interleave_ms": colil_ms, "peak_source": peak_src,
                    "algorithmic_bytes_per_launch": (GAUSS_BYTES_PER_PX_STEP if direct else GAUSS_BYTES_PER_PX_PASS) * pass_px,
                    "step_achieved_GBps": step_gbps, "step_frac": step_gbps / peak}
        if direct:
            roofline["note"] = ("algorithmic bytes = SURVEY 8d's 30 B/px of the two-pass pipeline; the fused kernel itself moves only the "
                                "compulsory 6 B/px (uint8 in + out) through HBM")
            roofline["compulsory_bytes_per_launch"] = 6.0 * pass_px

        # ---- end to end through the host entry point (pinned host buffers, copies inside)
        e2e = None
        if not args.no_e2e:
            if not multi:
                h_out = torch.empty_like(h_img).pin_memory()
                plan.run_host(h_img.numpy(), h_out.numpy())
                barrier()
                Ke = max(2, min(K, 5))
                t0 = time.perf_counter()
                for _ in range(Ke):
                    plan.run_host(h_img.numpy(), h_out.numpy())
                barrier()
                dt_single = (time.perf_counter() - t0) / Ke
                # the headline: Ks images streamed through blur_b200_gauss_run_host_batch, every image uploaded
                # from and downloaded to pinned host memory inside the timed region (upload of image i+1 /
                # kernels of image i / download of image i-1 overlap: PCIe is full duplex)
                Ks = max(4, min(K, 16))
                h_in2 = torch.from_numpy(np.ascontiguousarray(h_img.numpy()[::-1])).pin_memory()
                h_out2 = torch.empty_like(h_img).pin_memory()
                ins = [(h_img if i % 2 == 0 else h_in2).numpy() for i in range(Ks)]
                outs = [(h_out if i % 2 == 0 else h_out2).numpy() for i in range(Ks)]
                plan.run_host_batch(ins[:2], outs[:2])
                barrier()
                t0 = time.perf_counter()
                plan.run_host_batch(ins, outs)
                barrier()
                dt = (time.perf_counter() - t0) / Ks
                h2d = d2h = rows * cols * 3
                e2e = {"single_call_ms": dt_single * 1e3, "single_call_value": npx / 1e6 / dt_single,
                       "api": "blur_b200_gauss_run_host_batch (streamed: %d images, 2 distinct pinned inputs)" % Ks, "steps": Ks}
            else:
                g = sb.geom
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
                dt = allmax((time.perf_counter() - t0) / Ks)
                h2d, d2h = int(h_img.numel()), int(h_out.numel())
                e2e = {"api": "ShardedGaussBlur.run_host_stream (%d images, per-rank stripes from pinned host memory)" % Ks,
                       "steps": Ks, "bytes_are": "per rank"}
            e2e.update({"value": npx / 1e6 / dt, "unit": "megapixels/sec", "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                        "ms_per_step": dt * 1e3, "timing": "wall clock around synchronous C-ABI host calls"})
        dtype = "f32"
        line_extra["plan"] = {k: plan.info[k] for k in ("algo", "n_row", "n_col", "ols_L_row", "ols_L_col", "ols_V_row", "ols_V_col",
                                                         "ols_nblocks_row", "ols_nblocks_col")}

    # ================================================================================== FastBoxBlur batch
    else:
        from blur_algorithms_b200.dist import ShardedBatch
        batch, ksize, passes = wl["batch"], wl["ksize"], wl["passes"]
        sbat = ShardedBatch(batch, rank, world)
        mine = sbat.indices                               # images this rank owns (i mod world == rank)
        rng_imgs = [np.random.default_rng(i).integers(0, 256, (rows, cols, 3), dtype=np.uint8) for i in mine]
        h_imgs = torch.from_numpy(np.stack(rng_imgs)).pin_memory() if mine else torch.empty((0, rows, cols, 3), dtype=torch.uint8)
        d_src = h_imgs.to(dev)
        d_img = torch.empty_like(d_src)
        d_scr = torch.empty_like(d_src)
        npx = batch * rows * cols

        def step():
            d_img.copy_(d_src)                             # the blur is in place: restore the input (device copy, inside the step)
            if mine:
                bb.fastboxblur_device(d_img, cols, rows, 3, ksize, passes, len(mine), d_scr)

        def kernels_only():
            if mine:
                bb.fastboxblur_device(d_img, cols, rows, 3, ksize, passes, len(mine), d_scr)

        verified = None
        if not args.no_verify:
            from util import box_numpy
            step(); torch.cuda.synchronize()
            ok = 1
            for j in sorted({0, len(mine) - 1} if mine else set()):
                ok &= int(np.array_equal(d_img[j].cpu().numpy(), box_numpy(rng_imgs[j], ksize, passes)))
            verified = {"ok": bool(allmin_int(ok)), "check": "first and last image of every rank bit-exact against a numpy restatement "
                        "of the declared spec (SURVEY A.7)"}
        del rng_imgs
        for _ in range(W):
            kernels_only()
        barrier()
        sampler = ClockSampler(local)
        sampler.start()
        n0 = bb.kernel_launches()
        ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        ev0.record(stream)
        for _ in range(K):
            kernels_only()                                 # blurring the previous result again: same work per step
        ev1.record(stream)
        barrier()
        launches = bb.kernel_launches() - n0
        clocks = sampler.stop()
        ms_per_step = allmax(ev0.elapsed_time(ev1)) / K
        value = npx / 1e6 / (ms_per_step * 1e-3)
        achieved = BOX_BYTES_PER_PX * (len(mine) * rows * cols) / (ms_per_step * 1e-3) / 1e9
        roofline = {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": None,
                    "kernel": "boxblur kernels of one batch (%d fused 2-D passes)" % passes, "kernel_ms": ms_per_step,
                    "peak_source": peak_src, "algorithmic_bytes_per_launch": BOX_BYTES_PER_PX * len(mine) * rows * cols,
                    "note": "achieved = this rank's share (max over ranks of the time)"}
        e2e = None
        if not args.no_e2e:
            h_work = h_imgs.clone().pin_memory()
            if mine:
                bb.fastboxblur_batch(h_work[: min(4, len(mine))], ksize, passes)
            barrier()
            t0 = time.perf_counter()
            if mine:
                bb.fastboxblur_batch(h_work, ksize, passes)
            barrier()
            dt = allmax(time.perf_counter() - t0)
            e2e = {"value": npx / 1e6 / dt, "unit": "megapixels/sec", "h2d_bytes_per_step": int(h_work.numel()),
                   "d2h_bytes_per_step": int(h_work.numel()), "ms_per_step": dt * 1e3, "steps": 1, "bytes_are": "per rank",
                   "api": "blur_b200_fastboxblur_batch_host (grouped three-stream pipeline, pinned host batch in place)",
                   "timing": "wall clock around the synchronous C-ABI host call"}
        dtype = "u8"

    cpu = None
    if rank == 0 and not multi and not args.no_cpu_baseline:
        cpu, _, _ = cpu_reference_run(wl, 2, 1, args.nyquist)

    if rank == 0:
        line = {"metric": metric, "value": value, "unit": "megapixels/sec", "n_gpus": args.gpus, "steps": K, "warmup": W,
                "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": scaling, "vs_baseline": None,
                "dtype": dtype, "data": "synthetic", "config": config, "roofline": roofline, "clocks": clocks,
                "e2e": e2e, "gpu_launches": int(launches), "impl": "b200", "run": run_info, "verified": verified}
        line.update(line_extra)
        if cpu is not None:
            line["cpu_baseline"] = cpu
        print(json.dumps(line))
    if multi:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
