"""CPU: the C-ABI library loads and exports what include/blur_b200.h declares; host-side logic
(sizing, taps, planner) matches the oracle; the device code run by the CPU emulator matches the
oracle; the sharded host logic works over gloo with world_size 2.  No compute call needs a GPU."""
import os
import re
import subprocess
import sys

import numpy as np
import pytest

from util import Emulator, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def emu():
    return Emulator()


def test_cabi_exports_every_declared_symbol():
    import ctypes
    hdr = open(os.path.join(ROOT, "include", "blur_b200.h")).read()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    declared = sorted(set(re.findall(r"\b(blur_b200_[a-z0-9_]+)\s*\(", hdr)))
    assert len(declared) >= 25
    lib = ctypes.CDLL(os.path.join(ROOT, "blur_algorithms_b200", "libblur_b200.so"))
    missing = [n for n in declared if not hasattr(lib, n)]
    assert not missing, missing
    from blur_algorithms_b200 import _lib
    assert sorted(_lib.EXPORTED) == declared


def test_sizing_helpers_match_oracle(oracle):
    import blur_algorithms_b200 as bb
    for n in range(0, 40000, 3):
        assert bb.isValidSize(n) == oracle.is_valid_size(n)
    for n in [1, 31, 32, 33, 100, 1000, 1110, 1950, 2324, 4004, 4126, 16714, 20000, 33098]:
        assert bb.nearestTransformSize(n) == oracle.nearest_transform_size(n)
    for s in [0.3, 0.5, 1, 2, 3, 5, 10, 25, 38.729833, 50, 100, 106.770783]:
        assert bb.gaussian_window(s) == oracle.gaussian_window(s)
        for m in (100, 330, 331, 332):
            assert bb.gaussian_window(s, m) == oracle.gaussian_window(s, m)
    for s, w, N in [(1, 5, 0), (2, 11, 32), (5, 31, 64), (50, 331, 17280), (25, 165, 2400), (7.5, 0, 0)]:
        assert np.array_equal(bb.getGaussian(s, w, N), oracle.get_gaussian(s, w, N))
    for rows, cols, s in [(1080, 1920, 5), (2160, 3840, 25), (16384, 16384, 50), (467, 700, 7.5)]:
        g, o = bb.gauss_geometry(rows, cols, s), oracle.gauss_sizes(rows, cols, s)
        assert (g["kSize"], g["pad"], g["n_col"], g["n_row"]) == (o["kSize"], o["pad"], o["n_col"], o["n_row"])
    # SURVEY.md §8 table
    assert bb.gauss_geometry(16384, 16384, 50) == dict(kSize=331, pad=165, n_col=17280, n_row=17280)
    assert bb.gauss_geometry(2160, 3840, 25) == dict(kSize=165, pad=82, n_col=2400, n_row=4096)


def test_compute_entry_points_fail_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    import blur_algorithms_b200 as bb
    img = synth("noise", 64, 96)
    with pytest.raises(bb.BlurB200Error):
        bb.pffft_(img, 2.0)
    with pytest.raises(bb.BlurB200Error):
        bb.fastboxblur(img, 96, 64, 3, 9, 2)


def test_argument_validation():
    import blur_algorithms_b200 as bb
    with pytest.raises(bb.BlurB200Error):
        bb.pffft_(np.zeros((4, 4), np.uint8), 2.0)
    with pytest.raises(bb.BlurB200Error):
        bb.pocketfft_1D(np.zeros((4, 4, 3), np.float32), 2.0)
    # flag 1 = the reference's OpenCV comparison line (Source.cpp:596-603): host cv2 when importable, an error otherwise
    try:
        import cv2  # noqa: F401
        img = np.full((8, 8, 3), 7, np.uint8)
        bb.Test(img, 1, 2.0)
        assert (img == 7).all()
    except ImportError:
        with pytest.raises(bb.BlurB200Error):
            bb.Test(np.zeros((8, 8, 3), np.uint8), 1, 2.0)


def test_every_valid_fft_length_plans_and_transforms(emu, oracle):
    rng = np.random.default_rng(1)
    sizes = [n for n in range(32, 28161, 32) if oracle.is_valid_size(n)]
    assert len(sizes) == 82
    for N in sizes:
        a = rng.standard_normal(N).astype(np.float32)
        b = rng.standard_normal(N).astype(np.float32)
        code, outr, outi, spec = emu.fft_roundtrip(a, b)
        assert code == 0, N
        ref = np.fft.fft(a.astype(np.float64) + 1j * b.astype(np.float64))
        assert np.abs(spec - ref).max() / np.abs(ref).max() < 3e-6, N
        assert max(np.abs(outr - a).max(), np.abs(outi - b).max()) < 3e-5, N


def test_boxfft_device_code_on_cpu_matches_oracle(emu, oracle):
    """FFT box blur (the reference's `#define boxblur` build, Source.cpp:12,128-140): the device code runs the tent
    taps through the same overlap-save kernels (kernel kind = negative sigma, csrc/fft_plan.hpp)."""
    import blur_algorithms_b200 as bb
    for rows, cols, ns in [(120, 160, 3.0), (300, 200, 5.0), (200, 333, 7.0), (50, 400, 9.0)]:
        img = synth("noise", rows, cols, 3)
        klen = bb.boxfft_klen(rows, cols, ns)
        m = min(int(ns) * int(ns), min(rows - 1, cols - 1))
        assert klen == int(np.sqrt(m) * np.sqrt(m)) and klen % 2 == 1
        for mode in (0, 1):
            ref = oracle.boxfft_blur(img, ns, mode)
            out = emu.gauss(img, -float(klen), mode, 1)[0]
            assert np.abs(out.astype(int) - ref.astype(int)).max() <= 1, (rows, cols, ns, mode)
    assert bb.boxfft_klen(100, 100, 4.0) == 16            # even: the plan entry refuses it
    with pytest.raises(bb.BlurB200Error):
        bb.GaussPlan(100, 100, 4.0, boxblur=True)


@pytest.mark.parametrize("rows,cols,sigma", [(64, 96, 2.0), (100, 75, 3.0), (135, 241, 6.0), (240, 320, 5.0), (467, 700, 7.5)])
def test_device_code_on_cpu_matches_oracle(emu, oracle, rows, cols, sigma):
    img = synth("noise", rows, cols, 5)
    for mode in (0, 1):
        ref, rp, _ = oracle.gauss_blur_1d(img, sigma, mode, want_f32=True)
        for nparts in (1, 2, 3):
            out, erp, _, _ = emu.gauss(img, sigma, mode, nparts)
            assert np.abs(erp - rp).max() <= 255 * 1e-5          # float32 intermediate, 1e-5 relative
            assert np.abs(out.astype(int) - ref.astype(int)).max() <= 1   # +-1 LSB


@pytest.mark.parametrize("rows,cols,sigma", [(64, 96, 2.0), (100, 75, 3.0), (135, 241, 6.0), (240, 320, 5.0),
                                             (467, 700, 7.5), (300, 1300, 25.0), (1100, 90, 12.0)])
def test_overlap_save_device_code_on_cpu_matches_oracle(emu, oracle, rows, cols, sigma):
    """ols_core.cuh (the default algorithm for nyquist_mode 0) executed on the CPU: blocks,
    reflect patching, swizzled passes, planar Q, octet U, and the multi-rank cut."""
    img = synth("noise", rows, cols, 5)
    ref, rp, _ = oracle.gauss_blur_1d(img, sigma, 0, want_f32=True)
    single = None
    for nparts in (1, 2, 3, -2, -3):     # negative: fused exchange (row pass stores into the owners' buffers)
        out, erp, info = emu.ols_gauss(img, sigma, nparts)
        if nparts > 0:                                       # (the fused mode has no local copy of Q to return)
            assert np.abs(erp - rp).max() <= 255 * 1e-5      # float32 intermediate, 1e-5 relative
        assert np.abs(out.astype(int) - ref.astype(int)).max() <= 1   # +-1 LSB
        single = out if single is None else single
        assert np.array_equal(out, single)                   # the cut never changes a bit


def test_overlap_save_multi_block_lines(emu, oracle, monkeypatch):
    """Force the smallest block (L = 1024) so that lines span several blocks with both edge and
    interior blocks, and an odd width / height."""
    monkeypatch.setenv("BLUR_B200_OLS_R0", "4")
    for rows, cols, sigma in [(1500, 37, 4.0), (45, 2300, 9.0), (1201, 1111, 20.0)]:
        img = synth("noise", rows, cols, 11)
        ref = oracle.gauss_blur_1d(img, sigma, 0)
        out, _, info = emu.ols_gauss(img, sigma, 1)
        assert info[0] == 1024 and info[1] == 1024 and max(info[2], info[3]) >= 2
        assert np.abs(out.astype(int) - ref.astype(int)).max() <= 1, (rows, cols)
    out, _, _ = emu.ols_gauss(synth("const", 1300, 40), 3.0, 1)
    assert np.all(out == 200)


def test_device_code_edge_inputs(emu, oracle):
    for kind in ("const", "impulse", "stripes", "ramp"):
        img = synth(kind, 96, 128)
        ref = oracle.gauss_blur_1d(img, 2.5, 0)
        out, _, _, _ = emu.gauss(img, 2.5, 0, 1)
        assert np.abs(out.astype(int) - ref.astype(int)).max() <= 1, kind
    out, _, _, _ = emu.gauss(synth("const", 40, 56), 3.0, 0, 1)
    assert np.all(out == 200)


_GLOO_WORKER = r'''
import os, sys, ctypes as C
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.environ["BLUR_ROOT"]); sys.path.insert(0, os.path.join(os.environ["BLUR_ROOT"], "tests"))
from util import Emulator, synth
from oracle_lib import Oracle
from blur_algorithms_b200.dist import ShardedGaussBlur
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%s" % os.environ["BLUR_PORT"],
                        rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
rank, n = dist.get_rank(), dist.get_world_size()
emu = Emulator()
for rows, cols, sigma, mode in [(64, 96, 2.0, 0), (135, 241, 6.0, 1), (101, 75, 3.0, 0)]:
    # pffft mode sharded: the CPU emulator has the full-length kernel's Nyquist handling only (the overlap-save
    # corrections are CUDA kernels, covered by test_gauss_sharded_pffft_mode_emulated_on_one_gpu)
    if mode == 1: os.environ["BLUR_B200_QUIRK_SHARDED"] = "full"
    else: os.environ.pop("BLUR_B200_QUIRK_SHARDED", None)
    img = synth("noise", rows, cols, 9)
    def row_pass(stripe, q):
        assert emu.lib.emul_row_pass(stripe.data_ptr(), q.data_ptr(), rows, cols, sigma, mode, rank, n) == 0
    def col_pass(q, out):
        assert emu.lib.emul_col_pass(q.data_ptr(), out.data_ptr(), rows, cols, sigma, mode, rank, n) == 0
    sb = ShardedGaussBlur(rows, cols, sigma, mode, device=torch.device("cpu"), row_pass=row_pass, col_pass=col_pass)
    g = sb.geom
    stripe = torch.from_numpy(np.ascontiguousarray(img[g.row0:g.row0 + g.valid_rows]))
    mine = sb.run(stripe).contiguous()
    # gather column stripes on every rank and compare with the single-process result
    parts = [torch.zeros((rows, 2 * g.cpp, 3), dtype=torch.uint8) for _ in range(n)]
    dist.all_gather(parts, sb.out_cols)
    full = torch.cat(parts, 1)[:, :cols].numpy()
    single = emu.ols_gauss(img, sigma, 1)[0] if g.algo == 1 else emu.gauss(img, sigma, mode, 1)[0]
    assert np.array_equal(full, single), (rows, cols)
    ref = Oracle().gauss_blur_1d(img, sigma, mode)
    assert np.abs(full.astype(int) - ref.astype(int)).max() <= 1
    # return exchange: column stripes -> the row stripes the input came in
    back = sb.to_row_stripe()
    assert back.shape == (g.valid_rows, cols, 3) and np.array_equal(back.numpy(), single[g.row0:g.row0 + g.valid_rows]), (rows, cols)
# ranks whose BLUR_B200_* environment differs would lay the intermediate out differently: refused loudly on every rank
os.environ.pop("BLUR_B200_QUIRK_SHARDED", None)
if rank == 1: os.environ["BLUR_B200_ALGO"] = "full"
try:
    ShardedGaussBlur(64, 96, 2.0, 0, device=torch.device("cpu"), row_pass=lambda *a: None, col_pass=lambda *a: None)
    raise SystemExit("layout disagreement between the ranks was not detected")
except RuntimeError as e:
    assert "disagree" in str(e), e
os.environ.pop("BLUR_B200_ALGO", None)
dist.barrier()
print("rank", rank, "ok")
'''


def test_sharded_host_logic_gloo_world2(tmp_path):
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_WORKER)
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", BLUR_PORT=str(port), BLUR_ROOT=ROOT, OMP_NUM_THREADS="2")
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
    outs = [p.communicate(timeout=300)[0].decode() for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o


_GLOO_BATCH_WORKER = r'''
import os, sys
import numpy as np, torch, torch.distributed as dist
sys.path.insert(0, os.environ["BLUR_ROOT"]); sys.path.insert(0, os.path.join(os.environ["BLUR_ROOT"], "tests"))
from oracle_lib import Oracle
from blur_algorithms_b200.dist import ShardedBatch
dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%s" % os.environ["BLUR_PORT"],
                        rank=int(os.environ["RANK"]), world_size=int(os.environ["WORLD_SIZE"]))
rank, n = dist.get_rank(), dist.get_world_size()
o = Oracle()
batch = 7                                                    # odd: the shares are ragged
imgs = np.stack([np.random.default_rng(100 + i).integers(0, 256, (40, 56, 3), dtype=np.uint8) for i in range(batch)])
sb = ShardedBatch(batch)                                     # rank / world size from the process group
assert sb.indices == list(range(rank, batch, n)) and len(sb) == len(sb.indices)
mine = sb.take(imgs)
assert mine.shape[0] == len(sb) and all(np.array_equal(mine[j], imgs[i]) for j, i in enumerate(sb.indices))
# the per-rank work, with the CPU oracle standing in for fastboxblur_batch (no GPU here)
done = [o.fastboxblur(im, 9, 2) for im in mine]
full = sb.gather(done)
assert len(full) == batch
for i in range(batch):
    assert np.array_equal(full[i], o.fastboxblur(imgs[i], 9, 2)), i
assert ShardedBatch(0, 0, 2).indices == [] and ShardedBatch(1, 1, 2).indices == []
dist.barrier()
print("rank", rank, "ok")
'''


def test_sharded_batch_gloo_world2(tmp_path):
    """ShardedBatch (image i -> rank i mod P, no collective on the data path): shares, ragged batch,
    reassembly in image order, at world size 2 over gloo."""
    import socket
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    script = tmp_path / "worker.py"
    script.write_text(_GLOO_BATCH_WORKER)
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", BLUR_PORT=str(port), BLUR_ROOT=ROOT, OMP_NUM_THREADS="2")
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT))
    outs = [p.communicate(timeout=300)[0].decode() for p in procs]
    for p, o in zip(procs, outs):
        assert p.returncode == 0, o


def test_bench_reference_arm_full_size_and_threads():
    """bench.py --impl reference: times the workload's full image (no sub-image), forces the OpenMP
    thread count itself (torchrun exports OMP_NUM_THREADS=1) and labels config / sample accordingly."""
    import json
    env = dict(os.environ, OMP_NUM_THREADS="1", RANK="0", WORLD_SIZE="2")
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "0", "--workload", "gauss_c1"], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stderr[-2000:]
    line = json.loads(r.stdout.strip().splitlines()[-1])
    cores = len(os.sched_getaffinity(0))
    assert line["impl"] == "reference" and line["cpu_baseline"]["cores"] == cores and line["run"]["threads"] == cores
    # `config` holds only what names the workload (the b200 arm prints the same dict); exact mode = Test(flag=5) -> pocketfft_1D
    assert sorted(line["config"]) == ["cols", "l2_policy", "nyquist_mode", "rows", "sharding", "sigma", "workload"]
    assert line["config"]["nyquist_mode"] == "exact" and "pocketfft_1D" in line["cpu_baseline"]["sample"]
    assert line["config"]["rows"] == 1080 and "1080x1920" in line["cpu_baseline"]["sample"] and "whole workload image" in line["cpu_baseline"]["sample"]
    assert line["e2e"]["h2d_bytes_per_step"] == 0 and line["steps"] == 1 and line["warmup"] == 0
    # rank != 0 exits 0 without work
    env["RANK"] = "1"
    r = subprocess.run([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                        "--warmup", "0", "--workload", "gauss_c1"], env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and r.stdout.strip() == ""


def test_cli_ppm_roundtrip(tmp_path):
    """CLI image I/O (reference: cv::imread / cv::imwrite, Source.cpp:621,636): P6 round trip, BGR order."""
    from blur_algorithms_b200 import cli
    img = np.random.default_rng(3).integers(0, 256, (7, 5, 3), dtype=np.uint8)
    p = str(tmp_path / "a.ppm")
    cli.write_image(p, img)
    assert open(p, "rb").read(2) == b"P6"
    assert np.array_equal(cli.read_image(p), img)
    assert cli.resize(img, 14, 10).shape == (14, 10, 3)


def test_sweep_plot_tool(tmp_path):
    """tools/plot_sweep.py (the equivalent of the reference's py/bench.png) renders the committed sweep
    lines into a well-formed SVG with one device and one host series per file (+ the reference's four published
    series when /root/reference is mounted)."""
    import xml.dom.minidom
    out = tmp_path / "sweep.svg"
    files = [os.path.join(ROOT, "profiles", f"r02z_sweep_flag{f}.jsonl") for f in (3, 5)]
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "plot_sweep.py"), "--out", str(out)] + files,
                       capture_output=True, text=True, timeout=120)
    assert r.returncode == 0, r.stderr
    doc = xml.dom.minidom.parse(str(out))
    paths = doc.getElementsByTagName("path")
    assert len(paths) >= 4
    text = out.read_text()
    assert "B200 pffft_ device-resident" in text and "B200 pocketfft_1D host call" in text


def test_api_type_helpers_accept_plane_lists_and_numpy():
    """deinterleave_BGR / interleave_BGR take a [3, total] tensor or a list / tuple of three planes; flip_block takes
    anything with element_size() or itemsize; everything else is a BlurB200Error, never an AttributeError."""
    import torch
    from blur_algorithms_b200 import api
    planes = [torch.zeros(4, dtype=torch.float32) for _ in range(3)]
    assert api._type_code(planes) == 3 and api._type_code(tuple(planes)) == 3
    assert api._type_code(torch.zeros((3, 4), dtype=torch.uint8)) == 0
    assert api._type_code(np.zeros(3, np.int32)) == 2
    assert api._elem_bytes(np.zeros(3, np.float32)) == 4 and api._elem_bytes(torch.zeros(3, dtype=torch.uint16)) == 2
    for bad in ([object()], torch.zeros(3, dtype=torch.float64), 7):
        with pytest.raises(api.BlurB200Error):
            api._type_code(bad)
    with pytest.raises(api.BlurB200Error):
        api._elem_bytes(object())
