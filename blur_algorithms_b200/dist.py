"""Single large image sharded over the GPUs of one box.

Row pass on row stripes, one exchange of the float32 intermediate (the distributed form of the
reference's flip_block, Source.cpp:540), column pass on column stripes.  One process per GPU.

On GPUs every rank's receive buffer lives in symmetric memory (torch.distributed._symmetric_memory:
CUDA peer mappings over NVLink/NVSwitch) and no collective library call is on the data path.
Exchange modes (BLUR_B200_EXCHANGE):
  "staged" (default up to 4 GPUs where the plan supports it) every rank's part is cut into column sub-ranges; the row
          pass runs sub-range by sub-range, the copy engines push each finished sub-range to its owners
          under the next one's FFTs, and the column pass of a sub-range starts as soon as every rank's
          copies of it have landed (one device-side barrier per sub-range) while the later sub-ranges
          are still on the wire: the exchange hides behind both FFT passes
          (blur_b200_gauss_row_pass_staged / _stage_wait / _col_pass_stage);
  "sstore" (measured, not a default) the same sub-range pipeline without copy engines or send buffer: every row launch writes its
          sub-ranges straight into the owners' buffers with bulk shared->global copies over NVLink and is held to
          two CTAs per SM; the column pass of sub-range k runs on a second stream in the SMs' remaining
          room while row launch k+1 waits on the links;
  "push"  the row pass runs in wave-aligned chunks of row pairs; the kernel writes this
          rank's own part into its receive buffer and the rest into a local send buffer, and as soon
          as a chunk is done the GPU's COPY ENGINES move it into the owners' buffers (one 2-D peer
          copy per destination) underneath the next chunk's FFTs (blur_b200_gauss_row_pass_push);
  "store" (default beyond 4 GPUs) the row kernel's epilogue stores every result straight into the owner's buffer
          (blur_b200_gauss_row_pass_peer; bulk shared->global copies from the two-factor kernel): no send
          buffer, but the SMs stay busy until the last byte has crossed NVLink.
Device-side barriers order the ranks (push / store: two per blur).  Fallback (`fused=False`, or symmetric memory
unavailable): a local Q buffer + NCCL all_to_all.  The CPU tests run the same host logic over gloo
with the two passes supplied as callables.

Batches of images (BASELINE config 5, batched Gaussian) do not need any of this: ShardedBatch deals
image i to rank i mod P and every rank runs the single-GPU entry points on its share.

Layout contract (blur_b200_gauss_shard_geometry in include/blur_b200.h):
    algorithm 0 (full-length FFT per line)   q_send [nranks][3][cpp][stripe_rows][2]
    algorithm 1 (overlap-save)               q_send [3][nranks][stripe_rows][2*cpp]
block p of the sender goes to rank p and lands as block `rank` there; the column pass writes this
rank's column stripe [rows][2*cpp][3] uint8.
"""
from __future__ import annotations

import ctypes as C
import os
from dataclasses import dataclass

import torch
import torch.distributed as dist


@dataclass
class ShardGeometry:
    """Mirror of blur_b200_gauss_shard_geometry (host-only C call: the C side is the single
    source of truth for how the image is cut and which algorithm runs)."""
    rows: int
    cols: int
    nranks: int
    rank: int
    sigma: float = 1.0
    nyquist_mode: int = 0

    def __post_init__(self):
        from ._lib import check, lib
        o = (C.c_int64 * 8)()
        check(lib.blur_b200_gauss_shard_geometry(self.rows, self.cols, float(self.sigma), self.nyquist_mode,
                                                 self.rank, self.nranks, o))
        (self.algo, self.stripe_rows, self.cpp, self.row0, self.valid_rows, self.valid_cols,
         self.block_floats, self.qpitch) = (int(v) for v in o)

    @property
    def col0(self) -> int:
        return 2 * self.cpp * self.rank


def exchange(q_send: torch.Tensor, q_recv: torch.Tensor, group=None, algo: int = 0) -> None:
    """All-to-all of equal blocks: block p of q_send goes to rank p, lands as block `rank` there.
    algo 1 keeps the three channels outermost, so it is one all-to-all per channel."""
    n = dist.get_world_size(group) if dist.is_initialized() else 1
    if n == 1:
        q_recv.copy_(q_send)
        return
    if algo == 1:
        for s, r in zip(q_send.view(3, -1), q_recv.view(3, -1)):
            _all_to_all(s, r, n, group)
    else:
        _all_to_all(q_send.view(-1), q_recv.view(-1), n, group)


def _all_to_all(send: torch.Tensor, recv: torch.Tensor, n: int, group) -> None:
    try:
        dist.all_to_all_single(recv, send, group=group)
    except (RuntimeError, NotImplementedError):
        # backends without alltoall (older gloo): pairwise exchange
        rank = dist.get_rank(group)
        s, r = send.view(n, -1), recv.view(n, -1)
        r[rank].copy_(s[rank])
        reqs = []
        for p in range(n):
            if p == rank:
                continue
            reqs.append(dist.isend(s[p].contiguous(), p, group=group))
            reqs.append(dist.irecv(r[p], p, group=group))
        for q in reqs:
            q.wait()


class ShardedBatch:
    """A batch of independent images over the ranks: image i belongs to rank i mod nranks (SURVEY.md
    §8e row 2; BASELINE config 5).  No exchange, no collective: every rank calls the single-GPU entry
    points (fastboxblur_device / fastboxblur_batch, GaussPlan.run / run_host_batch) on its share.
    Host-side bookkeeping only, so it runs without a GPU (world-size-2 gloo test)."""

    def __init__(self, batch: int, rank: int | None = None, nranks: int | None = None, group=None):
        if rank is None:
            rank = dist.get_rank(group) if dist.is_initialized() else 0
        if nranks is None:
            nranks = dist.get_world_size(group) if dist.is_initialized() else 1
        if batch < 0 or nranks < 1 or not 0 <= rank < nranks:
            raise ValueError("bad batch / rank / nranks")
        self.batch, self.rank, self.nranks = batch, rank, nranks
        self.indices = list(range(rank, batch, nranks))

    def __len__(self) -> int:
        return len(self.indices)

    def take(self, images):
        """this rank's share of `images` (anything indexable by image number; arrays are gathered)"""
        if hasattr(images, "shape"):
            return images[self.rank::self.nranks]
        return [images[i] for i in self.indices]

    def fastboxblur(self, h_images, ksize: int, passes: int):
        """Box-blur this rank's share of a host batch [batch, h, w, C] in place through the streamed
        C-ABI host entry (blur_b200_fastboxblur_batch_host); returns the share (a copy if strided)."""
        import numpy as np
        from .api import fastboxblur_batch
        mine = np.ascontiguousarray(self.take(h_images))
        if len(mine):
            fastboxblur_batch(mine, ksize, passes)
        return mine

    def gauss(self, plan, h_images):
        """Gaussian-blur this rank's share of a list of host images in place through
        GaussPlan.run_host_batch (streamed uploads / kernels / downloads)."""
        mine = self.take(h_images)
        if len(mine):
            plan.run_host_batch(list(mine))
        return mine

    def gather(self, mine, group=None):
        """All ranks' shares back in image order (a list of length batch), via all_gather_object —
        convenience for tests and small batches; production callers keep results sharded."""
        if self.nranks == 1 or not dist.is_initialized():
            return list(mine)
        parts = [None] * self.nranks
        dist.all_gather_object(parts, list(mine), group=group)
        out = [None] * self.batch
        for r, p in enumerate(parts):
            for j, item in enumerate(p):
                out[r + j * self.nranks] = item
        return out


class ShardedGaussBlur:
    """One rank's half of the sharded blur.  Every rank's column stripe equals the single-GPU plan's output bit for
    bit when that plan runs the FFT kernels (always for half widths > 20; GaussPlan(..., direct=False) below that).  `row_pass(in_stripe, q_send)` and
    `col_pass(q_recv, out_cols)` default to the CUDA plan; tests may inject CPU stand-ins."""

    def __init__(self, rows: int, cols: int, sigma: float, nyquist_mode: int = 0, group=None,
                 device: torch.device | None = None, row_pass=None, col_pass=None, fused: bool | None = None,
                 exchange_mode: str | None = None, nchunks: int | None = None):
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.nranks = dist.get_world_size(group) if dist.is_initialized() else 1
        self.geom = ShardGeometry(rows, cols, self.nranks, self.rank, sigma, nyquist_mode)
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        g = self.geom
        nfl = self.nranks * g.block_floats
        self.plan = None
        self.symm = None
        # "push": row pass in chunks + copy engines over peer mappings; "store": the row kernel's epilogue stores
        # remotely itself; "staged": column sub-ranges pushed by the copy engines under the next sub-range's FFTs.
        # Which one wins depends on the box, not only on N (staged at N=4: 1.44 ms on an 8-GPU box, 2.4 ms on two
        # 4-GPU boxes; store 1.62, push 1.70), so the default "auto" times them on the first blur (tune_exchange)
        # and keeps the fastest.  An explicit exchange_mode / BLUR_B200_EXCHANGE skips the trial.
        self.exchange_mode = exchange_mode or os.environ.get("BLUR_B200_EXCHANGE") or "auto"
        self.exchange_trial = None                       # {mode: ms per blur, max over ranks} once tune_exchange has run
        self.nchunks = nchunks or int(os.environ.get("BLUR_B200_CHUNKS", "0")) or {2: 12}.get(self.nranks, 2)
        self.nstages = int(os.environ.get("BLUR_B200_STAGES", "0")) or 4
        # sstore: CTAs per SM of the row launches / of the column launches that run beside them
        self.row_ctas = int(os.environ.get("BLUR_B200_SSTORE_ROW_CTAS", "2"))
        self.col_ctas = int(os.environ.get("BLUR_B200_SSTORE_COL_CTAS", "1"))
        self._side = None
        self.q_send = None
        want_fused = (fused is not False and row_pass is None and col_pass is None and self.nranks > 1
                      and self.nranks <= 8 and g.algo == 1 and self.device.type == "cuda")
        if want_fused:
            try:
                import torch.distributed._symmetric_memory as symm_mem
                self.q_recv = symm_mem.empty(nfl, dtype=torch.float32, device=self.device)
                self.q_recv.zero_()
                self.symm = symm_mem.rendezvous(self.q_recv, group if group is not None else dist.group.WORLD)
                self.peer_ptrs = [int(p) for p in self.symm.buffer_ptrs]
            except Exception as e:                      # noqa: BLE001
                if fused is True:
                    raise
                self.symm = None
                self.fused_error = repr(e)
        if self.symm is None:
            self.q_send = torch.zeros(nfl, dtype=torch.float32, device=self.device)
            self.q_recv = torch.zeros_like(self.q_send)
        # [nranks * stripe_rows] rows so that the optional return exchange can send equal blocks; the blur uses the first `rows`
        self._out_store = torch.zeros((self.nranks * g.stripe_rows, 2 * g.cpp, 3), dtype=torch.uint8, device=self.device)
        self.out_cols = self._out_store[:rows]
        if row_pass is None or col_pass is None:
            from .api import GaussPlan
            self.plan = GaussPlan(rows, cols, sigma, nyquist_mode, device=self.device.index or 0,
                                  rank=self.rank, nranks=self.nranks)
            row_pass = self.plan.row_pass
            col_pass = self.plan.col_pass
        self._row_pass, self._col_pass = row_pass, col_pass
        self._check_geometry()
        # pffft mode on the overlap-save path: two small correction buffers are summed over the ranks per blur
        self._quirk = self.plan.quirk_buffers() if (self.plan is not None and nyquist_mode != 0 and g.algo == 1 and self.nranks > 1) else None
        if self._quirk is not None and self.exchange_mode == "sstore":
            self.exchange_mode = "staged"                # the second stream would not see quirk_finish
        # fewer stages if the part does not divide; the plan may not be able to run staged at all
        can_stage = self.plan is not None and self.symm is not None
        while can_stage and self.nstages > 1 and not self.plan.staged_supported(self.nstages):
            self.nstages //= 2
        self._can_stage = bool(can_stage and self.plan.staged_supported(self.nstages))
        if self.exchange_mode in ("staged", "sstore") and not self._can_stage:
            self.exchange_mode = "push"
        self._auto = self.exchange_mode == "auto"
        if self._auto:
            # until the trial has run (or where there is nothing to choose: no peer memory) the pair-major store
            self.exchange_mode = "store"
            self._auto = self.symm is not None and self.plan is not None

    def _check_geometry(self) -> None:
        """The layout of the exchanged intermediate is decided on the C side (shard_geometry), partly from
        BLUR_B200_* experiment knobs in the environment.  A rank whose environment differs, or a knob that changed
        between the geometry query and the plan's creation, would address the exchange buffers differently and
        corrupt them silently: compare the plan with the geometry, and the ranks with each other, once."""
        g = self.geom
        sig = (g.algo, g.stripe_rows, g.cpp, g.block_floats, g.qpitch)
        if self.plan is not None:
            i = self.plan.info
            mine = (i["algo"], i["stripe_rows"], i["cpp"], i["block_floats"])
            if mine != sig[:4]:
                raise RuntimeError(f"rank {self.rank}: plan layout {mine} differs from blur_b200_gauss_shard_geometry {sig[:4]} "
                                   "(BLUR_B200_* environment changed between the two calls?)")
            sig += (i["ols_L_row"], i["ols_V_row"], i["ols_L_col"], i["ols_V_col"], i["ols_nblocks_row"], i["ols_nblocks_col"])
        if self.nranks > 1 and dist.is_initialized():
            sigs = [None] * self.nranks
            dist.all_gather_object(sigs, sig, group=self.group)
            if any(tuple(x) != tuple(sigs[0]) for x in sigs):
                raise RuntimeError(f"sharded blur: the ranks disagree on the layout of the intermediate: {sigs} "
                                   "(the BLUR_B200_* environment must be the same on every rank)")

    @property
    def fused(self) -> bool:
        return self.symm is not None

    def describe_exchange(self) -> str:
        """What moves the float32 intermediate between the ranks (for bench.py's config line)."""
        if not self.fused:
            return "NCCL all_to_all of the float32 intermediate (local send buffer)"
        if self.exchange_mode == "sstore":
            return (f"staged store: {self.nstages} column sub-ranges per rank; every row launch writes its sub-ranges straight into "
                    f"the owners' symmetric-memory buffers (bulk shared->global copies over NVLink, {self.row_ctas} CTAs/SM); the column "
                    f"pass of a sub-range runs on a second stream ({self.col_ctas} CTA/SM beside the row launches) after a device-side "
                    "barrier for that sub-range")
        if self.exchange_mode == "staged":
            return (f"staged: {self.nstages} column sub-ranges per rank; row pass sub-range by sub-range, copy engines "
                    "(cudaMemcpy2DAsync peer copies, one per destination and sub-range) push each finished sub-range into the "
                    "owners' symmetric-memory buffers under the next one's FFTs; the column pass of a sub-range starts after a "
                    "device-side barrier for that sub-range while later ones are still on the wire")
        if self.exchange_mode == "store":
            return "row-kernel epilogue stores into the owners' symmetric-memory buffers over NVLink (bulk shared->global copies)"
        return (f"row pass in {self.nchunks} wave-aligned chunks; copy engines (cudaMemcpy2DAsync peer copies, one per "
                "destination and chunk) move each finished chunk from a local send buffer into the owners' "
                "symmetric-memory buffers under the next chunk's FFTs")

    def tune_exchange(self, in_stripe: torch.Tensor, candidates=("store", "staged", "push"), steps: int = 3) -> dict:
        """Collective: every rank times `steps` whole blurs of `in_stripe` per exchange variant (after one untimed
        blur each; CUDA events, MAX over ranks so that all ranks see the same numbers), keeps the fastest and
        returns {mode: ms per blur}.  Called by the first run() when the mode was left to "auto"; callers that
        build many plans of one geometry can pass the winner as exchange_mode to the next ones."""
        self._auto = False
        if self.symm is None or self.plan is None:
            return {}
        stream = torch.cuda.current_stream(self.device)
        trial = {}
        for mode in candidates:
            if mode in ("staged", "sstore") and (not self._can_stage or (mode == "sstore" and self._quirk is not None)):
                continue
            self.exchange_mode = mode
            self._rows(in_stripe); self._cols()
            ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            ev0.record(stream)
            for _ in range(steps):
                self._rows(in_stripe); self._cols()
            ev1.record(stream)
            ev1.synchronize()
            t = torch.tensor([ev0.elapsed_time(ev1) / steps], dtype=torch.float32, device=self.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
            trial[mode] = round(float(t.item()), 4)
        self.exchange_mode = min(trial, key=trial.get)      # same dict on every rank: same choice
        self.exchange_trial = trial
        return trial

    def row_pass_only(self) -> None:
        """The row phase as run() issues it, without the two barriers (per-pass timing in bench.py)."""
        if self.geom.valid_rows <= 0:
            return
        if not self.fused:
            self._row_pass(self._last_in, self.q_send)
        elif self.exchange_mode in ("staged", "sstore"):
            st = self.exchange_mode == "sstore"
            self.plan.row_pass_staged(self._last_in, self.peer_ptrs, self.nstages, store=st, row_ctas_per_sm=self.row_ctas if st else 0)
            for k in range(self.nstages):
                self.plan.stage_wait(k)
        elif self.exchange_mode == "store":
            self.plan.row_pass_peer(self._last_in, self.peer_ptrs)
        else:
            self.plan.row_pass_push(self._last_in, self.peer_ptrs, self.nchunks)

    def col_pass_only(self) -> None:
        """The column phase without barriers (per-pass timing in bench.py)."""
        if self.fused and self.exchange_mode in ("staged", "sstore"):
            for k in range(self.nstages):
                self.plan.col_pass_stage(self.q_recv, self.out_cols, k, self.nstages)
        else:
            self._col_pass(self.q_recv, self.out_cols)

    def run(self, in_stripe: torch.Tensor, row_sharded: bool = False) -> torch.Tensor:
        """in_stripe: uint8 [valid_rows, cols, 3] (this rank's rows).  Returns this rank's column
        stripe of the result, uint8 [rows, valid_cols, 3] (a view into an internal buffer) — or, with
        row_sharded=True, this rank's ROW stripe [valid_rows, cols, 3] after the return exchange."""
        if self._auto:
            self.tune_exchange(in_stripe)
        self._rows(in_stripe)
        out = self._cols()
        return self.to_row_stripe() if row_sharded else out

    def to_row_stripe(self) -> torch.Tensor:
        """Return exchange (optional): the uint8 result goes back from column stripes to the row stripes
        the input came in (3 B/px, a quarter of the float32 exchange), one all-to-all of equal blocks
        [stripe_rows][2*cpp][3]; the output of one sharded blur can then feed the next.  A collective call
        (NCCL / gloo all_to_all), not a hot-path kernel."""
        g, n = self.geom, self.nranks
        if n == 1:
            return self.out_cols[: g.valid_rows, : g.cols]
        send = self._out_store.view(n, g.stripe_rows, 2 * g.cpp, 3)
        if not hasattr(self, "_ret_recv"):
            self._ret_recv = torch.empty_like(send)
        _all_to_all(send.view(-1), self._ret_recv.view(-1), n, self.group)
        rows_here = max(g.valid_rows, 0)
        # block s holds this rank's rows x rank s' columns
        return self._ret_recv.permute(1, 0, 2, 3).reshape(g.stripe_rows, n * 2 * g.cpp, 3)[:rows_here, : g.cols].contiguous()

    def _quirk_sum(self) -> None:
        """pffft-mode sharded plans: the column corrections need the row corrections and the integer column sums of ALL
        rows — all-reduce the two small buffers (exact: disjoint non-zeros / integers), then finish on this rank's columns."""
        if self._quirk is None:
            return
        for t in self._quirk:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        self.plan.quirk_finish()

    def _rows(self, in_stripe: torch.Tensor) -> None:
        """row pass of this rank's stripe + the exchange (the distributed flip_block, Source.cpp:540)"""
        self._last_in = in_stripe
        if self._quirk is not None:
            self.plan.quirk_local(in_stripe if self.geom.valid_rows > 0 else None)
        if self.symm is not None:
            # peers may still be reading their buffers (previous blur's column pass)
            self.symm.barrier(channel=0)
            if self.exchange_mode == "sstore":
                if self._side is None:
                    self._side = torch.cuda.Stream(self.device, priority=-1)
                    self._ev_b0 = torch.cuda.Event()
                self._ev_b0.record(torch.cuda.current_stream(self.device))       # the second stream starts behind this barrier
            if self.exchange_mode in ("staged", "sstore"):
                if self.geom.valid_rows > 0:
                    st = self.exchange_mode == "sstore"
                    self.plan.row_pass_staged(in_stripe, self.peer_ptrs, self.nstages, store=st,
                                              row_ctas_per_sm=self.row_ctas if st else 0)
                return                                   # the per-stage barriers are in _cols()
            if self.geom.valid_rows > 0:
                if self.exchange_mode == "store":
                    self.plan.row_pass_peer(in_stripe, self.peer_ptrs)
                else:
                    self.plan.row_pass_push(in_stripe, self.peer_ptrs, self.nchunks)
            self.symm.barrier(channel=1)                 # every rank's stores have landed
        else:
            if self.geom.valid_rows > 0:
                self._row_pass(in_stripe, self.q_send)
            exchange(self.q_send, self.q_recv, self.group, self.geom.algo)

    def _cols(self, after=None) -> torch.Tensor:
        """`after`: an event the column passes must also wait for (the streamed host path: the previous image's
        download has left out_cols)."""
        self._quirk_sum()
        sstore = self.symm is not None and self.exchange_mode == "sstore"
        main = torch.cuda.current_stream(self.device) if (sstore or after is not None) else None
        if after is not None and not sstore:
            main.wait_event(after)
        if sstore:
            # second stream: barrier + column pass of sub-range k as soon as row launch k is done everywhere, beside
            # the row launches of the later sub-ranges (which mostly wait on NVLink)
            self._side.wait_event(self._ev_b0)
            if after is not None:
                self._side.wait_event(after)
            with torch.cuda.stream(self._side):
                for k in range(self.nstages):
                    if self.geom.valid_rows > 0:
                        self.plan.stage_wait(k, self._side.cuda_stream)
                    self.symm.barrier(channel=1 + k)
                    last = k == self.nstages - 1
                    self.plan.col_pass_stage(self.q_recv, self.out_cols, k, self.nstages, self._side.cuda_stream,
                                             col_ctas_per_sm=0 if last else self.col_ctas)
            main.wait_stream(self._side)
        elif self.symm is not None and self.exchange_mode == "staged":
            for k in range(self.nstages):
                if self.geom.valid_rows > 0:
                    self.plan.stage_wait(k)              # this rank's outgoing copies of sub-range k are complete ...
                self.symm.barrier(channel=1 + k)         # ... and so are everybody else's: sub-range k has landed here
                self.plan.col_pass_stage(self.q_recv, self.out_cols, k, self.nstages)
        else:
            self._col_pass(self.q_recv, self.out_cols)
        return self.out_cols[:, : self.geom.valid_cols]

    def run_host_stream(self, h_ins: list, h_outs: list) -> None:
        """Stream images through the sharded blur from host memory: h_ins[i] = this rank's row stripe of
        image i (pinned uint8 [valid_rows, cols, 3]), h_outs[i] = pinned uint8 [rows, 2*cpp, 3] receiving
        this rank's column stripe.  The upload of image i+1 runs under the kernels and the download of
        image i (separate streams, double-buffered input stripe); the column pass of image i+1 waits
        for the download of image i because the result buffer is single."""
        if len(h_ins) != len(h_outs):
            raise ValueError("h_ins and h_outs must have the same length")
        cur = torch.cuda.current_stream(self.device)
        if self._auto and len(h_ins):
            self.tune_exchange(h_ins[0].to(self.device))
        if not hasattr(self, "_s_up"):
            self._s_up = torch.cuda.Stream(self.device)
            self._s_dn = torch.cuda.Stream(self.device)
            self._d_in = [torch.empty((self.geom.valid_rows, self.geom.cols, 3), dtype=torch.uint8, device=self.device)
                          for _ in range(2)]
        ev_up = [torch.cuda.Event(), torch.cuda.Event()]
        ev_row = [torch.cuda.Event(), torch.cuda.Event()]
        ev_dn = None
        for i, (hi, ho) in enumerate(zip(h_ins, h_outs)):
            k = i & 1
            with torch.cuda.stream(self._s_up):
                if i >= 2:
                    self._s_up.wait_event(ev_row[k])     # the row pass of image i-2 has read this slot
                self._d_in[k].copy_(hi, non_blocking=True)
                ev_up[k].record(self._s_up)
            cur.wait_event(ev_up[k])
            self._rows(self._d_in[k])
            ev_row[k].record(cur)
            self._cols(after=ev_dn)                      # image i-1 has left out_cols
            ev_c = torch.cuda.Event()
            ev_c.record(cur)
            with torch.cuda.stream(self._s_dn):
                self._s_dn.wait_event(ev_c)
                ho.copy_(self.out_cols, non_blocking=True)
                ev_dn = torch.cuda.Event()
                ev_dn.record(self._s_dn)
        self._s_dn.synchronize()
        cur.synchronize()


class HaloShardedGaussBlur:
    """Row-stripe ("halo") sharding of one image: input AND output row-sharded (SURVEY.md 8e's halo alternative,
    reported beside the contract configuration, not instead of it).  Every rank keeps a stripe of whole column-pass
    blocks, runs the row pass on it, pushes its first / last `halo` rows of the float32 intermediate into the two
    neighbours' buffers (symmetric memory over NVLink, two copy-engine copies) and runs the column pass on its own
    blocks: 2 * halo * cols * 12 bytes per GPU on the wire instead of rows * cols * 12 * (P - 1) / P^2, no all-to-all,
    result bit-identical to the single-GPU plan's FFT kernels (GaussPlan(..., direct=False) for kernels of half width
    <= 20, whose whole-image entries otherwise take the direct kernel: +-1 LSB).  Exact mode, kernels narrow enough for the two-factor blocks."""

    def __init__(self, rows: int, cols: int, sigma: float, group=None, device: torch.device | None = None):
        from .api import GaussPlan
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.nranks = dist.get_world_size(group) if dist.is_initialized() else 1
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        # direct=False: plan.run() of the whole image then uses the same FFT kernels as the halo entries (bit-identical)
        self.plan = GaussPlan(rows, cols, sigma, 0, device=self.device.index or 0, direct=False)
        info = self.plan.halo_info(self.rank, self.nranks)
        if info is None:
            raise ValueError("this geometry cannot run row-stripe sharded (needs the two-factor kernels on both axes)")
        self.info, self.rows, self.cols = info, rows, cols
        self.row0, self.valid_rows = info["row0"], info["valid_rows"]
        import torch.distributed._symmetric_memory as symm_mem
        self.q_local = symm_mem.empty(info["q_floats"], dtype=torch.float32, device=self.device)
        self.q_local.zero_()
        self.symm = symm_mem.rendezvous(self.q_local, group if group is not None else dist.group.WORLD) if self.nranks > 1 else None
        ptrs = [int(p) for p in self.symm.buffer_ptrs] if self.symm is not None else [self.q_local.data_ptr()]
        self.q_up = ptrs[self.rank - 1] if self.rank > 0 else None
        self.q_down = ptrs[self.rank + 1] if self.rank + 1 < self.nranks else None
        self.out_rows = torch.zeros((max(self.valid_rows, 1), cols, 3), dtype=torch.uint8, device=self.device)

    def run(self, in_stripe: torch.Tensor) -> torch.Tensor:
        """in_stripe: uint8 [valid_rows, cols, 3] = image rows [row0, row0 + valid_rows).  Returns the same rows blurred."""
        r, n, p = self.rank, self.nranks, self.plan
        if self.symm is not None:
            self.symm.barrier(channel=0)                 # the neighbours' column passes of the previous blur have read my halos
        p.halo_row_pass(r, n, in_stripe if self.valid_rows > 0 else None, self.q_local)
        p.halo_push(r, n, self.q_local, self.q_up, self.q_down)
        if self.symm is not None:
            self.symm.barrier(channel=1)                 # both neighbours' halos have landed here
        p.halo_col_pass(r, n, self.q_local, self.out_rows)
        return self.out_rows[: self.valid_rows]
