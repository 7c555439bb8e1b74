"""Single large image sharded over the GPUs of one box.

Row pass on row stripes, one all-to-all of the float32 intermediate (the distributed form of the
reference's flip_block, Source.cpp:540), column pass on column stripes.  One process per GPU,
torch.distributed for the exchange (NCCL over NVLink on GPUs; gloo in the CPU tests, where the two
passes are supplied by the test as callables).

Layout contract (include/blur_b200.h, blur_io.cuh):
    row pass writes   q_send [nranks][3][cpp][stripe_rows][2]   block p is for rank p
    all-to-all gives  q_recv [nranks][3][cpp][stripe_rows][2]   block s came from rank s
    col pass reads q_recv and writes this rank's column stripe  [rows][2*cpp][3] uint8
"""
from __future__ import annotations

from dataclasses import dataclass

import torch
import torch.distributed as dist


@dataclass
class ShardGeometry:
    rows: int
    cols: int
    nranks: int
    rank: int

    @property
    def stripe_rows(self) -> int:
        sr = -(-self.rows // self.nranks)
        return sr + (sr & 1)

    @property
    def cpp(self) -> int:
        return -(-((self.cols + 1) // 2) // self.nranks)

    @property
    def row0(self) -> int:
        return self.rank * self.stripe_rows

    @property
    def valid_rows(self) -> int:
        return max(0, min(self.stripe_rows, self.rows - self.row0))

    @property
    def col0(self) -> int:
        return 2 * self.cpp * self.rank

    @property
    def valid_cols(self) -> int:
        return max(0, min(2 * self.cpp, self.cols - self.col0))

    @property
    def block_floats(self) -> int:
        return 3 * self.cpp * self.stripe_rows * 2


def exchange(q_send: torch.Tensor, q_recv: torch.Tensor, group=None) -> None:
    """All-to-all of equal blocks: block p of q_send goes to rank p, lands as block `rank` there."""
    n = dist.get_world_size(group)
    if n == 1:
        q_recv.copy_(q_send)
        return
    try:
        dist.all_to_all_single(q_recv.view(-1), q_send.view(-1), group=group)
    except (RuntimeError, NotImplementedError):
        # backends without alltoall (older gloo): pairwise exchange
        rank = dist.get_rank(group)
        s, r = q_send.view(n, -1), q_recv.view(n, -1)
        r[rank].copy_(s[rank])
        reqs = []
        for p in range(n):
            if p == rank:
                continue
            reqs.append(dist.isend(s[p].contiguous(), p, group=group))
            reqs.append(dist.irecv(r[p], p, group=group))
        for q in reqs:
            q.wait()


class ShardedGaussBlur:
    """One rank's half of the sharded blur.  `row_pass(in_stripe, q_send)` and
    `col_pass(q_recv, out_cols)` default to the CUDA plan; tests may inject CPU stand-ins."""

    def __init__(self, rows: int, cols: int, sigma: float, nyquist_mode: int = 0, group=None,
                 device: torch.device | None = None, row_pass=None, col_pass=None):
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.nranks = dist.get_world_size(group) if dist.is_initialized() else 1
        self.geom = ShardGeometry(rows, cols, self.nranks, self.rank)
        self.device = device if device is not None else torch.device("cuda", torch.cuda.current_device())
        g = self.geom
        self.q_send = torch.zeros(self.nranks * g.block_floats, dtype=torch.float32, device=self.device)
        self.q_recv = torch.zeros_like(self.q_send)
        self.out_cols = torch.zeros((rows, 2 * g.cpp, 3), dtype=torch.uint8, device=self.device)
        self.plan = None
        if row_pass is None or col_pass is None:
            from .api import GaussPlan
            self.plan = GaussPlan(rows, cols, sigma, nyquist_mode, device=self.device.index or 0,
                                  rank=self.rank, nranks=self.nranks)
            row_pass = self.plan.row_pass
            col_pass = self.plan.col_pass
        self._row_pass, self._col_pass = row_pass, col_pass

    def run(self, in_stripe: torch.Tensor) -> torch.Tensor:
        """in_stripe: uint8 [valid_rows, cols, 3] (this rank's rows).  Returns this rank's column
        stripe of the result, uint8 [rows, valid_cols, 3] (a view into an internal buffer)."""
        if self.geom.valid_rows > 0:
            self._row_pass(in_stripe, self.q_send)
        exchange(self.q_send, self.q_recv, self.group)
        self._col_pass(self.q_recv, self.out_cols)
        return self.out_cols[:, : self.geom.valid_cols]
