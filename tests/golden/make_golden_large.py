#!/usr/bin/env python
"""Golden vectors at BASELINE.json configs 1 and 2 (1080p sigma 5, 4K sigma 25), produced by the
reference itself: oracle/_ref = /root/reference/Source.cpp + Utils.hpp compiled unmodified
(oracle/Makefile `ref`).  Run in the build container (needs oracle/_ref):

    python tests/golden/make_golden_large.py      -> tests/golden/reference_large.npz

Per case and reference path (Test flag 5 = pocketfft_1D, flag 3 = pffft_): the SHA-256 of the full
uint8 output (pins the C restatement oracle/blur_oracle.c == reference at this size, CPU test) and
a lattice of sampled pixels plus the four borders' lattice (what the GPU output is compared with on
the GPU box, +-1 LSB).  Inputs: util.synth("noise", rows, cols, seed=0) = default_rng(0).
"""
import hashlib
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from oracle_lib import Reference  # noqa: E402
from util import LARGE_GOLDEN_CASES, lattice, synth  # noqa: E402


def main() -> None:
    ref = Reference()
    out = {}
    for rows, cols, sigma in LARGE_GOLDEN_CASES:
        img = synth("noise", rows, cols, 0)
        for flag, name in ((5, "pocketfft1d"), (3, "pffft")):
            res = ref.test(img, flag, sigma)
            key = f"noise_{rows}x{cols}_s{sigma}_{name}"
            out[key + "_sha256"] = np.frombuffer(hashlib.sha256(res.tobytes()).digest(), np.uint8)
            out[key + "_lattice"] = lattice(res)
            print(key, hashlib.sha256(res.tobytes()).hexdigest()[:16], out[key + "_lattice"].shape)
    np.savez_compressed(os.path.join(HERE, "reference_large.npz"), **out)


if __name__ == "__main__":
    main()
