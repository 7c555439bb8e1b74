"""Diagnostic: where does the GPU DFT_image differ from the golden spectrum by more than 1 LSB?"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import cv2
import blur_algorithms_b200 as bb
from oracle_lib import Oracle
o = Oracle()
G = os.path.join(ROOT, "tests", "golden")
im = cv2.imread(os.path.join(G, "spectrum_in_blur200.png"), cv2.IMREAD_COLOR)
gold = cv2.imread(os.path.join(G, "spectrum_out_blur200.png"), cv2.IMREAD_COLOR)
out, spec = bb.dft_image(im, 0.1, want_f32=True)
ref, rspec = o.dft_image(im, 0.1, want_f32=True)
d = np.abs(out.astype(int) - gold.astype(int))
print("hist gpu-vs-gold", np.bincount(d.ravel())[:8], "oracle-vs-gold", np.bincount(np.abs(ref.astype(int) - gold.astype(int)).ravel())[:8])
ys, xs, cs = np.nonzero(d > 1)
print("n>1:", len(ys))
for y, x, c in list(zip(ys, xs, cs))[:40]:
    print(y, x, c, "gpu", out[y, x, c], "gold", gold[y, x, c], "spec gpu", spec[c, y, x], "oracle", rspec[c, y, x],
          "|ReF| gpu", 10 ** (spec[c, y, x] / 20), "oracle", 10 ** (rspec[c, y, x] / 20))
