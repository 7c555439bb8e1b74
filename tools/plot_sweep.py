#!/usr/bin/env python3
"""Plot equivalent of the reference's py/bench.png (py/performance.ipynb cell 0) with the B200 series added.

The reference plots, for its 45-size sweep (Source.cpp:627-634, sigma = sqrt(rows)), the wall time in
milliseconds of pffft_, pocketfft_1D, pocketfft_2D and cv::GaussianBlur on an Apple M3 Pro.  This script
draws those four series (read at run time from the notebook under /root/reference when that tree is
mounted: nothing is copied into the repo) together with this repo's sweep reproducer output
(`python -m blur_algorithms_b200.cli FLAG 0 synthetic --sweep`, one JSON line per size) on a log axis.

matplotlib is not in this image, so the figure is written as a plain SVG.

    python tools/plot_sweep.py --out profiles/r02_sweep.svg profiles/r02_sweep_flag3_pffft.jsonl ...
"""
import argparse
import json
import math
import os
import re

REF_NB = "/root/reference/py/performance.ipynb"
REF_COLORS = {"pffft": "#1f77b4", "pocketfft_1d": "#ff7f0e", "pocketfft_2d": "#2ca02c", "opencv": "#d62728"}
B200_COLORS = ["#9467bd", "#8c564b", "#e377c2", "#17becf", "#7f7f7f", "#bcbd22"]


def reference_series():
    """{name: [ms] * 46} from notebook cell 0; entry 0 is the extra call at the source image's native size."""
    if not os.path.exists(REF_NB):
        return {}
    src = "".join(json.load(open(REF_NB))["cells"][0]["source"])
    out = {}
    for name in REF_COLORS:
        m = re.search(rf"^{name}\s*=\s*np\.array\(\[([^\]]*)\]\)", src, re.M)
        if m:
            out[name] = [float(v) for v in m.group(1).split(",")]
    return out


def load_jsonl(path):
    rows = []
    for line in open(path):
        line = line.strip()
        if line.startswith("{"):
            rows.append(json.loads(line))
    return rows


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("jsonl", nargs="+")
    ap.add_argument("--out", default="profiles/sweep.svg")
    a = ap.parse_args()
    series = []        # (label, colour, dash, [(index, ms)])
    ref = reference_series()
    for name, vals in ref.items():
        series.append((f"reference {name} (M3 Pro, 12 cores, published)", REF_COLORS[name], "", [(i, v) for i, v in enumerate(vals[1:46])]))
    ci = 0
    sizes = None
    for path in a.jsonl:
        rows = [r for r in load_jsonl(path) if "error" not in r]
        if not rows:
            continue
        nm = rows[0].get("name", os.path.basename(path))
        if sizes is None or len(rows) > len(sizes):
            sizes = [f"{r['rows']} x {r['cols']}" for r in rows]
        col = B200_COLORS[ci % len(B200_COLORS)]
        ci += 1
        if "device_ms" in rows[0]:
            series.append((f"B200 {nm} device-resident", col, "", [(r["i"], r["device_ms"]) for r in rows]))
        series.append((f"B200 {nm} host call (H2D + kernels + D2H)", col, "6,4", [(r["i"], r["host_call_ms"]) for r in rows]))
    if not series:
        raise SystemExit("nothing to plot")
    n = max(i for s in series for i, _ in s[3]) + 1
    lo = min(v for s in series for _, v in s[3] if v > 0)
    hi = max(v for s in series for _, v in s[3])
    l0, l1 = math.floor(math.log10(lo)), math.ceil(math.log10(hi))
    W, H, ml, mr, mt, mb = 1500, 900, 90, 430, 50, 150
    pw, ph = W - ml - mr, H - mt - mb

    def X(i):
        return ml + pw * i / max(n - 1, 1)

    def Y(v):
        return mt + ph * (1 - (math.log10(v) - l0) / (l1 - l0))

    o = [f'<svg xmlns="http://www.w3.org/2000/svg" width="{W}" height="{H}" font-family="sans-serif" font-size="13">',
         f'<rect width="{W}" height="{H}" fill="white"/>',
         f'<text x="{ml}" y="28" font-size="18">Gaussian blur sweep, sigma = sqrt(rows): reference (published, M3 Pro) and B200 (this repo), milliseconds, log scale</text>']
    for d in range(l0, l1 + 1):
        y = Y(10 ** d)
        o.append(f'<line x1="{ml}" y1="{y:.1f}" x2="{ml + pw}" y2="{y:.1f}" stroke="#ccc"/>')
        o.append(f'<text x="{ml - 8}" y="{y + 4:.1f}" text-anchor="end">{10 ** d:g} ms</text>')
        for m in (2, 5):
            if d < l1:
                ym = Y(m * 10 ** d)
                o.append(f'<line x1="{ml}" y1="{ym:.1f}" x2="{ml + pw}" y2="{ym:.1f}" stroke="#eee"/>')
    for i in range(0, n, 4):
        x = X(i)
        o.append(f'<line x1="{x:.1f}" y1="{mt}" x2="{x:.1f}" y2="{mt + ph}" stroke="#eee"/>')
        lab = sizes[i] if sizes and i < len(sizes) else str(i)
        o.append(f'<text transform="translate({x:.1f},{mt + ph + 12}) rotate(60)">{lab}</text>')
    o.append(f'<rect x="{ml}" y="{mt}" width="{pw}" height="{ph}" fill="none" stroke="black"/>')
    for k, (label, col, dash, pts) in enumerate(series):
        pts = [(i, v) for i, v in pts if v > 0]
        path = " ".join(f"{'M' if j == 0 else 'L'}{X(i):.1f},{Y(v):.1f}" for j, (i, v) in enumerate(pts))
        da = f' stroke-dasharray="{dash}"' if dash else ""
        o.append(f'<path d="{path}" fill="none" stroke="{col}" stroke-width="2"{da}/>')
        ly = mt + 18 * k + 10
        o.append(f'<line x1="{ml + pw + 15}" y1="{ly}" x2="{ml + pw + 50}" y2="{ly}" stroke="{col}" stroke-width="2"{da}/>')
        o.append(f'<text x="{ml + pw + 56}" y="{ly + 4}">{label}</text>')
    o.append("</svg>")
    os.makedirs(os.path.dirname(a.out) or ".", exist_ok=True)
    open(a.out, "w").write("\n".join(o) + "\n")
    print(f"wrote {a.out}: {len(series)} series, {n} sizes")


if __name__ == "__main__":
    main()
