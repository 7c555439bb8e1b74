"""Host model of box2_tile_kernel's geometry (box2_plan) and of its packed h/v arithmetic, checked
against a direct sliding mean.  Pure numpy, no GPU: run after touching box2_plan or box2_hpass."""
import sys
import numpy as np

M32 = 0xFFFFFFFF


def plan(w, h, C, r):
    k = 2 * r + 1; wb = w * C; rC = r * C; kC = k * C
    if C not in (3, 4) or k < 3 or k > 127 or wb % 16 or w <= r + 8 or h <= r:
        return None
    a = dict(w=w, h=h, r=r, wb=wb, C=C)
    tiles_x = (wb + 511) // 512
    TXB = (((wb + tiles_x - 1) // tiles_x) + 15) & ~15
    tiles_x = (wb + TXB - 1) // TXB
    bn = rC & 3; S = rC - bn; E = S & 15
    nw = (kC + 15) // 16; mcnt = 16 * nw - kC
    O = (max(rC + 16, 16 * nw - (S - E)) + 15) & ~15
    cN = (S - E + O) // 16
    So = O - E - C - rC
    ni = (TXB + E + 15) // 16
    chunks = max(ni + cN + 1, ni + (So >> 4) + 2, (O + TXB + rC + 15) // 16 + 1)
    if chunks % 2 == 0:
        chunks += 1
    a.update(tiles_x=tiles_x, TXB=TXB, bn=bn, E=E, nw=nw, mcnt=mcnt, O=O, cn0=cN - nw, co0=So >> 4, wo=(So >> 2) & 3,
             bo=So & 3, ni=ni, pitch=16 * chunks)
    assert So >= 0 and a["cn0"] >= 0
    inv = float(np.float32(1.0 / k))
    a["inv"] = np.float32(inv); a["cb"] = np.float32((r + 0.5) / k - 32768.0 * inv)
    return a


def funnel(lo, hi, sh):
    return ((lo | (hi << 32)) >> sh) & M32


def even(v): return v & 0x00FF00FF
def odd(v): return (v >> 8) & 0x00FF00FF


def round4(e, o, a):
    out = 0
    for idx, n in enumerate([e & 0xFFFF, o & 0xFFFF, e >> 16, o >> 16]):
        assert n < 32768
        t = np.float32(np.float32(32768 + n) * np.float64(a["inv"]) + np.float64(a["cb"]))   # fma: one rounding
        out |= (int(np.floor(t)) & 255) << (8 * idx)
    return out


def hrow(row_bytes, x0b, a):
    """row_bytes: reflect-extended image row as a function gb -> byte; returns h result bytes b in [0, TXB)."""
    C, r = a["C"], a["r"]; rC = r * C
    srow = np.zeros(a["pitch"] + 32, dtype=np.uint8)
    srow[:] = 0xAB                                                     # garbage
    for u in range(-(rC + 16), a["TXB"] + rC):
        srow[u + a["O"]] = row_bytes(x0b + u)
    words = srow.view(np.uint32).astype(np.uint64)

    def chunk(c): return [int(words[4 * c + j]) for j in range(4)]
    e = o = 0
    shn, sho = 8 * a["bn"], 8 * a["bo"]
    m = []
    for j in range(4):
        z = a["mcnt"] - 4 * j
        m.append(M32 if z <= 0 else (0 if z >= 4 else (M32 << (8 * z)) & M32))
    nc = chunk(a["cn0"])

    def step(e, o, nw_, ow_):
        de = (even(nw_) - even(ow_)) & M32; do = (odd(nw_) - odd(ow_)) & M32
        if C == 3:
            e2 = (o + de) & M32
            o2 = (((e >> 16) | ((e2 & 0xFFFF) << 16)) + do) & M32
            return e2, o2
        return (e + de) & M32, (o + do) & M32
    ci = a["cn0"]
    for i in range(a["nw"]):
        nn = chunk(ci + 1)
        ws = [funnel(nc[0], nc[1], shn), funnel(nc[1], nc[2], shn), funnel(nc[2], nc[3], shn), funnel(nc[3], nn[0], shn)]
        for j in range(4):
            e, o = step(e, o, ws[j] & m[j], 0)
        m = [M32] * 4
        nc = nn; ci += 1
    oc = chunk(a["co0"]); oi = a["co0"]
    out = np.zeros(16 * a["ni"], dtype=np.uint8)
    for i in range(a["ni"]):
        nn = chunk(ci + 1); on = chunk(oi + 1)
        ws = [funnel(nc[0], nc[1], shn), funnel(nc[1], nc[2], shn), funnel(nc[2], nc[3], shn), funnel(nc[3], nn[0], shn)]
        ow8 = oc + on
        for j in range(4):
            owd = funnel(ow8[a["wo"] + j], ow8[a["wo"] + j + 1], sho)
            e, o = step(e, o, ws[j], owd)
            if 16 * i + 4 * j - a["E"] >= 0 and 16 * i + 4 * j - a["E"] < a["TXB"]:
                v = round4(e, o, a)
            else:
                v = 0
            # in place: the store must not touch anything still to be read
            for t in range(4):
                out[16 * i + 4 * j + t] = (v >> (8 * t)) & 255
        # emulate the in-place write
        srow[16 * i:16 * i + 16] = out[16 * i:16 * i + 16]
        words = srow.view(np.uint32).astype(np.uint64)
        nc = nn; oc = on; ci += 1; oi += 1
    return out[a["E"]:a["E"] + a["TXB"]]


def check(w, C, r, seed=0):
    a = plan(w, 64, C, r)
    if a is None:
        return "skip"
    rng = np.random.default_rng(seed)
    row = rng.integers(0, 256, (w, C), dtype=np.uint8)
    if seed % 3 == 0:
        row[:] = 255

    def refl(j):
        j = abs(j)
        return 2 * (w - 1) - j if j >= w else j

    def rb(gb):
        j = gb // C; c = gb - j * C
        return row[min(max(refl(j), 0), w - 1), c]
    k = 2 * r + 1
    ext = np.stack([row[refl(j)] for j in range(-r, w + r)]).astype(np.int64)
    want = np.stack([(ext[j:j + k].sum(0) + r) // k for j in range(w)]).astype(np.uint8).reshape(-1)
    for tx in range(a["tiles_x"]):
        x0b = tx * a["TXB"]
        got = hrow(rb, x0b, a)
        n = min(a["TXB"], a["wb"] - x0b)
        if not np.array_equal(got[:n], want[x0b:x0b + n]):
            bad = np.nonzero(got[:n] != want[x0b:x0b + n])[0]
            return f"MISMATCH w={w} C={C} r={r} tile {tx}: first bad byte {bad[0]} got {got[bad[0]]} want {want[x0b + bad[0]]} ({len(bad)} bad) plan={a}"
    return "ok"


if __name__ == "__main__":
    bad = 0
    for C in (3, 4):
        for r in list(range(1, 22)) + [31, 40, 63]:
            for w in (64, 176, 1920 if C == 3 else 480):
                if (w * C) % 16:
                    w = (w * C + 15) // 16 * 16 // C if C == 4 else w + (16 - w % 16)
                res = check(w, C, r, seed=r + C)
                if res not in ("ok", "skip"):
                    print(res); bad += 1
    print("box2 model:", "FAILED" if bad else "all geometries ok")
    sys.exit(1 if bad else 0)
