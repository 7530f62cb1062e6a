#!/usr/bin/env python3
"""Development-time prototype of the data-parallel DistributeOctTree formulation used by csrc/k_quadtree.cu.

Each "sweep" of the reference's list walk (src/ORBextractor.cc:955-1021) is restated as: split every expandable
node independently, prefix-sum the size deltas in list order to find where `lNodes.size() >= N` (:959-963) would
stop the walk, then rebuild the list as reverse(pushed children) ++ survivors.  The final phase (:1030-1095)
is restated as an ordering by (count, creation order) descending plus a prefix sum.  This script checks the
formulation against the CPU oracle on random inputs; it is not part of the product.
"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))

f32 = np.float32


def c_round(v):  # C round(): half away from zero
    return int(np.floor(abs(v) + 0.5)) * (1 if v >= 0 else -1)


def distribute(keys, minX, maxX, minY, maxY, N):
    """keys: list of (x, y, resp) ints in candidate order -> list of kept keys in the reference's output order."""
    nIni = c_round(float(f32(maxX - minX) / f32(maxY - minY)))
    if nIni < 1:
        return []
    hX = f32(maxX - minX) / f32(nIni)
    H = maxY - minY
    # node = [x0,y0,x1,y1,keys(list),noMore]
    roots = []
    for i in range(nIni):
        roots.append([int(hX * f32(i)), 0, int(hX * f32(i + 1)), H, [], False])
    for k in keys:
        r = min(int(f32(k[0]) / hX), nIni - 1)
        roots[r][4].append(k)
    L = []
    for n in roots:
        if len(n[4]) == 0:
            continue
        n[5] = len(n[4]) == 1
        L.append(n)

    def split(n):
        x0, y0, x1, y1, ks, _ = n
        hx = -((-(x1 - x0)) // 2)
        hy = -((-(y1 - y0)) // 2)
        mx, my = x0 + hx, y0 + hy
        ch = [[x0, y0, mx, my, [], False], [mx, y0, x1, my, [], False], [x0, my, mx, y1, [], False], [mx, my, x1, y1, [], False]]
        for k in ks:
            c = (0 if k[0] < mx else 1) + (0 if k[1] < my else 2)
            ch[c][4].append(k)
        out = []
        for c in ch:
            if c[4]:
                c[5] = len(c[4]) == 1
                out.append(c)
        return out

    finish = False
    while not finish:
        prev = len(L)
        # --- sweep, parallel formulation ---
        kids = [split(n) if not n[5] else None for n in L]
        delta = np.array([(len(k) - 1) if k is not None else 0 for k in kids], dtype=np.int64)
        before = prev + np.concatenate([[0], np.cumsum(delta)[:-1]]) if prev else np.zeros(0, np.int64)
        hit = np.nonzero(before >= N)[0]
        cut = int(hit[0]) if hit.size else prev
        pushed = []
        for i in range(cut):
            if kids[i] is not None:
                pushed.extend(kids[i])
        survivors = [L[i] for i in range(cut) if kids[i] is None] + L[cut:]
        L = pushed[::-1] + survivors
        expandable = [n for n in pushed if len(n[4]) > 1]
        if len(L) >= N or len(L) == prev:
            finish = True
        elif len(L) + 3 * len(expandable) > N:
            while not finish:
                prev = len(L)
                P = expandable
                order = sorted(range(len(P)), key=lambda j: (len(P[j][4]), j), reverse=True)
                kids = [split(P[j]) for j in order]
                inc = np.cumsum([len(k) - 1 for k in kids]) if kids else np.zeros(0, np.int64)
                hit = np.nonzero(prev + inc >= N)[0]
                last = int(hit[0]) if hit.size else len(order) - 1
                pushed = []
                erased = set()
                for t in range(last + 1):
                    pushed.extend(kids[t])
                    erased.add(id(P[order[t]]))
                L = pushed[::-1] + [n for n in L if id(n) not in erased]
                expandable = [n for n in pushed if len(n[4]) > 1]
                if len(L) >= N or len(L) == prev:
                    finish = True
    out = []
    for n in L:
        best = n[4][0]
        for k in n[4][1:]:
            if k[2] > best[2]:
                best = k
        out.append(best)
    return out


def main():
    from oracle import oracle_lib as O
    rng = np.random.default_rng(1)
    nbad = 0
    for trial in range(400):
        W = int(rng.integers(40, 1300)); Hh = int(rng.integers(40, 500))
        n = int(rng.integers(0, 3000))
        N = int(rng.integers(1, 600))
        if trial % 3 == 0:  # clustered
            cx = rng.integers(0, W, size=8); cy = rng.integers(0, Hh, size=8)
            sel = rng.integers(0, 8, size=n)
            xs = np.clip(cx[sel] + rng.normal(0, W / 20, n), 0, W - 1).astype(int)
            ys = np.clip(cy[sel] + rng.normal(0, Hh / 20, n), 0, Hh - 1).astype(int)
        else:
            xs = rng.integers(0, W, size=n); ys = rng.integers(0, Hh, size=n)
        pts = np.unique(np.stack([ys, xs], 1), axis=0)
        rng.shuffle(pts) if trial % 2 else None
        resp = rng.integers(7, 60, size=len(pts))
        keys = [(int(p[1]), int(p[0]), int(r)) for p, r in zip(pts, resp)]
        kin = np.zeros(len(keys), dtype=O.KP_DTYPE)
        kin["x"] = [k[0] for k in keys]; kin["y"] = [k[1] for k in keys]; kin["response"] = [k[2] for k in keys]
        ref = O.distribute(kin, 16, 16 + W, 16, 16 + Hh, N)
        got = distribute(keys, 16, 16 + W, 16, 16 + Hh, N)
        r = [(int(a["x"]), int(a["y"]), int(a["response"])) for a in ref]
        if r != got:
            nbad += 1
            print("MISMATCH trial", trial, W, Hh, len(keys), N, len(r), len(got))
    print("done, mismatches:", nbad)


if __name__ == "__main__":
    main()
