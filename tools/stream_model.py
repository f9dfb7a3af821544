#!/usr/bin/env python3
"""Dev helper: lane-level numpy model of k_ibp_stream (csrc/ibp_stream.cu), checked against the oracle.

Transcribes the kernel's per-warp streaming algorithm literally (segments, strips, step/row indexing,
shuffles, lane constants, edge vectors, GEN/LEFT/top logic) in float64, so indexing mistakes show up
on the CPU before a GPU run.  Not part of the product or the test suite.
"""
import sys, os
import numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(__file__), ".."))
from oracle import sr_oracle as O

VALID = 112


def shfl_up(x):   # lane l gets lane l-1 (lane 0 keeps its own)
    return np.concatenate([x[:1], x[:-1]])


def shfl_down(x):
    return np.concatenate([x[1:], x[-1:]])


def row_pass(Q, b, C):  # C: [32,4]
    Cx, Cy, Cz, Cw = C[:, 0], C[:, 1], C[:, 2], C[:, 3]
    if Q == 0:
        f = [shfl_up(Cy), shfl_up(Cz), shfl_up(Cw), Cx, Cy, Cz, Cw, shfl_down(Cx), shfl_down(Cy)]
    else:
        f = [shfl_up(Cz), shfl_up(Cw), Cx, Cy, Cz, Cw, shfl_down(Cx), shfl_down(Cy), shfl_down(Cz)]
    B1 = sum(b[t] * f[6 - t] for t in range(7))
    B2 = sum(b[t] * f[8 - t] for t in range(7))
    return B1, B2


def back_row(Q, b, r1, r2):
    if Q == 0:
        L2, R1, R2 = shfl_up(r2), shfl_down(r1), shfl_down(r2)
        Wx = b[1] * L2 + b[3] * r1 + b[5] * r2
        Wy = b[0] * L2 + b[2] * r1 + b[4] * r2 + b[6] * R1
        Wz = b[1] * r1 + b[3] * r2 + b[5] * R1
        Ww = b[0] * r1 + b[2] * r2 + b[4] * R1 + b[6] * R2
    else:
        L1, L2, R1 = shfl_up(r1), shfl_up(r2), shfl_down(r1)
        Wx = b[0] * L1 + b[2] * L2 + b[4] * r1 + b[6] * r2
        Wy = b[1] * L2 + b[3] * r1 + b[5] * r2
        Wz = b[0] * L2 + b[2] * r1 + b[4] * r2 + b[6] * R1
        Ww = b[1] * r1 + b[3] * r2 + b[5] * R1
    return np.stack([Wx, Wy, Wz, Ww], 1)


def model_iteration(lrs, ints, psf, hr, step=0.5, lo=0.0, hi=255.0, nit=2):
    n = len(lrs); h, w = lrs[0].shape; H, W = 2 * h, 2 * w
    P, Q = ints[0][0] & 1, ints[0][1] & 1
    assert all((sy & 1, sx & 1) == (P, Q) for sy, sx in ints)
    # rank-1 factors as the plan builds them
    pu, pv = np.unravel_index(np.argmax(np.abs(psf)), psf.shape)
    acol = psf[:, pv].copy(); brow = psf[pu, :] / psf[pu, pv]
    oy = [(P + sy) >> 1 for sy, sx in ints]; ox = [(Q + sx) >> 1 for sy, sx in ints]
    # S1 plane and vsum (k_lattice_sums)
    S1 = np.zeros((h, w)); vsum = 0.0
    mtab = np.zeros(4)
    for ra in range(2):
        for cb in range(2):
            mtab[ra * 2 + cb] = sum(1 for k in range(n) if (not ra or oy[k] == 0) and (not cb or ox[k] == 0))
    for a in range(h):
        for b in range(w):
            xs = [lrs[k][a + oy[k], b + ox[k]] for k in range(n) if a + oy[k] < h and b + ox[k] < w]
            if xs:
                S1[a, b] = sum(xs); vsum += sum((x - np.mean(xs)) ** 2 for x in xs)
    # edge sums (k_edge_sums)
    colS1 = np.zeros(h); leftS1 = np.zeros(h); rowS1 = np.zeros(w); topS1 = np.zeros(w)
    for a in range(h):
        xs = []
        for k in range(n):
            i = a + oy[k]
            if i >= h: continue
            if ints[k][1] == -1: colS1[a] += lrs[k][i, 0]
            if ints[k][1] == +1: xs.append(lrs[k][i, 0])
        if xs: leftS1[a] = sum(xs); vsum += sum((x - np.mean(xs)) ** 2 for x in xs)
    for b in range(w):
        xs = []
        for k in range(n):
            j = b + ox[k]
            if j >= w: continue
            if ints[k][0] == -1: rowS1[b] += lrs[k][0, j]
            if ints[k][0] == +1: xs.append(lrs[k][0, j])
        if xs: topS1[b] = sum(xs); vsum += sum((x - np.mean(xs)) ** 2 for x in xs)
    scal0 = sum(lrs[k][0, 0] for k in range(n) if ints[k] == (-1, -1))
    xs = [lrs[k][0, 0] for k in range(n) if ints[k] == (1, 1)]
    scal1 = sum(xs)
    if xs: vsum += sum((x - np.mean(xs)) ** 2 for x in xs)
    mc = [sum(1 for k in range(n) if ints[k][1] == -1), sum(1 for k in range(n) if ints[k][1] == -1 and oy[k] == 0)]
    ml = [sum(1 for k in range(n) if ints[k][1] == +1), sum(1 for k in range(n) if ints[k][1] == +1 and oy[k] == 0)]
    mr = [sum(1 for k in range(n) if ints[k][0] == -1), sum(1 for k in range(n) if ints[k][0] == -1 and ox[k] == 0)]
    mt = [sum(1 for k in range(n) if ints[k][0] == +1), sum(1 for k in range(n) if ints[k][0] == +1 and ox[k] == 0)]
    m_dupc = sum(1 for k in range(n) if ints[k] == (-1, -1)); m_msec = sum(1 for k in range(n) if ints[k] == (1, 1))
    alpha = step / n

    def hrz(y, x0, ncol):   # zero-filled row fetch (TMA)
        out = np.zeros(ncol)
        if 0 <= y < H:
            xa, xb = max(x0, 0), min(x0 + ncol, W)
            if xb > xa: out[xa - x0:xb - x0] = hr[y, xa:xb]
        return out

    def s1z(a, b0, ncol):
        out = np.zeros(ncol)
        if 0 <= a < h:
            ba, bb = max(b0, 0), min(b0 + ncol, w)
            if bb > ba: out[ba - b0:bb - b0] = S1[a, ba:bb]
        return out

    seg_rows = 8 * nit - 6
    nseg = (H + seg_rows - 1) // seg_rows; nstrip = (W + VALID - 1) // VALID
    out = np.full((H, W), np.nan); e2tot = 0.0
    lane = np.arange(32)
    for seg in range(nseg):
        for strip in range(nstrip):
            x0 = strip * VALID; ya = seg * seg_rows; yb = min(ya + seg_rows, H)
            top = (P == 1) and ya == 0
            ys0 = ya - 3 + (1 - P); a0 = (ys0 - P) >> 1; xs0 = x0 - 8
            b1 = (xs0 >> 1) + 2 * lane; b2 = b1 + 1
            v1 = (b1 >= 0) & (b1 < w); v2 = (b2 >= 0) & (b2 < w)
            xs1 = 2 * b1 + Q; xs2 = 2 * b2 + Q
            o1 = v1 & (xs1 >= x0) & (xs1 < x0 + VALID); o2 = v2 & (xs2 >= x0) & (xs2 < x0 + VALID)
            l1 = (b1 == w - 1).astype(int); l2 = (b2 == w - 1).astype(int)
            mI1 = np.where(v1, mtab[l1], 0); mI2 = np.where(v2, mtab[l2], 0)
            mL1 = np.where(v1, mtab[2 + l1], 0); mL2 = np.where(v2, mtab[2 + l2], 0)
            inv = lambda o, m: np.where(o & (m > 0), 1.0 / np.maximum(m, 1e-30), 0.0)
            wI1, wI2, wL1, wL2 = inv(o1, mI1), inv(o2, mI2), inv(o1, mL1), inv(o2, mL2)
            left = (Q == 1) and x0 == 0
            lb = np.zeros((32, 4))
            if left: lb[2] = [brow[3], brow[2], brow[1], brow[0]]
            gx = xs0 + 4 * lane
            st_ok = (lane >= 2) & (lane < 30) & (gx < W)
            wr = [np.zeros((32, 4)) for _ in range(4)]; w0 = np.zeros((32, 4)); e2 = np.zeros(32)
            for j in range(4 * nit):
                T = j & 3; aj = a0 + j; ys = 2 * aj + P
                win = [hrz(ys + 3 - u, xs0, 128).reshape(32, 4) for u in range(7)]
                C = sum(acol[u] * win[u] for u in range(7))
                h1, h2 = win[6], win[5]
                B1, B2 = row_pass(Q, brow, C)
                s1 = s1z(aj, xs0 >> 1, 64).reshape(32, 2)
                valid = 0 <= aj < h; last = aj == h - 1; own = valid and ya <= ys < yb
                m1 = (mL1 if last else mI1) if valid else 0 * mI1; m2 = (mL2 if last else mI2) if valid else 0 * mI2
                w1 = (wL1 if last else wI1) if own else 0 * wI1; w2 = (wL2 if last else wI2) if own else 0 * wI2
                r1 = s1[:, 0] - m1 * B1; r2 = s1[:, 1] - m2 * B2
                e2 += r1 * w1 * r1 + r2 * w2 * r2
                Wv = back_row(Q, brow, r1, r2)
                if left:
                    cS = colS1[aj] if 0 <= aj < h else 0.0; lS = leftS1[aj] if 0 <= aj < h else 0.0
                    mcv = (mc[1] if last else mc[0]) if valid else 0; mlv = (ml[1] if last else ml[0]) if valid else 0
                    dcol = cS - mcv * B1
                    Wv = Wv + lb * dcol[:, None]
                    Bc0 = brow[0] * C[:, 3] + brow[1] * C[:, 2] + brow[2] * C[:, 1] + brow[3] * C[:, 0]
                    rl = lS - mlv * Bc0
                    wl = np.where((lane == 2) & own & (mlv > 0), 1.0 / max(mlv, 1e-30), 0.0)
                    e2 += rl * wl * rl
                wr[T] = Wv
                if top and aj == 0:
                    C0 = acol[0] * win[1] + acol[1] * win[2] + acol[2] * win[3] + acol[3] * win[4]
                    B01, B02 = row_pass(Q, brow, C0)
                    tS1 = np.where(v1, topS1[np.clip(b1, 0, w - 1)], 0); tS2 = np.where(v2, topS1[np.clip(b2, 0, w - 1)], 0)
                    rS1 = np.where(v1, rowS1[np.clip(b1, 0, w - 1)], 0); rS2 = np.where(v2, rowS1[np.clip(b2, 0, w - 1)], 0)
                    mt1 = np.array(mt)[l1]; mt2 = np.array(mt)[l2]
                    mr1 = np.where(v1, np.array(mr)[l1], 0); mr2 = np.where(v2, np.array(mr)[l2], 0)
                    e2 += np.where(o1 & (mt1 > 0), (tS1 - mt1 * B01) ** 2 / np.maximum(mt1, 1e-30), 0)
                    e2 += np.where(o2 & (mt2 > 0), (tS2 - mt2 * B02) ** 2 / np.maximum(mt2, 1e-30), 0)
                    if left and m_msec > 0:
                        B00 = brow[0] * C0[2, 3] + brow[1] * C0[2, 2] + brow[2] * C0[2, 1] + brow[3] * C0[2, 0]
                        e2[2] += (scal1 - m_msec * B00) ** 2 / m_msec
                    d1 = rS1 - mr1 * B1; d2 = rS2 - mr2 * B2
                    V = back_row(Q, brow, d1, d2)
                    if left:
                        d00 = scal0 - m_dupc * B1
                        V = V + lb * d00[:, None]
                    w0 = V
                wa, wb, wc, wd = wr[(T + 1) & 3], wr[(T + 2) & 3], wr[(T + 3) & 3], wr[T]
                q1 = acol[0] * wa + acol[2] * wb + acol[4] * wc + acol[6] * wd
                q2 = acol[1] * wb + acol[3] * wc + acol[5] * wd
                y1 = ys - 3
                if top:
                    if 0 <= y1 < 4: q1 = q1 + acol[3 - y1] * w0
                    if 0 <= y1 + 1 < 4: q2 = q2 + acol[2 - y1] * w0
                for (y, q, hh) in ((y1, q1, h1), (y1 + 1, q2, h2)):
                    if ya <= y < yb:
                        res = np.clip(hh + alpha * q, lo, hi)
                        for l in range(32):
                            if st_ok[l]:
                                assert np.isnan(out[y, gx[l]]), "written twice"
                                out[y, gx[l]:gx[l] + 4] = res[l]
            e2tot += e2.sum()
    assert not np.isnan(out).any(), "pixels never written"
    err = (e2tot + vsum) / (n * h * w)
    return out, err


def main():
    rng = np.random.default_rng(5)
    sets = [
        [(+0.5, -0.5), (+0.5, +0.5), (-0.5, -0.5), (-0.5, +0.5)],
        [(0.5, -0.5)] * 3 + [(-0.5, -0.5)] * 3,
        [(0.5, 0.5), (-0.5, -0.5)],
        [(-0.5, 0.5)], [(0.5, 0.5)], [(0.0, 0.0)], [(0.0, 0.5), (0.0, -0.5)], [(0.5, 0.0), (-0.5, 0.0), (-0.5, 0.0)],
    ]
    psf = O.make_gaussian_psf()
    worst = 0.0
    for shifts in sets:
        for (h, w) in [(20, 62), (7, 60), (2, 2), (33, 10)]:
            lrs = [rng.uniform(0, 255, (h, w)) for _ in shifts]
            ints = [(int(round(2 * dy)), int(round(2 * dx))) for dy, dx in shifts]
            hr0 = rng.uniform(-5, 260, (2 * h, 2 * w))
            ref, eref = O.ibp_integer_shifts(lrs, shifts, psf, hr0, n_iter=1)
            for nit in (2, 3):
                got, e = model_iteration(lrs, ints, psf, hr0, nit=nit)
                d = np.max(np.abs(got - ref)); de = abs(e / eref[0] - 1)
                worst = max(worst, d)
                flag = "" if (d < 1e-9 and de < 1e-12) else "   <-- MISMATCH"
                print(f"{str(shifts):70s} {h}x{w} nit={nit}: max|d|={d:.2e} mse rel={de:.2e}{flag}")
    print("worst", worst)


if __name__ == "__main__":
    main()
