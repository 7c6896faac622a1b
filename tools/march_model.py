#!/usr/bin/env python
"""Numpy model of applyMarchKernel (ndr_b200/csrc/apply_march.cuh): the element-marching Walsh-Hadamard K(rho)u.

Development aid: it executes the kernel's algorithm CTA by CTA with arrays shaped (RY, ZT) standing for the threads
of one block, np.roll for warp shuffles and explicit arrays for shared memory, so that the index / sign / ghost
logic can be checked on the CPU against a dense K0 element loop before the CUDA version goes to a GPU.
Run:  python tools/march_model.py
"""
import itertools
import sys

import numpy as np


def k0_hex(h, lam, mu):
    """Trilinear hex stiffness for unit modulus (same formula as Sim<3>::updateK0)."""
    gp = [0.5 - 0.5 / np.sqrt(3.0), 0.5 + 0.5 / np.sqrt(3.0)]
    K = np.zeros((24, 24))
    vol = h[0] * h[1] * h[2]
    for q in itertools.product(range(2), repeat=3):
        xi = [gp[i] for i in q]
        grad = np.zeros((8, 3))
        for p in range(8):
            bits = [(p >> (2 - d)) & 1 for d in range(3)]
            for a in range(3):
                v = 1.0
                for d in range(3):
                    v *= ((1.0 if bits[d] else -1.0) / h[d]) if d == a else (xi[d] if bits[d] else 1.0 - xi[d])
                grad[p, a] = v
        for p in range(8):
            for r in range(8):
                gg = grad[p] @ grad[r]
                for a in range(3):
                    for b in range(3):
                        K[p * 3 + a, r * 3 + b] += 0.125 * vol * (lam * grad[p, a] * grad[r, b] + mu * grad[p, b] * grad[r, a] + (mu * gg if a == b else 0.0))
    return 0.5 * (K + K.T)


def khat_blocks(K):
    """B[g][c][c'] = Khat[(g ^ e_c, c), (g ^ e_c', c')] with Khat = W K W^T / 64; also returns max off-block entry."""
    W = np.zeros((24, 24))
    for m in range(8):
        for j in range(8):
            s = (-1.0) ** bin(m & j).count("1")
            for c in range(3):
                W[m * 3 + c, j * 3 + c] = s
    Kh = W @ K @ W.T / 64.0
    e = [4, 2, 1]
    B = np.zeros((8, 3, 3))
    used = np.zeros((24, 24), bool)
    for g in range(8):
        for c in range(3):
            for c2 in range(3):
                a, b = (g ^ e[c]) * 3 + c, (g ^ e[c2]) * 3 + c2
                B[g, c, c2] = Kh[a, b]; used[a, b] = True
    return B, np.abs(Kh[~used]).max()


def reference_apply(E, K, u, Ey):
    """Dense element loop: out = sum_e Ey_e scatter(K0 gather(u))."""
    out = np.zeros_like(u)
    for x, y, z in itertools.product(range(E[0]), range(E[1]), range(E[2])):
        ue = np.concatenate([u[x + jx, y + jy, z + jz] for jx, jy, jz in itertools.product(range(2), repeat=3)])
        fe = Ey[x, y, z] * (K @ ue)
        for j, (jx, jy, jz) in enumerate(itertools.product(range(2), repeat=3)):
            out[x + jx, y + jy, z + jz] += fe[3 * j:3 * j + 3]
    return out


def march_apply(E, B, u, Ey, RY=4, ZT=64, XC=3, sparse=True):
    nx, ny, nz = E[0] + 1, E[1] + 1, E[2] + 1
    NT = -(-RY * ZT // 32) * 32
    out = np.full_like(u, np.nan)
    e_bit = [4, 2, 1]
    def ntiles(ne, t):  # tiles of t columns overlapping by one: 1 + ceil(max(0, ne - t) / (t - 1))
        return 1 + -(-max(0, ne - t) // (t - 1))
    ntz, nty, nch = ntiles(E[2], ZT), ntiles(E[1], RY), ntiles(E[0], XC)
    ry = np.arange(RY)[:, None] + np.zeros((1, ZT), int)
    tx = np.arange(ZT)[None, :] + np.zeros((RY, 1), int)
    lane = (ry * ZT + tx) & 31           # rows straddle warps when ZT is not a multiple of 32
    next_in_warp = (lane < 31) & (tx + 1 < ZT)
    prev_in_warp = (lane > 0) & (tx > 0)
    XS = ZT // 32 + 2

    def ld(p, yy, zz):
        """predicated node load: zeros outside the grid"""
        ok = (yy < ny) & (zz < nz) & (p < nx)
        v = np.zeros(yy.shape + (3,))
        v[ok] = u[p, yy[ok], zz[ok]]
        return v

    def _flat(a):
        f = np.full((NT,) + a.shape[2:], 1e300)  # idle threads / garbage
        f[:RY * ZT] = a.reshape((RY * ZT,) + a.shape[2:])
        return f.reshape((NT // 32, 32) + a.shape[2:])

    def shfl_down(a):  # lane i gets lane i+1 of its warp; lane 31 keeps its own (garbage for the algorithm)
        f = _flat(a); r = f.copy(); r[:, :31] = f[:, 1:]
        return r.reshape((NT,) + a.shape[2:])[:RY * ZT].reshape(a.shape)

    def shfl_up(a):
        f = _flat(a); r = f.copy(); r[:, 1:] = f[:, :31]
        return r.reshape((NT,) + a.shape[2:])[:RY * ZT].reshape(a.shape)

    for cz, cy, cx in itertools.product(range(ntz), range(nty), range(nch)):
        z0, y0, x0 = cz * (ZT - 1), cy * (RY - 1), cx * (XC - 1)
        x1 = min(x0 + XC, E[0])
        y, z = y0 + ry, z0 + tx
        col_ok = (y < ny - 1) & (z < nz - 1)
        last31 = ~next_in_warp

        def faces(p):
            a, b = ld(p, y, z), ld(p, y + 1, z)
            an, bn = shfl_down(a), shfl_down(b)
            an[last31] = ld(p, y, z + 1)[last31]; bn[last31] = ld(p, y + 1, z + 1)[last31]
            zs0, zd0, zs1, zd1 = a + an, a - an, b + bn, b - bn
            F = np.zeros((RY, ZT, 2, 2, 3))  # [my][mz][c]
            F[:, :, 0, 0], F[:, :, 1, 0], F[:, :, 0, 1], F[:, :, 1, 1] = zs0 + zs1, zs0 - zs1, zd0 + zd1, zd0 - zd1
            return F

        Fprev = faces(x0)
        Gprev = np.zeros((RY, ZT, 2, 2, 3))
        n_iter = (x1 - x0) + (1 if x1 == E[0] else 0)
        for it in range(n_iter):
            x = x0 + it
            if x < E[0]:
                Fnew = faces(x + 1)
                Ee = np.zeros((RY, ZT)); Ee[col_ok] = Ey[x, y[col_ok], z[col_ok]]
                uh = np.zeros((RY, ZT, 8, 3))
                for my, mz in itertools.product(range(2), repeat=2):
                    uh[:, :, 0 * 4 + my * 2 + mz] = Fprev[:, :, my, mz] + Fnew[:, :, my, mz]
                    uh[:, :, 1 * 4 + my * 2 + mz] = Fprev[:, :, my, mz] - Fnew[:, :, my, mz]
                yh = np.zeros((RY, ZT, 8, 3))
                for g in range(8):
                    for c in range(3):
                        for c2 in range(3):
                            if sparse and abs(B[g, c, c2]) < 1e-14 * np.abs(B).max():
                                continue
                            yh[:, :, g ^ e_bit[c], c] += B[g, c, c2] * uh[:, :, g ^ e_bit[c2], c2]
                G0 = yh[:, :, 0:4] + yh[:, :, 4:8]  # [my*2+mz]
                G1 = yh[:, :, 0:4] - yh[:, :, 4:8]
                Gtot = Gprev + Ee[:, :, None, None, None] * G0.reshape(RY, ZT, 2, 2, 3)
                Gprev = Ee[:, :, None, None, None] * G1.reshape(RY, ZT, 2, 2, 3)
                Fprev = Fnew
            else:
                Gtot = Gprev
            # ---- backward for plane x: z stage (shuffles) ----
            zl = Gtot[:, :, :, 0] + Gtot[:, :, :, 1]  # [my][c] part of node z
            zh = Gtot[:, :, :, 0] - Gtot[:, :, :, 1]  # part of node z+1
            r = shfl_up(zh); r[~prev_in_warp] = 0.0
            X = np.full((RY, XS, 2, 3), np.nan)        # smem: zh of a thread whose z+1 neighbour is in the next warp
            x_send = ~next_in_warp & (tx + 1 < ZT)
            x_recv = ~prev_in_warp & (tx > 0)
            for a, bq in np.argwhere(x_send):
                X[a, (bq + 1) >> 5] = zh[a, bq]
            Ye = zl + r
            yl = Ye[:, :, 0] + Ye[:, :, 1]
            yhh = Ye[:, :, 0] - Ye[:, :, 1]
            Yx = np.zeros((RY + 1, ZT, 3)); Yx[1:] = yhh   # smem: row ry writes slot ry+1
            # special column: node z+1 == nz-1 owned by thread tx == ZT-1
            spz = (tx == ZT - 1) & (z + 1 == nz - 1)
            spl = zh[:, :, 0] + zh[:, :, 1]
            sph = zh[:, :, 0] - zh[:, :, 1]
            Ysp = np.zeros((RY + 1, 3)); Ysp[1:] = sph[:, ZT - 1]
            # ---- barrier ----
            res = yl + Yx[:RY]
            # cross-warp correction: the part the shuffle could not deliver, for this row and for the row below
            for a, bq in np.argwhere(x_recv):
                res[a, bq] += X[a, bq >> 5, 0] + X[a, bq >> 5, 1]
                if a + 1 < RY:
                    res[a + 1, bq] += X[a, bq >> 5, 0] - X[a, bq >> 5, 1]
            store_plane = (x > x0) or (cx == 0)
            ok = (y < ny) & (z < nz) & ((ry > 0) | (cy == 0)) & ((tx > 0) | (cz == 0))
            # tiles own output nodes up to their last thread row / lane; the next tile re-computes the shared one as ghost
            if store_plane:
                out[x, y[ok], z[ok]] = res[ok]
                sres = spl + Ysp[:RY, None, :]
                oks = spz & (y < ny) & ((ry > 0) | (cy == 0))
                out[x, y[oks], z[oks] + 1] = sres[oks]
                # special row: node row y+1 == ny-1 owned by thread row RY-1
                if y0 + RY == ny - 1:
                    top = yhh[RY - 1].copy()
                    for a, bq in np.argwhere(x_recv):
                        if a == RY - 1:
                            top[bq] += X[a, bq >> 5, 0] - X[a, bq >> 5, 1]
                    zz = z[RY - 1]
                    okt = (zz < nz) & ((tx[RY - 1] > 0) | (cz == 0))
                    out[x, ny - 1, zz[okt]] = top[okt]
                    if spz[RY - 1, ZT - 1]:
                        out[x, ny - 1, nz - 1] = sph[RY - 1, ZT - 1]
    return out


def main():
    rng = np.random.default_rng(0)
    lam, mu = 0.3 / (1.3 * 0.4), 1.0 / 2.6
    worst = 0.0
    for E, h, RY, ZT, XC in (([5, 4, 6], [1 / 4, 1 / 4, 1 / 4], 4, 32, 3), ([4, 7, 40], [1 / 4, 1 / 4, 1 / 8], 3, 32, 4),
                             ([3, 6, 64], [1.0, 0.5, 0.25], 4, 64, 2), ([6, 9, 70], [1.0, 1.0, 1.0], 4, 64, 3),
                             ([2, 3, 31], [1.0, 1.0, 1.0], 4, 32, 5), ([7, 8, 32], [1.0, 2.0, 1.0], 3, 32, 3),
                             ([4, 2, 33], [1.0, 2.0, 1.0], 3, 32, 3), ([3, 3, 63], [1.0, 2.0, 1.0], 4, 64, 9),
                             ([3, 9, 31], [1.0, 2.0, 1.0], 7, 65, 2), ([3, 16, 70], [1.0, 1.0, 1.0], 5, 33, 3),
                             ([2, 11, 65], [1.0, 1.0, 1.0], 6, 65, 2), ([2, 30, 40], [1.0, 1.0, 1.0], 15, 33, 4)):
        K = k0_hex(h, lam, mu)
        B, off = khat_blocks(K)
        assert off < 1e-15 * np.abs(K).max(), off
        u = rng.standard_normal((E[0] + 1, E[1] + 1, E[2] + 1, 3))
        Ey = rng.uniform(0.01, 1, E)
        ref = reference_apply(E, K, u, Ey)
        got = march_apply(E, B, u, Ey, RY, ZT, XC)
        assert not np.isnan(got).any(), ("unwritten outputs", E, np.argwhere(np.isnan(got))[:5])
        err = np.abs(got - ref).max() / np.abs(ref).max()
        worst = max(worst, err)
        print(E, "RY", RY, "ZT", ZT, "XC", XC, "rel err %.2e" % err)
    assert worst < 1e-13
    print("OK")


if __name__ == "__main__":
    main()
