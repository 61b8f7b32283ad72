#!/usr/bin/env python
"""CPU prototype (numpy/scipy) used to choose the PCG preconditioner of the reduced pose system.

Linearises a synthetic graph at its initial state (same math as SURVEY.md Appendix B), forms the
matrix-free Schur operator S = Hpp - Hpl Hll^-1 Hpl^T and counts PCG iterations to a relative
preconditioned residual `tol` for
  * block-Jacobi  M^-1 = blockdiag(S)^-1                       (what round 1 started with)
  * two-level     M^-1 = blockdiag(S)^-1 + Z (Z^T S Z)^-1 Z^T   with Z = piecewise-constant
    (x, y, theta) indicator vectors over aggregates of K consecutive poses.
  python tools/precond_proto.py --workload c2 --scale 1.0 --K 16 32 64
"""
import argparse, os, sys, time
import numpy as np
import scipy.sparse as sp
import scipy.linalg as sl
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from toyslam_b200 import synth


def linearize(g):
    P, L = len(g["pose_id"]), len(g["lm_id"])
    x = g["pose_xyt"].astype(np.float64); lm = g["lm_xy"].astype(np.float64)
    c, s = np.cos(x[:, 2]), np.sin(x[:, 2])
    # ODOM
    i, j = g["od_ij"][:, 0].astype(np.int64), g["od_ij"][:, 1].astype(np.int64)
    Z = g["od_Z"].astype(np.float64).reshape(-1, 3, 3)
    Zi = np.linalg.inv(Z)
    Xi = np.zeros((len(i), 3, 3)); Xj = np.zeros((len(i), 3, 3))
    def mat(k):
        M = np.zeros((len(k), 3, 3)); M[:, 0, 0] = c[k]; M[:, 0, 1] = -s[k]; M[:, 0, 2] = x[k, 0]
        M[:, 1, 0] = s[k]; M[:, 1, 1] = c[k]; M[:, 1, 2] = x[k, 1]; M[:, 2, 2] = 1; return M
    D = Zi @ (np.linalg.inv(mat(i)) @ mat(j))
    e = np.stack([D[:, 0, 2], D[:, 1, 2], np.arctan2(D[:, 1, 0], D[:, 0, 0])], 1)
    w = g["od_w"].astype(np.float64)
    chi = np.sum(e * w * e, 1); hw = np.where(chi <= 2.25, 1.0, 1.5 / np.sqrt(np.maximum(chi, 1e-300)))
    W = w * hw[:, None]
    Hd = np.zeros((P, 3, 3)); bp = np.zeros((P, 3))
    for q in range(3):
        np.add.at(Hd[:, q, q], i, W[:, q]); np.add.at(Hd[:, q, q], j, W[:, q])
        np.add.at(bp[:, q], i, W[:, q] * e[:, q]); np.add.at(bp[:, q], j, -W[:, q] * e[:, q])
    # off-diagonal ODOM blocks: -W (diag)
    rows = (3 * i[:, None] + np.arange(3)).ravel(); cols = (3 * j[:, None] + np.arange(3)).ravel()
    Hoff = sp.coo_matrix((-W.ravel(), (rows, cols)), shape=(3 * P, 3 * P)).tocsr()
    # LM
    pi = g["lm_ij"][:, 0].astype(np.int64); li = g["lm_ij"][:, 1].astype(np.int64) - P
    m = g["lm_meas"].astype(np.float64); z = np.stack([m[:, 0] * np.cos(m[:, 1]), m[:, 0] * np.sin(m[:, 1])], 1)
    cc, ss, px, py = c[pi], s[pi], x[pi, 0], x[pi, 1]; lx, ly = lm[li, 0], lm[li, 1]
    e0 = cc * lx + ss * ly - (cc * px + ss * py) - z[:, 0]
    e1 = -ss * lx + cc * ly + (ss * px - cc * py) - z[:, 1]
    q0 = cc * ly - ss * lx - cc * py + ss * px; q1 = -ss * ly - cc * lx + ss * py + cc * px
    wl = g["lm_w"].astype(np.float64)
    chi = e0 * wl[:, 0] * e0 + e1 * wl[:, 1] * e1
    hw = np.where(chi <= 2.25, 1.0, 1.5 / np.sqrt(np.maximum(chi, 1e-300)))
    i0, i1 = wl[:, 0] * hw, wl[:, 1] * hw
    A = np.zeros((len(pi), 2, 3)); A[:, 0, 0] = -cc; A[:, 0, 1] = -ss; A[:, 0, 2] = q0; A[:, 1, 0] = ss; A[:, 1, 1] = -cc; A[:, 1, 2] = q1
    B = np.zeros((len(pi), 2, 2)); B[:, 0, 0] = cc; B[:, 0, 1] = ss; B[:, 1, 0] = -ss; B[:, 1, 1] = cc
    Wm = np.zeros((len(pi), 2, 2)); Wm[:, 0, 0] = i0; Wm[:, 1, 1] = i1
    ev = np.stack([e0, e1], 1)[:, :, None]
    AtWA = A.transpose(0, 2, 1) @ Wm @ A; BtWB = B.transpose(0, 2, 1) @ Wm @ B; AtWB = A.transpose(0, 2, 1) @ Wm @ B
    np.add.at(Hd, pi, AtWA)
    np.add.at(bp, pi, -(A.transpose(0, 2, 1) @ Wm @ ev)[:, :, 0])
    Hll = np.zeros((L, 2, 2)); bl = np.zeros((L, 2))
    np.add.at(Hll, li, BtWB); np.add.at(bl, li, -(B.transpose(0, 2, 1) @ Wm @ ev)[:, :, 0])
    Hd[0] += 1e6 * np.eye(3)
    rows = (3 * pi[:, None, None] + np.arange(3)[None, :, None] + 0 * np.arange(2)[None, None, :]).ravel()
    cols = (2 * li[:, None, None] + 0 * np.arange(3)[None, :, None] + np.arange(2)[None, None, :]).ravel()
    Hpl = sp.coo_matrix((AtWB.ravel(), (rows, cols)), shape=(3 * P, 2 * L)).tocsr()
    Hpp = sp.block_diag([sp.csr_matrix(b) for b in []]) if False else None
    # block diagonal as sparse
    r = (3 * np.arange(P)[:, None, None] + np.arange(3)[None, :, None] + 0 * np.arange(3)[None, None, :]).ravel()
    cidx = (3 * np.arange(P)[:, None, None] + 0 * np.arange(3)[None, :, None] + np.arange(3)[None, None, :]).ravel()
    Hpp = sp.coo_matrix((Hd.ravel(), (r, cidx)), shape=(3 * P, 3 * P)).tocsr() + Hoff + Hoff.T
    return Hpp, Hpl, Hll, bp.ravel(), bl, Hd


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c2"); ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--K", type=int, nargs="*", default=[16, 32, 64]); ap.add_argument("--tol", type=float, default=1e-12)
    ap.add_argument("--schwarz", action="store_true"); ap.add_argument("--rigid", action="store_true")
    ap.add_argument("--mult", action="store_true", help="also try the multiplicative (symmetric) combination")
    a = ap.parse_args()
    g = synth.make_config(a.workload, scale=a.scale)
    t0 = time.time()
    Hpp, Hpl, Hll, bp, bl, Hd = linearize(g)
    P, L = Hd.shape[0], Hll.shape[0]
    Hinv = np.linalg.inv(Hll)
    r_ = (2 * np.arange(L)[:, None, None] + np.arange(2)[None, :, None] + 0 * np.arange(2)[None, None, :]).ravel()
    c_ = (2 * np.arange(L)[:, None, None] + 0 * np.arange(2)[None, :, None] + np.arange(2)[None, None, :]).ravel()
    HinvS = sp.coo_matrix((Hinv.ravel(), (r_, c_)), shape=(2 * L, 2 * L)).tocsr()
    T = (Hpl @ HinvS).tocsr()
    HplT = Hpl.T.tocsr()
    Sop = lambda v: Hpp @ v - T @ (HplT @ v)
    rhs = bp - T @ bl.ravel()
    # blockdiag(S)
    Sd = Hd.copy()
    TH = T.multiply(Hpl).tocsr()      # not the block product; compute block diag explicitly instead
    pi = g["lm_ij"][:, 0].astype(np.int64); li = g["lm_ij"][:, 1].astype(np.int64) - P
    Hpl_b = np.zeros((len(pi), 3, 2))
    for rr in range(3):
        for cc in range(2):
            Hpl_b[:, rr, cc] = np.asarray(Hpl[3 * pi + rr, 2 * li + cc]).ravel()
    np.add.at(Sd, pi, -(Hpl_b @ Hinv[li] @ Hpl_b.transpose(0, 2, 1)))
    Minv = np.linalg.inv(Sd)
    jac = lambda r: (Minv @ r.reshape(P, 3, 1)).reshape(-1)
    print(f"P={P} L={L} El={len(pi)} setup {time.time() - t0:.1f}s")

    def pcg(apply_M, label):
        x = np.zeros(3 * P); r = rhs.copy(); z = apply_M(r); p = z.copy(); rz = r @ z; rz0 = rz; it = 0
        t0 = time.time()
        while it < 20000:
            Sp = Sop(p); alpha = rz / (p @ Sp); x += alpha * p; r -= alpha * Sp; z = apply_M(r); rzn = r @ z; it += 1
            if not (rzn > a.tol ** 2 * rz0): break
            p = z + (rzn / rz) * p; rz = rzn
        print(f"{label:40s} iters {it:6d}  true rel res {np.linalg.norm(rhs - Sop(x)) / np.linalg.norm(rhs):.2e}  ({time.time() - t0:.1f}s)")
        return x

    x0 = pcg(jac, "block-Jacobi")
    for K in a.K:
        nA = (P + K - 1) // K
        agg = np.arange(P) // K
        Z = sp.coo_matrix((np.ones(3 * P), (np.arange(3 * P), 3 * np.repeat(agg, 3) + np.tile(np.arange(3), P))), shape=(3 * P, 3 * nA)).tocsr()
        if a.rigid:
            # theta column = rigid rotation of the aggregate about its centroid (a near-null vector of S)
            xy = g["pose_xyt"].astype(np.float64)[:, :2]
            cen = np.stack([np.bincount(agg, xy[:, 0]), np.bincount(agg, xy[:, 1])], 1) / np.bincount(agg)[:, None]
            d = xy - cen[agg]
            rws = np.concatenate([np.arange(3 * P), 3 * np.arange(P), 3 * np.arange(P) + 1])
            cls = np.concatenate([3 * np.repeat(agg, 3) + np.tile(np.arange(3), P), 3 * agg + 2, 3 * agg + 2])
            vls = np.concatenate([np.ones(3 * P), -d[:, 1], d[:, 0]])
            Z = sp.coo_matrix((vls, (rws, cls)), shape=(3 * P, 3 * nA)).tocsr()
        ZtHplT = (HplT @ Z).tocsr()                         # 2L x 3nA
        Ac = (Z.T @ Hpp @ Z - ZtHplT.T @ HinvS @ ZtHplT).toarray()
        cf = sl.cho_factor(Ac)
        two = lambda r: jac(r) + Z @ sl.cho_solve(cf, Z.T @ r)
        x1 = pcg(two, f"two-level additive K={K} (coarse dim {3 * nA})")
        if a.mult:
            def mult(r):
                z1 = jac(r); r1 = r - Sop(z1); z2 = z1 + Z @ sl.cho_solve(cf, Z.T @ r1); r2 = r - Sop(z2); return z2 + jac(r2)
            pcg(mult, f"two-level multiplicative K={K}")
        print("   max |dx| vs block-Jacobi solution:", np.abs(x1 - x0).max())
        if a.schwarz:
            # exact aggregate blocks S_aa (non-overlapping additive Schwarz) as the fine-level part
            t0 = time.time()
            Tc = T.tocsr(); HT = HplT.tocsc(); Hppc = Hpp.tocsr()
            inv_blocks = []
            for q in range(nA):
                lo, hi = 3 * q * K, min(3 * (q + 1) * K, 3 * P)
                Saa = Hppc[lo:hi, lo:hi].toarray() - (Tc[lo:hi, :] @ HT[:, lo:hi]).toarray()
                inv_blocks.append(np.linalg.inv(Saa))
            def schw(r):
                z = np.empty_like(r)
                for q in range(nA):
                    lo, hi = 3 * q * K, min(3 * (q + 1) * K, 3 * P)
                    z[lo:hi] = inv_blocks[q] @ r[lo:hi]
                return z
            print(f"   aggregate blocks built in {time.time() - t0:.1f}s")
            pcg(schw, f"aggregate-block Jacobi K={K}")
            pcg(lambda r: schw(r) + Z @ sl.cho_solve(cf, Z.T @ r), f"two-level, exact aggregate blocks K={K}")


if __name__ == "__main__":
    main()
