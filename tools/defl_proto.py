#!/usr/bin/env python
"""Prototype: deflation of the slow modes of the BPX-preconditioned pose system with Ritz vectors recycled from
the previous Gauss-Newton iteration (round 2 experiment)."""
import argparse, os, sys, time
import numpy as np, scipy.sparse as sp, scipy.linalg as sl
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import mg_proto as mp
from toyslam_b200 import synth
from precond_proto import linearize

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="c3"); ap.add_argument("--scale", type=float, default=0.3)
ap.add_argument("--coarsest", type=int, default=130); ap.add_argument("--k", type=int, nargs="*", default=[10, 20, 40])
ap.add_argument("--lanczos", type=int, default=300); ap.add_argument("--gn", type=int, default=3)
a = ap.parse_args()
mp.LINEAR[0] = True
g = synth.make_config(a.workload, scale=a.scale)
tol = 1e-12


def setup(g):
    Hpp, Hpl, Hll, bp, bl, Hd = linearize(g)
    Hinv = np.linalg.inv(Hll)
    xy = g["pose_xyt"].astype(np.float64)[:, :2]
    l0 = mp.Level(Hpp, Hpl, Hinv, xy)
    rhs = bp - l0.Ccsr @ (l0.HinvS @ bl.ravel())
    levels = [l0]; Zs = []
    while levels[-1].n > a.coarsest:
        Z, lv = mp.coarsen(levels[-1], 4); lv.spacing = levels[-1].spacing * 4; Zs.append(Z); levels.append(lv)
    cf = sl.cho_factor(levels[-1].dense()); nl = len(levels)
    def M(r, l=0):
        if l == nl - 1: return sl.cho_solve(cf, r)
        return levels[l].jac(r) + Zs[l] @ M(Zs[l].T @ r, l + 1)
    return l0, rhs, M, (Hpl, Hinv, bl, bp)


def pcg(S, rhs, M, x0=None, maxit=5000, W=None, SW=None, G=None):
    """PCG; with W: deflated additive coarse term W G^-1 W^T added to M (G = W^T S W)"""
    def Md(r):
        z = M(r)
        if W is not None: z = z + W @ sl.cho_solve(G, W.T @ r)
        return z
    x = np.zeros_like(rhs) if x0 is None else x0.copy(); r = rhs - S(x) if x0 is not None else rhs.copy()
    z0 = M(rhs); rz0 = rhs @ z0
    z = Md(r); p = z.copy(); rz = r @ z; it = 0
    while it < maxit:
        Sp = S(p); al = rz / (p @ Sp); x += al * p; r -= al * Sp; z = Md(r); rzn = r @ z; it += 1
        if not (r @ M(r) > tol ** 2 * rz0): break
        p = z + (rzn / rz) * p; rz = rzn
    return x, it


def ritz(S, M, n, m, k):
    """m steps of Lanczos on M^-1 S (S-inner product ... done in the M-inner product of the symmetric pencil), full reorth"""
    rng = np.random.default_rng(1)
    # Lanczos for the pencil (S, M^-1): vectors v with v^T M^-1... use preconditioned Lanczos: q_j M-orthonormal residual-like
    r = rng.normal(size=n); z = M(r); beta = np.sqrt(r @ z); R = []; Zv = []
    T = np.zeros((m, m))
    for j in range(m):
        rj = r / beta; zj = z / beta; R.append(rj); Zv.append(zj)
        w = S(zj)
        alpha = w @ zj; T[j, j] = alpha
        w = w - alpha * rj - (beta_prev * R[j - 1] if j > 0 else 0)
        # full reorthogonalisation (w against R in the M-inner product: w^T Z_i)
        for _ in range(2):
            for i in range(j + 1): w -= (w @ Zv[i]) * R[i]
        r = w; z = M(r); beta_prev = np.sqrt(max(r @ z, 1e-300)); beta = beta_prev
        if j + 1 < m: T[j, j + 1] = T[j + 1, j] = beta
    ev, Y = np.linalg.eigh(T)
    Wm = np.stack(Zv, 1) @ Y[:, :k]          # Ritz vectors (in z-space): approximate eigenvectors of M^-1 S
    return ev, Wm


l0, rhs, M, _ = setup(g)
n = len(rhs)
x, it = pcg(l0.S, rhs, M); print(f"GN 0: plain BPX-PCG {it} iterations", flush=True)
t0 = time.time(); ev, Wall = ritz(l0.S, M, n, a.lanczos, max(a.k)); print(f"Lanczos {a.lanczos} steps {time.time() - t0:.0f}s: lowest Ritz values {ev[:6]}, 20th {ev[19]:.3g} 40th {ev[39]:.3g} max {ev[-1]:.3g}", flush=True)
for k in a.k:
    W = Wall[:, :k]; SW = np.stack([l0.S(W[:, j]) for j in range(k)], 1); G = sl.cho_factor(W.T @ SW)
    _, itd = pcg(l0.S, rhs, M, W=W, G=G)
    x0 = W @ sl.cho_solve(G, W.T @ rhs); _, iti = pcg(l0.S, rhs, M, x0=x0)
    print(f"   k={k:3d}: deflated preconditioner {itd} its; init-CG only {iti} its", flush=True)
# next GN iterations: state update with the exact solve, old W reused
Wold = Wall
for gn in range(1, a.gn + 1):
    # apply the damped update (poses and landmarks)
    Hpp, Hpl, Hll, bp, bl, Hd = linearize(g)
    Hinv = np.linalg.inv(Hll)
    dl = np.einsum("lij,lj->li", Hinv, bl - (Hpl.T @ x).reshape(-1, 2))
    g["pose_xyt"] = (g["pose_xyt"].astype(np.float64) + 0.2 * x.reshape(-1, 3)).astype(np.float64)
    g["lm_xy"] = (g["lm_xy"].astype(np.float64) + 0.2 * dl)
    l0, rhs, M, _ = setup(g)
    x, it = pcg(l0.S, rhs, M); msg = f"GN {gn}: plain {it}"
    for k in a.k:
        W = Wold[:, :k]; SW = np.stack([l0.S(W[:, j]) for j in range(k)], 1); G = sl.cho_factor(W.T @ SW)
        _, itd = pcg(l0.S, rhs, M, W=W, G=G); msg += f" | k={k} stale-W deflated {itd}"
    print(msg, flush=True)
