#!/usr/bin/env python
"""CPU prototype (numpy/scipy) of MULTILEVEL preconditioners for the reduced pose system (round 2).

Every level is a graph of the same shape as the fine one: 3-dof "poses" (rigid-body modes of an
aggregate of consecutive poses of the level below), the SAME 2-dof landmarks, a sparse pose-pose
part App_l and 3x2 pose-landmark blocks C_l = Z^T C_{l-1}.  S_l = App_l - C_l Hll^-1 C_l^T is then
exactly the Galerkin operator Z^T S_{l-1} Z, and can be applied matrix-free at every level.

  python tools/mg_proto.py --workload c3 --scale 0.3 --m 4 --variants add mult
"""
import argparse, os, sys, time
import numpy as np
import scipy.sparse as sp
import scipy.linalg as sl
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from toyslam_b200 import synth
from precond_proto import linearize


OMEGA = 1.0
WORK = [0.0]
UNIT = [1.0]


class Level:
    def __init__(self, App, C, Hinv, cen, dof=3):
        self.dof = dof
        self.App = App.tocsr(); self.n = App.shape[0] // dof
        self.C = C.tobsr(blocksize=(dof, 2)); self.C.sort_indices()
        self.CT = self.C.T.tocsr(); self.Ccsr = self.C.tocsr()
        self.Hinv = Hinv; self.cen = cen; self.spacing = 1
        L = Hinv.shape[0]
        r_ = (2 * np.arange(L)[:, None, None] + np.arange(2)[None, :, None] + 0 * np.arange(2)[None, None, :]).ravel()
        c_ = (2 * np.arange(L)[:, None, None] + 0 * np.arange(2)[None, :, None] + np.arange(2)[None, None, :]).ravel()
        self.HinvS = sp.coo_matrix((Hinv.ravel(), (r_, c_)), shape=(2 * L, 2 * L)).tocsr()
        # block diagonal of S_l
        n = self.n
        D = np.zeros((n, dof, dof))
        A = self.App.tobsr(blocksize=(dof, dof)); A.sort_indices()
        rows = np.repeat(np.arange(n), np.diff(A.indptr))
        dm = A.indices == rows
        D[rows[dm]] = A.data[dm]
        rowsC = np.repeat(np.arange(n), np.diff(self.C.indptr))
        Cb = self.C.data
        np.add.at(D, rowsC, -(Cb @ Hinv[self.C.indices] @ Cb.transpose(0, 2, 1)))
        self.D = D
        self.Dinv = np.linalg.inv(D)
        self.nnzC = len(Cb)

    def S(self, v):
        WORK[0] += self.nnzC + 5 * self.n
        return self.App @ v - self.Ccsr @ (self.HinvS @ (self.CT @ v))

    def jac(self, r, om=None):
        return (OMEGA if om is None else om) * (self.Dinv @ r.reshape(self.n, self.dof, 1)).reshape(-1)

    def dense(self):
        M = (self.Ccsr @ self.HinvS @ self.CT).toarray()
        return self.App.toarray() - M


LINEAR = [False]


def coarsen_linear(lev, m):
    """continuous piecewise-linear interpolation of rigid-body node motions along the chain"""
    n = lev.n; nA = (n + m - 1) // m
    k = np.arange(n) // m; t = (np.arange(n) % m) / m
    cen = lev.cen[::m][:nA].copy()
    k1 = np.minimum(k + 1, nA - 1)
    w0 = np.where(k1 == k, 1.0, 1.0 - t); w1 = np.where(k1 == k, 0.0, t)
    rws, cls, vls = [], [], []
    for par, w in ((k, w0), (k1, w1)):
        d = lev.cen - cen[par]
        rws += [np.arange(3 * n), 3 * np.arange(n), 3 * np.arange(n) + 1]
        cls += [3 * np.repeat(par, 3) + np.tile(np.arange(3), n), 3 * par + 2, 3 * par + 2]
        vls += [np.repeat(w, 3), -d[:, 1] * w, d[:, 0] * w]
    Z = sp.coo_matrix((np.concatenate(vls), (np.concatenate(rws), np.concatenate(cls))), shape=(3 * n, 3 * nA)).tocsr()
    Z.eliminate_zeros()
    App = (Z.T @ lev.App @ Z).tocsr()
    C = (Z.T @ lev.Ccsr).tocsr()
    return Z, Level(App, C, lev.Hinv, cen)


AFFINE = [False]


def coarsen_affine(lev, m):
    """linear interpolation along the chain of AFFINE node fields (a, A): 6 dofs per node"""
    n = lev.n; nA = (n + m - 1) // m; cd = lev.dof
    k = np.arange(n) // m; t = (np.arange(n) % m) / m
    cen = lev.cen[::m][:nA].copy()
    k1 = np.minimum(k + 1, nA - 1)
    w0 = np.where(k1 == k, 1.0, 1.0 - t); w1 = np.where(k1 == k, 0.0, t)
    rws, cls, vls = [], [], []
    def add(r, c, v): rws.append(r); cls.append(c); vls.append(v)
    ar = np.arange(n)
    for par, w in ((k, w0), (k1, w1)):
        d = lev.cen - cen[par]
        if cd == 3:
            # dx = ax + A11 dx + A12 dy ; dy = ay + A21 dx + A22 dy ; dth = (A21 - A12)/2 ; parent dofs (ax, ay, A11, A12, A21, A22)
            add(3 * ar, 6 * par + 0, w); add(3 * ar, 6 * par + 2, w * d[:, 0]); add(3 * ar, 6 * par + 3, w * d[:, 1])
            add(3 * ar + 1, 6 * par + 1, w); add(3 * ar + 1, 6 * par + 4, w * d[:, 0]); add(3 * ar + 1, 6 * par + 5, w * d[:, 1])
            add(3 * ar + 2, 6 * par + 4, 0.5 * w); add(3 * ar + 2, 6 * par + 3, -0.5 * w)
        else:
            add(6 * ar, 6 * par + 0, w); add(6 * ar, 6 * par + 2, w * d[:, 0]); add(6 * ar, 6 * par + 3, w * d[:, 1])
            add(6 * ar + 1, 6 * par + 1, w); add(6 * ar + 1, 6 * par + 4, w * d[:, 0]); add(6 * ar + 1, 6 * par + 5, w * d[:, 1])
            for q in range(2, 6): add(6 * ar + q, 6 * par + q, w)
    Z = sp.coo_matrix((np.concatenate(vls), (np.concatenate(rws), np.concatenate(cls))), shape=(cd * n, 6 * nA)).tocsr()
    App = (Z.T @ lev.App @ Z).tocsr()
    C = (Z.T @ lev.Ccsr).tocsr()
    return Z, Level(App, C, lev.Hinv, cen, dof=6)


GRAPH = [0]


def match_once(W, n):
    """greedy heavy-edge matching on symmetric weights W (csr); returns group id per node"""
    W = W.tocsr(); indptr, indices, data = W.indptr, W.indices, W.data
    grp = -np.ones(n, np.int64); ng = 0
    order = np.argsort(np.diff(indptr), kind="stable")       # low-degree nodes first
    for k in order:
        if grp[k] >= 0: continue
        best = -1; bw = 0.0
        for q in range(indptr[k], indptr[k + 1]):
            j = indices[q]
            if j != k and grp[j] < 0 and data[q] > bw: bw = data[q]; best = j
        grp[k] = ng
        if best >= 0: grp[best] = ng
        ng += 1
    return grp, ng


def coarsen_graph(lev, m):
    """aggregate strongly coupled nodes (shared landmarks), piecewise-constant transfer of AFFINE node fields"""
    assert lev.dof == 6
    n = lev.n
    B = sp.csr_matrix((np.ones(len(lev.C.indices)), lev.C.indices, lev.C.indptr), shape=(n, lev.Hinv.shape[0]))
    deg = np.asarray(B.sum(1)).ravel()
    grp = np.arange(n); ng = n
    G = sp.identity(n, format="csr")
    rounds = int(round(np.log2(m)))
    for _ in range(rounds):
        Bg = (G.T @ B); Bg.data[:] = 1.0
        W = (Bg @ Bg.T).tocsr()
        dg = np.asarray(Bg.sum(1)).ravel()
        W = sp.diags(1.0 / np.sqrt(np.maximum(dg, 1))) @ W @ sp.diags(1.0 / np.sqrt(np.maximum(dg, 1)))
        g2, n2 = match_once(W.tocsr(), ng)
        G = G @ sp.csr_matrix((np.ones(ng), (np.arange(ng), g2)), shape=(ng, n2))
        ng = n2
    par = np.asarray(G.argmax(1)).ravel()
    nA = ng
    cnt = np.bincount(par, minlength=nA)
    cen = np.stack([np.bincount(par, lev.cen[:, 0], nA), np.bincount(par, lev.cen[:, 1], nA)], 1) / cnt[:, None]
    d = lev.cen - cen[par]
    ar = np.arange(n); rws, cls, vls = [], [], []
    def add(r, c, v): rws.append(r); cls.append(c); vls.append(v)
    one = np.ones(n)
    add(6 * ar, 6 * par + 0, one); add(6 * ar, 6 * par + 2, d[:, 0]); add(6 * ar, 6 * par + 3, d[:, 1])
    add(6 * ar + 1, 6 * par + 1, one); add(6 * ar + 1, 6 * par + 4, d[:, 0]); add(6 * ar + 1, 6 * par + 5, d[:, 1])
    for q in range(2, 6): add(6 * ar + q, 6 * par + q, one)
    Z = sp.coo_matrix((np.concatenate(vls), (np.concatenate(rws), np.concatenate(cls))), shape=(6 * n, 6 * nA)).tocsr()
    App = (Z.T @ lev.App @ Z).tocsr()
    C = (Z.T @ lev.Ccsr).tocsr()
    return Z, Level(App, C, lev.Hinv, cen, dof=6)


def coarsen(lev, m):
    if GRAPH[0] and lev.spacing * m >= GRAPH[0]: return coarsen_graph(lev, m)
    if AFFINE[0] and lev.spacing * m >= AFFINE[1]: return coarsen_affine(lev, m)
    if LINEAR[0]: return coarsen_linear(lev, m)
    n = lev.n; nA = (n + m - 1) // m
    agg = np.arange(n) // m
    cnt = np.bincount(agg, minlength=nA)
    cen = np.stack([np.bincount(agg, lev.cen[:, 0]), np.bincount(agg, lev.cen[:, 1])], 1) / cnt[:, None]
    d = lev.cen - cen[agg]
    rws = np.concatenate([np.arange(3 * n), 3 * np.arange(n), 3 * np.arange(n) + 1])
    cls = np.concatenate([3 * np.repeat(agg, 3) + np.tile(np.arange(3), n), 3 * agg + 2, 3 * agg + 2])
    vls = np.concatenate([np.ones(3 * n), -d[:, 1], d[:, 0]])
    Z = sp.coo_matrix((vls, (rws, cls)), shape=(3 * n, 3 * nA)).tocsr()
    App = (Z.T @ lev.App @ Z).tocsr()
    C = (Z.T @ lev.Ccsr).tocsr()
    return Z, Level(App, C, lev.Hinv, cen)


def pcg(Sop, rhs, apply_M, tol, label, flexible=False, maxit=5000):
    WORK[0] = 0.0
    x = np.zeros_like(rhs); r = rhs.copy(); z = apply_M(r); p = z.copy(); rz = r @ z; rz0 = rz; it = 0
    t0 = time.time(); al = []; be = []
    while it < maxit:
        Sp = Sop(p); alpha = rz / (p @ Sp); x += alpha * p; al.append(alpha)
        rold = r.copy() if flexible else None
        r -= alpha * Sp; z = apply_M(r); rzn = r @ z; it += 1
        if not (rzn > tol ** 2 * rz0): break
        beta = (z @ (r - rold)) / rz if flexible else rzn / rz
        p = z + beta * p; rz = rzn; be.append(beta)
    if len(al) > 2 and not flexible:
        k = len(al); T = np.zeros((k, k))
        for j in range(k):
            T[j, j] = 1 / al[j] + (be[j - 1] / al[j - 1] if j > 0 else 0)
            if j + 1 < k and j < len(be): T[j, j + 1] = T[j + 1, j] = np.sqrt(be[j]) / al[j]
        ev = np.linalg.eigvalsh(T)
        label += f" [lam {ev[0]:.2e}..{ev[-1]:.2f}, 2nd {ev[1]:.2e} 5th {ev[4]:.2e} 20th {ev[min(19, k - 1)]:.2e}]"
    wk = WORK[0] / (len(rhs) // 3 * 5 + UNIT[0])
    tr = np.linalg.norm(rhs - Sop(x)) / np.linalg.norm(rhs)
    print(f"{label} iters {it:5d} work {wk:7.1f}  true rel res {tr:.1e}  ({time.time() - t0:.1f}s)", flush=True)
    return x, it


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--workload", default="c2"); ap.add_argument("--scale", type=float, default=1.0)
    ap.add_argument("--m", type=int, nargs="*", default=[4]); ap.add_argument("--tol", type=float, default=1e-12)
    ap.add_argument("--coarsest", type=int, default=400, help="stop coarsening at <= this many aggregates (dense solve)")
    ap.add_argument("--variants", nargs="*", default=["two", "add", "mult"])
    ap.add_argument("--first", type=int, default=0, help="first-level aggregate size (0 = m)")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--omega", type=float, default=1.0)
    ap.add_argument("--linear", action="store_true")
    ap.add_argument("--graph", type=int, default=0, help="levels with spacing >= this aggregate by shared landmarks (needs affine below)")
    ap.add_argument("--affine", type=int, default=0, help="levels with spacing >= this use affine node fields (6 dofs)")
    a = ap.parse_args()
    global OMEGA
    OMEGA = a.omega
    LINEAR[0] = a.linear
    if a.affine: AFFINE[0] = True; AFFINE.append(a.affine)
    GRAPH[0] = a.graph
    g = synth.make_config(a.workload, scale=a.scale, seed=a.seed)
    t0 = time.time()
    Hpp, Hpl, Hll, bp, bl, Hd = linearize(g)
    P, L = Hd.shape[0], Hll.shape[0]
    Hinv = np.linalg.inv(Hll)
    xy = g["pose_xyt"].astype(np.float64)[:, :2]
    l0 = Level(Hpp, Hpl, Hinv, xy)
    UNIT[0] = l0.nnzC
    rhs = bp - l0.Ccsr @ (l0.HinvS @ bl.ravel())
    print(f"P={P} L={L} El={l0.nnzC} setup {time.time() - t0:.1f}s", flush=True)

    if "jac" in a.variants:
        pcg(l0.S, rhs, l0.jac, a.tol, "block-Jacobi")
    for m in a.m:
        t0 = time.time()
        levels = [l0]; Zs = []
        first = a.first or m
        while levels[-1].n > a.coarsest:
            mm = first if len(levels) == 1 else m
            Z, lv = coarsen(levels[-1], mm); lv.spacing = levels[-1].spacing * mm
            Zs.append(Z); levels.append(lv)
        print(f"m={m} first={first}: levels " + " ".join(f"{lv.n}x{lv.dof}/{lv.nnzC}" for lv in levels) + f"  ({time.time() - t0:.1f}s)", flush=True)
        Ac = levels[-1].dense(); cf = sl.cho_factor(Ac)
        csolve = lambda r: sl.cho_solve(cf, r)
        nl = len(levels)

        if "two" in a.variants and nl >= 2:
            # exact two-level with the first coarse level solved exactly (sparse direct) -- the yardstick
            l1 = levels[1]
            if l1.n <= 12000:
                A1 = l1.dense(); cf1 = sl.cho_factor(A1)
                pcg(l0.S, rhs, lambda r: l0.jac(r) + Zs[0] @ sl.cho_solve(cf1, Zs[0].T @ r), a.tol, f"m={m} two-level additive, exact level-1 solve")

                def mult2(r):
                    z1 = l0.jac(r); r1 = r - l0.S(z1); z2 = z1 + Zs[0] @ sl.cho_solve(cf1, Zs[0].T @ r1); return z2 + l0.jac(r - l0.S(z2))
                pcg(l0.S, rhs, mult2, a.tol, f"m={m} two-level multiplicative V(1,1), exact level-1")

        if "add" in a.variants:
            def addml(r, l=0):
                if l == nl - 1: return csolve(r)
                return levels[l].jac(r) + Zs[l] @ addml(Zs[l].T @ r, l + 1)
            pcg(l0.S, rhs, addml, a.tol, f"m={m} additive multilevel (BPX-like), {nl} levels")

        for var in a.variants:
            if not var.startswith("clus"): continue
            # clus:<level>:<size>:<mode>  block-Jacobi with clusters of ~size nodes on that level (mode t = consecutive in time, g = grown on the co-visibility graph)
            _, cl, csz, mode = var.split(":"); cl = int(cl); csz = int(csz)
            lv = levels[cl]; n = lv.n; dof = lv.dof
            if mode == "t":
                cid = np.arange(n) // csz
            else:
                B = sp.csr_matrix((np.ones(len(lv.C.indices)), lv.C.indices, lv.C.indptr), shape=(n, lv.Hinv.shape[0]))
                W = (B @ B.T).tocsr(); W.setdiag(0); W.eliminate_zeros()
                cid = -np.ones(n, np.int64); nc_ = 0
                for seed in range(n):
                    if cid[seed] >= 0: continue
                    members = [seed]; cid[seed] = nc_
                    score = {}
                    def push(k):
                        for q in range(W.indptr[k], W.indptr[k + 1]):
                            j = W.indices[q]
                            if cid[j] < 0: score[j] = score.get(j, 0.0) + W.data[q]
                    push(seed)
                    while len(members) < csz and score:
                        j = max(score, key=score.get); del score[j]
                        if cid[j] >= 0: continue
                        cid[j] = nc_; members.append(j); push(j)
                    nc_ += 1
            ncl = int(cid.max()) + 1
            print(f"   level {cl}: {n} nodes in {ncl} clusters (mode {mode})", flush=True)
            Sd = lv.dense() if n * dof <= 12000 else None
            order = np.argsort(cid, kind="stable"); bounds = np.r_[0, np.cumsum(np.bincount(cid, minlength=ncl))]
            invs = []; idxs = []
            for c in range(ncl):
                nodes = order[bounds[c]:bounds[c + 1]]
                ix = (dof * nodes[:, None] + np.arange(dof)[None, :]).ravel()
                idxs.append(ix); invs.append(np.linalg.inv(Sd[np.ix_(ix, ix)]))
            def cjac(r):
                z = np.zeros_like(r)
                for ix, iv in zip(idxs, invs): z[ix] = iv @ r[ix]
                return z
            def addcl(r, l=0):
                if l == nl - 1: return csolve(r)
                sm = cjac(r) if l == cl else levels[l].jac(r)
                return sm + Zs[l] @ addcl(Zs[l].T @ r, l + 1)
            pcg(l0.S, rhs, addcl, a.tol, f"m={m} additive multilevel, level {cl} block-Jacobi over {mode}-clusters of {csz}")

        for var in a.variants:
            if not var.startswith("cheb"): continue
            # cheb:<levels>:<degree>:<ratio>  Chebyshev-Jacobi polynomial smoother (degree S_l applies) on those levels, eigenvalue window [lmax/ratio, lmax]
            _, lvs, deg, ratio = var.split(":"); lvs = set(int(x) for x in lvs.split(",")); deg = int(deg); ratio = float(ratio)
            lmax = {}
            for l in lvs:
                lv = levels[l]; v = np.random.default_rng(0).normal(size=lv.n * lv.dof)
                for _ in range(30): v = lv.jac(lv.S(v), 1.0); lam = np.linalg.norm(v); v /= lam
                lmax[l] = 1.1 * lam
            def cheb(l, r):
                lv = levels[l]; b_, a_ = lmax[l], lmax[l] / ratio
                theta = 0.5 * (b_ + a_); delta = 0.5 * (b_ - a_); sigma = theta / delta; rho = 1.0 / sigma
                x = np.zeros_like(r); res = r.copy(); dvec = lv.jac(res, 1.0) / theta; 
                for k in range(deg):
                    x = x + dvec
                    if k == deg - 1: break
                    res = res - lv.S(dvec)
                    rho_new = 1.0 / (2 * sigma - rho)
                    dvec = rho_new * rho * dvec + 2 * rho_new / delta * lv.jac(res, 1.0)
                    rho = rho_new
                return x
            def addch(r, l=0):
                if l == nl - 1: return csolve(r)
                sm = cheb(l, r) if l in lvs else levels[l].jac(r)
                return sm + Zs[l] @ addch(Zs[l].T @ r, l + 1)
            pcg(l0.S, rhs, addch, a.tol, f"m={m} additive multilevel, Chebyshev deg {deg} (window {ratio}) on levels {sorted(lvs)}")

        for var in a.variants:
            if not var.startswith("skip"): continue
            # skip:<list of level indices whose Jacobi term is dropped>, e.g. skip:2,3
            drop = set(int(x) for x in var.split(":")[1].split(","))
            def addsk(r, l=0):
                if l == nl - 1: return csolve(r)
                z = Zs[l] @ addsk(Zs[l].T @ r, l + 1)
                return z if l in drop else levels[l].jac(r) + z
            pcg(l0.S, rhs, addsk, a.tol, f"m={m} additive multilevel without Jacobi on levels {sorted(drop)}")

        if "mult" in a.variants:
            def vcyc(r, l=0):
                if l == nl - 1: return csolve(r)
                lv = levels[l]
                z = lv.jac(r)
                z = z + Zs[l] @ vcyc(Zs[l].T @ (r - lv.S(z)), l + 1)
                return z + lv.jac(r - lv.S(z))
            pcg(l0.S, rhs, vcyc, a.tol, f"m={m} V(1,1) multilevel, {nl} levels")

        if "hyb" in a.variants:
            # additive at the fine level (no extra fine S apply), V(1,1) below
            def vcyc(r, l):
                if l == nl - 1: return csolve(r)
                lv = levels[l]
                z = lv.jac(r)
                z = z + Zs[l] @ vcyc(Zs[l].T @ (r - lv.S(z)), l + 1)
                return z + lv.jac(r - lv.S(z))
            pcg(l0.S, rhs, lambda r: l0.jac(r) + Zs[0] @ vcyc(Zs[0].T @ r, 1), a.tol, f"m={m} fine additive + V(1,1) below, {nl} levels")

        if "hyb2" in a.variants:
            # additive at the fine level, V(2,2) below
            def vcyc2(r, l):
                if l == nl - 1: return csolve(r)
                lv = levels[l]
                z = lv.jac(r)
                z = z + lv.jac(r - lv.S(z))
                z = z + Zs[l] @ vcyc2(Zs[l].T @ (r - lv.S(z)), l + 1)
                z = z + lv.jac(r - lv.S(z))
                return z + lv.jac(r - lv.S(z))
            pcg(l0.S, rhs, lambda r: l0.jac(r) + Zs[0] @ vcyc2(Zs[0].T @ r, 1), a.tol, f"m={m} fine additive + V(2,2) below, {nl} levels")

        for var in a.variants:
            if not var.startswith("cyc"): continue
            # cyc:<nu>:<gamma>:<fineadd 0/1>  generic cycle: nu pre/post damped-Jacobi sweeps, gamma recursive calls
            _, nu, gamma, fa = var.split(":"); nu = int(nu); gamma = int(gamma); fa = int(fa)
            def cyc(r, l):
                if l == nl - 1: return csolve(r)
                lv = levels[l]
                z = lv.jac(r)
                for _ in range(nu - 1): z = z + lv.jac(r - lv.S(z))
                for _ in range(gamma if l < nl - 2 else 1):
                    z = z + Zs[l] @ cyc(Zs[l].T @ (r - lv.S(z)), l + 1)
                for _ in range(nu): z = z + lv.jac(r - lv.S(z))
                return z
            if fa:
                pcg(l0.S, rhs, lambda r: l0.jac(r, 1.0) + Zs[0] @ cyc(Zs[0].T @ r, 1), a.tol, f"m={m} fine additive + cyc(nu={nu},gamma={gamma}) below, {nl} levels")
            else:
                pcg(l0.S, rhs, lambda r: cyc(r, 0), a.tol, f"m={m} cyc(nu={nu},gamma={gamma}), {nl} levels")

        if "kcyc" in a.variants:
            # fine additive, below: K-cycle (2 flexible-CG steps per level preconditioned by the next level)
            def kc(r, l):
                if l == nl - 1: return csolve(r)
                lv = levels[l]
                def M(rr):
                    return lv.jac(rr) + Zs[l] @ kc(Zs[l].T @ rr, l + 1)
                # two steps of flexible CG on S_l e = r
                z1 = M(r); Sz1 = lv.S(z1); a1 = (z1 @ r) / (z1 @ Sz1)
                r1 = r - a1 * Sz1
                z2 = M(r1); Sz2 = lv.S(z2)
                # orthogonalise z2 against z1 in the S inner product
                g12 = (z2 @ Sz1) / (z1 @ Sz1)
                p2 = z2 - g12 * z1; Sp2 = Sz2 - g12 * Sz1
                a2 = (p2 @ r1) / (p2 @ Sp2)
                return a1 * z1 + a2 * p2
            pcg(l0.S, rhs, lambda r: l0.jac(r) + Zs[0] @ kc(Zs[0].T @ r, 1), a.tol, f"m={m} fine additive + K-cycle below, {nl} levels", flexible=True)


if __name__ == "__main__":
    main()
