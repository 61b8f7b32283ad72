"""CPU tests: the oracle (oracle/oracle.cpp) against the golden vectors generated from the
reference's Python implementation (tools/make_golden.py), and T1 (dense QR) against T2 (sparse)."""
import numpy as np
import pytest
from conftest import load_golden, GOLDEN_CASES, state_diff
from oracle.oracle import OracleGraph, solve_dense
from toyslam_b200 import synth, wire

# The reference's Python takes angles through scipy Rotation (python/tools.py:7-13), which
# re-orthonormalises the fp32-rounded ODOM measurement; the C++ (Helper.h:20-24) and this oracle use
# atan2(R10, R00).  The two differ by ~1e-8 rad on the ODOM theta residual, hence these tolerances.
TOL = {"c1": 1e-10}
DEFAULT_TOL = 2e-7


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_linearization_matches_reference_python(name):
    d = load_golden(name)
    o = OracleGraph(d["request"].tobytes())
    pi, li = o.ids()
    assert np.array_equal(pi, d["pose_ids"]) and np.array_equal(li, d["lm_ids"])
    for k in range(o.El):
        e, A, B = o.edge_lm(k)
        assert np.allclose(e, d["lm_e"][k], rtol=0, atol=1e-12)
        assert np.allclose(A, d["lm_A"][k], rtol=0, atol=1e-11)
        assert np.allclose(B, d["lm_B"][k], rtol=0, atol=1e-14)
    for k in range(o.Eo):
        e = o.edge_odom(k)
        assert np.allclose(e[:2], d["od_e"][k][:2], rtol=0, atol=1e-12)
        assert abs(e[2] - d["od_e"][k][2]) < 1e-7
    H, b, chi2 = o.assemble_dense()
    tol = TOL.get(name, DEFAULT_TOL)
    assert np.max(np.abs(H - d["H"])) <= tol * np.max(np.abs(d["H"]))
    assert np.max(np.abs(b - d["b"])) <= tol * np.max(np.abs(d["b"]))
    assert abs(chi2 - d["chi2"]) <= tol * d["chi2"]


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_gn_loop_matches_golden(name):
    d = load_golden(name)
    o = OracleGraph(d["request"].tobytes())
    traj, code = o.optimize_dense(int(d["iterations"]), bool(d["exits"]))
    assert len(traj) == len(d["traj"]) and code == int(d["exit_code"])
    tol = TOL.get(name, DEFAULT_TOL) * 10
    assert np.max(np.abs(traj - d["traj"]) / np.abs(d["traj"])) < tol
    ps, ls = o.state()
    assert state_diff(ps, ls, d["final_pose"], d["final_lm"]) < max(tol, 1e-9)


@pytest.mark.parametrize("name", ["g_small", "g_loops", "g_outliers", "g_odom_only"])
def test_sparse_tier_matches_dense_tier(name):
    d = load_golden(name)
    a, b = OracleGraph(d["request"].tobytes()), OracleGraph(d["request"].tobytes())
    ta, ca = a.optimize_dense(int(d["iterations"]), bool(d["exits"]))
    tb, its, cb = b.optimize_sparse(int(d["iterations"]), bool(d["exits"]), tol=1e-14, max_pcg=2000, threads=1)
    assert len(ta) == len(tb) and ca == cb
    assert np.max(np.abs(ta - tb) / np.abs(ta)) < 1e-10
    assert state_diff(*a.state(), *b.state()) < 1e-10


def test_qr_solve_against_lapack():
    rng = np.random.default_rng(0)
    A = rng.normal(size=(60, 60)); H = A @ A.T + np.diag(np.r_[np.full(3, 1e6), np.zeros(57)])
    b = rng.normal(size=60)
    x = solve_dense(H, b)
    assert np.max(np.abs(x - np.linalg.solve(H, b))) < 1e-9 * np.max(np.abs(x))


def test_reply_round_trip():
    d = load_golden("g_loops")
    o = OracleGraph(d["request"].tobytes())
    o.optimize_dense(3, True)
    r = wire.decode_reply(o.reply())
    ps, ls = o.state(); pi, li = o.ids()
    for i, vid in enumerate(pi):
        t, v = r["vertices"][int(vid)]
        assert t == 0 and np.allclose(v, ps[i].astype(np.float32))
    for i, vid in enumerate(li):
        t, v = r["vertices"][int(vid)]
        assert t == 1 and np.allclose(v, ls[i].astype(np.float32))
    assert len(r["edges"]) == o.Eo + o.El and len(r["fixed"]) == o.F
    assert all(e[3].shape == ((3, 3) if e[0] == 0 else (2, 1)) for e in r["edges"])


def test_malformed_payload_is_rejected():
    d = load_golden("g_small")
    buf = d["request"].tobytes()
    for cut in (0, 3, 17, len(buf) // 2, len(buf) - 1):
        with pytest.raises(ValueError):
            OracleGraph(buf[:cut])


def test_synthetic_generator_shapes():
    g = synth.make_config("c2", scale=0.02)
    c = synth.counts(g)
    assert c["P"] == 200 and c["Eo"] == 199 and c["El"] == 200 * 15
    o = OracleGraph(synth.to_payload(g))
    assert (o.P, o.Eo, o.El) == (200, 199, 3000)
    traj, its, code = o.optimize_sparse(5, False, 1e-12, 2000, 1)
    assert traj[-1] < traj[0]
