"""GPU parity tests (run on the B200 box): the CUDA path, called through the C ABI, against the
CPU oracle on the same wire payloads.  Tolerances are the north star's: state 1e-9 relative,
chi2 trajectory 1e-8 relative (fp64)."""
import numpy as np
import pytest
from conftest import load_golden, GOLDEN_CASES, state_diff

pytestmark = pytest.mark.gpu

STATE_TOL = 1e-9
CHI2_TOL = 1e-8


@pytest.fixture(scope="module")
def gpu():
    from toyslam_b200.capi import GraphOptimizerGpu
    import torch
    assert torch.cuda.is_available(), "GPU tests need a CUDA device"
    return GraphOptimizerGpu


def dense_from_blocks(sysd, od_i, od_j, lm_i, lm_l, P, L):
    n = 3 * P + 2 * L
    H = np.zeros((n, n)); b = np.zeros(n)
    tri = [(0, 0), (0, 1), (0, 2), (1, 1), (1, 2), (2, 2)]
    for p in range(P):
        for v, (r, c) in zip(sysd["hpp_diag"][p], tri):
            H[3 * p + r, 3 * p + c] = v; H[3 * p + c, 3 * p + r] = v
    b[:3 * P] = sysd["b_pose"].reshape(-1)
    for k, (i, j) in enumerate(zip(od_i, od_j)):
        for q in range(3):
            H[3 * i + q, 3 * j + q] += sysd["hpp_off"][k, q]; H[3 * j + q, 3 * i + q] += sysd["hpp_off"][k, q]
    for k, (i, l) in enumerate(zip(lm_i, lm_l)):
        H[3 * i:3 * i + 3, 3 * P + 2 * l:3 * P + 2 * l + 2] += sysd["hpl"][k]
        H[3 * P + 2 * l:3 * P + 2 * l + 2, 3 * i:3 * i + 3] += sysd["hpl"][k].T
    for l in range(L):
        o = 3 * P + 2 * l
        H[o, o], H[o, o + 1], H[o + 1, o], H[o + 1, o + 1] = sysd["hll"][l, 0], sysd["hll"][l, 1], sysd["hll"][l, 1], sysd["hll"][l, 2]
    b[3 * P:] = sysd["b_lm"].reshape(-1)
    return H, b


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_linearize_and_solve_match_oracle(gpu, name):
    from oracle.oracle import OracleGraph
    d = load_golden(name)
    payload = d["request"].tobytes()
    o = OracleGraph(payload)
    Ho, bo, chi2o = o.assemble_dense()
    g = gpu(pcg_tol=1e-14, pcg_max_iters=5000)
    c = g.load_wire(payload)
    assert (c.n_pose, c.n_lm, c.n_odom, c.n_lmedge) == (o.P, o.Lm, o.Eo, o.El)
    pi, li = g.ids(); opi, oli = o.ids()
    assert np.array_equal(pi, opi) and np.array_equal(li, oli)
    st = g.linearize()
    assert abs(st.chi2 - chi2o) <= 1e-12 * chi2o
    H, b = dense_from_blocks(g.system(), *o.edge_index(), o.P, o.Lm)
    assert np.max(np.abs(H - Ho)) <= 1e-12 * np.max(np.abs(Ho))
    assert np.max(np.abs(b - bo)) <= 1e-11 * np.max(np.abs(bo))
    # solve: landmark Schur + PCG against a dense LAPACK solve of the oracle's system
    s2 = g.solve()
    dp, dl = g.delta()
    x = np.linalg.solve(Ho, bo)
    got = np.r_[dp.reshape(-1), dl.reshape(-1)]
    assert np.max(np.abs(got - x)) <= 1e-9 * max(1.0, np.max(np.abs(x))), (s2.pcg_iters, s2.pcg_rel_residual)


@pytest.mark.parametrize("name", GOLDEN_CASES)
def test_optimize_matches_oracle(gpu, name):
    from oracle.oracle import OracleGraph
    d = load_golden(name)
    payload = d["request"].tobytes()
    iters, exits = int(d["iterations"]), bool(d["exits"])
    o = OracleGraph(payload)
    traj, code = o.optimize_dense(iters, exits)
    g = gpu(iterations=iters, pcg_tol=1e-14, pcg_max_iters=5000, enable_exits=exits)
    g.load_wire(payload)
    rep, per = g.optimize()
    got = np.array([p.chi2 for p in per])
    assert rep.iterations_run == len(traj) and rep.exit_code == code
    assert np.max(np.abs(got - traj) / np.abs(traj)) < CHI2_TOL
    assert state_diff(*g.state(), *o.state()) < STATE_TOL
    # and against the golden trajectory produced from the reference's Python linearisation
    assert len(d["traj"]) == len(got)
    assert np.max(np.abs(got - d["traj"]) / np.abs(d["traj"])) < (1e-9 if name == "c1" else 5e-6)


def test_structure_is_bit_exact(gpu):
    """The integer edge -> block index map held on the device equals an independent numpy derivation."""
    from oracle.oracle import OracleGraph
    from toyslam_b200 import synth
    for payload in (load_golden("c1")["request"].tobytes(), synth.to_payload(synth.make_config("c2", scale=0.05))):
        o = OracleGraph(payload)
        _, _, lm_i, lm_l = o.edge_index()
        g = gpu(); c = g.load_wire(payload); s = g.structure()
        order = np.lexsort((np.arange(len(lm_i)), lm_l, lm_i)).astype(np.uint32)
        assert np.array_equal(s["lm_perm"], order)
        ptr = np.r_[0, np.cumsum(np.bincount(lm_i, minlength=o.P))].astype(np.uint32)
        assert np.array_equal(s["pose_ptr"], ptr)
        slot_lm_expect = []
        for (e0, e1, p0, p1, s0, s1, blk, _) in s["chunk_desc"]:
            assert e1 - e0 <= 512 or p1 - p0 == 1
            assert ptr[p0] <= e0 and e1 <= ptr[p1]
            lms = lm_l[order[e0:e1]]; uniq, cnt = np.unique(lms, return_counts=True)
            by_len = np.lexsort((uniq, -cnt))                   # slots: descending edge count, then ascending landmark
            rank = np.empty(len(uniq), np.uint32); rank[by_len] = np.arange(len(uniq), dtype=np.uint32)
            assert s1 - s0 == len(uniq); slot_lm_expect.append(uniq[by_len])
            pose_local = lm_i[order[e0:e1]] - p0
            packed = (pose_local.astype(np.uint32) << 16) | rank[np.searchsorted(uniq, lms)]
            assert np.array_equal(s["lm_packed"][e0:e1], packed)
        slot_lm = np.concatenate(slot_lm_expect) if slot_lm_expect else np.zeros(0, np.uint32)
        assert np.array_equal(s["slot_lm"], slot_lm)
        assert np.array_equal(s["lm_part_ptr"], np.r_[0, np.cumsum(np.bincount(slot_lm, minlength=o.Lm))])
        assert np.array_equal(s["lm_part_idx"], np.argsort(slot_lm, kind="stable"))
        # chunks tile the edge range and the blocks tile the pose range
        cd = s["chunk_desc"]
        assert cd[0, 0] == 0 and cd[-1, 1] == c.n_lmedge and np.array_equal(cd[1:, 0], cd[:-1, 1])


def test_runs_are_bit_reproducible(gpu):
    from toyslam_b200 import synth
    payload = synth.to_payload(synth.make_config("c2", scale=0.2, seed=3))
    outs = []
    for _ in range(2):
        g = gpu(iterations=4, enable_exits=False); g.load_wire(payload)
        rep, per = g.optimize()
        outs.append((np.array([p.chi2 for p in per]), *g.state(), [p.pcg_iters for p in per]))
    assert np.array_equal(outs[0][0], outs[1][0]) and np.array_equal(outs[0][1], outs[1][1])
    assert np.array_equal(outs[0][2], outs[1][2]) and outs[0][3] == outs[1][3]


@pytest.mark.parametrize("cfg,scale", [("c2", 0.1), ("c3", 0.01), ("c5", 0.002)])
def test_synthetic_matches_sparse_oracle(gpu, cfg, scale):
    from oracle.oracle import OracleGraph
    from toyslam_b200 import synth
    payload = synth.to_payload(synth.make_config(cfg, scale=scale, seed=1))
    o = OracleGraph(payload)
    traj, its, code = o.optimize_sparse(8, False, tol=1e-14, max_pcg=20000, threads=0)
    g = gpu(iterations=8, pcg_tol=1e-14, pcg_max_iters=20000, enable_exits=False); g.load_wire(payload)
    rep, per = g.optimize()
    got = np.array([p.chi2 for p in per])
    assert np.max(np.abs(got - traj) / np.abs(traj)) < CHI2_TOL
    assert state_diff(*g.state(), *o.state()) < STATE_TOL


@pytest.mark.parametrize("path", ["default", "no_resident", "streamed"])
def test_pose_with_more_edges_than_a_chunk(gpu, monkeypatch, path):
    """A pose that observes more landmarks than a chunk holds (> 512 LM edges) is split over several chunks: its
    row sums are carried from chunk to chunk in every kernel (unit, streamed, resident).  24 poses x 650 edges."""
    from oracle.oracle import OracleGraph
    from toyslam_b200 import synth
    payload = synth.to_payload(synth.make_graph(24, 800, 650, "manhattan", seed=3))
    for k in ("TSB_NO_RESIDENT", "TSB_STREAM_MIN_UNITS", "TSB_NO_STREAM"):
        monkeypatch.delenv(k, raising=False)
    if path != "default":
        monkeypatch.setenv("TSB_NO_RESIDENT", "1")
    if path == "streamed":
        monkeypatch.setenv("TSB_STREAM_MIN_UNITS", "1")
    o = OracleGraph(payload)
    traj, its, code = o.optimize_sparse(5, False, tol=1e-14, max_pcg=20000, threads=0)
    g = gpu(iterations=5, pcg_tol=1e-14, pcg_max_iters=20000, enable_exits=False); c = g.load_wire(payload)
    assert c.n_chunks > c.n_blocks >= 24                 # every pose is a unit of two chunks
    rep, per = g.optimize()
    got = np.array([p.chi2 for p in per])
    assert np.max(np.abs(got - traj) / np.abs(traj)) < CHI2_TOL
    assert state_diff(*g.state(), *o.state()) < STATE_TOL
    if path == "streamed":
        assert g.refresh_counts().pcg_path == 3


@pytest.mark.parametrize("kind", ["empty", "one_fixed_pose", "odom_only_streamed"])
def test_degenerate_graphs(gpu, monkeypatch, kind):
    """Empty request, a single fixed pose without edges, and an ODOM-only chain through the streamed kernels: no
    crash, state and reply as the oracle's."""
    from oracle.oracle import OracleGraph
    from toyslam_b200 import synth, wire
    for k in ("TSB_NO_RESIDENT", "TSB_STREAM_MIN_UNITS", "TSB_NO_STREAM"):
        monkeypatch.delenv(k, raising=False)
    z = lambda *shape, dt=np.float32: np.zeros(shape, dt)
    if kind == "empty":
        payload = wire.encode_request(z(0, dt=np.uint32), z(0, 3), z(0, dt=np.uint32), z(0, 2), z(0, 2, dt=np.uint32), z(0, 9), z(0, 3),
                                      z(0, 2, dt=np.uint32), z(0, 2), z(0, 2), z(0, dt=np.uint32))
    elif kind == "one_fixed_pose":
        payload = wire.encode_request(np.array([7], np.uint32), np.array([[1.0, 2.0, 0.5]], np.float32), z(0, dt=np.uint32), z(0, 2),
                                      z(0, 2, dt=np.uint32), z(0, 9), z(0, 3), z(0, 2, dt=np.uint32), z(0, 2), z(0, 2), np.array([7], np.uint32))
    else:
        payload = load_golden("g_odom_only")["request"].tobytes()
        monkeypatch.setenv("TSB_NO_RESIDENT", "1"); monkeypatch.setenv("TSB_STREAM_MIN_UNITS", "1")
    o = OracleGraph(payload); o.optimize_dense(3, True)
    g = gpu(iterations=3, pcg_tol=1e-14); g.load_wire(payload); g.optimize()
    a, b = wire.decode_reply(g.reply()), wire.decode_reply(o.reply())
    assert set(a["vertices"]) == set(b["vertices"]) and a["fixed"] == b["fixed"] and len(a["edges"]) == len(b["edges"])
    for vid, (t, v) in a["vertices"].items():
        assert t == b["vertices"][vid][0] and np.allclose(v, b["vertices"][vid][1], rtol=0, atol=2e-6)


def test_reply_matches_oracle_reply(gpu):
    from oracle.oracle import OracleGraph
    from toyslam_b200 import wire
    d = load_golden("g_loops"); payload = d["request"].tobytes()
    o = OracleGraph(payload); o.optimize_dense(5, True)
    g = gpu(iterations=5, pcg_tol=1e-14); g.load_wire(payload); g.optimize()
    a, b = wire.decode_reply(g.reply()), wire.decode_reply(o.reply())
    assert set(a["vertices"]) == set(b["vertices"]) and a["fixed"] == b["fixed"] and len(a["edges"]) == len(b["edges"])
    for vid, (t, v) in a["vertices"].items():
        assert t == b["vertices"][vid][0] and np.allclose(v, b["vertices"][vid][1], rtol=0, atol=2e-6)
    key = lambda e: (e[0], e[1], e[2], e[3].tobytes(), e[4].tobytes())
    assert sorted(map(key, a["edges"])) == sorted(map(key, b["edges"]))


@pytest.mark.parametrize("name", ["c1", "g_loops"])
def test_reply_matches_reference_decoded_fixture(gpu, name):
    """The reply the CUDA path writes against tests/golden/*_reply.npz: an oracle reply as decoded by the REFERENCE's own
    decoder (python/remote/bytes_to_graph.py:49-109, tests/make_reply_fixture.py).  Same vertex ids in the same order,
    same types, every edge and fixed id echoed bit for bit, values within the fp32 wire precision."""
    from toyslam_b200 import wire
    f = load_golden(name + "_reply"); d = load_golden(name)
    g = gpu(iterations=int(d["iterations"]), pcg_tol=1e-14, pcg_max_iters=5000, enable_exits=bool(d["exits"]))
    g.load_wire(d["request"].tobytes()); g.optimize()
    r = wire.decode_reply(g.reply())
    assert np.array_equal(np.array(r["order"], np.uint32), f["vertex_id"]) and sorted(r["fixed"]) == list(f["fixed"])
    for vid, vt, val in zip(f["vertex_id"], f["vertex_type"], f["vertex_value"]):
        t, v = r["vertices"][int(vid)]
        assert t == vt and np.allclose(v[:2], val[:2], rtol=0, atol=2e-6 * max(1.0, np.max(np.abs(val[:2]))))
        if t == 0:
            assert abs(np.arctan2(np.sin(v[2] - val[2]), np.cos(v[2] - val[2]))) < 2e-6
    assert len(r["edges"]) == len(f["edge_type"])
    for k, (et, a, b, meas, inf) in enumerate(r["edges"]):
        assert (et, a, b) == (f["edge_type"][k], f["edge_ij"][k, 0], f["edge_ij"][k, 1])
        assert np.array_equal(meas.ravel(), f["edge_meas"][k, :meas.size]) and np.array_equal(inf, f["edge_inf"][k, :len(inf)])


def test_load_soa_equals_load_wire(gpu):
    from toyslam_b200 import synth
    gr = synth.make_config("c2", scale=0.03, seed=5)
    g1 = gpu(iterations=3, enable_exits=False); g1.load_wire(synth.to_payload(gr)); g1.optimize()
    g2 = gpu(iterations=3, enable_exits=False); g2.load_soa(gr); g2.optimize()
    assert np.array_equal(g1.state()[0], g2.state()[0]) and np.array_equal(g1.state()[1], g2.state()[1])


def test_malformed_payload_returns_parse_error(gpu):
    from toyslam_b200.capi import TsbError
    buf = load_golden("g_small")["request"].tobytes()
    g = gpu()
    for cut in (0, 3, 17, len(buf) // 2, len(buf) - 1):
        with pytest.raises(TsbError) as ei:
            g.load_wire(buf[:cut])
        assert ei.value.code == 2
    g.load_wire(buf)                      # the context survives a rejected payload
    assert g.linearize().chi2 > 0


def test_full_size_c2_properties(gpu, monkeypatch):
    """BASELINE config 2 at full size: chi2 falls monotonically over the first iterations, the PCG
    reaches its tolerance, and the whole run agrees with the sparse oracle on chi2."""
    from oracle.oracle import OracleGraph
    from toyslam_b200 import synth
    payload = synth.to_payload(synth.make_config("c2"))
    g = gpu(iterations=6, pcg_tol=1e-12, pcg_max_iters=20000, enable_exits=False); g.load_wire(payload)
    rep, per = g.optimize()
    chi = np.array([p.chi2 for p in per])
    assert np.all(np.diff(chi) < 0)
    assert all(p.pcg_rel_residual <= 1e-12 for p in per)
    o = OracleGraph(payload)
    traj, _, _ = o.optimize_sparse(6, False, tol=1e-13, max_pcg=40000, threads=0)
    assert np.max(np.abs(chi - traj) / traj) < CHI2_TOL
    # state as well: the same request solved to 1e-13 by both sides (GPU: two-level / smoother-block PCG, oracle: block-
    # Jacobi PCG).  Each stops on its own preconditioned norm, so at 34 000 unknowns the two states agree to ~1e-9, not to
    # the last bit of the 1e-9 bar that the golden graphs (direct solve) are held to: asserted at 1e-8 here, and the
    # GPU step is checked against the oracle's own full system below (no solver on the oracle side at all)
    g13 = gpu(iterations=6, pcg_tol=1e-13, pcg_max_iters=40000, enable_exits=False); g13.load_wire(payload); g13.optimize()
    assert state_diff(*g13.state(), *o.state()) < 10 * STATE_TOL
    o.set_state(*g13.state())
    g13.linearize(); g13.solve()
    res, chi2_o = o.residual_sparse(*g13.delta())
    assert res <= 1e-10, res
    # fine level of the resident kernel: exact 34-pose Schur blocks (the default starts a request with them and
    # falls back to the 3x3 pose blocks once a probe solve says they stopped paying) against 3x3 blocks throughout
    # (TSB_SMOOTHER=0) and smoother blocks throughout (=1): same trajectory, well under 60 % of the PCG iterations on
    # the first solve (247 -> 101 measured)
    runs = {}
    for mode in ("0", "1"):
        monkeypatch.setenv("TSB_SMOOTHER", mode)
        g2 = gpu(iterations=6, pcg_tol=1e-12, pcg_max_iters=20000, enable_exits=False); g2.load_wire(payload)
        rep2, per2 = g2.optimize()
        assert np.max(np.abs(np.array([p.chi2 for p in per2]) - chi) / chi) < CHI2_TOL
        assert state_diff(*g2.state(), *g.state()) < STATE_TOL
        runs[mode] = [p.pcg_iters for p in per2]
    assert runs["1"][0] < 0.6 * runs["0"][0], runs
    assert per[0].pcg_iters == runs["1"][0], (per[0].pcg_iters, runs)


@pytest.mark.parametrize("case", ["g_small", "c3_small", "c5_small"])
def test_streaming_linearize_is_bit_identical(gpu, monkeypatch, case):
    """HBM-sized graphs take the TMA-streamed persistent linearize kernel; forced on small graphs here it must
    reproduce the one-unit-per-block kernel bit for bit (same device functions, same summation orders)."""
    from toyslam_b200 import synth
    if case == "g_small":
        payload = load_golden("g_small")["request"].tobytes()
    else:
        payload = synth.to_payload(synth.make_config(case[:2], scale=0.04, seed=5))
    out = []
    for env in ({"TSB_NO_STREAM": "1"}, {"TSB_STREAM_MIN_UNITS": "1"}):
        for k in ("TSB_NO_STREAM", "TSB_STREAM_MIN_UNITS"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        g = gpu(iterations=1); g.load_wire(payload)
        st = g.linearize()
        out.append((st.chi2, g.system()))
    for other in out[1:]:
        assert abs(out[0][0] - other[0]) <= 1e-12 * abs(out[0][0])
        for k in out[0][1]:
            assert np.array_equal(out[0][1][k], other[1][k]), k


def test_resident_pcg_matches_streaming_pcg(gpu, monkeypatch):
    """Small graphs run the shared-memory-resident PCG kernel (3 grid barriers per iteration); it must agree with
    the streamed kernel with the multilevel hierarchy (TSB_NO_RESIDENT=1), with block-Jacobi only (TSB_COARSE=0), with
    the block-synchronous persistent kernel and two-level aggregates (TSB_NO_RESIDENT=1 TSB_NO_STREAM=1) and with
    dense smoother blocks / 3x3 pose blocks throughout as the fine level (TSB_SMOOTHER=1 / 0; the default adapts)."""
    from toyslam_b200 import synth
    payload = synth.to_payload(synth.make_config("c2", scale=0.3, seed=4))
    res = []
    for env in ({}, {"TSB_NO_RESIDENT": "1"}, {"TSB_COARSE": "0"}, {"TSB_NO_RESIDENT": "1", "TSB_NO_STREAM": "1"}, {"TSB_SMOOTHER": "1"}, {"TSB_SMOOTHER": "0"}):
        for k in ("TSB_NO_RESIDENT", "TSB_COARSE", "TSB_NO_STREAM", "TSB_SMOOTHER"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        g = gpu(iterations=4, pcg_tol=1e-13, pcg_max_iters=20000, enable_exits=False); g.load_wire(payload)
        rep, per = g.optimize()
        assert all(p.pcg_rel_residual <= 1e-13 for p in per)
        res.append((np.array([p.chi2 for p in per]), g.state(), rep.pcg_iters_total))
    for other in res[1:]:
        assert np.max(np.abs(res[0][0] - other[0]) / res[0][0]) < CHI2_TOL
        assert state_diff(*res[0][1], *other[1]) < STATE_TOL
    assert res[0][2] < res[2][2]            # the coarse level pays for itself
    assert res[1][2] < res[2][2] and res[3][2] < res[2][2]      # so do the hierarchy and the uniform aggregates
    # dense smoother blocks (k_sb_build, TSB_SMOOTHER=1): at this size a block is one 20-pose aggregate and the count
    # stays where it was (the full-size check is in test_full_size_c2_properties)
    assert res[4][2] <= 1.1 * res[5][2], (res[4][2], res[5][2])


@pytest.mark.parametrize("scale", [0.45, 0.65])
def test_multilevel_pcg_matches_two_level_and_block_jacobi(gpu, monkeypatch, scale):
    """Graphs beyond the resident kernel run the streamed PCG with the multilevel hierarchy (mlevel.cuh: hat-interpolated
    rigid-body nodes every 4 / 16 / 64 ... poses, dense top level).  It must agree with the two-level method on uniform
    aggregates (TSB_ML=0: streamed; TSB_NO_STREAM=1: block-synchronous persistent kernel; TSB_PCG_PHASED=1: one launch
    per phase) and with plain block-Jacobi (TSB_COARSE=0), and need fewer iterations than any of them."""
    from toyslam_b200 import synth
    payload = synth.to_payload(synth.make_config("c3", scale=scale, seed=3))
    envs = [{}, {"TSB_ML": "0"}, {"TSB_NO_STREAM": "1"}, {"TSB_PCG_PHASED": "1"}, {"TSB_COARSE": "0"}]
    res = []
    for env in envs:
        for k in ("TSB_ML", "TSB_NO_STREAM", "TSB_PCG_PHASED", "TSB_COARSE"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        g = gpu(iterations=2, pcg_tol=1e-13, pcg_max_iters=60000, enable_exits=False); c = g.load_wire(payload)
        if not env:
            assert c.n_ml_levels >= 3 and 512 < c.n_ml_top <= 1024 and c.n_aggregates == 0, (c.n_ml_levels, c.n_ml_top)
        elif "TSB_COARSE" not in env:
            assert c.n_ml_levels == 0 and 3 * c.n_aggregates > 1536, c.n_aggregates
        rep, per = g.optimize()
        assert all(p.pcg_rel_residual <= 1e-13 for p in per)
        assert rep.pcg_not_converged == 0 and rep.coarse_failures == 0
        res.append((np.array([p.chi2 for p in per]), g.state(), rep.pcg_iters_total))
        g.close()
    for i, other in enumerate(res[1:], 1):
        assert np.max(np.abs(res[0][0] - other[0]) / res[0][0]) < CHI2_TOL
        # Two solves that both reach 1e-13 in their own preconditioned residual still differ by cond(S) x rounding: at
        # 45-65 k poses the kernels agree to ~2e-9 among themselves (measured 1.6e-9) and to ~1e-7 with block-Jacobi,
        # which stops on a different norm.  The 1e-9 parity bar is asserted where a direct solve exists (golden graphs).
        assert state_diff(*res[0][1], *other[1]) < (10 * STATE_TOL if i < 4 else 100 * STATE_TOL), i
    assert abs(res[1][2] - res[2][2]) <= max(4, res[1][2] // 50)      # same aggregates, streamed vs block-synchronous
    assert res[0][2] < res[1][2] and 4 * res[0][2] < res[4][2], [r[2] for r in res]


@pytest.mark.parametrize("cfg", ["c3", "c5", "c4"])
def test_full_size_solve_has_a_small_true_residual(gpu, cfg):
    """Configs 3, 4, 5 at FULL size, where no dense solve and no converged CPU run exist: the step the streamed PCG
    kernel returns (poses and back-substituted landmarks) is fed to the oracle's own sparse linearisation of the same
    state, which evaluates the true residual || H delta - b || / || b || of the full system (one CPU mat-vec) and chi2.
    Checked at the loaded state and again at the state three GN iterations later."""
    from oracle.oracle import OracleGraph
    from toyslam_b200 import synth
    payload = synth.to_payload(synth.make_config(cfg, seed=0))
    g = gpu(iterations=3, pcg_tol=1e-12, pcg_max_iters=30000, enable_exits=False)
    c = g.load_wire(payload)
    assert c.pcg_path in (0, 3) and c.n_ml_levels >= 3
    o = OracleGraph(payload)
    for stage in range(2):
        st = g.linearize(); s2 = g.solve()
        dp, dl = g.delta()
        res, chi2 = o.residual_sparse(dp, dl)
        assert abs(st.chi2 - chi2) <= 1e-11 * chi2, (cfg, stage, st.chi2, chi2)
        assert res <= 1e-10, (cfg, stage, res, s2.pcg_iters, s2.pcg_rel_residual)
        if stage == 0:
            rep, per = g.optimize(3)
            assert rep.pcg_not_converged == 0 and rep.coarse_failures == 0
            assert per[-1].chi2 < per[0].chi2
            o.set_state(*g.state())
    g.close()


def test_streamed_pcg_matches_persistent_kernel(gpu, monkeypatch):
    """HBM-sized graphs on one GPU run k_pcg_stream (edge passes through the TMA ring, register prefetch); forced on
    a small graph it must agree with the block-synchronous persistent kernel: same edge-pass arithmetic, only p'Sp
    is summed per block instead of per unit."""
    from toyslam_b200 import synth
    payload = synth.to_payload(synth.make_config("c3", scale=0.04, seed=9))
    res = []
    for env in ({"TSB_NO_STREAM": "1"}, {"TSB_STREAM_MIN_UNITS": "1", "TSB_ML": "0"}):
        for k in ("TSB_NO_STREAM", "TSB_STREAM_MIN_UNITS", "TSB_ML"):
            monkeypatch.delenv(k, raising=False)
        monkeypatch.setenv("TSB_NO_RESIDENT", "1")
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        g = gpu(iterations=4, pcg_tol=1e-13, pcg_max_iters=20000, enable_exits=False); g.load_wire(payload)
        rep, per = g.optimize()
        assert all(p.pcg_rel_residual <= 1e-13 for p in per)
        res.append((np.array([p.chi2 for p in per]), g.state(), rep.pcg_iters_total, g.refresh_counts().pcg_path))
    assert res[0][3] == 0 and res[1][3] == 3
    assert np.max(np.abs(res[0][0] - res[1][0]) / res[0][0]) < CHI2_TOL
    assert state_diff(*res[0][1], *res[1][1]) < STATE_TOL
    assert abs(res[0][2] - res[1][2]) <= max(4, res[0][2] // 50)


def test_context_reuse_across_requests_is_bit_reproducible(gpu):
    """The server keeps one context for every request (ConnectionHandler.h:18-20 rebuilds everything instead); device
    memory is recycled between graphs.  A -> B (larger) -> A -> C (smaller) -> A on one context: every run of A
    must reproduce the first one bit for bit (nothing may depend on what a buffer held before)."""
    from toyslam_b200 import synth
    pa = synth.to_payload(synth.make_config("c2", scale=0.05, seed=11))
    pb = synth.to_payload(synth.make_config("c3", scale=0.02, seed=12))
    pc = load_golden("g_small")["request"].tobytes()
    g = gpu(iterations=4, pcg_tol=1e-13, pcg_max_iters=20000, enable_exits=False)
    runs = []
    for payload in (pa, pb, pa, pc, pa):
        g.load_wire(payload)
        rep, per = g.optimize()
        state = g.state()
        g.linearize()                                   # H, b at the final state
        runs.append((np.array([p.chi2 for p in per]), state, g.system(), payload is pa))
    first = runs[0]
    for r in runs[1:]:
        if not r[3]:
            continue
        assert np.array_equal(first[0], r[0])
        assert np.array_equal(first[1][0], r[1][0]) and np.array_equal(first[1][1], r[1][1])
        for k in first[2]:
            assert np.array_equal(first[2][k], r[2][k]), k
    fresh = gpu(iterations=4, pcg_tol=1e-13, pcg_max_iters=20000, enable_exits=False); fresh.load_wire(pa)
    rep, per = fresh.optimize()
    assert np.array_equal(first[0], np.array([p.chi2 for p in per]))


def test_structure_reuse_for_repeated_topology(gpu):
    """A request with the topology of the previous one on the same context (same ids, edge lists and fixed ids; any
    values) keeps the assembly structure and the device index arrays: tsb_load_wire is several times faster and the
    results are bit-identical to a fresh context.  (The reference rebuilds everything per request: ConnectionHandler.h:18-20.)"""
    import time
    from toyslam_b200 import synth
    ga = synth.make_config("c3", scale=0.3, seed=21)
    pa = synth.to_payload(ga)
    gb = dict(ga); gb["pose_xyt"] = (ga["pose_xyt"] + np.float32(0.01)).astype(np.float32); gb["lm_meas"] = (ga["lm_meas"] * np.float32(1.001)).astype(np.float32)
    pb = synth.to_payload(gb)                                    # same topology, different values
    g = gpu(iterations=3, pcg_tol=1e-13, pcg_max_iters=20000, enable_exits=False)
    t = []
    outs = []
    for payload in (pa, pa, pa, pb):
        t0 = time.perf_counter(); c = g.load_wire(payload); t.append(time.perf_counter() - t0)
        rep, per = g.optimize()
        outs.append((np.array([p.chi2 for p in per]), *g.state(), c.structure_reuses))
    assert [o[3] for o in outs] == [0, 1, 2, 3] or [o[3] for o in outs] == [0, 0, 1, 2], [o[3] for o in outs]     # the arena may settle on the second load
    assert t[2] * 2 < t[0], t            # 47 -> 17 ms at this size (parse + value upload remain); C3: ~250 -> ~65 ms per request incl. the reply
    for o in outs[1:3]:
        assert np.array_equal(o[0], outs[0][0]) and np.array_equal(o[1], outs[0][1]) and np.array_equal(o[2], outs[0][2])
    fresh = gpu(iterations=3, pcg_tol=1e-13, pcg_max_iters=20000, enable_exits=False); fresh.load_wire(pb)
    rep, per = fresh.optimize()
    assert np.array_equal(outs[3][0], np.array([p.chi2 for p in per]))
    ps, ls = fresh.state()
    assert np.array_equal(outs[3][1], ps) and np.array_equal(outs[3][2], ls)
    assert not np.array_equal(outs[3][0], outs[0][0])


def test_phased_pcg_path_matches_persistent(gpu, monkeypatch):
    """world == 1 normally runs the persistent cooperative PCG kernel; the one-launch-per-phase path
    (used when collectives sit between the phases) must give the same answer."""
    from toyslam_b200 import synth
    payload = synth.to_payload(synth.make_config("c2", scale=0.1, seed=2))
    res = []
    for phased, stream in (("0", None), ("1", None), ("1", "1")):      # the last: phased with the streamed edge-pass kernels
        monkeypatch.setenv("TSB_PCG_PHASED", phased)
        monkeypatch.delenv("TSB_STREAM_MIN_UNITS", raising=False)
        if stream:
            monkeypatch.setenv("TSB_STREAM_MIN_UNITS", stream)
        g = gpu(iterations=5, pcg_tol=1e-14, pcg_max_iters=20000, enable_exits=False); g.load_wire(payload)
        rep, per = g.optimize()
        res.append((np.array([p.chi2 for p in per]), g.state(), rep.kernel_launches))
    assert res[0][2] < res[1][2]
    for other in res[1:]:
        assert np.max(np.abs(res[0][0] - other[0]) / res[0][0]) < CHI2_TOL
        assert state_diff(*res[0][1], *other[1]) < STATE_TOL


def test_two_ranks_match_one_rank(gpu):
    """Edge-sharded run over 2 GPUs (NCCL) against the single-GPU run and the oracle; skipped on a 1-GPU box."""
    import os, subprocess, sys, torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from conftest import ROOT
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
                        "127.0.0.1", "--master-port", "29741", os.path.join(ROOT, "tools", "mgpu_check.py")],
                       capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MGPU OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
    # the ODOM-dominated chain: the two ranks' hierarchies are joined by the cut edge in the replicated top operator, so the sharded
    # solve needs (nearly) the iterations of one GPU -- 1.3x without it.  Ill-conditioned: state to 1e-7 (two solves at 1e-13)
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
                        "127.0.0.1", "--master-port", "29742", os.path.join(ROOT, "tools", "mgpu_check.py"), "--workload", "c5", "--scale", "0.02",
                        "--iters", "3", "--state-tol", "1e-7", "--oracle", "0", "--max-iter-ratio", "1.1"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "MGPU OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


def test_graph_optimizer_server_end_to_end(gpu):
    """`graph_optimizer 127.0.0.1 PORT 50 gpu cuda` serves the reference client's request for the default
    simulation (C1) over the reference's framing; the reply matches the oracle's `cpu eigen` result."""
    import os, socket, subprocess, time
    from conftest import ROOT
    from oracle.oracle import OracleGraph
    from toyslam_b200 import wire
    exe = os.path.join(ROOT, "toyslam_b200", "graph_optimizer")
    assert os.path.exists(exe), "build it with make -C toyslam_b200/csrc all"
    payload = load_golden("c1")["request"].tobytes()
    port = 18000 + os.getpid() % 2000
    srv = subprocess.Popen([exe, "127.0.0.1", str(port), "50", "gpu", "cuda"], stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    try:
        s = None
        for _ in range(150):
            try:
                s = socket.create_connection(("127.0.0.1", port), timeout=1); break
            except OSError:
                time.sleep(0.2)
        assert s is not None, "server did not come up"
        s.settimeout(120)
        replies = []
        for _ in range(2):                                   # the connection stays open for the next request
            s.sendall(wire.frame(payload))
            n = int(np.frombuffer(s.recv(4, socket.MSG_WAITALL), "<u4")[0])
            buf = b""
            while len(buf) < n:
                buf += s.recv(n - len(buf))
            replies.append(buf)
        s.close()
    finally:
        srv.terminate(); out = srv.communicate(timeout=10)[0]
    assert replies[0] == replies[1]
    assert "Plateau: NO MORE OPT" in out and "Summary() error" in out      # the reference's stdout lines
    o = OracleGraph(payload); o.optimize_dense(50, True)
    a, b = wire.decode_reply(replies[0]), wire.decode_reply(o.reply())
    assert set(a["vertices"]) == set(b["vertices"]) and len(a["edges"]) == len(b["edges"])
    for vid, (t, v) in a["vertices"].items():
        assert np.allclose(v, b["vertices"][vid][1], rtol=0, atol=2e-6)


def test_graph_optimizer_server_drives_two_gpus(gpu):
    """TSB_GPUS=2: the one server process creates a context per GPU and shards a 10 k-pose request over them; the reply
    equals the single-GPU result (tools/server_check.py).  Skipped on a 1-GPU box."""
    import os, subprocess, sys, torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from conftest import ROOT
    r = subprocess.run([sys.executable, os.path.join(ROOT, "tools", "server_check.py"), "--gpus", "2"], capture_output=True, text=True, timeout=600)
    assert r.returncode == 0 and "SERVER MGPU OK" in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]
