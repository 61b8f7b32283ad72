#!/usr/bin/env python
"""Generates tests/golden/*.npz|bin by IMPORTING the reference's Python implementation.

Run in the build container only (needs /root/reference; the GPU box never reads it):
    python tools/make_golden.py

What comes from the reference (imported, unmodified, with an empty `matplotlib` stub because
python/tools.py:2 imports it without using it on the math path):
  * request bytes            python/remote/graph_to_bytes.py:32-67
  * per-edge e, A, B         python/optimizer/edges2d.py:21-53,65-78
  * H, b, chi2               python/optimizer/graph_optimizer.py:94-156  (b sign flipped and the
                             fixed-vertex b zeroing undone to match OptimizerCpu.h:109-138)
  * reply decoding           python/remote/bytes_to_graph.py:49-109 -- pinned by tests/make_reply_fixture.py (*_reply.npz)
  * C1 default-sim graph     python/slam_main.py:26-187 re-stated headlessly (the script opens a
                             matplotlib window at import), calling the reference's own lidar /
                             environment / slam helpers with np.random.seed(0)
What is restated here (no runnable reference artefact exists): the OptimizerCpu.h:70-182 outer
loop (step 0.2, exits), driven by the reference's Python linearisation and numpy.linalg.solve.
"""
import os
import sys
import types
import io
import contextlib
import numpy as np

REF = "/root/reference/python"
OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")

for name in ("matplotlib", "matplotlib.pyplot"):
    sys.modules.setdefault(name, types.ModuleType(name))
sys.path.insert(0, REF)

from optimizer.opt_graph import OptGraph                       # noqa: E402
from optimizer.vertices import VertexPose2d, Vertex2d          # noqa: E402
from optimizer.edges2d import EdgeOdometry2d, EdgeLandmark2d   # noqa: E402
from optimizer.graph_optimizer import GraphOptimizer           # noqa: E402
from remote.graph_to_bytes import graph_to_bytes               # noqa: E402
from tools import mat_to_angle_2d, angle_to_mat_2d             # noqa: E402


def quiet(fn, *a, **k):
    with contextlib.redirect_stdout(io.StringIO()):
        return fn(*a, **k)


def f32(x):
    return np.asarray(x, np.float32).astype(np.float64)


def pose_mat(x, y, th):
    c, s = np.cos(th), np.sin(th)
    return np.array([[c, -s, x], [s, c, y], [0, 0, 1.0]])


def random_graph(seed, P, L, n_lm_edges, loop_closures=0, fix_lm=False, scrambled_ids=False, outliers=False,
                 est_noise=0.15, lm_noise=0.3, info_scale=1.0, meas_noise=1.0):
    """A random OptGraph whose every value is already fp32-representable, so the wire round trip
    is lossless and reference-Python and wire consumers see identical numbers."""
    rng = np.random.default_rng(seed)
    ids = np.arange(P + L)
    if scrambled_ids:
        ids = rng.permutation(10 * (P + L))[: P + L]
    pid, lid = ids[:P], ids[P:]
    gt = np.zeros((P, 3))
    for i in range(1, P):
        th = gt[i - 1, 2]
        gt[i] = [gt[i - 1, 0] + np.cos(th), gt[i - 1, 1] + np.sin(th), th + rng.normal(0, 0.4)]
    lm_gt = gt[rng.integers(0, P, L), :2] + rng.uniform(-4, 4, (L, 2))
    g = OptGraph()
    est = gt + rng.normal(0, est_noise, gt.shape)
    big = np.abs(est[:, 2]) > np.pi          # the client sends atan2(R10, R00): keep theta in (-pi, pi] so that fp32(theta) is what travels
    est[big, 2] = np.arctan2(np.sin(est[big, 2]), np.cos(est[big, 2]))
    est = f32(est)
    est[0] = f32(gt[0])
    order = rng.permutation(P + L) if scrambled_ids else np.arange(P + L)
    lm_est = f32(lm_gt + rng.normal(0, lm_noise, lm_gt.shape))
    for k in order:
        if k < P:
            g.add_vertex(int(pid[k]), VertexPose2d(pose_mat(*est[k])), k == 0)
        else:
            g.add_vertex(int(lid[k - P]), Vertex2d(lm_est[k - P].copy()), False)
    pairs = [(i, i + 1) for i in range(P - 1)]
    for _ in range(loop_closures):
        a, b = sorted(rng.choice(P, 2, replace=False)); pairs.append((int(a), int(b)))
    odom_inf = f32(np.diag([4.0, 4.0, 65.1]) * info_scale)
    for (a, b) in pairs:
        rel = np.linalg.inv(pose_mat(*gt[a])) @ pose_mat(*gt[b])
        th = np.arctan2(rel[1, 0], rel[0, 0]) + rng.normal(0, 0.02) * meas_noise
        Z = pose_mat(rel[0, 2] + rng.normal(0, 0.05) * meas_noise, rel[1, 2] + rng.normal(0, 0.05) * meas_noise, th)
        if outliers and rng.random() < 0.2:
            Z[0, 2] += 3.0
        g.add_edge(EdgeOdometry2d(int(pid[a]), int(pid[b]), f32(Z), odom_inf))
    lm_inf = f32(np.diag([44.4, 30.0]) * info_scale)
    for _ in range(n_lm_edges):
        a, l = int(rng.integers(0, P)), int(rng.integers(0, L))
        loc = (np.linalg.inv(pose_mat(*gt[a])) @ np.array([lm_gt[l, 0], lm_gt[l, 1], 1.0]))[:2] + rng.normal(0, 0.05, 2) * meas_noise
        if outliers and rng.random() < 0.2:
            loc += 2.0
        meas = f32([np.hypot(*loc), np.arctan2(loc[1], loc[0])])
        g.add_edge(EdgeLandmark2d(int(pid[a]), int(lid[l]), meas, lm_inf))
    used = {e.id_2 for e in g.edges if e.get_type() == 1}
    for l in range(L):                      # every landmark observed at least once (H must be regular)
        if int(lid[l]) not in used:
            a = int(rng.integers(0, P))
            loc = (np.linalg.inv(pose_mat(*gt[a])) @ np.array([lm_gt[l, 0], lm_gt[l, 1], 1.0]))[:2]
            g.add_edge(EdgeLandmark2d(int(pid[a]), int(lid[l]), f32([np.hypot(*loc), np.arctan2(loc[1], loc[0])]), lm_inf))
    if fix_lm:
        g.fix_vertex(int(lid[0]))
    return g


def canonical(g):
    """canonical vertex order used by oracle and product: poses by id, then landmarks by id."""
    p = sorted(v for v in g.vertices if g.vertices[v].get_type() == 0)
    l = sorted(v for v in g.vertices if g.vertices[v].get_type() == 1)
    off, o = {}, 0
    for v in p:
        off[v] = o; o += 3
    for v in l:
        off[v] = o; o += 2
    return p, l, off, o


def ref_linearize(g):
    """reference calculate_H_b -> (H, b, chi2) in canonical order with OptimizerCpu sign/prior rules."""
    opt = GraphOptimizer(g)
    err = quiet(opt.calculate_H_b)
    H, b = opt.H.copy(), opt.b.copy()
    # undo the Python-only `b[fixed] = 0` (graph_optimizer.py:150): recompute those rows from edges
    bfix = np.zeros_like(b)
    per_edge = []
    for e in g.edges:
        er, A, B = e.calc_error(g)
        per_edge.append((er, A, B))
    if g.fixed_vertices:
        delta = 1.5
        for e, (er, A, B) in zip(g.edges, per_edge):
            chi2 = er @ e.information @ er
            w = 1.0 if chi2 <= delta ** 2 else delta / np.sqrt(chi2)
            infw = (e.information * w) @ er
            for vid, J in ((e.id_1, A), (e.id_2, B)):
                if vid in g.fixed_vertices:
                    i = opt.vertex_ids_map[vid]
                    bfix[i:i + J.shape[1]] += J.T @ infw
        for vid in g.fixed_vertices:
            i = opt.vertex_ids_map[vid]; n = g.vertices[vid].get_dims()
            b[i:i + n] = bfix[i:i + n]
    p, l, off, n = canonical(g)
    perm = np.zeros(n, np.int64)
    for vid, i in opt.vertex_ids_map.items():
        d = g.vertices[vid].get_dims()
        perm[off[vid]:off[vid] + d] = np.arange(i, i + d)
    return H[np.ix_(perm, perm)], -b[perm], err, per_edge      # reference C++: b -= J^T INF e


def gn_loop(g, iterations, exits=True):
    """OptimizerCpu.h:70-182 driven by the reference Python linearisation."""
    p, l, off, n = canonical(g)
    prev, penalty, traj, code = -1.0, 0, [], 0
    for _ in range(iterations):
        H, b, err, _ = ref_linearize(g)
        traj.append(err)
        if prev > 0 and err > prev:
            penalty += 1
            if exits and penalty > 2:
                code = 1; break
        else:
            penalty = 0
        d = np.linalg.solve(H, b)
        for vid, o in off.items():
            v = g.vertices[vid]
            v.update(d[o:o + v.get_dims()] * 0.2)
        if exits and abs(err - prev) < 1e-3:
            code = 2; break
        if exits and np.linalg.norm(d) < 1e-3:
            code = 3; break
        prev = err
    pose = np.array([[g.vertices[v].position[0, 2], g.vertices[v].position[1, 2],
                      np.arctan2(g.vertices[v].position[1, 0], g.vertices[v].position[0, 0])] for v in p])
    lm = np.array([g.vertices[v].position for v in l]).reshape(-1, 2)
    return np.array(traj), code, pose, lm


def save_case(name, g, iterations, exits=True):
    req = quiet(graph_to_bytes, g)[4:]                # strip the size prefix: payload only
    H, b, err, per_edge = ref_linearize(g)
    p, l, off, n = canonical(g)
    od = [(e, pe) for e, pe in zip(g.edges, per_edge) if e.get_type() == 0]
    lm = [(e, pe) for e, pe in zip(g.edges, per_edge) if e.get_type() == 1]
    traj, code, pose, lmxy = gn_loop(g, iterations, exits)
    np.savez_compressed(
        os.path.join(OUT, name + ".npz"), request=np.frombuffer(req, np.uint8), H=H, b=b, chi2=err,
        od_e=np.array([pe[0] for _, pe in od]).reshape(-1, 3),
        lm_e=np.array([pe[0] for _, pe in lm]).reshape(-1, 2),
        lm_A=np.array([pe[1] for _, pe in lm]).reshape(-1, 2, 3),
        lm_B=np.array([pe[2] for _, pe in lm]).reshape(-1, 2, 2),
        pose_ids=np.array(p, np.uint32), lm_ids=np.array(l, np.uint32),
        traj=traj, exit_code=code, final_pose=pose, final_lm=lmxy, iterations=iterations, exits=int(exits))
    print(f"{name}: dim={n} edges={len(g.edges)} iters_run={len(traj)} exit={code} chi2[0]={err:.6g} chi2[-1]={traj[-1]:.6g}")


def default_sim_graph():
    """python/slam_main.py:26-187 without the viewer (np.random.seed(0), 150 robot steps)."""
    from lidar_sensor import calc_lidar_measurements
    from environment import load_env
    from slam.graph2d import Graph2d
    from slam.slam_helper import add_to_graph, motion_model
    np.random.seed(0)
    ROBOT_STEPS = 150
    lidar_fov, lidar_ray_step = np.deg2rad(120), np.deg2rad(6)
    LIDAR_NOISE = np.identity(2) * 0.15 ** 2
    ODOMETRY_NOISE = np.identity(3); ODOMETRY_NOISE[:2, :2] *= 0.5 ** 2; ODOMETRY_NOISE[2, 2] *= np.deg2rad(7.1) ** 2
    LIDAR_INF, ODOM_INF = np.linalg.inv(LIDAR_NOISE), np.linalg.inv(ODOMETRY_NOISE)
    environment, radius = load_env()
    state_mat = np.eye(3); state_mat[0, 2] = 5; state_mat[1, 2] = 15
    state_mat_gt = np.copy(state_mat)
    graph = Graph2d()
    lms, ids = calc_lidar_measurements(state_mat, environment, radius, lidar_fov, lidar_ray_step)
    pos_id = add_to_graph(graph, state_mat, lms, ids, LIDAR_NOISE, True)
    schedule = [(10, 3., 2.0), (20, 6., 0.9), (40, -6., 0.9), (60, 5., 0.8), (10 ** 9, 3., 0.7)]
    while graph.get_size() < ROBOT_STEPS:
        deg, fwd = next((d, f) for lim, d, f in schedule if pos_id < lim)
        T = np.eye(3); T[:2, :2] = angle_to_mat_2d(np.deg2rad(deg)); T[:2, 2] = [fwd, 0.0]
        state_mat_gt = motion_model(state_mat_gt, T)
        lms, ids = calc_lidar_measurements(state_mat_gt, environment, radius, lidar_fov, lidar_ray_step)
        if lms is None or ids is None:
            continue
        RT = np.copy(T)
        RT[0, 2] += np.random.normal(0, ODOMETRY_NOISE[0, 0])
        RT[1, 2] += np.random.normal(0, ODOMETRY_NOISE[1, 1])
        RT[:2, :2] = angle_to_mat_2d(mat_to_angle_2d(T[:2, :2]) + np.random.normal(0, ODOMETRY_NOISE[2, 2]))
        graph.get_pose(pos_id).set_odometry(RT)
        state_mat = motion_model(state_mat, RT)
        pos_id = add_to_graph(graph, state_mat, lms, ids, LIDAR_NOISE, False)
    # construct_optimizer_graph, slam_main.py:157-187
    g = OptGraph()
    positions = graph.get_positions()
    for i in range(len(positions)):
        g.add_vertex(positions[i].id, VertexPose2d(positions[i].position), positions[i].is_fixed)
    for i in range(len(positions)):
        if positions[i].odometry is not None:
            g.add_edge(EdgeOdometry2d(positions[i].id, positions[i].id + 1, positions[i].odometry, ODOM_INF))
    lm_glob, idmap = len(positions), {}
    for i in range(len(positions)):
        for lm_id in positions[i].landmark_measurements:
            if lm_id not in idmap:
                idmap[lm_id] = lm_glob; lm_glob += 1
            g.add_edge(EdgeLandmark2d(positions[i].id, idmap[lm_id], positions[i].landmark_measurements[lm_id], LIDAR_INF))
    for lm_id, pos in graph.get_landmarks().items():
        if lm_id in idmap:
            g.add_vertex(idmap[lm_id], Vertex2d(pos), False)
    return g


def roundtrip_f32(g):
    """The same graph with every value rounded through fp32, i.e. what a wire consumer sees."""
    r = OptGraph()
    for vid, v in g.vertices.items():
        if v.get_type() == 0:
            th = float(np.float32(mat_to_angle_2d(v.position[:2, :2])))
            r.add_vertex(vid, VertexPose2d(pose_mat(float(np.float32(v.position[0, 2])), float(np.float32(v.position[1, 2])), th)))
        else:
            r.add_vertex(vid, Vertex2d(f32(v.position)))
    for e in g.edges:
        cls = EdgeOdometry2d if e.get_type() == 0 else EdgeLandmark2d
        r.add_edge(cls(e.id_1, e.id_2, f32(e.measurement), f32(np.diag(np.diag(e.information)))))
    for vid in g.fixed_vertices:
        r.fix_vertex(vid)
    return r


def main():
    os.makedirs(OUT, exist_ok=True)
    save_case("g_tiny", random_graph(1, P=3, L=2, n_lm_edges=4), 10)
    save_case("g_small", random_graph(2, P=12, L=8, n_lm_edges=60), 30)
    save_case("g_loops", random_graph(3, P=30, L=15, n_lm_edges=200, loop_closures=6, scrambled_ids=True), 30)
    save_case("g_outliers", random_graph(4, P=25, L=12, n_lm_edges=150, loop_closures=3, outliers=True, fix_lm=True), 25, exits=False)
    save_case("g_odom_only", random_graph(5, P=20, L=0, n_lm_edges=0, loop_closures=4), 20)
    # the two early exits no other fixture reaches (OptimizerCpu.h:140-149 "Error is getting worse", :173-177 "CONVERGED"):
    # a start far from the optimum with loop closures (the error rises three times in a row), once ODOM-only and once
    # with landmarks, and a near-converged start with stiff information (|delta| < 1e-3 while the error still moves)
    save_case("g_worse", random_graph(102, P=8, L=0, n_lm_edges=0, loop_closures=6, outliers=True, est_noise=3.0), 40)
    save_case("g_worse_lm", random_graph(100, P=8, L=3, n_lm_edges=8, loop_closures=5, outliers=True, est_noise=3.0, lm_noise=3.0), 40)
    save_case("g_converged", random_graph(20, P=10, L=6, n_lm_edges=40, est_noise=2e-4, lm_noise=2e-4, info_scale=1e3, meas_noise=0.02), 30)
    # C1: the reference's default simulation, as sent by its own client
    g = default_sim_graph()
    req = quiet(graph_to_bytes, g)
    with open(os.path.join(OUT, "c1_request.bin"), "wb") as f:
        f.write(req[4:])
    g32 = roundtrip_f32(g)
    assert quiet(graph_to_bytes, g32) == req, "fp32 round trip changed the request bytes"
    save_case("c1", g32, 50)
    nl = sum(1 for v in g.vertices.values() if v.get_type() == 1)
    print(f"c1: vertices={len(g.vertices)} (landmarks {nl}) edges={len(g.edges)} bytes={len(req) - 4}")


if __name__ == "__main__":
    main()
