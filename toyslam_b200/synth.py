"""Synthetic Manhattan-world / chain graphs as ToySlam request payloads.

Not in the reference (it has a 150-step simulator only, python/slam_main.py).  Noise and
information values are the reference simulator's effective ones: it passes *variances* as
std-devs to np.random.normal (slam_main.py:41-51,138-141; slam_helper.py:9-14), so
  odometry noise  N(0, 0.25) m on x,y and N(0, 0.01536) rad on theta, INF = diag(4, 4, 65.1)
  lidar noise     N(0, 0.0225) m on the local x,y,                   INF = diag(44.4, 44.4)
Initial poses: dead-reckoned noisy odometry (slam_main.py:146); landmark initial value: its first
observation mapped through the observing pose's initial estimate (slam/graph2d.py:30-32).
Every value is rounded through fp32 (the wire type) before it is used anywhere.
"""
import numpy as np
from . import wire

CONFIGS = {
    # name: (P, L, lm_per_pose, kind)
    "c2": (10_000, 2_000, 15, "manhattan"),
    "c3": (100_000, 20_000, 20, "manhattan"),
    "c4": (1_000_000, 200_000, 19, "manhattan"),
    "c5": (1_000_000, 10_000, 10, "chain"),
}

ODOM_VAR = np.array([0.25, 0.25, np.deg2rad(7.1) ** 2])
LIDAR_VAR = 0.15 ** 2


def _compose(x, y, t, dx, dy, dt):
    c, s = np.cos(t), np.sin(t)
    return x + c * dx - s * dy, y + s * dx + c * dy, t + dt


def _dead_reckon(steps):
    """steps (n,3) local motions -> (n+1,3) poses starting at the origin."""
    n = len(steps)
    th = np.concatenate([[0.0], np.cumsum(steps[:, 2])])
    c, s = np.cos(th[:-1]), np.sin(th[:-1])
    x = np.concatenate([[0.0], np.cumsum(c * steps[:, 0] - s * steps[:, 1])])
    y = np.concatenate([[0.0], np.cumsum(s * steps[:, 0] + c * steps[:, 1])])
    return np.stack([x, y, th], 1)


def make_graph(P, L, lm_per_pose, kind="manhattan", seed=0, noise_scale=1.0):
    """Returns a dict of SoA arrays (fp32-rounded values held as fp32) describing one graph."""
    from scipy.spatial import cKDTree
    rng = np.random.default_rng(seed)
    # ground-truth motion: 1 m forward then turn 0 / +90 / -90 deg with p = .8 / .1 / .1
    if kind == "manhattan":
        turn = rng.choice([0.0, np.pi / 2, -np.pi / 2], size=P - 1, p=[0.8, 0.1, 0.1])
    else:
        turn = np.zeros(P - 1)
    gt_steps = np.stack([np.ones(P - 1), np.zeros(P - 1), turn], 1)
    gt = _dead_reckon(gt_steps)
    noisy = gt_steps + rng.normal(0.0, 1.0, (P - 1, 3)) * (ODOM_VAR * noise_scale)
    noisy = noisy.astype(np.float32)                                  # the measurement as sent
    init = _dead_reckon(noisy.astype(np.float64))
    init[:, 2] = np.arctan2(np.sin(init[:, 2]), np.cos(init[:, 2]))
    # measurement matrix [[c,-s,x],[s,c,y],[0,0,1]] in fp32 (graph_to_bytes.py:25-28)
    c, s = np.cos(noisy[:, 2].astype(np.float64)), np.sin(noisy[:, 2].astype(np.float64))
    Z = np.zeros((P - 1, 9), np.float32)
    Z[:, 0], Z[:, 1], Z[:, 2] = c, -s, noisy[:, 0]
    Z[:, 3], Z[:, 4], Z[:, 5] = s, c, noisy[:, 1]
    Z[:, 8] = 1.0
    od_ij = np.stack([np.arange(P - 1), np.arange(1, P)], 1).astype(np.uint32)
    od_w = np.broadcast_to((1.0 / ODOM_VAR).astype(np.float32), (P - 1, 3))

    if kind == "manhattan":
        # landmarks scattered within 3 m of the trajectory (SURVEY said "uniform over the bounding
        # box"; a random walk covers little of its box, which left half the landmarks unobserved)
        lm_gt = gt[rng.integers(0, P, L), :2] + rng.uniform(-3.0, 3.0, (L, 2))
        _, nn = cKDTree(lm_gt).query(gt[:, :2], k=lm_per_pose)
        nn = np.sort(nn.reshape(P, lm_per_pose), axis=1)
        pose_of = np.repeat(np.arange(P), lm_per_pose)
        lm_of = nn.reshape(-1)
    else:
        # chain: every landmark is seen from `lm_per_pose` poses spread >= P/(2*lm_per_pose) apart
        stride = P // lm_per_pose
        base = rng.integers(0, stride, L)
        pose_of = (base[:, None] + stride * np.arange(lm_per_pose)[None, :]).reshape(-1)
        lm_of = np.repeat(np.arange(L), lm_per_pose)
        lm_gt = gt[base, :2] + rng.uniform(-5.0, 5.0, (L, 2))
        o = np.lexsort((lm_of, pose_of)); pose_of, lm_of = pose_of[o], lm_of[o]
    # observation: landmark in the (true) pose frame + noise, sent as (range, bearing)
    d = lm_gt[lm_of] - gt[pose_of, :2]
    ct, st = np.cos(gt[pose_of, 2]), np.sin(gt[pose_of, 2])
    loc = np.stack([ct * d[:, 0] + st * d[:, 1], -st * d[:, 0] + ct * d[:, 1]], 1)
    loc += rng.normal(0.0, LIDAR_VAR * noise_scale, loc.shape)
    meas = np.stack([np.hypot(loc[:, 0], loc[:, 1]), np.arctan2(loc[:, 1], loc[:, 0])], 1).astype(np.float32)
    # landmark initial value from its first observation through the initial pose estimate
    first = np.full(L, -1, np.int64)
    e_idx = np.arange(len(lm_of))[::-1]
    first[lm_of[::-1]] = e_idx                                        # smallest edge index wins
    seen = first >= 0
    m = meas[first[seen]].astype(np.float64); pi = init[pose_of[first[seen]]]
    lx, ly = m[:, 0] * np.cos(m[:, 1]), m[:, 0] * np.sin(m[:, 1])
    lm_init = np.zeros((L, 2))
    lm_init[seen, 0] = pi[:, 0] + np.cos(pi[:, 2]) * lx - np.sin(pi[:, 2]) * ly
    lm_init[seen, 1] = pi[:, 1] + np.sin(pi[:, 2]) * lx + np.cos(pi[:, 2]) * ly
    # unseen landmarks are dropped (the reference client only sends observed ones, slam_main.py:181-185)
    remap = np.cumsum(seen) - 1
    Lk = int(seen.sum())
    lm_ij = np.stack([pose_of, P + remap[lm_of]], 1).astype(np.uint32)
    return dict(
        pose_id=np.arange(P, dtype=np.uint32), pose_xyt=init.astype(np.float32),
        lm_id=(P + np.arange(Lk)).astype(np.uint32), lm_xy=lm_init[seen].astype(np.float32),
        od_ij=od_ij, od_Z=Z, od_w=np.ascontiguousarray(od_w),
        lm_ij=lm_ij, lm_meas=meas,
        lm_w=np.full((len(lm_ij), 2), np.float32(1.0 / LIDAR_VAR), np.float32),
        fixed_ids=np.array([0], np.uint32),
    )


def to_payload(g) -> bytes:
    return wire.encode_request(g["pose_id"], g["pose_xyt"], g["lm_id"], g["lm_xy"], g["od_ij"], g["od_Z"],
                               g["od_w"], g["lm_ij"], g["lm_meas"], g["lm_w"], g["fixed_ids"])


def make_config(name, seed=0, scale=1.0, noise_scale=1.0):
    """Named config (c2..c5), optionally shrunk by `scale` (same edges-per-pose ratio)."""
    P, L, k, kind = CONFIGS[name]
    P = max(4, int(round(P * scale))); L = max(k, int(round(L * scale)))
    return make_graph(P, L, k, kind, seed, noise_scale)


def counts(g):
    return dict(P=len(g["pose_id"]), L=len(g["lm_id"]), Eo=len(g["od_ij"]), El=len(g["lm_ij"]))
