"""ToySlam wire format, vectorised (numpy structured records).

Byte layout (little-endian, fp32 payload) follows the reference codec:
  request  : python/remote/graph_to_bytes.py:32-67, parsed by remote/serialization/DeserializeGraph.h:9-174
  reply    : remote/serialization/SerializeGraph.h:8-72 + SerializeGraphFuncCpu.h:10-65,
             parsed by python/remote/bytes_to_graph.py:49-109
  framing  : [u32 nbytes][payload] (remote/app/ConnectionHandlerGraph.h:16-29)

payload := u32 V, V x vertex, u32 E, E x edge, u32 F, F x u32
vertex  := u32 id, u32 type (0 Se2: x y theta | 1 Point2: x y)
edge    := u32 type, u32 id1, u32 id2, meas, info
  ODOM request/reply meas: u32 3, u32 3, 9 f32       info: u32 0, u32 3, 3 f32   (76 B)
  LM   request meas: u32 0, u32 2, range, bearing    info: u32 0, u32 2, 2 f32   (44 B)
  LM   reply   meas: u32 2, u32 1, range, bearing

This module is host-side tooling (bench / tests / synthetic generator).  The product parser
and encoder are the C++ ones behind tsb_load_wire / tsb_write_reply.
"""
import numpy as np

SE2_REC = np.dtype([("id", "<u4"), ("type", "<u4"), ("x", "<f4"), ("y", "<f4"), ("t", "<f4")])
PT2_REC = np.dtype([("id", "<u4"), ("type", "<u4"), ("x", "<f4"), ("y", "<f4")])
ODOM_REC = np.dtype([("type", "<u4"), ("id1", "<u4"), ("id2", "<u4"), ("rows", "<u4"), ("cols", "<u4"),
                     ("Z", "<f4", (9,)), ("irows", "<u4"), ("icols", "<u4"), ("w", "<f4", (3,))])
LM_REC = np.dtype([("type", "<u4"), ("id1", "<u4"), ("id2", "<u4"), ("rows", "<u4"), ("cols", "<u4"),
                   ("m", "<f4", (2,)), ("irows", "<u4"), ("icols", "<u4"), ("w", "<f4", (2,))])
assert SE2_REC.itemsize == 20 and PT2_REC.itemsize == 16 and ODOM_REC.itemsize == 76 and LM_REC.itemsize == 44


def encode_request(pose_id, pose_xyt, lm_id, lm_xy, od_ij, od_Z, od_w, lm_ij, lm_meas, lm_w, fixed_ids):
    """SoA -> request payload bytes (no size prefix).  Vertices: poses then landmarks; edges:
    ODOM then LM.  od_Z is (Eo, 9) row-major 3x3; lm_meas is (El, 2) = (range, bearing)."""
    P, L, Eo, El = len(pose_id), len(lm_id), len(od_ij), len(lm_ij)
    v0 = np.zeros(P, SE2_REC)
    v0["id"] = pose_id; v0["type"] = 0
    v0["x"], v0["y"], v0["t"] = pose_xyt[:, 0], pose_xyt[:, 1], pose_xyt[:, 2]
    v1 = np.zeros(L, PT2_REC)
    v1["id"] = lm_id; v1["type"] = 1
    if L:
        v1["x"], v1["y"] = lm_xy[:, 0], lm_xy[:, 1]
    e0 = np.zeros(Eo, ODOM_REC)
    if Eo:
        e0["type"] = 0; e0["id1"] = od_ij[:, 0]; e0["id2"] = od_ij[:, 1]
        e0["rows"] = 3; e0["cols"] = 3; e0["Z"] = od_Z; e0["irows"] = 0; e0["icols"] = 3; e0["w"] = od_w
    e1 = np.zeros(El, LM_REC)
    if El:
        e1["type"] = 1; e1["id1"] = lm_ij[:, 0]; e1["id2"] = lm_ij[:, 1]
        e1["rows"] = 0; e1["cols"] = 2; e1["m"] = lm_meas; e1["irows"] = 0; e1["icols"] = 2; e1["w"] = lm_w
    fx = np.asarray(fixed_ids, "<u4")
    u32 = lambda n: np.uint32(n).tobytes()
    return b"".join([u32(P + L), v0.tobytes(), v1.tobytes(), u32(Eo + El), e0.tobytes(), e1.tobytes(),
                     u32(len(fx)), fx.tobytes()])


def frame(payload: bytes) -> bytes:
    return np.uint32(len(payload)).tobytes() + payload


def decode_reply(buf: bytes):
    """Reply payload -> dict(id -> (type, values)) plus edge / fixed counts.  Scalar loop:
    meant for tests on small graphs (mirrors python/remote/bytes_to_graph.py:49-109)."""
    off = 0
    def u32():
        nonlocal off
        v = int(np.frombuffer(buf, "<u4", 1, off)[0]); off += 4; return v
    def f32(n):
        nonlocal off
        v = np.frombuffer(buf, "<f4", n, off).copy(); off += 4 * n; return v
    verts = {}
    order = []
    for _ in range(u32()):
        vid, vt = u32(), u32()
        verts[vid] = (vt, f32(3 if vt == 0 else 2)); order.append(vid)
    edges = []
    for _ in range(u32()):
        et, a, b = u32(), u32(), u32()
        rows, cols = u32(), u32()
        meas = f32(rows * cols).reshape(rows, cols)
        ir, ic = u32(), u32()
        assert ir == 0
        edges.append((et, a, b, meas, f32(ic)))
    fixed = [u32() for _ in range(u32())]
    assert off == len(buf), (off, len(buf))
    return dict(vertices=verts, order=order, edges=edges, fixed=fixed)


def reply_vertices_fast(buf: bytes, P: int, L: int):
    """Vectorised reply-vertex reader for the layout the product emits when the request listed
    poses first and then landmarks (the synthetic generator's layout)."""
    n = int(np.frombuffer(buf, "<u4", 1, 0)[0])
    assert n == P + L
    v0 = np.frombuffer(buf, SE2_REC, P, 4)
    v1 = np.frombuffer(buf, PT2_REC, L, 4 + 20 * P)
    return v0, v1
