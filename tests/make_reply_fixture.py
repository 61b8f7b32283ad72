#!/usr/bin/env python
"""Generates tests/golden/*_reply.npz: a REPLY payload decoded by the reference's own decoder.

Run in the build container only (needs /root/reference; the GPU box never reads it):
    python tests/make_reply_fixture.py
The reply bytes come from the CPU oracle (oracle/ is test infrastructure, which is why this script lives under tests/):
it optimises the golden request exactly like the golden test does and encodes the reply in the shape of the reference's
SerializeGraph (remote/serialization/SerializeGraph.h:8-72).  The decoding is the reference's, unmodified:
python/remote/bytes_to_graph.py:49-109 (imported with an empty `matplotlib` stub, as tools/make_golden.py does).
The tests then require (CPU) that toyslam_b200/wire.py decodes the same bytes to the same graph, and (GPU) that the reply
the CUDA path writes decodes to the same ids, types, edges and -- within the fp32 wire precision -- values."""
import os
import sys
import types
import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
for name in ("matplotlib", "matplotlib.pyplot"):
    sys.modules.setdefault(name, types.ModuleType(name))
sys.path.insert(0, "/root/reference/python")
from remote.bytes_to_graph import bytes_to_graph      # noqa: E402
from oracle.oracle import OracleGraph                  # noqa: E402

GOLDEN = os.path.join(ROOT, "tests", "golden")


def main():
    for name in ("c1", "g_loops"):
        d = np.load(os.path.join(GOLDEN, name + ".npz"))
        o = OracleGraph(d["request"].tobytes())
        o.optimize_dense(int(d["iterations"]), bool(d["exits"]))
        reply = o.reply()
        g = bytes_to_graph(reply)                      # the reference's decoder
        ids = np.array(list(g.vertices.keys()), np.uint32)
        vt = np.array([g.vertices[i].get_type() for i in ids], np.uint32)
        val = np.zeros((len(ids), 3), np.float64)
        for k, i in enumerate(ids):
            p = np.asarray(g.vertices[i].position, np.float64)
            val[k] = [p[0, 2], p[1, 2], np.arctan2(p[1, 0], p[0, 0])] if vt[k] == 0 else [p[0], p[1], 0.0]
        et = np.array([e.get_type() for e in g.edges], np.uint32)
        eij = np.array([[e.id_1, e.id_2] for e in g.edges], np.uint32).reshape(-1, 2)
        emeas = np.zeros((len(g.edges), 9), np.float32); einf = np.zeros((len(g.edges), 3), np.float32)
        for k, e in enumerate(g.edges):
            m = np.asarray(e.measurement, np.float32).ravel(); emeas[k, :len(m)] = m
            di = np.diag(np.asarray(e.information, np.float32)); einf[k, :len(di)] = di
        np.savez_compressed(os.path.join(GOLDEN, name + "_reply.npz"), reply=np.frombuffer(reply, np.uint8), vertex_id=ids, vertex_type=vt,
                            vertex_value=val, edge_type=et, edge_ij=eij, edge_meas=emeas, edge_inf=einf,
                            fixed=np.array(sorted(g.fixed_vertices), np.uint32))
        print(f"{name}: reply {len(reply)} bytes -> {len(ids)} vertices, {len(et)} edges, {len(g.fixed_vertices)} fixed (reference decoder)")


if __name__ == "__main__":
    main()
