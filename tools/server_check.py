#!/usr/bin/env python
"""`TSB_GPUS=N graph_optimizer 127.0.0.1 PORT ITERS gpu cuda` against an in-process single-GPU run of the same request.

The ONE server process drives N GPUs (a context per GPU, host threads in toyslam_b200/host/graph_gpu.h); the request
travels over the reference's framing ([int32 size][payload], remote/app/ConnectionHandlerGraph.h:16-52).  The graph is
large enough to be edge-sharded (it does not fit the shared-memory-resident kernel).  Checks: the server reports N GPUs,
two requests on one connection give identical replies, and the reply equals the single-GPU result within the fp32 wire
precision.    python tools/server_check.py --gpus 2"""
import argparse, os, socket, subprocess, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from toyslam_b200 import synth, wire
from toyslam_b200.capi import GraphOptimizerGpu

ap = argparse.ArgumentParser()
ap.add_argument("--gpus", type=int, default=2); ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--workload", default="c3"); ap.add_argument("--scale", type=float, default=0.1)
a = ap.parse_args()
g = synth.make_config(a.workload, scale=a.scale, seed=5)
payload = synth.to_payload(g)
P, L = len(g["pose_id"]), len(g["lm_id"])
port = 19000 + os.getpid() % 2000
env = dict(os.environ, TSB_GPUS=str(a.gpus))
srv = subprocess.Popen([os.path.join(ROOT, "toyslam_b200", "graph_optimizer"), "127.0.0.1", str(port), str(a.iters), "gpu", "cuda"],
                       stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, env=env)
replies = []
try:
    s = None
    for _ in range(300):
        try:
            s = socket.create_connection(("127.0.0.1", port), timeout=1); break
        except OSError:
            time.sleep(0.2)
    assert s is not None, "server did not come up"
    s.settimeout(300)
    for _ in range(2):
        t0 = time.perf_counter()
        s.sendall(wire.frame(payload))
        n = int(np.frombuffer(s.recv(4, socket.MSG_WAITALL), "<u4")[0])
        buf = bytearray()
        while len(buf) < n:
            buf += s.recv(min(1 << 20, n - len(buf)))
        replies.append(bytes(buf)); print(f"request {len(replies)}: {len(payload)} bytes -> {n} bytes in {time.perf_counter() - t0:.3f} s")
    s.close()
finally:
    srv.terminate(); out = srv.communicate(timeout=30)[0]
print(out[-1500:])
assert f"GPUs: {a.gpus}" in out, "the server did not create one context per GPU"
assert replies[0] == replies[1], "two identical requests gave different replies"
one = GraphOptimizerGpu(iterations=a.iters, device=0, pcg_tol=1e-12, pcg_max_iters=20000, enable_exits=True)
one.load_wire(payload); one.optimize(); ref = one.reply(); one.close()
v0, v1 = wire.reply_vertices_fast(replies[0], P, L); r0, r1 = wire.reply_vertices_fast(ref, P, L)
names0 = [n for n in v0.dtype.names]; names1 = [n for n in v1.dtype.names]
err = 0.0
for va, vb, names in ((v0, r0, names0), (v1, r1, names1)):
    for nm in names:
        if va[nm].dtype.kind == "f":
            d = np.abs(va[nm].astype(np.float64) - vb[nm].astype(np.float64)); 
            if nm in ("theta", "t"): d = np.abs(np.arctan2(np.sin(d), np.cos(d)))
            err = max(err, float(np.max(d / np.maximum(1.0, np.abs(vb[nm].astype(np.float64))))))
        else:
            assert np.array_equal(va[nm], vb[nm]), nm
print(f"server on {a.gpus} GPUs vs 1 GPU in process: max relative difference of the reply values {err:.2e}")
assert err < 5e-6, err
print("SERVER MGPU OK")
