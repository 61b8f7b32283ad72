import os, sys, numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from toyslam_b200 import synth
from toyslam_b200.capi import GraphOptimizerGpu
scale = float(sys.argv[1]) if len(sys.argv) > 1 else 0.04
payload = synth.to_payload(synth.make_config("c3", scale=scale, seed=5))
out = []
for env in ({"TSB_NO_STREAM": "1"}, {"TSB_STREAM_MIN_UNITS": "1"}):
    os.environ.pop("TSB_NO_STREAM", None); os.environ.pop("TSB_STREAM_MIN_UNITS", None)
    os.environ.update(env)
    g = GraphOptimizerGpu(iterations=1); c = g.load_wire(payload)
    st = g.linearize(); out.append((st.chi2, g.system(), g.structure()))
    print(env, "chi2", st.chi2, "units", c.n_blocks)
a, b = out[0][1], out[1][1]
for k in a:
    bad = np.argwhere(a[k] != b[k])
    print(k, a[k].shape, "mismatches", len(bad), bad[:6].tolist())
pp = out[0][2]["pose_ptr"]; cd = out[0][2]["chunk_desc"]
bad = np.unique(np.argwhere(a["hpp_diag"] != b["hpp_diag"])[:, 0])
print("bad poses", bad[:40], "n", len(bad))
print("chunks", cd[:4].tolist())
if len(bad):
    p = bad[0]; print("pose", p, "edges", pp[p], pp[p + 1], a["hpp_diag"][p], b["hpp_diag"][p])
