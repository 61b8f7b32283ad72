# round 2: first GPU check of the multilevel streamed PCG
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
python - <<'PY'
import os, numpy as np, time
from toyslam_b200 import synth
from toyslam_b200.capi import GraphOptimizerGpu
from oracle.oracle import OracleGraph
for name, scale in (("c2", 0.3), ("c3", 0.05), ("c5", 0.01)):
    payload = synth.to_payload(synth.make_config(name, scale=scale, seed=1))
    os.environ["TSB_NO_RESIDENT"] = "1"; os.environ["TSB_STREAM_MIN_UNITS"] = "1"
    res = {}
    for ml in ("1", "0"):
        os.environ["TSB_ML"] = ml
        g = GraphOptimizerGpu(iterations=3, pcg_tol=1e-13, pcg_max_iters=20000, enable_exits=False)
        c = g.load_wire(payload); rep, per = g.optimize()
        res[ml] = (np.array([p.chi2 for p in per]), g.state(), [p.pcg_iters for p in per]); g.close()
    o = OracleGraph(payload); traj = np.array(o.optimize_sparse(3, False, tol=1e-13, max_pcg=20000)[0])
    ps, ls = res["1"][1]; qs, ms = res["0"][1]
    print(name, scale, "P", c.n_pose, "its ML", res["1"][2], "two-level", res["0"][2],
          "chi2 rel ML-vs-2lvl", np.max(np.abs(res["1"][0] - res["0"][0]) / res["0"][0]),
          "vs oracle", np.max(np.abs(res["1"][0] - traj) / traj),
          "state diff", max(np.max(np.abs(ps - qs)), np.max(np.abs(ls - ms))), flush=True)
PY
for w in c3 c4 c5; do echo "--- $w"; TSB_PCG_TIMING=1 timeout 900 python tools/prof.py --workload $w --lin 2 --pcg 20000 --gn 3 2>&1 | grep -v "^$" | cut -c1-600 | tail -8; done
