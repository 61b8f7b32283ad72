#!/usr/bin/env python
"""Small driver for ncu captures: linearize `--lin` times and run one solve capped at `--pcg` PCG
iterations on a synthetic workload.  python tools/prof.py --workload c4 --lin 3 --pcg 10"""
import argparse, os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from toyslam_b200 import synth
from toyslam_b200.capi import GraphOptimizerGpu

ap = argparse.ArgumentParser()
ap.add_argument("--workload", default="c3"); ap.add_argument("--scale", type=float, default=1.0)
ap.add_argument("--lin", type=int, default=3); ap.add_argument("--pcg", type=int, default=10)
ap.add_argument("--gn", type=int, default=0)
a = ap.parse_args()
t0 = time.time()
payload = synth.to_payload(synth.make_config(a.workload, scale=a.scale))
t1 = time.time()
rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
nccl_id = None
if world > 1:                      # torchrun: one process per GPU, the graph is edge-sharded over the ranks
    import torch, torch.distributed as dist
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0"))))
    from toyslam_b200 import dist as tdist
    nccl_id = tdist.broadcast_unique_id(rank)
g = GraphOptimizerGpu(iterations=a.gn, device=int(os.environ.get("LOCAL_RANK", "0")), rank=rank, world=world, nccl_id=nccl_id,
                      pcg_tol=1e-12, pcg_max_iters=a.pcg, enable_exits=False)
c = g.load_wire(payload)
t2 = time.time()
print(f"gen {t1 - t0:.1f}s load {t2 - t1:.1f}s P={c.n_pose} L={c.n_lm} El={c.n_lmedge} chunks={c.n_chunks} blocks={c.n_blocks} slots={c.n_slots}")
for _ in range(a.lin):
    st = g.linearize(); print(f"linearize {st.ms_linearize:.4f} ms chi2 {st.chi2:.6g}")
if a.pcg:
    st = g.solve(); print(f"solve {st.ms_solve:.3f} ms pcg_iters {st.pcg_iters} rel {st.pcg_rel_residual:.2e} -> {st.ms_solve / max(st.pcg_iters, 1) * 1e3:.1f} us/iter")
if a.gn:
    rep, per = g.optimize(a.gn); print(f"optimize {rep.ms_total:.2f} ms, pcg total {rep.pcg_iters_total}; per GN: " + " ".join(f"{q.pcg_iters}/{q.ms_solve:.2f}ms" for q in per))
