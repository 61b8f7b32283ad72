#!/usr/bin/env python
"""bench.py -- GN iterations/s (and edges linearized/s) of the B200 path on BASELINE.json's configs.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c4|c5] [--impl ours|reference]

A "step" is one optimisation request: 50 Gauss-Newton iterations (early exits disabled so all 50
run) on one synthetic Manhattan-world graph (toyslam_b200/synth.py, seed 0).  The default workload at
every N is BASELINE configs[2] (C3: 100k poses / 20k landmarks / 2M LM edges), the largest config a
50-iteration request finishes in seconds on; with --gpus N the ONE request is edge-sharded over the N
ranks ("scaling": "strong") and the line carries parity_vs_n1.  C4 (1M poses / 20M edges, sharded the
same way) and C2 (N independent replicas) are reported as side keys.
  value : GN iterations/s with the graph resident in HBM (state reset device-to-device each step)
  e2e   : the same through the C ABI with host buffers: tsb_load_wire (parse + H2D) ->
          tsb_optimize -> tsb_write_reply (D2H + encode), payload in pinned host memory
  --impl reference : the CPU restatement of the reference's `cpu eigen` algorithm (oracle/, sparse
          tier, all host threads) on a bounded sample of the same workload.  The reference C++ itself
          needs Eigen + Boost (absent) and a dense dim x dim H (9 GB at this size): see DESIGN.md.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
GN_ITERS = 50


def algorithmic_bytes(c):
    """fp64 SoA algorithmic minimum per GN linearisation / per PCG iteration."""
    lin = 88 * c["El"] + 152 * c["Eo"] + 120 * c["P"] + 64 * c["L"]
    pcg = 104 * c["El"] + 144 * c["Eo"] + 384 * c["P"] + 64 * c["L"]
    return lin, pcg


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"], "measured"
    except Exception:
        return 6650.0, "fallback"


class ClockSampler:
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index=0):
        self.samples, self.stop, self.index = [], False, index
        self.t = threading.Thread(target=self.run, daemon=True)

    def run(self):
        while not self.stop:
            try:
                out = subprocess.run(["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits"],
                                     capture_output=True, text=True, timeout=5).stdout.strip()
                if out:
                    self.samples.append([x.strip() for x in out.split(",")])
            except Exception:
                pass
            time.sleep(0.2)

    def __enter__(self):
        self.t.start(); return self

    def __exit__(self, *a):
        self.stop = True; self.t.join(timeout=6)

    def summary(self):
        if not self.samples:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        sm = sorted(float(s[0]) for s in self.samples)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = [n for k, n in enumerate(names) if any(s[2 + k].lower().startswith("active") for s in self.samples)]
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(self.samples[0][1]), "reasons": reasons, "samples": len(sm)}


def make_workload(name, seed=0):
    from toyslam_b200 import synth
    g = synth.make_config(name, seed=seed)
    return synth.to_payload(g), synth.counts(g)


def host_threads(omp_default):
    """Threads for the CPU arm: every core this process may run on.  torchrun exports OMP_NUM_THREADS=1 when the
    variable is unset, which would otherwise cut the reference arm to one thread at N > 1; TSB_REF_THREADS overrides."""
    env = os.environ.get("TSB_REF_THREADS")
    if env:
        return max(1, int(env))
    try:
        return max(omp_default, len(os.sched_getaffinity(0)))
    except AttributeError:
        return max(omp_default, os.cpu_count() or 1)


def run_reference(args):
    """CPU arm: sparse restatement of the reference algorithm (oracle tier T2), all host threads, on the SAME workload.
    A step is a bounded sample: `--ref-iters` GN iterations from the start of the request, with the block-Jacobi PCG of
    each stopped after `--ref-pcg-cap` iterations (to convergence it needs thousands at C3: minutes per GN iteration).
    The value is therefore an UPPER bound on the CPU arm's throughput, i.e. GPU / CPU ratios built on it are lower bounds."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle.oracle import OracleGraph, max_threads
    payload, c = make_workload(args.workload)
    threads = host_threads(max_threads())
    sample_iters = args.ref_iters
    times, its_seen = [], []
    for s in range(args.warmup + args.steps):
        o = OracleGraph(payload)
        t0 = time.perf_counter()
        traj, its, _ = o.optimize_sparse(sample_iters, False, args.pcg_tol, args.ref_pcg_cap, threads)
        dt = time.perf_counter() - t0
        if s >= args.warmup:
            times.append(dt); its_seen.append(float(np.mean(its)))
    t = float(np.mean(times))
    val = sample_iters / t
    line = {"impl": "reference", "metric": "gn_iterations_per_s", "value": val, "unit": "iter/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"{args.workload}: P={c['P']} L={c['L']} Eo={c['Eo']} El={c['El']}, {GN_ITERS} GN iterations/step, exits off",
                       "gn_iterations_per_step_sampled": sample_iters, "pcg_tol": args.pcg_tol, "pcg_cap": args.ref_pcg_cap,
                       "pcg_iters_per_gn": float(np.mean(its_seen))},
            "cpu_baseline": {"value": val, "unit": "iter/s", "cores": threads, "kind": "port",
                             "sample": f"first {sample_iters} of {GN_ITERS} GN iterations per step, PCG stopped after {args.ref_pcg_cap} iterations per GN iteration "
                                       "(upper bound on the CPU throughput: block-Jacobi needs thousands to reach 1e-12 at this size); sparse Schur+PCG restatement "
                                       "of OptimizerCpu (reference C++ unbuildable here: Eigen/Boost absent; dense H impossible at this size)"},
            "e2e": {"value": val, "unit": "iter/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "edges_linearized_per_s_incl_solve": (c["Eo"] + c["El"]) * val}
    print(json.dumps(line), flush=True)


def measure_request(opt, torch, dist, world, steps, warmup, flush, local):
    """`warmup` untimed + `steps` timed requests of GN_ITERS iterations on the loaded graph (state reset on the device
    before each).  Device time by the library's CUDA events on its own stream, MAX over ranks; clocks sampled meanwhile."""
    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step():
        opt.reset_state()
        return opt.optimize(GN_ITERS)[0]

    for _ in range(warmup):
        step()
    reps = []
    barrier()
    with ClockSampler(local) as clk:
        t0 = time.perf_counter()
        for _ in range(steps):
            flush.fill_(1)                                   # L2 flush between timed steps
            torch.cuda.synchronize()
            reps.append(step())
        barrier()
        wall = time.perf_counter() - t0
    n = len(reps)
    out = {"ms": sum(r.ms_total for r in reps) / n, "launches": sum(r.kernel_launches for r in reps),
           "pcg": sum(r.pcg_iters_total for r in reps) / n, "ms_lin": sum(r.ms_linearize_total for r in reps) / n,
           "ms_sol": sum(r.ms_solve_total for r in reps) / n, "ms_pcg": sum(r.ms_pcg_total for r in reps) / n,
           "refreshes": sum(r.coarse_refreshes for r in reps) / n, "not_converged": sum(r.pcg_not_converged for r in reps),
           "coarse_failures": sum(r.coarse_failures for r in reps), "wall_ms": 1e3 * wall / n, "clocks": clk.summary()}
    t = torch.tensor([out["ms"]], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    out["ms"] = float(t.item())
    return out


def parity_vs_one_rank(args, payload, torch, dist, rank, world, local, nccl_id, iters=4):
    """The sharded request against the same request on ONE GPU (rank 0 runs it), `iters` GN iterations from the loaded
    state: relative chi2 trajectory error and state error (relative to max(1, |value|)).  Both runs solve to a PCG
    tolerance of 1e-13: at 1e-12 two correct solves of a system this size already differ by ~2e-8 in the state."""
    from toyslam_b200.capi import GraphOptimizerGpu
    tol = min(args.pcg_tol, 1e-13)
    opt = GraphOptimizerGpu(iterations=iters, device=local, rank=rank, world=world, nccl_id=nccl_id, pcg_tol=tol,
                            pcg_max_iters=args.pcg_max, enable_exits=False)
    opt.load_wire(payload)
    _, per = opt.optimize(iters)
    chi = np.array([p.chi2 for p in per]); ps, ls = opt.state()
    opt.close()
    res = torch.zeros(3, dtype=torch.float64, device="cuda")
    if rank == 0:
        one = GraphOptimizerGpu(iterations=iters, device=local, pcg_tol=tol, pcg_max_iters=args.pcg_max, enable_exits=False)
        one.load_wire(payload)
        rep1, per1 = one.optimize(iters)
        chi1 = np.array([p.chi2 for p in per1]); ps1, ls1 = one.state()
        d = ps - ps1; d[:, 2] = np.arctan2(np.sin(d[:, 2]), np.cos(d[:, 2]))
        e_st = max(float((np.abs(d) / np.maximum(1.0, np.abs(ps1))).max()), float((np.abs(ls - ls1) / np.maximum(1.0, np.abs(ls1))).max()) if len(ls) else 0.0)
        res[0] = float(np.max(np.abs(chi - chi1) / chi1)); res[1] = e_st
        res[2] = float(sum(p.pcg_iters for p in per1)) / iters
        one.close()
    dist.broadcast(res, 0)
    return {"chi2_rel": float(res[0].item()), "state_rel": float(res[1].item()), "gn_iterations": iters, "pcg_tol": tol,
            "pcg_iters_per_gn_n1": float(res[2].item()), "pcg_iters_per_gn": float(sum(p.pcg_iters for p in per)) / iters,
            "tolerance": {"chi2_rel": 1e-8, "state_rel": 1e-8}}


def kernel_sample(args, name, torch, dist, rank, world, local, nccl_id_fn, gn=2, lin_reps=6):
    """Side measurement on another config (sharded over the ranks like the headline): linearize time (K1+K2, CUDA events in
    the library), time per PCG iteration inside the PCG kernel, and `gn` full GN iterations to the PCG tolerance."""
    from toyslam_b200.capi import GraphOptimizerGpu
    payload, c = make_workload(name)
    opt = GraphOptimizerGpu(iterations=gn, device=local, rank=rank, world=world, nccl_id=nccl_id_fn(), pcg_tol=args.pcg_tol,
                            pcg_max_iters=args.pcg_max, enable_exits=False)
    buf = torch.frombuffer(bytearray(payload), dtype=torch.uint8)
    t0 = time.perf_counter(); cnt = opt.load_wire_ptr(buf.data_ptr(), buf.numel()); load_s = time.perf_counter() - t0
    del payload
    lin = []
    for i in range(lin_reps + 2):
        st = opt.linearize()
        if i >= 2:
            lin.append(st.ms_linearize)
    rep, per = opt.optimize(gn)
    vals = torch.tensor([float(np.median(lin)), rep.ms_total, rep.ms_pcg_total], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(vals, op=dist.ReduceOp.MAX)
    ms_lin, ms_tot, ms_pcg = (float(v) for v in vals.tolist())
    B_lin, B_pcg = algorithmic_bytes(c)
    peak, _ = peaks()
    its = max(1, rep.pcg_iters_total)
    path = {0: "k_pcg_persistent", 1: "k_pcg_resident", 2: "k_pcg_* phases + NCCL", 3: "k_pcg_stream"}[cnt.pcg_path]
    out = {"workload": f"{name}: P={c['P']} L={c['L']} Eo={c['Eo']} El={c['El']}", "n_gpus": world, "sharded": world > 1,
           "gn_iterations_per_s": gn / (ms_tot * 1e-3), "gn_iterations_sampled": gn, "ms_per_gn_iteration": ms_tot / gn,
           "pcg_iters_per_gn": rep.pcg_iters_total / gn, "pcg_not_converged": rep.pcg_not_converged,
           "ml_levels": cnt.n_ml_levels, "ml_top_nodes": cnt.n_ml_top, "load_wire_s": load_s,
           "linearize": {"kernel": "k_odom + k_linearize_stream + landmark sums", "bound": "hbm", "ms": ms_lin, "algorithmic_bytes": B_lin,
                         "achieved": B_lin / (ms_lin * 1e-3) / 1e9, "peak": peak * world, "unit": "GB/s",
                         "frac": B_lin / (ms_lin * 1e-3) / 1e9 / (peak * world), "traffic": None,
                         "edges_linearized_per_s": (c["Eo"] + c["El"]) / (ms_lin * 1e-3)},
           "pcg_iteration": {"kernel": path, "bound": "hbm", "us": 1e3 * ms_pcg / its, "algorithmic_bytes": B_pcg,
                             "achieved": B_pcg / (ms_pcg / its * 1e-3) / 1e9, "peak": peak * world, "unit": "GB/s",
                             "frac": B_pcg / (ms_pcg / its * 1e-3) / 1e9 / (peak * world), "traffic": None}}
    opt.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default="c3", help="headline workload (c2..c5); one request, edge-sharded over the ranks at N > 1")
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--pcg-tol", type=float, default=1e-12)
    ap.add_argument("--pcg-max", type=int, default=20000)
    ap.add_argument("--ref-iters", type=int, default=1, help="GN iterations per sample of the CPU arm")
    ap.add_argument("--ref-pcg-cap", type=int, default=300, help="PCG iterations per GN iteration after which the CPU arm's solve is stopped")
    ap.add_argument("--no-c4", action="store_true", help="skip the side measurement at the north-star size (1M poses / 20M edges)")
    ap.add_argument("--no-c2", action="store_true", help="skip the C2 side measurement (N independent replicas)")
    ap.add_argument("--lin-reps", type=int, default=20)
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)

    import torch
    import torch.distributed as dist
    from toyslam_b200.capi import GraphOptimizerGpu
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (no CPU fallback in the product path)")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"

    def new_id():
        if world == 1:
            return None
        from toyslam_b200 import dist as tdist
        return tdist.broadcast_unique_id(rank)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    os.environ["TSB_FORCE_SHARD"] = "1"          # the headline request is ONE request on N GPUs, never N replicas
    payload, c = make_workload(args.workload)
    n_edges = c["Eo"] + c["El"]
    B_lin, B_pcg = algorithmic_bytes(c)
    peak, peak_kind = peaks()
    opt = GraphOptimizerGpu(iterations=GN_ITERS, device=local, rank=rank, world=world, nccl_id=new_id(),
                            pcg_tol=args.pcg_tol, pcg_max_iters=args.pcg_max, enable_exits=False)
    req = torch.frombuffer(bytearray(payload), dtype=torch.uint8).pin_memory()       # pinned request / reply for the end-to-end leg
    counts = opt.load_wire_ptr(req.data_ptr(), req.numel())
    reply = torch.empty(opt.reply_size(), dtype=torch.uint8).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    # ---- resident leg: value ------------------------------------------------------------------------------------
    m = measure_request(opt, torch, dist, world, args.steps, args.warmup, flush, local)
    value = GN_ITERS / (m["ms"] * 1e-3)
    pcg_ms = m["ms_pcg"] / max(m["pcg"], 1)                  # device time per PCG iteration inside the PCG kernel

    # ---- linearize (K1 + K2) at this workload, CUDA events inside the library ---------------------------------------
    lin = []
    for i in range(args.lin_reps + 3):
        flush.fill_(1); torch.cuda.synchronize()
        st = opt.linearize()
        if i >= 3:
            lin.append(st.ms_linearize)
    tl = torch.tensor([float(np.median(lin))], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(tl, op=dist.ReduceOp.MAX)
    lin_ms = float(tl.item())

    # ---- end-to-end leg: host request bytes -> tsb_load_wire (parse, structure, H2D) -> optimize -> reply bytes (D2H) ----
    def step_e2e():
        opt.load_wire_ptr(req.data_ptr(), req.numel())
        rep, _ = opt.optimize(GN_ITERS)
        opt.write_reply_into(reply.data_ptr(), reply.numel())
        return rep

    step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step_e2e()
    barrier()
    e2e_s = (time.perf_counter() - t0) / args.steps
    h2d, d2h = opt.transfer_bytes()
    te = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(te, op=dist.ReduceOp.MAX)
    e2e_s = float(te.item())

    path = {0: "k_pcg_persistent (whole PCG solve in one cooperative launch)", 1: "k_pcg_resident (whole solve in one cooperative launch, Hpl resident in shared memory)",
            2: "k_pcg_* phases + NCCL all-reduces", 3: ("k_pcg_stream<true>" if world > 1 else "k_pcg_stream1<false>") + " (whole PCG solve in one persistent cooperative launch: edge passes through a 3-stage TMA ring, "
            "multilevel preconditioner" + (", landmark sums / top-level residual / dot products exchanged over NVLink peer memory" if world > 1 else "") + ")"}[counts.pcg_path]
    line = {
        "metric": "gn_iterations_per_s", "value": value, "unit": "iter/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": m["ms"], "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{args.workload}: P={c['P']} L={c['L']} Eo={c['Eo']} El={c['El']}, {GN_ITERS} GN iterations/step, exits off",
                   "pcg_tol": args.pcg_tol, "pcg_iters_per_gn": m["pcg"] / GN_ITERS, "l2": "flushed between timed steps (256 MiB write)",
                   "wall_ms_per_step": m["wall_ms"], "pcg_not_converged": m["not_converged"], "coarse_failures": m["coarse_failures"],
                   "preconditioner": f"{counts.n_ml_levels}-level hierarchy, dense top level of {counts.n_ml_top} nodes per rank" if counts.n_ml_levels else f"two-level, {counts.n_aggregates} aggregates",
                   "parallelism": "single GPU" if world == 1 else
                   f"ONE request edge-sharded over {world} ranks (contiguous pose ranges, one process per GPU): landmark sums by owner-computes "
                   "pushes over NVLink peer memory inside the persistent PCG kernel, global dense top level of the preconditioner replicated"},
        "edges_linearized_per_s": n_edges / (lin_ms * 1e-3),
        "e2e": {"value": GN_ITERS / e2e_s, "unit": "iter/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h), "ms_per_step": 1e3 * e2e_s},
        "gpu_launches": int(m["launches"]),
        "clocks": m["clocks"],
        # dominant kernel of the step: the PCG kernel, one launch per solve.  Algorithmic bytes per launch = B_pcg x the PCG
        # iterations of that solve; duration from CUDA events around the launch on the library's stream.
        "roofline": {"kernel": path, "bound": "hbm", "achieved": B_pcg / (pcg_ms * 1e-3) / 1e9, "peak": peak * world, "peak_kind": peak_kind,
                     "unit": "GB/s", "frac": B_pcg / (pcg_ms * 1e-3) / 1e9 / (peak * world), "traffic": None,
                     "traffic_note": "not measured in this run; ncu --set full summaries of the same kernel: profiles/",
                     "algorithmic_bytes_per_launch": B_pcg * m["pcg"] / GN_ITERS, "algorithmic_bytes_per_pcg_iteration": B_pcg,
                     "launch_ms": m["ms_pcg"] / GN_ITERS, "pcg_iterations_per_launch": m["pcg"] / GN_ITERS, "us_per_pcg_iteration": 1e3 * pcg_ms,
                     "share_of_step": m["ms_pcg"] / m["ms"]},
        "roofline_linearize": {"kernel": "k_odom + k_linearize[_stream] + landmark sums", "bound": "hbm",
                               "achieved": B_lin / (lin_ms * 1e-3) / 1e9, "peak": peak * world, "unit": "GB/s",
                               "frac": B_lin / (lin_ms * 1e-3) / 1e9 / (peak * world), "algorithmic_bytes": B_lin, "ms": lin_ms, "traffic": None},
        "solve_breakdown_ms_per_gn": {"linearize": m["ms_lin"] / GN_ITERS, "solve_total": m["ms_sol"] / GN_ITERS, "pcg": m["ms_pcg"] / GN_ITERS,
                                      "schur_prep_and_coarse_setup": (m["ms_sol"] - m["ms_pcg"]) / GN_ITERS, "coarse_refreshes_per_step": m["refreshes"]},
    }
    # ---- N > 1: the sharded request against the same request on one GPU, in this run -----------------------------------
    opt.close()
    del opt
    if world > 1:
        try:
            line["parity_vs_n1"] = parity_vs_one_rank(args, payload, torch, dist, rank, world, local, new_id())
        except Exception as e:
            line["parity_vs_n1"] = {"error": str(e)}

    # ---- the north-star size (C4: 1M poses / 20M edges), sharded the same way: kernels against the roofline and full GN iterations
    if not args.no_c4 and args.workload != "c4":
        try:
            line["c4"] = kernel_sample(args, "c4", torch, dist, rank, world, local, new_id, gn=2)
        except Exception as e:                                # never lose the main line to a side measurement
            line["c4"] = {"error": str(e)}
    # ---- C2 (fits one GPU's shared-memory-resident kernel; latency-bound): N independent requests, one per GPU ------------
    if not args.no_c2 and args.workload != "c2":
        try:
            os.environ.pop("TSB_FORCE_SHARD", None)
            p2, c2 = make_workload("c2", seed=rank)
            o2 = GraphOptimizerGpu(iterations=GN_ITERS, device=local, pcg_tol=args.pcg_tol, pcg_max_iters=args.pcg_max, enable_exits=False)
            o2.load_wire(p2)
            m2 = measure_request(o2, torch, dist, world, 2, 1, flush, local)
            line["c2_replicas"] = {"workload": f"c2: P={c2['P']} L={c2['L']} El={c2['El']}, one request per GPU (seed = rank), no collective",
                                   "gn_iterations_per_s_aggregate": world * GN_ITERS / (m2["ms"] * 1e-3), "ms_per_request": m2["ms"],
                                   "pcg_iters_per_gn": m2["pcg"] / GN_ITERS, "scaling": "weak",
                                   "note": "L2/shared-memory-resident working set: latency-bound, no HBM roofline fraction is quoted for it"}
            o2.close()
        except Exception as e:
            line["c2_replicas"] = {"error": str(e)}

    # ---- CPU baseline on the host cores of this box (rank 0, N=1 only): the oracle's sparse tier on the SAME graph,
    # sampled at the start of the request and from the state the GPU run reached after half of it --------------------------
    if world == 1 and rank == 0:
        try:
            from oracle.oracle import OracleGraph, max_threads
            th = host_threads(max_threads())
            o = OracleGraph(payload)
            t0 = time.perf_counter(); _, its0, _ = o.optimize_sparse(args.ref_iters, False, args.pcg_tol, args.ref_pcg_cap, th); dt0 = time.perf_counter() - t0
            g = GraphOptimizerGpu(iterations=GN_ITERS // 2, device=local, pcg_tol=args.pcg_tol, pcg_max_iters=args.pcg_max, enable_exits=False)
            g.load_wire(payload); g.optimize(GN_ITERS // 2); ps, ls = g.state(); g.close()
            o.set_state(ps, ls)
            t0 = time.perf_counter(); _, its1, _ = o.optimize_sparse(args.ref_iters, False, args.pcg_tol, args.ref_pcg_cap, th); dt1 = time.perf_counter() - t0
            line["cpu_baseline"] = {"value": 2 * args.ref_iters / (dt0 + dt1), "unit": "iter/s", "cores": th, "kind": "port",
                                    "sample": f"GN iterations 0..{args.ref_iters - 1} and {GN_ITERS // 2}..{GN_ITERS // 2 + args.ref_iters - 1} of the same {args.workload} request "
                                              f"(the second sample starts from the state the GPU run reached), PCG stopped after {args.ref_pcg_cap} iterations per GN iteration (an upper bound on the CPU "
                                              "throughput: to 1e-12 block-Jacobi needs thousands), sparse Schur + block-Jacobi PCG restatement of OptimizerCpu "
                                              "(oracle tier T2, OpenMP); the reference C++ needs Eigen/Boost (absent) and a dense dim x dim H",
                                    "s_per_gn_iteration": {"start": dt0 / args.ref_iters, "mid": dt1 / args.ref_iters},
                                    "pcg_iters_per_gn": {"start": float(np.mean(its0)), "mid": float(np.mean(its1))}}
        except Exception as e:
            line["cpu_baseline"] = {"error": str(e)}
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
