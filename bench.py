#!/usr/bin/env python
"""bench.py -- GN iterations/s (and edges linearized/s) of the B200 path on BASELINE.json's configs.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--workload c2|c3|c4|c5] [--impl ours|reference]

A "step" is one optimisation request: 50 Gauss-Newton iterations (early exits disabled so all 50
run) on one synthetic Manhattan-world graph (toyslam_b200/synth.py, seed 0).  At N=1 the default
workload is BASELINE configs[1] (10k poses / 2k landmarks / 150k LM edges).
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
    """CPU arm: sparse restatement of the reference algorithm (oracle tier T2), all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from oracle.oracle import OracleGraph, max_threads
    payload, c = make_workload(args.workload)
    threads = host_threads(max_threads())
    sample_iters = args.ref_iters
    times = []
    for s in range(args.warmup + args.steps):
        o = OracleGraph(payload)
        t0 = time.perf_counter()
        traj, its, _ = o.optimize_sparse(sample_iters, False, args.pcg_tol, args.pcg_max, threads)
        dt = time.perf_counter() - t0
        if s >= args.warmup:
            times.append(dt)
    t = float(np.mean(times))
    val = sample_iters / t
    line = {"impl": "reference", "metric": "gn_iterations_per_s", "value": val, "unit": "iter/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": {"workload": f"{args.workload}: P={c['P']} L={c['L']} Eo={c['Eo']} El={c['El']}",
                       "gn_iterations_per_step_sampled": sample_iters, "pcg_tol": args.pcg_tol},
            "cpu_baseline": {"value": val, "unit": "iter/s", "cores": threads, "kind": "port",
                             "sample": f"first {sample_iters} of {GN_ITERS} GN iterations per step, sparse Schur+PCG restatement "
                                       "of OptimizerCpu (reference C++ unbuildable here: Eigen/Boost absent; dense H impossible at this size)"},
            "e2e": {"value": val, "unit": "iter/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "edges_linearized_per_s_incl_solve": (c["Eo"] + c["El"]) * val}
    print(json.dumps(line), flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--workload", default=None)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--pcg-tol", type=float, default=1e-12)
    ap.add_argument("--pcg-max", type=int, default=20000)
    ap.add_argument("--ref-iters", type=int, default=2, help="GN iterations per step sampled by the CPU arm")
    ap.add_argument("--no-c4", action="store_true", help="skip the extra 1M-pose linearize roofline measurement")
    ap.add_argument("--lin-reps", type=int, default=20)
    args = ap.parse_args()
    if args.workload is None:
        args.workload = "c2"          # BASELINE configs[1] at every N, so the per-N lines are comparable
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
    nccl_id = None
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
        from toyslam_b200 import dist as tdist
        nccl_id = tdist.broadcast_unique_id(rank)
    assert world == args.gpus, f"--gpus {args.gpus} but WORLD_SIZE={world}"

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    payload, c = make_workload(args.workload)
    n_edges = c["Eo"] + c["El"]
    B_lin, B_pcg = algorithmic_bytes(c)
    peak, peak_kind = peaks()
    opt = GraphOptimizerGpu(iterations=GN_ITERS, device=local, rank=rank, world=world, nccl_id=nccl_id,
                            pcg_tol=args.pcg_tol, pcg_max_iters=args.pcg_max, enable_exits=False)
    # pinned host copies of request and reply for the end-to-end leg
    req = torch.frombuffer(bytearray(payload), dtype=torch.uint8).pin_memory()
    counts = opt.load_wire_ptr(req.data_ptr(), req.numel())
    # N > 1 and the request fits ONE GPU (the library would run it replicated: same answer on every rank, no speed-up):
    # the N GPUs of the box then serve N requests at once -- one process per GPU, each with its own graph of this
    # config (seed = rank) and a context without a communicator; no data-path collective.  `opt_n` (the world = N
    # context) is kept for the single-request numbers reported beside the aggregate.
    opt_n, req_n, independent = None, None, False
    if world > 1 and counts.replicated == 1:
        independent = True
        opt_n, req_n = opt, req
        payload, c = make_workload(args.workload, seed=rank)
        opt = GraphOptimizerGpu(iterations=GN_ITERS, device=local, rank=0, world=1, nccl_id=None,
                                pcg_tol=args.pcg_tol, pcg_max_iters=args.pcg_max, enable_exits=False)
        req = torch.frombuffer(bytearray(payload), dtype=torch.uint8).pin_memory()
        counts = opt.load_wire_ptr(req.data_ptr(), req.numel())
    reply = torch.empty(opt.reply_size(), dtype=torch.uint8).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")

    # ---- resident leg: value ------------------------------------------------------------------------
    def step_resident():
        opt.reset_state()
        rep, per = opt.optimize(GN_ITERS)
        return rep, per

    for _ in range(args.warmup):
        step_resident()
    reps = []
    barrier()
    with ClockSampler(local) as clk:
        t0 = time.perf_counter()
        for _ in range(args.steps):
            flush.fill_(1)                                   # L2 flush between timed steps
            torch.cuda.synchronize()
            reps.append(step_resident())
        barrier()
        wall = time.perf_counter() - t0
    dev_ms = sum(r.ms_total for r, _ in reps) / args.steps
    launches = sum(r.kernel_launches for r, _ in reps)
    pcg_total = sum(r.pcg_iters_total for r, _ in reps) / args.steps
    ms_lin = sum(r.ms_linearize_total for r, _ in reps) / args.steps / GN_ITERS
    ms_sol = sum(r.ms_solve_total for r, _ in reps) / args.steps
    ms_pcg = sum(r.ms_pcg_total for r, _ in reps) / args.steps       # the PCG proper (one persistent kernel per solve at N=1)
    refreshes = sum(r.coarse_refreshes for r, _ in reps) / args.steps
    t = torch.tensor([dev_ms], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dev_ms = float(t.item())
    n_req = world if independent else 1                      # requests in flight: units = n_req x GN_ITERS per step
    value = n_req * GN_ITERS / (dev_ms * 1e-3)

    # ---- linearize roofline at this workload (K1 + K2a + K2b), CUDA events inside the library ------------
    lin = []
    for i in range(args.lin_reps + 3):
        flush.fill_(1); torch.cuda.synchronize()
        st = opt.linearize()
        if i >= 3:
            lin.append(st.ms_linearize)
    lin_ms = float(np.median(lin))
    pcg_ms = ms_pcg / max(pcg_total, 1)                      # device time per PCG iteration inside the PCG kernel(s)

    # ---- end-to-end leg -----------------------------------------------------------------------------
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

    resident_kernel = counts.pcg_path == 1
    line = {
        "metric": "gn_iterations_per_s", "value": value, "unit": "iter/s", "n_gpus": world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dev_ms, "higher_is_better": True,
        "scaling": "strong" if world > 1 and not independent else "weak",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"{args.workload}: P={c['P']} L={c['L']} Eo={c['Eo']} El={c['El']}, {GN_ITERS} GN iterations/step, exits off",
                   "pcg_tol": args.pcg_tol, "pcg_iters_per_gn": pcg_total / GN_ITERS, "l2": "flushed between timed steps (256 MiB write)",
                   "wall_ms_per_step": 1e3 * wall / args.steps},
        "edges_linearized_per_s": n_req * n_edges / (lin_ms * 1e-3),
        "e2e": {"value": n_req * GN_ITERS / e2e_s, "unit": "iter/s", "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "ms_per_step": 1e3 * e2e_s},
        "gpu_launches": int(launches),
        "clocks": clk.summary(),
        # dominant kernel of the step (profiles/: > 90 % of the C2 step): the persistent PCG kernel, one launch per solve.
        # algorithmic bytes per launch = B_pcg x the PCG iterations of that solve; duration by CUDA events around the launch.
        "roofline": {"kernel": ("k_pcg_resident (whole PCG solve in one cooperative launch, Hpl resident in shared memory)" if resident_kernel else
                                "k_pcg_persistent (whole PCG solve in one cooperative launch)" if world == 1 else "k_pcg_* phases + NCCL all-reduces"),
                     "bound": "hbm", "achieved": B_pcg / (pcg_ms * 1e-3) / 1e9, "peak": peak, "peak_kind": peak_kind, "unit": "GB/s",
                     "frac": B_pcg / (pcg_ms * 1e-3) / 1e9 / peak,
                     # DRAM bytes of one k_pcg_resident launch (= one solve) from the ncu --set full captures summarised in
                     # profiles/r1_pcg_resident_c2_v3_ncu_summary.csv (first solve, smoother blocks: 44.3 MB read + 1.2 MB
                     # written, 24.5 MB of it the smoother inverses read once into L2) and ..._v2_... (3x3 blocks: 17.6 MB):
                     # far below the algorithmic bytes because both orders of Hpl stay in shared memory for the whole solve
                     "traffic": 45.5e6 if (resident_kernel and args.workload == "c2") else None,
                     "traffic_source": "profiles/r1_pcg_resident_c2_v3_ncu_summary.csv" if (resident_kernel and args.workload == "c2") else None,
                     "algorithmic_bytes_per_launch": B_pcg * pcg_total / GN_ITERS, "algorithmic_bytes_per_pcg_iteration": B_pcg,
                     "launch_ms": ms_pcg / GN_ITERS, "pcg_iterations_per_launch": pcg_total / GN_ITERS,
                     "share_of_step": ms_pcg / dev_ms,
                     "note": "working set (%.0f MB) is L2-resident at this size: latency-bound, not HBM-bound" % (B_pcg / 1e6) if B_pcg < 100e6 else ""},
        "roofline_linearize": {"kernel": "k_odom + k_linearize[_stream] + k_lm_sum", "bound": "hbm",
                               "achieved": B_lin / (lin_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s",
                               "frac": B_lin / (lin_ms * 1e-3) / 1e9 / peak, "algorithmic_bytes": B_lin, "ms": lin_ms,
                               "note": "working set is L2-resident at this size" if B_lin < 100e6 else ""},
        "solve_breakdown_ms_per_gn": {"linearize": ms_lin, "solve_total": ms_sol / GN_ITERS, "pcg": ms_pcg / GN_ITERS,
                                      "schur_prep_and_coarse_setup": (ms_sol - ms_pcg) / GN_ITERS, "coarse_refreshes_per_step": refreshes},
    }

    line["config"]["parallelism"] = (
        "single GPU" if world == 1 else
        f"{world} independent requests, one per GPU (this config fits one GPU's shared-memory-resident PCG kernel and is "
        "latency-bound: sharding it over NVLink only adds collective latency, see `single_request`); one process per GPU, "
        "graphs of the same config with seed = rank, no data-path collective; value = all ranks' GN iterations / slowest rank"
        if independent else f"one request edge-sharded over {world} ranks, NCCL all-reduces")
    # ---- world > 1: ONE request on the N GPUs, for the record: replicated (what the library does by default for a
    # graph of this size) and forced through the edge-sharded NCCL path ---------------------------------------------
    if independent:
        def step_n():
            opt_n.reset_state()
            return opt_n.optimize(GN_ITERS)[0]
        single = {}
        for mode in ("replicated", "sharded"):
            if mode == "sharded":
                os.environ["TSB_FORCE_SHARD"] = "1"
                opt_n.load_wire_ptr(req_n.data_ptr(), req_n.numel())
            step_n()
            barrier()
            sh = [step_n() for _ in range(max(1, args.steps - 1))]
            barrier()
            ts = torch.tensor([sum(r.ms_total for r in sh) / len(sh)], dtype=torch.float64, device="cuda")
            dist.all_reduce(ts, op=dist.ReduceOp.MAX)
            single[mode] = {"value": GN_ITERS / (float(ts.item()) * 1e-3), "unit": "iter/s", "ms_per_step": float(ts.item()),
                            "pcg_iters_per_gn": sum(r.pcg_iters_total for r in sh) / len(sh) / GN_ITERS}
        os.environ.pop("TSB_FORCE_SHARD", None)
        single["replicated"]["note"] = "every rank solves the whole graph, bit-identical results, no collective"
        single["sharded"]["note"] = "contiguous pose ranges per rank, per-rank coarse spaces, 3 NCCL all-reduces per PCG iteration"
        line["single_request"] = single
    # ---- the north-star size (1M poses / 20M edges): linearize roofline and a bounded PCG sample --------------------
    # (its own context, edge-sharded over the N ranks, PCG capped at 30 iterations per solve so the cost is fixed)
    if not args.no_c4 and args.workload != "c4":
        try:
            _, B_pcg4 = 0, 0
            id4 = None
            if world > 1:
                from toyslam_b200 import dist as tdist
                id4 = tdist.broadcast_unique_id(rank)
            opt4 = GraphOptimizerGpu(iterations=1, device=local, rank=rank, world=world, nccl_id=id4,
                                     pcg_tol=args.pcg_tol, pcg_max_iters=30, enable_exits=False)
            p4, c4 = make_workload("c4")
            r4 = torch.frombuffer(bytearray(p4), dtype=torch.uint8)
            opt4.load_wire_ptr(r4.data_ptr(), r4.numel())
            del p4
            l4 = []
            for i in range(8):
                st = opt4.linearize()
                if i >= 3:
                    l4.append(st.ms_linearize)
            ms4 = float(np.median(l4)); B4, B_pcg4 = algorithmic_bytes(c4)
            rep4 = [opt4.optimize(1)[0] for _ in range(3)][1:]            # GN iterations with 30 PCG iterations each
            us_it = 1e3 * sum(r.ms_pcg_total for r in rep4) / max(1, sum(r.pcg_iters_total for r in rep4))
            if world > 1:                                     # edge-sharded over the ranks: the slowest rank counts
                t4 = torch.tensor([ms4, us_it], dtype=torch.float64, device="cuda"); dist.all_reduce(t4, op=dist.ReduceOp.MAX)
                ms4, us_it = float(t4[0].item()), float(t4[1].item())
            wl4 = f"c4: P={c4['P']} L={c4['L']} Eo={c4['Eo']} El={c4['El']}"
            line["roofline_linearize_c4"] = {"workload": wl4, "bound": "hbm",
                                             "achieved": B4 / (ms4 * 1e-3) / 1e9, "peak": peak * world, "unit": "GB/s", "n_gpus": world,
                                             "frac": B4 / (ms4 * 1e-3) / 1e9 / (peak * world), "algorithmic_bytes": B4, "ms": ms4,
                                             # dram read + write of k_linearize_stream + k_lm_sum + k_odom at N=1, ncu --set full
                                             # (profiles/r1_lin_c4_planes_ncu_summary.csv): 1 865 + 113 + 128 MB
                                             "traffic": 2.106e9 if world == 1 else None,
                                             "edges_linearized_per_s": (c4["Eo"] + c4["El"]) / (ms4 * 1e-3)}
            path4 = {0: "k_pcg_persistent", 1: "k_pcg_resident", 2: "k_pcg_* phases + NCCL all-reduces",
                     3: "k_pcg_stream (persistent cooperative kernel, edge passes through a 3-stage TMA ring)"}[opt4.refresh_counts().pcg_path]
            line["roofline_pcg_c4"] = {"workload": wl4, "kernel": path4,
                                       "bound": "hbm", "achieved": B_pcg4 / (us_it * 1e-6) / 1e9, "peak": peak * world, "unit": "GB/s",
                                       "n_gpus": world, "frac": B_pcg4 / (us_it * 1e-6) / 1e9 / (peak * world),
                                       "algorithmic_bytes_per_pcg_iteration": B_pcg4, "us_per_pcg_iteration": us_it,
                                       # k_pcg_stream at N=1, ncu --set full (profiles/r1_pcg_stream_c4_v2_ncu_summary.csv):
                                       # 17.11 GB read + 1.19 GB written in a 6-iteration launch; B_pcg + the 275 MB dense
                                       # coarse inverse = 2.79 GB per iteration
                                       "traffic_per_pcg_iteration": 3.05e9 if world == 1 else None,
                                       "sample": "2 GN iterations x 30 PCG iterations (capped), PCG kernel time only"}
            opt4.close()
        except Exception as e:                                # never lose the main line to the extra measurement
            line["roofline_linearize_c4"] = {"error": str(e)}

    # ---- CPU baseline on the host cores of this box (rank 0, N=1 only) ----------------------------------------
    if world == 1 and rank == 0:
        from oracle.oracle import OracleGraph, max_threads
        o = OracleGraph(payload)
        th = host_threads(max_threads())
        t0 = time.perf_counter()
        o.optimize_sparse(args.ref_iters, False, args.pcg_tol, args.pcg_max, th)
        dt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": args.ref_iters / dt, "unit": "iter/s", "cores": th, "kind": "port",
                                "sample": f"first {args.ref_iters} of {GN_ITERS} GN iterations of the same graph, sparse Schur+PCG "
                                          "restatement of OptimizerCpu (oracle tier T2, OpenMP)"}
    if rank == 0:
        print(json.dumps(line), flush=True)
    opt.close()
    if opt_n is not None:
        opt_n.close()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
