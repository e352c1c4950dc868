#!/usr/bin/env python3
"""bench.py -- headline benchmark of the batched DDP hot path (BASELINE.json).

Metric: batched FDDP OCP-iterations/s (sum over OCPs of solver iterations / time) on BASELINE.json
configs[3]: 4096 independent 2-DOF SEA reaching OCPs (T = 100, random initial posture / joint
stiffness, default_rng(20240001)) solved to convergence (th_stop 1e-7, maxiter 100).  A "step" is one
complete solve of the batch from the same initial guess.  Also reported: fused calc+calcDiff knots/s
(the metric's second half) with its HBM roofline, an end-to-end number through the public API with
host buffers, and the CPU oracle timed on the host cores.

  python bench.py [--gpus N] [--steps K] [--warmup W]          our CUDA path
  python bench.py --impl reference ...                          the reference-equivalent CPU path

Multi-GPU (STRONG scaling, the configuration BASELINE.json configs[3] names): one process per GPU under torchrun,
the ONE 4096-OCP batch is cut into contiguous shards of 4096 / N OCPs (aslr_to_b200/sharding.py); there is no
per-iteration collective, only a final NCCL all-gather of costs, iteration counts and trajectories to every rank
(inside the timed region).  `--scaling weak` keeps round 1's replica mode (4096 OCPs per GPU, seed + rank).
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
sys.path.insert(0, os.path.join(ROOT, "tests"))

METRIC = "batched_fddp_ocp_iters_per_s"
UNIT = "OCP-iterations/s"
B_TOTAL = 4096  # BASELINE.json configs[3]
B_PER_GPU = B_TOTAL  # weak mode: every rank solves a batch of this size
SEED = 20240001
ALGO_BYTES_PER_KNOT = 1544  # 2-DOF SEA, read x,u; write xnext,cost,Fx,Fu,Lx,Lu,Lxx,Lxu,Luu


def workload_config(n_gpus, solver="fddp", scaling="strong"):
    per = B_TOTAL // n_gpus if scaling == "strong" else B_PER_GPU
    total = B_TOTAL if scaling == "strong" else B_PER_GPU * n_gpus
    return {"workload": "%d x 2-DOF SEA %s reaching OCPs in total, %d per GPU (BASELINE.json configs[3]), T=100, "
                        "th_stop=1e-7, maxiter=100, solved to convergence" % (total, solver.upper(), per),
            "ocps_total": total, "ocps_per_gpu": per, "T": 100, "nx": 8, "nu": 2, "solver": solver,
            "parallelism": "one batch of independent OCPs cut into %d contiguous shard(s), no per-iteration collective, "
                           "final NCCL all-gather of trajectories and solver state inside the timed region" % n_gpus
                           if scaling == "strong" else
                           "replicas: every GPU solves its own %d-OCP batch (seed + rank)" % B_PER_GPU,
            "l2": "per-iteration working set of a 4096-OCP shard (derivative records + trial trajectories, > 500 MB) "
                  "exceeds the 126 MB L2, no explicit flush; smaller shards (N >= 4) fit in L2 and say so by being "
                  "timed as they are: the same bytes are re-written by every iteration"}


# ------------------------------------------------------------------------------------ clocks
class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        self.index, self.proc, self.lines = index, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap")
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 7:
                continue
            try:
                sm.append(float(f[0])), mx.append(float(f[1]))
            except ValueError:
                continue
            for nm, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------------------ CPU baselines
def _cpu_worker(args):
    """Solves the OCPs `bs` of the batch with the oracle; returns (iterations, seconds spent in the solves alone:
    process start-up, imports and building the workload are not part of the rate)."""
    bs, seed = args
    import oracle_build as ob
    from aslr_to_b200 import workloads as W
    w = W.sea_batch(B_TOTAL, seed=seed)
    iters, t = 0, 0.0
    for b in bs:
        t0 = time.perf_counter()
        solver, log = ob.solve(w, b, "port")
        t += time.perf_counter() - t0
        iters += len(log.costs)
    return iters, t


def cpu_oracle_rate(n_ocps, procs, seed=SEED, pool=None):
    """OCP-iterations/s of the NumPy oracle (reference-equivalent CPU path) on `procs` host cores: OCPs 0..n_ocps-1
    of the batch dealt round-robin to the workers; rate = iterations / the busiest worker's solve time."""
    import multiprocessing as mp
    jobs = [(list(range(k, n_ocps, procs)), seed) for k in range(procs)]
    jobs = [j for j in jobs if j[0]]
    if procs == 1:
        res = [_cpu_worker(jobs[0])]
    elif pool is not None:
        res = pool.map(_cpu_worker, jobs)
    else:
        with mp.get_context("spawn").Pool(procs) as pl:
            res = pl.map(_cpu_worker, jobs)
    iters, wall = sum(r[0] for r in res), max(r[1] for r in res)
    return iters / wall, iters, wall


def _cxx_worker(args):
    lo, hi, seed = args
    import ctypes
    sys.path.insert(0, os.path.join(ROOT, "tests", "host_harness"))
    import build as hb
    import util as U
    from aslr_to_b200 import workloads as W
    hm = ctypes.CDLL(hb.LIB)
    w = W.sea_batch(B_PER_GPU, seed=seed)
    t0 = time.perf_counter()
    res = U.host_solve(hm, w, ocps=list(range(lo, hi)))
    st = res["states"]
    return int(np.sum(st["iter"] + (st["status"] == 1))), time.perf_counter() - t0


def cpu_cxx_rate(n_ocps, procs=1, seed=SEED):
    """Second CPU figure (baseline B): the very same __host__ __device__ solver source
    compiled for the host with g++ -O2 (tests/host_harness, test infrastructure) on `n_ocps` OCPs of the batch
    per process, `procs` processes over disjoint OCP ranges.  Not the reference's code path -- it shows what
    compiled FP64 code does on the host cores.  Returns (rate over the workers' own solve time, iterations,
    slowest worker's solve time)."""
    import multiprocessing as mp
    sys.path.insert(0, os.path.join(ROOT, "tests", "host_harness"))
    import build as hb
    hb.build()
    jobs = [(i * n_ocps, (i + 1) * n_ocps, seed) for i in range(procs)]
    if procs == 1:
        res = [_cxx_worker(jobs[0])]
    else:
        with mp.get_context("spawn").Pool(procs) as pool:
            res = pool.map(_cxx_worker, jobs)
    iters, wall = sum(r[0] for r in res), max(r[1] for r in res)
    return iters / wall, iters, wall


def run_reference(args):
    """--impl reference: the reference's CPU algorithm (oracle port: the reference's model maths on the
    NumPy restatement of crocoddyl/pinocchio -- the originals cannot be built here) on all
    host cores; each step is a bounded sample (one OCP per core) of the same workload.  The workers are started
    once; a step's time is the busiest worker's time inside its solves."""
    import multiprocessing as mp
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    t_total, it_total = 0.0, 0
    with mp.get_context("spawn").Pool(cores) as pool:
        for _ in range(args.warmup):
            cpu_oracle_rate(cores, cores, pool=pool)
        for _ in range(args.steps):
            rate, iters, wall = cpu_oracle_rate(cores, cores, pool=pool)
            t_total += wall
            it_total += iters
    value = it_total / t_total
    sample = "%d OCPs (one per core) of the 4096-OCP batch per step, full FDDP solve each" % cores
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * t_total / max(args.steps, 1),
        "higher_is_better": True, "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": workload_config(args.gpus, scaling=args.scaling),
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))


# -------------------------------------------------------------------------------- our CUDA arm
def run_ours(args):
    import ctypes as C
    import torch
    import torch.distributed as dist
    from aslr_to_b200 import _abi as A, batched as BT, sharding as S, workloads as W
    initial_guess_arrays = W.initial_guess

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback for the product path")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        # stdout carries exactly one JSON line: keep NCCL's version banner (NCCL_DEBUG=VERSION on this pool) off it
        if os.environ.get("NCCL_DEBUG", "").upper() in ("", "VERSION"):
            os.environ["NCCL_DEBUG"] = "WARN"
        dist.init_process_group("nccl", device_id=dev)

    strong = args.scaling == "strong"
    if strong:  # BASELINE.json configs[3]: ONE 4096-OCP batch, contiguous shards of 4096 / N (sharding.py)
        w = S.shard_workload(W.sea_batch(B_TOTAL, seed=SEED, solver=args.solver), world, rank)
        B_all = B_TOTAL
    else:       # replicas: every rank its own batch
        w = W.sea_batch(B_PER_GPU, seed=SEED + rank, solver=args.solver)
        B_all = B_PER_GPU * world
    B, T, nx, nu = int(np.asarray(w["x0"]).shape[0]), w["T"], 8, 2
    desc = W.to_desc(w)
    kind = W.SOLVER_KIND[args.solver]
    prob = BT.BatchedShootingProblem(desc, w["x0"], T, K_ocp=w["K_ocp"], device=dev)
    solver = BT.BatchedSolver(prob, kind, maxiter=w["maxiter"], th_stop=w["th_stop"], store_vxx=False,
                              store_quu=False, store_log=False)
    xs0_h, us0_h = initial_guess_arrays(w, np.asarray(w["x0"]), T, nu)
    # pinned host buffers = what a caller of the host-pointer C ABI holds
    pin = lambda a: torch.from_numpy(np.ascontiguousarray(a, dtype=np.float64)).pin_memory()
    h_x0, h_K, h_xs0, h_us0 = pin(w["x0"]), pin(np.asarray(w["K_ocp"]).reshape(B, 4)), pin(xs0_h), pin(us0_h)
    d_xs0, d_us0 = h_xs0.to(dev), h_us0.to(dev)
    h_xs = torch.empty((B, T + 1, nx), dtype=torch.float64).pin_memory()
    h_us = torch.empty((B, T, nu), dtype=torch.float64).pin_memory()
    h_state = torch.empty((B, A.OCP_STATE_DOUBLES), dtype=torch.float64).pin_memory()
    # second handle for the end-to-end path: aslr_solve_host owns its device workspace and does every copy itself
    lib = prob.lib
    e2e_handle = C.c_void_p()
    A.check(lib.aslr_problem_create(C.byref(desc), C.byref(e2e_handle)))
    e2e_result = {}

    def step_device():
        solver.t["xs"].copy_(d_xs0)
        solver.t["us"].copy_(d_us0)
        solver.solve(set_candidate=False)
        if world > 1:  # the only collective of the path: final gather of costs / iterations / trajectories
            S.gather_results({k: solver.t[k] for k in ("xs", "us", "state")}, B_all)

    def step_e2e():
        # the reference-facing call with HOST buffers: x0 / K / warm start in, xs / us / solver state out, the
        # host <-> device copies inside the call (include/aslr_b200.h: aslr_solve_host)
        e2e_result["r"] = BT.solve_host(desc, h_x0, T, kind, xs=h_xs0, us=h_us0, maxiter=w["maxiter"],
                                        th_stop=w["th_stop"], K_ocp=h_K, out={"xs": h_xs, "us": h_us, "state": h_state},
                                        handle=e2e_handle, lib=lib)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def timed(fn, steps):
        barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(steps):
            fn()
        e1.record()
        barrier()
        ms = torch.tensor([e0.elapsed_time(e1)], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    def timed_wall(fn, steps):
        """host-side clock around a blocking host-pointer call (its copies and kernels run on the library's own
        stream, which torch's events do not see); barrier + synchronize on both sides, max over ranks"""
        barrier()
        t0 = time.perf_counter()
        for _ in range(steps):
            fn()
        torch.cuda.synchronize()
        ms = torch.tensor([(time.perf_counter() - t0) * 1e3], dtype=torch.float64, device=dev)
        barrier()
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        return float(ms.item())

    if args.profile:  # for ncu: one warm-up solve, one profiled solve, nothing else
        step_device()
        torch.cuda.synchronize()
        step_device()
        torch.cuda.synchronize()
        for _ in range(2):  # the stand-alone fused calc+calcDiff sweep the roofline is quoted on
            prob.calc_diff(solver.t["xs"], solver.t["us"])
        torch.cuda.synchronize()
        print(json.dumps({"profile": True, "kernel_launches_per_solve": int(solver.result.kernel_launches)}))
        return
    if args.profile_floor:  # for ncu: the latency floor = a batch holding only the OCPs that never converge
        step_device()
        torch.cuda.synchronize()
        idx = np.nonzero(solver.states()["status"] != A.OCP_CONVERGED)[0]
        x0f, Kf = np.asarray(w["x0"])[idx], np.asarray(w["K_ocp"])[idx]
        probf = BT.BatchedShootingProblem(W.to_desc(w), x0f, T, K_ocp=Kf, device=dev)
        solf = BT.BatchedSolver(probf, kind, maxiter=w["maxiter"], th_stop=w["th_stop"],
                                store_vxx=False, store_quu=False, store_log=False)
        xf, uf = initial_guess_arrays(w, x0f, T, nu)
        for _ in range(2):
            solf.solve(xf, uf)
            torch.cuda.synchronize()
        print(json.dumps({"profile_floor": True, "ocps": idx.tolist(),
                          "iters": solf.states()["iter"].tolist()}))
        return
    for _ in range(max(args.warmup, 3)):
        step_device()
    torch.cuda.synchronize()
    st = solver.states()
    iters_per_solve = int(np.sum(st["iter"] + (st["status"] == A.OCP_CONVERGED)))
    n_conv = int(np.sum(st["status"] == A.OCP_CONVERGED))
    launches_per_step = int(solver.result.kernel_launches)
    tot = torch.tensor([float(iters_per_solve), float(n_conv)], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(tot)
    total_iters, total_conv = float(tot[0].item()), int(tot[1].item())
    step_e2e()  # warm-up of the host-pointer path (allocates its workspace)
    step_e2e()
    st_e2e = e2e_result["r"]["states"]
    assert np.array_equal(st_e2e["iter"], st["iter"]) and np.array_equal(st_e2e["cost"], st["cost"]), \
        "the host-pointer path and the device-pointer path disagree"

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = timed(step_device, args.steps)
    ms_e2e = timed_wall(step_e2e, args.steps)
    value = total_iters * args.steps / (ms * 1e-3)
    e2e_value = total_iters * args.steps / (ms_e2e * 1e-3)

    # ---- the metric's second half: the fused calc+calcDiff sweep over the shard's B*(T+1) knots
    xs_sol, us_sol = solver.t["xs"].clone(), solver.t["us"].clone()
    out = A.KnotOut()
    out.xnext, out.cost, out.deriv = solver.t["xnext"].data_ptr(), solver.t["knot_cost"].data_ptr(), \
        solver.t["deriv"].data_ptr()
    bs = prob.batch_struct()
    stream = prob._stream()

    def calcdiff_sweep(xs_t=xs_sol, us_t=us_sol):
        A.check(lib.aslr_calc_diff(prob.handle, C.byref(bs), xs_t.data_ptr(), us_t.data_ptr(), C.byref(out), stream))
    for _ in range(3):
        calcdiff_sweep()
    reps = 20
    ms_k1 = timed(lambda: calcdiff_sweep(), reps) / reps
    # the same sweep on random states: x ~ U(-pi, pi)^2n x U(-2, 2)^2n, u ~ U(-1, 1)
    g = torch.Generator(device=dev)
    g.manual_seed(SEED + rank)
    xs_rnd = torch.cat([(torch.rand((B, T + 1, 4), generator=g, device=dev, dtype=torch.float64) * 2 - 1) * np.pi,
                        (torch.rand((B, T + 1, 4), generator=g, device=dev, dtype=torch.float64) * 2 - 1) * 2.0], dim=2)
    us_rnd = torch.rand((B, T, nu), generator=g, device=dev, dtype=torch.float64) * 2 - 1
    for _ in range(3):
        calcdiff_sweep(xs_rnd, us_rnd)
    ms_k1_rnd = timed(lambda: calcdiff_sweep(xs_rnd, us_rnd), reps) / reps
    clocks = sampler.stop() if rank == 0 else None
    knots = B * (T + 1)
    knots_per_s = knots * world / (ms_k1 * 1e-3)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    achieved = ALGO_BYTES_PER_KNOT * knots / (ms_k1 * 1e-3) / 1e9

    # ---- where one solve spends its time: CUDA events around every stage of one more solve (untimed extra)
    stages = None
    if rank == 0:
        rows = np.array(stage_event_times(prob, solver, w, w["maxiter"], kind))
        stage_ms = rows.sum(0) * 1e-3
        stages = {"ms_per_solve": dict(zip(STAGES, [round(float(v), 3) for v in stage_ms])),
                  "share": dict(zip(STAGES, [round(float(v / stage_ms.sum()), 3) for v in stage_ms])),
                  "full_batch_us": dict(zip(STAGES, [float(v) for v in rows[1]])),
                  "floor_us": dict(zip(STAGES, [float(v) for v in rows[-1]])),
                  "note": "stepwise ABI with one event pair and one sync per stage, after the timed region; full_batch = "
                          "iteration 1, floor = last iteration (stragglers only)"}
    if world > 1:
        barrier()

    if rank == 0:
        line = {
            "metric": METRIC.replace("fddp", args.solver), "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": args.scaling, "vs_baseline": None, "dtype": "f64", "data": "synthetic",
            "config": workload_config(world, args.solver, args.scaling),
            "ocp_iters_per_solve_all_gpus": total_iters, "converged_all_gpus": total_conv,
            "calcdiff_knots_per_s": knots_per_s, "calcdiff_ms": ms_k1,
            "calcdiff_random_states": {"knots_per_s": knots * world / (ms_k1_rnd * 1e-3), "ms": ms_k1_rnd,
                                       "inputs": "x ~ U(-pi,pi)^4 x U(-2,2)^4, u ~ U(-1,1)^2 (SURVEY.md 8d config 4)"},
            "e2e": {"value": e2e_value, "unit": UNIT, "ms_per_step": ms_e2e / args.steps,
                    "api": "aslr_solve_host (C ABI, host pointers; pinned caller memory used in place)",
                    "h2d_bytes_per_step": int(h_x0.nbytes + h_K.nbytes + h_xs0.nbytes + h_us0.nbytes),
                    "d2h_bytes_per_step": int(h_xs.nbytes + h_us.nbytes + h_state.nbytes)},
            "gpu_launches": launches_per_step * args.steps,
            "roofline": roofline_block(rank, knots, ms_k1, achieved, peak, peaks, stages, B, T),
            "clocks": clocks, "stages": stages,
        }
        if world == 1 and not args.no_cpu:
            rate, iters, wall = cpu_oracle_rate(args.cpu_ocps, 1)
            line["cpu_baseline"] = {"value": rate, "unit": UNIT, "cores": 1, "kind": "port",
                                    "sample": "%d OCPs of the same batch, full FDDP solve each (%d iterations, %.1f s)"
                                              % (args.cpu_ocps, iters, wall)}
            try:
                what = "this repository's solver source compiled for the host (g++ -O2), not the reference's code path"
                crate, citers, cwall = cpu_cxx_rate(32)
                line["cpu_cxx"] = {"value": crate, "unit": UNIT, "cores": 1, "kind": what,
                                   "sample": "32 OCPs of the same batch (%d iterations, %.1f s)" % (citers, cwall)}
                cores = os.cpu_count() or 1
                if cores > 1:
                    crate, citers, cwall = cpu_cxx_rate(32, cores)
                    line["cpu_cxx_all_cores"] = {
                        "value": crate, "unit": UNIT, "cores": cores, "kind": what,
                        "sample": "32 OCPs per core, %d processes (%d iterations, slowest worker %.1f s)"
                                  % (cores, citers, cwall)}
            except Exception as e:  # no host compiler on the box: the figure is optional
                line["cpu_cxx"] = {"unavailable": str(e)[:120]}
        if world == 1 and not args.no_extras:
            # SURVEY.md section 8d extras, bounded: the SolverDDP line of config 4 and the config-5 MPC line
            try:
                line["extras"] = {"sea4096_ddp": extra_ddp_line(prob, w, d_xs0, d_us0),
                                  "chain7_mpc": extra_chain7_line()}
            except Exception as e:
                line["extras"] = {"error": str(e)[:200]}
        print(json.dumps(line))
    lib.aslr_problem_destroy(e2e_handle)
    if world > 1:
        dist.destroy_process_group()


# Algorithmic bytes per knot of the three solver stages at (nx, nu) = (8, 2), 8-byte words:
#   calcdiff (K1 in the solver): read x, u (10); write the compact record (56), xnext (8), cost (1), gap (8)
#   backward (K2): read the compact record (56), gap (8); write K (16), k, Qu, Quuk (6), Vx, Vxx fs (16)
#   rollout (K3, per trial and knot): read K (16), xs (8), us (2), k (2); write xs_try (8), us_try (2)
STAGE_BYTES_PER_KNOT = {"calcdiff": 8 * 83, "backward": 8 * 102, "rollout": 8 * 38}
STAGE_FLOP_KEY = {"calcdiff": "sea2.calc_diff_knot", "backward": "sea2.backward_knot", "rollout": "sea2.rollout_knot"}


def roofline_block(rank, knots, ms_k1, achieved, peak, peaks, stages, B=None, T=None):
    """`roofline` of the bench line = the stage that takes the largest share of a solve (measured live, CUDA events),
    at full batch, against BOTH roofs: achieved = max(algorithmic bytes/s / HBM peak, flop/s / FP64 peak), flop counts
    from profiles/flop_counts.json (tools/flop_count.cpp: the product sources compiled over a counting scalar type).
    `calcdiff_sweep` keeps the HBM figure of the public fused calc+calcDiff sweep on SURVEY.md's 1544 B / knot.
    ncu-derived figures (DRAM traffic per launch, FP64 pipe utilisation) are read from profiles/ncu_figures.json, written
    by tools/ncu_summary.py together with the hash of the kernel sources they were captured from; when the sources
    have changed since, they are withheld rather than quoted stale."""
    figs, stale = {}, None
    try:
        figs = json.load(open(os.path.join(ROOT, "profiles", "ncu_figures.json")))
        import __graft_entry__ as ge
        if figs.get("source_hash") != ge._source_hash():
            stale = "profiles/ncu_figures.json was captured from other kernel sources (%s): figures withheld" % \
                    str(figs.get("source_hash"))[:12]
    except Exception as e:
        stale = "profiles/ncu_figures.json unavailable (%s)" % type(e).__name__
    kern = (figs.get("kernels", {}) or {}) if not stale else {}
    flops, fp64_peak, fp64_src = {}, None, None
    try:
        flops = json.load(open(os.path.join(ROOT, "profiles", "flop_counts.json")))
    except Exception:
        pass
    try:
        fp = json.load(open(os.path.join(ROOT, "profiles", "r01b_fp64_peak.json")))
        fp64_peak, fp64_src = float(fp["fp64_tflops"]), "tools/fp64_peak.cu on this pool (profiles/r01b_fp64_peak.json)"
    except Exception:
        pass
    peak_src = "MEASURED_PEAKS.json hbm_gbs (of measured)" if peaks else "fallback 6650 (of fallback)"
    sweep = {"kernel": "knot_kernel<2,2,SEA,DIFF> (public fused calc+calcDiff sweep, %d knots)" % knots, "bound": "hbm",
             "achieved": achieved, "peak": peak, "peak_source": peak_src, "unit": "GB/s", "frac": achieved / peak,
             "traffic": (kern.get("knot_kernel") or {}).get("dram_bytes_per_launch"),
             "algorithmic_bytes_per_knot": ALGO_BYTES_PER_KNOT}
    if fp64_peak and "sea2.calc_diff_knot" in flops:
        tf = flops["sea2.calc_diff_knot"]["flops"] * knots / (ms_k1 * 1e-3) / 1e12
        sweep["fp64"] = {"flops_per_knot": flops["sea2.calc_diff_knot"]["flops"], "achieved_tflops": tf,
                         "peak_tflops": fp64_peak, "frac": tf / fp64_peak}
    if not stages or B is None:
        sweep["ncu_figures"] = stale or {"source": "profiles/ncu_figures.json", "capture": figs.get("capture")}
        return sweep
    # ---- the dominant stage of the solve
    per_stage = {}
    names = {"calcdiff": "solver_calcdiff_kernel", "backward": "backward_cols_kernel", "rollout": "rollout_kernel"}
    for st in ("calcdiff", "backward", "rollout"):
        us = stages["full_batch_us"][st]
        n_knots = B * (T + 1) if st == "calcdiff" else B * T * (10 if st == "rollout" else 1)
        gbs = STAGE_BYTES_PER_KNOT[st] * n_knots / (us * 1e-6) / 1e9
        e = {"share_of_solve": stages["share"][st], "full_batch_us": us, "knots_per_launch": n_knots,
             "algorithmic_bytes_per_knot": STAGE_BYTES_PER_KNOT[st], "achieved_gbs": gbs, "frac_hbm": gbs / peak,
             "floor_us": stages["floor_us"][st]}
        fk = STAGE_FLOP_KEY[st]
        if fp64_peak and fk in flops:
            tf = flops[fk]["flops"] * n_knots / (us * 1e-6) / 1e12
            e.update(flops_per_knot=flops[fk]["flops"], achieved_tflops=tf, frac_fp64=tf / fp64_peak)
        if names[st] in kern:
            e["ncu"] = kern[names[st]]
        per_stage[st] = e
    dom = max(per_stage, key=lambda k: per_stage[k]["share_of_solve"])
    d = per_stage[dom]
    by_flops = d.get("frac_fp64", 0.0) > d["frac_hbm"]
    block = {"kernel": "%s (stage '%s': %.0f %% of the solve, full batch = iteration 1: all OCPs running%s)"
                       % (names[dom], dom, 100 * d["share_of_solve"],
                          ", all ten step lengths rolled out" if dom == "rollout" else ""),
             "bound": "hbm",
             "achieved": d["achieved_gbs"], "peak": peak, "peak_source": peak_src, "unit": "GB/s",
             "frac": max(d["frac_hbm"], d.get("frac_fp64", 0.0)),
             "frac_is": "flop/s over the measured FP64 peak" if by_flops else "algorithmic bytes/s over the measured HBM peak",
             "traffic": (d.get("ncu") or {}).get("dram_bytes_per_launch"),
             "fp64": {"peak_tflops": fp64_peak, "peak_source": fp64_src, "achieved_tflops": d.get("achieved_tflops"),
                      "flop_counts": "profiles/flop_counts.json (dense serial form of the stage's arithmetic)"},
             "note": "K2 / K3 are time-sequential dependent chains (one knot after the other) with one to seven warps "
                     "per SM: neither roof binds, the limiter is dependent-issue latency (stages.floor_us is what a "
                     "launch costs when two OCPs are left); both fractions are given so the distance is visible",
             "stages": per_stage, "calcdiff_sweep": sweep}
    block["ncu_figures"] = stale or {"source": "profiles/ncu_figures.json", "capture": figs.get("capture")}
    return block


def extra_ddp_line(prob, w, d_xs0, d_us0, steps=3):
    """SolverDDP on the same 4096-OCP batch ('Solver FDDP and DDP')."""
    import torch
    from aslr_to_b200 import _abi as A, batched as BT
    solver = BT.BatchedSolver(prob, A.SOLVER_DDP, maxiter=w["maxiter"], th_stop=w["th_stop"], store_vxx=False,
                              store_quu=False, store_log=False)

    def step():
        solver.t["xs"].copy_(d_xs0)
        solver.t["us"].copy_(d_us0)
        solver.solve(set_candidate=False)
    for _ in range(2):
        step()
    torch.cuda.synchronize()
    st = solver.states()
    iters = int(np.sum(st["iter"] + (st["status"] == A.OCP_CONVERGED)))
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    return {"metric": "batched_ddp_ocp_iters_per_s", "value": iters / (ms * 1e-3), "unit": UNIT, "ms_per_step": ms,
            "ocp_iters_per_solve": iters, "converged": int(np.sum(st["status"] == A.OCP_CONVERGED)), "steps": steps,
            "warmup": 2}


def extra_chain7_line(steps=5, warmup=2):
    """BASELINE.json configs[4] as an extra key of the default line (bounded: 2 warm-up + 5 MPC steps)."""
    import argparse as ap
    import io
    import contextlib
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        run_chain7_mpc(ap.Namespace(chain7_ocps=1024, steps=steps, warmup=warmup))
    return json.loads(buf.getvalue().strip().splitlines()[-1])


def run_chain7_mpc(args):
    """Secondary workload (BASELINE.json configs[4]): 1024 x 7-link SEA arm (nx = 28), T = 200.  One untimed
    solve to (near) convergence, then timed MPC steps: shift the solution by one knot, perturb x0 by N(0, 1e-2)
    and re-solve with maxiter = 10 from the warm start.  Single GPU; prints its own JSON line."""
    import torch
    from aslr_to_b200 import _abi as A, batched as BT, mpc as M, workloads as W
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: there is no CPU fallback for the product path")
    torch.cuda.set_device(0)
    B, T = args.chain7_ocps, 200
    w = W.chain7_batch(B=B, T=T)
    prob = BT.BatchedShootingProblem(W.to_desc(w), w["x0"], T, mref_ocp=w["mref_ocp"])
    mpc = M.BatchedMPC(prob, A.SOLVER_FDDP, maxiter=10, th_stop=w["th_stop"], store_log=False)
    xs0, us0 = W.initial_guess(w, w["x0"], T, prob.nu)
    t0 = time.perf_counter()
    first = mpc.initialise(xs0, us0, maxiter=100)
    torch.cuda.synchronize()
    first_s = time.perf_counter() - t0
    st0 = first.states()
    rng = np.random.default_rng(5)
    noise = [torch.as_tensor(1e-2 * rng.standard_normal((B, prob.nx)), device=prob.device)
             for _ in range(args.warmup + args.steps)]
    iters = 0

    def one(k):
        x_meas = mpc.solver.t["xs"][:, 1, :] + noise[k]
        mpc.step(x_meas)
        st = mpc.solver.states()
        return int(np.sum(st["iter"] + (st["status"] == A.OCP_CONVERGED)))
    for k in range(args.warmup):
        one(k)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for k in range(args.steps):
        iters += one(args.warmup + k)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    print(json.dumps({
        "metric": "batched_fddp_ocp_iters_per_s", "value": iters / (ms * 1e-3), "unit": UNIT, "n_gpus": 1,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": "%d x 7-link SEA arm MPC warm-start steps (BASELINE.json configs[4]), T=200, nx=28, "
                               "maxiter=10 per step" % B, "ocps": B, "T": T, "nx": 28, "nu": 7},
        "first_solve_s": first_s, "first_solve_converged": int(np.sum(st0["status"] == A.OCP_CONVERGED)),
        "first_solve_iters": int(np.sum(st0["iter"])), "ocp_iters_per_step": iters / args.steps,
        "gpu_launches": int(mpc.solver.result.kernel_launches) * args.steps}))


STAGES = ("calcdiff", "backward", "rollout", "accept")


def stage_event_times(prob, solver, w, n_iters, kind):
    """CUDA-event time (us) of the four stages of the first `n_iters` solver iterations, driven through the
    stepwise ABI from the workload's initial guess: rows of [calcdiff, backward, rollout + trial costs, accept]."""
    import ctypes as C
    import torch
    from aslr_to_b200 import _abi as A, batched as BT, workloads as W
    xs0, us0 = W.initial_guess(w, np.asarray(w["x0"]), prob.T, prob.nu)
    solver.set_candidate(xs0, us0)
    lib, bs, st = prob.lib, prob.batch_struct(), prob._stream()
    opts = BT.make_opts(kind, w["maxiter"], w["th_stop"], lib=lib)
    A.check(lib.aslr_solver_init(prob.handle, C.byref(bs), C.byref(opts), C.byref(solver.buf), st))
    rows = []
    for it in range(n_iters):
        row = []
        for fn, extra in ((lib.aslr_solver_calc_diff, (it,)), (lib.aslr_backward_pass, ()),
                          (lib.aslr_forward_pass, ()), (lib.aslr_accept_step, (it,))):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            A.check(fn(prob.handle, C.byref(bs), C.byref(opts), C.byref(solver.buf), *extra, st))
            e1.record()
            torch.cuda.synchronize()
            row.append(round(e0.elapsed_time(e1) * 1e3, 1))
        rows.append(row)
    return rows


def run_kernel_times(args):
    """--kernels N: CUDA-event time of each solver kernel in the first N iterations (full batch active)."""
    import torch
    from aslr_to_b200 import _abi as A, batched as BT, workloads as W
    torch.cuda.set_device(0)
    if args.workload == "chain7_mpc":
        w = W.chain7_batch(B=args.chain7_ocps, T=200)
        prob = BT.BatchedShootingProblem(W.to_desc(w), w["x0"], 200, mref_ocp=w["mref_ocp"])
    else:
        w = W.sea_batch(B_PER_GPU, seed=SEED)
        prob = BT.BatchedShootingProblem(W.to_desc(w), w["x0"], w["T"], K_ocp=w["K_ocp"])
    solver = BT.BatchedSolver(prob, A.SOLVER_FDDP, maxiter=w["maxiter"], th_stop=w["th_stop"], store_vxx=False,
                              store_quu=False, store_log=False)
    rows = stage_event_times(prob, solver, w, args.kernels, A.SOLVER_FDDP)
    print(json.dumps({"workload": args.workload, "kernel_us_per_iteration[calcdiff,backward,rollout,accept]": rows}))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--cpu-ocps", type=int, default=2, help="OCPs in the bounded CPU-oracle sample")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--kernels", type=int, default=0, help="time each kernel of the first N iterations")
    ap.add_argument("--profile", action="store_true", help="two solves only (for ncu launch lists)")
    ap.add_argument("--profile-floor", action="store_true",
                    help="one full solve, then two solves of only the never-converging OCPs (for ncu: the latency floor)")
    ap.add_argument("--workload", default="sea4096", choices=["sea4096", "chain7_mpc"])
    ap.add_argument("--chain7-ocps", type=int, default=1024)
    ap.add_argument("--solver", default="fddp", choices=["fddp", "ddp"], help="sea4096 workload: SolverFDDP (headline) or SolverDDP")
    ap.add_argument("--scaling", default="strong", choices=["strong", "weak"],
                    help="strong (default, BASELINE.json configs[3]): one 4096-OCP batch cut into 4096 / N shards; "
                         "weak: 4096 OCPs per GPU (round 1's replica mode)")
    ap.add_argument("--no-extras", action="store_true", help="skip the SolverDDP and chain7 MPC extra lines (N = 1 only)")
    args = ap.parse_args()
    # stdout carries exactly one JSON line.  Native libraries write banners to file descriptor 1 (NCCL prints its
    # version there whatever NCCL_DEBUG says on this pool): point fd 1 at stderr and keep Python's stdout on a
    # duplicate of the real one.
    sys.stdout.flush()
    real_stdout = os.fdopen(os.dup(1), "w", buffering=1)
    os.dup2(2, 1)
    sys.stdout = real_stdout
    if args.kernels:
        run_kernel_times(args)
    elif args.workload == "chain7_mpc":
        run_chain7_mpc(args)
    elif args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
