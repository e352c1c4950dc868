#!/usr/bin/env python3
"""Generates the golden fixtures in tests/golden/ by running the REFERENCE'S OWN, UNMODIFIED model
classes (/root/reference/python/aslr_to) on the NumPy shims of pinocchio / crocoddyl.

Must be run in the build container (where /root/reference exists):
    python tests/golden/make_golden.py
The fixtures travel to the GPU box; the reference does not.

  knots_<workload>.npz    seeded (x, u) -> xnext, cost, xout, dFx, dFu, Fx, Fu, Lx, Lu, Lxx, Lxu, Luu for the
                          running model, and the terminal model evaluated with u = None
  solve_<workload>.npz    full solver trace of the example (per-iteration cost, stop, x_reg, step, dV,
                          dVexp; final xs, us, K, k, Vx, Vxx, Qu; iteration count; converged flag)
  solve_sea_batch_fddp_b924.npz   one of the two members of the 4096-OCP headline batch that never converge
                          (924 and 3415: all 100 iterations, 43 % of the timed region of bench.py)
  solve_chain7_T200_b<k>.npz      7-link arm at the bench's horizon T = 200: OCP 2 (converges, 50 iterations)
                          and OCP 1 (does not converge in the bench's first solve; trace cut at 25 iterations)
"""
import os
import sys
import time

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))

import oracle_build as ob  # noqa: E402
from aslr_to_b200 import workloads as W  # noqa: E402

KNOT_FIELDS = ("xnext", "cost", "xout", "dFx", "dFu", "Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu")


def knot_samples(w, nsamples, seed):
    rng = np.random.default_rng(seed)
    n = len(ob.robot_model(w).joints)
    nu = n if w["kind"] == "sea" else 2 * n
    xs = np.concatenate([rng.uniform(-np.pi, np.pi, (nsamples, 2 * n)), rng.uniform(-2, 2, (nsamples, 2 * n))], axis=1)
    us = rng.uniform(-1, 1, (nsamples, nu))
    if w["kind"] == "vsa":
        us[:, n:] = rng.uniform(0.05, 5.0, (nsamples, n))
    return xs, us


def eval_knots(models, xs, us):
    run, term, _ = models
    out = dict((("run_" + k), []) for k in KNOT_FIELDS)
    out.update(dict((("term_" + k), []) for k in KNOT_FIELDS))
    for tag, m in (("run_", run), ("term_", term)):
        d = m.createData()
        for x, u in zip(xs, us):
            if tag == "run_":
                m.calc(d, x, u)
                m.calcDiff(d, x, u)
            else:
                m.calc(d, x)
                m.calcDiff(d, x)
            diff = getattr(d, "differential", d)
            vals = dict(xnext=d.xnext, cost=d.cost, xout=diff.xout, dFx=getattr(diff, "dFx", diff.Fx),
                        dFu=getattr(diff, "dFu", diff.Fu), Fx=d.Fx, Fu=d.Fu, Lx=d.Lx, Lu=d.Lu, Lxx=d.Lxx, Lxu=d.Lxu,
                        Luu=d.Luu)
            for k in KNOT_FIELDS:
                out[tag + k].append(np.array(vals[k], float))
    return dict((k, np.array(v)) for k, v in out.items())


def solver_trace(solver, log):
    return dict(costs=np.array(log.costs), stops=np.array(log.stops), x_regs=np.array(log.x_regs),
                steps=np.array(log.steps), dVs=np.array(log.dVs), dVexps=np.array(log.dVexps),
                xs=np.array(solver.xs), us=np.array(solver.us), K=np.array(solver.K), k=np.array(solver.k),
                Vx=np.array(solver.Vx), Vxx=np.array(solver.Vxx), Qu=np.array(solver.Qu),
                iter=np.array(solver.iter), converged=np.array(bool(solver.converged)),
                cost=np.array(solver.cost), stop=np.array(solver.stop))


def _have(path):
    return os.path.isfile(path) and "--force" not in sys.argv


def main():
    if not ob.refenv.have_reference():
        raise SystemExit("/root/reference is not available: golden vectors can only be made in the build container")
    singles = [W.double_pendulum(), W.double_pendulum_residual(), W.two_dof_sea(), W.two_dof_vsa_boxddp(),
               W.two_dof_vsa_modified()]
    c7 = W.chain7_batch(B=2, T=20)
    c7v = dict(c7, name="chain7_vsa", kind="vsa", K=None,
               running=[t for t in c7["running"] if t["kind"] != "control"] +
               [dict(name="uReg", kind="control", weight=1e-2, ref=np.zeros(14), act=np.ones(14), frame=None)])
    for w, ns in [(s, 16) for s in singles] + [(c7, 6), (c7v, 6)]:
        if _have(os.path.join(HERE, "knots_%s.npz" % w["name"].split("_batch")[0])):
            continue
        xs, us = knot_samples(w, ns, 1234)
        res = eval_knots(ob.reference_models(w, 0), xs, us)
        np.savez_compressed(os.path.join(HERE, "knots_%s.npz" % w["name"].split("_batch")[0]), xs=xs, us=us, **res)
        print("knots", w["name"], ns)
    for w in singles:
        t0 = time.time()
        if _have(os.path.join(HERE, "solve_%s.npz" % w["name"])):
            continue
        solver, log = ob.solve(w, 0, "reference")
        np.savez_compressed(os.path.join(HERE, "solve_%s.npz" % w["name"]), **solver_trace(solver, log))
        print("solve", w["name"], "iters", solver.iter, "converged", solver.converged, "cost", solver.cost,
              "%.1fs" % (time.time() - t0))
    # two members of the 4096-OCP batch (random x0 / stiffness), FDDP and DDP
    for solver_name in ("fddp", "ddp"):
        wb = W.sea_batch(4096, solver=solver_name)
        for b in (0, 1):
            t0 = time.time()
            if _have(os.path.join(HERE, "solve_sea_batch_%s_b%d.npz" % (solver_name, b))):
                continue
            solver, log = ob.solve(wb, b, "reference")
            np.savez_compressed(os.path.join(HERE, "solve_sea_batch_%s_b%d.npz" % (solver_name, b)),
                                **solver_trace(solver, log))
            print("solve sea_batch", solver_name, b, "iters", solver.iter, solver.converged,
                  "%.1fs" % (time.time() - t0))
    # a member of the headline batch that stops at maxiter (found with the g++ build of the solver: 924, 3415)
    path = os.path.join(HERE, "solve_sea_batch_fddp_b924.npz")
    if not _have(path):
        t0 = time.time()
        solver, log = ob.solve(W.sea_batch(4096), 924, "reference")
        np.savez_compressed(path, **solver_trace(solver, log))
        print("solve sea_batch fddp 924 iters", solver.iter, solver.converged, "%.1fs" % (time.time() - t0))
    # the 7-link arm at the bench's horizon
    for b, maxiter in ((2, 100), (1, 25)):
        path = os.path.join(HERE, "solve_chain7_T200_b%d.npz" % b)
        if _have(path):
            continue
        t0 = time.time()
        solver, log = ob.solve(W.chain7_batch(B=1024, T=200), b, "reference", maxiter=maxiter)
        tr = solver_trace(solver, log)
        tr.pop("Vxx")  # 201 x 28 x 28: three quarters of the file; K, k, Vx, Qu stay
        tr["maxiter"] = np.array(maxiter)
        np.savez_compressed(path, **tr)
        print("solve chain7 T=200", b, "iters", solver.iter, solver.converged, "%.1fs" % (time.time() - t0))


if __name__ == "__main__":
    main()
