"""Compiles the compat layer's parameter-holder objects into an `AslrProblemDesc` and runs them on the
GPU through `BatchedShootingProblem` / `BatchedSolver`.  Host-side bookkeeping only."""
import ctypes as C

import numpy as np

from .. import _abi as A
from .. import desc as D
from ..batched import BatchedShootingProblem

_KIND = {"placement": A.COST_FRAME_PLACEMENT, "state": A.COST_STATE, "control": A.COST_CONTROL,
         "stiffness": A.COST_STIFFNESS, "pendulum": A.COST_DOUBLE_PENDULUM,
         "pendulum_residual": A.COST_RESIDUAL_DOUBLE_PENDULUM}


def term(kind, weight, ref=None, act=None, frame=None):
    """One entry of a cost sum: kind in _KIND, `frame` = (parent joint, R, p) for placement terms."""
    return dict(kind=kind, weight=float(weight), ref=None if ref is None else np.asarray(ref, float).ravel(),
                act=None if act is None else np.asarray(act, float).ravel(), frame=frame)


def _abi_terms(terms):
    out = []
    for t in terms:
        if t["kind"] == "placement":
            parent, R, p = t["frame"]
            out.append(D.cost_term(A.COST_FRAME_PLACEMENT, t["weight"], t["ref"], t["act"], parent, R, p))
        else:
            out.append(D.cost_term(_KIND[t["kind"]], t["weight"], t["ref"], t["act"]))
    return out


class ModelSpec(object):
    """What one action model contributes to the description: differential model + time step."""

    def __init__(self, differential, dt):
        self.diff, self.dt = differential, float(dt)

    def key(self):
        d = self.diff
        return (id(d.state.pinocchio), d._kind, self.dt)

    def cost_terms(self):
        costs = self.diff.costs
        if not hasattr(costs, "_terms"):
            raise NotImplementedError("cost container %s is not supported (use crocoddyl.CostModelSum)" % type(costs).__name__)
        return costs._terms()


def _check_same_dynamics(a, b):
    da, db = a.diff, b.diff
    same = (da._kind == db._kind and da.state.pinocchio is db.state.pinocchio and np.array_equal(da.B, db.B) and
            (da.K is None) == (db.K is None) and (da.K is None or np.array_equal(da.K, db.K)) and
            np.array_equal(da.actuation._Sm(), db.actuation._Sm()))
    if not same:
        raise NotImplementedError("running and terminal models must share robot, dynamics kind, K, B and actuation "
                                  "(they do in every example of the reference)")


def compile_desc(running, terminal=None):
    """ModelSpec (running) [+ ModelSpec (terminal)] -> AslrProblemDesc."""
    terminal = running if terminal is None else terminal
    _check_same_dynamics(running, terminal)
    d = running.diff
    model = d.state.pinocchio
    n = model.nv
    dyn = D.make_dynamics(d._kind, n, d.B, d.K, d.actuation._Sm(), d._write_fu(),
                          None if not np.all(np.isfinite(d.u_lb)) else d.u_lb,
                          None if not np.all(np.isfinite(d.u_ub)) else d.u_ub)
    if dyn.nu != d.nu:
        raise ValueError("control dimension %d does not match the dynamics (%d)" % (d.nu, dyn.nu))
    robot = D.make_robot(model._dict, model.gravity.linear)
    return D.make_problem_desc(robot, dyn, _abi_terms(running.cost_terms()), _abi_terms(terminal.cost_terms()),
                               running.dt, terminal.dt if terminal is not running else running.dt)


def eval_knot(spec, x, u, diff):
    """calc (+ calcDiff) of one action model at one (x, u) on the GPU.  u = None -> the model's `u is None`
    default (free_fwddyn_asr.py:21-22, free_fwddyn_vsa.py:21-23)."""
    desc = compile_desc(spec)
    x = np.asarray(x, float).ravel()
    prob = BatchedShootingProblem(desc, x[None, :], 1)
    if x.shape[0] != prob.nx:
        raise ValueError("x has wrong dimension (it should be %d)" % prob.nx)
    xs = np.stack([x, x])[None]
    us = None
    if u is not None:
        u = np.asarray(u, float).ravel()
        if u.shape[0] != prob.nu:
            raise ValueError("u has wrong dimension (it should be %d)" % prob.nu)
        us = u[None, None, :]
    want = ("xnext", "cost", "xout") + (("deriv", "dFx", "dFu") if diff else ())
    res = (prob.calc_diff if diff else prob.calc)(xs, us, want=want)
    out = {"xnext": res["xnext"][0, 0].cpu().numpy(), "cost": float(res["cost"][0, 0].item()),
           "xout": res["xout"][0, 0].cpu().numpy()}
    if diff:
        out.update({k: v[0, 0] for k, v in prob.split_record(res["deriv"].cpu().numpy()).items()})
        out["dFx"] = res["dFx"][0, 0].cpu().numpy()
        out["dFu"] = res["dFu"][0, 0].cpu().numpy()
    return out


def quasi_static(spec, xs, maxiter=100, tol=1e-9):
    """ActionModelAbstract::quasiStatic for every row of xs (K, nx) -> (K, nu), on the GPU."""
    desc = compile_desc(spec)
    xs = np.ascontiguousarray(xs, float)
    K = xs.shape[0]
    prob = BatchedShootingProblem(desc, xs[:1], K)  # one OCP whose K running knots are the K states
    return prob.quasi_static(xs[None], maxiter, tol)[0].cpu().numpy()
