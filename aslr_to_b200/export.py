"""Result export: what the reference's example scripts write at their end
(`savemat` of the optimised trajectory and controls, examples/two_dof_vsa_boxddp.py:104-127,
examples/two_dof_sea.py:121-143) plus a CallbackLogger-compatible per-iteration log, from a batched solve.
Host-side bookkeeping only: the arrays come from the device buffers of a finished `BatchedSolver`."""
import numpy as np

from . import _abi as A

LOG_COLUMNS = ("costs", "stops", "x_regs", "steps", "dVs", "dVexps", "ffeass", "is_feasible")


def iterations_logged(states, b):
    """Number of per-iteration log rows OCP b wrote (crocoddyl calls the callbacks once per completed iteration)."""
    it, status = int(states["iter"][b]), int(states["status"][b])
    return it + 1 if status == A.OCP_CONVERGED else it


def callback_log(logs, states, b):
    """dict with the attribute names of crocoddyl.CallbackLogger for OCP b (`logs`: BatchedSolver.logs())."""
    n = iterations_logged(states, b)
    rows = np.asarray(logs)[b, :n]
    out = {name: rows[:, k].copy() for k, name in enumerate(LOG_COLUMNS)}
    out["u_regs"] = out["x_regs"].copy()  # crocoddyl keeps x_reg == u_reg
    out["iters"] = np.arange(n)
    return out


def trajectory_dict(xs, us, dt, n):
    """The variables the examples save: link positions q1.., motor positions th1.., controls u1.., time."""
    xs, us = np.asarray(xs, float), np.asarray(us, float)
    T = us.shape[0]
    d = {"t": np.arange(T) * float(dt)}
    for i in range(n):
        d["q%d" % (i + 1)] = xs[:, i].copy()
        d["th%d" % (i + 1)] = xs[:, n + i].copy()
    for j in range(us.shape[1]):
        d["u%d" % (j + 1)] = us[:, j].copy()
    return d


def save_solution(path, solver, b=0, dt=None):
    """Write OCP b of a finished BatchedSolver to `path` (.npz, or .mat through scipy.io.savemat)."""
    p = solver.problem
    st = solver.states()
    data = trajectory_dict(solver.t["xs"][b].cpu().numpy(), solver.t["us"][b].cpu().numpy(),
                           p.desc.dt if dt is None else dt, p.n)
    data.update(xs=solver.t["xs"][b].cpu().numpy(), us=solver.t["us"][b].cpu().numpy(),
                K=solver.t["Kgain"][b].cpu().numpy(), k=solver.t["kff"][b].cpu().numpy(),
                cost=float(st["cost"][b]), iterations=iterations_logged(st, b), status=int(st["status"][b]))
    if "log" in solver.t:
        data.update({"log_" + k: v for k, v in callback_log(solver.logs(), st, b).items()})
    if str(path).endswith(".mat"):
        from scipy.io import savemat
        savemat(str(path), data)
    else:
        np.savez(str(path), **data)
    return data
