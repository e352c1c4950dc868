"""The crocoddyl calls the reference's examples make (examples/two_dof_sea.py:27-81,
examples/two_dof_vsa_boxddp.py:27-81, examples/double_pendulum.py:22-70), as parameter holders in front of
the CUDA library.  `ShootingProblem` compiles its models into one `AslrProblemDesc`;
`SolverDDP / SolverFDDP / SolverBoxDDP.solve` run the batched solver with B = 1."""
import types

import numpy as np

from .. import _abi as A
from ..batched import BatchedShootingProblem, BatchedSolver
from . import _engine as E


# ------------------------------------------------------------------------------------ activations / residuals
class ActivationModelQuad(object):
    def __init__(self, nr):
        self.nr = int(nr)
        self.weights = np.ones(self.nr)


class ActivationModelWeightedQuad(object):
    def __init__(self, weights):
        self.weights = np.asarray(weights, float).ravel().copy()
        self.nr = self.weights.shape[0]


class ResidualModelState(object):
    """ResidualModelState(state, xref, nu) (examples/two_dof_sea.py:31): r = diff(xref, x)."""

    def __init__(self, state, xref=None, nu=None):
        if np.ndim(xref) == 0 and xref is not None:  # (state, nu) overload
            xref, nu = None, xref
        self.state = state
        self.reference = state.zero() if xref is None else np.asarray(xref, float).ravel().copy()
        self.nu = state.nv if nu is None else int(nu)
        self.nr = state.ndx


class ResidualModelControl(object):
    """ResidualModelControl(state, nu) / (state, uref) (examples/two_dof_sea.py:33): r = u - uref."""

    def __init__(self, state, uref_or_nu=None):
        self.state = state
        if uref_or_nu is None:
            uref_or_nu = state.nv
        if np.ndim(uref_or_nu) == 0:
            self.nu = int(uref_or_nu)
            self.reference = np.zeros(self.nu)
        else:
            self.reference = np.asarray(uref_or_nu, float).ravel().copy()
            self.nu = self.reference.shape[0]
        self.nr = self.nu


class CostModelResidual(object):
    """CostModelResidual(state, residual) / (state, activation, residual)."""

    def __init__(self, state, a, b=None):
        self.state = state
        activation, residual = (None, a) if b is None else (a, b)
        if activation is None:
            activation = ActivationModelQuad(residual.nr)
        if activation.nr != residual.nr:
            raise ValueError("nr is equals to %d" % residual.nr)
        self.activation, self.residual = activation, residual
        self.nu, self.nr = residual.nu, residual.nr

    def _term(self, weight):
        r, w = self.residual, self.activation.weights
        if isinstance(r, ResidualModelState):
            return E.term("state", weight, ref=r.reference, act=w)
        if isinstance(r, ResidualModelControl):
            return E.term("control", weight, ref=r.reference, act=w)
        if type(r).__name__ == "ResidualModelFramePlacementASR":
            M = r.reference
            ref = np.concatenate([np.asarray(M.rotation, float).reshape(9), np.asarray(M.translation, float).reshape(3)])
            return E.term("placement", weight, ref=ref, act=w, frame=r.state.pinocchio._frame_desc(r.id))
        if type(r).__name__ == "ResidualModelDoublePendulum":
            return E.term("pendulum_residual", weight, act=w)
        raise NotImplementedError("residual %s has no CUDA implementation" % type(r).__name__)


class CostModelSum(object):
    """CostModelSum(state, nu).addCost(name, cost, weight) (examples/two_dof_sea.py:27-47).  Terms are
    evaluated in name order, as crocoddyl's std::map iterates."""

    def __init__(self, state, nu=None):
        self.state = state
        self.nu = state.nv if nu is None else int(nu)
        self.costs = {}

    def addCost(self, name, cost, weight, active=True):
        if name in self.costs:
            raise ValueError("cost %r already exists" % name)
        if not hasattr(cost, "_term"):
            raise NotImplementedError("cost %s has no CUDA implementation" % type(cost).__name__)
        self.costs[name] = types.SimpleNamespace(name=name, cost=cost, weight=float(weight), active=bool(active))

    def removeCost(self, name):
        del self.costs[name]

    def changeCostStatus(self, name, active):
        self.costs[name].active = bool(active)

    @property
    def nr(self):
        return sum(c.cost.nr for c in self.costs.values() if c.active)

    def _terms(self):
        return [self.costs[k].cost._term(self.costs[k].weight) for k in sorted(self.costs) if self.costs[k].active]


# ------------------------------------------------------------------------------------ shooting problem
class ShootingProblem(object):
    """ShootingProblem(x0, [runningModel] * T, terminalModel) (examples/two_dof_sea.py:66).  The kernels
    share one running model over the horizon, so the running models must be one object (or equal)."""

    def __init__(self, x0, runningModels, terminalModel):
        runningModels = list(runningModels)
        if not runningModels:
            raise ValueError("at least one running model is required")
        first = runningModels[0]
        if any(m is not first for m in runningModels):
            raise NotImplementedError("time-varying running models are not supported: pass [model] * T")
        self.x0 = np.asarray(x0, float).ravel().copy()
        self.runningModels, self.terminalModel = runningModels, terminalModel
        self.T = len(runningModels)
        self.nthreads = 1
        self.runningDatas = [m.createData() for m in runningModels]
        self.terminalData = terminalModel.createData()
        self._batched = None

    # --- compilation (re-done when a model parameter such as u_lb changed after construction)
    def _specs(self):
        return (E.ModelSpec(self.runningModels[0].differential, self.runningModels[0].timeStep),
                E.ModelSpec(self.terminalModel.differential, self.terminalModel.timeStep))

    def _problem(self):
        run, term = self._specs()
        desc = E.compile_desc(run, term)
        blob = bytes(desc)
        if self._batched is None or self._batched[0] != blob:
            self._batched = (blob, BatchedShootingProblem(desc, self.x0[None, :], self.T))
        return self._batched[1]

    def _arrays(self, xs, us):
        p = self._problem()
        xs = np.asarray(xs, float).reshape(1, self.T + 1, p.nx)
        us = np.asarray(us, float).reshape(1, self.T, p.nu)
        return p, xs, us

    def _fill(self, res, p, diff):
        xn, cost = res["xnext"][0].cpu().numpy(), res["cost"][0].cpu().numpy()
        rec = p.split_record(res["deriv"][0].cpu().numpy()) if diff else None
        for t, d in enumerate(self.runningDatas + [self.terminalData]):
            d.cost = float(cost[t])
            if t < self.T:
                d.xnext[:] = xn[t]
            if diff:
                for k in ("Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu"):
                    getattr(d, k)[...] = rec[k][t]
        return float(np.sum(cost))

    def calc(self, xs, us):
        p, xs, us = self._arrays(xs, us)
        self.terminalData.xnext[:] = xs[0, -1]
        return self._fill(p.calc(xs, us), p, False)

    def calcDiff(self, xs, us):
        p, xs, us = self._arrays(xs, us)
        return self._fill(p.calc_diff(xs, us), p, True)

    def rollout(self, us):
        """xs[t + 1] = runningModel.calc(xs[t], us[t]).xnext from x0 (sequential: one knot launch per step)."""
        p = self._problem()
        xs = [self.x0.copy()]
        for t in range(self.T):
            out = E.eval_knot(self._specs()[0], xs[-1], us[t], diff=False)
            xs.append(out["xnext"])
        return xs

    def quasiStatic(self, xs):
        xs = np.asarray(xs, float).reshape(self.T, -1)
        us = E.quasi_static(self._specs()[0], xs)
        return [u.copy() for u in us]


# ------------------------------------------------------------------------------------ callbacks
class CallbackLogger(object):
    """Filled from the device-side iteration log after the solve.  `xs`, `us`, `fs` hold the final
    trajectory (per-iteration trajectories are not kept on the device)."""

    def __init__(self):
        self.xs, self.us, self.fs = [], [], []
        self.steps, self.iters, self.costs = [], [], []
        self.u_regs, self.x_regs, self.stops, self.grads = [], [], [], []
        self.dVs, self.dVexps, self.ffeass = [], [], []

    def __call__(self, solver):
        self.xs, self.us, self.fs = list(solver.xs), list(solver.us), list(solver.fs)
        self.steps.append(solver.stepLength)
        self.iters.append(solver.iter)
        self.costs.append(solver.cost)
        self.u_regs.append(solver.u_reg)
        self.x_regs.append(solver.x_reg)
        self.stops.append(solver.stop)
        self.grads.append(float("nan"))
        self.dVs.append(solver.dV)
        self.dVexps.append(solver.dVexp)
        self.ffeass.append(solver.ffeas)


class CallbackVerbose(object):
    def __init__(self, level=0):
        self.level = level

    def __call__(self, solver):
        if solver.iter % 10 == 0:
            print("iter     cost         stop         grad         xreg         ureg       step    ||ffeas||")
        print("%4i  %.5e  %.5e  %.5e  %.5e  %.5e  %.4f  %.5e" % (solver.iter, solver.cost, solver.stop, -solver.dVexp,
                                                                   solver.x_reg, solver.u_reg, solver.stepLength,
                                                                   solver.ffeas))


# ------------------------------------------------------------------------------------ solvers
class _Solver(object):
    _kind = A.SOLVER_DDP
    _th_stop_default = 1e-9

    def __init__(self, problem):
        self.problem = problem
        self.th_stop = self._th_stop_default
        self._callbacks = []
        T = problem.T
        nx = problem.runningModels[0].state.nx
        nu = problem.runningModels[0].nu
        self.xs = [np.zeros(nx) for _ in range(T + 1)]
        self.us = [np.zeros(nu) for _ in range(T)]
        self.fs = [np.zeros(nx) for _ in range(T + 1)]
        self.K, self.k, self.Vx, self.Vxx, self.Qu, self.Quu = [], [], [], [], [], []
        self.cost = self.stop = self.dV = self.dVexp = self.ffeas = 0.0
        self.x_reg = self.u_reg = float("nan")
        self.stepLength, self.iter, self.isFeasible = 1.0, 0, False
        self._batched_solver = None

    def setCallbacks(self, callbacks):
        self._callbacks = list(callbacks)

    def getCallbacks(self):
        return self._callbacks

    def setCandidate(self, xs=None, us=None, isFeasible=False):
        T = self.problem.T
        self.xs = [self.problem.runningModels[0].state.zero() for _ in range(T + 1)] if xs is None or len(xs) == 0 \
            else [np.asarray(x, float).copy() for x in xs]
        self.us = [np.zeros(self.problem.runningModels[0].nu) for _ in range(T)] if us is None or len(us) == 0 \
            else [np.asarray(u, float).copy() for u in us]
        self.isFeasible = bool(isFeasible)

    def solve(self, init_xs=None, init_us=None, maxiter=100, isFeasible=False, regInit=None):
        """solver.solve(xs, us, maxiter, isFeasible, regInit) -> bool (examples/two_dof_sea.py:81)."""
        p = self.problem._problem()
        T = p.T
        init_xs = [] if init_xs is None else init_xs
        init_us = [] if init_us is None else init_us
        if len(init_xs) not in (0, T + 1):
            raise ValueError("xs has wrong dimension (it should be %d)" % (T + 1))
        if len(init_us) not in (0, T):
            raise ValueError("us has wrong dimension (it should be %d)" % T)
        xs = None if len(init_xs) == 0 else np.asarray(init_xs, float).reshape(1, T + 1, p.nx)
        us = None if len(init_us) == 0 else np.asarray(init_us, float).reshape(1, T, p.nu)
        bs = BatchedSolver(p, self._kind, maxiter=max(int(maxiter), 1), th_stop=self.th_stop)
        bs.solve(xs, us, maxiter=int(maxiter), is_feasible=isFeasible, reg_init=regInit)
        self._batched_solver = bs
        st = {k: v[0] for k, v in bs.states().items()}
        g = lambda k: bs.get(k)[0].cpu().numpy()
        self.xs, self.us, self.fs = list(g("xs")), list(g("us")), list(g("fs"))
        self.K, self.k = list(g("Kgain")), list(g("kff"))
        self.Vx, self.Vxx = list(g("Vx")), list(g("Vxx"))
        self.Qu, self.Quu = list(g("Qu")), list(g("Quu"))
        log = bs.logs()[0]
        # exactly the iterations whose callbacks crocoddyl runs: an iteration that ends at reg_max (in the
        # backward pass or after the line search) returns before the callbacks and has no log row
        n_it = min(int(st["n_logged"]), int(maxiter))
        for it in range(n_it):  # replay the per-iteration callbacks from the device log
            c, stop, xreg, step, dV, dVexp, ffeas, feas = log[it]
            self.cost, self.stop, self.x_reg, self.u_reg, self.stepLength = c, stop, xreg, xreg, step
            self.dV, self.dVexp, self.ffeas, self.isFeasible, self.iter = dV, dVexp, ffeas, bool(feas), it
            for cb in self._callbacks:
                cb(self)
        self.cost, self.stop, self.x_reg = float(st["cost"]), float(st["stop"]), float(st["xreg"])
        self.u_reg, self.stepLength = self.x_reg, float(st["steplength"])
        self.dV, self.dVexp, self.ffeas = float(st["dV"]), float(st["dVexp"]), float(st["ffeas"])
        self.isFeasible, self.iter = bool(st["is_feasible"]), int(st["iter"])
        return bool(st["status"] == A.OCP_CONVERGED)


class SolverDDP(_Solver):
    _kind = A.SOLVER_DDP


class SolverFDDP(_Solver):
    _kind = A.SOLVER_FDDP


class SolverBoxDDP(_Solver):
    _kind = A.SOLVER_BOXDDP
    _th_stop_default = 5e-5
