"""ORACLE / TEST INFRASTRUCTURE ONLY -- crocoddyl SolverDDP / SolverFDDP / SolverBoxDDP (+BoxQP)
restated in NumPy from the published algorithm of crocoddyl `devel` ~ v1.9.0+ (mid-2022), the
snapshot the reference was built against.  Call sites:
examples/double_pendulum.py:55,70; two_dof_sea.py:69,79-81; two_dof_vsa_boxddp.py:69,79-81.
PARITY UNPINNED: the reference holds no solver-level golden vectors and crocoddyl itself cannot
be built here; the only independent pins are the LQR known-answer tests in tests/.
"""
import numpy as np


class _SolverError(Exception):
    pass


def _bad(v):
    """crocoddyl::raiseIfNaN"""
    v = float(v)
    return np.isnan(v) or np.isinf(v) or v >= 1e30


def _llt(A):
    """Eigen::LLT semantics: fail on a non-positive pivot (NaN pivots pass through)."""
    n = A.shape[0]
    L = np.zeros((n, n))
    for j in range(n):
        x = A[j, j] - np.dot(L[j, :j], L[j, :j])
        if x <= 0.0:
            raise _SolverError("backward_error")
        L[j, j] = np.sqrt(x)
        for i in range(j + 1, n):
            L[i, j] = (A[i, j] - np.dot(L[i, :j], L[j, :j])) / L[j, j]
    return L


def _llt_solve(L, B):
    n = L.shape[0]
    Y = np.array(B, float)
    for i in range(n):
        Y[i] = (Y[i] - L[i, :i] @ Y[:i]) / L[i, i]
    for i in range(n - 1, -1, -1):
        Y[i] = (Y[i] - L[i + 1:, i] @ Y[i + 1:]) / L[i, i]
    return Y


class SolverAbstract(object):
    def __init__(self, problem):
        self.problem = problem
        T, ndx = problem.T, problem.terminalModel.state.ndx
        self.xs = [np.array(problem.terminalModel.state.zero(), float) for _ in range(T + 1)]
        self.us = [np.zeros(m.nu) for m in problem.runningModels]
        self.fs = [np.zeros(ndx) for _ in range(T + 1)]
        self.callbacks = []
        self.isFeasible = False
        self.wasFeasible = False
        self.cost = 0.0
        self.stop = 0.0
        self.d = np.zeros(2)
        self.x_reg = np.nan
        self.u_reg = np.nan
        self.stepLength = 1.0
        self.dV = 0.0
        self.dVexp = 0.0
        self.th_acceptstep = 0.1
        self.th_stop = 1e-9
        self.th_gaptol = 1e-16
        self.iter = 0
        self.ffeas = 0.0
        self.inffeas = True  # LInf feasibility norm

    def setCallbacks(self, cbs):
        self.callbacks = list(cbs)

    def getCallbacks(self):
        return self.callbacks

    def setCandidate(self, xs=None, us=None, isFeasible=False):
        T = self.problem.T
        if xs is None or len(xs) == 0:
            self.xs = [np.array(self.problem.terminalModel.state.zero(), float) for _ in range(T + 1)]
        else:
            if len(xs) != T + 1:
                raise ValueError("Warm start state has wrong dimension, got %d expecting %d" % (len(xs), T + 1))
            self.xs = [np.array(x, float) for x in xs]
        if us is None or len(us) == 0:
            self.us = [np.zeros(m.nu) for m in self.problem.runningModels]
        else:
            if len(us) != T:
                raise ValueError("Warm start control has wrong dimension, got %d expecting %d" % (len(us), T))
            self.us = [np.array(u, float) for u in us]
        self.isFeasible = isFeasible

    def computeDynamicFeasibility(self):
        p = self.problem
        if not self.isFeasible:
            st = p.terminalModel.state
            self.fs[0] = np.asarray(st.diff(self.xs[0], p.x0)).ravel()
            for t in range(p.T):
                self.fs[t + 1] = np.asarray(st.diff(self.xs[t + 1], p.runningDatas[t].xnext)).ravel()
            if self.inffeas:
                self.ffeas = max(np.max(np.abs(f)) for f in self.fs)
            else:
                self.ffeas = sum(np.sum(np.abs(f)) for f in self.fs)
        elif not self.wasFeasible:
            for f in self.fs:
                f[:] = 0.0
            self.ffeas = 0.0
        return self.ffeas


class SolverDDP(SolverAbstract):
    def __init__(self, problem):
        SolverAbstract.__init__(self, problem)
        self.reg_incfactor = 10.0
        self.reg_decfactor = 10.0
        self.reg_min = 1e-9
        self.reg_max = 1e9
        self.cost_try = 0.0
        self.th_grad = 1e-12
        self.th_stepdec = 0.5
        self.th_stepinc = 0.01
        self.alphas = [1.0 / 2.0 ** n for n in range(10)]
        if self.th_stepinc < self.alphas[-1]:
            self.th_stepinc = self.alphas[-1]
        T = problem.T
        ndx = problem.terminalModel.state.ndx
        self.xs_try = [np.zeros_like(self.xs[0]) for _ in range(T + 1)]
        self.xs_try[0] = problem.x0.copy()
        self.us_try = [np.zeros(m.nu) for m in problem.runningModels]
        self.Vxx = [np.zeros((ndx, ndx)) for _ in range(T + 1)]
        self.Vx = [np.zeros(ndx) for _ in range(T + 1)]
        self.Qxx = [np.zeros((ndx, ndx)) for _ in range(T)]
        self.Qx = [np.zeros(ndx) for _ in range(T)]
        self.Qxu = [np.zeros((ndx, m.nu)) for m in problem.runningModels]
        self.Quu = [np.zeros((m.nu, m.nu)) for m in problem.runningModels]
        self.Qu = [np.zeros(m.nu) for m in problem.runningModels]
        self.K = [np.zeros((m.nu, ndx)) for m in problem.runningModels]
        self.k = [np.zeros(m.nu) for m in problem.runningModels]
        self.Quuk = [np.zeros(m.nu) for m in problem.runningModels]

    # ------------------------------------------------------------------ main loop (Appendix A.1)
    def _accept(self, a):
        return self.dVexp >= 0 and (self.d[0] < self.th_grad or not self.isFeasible
                                    or self.dV > self.th_acceptstep * self.dVexp)

    def _feasible_after_accept(self, a):
        return True

    def _pre_linesearch(self):
        self.expectedImprovement()

    def _post_trystep(self):
        pass

    def solve(self, init_xs=None, init_us=None, maxiter=100, isFeasible=False, regInit=None):
        self.xs_try[0] = self.problem.x0.copy()
        self.setCandidate(init_xs, init_us, isFeasible)
        if regInit is None or np.isnan(regInit):
            self.x_reg = self.u_reg = self.reg_min
        else:
            self.x_reg = self.u_reg = float(regInit)
        self.wasFeasible = False
        recalc = True
        for self.iter in range(maxiter):
            while True:
                try:
                    self.computeDirection(recalc)
                except _SolverError:
                    recalc = False
                    self.increaseRegularization()
                    if self.x_reg == self.reg_max:
                        return False
                    continue
                break
            self._pre_linesearch()
            recalc = False
            for a in self.alphas:
                self.stepLength = a
                try:
                    self.dV = self.tryStep(a)
                except _SolverError:
                    continue
                self._post_trystep()
                self.dVexp = a * (self.d[0] + 0.5 * a * self.d[1])
                if self._accept(a):
                    self.wasFeasible = self.isFeasible
                    self.setCandidate(self.xs_try, self.us_try, self._feasible_after_accept(a))
                    self.cost = self.cost_try
                    recalc = True
                    break
            if self.stepLength > self.th_stepdec:
                self.decreaseRegularization()
            if self.stepLength <= self.th_stepinc:
                self.increaseRegularization()
                if self.x_reg == self.reg_max:
                    return False
            self.stoppingCriteria()
            for cb in self.callbacks:
                cb(self)
            if self.wasFeasible and self.stop < self.th_stop:
                return True
        self.iter = maxiter
        return False

    def computeDirection(self, recalc=True):
        if recalc:
            self.calcDiff()
        self.backwardPass()

    def tryStep(self, a=1.0):
        self.forwardPass(a)
        return self.cost - self.cost_try

    def stoppingCriteria(self):
        self.stop = 0.0
        for q in self.Qu:
            self.stop += float(q @ q)
        return self.stop

    def expectedImprovement(self):
        self.d[:] = 0.0
        for t in range(self.problem.T):
            self.d[0] += float(self.Qu[t] @ self.k[t])
            self.d[1] -= float(self.k[t] @ self.Quuk[t])
        return self.d

    def calcDiff(self):
        if self.iter == 0:
            self.problem.calc(self.xs, self.us)
        self.cost = self.problem.calcDiff(self.xs, self.us)
        self.ffeas = self.computeDynamicFeasibility()
        return self.cost

    # ---------------------------------------------------------------- backward pass (Appendix A.3)
    def backwardPass(self):
        p = self.problem
        dT = p.terminalData
        self.Vxx[-1] = np.array(dT.Lxx, float)
        self.Vx[-1] = np.array(dT.Lx, float)
        ndx = self.Vx[-1].shape[0]
        if not np.isnan(self.x_reg):
            self.Vxx[-1][range(ndx), range(ndx)] += self.x_reg
        if not self.isFeasible:
            self.Vx[-1] = self.Vx[-1] + self.Vxx[-1] @ self.fs[-1]
        for t in range(p.T - 1, -1, -1):
            d = p.runningDatas[t]
            Vxx_p, Vx_p = self.Vxx[t + 1], self.Vx[t + 1]
            Fx, Fu = np.asarray(d.Fx), np.asarray(d.Fu)
            FxTVxx = Fx.T @ Vxx_p
            self.Qxx[t] = np.asarray(d.Lxx) + FxTVxx @ Fx
            self.Qx[t] = np.asarray(d.Lx) + Fx.T @ Vx_p
            self.Qxu[t] = np.asarray(d.Lxu) + FxTVxx @ Fu
            self.Quu[t] = np.asarray(d.Luu) + (Fu.T @ Vxx_p) @ Fu
            self.Qu[t] = np.asarray(d.Lu) + Fu.T @ Vx_p
            nu = self.Qu[t].shape[0]
            if not np.isnan(self.u_reg):
                self.Quu[t][range(nu), range(nu)] += self.u_reg
            self.computeGains(t)
            self.Quuk[t] = self.Quu[t] @ self.k[t]
            Vx = self.Qx[t] + self.K[t].T @ self.Quuk[t] - 2.0 * (self.K[t].T @ self.Qu[t])
            Vxx = self.Qxx[t] - self.Qxu[t] @ self.K[t]
            Vxx = 0.5 * (Vxx + Vxx.T)
            if not np.isnan(self.x_reg):
                Vxx[range(ndx), range(ndx)] += self.x_reg
            if not self.isFeasible:
                Vx = Vx + Vxx @ self.fs[t]
            self.Vx[t], self.Vxx[t] = Vx, Vxx
            if _bad(np.max(np.abs(Vx))) or _bad(np.max(np.abs(Vxx))):
                raise _SolverError("backward_error")

    def computeGains(self, t):
        L = _llt(self.Quu[t])
        self.K[t] = _llt_solve(L, self.Qxu[t].T)
        self.k[t] = _llt_solve(L, self.Qu[t])

    # ----------------------------------------------------------------- forward pass (Appendix A.4)
    def _clamp(self, m, u):
        return u

    def forwardPass(self, a):
        p = self.problem
        self.cost_try = 0.0
        self.xs_try[0] = p.x0.copy()
        for t in range(p.T):
            m, d = p.runningModels[t], p.runningDatas[t]
            dx = np.asarray(m.state.diff(self.xs[t], self.xs_try[t])).ravel()
            self.us_try[t] = self._clamp(m, self.us[t] - self.k[t] * a - self.K[t] @ dx)
            m.calc(d, self.xs_try[t], self.us_try[t])
            self.xs_try[t + 1] = np.array(d.xnext, float)
            self.cost_try += d.cost
            if _bad(self.cost_try) or _bad(np.max(np.abs(self.xs_try[t + 1]))):
                raise _SolverError("forward_error")
        p.terminalModel.calc(p.terminalData, self.xs_try[-1])
        self.cost_try += p.terminalData.cost
        if _bad(self.cost_try):
            raise _SolverError("forward_error")

    def increaseRegularization(self):
        self.x_reg *= self.reg_incfactor
        if self.x_reg > self.reg_max:
            self.x_reg = self.reg_max
        self.u_reg = self.x_reg

    def decreaseRegularization(self):
        self.x_reg /= self.reg_decfactor
        if self.x_reg < self.reg_min:
            self.x_reg = self.reg_min
        self.u_reg = self.x_reg


class SolverFDDP(SolverDDP):
    def __init__(self, problem):
        SolverDDP.__init__(self, problem)
        self.dg = self.dq = self.dv = 0.0
        self.th_acceptnegstep = 2.0

    def _accept(self, a):
        if self.dVexp >= 0:
            return self.d[0] < self.th_grad or self.dV > self.th_acceptstep * self.dVexp
        return self.dV > self.th_acceptnegstep * self.dVexp

    def _feasible_after_accept(self, a):
        return self.wasFeasible or a == 1.0

    def _pre_linesearch(self):
        self.updateExpectedImprovement()

    def _post_trystep(self):
        self.expectedImprovement()

    def updateExpectedImprovement(self):
        self.dg = self.dq = 0.0
        T = self.problem.T
        if not self.isFeasible:
            self.dg -= float(self.Vx[-1] @ self.fs[-1])
            self.dq += float(self.fs[-1] @ (self.Vxx[-1] @ self.fs[-1]))
        for t in range(T):
            self.dg += float(self.Qu[t] @ self.k[t])
            self.dq -= float(self.k[t] @ self.Quuk[t])
            if not self.isFeasible:
                self.dg -= float(self.Vx[t] @ self.fs[t])
                self.dq += float(self.fs[t] @ (self.Vxx[t] @ self.fs[t]))

    def expectedImprovement(self):
        self.dv = 0.0
        if not self.isFeasible:
            st = self.problem.terminalModel.state
            dx = np.asarray(st.diff(self.xs_try[-1], self.xs[-1])).ravel()
            self.dv -= float(self.fs[-1] @ (self.Vxx[-1] @ dx))
            for t in range(self.problem.T):
                dx = np.asarray(st.diff(self.xs_try[t], self.xs[t])).ravel()
                self.dv -= float(self.fs[t] @ (self.Vxx[t] @ dx))
        self.d[0] = self.dg + self.dv
        self.d[1] = self.dq - 2.0 * self.dv
        return self.d

    def forwardPass(self, a):
        p = self.problem
        self.cost_try = 0.0
        xnext = p.x0.copy()
        gaps = not (self.isFeasible or a == 1.0)
        for t in range(p.T):
            m, d = p.runningModels[t], p.runningDatas[t]
            if gaps:
                self.xs_try[t] = np.asarray(m.state.integrate(xnext, self.fs[t] * (a - 1.0))).ravel()
            else:
                self.xs_try[t] = xnext
            dx = np.asarray(m.state.diff(self.xs[t], self.xs_try[t])).ravel()
            self.us_try[t] = self.us[t] - self.k[t] * a - self.K[t] @ dx
            m.calc(d, self.xs_try[t], self.us_try[t])
            xnext = np.array(d.xnext, float)
            self.cost_try += d.cost
            if _bad(self.cost_try) or _bad(np.max(np.abs(xnext))):
                raise _SolverError("forward_error")
        m = p.terminalModel
        if gaps:
            self.xs_try[-1] = np.asarray(m.state.integrate(xnext, self.fs[-1] * (a - 1.0))).ravel()
        else:
            self.xs_try[-1] = xnext
        m.calc(p.terminalData, self.xs_try[-1])
        self.cost_try += p.terminalData.cost
        if _bad(self.cost_try):
            raise _SolverError("forward_error")


class BoxQPSolution(object):
    def __init__(self):
        self.Hff_inv = np.zeros((0, 0))
        self.x = np.zeros(0)
        self.free_idx = []
        self.clamped_idx = []


class BoxQP(object):
    """Projected-Newton box QP (Tassa et al. 2014) as in crocoddyl/core/solvers/box-qp.cpp
    (SURVEY.md Appendix A.4).  min 0.5 x'Hx + q'x  s.t. lb <= x <= ub."""

    def __init__(self, nx, maxiter=100, th_acceptstep=0.1, th_grad=1e-9, reg=1e-9):
        self.nx, self.maxiter, self.th_acceptstep, self.th_grad, self.reg = nx, maxiter, th_acceptstep, th_grad, reg
        self.alphas = [1.0 / 2.0 ** n for n in range(10)]
        self.solution = BoxQPSolution()
        self._fact_set = None  # free set of the factorisation currently held in solution.Hff_inv

    def _factor(self, H, free):
        Hff = H[np.ix_(free, free)].copy()
        if self.reg != 0.0:
            Hff[range(len(free)), range(len(free))] += self.reg
        L = _llt(Hff)
        self.solution.Hff_inv = _llt_solve(L, np.eye(len(free))) if free else np.zeros((0, 0))
        self._fact_set = list(free)
        return L

    def solve(self, H, q, lb, ub, xinit):
        sol = self.solution
        x = np.maximum(np.minimum(np.asarray(xinit, float), ub), lb)
        self._fact_set = None
        for k in range(self.maxiter):
            g = q + H @ x
            sol.clamped_idx, sol.free_idx = [], []
            for j in range(self.nx):
                if (x[j] == lb[j] and g[j] > 0.0) or (x[j] == ub[j] and g[j] < 0.0):
                    sol.clamped_idx.append(j)
                else:
                    sol.free_idx.append(j)
            free, clamped = sol.free_idx, sol.clamped_idx
            nf = len(free)
            gnorm = np.max(np.abs(g[free])) if nf else 0.0
            if gnorm <= self.th_grad or nf == 0:
                # upstream refactors only when k == 0 and otherwise returns the inverse of the
                # last Newton step; when that step's free set differs from the final one the
                # upstream indexing is ill-defined, so the inverse is recomputed.
                if k == 0 or self._fact_set != free:
                    self._factor(H, free)
                sol.x = x
                return sol
            L = self._factor(H, free)
            dxf = -q[free]
            if clamped:
                dxf = dxf - H[np.ix_(free, clamped)] @ x[clamped]
            dxf = _llt_solve(L, dxf) - x[free]
            dx = np.zeros(self.nx)
            dx[free] = dxf
            fold = 0.5 * float(x @ (H @ x)) + float(q @ x)
            for a in self.alphas:
                xnew = np.maximum(np.minimum(x + a * dx, ub), lb)
                fnew = 0.5 * float(xnew @ (H @ xnew)) + float(q @ xnew)
                if fold - fnew >= self.th_acceptstep * float(g @ (x - xnew)):
                    x = xnew
                    break
        sol.x = x
        return sol


class SolverBoxDDP(SolverDDP):
    def __init__(self, problem):
        SolverDDP.__init__(self, problem)
        self.qp = BoxQP(problem.runningModels[0].nu, 100, 0.1, 1e-5, 0.0)
        self.Quu_inv = [np.zeros((m.nu, m.nu)) for m in problem.runningModels]
        self.th_stop = 5e-5

    def computeGains(self, t):
        m = self.problem.runningModels[t]
        if not m.has_control_limits or not self.isFeasible:
            SolverDDP.computeGains(self, t)
            return
        du_lb = m.u_lb - self.us[t]
        du_ub = m.u_ub - self.us[t]
        sol = self.qp.solve(self.Quu[t], self.Qu[t], du_lb, du_ub, self.k[t])
        Qi = np.zeros_like(self.Quu[t])
        for i, fi in enumerate(sol.free_idx):
            for j, fj in enumerate(sol.free_idx):
                Qi[fi, fj] = sol.Hff_inv[i, j]
        self.Quu_inv[t] = Qi
        self.K[t] = Qi @ self.Qxu[t].T
        self.k[t] = -sol.x
        for i in sol.clamped_idx:
            self.Qu[t][i] = 0.0

    def _clamp(self, m, u):
        if m.has_control_limits:
            return np.minimum(np.maximum(u, m.u_lb), m.u_ub)
        return u
