"""Independent pins of the solver semantics (CPU tier).  The crocoddyl solvers are not in /root/reference and
cannot be built here, so the oracle's solver shim is a restatement; these tests check it --
and, through the g++ build of the device code, the product's box QP -- against things that do NOT come from that
restatement:
  * the box QP against brute-force enumeration of the active sets (KKT conditions) on random SPD problems,
    including degenerate free sets, start points on the bounds and collapsed boxes;
  * a box-constrained solve whose bounds never bind against the unconstrained DDP solve, bit for bit;
  * the backward pass + FDDP full step on an infeasible linear-quadratic problem against one dense KKT solve of
    the whole equality-constrained QP;
  * two scalar problems whose iterates are derived by hand in the test body: one that needs an alpha < 1 step
    (Newton on sqrt(1 + u^2) overshoots) and one that needs nine regularisation increases (negative curvature)."""
import ctypes as C
import itertools

import numpy as np
import pytest

import oracle_build as ob
from util import rel_err

croc = ob.crocoddyl
from crocoddyl import solvers as S  # noqa: E402  (the NumPy shim)


# ------------------------------------------------------------------------------------------ box QP
def brute_force_boxqp(H, q, lb, ub):
    """Unique minimiser of 0.5 x'Hx + q'x over the box (H SPD) by enumerating which bound every variable sits on:
    the candidate that is feasible and satisfies the sign conditions of the multipliers is the solution."""
    n = len(q)
    best = None
    for pattern in itertools.product((-1, 0, 1), repeat=n):
        x = np.zeros(n)
        free = [i for i in range(n) if pattern[i] == 0]
        for i in range(n):
            if pattern[i] == -1:
                x[i] = lb[i]
            elif pattern[i] == 1:
                x[i] = ub[i]
        if free:
            fixed = [i for i in range(n) if pattern[i] != 0]
            rhs = -q[free] - (H[np.ix_(free, fixed)] @ x[fixed] if fixed else 0.0)
            x[free] = np.linalg.solve(H[np.ix_(free, free)], rhs)
        if np.any(x < lb - 1e-12) or np.any(x > ub + 1e-12):
            continue
        g = H @ x + q
        ok = all((pattern[i] == 0 and abs(g[i]) < 1e-9) or (pattern[i] == -1 and g[i] >= -1e-12) or
                 (pattern[i] == 1 and g[i] <= 1e-12) for i in range(n))
        if ok:
            f = 0.5 * x @ H @ x + q @ x
            if best is None or f < best[0]:
                best = (f, x)
    assert best is not None
    return best[1]


def random_boxqps(count, seed):
    rng = np.random.default_rng(seed)
    out = []
    for k in range(count):
        n = int(rng.integers(1, 5))
        A = rng.standard_normal((n, n))
        H = A @ A.T + (0.05 + rng.uniform(0, 1)) * np.eye(n)
        q = rng.standard_normal(n) * rng.choice([0.1, 1.0, 10.0])
        lb = -rng.uniform(0.05, 2.0, n)
        ub = rng.uniform(0.05, 2.0, n)
        xinit = rng.uniform(-3, 3, n)
        mode = k % 6
        if mode == 1:  # start on the bounds
            xinit = np.where(rng.uniform(size=n) < 0.5, lb, ub)
        elif mode == 2:  # unconstrained minimiser strictly inside: nothing clamps
            xs = rng.uniform(lb * 0.5, ub * 0.5)
            q = -H @ xs
        elif mode == 3:  # degenerate: the minimiser sits exactly on a bound with zero multiplier there
            xs = rng.uniform(lb * 0.5, ub * 0.5)
            j = int(rng.integers(0, n))
            xs[j] = ub[j]
            q = -H @ xs
        elif mode == 4:  # collapsed box in one coordinate
            j = int(rng.integers(0, n))
            lb[j] = ub[j] = rng.uniform(-0.5, 0.5)
        elif mode == 5:  # everything clamps
            q = np.sign(rng.standard_normal(n)) * 1e3
        out.append((H, q, lb, ub, xinit))
    return out


@pytest.mark.parametrize("seed", [11, 12])
def test_boxqp_shim_and_device_code_against_brute_force_kkt(hostmath, seed):
    """>= 100 random SPD box QPs per seed, nu = 1..4: the oracle's BoxQP (restated crocoddyl) and the product's
    boxqp_solve (g++ build of backward_dev.cuh) both reach the KKT point found by enumeration, and the returned
    free-block inverse is the inverse of H on the FINAL free set (the case of the documented deviation from
    upstream)."""
    hostmath.hh_boxqp.argtypes = [C.c_int] + [C.c_void_p] * 8
    for H, q, lb, ub, xinit in random_boxqps(120, seed):
        n = len(q)
        ref = brute_force_boxqp(H, q, lb, ub)
        qp = S.BoxQP(n, 100, 0.1, 1e-5, 0.0)  # SolverBoxDDP's constructor arguments
        sol = qp.solve(H, q, lb, ub, xinit)
        # th_grad = 1e-5 on the free gradient: |x - x*| <= 1e-5 / lambda_min(H)
        tol = 2e-5 / np.linalg.eigvalsh(H).min()
        assert np.abs(sol.x - ref).max() <= tol, (sol.x, ref)
        assert np.all(sol.x >= lb) and np.all(sol.x <= ub)
        g = H @ sol.x + q
        for j in sol.clamped_idx:
            assert (sol.x[j] == lb[j] and g[j] > 0) or (sol.x[j] == ub[j] and g[j] < 0)
        if sol.free_idx:
            assert np.abs(g[sol.free_idx]).max() <= 1e-5
            Hff = H[np.ix_(sol.free_idx, sol.free_idx)]
            assert rel_err(sol.Hff_inv @ Hff, np.eye(len(sol.free_idx))) < 1e-9
        # the product's device code (host build): same point, same free set, same inverse
        Hc, qc, lc, uc, xc = (np.ascontiguousarray(a, float) for a in (H, q, lb, ub, xinit))
        x, Hinv, cl = np.zeros(n), np.zeros((n, n)), np.zeros(n, np.int32)
        rc = hostmath.hh_boxqp(n, Hc.ctypes.data, qc.ctypes.data, lc.ctypes.data, uc.ctypes.data, xc.ctypes.data,
                               x.ctypes.data, Hinv.ctypes.data, cl.ctypes.data)
        assert rc == 0
        assert np.abs(x - sol.x).max() <= 1e-12 * max(1.0, np.abs(sol.x).max())
        # a variable that sits on a bound with a multiplier that is zero up to rounding (the degenerate problems
        # above) is "clamped" or "free" by the sign of that rounding noise: both are KKT-valid, so such indices are
        # left out of the set comparison
        ambiguous = set(j for j in range(n) if min(abs(sol.x[j] - lb[j]), abs(sol.x[j] - ub[j])) < 1e-12 and
                        abs(g[j]) < 1e-9)
        ours = set(np.nonzero(cl)[0].tolist())
        assert ours - ambiguous == set(sol.clamped_idx) - ambiguous
        free = [j for j in range(n) if j not in ours]
        if free:  # Hinv is the inverse of H on the final free set of THIS solve, zero elsewhere
            assert rel_err(Hinv[np.ix_(free, free)] @ H[np.ix_(free, free)], np.eye(len(free))) < 1e-9
        mask = np.ones((n, n), bool)
        mask[np.ix_(free, free)] = False
        assert np.all(Hinv[mask] == 0.0)
        if ours == set(sol.clamped_idx):
            full = np.zeros((n, n))
            for a, fa in enumerate(sol.free_idx):
                for b, fb in enumerate(sol.free_idx):
                    full[fa, fb] = sol.Hff_inv[a, b]
            assert rel_err(Hinv, full) < 1e-10


# ------------------------------------------------------------------- LQ problems (known structure)
class _VecState(croc.StateAbstract):
    def __init__(self, n):
        croc.StateAbstract.__init__(self, n, n)

    def zero(self):
        return np.zeros(self.nx)

    def diff(self, a, b):
        return np.asarray(b, float) - np.asarray(a, float)

    def integrate(self, x, dx):
        return np.asarray(x, float) + np.asarray(dx, float)


class _LQ(croc.ActionModelAbstract):
    """x' = A x + B u + c,  cost = 0.5 [x;u]' [[Q, N],[N', R]] [x;u] + q'x + r'u  (convex)."""

    def __init__(self, state, A_, B_, c, Q, N, R, q, r, u_lb=None, u_ub=None):
        croc.ActionModelAbstract.__init__(self, state, B_.shape[1], 0)
        self.A, self.Bm, self.c, self.Q, self.N, self.R, self.q, self.r = A_, B_, c, Q, N, R, q, r
        if u_lb is not None:
            self.u_lb, self.u_ub = np.asarray(u_lb, float), np.asarray(u_ub, float)

    def calc(self, data, x, u=None):
        u = np.zeros(self.nu) if u is None else u
        data.xnext = self.A @ x + self.Bm @ u + self.c
        data.cost = 0.5 * x @ self.Q @ x + 0.5 * u @ self.R @ u + x @ self.N @ u + self.q @ x + self.r @ u

    def calcDiff(self, data, x, u=None):
        u = np.zeros(self.nu) if u is None else u
        data.Fx, data.Fu = self.A, self.Bm
        data.Lx, data.Lu = self.Q @ x + self.N @ u + self.q, self.R @ u + self.N.T @ x + self.r
        data.Lxx, data.Luu, data.Lxu = self.Q, self.R, self.N


def _random_lq(rng, n, m, T, time_varying=True, bounds=None):
    st = _VecState(n)
    models = []
    for t in range(T + 1):
        if t == 0 or time_varying:
            A_ = np.eye(n) + 0.2 * rng.standard_normal((n, n))
            B_ = rng.standard_normal((n, m))
            W_ = rng.standard_normal((n + m, n + m))
            W_ = W_ @ W_.T + 0.5 * np.eye(n + m)
            args = (A_, B_, 0.1 * rng.standard_normal(n), W_[:n, :n], W_[:n, n:], W_[n:, n:], rng.standard_normal(n),
                    rng.standard_normal(m))
        models.append(_LQ(st, *args, **(bounds or {})))
    return models[:T], models[T]


def test_backward_pass_and_fddp_step_against_dense_kkt():
    """One FDDP iteration from an INFEASIBLE candidate on a time-varying convex LQ problem.  The Newton step of the
    whole problem is the solution of one dense KKT system (stationarity + linearised dynamics with the gaps as
    right-hand side); the backward pass's (K, k, Vx) and the alpha = 1 forward pass must reproduce it:
    du_t = -k_t - K_t dx_t, dx_{t+1} = Fx dx_t + Fu du_t + fs_{t+1}, Vx[0] = gradient of the optimal cost-to-go."""
    rng = np.random.default_rng(21)
    n, m, T = 3, 2, 7
    run, term = _random_lq(rng, n, m, T)
    x0 = rng.standard_normal(n)
    problem = croc.ShootingProblem(x0, run, term)
    solver = croc.SolverFDDP(problem)
    xs = [rng.standard_normal(n) for _ in range(T + 1)]  # infeasible: gaps everywhere, xs[0] != x0
    us = [rng.standard_normal(m) for _ in range(T)]
    solver.setCandidate(xs, us, False)
    solver.x_reg = solver.u_reg = 0.0  # the exact Newton step (the solver's 1e-9 would show at the 1e-9 level)
    solver.iter = 0
    solver.calcDiff()
    solver.backwardPass()
    # ---- dense KKT system in z = [dx_0 .. dx_T, du_0 .. du_{T-1}, lambda_0 .. lambda_T]
    nz = (T + 1) * n + T * m
    nc = (T + 1) * n
    Hm, gv = np.zeros((nz, nz)), np.zeros(nz)
    Cm, dv = np.zeros((nc, nz)), np.zeros(nc)
    ix = lambda t: slice(t * n, (t + 1) * n)
    iu = lambda t: slice((T + 1) * n + t * m, (T + 1) * n + (t + 1) * m)
    for t in range(T):
        d = problem.runningDatas[t]
        Hm[ix(t), ix(t)] += d.Lxx
        Hm[iu(t), iu(t)] += d.Luu
        Hm[ix(t), iu(t)] += d.Lxu
        Hm[iu(t), ix(t)] += d.Lxu.T
        gv[ix(t)] += d.Lx
        gv[iu(t)] += d.Lu
        # dx_{t+1} - Fx dx_t - Fu du_t = fs_{t+1}
        Cm[ix(t + 1), ix(t + 1)] += np.eye(n)
        Cm[ix(t + 1), ix(t)] -= d.Fx
        Cm[ix(t + 1), iu(t)] -= d.Fu
        dv[ix(t + 1)] = solver.fs[t + 1]
    Hm[ix(T), ix(T)] += problem.terminalData.Lxx
    gv[ix(T)] += problem.terminalData.Lx
    Cm[ix(0), ix(0)] = np.eye(n)
    dv[ix(0)] = solver.fs[0]
    KKT = np.block([[Hm, Cm.T], [Cm, np.zeros((nc, nc))]])
    sol = np.linalg.solve(KKT, np.concatenate([-gv, dv]))
    for t in range(T):
        dx, du = sol[ix(t)], sol[iu(t)]
        assert rel_err(-solver.k[t] - solver.K[t] @ dx, du) < 1e-10, t
    # the multiplier of the initial-state constraint is minus the gradient of the optimal cost in dx_0; Vx[0]
    # already contains Vxx[0] fs[0] (the gap is folded in), i.e. it is that gradient evaluated at dx_0 = fs[0]
    lam0 = sol[nz:nz + n]
    assert rel_err(solver.Vx[0], -lam0) < 1e-9
    # alpha = 1 forward pass closes every gap and lands on xs + dx (linear dynamics: exactly the KKT point)
    solver.updateExpectedImprovement()
    solver.forwardPass(1.0)
    for t in range(T + 1):
        assert rel_err(solver.xs_try[t], xs[t] + sol[ix(t)]) < 1e-10, t
    for t in range(T):
        assert rel_err(solver.us_try[t], us[t] + sol[iu(t)]) < 1e-10, t
    # and the model of the improvement is exact on an LQ problem: dV == dVexp at alpha = 1
    dV = solver.cost - solver.cost_try
    d = solver.expectedImprovement()
    assert abs(dV - (d[0] + 0.5 * d[1])) < 1e-9 * max(1.0, abs(dV))


def test_boxddp_with_inactive_bounds_reproduces_ddp_bit_for_bit():
    """SolverBoxDDP on a problem whose bounds never bind: the box QP returns the unconstrained Newton point, so the
    whole trace (costs, steps, regularisation) must equal plain SolverDDP's, and the solutions agree to rounding
    (Quu^-1 Qxu' through an explicit inverse instead of two triangular solves)."""
    rng = np.random.default_rng(22)
    n, m, T = 3, 2, 9
    x0 = rng.standard_normal(n)
    traces = {}
    for name in ("ddp", "boxddp"):
        rng2 = np.random.default_rng(23)
        bounds = dict(u_lb=-1e6 * np.ones(m), u_ub=1e6 * np.ones(m)) if name == "boxddp" else None
        run, term = _random_lq(rng2, n, m, T, bounds=bounds)
        problem = croc.ShootingProblem(x0, run, term)
        solver = (croc.SolverBoxDDP if name == "boxddp" else croc.SolverDDP)(problem)
        solver.th_stop = 1e-12
        log = croc.CallbackLogger()
        solver.setCallbacks([log])
        solver.solve([], [], 20)
        traces[name] = (solver, log)
    (sd, ld), (sb, lb) = traces["ddp"], traces["boxddp"]
    assert sd.iter == sb.iter and ld.steps == lb.steps and ld.x_regs == lb.x_regs
    assert rel_err(lb.costs, ld.costs) < 1e-12
    assert rel_err(np.array(sb.xs), np.array(sd.xs)) < 1e-10 and rel_err(np.array(sb.us), np.array(sd.us)) < 1e-10
    assert rel_err(np.array(sb.K), np.array(sd.K)) < 1e-9


# ----------------------------------------------------------------------- scalar problems, by hand
class _Scalar(croc.ActionModelAbstract):
    """x' = x + u, cost = l(u) with user-supplied l, l', l''."""

    def __init__(self, l, dl, ddl):
        croc.ActionModelAbstract.__init__(self, _VecState(1), 1, 0)
        self.l, self.dl, self.ddl = l, dl, ddl

    def calc(self, data, x, u=None):
        u = np.zeros(1) if u is None else u
        data.xnext = x + u
        data.cost = float(self.l(u[0]))

    def calcDiff(self, data, x, u=None):
        u = np.zeros(1) if u is None else u
        data.Fx, data.Fu = np.eye(1), np.eye(1)
        data.Lx, data.Lu = np.zeros(1), np.array([self.dl(u[0])])
        data.Lxx, data.Luu, data.Lxu = np.zeros((1, 1)), np.array([[self.ddl(u[0])]]), np.zeros((1, 1))


class _Zero(croc.ActionModelAbstract):
    def __init__(self):
        croc.ActionModelAbstract.__init__(self, _VecState(1), 1, 0)

    def calc(self, data, x, u=None):
        data.xnext, data.cost = x, 0.0

    def calcDiff(self, data, x, u=None):
        data.Fx, data.Fu, data.Lx, data.Lu = np.eye(1), np.zeros((1, 1)), np.zeros(1), np.zeros(1)
        data.Lxx, data.Luu, data.Lxu = np.zeros((1, 1)), np.zeros((1, 1)), np.zeros((1, 1))


@pytest.mark.parametrize("cls", [croc.SolverDDP, croc.SolverFDDP])
def test_line_search_takes_the_hand_derived_partial_step(cls):
    """l(u) = sqrt(1 + u^2): Newton's step from u is -u (1 + u^2), so from u = 2 the full step lands on -8 and
    alpha = 1/2 on -3 (both increase the cost: rejected) and alpha = 1/4 on -0.5, whose decrease 2.236 - 1.118
    exceeds th_acceptstep x the model's 1.956: accepted.  After that |u| < 1 and every full step is taken:
    u -> -u^3 up to the 2e-9 of regularisation."""
    l = lambda u: np.sqrt(1 + u * u)
    dl = lambda u: u / np.sqrt(1 + u * u)
    ddl = lambda u: (1 + u * u) ** -1.5
    problem = croc.ShootingProblem(np.zeros(1), [_Scalar(l, dl, ddl)], _Zero())
    solver = cls(problem)
    log = croc.CallbackLogger()
    solver.setCallbacks([log])
    assert solver.solve([np.zeros(1), np.array([2.0])], [np.array([2.0])], 20, True)
    # by hand: Quu = l''(2) + Vxx'(= x_reg) + u_reg, k = l'(2) / Quu
    k0 = dl(2.0) / (ddl(2.0) + 2e-9)
    u1 = 2.0 - 0.25 * k0
    assert abs(u1 - (-0.5)) < 1e-6
    assert log.steps[0] == 0.25 and log.steps[1:] == [1.0] * (len(log.steps) - 1)
    assert abs(log.costs[0] - l(u1)) < 1e-15
    # regularisation: 0.25 is neither > th_stepdec (0.5) nor <= th_stepinc (0.01): unchanged; then it stays at reg_min
    assert log.x_regs == [1e-9] * len(log.x_regs)
    u2 = u1 - dl(u1) / (ddl(u1) + 2e-9)
    assert abs(u2 - (-u1 ** 3)) < 1e-8 and abs(log.costs[1] - l(u2)) < 1e-15
    # u: 2 -> -0.5 -> 0.125 -> -0.00195 -> 7.5e-9: the stop criterion Qu^2 < 1e-9 first holds in the backward pass
    # at u = 7.5e-9, i.e. in the fifth iteration (iter = 4)
    assert solver.iter == 4 and abs(solver.us[0][0]) < 1e-12


def test_negative_curvature_needs_nine_regularisation_increases():
    """l(u) = -u^2/4 + u^4/4 at u = 0.1: l'' = -0.47, so Quu = l'' + x_reg + u_reg is negative until the
    regularisation, multiplied by 10 from 1e-9 after every failed Cholesky, reaches 1 (nine increases):
    Quu = 1.53, k = l'(0.1) / 1.53, the full step decreases the cost and is taken, regularisation drops to 0.1;
    in the next iteration l''(u1) + 0.2 < 0 again: one more increase, back to 1, and so on."""
    l = lambda u: -0.25 * u * u + 0.25 * u ** 4
    dl = lambda u: -0.5 * u + u ** 3
    ddl = lambda u: -0.5 + 3 * u * u
    problem = croc.ShootingProblem(np.zeros(1), [_Scalar(l, dl, ddl)], _Zero())
    solver = croc.SolverDDP(problem)
    log = croc.CallbackLogger()
    solver.setCallbacks([log])
    solver.solve([np.zeros(1), np.array([0.1])], [np.array([0.1])], 2, True)
    quu = ddl(0.1) + 2.0
    k0 = dl(0.1) / quu
    u1 = 0.1 - k0
    assert abs(quu - 1.53) < 1e-12 and abs(log.costs[0] - l(u1)) < 1e-15
    assert log.steps[0] == 1.0
    assert abs(log.x_regs[0] - 0.1) < 1e-12  # 1e-9 x 10^9 = 1, then one decrease after the accepted full step
    # second iteration: l''(u1) + 2 * 0.1 < 0 -> one failed Cholesky -> 1.0 -> step taken -> back to 0.1
    assert ddl(u1) + 0.2 < 0 < ddl(u1) + 2.0
    u2 = u1 - dl(u1) / (ddl(u1) + 2.0)
    assert abs(log.costs[1] - l(u2)) < 1e-15 and abs(log.x_regs[1] - 0.1) < 1e-12
