"""Pins the oracle (CPU tier, no GPU):
  * the pinocchio shim against independent checks (closed-form planar 2-link Lagrangian dynamics,
    central differences, log6/exp6 round trips) -- the reference has no golden vectors for it;
  * the crocoddyl solver shim against the discrete-time LQR known answer;
  * the port (oracle/aslr_port.py) against the committed golden vectors, which were generated from the
    reference's own unmodified classes (tests/golden/make_golden.py), and -- where /root/reference
    exists -- against the reference itself on fresh seeded inputs."""
import numpy as np
import pytest

import oracle_build as ob
from aslr_to_b200 import workloads as W
from util import golden, rel_err

pin = ob.pinocchio
croc = ob.crocoddyl

SINGLES = {"double_pendulum": W.double_pendulum, "double_pendulum_residual": W.double_pendulum_residual,
           "two_dof_sea": W.two_dof_sea,
           "two_dof_vsa_boxddp": W.two_dof_vsa_boxddp, "two_dof_vsa_modified": W.two_dof_vsa_modified}
KNOT_FIELDS = ("xnext", "cost", "xout", "dFx", "dFu", "Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu")


def chain7_workloads():
    c7 = W.chain7_batch(B=2, T=20)
    c7v = dict(c7, name="chain7_vsa", kind="vsa", K=None,
               running=[t for t in c7["running"] if t["kind"] != "control"] +
               [dict(name="uReg", kind="control", weight=1e-2, ref=np.zeros(14), act=np.ones(14), frame=None)])
    return {"chain7": c7, "chain7_vsa": c7v}


def all_knot_workloads():
    d = dict((k, f()) for k, f in SINGLES.items())
    d.update(chain7_workloads())
    return d


# ------------------------------------------------------------------------------ pinocchio shim
def test_planar_two_link_closed_form():
    """M and nle of the asr_twodof fixture against the textbook planar 2R manipulator equations."""
    m = ob.example_robot_data.load("asr_twodof").model
    m.gravity.linear = np.array([9.81, 0.0, 0.0])
    j1, j2 = m.joints
    l1 = j2.placement.translation[0]
    m1, m2, lc1, lc2 = j1.mass, j2.mass, j1.com[0], j2.com[0]
    I1, I2 = j1.inertia[2, 2], j2.inertia[2, 2]
    rng = np.random.default_rng(0)
    d = m.createData()
    for _ in range(5):
        q, v = rng.uniform(-3, 3, 2), rng.uniform(-2, 2, 2)
        pin.computeAllTerms(m, d, q, v)
        c2, s2 = np.cos(q[1]), np.sin(q[1])
        M11 = I1 + I2 + m1 * lc1 ** 2 + m2 * (l1 ** 2 + lc2 ** 2 + 2 * l1 * lc2 * c2)
        M12 = I2 + m2 * (lc2 ** 2 + l1 * lc2 * c2)
        M22 = I2 + m2 * lc2 ** 2
        h = m2 * l1 * lc2 * s2
        # potential U = -m g . r_com with g = (9.81, 0, 0): generalized gravity torque = dU/dq
        g = 9.81
        G1 = (m1 * lc1 + m2 * l1) * g * np.sin(q[0]) + m2 * lc2 * g * np.sin(q[0] + q[1])
        G2 = m2 * lc2 * g * np.sin(q[0] + q[1])
        nle = np.array([-h * (2 * v[0] * v[1] + v[1] ** 2) + G1, h * v[0] ** 2 + G2])
        assert rel_err(d.M, np.array([[M11, M12], [M12, M22]])) < 1e-13
        assert rel_err(d.nle, nle) < 1e-12


@pytest.mark.parametrize("robot", ["asr_twodof", "double_pendulum", "chain7"])
def test_rnea_derivatives_and_frame_jacobian_vs_central_differences(robot):
    m = ob.example_robot_data.load(robot).model
    d = m.createData()
    rng = np.random.default_rng(1)
    q, v, a = rng.uniform(-2, 2, m.nv), rng.uniform(-1, 1, m.nv), rng.uniform(-1, 1, m.nv)
    dq, dv = pin.computeRNEADerivatives(m, d, q, v, a)
    h = 1e-6
    for j in range(m.nv):
        e = np.zeros(m.nv)
        e[j] = h
        fq = (pin.rnea(m, d, q + e, v, a) - pin.rnea(m, d, q - e, v, a)) / (2 * h)
        fv = (pin.rnea(m, d, q, v + e, a) - pin.rnea(m, d, q, v - e, a)) / (2 * h)
        assert np.abs(fq - dq[:, j]).max() < 2e-7 * max(1, np.abs(dq).max())
        assert np.abs(fv - dv[:, j]).max() < 2e-7 * max(1, np.abs(dv).max())
    pin.computeAllTerms(m, d, q, v)
    assert np.abs(pin.rnea(m, d, q, v, a) - (d.M @ a + d.nle)).max() < 1e-12
    assert np.linalg.eigvalsh(d.M).min() > 0
    fid = len(m.frames) - 1
    pin.forwardKinematics(m, d, q)
    pin.updateFramePlacements(m, d)
    M0 = d.oMf[fid].copy()
    J = pin.getFrameJacobian(m, d, fid, pin.ReferenceFrame.LOCAL)
    for j in range(m.nv):
        e = np.zeros(m.nv)
        e[j] = h
        pin.forwardKinematics(m, d, q + e)
        pin.updateFramePlacements(m, d)
        Jp = pin.log6(M0.inverse() * d.oMf[fid]).vector
        pin.forwardKinematics(m, d, q - e)
        pin.updateFramePlacements(m, d)
        Jm = pin.log6(M0.inverse() * d.oMf[fid]).vector
        assert np.abs((Jp - Jm) / (2 * h) - J[:, j]).max() < 1e-8


def test_log6_round_trip_and_right_jacobian():
    rng = np.random.default_rng(2)
    for scale in (1e-7, 1e-3, 0.5, 2.5):
        nu = rng.uniform(-1, 1, 6) * scale
        M = pin.exp6(nu)
        assert np.abs(pin.log6(M).vector - nu).max() < 1e-12 * max(1, scale)
        J = pin.Jlog6(M)
        h = 1e-6
        for k in range(6):
            e = np.zeros(6)
            e[k] = h
            fd = (pin.log6(M * pin.exp6(e)).vector - pin.log6(M * pin.exp6(-e)).vector) / (2 * h)
            assert np.abs(fd - J[:, k]).max() < 1e-8


# ------------------------------------------------------------------------ crocoddyl solver shim
class _LqrState(croc.StateAbstract):
    def __init__(self, n):
        croc.StateAbstract.__init__(self, n, n)

    def zero(self):
        return np.zeros(self.nx)

    def diff(self, a, b):
        return np.asarray(b) - np.asarray(a)

    def integrate(self, x, dx):
        return np.asarray(x) + np.asarray(dx)


class _Lqr(croc.ActionModelAbstract):
    def __init__(self, A_, B_, Q, R, q, r):
        croc.ActionModelAbstract.__init__(self, _LqrState(A_.shape[0]), B_.shape[1], 0)
        self.A, self.Bm, self.Q, self.R, self.q, self.r = A_, B_, Q, R, q, r

    def calc(self, data, x, u=None):
        u = np.zeros(self.nu) if u is None else u
        data.xnext = self.A @ x + self.Bm @ u
        data.cost = 0.5 * x @ self.Q @ x + 0.5 * u @ self.R @ u + self.q @ x + self.r @ u

    def calcDiff(self, data, x, u=None):
        u = np.zeros(self.nu) if u is None else u
        data.Fx, data.Fu = self.A, self.Bm
        data.Lx, data.Lu = self.Q @ x + self.q, self.R @ u + self.r
        data.Lxx, data.Luu, data.Lxu = self.Q, self.R, np.zeros((self.A.shape[0], self.nu))


@pytest.mark.parametrize("solver_cls", [croc.SolverDDP, croc.SolverFDDP])
def test_solver_shim_solves_lqr_in_one_iteration(solver_cls):
    """Known answer: on an LQR problem DDP/FDDP converge after one full step and the gains equal the
    discrete Riccati recursion."""
    rng = np.random.default_rng(3)
    n, m, T = 4, 2, 12
    A_ = np.eye(n) + 0.1 * rng.standard_normal((n, n))
    B_ = rng.standard_normal((n, m))
    Q = np.eye(n)
    R = 0.1 * np.eye(m)
    model = _Lqr(A_, B_, Q, R, rng.standard_normal(n), rng.standard_normal(m))
    term = _Lqr(A_, B_, 10 * Q, R, np.zeros(n), np.zeros(m))
    x0 = rng.standard_normal(n)
    problem = croc.ShootingProblem(x0, [model] * T, term)
    solver = solver_cls(problem)
    log = croc.CallbackLogger()
    solver.setCallbacks([log])
    solver.th_stop = 1e-12
    assert solver.solve([], [], 10)
    assert solver.iter == 1 and log.steps[0] == 1.0
    P = 10 * Q
    for t in range(T - 1, -1, -1):
        Kt = np.linalg.solve(R + B_.T @ P @ B_, B_.T @ P @ A_)
        # x_reg = u_reg = 1e-9 perturb the recursion at the 1e-9 level
        assert np.abs(Kt - solver.K[t]).max() < 1e-6
        P = Q + A_.T @ P @ A_ - A_.T @ P @ B_ @ Kt
    if solver_cls is croc.SolverFDDP:  # plain DDP's expected improvement ignores the initial gaps
        assert abs(log.dVs[0] - log.dVexps[0]) < 1e-8 * abs(log.dVs[0])


# --------------------------------------------------------------------------------------- the port
@pytest.mark.parametrize("name", ["double_pendulum", "double_pendulum_residual", "two_dof_sea", "two_dof_vsa_boxddp",
                                  "two_dof_vsa_modified", "chain7", "chain7_vsa"])
def test_port_matches_golden_knots(name):
    w = all_knot_workloads()[name]
    g = golden("knots_" + name)
    run, term, _ = ob.port_models(w, 0)
    for tag, m in (("run_", run), ("term_", term)):
        d = m.createData()
        for i, (x, u) in enumerate(zip(g["xs"], g["us"])):
            if tag == "run_":
                m.calc(d, x, u)
                m.calcDiff(d, x, u)
            else:
                m.calc(d, x)
                m.calcDiff(d, x)
            vals = dict(xnext=d.xnext, cost=d.cost, xout=d.xout, dFx=d.dFx, dFu=d.dFu, Fx=d.Fx, Fu=d.Fu, Lx=d.Lx,
                        Lu=d.Lu, Lxx=d.Lxx, Lxu=d.Lxu, Luu=d.Luu)
            for k in KNOT_FIELDS:
                assert rel_err(vals[k], g[tag + k][i]) < 1e-12, (tag, k, i)


@pytest.mark.parametrize("name", ["double_pendulum", "two_dof_sea"])
def test_port_solver_trace_matches_golden(name):
    w = SINGLES[name]()
    g = golden("solve_" + name)
    solver, log = ob.solve(w, 0, "port")
    assert solver.iter == int(g["iter"]) and bool(solver.converged) == bool(g["converged"])
    assert rel_err(log.steps, g["steps"]) == 0 and rel_err(log.x_regs, g["x_regs"]) == 0
    if name == "double_pendulum":
        # ill-conditioned, non-converging swing-up (B = 1e-3, q_m far from q_l): rounding differences are
        # amplified ~10x every few iterations, so only the first iterations can agree to 1e-9
        assert rel_err(log.costs[:20], g["costs"][:20]) < 1e-9
        assert rel_err(log.costs, g["costs"]) < 1e-5
    else:
        assert rel_err(log.costs, g["costs"]) < 1e-10
        assert rel_err(np.array(solver.xs), g["xs"]) < 1e-9
        assert rel_err(np.array(solver.us), g["us"]) < 1e-9


@pytest.mark.skipif(not ob.refenv.have_reference(), reason="/root/reference only exists in the build container")
@pytest.mark.parametrize("name", ["double_pendulum", "double_pendulum_residual", "two_dof_sea", "two_dof_vsa_modified",
                                  "chain7"])
def test_port_matches_reference_on_fresh_inputs(name):
    w = all_knot_workloads()[name]
    rng = np.random.default_rng(99)
    pr, pt, _ = ob.port_models(w, 0)
    rr, rt, _ = ob.reference_models(w, 0)
    n = pr.spec.n
    for _ in range(3):
        x = np.concatenate([rng.uniform(-3, 3, 2 * n), rng.uniform(-2, 2, 2 * n)])
        u = rng.uniform(0.1, 2, pr.nu)
        for pm, rm_, uu in ((pr, rr, u), (pt, rt, None)):
            dp, dr = pm.createData(), rm_.createData()
            if uu is None:
                pm.calc(dp, x), pm.calcDiff(dp, x), rm_.calc(dr, x), rm_.calcDiff(dr, x)
            else:
                pm.calc(dp, x, uu), pm.calcDiff(dp, x, uu), rm_.calc(dr, x, uu), rm_.calcDiff(dr, x, uu)
            for k in ("xnext", "cost", "Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu"):
                assert rel_err(getattr(dp, k), getattr(dr, k)) < 1e-12, k


def test_reference_unit_test_pattern_numdiff():
    """The reference's own test strategy (unittest/test_stiffness_residual.py:36-56): analytic
    derivatives against crocoddyl's NumDiff -- re-run on the port-equivalent reference classes at a much
    tighter tolerance than the reference's 6e-3."""
    if not ob.refenv.have_reference():
        pytest.skip("/root/reference only exists in the build container")
    aslr_to = ob.refenv.import_reference()
    rm = ob.example_robot_data.load("asr_twodof").model
    state = aslr_to.StateMultibodyASR(rm)
    act = aslr_to.VSAASRActuation(state)
    nu = 2 * act.nu
    costs = croc.CostModelSum(state, nu)
    res = aslr_to.ResidualModelFramePlacementASR(state, rm.getFrameId("EE"),
                                                 pin.SE3(np.eye(3), np.array([.0, .0, .4])), nu)
    costs.addCost("gripperPose", croc.CostModelResidual(state, res), nu)
    costs.addCost("xReg", croc.CostModelResidual(state, croc.ResidualModelControl(state, nu)), 1e-2)
    costs.addCost("vsa", aslr_to.CostModelStiffness(state, nu, .1, np.zeros(2)), 1.0)
    model = aslr_to.DifferentialFreeFwdDynamicsModelVSA(state, act, costs)
    np.random.seed(5)
    x, u = state.rand(), np.random.rand(nu)
    data = model.createData()
    nd = croc.DifferentialActionModelNumDiff(model)
    dnd = nd.createData()
    model.calc(data, x, u), model.calcDiff(data, x, u), nd.calc(dnd, x, u), nd.calcDiff(dnd, x, u)
    for k in ("Fx", "Fu", "Lx", "Lu"):
        a, b = np.asarray(getattr(data, k)), np.asarray(getattr(dnd, k))
        assert np.abs(a - b).max() < 2e-5 * max(1.0, np.abs(a).max()), k
