"""The drop-in Python layer (aslr_to_b200.compat): the reference's example scripts, rebuilt with the
reference's own class names, must compile to exactly the problem description the workload tables give
(CPU tier, byte comparison) and reproduce the oracle's golden solver traces on the GPU (-m gpu tier)."""
import ctypes as C

import numpy as np
import pytest

import oracle_build as ob
import util as U
from aslr_to_b200 import _abi as A
from aslr_to_b200 import workloads as W
from aslr_to_b200.compat import aslr_to, crocoddyl, example_robot_data, pinocchio


def build_two_dof_sea():
    """examples/two_dof_sea.py:19-69, import lines changed only."""
    two_dof = example_robot_data.load('asr_twodof')
    robot_model = two_dof.model
    robot_model.gravity.linear = np.array([9.81, 0, 0])
    state = aslr_to.StateMultibodyASR(robot_model)
    actuation = aslr_to.ASRActuation(state)
    nu = actuation.nu
    runningCostModel = crocoddyl.CostModelSum(state, nu)
    terminalCostModel = crocoddyl.CostModelSum(state, nu)
    xActivation = crocoddyl.ActivationModelWeightedQuad(np.array([1e0] * 2 + [0] * 2 + [1e0] * robot_model.nv + [0] * robot_model.nv))
    xResidual = crocoddyl.ResidualModelState(state, state.zero(), nu)
    xRegCost = crocoddyl.CostModelResidual(state, xActivation, xResidual)
    uResidual = crocoddyl.ResidualModelControl(state, nu)
    uRegCost = crocoddyl.CostModelResidual(state, uResidual)
    target = np.array([0.01, 2.03063311e-01, 1.80000000e-01])
    framePlacementResidual = aslr_to.ResidualModelFramePlacementASR(state, robot_model.getFrameId("EE"),
                                                                   pinocchio.SE3(np.eye(3), target), nu)
    goalTrackingCost = crocoddyl.CostModelResidual(state, framePlacementResidual)
    runningCostModel.addCost("gripperPose", goalTrackingCost, 1e-1)
    runningCostModel.addCost("xReg", xRegCost, 1e-3)
    runningCostModel.addCost("uReg", uRegCost, 1e-2)
    terminalCostModel.addCost("gripperPose", goalTrackingCost, 1e4)
    K = 1 * np.eye(int(state.nv / 2))
    B = .01 * np.eye(int(state.nv / 2))
    dt = 1e-2
    runningModel = aslr_to.IntegratedActionModelEulerASR(
        aslr_to.DifferentialFreeASRFwdDynamicsModel(state, actuation, runningCostModel, K, B), dt)
    terminalModel = aslr_to.IntegratedActionModelEulerASR(
        aslr_to.DifferentialFreeASRFwdDynamicsModel(state, actuation, terminalCostModel, K, B), 0)
    T = 100
    q0 = np.array([.0, 0])
    x0 = np.concatenate([q0, q0, pinocchio.utils.zero(state.nv)])
    problem = crocoddyl.ShootingProblem(x0, [runningModel] * T, terminalModel)
    return problem, x0


def build_two_dof_vsa(modified):
    """examples/two_dof_vsa_boxddp.py:14-69 and two_dof_vsa_modified.py:14-66."""
    two_dof = example_robot_data.load('asr_twodof')
    robot_model = two_dof.model
    robot_model.gravity.linear = np.array([9.81, 0, 0])
    state = aslr_to.StateMultibodyASR(robot_model)
    actuation = aslr_to.VSAASRActuation(state)
    nu = 2 * actuation.nu
    runningCostModel = crocoddyl.CostModelSum(state, nu)
    terminalCostModel = crocoddyl.CostModelSum(state, nu)
    xResidual = crocoddyl.ResidualModelState(state, state.zero(), nu)
    xRegCost = crocoddyl.CostModelResidual(state, xResidual)
    uResidual = crocoddyl.ResidualModelControl(state, nu)
    if modified:
        uRegCost = crocoddyl.CostModelResidual(state, crocoddyl.ActivationModelWeightedQuad(np.array([1., 1., 0, 0])), uResidual)
    else:
        uRegCost = crocoddyl.CostModelResidual(state, uResidual)
    target = np.array([.01, .2, .18])
    framePlacementResidual = aslr_to.ResidualModelFramePlacementASR(state, robot_model.getFrameId("EE"),
                                                                   pinocchio.SE3(np.eye(3), target), nu)
    goalTrackingCost = crocoddyl.CostModelResidual(state, framePlacementResidual)
    Kref = 0.002 * np.ones(int(nu / 2))
    vsaCost = aslr_to.CostModelStiffness(state, nu, 10, Kref)
    if modified:
        runningCostModel.addCost("gripperPose", goalTrackingCost, 1e0)
        runningCostModel.addCost("xReg", xRegCost, 1e-3)
        runningCostModel.addCost("uReg", uRegCost, 1e-2)
        runningCostModel.addCost("vsa", vsaCost, 1e-2)
        terminalCostModel.addCost("gripperPose", goalTrackingCost, 1e4)
    else:
        runningCostModel.addCost("gripperPose", goalTrackingCost, 1e0)
        runningCostModel.addCost("xReg", xRegCost, 1e-1)
        runningCostModel.addCost("uReg", uRegCost, 1e-1)
        terminalCostModel.addCost("gripperPose", goalTrackingCost, 4e4)
    B = .001 * np.eye(int(state.nv / 2))
    dt = 1e-2
    runningModel = aslr_to.IntegratedActionModelEulerASR(
        aslr_to.DifferentialFreeFwdDynamicsModelVSA(state, actuation, runningCostModel, B), dt)
    terminalModel = aslr_to.IntegratedActionModelEulerASR(
        aslr_to.DifferentialFreeFwdDynamicsModelVSA(state, actuation, terminalCostModel, B), 0)
    lo = 0.002 if modified else 0.0
    runningModel.u_lb = np.array([-100, -100, lo, lo])
    runningModel.u_ub = np.array([100, 100, 100, 100])
    T = 200
    q0 = np.array([.0, .0])
    x0 = np.concatenate([q0, q0, pinocchio.utils.zero(state.nv)])
    return crocoddyl.ShootingProblem(x0, [runningModel] * T, terminalModel), x0


def build_double_pendulum():
    """examples/double_pendulum.py:13-53 (nu = 2)."""
    pendulum = example_robot_data.load('double_pendulum')
    model = pendulum.model
    state = aslr_to.StateMultibodyASR(model)
    actuation = aslr_to.ActuationModelDoublePendulum(state, actLink=0, nu=2)
    nu = actuation.nu
    runningCostModel = crocoddyl.CostModelSum(state, nu)
    terminalCostModel = crocoddyl.CostModelSum(state, nu)
    xResidual = crocoddyl.ResidualModelState(state, state.zero(), nu)
    xActivation = crocoddyl.ActivationModelWeightedQuad(np.array([1.] * 2 + [0] * 2 + [1.] * 2 + [0] * 2))
    xRegCost = crocoddyl.CostModelResidual(state, xActivation, xResidual)
    uResidual = crocoddyl.ResidualModelControl(state, nu)
    uActivation = crocoddyl.ActivationModelWeightedQuad(np.array([1., 0.]))
    uRegCost = crocoddyl.CostModelResidual(state, uActivation, uResidual)
    xPendCost = aslr_to.CostModelDoublePendulum(state, crocoddyl.ActivationModelWeightedQuad(np.array([1.] * 4 + [0.1] * 2)), nu)
    dt = 1e-2
    runningCostModel.addCost("uReg", uRegCost, 1e-1)
    runningCostModel.addCost("xReg", xRegCost, 1e-2)
    runningCostModel.addCost("xGoalR", xPendCost, 1e-1)
    terminalCostModel.addCost("xGoal", xPendCost, 1e4)
    K = 1 * np.eye(int(state.nv / 2))
    B = .001 * np.eye(int(state.nv / 2))
    runningModel = aslr_to.IntegratedActionModelEulerASR(
        aslr_to.DifferentialFreeASRFwdDynamicsModel(state, actuation, runningCostModel, K, B), dt)
    terminalModel = aslr_to.IntegratedActionModelEulerASR(
        aslr_to.DifferentialFreeASRFwdDynamicsModel(state, actuation, terminalCostModel, K, B), 0)
    T = 10
    x0 = np.array([3.14, 0, 0, 0, 0., 0., 0., 0.])
    return crocoddyl.ShootingProblem(x0, [runningModel] * T, terminalModel), x0


BUILDERS = {"two_dof_sea": (build_two_dof_sea, W.two_dof_sea), "double_pendulum": (build_double_pendulum, W.double_pendulum),
            "two_dof_vsa_boxddp": (lambda: build_two_dof_vsa(False), W.two_dof_vsa_boxddp),
            "two_dof_vsa_modified": (lambda: build_two_dof_vsa(True), W.two_dof_vsa_modified)}
SOLVERS = {"fddp": crocoddyl.SolverFDDP, "ddp": crocoddyl.SolverDDP, "boxddp": crocoddyl.SolverBoxDDP}


# ------------------------------------------------------------------------------------------ CPU tier
@pytest.mark.parametrize("name", sorted(BUILDERS))
def test_example_built_with_reference_names_compiles_to_the_workload_description(name):
    from aslr_to_b200.compat import _engine as E
    build, workload = BUILDERS[name]
    problem, x0 = build()
    w = workload()
    run, term = problem._specs()
    desc = E.compile_desc(run, term)
    assert bytes(desc) == bytes(W.to_desc(w)), "AslrProblemDesc differs from the workload table"
    assert problem.T == w["T"] and np.array_equal(x0, w["x0"])


def test_reference_signatures_and_errors():
    model = example_robot_data.load('asr_twodof').model
    state = aslr_to.StateMultibodyASR(model)
    assert (state.nx, state.ndx, state.nq, state.nv) == (8, 8, 4, 4)  # statemultibody_aslr.py:14-16
    assert np.array_equal(state.zero(), np.zeros(8))
    x0, x1 = state.rand(), state.rand()
    assert np.allclose(state.integrate(x0, state.diff(x0, x1)), x1)
    vsa = aslr_to.DifferentialFreeFwdDynamicsModelVSA(state, aslr_to.VSAASRActuation(state), crocoddyl.CostModelSum(state, 4))
    assert vsa.nu == 4 and np.array_equal(vsa.B, 1e-3 * np.eye(2))  # free_fwddyn_vsa.py:8,12-15
    sea = aslr_to.DifferentialFreeASRFwdDynamicsModel(state, aslr_to.ASRActuation(state), crocoddyl.CostModelSum(state, 2))
    assert sea.nu == 2 and np.array_equal(sea.K, 1e-1 * np.eye(2))  # free_fwddyn_asr.py:12-15
    with pytest.raises(ValueError):
        crocoddyl.CostModelResidual(state, crocoddyl.ActivationModelWeightedQuad(np.ones(3)), crocoddyl.ResidualModelControl(state, 2))
    m = aslr_to.IntegratedActionModelEulerASR(sea, 1e-2)
    with pytest.raises(NotImplementedError):
        crocoddyl.ShootingProblem(np.zeros(8), [m, aslr_to.IntegratedActionModelEulerASR(sea, 2e-2)], m)
    log = crocoddyl.CallbackLogger()
    log.us = [np.array([1.0, 2.0]), np.array([3.0, -1.0])]
    assert np.array_equal(aslr_to.u_squared(log), [10.0, 5.0])  # __init__.py:63-68


@pytest.mark.parametrize("name", ["two_dof_sea", "double_pendulum", "two_dof_vsa_modified", "chain7"])
def test_quasi_static_device_body_matches_the_oracle(hostmath, name):
    """crocoddyl's base-class quasiStatic (oracle shim, np.linalg.pinv) against the device body compiled for
    the host, at states with q_l != q_m where the answer is non-trivial (u = -K (q_l - q_m) for SEA)."""
    rng = np.random.default_rng(7)
    if name == "chain7":
        w = W.chain7_batch(B=2, T=5)
    else:
        w = getattr(W, name)()
    desc = W.to_desc(w)
    run, _, _ = ob.port_models(w, 0)
    n = desc.robot.n
    for trial in range(3):
        x = np.concatenate([rng.uniform(-0.5, 0.5, 2 * n), np.zeros(2 * n)])
        ref = np.asarray(run.quasiStatic(run.createData(), x), float)
        u = np.zeros(desc.dyn.nu)
        rc = hostmath.hh_quasi_static(C.byref(desc), x.ctypes.data_as(C.c_void_p), None, 100, C.c_double(1e-9),
                                      u.ctypes.data_as(C.c_void_p))
        assert rc == 0
        assert U.rel_err(u, ref) <= U.RTOL, (name, u, ref)
        if name == "two_dof_sea":  # closed form
            assert U.rel_err(u, -np.asarray(w["K"]) @ (x[:n] - x[n:2 * n])) <= U.RTOL


# ------------------------------------------------------------------------------------------ GPU tier
@pytest.mark.gpu
@pytest.mark.parametrize("name", ["two_dof_sea", "double_pendulum", "two_dof_vsa_boxddp"])
def test_example_through_the_drop_in_layer_matches_the_golden_trace(built_library, name):
    build, workload = BUILDERS[name]
    problem, x0 = build()
    w = workload()
    solver = SOLVERS[w["solver"]](problem)
    solver.setCallbacks([crocoddyl.CallbackLogger()])
    if w["th_stop"] is not None:
        solver.th_stop = w["th_stop"]
    if w["init"] == "quasistatic":
        xs = [x0] * (problem.T + 1)
        us = problem.quasiStatic([x0] * problem.T)
        assert np.array_equal(np.asarray(us), np.zeros((problem.T, 2)))
        converged = solver.solve(xs, us, w["maxiter"])
    else:
        converged = solver.solve([], [], w["maxiter"])
    gold = U.golden("solve_" + name)
    log = solver.getCallbacks()[0]
    assert converged == bool(gold["converged"])
    assert len(log.costs) == len(gold["costs"])
    assert np.array_equal(log.steps, gold["steps"])
    loose = 1e-9 if name != "double_pendulum" else None
    if loose:
        assert U.rel_err(log.costs, gold["costs"]) <= 1e-9
        assert U.rel_err(np.asarray(solver.xs), gold["xs"]) <= 1e-9
        assert U.rel_err(np.asarray(solver.us), gold["us"]) <= 1e-9
        assert U.rel_err(np.asarray(solver.K), gold["K"]) <= 1e-9
    else:
        # chaotic, non-converging example: identical decisions, values drift
        assert U.rel_err(log.costs[:2], gold["costs"][:2]) <= 1e-9
        assert U.rel_err(log.costs, gold["costs"]) <= 5e-2
    assert np.sum(aslr_to.u_squared(log)) >= 0.0


@pytest.mark.gpu
def test_model_level_calc_and_calcdiff_match_the_oracle(built_library):
    """unittest/test_asr_free_fwddyn.py pattern: model.calc / calcDiff called directly on one (x, u)."""
    rng = np.random.default_rng(3)
    problem, _ = build_two_dof_sea()
    model = problem.runningModels[0]
    run, _, _ = ob.port_models(W.two_dof_sea(), 0)
    x, u = rng.uniform(-1, 1, 8), rng.uniform(-1, 1, 2)
    data, odata = model.createData(), run.createData()
    model.calc(data, x, u)
    model.calcDiff(data, x, u)
    run.calc(odata, x, u)
    run.calcDiff(odata, x, u)
    for k in ("xnext", "Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu"):
        assert U.rel_err(getattr(data, k), np.asarray(getattr(odata, k)).reshape(np.shape(getattr(data, k)))) <= U.RTOL, k
    assert abs(data.cost - odata.cost) <= U.RTOL * max(1.0, abs(odata.cost))
    # differential level + `u is None` default
    diff = model.differential
    dd = diff.createData()
    diff.calcDiff(dd, x)
    run.calc(odata, x, np.zeros(2))
    assert U.rel_err(dd.xout, (np.asarray(odata.xnext)[4:] - x[4:]) / 1e-2) <= 1e-7
    # problem-level sweep and sequential rollout
    xs = problem.rollout([u] * problem.T)
    total = problem.calc(xs, [u] * problem.T)
    assert np.isfinite(total) and U.rel_err(problem.runningDatas[3].xnext, xs[4]) <= 1e-12
    us = problem.quasiStatic([x] * problem.T)
    assert U.rel_err(us[0], np.asarray(run.quasiStatic(run.createData(), x), float)) <= U.RTOL
