"""Workload definitions = the constants of the reference's example scripts as plain data
(SURVEY.md section 8d).  The example scripts themselves are not ported; these dictionaries are
what bench.py, the tests and the oracle builders all start from, so both sides of every parity
check see identical numbers.

Cost terms are listed in crocoddyl's CostModelSum iteration order (a std::map: sorted by name).
"""
import numpy as np

from . import _abi as A
from . import desc as D

SOLVER_KIND = {"ddp": A.SOLVER_DDP, "fddp": A.SOLVER_FDDP, "boxddp": A.SOLVER_BOXDDP}


def _term(name, kind, weight, ref=None, act=None, frame=None):
    return dict(name=name, kind=kind, weight=float(weight), ref=None if ref is None else np.asarray(ref, float),
                act=None if act is None else np.asarray(act, float), frame=frame)


def _placement(name, weight, frame, target_p, target_R=None, act=None):
    R = np.eye(3) if target_R is None else np.asarray(target_R, float)
    ref = np.concatenate([R.reshape(9), np.asarray(target_p, float)])
    return _term(name, "placement", weight, ref, np.ones(6) if act is None else act, frame)


def _sorted(terms):
    return sorted(terms, key=lambda t: t["name"])


def double_pendulum():
    """examples/double_pendulum.py:13-70 (with nu = 2)."""
    n, nu = 2, 2
    Sm = np.zeros((n, n))
    Sm[0, 0] = 1.0  # ActuationModelDoublePendulum(actLink=0): S[nv/2, 0] = 1 (__init__.py:288)
    pend = np.array([1.0] * 4 + [0.1] * 2)
    running = _sorted([
        _term("uReg", "control", 1e-1, np.zeros(nu), np.array([1.0, 0.0])),
        _term("xReg", "state", 1e-2, np.zeros(8), np.array([1.0] * 2 + [0] * 2 + [1.0] * 2 + [0] * 2)),
        _term("xGoalR", "pendulum", 1e-1, None, pend)])
    terminal = [_term("xGoal", "pendulum", 1e4, None, pend)]
    return dict(name="double_pendulum", robot="double_pendulum", gravity=None, kind="sea", K=np.eye(n),
                B=1e-3 * np.eye(n), Sm=Sm, dt=1e-2, T=10, x0=np.array([3.14, 0, 0, 0, 0, 0, 0, 0.0]),
                running=running, terminal=terminal, solver="fddp", th_stop=None, maxiter=100, init="empty",
                u_lb=None, u_ub=None)


def double_pendulum_residual():
    """examples/double_pendulum.py with the swing-up cost expressed through the package's residual model
    (python/aslr_to/residual_acrobot.py:5-29, exported at __init__.py:9; no example uses it) under crocoddyl's
    CostModelResidual + weighted-quad activation: residual `1 - c2`, Gauss-Newton Hessian."""
    w = double_pendulum()
    for terms in (w["running"], w["terminal"]):
        for t in terms:
            if t["kind"] == "pendulum":
                t["kind"] = "pendulum_residual"
    w["name"] = "double_pendulum_residual"
    return w


def two_dof_sea():
    """examples/two_dof_sea.py:18-81."""
    n, nu = 2, 2
    target = np.array([0.01, 2.03063311e-01, 1.80000000e-01])
    running = _sorted([
        _placement("gripperPose", 1e-1, "EE", target),
        _term("xReg", "state", 1e-3, np.zeros(8), np.array([1.0] * 2 + [0] * 2 + [1.0] * 2 + [0] * 2)),
        _term("uReg", "control", 1e-2, np.zeros(nu), np.ones(nu))])
    terminal = [_placement("gripperPose", 1e4, "EE", target)]
    return dict(name="two_dof_sea", robot="asr_twodof", gravity=[9.81, 0, 0], kind="sea", K=np.eye(n),
                B=0.01 * np.eye(n), Sm=None, dt=1e-2, T=100, x0=np.zeros(8), running=running, terminal=terminal,
                solver="fddp", th_stop=1e-7, maxiter=100, init="quasistatic", u_lb=None, u_ub=None)


def two_dof_vsa_boxddp():
    """examples/two_dof_vsa_boxddp.py:14-81 (stiffness cost constructed but not added, :38,47)."""
    n, nu = 2, 4
    target = np.array([0.01, 0.2, 0.18])
    running = _sorted([
        _placement("gripperPose", 1e0, "EE", target),
        _term("xReg", "state", 1e-1, np.zeros(8), np.ones(8)),
        _term("uReg", "control", 1e-1, np.zeros(nu), np.ones(nu))])
    terminal = [_placement("gripperPose", 4e4, "EE", target)]
    return dict(name="two_dof_vsa_boxddp", robot="asr_twodof", gravity=[9.81, 0, 0], kind="vsa", K=None,
                B=1e-3 * np.eye(n), Sm=None, dt=1e-2, T=200, x0=np.zeros(8), running=running, terminal=terminal,
                solver="boxddp", th_stop=1e-7, maxiter=400, init="empty", u_lb=np.array([-100.0, -100, 0, 0]),
                u_ub=np.array([100.0, 100, 100, 100]))


def two_dof_vsa_modified():
    """examples/two_dof_vsa_modified.py:14-78 (active linear stiffness cost, sigma_lb = 0.002)."""
    n, nu = 2, 4
    target = np.array([0.01, 0.2, 0.18])
    running = _sorted([
        _placement("gripperPose", 1e0, "EE", target),
        _term("xReg", "state", 1e-3, np.zeros(8), np.ones(8)),
        _term("uReg", "control", 1e-2, np.zeros(nu), np.array([1.0, 1.0, 0, 0])),
        _term("vsa", "stiffness", 1e-2, 0.002 * np.ones(n), np.array([10.0]))])
    terminal = [_placement("gripperPose", 1e4, "EE", target)]
    return dict(name="two_dof_vsa_modified", robot="asr_twodof", gravity=[9.81, 0, 0], kind="vsa", K=None,
                B=1e-3 * np.eye(n), Sm=None, dt=1e-2, T=200, x0=np.zeros(8), running=running, terminal=terminal,
                solver="boxddp", th_stop=1e-7, maxiter=400, init="empty",
                u_lb=np.array([-100.0, -100, 0.002, 0.002]), u_ub=np.array([100.0, 100, 100, 100]))


def sea_batch(B=4096, seed=20240001, solver="fddp"):
    """BASELINE.json configs[3]: B independent copies of two_dof_sea with random initial posture and
    joint stiffness."""
    w = two_dof_sea()
    rng = np.random.default_rng(seed)
    q0 = rng.uniform(-np.pi / 2, np.pi / 2, size=(B, 2))
    k = np.exp(rng.uniform(np.log(0.5), np.log(5.0), size=(B, 2)))
    x0 = np.zeros((B, 8))
    x0[:, :2] = q0
    x0[:, 2:4] = q0
    K = np.zeros((B, 2, 2))
    K[:, 0, 0] = k[:, 0]
    K[:, 1, 1] = k[:, 1]
    w.update(name="sea_batch_%d" % B, x0=x0, K_ocp=K, solver=solver, init="x0-zero-u")
    return w


def chain7_terms(target_p):
    running = _sorted([
        _placement("gripperPose", 1e-1, "tip", target_p),
        _term("xReg", "state", 1e-3, np.zeros(28), np.array([1.0] * 7 + [0] * 7 + [1.0] * 7 + [0] * 7)),
        _term("uReg", "control", 1e-2, np.zeros(7), np.ones(7))])
    terminal = [_placement("gripperPose", 1e4, "tip", target_p)]
    return running, terminal


def chain7_batch(B=1024, T=200, seed=20240002):
    """BASELINE.json configs[4]: synthetic 7-link SEA arm, K = 50 I, B = 0.05 I, per-OCP targets."""
    rng = np.random.default_rng(seed)
    q0 = rng.uniform(-1.0, 1.0, size=(B, 7))
    x0 = np.zeros((B, 28))
    x0[:, :7] = q0
    x0[:, 7:14] = q0
    targets = np.stack([rng.uniform(0.3, 0.9, B), rng.uniform(-0.5, 0.5, B), rng.uniform(0.6, 1.4, B)], axis=1)
    mref = np.zeros((B, 12))
    mref[:, :9] = np.eye(3).reshape(9)
    mref[:, 9:] = targets
    running, terminal = chain7_terms(targets[0])
    return dict(name="chain7_batch_%d" % B, robot="chain7", gravity=None, kind="sea", K=50.0 * np.eye(7),
                B=0.05 * np.eye(7), Sm=None, dt=1e-2, T=T, x0=x0, mref_ocp=mref, running=running, terminal=terminal,
                solver="fddp", th_stop=1e-7, maxiter=100, init="x0-zero-u", u_lb=None, u_ub=None)


def chain_batch(n, B=2, T=12, seed=20240004, kind="sea"):
    """The first n links of the 7-link arm (fixtures chain1 ... chain6, tools/make_robots.py) with chain7's cost
    structure: one workload per link DOF the kernels are instantiated for.  kind = "vsa": SolverBoxDDP with an
    initially active stiffness bound, as chain7_vsa_batch."""
    if n == 7:
        return chain7_batch(B, T, seed) if kind == "sea" else chain7_vsa_batch(B, T, seed)
    rng = np.random.default_rng(seed + n)
    q0 = rng.uniform(-1.0, 1.0, size=(B, n))
    x0 = np.zeros((B, 4 * n))
    x0[:, :n] = q0
    x0[:, n:2 * n] = q0
    reach = 0.3 * n
    targets = np.stack([rng.uniform(0.2, 0.6, B) * reach, rng.uniform(-0.4, 0.4, B) * reach,
                        rng.uniform(0.4, 0.9, B) * reach], axis=1)
    mref = np.zeros((B, 12))
    mref[:, :9] = np.eye(3).reshape(9)
    mref[:, 9:] = targets
    nu = n if kind == "sea" else 2 * n
    running = _sorted([
        _placement("gripperPose", 1e-1, "tip", targets[0]),
        _term("xReg", "state", 1e-3, np.zeros(4 * n), np.array([1.0] * n + [0] * n + [1.0] * n + [0] * n)),
        _term("uReg", "control", 1e-2, np.zeros(nu), np.ones(nu))])
    terminal = [_placement("gripperPose", 1e4, "tip", targets[0])]
    w = dict(name="chain%d_batch_%d" % (n, B), robot="chain%d" % n, gravity=None, kind=kind,
             K=50.0 * np.eye(n) if kind == "sea" else None, B=0.05 * np.eye(n), Sm=None, dt=1e-2, T=T, x0=x0,
             mref_ocp=mref, running=running, terminal=terminal, solver="fddp" if kind == "sea" else "boxddp",
             th_stop=1e-7, maxiter=100, init="x0-zero-u", u_lb=None, u_ub=None)
    if kind == "vsa":
        w.update(u_lb=np.array([-50.0] * n + [5.0] * n), u_ub=np.array([50.0] * n + [100.0] * n))
    return w


def chain7_vsa_batch(B=2, T=20, seed=20240003):
    """7-link variable-stiffness arm under SolverBoxDDP (nx = 28, nu = 14): the 7-DOF counterpart of
    examples/two_dof_vsa_boxddp.py, with the stiffness bound starting active (sigma_lb > initial guess)."""
    w = chain7_batch(B, T, seed)
    running = _sorted([t for t in w["running"] if t["kind"] != "control"] +
                      [_term("uReg", "control", 1e-2, np.zeros(14), np.ones(14))])
    return dict(w, name="chain7_vsa_batch_%d" % B, kind="vsa", K=None, running=running, solver="boxddp",
                u_lb=np.array([-50.0] * 7 + [5.0] * 7), u_ub=np.array([50.0] * 7 + [100.0] * 7))


# ------------------------------------------------------------------------------------------------
_KIND = {"placement": A.COST_FRAME_PLACEMENT, "state": A.COST_STATE, "control": A.COST_CONTROL,
         "stiffness": A.COST_STIFFNESS, "pendulum": A.COST_DOUBLE_PENDULUM,
         "pendulum_residual": A.COST_RESIDUAL_DOUBLE_PENDULUM}


def frame_of(robot_dict, name):
    for i, f in enumerate(robot_dict["frames"]):
        if f["name"] == name:
            return i, f
    raise KeyError(name)


def _abi_terms(robot_dict, terms):
    out = []
    for t in terms:
        if t["kind"] == "placement":
            _, f = frame_of(robot_dict, t["frame"])
            out.append(D.cost_term(A.COST_FRAME_PLACEMENT, t["weight"], t["ref"], t["act"], f["parent"],
                                   f["placement"]["R"], f["placement"]["p"]))
        else:
            out.append(D.cost_term(_KIND[t["kind"]], t["weight"], t["ref"], t["act"]))
    return out


def to_desc(w):
    """workload dict -> AslrProblemDesc."""
    rd = D.load_robot_dict(w["robot"])
    n = len(rd["joints"])
    kind = A.DYN_SEA if w["kind"] == "sea" else A.DYN_VSA
    dyn = D.make_dynamics(kind, n, w["B"], w["K"], w.get("Sm"), True, w.get("u_lb"), w.get("u_ub"))
    return D.make_problem_desc(D.make_robot(rd, w["gravity"]), dyn, _abi_terms(rd, w["running"]),
                               _abi_terms(rd, w["terminal"]), w["dt"], 0.0)


def initial_guess(w, x0, T, nu):
    """(xs, us) arrays the examples hand to solver.solve: `empty` = ([], []) i.e. state.zero() / zeros
    (examples/two_dof_vsa_boxddp.py:81); otherwise xs = [x0] * (T + 1) and us = quasiStatic(x0), which is
    zero for every example because q_l = q_m at x0."""
    x0 = np.atleast_2d(np.asarray(x0, float))
    B, nx = x0.shape
    if w["init"] == "empty":
        return np.zeros((B, T + 1, nx)), np.zeros((B, T, nu))
    return np.repeat(x0[:, None, :], T + 1, axis=1), np.zeros((B, T, nu))
