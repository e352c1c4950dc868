"""Builds the ORACLE side of a workload (aslr_to_b200/workloads.py): either the port
(oracle/aslr_port.py, available everywhere) or the reference's own unmodified classes on the
NumPy shims (only where /root/reference exists).  Test infrastructure only."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from oracle import refenv  # noqa: E402

refenv.activate_shims()
import crocoddyl  # noqa: E402
import example_robot_data  # noqa: E402
import pinocchio  # noqa: E402

from oracle import aslr_port as port  # noqa: E402

_PORT_KIND = {"placement": port.FRAME_PLACEMENT, "state": port.STATE, "control": port.CONTROL,
              "stiffness": port.STIFFNESS, "pendulum": port.DOUBLE_PENDULUM,
              "pendulum_residual": port.RESIDUAL_DOUBLE_PENDULUM}
SOLVERS = {"ddp": crocoddyl.SolverDDP, "fddp": crocoddyl.SolverFDDP, "boxddp": crocoddyl.SolverBoxDDP}


def robot_model(w):
    m = example_robot_data.load(w["robot"]).model
    if w.get("gravity") is not None:
        m.gravity.linear = np.array(w["gravity"], float)
    return m


def per_ocp(w, b):
    """(x0, K, mref) of OCP b (handles both single-OCP and batched workloads)."""
    x0 = np.asarray(w["x0"], float)
    x0 = x0[b] if x0.ndim == 2 else x0
    K = w.get("K_ocp")
    K = np.asarray(K[b], float) if K is not None else w["K"]
    mref = w.get("mref_ocp")
    mref = np.asarray(mref[b], float) if mref is not None else None
    return x0, K, mref


def _port_costs(rm, terms, mref):
    out = []
    for t in terms:
        if t["kind"] == "placement":
            ref = mref if mref is not None else t["ref"]
            out.append(port.cost(port.FRAME_PLACEMENT, t["weight"], ref, t["act"], rm.getFrameId(t["frame"])))
        else:
            out.append(port.cost(_PORT_KIND[t["kind"]], t["weight"], t["ref"], t["act"]))
    return out


def port_models(w, b=0):
    rm = robot_model(w)
    x0, K, mref = per_ocp(w, b)
    kind = port.SEA if w["kind"] == "sea" else port.VSA
    state = port.PortState(rm)
    run = port.PortActionModel(port.Spec(rm, kind, w["B"], K, w.get("Sm"), True, _port_costs(rm, w["running"], mref),
                                         w["dt"]), state)
    term = port.PortActionModel(port.Spec(rm, kind, w["B"], K, w.get("Sm"), True,
                                          _port_costs(rm, w["terminal"], mref), 0.0), state)
    if w.get("u_lb") is not None:
        run.u_lb = np.asarray(w["u_lb"], float)
        run.u_ub = np.asarray(w["u_ub"], float)
    return run, term, x0


def reference_models(w, b=0):
    """The reference's own classes (unmodified, /root/reference/python) on the shims."""
    aslr_to = refenv.import_reference()
    rm = robot_model(w)
    x0, K, mref = per_ocp(w, b)
    n = rm.nv
    state = aslr_to.StateMultibodyASR(rm)
    vsa = w["kind"] == "vsa"
    if w.get("Sm") is not None:
        act = aslr_to.ActuationModelDoublePendulum(state, 0, n)  # actLink=0, nu=2
    else:
        act = aslr_to.VSAASRActuation(state) if vsa else aslr_to.ASRActuation(state)
    nu = 2 * act.nu if vsa else act.nu

    def cost_sum(terms):
        cs = crocoddyl.CostModelSum(state, nu)
        for t in terms:
            if t["kind"] == "placement":
                ref = mref if mref is not None else t["ref"]
                res = aslr_to.ResidualModelFramePlacementASR(state, rm.getFrameId(t["frame"]),
                                                             pinocchio.SE3(ref[:9].reshape(3, 3), ref[9:12]), nu)
                c = crocoddyl.CostModelResidual(state, crocoddyl.ActivationModelWeightedQuad(t["act"]), res)
            elif t["kind"] == "state":
                c = crocoddyl.CostModelResidual(state, crocoddyl.ActivationModelWeightedQuad(t["act"]),
                                                crocoddyl.ResidualModelState(state, t["ref"], nu))
            elif t["kind"] == "control":
                assert np.all(t["ref"] == 0)
                c = crocoddyl.CostModelResidual(state, crocoddyl.ActivationModelWeightedQuad(t["act"]),
                                                crocoddyl.ResidualModelControl(state, nu))
            elif t["kind"] == "stiffness":
                c = aslr_to.CostModelStiffness(state, nu, t["act"][0], t["ref"])
            elif t["kind"] == "pendulum":
                c = aslr_to.CostModelDoublePendulum(state, crocoddyl.ActivationModelWeightedQuad(t["act"]), nu)
            elif t["kind"] == "pendulum_residual":
                c = crocoddyl.CostModelResidual(state, crocoddyl.ActivationModelWeightedQuad(t["act"]),
                                                aslr_to.ResidualModelDoublePendulum(state, nu))
            cs.addCost(t["name"], c, t["weight"])
        return cs

    def diff_model(costs):
        if vsa:
            return aslr_to.DifferentialFreeFwdDynamicsModelVSA(state, act, costs, w["B"])
        return aslr_to.DifferentialFreeASRFwdDynamicsModel(state, act, costs, K, w["B"])

    run = aslr_to.IntegratedActionModelEulerASR(diff_model(cost_sum(w["running"])), w["dt"])
    term = aslr_to.IntegratedActionModelEulerASR(diff_model(cost_sum(w["terminal"])), 0)
    if w.get("u_lb") is not None:
        run.u_lb = np.asarray(w["u_lb"], float)
        run.u_ub = np.asarray(w["u_ub"], float)
    return run, term, x0


def initial_guess(w, problem, x0):
    """(xs, us) exactly as the example scripts pass them to solver.solve."""
    T = problem.T
    if w["init"] == "empty":
        return [], []
    if w["init"] == "quasistatic":
        return [x0] * (T + 1), problem.quasiStatic([x0] * T)
    if w["init"] == "x0-zero-u":
        return [x0] * (T + 1), [np.zeros(problem.runningModels[0].nu)] * T
    raise ValueError(w["init"])


def solve(w, b=0, which="port", maxiter=None, T=None):
    """Runs the oracle solver on OCP b of workload w.  Returns (solver, log)."""
    run, term, x0 = (port_models if which == "port" else reference_models)(w, b)
    T = w["T"] if T is None else T
    problem = crocoddyl.ShootingProblem(x0, [run] * T, term)
    solver = SOLVERS[w["solver"]](problem)
    log = crocoddyl.CallbackLogger()
    solver.setCallbacks([log])
    if w["th_stop"] is not None:
        solver.th_stop = w["th_stop"]
    xs, us = initial_guess(w, problem, x0)
    solver.converged = solver.solve(xs, us, w["maxiter"] if maxiter is None else maxiter)
    return solver, log
