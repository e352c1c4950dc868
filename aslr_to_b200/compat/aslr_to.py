"""The reference package's public classes (python/aslr_to/__init__.py:1-9), same names and constructor
signatures, as parameter holders whose calc / calcDiff run the CUDA kernels (B = 1, one knot).

reference class (file:line)                                          -> what the kernels read from it
  StateMultibodyASR(pinModel)            statemultibody_aslr.py:13-16  -> robot, nx = 4 nv
  ASRActuation / VSAASRActuation(state)  actuation_asr.py:5-13         -> motor map Sm = I
  ActuationModelDoublePendulum(state, actLink, nu)  __init__.py:262-289 -> one-hot Sm
  DifferentialFreeASRFwdDynamicsModel(state, actuation, costs, K, B)   free_fwddyn_asr.py:6-19
  DifferentialFreeFwdDynamicsModelVSA(state, actuation, costs, B)      free_fwddyn_vsa.py:6-19
  IntegratedActionModelEulerASR(diff, timeStep, withCostResiduals)     integrated_action.py:6-12
  ResidualModelFramePlacementASR(state, id, placement, nu)             residual_frame_placement.py:7-11
  CostModelStiffness(state, nu, lamda, Kref)                           stiffness_cost.py:6-11
  CostModelDoublePendulum(state, activation, nu)                       __init__.py:223-228
"""
import types

import numpy as np

from .. import _abi as A
from . import pinocchio
from . import _engine as E


class StateMultibodyASR(object):
    """statemultibody_aslr.py:13-109.  x = [q_l, q_m, v_l, v_m]; the manifold operations are the
    vector-space ones because every supported joint is revolute (nq == nv, statemultibody_aslr.py:52)."""

    def __init__(self, pinocchioModel):
        self.pinocchio = pinocchioModel
        self.nq = 2 * pinocchioModel.nq
        self.nv = 2 * pinocchioModel.nv
        self.nx = self.nq + self.nv
        self.ndx = 2 * self.nv
        self.nv_m = self.nq_m = pinocchioModel.nv
        self.nv_l = self.nq_l = pinocchioModel.nv
        self.lb = -np.inf * np.ones(self.nx)
        self.ub = np.inf * np.ones(self.nx)

    def zero(self):
        return np.zeros(self.nx)

    def rand(self):
        n = self.pinocchio.nq
        q = pinocchio.randomConfiguration(self.pinocchio)
        return np.concatenate([q, pinocchio.utils.rand(n), pinocchio.utils.rand(2 * n)])

    def diff(self, x0, x1):
        return np.asarray(x1, float) - np.asarray(x0, float)

    def integrate(self, x, dx):
        return np.asarray(x, float) + np.asarray(dx, float)

    def Jdiff(self, x1, x2, firstsecond="both"):
        I = np.eye(self.ndx)
        if firstsecond == "first":
            return -I
        if firstsecond == "second":
            return I
        return [-I, I]

    def Jintegrate(self, x, dx, firstsecond="both"):
        I = np.eye(self.ndx)
        return I if firstsecond in ("first", "second") else [I, I.copy()]


# ------------------------------------------------------------------------------------ actuation
class _Actuation(object):
    def __init__(self, state, nu):
        self.state, self.nu = state, int(nu)

    def _Sm(self):
        return np.eye(self.state.pinocchio.nv)

    def createData(self):
        n2 = self.state.nv
        return types.SimpleNamespace(tau=np.zeros(n2), dtau_dx=np.zeros((n2, self.state.ndx)),
                                     dtau_du=np.zeros((n2, self.nu)))

    def calc(self, data, x, u):
        n = self.state.pinocchio.nv
        data.tau[:] = 0.0
        data.tau[n:] = self._Sm() @ np.asarray(u, float)[:n]

    def calcDiff(self, data, x, u):
        n = self.state.pinocchio.nv
        data.dtau_du[:] = 0.0
        data.dtau_du[n:, :n] = self._Sm()


class ASRActuation(_Actuation):
    """actuation_asr.py:5-13: tau[n:] = u."""

    def __init__(self, state):
        _Actuation.__init__(self, state, state.pinocchio.nv)


class VSAASRActuation(_Actuation):
    """actuation_vsa.py:5-13: tau[n:] = u[:n]; the VSA differential model doubles nu (free_fwddyn_vsa.py:8)."""

    def __init__(self, state):
        _Actuation.__init__(self, state, state.pinocchio.nv)


class ActuationModelDoublePendulum(_Actuation):
    """__init__.py:262-289: S has a single 1 at S[n, 0] (actLink != 1) or S[-1, -1] (actLink == 1)."""

    def __init__(self, state, actLink, nu):
        _Actuation.__init__(self, state, nu)
        self.actLink = actLink

    def _Sm(self):
        n = self.state.pinocchio.nv
        S = np.zeros((n, n))
        if self.actLink == 1:
            S[-1, -1] = 1.0
        else:
            S[0, 0] = 1.0
        return S


# ------------------------------------------------------------------------------------ residuals / costs
class ResidualModelFramePlacementASR(object):
    """residual_frame_placement.py:7-24."""

    def __init__(self, state, id, placement, nu):
        self.state, self.id, self.reference, self.nu, self.nr = state, int(id), placement, int(nu), 6
        self.q_dependent, self.v_dependent, self.u_dependent = True, False, False


class ResidualModelDoublePendulum(object):
    """residual_acrobot.py:5-29: r = [s1, s2, 1 + c1, 1 - c2, v1, v2] (q- and v-dependent, not u-dependent);
    used through crocoddyl.CostModelResidual(state, activation, residual)."""

    def __init__(self, state, nu):
        if state.pinocchio.nv != 2:
            raise ValueError("ResidualModelDoublePendulum indexes x[4], x[5] as the link velocities: 2 link DOF only")
        self.state, self.nu, self.nr = state, int(nu), 6
        self.q_dependent, self.v_dependent, self.u_dependent = True, True, False


class CostModelStiffness(object):
    """stiffness_cost.py:6-20: cost = sum(lamda (sigma - Kref)), linear in the stiffness controls."""

    def __init__(self, state, nu, lamda, Kref):
        self.state, self.nu, self.lamda = state, int(nu), float(lamda)
        self.Kref = np.asarray(Kref, float).ravel()
        self.nr = self.nu // 2

    def _term(self, weight):
        return E.term("stiffness", weight, ref=self.Kref, act=[self.lamda])


class CostModelDoublePendulum(object):
    """__init__.py:223-259."""

    def __init__(self, state, activation, nu):
        self.state, self.activation, self.nu, self.nr = state, activation, int(nu), 6

    def _term(self, weight):
        return E.term("pendulum", weight, act=self.activation.weights)


def u_squared(log):
    """__init__.py:63-68: per-control sum over the horizon of u^2."""
    out = np.zeros(len(log.us[0]))
    for u in log.us:
        out += np.asarray(u, float) ** 2
    return out


# ------------------------------------------------------------------------------------ differential models
class _DifferentialBase(object):
    def __init__(self, state, actuation, costs, nu, B):
        n = state.pinocchio.nv
        self.state, self.actuation, self.costs = state, actuation, costs
        self.nu, self.nr = int(nu), getattr(costs, "nr", 0)
        self.B = 1e-3 * np.eye(n) if B is None else np.asarray(B, float).reshape(n, n)
        self.u_lb = -np.inf * np.ones(self.nu)
        self.u_ub = np.inf * np.ones(self.nu)
        self.enable_force = True

    def createData(self):
        n, ndx, nu = self.state.pinocchio.nv, self.state.ndx, self.nu
        return types.SimpleNamespace(cost=0.0, xout=np.zeros(2 * n), Fx=np.zeros((2 * n, ndx)), Fu=np.zeros((2 * n, nu)),
                                     Lx=np.zeros(ndx), Lu=np.zeros(nu), Lxx=np.zeros((ndx, ndx)),
                                     Lxu=np.zeros((ndx, nu)), Luu=np.zeros((nu, nu)), r=np.zeros(self.nr))

    def calc(self, data, x, u=None):
        out = E.eval_knot(E.ModelSpec(self, 0.0), x, u, diff=False)
        data.xout[:], data.cost = out["xout"], out["cost"]

    def calcDiff(self, data, x, u=None):
        out = E.eval_knot(E.ModelSpec(self, 0.0), x, u, diff=True)
        data.xout[:], data.cost = out["xout"], out["cost"]
        data.Fx[:], data.Fu[:] = out["dFx"], out["dFu"]
        for k in ("Lx", "Lu", "Lxx", "Lxu", "Luu"):
            getattr(data, k)[...] = out[k]


class DifferentialFreeASRFwdDynamicsModel(_DifferentialBase):
    """free_fwddyn_asr.py:6-92 (series-elastic actuation, constant stiffness K)."""
    _kind = A.DYN_SEA

    def __init__(self, state, actuation, costs, K=None, B=None):
        n = state.pinocchio.nv
        _DifferentialBase.__init__(self, state, actuation, costs, actuation.nu, B)
        self.K = 1e-1 * np.eye(n) if K is None else np.asarray(K, float).reshape(n, n)  # free_fwddyn_asr.py:12-15

    def _write_fu(self):
        return self.actuation.nu > 1  # free_fwddyn_asr.py:88


class DifferentialFreeFwdDynamicsModelVSA(_DifferentialBase):
    """free_fwddyn_vsa.py:6-94 (variable stiffness: u = [tau_m, sigma], K = diag(sigma))."""
    _kind = A.DYN_VSA

    def __init__(self, state, actuation, costs, B=None):
        _DifferentialBase.__init__(self, state, actuation, costs, 2 * actuation.nu, B)  # free_fwddyn_vsa.py:8
        self.K = None

    def _write_fu(self):
        return True


# ------------------------------------------------------------------------------------ action model
class IntegratedActionModelEulerASR(object):
    """integrated_action.py:6-42: semi-implicit Euler, cost NOT scaled by dt."""

    def __init__(self, diffModel, timeStep=1e-3, withCostResiduals=True):
        self.differential = diffModel
        self.state = diffModel.state
        self.nu, self.nr = diffModel.nu, diffModel.nr
        self.timeStep = float(timeStep)
        self.withCostResiduals = withCostResiduals
        self.enable_integration_ = self.timeStep != 0.0

    # crocoddyl exposes the control bounds on the action model (examples/two_dof_vsa_boxddp.py:59-60)
    @property
    def u_lb(self):
        return self.differential.u_lb

    @u_lb.setter
    def u_lb(self, v):
        self.differential.u_lb = np.asarray(v, float).ravel()

    @property
    def u_ub(self):
        return self.differential.u_ub

    @u_ub.setter
    def u_ub(self, v):
        self.differential.u_ub = np.asarray(v, float).ravel()

    def createData(self):
        ndx, nu = self.state.ndx, self.nu
        return types.SimpleNamespace(cost=0.0, xnext=np.zeros(self.state.nx), Fx=np.zeros((ndx, ndx)),
                                     Fu=np.zeros((ndx, nu)), Lx=np.zeros(ndx), Lu=np.zeros(nu), Lxx=np.zeros((ndx, ndx)),
                                     Lxu=np.zeros((ndx, nu)), Luu=np.zeros((nu, nu)), r=np.zeros(self.nr),
                                     differential=self.differential.createData())

    def calc(self, data, x, u=None):
        out = E.eval_knot(E.ModelSpec(self.differential, self.timeStep), x, u, diff=False)
        data.xnext[:], data.cost = out["xnext"], out["cost"]
        data.differential.xout[:], data.differential.cost = out["xout"], out["cost"]

    def calcDiff(self, data, x, u=None):
        out = E.eval_knot(E.ModelSpec(self.differential, self.timeStep), x, u, diff=True)
        data.xnext[:], data.cost = out["xnext"], out["cost"]
        for k in ("Fx", "Fu", "Lx", "Lu", "Lxx", "Lxu", "Luu"):
            getattr(data, k)[...] = out[k]
        d = data.differential
        d.xout[:], d.cost, d.Fx[:], d.Fu[:] = out["xout"], out["cost"], out["dFx"], out["dFu"]

    def quasiStatic(self, data, x, maxiter=100, tol=1e-9):
        return E.quasi_static(E.ModelSpec(self.differential, self.timeStep), np.asarray(x, float)[None, :], maxiter, tol)[0]
