"""ORACLE / TEST INFRASTRUCTURE ONLY -- NumPy restatement of crocoddyl's model-side classes
(states, actuation, action models, costs, activations, residuals, NumDiff).

crocoddyl (https://github.com/loco-3d/crocoddyl) is an un-vendored, un-pinned dependency of the
reference (/root/reference/CMakeLists.txt:22); forensic evidence puts the author's build at
`devel` between v1.9.0 and v2.0.0.  The classes below restate that
snapshot's published behaviour for exactly the API surface the reference touches, so that
/root/reference/python/aslr_to/*.py can be imported UNMODIFIED on top of them.
PARITY UNPINNED for this file: the reference holds no golden vectors for crocoddyl itself.
"""
import numpy as np


class Jcomponent(object):
    both = 0
    first = 1
    second = 2


# ------------------------------------------------------------------------------------------ state
class StateAbstract(object):
    def __init__(self, nx, ndx):
        self.nx = int(nx)
        self.ndx = int(ndx)
        self.nv = self.ndx // 2
        self.nq = self.nx - self.nv
        self.lb = -np.inf * np.ones(self.nx)
        self.ub = np.inf * np.ones(self.nx)
        self.has_limits = False


# -------------------------------------------------------------------------------------- actuation
class ActuationModelAbstract(object):
    def __init__(self, state, nu):
        self.state = state
        self.nu = int(nu)

    def createData(self):
        return ActuationDataAbstract(self)


class ActuationDataAbstract(object):
    def __init__(self, model):
        nv, ndx, nu = model.state.nv, model.state.ndx, model.nu
        self.tau = np.zeros(nv)
        self.dtau_dx = np.zeros((nv, ndx))
        self.dtau_du = np.zeros((nv, nu))


# ---------------------------------------------------------------------------------- action models
class _ModelBase(object):
    def _init_base(self, state, nu, nr):
        self.state = state
        self.nu = int(nu)
        self.nr = int(nr)
        self.unone = np.zeros(self.nu)
        self._u_lb = -np.inf * np.ones(self.nu)
        self._u_ub = np.inf * np.ones(self.nu)
        self.has_control_limits = False

    def _upd(self):
        self.has_control_limits = bool(np.all(np.isfinite(self._u_lb)) and np.all(np.isfinite(self._u_ub)))

    @property
    def u_lb(self):
        return self._u_lb

    @u_lb.setter
    def u_lb(self, v):
        v = np.asarray(v, float)
        if v.shape != (self.nu,):
            raise ValueError("lower bound has wrong dimension (it should be %d)" % self.nu)
        self._u_lb = v
        self._upd()

    @property
    def u_ub(self):
        return self._u_ub

    @u_ub.setter
    def u_ub(self, v):
        v = np.asarray(v, float)
        if v.shape != (self.nu,):
            raise ValueError("upper bound has wrong dimension (it should be %d)" % self.nu)
        self._u_ub = v
        self._upd()


class DifferentialActionModelAbstract(_ModelBase):
    def __init__(self, state, nu, nr=0):
        self._init_base(state, nu, nr)

    def createData(self):
        return DifferentialActionDataAbstract(self)


class DifferentialActionDataAbstract(object):
    def __init__(self, model):
        nv, ndx, nu, nr = model.state.nv, model.state.ndx, model.nu, model.nr
        self.cost = 0.0
        self.xout = np.zeros(nv)
        self.Fx = np.zeros((nv, ndx))
        self.Fu = np.zeros((nv, nu))
        self.r = np.zeros(nr)
        self.Lx = np.zeros(ndx)
        self.Lu = np.zeros(nu)
        self.Lxx = np.zeros((ndx, ndx))
        self.Lxu = np.zeros((ndx, nu))
        self.Luu = np.zeros((nu, nu))


class ActionModelAbstract(_ModelBase):
    def __init__(self, state, nu, nr=0):
        self._init_base(state, nu, nr)

    def createData(self):
        return ActionDataAbstract(self)

    def quasiStatic(self, data, x, maxiter=100, tol=1e-9):
        """Base-class Newton / pseudo-inverse iteration."""
        u = np.zeros(self.nu)
        if self.nu == 0:
            return u
        for _ in range(maxiter):
            self.calc(data, x, u)
            self.calcDiff(data, x, u)
            dx = np.asarray(self.state.diff(x, data.xnext)).ravel()
            du = -np.linalg.pinv(np.asarray(data.Fu)) @ dx
            u = u + du
            if np.linalg.norm(du) <= tol:
                break
        return u


class ActionDataAbstract(object):
    def __init__(self, model):
        ndx, nu, nr = model.state.ndx, model.nu, model.nr
        self.cost = 0.0
        self.xnext = np.array(model.state.zero(), float)
        self.Fx = np.zeros((ndx, ndx))
        self.Fu = np.zeros((ndx, nu))
        self.r = np.zeros(nr)
        self.Lx = np.zeros(ndx)
        self.Lu = np.zeros(nu)
        self.Lxx = np.zeros((ndx, ndx))
        self.Lxu = np.zeros((ndx, nu))
        self.Luu = np.zeros((nu, nu))


class DataCollectorMultibody(object):
    def __init__(self, pinocchio_data):
        self.pinocchio = pinocchio_data


# ------------------------------------------------------------------------------------ activations
class ActivationDataAbstract(object):
    def __init__(self, model):
        self.a_value = 0.0
        self.Ar = np.zeros(model.nr)
        self.Arr = np.zeros((model.nr, model.nr))


class ActivationModelQuad(object):
    def __init__(self, nr):
        self.nr = int(nr)

    def createData(self):
        d = ActivationDataAbstract(self)
        d.Arr = np.eye(self.nr)
        return d

    def calc(self, data, r):
        data.a_value = 0.5 * float(np.dot(r, r))

    def calcDiff(self, data, r):
        data.Ar = np.array(r, float)


class ActivationModelWeightedQuad(object):
    def __init__(self, weights):
        self.weights = np.array(weights, float)
        self.nr = self.weights.shape[0]

    def createData(self):
        d = ActivationDataAbstract(self)
        d.Arr = np.diag(self.weights)
        return d

    def calc(self, data, r):
        data.a_value = 0.5 * float(np.dot(self.weights * r, r))

    def calcDiff(self, data, r):
        data.Ar = self.weights * r
        data.Arr = np.diag(self.weights)


# -------------------------------------------------------------------------------------- residuals
class ResidualModelAbstract(object):
    def __init__(self, state, nr, nu=None, q_dependent=True, v_dependent=True, u_dependent=True):
        self.state = state
        self.nr = int(nr)
        self.nu = int(state.nv if nu is None else nu)
        self.q_dependent = q_dependent
        self.v_dependent = v_dependent
        self.u_dependent = u_dependent

    def calc(self, data, x, u):
        pass

    def calcDiff(self, data, x, u):
        pass

    def createData(self, collector):
        return ResidualDataAbstract(self, collector)


class ResidualDataAbstract(object):
    def __init__(self, model, collector):
        self.shared = collector
        self.r = np.zeros(model.nr)
        self.Rx = np.zeros((model.nr, model.state.ndx))
        self.Ru = np.zeros((model.nr, model.nu))


class ResidualModelState(ResidualModelAbstract):
    def __init__(self, state, xref=None, nu=None):
        ResidualModelAbstract.__init__(self, state, state.ndx, nu, True, True, False)
        self.reference = np.array(state.zero() if xref is None else xref, float)

    def calc(self, data, x, u):
        data.r[:] = np.asarray(self.state.diff(self.reference, x)).ravel()

    def calcDiff(self, data, x, u):
        data.Rx[:, :] = np.asarray(self.state.Jdiff(self.reference, x, Jcomponent.second))


class ResidualModelControl(ResidualModelAbstract):
    def __init__(self, state, uref_or_nu=None):
        if uref_or_nu is None:
            nu, uref = state.nv, None
        elif np.ndim(uref_or_nu) == 0:
            nu, uref = int(uref_or_nu), None
        else:
            uref = np.array(uref_or_nu, float)
            nu = uref.shape[0]
        ResidualModelAbstract.__init__(self, state, nu, nu, False, False, True)
        self.reference = np.zeros(nu) if uref is None else uref

    def calc(self, data, x, u):
        data.r[:] = u - self.reference

    def calcDiff(self, data, x, u):
        data.Ru[:, :] = np.eye(self.nu)
