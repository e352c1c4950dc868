"""ORACLE / TEST INFRASTRUCTURE ONLY -- crocoddyl cost classes restated in NumPy
(CostModelAbstract / CostModelResidual / CostModelSum).
Call sites in the reference: examples/two_dof_sea.py:27-47, stiffness_cost.py:8,
python/aslr_to/__init__.py:226.  PARITY UNPINNED (no crocoddyl golden vectors exist)."""
import numpy as np

from .core import ActivationModelQuad, ResidualModelAbstract


class CostModelAbstract(object):
    """Overloads served: (state, activation, residual), (state, activation, nu=),
    (state, activation), (state, residual), (state, nr, nu), (state, nr)."""

    def __init__(self, state, a=None, b=None, nu=None):
        self.state = state
        activation = residual = None
        if isinstance(a, ResidualModelAbstract):
            residual = a
        elif np.ndim(a) == 0 and isinstance(a, (int, np.integer)):
            activation = ActivationModelQuad(int(a))
        else:
            activation = a
        if isinstance(b, ResidualModelAbstract):
            residual = b
        elif b is not None:
            nu = b
        if residual is None:
            residual = ResidualModelAbstract(state, activation.nr, state.nv if nu is None else int(nu))
        if activation is None:
            activation = ActivationModelQuad(residual.nr)
        if activation.nr != residual.nr:
            raise ValueError("nr is equals to %d" % residual.nr)
        self.activation = activation
        self.residual = residual
        self.nu = residual.nu

    def calc(self, data, x, u):
        raise NotImplementedError

    def calcDiff(self, data, x, u):
        raise NotImplementedError

    def createData(self, collector):
        return CostDataAbstract(self, collector)


class CostDataAbstract(object):
    def __init__(self, model, collector):
        ndx, nu = model.state.ndx, model.nu
        self.shared = collector
        self.activation = model.activation.createData()
        self.residual = model.residual.createData(collector)
        self.cost = 0.0
        self.Lx = np.zeros(ndx)
        self.Lu = np.zeros(nu)
        self.Lxx = np.zeros((ndx, ndx))
        self.Lxu = np.zeros((ndx, nu))
        self.Luu = np.zeros((nu, nu))

    @property
    def r(self):
        return self.residual.r

    @property
    def Rx(self):
        return self.residual.Rx

    @property
    def Ru(self):
        return self.residual.Ru


class CostModelResidual(CostModelAbstract):
    def __init__(self, state, a, b=None):
        CostModelAbstract.__init__(self, state, a, b)

    def calc(self, data, x, u):
        self.residual.calc(data.residual, x, u)
        self.activation.calc(data.activation, data.residual.r)
        data.cost = data.activation.a_value

    def calcDiff(self, data, x, u):
        """Gauss-Newton derivatives restricted to the blocks the residual's (q, v, u) dependency
        flags enable; the q block is the first state.nv columns (Appendix A.5)."""
        res, rd, ad = self.residual, data.residual, data.activation
        res.calcDiff(rd, x, u)
        self.activation.calcDiff(ad, rd.r)
        nv = self.state.nv
        w = np.diag(ad.Arr)
        is_ru = res.u_dependent and self.nu != 0
        if is_ru:
            data.Lu[:] = rd.Ru.T @ ad.Ar
            data.Luu[:, :] = rd.Ru.T @ (w[:, None] * rd.Ru)
        if res.q_dependent and res.v_dependent:
            data.Lx[:] = rd.Rx.T @ ad.Ar
            data.Lxx[:, :] = rd.Rx.T @ (w[:, None] * rd.Rx)
            if is_ru:
                data.Lxu[:, :] = rd.Rx.T @ (w[:, None] * rd.Ru)
        elif res.q_dependent:
            Rq = rd.Rx[:, :nv]
            data.Lx[:nv] = Rq.T @ ad.Ar
            data.Lxx[:nv, :nv] = Rq.T @ (w[:, None] * Rq)
            if is_ru:
                data.Lxu[:nv, :] = Rq.T @ (w[:, None] * rd.Ru)
        elif res.v_dependent:
            Rv = rd.Rx[:, -nv:]
            data.Lx[-nv:] = Rv.T @ ad.Ar
            data.Lxx[-nv:, -nv:] = Rv.T @ (w[:, None] * Rv)
            if is_ru:
                data.Lxu[-nv:, :] = Rv.T @ (w[:, None] * rd.Ru)


class _CostItem(object):
    def __init__(self, name, cost, weight, active=True):
        self.name, self.cost, self.weight, self.active = name, cost, float(weight), active


class CostModelSum(object):
    def __init__(self, state, nu=None):
        self.state = state
        self.nu = int(state.nv if nu is None else nu)
        self.costs = {}
        self._order = []
        self.nr = 0

    def addCost(self, name, cost, weight, active=True):
        if name in self.costs:
            raise ValueError("cost %s already exists" % name)
        if cost.nu != self.nu:
            raise ValueError("%s cost item doesn't have the same control dimension (it should be %d)" % (name, self.nu))
        self.costs[name] = _CostItem(name, cost, weight, active)
        # crocoddyl keeps a std::map: iteration is by sorted name
        self._order = sorted(self.costs)
        if active:
            self.nr += cost.activation.nr

    def calc(self, data, x, u):
        data.cost = 0.0
        for name in self._order:
            it = self.costs[name]
            if it.active:
                d = data.costs[name]
                it.cost.calc(d, x, u)
                data.cost += it.weight * d.cost

    def calcDiff(self, data, x, u):
        data.Lx[:] = 0
        data.Lu[:] = 0
        data.Lxx[:, :] = 0
        data.Lxu[:, :] = 0
        data.Luu[:, :] = 0
        for name in self._order:
            it = self.costs[name]
            if it.active:
                d = data.costs[name]
                it.cost.calcDiff(d, x, u)
                data.Lx += it.weight * d.Lx
                data.Lu += it.weight * d.Lu
                data.Lxx += it.weight * d.Lxx
                data.Lxu += it.weight * d.Lxu
                data.Luu += it.weight * d.Luu

    def createData(self, collector):
        return CostDataSum(self, collector)


class CostDataSum(object):
    def __init__(self, model, collector):
        ndx, nu = model.state.ndx, model.nu
        self.shared = collector
        self.costs = dict((n, it.cost.createData(collector)) for n, it in model.costs.items())
        self.cost = 0.0
        self.Lx = np.zeros(ndx)
        self.Lu = np.zeros(nu)
        self.Lxx = np.zeros((ndx, ndx))
        self.Lxu = np.zeros((ndx, nu))
        self.Luu = np.zeros((nu, nu))

    def shareMemory(self, data):
        self.Lx, self.Lu, self.Lxx, self.Lxu, self.Luu = data.Lx, data.Lu, data.Lxx, data.Lxu, data.Luu


class DifferentialActionModelNumDiff(object):
    """Forward-difference wrapper used by the reference's unit tests
    (unittest/test_stiffness_residual.py:40-46)."""

    def __init__(self, model, with_gauss_approx=False):
        self.model = model
        self.state, self.nu, self.nr = model.state, model.nu, model.nr
        self.disturbance = float(np.sqrt(2.0 * np.finfo(float).eps))

    def createData(self):
        return _NumDiffData(self)

    def calc(self, data, x, u):
        self.model.calc(data.data_0, x, u)
        data.cost = data.data_0.cost
        data.xout = np.array(data.data_0.xout)

    def calcDiff(self, data, x, u):
        h = self.disturbance
        m, st = self.model, self.state
        x = np.asarray(x, float)
        u = np.asarray(u, float)
        c0, xo0 = data.data_0.cost, np.array(data.data_0.xout)
        d = m.createData()
        for i in range(st.ndx):
            dx = np.zeros(st.ndx)
            dx[i] = h
            m.calc(d, np.asarray(st.integrate(x, dx)).ravel(), u)
            data.Fx[:, i] = (np.array(d.xout) - xo0) / h
            data.Lx[i] = (d.cost - c0) / h
        for i in range(self.nu):
            du = np.zeros(self.nu)
            du[i] = h
            m.calc(d, x, u + du)
            data.Fu[:, i] = (np.array(d.xout) - xo0) / h
            data.Lu[i] = (d.cost - c0) / h


class _NumDiffData(object):
    def __init__(self, nd):
        m = nd.model
        nv, ndx, nu = m.state.nv, m.state.ndx, m.nu
        self.data_0 = m.createData()
        self.cost = 0.0
        self.xout = np.zeros(nv)
        self.Fx = np.zeros((nv, ndx))
        self.Fu = np.zeros((nv, nu))
        self.Lx = np.zeros(ndx)
        self.Lu = np.zeros(nu)
