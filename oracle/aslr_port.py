"""ORACLE / TEST INFRASTRUCTURE ONLY -- NumPy restatement of the reference's own model code
 for machines where /root/reference is absent (the GPU box).
Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may
import this; the product package never does.

Restated functions (file:line under /root/reference/python/aslr_to):
  sea_accelerations / sea_jacobians   free_fwddyn_asr.py:20-56 / :58-92
  vsa_accelerations / vsa_jacobians   free_fwddyn_vsa.py:20-57 / :59-94
  euler_step / euler_jacobians        integrated_action.py:13-26 / :28-42
  PortState                           statemultibody_aslr.py:13-109 (vector-space joints only)
  cost terms                          residual_frame_placement.py:13-24, stiffness_cost.py:13-20,
                                      __init__.py:229-249, residual_acrobot.py:9-21, + crocoddyl CostModelResidual over
                                      ResidualModelState / ResidualModelControl (Appendix A.5)
  motor-side actuation map Sm         actuation_asr.py:9-13, actuation_vsa.py:9-13, __init__.py:276-289
It is written as plain functions over one flat problem spec (not class-by-class like the
reference) and is PINNED against the reference itself: tests/test_oracle_pinning.py runs the
unmodified reference classes on the shims next to this port on seeded inputs (in the build
container, where /root/reference exists) and against the committed golden vectors (everywhere).
The rigid-body terms come from the pinocchio shim (oracle/shims/pinocchio); the solvers from the
crocoddyl shim -- both "parity unpinned" against the real libraries, see their headers.
"""
import numpy as np

from . import refenv

refenv.activate_shims()
import crocoddyl  # noqa: E402  (the NumPy shim)
import pinocchio  # noqa: E402  (the NumPy shim)

SEA, VSA = 0, 1
FRAME_PLACEMENT, STATE, CONTROL, STIFFNESS, DOUBLE_PENDULUM, RESIDUAL_DOUBLE_PENDULUM = 1, 2, 3, 4, 5, 6


class Spec(object):
    """Flat description of one action model (running or terminal)."""

    def __init__(self, robot, kind, B, K=None, Sm=None, write_fu=True, costs=(), dt=1e-2, u_none=None):
        self.robot = robot  # pinocchio-shim Model
        self.kind = kind
        n = robot.nv
        self.n = n
        self.nu = n if kind == SEA else 2 * n
        self.B = np.asarray(B, float).reshape(n, n)
        self.K = None if K is None else np.asarray(K, float).reshape(n, n)
        self.Sm = np.eye(n) if Sm is None else np.asarray(Sm, float).reshape(n, n)
        self.write_fu = write_fu
        self.costs = list(costs)  # dicts: kind, weight, ref, act, frame_id
        self.dt = float(dt)
        if u_none is None:
            u_none = np.zeros(self.nu)
            if kind == VSA:
                u_none[n:] = 3.0
        self.u_none = np.asarray(u_none, float)


def cost(kind, weight, ref=None, act=None, frame_id=None):
    return dict(kind=kind, weight=float(weight), ref=None if ref is None else np.asarray(ref, float),
                act=None if act is None else np.asarray(act, float), frame_id=frame_id)


# ------------------------------------------------------------------------------------- dynamics
def _split(spec, x):
    n = spec.n
    return x[:n], x[n:2 * n], x[2 * n:3 * n], x[3 * n:]


def accelerations(spec, pdata, x, u):
    """[a_l; a_m], Minv: free_fwddyn_asr.py:35-46 / free_fwddyn_vsa.py:34-47."""
    n = spec.n
    ql, qm, vl, vm = _split(spec, x)
    Kmat = spec.K if spec.kind == SEA else np.diag(u[n:])
    tau_m = spec.Sm @ u[:n]
    tau_c = Kmat @ (ql - qm)
    pinocchio.computeAllTerms(spec.robot, pdata, ql, vl)
    Minv = np.linalg.inv(pdata.M)
    Binv = np.linalg.inv(spec.B)
    a_l = Minv @ (-pdata.nle - tau_c)
    a_m = Binv @ (tau_m + tau_c)
    return np.concatenate([a_l, a_m]), Minv, Binv, Kmat


def acceleration_jacobians(spec, pdata, x, u, a, Minv, Binv, Kmat):
    """d[a_l; a_m]/dx (2n x 4n) and /du (2n x nu): free_fwddyn_asr.py:75-89, free_fwddyn_vsa.py:78-92."""
    n, nu = spec.n, spec.nu
    ql, qm, vl, vm = _split(spec, x)
    pinocchio.computeRNEADerivatives(spec.robot, pdata, ql, vl, a[:n])
    Ax = np.zeros((2 * n, 4 * n))
    Au = np.zeros((2 * n, nu))
    Ax[:n, :n] = Minv @ (-pdata.dtau_dq - Kmat)
    Ax[:n, n:2 * n] = Minv @ Kmat
    Ax[:n, 2 * n:3 * n] = -Minv @ pdata.dtau_dv
    Ax[n:, :n] = Binv @ Kmat
    Ax[n:, n:2 * n] = -Binv @ Kmat
    if spec.kind == SEA:
        if spec.write_fu:
            Au[n:, :] = Binv @ spec.Sm
    else:
        Au[:n, n:] = Minv * (qm - ql)[None, :]
        Au[n:, n:] = Binv * (ql - qm)[None, :]
        Au[n:, :n] = Binv @ spec.Sm
    return Ax, Au


# ---------------------------------------------------------------------------------------- costs
def cost_value_and_derivatives(spec, pdata, x, u, diff):
    n, nx, nu = spec.n, 4 * spec.n, spec.nu
    total = 0.0
    Lx, Lu = np.zeros(nx), np.zeros(nu)
    Lxx, Lxu, Luu = np.zeros((nx, nx)), np.zeros((nx, nu)), np.zeros((nu, nu))
    for c in spec.costs:
        w, kind = c["weight"], c["kind"]
        if kind == FRAME_PLACEMENT:
            fid = c["frame_id"]
            Mref = pinocchio.SE3(c["ref"][:9].reshape(3, 3), c["ref"][9:12])
            pinocchio.forwardKinematics(spec.robot, pdata, x[:n])
            pinocchio.updateFramePlacements(spec.robot, pdata)
            rMf = Mref.inverse() * pdata.oMf[fid]
            r = pinocchio.log(rMf).vector
            act = c["act"]
            total += w * 0.5 * float(np.dot(act * r, r))
            if diff:
                J = pinocchio.Jlog6(rMf) @ pinocchio.getFrameJacobian(spec.robot, pdata, fid,
                                                                      pinocchio.ReferenceFrame.LOCAL)
                Lx[:n] += w * (J.T @ (act * r))
                Lxx[:n, :n] += w * (J.T @ (act[:, None] * J))
        elif kind == STATE:
            r = x - c["ref"]
            total += w * 0.5 * float(np.dot(c["act"] * r, r))
            if diff:
                Lx += w * c["act"] * r
                Lxx[range(nx), range(nx)] += w * c["act"]
        elif kind == CONTROL:
            r = u - c["ref"]
            total += w * 0.5 * float(np.dot(c["act"] * r, r))
            if diff:
                Lu += w * c["act"] * r
                Luu[range(nu), range(nu)] += w * c["act"]
        elif kind == STIFFNESS:
            lam = c["act"][0]
            total += w * float(np.sum(lam * (u[n:] - c["ref"][:n])))
            if diff:
                Lu[n:] += w * lam
        elif kind == DOUBLE_PENDULUM:
            s1, s2, c1, c2 = np.sin(x[0]), np.sin(x[1]), np.cos(x[0]), np.cos(x[1])
            a = c["act"]
            r = np.array([s1, s2, 1 + c1, 1 + c2, x[4], x[5]])
            total += w * 0.5 * float(np.dot(a * r, r))
            if diff:
                Ar = a * r
                Lx[0] += w * (c1 * Ar[0] - s1 * Ar[2])
                Lx[1] += w * (c2 * Ar[1] - s2 * Ar[3])
                Lx[4] += w * Ar[4]
                Lx[5] += w * Ar[5]
                Lxx[0, 0] += w * ((c1 ** 2 - s1 ** 2) * a[0] + (s1 ** 2 + (1 - c1) * c1) * a[2])
                Lxx[1, 1] += w * ((c2 ** 2 - s2 ** 2) * a[1] + (s2 ** 2 + (1 - c2) * c2) * a[3])
                Lxx[4, 4] += w * a[4]
                Lxx[5, 5] += w * a[5]
        elif kind == RESIDUAL_DOUBLE_PENDULUM:
            # CostModelResidual(activation, ResidualModelDoublePendulum): residual_acrobot.py:9-21
            s1, s2, c1, c2 = np.sin(x[0]), np.sin(x[1]), np.cos(x[0]), np.cos(x[1])
            a = c["act"]
            r = np.array([s1, s2, 1 + c1, 1 - c2, x[4], x[5]])
            total += w * 0.5 * float(np.dot(a * r, r))
            if diff:
                Rx = np.zeros((6, nx))
                Rx[:2, :2] = np.diag([c1, c2])
                Rx[2:4, :2] = np.diag([-s1, s2])
                Rx[4:6, 4:6] = np.eye(2)
                Lx += w * (Rx.T @ (a * r))
                Lxx += w * (Rx.T @ (a[:, None] * Rx))
        else:
            raise ValueError("unknown cost kind %r" % kind)
    return total, Lx, Lu, Lxx, Lxu, Luu


# ------------------------------------------------------------------------------ integrated model
class PortState(crocoddyl.StateAbstract):
    """x = [q_l, q_m, v_l, v_m] on a vector space (revolute chains): statemultibody_aslr.py."""

    def __init__(self, robot):
        crocoddyl.StateAbstract.__init__(self, 4 * robot.nv, 4 * robot.nv)
        self.pinocchio = robot

    def zero(self):
        return np.zeros(self.nx)

    def diff(self, x0, x1):
        return np.asarray(x1, float) - np.asarray(x0, float)

    def integrate(self, x, dx):
        return np.asarray(x, float) + np.asarray(dx, float)

    def Jdiff(self, x0, x1, firstsecond=crocoddyl.Jcomponent.both):
        I = np.eye(self.ndx)
        if firstsecond == crocoddyl.Jcomponent.first:
            return -I
        if firstsecond == crocoddyl.Jcomponent.second:
            return I
        return [-I, I]

    def Jintegrate(self, x, dx, firstsecond=crocoddyl.Jcomponent.both):
        I = np.eye(self.ndx)
        return [I, I.copy()] if firstsecond == crocoddyl.Jcomponent.both else I


class PortActionData(crocoddyl.ActionDataAbstract):
    def __init__(self, model):
        crocoddyl.ActionDataAbstract.__init__(self, model)
        self.pinocchio = model.spec.robot.createData()
        self.xout = np.zeros(2 * model.spec.n)
        self.dFx = np.zeros((2 * model.spec.n, 4 * model.spec.n))
        self.dFu = np.zeros((2 * model.spec.n, model.nu))


class PortActionModel(crocoddyl.ActionModelAbstract):
    """IntegratedActionModelEulerASR(Differential{ASR,VSA}...) folded into one object."""

    def __init__(self, spec, state=None):
        self.spec = spec
        crocoddyl.ActionModelAbstract.__init__(self, state or PortState(spec.robot), spec.nu, 0)

    def createData(self):
        return PortActionData(self)

    def calc(self, data, x, u=None):
        s = self.spec
        x = np.asarray(x, float)
        u = s.u_none if u is None else np.asarray(u, float)
        a, data.Minv, data.Binv, data.Kmat = accelerations(s, data.pinocchio, x, u)
        data.xout = a
        data.cost = cost_value_and_derivatives(s, data.pinocchio, x, u, False)[0]  # not scaled by dt
        n, dt = s.n, s.dt
        dx = np.concatenate([x[2 * n:] * dt + a * dt ** 2, a * dt])  # integrated_action.py:23
        data.xnext = x + dx
        return data.xnext, data.cost

    def calcDiff(self, data, x, u=None):
        s = self.spec
        x = np.asarray(x, float)
        u = s.u_none if u is None else np.asarray(u, float)
        n, dt = s.n, s.dt
        Ax, Au = acceleration_jacobians(s, data.pinocchio, x, u, data.xout, data.Minv, data.Binv, data.Kmat)
        data.dFx, data.dFu = Ax, Au
        ddx_dx = np.vstack([Ax * dt, Ax])
        ddx_dx[range(2 * n), range(2 * n, 4 * n)] += 1.0
        data.Fx = np.eye(4 * n) + dt * ddx_dx  # integrated_action.py:35 with identity Jintegrate
        data.Fu = dt * np.vstack([Au * dt, Au])
        _, data.Lx, data.Lu, data.Lxx, data.Lxu, data.Luu = cost_value_and_derivatives(s, data.pinocchio, x, u, True)


def u_squared(us):
    """python/aslr_to/__init__.py:63-68 on a list of controls."""
    return np.sum(np.asarray(us, float) ** 2, axis=0)
