"""ORACLE / TEST INFRASTRUCTURE ONLY -- NumPy restatement of the slice of the pinocchio API that
spykspeigel/aslr_to calls.  Never imported by the product package `aslr_to_b200`.

pinocchio (https://github.com/stack-of-tasks/pinocchio) is an un-vendored, un-pinned dependency
of the reference (/root/reference/CMakeLists.txt:23) and is absent from this build, so the
published algorithms are restated here (Featherstone RNEA in body coordinates; SE(3) log and its
right Jacobian as in pinocchio's explog.hpp).  Call sites served:
  free_fwddyn_asr.py:38,49-50,75,111   computeAllTerms / forwardKinematics /
                                       updateFramePlacements / computeRNEADerivatives / rnea
  residual_frame_placement.py:14-22    SE3, log, Jlog6, getFrameJacobian(LOCAL)
  statemultibody_aslr.py:19-28,43,58,77,104  neutral / randomConfiguration / difference /
                                       integrate / dIntegrate / utils.zero / utils.rand
Conventions follow pinocchio: spatial vectors are [linear; angular]; only revolute joints on a
serial chain (nq == nv), which is all the reference supports (statemultibody_aslr.py:52).

`computeRNEADerivatives` is evaluated by complex-step differentiation of the RNEA recursion
(exact to rounding, independent of the analytic tangent recursion used by the CUDA kernels).
PARITY: no pinocchio golden vectors exist in the reference; this shim is
pinned instead by independent checks in tests/ (sympy Lagrangian closed form, central
differences, log6(exp6) round trips).
"""
import json

import numpy as np

from . import utils  # noqa: F401


def _x3(a, b):
    """3-vector cross product (np.cross is ~50x slower on tiny arrays; complex-step safe)."""
    return np.array([a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0]])


def skew(v):
    return np.array([[0, -v[2], v[1]], [v[2], 0, -v[0]], [-v[1], v[0], 0]], dtype=np.result_type(v, float))


def _rodrigues(axis, ang):
    K = skew(np.asarray(axis, float))
    return np.eye(3) + np.sin(ang) * K + (1.0 - np.cos(ang)) * (K @ K)


class SE3(object):
    def __init__(self, R=None, p=None):
        self.rotation = np.eye(3) if R is None else np.array(R)
        self.translation = np.zeros(3) if p is None else np.array(p).reshape(3)

    @staticmethod
    def Identity():
        return SE3(np.eye(3), np.zeros(3))

    def inverse(self):
        Rt = self.rotation.T
        return SE3(Rt, -Rt @ self.translation)

    def __mul__(self, other):
        return SE3(self.rotation @ other.rotation, self.rotation @ other.translation + self.translation)

    def copy(self):
        return SE3(self.rotation.copy(), self.translation.copy())

    # spatial motion [lin; ang]
    def act_motion(self, m):
        w = self.rotation @ m[3:]
        return np.concatenate([self.rotation @ m[:3] + _x3(self.translation, w), w])

    def actInv_motion(self, m):
        Rt = self.rotation.T
        return np.concatenate([Rt @ (m[:3] - _x3(self.translation, m[3:])), Rt @ m[3:]])

    # spatial force [lin; ang]
    def act_force(self, f):
        fl = self.rotation @ f[:3]
        return np.concatenate([fl, self.rotation @ f[3:] + _x3(self.translation, fl)])

    @property
    def homogeneous(self):
        H = np.eye(4)
        H[:3, :3] = self.rotation
        H[:3, 3] = self.translation
        return H

    def __repr__(self):
        return "SE3(R=%r, p=%r)" % (self.rotation, self.translation)


class Motion(object):
    def __init__(self, vec):
        self.vector = np.asarray(vec)

    @property
    def linear(self):
        return self.vector[:3]

    @property
    def angular(self):
        return self.vector[3:]


class ReferenceFrame(object):
    WORLD = 0
    LOCAL = 1
    LOCAL_WORLD_ALIGNED = 2


class _Gravity(object):
    def __init__(self, lin):
        self.linear = np.array(lin, float)
        self.angular = np.zeros(3)


class _Joint(object):
    __slots__ = ("name", "placement", "axis", "mass", "com", "inertia", "q_min", "q_max")


class _Frame(object):
    __slots__ = ("name", "parent", "placement")


class Model(object):
    """Serial chain of revolute joints built from a fixture dict (tools/make_robots.py)."""

    def __init__(self, desc):
        if isinstance(desc, str):
            with open(desc) as f:
                desc = json.load(f)
        self.name = desc["name"]
        self.gravity = _Gravity(desc["gravity"])
        self.joints = []
        for jd in desc["joints"]:
            j = _Joint()
            j.name = jd["name"]
            j.placement = SE3(np.array(jd["placement"]["R"]), np.array(jd["placement"]["p"]))
            j.axis = np.array(jd["axis"], float)
            j.mass = float(jd["mass"])
            j.com = np.array(jd["com"], float)
            j.inertia = np.array(jd["inertia"], float)
            j.q_min, j.q_max = jd.get("q_min", -np.pi), jd.get("q_max", np.pi)
            self.joints.append(j)
        self.frames = []
        for fd in desc["frames"]:
            fr = _Frame()
            fr.name = fd["name"]
            fr.parent = int(fd["parent"])
            fr.placement = SE3(np.array(fd["placement"]["R"]), np.array(fd["placement"]["p"]))
            self.frames.append(fr)
        self.nq = self.nv = len(self.joints)
        self.njoints = self.nq + 1
        self.lowerPositionLimit = np.array([j.q_min for j in self.joints])
        self.upperPositionLimit = np.array([j.q_max for j in self.joints])

    def getFrameId(self, name):
        for i, f in enumerate(self.frames):
            if f.name == name:
                return i
        return len(self.frames)

    def createData(self):
        return Data(self)


class Data(object):
    def __init__(self, model):
        n = model.nv
        self.M = np.zeros((n, n))
        self.nle = np.zeros(n)
        self.tau = np.zeros(n)
        self.dtau_dq = np.zeros((n, n))
        self.dtau_dv = np.zeros((n, n))
        self.oMi = [SE3.Identity() for _ in range(n)]
        self.oMf = [SE3.Identity() for _ in model.frames]


# ----------------------------------------------------------------------------------------------
# algorithms
# ----------------------------------------------------------------------------------------------
def _cross_motion(v, m):
    """v x m (motion cross motion), [lin; ang]."""
    return np.concatenate([_x3(v[3:], m[:3]) + _x3(v[:3], m[3:]), _x3(v[3:], m[3:])])


def _cross_force(v, f):
    """v x* f (motion cross force)."""
    return np.concatenate([_x3(v[3:], f[:3]), _x3(v[3:], f[3:]) + _x3(v[:3], f[:3])])


def _inertia_mul(j, v):
    fl = j.mass * (v[:3] - _x3(j.com, v[3:]))
    return np.concatenate([fl, j.inertia @ v[3:] + _x3(j.com, fl)])


def _rnea(model, q, v, a, gravity):
    """Recursive Newton-Euler in body coordinates; complex-step safe (no abs/conj)."""
    n = model.nv
    dt = np.result_type(q, v, a, float)
    vs, as_, fs, Xs = [], [], [], []
    vp = np.zeros(6, dt)
    ap = np.concatenate([-np.asarray(gravity, float), np.zeros(3)]).astype(dt)
    for i, j in enumerate(model.joints):
        R = j.placement.rotation @ _rodrigues(j.axis, q[i])
        X = SE3(R, j.placement.translation)
        S = np.concatenate([np.zeros(3), j.axis])
        vi = X.actInv_motion(vp) + S * v[i]
        ai = X.actInv_motion(ap) + S * a[i] + _cross_motion(vi, S * v[i])
        fi = _inertia_mul(j, ai) + _cross_force(vi, _inertia_mul(j, vi))
        vs.append(vi), as_.append(ai), fs.append(fi), Xs.append(X)
        vp, ap = vi, ai
    tau = np.zeros(n, dt)
    for i in range(n - 1, -1, -1):
        tau[i] = model.joints[i].axis @ fs[i][3:]
        if i > 0:
            fs[i - 1] = fs[i - 1] + Xs[i].act_force(fs[i])
    return tau


def rnea(model, data, q, v, a):
    data.tau = _rnea(model, np.asarray(q), np.asarray(v), np.asarray(a), model.gravity.linear)
    return data.tau


def forwardKinematics(model, data, q, v=None, a=None):
    M = SE3.Identity()
    for i, j in enumerate(model.joints):
        M = M * j.placement * SE3(_rodrigues(j.axis, q[i]), np.zeros(3))
        data.oMi[i] = M


def updateFramePlacements(model, data):
    for k, f in enumerate(model.frames):
        data.oMf[k] = f.placement.copy() if f.parent < 0 else data.oMi[f.parent] * f.placement


def framesForwardKinematics(model, data, q):
    forwardKinematics(model, data, q)
    updateFramePlacements(model, data)


def crba(model, data, q):
    n = model.nv
    z = np.zeros(n)
    M = np.zeros((n, n))
    for j in range(n):
        e = np.zeros(n)
        e[j] = 1.0
        M[:, j] = _rnea(model, np.asarray(q, float), z, e, np.zeros(3))
    data.M = 0.5 * (M + M.T)
    return data.M


def nonLinearEffects(model, data, q, v):
    data.nle = _rnea(model, np.asarray(q, float), np.asarray(v, float), np.zeros(model.nv), model.gravity.linear)
    return data.nle


def computeAllTerms(model, data, q, v):
    """Only the outputs the reference reads: data.M (full symmetric), data.nle, and kinematics
    (free_fwddyn_asr.py:38-43)."""
    crba(model, data, q)
    nonLinearEffects(model, data, q, v)
    forwardKinematics(model, data, q, v)


def computeRNEADerivatives(model, data, q, v, a):
    n = model.nv
    h = 1e-30
    q = np.asarray(q, float)
    v = np.asarray(v, float)
    a = np.asarray(a, float)
    g = model.gravity.linear
    dq = np.zeros((n, n))
    dv = np.zeros((n, n))
    for j in range(n):
        e = np.zeros(n, complex)
        e[j] = 1j * h
        dq[:, j] = _rnea(model, q + e, v.astype(complex), a.astype(complex), g).imag / h
        dv[:, j] = _rnea(model, q.astype(complex), v + e, a.astype(complex), g).imag / h
    data.dtau_dq = dq
    data.dtau_dv = dv
    data.tau = _rnea(model, q, v, a, g)
    return dq, dv


def getFrameJacobian(model, data, frame_id, rf=ReferenceFrame.LOCAL):
    """6 x nv frame Jacobian, [lin; ang] rows (residual_frame_placement.py:21-22).  Uses
    data.oMi / data.oMf from the last forward kinematics."""
    f = model.frames[frame_id]
    J = np.zeros((6, model.nv))
    oMf = data.oMi[f.parent] * f.placement if f.parent >= 0 else f.placement
    if rf == ReferenceFrame.LOCAL:
        ref = oMf.inverse()
    elif rf == ReferenceFrame.WORLD:
        ref = SE3.Identity()
    else:
        ref = SE3(np.eye(3), -oMf.translation)
    for j in range(f.parent + 1):
        S = np.concatenate([np.zeros(3), model.joints[j].axis])
        J[:, j] = (ref * data.oMi[j]).act_motion(S)
    return J


# ----------------------------------------------------------------------------------------------
# SE(3) log map and its right Jacobian (pinocchio explog.hpp)
# ----------------------------------------------------------------------------------------------
_TAYLOR3 = np.finfo(float).eps ** 0.25


def log3(R):
    tr = np.trace(R)
    if tr >= 3.0:
        theta = 0.0
    elif tr <= -1.0:
        theta = np.pi
    else:
        theta = np.arccos(0.5 * (tr - 1.0))
    if theta >= np.pi - 1e-2:
        # near pi: recover the axis from the diagonal, signs from the antisymmetric part
        cphi = -0.5 * (tr - 1.0)
        beta = theta * theta / (1.0 + cphi)
        tmp = (np.diag(R) + cphi) * beta
        sg = [1.0 if R[2, 1] > R[1, 2] else -1.0, 1.0 if R[0, 2] > R[2, 0] else -1.0,
              1.0 if R[1, 0] > R[0, 1] else -1.0]
        w = np.array([sg[k] * (np.sqrt(tmp[k]) if tmp[k] > 0 else 0.0) for k in range(3)])
    else:
        t = (theta / np.sin(theta) if theta > _TAYLOR3 else 1.0) / 2.0
        w = t * np.array([R[2, 1] - R[1, 2], R[0, 2] - R[2, 0], R[1, 0] - R[0, 1]])
    return w, theta


def log6(M):
    R, p = M.rotation, M.translation
    w, theta = log3(R)
    t2 = theta * theta
    if theta < _TAYLOR3:
        alpha = 1.0 - t2 / 12.0 - t2 * t2 / 720.0
        beta = 1.0 / 12.0 + t2 / 720.0
    else:
        st, ct = np.sin(theta), np.cos(theta)
        alpha = theta * st / (2.0 * (1.0 - ct))
        beta = 1.0 / t2 - st / (2.0 * theta * (1.0 - ct))
    lin = alpha * p - 0.5 * _x3(w, p) + beta * (w @ p) * w
    return Motion(np.concatenate([lin, w]))


def log(M):
    if isinstance(M, SE3):
        return log6(M)
    return log3(np.asarray(M))[0]


def Jlog3(theta, w):
    t2 = theta * theta
    if theta < _TAYLOR3:
        alpha = 1.0 / 12.0 + t2 / 720.0
        diag = 0.5 * (2.0 - t2 / 6.0)
    else:
        st, ct = np.sin(theta), np.cos(theta)
        alpha = 1.0 / t2 - st / (2.0 * theta * (1.0 - ct))
        diag = 0.5 * (theta * st / (1.0 - ct))
    return alpha * np.outer(w, w) + diag * np.eye(3) + 0.5 * skew(w)


def Jlog6(M):
    R, p = M.rotation, M.translation
    w, theta = log3(R)
    t2 = theta * theta
    A = Jlog3(theta, w)
    if theta < _TAYLOR3:
        beta = 1.0 / 12.0 + t2 / 720.0
        bdot = 1.0 / 360.0
    else:
        st, ct = np.sin(theta), np.cos(theta)
        tinv = 1.0 / theta
        t2inv = tinv * tinv
        inv_2_2ct = 1.0 / (2.0 * (1.0 - ct))
        beta = t2inv - (st / (1.0 - ct)) * tinv * 0.5
        bdot = -2.0 * t2inv * t2inv + (1.0 + st * tinv) * t2inv * inv_2_2ct
    wTp = w @ p
    v3 = (bdot * wTp) * w - (t2 * bdot + 2.0 * beta) * p
    C = np.outer(v3, w) + beta * np.outer(w, p) + beta * wTp * np.eye(3) + 0.5 * skew(p)
    J = np.zeros((6, 6))
    J[:3, :3] = A
    J[3:, 3:] = A
    J[:3, 3:] = C @ A
    return J


def exp6(nu):
    """Inverse of log6 (test helper)."""
    v, w = np.asarray(nu[:3], float), np.asarray(nu[3:], float)
    th = np.linalg.norm(w)
    W = skew(w)
    if th < 1e-10:
        return SE3(np.eye(3) + W, v + 0.5 * W @ v)
    R = np.eye(3) + np.sin(th) / th * W + (1 - np.cos(th)) / th ** 2 * (W @ W)
    V = np.eye(3) + (1 - np.cos(th)) / th ** 2 * W + (th - np.sin(th)) / th ** 3 * (W @ W)
    return SE3(R, V @ v)


# ----------------------------------------------------------------------------------------------
# configuration-space helpers (vector space for revolute chains)
# ----------------------------------------------------------------------------------------------
def neutral(model):
    return np.zeros(model.nq)


def randomConfiguration(model, lo=None, hi=None):
    lo = model.lowerPositionLimit if lo is None else lo
    hi = model.upperPositionLimit if hi is None else hi
    return lo + (hi - lo) * np.random.rand(model.nq)


def difference(model, q0, q1):
    return np.asarray(q1) - np.asarray(q0)


def integrate(model, q, dq):
    return np.asarray(q) + np.asarray(dq)


def dIntegrate(model, q, dq):
    return np.eye(model.nv), np.eye(model.nv)
