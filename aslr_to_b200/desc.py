"""Builders for the plain-data problem description handed to the C ABI (include/aslr_b200.h).

Everything here is host-side bookkeeping: turning the robot fixture and the cost / dynamics
parameters of the reference's Python objects into an `AslrProblemDesc`.
"""
import json
import os

import numpy as np

from . import _abi as A

ROBOT_DIR = os.path.join(os.path.dirname(os.path.abspath(__file__)), "robots")


def load_robot_dict(name_or_path):
    path = name_or_path if os.path.isfile(name_or_path) else os.path.join(ROBOT_DIR, name_or_path + ".json")
    with open(path) as f:
        return json.load(f)


def _fill(dst, values):
    values = np.asarray(values, float).ravel()
    for i, v in enumerate(values):
        dst[i] = float(v)


def _se3(R, p):
    s = A.SE3()
    _fill(s.R, np.asarray(R, float).reshape(3, 3))
    _fill(s.p, p)
    return s


def make_robot(robot_dict, gravity=None):
    """AslrRobot from a fixture dict (tools/make_robots.py); `gravity` overrides model.gravity.linear
    (examples/two_dof_sea.py:20)."""
    joints = robot_dict["joints"]
    n = len(joints)
    if not 1 <= n <= A.MAX_DOF:
        raise ValueError("unsupported number of joints %d (1..%d)" % (n, A.MAX_DOF))
    r = A.Robot()
    r.n = n
    _fill(r.gravity, robot_dict["gravity"] if gravity is None else gravity)
    for i, jd in enumerate(joints):
        j = r.joints[i]
        j.placement = _se3(jd["placement"]["R"], jd["placement"]["p"])
        axis = np.asarray(jd["axis"], float)
        _fill(j.axis, axis / np.linalg.norm(axis))
        j.mass = float(jd["mass"])
        _fill(j.com, jd["com"])
        _fill(j.inertia, np.asarray(jd["inertia"], float).reshape(3, 3))
    return r


def make_dynamics(kind, n, B, K=None, Sm=None, write_fu=True, u_lb=None, u_ub=None, u_none=None):
    d = A.Dynamics()
    d.kind = kind
    nu = n if kind == A.DYN_SEA else 2 * n
    d.nu = nu
    d.write_fu = 1 if write_fu else 0
    Kmat = np.zeros((n, n)) if K is None else np.asarray(K, float).reshape(n, n)
    _fill(d.K, Kmat)
    _fill(d.Binv, np.linalg.inv(np.asarray(B, float).reshape(n, n)))  # free_fwddyn_asr.py:41
    _fill(d.Sm, np.eye(n) if Sm is None else np.asarray(Sm, float).reshape(n, n))
    if u_none is None:
        u_none = np.zeros(nu)
        if kind == A.DYN_VSA:
            u_none[n:] = 3.0  # free_fwddyn_vsa.py:21-23
    _fill(d.u_none, u_none)
    lb = -np.inf * np.ones(nu) if u_lb is None else np.asarray(u_lb, float)
    ub = np.inf * np.ones(nu) if u_ub is None else np.asarray(u_ub, float)
    if lb.shape != (nu,) or ub.shape != (nu,):
        raise ValueError("control bounds have wrong dimension (it should be %d)" % nu)
    _fill(d.u_lb, lb)
    _fill(d.u_ub, ub)
    d.has_control_limits = int(bool(np.all(np.isfinite(lb)) and np.all(np.isfinite(ub))))
    return d


def cost_term(kind, weight, ref=None, act=None, frame_parent=0, frame_R=None, frame_p=None):
    t = A.CostTerm()
    t.kind = kind
    t.weight = float(weight)
    t.frame_parent = int(frame_parent)
    if ref is not None:
        _fill(t.ref, ref)
    if act is not None:
        _fill(t.act, act)
    t.frame = _se3(np.eye(3) if frame_R is None else frame_R, np.zeros(3) if frame_p is None else frame_p)
    return t


def placement_term(weight, frame_parent, frame_R, frame_p, ref_R, ref_p, act=None):
    ref = np.concatenate([np.asarray(ref_R, float).reshape(9), np.asarray(ref_p, float).reshape(3)])
    return cost_term(A.COST_FRAME_PLACEMENT, weight, ref, np.ones(6) if act is None else act, frame_parent, frame_R,
                     frame_p)


def make_cost_sum(terms):
    if len(terms) > A.MAX_COST_TERMS:
        raise ValueError("at most %d cost terms per cost sum are supported" % A.MAX_COST_TERMS)
    s = A.CostSum()
    s.nterms = len(terms)
    for i, t in enumerate(terms):
        s.terms[i] = t
    return s


def make_problem_desc(robot, dyn, running_terms, terminal_terms, dt, dt_terminal=0.0):
    d = A.ProblemDesc()
    d.robot = robot
    d.dyn = dyn
    d.running = make_cost_sum(running_terms)
    d.terminal = make_cost_sum(terminal_terms)
    d.dt = float(dt)
    d.dt_terminal = float(dt_terminal)
    return d
