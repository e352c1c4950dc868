"""Shared helpers of the test-suite (test infrastructure only)."""
import ctypes as C
import os

import numpy as np

from aslr_to_b200 import _abi as A
from aslr_to_b200 import batched as BT
from aslr_to_b200 import workloads as W

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
# BASELINE.json north_star: "within 1e-9 relative error on cost, gradient and gains"
RTOL = 1e-9


def golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


def rel_err(a, b):
    """max |a - b| / max(1, max |b|): relative to the array scale, absolute below 1."""
    a, b = np.asarray(a, float), np.asarray(b, float)
    assert a.shape == b.shape, (a.shape, b.shape)
    if a.size == 0:
        return 0.0
    return float(np.abs(a - b).max() / max(1.0, np.abs(b).max()))


def split_record(rec, nx, nu):
    o, out = 0, {}
    lead = tuple(rec.shape[:-1])
    for k, shp in (("Fx", (nx, nx)), ("Fu", (nx, nu)), ("Lxx", (nx, nx)), ("Lxu", (nx, nu)), ("Luu", (nu, nu)),
                   ("Lx", (nx,)), ("Lu", (nu,))):
        sz = int(np.prod(shp))
        out[k] = rec[..., o:o + sz].reshape(lead + shp)
        o += sz
    return out


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def host_knot(hostmath, desc, x, u, terminal=False, diff=True, Kocp=None, mref=None):
    """One knot through the g++ build of the device arithmetic."""
    n = desc.robot.n
    nx, nu = 4 * n, desc.dyn.nu
    rec = np.zeros(A.record_doubles(nx, nu))
    xnext, xout = np.zeros(nx), np.zeros(2 * n)
    dFx, dFu = np.zeros((2 * n, nx)), np.zeros((2 * n, nu))
    cost = C.c_double()
    x = np.ascontiguousarray(x, float)
    u = None if u is None else np.ascontiguousarray(u, float)
    Kocp = None if Kocp is None else np.ascontiguousarray(Kocp, float)
    mref = None if mref is None else np.ascontiguousarray(mref, float)
    rc = hostmath.hh_knot_eval(C.byref(desc), int(terminal), int(diff), _p(x), _p(u), _p(Kocp), _p(mref), _p(xnext),
                               C.byref(cost), _p(rec), _p(xout), _p(dFx), _p(dFu))
    assert rc == 0
    out = dict(xnext=xnext, cost=cost.value, xout=xout, dFx=dFx, dFu=dFu)
    out.update(split_record(rec, nx, nu))
    return out


def default_opts(kind, maxiter, th_stop=None):
    """AslrSolverOpts with crocoddyl's defaults, without needing the CUDA library."""
    o = A.SolverOpts()
    o.kind, o.maxiter, o.is_feasible, o.poll_every = kind, maxiter, 0, 1
    o.th_stop = (5e-5 if kind == A.SOLVER_BOXDDP else 1e-9) if th_stop is None else th_stop
    o.reg_init = float("nan")
    o.reg_min, o.reg_max, o.reg_incfactor, o.reg_decfactor = 1e-9, 1e9, 10.0, 10.0
    o.th_grad, o.th_stepdec, o.th_stepinc, o.th_acceptstep, o.th_acceptnegstep = 1e-12, 0.5, 0.01, 0.1, 2.0
    return o


initial_guess_arrays = W.initial_guess


def host_solve(hostmath, w, ocps=None, maxiter=None, T=None):
    """Serial CPU emulation of aslr_solve (same device bodies, G = 1) on a subset of OCPs."""
    desc = W.to_desc(w)
    n = desc.robot.n
    nx, nu = 4 * n, desc.dyn.nu
    T = w["T"] if T is None else T
    x0 = np.atleast_2d(np.asarray(w["x0"], float))
    K = w.get("K_ocp")
    mref = w.get("mref_ocp")
    if ocps is not None:
        x0 = x0[ocps]
        K = None if K is None else np.asarray(K)[ocps]
        mref = None if mref is None else np.asarray(mref)[ocps]
    x0 = np.ascontiguousarray(x0)
    B = x0.shape[0]
    maxiter = w["maxiter"] if maxiter is None else maxiter
    shapes = BT.solver_buffer_shapes(B, T, nx, nu, maxiter)
    arr = dict((k, np.zeros(s)) for k, s in shapes.items())
    arr["xs"][:], arr["us"][:] = initial_guess_arrays(w, x0, T, nu)
    buf = A.SolverBuffers()
    for k in A.SolverBuffers._names:
        setattr(buf, k, arr[k].ctypes.data if k in arr else None)
    batch = A.Batch()
    batch.B, batch.T, batch.x0 = B, T, x0.ctypes.data
    if K is not None:
        K = np.ascontiguousarray(np.asarray(K, float).reshape(B, n * n))
        batch.K_ocp = K.ctypes.data
    if mref is not None:
        mref = np.ascontiguousarray(np.asarray(mref, float).reshape(B, 12))
        batch.mref_ocp = mref.ctypes.data
    opts = default_opts(W.SOLVER_KIND[w["solver"]], maxiter, w["th_stop"])
    iters = C.c_int()
    rc = hostmath.hh_solve(C.byref(desc), C.byref(batch), C.byref(opts), C.byref(buf), C.byref(iters))
    assert rc == 0
    arr["states"] = BT.decode_states(arr["state"])
    arr["iterations_launched"] = iters.value
    return arr


def compare_solver_trace(res, b, gold, rtol=RTOL, check_gains=True, n_tight=None, loose=None, trace_rtol=None):
    """res: arrays of a batched solve (host emulation or GPU); gold: golden npz of the oracle solver.
    Iteration counts, accepted step lengths and regularisation must be identical; values within rtol.
    n_tight/loose: for the chaotic double-pendulum example only the first n_tight iterations are held to
    rtol and everything after (including the final trajectory and gains) to `loose`.
    trace_rtol: tolerance of the per-iteration cost / stop logs when it differs from rtol (solves of several
    hundred iterations amplify rounding differences in the intermediate iterates; the final solution and
    gains are still held to rtol)."""
    st = res["states"]
    n_it = len(gold["costs"])
    it_ours = int(st["iter"][b]) + (1 if st["status"][b] == A.OCP_CONVERGED else 0)
    assert bool(gold["converged"]) == (st["status"][b] == A.OCP_CONVERGED), (st["status"][b], gold["converged"])
    assert it_ours == n_it, "iteration count differs: ours %d, oracle %d" % (it_ours, n_it)
    lg = res["log"][b][:n_it]
    assert rel_err(lg[:, 3], gold["steps"]) == 0, "accepted step lengths differ"
    assert rel_err(lg[:, 2], gold["x_regs"]) == 0, "regularisation sequence differs"
    if n_tight is not None:
        assert rel_err(lg[:n_tight, 0], gold["costs"][:n_tight]) <= rtol
        rtol = loose
    trace_rtol = rtol if trace_rtol is None else trace_rtol
    assert rel_err(lg[:, 0], gold["costs"]) <= trace_rtol, rel_err(lg[:, 0], gold["costs"])
    assert rel_err(lg[:, 1], gold["stops"]) <= trace_rtol, rel_err(lg[:, 1], gold["stops"])
    errs = dict(cost=rel_err(lg[-1:, 0], gold["costs"][-1:]), xs=rel_err(res["xs"][b], gold["xs"]),
                us=rel_err(res["us"][b], gold["us"]))
    if check_gains:
        errs.update(K=rel_err(res["Kgain"][b], gold["K"]), k=rel_err(res["kff"][b], gold["k"]),
                    Vx=rel_err(res["Vx"][b], gold["Vx"]), Vxx=rel_err(res["Vxx"][b], gold["Vxx"]),
                    Qu=rel_err(res["Qu"][b], gold["Qu"]))
    bad = dict((k, v) for k, v in errs.items() if not v <= rtol)
    assert not bad, "beyond %g: %r (all: %r)" % (rtol, bad, errs)
    return errs
