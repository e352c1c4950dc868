"""Batched shooting problems and solvers on the GPU: the host side of the C ABI.

`BatchedShootingProblem` is the new entry point the B200 build adds next to the reference's
single-OCP API: B independent OCPs that share one running and one terminal
action model (crocoddyl.ShootingProblem(x0, [runningModel] * T, terminalModel),
examples/two_dof_sea.py:66) and differ in x0 and, optionally, the SEA stiffness matrix and the
frame-placement target.  `BatchedSolver` is solver.solve(xs, us, maxiter) for all of them at once
(examples/two_dof_sea.py:69-81).  PyTorch is used for device memory and streams only.
"""
import ctypes as C

import numpy as np

from . import _abi as A


def solver_buffer_shapes(B, T, nx, nu, maxiter, store_vxx=True, store_quu=True, store_log=True):
    """name -> shape (in doubles) of every array of AslrSolverBuffers."""
    rec = A.record_doubles(nx, nu)
    NA = A.NUM_ALPHAS
    shapes = {
        "xs": (B, T + 1, nx), "us": (B, T, nu), "xnext": (B, T, nx), "knot_cost": (B, T + 1),
        "deriv": (B, T + 1, rec), "fs": (B, T + 1, nx), "Kgain": (B, T, nu, nx), "kff": (B, T, nu),
        "Vx": (B, T + 1, nx), "Vxxfs": (B, T + 1, nx), "Qu": (B, T, nu), "Quuk": (B, T, nu),
        "xs_try": (B, NA, T + 1, nx), "us_try": (B, NA, T, nu), "trial": (B, NA, 4),
        "state": (B, A.OCP_STATE_DOUBLES), "counters": ((A.COUNTER_INTS + 2 * NA * B + 1) // 2,),  # int32: running words + line-search work lists
    }
    if store_vxx:
        shapes["Vxx"] = (B, T + 1, nx, nx)
    if store_quu:
        shapes["Quu"] = (B, T, nu, nu)
    if store_log:
        shapes["log"] = (B, max(maxiter, 1), A.LOG_FIELDS)
    return shapes


def decode_states(state_array):
    """(B, OCP_STATE_DOUBLES) float64 array (raw AslrOcpState bytes) -> dict of per-OCP arrays."""
    raw = np.ascontiguousarray(state_array)
    B = raw.shape[0]
    out = {}
    names_d = ("cost", "cost_try", "stop", "xreg", "steplength", "dV", "dVexp", "d1", "d2", "dg", "dq", "ffeas")
    for i, k in enumerate(names_d):
        out[k] = raw[:, i].copy()
    ints = raw[:, len(names_d):].copy().view(np.int32).reshape(B, -1)
    for i, k in enumerate(("iter", "status", "is_feasible", "was_feasible", "recalc", "accepted", "src_slot",
                           "n_backward_retries", "full_steps", "n_logged")):
        out[k] = ints[:, i].copy()
    return out


def make_opts(kind, maxiter=100, th_stop=None, is_feasible=False, reg_init=None, poll_every=1, lib=None):
    o = A.SolverOpts()
    (lib or A.load_library()).aslr_default_solver_opts(kind, C.byref(o))
    o.maxiter = int(maxiter)
    o.is_feasible = int(bool(is_feasible))
    o.poll_every = int(poll_every)
    if th_stop is not None:
        o.th_stop = float(th_stop)
    if reg_init is not None:
        o.reg_init = float(reg_init)
    return o


def _torch():
    import torch
    if not torch.cuda.is_available():
        raise A.AslrError("aslr_to_b200 needs a CUDA device (B200); there is no CPU fallback")
    return torch


class BatchedShootingProblem(object):
    """B independent OCPs sharing one AslrProblemDesc.

    x0: (B, nx); K_ocp: optional (B, n, n) per-OCP SEA stiffness; mref_ocp: optional (B, 12) per-OCP
    placement target (R row-major, p)."""

    def __init__(self, desc, x0, T, K_ocp=None, mref_ocp=None, device=None):
        torch = _torch()
        self.lib = A.load_library()
        self.desc = desc
        self.handle = C.c_void_p()
        # the handle belongs to the device that is current when it is created (include/aslr_b200.h)
        dev_index = torch.cuda.current_device() if device is None else (torch.device(device).index
                                                                         if torch.device(device).index is not None
                                                                         else torch.cuda.current_device())
        with torch.cuda.device(dev_index):
            A.check(self.lib.aslr_problem_create(C.byref(desc), C.byref(self.handle)))
        n, nx, nu, rec = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
        A.check(self.lib.aslr_problem_dims(self.handle, C.byref(n), C.byref(nx), C.byref(nu), C.byref(rec)))
        self.n, self.nx, self.nu, self.rec = n.value, nx.value, nu.value, rec.value
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.T = int(T)
        self.x0 = self._dev(x0, (-1, self.nx))
        self.B = self.x0.shape[0]
        self.K_ocp = None if K_ocp is None else self._dev(K_ocp, (self.B, self.n * self.n))
        self.mref_ocp = None if mref_ocp is None else self._dev(mref_ocp, (self.B, 12))

    def _dev(self, a, shape):
        torch = _torch()
        if isinstance(a, torch.Tensor):
            t = a.to(device=self.device, dtype=torch.float64)
        else:
            t = torch.as_tensor(np.ascontiguousarray(a, dtype=np.float64), device=self.device)
        return t.reshape(shape).contiguous()

    def batch_struct(self):
        b = A.Batch()
        b.B, b.T = self.B, self.T
        b.x0 = self.x0.data_ptr()
        b.K_ocp = self.K_ocp.data_ptr() if self.K_ocp is not None else None
        b.mref_ocp = self.mref_ocp.data_ptr() if self.mref_ocp is not None else None
        return b

    def _stream(self):
        return C.c_void_p(_torch().cuda.current_stream(self.device).cuda_stream)

    def on_device(self):
        """Context manager that makes this problem's device current (allocations, streams and every ABI call of a
        handle happen on the handle's device, whatever the caller's current device is)."""
        return _torch().cuda.device(self.device)

    def _knots(self, fn, xs, us, want):
        torch = _torch()
        xs = self._dev(xs, (self.B, self.T + 1, self.nx))
        us = None if us is None else self._dev(us, (self.B, self.T, self.nu))
        shapes = {"xnext": (self.B, self.T, self.nx), "cost": (self.B, self.T + 1),
                  "deriv": (self.B, self.T + 1, self.rec), "xout": (self.B, self.T + 1, 2 * self.n),
                  "dFx": (self.B, self.T + 1, 2 * self.n, self.nx), "dFu": (self.B, self.T + 1, 2 * self.n, self.nu)}
        out = A.KnotOut()
        res = {}
        for k in want:
            res[k] = torch.zeros(shapes[k], dtype=torch.float64, device=self.device)
            setattr(out, k, res[k].data_ptr())
        b = self.batch_struct()
        A.check(fn(self.handle, C.byref(b), xs.data_ptr(), us.data_ptr() if us is not None else None, C.byref(out),
                   self._stream()))
        return res

    def calc(self, xs, us, want=("xnext", "cost")):
        """ShootingProblem.calc for every OCP: returns device tensors."""
        return self._knots(self.lib.aslr_calc, xs, us, want)

    def calc_diff(self, xs, us, want=("xnext", "cost", "deriv")):
        """ShootingProblem.calc + calcDiff for every OCP (one fused kernel)."""
        return self._knots(self.lib.aslr_calc_diff, xs, us, want)

    def quasi_static(self, xs, maxiter=100, tol=1e-9):
        """ShootingProblem.quasiStatic for every OCP: xs (B, T, nx) -> us (B, T, nu), device tensor."""
        torch = _torch()
        xs = self._dev(xs, (self.B, self.T, self.nx))
        us = torch.zeros((self.B, self.T, self.nu), dtype=torch.float64, device=self.device)
        b = self.batch_struct()
        A.check(self.lib.aslr_quasi_static(self.handle, C.byref(b), xs.data_ptr(), us.data_ptr(), int(maxiter),
                                           float(tol), self._stream()))
        return us

    def split_record(self, deriv):
        """(…, rec) tensor/array -> dict Fx, Fu, Lxx, Lxu, Luu, Lx, Lu."""
        nx, nu = self.nx, self.nu
        lead = tuple(deriv.shape[:-1])
        o, out = 0, {}
        for k, shp in (("Fx", (nx, nx)), ("Fu", (nx, nu)), ("Lxx", (nx, nx)), ("Lxu", (nx, nu)), ("Luu", (nu, nu)),
                       ("Lx", (nx,)), ("Lu", (nu,))):
            sz = int(np.prod(shp))
            out[k] = deriv[..., o:o + sz].reshape(lead + shp)
            o += sz
        return out

    def __del__(self):
        try:
            if self.handle:
                self.lib.aslr_problem_destroy(self.handle)
                self.handle = None
        except Exception:
            pass


class BatchedSolver(object):
    """crocoddyl.SolverDDP / SolverFDDP / SolverBoxDDP over a BatchedShootingProblem."""

    def __init__(self, problem, kind=A.SOLVER_FDDP, maxiter=100, th_stop=None, store_vxx=True, store_quu=True,
                 store_log=True):
        torch = _torch()
        self.problem = problem
        self.kind = kind
        self.maxiter = int(maxiter)
        self.th_stop = th_stop
        p = problem
        self.shapes = solver_buffer_shapes(p.B, p.T, p.nx, p.nu, self.maxiter, store_vxx, store_quu, store_log)
        self.t = dict((k, torch.zeros(s, dtype=torch.float64, device=p.device)) for k, s in self.shapes.items())
        self.buf = A.SolverBuffers()
        for k in A.SolverBuffers._names:
            setattr(self.buf, k, self.t[k].data_ptr() if k in self.t else None)
        self.result = A.SolveResult()
        self.opts = None

    def set_candidate(self, xs=None, us=None):
        """setCandidate: None -> state.zero() / zeros."""
        p = self.problem
        if xs is None:
            self.t["xs"].zero_()
        else:
            self.t["xs"].copy_(p._dev(xs, (p.B, p.T + 1, p.nx)))
        if us is None:
            self.t["us"].zero_()
        else:
            self.t["us"].copy_(p._dev(us, (p.B, p.T, p.nu)))

    def solve(self, xs=None, us=None, maxiter=None, is_feasible=False, reg_init=None, poll_every=1,
              set_candidate=True):
        p = self.problem
        maxiter = self.maxiter if maxiter is None else int(maxiter)
        if maxiter > self.maxiter:
            raise ValueError("maxiter exceeds the log capacity this solver was created with")
        if set_candidate:
            self.set_candidate(xs, us)
        self.opts = make_opts(self.kind, maxiter, self.th_stop, is_feasible, reg_init, poll_every, p.lib)
        b = p.batch_struct()
        A.check(p.lib.aslr_solve(p.handle, C.byref(b), C.byref(self.opts), C.byref(self.buf), C.byref(self.result),
                                 p._stream()))
        return self

    # ---- results (host copies on demand)
    def states(self):
        return decode_states(self.t["state"].cpu().numpy())

    def get(self, name):
        return self.t[name]

    def logs(self):
        """(B, maxiter, LOG_FIELDS): cost, stop, xreg, steplength, dV, dVexp, ffeas, is_feasible per iteration."""
        B, m = self.problem.B, self.opts.maxiter
        return self.t["log"].reshape(-1)[:B * m * A.LOG_FIELDS].reshape(B, m, A.LOG_FIELDS).cpu().numpy()


def solve_host(problem_desc, x0, T, kind=A.SOLVER_FDDP, xs=None, us=None, maxiter=100, th_stop=None, K_ocp=None,
               mref_ocp=None, is_feasible=False, reg_init=None, want_gains=False, want_log=False, out=None,
               handle=None, lib=None):
    """solver.solve(xs, us, maxiter) for a batch of OCPs held in HOST memory (numpy arrays or pinned torch tensors),
    through aslr_solve_host: the library owns the device workspace and does every copy itself.  Returns a dict of
    host arrays (xs, us, states, optionally Kgain / kff / log) plus the AslrSolveResult.  `out`: optional dict of
    preallocated host arrays (e.g. pinned) to receive xs / us / state; `handle`: reuse an AslrProblem handle
    (keeps the device workspace across calls)."""
    lib = lib or A.load_library()
    own = handle is None
    if own:
        handle = C.c_void_p()
        A.check(lib.aslr_problem_create(C.byref(problem_desc), C.byref(handle)))
    try:
        n = problem_desc.robot.n
        nx, nu = 4 * n, problem_desc.dyn.nu

        def host(a, shape):
            if a is None:
                return None, None
            if hasattr(a, "data_ptr"):  # torch CPU tensor (pinned or not)
                assert a.dtype.itemsize == 8 and a.is_contiguous() and tuple(a.shape) == tuple(shape)
                return a, a.data_ptr()
            arr = np.ascontiguousarray(a, dtype=np.float64).reshape(shape)
            return arr, arr.ctypes.data

        x0a, x0p = host(x0 if hasattr(x0, "data_ptr") else np.atleast_2d(np.asarray(x0, float)),
                        (len(x0) if np.ndim(x0) == 2 else 1, nx))
        B = x0a.shape[0]
        keep = [x0a]
        io = A.HostSolveIO()
        io.B, io.T, io.x0 = B, int(T), x0p
        for name, a, shape in (("K_ocp", K_ocp, (B, n * n)), ("mref_ocp", mref_ocp, (B, 12)),
                               ("xs_init", xs, (B, T + 1, nx)), ("us_init", us, (B, T, nu))):
            arr, ptr = host(a, shape)
            keep.append(arr)
            setattr(io, name, ptr)
        out = {} if out is None else out
        res = {}
        wants = [("xs", (B, T + 1, nx)), ("us", (B, T, nu)), ("state", (B, A.OCP_STATE_DOUBLES))]
        if want_gains:
            wants += [("Kgain", (B, T, nu, nx)), ("kff", (B, T, nu))]
        if want_log:
            wants += [("log", (B, max(int(maxiter), 1), A.LOG_FIELDS))]
        for name, shape in wants:
            arr = out.get(name)
            if arr is None:
                arr = np.zeros(shape)
            a2, ptr = host(arr, shape)
            res[name] = a2
            setattr(io, name, ptr)
        opts = make_opts(kind, maxiter, th_stop, is_feasible, reg_init, 1, lib)
        result = A.SolveResult()
        A.check(lib.aslr_solve_host(handle, C.byref(io), C.byref(opts), C.byref(result)))
        st = res["state"]
        res["states"] = decode_states(st.numpy() if hasattr(st, "numpy") else st)
        res["result"] = result
        return res
    finally:
        if own:
            lib.aslr_problem_destroy(handle)
