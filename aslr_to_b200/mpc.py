"""Receding-horizon (MPC) loop around the batched solver: the caller side of
`solver.solve(xs, us, maxiter)` (examples/two_dof_sea.py:77-81; SURVEY.md section 8f rank 1,
BASELINE.json configs[4]).  Trajectories, gains and solver state stay on the GPU between steps; one
step = shift the previous solution by one knot, adopt the new measured state as x0 and xs[0], and
re-solve for a few iterations from that warm start."""
import ctypes as C

import numpy as np

from . import _abi as A
from .batched import BatchedSolver


class BatchedMPC(object):
    def __init__(self, problem, kind=A.SOLVER_FDDP, maxiter=10, th_stop=None, store_log=True):
        self.problem = problem
        self.solver = BatchedSolver(problem, kind, maxiter=maxiter, th_stop=th_stop, store_vxx=False, store_quu=False,
                                    store_log=store_log)
        self.maxiter = int(maxiter)

    def initialise(self, xs, us, maxiter=100):
        """First solve (to convergence) from an arbitrary guess; needs a solver with enough log capacity."""
        first = BatchedSolver(self.problem, self.solver.kind, maxiter=maxiter, th_stop=self.solver.th_stop,
                              store_vxx=False, store_quu=False, store_log=False)
        first.solve(xs, us, maxiter=maxiter)
        self.solver.t["xs"].copy_(first.t["xs"])
        self.solver.t["us"].copy_(first.t["us"])
        return first

    def shift(self, x0_new=None):
        """xs[t] <- xs[t+1], us[t] <- us[t+1] (last knot repeated); x0 and xs[0] <- x0_new."""
        p = self.problem
        if x0_new is not None:
            p.x0.copy_(p._dev(x0_new, (p.B, p.nx)))
        b = p.batch_struct()
        A.check(p.lib.aslr_shift_warm_start(p.handle, C.byref(b), self.solver.t["xs"].data_ptr(),
                                            self.solver.t["us"].data_ptr(),
                                            p.x0.data_ptr() if x0_new is not None else None, p._stream()))

    def step(self, x0_new, maxiter=None):
        """One MPC step: shift, then solve(xs_shift, us_shift, maxiter) on the device.  Returns the first
        control of every OCP (device tensor (B, nu))."""
        self.shift(x0_new)
        self.solver.solve(maxiter=self.maxiter if maxiter is None else maxiter, set_candidate=False)
        return self.solver.t["us"][:, 0, :]

