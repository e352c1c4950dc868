"""GPU tier, collected last on purpose: aslr_solve_static, the entry point that
allocates nothing (its solver scratch is a file-scope device array of the library, include/aslr_b200.h), called through
the C ABI and compared with the existing entry point aslr_solve on caller-allocated buffers.  OCPs are independent and
the launch sequence is the same, so the results must be identical bit for bit -- the 1e-9 tolerance of the other GPU
tests (util.RTOL) is the stated bound, equality is what is observed.  Two calls in a row share the static array, which
keeps the first call's contents: the second call (another batch size and horizon) must not see them.  No device memory
is allocated between the two calls: every tensor is created beforehand."""
import ctypes as C

import numpy as np
import pytest

from aslr_to_b200 import _abi as A
from aslr_to_b200 import workloads as W
from util import RTOL, initial_guess_arrays, rel_err

pytestmark = pytest.mark.gpu


def _case(BT, torch, B, T, seed, maxiter):
    """Problem, inputs and preallocated outputs of one call; the reference result from aslr_solve."""
    w = W.sea_batch(B, seed=seed)
    w["T"] = T
    prob = BT.BatchedShootingProblem(W.to_desc(w), w["x0"], T, K_ocp=w["K_ocp"])
    xs0, us0 = initial_guess_arrays(w, w["x0"], T, prob.nu)
    ref = BT.BatchedSolver(prob, A.SOLVER_FDDP, maxiter=maxiter, th_stop=w["th_stop"])
    ref.solve(xs0, us0, maxiter=maxiter)
    torch.cuda.synchronize()
    dev = prob.device
    t = dict(xs=torch.as_tensor(xs0, device=dev).contiguous(), us=torch.as_tensor(us0, device=dev).contiguous(),
             state=torch.zeros((B, A.OCP_STATE_DOUBLES), dtype=torch.float64, device=dev),
             Kgain=torch.zeros((B, T, prob.nu, prob.nx), dtype=torch.float64, device=dev),
             kff=torch.zeros((B, T, prob.nu), dtype=torch.float64, device=dev))
    opts = BT.make_opts(A.SOLVER_FDDP, maxiter, w["th_stop"], lib=prob.lib)
    return prob, ref, t, opts


def _call(prob, t, opts, with_gains=True):
    b = prob.batch_struct()
    return prob.lib.aslr_solve_static(prob.handle, C.byref(b), C.byref(opts), t["xs"].data_ptr(), t["us"].data_ptr(),
                                      t["state"].data_ptr(), t["Kgain"].data_ptr() if with_gains else None,
                                      t["kff"].data_ptr() if with_gains else None, prob._stream())


def _compare(BT, torch, ref, t, with_gains=True):
    torch.cuda.synchronize()
    st = BT.decode_states(t["state"].cpu().numpy())
    rs = ref.states()
    for k in ("iter", "status", "n_logged"):
        assert np.array_equal(st[k], rs[k]), k
    for k in ("cost", "stop", "xreg", "steplength"):
        assert rel_err(st[k], rs[k]) <= RTOL, k
        assert np.array_equal(st[k], rs[k]), k  # observed: identical
    names = ("xs", "us") + (("Kgain", "kff") if with_gains else ())
    for k in names:
        ours, theirs = t[k].cpu().numpy(), ref.get(k).cpu().numpy()
        assert rel_err(ours, theirs) <= RTOL, k
        assert np.array_equal(ours, theirs), k


def test_no_alloc_solve_matches_aslr_solve_on_two_calls_in_a_row(built_library):
    import torch
    assert torch.cuda.is_available(), "the gpu tier needs a CUDA device"
    from aslr_to_b200 import batched as BT
    # everything either call needs is created before the first one
    prob1, ref1, t1, opts1 = _case(BT, torch, B=37, T=20, seed=11, maxiter=12)
    prob2, ref2, t2, opts2 = _case(BT, torch, B=5, T=33, seed=12, maxiter=9)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    A.check(_call(prob1, t1, opts1))
    A.check(_call(prob2, t2, opts2, with_gains=False))  # same stream: ordered after the first call
    assert torch.cuda.memory_allocated() == before
    _compare(BT, torch, ref1, t1)
    _compare(BT, torch, ref2, t2, with_gains=False)
    # and once more over the first problem, now that the array holds the second call's contents
    xs0, us0 = initial_guess_arrays(dict(init="x0-zero-u"), prob1.x0.cpu().numpy(), prob1.T, prob1.nu)
    t1["xs"].copy_(torch.as_tensor(xs0, device=prob1.device))
    t1["us"].copy_(torch.as_tensor(us0, device=prob1.device))
    A.check(_call(prob1, t1, opts1))
    _compare(BT, torch, ref1, t1)


def test_no_alloc_solve_rejects_batches_beyond_the_static_array(built_library):
    import torch
    assert torch.cuda.is_available(), "the gpu tier needs a CUDA device"
    from aslr_to_b200 import batched as BT
    w = W.sea_batch(4096)
    prob = BT.BatchedShootingProblem(W.to_desc(w), w["x0"], w["T"], K_ocp=w["K_ocp"])
    opts = BT.make_opts(A.SOLVER_FDDP, 3, lib=prob.lib)
    dummy = torch.zeros(8, dtype=torch.float64, device=prob.device)
    b = prob.batch_struct()
    rc = prob.lib.aslr_solve_static(prob.handle, C.byref(b), C.byref(opts), dummy.data_ptr(), dummy.data_ptr(),
                                    dummy.data_ptr(), None, None, prob._stream())
    assert rc == -4
    msg = prob.lib.aslr_last_error().decode()
    assert "ASLR_STATIC_SCRATCH_BYTES" in msg and str(A.STATIC_SCRATCH_BYTES) in msg
