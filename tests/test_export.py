"""Result export: host-side bookkeeping on fake solver arrays (CPU tier) and a round trip on the GPU."""
import numpy as np
import pytest

from aslr_to_b200 import _abi as A
from aslr_to_b200 import export as E


def test_callback_log_and_trajectory_dict():
    logs = np.arange(2 * 5 * A.LOG_FIELDS, dtype=float).reshape(2, 5, A.LOG_FIELDS)
    states = {"iter": np.array([2, 5]), "status": np.array([A.OCP_CONVERGED, A.OCP_MAXITER])}
    l0, l1 = E.callback_log(logs, states, 0), E.callback_log(logs, states, 1)
    assert len(l0["costs"]) == 3 and len(l1["costs"]) == 5  # converged at index 2 -> 3 rows; maxiter 5 -> 5 rows
    assert np.array_equal(l0["steps"], logs[0, :3, 3]) and np.array_equal(l0["u_regs"], l0["x_regs"])
    xs = np.arange(3 * 8, dtype=float).reshape(3, 8)
    us = np.arange(2 * 2, dtype=float).reshape(2, 2)
    d = E.trajectory_dict(xs, us, 1e-2, 2)
    assert np.array_equal(d["q2"], xs[:, 1]) and np.array_equal(d["th1"], xs[:, 2]) and np.array_equal(d["u2"], us[:, 1])
    assert np.allclose(d["t"], [0.0, 0.01])


@pytest.mark.gpu
def test_save_solution_round_trip(built_library, tmp_path):
    from aslr_to_b200 import batched as BT, workloads as W
    w = W.sea_batch(3)
    prob = BT.BatchedShootingProblem(W.to_desc(w), w["x0"], 10, K_ocp=w["K_ocp"])
    solver = BT.BatchedSolver(prob, A.SOLVER_FDDP, maxiter=4, th_stop=1e-7)
    solver.solve(*W.initial_guess(w, w["x0"], 10, 2))
    for name in ("sol.npz", "sol.mat"):
        data = E.save_solution(tmp_path / name, solver, b=1)
        if name.endswith(".npz"):
            back = np.load(tmp_path / name)
        else:
            from scipy.io import loadmat
            back = loadmat(str(tmp_path / name))
        assert np.allclose(np.asarray(back["xs"]), data["xs"]) and np.allclose(np.ravel(back["q1"]), data["q1"])
        assert len(np.ravel(back["log_costs"])) == data["iterations"] == 4
