#!/usr/bin/env python3
"""Time per iteration of the straggler regime: the two never-converging OCPs of the headline batch (924, 3415) solved
alone, 100 iterations, with the persistent straggler kernel (ASLR_TAIL=128) and with the batch kernels (ASLR_TAIL=0).
usage: ASLR_TAIL=0|128 python tools/tail_time.py"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from aslr_to_b200 import _abi as A, batched as BT, workloads as W  # noqa: E402

w = W.sea_batch(4096)
idx = [924, 3415]
x0, K = np.asarray(w["x0"])[idx], np.asarray(w["K_ocp"])[idx]
prob = BT.BatchedShootingProblem(W.to_desc(w), x0, w["T"], K_ocp=K)
sol = BT.BatchedSolver(prob, A.SOLVER_FDDP, maxiter=100, th_stop=1e-7, store_vxx=False, store_quu=False, store_log=False)
xs, us = W.initial_guess(w, x0, w["T"], 2)
for _ in range(3):
    sol.solve(xs, us)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    sol.solve(xs, us)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(json.dumps({"ASLR_TAIL": os.environ.get("ASLR_TAIL", "default"), "ms_per_solve": ms, "us_per_iteration": ms * 10,
                  "launches": int(sol.result.kernel_launches), "iters": sol.states()["iter"].tolist()}))
