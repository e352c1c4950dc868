"""aslr_to_b200 -- B200-native (sm_100a) batched DDP / FDDP / BoxDDP hot path for soft-actuated
(SEA / VSA) robots, a from-scratch drop-in for that path of spykspeigel/aslr_to.

Layers (host side is Python; all arithmetic runs in hand-written CUDA kernels behind the C ABI of
include/aslr_b200.h -- there is no CPU fallback):
  _abi       ctypes mirror of the C ABI + library loader
  desc       builders for the plain-data problem description
  workloads  the reference's example constants as data (BASELINE.json configs)
  batched    BatchedShootingProblem / BatchedSolver (device buffers, streams)
"""
from . import _abi, desc, workloads  # noqa: F401
from ._abi import AslrError  # noqa: F401
from .batched import BatchedShootingProblem, BatchedSolver  # noqa: F401

__version__ = "0.1.0"
