"""Drop-in Python layer: the reference's class names on top of the CUDA library.

    from aslr_to_b200.compat import crocoddyl, pinocchio, example_robot_data, aslr_to

is the only change an example script of the reference needs (examples/two_dof_sea.py:6-10).  The
classes below hold parameters only; every `calc`, `calcDiff`, `quasiStatic` and `solve` compiles
them into an `AslrProblemDesc` (include/aslr_b200.h) and runs the CUDA kernels through the C ABI.
There is no CPU implementation of any of the arithmetic here: without the library or without a
CUDA device these calls raise `AslrError`.

Supported surface = what /root/reference's package, examples and unit tests call (SURVEY.md
section 8b).  Anything else raises `NotImplementedError` with the name of the missing feature.
"""
from . import pinocchio, example_robot_data, crocoddyl, aslr_to  # noqa: F401
