"""CPU tier: the C-ABI library builds for sm_100a, loads, and exports every symbol that
include/aslr_b200.h declares; struct layouts of the ctypes mirror match the C header; error paths
that need no GPU behave (no compute calls here)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

from aslr_to_b200 import _abi as A
from aslr_to_b200 import workloads as W

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "aslr_b200.h")


def declared_functions():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(aslr_[a-z_0-9]+)\s*\(", src)))


def test_header_and_ctypes_mirror_agree_on_the_symbol_list():
    assert declared_functions() == sorted(A.EXPORTED_SYMBOLS)


def test_library_exports_every_declared_symbol(built_library):
    lib = C.CDLL(built_library)
    for name in declared_functions():
        assert hasattr(lib, name), name
    A.bind(lib)
    assert lib.aslr_version() >= 100
    assert lib.aslr_record_doubles(8, 2) == 174 and lib.aslr_record_doubles(28, 7) == 2044


def test_struct_sizes_match_the_c_header(built_library, tmp_path):
    """Compile a tiny C program against the header and compare sizeof() with ctypes."""
    names = ["AslrSE3", "AslrJoint", "AslrRobot", "AslrDynamics", "AslrCostTerm", "AslrCostSum", "AslrProblemDesc",
             "AslrBatch", "AslrKnotOut", "AslrSolverOpts", "AslrOcpState", "AslrSolverBuffers", "AslrSolveResult",
             "AslrHostSolveIO"]
    mirror = [A.SE3, A.Joint, A.Robot, A.Dynamics, A.CostTerm, A.CostSum, A.ProblemDesc, A.Batch, A.KnotOut,
              A.SolverOpts, A.OcpState, A.SolverBuffers, A.SolveResult, A.HostSolveIO]
    src = tmp_path / "sz.c"
    src.write_text('#include <stdio.h>\n#include "aslr_b200.h"\nint main(void){' +
                   "".join('printf("%%zu\\n", sizeof(%s));' % n for n in names) + "return 0;}\n")
    exe = tmp_path / "sz"
    subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), str(src), "-o", str(exe)])
    sizes = [int(v) for v in subprocess.check_output([str(exe)]).split()]
    assert sizes == [C.sizeof(m) for m in mirror]


def test_problem_create_validates_without_a_gpu(built_library):
    lib = A.bind(C.CDLL(built_library))
    desc = W.to_desc(W.two_dof_sea())
    h = C.c_void_p()
    assert lib.aslr_problem_create(C.byref(desc), C.byref(h)) == 0
    n, nx, nu, rec = C.c_int32(), C.c_int32(), C.c_int32(), C.c_int32()
    assert lib.aslr_problem_dims(h, C.byref(n), C.byref(nx), C.byref(nu), C.byref(rec)) == 0
    assert (n.value, nx.value, nu.value, rec.value) == (2, 8, 2, 174)
    lib.aslr_problem_destroy(h)
    bad = W.to_desc(W.two_dof_sea())
    bad.dyn.nu = 3
    assert lib.aslr_problem_create(C.byref(bad), C.byref(h)) == -1
    assert b"control dimension" in lib.aslr_last_error()
    bad = W.to_desc(W.two_dof_sea())
    bad.running.terms[0].kind = A.COST_STIFFNESS
    assert lib.aslr_problem_create(C.byref(bad), C.byref(h)) == -1
    # every link DOF 1 .. ASLR_MAX_DOF is instantiated (the reference's classes take any pinModel.nv)
    for n_dof in range(1, 8):
        ok = W.to_desc(W.chain_batch(n_dof, B=1, T=2))
        assert lib.aslr_problem_create(C.byref(ok), C.byref(h)) == 0
        assert lib.aslr_problem_dims(h, C.byref(n), C.byref(nx), C.byref(nu), C.byref(rec)) == 0
        assert (n.value, nx.value, nu.value) == (n_dof, 4 * n_dof, n_dof)
        lib.aslr_problem_destroy(h)
    bad = W.to_desc(W.two_dof_sea())
    bad.robot.n = 8
    bad.dyn.nu = 8
    assert lib.aslr_problem_create(C.byref(bad), C.byref(h)) == -4  # ASLR_ERR_UNSUPPORTED beyond ASLR_MAX_DOF
    # the residual-model form of the pendulum cost needs two link DOF, like the cost-model form
    bad = W.to_desc(W.chain_batch(3, B=1, T=2))
    bad.running.terms[0].kind = A.COST_RESIDUAL_DOUBLE_PENDULUM
    assert lib.aslr_problem_create(C.byref(bad), C.byref(h)) == -1


def test_product_refuses_to_run_without_cuda():
    """No CPU fallback: without a device the host layer raises instead of computing."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from aslr_to_b200 import batched
    with pytest.raises(A.AslrError):
        batched.BatchedShootingProblem(W.to_desc(W.two_dof_sea()), np.zeros((1, 8)), 5)
