"""ctypes mirror of include/aslr_b200.h (the C ABI of libaslr_b200.so) and the library loader.

There is deliberately no CPU fallback: if the CUDA library is missing or no B200 is visible the
product raises (BASELINE.json north_star: "no CPU fallback").
"""
import ctypes as C
import os

MAX_DOF = 7
MAX_NX = 4 * MAX_DOF
MAX_NU = 2 * MAX_DOF
MAX_COST_TERMS = 6
NUM_ALPHAS = 10
LOG_FIELDS = 8
COUNTER_INTS = 40
STATIC_SCRATCH_BYTES = 64 << 20

(COST_NONE, COST_FRAME_PLACEMENT, COST_STATE, COST_CONTROL, COST_STIFFNESS, COST_DOUBLE_PENDULUM,
 COST_RESIDUAL_DOUBLE_PENDULUM) = range(7)
DYN_SEA, DYN_VSA = 0, 1
SOLVER_DDP, SOLVER_FDDP, SOLVER_BOXDDP = 0, 1, 2
OCP_RUNNING, OCP_CONVERGED, OCP_MAXITER, OCP_REG_MAX = 0, 1, 2, 3

c_double_p = C.POINTER(C.c_double)


class SE3(C.Structure):
    _fields_ = [("R", C.c_double * 9), ("p", C.c_double * 3)]


class Joint(C.Structure):
    _fields_ = [("placement", SE3), ("axis", C.c_double * 3), ("mass", C.c_double), ("com", C.c_double * 3),
                ("inertia", C.c_double * 9)]


class Robot(C.Structure):
    _fields_ = [("n", C.c_int32), ("reserved", C.c_int32), ("gravity", C.c_double * 3), ("joints", Joint * MAX_DOF)]


class Dynamics(C.Structure):
    _fields_ = [("kind", C.c_int32), ("nu", C.c_int32), ("write_fu", C.c_int32), ("has_control_limits", C.c_int32),
                ("K", C.c_double * (MAX_DOF * MAX_DOF)), ("Binv", C.c_double * (MAX_DOF * MAX_DOF)),
                ("Sm", C.c_double * (MAX_DOF * MAX_DOF)), ("u_none", C.c_double * MAX_NU),
                ("u_lb", C.c_double * MAX_NU), ("u_ub", C.c_double * MAX_NU)]


class CostTerm(C.Structure):
    _fields_ = [("kind", C.c_int32), ("frame_parent", C.c_int32), ("weight", C.c_double),
                ("ref", C.c_double * MAX_NX), ("act", C.c_double * MAX_NX), ("frame", SE3)]


class CostSum(C.Structure):
    _fields_ = [("nterms", C.c_int32), ("reserved", C.c_int32), ("terms", CostTerm * MAX_COST_TERMS)]


class ProblemDesc(C.Structure):
    _fields_ = [("robot", Robot), ("dyn", Dynamics), ("running", CostSum), ("terminal", CostSum),
                ("dt", C.c_double), ("dt_terminal", C.c_double)]


class Batch(C.Structure):
    _fields_ = [("B", C.c_int32), ("T", C.c_int32), ("x0", C.c_void_p), ("K_ocp", C.c_void_p),
                ("mref_ocp", C.c_void_p)]


class KnotOut(C.Structure):
    _fields_ = [("xnext", C.c_void_p), ("cost", C.c_void_p), ("deriv", C.c_void_p), ("xout", C.c_void_p),
                ("dFx", C.c_void_p), ("dFu", C.c_void_p)]


class SolverOpts(C.Structure):
    _fields_ = [("kind", C.c_int32), ("maxiter", C.c_int32), ("is_feasible", C.c_int32), ("poll_every", C.c_int32),
                ("th_stop", C.c_double), ("reg_init", C.c_double), ("reg_min", C.c_double), ("reg_max", C.c_double),
                ("reg_incfactor", C.c_double), ("reg_decfactor", C.c_double), ("th_grad", C.c_double),
                ("th_stepdec", C.c_double), ("th_stepinc", C.c_double), ("th_acceptstep", C.c_double),
                ("th_acceptnegstep", C.c_double)]


class OcpState(C.Structure):
    _fields_ = [(k, C.c_double) for k in ("cost", "cost_try", "stop", "xreg", "steplength", "dV", "dVexp", "d1", "d2",
                                          "dg", "dq", "ffeas")] + \
               [(k, C.c_int32) for k in ("iter", "status", "is_feasible", "was_feasible", "recalc", "accepted",
                                         "src_slot", "n_backward_retries", "full_steps", "n_logged", "ls_done",
                                         "reserved0")]


OCP_STATE_DOUBLES = C.sizeof(OcpState) // 8
assert C.sizeof(OcpState) == 12 * 8 + 12 * 4


class SolverBuffers(C.Structure):
    _names = ("xs", "us", "xnext", "knot_cost", "deriv", "fs", "Kgain", "kff", "Vx", "Vxx", "Vxxfs", "Qu", "Quu",
              "Quuk", "xs_try", "us_try", "trial", "log", "state", "counters")
    _fields_ = [(k, C.c_void_p) for k in _names]


class SolveResult(C.Structure):
    _fields_ = [("iterations_launched", C.c_int32), ("n_converged", C.c_int32), ("n_running", C.c_int32),
                ("kernel_launches", C.c_int32)]


class HostSolveIO(C.Structure):
    """AslrHostSolveIO: host pointers of aslr_solve_host."""
    _fields_ = [("B", C.c_int32), ("T", C.c_int32)] + \
               [(k, C.c_void_p) for k in ("x0", "K_ocp", "mref_ocp", "xs_init", "us_init", "xs", "us", "state", "Kgain",
                                          "kff", "log")]


def record_doubles(nx, nu):
    return 2 * nx * nx + 2 * nx * nu + nu * nu + nx + nu


_PROTOS = {
    "aslr_last_error": (C.c_char_p, []),
    "aslr_version": (C.c_int, []),
    "aslr_device_count": (C.c_int, []),
    "aslr_problem_create": (C.c_int, [C.POINTER(ProblemDesc), C.POINTER(C.c_void_p)]),
    "aslr_problem_destroy": (C.c_int, [C.c_void_p]),
    "aslr_problem_dims": (C.c_int, [C.c_void_p] + [C.POINTER(C.c_int32)] * 4),
    "aslr_record_doubles": (C.c_int, [C.c_int32, C.c_int32]),
    "aslr_default_solver_opts": (None, [C.c_int32, C.POINTER(SolverOpts)]),
    "aslr_calc": (C.c_int, [C.c_void_p, C.POINTER(Batch), C.c_void_p, C.c_void_p, C.POINTER(KnotOut), C.c_void_p]),
    "aslr_calc_diff": (C.c_int, [C.c_void_p, C.POINTER(Batch), C.c_void_p, C.c_void_p, C.POINTER(KnotOut),
                                 C.c_void_p]),
    "aslr_quasi_static": (C.c_int, [C.c_void_p, C.POINTER(Batch), C.c_void_p, C.c_void_p, C.c_int32, C.c_double,
                                    C.c_void_p]),
    "aslr_shift_warm_start": (C.c_int, [C.c_void_p, C.POINTER(Batch), C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p]),
    "aslr_solver_init": (C.c_int, [C.c_void_p, C.POINTER(Batch), C.POINTER(SolverOpts), C.POINTER(SolverBuffers),
                                   C.c_void_p]),
    "aslr_solver_calc_diff": (C.c_int, [C.c_void_p, C.POINTER(Batch), C.POINTER(SolverOpts),
                                        C.POINTER(SolverBuffers), C.c_int32, C.c_void_p]),
    "aslr_backward_pass": (C.c_int, [C.c_void_p, C.POINTER(Batch), C.POINTER(SolverOpts), C.POINTER(SolverBuffers),
                                     C.c_void_p]),
    "aslr_forward_pass": (C.c_int, [C.c_void_p, C.POINTER(Batch), C.POINTER(SolverOpts), C.POINTER(SolverBuffers),
                                    C.c_void_p]),
    "aslr_accept_step": (C.c_int, [C.c_void_p, C.POINTER(Batch), C.POINTER(SolverOpts), C.POINTER(SolverBuffers),
                                   C.c_int32, C.c_void_p]),
    "aslr_solve": (C.c_int, [C.c_void_p, C.POINTER(Batch), C.POINTER(SolverOpts), C.POINTER(SolverBuffers),
                             C.POINTER(SolveResult), C.c_void_p]),
    "aslr_solve_host": (C.c_int, [C.c_void_p, C.POINTER(HostSolveIO), C.POINTER(SolverOpts), C.POINTER(SolveResult)]),
    "aslr_solve_static": (C.c_int, [C.c_void_p, C.POINTER(Batch), C.POINTER(SolverOpts)] + [C.c_void_p] * 6),
}

EXPORTED_SYMBOLS = tuple(sorted(_PROTOS))
LIB_NAME = "libaslr_b200.so"
LIB_PATH = os.environ.get("ASLR_B200_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), LIB_NAME)
# (ASLR_B200_LIB: tuning builds of the same library, tools/build_variant.sh)

_lib = None


class AslrError(RuntimeError):
    pass


def bind(lib):
    for name, (res, args) in _PROTOS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    return lib


def load_library():
    """Loads libaslr_b200.so (built in-tree by __graft_entry__.build()).  Raises if it is missing."""
    global _lib
    if _lib is None:
        if not os.path.isfile(LIB_PATH):
            raise AslrError("%s is not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                            "(there is no CPU fallback)" % LIB_PATH)
        _lib = bind(C.CDLL(LIB_PATH))
    return _lib


def check(rc):
    if rc != 0:
        msg = load_library().aslr_last_error()
        raise AslrError("libaslr_b200 error %d: %s" % (rc, msg.decode() if msg else "?"))
