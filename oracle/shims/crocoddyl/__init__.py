"""ORACLE / TEST INFRASTRUCTURE ONLY -- NumPy stand-in for the `crocoddyl` Python module, exposing
exactly the names spykspeigel/aslr_to uses.  Never imported by the product
package.  PARITY UNPINNED for the solver side (see solvers.py)."""
from .core import (Jcomponent, StateAbstract, ActuationModelAbstract, ActuationDataAbstract,  # noqa: F401
                   DifferentialActionModelAbstract, DifferentialActionDataAbstract, ActionModelAbstract,
                   ActionDataAbstract, DataCollectorMultibody, ActivationDataAbstract, ActivationModelQuad,
                   ActivationModelWeightedQuad, ResidualModelAbstract, ResidualDataAbstract, ResidualModelState,
                   ResidualModelControl)
from .costs import (CostModelAbstract, CostDataAbstract, CostModelResidual, CostModelSum, CostDataSum,  # noqa: F401
                    DifferentialActionModelNumDiff)
from .problem import ShootingProblem, CallbackLogger, CallbackVerbose  # noqa: F401
from .solvers import SolverAbstract, SolverDDP, SolverFDDP, SolverBoxDDP, BoxQP, BoxQPSolution  # noqa: F401
