"""example_robot_data.load(name) (examples/two_dof_sea.py:19) on top of the robot fixtures.

The reference's URDFs (`asr_twodof`, `double_pendulum`) are not in its tree;
the fixtures under aslr_to_b200/robots/ are this repository's stand-ins (tools/make_robots.py)."""
import types

from . import pinocchio

_ALIASES = {"asr_twodof": "asr_twodof", "double_pendulum": "double_pendulum", "chain7": "chain7"}
_ALIASES.update(("chain%d" % n, "chain%d" % n) for n in range(1, 7))


def load(name):
    if name not in _ALIASES:
        raise ValueError("unknown robot %r (available: %s)" % (name, ", ".join(sorted(_ALIASES))))
    model = pinocchio.model_from_fixture(_ALIASES[name])
    return types.SimpleNamespace(model=model, nq=model.nq, nv=model.nv, q0=pinocchio.neutral(model))
