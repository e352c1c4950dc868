"""ORACLE / TEST INFRASTRUCTURE ONLY -- puts the NumPy shims (oracle/shims) and, when it exists,
the reference's own Python package (/root/reference/python) on sys.path so that the UNMODIFIED
reference model code runs on top of the shims.  /root/reference is absent on the GPU box: callers
must check `have_reference()`; everything that runs there uses oracle/aslr_port.py instead."""
import os
import sys

_HERE = os.path.dirname(os.path.abspath(__file__))
SHIMS = os.path.join(_HERE, "shims")
REFERENCE_PY = "/root/reference/python"


def activate_shims():
    if SHIMS not in sys.path:
        sys.path.insert(0, SHIMS)


def have_reference():
    return os.path.isfile(os.path.join(REFERENCE_PY, "aslr_to", "__init__.py"))


def import_reference():
    """Returns the reference's `aslr_to` module running on the shims (this container only)."""
    activate_shims()
    if not have_reference():
        raise ImportError("/root/reference is not present on this machine")
    if REFERENCE_PY not in sys.path:
        sys.path.insert(0, REFERENCE_PY)
    import aslr_to
    return aslr_to
