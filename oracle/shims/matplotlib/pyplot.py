"""ORACLE / TEST INFRASTRUCTURE ONLY -- see matplotlib/__init__.py."""


def __getattr__(name):
    def _unavailable(*a, **k):
        raise RuntimeError("matplotlib is not available in this build (plotting is out of scope)")
    return _unavailable
