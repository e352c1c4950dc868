"""ORACLE / TEST INFRASTRUCTURE ONLY -- empty stand-in so that the reference package's
module-level `import matplotlib.pyplot as plt` (python/aslr_to/__init__.py:15) succeeds.
Plotting is out of scope (no graphics stack: unavailable on this box)."""
