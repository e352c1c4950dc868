"""ORACLE / TEST INFRASTRUCTURE ONLY -- pinocchio.utils.{zero,rand} as 1-D arrays
(statemultibody_aslr.py:20-29 concatenates them with 1-D configurations)."""
import numpy as np


def zero(n):
    return np.zeros(n)


def rand(n):
    return np.random.rand(n)
