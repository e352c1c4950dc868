"""The handful of pinocchio names the reference's examples touch, as plain data holders.

The rigid-body algorithms themselves (what pinocchio computes for the reference at
free_fwddyn_asr.py:38,75 and residual_frame_placement.py:14-22) run inside the CUDA kernels
(aslr_to_b200/csrc/knot.cuh); nothing here does arithmetic beyond composing constant transforms.
"""
import types

import numpy as np

from .. import desc as D


class SE3(object):
    """pinocchio.SE3(R, p) (examples/two_dof_sea.py:38)."""

    def __init__(self, rotation=None, translation=None):
        self.rotation = np.eye(3) if rotation is None else np.array(rotation, float).reshape(3, 3)
        self.translation = np.zeros(3) if translation is None else np.array(translation, float).reshape(3)

    @staticmethod
    def Identity():
        return SE3()

    def inverse(self):
        return SE3(self.rotation.T, -self.rotation.T @ self.translation)

    def __mul__(self, other):
        return SE3(self.rotation @ other.rotation, self.translation + self.rotation @ other.translation)

    def copy(self):
        return SE3(self.rotation, self.translation)

    def __repr__(self):
        return "SE3(R=%s, p=%s)" % (self.rotation.tolist(), self.translation.tolist())


class _Frame(object):
    def __init__(self, name, parent, placement):
        self.name, self.parent, self.placement = name, parent, placement


class Model(object):
    """pinocchio.Model of a serial chain of revolute joints, built from a robot fixture
    (aslr_to_b200/robots/*.json)."""

    def __init__(self, robot_dict):
        self._dict = robot_dict
        self.name = robot_dict.get("name", "robot")
        self.nq = self.nv = self.njoints_actuated = len(robot_dict["joints"])
        self.gravity = types.SimpleNamespace(linear=np.array(robot_dict["gravity"], float))
        # the fixtures list "universe" first, like pinocchio; `parent` is the 0-based joint (-1 = universe)
        self.frames = [_Frame(f["name"], int(f["parent"]), SE3(f["placement"]["R"], f["placement"]["p"]))
                       for f in robot_dict["frames"]]

    def getFrameId(self, name):
        for i, f in enumerate(self.frames):
            if f.name == name:
                return i
        return len(self.frames)  # pinocchio returns nframes for an unknown name

    def existFrame(self, name):
        return self.getFrameId(name) < len(self.frames)

    def createData(self):
        return types.SimpleNamespace()

    # --- used by the compat layer
    def _frame_desc(self, frame_id):
        if frame_id < 0 or frame_id >= len(self.frames) or self.frames[frame_id].parent < 0:
            raise ValueError("frame id %d does not name a body frame" % frame_id)
        f = self.frames[frame_id]
        return f.parent, f.placement.rotation, f.placement.translation


def neutral(model):
    return np.zeros(model.nq)


def randomConfiguration(model):
    return np.random.uniform(-np.pi, np.pi, model.nq)


def difference(model, q0, q1):
    return np.asarray(q1, float) - np.asarray(q0, float)


def integrate(model, q, dq):
    return np.asarray(q, float) + np.asarray(dq, float)


class _Utils(object):
    @staticmethod
    def zero(n):
        return np.zeros(n)

    @staticmethod
    def rand(n):
        return np.random.rand(n) * 2.0 - 1.0


utils = _Utils()


def model_from_fixture(name_or_path):
    return Model(D.load_robot_dict(name_or_path))
