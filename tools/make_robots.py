#!/usr/bin/env python3
"""Generate the robot fixture files under aslr_to_b200/robots/.

The reference loads its robots from URDFs shipped by example-robot-data
(`asr_twodof`, `double_pendulum`, `talos_arm`: /root/reference/examples/two_dof_sea.py:18,
examples/double_pendulum.py:13, unittest/test_asr_free_fwddyn.py:14).  None of those URDFs is
available in this build, so the robots are defined here as plain data:
serial chains of revolute joints.  The numbers are OURS, not the authors'; every comparison in
this repository is oracle-vs-kernel on these same fixtures.

Per joint (joint i's parent is joint i-1, joint 0's parent is the fixed universe):
  placement  R (3x3), p (3)   pose of the joint frame in the parent joint frame at q_i = 0
  axis       unit 3-vector    rotation axis in the joint frame
  mass, com (3), inertia (3x3 about the COM, joint-frame axes)
Per frame: parent joint index (0-based, chain order) + placement in that joint's frame.
"""
import json
import os

import numpy as np

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "aslr_to_b200", "robots")


def rot(axis, ang):
    axis = np.asarray(axis, float)
    axis = axis / np.linalg.norm(axis)
    K = np.array([[0, -axis[2], axis[1]], [axis[2], 0, -axis[0]], [-axis[1], axis[0], 0]])
    return np.eye(3) + np.sin(ang) * K + (1 - np.cos(ang)) * (K @ K)


def joint(name, R, p, axis, mass, com, inertia, lo=-np.pi, hi=np.pi):
    inertia = np.asarray(inertia, float)
    assert np.allclose(inertia, inertia.T) and np.all(np.linalg.eigvalsh(inertia) > 0)
    return dict(name=name, placement=dict(R=np.asarray(R, float).tolist(), p=list(map(float, p))),
                axis=list(map(float, axis)), mass=float(mass), com=list(map(float, com)),
                inertia=inertia.tolist(), q_min=float(lo), q_max=float(hi))


def frame(name, parent, R, p):
    return dict(name=name, parent=int(parent), placement=dict(R=np.asarray(R, float).tolist(), p=list(map(float, p))))


def dump(robot):
    path = os.path.join(OUT, robot["name"] + ".json")
    with open(path, "w") as f:
        json.dump(robot, f, indent=1)
    print("wrote", path)


def asr_twodof():
    # planar 2-link arm in the plane z = 0.18, joints about z; reach 0.255 m (the reference's
    # commented-out target x = -0.255 suggests that reach, examples/two_dof_sea.py:35).
    I3 = np.eye(3)
    j = [
        joint("joint1", I3, [0, 0, 0.18], [0, 0, 1], 0.5, [0.065, 0.0, 0.0],
              [[1.0e-4, 1e-5, 0], [1e-5, 7.5e-4, 0], [0, 0, 7.5e-4]]),
        joint("joint2", I3, [0.13, 0, 0], [0, 0, 1], 0.3, [0.06, 0.0, 0.0],
              [[5.0e-5, 0, 0], [0, 4.0e-4, 2e-6], [0, 2e-6, 4.0e-4]]),
    ]
    f = [frame("universe", -1, I3, [0, 0, 0]), frame("link1", 0, I3, [0, 0, 0]),
         frame("link2", 1, I3, [0, 0, 0]), frame("EE", 1, I3, [0.125, 0, 0])]
    return dict(name="asr_twodof", gravity=[0, 0, -9.81], joints=j, frames=f)


def double_pendulum():
    # two links hanging along -z... joints about y; q = 0 points the links along +z.
    I3 = np.eye(3)
    j = [
        joint("joint1", I3, [0, 0, 0], [0, 1, 0], 0.2, [0, 0, 0.05],
              [[1.8e-4, 0, 0], [0, 1.8e-4, 0], [0, 0, 2e-5]]),
        joint("joint2", I3, [0, 0, 0.1], [0, 1, 0], 0.3, [0, 0, 0.1],
              [[1.0e-3, 0, 0], [0, 1.0e-3, 0], [0, 0, 3e-5]]),
    ]
    f = [frame("universe", -1, I3, [0, 0, 0]), frame("link1", 0, I3, [0, 0, 0]),
         frame("link2", 1, I3, [0, 0, 0]), frame("tip", 1, I3, [0, 0, 0.2])]
    return dict(name="double_pendulum", gravity=[0, 0, -9.81], joints=j, frames=f)


def chain7():
    # synthetic 7-link serial arm (BASELINE.json configs[4]; mirrors talos_arm's nq = 7, nx = 28)
    # roll-pitch-roll-pitch...: axes alternate z / y, links 0.3 m along z, off-axis COMs and a
    # couple of tilted joint placements so no term of the spatial algebra is trivially zero.
    masses = [3.0, 2.7, 2.4, 2.0, 1.6, 1.3, 1.0]
    j = []
    for i in range(7):
        axis = [0, 0, 1] if i % 2 == 0 else [0, 1, 0]
        R = np.eye(3)
        if i == 2:
            R = rot([1, 0, 0], 0.1)
        if i == 5:
            R = rot([1, 1, 0], -0.15)
        p = [0, 0, 0.0] if i == 0 else [0.01 * (i % 3), 0, 0.3]
        m = masses[i]
        com = [0.02, -0.01 * (1 + i % 2), 0.15]
        a, b, c = m * 0.3 ** 2 / 12, m * 0.3 ** 2 / 12 * 0.9, m * 0.09 ** 2
        inertia = np.array([[a, 0.02 * a, 0.01 * a], [0.02 * a, b, -0.015 * a], [0.01 * a, -0.015 * a, c]])
        j.append(joint("joint%d" % (i + 1), R, p, axis, m, com, inertia, -2.5, 2.5))
    f = [frame("universe", -1, np.eye(3), [0, 0, 0])]
    f += [frame("link%d" % (i + 1), i, np.eye(3), [0, 0, 0]) for i in range(7)]
    f.append(frame("tip", 6, rot([0, 1, 0], 0.2), [0.0, 0.02, 0.3]))
    return dict(name="chain7", gravity=[0, 0, -9.81], joints=j, frames=f)


def chain_n(n):
    """The first n joints of chain7 with the tip frame moved to the last kept joint: fixtures for every link
    DOF the kernels are instantiated for (the reference's classes work for any pinModel.nv,
    python/aslr_to/statemultibody_aslr.py:14-16)."""
    c = chain7()
    tip = c["frames"][-1]
    f = c["frames"][:n + 1] + [dict(tip, parent=n - 1)]
    return dict(name="chain%d" % n, gravity=c["gravity"], joints=c["joints"][:n], frames=f)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    for r in [asr_twodof(), double_pendulum(), chain7()] + [chain_n(n) for n in range(1, 7)]:
        dump(r)
