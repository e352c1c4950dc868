"""ORACLE / TEST INFRASTRUCTURE ONLY -- crocoddyl.ShootingProblem and callbacks in NumPy
(call sites examples/two_dof_sea.py:66,75-78,90).
PARITY UNPINNED (no crocoddyl golden vectors exist in the reference)."""
import numpy as np


class _DataList(list):
    def tolist(self):
        return list(self)


class ShootingProblem(object):
    def __init__(self, x0, runningModels, terminalModel):
        self.x0 = np.array(x0, float)
        self.runningModels = list(runningModels)
        self.terminalModel = terminalModel
        self.T = len(self.runningModels)
        self.runningDatas = _DataList(m.createData() for m in self.runningModels)
        self.terminalData = terminalModel.createData()
        self.nthreads = 1
        self.cost = 0.0

    def calc(self, xs, us):
        for m, d, x, u in zip(self.runningModels, self.runningDatas, xs, us):
            m.calc(d, x, u)
        # the terminal node is evaluated WITHOUT a control: the Python override sees u=None
        self.terminalModel.calc(self.terminalData, xs[-1])
        self.cost = sum(d.cost for d in self.runningDatas) + self.terminalData.cost
        return self.cost

    def calcDiff(self, xs, us):
        for m, d, x, u in zip(self.runningModels, self.runningDatas, xs, us):
            m.calcDiff(d, x, u)
        self.terminalModel.calcDiff(self.terminalData, xs[-1])
        c = 0.0
        for d in self.runningDatas:
            c += d.cost
        c += self.terminalData.cost
        self.cost = c
        return c

    def rollout(self, us):
        xs = [self.x0.copy()]
        for m, d, u in zip(self.runningModels, self.runningDatas, us):
            m.calc(d, xs[-1], u)
            xs.append(np.array(d.xnext, float))
        self.terminalModel.calc(self.terminalData, xs[-1])
        return xs

    def quasiStatic(self, xs):
        return [np.array(m.quasiStatic(d, x), float) for m, d, x in zip(self.runningModels, self.runningDatas, xs)]


class CallbackLogger(object):
    def __init__(self):
        self.xs, self.us, self.fs = [], [], []
        self.steps, self.iters, self.costs = [], [], []
        self.u_regs, self.x_regs, self.stops, self.grads = [], [], [], []
        self.dVs, self.dVexps, self.ffeass = [], [], []

    def __call__(self, solver):
        self.xs = [np.array(x) for x in solver.xs]
        self.us = [np.array(u) for u in solver.us]
        self.fs = [np.array(f) for f in solver.fs]
        self.steps.append(solver.stepLength)
        self.iters.append(solver.iter)
        self.costs.append(solver.cost)
        self.u_regs.append(solver.u_reg)
        self.x_regs.append(solver.x_reg)
        self.stops.append(solver.stop)
        self.grads.append(-solver.d[1])
        self.dVs.append(solver.dV)
        self.dVexps.append(solver.dVexp)
        self.ffeass.append(solver.ffeas)


class CallbackVerbose(object):
    def __init__(self, level=0):
        self.level = level

    def __call__(self, solver):
        if self.level:
            print("%4d  %.5e  %.5e  %.5e  %.5e  %.4f  %d" % (solver.iter, solver.cost, solver.stop, -solver.d[1],
                                                             solver.x_reg, solver.stepLength, solver.isFeasible))
