"""Uniform-flow sanity case (reference: src/cases/uniform.py): zero vorticity, constant velocity on the boundary,
the KLE solve must return the constant field (src/tests/test_solver.py:20-27)."""
import numpy as np

from .. import backend as be
from .base_problem import FreeSlip


class UniformFlow(FreeSlip):
    def setUp(self):
        self.setUpGeneral()
        self.setUpBoundaryConditions()
        self.setUpEmptyMats()
        self.buildKLEMats()
        self.buildOperators()
        if self.dim == 2:
            self.cteValue = [1, 0]
        else:
            raise Exception("Wrong dim")
        # (the reference builds everything a second time here, uniform.py:26-30; the result is identical)
        self._bcDofs = None

    def computeInitialCondition(self, startTime):
        self.vort.set(0.0)

    def applyBoundaryConditions(self, time):
        self.vel.set(0.0)
        if self._bcDofs is None:
            nodes = np.array(sorted(self.bcNodes), dtype=np.int64)
            dofs = (nodes[:, None] * self.dim + np.arange(self.dim)).reshape(-1)
            self._bcDofs = (be.to_device(dofs), be.to_device(np.tile(np.asarray(self.cteValue, float), len(nodes))))
        be.vec_scatter_set(self.vel.data, *self._bcDofs)

    def generateExactVecs(self, time=None):
        exactVel = self.mat.K.createVecRight()
        exactVort = self.mat.Rw.createVecRight()
        exactVel.setName(f"{self.caseName}-exact-vel")
        exactVort.setName(f"{self.caseName}-exact-vort")
        exactVel = self.dom.applyValuesToVec(self.dom.getAllNodes(), self.cteValue, exactVel)
        exactVort.set(0.0)
        return exactVel, exactVort
