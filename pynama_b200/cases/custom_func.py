"""Analytic-solution cases (reference: src/cases/custom_func.py): Taylor-Green vortex in 2-D on the device path.

The analytic fields are evaluated by a kernel over the device node coordinates (`pyn_taylor_green`) instead of the
reference's Python `map` over coordinates (custom_func.py:57-83, dmplex.py:263-272).  The other analytic families of
the reference (3-D Taylor-Green, senoidal, flat plate) and its plotting studies are outside the hot path.
"""
from math import cos, exp, pi, sin

import numpy as np

from .. import backend as be
from .base_problem import FreeSlip


class CustomFuncCase(FreeSlip):
    def setUp(self):
        self.setUpGeneral()
        self.setUpBoundaryConditions()
        self.setUpEmptyMats()
        self.buildKLEMats()
        self.buildOperators()
        self.nu = self.mu / self.rho
        if self.case == 'taylor-green':
            if self.dim != 2:
                raise Exception("3-D Taylor-Green is outside the B200 hot path (SURVEY.md 8f item 3)")
            self.velFunction = self.taylorGreenVel_2D
            self.vortFunction = self.taylorGreenVort_2D
        else:
            print(self.case)
            raise Exception("Case not found")
        self._bcIdx = be.to_device(np.array(sorted(self.bcNodes), dtype=np.int64))

    def _fields(self, time, vel=None, vort=None, idx=None):
        be.taylor_green(self.dom.deviceNodeCoords(), self.nu, time, None if vel is None else vel.data,
                        None if vort is None else vort.data, idx)

    def computeInitialCondition(self, startTime):
        self._fields(startTime, vort=self.vort)

    def generateExactVecs(self, time):
        exactVel = self.mat.K.createVecRight()
        exactVort = self.mat.Rw.createVecRight()
        exactVel.setName(f"{self.caseName}-exact-vel")
        exactVort.setName(f"{self.caseName}-exact-vort")
        self._fields(time, exactVel, exactVort)
        return exactVel, exactVort

    def applyBoundaryConditions(self, time):
        """vel = 0 then exact velocity on the boundary nodes; the boundary entries of `self.vort` (the time
        integrator's solution vector) are overwritten with the exact vorticity (custom_func.py:76-83)."""
        self.vel.set(0.0)
        self._fields(time, self.vel, self.vort, self._bcIdx)

    @staticmethod
    def taylorGreenVel_2D(coord, nu, t=None):
        x_, y_ = 2 * pi * coord[0], 2 * pi * coord[1]
        expon = exp(-4 * (pi ** 2) * nu * t * 2.0)
        return [cos(x_) * sin(y_) * expon, -sin(x_) * cos(y_) * expon]

    @staticmethod
    def taylorGreenVort_2D(coord, nu, t=None):
        x_, y_ = 2 * pi * coord[0], 2 * pi * coord[1]
        expon = exp(-4 * (pi ** 2) * nu * t * 2.0)
        return [-2 * pi * 2.0 * cos(x_) * cos(y_) * expon]
