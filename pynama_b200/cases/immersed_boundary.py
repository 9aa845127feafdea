"""Immersed-boundary cases: flow past bodies on a FreeSlip mesh with a velocity correction step.

Reference: src/cases/immersed_boundary.py -- `ImmersedBoundaryStatic` (:19-345) and `ImmersedBoundaryDynamic`
(:347-433).  Per time step the reference solves the KLE, then corrects the velocity so that it matches the body's at
the Lagrangian markers (computeVelocityCorrection, :205-211):

    rhs = H v - V_body ;  A q = -rhs ;  v += S q           H = delta-function interpolation (h^2), S = H^T dl h, A = H S

and takes the curl of the corrected velocity as the new vorticity.  Here the coupling matrices are built on the device
(csrc/ibm.cu: which Euler nodes are close to which marker, the delta values, and A as a sparse Gram matrix -- the
reference's Python double loops, which the dynamic case repeats every step), H, S, A are device CSR matrices applied
through the stream SpMV, and A is solved by the same KspSolver as everything else (the reference's default
`preonly + lu`, served by device CG to round-off: A is small, symmetric positive definite and well conditioned).
"""
from math import cos, radians, sin, sqrt, pi, ceil, erf, exp  # noqa: F401  (`longRef` strings are eval'ed: 'sqrt(2)')

import numpy as np
import torch

from .. import backend as be
from ..domain.immersed_body import BodiesContainer
from ..petsc_like import Mat
from ..solver.ksp_solver import KspSolver
from .base_problem import FreeSlip


class ImmersedBoundaryStatic(FreeSlip):
    closeMode = 0      # computeClose of the static case (immersed_boundary.py:325-336)

    def setUp(self):
        if self.comm.size > 1:
            raise Exception("immersed-boundary cases run on one GPU (the markers are not partitioned)")
        super().setUp()
        self.setUpBodies()
        cells = self.getAffectedCells(10)
        self.collectedNodes, self.maxNodesPerLag = self.collectNodes(cells)
        self.totalEulerNodes = self.dom.getTotalNodes()
        self.createEmptyIBMMatrix()
        self.mNodes = self.buildIBMMatrix()

    def readBoundaryCondition(self, bc):
        self.nu = self.mu / self.rho
        try:
            re = bc['constant']['re']
            self.logger.info(f"Material Properties: density={self.rho} ; viscosity={self.mu}")
            directionAngle = bc['constant']['direction']
            angleRadian = radians(directionAngle)
            L = eval(bc['constant']['longRef'])
            velRef = re * (self.mu / self.rho) / L
            self.U_ref = velRef
            self.cteValue = [cos(angleRadian) * velRef, sin(angleRadian) * velRef]
            self.re = re
        except Exception:
            vel = bc['constant']['vel']
            self.U_ref = vel[0]
            self.cteValue = [self.U_ref, 0]
            self.re = self.U_ref / self.nu
        self.logger.info(f"Velocity Free Stream: {self.cteValue}")

    def setUpDomain(self):
        super().setUpDomain()
        if self.meshType == 'box-mesh':
            nelem = self.opts['nelem'] if 'nelem' in self.opts else self.config['domain']['box-mesh']['nelem']
            lower, upper = self.dom.getBoundingBox()
            height = upper[1] - lower[0]                 # (sic: immersed_boundary.py:52)
            self.h = (height / nelem[0]) / (self.ngl - 1)
        elif self.meshType == 'gmsh':
            self.h = self.config['domain']['h-min'] / (self.ngl - 1)
        else:
            raise Exception("Mesh not defined")

    def setUpBodies(self):
        if 'bodies' not in self.config:
            raise Exception("Bodies not defined")
        self.body = BodiesContainer(self.config['bodies'])
        self.logger.info(f"Node separation: {self.h}")
        self.body.createBodies(self.h)
        self.body.setVelRef(self.U_ref)

    # -- coupling matrices ------------------------------------------------------------------------------------
    def getAffectedCells(self, xSide, ySide=None, center=[0, 0]):
        """Cells whose centroid lies inside the box |x - cx| < xSide, |y - cy| < ySide (immersed_boundary.py:279-292):
        a device mask over all cells."""
        ySide = ySide or xSide
        corners = self.dom.deviceCellCorners().view(-1, 4, 2)
        cen = corners.mean(dim=1)
        keep = ((cen[:, 0] - center[0]).abs() < xSide) & ((cen[:, 1] - center[1]).abs() < ySide)
        return torch.nonzero(keep).view(-1)

    def collectNodes(self, cells):
        """Close Euler nodes and delta values of every marker (immersed_boundary.py:258-277, 318-338) on the device.
        Returns ((ptr, nodes, dirac) device arrays, the largest number of Euler nodes of a marker)."""
        e2n = self.dom.deviceCellNodes().view(self.dom.getTotalCells(), -1)
        cand = torch.unique(e2n[cells].reshape(-1)).to(torch.int32)          # ascending node ids
        self._ibmCandidates = cand
        lagxy = be.to_device(self.body.coordinatesArray().reshape(-1))
        p, nodes, dirac = be.ibm_collect(lagxy, cand, self.dom.deviceNodeCoords().view(-1), self.body.getElementLong(),
                                         self.h, mode=self.closeMode, kind=self.body.deltaName())
        counts = p[1:] - p[:-1]
        if int(counts.min()) == 0:
            raise Exception("Lag Node without Euler")
        return (p, nodes, dirac), int(counts.max())

    def createEmptyIBMMatrix(self):
        self.H = self.S = None           # sized by buildIBMMatrix (the device lists already know their lengths)

    def buildIBMMatrix(self):
        """H, S = H^T dl h, H scaled by h^2, A = H S and its solver (immersed_boundary.py:222-251).  Returns the Euler
        nodes that carry a non-zero delta value."""
        p, nodes, dirac = self.collectedNodes
        dl, h, dim = self.body.getElementLong(), self.h, self.dim
        H = be.dof_blocked_csr(p, nodes, dirac, self.totalEulerNodes, dim, name="H")
        rows = torch.repeat_interleave(torch.arange(H.shape[0], device=nodes.device), H.indptr[1:] - H.indptr[:-1])
        S = be.coo_to_csr(H.indices, rows.to(torch.int32), H.data * (dl * h), (H.shape[1], H.shape[0]), name="S")
        H.data.mul_(h ** 2)
        ap, aj, av = be.ibm_gram(p, nodes, dirac, h ** 2, dl * h)
        A = be.dof_blocked_csr(ap, aj, av, p.numel() - 1, dim, name="A")
        self.H, self.S, self.A = Mat.fromCSR(H), Mat.fromCSR(S), Mat.fromCSR(A)
        self.ksp = KspSolver()
        self.ksp.createSolver(self.A, self.comm)
        self.logger.info("IBM Matrices builded")
        return torch.unique(nodes[dirac > 0])

    def getMarkedNodes(self):
        return torch.unique(self.collectedNodes[1])

    # -- solver -----------------------------------------------------------------------------------------------
    def setUpSolver(self):
        super().setUpSolver()
        self.vel_correction = self.vel.duplicate()
        self.vel_correction.setName("velocity_correction")
        self.vort_correction = self.vort.duplicate()
        self.vort_correction.setName("vorticity_correction")
        self.virtualFlux = self.S.createVecRight()
        self.virtualFlux.setName("virtual_flux")
        self.ibm_rhs = self.S.createVecRight()
        self._bcDofs = None

    def applyBoundaryConditions(self, time):
        self.vel.set(0.0)
        if self._bcDofs is None:
            nodes = np.array(sorted(self.bcNodes), dtype=np.int64)
            dofs = (nodes[:, None] * self.dim + np.arange(self.dim)).reshape(-1)
            self._bcDofs = (be.to_device(dofs), be.to_device(np.tile(np.asarray(self.cteValue, float), len(nodes))),
                            be.to_device(nodes))
        be.vec_scatter_set(self.vel.data, self._bcDofs[0], self._bcDofs[1])
        be.vec_scatter_const(self.vort.data, self._bcDofs[2], 0.0)

    def computeVelocityCorrection(self, **kwargs):
        bodyVel = self.body.getVelocity()
        self.H.mult(self.vel, self.ibm_rhs)
        self.ibm_rhs.axpy(-1, bodyVel)
        self.ibm_rhs.scale(-1.0)                              # ksp.solve(-ibm_rhs, ...)
        self.ksp.solve(self.ibm_rhs, self.virtualFlux)
        self.ibm_rhs.scale(-1.0)
        self.S.mult(self.virtualFlux, self.vel_correction)
        self.vel.axpy(1.0, self.vel_correction)

    def computeDragForce(self, dt):
        denom = 0.5 * (self.U_ref ** 2)
        return self.body.computeForce(self.virtualFlux, denom * dt)

    def computeInitialCondition(self, startTime=None):
        startTime = self.ts.getTime() if startTime is None else startTime
        self.vort.set(0.0)
        self.ibmZone = self.vort.duplicate()
        self.ibmZone.set(0.0)
        self.ibmZone.setName("ibm-zone")
        self.affectedNodes = self.ibmZone.duplicate()
        self.affectedNodes.set(0.0)
        self.affectedNodes.setName("affected-nodes")
        self.body.setVelRef(self.U_ref)
        self.solveKLE(startTime, self.vort)
        self.computeVelocityCorrection(t=startTime)
        self.operator.Curl.mult(self.vel, self.vort)

    def markAffectedNodes(self, nodes):
        self.affectedNodes.set(0.0)
        be.vec_scatter_const(self.affectedNodes.data, nodes.to(torch.int64), 1.0)

    def markZone(self, nodes):
        be.vec_scatter_const(self.ibmZone.data, nodes.to(torch.int64), 1.0)

    def startSolver(self):
        self.computeInitialCondition()
        self.ts.setSolution(self.vort)
        saveSteps = self.config['save-n-steps']
        self.body.updateVelocity()
        maxSteps = self.ts.getMaxSteps()
        self.markAffectedNodes(self.mNodes)
        self.markZone(self.getMarkedNodes())
        for i in range(maxSteps + 1):
            self.ts.step()
            step = self.ts.getStepNumber()
            time = self.ts.time
            sol = self.ts.getSolution()
            self.solveKLE(time, sol)
            self.computeVelocityCorrection()
            self.operator.Curl.mult(self.vel, self.vort)
            dt = self.ts.getTimeStep()
            self.ts.setSolution(self.vort)
            self.logger.info(f"Converged: Step {step:4} | Time {time:.4e} | DT: {dt:.4e}  ")
            if time > self.ts.getMaxTime():
                break
            elif i % saveSteps == 0:
                self.viewer.saveData(step, time, self.vort, self.vel, self.ibmZone, self.affectedNodes)
                self.viewer.writeXmf(self.caseName)
                self.H.mult(self.vel, self.ibm_rhs)

    @staticmethod
    def computeCentroid(corners):
        return np.mean(corners, axis=0)


class ImmersedBoundaryDynamic(ImmersedBoundaryStatic):
    closeMode = 1      # computeClose of the dynamic case (immersed_boundary.py:391-402)

    def startSolver(self):
        self.computeInitialCondition()
        self.ts.setSolution(self.vort)
        maxSteps = self.ts.getMaxSteps()
        self.body.viewBodies()
        saveSteps = self.config['save-n-steps']
        self.markZone(self.getMarkedNodes())
        self.history = {"times": [], "cd": [], "cl": [], "dt": [], "steps": []}
        for step in range(1, maxSteps + 1):
            self.ts.step()
            time = self.ts.time
            dt = self.ts.getTimeStep()
            nods = self.computeVelocityCorrection(time)
            self.markAffectedNodes(nods)
            self.operator.Curl.mult(self.vel, self.vort)
            self.ts.setSolution(self.vort)
            self.logger.info(f"Converged: Step {step:4} | Time {time:.4e} | Increment {dt:.3e}")
            if time > self.ts.getMaxTime():
                break
            elif step % saveSteps == 0:
                self.viewer.saveData(step, time, self.vort, self.vel, self.ibmZone, self.affectedNodes)
                self.viewer.writeXmf(self.caseName)
                self.H.mult(self.vel, self.ibm_rhs)
            if step % max(1, int(saveSteps / 10)) == 0:
                cd, cl = self.computeDragForce(dt)
                for k, v in (("times", time), ("cd", cd), ("cl", cl), ("dt", dt), ("steps", step)):
                    self.history[k].append(v)

    def computeInitialCondition(self, startTime=None):
        startTime = self.ts.getTime() if startTime is None else startTime
        super().computeInitialCondition(startTime)

    def computeVelocityCorrection(self, t=None, **kwargs):
        """The body moves: new marker positions, new coupling matrices, then the correction (:405-414)."""
        if t is not None:
            self.body.updateBodyParameters(t)
        affNodes = self.rebuildMatrix()
        super().computeVelocityCorrection()
        return affNodes

    def rebuildMatrix(self):
        """immersed_boundary.py:416-433: everything that depends on the marker positions, on the device."""
        lagxy = be.to_device(self.body.coordinatesArray().reshape(-1))
        p, nodes, dirac = be.ibm_collect(lagxy, self._ibmCandidates, self.dom.deviceNodeCoords().view(-1),
                                         self.body.getElementLong(), self.h, mode=self.closeMode,
                                         kind=self.body.deltaName())
        counts = p[1:] - p[:-1]
        if int(counts.min()) == 0:
            raise Exception("Lag Node without Euler")
        self.collectedNodes, self.maxNodesPerLag = (p, nodes, dirac), int(counts.max())
        return self.buildIBMMatrix()
