"""Case layer: `BaseProblem`, `NoSlipFreeSlip`, `FreeSlip` (reference: src/cases/base_problem.py).

The public surface is the reference's -- constructor `(config, **kwargs)` with the same YAML schema and keyword
overrides (`ngl`, `nelem`, `lower`, `upper`, `case`, `chart`), `setUp()`, `setUpSolver()`, `startSolver()`,
`solveKLE(time, vort)`, `evalRHS(ts, t, Vort, f)`, `computeVtensV(vec=None)`, attributes `vel, vort, velFS, mat,
operator, solver, solverFS, dom, elemType, _VtensV, _Aux1, comm, logger, timer, timerTotal` -- but nothing here loops
over cells in Python: the three hot loops of the reference (`buildKLEMats`, `buildOperators`, `evalRHS`) are each a
handful of kernel launches on device-resident data (pynama_b200/csrc), and every vector stays in HBM.

    buildKLEMats   -> Mat/MatNS.assembleFromMesh      batched element integration + deterministic gather assembly
    buildOperators -> Operators.assembleFromMesh      same kernels, 1/W row scaling folded in
    solveKLE       -> SpMV(+add) launches + device CG with PETSc's KSP semantics
    evalRHS        -> 3 SpMV launches; the convective tensor, the 2*mu / -rho combination and the 1/rho scaling are
                      fused into the SpMV epilogues (base_problem.py:217-232 is 3 SpMV + 5 vector passes)
"""
import logging

import numpy as np

from .. import backend as be
from ..common.options import Options
from ..common.timer import Timer
from ..domain.dmplex import DMPlexDom
from ..elements.spectral import Spectral
from ..matrices.mat_generator import Mat, Operators
from ..matrices.mat_ns import MatNS
from ..petsc_like import Vec
from ..solver.ksp_solver import KspSolver
from ..solver.ts_solver import TsSolver
from ..viewer.paraviewer import Paraviewer

SPMV_SET, SPMV_ADD, SPMV_SCALE, SPMV_SRT_VTENSV = 0, 1, 2, 3


class _Comm(object):
    """Stand-in for MPI.COMM_WORLD: rank/size of the torch.distributed job when there is one, else a single rank."""

    def __init__(self, serial=False):
        self.rank, self.size = 0, 1
        self.dist = None            # pynama_b200.dist.DistContext once the mesh has been partitioned
        if serial:                  # a case replicated on every rank (the coarse levels of the multigrid hierarchy)
            return
        try:
            import torch.distributed as dist
            if dist.is_available() and dist.is_initialized():
                self.rank, self.size = dist.get_rank(), dist.get_world_size()
        except Exception:
            pass

    def Get_rank(self):
        return self.rank

    def Get_size(self):
        return self.size


class BaseProblem(object):
    def __init__(self, config, **kwargs):
        self.comm = _Comm(serial=bool(kwargs.pop("serial", False)))
        self.timerTotal = Timer()
        self.timerTotal.tic()
        self.timer = Timer()
        case = kwargs['case'] if 'case' in kwargs else Options().getString('case', 'uniform')
        self.config = config
        self.logger = logging.getLogger(f"[{self.comm.rank}] {self.config.get('name')}")
        self.case = case
        self.caseName = self.config.get("name")
        self.readDomainData(kwargs)
        self.readMaterialData()
        self.opts = kwargs
        if "chart" in kwargs:
            self.setUpTimeSolverTest()
        elif 'time-solver' in self.config:
            self.setUpTimeSolver()
        if 'boundary-conditions' in self.config:
            self.readBoundaryCondition(self.config.get("boundary-conditions"))

    # -- set-up -----------------------------------------------------------------------------------------------
    def setUp(self):
        self.setUpGeneral()
        self.setUpBoundaryConditions()
        self.setUpEmptyMats()
        self.buildKLEMats()
        self.buildOperators()

    def setUpDomain(self):
        domain = self.config.get("domain")
        self.dom = None
        if "box-mesh" in domain:
            self.logger.info("Creating dom with box Mesh")
            self.meshType = "box-mesh"
            ngl = self.opts['ngl'] if "ngl" in self.opts else domain['ngl']
            if self.comm.size > 1:      # one strip of cell rows per GPU (domain/partition.py)
                from ..domain.partition import StripPartition
                faces = self.opts['nelem'] if 'nelem' in self.opts else domain['box-mesh']['nelem']
                part = StripPartition(faces[0], faces[1], ngl, self.comm.rank, self.comm.size)
                self.dom = DMPlexDom(boxMesh=domain.get('box-mesh'), partition=part, **self.opts)
            else:
                self.dom = DMPlexDom(boxMesh=domain.get('box-mesh'), **self.opts)
        elif "gmsh-file" in domain:
            # unstructured quadrilaterals (the reference's createFromFile, src/domain/dmplex.py:22-23)
            from ..domain.quadmesh import QuadMeshDom
            if self.comm.size > 1:
                raise Exception("gmsh meshes run on one GPU (the strip partition needs a box mesh)")
            self.meshType = "gmsh"
            self.dom = QuadMeshDom(fileName=domain["gmsh-file"], comm=None)
        self.dim = self.dom.getDimension()
        self.dim_w = 1 if self.dim == 2 else 3
        self.dim_s = 3 if self.dim == 2 else 6
        self.ngl = self.opts['ngl'] if "ngl" in self.opts else domain['ngl']
        self.dom.setFemIndexing(self.ngl)
        if self.dom.partition is not None:
            from ..dist import DistContext
            self.comm.dist = DistContext(self.dom.partition, parent=self.opts.get("dist_parent"))
            if self.opts.get("dist_parent") is None:     # (a coarse level of a multigrid hierarchy leaves the case's own)
                from .. import petsc_like
                petsc_like.set_dist(self.comm.dist)      # Vec.norm / Vec.dot become global (owned entries, all-reduced)
            self.logger.info(self.dom.partition.describe())
        if not self.comm.rank:
            self.logger.info("DMPlex dom created")

    def setUpElement(self):
        self.elemType = Spectral(self.ngl, self.dim)
        if not self.comm.rank:
            self.logger.info(f"{self.dim}-D ngl:{self.ngl} Spectral element created")

    def setUpBoundaryConditions(self):
        self.dom.setLabelToBorders()
        self.dom.setBoundaryCondition()
        if not self.comm.rank:
            self.logger.info("Boundary Conditions setted up")

    def readMaterialData(self):
        materialData = self.config.get("material-properties")
        self.rho = materialData['rho']
        self.mu = materialData['mu']
        self.nu = self.mu / self.rho

    def readDomainData(self, kwargs):
        domain = self.config.get("domain")
        if "nelem" in kwargs:
            self.nelem = kwargs['nelem']
        elif "box-mesh" in domain:
            self.nelem = domain['box-mesh']['nelem']
            self.lower = domain['box-mesh']['lower']
            self.upper = domain['box-mesh']['upper']
        elif "gmsh-file" in domain:
            self.nelem = [0, 0]            # 2-D; the cell count comes from the file
            self.lower = self.upper = None
        else:
            raise Exception("No Gmsh Implemented")
        self.dim = len(self.nelem)
        self.dim_w = 1 if self.dim == 2 else 3
        self.dim_s = 3 if self.dim == 2 else 6
        self.ngl = kwargs['ngl'] if "ngl" in kwargs else domain['ngl']

    def createMesh(self, saveMesh=True):
        saveDir = self.config.get("save-dir")
        self.viewer = Paraviewer(self.dim, self.comm, saveDir)
        self.dom.computeFullCoordinates(self.elemType)
        if saveMesh:
            self.viewer.saveMesh(self.dom.fullCoordVec)
        if not self.comm.rank:
            self.logger.info("Mesh created")

    def setUpGeneral(self):
        self.setUpDomain()
        self.setUpElement()
        self.createMesh()
        self.bcNodes = self.dom.getNodesFromLabel("External Boundary")

    def buildOperators(self):
        """Curl, DivSrT, SrT with their lumped-mass row scaling (base_problem.py:132-140): cell 0's element
        operators for every cell, assembled on the device."""
        self.operator.assembleFromMesh(self.dom, self.elemType)
        self.operator.assembleAll()
        if not self.comm.rank:
            self.logger.info("Operators Matrices builded")

    # -- time solver ------------------------------------------------------------------------------------------
    def setUpTimeSolver(self):
        options = self.config.get("time-solver")
        self.ts = TsSolver(self.comm)
        self.ts.setUpTimes(options['start-time'], options['end-time'], options['max-steps'])
        self.ts.initSolver(self.evalRHS, self.convergedStepFunction)

    def setUpTimeSolverTest(self):
        options = self.config.get("time-solver")
        self.ts = TsSolver(self.comm)
        self.ts.setUpTimes(options['start-time'], options['end-time'], options['max-steps'])
        self.saveError2 = []
        self.saveError8 = []
        self.saveStep = []
        self.saveTime = []
        self.ts.initSolver(self.evalRHS, self.convergedStepFunctionKLET)

    def convergedStepFunction(self, ts):
        time, step, incr = ts.time, ts.step_number, ts.getTimeStep()
        self.viewer.saveData(step, time, self.vel, self.vort)
        self.viewer.writeXmf(self.caseName)
        if not self.comm.rank:
            self.logger.info(f"Converged: Step {step:4} | Time {time:.4e} | Increment Time: {incr:.2e} ")

    def convergedStepFunctionKLET(self, ts):
        time, step = ts.time, ts.step_number
        self.solveKLE(time, self.vort)
        incr = ts.getTimeStep()
        exactVel, exactVort = self.generateExactVecs(time)
        errorVel = exactVel - self.vel
        errorVort = exactVort - self.vort
        self.saveError2.append(errorVel.norm(norm_type=2))
        self.saveError8.append(errorVel.norm(norm_type=3))
        self.saveTime.append(time)
        self.saveStep.append(step)
        errorVort.setName("ErrorVort")
        errorVel.setName("ErrorVel")
        exactVort.setName("ExactVort")
        exactVel.setName("ExactVel")
        self.viewer.saveData(step, time, self.vel, self.vort, exactVel, exactVort, errorVel, errorVort)
        self.viewer.writeXmf(self.caseName)
        if not self.comm.rank:
            self.logger.info(f"Converged: Step {step:4} | Time {time:.4e} | Increment Time: {incr:.2e} ")

    # -- the vorticity right-hand side --------------------------------------------------------------------------
    def evalRHS(self, ts, t, Vort, f):
        """f = Curl (1/rho) DivSrT (2 mu SrT v - rho v(x)v)  with v from the KLE solve (base_problem.py:212-232).
        Three launches: the SrT product forms v(x)v from `vel` in its epilogue and combines, the DivSrT product scales
        by 1/rho, the Curl product writes f."""
        self.solveKLE(t, Vort)
        op = self.operator.spmv
        op["SrT"].csr.mult(self.vel.data, self._Aux1.data, mode=SPMV_SRT_VTENSV, p0=2.0 * self.mu, p1=self.rho,
                           aux=self.vel.data)
        self._halo(self._Aux1, self.dim_s)
        op["DivSrT"].csr.mult(self._Aux1.data, self._rhs.data, mode=SPMV_SCALE, p0=1 / self.rho)
        self._halo(self._rhs, self.dim)
        op["Curl"].csr.mult(self._rhs.data, f.data)

    def _halo(self, vec, ncomp):
        """Multi-GPU: refresh the ghost entries of a product result before it is multiplied again (the products only
        fill owned rows).  No-op on one GPU."""
        if self.comm.dist is not None:
            self.comm.dist.exchange(vec, ncomp)

    def computeVtensV(self, vec=None):
        """Node-wise convective tensor [vx^2, vx vy, vy^2] into `_VtensV` (base_problem.py:234-252).  The time loop
        does not call this (fused into evalRHS); it is kept for drivers/tests that do (src/tests/test_solver.py:66-87)."""
        src = self.vel if vec is None else vec
        be.vtensv(src.data, self._VtensV.data)

    def startSolver(self):
        initTime = self.ts.getTime()
        self.computeInitialCondition(startTime=initTime)
        self.ts.solve(self.vort)

    def solveKLE(self, time, vort):
        pass

    def buildKLEMats(self):
        pass

    def computeInitialCondition(self, startTime):
        pass

    def applyBoundaryConditions(self, time):
        pass

    def readBoundaryCondition(self, inputData=None):
        pass

    def setUpSolver(self):
        self.solver = KspSolver()
        self.solver.createSolver(self.mat.K, self.comm)
        if self.solver.wantsMultigrid():
            self.solver.setUpMultigrid(self, "K")
        self.vel = self.mat.K.createVecRight()
        self.vel.setName("velocity")
        self.vort = self.mat.Rw.createVecRight()
        self.vort.setName("vorticity")
        self.vort.set(0.0)
        nvel = self.vel.getSize()
        nstr = nvel * self.dim_s // self.dim
        self._VtensV = Vec(nstr)
        self._Aux1 = Vec(nstr)
        self._rhs = Vec(nvel)           # the reference allocates this inside every evalRHS call (:226)
        self._kleRhs = Vec(nvel)        # right-hand side of the KLE solves

    def setUpEmptyMats(self):
        self.mat = None
        self.operator = None

    def view(self):
        print(f"Case: {self.case}")
        print(f"Domain: {self.dom.nx} x {self.dom.ny} cells")
        print(f"NGL: {self.dom.getNGL()}")


class NoSlipFreeSlip(BaseProblem):
    """Walls with prescribed velocity: free-slip solve, vorticity correction, no-slip solve (base_problem.py:300-454)."""

    def setUpEmptyMats(self):
        self.mat = MatNS(self.dim, self.comm)
        self.operator = Operators(self.dim, self.comm)
        rStart, rEnd, d_nnz_ind, o_nnz_ind, ind_d, ind_o = self.dom.getMatIndices()
        self.globalNodesDIR = self.dom.getGlobalIndicesDirichlet()
        globalNodesNS = self.dom.getGlobalIndicesNoSlip()
        self.mat.createEmptyKLEMats(rStart, rEnd, d_nnz_ind, o_nnz_ind, ind_d, ind_o, self.globalNodesDIR,
                                    globalNodesNS)
        if not self.comm.rank:
            self.logger.info("Empty KLE Matrices created")
        self.operator.createAll(rStart, rEnd, d_nnz_ind, o_nnz_ind)
        if not self.comm.rank:
            self.logger.info("Empty Operators created")

    def setUpSolver(self):
        super().setUpSolver()
        self.solverFS = KspSolver()
        KFS = self.mat.K + self.mat.Kfs
        # Kfs only touches boundary-affected rows, so K + Kfs keeps K's row templates on the interior
        if self.mat.K.csr._tmat_asm is not None:
            KFS.csr.enable_templates(like=self.mat.K.csr)
        self.solverFS.createSolver(KFS, self.comm)
        if self.solverFS.wantsMultigrid():
            self.solverFS.setUpMultigrid(self, "KFS")
        self.velFS = self.vel.copy()
        self._vortFS = self.vort.duplicate()

    def solveKLE(self, time, vort):
        """base_problem.py:321-327: rhs = Rw w + Rwfs w + Krhsfs v_bc -> velFS ; wall values ; w' = Curl velFS ;
        rhs = Rw w' + Krhs v_bc -> vel.  Six SpMV launches (accumulating in place, same order of additions) and two
        device Krylov solves."""
        self.applyBoundaryConditions()
        self._halo(vort, self.dim_w)
        m, b = self.mat, self._kleRhs.data
        m.Rw.csr.mult(vort.data, b)
        m.Rwfs.csr.mult(vort.data, b, mode=SPMV_ADD)
        m.Krhsfs.csr.mult(self.vel.data, b, mode=SPMV_ADD)
        self.solverFS(self._kleRhs, self.velFS)
        self.applyBoundaryConditionsFS()
        self.operator.spmv["Curl"].csr.mult(self.velFS.data, self._vortFS.data)
        self._halo(self._vortFS, self.dim_w)
        m.Rw.csr.mult(self._vortFS.data, b)
        m.Krhs.csr.mult(self.vel.data, b, mode=SPMV_ADD)
        self.solver(self._kleRhs, self.vel)

    def buildKLEMats(self):
        """base_problem.py:329-454 for the whole mesh in eight assembly launches (matrices/mat_ns.py)."""
        self.mat.assembleFromMesh(self.dom, self.elemType)
        self.mat.assembleAll()
        if not self.comm.rank:
            self.logger.info("KLE Matrices builded")


class FreeSlip(BaseProblem):
    """All boundary nodes Dirichlet: one velocity solve (base_problem.py:456-552)."""
    integratePerCell = True      # the reference integrates every cell here (:505-506)

    def generateExactVecs(self, time):
        return 0, 0

    def buildMatrices(self):
        pass

    def setUpEmptyMats(self):
        self.mat = Mat(self.dim, self.comm)
        self.operator = Operators(self.dim, self.comm)
        rStart, rEnd, d_nnz_ind, o_nnz_ind, ind_d, ind_o = self.dom.getMatIndices()
        globalIndicesDIR = self.dom.getGlobalIndicesDirichlet()
        self.mat.createEmptyKLEMats(rStart, rEnd, d_nnz_ind, o_nnz_ind, ind_d, ind_o, globalIndicesDIR)
        if not self.comm.rank:
            self.logger.info("Empty KLE Matrices created")
        self.operator.createAll(rStart, rEnd, d_nnz_ind, o_nnz_ind)
        if not self.comm.rank:
            self.logger.info("Empty Operators created")

    def solveKLE(self, time, vort):
        self.applyBoundaryConditions(time)
        self._solveVelocity(vort)

    def _solveVelocity(self, vort):
        """solver(Rw w + Krhs v_bc, vel)  (base_problem.py:481)"""
        b = self._kleRhs.data
        self._halo(vort, self.dim_w)
        self.mat.Rw.csr.mult(vort.data, b)
        self.mat.Krhs.csr.mult(self.vel.data, b, mode=SPMV_ADD)
        self.solver(self._kleRhs, self.vel)

    def getKLEError(self, viscousTimes=None, startTime=0.0, endTime=1.0, steps=10):
        if viscousTimes is None:
            viscousTimes = np.arange(startTime, endTime, (endTime - startTime) / steps)
        times = [(tau ** 2) / (4 * self.nu) for tau in viscousTimes]
        errors = list()
        for time in times:
            exactVel, exactVort = self.generateExactVecs(time)
            self.applyBoundaryConditions(time)
            self._solveVelocity(exactVort)
            errors.append((exactVel - self.vel).norm(norm_type=2))
        return errors

    def buildKLEMats(self):
        """base_problem.py:499-552 for the whole mesh: K, Krhs, Rw in three assembly launches."""
        self.mat.assembleFromMesh(self.dom, self.elemType, perCell=self.integratePerCell)
        self.mat.setIndices2One(self.mat.globalIndicesDIR)
        self.mat.assembleAll()
        if not self.comm.rank:
            self.logger.info("KLE Matrices builded")
