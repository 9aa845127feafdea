"""`KspSolver`: the linear solver of the velocity-from-vorticity problem (reference: src/solver/ksp_solver.py:9-19;
use sites src/cases/base_problem.py:276,318,323-327,481).

Same surface -- `KspSolver()`, `createSolver(mat, comm)`, `solver(b, x)` -- and the same precedence of settings: the
reference's defaults (`preonly` + `lu`) first, then whatever the options database says (`-ksp_type`, `-pc_type`,
`-ksp_rtol`, `-ksp_atol`, `-ksp_divtol`, `-ksp_max_it`, `-ksp_gmres_restart`, `-ksp_initial_guess_nonzero`), then `setOperators` / `setUp`.
The iterations run on the device (csrc/ksp.cu) with PETSc's semantics: zero initial guess (or the contents
of `x` under `-ksp_initial_guess_nonzero`: r0 = b - A x0, rtol measured against |M^-1 b|), left preconditioning,
preconditioned residual norm, KSPConvergedDefault.

`preonly` + `lu` (a sparse direct factorisation) has no device counterpart in this build; it is served by
Jacobi-preconditioned CG driven to `DIRECT_RTOL` followed by `DIRECT_REFINE` passes of iterative refinement
(r = b - A x, A d = r, x += d), i.e. "solve to round-off" like a direct solver would, and says so in the log.  Parity statements about KSP are made for `-ksp_type cg -pc_type jacobi` at matched tolerances.

`-pc_type mg` preconditions CG with a geometric p/h-multigrid V-cycle (solver/multigrid.py, csrc/mg.cu):
rediscretised K on the same mesh at lower orders and on coarsened bilinear meshes, Chebyshev smoothing
(`-pc_mg_degree`, `-pc_mg_ratio`), dense coarse inverse (`-pc_mg_coarse_cells`).
"""
import logging

from .. import backend as be
from ..common.options import Options


class ConvergedReason:
    CONVERGED_RTOL, CONVERGED_ATOL, CONVERGED_ITS = 2, 3, 4
    DIVERGED_MAX_IT, DIVERGED_DTOL, DIVERGED_BREAKDOWN, DIVERGED_INDEFINITE_MAT = -3, -4, -5, -8


class KspSolver(object):
    DIRECT_RTOL = 1e-10          # stand-in for preonly + lu: CG to this tolerance ...
    DIRECT_REFINE = 2            # ... then this many passes of iterative refinement
    DIRECT_MAX_IT = 200000

    def __init__(self):
        self.logger = None
        self._ksp = None
        self._direct = False
        self._type, self._pc = "preonly", "lu"
        self.rtol, self.atol, self.divtol, self.max_it = 1e-5, 1e-50, 1e4, 10000
        self.restart = 30
        self.check_every = 0
        self.guess_nonzero = False
        self.error_if_not_converged = False
        self.its, self.reason, self.rnorm = 0, 0, 0.0
        self.history = []

    # -- petsc4py KSP surface used by the reference -------------------------------------------------------------
    def create(self, comm=None):
        self.comm = comm
        return self

    def setType(self, t):
        self._type = t

    def getType(self):
        return self._type

    def setPCType(self, t):
        self._pc = t

    def setTolerances(self, rtol=None, atol=None, divtol=None, max_it=None):
        if rtol is not None:
            self.rtol = rtol
        if atol is not None:
            self.atol = atol
        if divtol is not None:
            self.divtol = divtol
        if max_it is not None:
            self.max_it = max_it

    def setInitialGuessNonzero(self, flag):
        self.guess_nonzero = bool(flag)

    def getInitialGuessNonzero(self):
        return self.guess_nonzero

    def getTolerances(self):
        return self.rtol, self.atol, self.divtol, self.max_it

    def setFromOptions(self):
        o = Options()
        self._type = o.getString("ksp_type", self._type)
        self._pc = o.getString("pc_type", self._pc)
        self.rtol = o.getReal("ksp_rtol", self.rtol)
        self.atol = o.getReal("ksp_atol", self.atol)
        self.divtol = o.getReal("ksp_divtol", self.divtol)
        self.max_it = o.getInt("ksp_max_it", self.max_it)
        self.restart = o.getInt("ksp_gmres_restart", self.restart)
        self.check_every = o.getInt("ksp_check_every", self.check_every)
        self.guess_nonzero = o.getBool("ksp_initial_guess_nonzero", self.guess_nonzero)
        self.error_if_not_converged = o.getBool("ksp_error_if_not_converged", self.error_if_not_converged)
        self.mg_degree = o.getInt("pc_mg_degree", 3)                 # Chebyshev smoother degree
        self.mg_ratio = o.getReal("pc_mg_ratio", 30.0)               # ... on [lmax / ratio, lmax]; 30: see profiles/README.md
        self.mg_coarse_cells = o.getInt("pc_mg_coarse_cells", 1024)  # coarsest mesh: dense inverse

    def setOperators(self, mat):
        self.mat = mat

    def getOperators(self):
        return self.mat, self.mat

    def setUp(self):
        ksp_type, pc = self._type, self._pc
        rtol, max_it = self.rtol, self.max_it
        if ksp_type == "preonly":
            if pc not in ("lu", "cholesky"):
                raise Exception(f"KSP preonly with PC {pc} is not available on the device path")
            self.logger.info("preonly+lu requested: served by device CG (multigrid or Jacobi) to rtol %g + iterative "
                             "refinement (no sparse direct factorisation on the device path)", self.DIRECT_RTOL)
            # the multigrid V-cycle makes "solve to round-off" affordable (tens of iterations instead of tens of
            # thousands); it needs the case's mesh, which the case layer supplies right after createSolver
            # (setUpMultigrid) -- a solver used on its own falls back to Jacobi at the first solve
            pc = "mg" if Options().getString("pc_direct_standin", "mg") == "mg" else "jacobi"
            ksp_type, rtol, max_it = "cg", self.DIRECT_RTOL, self.DIRECT_MAX_IT
            self._direct = True
        if ksp_type not in be.Ksp.TYPES:
            raise Exception(f"KSP type {ksp_type} is not available on the device path (cg, gmres)")
        if pc not in be.Ksp.PCS:
            raise Exception(f"PC type {pc} is not available on the device path (none, jacobi, mg)")
        csr = self.mat.csr if hasattr(self.mat, "csr") else self.mat
        self._ksp = be.Ksp(csr, ksp_type, pc, rtol=rtol, atol=self.atol, dtol=self.divtol, max_it=max_it,
                           restart=self.restart, check_every=self.check_every,
                           nonzero_guess=self.guess_nonzero and not self._direct)
        dist = getattr(self.comm, "dist", None)
        if dist is not None:       # multi-GPU: halo exchange of the search direction + all-reduced scalars
            dist.attach(self._ksp, ncomp=csr.shape[0] // dist.part.nnodes)

    def wantsMultigrid(self):
        return self._ksp is not None and self._ksp.opts.pc == be.Ksp.PCS["mg"] and getattr(self._ksp, "_mg", None) is None

    def setUpMultigrid(self, fem, kind):
        """`-pc_type mg`: build the V-cycle levels of this solve from the case (solver/multigrid.py) and attach them.
        The case layer calls this right after createSolver (the hierarchy needs the mesh and the boundary classes, which
        the reference's createSolver(mat, comm) signature does not carry)."""
        from .multigrid import Hierarchy
        h = Hierarchy(fem, kind, self._ksp.A, degree=self.mg_degree, ratio=self.mg_ratio,
                      coarse_cells=self.mg_coarse_cells)
        h.attach(self._ksp)
        self.logger.info("multigrid: %s", h.describe())
        return h

    def createSolver(self, mat, comm=None):
        self.logger = logging.getLogger("KSP Solver")
        self.logger.debug("setupKSP")
        self.create(comm)
        self.setType('preonly')
        self.setPCType('lu')
        self.setFromOptions()
        self.setOperators(mat)
        self.setUp()

    def solve(self, b, x):
        if self.wantsMultigrid():      # -pc_type mg (or the direct stand-in) without a hierarchy: Jacobi instead
            self.logger.warning("multigrid hierarchy not set up for this solver (no case attached): using Jacobi")
            self._pc = "jacobi"
            Options.setValue("pc_direct_standin", "jacobi")
            try:
                self.setUp()
            finally:
                Options.setValue("pc_direct_standin", "mg")
        r = self._ksp.solve(b.data, x.data)
        if self._direct and r.reason > 0:
            csr = self._ksp.A
            if getattr(self, "_res", None) is None or self._res.numel() != b.data.numel():
                self._res, self._cor = be.empty(b.data.numel()), be.empty(b.data.numel())
            for _ in range(self.DIRECT_REFINE):
                csr.mult(x.data, self._res)                       # A x
                be.vec_maxpy(self._res, b.data, [-1.0], [self._res])   # r = b - A x
                rr = self._ksp.solve(self._res, self._cor)
                if rr.reason < 0:
                    r = rr          # a failed refinement pass fails the "direct" solve
                    break
                be.vec_axpy(x.data, 1.0, self._cor)
        self.its, self.reason, self.rnorm = r.its, r.reason, r.rnorm
        self.history.append((r.its, r.reason))
        if r.reason < 0:
            msg = f"KSP did not converge: reason {r.reason} after {r.its} iterations (rnorm {r.rnorm:.3e})"
            # the reference's default is a direct solve (exact or an error): its iterative stand-in must not hand an
            # unconverged velocity to the time loop; an iterative KSP only raises under -ksp_error_if_not_converged
            if self._direct or self.error_if_not_converged:
                raise Exception(msg)
            self.logger.warning(msg)
        return r

    __call__ = solve

    def getIterationNumber(self):
        return self.its

    def getConvergedReason(self):
        return self.reason

    def getResidualNorm(self):
        return self.rnorm

    def destroy(self):
        self._ksp = None
