"""TEST INFRASTRUCTURE ONLY -- import shim for the *unmodified* reference element layer.

Only usable in the build container, where /root/reference exists (it does not exist on the GPU
box).  Used by oracle/gen_golden.py to produce tests/golden/*.npz and by the CPU tests that
re-validate the numpy restatement (oracle/kle_oracle.py) against the real reference code.

The reference element layer (src/elements/{element,utilities,spectral}.py) only touches
``mpi4py.MPI.COMM_WORLD.rank`` (element.py:3,7,13), so a 5-line stub is enough.
"""
import os
import sys
import types

REFERENCE_SRC = "/root/reference/src"


def reference_available():
    return os.path.isdir(os.path.join(REFERENCE_SRC, "elements"))


def import_reference_elements():
    """Return (Spectral, utilities module) of the reference, imported unmodified."""
    if not reference_available():
        raise RuntimeError("reference tree not present (expected only in the build container)")
    if "mpi4py" not in sys.modules:
        stub = types.ModuleType("mpi4py")

        class _Comm:
            rank = 0
            size = 1

        class _MPI:
            COMM_WORLD = _Comm()

        stub.MPI = _MPI
        sys.modules["mpi4py"] = stub
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    from elements.spectral import Spectral  # noqa: E402  (reference code)
    from elements import utilities  # noqa: E402
    return Spectral, utilities


def _install_petsc_stub():
    """petsc4py as an importable PACKAGE whose PETSc module hands out an empty base class for every name: enough for the
    reference's modules to be imported (they subclass PETSc.DMPlex / KSP / TS and import names from petsc4py.PETSc), not to
    run anything that needs PETSc."""
    if "petsc4py.PETSc" in sys.modules and getattr(sys.modules["petsc4py.PETSc"], "_pyn_stub", False):
        return

    class _Base(object):
        def __init__(self, *a, **k):
            pass

    class _Mod(types.ModuleType):
        _pyn_stub = True

        def __getattr__(self, k):
            if k.startswith("__"):
                raise AttributeError(k)
            return _Base

    pkg = types.ModuleType("petsc4py")
    pkg.__path__ = []
    pkg.init = lambda *a, **k: None
    petsc = _Mod("petsc4py.PETSc")
    pkg.PETSc = petsc
    sys.modules["petsc4py"] = pkg
    sys.modules["petsc4py.PETSc"] = petsc


def import_reference_immersed_body():
    """The reference's src/domain/immersed_body.py, imported unmodified.  It imports petsc4py at module level (the
    bodies are stored as 1-D DMPlex objects) but its discrete delta functions (threeGrid / linear / fourGrid, :398-421),
    ImmersedBody.getDiracs (:259-264) and the generateBody methods of Line / OpenBox / Circle (:293-389) are plain
    Python + NumPy: an empty petsc4py stand-in is enough to reach them."""
    if not reference_available():
        raise RuntimeError("reference tree not present (expected only in the build container)")
    import_reference_elements()          # (the mpi4py stand-in)
    _install_petsc_stub()
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    import domain.immersed_body as ib  # noqa: E402  (reference code)
    return ib


# ---- the reference's immersed-boundary CASE code on scipy stand-ins for PETSc Mat / Vec ------------------------------
class _FakeVec(object):
    """The slice of PETSc.Vec that computeVelocityCorrection touches (immersed_boundary.py:205-211)."""

    def __init__(self, arr):
        import numpy as np
        self.a = np.array(arr, dtype=float)

    def axpy(self, alpha, x):
        self.a = self.a + alpha * x.a

    def __neg__(self):
        return _FakeVec(-self.a)

    def __iadd__(self, other):
        self.a = self.a + other.a
        return self


class _FakeMat(object):
    """The slice of PETSc.Mat that buildIBMMatrix touches (:222-251): INSERT setValues, assemble, copy, in-place
    transpose, scale, matMult (row times column over the sorted rows), mult."""

    def __init__(self, shape, csr=None):
        self.shape, self.entries, self.csr = shape, {}, csr

    def setValues(self, row, cols, vals, addv=False):
        for c, v in zip(cols, vals):
            self.entries[(int(row), int(c))] = float(v)

    def assemble(self):
        import numpy as np
        import scipy.sparse as sp
        keys = sorted(self.entries)
        r = np.array([k[0] for k in keys], dtype=np.int64)
        c = np.array([k[1] for k in keys], dtype=np.int64)
        v = np.array([self.entries[k] for k in keys])
        self.csr = sp.csr_matrix((v, (r, c)), shape=self.shape)
        self.csr.sort_indices()

    def copy(self):
        return _FakeMat(self.shape, self.csr.copy())

    def transpose(self):
        import scipy.sparse as sp
        self.csr = sp.csr_matrix(self.csr.T)
        self.csr.sort_indices()
        self.shape = self.csr.shape
        return self

    def scale(self, alpha):
        self.csr.data = self.csr.data * alpha

    def matMult(self, other):
        # sequential row-times-column accumulation in ascending column order of the left row (MatMatMult on sorted AIJ)
        import numpy as np
        import scipy.sparse as sp
        A, B = self.csr, sp.csc_matrix(other.csr)
        B.sort_indices()
        rows, cols, vals = [], [], []
        for i in range(A.shape[0]):
            ci, vi = A.indices[A.indptr[i]:A.indptr[i + 1]], A.data[A.indptr[i]:A.indptr[i + 1]]
            if len(ci) == 0:
                continue
            pos = {int(c): k for k, c in enumerate(ci)}
            for j in range(B.shape[1]):
                rj, vj = B.indices[B.indptr[j]:B.indptr[j + 1]], B.data[B.indptr[j]:B.indptr[j + 1]]
                s, hit = 0.0, False
                for r_, v_ in zip(rj, vj):
                    k = pos.get(int(r_))
                    if k is not None:
                        s, hit = s + vi[k] * v_, True
                if hit:
                    rows.append(i); cols.append(j); vals.append(s)
        out = sp.csr_matrix((np.array(vals), (np.array(rows), np.array(cols))), shape=(A.shape[0], B.shape[1]))
        out.sort_indices()
        return _FakeMat(out.shape, out)

    def mult(self, x, y):
        y.a = self.csr @ x.a


def run_reference_ibm(nodexy, candidate_nodes, lagxy, dl, h, dynamic, vel, body_vel):
    """collectNodes + buildIBMMatrix + computeVelocityCorrection of the UNMODIFIED reference case classes
    (src/cases/immersed_boundary.py) on a bare instance: the mesh, the body and the PETSc objects are the stand-ins
    above, every line that decides which Euler nodes belong to a marker, what goes into H, how S and A are formed and
    how the velocity is corrected is the reference's own.  Returns (H, S, A as scipy CSR, corrected velocity, flux)."""
    import logging
    import numpy as np
    import scipy.sparse.linalg as spla
    import scipy.sparse as sp
    ib = import_reference_immersed_body()
    import cases.immersed_boundary as cib  # noqa: E402  (reference code)
    nodexy, lagxy = np.asarray(nodexy, dtype=float), np.asarray(lagxy, dtype=float)

    class Dom(object):
        def getGlobalNodesFromEntities(self, cells, shared):
            return set(int(n) for n in candidate_nodes)

        def getNodesCoordinates(self, nodes):
            return np.array([nodexy[n] for n in nodes])

    class Body(object):
        def getTotalNodes(self):
            return len(lagxy)

        def getNodeCoordinates(self, l):
            return lagxy[l]

        def getElementLong(self):
            return dl

        def getDiracs(self, dist, hh):
            return ib.ImmersedBody.getDiracs(types.SimpleNamespace(dirac=ib.fourGrid), dist, hh)

        def setEulerNodes(self, l, n):
            pass

        def getVelocity(self):
            return _FakeVec(body_vel)

    class Ksp(object):
        def createSolver(self, A, comm):
            self.A = A

        def solve(self, b, x):
            x.a = spla.spsolve(sp.csc_matrix(self.A.csr), b.a)

    base = cib.ImmersedBoundaryDynamic if dynamic else cib.ImmersedBoundaryStatic

    class Bare(base):
        def __init__(self):
            pass

    fem = Bare()
    fem.dom, fem.body, fem.dim, fem.h, fem.comm = Dom(), Body(), 2, h, None
    fem.logger = logging.getLogger("reference-ibm")
    fem.collectedNodes, fem.maxNodesPerLag = fem.collectNodes(None)
    fem.H = _FakeMat((len(lagxy) * 2, len(nodexy) * 2))
    saved = cib.KspSolver
    cib.KspSolver = Ksp
    try:
        fem.buildIBMMatrix()
    finally:
        cib.KspSolver = saved
    fem.vel, fem.ibm_rhs = _FakeVec(vel), _FakeVec(np.zeros(2 * len(lagxy)))
    fem.virtualFlux, fem.vel_correction = _FakeVec(np.zeros(2 * len(lagxy))), _FakeVec(np.zeros(len(vel)))
    cib.ImmersedBoundaryStatic.computeVelocityCorrection(fem)
    return fem.H.csr, fem.S.csr, fem.ksp.A.csr, fem.vel.a, fem.virtualFlux.a



# ---- the reference's buildKLEMats loops (A1 / A2 of SURVEY 8) on stand-ins -------------------------------------------
class _FakeAijMat(object):
    """MatSetValues / MatAssembly of a serial AIJ matrix: every (row, col) ever set is stored -- explicit zeros included
    --, ADD_VALUES accumulate in call order, INSERT_VALUES replace."""

    def __init__(self, shape, name=""):
        self.shape, self.name, self.entries = shape, name, {}

    def setValues(self, rows, cols, vals, addv=False):
        import numpy as np
        rows, cols = np.atleast_1d(np.asarray(rows, dtype=np.int64)), np.atleast_1d(np.asarray(cols, dtype=np.int64))
        vals = np.asarray(vals, dtype=float)
        if vals.ndim == 0:
            vals = np.full((len(rows), len(cols)), float(vals))
        vals = vals.reshape(len(rows), len(cols))
        for i, r in enumerate(rows):
            for j, c in enumerate(cols):
                k = (int(r), int(c))
                if addv and k in self.entries:
                    self.entries[k] = self.entries[k] + vals[i, j]
                else:
                    self.entries[k] = float(vals[i, j])

    def assemble(self):
        pass

    def getName(self):
        return self.name

    def to_csr(self):
        import numpy as np
        import scipy.sparse as sp
        keys = sorted(self.entries)
        indptr = np.zeros(self.shape[0] + 1, dtype=np.int64)
        for r, _ in keys:
            indptr[r + 1] += 1
        return sp.csr_matrix((np.array([self.entries[k] for k in keys]), np.array([k[1] for k in keys], dtype=np.int32),
                              np.cumsum(indptr)), shape=self.shape)


def run_reference_build_kle_mats(mesh, noslip):
    """NoSlipFreeSlip.buildKLEMats (src/cases/base_problem.py:329-454) or FreeSlip.buildKLEMats (:499-552), UNMODIFIED,
    on a bare case instance: the element layer is the reference's own Spectral, the matrix container its own MatNS / Mat
    (assembleAll, setIndices2One), the PETSc matrices are _FakeAijMat, and the mesh queries (cell nodes, corner
    coordinates, border nodes with their wall normal: src/domain/dmplex.py:245-261) are answered from the oracle's
    BoxMesh `mesh`.  Every decision about which dof is free / tangential / set in a cell, the thirteen setValues calls
    and the post-assembly loops are the reference's.  Returns {name: scipy CSR}."""
    import logging
    import numpy as np
    Spectral, _ = import_reference_elements()
    _install_petsc_stub()
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    import cases.base_problem as cbp  # noqa: E402  (reference code)
    from matrices.mat_generator import Mat  # noqa: E402
    from matrices.mat_ns import MatNS  # noqa: E402
    cell_nodes = mesh.cell_nodes()
    corner = mesh.cell_corner_coords()
    names = ["down", "right", "up", "left"]                      # dmplex.py:38
    face_nodes = {nm: set(int(n) for n in mesh.border_nodes(nm)) for nm in names}
    normal = {"down": 1, "up": 1, "left": 0, "right": 0}         # |normal| of the face (computeCellGeometryFVM)
    boundary = set(int(n) for n in mesh.boundary_nodes())

    class Dom(object):
        cellStart, cellEnd = 0, len(cell_nodes)

        def getCellCornersCoords(self, cell):
            return corner[cell].reshape(-1).copy()

        def getGlobalNodesFromCell(self, cell, shared):
            return [int(n) for n in cell_nodes[cell]]

        def getVelocityIndex(self, nodes):
            return [n * 2 + d for n in nodes for d in range(2)]

        def getVorticityIndex(self, nodes):
            return [n for n in nodes]

        def getNodesFromLabel(self, label, shared=False):
            assert label == "External Boundary"
            return set(boundary)

        def getBorderNodesWithNormal(self, cell, intersect):
            nodes, normals = [], []
            mine = [int(n) for n in cell_nodes[cell]]
            for nm in names:
                on_face = [n for n in mine if n in face_nodes[nm]]
                # (a cell of a box mesh that touches a border does so with a whole side)
                if len(on_face) >= mesh.ngl and not (set(on_face) - intersect):
                    normals.append(normal[nm])
                    nodes.append(on_face)
            return nodes, normals

    N = mesh.nnodes
    comm = types.SimpleNamespace(rank=0, size=1)
    mat = (MatNS if noslip else Mat).__new__(MatNS if noslip else Mat)
    mat.dim, mat.dim_w, mat.dim_s, mat.comm = 2, 1, 3, comm
    mat.logger = logging.getLogger("reference-mat")
    sq, rc = (2 * N, 2 * N), (2 * N, N)
    for nm, shape in (("K", sq), ("Krhs", sq), ("Rw", rc), ("Rd", rc), ("Kfs", sq), ("Krhsfs", sq), ("Rwfs", rc),
                      ("Rdfs", rc)):
        setattr(mat, nm, _FakeAijMat(shape, nm))
    mat.mats = [mat.K, mat.Rw, mat.Rd, mat.Krhs]
    if noslip:
        mat.globalIndicesNS, mat.globalIndicesDIR = set(boundary), set()
    else:
        mat.globalIndicesDIR = set(2 * n + d for n in boundary for d in range(2))
    base = cbp.NoSlipFreeSlip if noslip else cbp.FreeSlip

    class Bare(base):
        def __init__(self):
            pass

    fem = Bare()
    fem.dim, fem.dom, fem.mat, fem.comm = 2, Dom(), mat, comm
    fem.elemType = Spectral(mesh.ngl, 2)
    fem.logger = logging.getLogger("reference-case")
    fem.buildKLEMats()
    keep = ("K", "Krhs", "Rw", "Rd", "Kfs", "Krhsfs", "Rwfs", "Rdfs") if noslip else ("K", "Krhs", "Rw")
    return {nm: getattr(mat, nm).to_csr() for nm in keep}


# ---- buildOperators / Operators.assembleAll (A3) and the evalRHS chain (R2 / R3) on stand-ins --------------------------
class _FakeNpVec(object):
    """PETSc.Vec calls of mat_generator.py:157-190 and base_problem.py:212-252 on a NumPy array."""

    def __init__(self, n=0, arr=None):
        import numpy as np
        self.a = np.zeros(n) if arr is None else np.array(arr, dtype=float)

    def setValues(self, idx, vals, addv=False):
        import numpy as np
        idx = np.asarray(idx, dtype=np.int64).reshape(-1)
        vals = np.broadcast_to(np.asarray(vals, dtype=float).reshape(-1), idx.shape)
        if addv:
            for i, v in zip(idx, vals):          # sequential: the order of the ADD_VALUES stash
                self.a[i] += v
        else:
            self.a[idx] = vals

    def assemble(self):
        pass

    def destroy(self):
        pass

    def reciprocal(self):
        self.a = 1.0 / self.a

    def getArray(self):
        return self.a

    def duplicate(self):
        return _FakeNpVec(len(self.a))

    def scale(self, alpha):
        self.a = self.a * alpha

    def axpy(self, alpha, x):
        self.a = self.a + alpha * x.a

    def __imul__(self, alpha):
        self.a = self.a * alpha
        return self


def _aij_extras():
    """diagonalScale / mult / getOwnershipRange for _FakeAijMat (kept apart from the assembly stand-in above)."""
    import numpy as np

    def diagonalScale(self, L=None, R=None):
        assert R is None
        for (r, c) in self.entries:
            self.entries[(r, c)] = L.a[r] * self.entries[(r, c)]

    def mult(self, x, y):
        y.a = self.to_csr() @ np.asarray(x.a)

    def getOwnershipRange(self):
        return 0, self.shape[0]

    _FakeAijMat.diagonalScale, _FakeAijMat.mult, _FakeAijMat.getOwnershipRange = diagonalScale, mult, getOwnershipRange


def run_reference_operators_and_rhs(mesh, vel, rho, mu):
    """BaseProblem.buildOperators (src/cases/base_problem.py:132-140) with Operators.setValues / assembleAll
    (src/matrices/mat_generator.py:157-190), then evalRHS (:212-232, computeVtensV :234-252) on a given velocity --
    UNMODIFIED, solveKLE replaced by a no-op.  Returns ({SrT, DivSrT, Curl: scipy CSR}, f)."""
    import logging
    import numpy as np
    Spectral, _ = import_reference_elements()
    _install_petsc_stub()
    _aij_extras()
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    import cases.base_problem as cbp  # noqa: E402  (reference code)
    from matrices.mat_generator import Operators  # noqa: E402
    cell_nodes, corner, N = mesh.cell_nodes(), mesh.cell_corner_coords(), mesh.nnodes

    class Dom(object):
        cellStart, cellEnd = 0, len(cell_nodes)

        def getCellCornersCoords(self, cell):
            return corner[cell].reshape(-1).copy()

        def getGlobalNodesFromCell(self, cell, shared):
            return [int(n) for n in cell_nodes[cell]]

    op = Operators.__new__(Operators)
    op.dim, op.dim_w, op.dim_s = 2, 1, 3
    op.Curl, op.SrT, op.DivSrT = (_FakeAijMat((N, 2 * N), "Curl"), _FakeAijMat((3 * N, 2 * N), "SrT"),
                                  _FakeAijMat((2 * N, 3 * N), "DivSrT"))
    op.weigCurl, op.weigSrT, op.weigDivSrT = _FakeNpVec(N), _FakeNpVec(3 * N), _FakeNpVec(2 * N)

    class Bare(cbp.BaseProblem):
        def __init__(self):
            pass

        def solveKLE(self, t, vort):
            pass

    fem = Bare()
    fem.dim, fem.dim_w, fem.dim_s = 2, 1, 3
    fem.dom, fem.operator, fem.comm = Dom(), op, types.SimpleNamespace(rank=0, size=1)
    fem.elemType, fem.logger = Spectral(mesh.ngl, 2), logging.getLogger("reference-case")
    fem.buildOperators()
    fem.rho, fem.mu = rho, mu
    fem.mat = types.SimpleNamespace(K=_FakeAijMat((2 * N, 2 * N)))
    fem.vel = _FakeNpVec(arr=vel)
    fem._VtensV, fem._Aux1 = _FakeNpVec(3 * N), _FakeNpVec(3 * N)
    f = _FakeNpVec(N)
    fem.evalRHS(None, 0.0, None, f)
    return {nm: getattr(op, nm).to_csr() for nm in ("SrT", "DivSrT", "Curl")}, np.array(f.a)


# ---- Cavity.solveKLE (R1) on stand-ins -----------------------------------------------------------------------------
def run_reference_cavity_solve_kle(mesh, vort, no_slip={"up": [1, 0]}):
    """NoSlipFreeSlip.solveKLE (src/cases/base_problem.py:321-327) with Cavity.applyBoundaryConditions /
    applyBoundaryConditionsFS (src/cases/cavity.py:53-81) and the reference's own NoSlipWalls (src/common/nswalls.py),
    UNMODIFIED, on the matrices its own buildKLEMats / buildOperators produced (the stand-ins above); the two KSP solves
    are SuperLU, the stand-in for the reference's default preonly + lu.  Returns (velocity, free-slip velocity)."""
    import logging
    import numpy as np
    import scipy.sparse as sp
    import scipy.sparse.linalg as spla
    Spectral, _ = import_reference_elements()
    _install_petsc_stub()
    _aij_extras()
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    import cases.base_problem as cbp  # noqa: E402  (reference code)
    import cases.cavity as ccav  # noqa: E402
    from matrices.mat_generator import Operators  # noqa: E402
    from matrices.mat_ns import MatNS  # noqa: E402
    N = mesh.nnodes
    mats = run_reference_build_kle_mats(mesh, True)
    ops, _ = run_reference_operators_and_rhs(mesh, np.zeros(2 * N), 1.0, 1.0)

    class M(object):                      # Mat * Vec and Mat + Mat of petsc4py on an assembled matrix
        def __init__(self, csr):
            self.csr = csr

        def __mul__(self, v):
            return V(self.csr @ v.a)

        def __add__(self, other):         # MatAXPY(1.0, DIFFERENT_NONZERO_PATTERN): union pattern
            A, B = self.csr.tocoo(), other.csr.tocoo()
            import kle_oracle as o
            return M(o.coo_to_csr(np.concatenate([A.row, B.row]), np.concatenate([A.col, B.col]),
                                  np.concatenate([A.data, B.data]), A.shape))

    class V(_FakeNpVec):
        def __init__(self, arr):
            super().__init__(arr=arr)

        def __add__(self, other):
            return V(self.a + other.a)

        def set(self, value):
            self.a[:] = value

    def lu(A):
        fac = spla.splu(sp.csc_matrix(A.csr))

        def solve(b, x):
            x.a = fac.solve(b.a)
        return solve

    border = {nm: [int(n) for n in mesh.border_nodes(nm)] for nm in ("down", "right", "up", "left")}

    class Dom(object):
        def getBorderNodes(self, name):
            return border[name]

    class Bare(ccav.Cavity):
        def __init__(self):
            pass

    fem = Bare()
    fem.dim, fem.dom, fem.comm = 2, Dom(), types.SimpleNamespace(rank=0, size=1)
    fem.logger = logging.getLogger("reference-case")
    fem.lower, fem.upper = list(mesh.lower), list(mesh.upper)
    fem.readBoundaryCondition({"no-slip": dict(no_slip)})
    fem.mat = types.SimpleNamespace(**{k: M(v) for k, v in mats.items()})
    fem.operator = types.SimpleNamespace(**{k: M(v) for k, v in ops.items()})
    fem.solver, fem.solverFS = lu(fem.mat.K), lu(fem.mat.K + fem.mat.Kfs)
    fem.vel, fem.velFS = V(np.zeros(2 * N)), V(np.zeros(2 * N))
    fem.solveKLE(0.0, V(np.asarray(vort, dtype=float)))
    return np.array(fem.vel.a), np.array(fem.velFS.a)


# ---- CustomFuncCase (Taylor-Green): boundary data and FreeSlip.solveKLE on stand-ins -----------------------------------
def run_reference_taylor_green_solve_kle(mesh, nu, t):
    """FreeSlip.solveKLE (src/cases/base_problem.py:479-481) with CustomFuncCase.applyBoundaryConditions /
    generateExactVecs (src/cases/custom_func.py:62-83) and the Taylor-Green functions (:173-193), DMPlexDom's
    applyFunctionVecToVec (src/domain/dmplex.py:263-272) included, UNMODIFIED, on the matrices the reference's own
    FreeSlip.buildKLEMats produced; the solve is SuperLU.  Returns (velocity from the exact vorticity, exact velocity,
    exact vorticity)."""
    import logging
    import numpy as np
    import scipy.sparse as sp
    import scipy.sparse.linalg as spla
    import_reference_elements()
    _install_petsc_stub()
    _aij_extras()
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    import cases.custom_func as ccf  # noqa: E402  (reference code)
    from domain.dmplex import DMPlexDom  # noqa: E402
    N = mesh.nnodes
    mats = run_reference_build_kle_mats(mesh, False)
    xy = mesh.node_coords()

    class V(_FakeNpVec):
        def __init__(self, arr):
            super().__init__(arr=arr)

        def __add__(self, other):
            return V(self.a + other.a)

        def set(self, value):
            self.a[:] = value

        def setName(self, name):
            pass

    class M(object):
        def __init__(self, csr, ncols):
            self.csr, self.ncols = csr, ncols

        def __mul__(self, v):
            return V(self.csr @ v.a)

        def createVecRight(self):
            return V(np.zeros(self.ncols))

    class Dom(object):
        applyFunctionVecToVec = DMPlexDom.applyFunctionVecToVec      # the reference's own method, unbound

        def getNodesCoordinates(self, nodes):
            return np.array([xy[int(n)] for n in nodes])

        def getAllNodes(self):
            return list(range(N))

    class Bare(ccf.CustomFuncCase):
        def __init__(self):
            pass

    fem = Bare()
    fem.dim, fem.dim_w, fem.dom, fem.nu, fem.caseName = 2, 1, Dom(), nu, "taylor-green"
    fem.comm, fem.logger = types.SimpleNamespace(rank=0, size=1), logging.getLogger("reference-case")
    fem.velFunction, fem.vortFunction = fem.taylorGreenVel_2D, fem.taylorGreenVort_2D
    fem.bcNodes = set(int(n) for n in mesh.boundary_nodes())
    fem.mat = types.SimpleNamespace(K=M(mats["K"], 2 * N), Krhs=M(mats["Krhs"], 2 * N), Rw=M(mats["Rw"], N))
    fac = spla.splu(sp.csc_matrix(mats["K"]))

    def solver(b, x):
        x.a = fac.solve(b.a)
    fem.solver = solver
    fem.vel, fem.vort = V(np.zeros(2 * N)), V(np.zeros(N))
    exactVel, exactVort = fem.generateExactVecs(t)
    fem.solveKLE(t, exactVort)
    return np.array(fem.vel.a), np.array(exactVel.a), np.array(exactVort.a)


# ---- node numbering: the reference's IndicesManager over a DMPlex-style closure (D2 / D3) --------------------------------
def run_reference_cell_nodes(mesh):
    """IndicesManager.mapEntitiesToNodes / getGlobalNodes / getSectionOffset / reorderEntities2D of the UNMODIFIED
    src/domain/indices.py on every cell of `mesh`.  What PETSc would supply is stated here: the transitive closure of a
    cell of DMPlexCreateBoxMesh -- [cell, bottom, right, top, left edge, 4 vertices counter-clockwise from (x0, y0)],
    top and left edges with a negative orientation -- and the section of setFemIndexing (src/domain/dmplex.py:42-61:
    (ngl-2)^2 nodes per cell, 1 per vertex, ngl-2 per edge, offsets in point order).  Returns (ncell, ngl^2) node ids."""
    import numpy as np
    import_reference_elements()
    _install_petsc_stub()
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    from domain.indices import IndicesManager  # noqa: E402  (reference code)
    nx, ny, m = mesh.nx, mesh.ny, mesh.ngl - 2
    nc, nv = nx * ny, (nx + 1) * (ny + 1)
    nxe = nx * (ny + 1)

    class Section(object):
        def getDof(self, p):
            return m * m if p < nc else (1 if p < nc + nv else m)

        def getOffset(self, p):
            if p < nc:
                return p * m * m
            if p < nc + nv:
                return nc * m * m + (p - nc)
            return nc * m * m + nv + (p - nc - nv) * m

    im = IndicesManager(2, mesh.ngl, types.SimpleNamespace(rank=0, size=1))
    im.setGlobalIndicesSection(Section())
    out = []
    for c in range(nc):
        fx, fy = c % nx, c // nx
        vert = lambda vx, vy: nc + vy * (nx + 1) + vx                    # noqa: E731
        xedge = lambda ex, vy: nc + nv + vy * nx + ex                    # noqa: E731
        yedge = lambda vx, ey: nc + nv + nxe + vx * ny + ey              # noqa: E731
        closure = np.array([c, xedge(fx, fy), yedge(fx + 1, fy), xedge(fx, fy + 1), yedge(fx, fy),
                            vert(fx, fy), vert(fx + 1, fy), vert(fx + 1, fy + 1), vert(fx, fy + 1)])
        orient = np.array([0, 0, 0, -1, -1, 0, 0, 0, 0])
        out.append(im.mapEntitiesToNodes(closure, orient, True))
    return np.array(out, dtype=np.int64)
