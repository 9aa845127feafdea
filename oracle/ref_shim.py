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

