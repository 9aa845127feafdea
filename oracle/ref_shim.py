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


def import_reference_immersed_body():
    """The reference's src/domain/immersed_body.py, imported unmodified.  It imports petsc4py at module level (the
    bodies are stored as 1-D DMPlex objects) but its discrete delta functions (threeGrid / linear / fourGrid, :398-421),
    ImmersedBody.getDiracs (:259-264) and the generateBody methods of Line / OpenBox / Circle (:293-389) are plain
    Python + NumPy: an empty petsc4py stand-in is enough to reach them."""
    if not reference_available():
        raise RuntimeError("reference tree not present (expected only in the build container)")
    if "petsc4py" not in sys.modules:
        stub = types.ModuleType("petsc4py")
        stub.PETSc = types.SimpleNamespace()
        stub.init = lambda *a, **k: None
        sys.modules["petsc4py"] = stub
    if REFERENCE_SRC not in sys.path:
        sys.path.insert(0, REFERENCE_SRC)
    import domain.immersed_body as ib  # noqa: E402  (reference code)
    return ib

