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
