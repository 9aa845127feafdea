"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/*.npz by running the UNMODIFIED reference element layer
(/root/reference/src/elements, imported through oracle/ref_shim.py).  Runs only in the build container; the
vectors it writes are committed so that the GPU box (which has no /root/reference) can check against them.

    python oracle/gen_golden.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from ref_shim import import_reference_elements  # noqa: E402

OUT = os.path.join(HERE, "..", "tests", "golden")

# corner coordinates in DMPlex closure order (x0,y0),(x1,y0),(x1,y1),(x0,y1); `skew` is a non-affine quad
QUADS = {
    "affine": np.array([[0.0, 0.0], [0.3, 0.0], [0.3, 0.17], [0.0, 0.17]]),
    "skew": np.array([[0.02, -0.01], [0.31, 0.03], [0.27, 0.2], [-0.03, 0.15]]),
    # the unit square in the corner order used by the reference's own test (src/tests/test_element.py:280)
    "reftest": np.array([[1.0, 1.0], [0.0, 1.0], [0.0, 0.0], [1.0, 0.0]]),
}


def main():
    Spectral, util = import_reference_elements()
    os.makedirs(OUT, exist_ok=True)
    out = {}
    for N in range(1, 12):
        x, w = util.gaussPoints(N)
        out[f"gauss_x_{N}"], out[f"gauss_w_{N}"] = x, w
        if N >= 2:
            x, w = util.lobattoPoints(N)
            out[f"lobatto_x_{N}"], out[f"lobatto_w_{N}"] = x, w
    for n in (2, 3, 4, 5):
        s = Spectral(n, 2)
        for name in ("H", "Hrs", "HRed", "HrsRed", "HOp", "HrsOp", "HCoo", "HrsCoo", "HCooRed", "HrsCooRed",
                     "HCooOp", "HrsCooOp"):
            out[f"tab_{name}_{n}"] = np.array(getattr(s, name))
        for name in ("gps", "gpsRed", "gpsOp"):
            out[f"tab_{name}_{n}"] = np.array([[g.r, g.s, g.w] for g in getattr(s, name)])
    np.savez_compressed(os.path.join(OUT, "quadrature_tables.npz"), **out)

    out = {}
    for n in range(2, 10):
        s = Spectral(n, 2)
        for qname, quad in QUADS.items():
            if n >= 8 and qname != "skew":
                continue
            K, Rw, Rd = s.getElemKLEMatrices(quad.reshape(-1).copy())
            SrT, Div, Curl, W = s.getElemKLEOperators(quad.reshape(-1).copy())
            for nm, a in zip(("K", "Rw", "Rd", "SrT", "DivSrT", "Curl", "W"), (K, Rw, Rd, SrT, Div, Curl, W)):
                out[f"{nm}_{n}_{qname}"] = a
    for qname, quad in QUADS.items():
        out[f"coords_{qname}"] = quad
    np.savez_compressed(os.path.join(OUT, "elem_ref.npz"), **out)
    print("wrote", sorted(os.listdir(OUT)))


if __name__ == "__main__":
    main()
