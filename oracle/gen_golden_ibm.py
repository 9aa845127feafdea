"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/ibm_ref.npz by running the UNMODIFIED reference module
src/domain/immersed_body.py (imported through oracle/ref_shim.py with an empty petsc4py stand-in): its discrete delta
functions, ImmersedBody.getDiracs and the marker coordinates of Circle / Line / OpenBox.  Runs only in the build
container; the vectors are committed so that the GPU box (no /root/reference) can check the oracle against them.

    python oracle/gen_golden_ibm.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from ref_shim import import_reference_immersed_body  # noqa: E402

OUT = os.path.join(HERE, "..", "tests", "golden", "ibm_ref.npz")


def main():
    ib = import_reference_immersed_body()
    out = {}
    r = np.linspace(0.0, 2.5, 501)
    out["r"] = r
    for name in ("fourGrid", "threeGrid", "linear"):
        out[f"phi_{name}"] = np.array([getattr(ib, name)(x) for x in r], dtype=float)
    rng = np.random.default_rng(11)
    dist = rng.uniform(-0.15, 0.15, size=(400, 2))
    out["dirac_dist"], out["dirac_h"] = dist, np.array(0.06)
    body = ib.Circle(vel=[0, 0], center=[0, 0], radius=0.5)
    out["dirac"] = np.array([body.getDiracs(d, 0.06) for d in dist])
    for i, (radius, h) in enumerate(((0.5, 0.06), (0.5, 0.01), (0.37, 0.0431))):
        coords, cone, dl = ib.Circle(vel=[0, 0], center=[0, 0], radius=radius).generateBody(h)
        out[f"circle{i}_args"], out[f"circle{i}_xy"], out[f"circle{i}_dl"] = np.array([radius, h]), np.array(coords), np.array(dl)
        assert cone[-1] == [len(coords) - 1, 0]
    for i, dl in enumerate((0.1, 0.0431)):
        coords, cone, d = ib.Line(vel=[0, 0], center=[0, 0]).generateBody(dl)
        out[f"line{i}_args"], out[f"line{i}_xy"] = np.array([dl]), np.array(coords)
        coords, cone, d = ib.OpenBox(vel=[0, 0], center=[0, 0]).generateBody(dl)
        out[f"box{i}_args"], out[f"box{i}_xy"] = np.array([dl]), np.array(coords)
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
