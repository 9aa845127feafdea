"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/ibm_ref.npz by running the UNMODIFIED reference code (imported
through oracle/ref_shim.py with a petsc4py stand-in):
  * src/domain/immersed_body.py: the discrete delta functions, ImmersedBody.getDiracs, the marker coordinates of Circle /
    Line / OpenBox;
  * src/cases/immersed_boundary.py: collectNodes / computeClose / computeDirac / buildIBMMatrix /
    computeVelocityCorrection of the static and the dynamic case on a bare instance whose mesh, body and PETSc Mat / Vec
    are scipy stand-ins (ref_shim.run_reference_ibm): H, S, A, the corrected velocity and the virtual flux.
Runs only in the build container; the vectors are committed so that the GPU box (no /root/reference) can check the
oracle against them.

    python oracle/gen_golden_ibm.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from ref_shim import import_reference_immersed_body, run_reference_ibm  # noqa: E402

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
    # the case code: a perturbed 26 x 26 grid of Euler nodes, a circle of markers off the grid lines
    h = 0.12
    gx = np.linspace(-1.5, 1.5, 26)
    nodexy = np.stack(np.meshgrid(gx, gx, indexing="xy"), -1).reshape(-1, 2)
    nodexy = nodexy + 0.1 * h * np.random.default_rng(1).standard_normal(nodexy.shape)
    coords, _, dl = ib.Circle(vel=[0, 0], center=[0, 0], radius=0.5).generateBody(h)
    lag = np.array(coords) + np.array([0.03, -0.07])
    cand = np.sort(np.random.default_rng(4).choice(len(nodexy), size=600, replace=False))
    rng = np.random.default_rng(2)
    vel, vb = rng.standard_normal(2 * len(nodexy)), np.tile([0.3, -0.1], len(lag))
    out.update(case_nodexy=nodexy, case_lag=lag, case_cand=cand, case_dl=np.array(dl), case_h=np.array(h), case_vel=vel,
               case_body_vel=vb)
    for tag, dyn in (("static", False), ("dynamic", True)):
        H, S, A, vnew, q = run_reference_ibm(nodexy, cand, lag, dl, h, dyn, vel, vb)
        for nm, M in (("H", H), ("S", S), ("A", A)):
            out[f"{tag}_{nm}_indptr"], out[f"{tag}_{nm}_indices"], out[f"{tag}_{nm}_data"] = M.indptr, M.indices, M.data
        out[f"{tag}_vel"], out[f"{tag}_flux"] = vnew, q
    np.savez_compressed(OUT, **out)
    print("wrote", OUT, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main()
