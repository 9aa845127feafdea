"""Immersed-boundary coupling on the device (csrc/ibm.cu, pynama_b200/cases/immersed_boundary.py) against the oracle
restatement of src/cases/immersed_boundary.py:205-277, 318-338 and src/domain/immersed_body.py.

  * close-node lists: identical; delta values, H, S, A: bit pattern of the CSR arrays exact, values to 1e-15
    (every operation is a round-to-nearest intrinsic in the reference's order)
  * the velocity correction through the case class == the oracle's on the same velocity; afterwards the interpolated
    velocity equals the body's at every marker
  * the static case runs time steps; the dynamic case rebuilds its matrices where the body has moved to
"""
import numpy as np
import pytest
import yaml

import kle_oracle as o
from conftest import ROOT, rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def be():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("needs a CUDA device")
    from pynama_b200 import backend
    return backend


def _grid(n, lo, hi):
    gx = np.linspace(lo, hi, n)
    return np.stack(np.meshgrid(gx, gx, indexing="xy"), -1).reshape(-1, 2)


@pytest.mark.parametrize("mode,kind", [(0, "four"), (1, "four"), (0, "three"), (0, "linear")])
def test_collect_and_gram_vs_oracle(be, mode, kind):
    rng = np.random.default_rng(3)
    h = 3.0 / 40
    nodexy = _grid(41, -1.5, 1.5) + 0.2 * h * rng.standard_normal((41 * 41, 2))       # a perturbed grid
    lag, dl = o.ibm_circle(0.5, h, center=(0.05, -0.11))
    cand = np.sort(rng.choice(len(nodexy), size=1500, replace=False)).astype(np.int32)
    ref = o.ibm_collect(lag, cand, nodexy, dl, h, mode=mode, kind=kind)
    p, nodes, dirac = be.ibm_collect(be.to_device(lag.reshape(-1)), be.to_device(cand), be.to_device(nodexy.reshape(-1)),
                                     dl, h, mode=mode, kind=kind)
    p, nodes, dirac = p.cpu().numpy(), nodes.cpu().numpy(), dirac.cpu().numpy()
    np.testing.assert_array_equal(np.diff(p), [len(r[0]) for r in ref])
    np.testing.assert_array_equal(nodes, np.concatenate([r[0] for r in ref]))
    dref = np.concatenate([r[1] for r in ref])
    assert rel_err(dirac, dref) < 1e-15
    # matrices
    Href, Sref, Aref = o.ibm_matrices(ref, len(nodexy), dl, h)
    pd, nd, dd = be.to_device(p), be.to_device(nodes), be.to_device(dirac)
    H = be.dof_blocked_csr(pd, nd, dd * (h ** 2), len(nodexy), 2)
    ip, ix, dv = H.getValuesCSR()
    np.testing.assert_array_equal(ip, Href.indptr)
    np.testing.assert_array_equal(ix, Href.indices)
    assert rel_err(dv, Href.data) < 1e-15
    ap, aj, av = be.ibm_gram(pd, nd, dd, h ** 2, dl * h)
    A = be.dof_blocked_csr(ap, aj, av, len(lag), 2)
    ip, ix, dv = A.getValuesCSR()
    np.testing.assert_array_equal(ip, Aref.indptr)
    np.testing.assert_array_equal(ix, Aref.indices)
    assert rel_err(dv, Aref.data) < 1e-15


def _case(cls_name, nelem=24, ngl=3, moving=False):
    from pynama_b200.cases import immersed_boundary as ib
    from pynama_b200.common.options import Options
    Options.clear()
    for k, v in dict(ksp_type="cg", pc_type="jacobi", ksp_rtol=1e-12, ksp_max_it=20000).items():
        Options.setValue(k, v)
    with open(f"{ROOT}/pynama_b200/cases/ibm-static.yaml") as f:
        cfg = yaml.load(f, Loader=yaml.Loader)
    cfg["domain"]["box-mesh"]["nelem"] = [nelem, nelem]
    cfg["domain"]["ngl"] = ngl
    if moving:
        cfg["bodies"][0]["vel"] = "dynamic"
    fem = getattr(ib, cls_name)(cfg, case="ibm-static")
    fem.setUp()
    fem.setUpSolver()
    return fem


def _oracle_matrices(fem, mode):
    nodexy = fem.dom.deviceNodeCoords().cpu().numpy().reshape(-1, 2)
    lag = fem.body.coordinatesArray()
    cand = fem._ibmCandidates.cpu().numpy()
    col = o.ibm_collect(lag, cand, nodexy, fem.body.getElementLong(), fem.h, mode=mode, kind="four")
    return o.ibm_matrices(col, len(nodexy), fem.body.getElementLong(), fem.h)


def test_static_case_correction_vs_oracle(be):
    import torch
    fem = _case("ImmersedBoundaryStatic")
    assert fem.body.getTotalNodes() == int(np.ceil(2 * np.pi * 0.5 / fem.h))
    H, S, A = _oracle_matrices(fem, 0)
    for mine, ref in ((fem.H, H), (fem.S, S), (fem.A, A)):
        ip, ix, dv = mine.getValuesCSR()
        np.testing.assert_array_equal(ip, ref.indptr)
        np.testing.assert_array_equal(ix, ref.indices)
        assert rel_err(dv, ref.data) < 1e-15
    # KLE solve with the free stream on the boundary, then the correction: the same numbers as the oracle's direct solve
    fem.vort.set(0.0)
    fem.solveKLE(0.0, fem.vort)
    v0 = fem.vel.getArray().copy()
    fem.computeVelocityCorrection()
    vbody = fem.body.getVelocity().getArray()
    vref, qref = o.ibm_velocity_correction(H, S, A, v0, vbody)
    assert rel_err(fem.vel.getArray(), vref) < 1e-9
    assert rel_err(fem.virtualFlux.getArray(), qref) < 1e-8
    # the corrected field has the body's velocity (zero) at every marker
    assert np.abs(H @ fem.vel.getArray() - vbody).max() < 1e-9 * fem.U_ref
    assert np.abs(H @ v0 - vbody).max() > 0.1 * fem.U_ref          # ... which the uncorrected one had not
    # two time steps of the reference's loop
    fem.ts.setMaxSteps(1)
    fem.config["save-n-steps"] = 1000
    fem.startSolver()
    assert fem.ts.getStepNumber() == 2 and fem.ts.time > 0
    assert bool(torch.isfinite(fem.vort.data).all()) and float(fem.vort.data.abs().max()) > 0
    cd, cl = fem.computeDragForce(fem.ts.getTimeStep())
    assert len(cd) == 1 and np.isfinite(cd[0]) and np.isfinite(cl[0])


def test_dynamic_case_rebuilds_where_the_body_is(be):
    fem = _case("ImmersedBoundaryDynamic", moving=True)
    y0 = fem.body.getCenters()[0][1]
    fem.vort.set(0.0)
    fem.solveKLE(0.0, fem.vort)
    t = 0.37 * (5 / fem.U_ref)
    fem.computeVelocityCorrection(t)
    y1 = fem.body.getCenters()[0][1]
    assert abs(y1 - (0.3 * np.sin(2 * np.pi * 0.37))) < 1e-14 and y1 != y0
    H, S, A = _oracle_matrices(fem, 1)                   # at the new position, the dynamic closeness rule
    for mine, ref in ((fem.H, H), (fem.S, S), (fem.A, A)):
        ip, ix, dv = mine.getValuesCSR()
        np.testing.assert_array_equal(ip, ref.indptr)
        np.testing.assert_array_equal(ix, ref.indices)
        assert rel_err(dv, ref.data) < 1e-15
    vbody = fem.body.getVelocity().getArray()
    assert np.allclose(vbody[1::2], 2 * np.pi * 0.3 * np.cos(2 * np.pi * 0.37) / (5 / fem.U_ref))
    assert np.abs(H @ fem.vel.getArray() - vbody).max() < 1e-9 * max(fem.U_ref, np.abs(vbody).max())


def test_two_bodies_share_one_system(be):
    """The reference's two-cylinder set-up (commented in ibm-static.yaml): markers of both bodies in one H / S / A,
    global marker numbers body after body (BodiesContainer.mapGlobToLocal, immersed_body.py:50-67)."""
    from pynama_b200.cases import immersed_boundary as ib
    from pynama_b200.common.options import Options
    Options.clear()
    for k, v in dict(ksp_type="cg", pc_type="jacobi", ksp_rtol=1e-12, ksp_max_it=20000).items():
        Options.setValue(k, v)
    with open(f"{ROOT}/pynama_b200/cases/ibm-static.yaml") as f:
        cfg = yaml.load(f, Loader=yaml.Loader)
    cfg["domain"]["box-mesh"]["nelem"] = [32, 32]
    cfg["bodies"].append({"type": "circle", "vel": "static", "radius": 0.5, "center": [0, -1.5]})
    fem = ib.ImmersedBoundaryStatic(cfg, case="ibm-static")
    fem.setUp()
    fem.setUpSolver()
    n1 = fem.body.bodies[0].getTotalNodes()
    assert fem.body.getTotalNodes() == 2 * n1 and fem.H.getSize()[0] == 4 * n1
    assert fem.body.mapGlobToLocal(n1 + 3) == (3, 1) and fem.body.mapGlobToLocal(n1 - 1) == (n1 - 1, 0)
    np.testing.assert_allclose(fem.body.getNodeCoordinates(n1 + 3),
                               fem.body.bodies[1].getNodeCoordinates(3), rtol=0, atol=0)
    H, S, A = _oracle_matrices(fem, 0)
    for mine, ref in ((fem.H, H), (fem.A, A)):
        ip, ix, dv = mine.getValuesCSR()
        np.testing.assert_array_equal(ip, ref.indptr)
        np.testing.assert_array_equal(ix, ref.indices)
        assert rel_err(dv, ref.data) < 1e-15
    fem.vort.set(0.0)
    fem.solveKLE(0.0, fem.vort)
    fem.computeVelocityCorrection()
    assert np.abs(H @ fem.vel.getArray()).max() < 1e-9 * fem.U_ref
    fx, fy = fem.computeDragForce(0.1)
    assert len(fx) == 2 and len(fy) == 2


@pytest.mark.parametrize("tag,mode", [("static", 0), ("dynamic", 1)])
def test_device_matrices_vs_reference_golden(be, tag, mode):
    """The device lists and matrices against what the UNMODIFIED reference case code produced for the same markers and
    Euler nodes (tests/golden/ibm_ref.npz, oracle/gen_golden_ibm.py): patterns exact, values to 1e-15."""
    import os
    from conftest import GOLDEN
    g = np.load(os.path.join(GOLDEN, "ibm_ref.npz"))
    nodexy, lag, dl, h = g["case_nodexy"], g["case_lag"], float(g["case_dl"]), float(g["case_h"])
    p, nodes, dirac = be.ibm_collect(be.to_device(lag.reshape(-1)), be.to_device(g["case_cand"].astype(np.int32)),
                                     be.to_device(nodexy.reshape(-1)), dl, h, mode=mode, kind="four")
    H = be.dof_blocked_csr(p, nodes, dirac * (h ** 2), len(nodexy), 2)
    ap, aj, av = be.ibm_gram(p, nodes, dirac, h ** 2, dl * h)
    A = be.dof_blocked_csr(ap, aj, av, len(lag), 2)
    for nm, M in (("H", H), ("A", A)):
        ip, ix, dv = M.getValuesCSR()
        np.testing.assert_array_equal(ip, g[f"{tag}_{nm}_indptr"])
        np.testing.assert_array_equal(ix, g[f"{tag}_{nm}_indices"])
        assert rel_err(dv, g[f"{tag}_{nm}_data"]) < 1e-15
