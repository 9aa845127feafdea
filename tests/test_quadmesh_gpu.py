"""an unstructured, skewed quadrilateral mesh read from a gmsh file through the case class -- element
matrices AND operators integrated for every cell -- against the oracle's restatement on the same mesh."""
import math

import numpy as np
import pytest
import yaml

import kle_oracle as o
from conftest import ROOT, rel_err
from quadmesh_util import skewed_square, write_msh

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def skew(tmp_path_factory):
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from pynama_b200.cases.custom_func import CustomFuncCase
    from pynama_b200.common.options import Options
    Options.clear()
    for k, v in dict(ksp_type="cg", pc_type="jacobi", ksp_rtol=1e-12, ksp_max_it=100000).items():
        Options.setValue(k, v)
    n, ngl = 8, 4
    P, Q = skewed_square(n)
    path = str(tmp_path_factory.mktemp("msh") / "skew8.msh")
    write_msh(path, P, Q)
    with open(f"{ROOT}/pynama_b200/cases/taylor-green.yaml") as f:
        cfg = yaml.load(f, Loader=yaml.Loader)
    cfg["domain"] = {"ngl": ngl, "gmsh-file": path}
    fem = CustomFuncCase(cfg, case="taylor-green")
    fem.setUp()
    fem.setUpSolver()
    m = o.QuadMesh(P, Q, ngl)
    return fem, m, o.assemble_freeslip(m, per_cell=True), o.assemble_operators(m, per_cell=True)


def test_unstructured_matrices_vs_oracle(skew):
    fem, m, mats, ops = skew
    assert not fem.mat.fusedAssembly                     # skewed cells are not parallelograms: general kernels
    assert rel_err(fem.dom.fullCoordVec.getArray().reshape(-1, 2), m.node_coords()) < 1e-15
    for name in ("K", "Krhs", "Rw"):
        ip, ix, dv = getattr(fem.mat, name).getValuesCSR()
        np.testing.assert_array_equal(ip, mats[name].indptr)
        np.testing.assert_array_equal(ix, mats[name].indices)
        assert rel_err(dv, mats[name].data) < 1e-12, name
    for name in ("SrT", "DivSrT", "Curl"):
        ip, ix, dv = getattr(fem.operator, name).getValuesCSR()
        np.testing.assert_array_equal(ip, ops[name].indptr)
        np.testing.assert_array_equal(ix, ops[name].indices)
        assert rel_err(dv, ops[name].data) < 1e-12, name
    assert rel_err(fem.operator.weights.cpu().numpy(), ops["W"]) < 1e-13
    assert abs(ops["W"].sum() - 1.0) < 1e-12             # lumped weights add up to the area of the unit square


def test_unstructured_solve_and_rhs_vs_oracle(skew):
    fem, m, mats, ops = skew
    xy, bc, nu = m.node_coords(), mats["bc_nodes"], fem.nu
    vel_ex, w0 = o.taylor_green(xy, nu, 0.0)
    vb = np.zeros(2 * m.nnodes)
    vb.reshape(-1, 2)[bc] = vel_ex[bc]
    x_d = o.direct_solve(mats["K"], mats["Rw"] @ w0 + mats["Krhs"] @ vb)
    fem.computeInitialCondition(0.0)
    fem.solveKLE(0.0, fem.vort)
    assert fem.solver.reason == 2
    assert rel_err(fem.vel.getArray(), x_d) < 1e-10
    assert np.abs(fem.vel.getArray() - vel_ex.reshape(-1)).max() < 5e-3      # discretisation error at 8x8, ngl 4
    fv = fem.vort.duplicate()
    fem.solveKLE = lambda *a, **k: None
    fem.vel.setArray(x_d)
    fem.evalRHS(fem.ts, 0.0, fem.vort, fv)
    f_o = o.eval_rhs_from_vel(ops, x_d, fem.rho, fem.mu)
    assert rel_err(fv.getArray(), f_o) < 1e-9
