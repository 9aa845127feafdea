"""`-pc_type mg`: the device V-cycle (csrc/mg.cu, solver/multigrid.py) against the oracle's restatement
of the same algorithm (oracle/kle_oracle.py MultigridPC), and MG-preconditioned CG against the oracle / a direct solve."""
import copy

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
        pytest.skip("no CUDA device")
    from pynama_b200 import backend
    return backend


def _config(case):
    with open(f"{ROOT}/pynama_b200/cases/{case}.yaml") as f:
        return yaml.load(f, Loader=yaml.Loader)


def _options(**kw):
    from pynama_b200.common.options import Options
    Options.clear()
    for k, v in kw.items():
        Options.setValue(k, v)


@pytest.fixture(scope="module")
def cavity_mg(be):
    from pynama_b200.cases.cavity import Cavity
    nx, n = 32, 5
    _options(ksp_type="cg", pc_type="mg", ksp_rtol=1e-12, ksp_max_it=500, pc_mg_coarse_cells=64)
    cfg = copy.deepcopy(_config("cavity"))
    cfg["domain"]["box-mesh"]["nelem"] = [nx, nx]
    cfg["domain"]["ngl"] = n
    cfg["material-properties"] = {"rho": 1.0, "mu": 0.01}
    fem = Cavity(cfg, case="cavity")
    fem.setUp()
    fem.setUpSolver()
    m = o.BoxMesh([nx, nx], [0, 0], [1, 1], n)
    return fem, m, o.assemble_noslip(m), o.assemble_operators(m)


def test_level_sequence_and_transfers(cavity_mg):
    fem, m, mats, ops = cavity_mg
    h = fem.solver._ksp._mg
    assert h.sequence == o.mg_coarsening_sequence(32, 32, 5, 64) == [(32, 32, 5), (32, 32, 3), (32, 32, 2), (16, 16, 2),
                                                                      (8, 8, 2)]
    pc = o.MultigridPC(32, 32, 5, "K", coarse_cells=64, lmax=[L.lmax for L in h.levels])
    for Ld, Lo in zip(h.levels, pc.levels):
        np.testing.assert_array_equal(Ld.free.cpu().numpy(), Lo["free"])
        assert rel_err(Ld.dinv.cpu().numpy(), Lo["dinv"]) < 1e-12
    for Ld, Lo in zip(h.levels[:-1], pc.levels[:-1]):
        P = Ld.P.to_scipy()
        assert abs(P - Lo["P"]).max() < 1e-12 and P.nnz == Lo["P"].nnz
        assert abs(Ld.R.to_scipy() - Lo["R"]).max() < 1e-12
    # the device's own spectral bounds agree with the oracle's power iteration to the accuracy of 30 steps
    own = o.MultigridPC(32, 32, 5, "K", coarse_cells=64)
    for Ld, Lo in zip(h.levels, own.levels):
        assert abs(Ld.lmax / Lo["lmax"] - 1.0) < 0.05


@pytest.mark.parametrize("kind", ["K", "KFS"])
def test_vcycle_and_cg_vs_oracle(cavity_mg, be, kind):
    fem, m, mats, ops = cavity_mg
    sv = fem.solver if kind == "K" else fem.solverFS
    h = sv._ksp._mg
    A = mats["K"] if kind == "K" else o.mat_add(mats["K"], mats["Kfs"])
    pc = o.MultigridPC(32, 32, 5, kind, coarse_cells=64, lmax=[L.lmax for L in h.levels])
    r = np.random.default_rng(11).standard_normal(A.shape[0])
    z = be.empty(A.shape[0])
    h.apply(be.to_device(r), z)
    z_o = pc._cycle(0, r)
    assert rel_err(z.cpu().numpy(), z_o) < 1e-10
    # symmetric positive definite cycle: <V r, s> = <r, V s>
    s = np.random.default_rng(12).standard_normal(A.shape[0])
    zs = be.empty(A.shape[0])
    h.apply(be.to_device(s), zs)
    a, b = float(z.cpu().numpy() @ s), float(zs.cpu().numpy() @ r)
    assert abs(a - b) < 1e-10 * max(abs(a), abs(b))
    # MG-preconditioned CG, PETSc semantics, matched rtol
    b = np.random.default_rng(13).standard_normal(A.shape[0])
    x_o, its_o, reason_o, _ = o.ksp_cg_jacobi(A, b, rtol=1e-12, pc=pc)
    x = be.empty(A.shape[0])
    res = sv._ksp.solve(be.to_device(b), x)
    assert res.reason == reason_o == 2 and abs(res.its - its_o) <= 1, (res.its, its_o)
    assert rel_err(x.cpu().numpy(), x_o) < 1e-10
    x_j, its_j, _, _ = o.ksp_cg_jacobi(A, b, rtol=1e-12)
    assert res.its * 8 < its_j                     # what the cycle buys: iterations, not accuracy
    assert rel_err(x.cpu().numpy(), o.direct_solve(A, b)) < 1e-9


def test_cavity_solve_kle_with_mg_vs_oracle(cavity_mg):
    fem, m, mats, ops = cavity_mg
    vb, fd, fv = o.cavity_wall_vectors(m)
    vort = np.random.default_rng(3).standard_normal(m.nnodes)
    vel_o, velfs_o, _ = o.solve_kle_noslip(mats, ops, vb, fd, fv, vort)
    w = fem.vort.duplicate()
    w.setArray(vort)
    fem.solveKLE(0.0, w)
    assert fem.solver.reason > 0 and fem.solverFS.reason > 0
    assert fem.solver.its < 150 and fem.solverFS.its < 150      # rtol 1e-12; Jacobi needs 1500-2500
    assert rel_err(fem.velFS.getArray(), velfs_o) < 1e-10
    assert rel_err(fem.vel.getArray(), vel_o) < 1e-10


def test_taylor_green_freeslip_with_mg(be):
    """FreeSlip case (element matrices integrated per cell on the fine level, one element matrix on the coarse ones)."""
    from pynama_b200.cases.custom_func import CustomFuncCase
    _options(ksp_type="cg", pc_type="mg", ksp_rtol=1e-12, ksp_max_it=500, pc_mg_coarse_cells=16)
    nx, n = 16, 5
    fem = CustomFuncCase(_config("taylor-green"), case="taylor-green", ngl=n, nelem=[nx, nx])
    fem.setUp()
    fem.setUpSolver()
    m = o.BoxMesh([nx, nx], [0, 0], [1, 1], n)
    mats = o.assemble_freeslip(m)
    xy, bc = m.node_coords(), mats["bc_nodes"]
    vel_ex, w0 = o.taylor_green(xy, fem.nu, 0.0)
    vb = np.zeros(2 * m.nnodes)
    vb.reshape(-1, 2)[bc] = vel_ex[bc]
    x_d = o.direct_solve(mats["K"], mats["Rw"] @ w0 + mats["Krhs"] @ vb)
    fem.computeInitialCondition(0.0)
    fem.solveKLE(0.0, fem.vort)
    assert fem.solver.reason == 2 and fem.solver.its < 60
    assert rel_err(fem.vel.getArray(), x_d) < 1e-10
