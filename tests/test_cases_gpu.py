"""Case-level parity on the GPU: the reference's own known answers (src/tests/test_solver.py) run through the
drop-in classes, and the cavity / Taylor-Green flows (solveKLE, evalRHS, RK steps) against the CPU oracle."""
import copy
import math

import numpy as np
import pytest
import yaml

import kle_oracle as o
from conftest import ROOT, rel_err

pytestmark = pytest.mark.gpu

TOL_ASM, TOL_KSP = 1e-12, 1e-10


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


def _csr_equal(A, ref, exact=True):
    ip, ix, dv = A.getValuesCSR()
    np.testing.assert_array_equal(ip, ref.indptr)
    np.testing.assert_array_equal(ix, ref.indices)
    if exact:
        np.testing.assert_array_equal(dv, ref.data)
    else:
        assert rel_err(dv, ref.data) < TOL_ASM


# ---- the reference's own known answers (src/tests/test_solver.py) ------------------------------------------------
def test_uniform_flow_known_answer(be):
    """test_solver.py:20-27: 10x10, ngl 3, zero vorticity -> the constant field to 1e-12 (default preonly+lu)."""
    from pynama_b200.cases.uniform import UniformFlow
    _options()
    fem = UniformFlow(_config("uniform"), case="uniform")
    fem.setUp()
    fem.setUpSolver()
    exactVel, exactVort = fem.generateExactVecs()
    fem.solveKLE(0.0, exactVort)
    error = (exactVel - fem.vel).norm(norm_type=2)
    assert error < 1e-12


def test_taylor_green_known_answer(be):
    """test_solver.py:29-37: 2x2 elements, ngl 11, ||v - v_exact||_2 < 2e-8."""
    from pynama_b200.cases.custom_func import CustomFuncCase
    _options()
    fem = CustomFuncCase(_config("taylor-green"), case="taylor-green", ngl=11, nelem=[2, 2])
    fem.setUp()
    fem.setUpSolver()
    exactVel, exactVort = fem.generateExactVecs(0.0)
    fem.solveKLE(0.0, exactVort)
    error = (exactVel - fem.vel).norm(norm_type=2)
    assert error < 2e-8


def test_vtensv_known_answer(be):
    """test_solver.py:66-87: computeVtensV(vec) = [vx^2, vx vy, vy^2] per node."""
    from pynama_b200.cases.uniform import UniformFlow
    _options()
    fem = UniformFlow(_config("uniform"), case="uniform", nelem=[2, 2], ngl=3)
    fem.setUp()
    fem.setUpSolver()
    n = fem.vel.getSize() // 2
    vec_init = np.arange(2 * n, dtype=float)
    ref = np.empty(3 * n)
    ref[0::3], ref[1::3], ref[2::3] = vec_init[0::2] ** 2, vec_init[0::2] * vec_init[1::2], vec_init[1::2] ** 2
    vec = fem.vel.duplicate()
    vec.setArray(vec_init)
    fem.computeVtensV(vec=vec)
    np.testing.assert_array_equal(fem._VtensV.getArray(), ref)


# ---- cavity (configs[0]: 8x8, ngl 3) vs the oracle ------------------------------------------------------------------
@pytest.fixture(scope="module")
def cavity(be):
    from pynama_b200.cases.cavity import Cavity
    _options(ksp_type="cg", pc_type="jacobi", ksp_rtol=1e-13, ksp_max_it=20000, ts_adapt_type="none", ts_dt=0.005)
    cfg = copy.deepcopy(_config("cavity"))
    cfg["domain"]["box-mesh"]["nelem"] = [8, 8]
    cfg["material-properties"] = {"rho": 1.0, "mu": 0.01}      # Re = 100
    fem = Cavity(cfg, case="cavity")
    fem.setUp()
    fem.setUpSolver()
    m = o.BoxMesh([8, 8], [0, 0], [1, 1], 3)
    return fem, m, o.assemble_noslip(m), o.assemble_operators(m)


def test_cavity_matrices_vs_oracle(cavity):
    fem, m, mats, ops = cavity
    for name in ("K", "Krhs", "Kfs", "Krhsfs", "Rw", "Rwfs", "Rd", "Rdfs"):
        _csr_equal(getattr(fem.mat, name), mats[name], exact=False)
    for name in ("SrT", "DivSrT", "Curl"):
        _csr_equal(getattr(fem.operator, name), ops[name], exact=False)
    KFS = o.mat_add(mats["K"], mats["Kfs"])
    _csr_equal(fem.solverFS.mat, KFS, exact=False)
    np.testing.assert_array_equal(fem.mat.dofClasses.cpu().numpy(), mats["cls"])


def test_cavity_solve_kle_and_rhs_vs_oracle(cavity):
    fem, m, mats, ops = cavity
    vb, fd, fv = o.cavity_wall_vectors(m)
    rng = np.random.default_rng(3)
    vort = rng.standard_normal(m.nnodes)
    vel_o, velfs_o, _ = o.solve_kle_noslip(mats, ops, vb, fd, fv, vort)
    w = fem.vort.duplicate()
    w.setArray(vort)
    fem.solveKLE(0.0, w)
    assert fem.solver.reason > 0 and fem.solverFS.reason > 0
    assert rel_err(fem.velFS.getArray(), velfs_o) < TOL_KSP
    assert rel_err(fem.vel.getArray(), vel_o) < TOL_KSP
    # wall values: lid tangential velocity 1 (corners included), other wall dofs 0
    np.testing.assert_allclose(fem.vel.getArray()[mats["cls"] > 0], vb[mats["cls"] > 0], rtol=0, atol=1e-12)
    f = fem.vort.duplicate()
    fem.evalRHS(fem.ts, 0.0, w, f)
    f_o = o.eval_rhs_from_vel(ops, vel_o, fem.rho, fem.mu)
    assert rel_err(f.getArray(), f_o) < 1e-9           # Curl . DivSrT . SrT amplifies the 1e-10 of vel by ~|D|^3 h


def test_cavity_fixed_step_rk_vs_oracle(cavity):
    """A few fixed-dt RK5BS steps of the cavity from rest (TS with -ts_adapt_type none): vorticity vs the oracle's
    restated loop.  Short horizon only (trajectories are not a parity target)."""
    fem, m, mats, ops = cavity
    vb, fd, fv = o.cavity_wall_vectors(m)
    KFS = o.mat_add(mats["K"], mats["Kfs"])

    def f(t, w):
        vel, _, _ = o.solve_kle_noslip(mats, ops, vb, fd, fv, w, KFS=KFS)
        return o.eval_rhs_from_vel(ops, vel, fem.rho, fem.mu)

    nsteps, dt = 3, 0.005
    _, w_o = o.rk_fixed_steps(f, 0.0, np.zeros(m.nnodes), dt, nsteps)
    fem.ts.setTime(0.0)
    fem.ts.step_number = 0
    fem.ts.setMaxSteps(nsteps)
    fem.ts.setTimeStep(dt)
    fem.startSolver()
    assert fem.ts.step_number == nsteps and abs(fem.ts.time - nsteps * dt) < 1e-14
    assert fem.ts.rhs_evaluations == 7 * nsteps + 1       # FSAL: 8 evaluations for the first step, 7 afterwards
    assert rel_err(fem.vort.getArray(), w_o) < 1e-8


# ---- Taylor-Green (configs[1] shape, reduced to 16x16 for the oracle) ------------------------------------------------
def test_taylor_green_rhs_and_steps_vs_oracle(be):
    from pynama_b200.cases.custom_func import CustomFuncCase
    _options(ksp_type="cg", pc_type="jacobi", ksp_rtol=1e-13, ksp_max_it=50000, ts_adapt_type="none", ts_dt=0.01)
    nx, n = 8, 5
    fem = CustomFuncCase(_config("taylor-green"), case="taylor-green", ngl=n, nelem=[nx, nx])
    fem.setUp()
    fem.setUpSolver()
    m = o.BoxMesh([nx, nx], [0, 0], [1, 1], n)
    mats, ops = o.assemble_freeslip(m), o.assemble_operators(m)
    for name in ("K", "Krhs", "Rw"):
        _csr_equal(getattr(fem.mat, name), mats[name], exact=False)
    assert rel_err(fem.dom.fullCoordVec.getArray().reshape(-1, 2), m.node_coords()) < 1e-15
    xy, bc, nu = m.node_coords(), mats["bc_nodes"], fem.nu

    def f(t, w):
        vel_ex, _ = o.taylor_green(xy, nu, t)
        vb = np.zeros(2 * m.nnodes)
        vb.reshape(-1, 2)[bc] = vel_ex[bc]
        vel = o.solve_kle_freeslip(mats, vb, w)
        return o.eval_rhs_from_vel(ops, vel, fem.rho, fem.mu)

    _, w0 = o.taylor_green(xy, nu, 0.0)
    fem.computeInitialCondition(0.0)
    assert rel_err(fem.vort.getArray(), w0) < 1e-14
    fv = fem.vort.duplicate()
    fem.evalRHS(fem.ts, 0.0, fem.vort, fv)
    f_o = f(0.0, w0)
    assert rel_err(fv.getArray(), f_o) < 1e-9
    # analytic decay d(omega)/dt = -8 pi^2 nu omega, interior nodes, spectral accuracy at this resolution
    interior = np.setdiff1d(np.arange(m.nnodes), bc)
    assert np.abs(fv.getArray()[interior] + 8 * math.pi ** 2 * nu * w0[interior]).max() < 0.05 * np.abs(w0).max()

    # three fixed steps; the oracle mimics CustomFuncCase.applyBoundaryConditions overwriting the boundary entries of
    # the SOLUTION vector at every stage time (custom_func.py:81), the stage vectors being separate work vectors
    A, b, bh, c = o.rk5bs_tableau()
    dt, nsteps = 0.01, 3
    y, t, k1 = w0.copy(), 0.0, None
    for _ in range(nsteps):
        k = []
        for i in range(8):
            if i == 0 and k1 is not None:
                k.append(k1)
                continue
            yi = y.copy()
            for j in range(i):
                if A[i, j] != 0.0:
                    yi = yi + (dt * A[i, j]) * k[j]
            ti = t + c[i] * dt
            k.append(f(ti, yi))
            y[bc] = o.taylor_green(xy, nu, ti)[1][bc]          # side effect on the solution vector
        for j in range(8):
            if b[j] != 0.0:
                y = y + (dt * b[j]) * k[j]
        k1, t = k[7], t + dt
    fem.ts.setTime(0.0)
    fem.ts.step_number = 0
    fem.ts.setMaxSteps(nsteps)
    fem.ts.setTimeStep(dt)
    fem.startSolver()
    assert rel_err(fem.vort.getArray(), y) < 1e-8


# ---- the integrator alone ---------------------------------------------------------------------------------------------
def test_ts_adaptive_matches_restated_adaptor(be):
    """y' = lambda*y on device vectors: same accepted step sizes / rejections as the oracle's restatement of TSRK 5bs
    + TSAdapt basic + MATCHSTEP, final time hit exactly."""
    from pynama_b200.petsc_like import Vec
    from pynama_b200.solver.ts_solver import TsSolver
    _options()
    lam = np.array([-1.0, -25.0, -3.0, 0.5])

    class Prob:
        def __init__(self):
            self.l = Vec.createWithArray(lam)

        def rhs(self, ts, t, X, Fv):
            Fv.data.copy_(X.data * self.l.data)

    p = Prob()
    ts = TsSolver(None)
    ts.setUpTimes(0.0, 2.0, 1000)
    hs = []
    ts.initSolver(p.rhs, lambda ts_: hs.append(ts_.time))
    y = Vec.createWithArray(np.ones(4))
    ts.solve(y)
    t_o, y_o, hs_o, nrej_o = o.rk_adaptive(lambda t, yy: lam * yy, 0.0, np.ones(4), 2.0)
    assert ts.time == 2.0 and t_o == 2.0
    assert ts.step_number == len(hs_o) and ts.rejections == nrej_o
    np.testing.assert_allclose(np.diff([0.0] + hs), hs_o, rtol=1e-12)
    assert rel_err(y.getArray(), y_o) < 1e-13
    assert rel_err(y.getArray(), np.exp(2.0 * lam)) < 1e-3


def test_ts_order_of_accuracy(be):
    from pynama_b200.petsc_like import Vec
    from pynama_b200.solver.ts_solver import TsSolver
    errs = {}
    for rk, order in (("5bs", 5), ("5dp", 5), ("3bs", 3), ("4", 4), ("1fe", 1)):
        e = []
        for dt in (0.1, 0.05):
            _options(ts_adapt_type="none", ts_dt=dt, ts_rk_type=rk)
            ts = TsSolver(None)
            ts.setUpTimes(0.0, 1.0, 100000)
            ts.initSolver(lambda ts_, t, X, Fv: Fv.data.copy_(-X.data * (1.0 + t)), None)
            y = Vec.createWithArray(np.ones(2))
            ts.solve(y)
            assert abs(ts.time - 1.0) < 1e-12
            e.append(abs(y.getArray()[0] - math.exp(-1.5)))
        errs[rk] = math.log2(e[0] / e[1])
        assert abs(errs[rk] - order) < 0.6, errs
