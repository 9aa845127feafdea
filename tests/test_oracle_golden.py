"""CPU tests: the oracle (oracle/kle_oracle.py) and the product's host tables against the golden vectors made by the
UNMODIFIED reference (oracle/gen_golden.py) and against the reference's own known answers."""
import math
import os

import numpy as np
import pytest

import kle_oracle as o
from conftest import GOLDEN, rel_err


# ---- closed-form rules pinned by the reference (src/tests/test_element.py:181-229) -----------------------
def test_gauss_closed_form():
    x, w = o.gauss_points(2)
    np.testing.assert_array_equal(x, [-1 / math.sqrt(3), 1 / math.sqrt(3)])
    np.testing.assert_array_equal(w, [1, 1])
    x, w = o.gauss_points(3)
    np.testing.assert_allclose(x, [-math.sqrt(3 / 5), 0, math.sqrt(3 / 5)], atol=1e-12)
    np.testing.assert_allclose(w, [5 / 9, 8 / 9, 5 / 9], atol=1e-12)


def test_lobatto_closed_form():
    exp = {2: ([-1, 1], [1, 1]), 3: ([-1, 0, 1], [1 / 3, 4 / 3, 1 / 3]),
           4: ([-1, -1 / math.sqrt(5), 1 / math.sqrt(5), 1], [1 / 6, 5 / 6, 5 / 6, 1 / 6])}
    for n, (x, w) in exp.items():
        xx, ww = o.lobatto_points(n)
        np.testing.assert_allclose(xx, x, atol=1e-12)
        np.testing.assert_allclose(ww, w, atol=1e-12)


def test_rules_match_reference(golden_tables):
    from pynama_b200.elements.utilities import gaussPoints, lobattoPoints
    for N in range(1, 12):
        for fn in (o.gauss_points, gaussPoints):
            x, w = fn(N)
            np.testing.assert_array_equal(x, golden_tables[f"gauss_x_{N}"])
            np.testing.assert_array_equal(w, golden_tables[f"gauss_w_{N}"])
        if N >= 2:
            for fn in (o.lobatto_points, lobattoPoints):
                x, w = fn(N)
                np.testing.assert_array_equal(x, golden_tables[f"lobatto_x_{N}"])
                np.testing.assert_array_equal(w, golden_tables[f"lobatto_w_{N}"])


@pytest.mark.parametrize("n", [2, 3, 4, 5])
def test_tabulations_match_reference(golden_tables, n):
    """H / Hrs / Gauss-point lists of Spectral(n,2): oracle and product host tables vs the reference, bit for bit."""
    from pynama_b200.elements.spectral import Spectral
    t = o.tables(n)
    s = Spectral(n, 2)
    pairs = {"H": (t.full.H, s.H), "Hrs": (t.full.Hrs, s.Hrs), "HRed": (t.red.H, s.HRed),
             "HrsRed": (t.red.Hrs, s.HrsRed), "HOp": (t.op.H, s.HOp), "HrsOp": (t.op.Hrs, s.HrsOp),
             "HCoo": (t.coo_full.H, s.HCoo), "HrsCoo": (t.coo_full.Hrs, s.HrsCoo),
             "HCooRed": (t.coo_red.H, s.HCooRed), "HrsCooRed": (t.coo_red.Hrs, s.HrsCooRed),
             "HCooOp": (t.coo_op.H, s.HCooOp), "HrsCooOp": (t.coo_op.Hrs, s.HrsCooOp)}
    for name, (orc, prod) in pairs.items():
        ref = golden_tables[f"tab_{name}_{n}"]
        np.testing.assert_array_equal(orc, ref, err_msg=f"oracle {name}")
        np.testing.assert_array_equal(np.array(prod), ref, err_msg=f"product {name}")
    for name, tab, plist in (("gps", t.full, s.gps), ("gpsRed", t.red, s.gpsRed), ("gpsOp", t.op, s.gpsOp)):
        ref = golden_tables[f"tab_{name}_{n}"]
        np.testing.assert_array_equal(np.stack([tab.r, tab.s, tab.w], 1), ref)
        np.testing.assert_array_equal(np.array([[g.r, g.s, g.w] for g in plist]), ref)


def test_node_count_formulas():
    """src/tests/test_element.py:231-266"""
    from pynama_b200.elements.spectral import Spectral
    for i, n in enumerate(range(2, 5)):
        s = Spectral(n, 2)
        assert s.nnode == n ** 2 and s.nnodedge == i and s.nnodcell == i ** 2


@pytest.mark.parametrize("n", range(2, 10))
def test_oracle_element_matrices_vs_reference(golden_elem, n):
    for q in ("affine", "skew", "reftest"):
        if f"K_{n}_{q}" not in golden_elem:
            continue
        c = golden_elem[f"coords_{q}"][None]
        K, Rw, Rd = o.elem_kle_matrices(n, c)
        S, D, C, W = o.elem_kle_operators(n, c)
        for nm, a in zip(("K", "Rw", "Rd", "SrT", "DivSrT", "Curl", "W"), (K, Rw, Rd, S, D, C, W)):
            assert rel_err(a[0], golden_elem[f"{nm}_{n}_{q}"]) < 1e-13, (nm, n, q)


# ---- numbering: what the reference's test_domain.py pins ------------------------------------------------------
def test_box_numbering_pins():
    m = o.BoxMesh([3, 4], [0, 0], [0.6, 0.8], 3)
    assert m.ncell == 12                                                        # test_domain.py:22-24
    np.testing.assert_allclose(m.cell_corner_coords([0])[0].reshape(-1),
                               [0, 0, 0.2, 0, 0.2, 0.2, 0, 0.2], atol=1e-13)      # :26-30
    assert len(m.boundary_nodes()) == 28                                        # :32-42
    for b in ("up", "down", "left", "right"):
        assert len(m.border_nodes(b)) == (7 if b in ("up", "down") else 9)      # :44-50
    for n in range(2, 10, 2):                                                   # :61-78
        m = o.BoxMesh([2, 3], [0, 0], [0.6, 0.8], n)
        assert len(m.boundary_nodes()) == 10 + 10 * (n - 2)
        for b in ("up", "down", "left", "right"):
            assert len(m.border_nodes(b)) == ((3 + 2 * (n - 2)) if b in ("up", "down") else (4 + 3 * (n - 2)))
    for n in range(2, 14):                                                      # :187-195
        m = o.BoxMesh([2, 3], [0, 0], [0.6, 0.8], n)
        assert m.nnodes == 12 + 17 * (n - 2) + 6 * (n - 2) ** 2
        assert len(np.unique(m.cell_nodes())) == m.nnodes
    m = o.BoxMesh([2, 2], [0, 0], [1, 1], 2)                                    # :197-201 row-major numbering
    np.testing.assert_allclose(m.node_coords(), [[0, 0], [.5, 0], [1, 0], [0, .5], [.5, .5], [1, .5],
                                                 [0, 1], [.5, 1], [1, 1]], atol=1e-14)


# ---- known answers of src/tests/test_solver.py ------------------------------------------------------------------
def _direct(A, b):
    import scipy.sparse.linalg as spl
    return spl.spsolve(A.tocsc(), b)


def test_uniform_flow_known_answer():
    """test_solver.py:20-27: ||v - v_exact||_2 < 1e-12 (10x10, ngl 3)."""
    m = o.BoxMesh([10, 10], [0, 0], [1, 1], 3)
    A = o.assemble_freeslip(m)
    vel = np.zeros(2 * m.nnodes)
    vel.reshape(-1, 2)[A["bc_nodes"]] = [1.0, 0.0]
    x = _direct(A["K"], A["Rw"] @ np.zeros(m.nnodes) + A["Krhs"] @ vel)
    assert np.linalg.norm(x - np.tile([1.0, 0.0], m.nnodes)) < 1e-12


def test_taylor_green_known_answer():
    """test_solver.py:29-37: 2x2 elements, ngl 11: ||v - v_exact||_2 < 2e-8."""
    m = o.BoxMesh([2, 2], [0, 0], [1, 1], 11)
    A = o.assemble_freeslip(m)
    vel_ex, vort_ex = o.taylor_green(m.node_coords(), 0.01 / 0.5, 0.0)
    vel = np.zeros(2 * m.nnodes)
    vel.reshape(-1, 2)[A["bc_nodes"]] = vel_ex[A["bc_nodes"]]
    x = _direct(A["K"], A["Rw"] @ vort_ex + A["Krhs"] @ vel)
    assert np.linalg.norm(x - vel_ex.reshape(-1)) < 2e-8


def test_vtensv_known_answer():
    """test_solver.py:66-87"""
    v = np.arange(1.0, 19.0)
    ref = np.array([[a * a, a * b, b * b] for a, b in v.reshape(-1, 2)]).reshape(-1)
    np.testing.assert_array_almost_equal(o.vtensv(v), ref, decimal=10)


def test_cavity_matrices_structure():
    """No reference-side known answer exists for the cavity: structural identities only."""
    m = o.BoxMesh([6, 6], [0, 0], [1, 1], 3)
    B = o.assemble_noslip(m)
    KK = o.mat_add(B["K"], B["Kfs"])
    assert abs(B["K"] - B["K"].T).max() == 0 and abs(KK - KK.T).max() < 1e-12
    assert np.linalg.eigvalsh(KK.toarray()).min() > 0 and np.linalg.eigvalsh(B["K"].toarray()).min() > 0
    cls = B["cls"]
    d = B["K"].diagonal()
    np.testing.assert_array_equal(d[cls > 0], 1.0)
    ops = o.assemble_operators(m)
    assert abs(ops["W"].sum() - 1.0) < 1e-13                       # sum of lumped weights = area
    vel = np.tile([0.3, -0.2], m.nnodes)                            # rigid translation
    assert np.abs(ops["Curl"] @ vel).max() < 1e-12 and np.abs(ops["SrT"] @ vel).max() < 1e-12


def test_rk5bs_tableau_order_conditions():
    A, b, bh, c = o.rk5bs_tableau()
    np.testing.assert_allclose(A.sum(1), c, atol=1e-15)
    np.testing.assert_allclose(A[7, :7], b[:7], atol=0)              # FSAL
    for k in range(1, 6):                                           # quadrature conditions up to order 5 / 4
        assert abs(b @ c ** (k - 1) - 1.0 / k) < 1e-14
        if k <= 4:
            assert abs(bh @ c ** (k - 1) - 1.0 / k) < 1e-14
    y, err, _ = o.rk5bs_step(lambda t, y: -y, 0.0, np.array([1.0]), 0.1)
    assert abs(y[0] - math.exp(-0.1)) < 1e-9


def test_rhs_chain_taylor_green_decay():
    """d(omega)/dt = -8 pi^2 nu omega for Taylor-Green: spectral convergence of the operator chain."""
    m = o.BoxMesh([4, 4], [0, 0], [1, 1], 9)
    ops = o.assemble_operators(m)
    nu, rho, mu = 0.02, 0.5, 0.01
    vel, vort = o.taylor_green(m.node_coords(), nu, 0.0)
    f = o.eval_rhs_from_vel(ops, vel.reshape(-1), rho, mu)
    interior = np.setdiff1d(np.arange(m.nnodes), m.boundary_nodes())
    assert np.abs(f[interior] - (-8 * math.pi ** 2 * nu) * vort[interior]).max() < 5e-2 * np.abs(vort).max()


def test_cg_initial_guess_nonzero_semantics():
    """KSPCG with -ksp_initial_guess_nonzero (restated): r0 = b - A x0, the relative tolerance is measured against the
    preconditioned right-hand side (KSPConvergedDefault), so a close guess saves iterations and an exact one stops at
    iteration 0; with b = 0 the initial residual is the yardstick.  x0 = 0 reproduces the default path bit for bit."""
    m = o.BoxMesh([8, 8], [0, 0], [1, 1], 4)
    ref = o.assemble_freeslip(m, per_cell=False)
    K = ref["K"]
    b = ref["Rw"] @ np.random.default_rng(0).standard_normal(m.nnodes)
    xs = o.direct_solve(K, b)
    x0 = xs * (1 + 1e-3 * np.random.default_rng(5).standard_normal(xs.size))
    xa, ita, ra, ha = o.ksp_cg_jacobi(K, b, rtol=1e-10)
    xg, itg, rg, hg = o.ksp_cg_jacobi(K, b, rtol=1e-10, x0=x0)
    assert (ra, rg) == (2, 2) and 0 < itg < ita
    assert rel_err(xg, xs) < 1e-8 and hg[-1] <= 1e-10 * ha[0]
    xz, itz, rz, hz = o.ksp_cg_jacobi(K, b, rtol=1e-10, x0=np.zeros_like(b))
    assert itz == ita and np.array_equal(xz, xa)
    assert o.ksp_cg_jacobi(K, b, rtol=1e-6, x0=xs)[1:3] == (0, 2)
    xb, itb, rb, hb = o.ksp_cg_jacobi(K, np.zeros_like(b), rtol=1e-6, x0=x0)
    assert rb == 2 and itb > 0 and hb[-1] <= 1e-6 * hb[0] and np.abs(xb).max() < 1e-4 * np.abs(x0).max()


# ---- immersed bodies: the reference's own delta functions and marker coordinates (src/domain/immersed_body.py) ---------
def test_ibm_oracle_vs_reference_golden():
    """tests/golden/ibm_ref.npz holds outputs of the UNMODIFIED reference module (oracle/gen_golden_ibm.py): the three
    discrete delta functions on 501 radii, ImmersedBody.getDiracs on 400 offsets, the markers of Circle / Line /
    OpenBox.  The oracle -- and the host functions of the product -- reproduce them bit for bit."""
    from pynama_b200.domain import immersed_body as ib
    g = np.load(os.path.join(GOLDEN, "ibm_ref.npz"))
    for kind, name, host in (("four", "fourGrid", ib.fourGrid), ("three", "threeGrid", ib.threeGrid),
                             ("linear", "linear", ib.linear)):
        np.testing.assert_array_equal([o.ibm_phi(kind, r) for r in g["r"]], g[f"phi_{name}"])
        np.testing.assert_array_equal([float(host(r)) for r in g["r"]], g[f"phi_{name}"])
    h = float(g["dirac_h"])
    np.testing.assert_array_equal([o.ibm_dirac("four", d, h) for d in g["dirac_dist"]], g["dirac"])
    for i in range(3):
        radius, hh = g[f"circle{i}_args"]
        xy, dl = o.ibm_circle(radius, hh)
        np.testing.assert_array_equal(xy, g[f"circle{i}_xy"])
        assert dl == float(g[f"circle{i}_dl"])
    for i in range(2):
        np.testing.assert_array_equal(o.ibm_line(float(g[f"line{i}_args"][0]))[0], g[f"line{i}_xy"])
        np.testing.assert_array_equal(o.ibm_open_box(float(g[f"box{i}_args"][0]))[0], g[f"box{i}_xy"])


def test_ibm_oracle_vs_live_reference():
    """Where /root/reference exists (the build container): the same comparison against the module itself."""
    import ref_shim
    if not ref_shim.reference_available():
        pytest.skip("reference tree not present")
    ref = ref_shim.import_reference_immersed_body()
    rng = np.random.default_rng(5)
    for r in rng.uniform(0, 2.5, 200):
        assert o.ibm_phi("four", r) == ref.fourGrid(r) and o.ibm_phi("three", r) == ref.threeGrid(r)
        assert o.ibm_phi("linear", r) == ref.linear(r)
    body = ref.Circle(vel=[0, 0], center=[0, 0], radius=0.5)
    for d in rng.uniform(-0.2, 0.2, (50, 2)):
        assert o.ibm_dirac("four", d, 0.07) == body.getDiracs(d, 0.07)
    coords, _, dl = ref.Circle(vel=[0, 0], center=[0, 0], radius=0.41).generateBody(0.033)
    xy, dl2 = o.ibm_circle(0.41, 0.033)
    np.testing.assert_array_equal(xy, np.array(coords))
    assert dl == dl2


def _ibm_case_from_golden(g, dynamic):
    col = o.ibm_collect(g["case_lag"], g["case_cand"], g["case_nodexy"], float(g["case_dl"]), float(g["case_h"]),
                        mode=1 if dynamic else 0, kind="four")
    H, S, A = o.ibm_matrices(col, len(g["case_nodexy"]), float(g["case_dl"]), float(g["case_h"]))
    vnew, q = o.ibm_velocity_correction(H, S, A, g["case_vel"], g["case_body_vel"])
    return {"H": H, "S": S, "A": A}, vnew, q


@pytest.mark.parametrize("tag", ["static", "dynamic"])
def test_ibm_case_code_vs_reference_golden(tag):
    """H, S, A, the corrected velocity and the virtual flux of the UNMODIFIED collectNodes / computeClose / computeDirac /
    buildIBMMatrix / computeVelocityCorrection (src/cases/immersed_boundary.py, run on scipy stand-ins for PETSc's Mat
    and Vec: oracle/ref_shim.run_reference_ibm): the oracle's restatement gives the same patterns and the same bits."""
    g = np.load(os.path.join(GOLDEN, "ibm_ref.npz"))
    mats, vnew, q = _ibm_case_from_golden(g, tag == "dynamic")
    for nm, M in mats.items():
        np.testing.assert_array_equal(M.indptr, g[f"{tag}_{nm}_indptr"])
        np.testing.assert_array_equal(M.indices, g[f"{tag}_{nm}_indices"])
        np.testing.assert_array_equal(M.data, g[f"{tag}_{nm}_data"])
    assert rel_err(vnew, g[f"{tag}_vel"]) < 1e-13 and rel_err(q, g[f"{tag}_flux"]) < 1e-12     # (two sparse LU solves)


def test_ibm_case_code_vs_live_reference():
    import ref_shim
    if not ref_shim.reference_available():
        pytest.skip("reference tree not present")
    g = np.load(os.path.join(GOLDEN, "ibm_ref.npz"))
    H, S, A, vnew, q = ref_shim.run_reference_ibm(g["case_nodexy"], g["case_cand"], g["case_lag"], float(g["case_dl"]),
                                                  float(g["case_h"]), False, g["case_vel"], g["case_body_vel"])
    np.testing.assert_array_equal(H.data, g["static_H_data"])
    np.testing.assert_array_equal(A.indices, g["static_A_indices"])
    np.testing.assert_array_equal(A.data, g["static_A_data"])
    np.testing.assert_array_equal(vnew, g["static_vel"])


# ---- the assembly loops themselves: the reference's buildKLEMats run unmodified on stand-ins ---------------------------
def test_assembly_oracle_vs_reference_golden():
    """tests/golden/asm_ref.npz: every matrix that NoSlipFreeSlip.buildKLEMats / FreeSlip.buildKLEMats of the UNMODIFIED
    reference (src/cases/base_problem.py:329-454, 499-552) produce on three box meshes, with PETSc's AIJ matrices and
    the mesh queries replaced by stand-ins (oracle/ref_shim.run_reference_build_kle_mats, oracle/gen_golden_asm.py).
    The oracle's vectorised restatement has the same stored pattern -- explicit zeros included -- and the same values up
    to the round-off of its own element matrices (1e-13)."""
    g = np.load(os.path.join(GOLDEN, "asm_ref.npz"))
    for nx, ny, n in g["meshes"]:
        m = o.BoxMesh([int(nx), int(ny)], [0, 0], [1.0, 0.7], int(n))
        for tag, mats in (("ns", o.assemble_noslip(m)), ("fs", o.assemble_freeslip(m))):
            names = ("K", "Krhs", "Rw", "Rd", "Kfs", "Krhsfs", "Rwfs", "Rdfs") if tag == "ns" else ("K", "Krhs", "Rw")
            for nm in names:
                key = f"{tag}_{nx}x{ny}_{n}_{nm}"
                A = mats[nm]
                np.testing.assert_array_equal(A.indptr, g[key + "_indptr"], err_msg=key)
                np.testing.assert_array_equal(A.indices, g[key + "_indices"], err_msg=key)
                assert rel_err(A.data, g[key + "_data"]) < 1e-13, key


@pytest.mark.parametrize("nx", [8, 16, 30])
def test_cavity_stored_entry_counts_are_the_reference_codes(nx):
    """The stored-entry counts of the cavity at the sizes SURVEY.md 8(c) quotes, as the reference's own loops produce
    them (K 12 228 / 56 900 / 213 924 ...): the survey's K-family figures (10 064 ...) were counts of numerically
    non-zero entries."""
    g = np.load(os.path.join(GOLDEN, "asm_ref.npz"))
    mats = o.assemble_noslip(o.BoxMesh([nx, nx], [0, 0], [1, 1], 3))
    assert [mats[k].nnz for k in ("K", "Kfs", "Krhs", "Krhsfs", "Rw", "Rwfs")] == list(g[f"cavity_counts_{nx}"])


def test_assembly_oracle_vs_live_reference():
    import ref_shim
    if not ref_shim.reference_available():
        pytest.skip("reference tree not present")
    m = o.BoxMesh([4, 5], [0, 0], [2.0, 0.9], 4)
    for noslip, mine in ((True, o.assemble_noslip(m)), (False, o.assemble_freeslip(m))):
        for nm, A in ref_shim.run_reference_build_kle_mats(m, noslip).items():
            np.testing.assert_array_equal(mine[nm].indptr, A.indptr, err_msg=nm)
            np.testing.assert_array_equal(mine[nm].indices, A.indices, err_msg=nm)
            assert rel_err(mine[nm].data, A.data) < 1e-13, nm


@pytest.mark.parametrize("nx,ny,n", [(4, 3, 3), (3, 4, 5)])
def test_operators_rhs_chain_and_solve_kle_vs_reference_golden(nx, ny, n):
    """buildOperators + Operators.assembleAll, the evalRHS chain and the cavity's two-stage solveKLE of the UNMODIFIED
    reference (src/cases/base_problem.py:132-140, 212-252, 321-327; src/matrices/mat_generator.py:157-190;
    src/cases/cavity.py:53-81; src/common/nswalls.py) on stand-ins for PETSc (oracle/ref_shim.py): same operator
    patterns, values / right-hand side / velocities to round-off."""
    g = np.load(os.path.join(GOLDEN, "asm_ref.npz"))
    m = o.BoxMesh([nx, ny], [0, 0], [1.0, 0.7], n)
    ops = o.assemble_operators(m)
    key = f"ops_{nx}x{ny}_{n}"
    for nm in ("SrT", "DivSrT", "Curl"):
        np.testing.assert_array_equal(ops[nm].indptr, g[f"{key}_{nm}_indptr"])
        np.testing.assert_array_equal(ops[nm].indices, g[f"{key}_{nm}_indices"])
        assert rel_err(ops[nm].data, g[f"{key}_{nm}_data"]) < 1e-13
    assert rel_err(o.eval_rhs_from_vel(ops, g[f"{key}_vel"], 0.5, 0.01), g[f"{key}_rhs"]) < 1e-13
    mc = o.BoxMesh([nx, ny], [0, 0], [1.0, 1.0], n)
    vb, fd, fv = o.cavity_wall_vectors(mc)
    vel, velfs, _ = o.solve_kle_noslip(o.assemble_noslip(mc), o.assemble_operators(mc), vb, fd, fv,
                                       g[f"kle_{nx}x{ny}_{n}_vort"])
    assert rel_err(vel, g[f"kle_{nx}x{ny}_{n}_vel"]) < 1e-12 and rel_err(velfs, g[f"kle_{nx}x{ny}_{n}_velfs"]) < 1e-12


@pytest.mark.parametrize("nx,n", [(4, 3), (2, 7)])
def test_taylor_green_solve_kle_vs_reference_golden(nx, n):
    """CustomFuncCase (src/cases/custom_func.py:62-83, 173-193) + FreeSlip.solveKLE (base_problem.py:479-481) of the
    UNMODIFIED reference on stand-ins: exact fields bit for bit, the velocity recovered from the exact vorticity to
    round-off."""
    g = np.load(os.path.join(GOLDEN, "asm_ref.npz"))
    m = o.BoxMesh([nx, nx], [0, 0], [1.0, 1.0], n)
    ve, we = o.taylor_green(m.node_coords(), 0.02, 0.3)
    np.testing.assert_array_equal(ve.reshape(-1), g[f"tg_{nx}_{n}_exact_vel"])
    np.testing.assert_array_equal(we, g[f"tg_{nx}_{n}_exact_vort"])
    mats = o.assemble_freeslip(m)
    vb = np.zeros(2 * m.nnodes)
    vb.reshape(-1, 2)[mats["bc_nodes"]] = ve[mats["bc_nodes"]]
    assert rel_err(o.solve_kle_freeslip(mats, vb, we), g[f"tg_{nx}_{n}_vel"]) < 1e-12


@pytest.mark.parametrize("nx,ny,n", [(2, 2, 2), (3, 2, 3), (4, 3, 5), (2, 5, 8)])
def test_cell_node_numbering_vs_reference_indices_manager(nx, ny, n):
    """The element-local node order of every cell as the UNMODIFIED IndicesManager (src/domain/indices.py:66-118)
    produces it from a DMPlex-style closure and the section of setFemIndexing (oracle/ref_shim.run_reference_cell_nodes;
    the closure convention itself -- PETSc's -- is stated there, pinned by src/tests/test_domain.py for ngl 2)."""
    g = np.load(os.path.join(GOLDEN, "asm_ref.npz"))
    np.testing.assert_array_equal(o.BoxMesh([nx, ny], [0, 0], [1, 1], n).cell_nodes(), g[f"cellnodes_{nx}x{ny}_{n}"])
    from pynama_b200.domain.dmplex import DMPlexDom  # the product's host numbering: the same table
    try:
        dom = DMPlexDom(boxMesh={"nelem": [nx, ny], "lower": [0, 0], "upper": [1, 1]})
        dom.setFemIndexing(n)
        got = np.array([dom.getGlobalNodesFromCell(c, shared=True) for c in range(nx * ny)])
    except Exception as e:              # (a device is needed for some of the domain's tables)
        pytest.skip(f"domain object needs the device: {e}")
    np.testing.assert_array_equal(got, g[f"cellnodes_{nx}x{ny}_{n}"])


def test_noslip_walls_vs_live_reference():
    """The product's NoSlipWalls (pynama_b200/common/nswalls.py) against the reference's own class
    (src/common/nswalls.py), where /root/reference exists: wall names, moving / static walls, wall velocities and
    their tangential dofs, static dofs."""
    import sys
    import ref_shim
    if not ref_shim.reference_available():
        pytest.skip("reference tree not present")
    ref_shim.import_reference_elements()
    ref_shim._install_petsc_stub()
    if ref_shim.REFERENCE_SRC not in sys.path:
        sys.path.insert(0, ref_shim.REFERENCE_SRC)
    from common.nswalls import NoSlipWalls as RefWalls
    from pynama_b200.common.nswalls import NoSlipWalls as MyWalls
    for excl, moving in (([], "up"), (["left"], "right"), (["down", "up"], "right")):
        a, b = RefWalls([0, 0], [1, 2], exclude=excl), MyWalls([0, 0], [1, 2], exclude=excl)
        a.setWallVelocity(moving, [0.5, 2])
        b.setWallVelocity(moving, [0.5, 2])
        assert list(a.getWallsNames()) == list(b.getWallsNames())
        assert a.getWallsWithVelocity() == b.getWallsWithVelocity() and a.getStaticWalls() == b.getStaticWalls()
        for wname in a.getWallsWithVelocity():
            (va, da), (vb, db) = a.getWallVelocity(wname), b.getWallVelocity(wname)
            assert list(np.atleast_1d(va)) == list(np.atleast_1d(vb)) and list(da) == list(db)
        for wname in a.getStaticWalls():
            assert list(a.getStaticDofsByName(wname)) == list(b.getStaticDofsByName(wname))
