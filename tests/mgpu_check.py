"""Multi-GPU parity check, run under torchrun (one rank per GPU):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 \
        tests/mgpu_check.py

Every rank builds its strip of the cavity (8x9 cells, ngl 3) and of a Taylor-Green case (6x8 cells, ngl 5) through
the drop-in classes, and compares -- in the serial numbering, against the CPU oracle of the WHOLE mesh -- the owned
rows of every matrix (bit-exact pattern per row, values to 1e-12), the velocity of solveKLE (1e-10), the vorticity
right-hand side and three fixed RK steps.  Prints `MGPU_OK <rank>` per rank on success."""
import copy
import os
import sys

import numpy as np
import scipy.sparse as sp
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "oracle")):
    if p not in sys.path:
        sys.path.insert(0, p)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

import kle_oracle as o  # noqa: E402


def rel(a, b):
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


def cfg(case):
    with open(f"{ROOT}/pynama_b200/cases/{case}.yaml") as f:
        return yaml.load(f, Loader=yaml.Loader)


def check_rows(A, ref, part, rdof, cdof, name):
    """owned rows of the local CSR, mapped to global numbering, equal the reference rows (same column set, values)."""
    ip, ix, dv = A.getValuesCSR()
    l2g = part.local_to_global()
    own = part.owned_mask()
    rows_l = np.repeat(np.arange(len(ip) - 1), np.diff(ip))
    gr = l2g[rows_l // rdof] * rdof + rows_l % rdof
    gc = l2g[ix // cdof] * cdof + ix % cdof
    assert np.repeat(own, rdof)[rows_l].all(), f"{name}: entries stored on ghost rows"
    loc = sp.csr_matrix((dv, (gr, gc)), shape=ref.shape)
    # pattern incl. explicit zeros: compare structure through a matrix of ones
    pat = sp.csr_matrix((np.ones_like(dv), (gr, gc)), shape=ref.shape)
    rpat = sp.csr_matrix((np.ones_like(ref.data), ref.indices, ref.indptr), shape=ref.shape)
    owned_rows = (l2g[own][:, None] * rdof + np.arange(rdof)).reshape(-1)
    d = (pat[owned_rows] - rpat[owned_rows])
    assert d.nnz == 0 or np.abs(d.data).max() == 0, f"{name}: pattern of owned rows differs"
    dv_ = (loc[owned_rows] - ref[owned_rows])
    err = (np.abs(dv_.data).max() if dv_.nnz else 0.0) / max(np.abs(ref.data).max(), 1e-300)
    assert err < 1e-12, (name, err)


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from pynama_b200.common.options import Options
    from pynama_b200.cases.cavity import Cavity
    from pynama_b200.cases.custom_func import CustomFuncCase

    # ---- cavity ------------------------------------------------------------------------------------------------
    Options.clear()
    for k, v in dict(ksp_type="cg", pc_type="jacobi", ksp_rtol=1e-13, ksp_max_it=20000, ts_adapt_type="none",
                     ts_dt=0.005).items():
        Options.setValue(k, v)
    nx, ny, n = 8, 9, 3
    c = copy.deepcopy(cfg("cavity"))
    c["domain"]["box-mesh"]["nelem"] = [nx, ny]
    c["material-properties"] = {"rho": 1.0, "mu": 0.01}
    fem = Cavity(c, case="cavity")
    fem.setUp()
    fem.setUpSolver()
    part, dc = fem.dom.partition, fem.comm.dist
    assert part is not None and part.world == world
    m = o.BoxMesh([nx, ny], [0, 0], [1, 1], n)
    mats, ops = o.assemble_noslip(m), o.assemble_operators(m)
    for name in ("K", "Krhs", "Kfs", "Krhsfs", "Rw", "Rwfs", "Rd", "Rdfs"):
        check_rows(getattr(fem.mat, name), mats[name], part, 2, 2 if name.startswith("K") else 1, name)
    for name, rd, cd in (("SrT", 3, 2), ("DivSrT", 2, 3), ("Curl", 1, 2)):
        check_rows(getattr(fem.operator, name), ops[name], part, rd, cd, name)
    xy = fem.dom.fullCoordVec.getArray().reshape(-1, 2)
    assert np.array_equal(xy, m.node_coords()[part.local_to_global()]), "node coordinates differ from the serial mesh"
    vb, fd, fv = o.cavity_wall_vectors(m)
    vort_g = np.random.default_rng(3).standard_normal(m.nnodes)
    vel_o, velfs_o, _ = o.solve_kle_noslip(mats, ops, vb, fd, fv, vort_g)
    w = fem.vort.duplicate()
    w.setArray(vort_g[part.local_to_global()] * part.owned_mask())       # ghosts deliberately stale (zero)
    fem.solveKLE(0.0, w)
    assert fem.solver.reason > 0 and fem.solverFS.reason > 0
    dofs, vals = dc.gather_owned(fem.vel, 2)
    assert rel(vals, vel_o[dofs]) < 1e-10, rel(vals, vel_o[dofs])
    l2g2 = (part.local_to_global()[:, None] * 2 + np.arange(2)).reshape(-1)
    assert rel(fem.vel.getArray(), vel_o[l2g2]) < 1e-10, "ghost entries of the solution are not the owners' values"
    if os.environ.get("MGPU_DUMP"):        # peer-memory path vs NCCL path: compared by tests/test_multigpu.py
        np.save(f"{os.environ['MGPU_DUMP']}.{rank}.npy", np.concatenate([[fem.solver.its, fem.solverFS.its],
                                                                          fem.vel.getArray()]))
    f = fem.vort.duplicate()
    fem.evalRHS(fem.ts, 0.0, w, f)
    f_o = o.eval_rhs_from_vel(ops, vel_o, fem.rho, fem.mu)
    dofs, vals = dc.gather_owned(f, 1)
    assert rel(vals, f_o[dofs]) < 1e-9, rel(vals, f_o[dofs])
    KFS = o.mat_add(mats["K"], mats["Kfs"])

    def rhs(t, ww):
        vel, _, _ = o.solve_kle_noslip(mats, ops, vb, fd, fv, ww, KFS=KFS)
        return o.eval_rhs_from_vel(ops, vel, fem.rho, fem.mu)

    nsteps, dt = 3, 0.005
    _, w_o = o.rk_fixed_steps(rhs, 0.0, np.zeros(m.nnodes), dt, nsteps)
    fem.ts.setTime(0.0)
    fem.ts.step_number = 0
    fem.ts.setMaxSteps(nsteps)
    fem.ts.setTimeStep(dt)
    fem.startSolver()
    dofs, vals = dc.gather_owned(fem.vort, 1)
    assert rel(vals, w_o[dofs]) < 1e-8, rel(vals, w_o[dofs])

    # ---- -pc_type mg on the partitioned mesh: distributed fine level, replicated coarse levels --------------------
    Options.clear()
    for k, v in dict(ksp_type="cg", pc_type="mg", ksp_rtol=1e-12, ksp_max_it=500, pc_mg_coarse_cells=16).items():
        Options.setValue(k, v)
    nxm, nm = 16, 5
    cm = copy.deepcopy(cfg("cavity"))
    cm["domain"]["box-mesh"]["nelem"] = [nxm, nxm]
    cm["domain"]["ngl"] = nm
    cm["material-properties"] = {"rho": 1.0, "mu": 0.01}
    fm_ = Cavity(cm, case="cavity")
    fm_.setUp()
    fm_.setUpSolver()
    partm, dcm = fm_.dom.partition, fm_.comm.dist
    mm = o.BoxMesh([nxm, nxm], [0, 0], [1, 1], nm)
    matsm, opsm = o.assemble_noslip(mm), o.assemble_operators(mm)
    hK = fm_.solver._ksp._mg
    assert hK.sequence == o.mg_coarsening_sequence(nxm, nxm, nm, 16)
    pcK = o.MultigridPC(nxm, nxm, nm, "K", coarse_cells=16, lmax=[L.lmax for L in hK.levels])
    bglob = np.random.default_rng(21).standard_normal(2 * mm.nnodes)
    x_o, its_o, reason_o, _ = o.ksp_cg_jacobi(matsm["K"], bglob, rtol=1e-12, pc=pcK)
    l2g2m = (partm.local_to_global()[:, None] * 2 + np.arange(2)).reshape(-1)
    from pynama_b200 import backend as be
    xl = be.empty(len(l2g2m))
    res = fm_.solver._ksp.solve(be.to_device(bglob[l2g2m]), xl)
    assert res.reason == reason_o == 2 and abs(res.its - its_o) <= 1, (res.its, its_o)
    dofs, vals = dcm.gather_owned(xl, 2)
    assert rel(vals, x_o[dofs]) < 1e-10, rel(vals, x_o[dofs])
    vbm, fdm, fvm = o.cavity_wall_vectors(mm)
    vortm = np.random.default_rng(5).standard_normal(mm.nnodes)
    velm_o, _, _ = o.solve_kle_noslip(matsm, opsm, vbm, fdm, fvm, vortm)
    wm = fm_.vort.duplicate()
    wm.setArray(vortm[partm.local_to_global()] * partm.owned_mask())
    fm_.solveKLE(0.0, wm)
    assert fm_.solver.reason > 0 and fm_.solverFS.reason > 0 and fm_.solver.its < 150 and fm_.solverFS.its < 150
    dofs, vals = dcm.gather_owned(fm_.vel, 2)
    assert rel(vals, velm_o[dofs]) < 1e-10, rel(vals, velm_o[dofs])
    del fm_

    # ---- Taylor-Green, adaptive time step (the error norm is all-reduced) -----------------------------------------
    Options.clear()
    for k, v in dict(ksp_type="cg", pc_type="jacobi", ksp_rtol=1e-12, ksp_max_it=50000, ts_dt=0.01).items():
        Options.setValue(k, v)
    nx, ny, n = 6, 8, 5
    t = copy.deepcopy(cfg("taylor-green"))
    t["time-solver"] = {"max-steps": 4, "start-time": 0, "end-time": 0.05}
    tg = CustomFuncCase(t, case="taylor-green", ngl=n, nelem=[nx, ny])
    tg.setUp()
    tg.setUpSolver()
    part, dc = tg.dom.partition, tg.comm.dist
    m = o.BoxMesh([nx, ny], [0, 0], [1, 1], n)
    fm = o.assemble_freeslip(m)
    for name in ("K", "Krhs", "Rw"):
        check_rows(getattr(tg.mat, name), fm[name], part, 2, 2 if name != "Rw" else 1, "TG " + name)
    exactVel, exactVort = tg.generateExactVecs(0.0)
    tg.solveKLE(0.0, exactVort)
    vel_ex, vort_ex = o.taylor_green(m.node_coords(), tg.nu, 0.0)
    vbc = np.zeros(2 * m.nnodes)
    vbc.reshape(-1, 2)[fm["bc_nodes"]] = vel_ex[fm["bc_nodes"]]
    v_o = o.solve_kle_freeslip(fm, vbc, vort_ex)
    dofs, vals = dc.gather_owned(tg.vel, 2)
    assert rel(vals, v_o[dofs]) < 1e-10, rel(vals, v_o[dofs])
    # -ksp_initial_guess_nonzero on N ranks: r0 = b - A x0 needs the halo exchange of x0 and three all-reduced scalars;
    # started from the converged velocity (owned entries perturbed by 1e-6, ghosts left stale) it needs far fewer
    # iterations and lands on the same solution
    its_zero = tg.solver.its
    Options.setValue("ksp_initial_guess_nonzero", True)
    tg.solver.createSolver(tg.mat.K, tg.comm)
    pert = 1.0 + 1e-6 * np.random.default_rng(17 + rank).standard_normal(tg.vel.getArray().size)
    tg.vel.setArray(tg.vel.getArray() * np.where(np.repeat(part.owned_mask(), 2) > 0, pert, 0.0))
    tg.solveKLE(0.0, exactVort)
    assert tg.solver.reason > 0 and 0 < tg.solver.its < its_zero, (tg.solver.its, its_zero)
    dofs, vals = dc.gather_owned(tg.vel, 2)
    assert rel(vals, v_o[dofs]) < 1e-10, rel(vals, v_o[dofs])
    l2g2 = (part.local_to_global()[:, None] * 2 + np.arange(2)).reshape(-1)
    assert rel(tg.vel.getArray(), v_o[l2g2]) < 1e-10, "ghost entries after a nonzero-guess solve"
    Options.setValue("ksp_initial_guess_nonzero", False)
    tg.solver.createSolver(tg.mat.K, tg.comm)
    tg.startSolver()
    steps = torch.tensor([tg.ts.step_number, tg.ts.rejections], device="cuda")
    gathered = [torch.zeros_like(steps) for _ in range(world)]
    dist.all_gather(gathered, steps)
    assert all(torch.equal(g, steps) for g in gathered), "ranks disagree on accepted/rejected steps"
    assert tg.ts.step_number >= 1
    dofs, vals = dc.gather_owned(tg.vort, 1)
    _, w_ex = o.taylor_green(m.node_coords(), tg.nu, tg.ts.time)
    assert rel(vals, w_ex[dofs]) < 0.05
    torch.cuda.synchronize()
    dist.barrier()
    print(f"MGPU_OK {rank} (halo bytes/exchange vel: {dc.halo_bytes(2)}, peer_memory={dc.peer_memory})", flush=True)
    sys.stdout.flush()
    # leave without running destructors: NCCL communicators captured in CUDA graphs are not torn down rank by rank
    os._exit(0)


if __name__ == "__main__":
    main()
