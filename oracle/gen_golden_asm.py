"""TEST INFRASTRUCTURE ONLY -- generates tests/golden/asm_ref.npz by running the UNMODIFIED reference assembly loops
NoSlipFreeSlip.buildKLEMats (src/cases/base_problem.py:329-454) and FreeSlip.buildKLEMats (:499-552) on stand-ins
for PETSc's AIJ matrices and for the mesh queries (oracle/ref_shim.run_reference_build_kle_mats): the CSR arrays of all
the matrices on three small box meshes, and the STORED-entry counts of the lid-driven cavity at 8x8, 16x16 and 30x30
elements, ngl 3 (the sizes SURVEY.md 8(c) quotes); the operators of buildOperators / Operators.assembleAll
(base_problem.py:132-140, mat_generator.py:157-190), the evalRHS chain (:212-252) on a random velocity and the cavity's
two-stage solveKLE (:321-327, cavity.py:53-81) on a random vorticity.  Runs only in the build container.

    python oracle/gen_golden_asm.py
"""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
import kle_oracle as o  # noqa: E402
from ref_shim import (run_reference_build_kle_mats, run_reference_cavity_solve_kle,  # noqa: E402
                      run_reference_cell_nodes, run_reference_operators_and_rhs,
                      run_reference_taylor_green_solve_kle)

OUT = os.path.join(HERE, "..", "tests", "golden", "asm_ref.npz")
MESHES = ((4, 3, 3), (5, 5, 2), (3, 4, 5))         # (nx, ny, ngl) on [0, 1] x [0, 0.7]


def main():
    out = {"meshes": np.array(MESHES)}
    for nx, ny, n in MESHES:
        m = o.BoxMesh([nx, ny], [0, 0], [1.0, 0.7], n)
        for tag, noslip in (("ns", True), ("fs", False)):
            for nm, A in run_reference_build_kle_mats(m, noslip).items():
                key = f"{tag}_{nx}x{ny}_{n}_{nm}"
                out[key + "_indptr"], out[key + "_indices"], out[key + "_data"] = A.indptr, A.indices, A.data
    for nx in (8, 16, 30):
        m = o.BoxMesh([nx, nx], [0, 0], [1, 1], 3)
        mats = run_reference_build_kle_mats(m, True)
        out[f"cavity_counts_{nx}"] = np.array([mats[k].nnz for k in ("K", "Kfs", "Krhs", "Krhsfs", "Rw", "Rwfs")])
        print(nx, out[f"cavity_counts_{nx}"])
    for nx, ny, n in ((4, 3, 3), (3, 4, 5)):
        m = o.BoxMesh([nx, ny], [0, 0], [1.0, 0.7], n)
        vel = np.random.default_rng(1).standard_normal(2 * m.nnodes)
        ops, f = run_reference_operators_and_rhs(m, vel, 0.5, 0.01)
        key = f"ops_{nx}x{ny}_{n}"
        for nm, A in ops.items():
            out[f"{key}_{nm}_indptr"], out[f"{key}_{nm}_indices"], out[f"{key}_{nm}_data"] = A.indptr, A.indices, A.data
        out[f"{key}_vel"], out[f"{key}_rhs"] = vel, f
        mc = o.BoxMesh([nx, ny], [0, 0], [1.0, 1.0], n)
        vort = np.random.default_rng(3).standard_normal(mc.nnodes)
        v, vfs = run_reference_cavity_solve_kle(mc, vort)
        out[f"kle_{nx}x{ny}_{n}_vort"], out[f"kle_{nx}x{ny}_{n}_vel"], out[f"kle_{nx}x{ny}_{n}_velfs"] = vort, v, vfs
    for nx, n in ((4, 3), (2, 7)):               # Taylor-Green (configs[1]'s case): FreeSlip.solveKLE from the exact vorticity
        m = o.BoxMesh([nx, nx], [0, 0], [1.0, 1.0], n)
        v, ve, we = run_reference_taylor_green_solve_kle(m, 0.02, 0.3)
        out[f"tg_{nx}_{n}_vel"], out[f"tg_{nx}_{n}_exact_vel"], out[f"tg_{nx}_{n}_exact_vort"] = v, ve, we
    for nx, ny, n in ((2, 2, 2), (3, 2, 3), (4, 3, 5), (2, 5, 8)):     # IndicesManager over DMPlex-style closures
        out[f"cellnodes_{nx}x{ny}_{n}"] = run_reference_cell_nodes(o.BoxMesh([nx, ny], [0, 0], [1, 1], n))
    np.savez_compressed(OUT, **out)
    print("wrote", OUT)


if __name__ == "__main__":
    main()
