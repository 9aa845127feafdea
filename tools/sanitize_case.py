"""Small end-to-end pass over every kernel family, for `compute-sanitizer --tool memcheck|racecheck` (one tool per run):

    compute-sanitizer --tool memcheck python tools/sanitize_case.py

Cavity 24x24, ngl 5: first assembly, values-only re-assembly with row templates on and off, CG through the element
operator, CG on the CSR stream (register-staged and bulk-copy kernels), GMRES, CG + multigrid, the RHS chain; Taylor-Green
12x12, ngl 4 on the per-cell path, fused and materialised."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import yaml  # noqa: E402

from bench import cavity_config  # noqa: E402
from pynama_b200 import backend as be  # noqa: E402
from pynama_b200.cases.cavity import Cavity  # noqa: E402
from pynama_b200.cases.custom_func import CustomFuncCase  # noqa: E402
from pynama_b200.common.options import Options  # noqa: E402


def cavity(**opts):
    Options.clear()
    for k, v in opts.items():
        Options.setValue(k, v)
    fem = Cavity(cavity_config(24, 5), case="cavity")
    fem.setUp()
    fem.setUpSolver()
    return fem


def solve(fem):
    g = torch.Generator(device="cuda").manual_seed(3)
    fem.vort.data.copy_(torch.randn(fem.vort.data.numel(), dtype=torch.float64, device="cuda", generator=g))
    fem.solveKLE(0.0, fem.vort)
    f = fem.vort.duplicate()
    fem.evalRHS(fem.ts, 0.0, fem.vort, f)
    torch.cuda.synchronize()
    return fem.solver.its, fem.solverFS.its, float(f.data.abs().max())


fem = cavity(ksp_type="cg", pc_type="jacobi", ksp_rtol=1e-8, ksp_max_it=5000)
for on in (True, False):
    be.set_templates(on)
    fem.mat.assembleFromMesh(fem.dom, fem.elemType)
be.set_templates(True)
print("cg + element operator", solve(fem))
fem = cavity(ksp_type="cg", pc_type="jacobi", ksp_rtol=1e-8, ksp_max_it=5000, mat_templates=0)
print("cg + CSR stream (bulk copy)", solve(fem))
fem = cavity(ksp_type="gmres", pc_type="jacobi", ksp_rtol=1e-6, ksp_max_it=3000)
print("gmres", solve(fem))
fem = cavity(ksp_type="cg", pc_type="mg", ksp_rtol=1e-8, pc_mg_coarse_cells=36)
print("cg + multigrid", solve(fem))
for fused in (1, 0):
    Options.clear()
    for k, v in dict(ksp_type="cg", pc_type="jacobi", ksp_rtol=1e-8, ksp_max_it=5000, mat_fused=fused).items():
        Options.setValue(k, v)
    with open(f"{ROOT}/pynama_b200/cases/taylor-green.yaml") as f:
        cfg = yaml.load(f, Loader=yaml.Loader)
    tg = CustomFuncCase(cfg, case="taylor-green", ngl=4, nelem=[12, 12])
    tg.setUp()
    tg.setUpSolver()
    tg.mat.assembleFromMesh(tg.dom, tg.elemType, perCell=True)      # values-only re-assembly
    tg.computeInitialCondition(0.0)
    tg.solveKLE(0.0, tg.vort)
    torch.cuda.synchronize()
    print("per-cell, fused" if fused else "per-cell, materialised", tg.mat.fusedAssembly, tg.solver.its)
print("SANITIZE_CASE_OK")
