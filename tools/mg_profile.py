"""One MG-preconditioned CG solve of the cavity K (for `ncu --metrics gpu__time_duration.sum` launch lists).

    python tools/mg_profile.py --nx 1024 --its 3
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from bench import cavity_config  # noqa: E402
from pynama_b200.cases.cavity import Cavity  # noqa: E402
from pynama_b200.common.options import Options  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--nx", type=int, default=1024)
ap.add_argument("--ngl", type=int, default=5)
ap.add_argument("--its", type=int, default=3)
ap.add_argument("--degree", type=int, default=3)
a = ap.parse_args()
Options.clear()
for k, v in dict(ksp_type="cg", pc_type="mg", ksp_max_it=a.its, ksp_check_every=a.its, pc_mg_degree=a.degree).items():
    Options.setValue(k, v)
fem = Cavity(cavity_config(a.nx, a.ngl), case="cavity")
fem.setUpGeneral()
fem.setUpBoundaryConditions()
fem.setUpEmptyMats()
fem.buildKLEMats()
from pynama_b200.solver.ksp_solver import KspSolver  # noqa: E402
sv = KspSolver()
sv.createSolver(fem.mat.K, fem.comm)
sv.setUpMultigrid(fem, "K")
b = fem.mat.K.createVecRight()
x = b.duplicate()
b.data.copy_(torch.randn(b.data.numel(), dtype=torch.float64, device="cuda",
                         generator=torch.Generator(device="cuda").manual_seed(1)))
sv.solve(b, x)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
sv.solve(b, x)
e1.record()
torch.cuda.synchronize()
print(f"its {sv.its} reason {sv.reason} ms/iteration {e0.elapsed_time(e1) / max(sv.its, 1):.3f}")
