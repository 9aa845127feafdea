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
world, rank = int(os.environ.get("WORLD_SIZE", "1")), int(os.environ.get("RANK", "0"))
if world > 1:
    import torch.distributed as dist
    torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
    dist.init_process_group("nccl", device_id=torch.device("cuda", int(os.environ["LOCAL_RANK"])))
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
if rank == 0:
    print(f"its {sv.its} reason {sv.reason} ms/iteration {e0.elapsed_time(e1) / max(sv.its, 1):.3f}")

# per-level cost of one operator application and of the transfers (a V-cycle does 2 * degree of the former per level)
h = sv._ksp._mg


def timed(fn, reps=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


z = torch.empty_like(b.data)
t_v = timed(lambda: h.apply(b.data, z), 10)
if rank == 0:
    print(f"V-cycle alone: {t_v:.3f} ms")
if world > 1:
    from pynama_b200 import backend as be
    from pynama_b200._lib import call
    dc = fem.comm.dist
    big = torch.randn(h.levels[1].n, dtype=torch.float64, device="cuda")
    t_ar = timed(lambda: call("pyn_comm_allreduce_sum", dc._comm, be.ptr(big), big.numel(), be.stream()), 10)
    t_h = timed(lambda: dc.exchange(z, 2), 20)
    if rank == 0:
        print(f"all-reduce of the level-1 residual ({big.numel() * 8 / 1e6:.0f} MB): {t_ar:.3f} ms; halo exchange {t_h * 1e3:.1f} us")
    import sys as _s
    torch.cuda.synchronize(); dist.barrier()
    if rank != 0:
        os._exit(0)
for i, L in enumerate(h.levels):
    xx = torch.randn(L.n, dtype=torch.float64, device="cuda")
    yy = torch.empty_like(xx)
    t_op = timed(lambda: L.csr.mult(xx, yy))
    t_csr = timed(lambda: L.csr.mult(xx, yy, planned=False)) if L.n < 5_000_000 else float("nan")
    line = f"level {i} {h.sequence[i]} n {L.n} operator {'eop' if L.tm is not None else 'csr'} {t_op * 1e3:.1f} us (plan-less CSR {t_csr * 1e3:.1f} us)"
    if i + 1 < len(h.levels):
        xc = torch.randn(h.levels[i + 1].n, dtype=torch.float64, device="cuda")
        yc = torch.empty_like(xc)
        line += f"  P {timed(lambda: L.P.mult(xc, yy)) * 1e3:.1f} us  R {timed(lambda: L.R.mult(xx, yc)) * 1e3:.1f} us"
    print(line)
if world > 1:
    os._exit(0)
