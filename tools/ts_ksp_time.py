"""KSP solve time per time step of the lid-driven cavity (BASELINE metric 2: "FP64 KSP solve time per step") through the
public classes: Cavity.setUp / setUpSolver / startSolver with -ksp_type cg -pc_type jacobi at a fixed rtol, fixed dt,
a few RK 5(4) steps (7 right-hand sides = 14 solves per step: free-slip stage + no-slip stage, base_problem.py:321-327).
Run once from a zero initial guess (PETSc's default) and once with -ksp_initial_guess_nonzero (the previous stage's
velocity).  One JSON line per run.

    python tools/ts_ksp_time.py --nx 256 --ngl 5 --steps 3 --rtol 1e-5
"""
import argparse
import copy
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def run(a, guess):
    import torch
    import yaml
    from pynama_b200.common.options import Options
    from pynama_b200.cases.cavity import Cavity
    Options.clear()
    for k, v in dict(ksp_type="cg", pc_type="jacobi", ksp_rtol=a.rtol, ksp_max_it=200000, ts_adapt_type="none",
                     ts_dt=a.dt, ksp_check_every=50).items():
        Options.setValue(k, v)
    if guess:
        Options.setValue("ksp_initial_guess_nonzero", True)
    with open(f"{ROOT}/pynama_b200/cases/cavity.yaml") as f:
        cfg = yaml.load(f, Loader=yaml.Loader)
    cfg = copy.deepcopy(cfg)
    cfg["domain"]["box-mesh"]["nelem"] = [a.nx, a.nx]
    cfg["domain"]["ngl"] = a.ngl
    cfg["material-properties"] = {"rho": 1.0, "mu": 0.01}
    fem = Cavity(cfg, case="cavity")
    fem.setUp()
    fem.setUpSolver()
    its, solves, t_ksp = [0], [0], [0.0]
    for s in (fem.solver, fem.solverFS):
        inner = s._ksp.solve                 # be.Ksp.solve(b, x) -> result with .its

        def timed(b, x, _inner=inner):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            r = _inner(b, x)
            torch.cuda.synchronize()
            t_ksp[0] += time.perf_counter() - t0
            its[0] += r.its
            solves[0] += 1
            return r
        s._ksp.solve = timed
    fem.ts.setTime(0.0)
    fem.ts.step_number = 0
    fem.ts.setMaxSteps(a.steps)
    fem.ts.setTimeStep(a.dt)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    fem.startSolver()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    steps = max(fem.ts.step_number, 1)
    out = {"case": f"cavity {a.nx}x{a.nx} ngl {a.ngl}", "vel_dofs": int(fem.vel.getSize()), "rtol": a.rtol, "dt": a.dt,
           "initial_guess_nonzero": bool(guess), "steps": steps, "solves": solves[0], "iterations": its[0],
           "iterations_per_solve": its[0] / max(solves[0], 1), "ksp_ms_per_step": 1e3 * t_ksp[0] / steps,
           "ms_per_step": 1e3 * wall / steps, "ksp_ms_per_iteration": 1e3 * t_ksp[0] / max(its[0], 1),
           "operator": "element" if fem.solver._ksp._tm is not None else "csr-stream",
           "vort_max": float(fem.vort.data.abs().max())}
    print(json.dumps(out), flush=True)
    del fem
    torch.cuda.empty_cache()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--nx", type=int, default=256)
    ap.add_argument("--ngl", type=int, default=5)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--dt", type=float, default=1e-3)
    ap.add_argument("--rtol", type=float, default=1e-5)
    ap.add_argument("--only", choices=["zero", "guess", "both"], default="both")
    a = ap.parse_args()
    if a.only in ("zero", "both"):
        run(a, False)
    if a.only in ("guess", "both"):
        run(a, True)


if __name__ == "__main__":
    main()
