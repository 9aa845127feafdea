#!/usr/bin/env python
"""bench.py -- the hot path of BASELINE.json on B200: cavity nx x nx spectral elements, ngl 5 (configs[3]).

One "step" = one assembly pass of the matrices the velocity-from-vorticity solve uses (element integration from the
corner coordinates + deterministic CSR assembly of the eight MatNS matrices K, Krhs, Kfs, Krhsfs, Rw, Rwfs, Rd, Rdfs on the prebuilt pattern;
NoSlipFreeSlip.buildKLEMats, reference src/cases/base_problem.py:329-454).  `value` = velocity DOF rows assembled per
second with every input resident in HBM.  The second half of the metric ("FP64 KSP solve time/step") is the `ksp`
object of the same JSON line: CG + Jacobi (PETSc semantics) on the assembled K, timed over a fixed iteration budget
with CUDA events.  `e2e` is the same assembly step driven through the public Python API from HOST buffers (pinned
host arrays -> device, result diagonal -> host).  `cpu_baseline` / `--impl reference` time the CPU oracle (a
NumPy/SciPy port of the reference path; petsc4py is not installable here) on a bounded sample of the same workload.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--nx 1024] [--ngl 5] [--ksp-its 50] [--impl reference]
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "assembled DOF/s and FP64 KSP solve time/step"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--nx", type=int, default=1024)
    ap.add_argument("--ngl", type=int, default=5)
    ap.add_argument("--ksp-its", type=int, default=50, dest="ksp_its")
    ap.add_argument("--ksp-converge", type=float, default=0.0, dest="ksp_converge", metavar="RTOL",
                    help="also run one CG+Jacobi solve to this rtol (PETSc default 1e-5) and report its / seconds")
    ap.add_argument("--impl", default="native", choices=["native", "reference"])
    ap.add_argument("--cpu-sample-nx", type=int, default=192, dest="cpu_nx")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-time-loop", action="store_true", dest="no_time_loop")
    ap.add_argument("--mg", default="", help="multigrid tuning for the time-loop solves: degree[,ratio] of the "
                    "Chebyshev smoother (-pc_mg_degree, -pc_mg_ratio; default 3,10)")
    ap.add_argument("--no-percell", action="store_true", dest="no_percell",
                    help="skip the per-cell record (configs[2] at the headline order) of the default single-GPU line")
    ap.add_argument("--no-configs4", action="store_true", dest="no_configs4",
                    help="N >= 4: skip the second record on BASELINE configs[4] (4096^2, ngl 4)")
    ap.add_argument("--workload", default="cavity", choices=["cavity", "percell"],
                    help="cavity: configs[3] (default); percell: FreeSlip, element matrices of every cell (configs[1]/[2])")
    ap.add_argument("--fused", default=None, help="percell: -mat_fused 0|1 (fused integrate+assemble kernel)")
    ap.add_argument("--pc", default=None, help="-pc_type of the converged solves in the time-loop phase "
                                               "(default: mg)")
    return ap.parse_args()


# ------------------------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    """nvidia-smi clocks / throttle reasons sampled every 200 ms while the timed region runs."""
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index=0):
        super().__init__(daemon=True)
        self.index, self.rows, self.proc = index, [], None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                          "--format=csv,noheader,nounits", "-lms", "200"],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc is not None:
            self.proc.terminate()
        self.join(timeout=2)
        sm, mx, reasons = [], 0.0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                for nm, v in zip(names, r[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(nm)
            except (ValueError, IndexError):
                continue
        sm.sort()
        # median over the upper half = samples under load (the sampler also sees idle gaps between phases)
        load = sm[len(sm) // 2:] if sm else []
        return {"sm_mhz": load[len(load) // 2] if load else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


def measured_peak_hbm():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ------------------------------------------------------------------------------------------------------------------
# CPU arm: the oracle port of the reference path on host cores (test infrastructure used ONLY as the timed baseline)
# ------------------------------------------------------------------------------------------------------------------
def _cpu_strip(args):
    """Assemble the cavity matrices of one strip of cell rows with the oracle (worker of the process pool)."""
    nx, ny_strip, ngl = args
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import kle_oracle as o
    m = o.BoxMesh([nx, ny_strip], [0.0, 0.0], [1.0, ny_strip / nx], ngl)
    t0 = time.perf_counter()
    r = o.assemble_noslip(m)
    return time.perf_counter() - t0, 2 * m.nnodes, int(r["K"].nnz)


def cpu_reference_step(nx, ngl, cores):
    """One bounded sample of the workload on the host: `cores` independent strips (nx x nx/cores cells each) of the
    cavity assembly, one process per core; returns (seconds, dof rows, nnz)."""
    from concurrent.futures import ProcessPoolExecutor
    strip = max(1, nx // cores)
    t0 = time.perf_counter()
    if cores == 1:
        res = [_cpu_strip((nx, strip, ngl))]
    else:
        with ProcessPoolExecutor(cores) as ex:
            res = list(ex.map(_cpu_strip, [(nx, strip, ngl)] * cores))
    wall = time.perf_counter() - t0
    # the strips run concurrently: the step takes as long as the slowest worker's assembly (process start-up and the
    # NumPy/SciPy imports of the pool are not part of the workload); fall back to the wall clock if that is shorter
    busy = min(wall, max(r[0] for r in res))
    return busy, sum(r[1] for r in res), sum(r[2] for r in res)


def cpu_ksp_iteration(nx, ngl, its=20):
    """ms per CG+Jacobi iteration of the oracle (SciPy SpMV, 1 core) on an nx x nx sample."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import numpy as np
    import kle_oracle as o
    m = o.BoxMesh([nx, nx], [0, 0], [1, 1], ngl)
    K = o.assemble_noslip(m)["K"]
    b = np.random.default_rng(0).standard_normal(K.shape[0])
    t0 = time.perf_counter()
    o.ksp_cg_jacobi(K, b, rtol=1e-300, max_it=its)
    return (time.perf_counter() - t0) / its * 1e3, K.shape[0], int(K.nnz)


def run_reference(a):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    nx = a.cpu_nx
    times, dof = [], 0
    for i in range(a.warmup + a.steps):
        wall, dof, _ = cpu_reference_step(nx, a.ngl, cores)
        if i >= a.warmup:
            times.append(wall)
    t = sum(times) / len(times)
    ksp_ms, krows, knnz = cpu_ksp_iteration(min(nx, 64), a.ngl)
    val = dof / t
    sample = (f"{cores} strips of {nx}x{max(1, nx // cores)} cells, ngl {a.ngl}, oracle/kle_oracle.py "
              f"assemble_noslip (8 matrices incl. Rd/Rdfs), one process per core")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": "DOF/s", "n_gpus": a.gpus, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": t * 1e3, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"cavity {a.nx}x{a.nx} elements ngl {a.ngl} (configs[3]); CPU arm times a bounded "
                               f"sample: {sample}"},
        "cpu_baseline": {"value": val, "unit": "DOF/s", "cores": cores, "kind": "port", "sample": sample},
        "ksp": {"ms_per_iteration": ksp_ms, "rows": krows, "nnz": knnz, "cores": 1,
                "note": "oracle CG+Jacobi (SciPy CSR SpMV) on a 64x64-cell sample"},
        "e2e": {"value": val, "unit": "DOF/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }), flush=True)


# ------------------------------------------------------------------------------------------------------------------
# native arm
# ------------------------------------------------------------------------------------------------------------------
def cavity_config(nx, ngl):
    """cases/cavity.yaml with the mesh of BASELINE configs[3] and Re = 100 (rho 1, mu 0.01, lid `up: [1, 0]`)."""
    return {"name": "Cavity", "material-properties": {"rho": 1.0, "mu": 0.01},
            "domain": {"ngl": ngl, "box-mesh": {"nelem": [nx, nx], "lower": [0, 0], "upper": [1, 1]}},
            "time-solver": {"max-steps": 1, "start-time": 0, "end-time": 1},
            "boundary-conditions": {"no-slip": {"up": [1, 0]}}}


class CavityAssembly:
    """The cavity through the drop-in classes (pynama_b200.cases.cavity.Cavity): on N ranks every rank holds its strip
    of cell rows (domain/partition.py) and assembles the rows of the nodes it owns.  One step = element integration +
    values of the eight MatNS matrices on the prebuilt pattern (NoSlipFreeSlip.buildKLEMats)."""
    NAMES = ("K", "Krhs", "Kfs", "Krhsfs", "Rw", "Rwfs", "Rd", "Rdfs")

    def __init__(self, nx, ngl, **extra_options):
        from pynama_b200 import backend as be
        from pynama_b200.cases.cavity import Cavity
        from pynama_b200.common.options import Options
        Options.clear()
        for k, v in (("ksp_type", "cg"), ("pc_type", "jacobi")):
            Options.setValue(k, v)
        import torch
        self.be = be
        for k, v in extra_options.items():
            Options.setValue(k, v)

        def lap(t0=[None]):
            torch.cuda.synchronize()
            t1 = time.perf_counter()
            dt, t0[0] = (None if t0[0] is None else (t1 - t0[0]) * 1e3), t1
            return dt

        lap()
        self.fem = Cavity(cavity_config(nx, ngl), case="cavity")
        fem = self.fem
        fem.setUpGeneral()
        fem.setUpBoundaryConditions()
        fem.setUpEmptyMats()
        fem.dom.topology()
        t_topo = lap()
        fem.buildKLEMats()                      # pattern build (count + index fill) + first assembly of the values
        t_first = lap()
        self.dom, self.mat = fem.dom, fem.mat
        # the same values through the generic (non-template) kernels on the now existing pattern: what the first
        # assembly spends on values; the difference to `t_first` is the pattern build (counts, scans, index fill)
        be.set_templates(False)
        self.step()
        lap()
        self.step()
        t_generic = lap()
        be.set_templates(True)
        self.setup = {"topology_ms": t_topo, "pattern_build_ms": max(t_first - t_generic, 0.0),
                      "first_assembly_ms": t_first, "generic_values_ms": t_generic,
                      "note": "once per mesh, outside the timed step: topology = mesh numbering, node->element "
                              "incidence, adjacency, inverse maps, row signatures; first_assembly = "
                              "NoSlipFreeSlip.buildKLEMats from nothing (pattern counts + index fill + values of the "
                              "8 matrices, generic gather kernels); generic_values = the values alone through the "
                              "generic kernels; pattern_build = the difference.  Host wall clock around "
                              "synchronised launches."}
        own = fem.dom.ownedNodeMask()
        self.dof_owned = 2 * int(own.sum())
        self.nrows = 2 * fem.dom.getTotalNodes()

    @property
    def mats(self):
        return {n: getattr(self.mat, n).csr for n in self.NAMES}

    def step(self, events=None, classes=None, row_classes=None, corners=None):
        m, fem = self.mat, self.fem
        if events is not None:
            events[0].record()
        # one element integration, then the eight matrices; events[1] closes the K launches (template + stream on this
        # stream; K's boundary list runs beside them on the side stream)
        m.assembleFromMesh(fem.dom, fem.elemType, names=self.NAMES, classes=classes if classes is not None else m.dofClasses,
                           rowClassesDev=row_classes, corners=corners,
                           after={"K": events[1].record} if events is not None else None)
        return self.mats


def run_native(a):
    import torch
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (B200); there is no CPU fallback. Use --impl reference for "
                         "the CPU arm.")
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if a.pc is None:
        a.pc = "mg"
    from pynama_b200 import backend as be
    ev = lambda: torch.cuda.Event(enable_timing=True)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    nx, n = a.nx, a.ngl
    # strong scaling: the fixed nx x nx mesh is cut into `world` strips of cell rows, one per GPU; a rank assembles the
    # rows of the nodes it owns (complete thanks to one ghost row of cells: no matrix entry crosses NVLink)
    mg_opts = {}
    if a.mg:
        parts = a.mg.split(",")
        mg_opts["pc_mg_degree"] = int(parts[0])
        if len(parts) > 1:
            mg_opts["pc_mg_ratio"] = float(parts[1])
    prob = CavityAssembly(nx, n, pc_type=a.pc, ksp_max_it=200000, **mg_opts,
                          **({"ksp_check_every": 50} if a.pc == "jacobi" else {}))
    prob.step()
    barrier()
    nnz = {k: v.nnz for k, v in prob.mats.items()}
    dof_local = prob.dof_owned
    for _ in range(a.warmup):
        prob.step()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
        time.sleep(0.25)
    l0 = be.launch_count()
    kev = [(ev(), ev()) for _ in range(a.steps)]
    t0, t1 = ev(), ev()
    barrier()
    t0.record()
    for i in range(a.steps):
        prob.step(events=kev[i])
    t1.record()
    barrier()
    launches = be.launch_count() - l0
    ms_total = t0.elapsed_time(t1)
    k_ms = sum(s.elapsed_time(e) for s, e in kev) / a.steps
    tt = torch.tensor([ms_total, float(dof_local)], dtype=torch.float64, device="cuda")
    if world > 1:
        tmax = tt.clone()
        dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
        dist.all_reduce(tt, op=dist.ReduceOp.SUM)
        ms_total, dof_total = float(tmax[0]), float(tt[1])
    else:
        dof_total = float(dof_local)
    ms_step = ms_total / a.steps
    value = dof_total / (ms_step * 1e-3)

    peak, peak_src = measured_peak_hbm()
    K = prob.mats["K"]
    # algorithmic bytes of the dominant launch (assembly of K, values only): 8 nnz written + int64 row pointers read
    # + 64 B of corner coordinates per cell; DESIGN.md "Roofline accounting"
    k_bytes = 8 * K.nnz + 8 * (K.shape[0] + 1) + 64 * prob.dom.ncell
    roof = {"bound": "hbm", "kernel": "asm_stream_kernel<2,2> (+ asm_template_kernel, boundary list): K values on the prebuilt pattern",
            "achieved": k_bytes / (k_ms * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "peak_source": peak_src,
            "frac": k_bytes / (k_ms * 1e-3) / 1e9 / peak, "ms_per_launch": k_ms, "bytes_per_launch": k_bytes,
            "traffic": load_traffic("asm_stream_kernel<2, 2>") if world == 1 else None}
    out = {
        "metric": METRIC, "value": value, "unit": "DOF/s", "n_gpus": world, "steps": a.steps, "warmup": a.warmup,
        "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
        "data": "synthetic",
        "config": {"workload": f"cavity {nx}x{nx} elements ngl {n} (BASELINE configs[3]): integrate + assemble "
                               f"K,Krhs,Kfs,Krhsfs,Rw,Rwfs,Rd,Rdfs values on the prebuilt CSR pattern",
                   "partition": f"{world} strip(s) of cell rows of the one {nx}x{nx} mesh, one per GPU; assembly needs no "
                                f"collective, CG exchanges interface dofs + 3 scalars per iteration over NCCL",
                   "dof_rows": int(dof_total), "nnz_rank0": nnz,
                   "l2": "outputs (>= 8*nnz(K) bytes per step) exceed the 126 MB L2 for nx >= 128; no flush needed"},
        "roofline": roof, "gpu_launches": int(launches),
    }
    # ---- KSP phase: CG + Jacobi on K, fixed iteration budget ---------------------------------------------------
    out["ksp"] = ksp_phase(a, prob, be, torch, peak, barrier)
    if world > 1:
        t = torch.tensor([out["ksp"]["ms_per_solve"]], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        out["ksp"]["ms_per_solve"] = float(t[0])
        out["ksp"]["ms_per_iteration"] = float(t[0]) / max(out["ksp"]["its"], 1)
    barrier()
    if rank == 0:
        out["clocks"] = sampler.stop()
    out["setup"] = prob.setup
    out["parity_checks"] = parity_checks(prob, be, torch, world, nx, n, dof_total)
    if not a.no_time_loop:
        tl = time_loop_phase(a, prob, be, torch, peak, barrier, world)
        out["time_loop"] = tl
        out["rhs_chain"] = tl["rhs_chain"]
        out["ksp"]["converged"] = dict(tl["solve_kle"]["no_slip_solve"], rtol=tl["rtol"], pc=tl["pc"],
                                       seconds_per_time_step=tl["seconds_per_time_step"],
                                       note="the K solve of Cavity.solveKLE (second stage); a time step is 14 such "
                                            "solves + 7 RHS chains (time_loop)")
    # ---- end to end through the public API from host buffers (rank-local; max over ranks) ----------------------
    if not a.no_e2e:
        e = e2e_phase(a, prob, be, torch, barrier)
        if world > 1:
            t = torch.tensor([e["ms"]], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            e["ms"] = float(t[0])
        out["e2e"] = {"value": dof_total / (e["ms"] * 1e-3), "unit": "DOF/s", "ms_per_step": e["ms"],
                      "h2d_bytes_per_step": e["h2d"] * world, "d2h_bytes_per_step": e["d2h"] * world,
                      "what": e["what"]}
    if rank == 0 and world == 1 and not a.no_cpu_baseline:
        cores = os.cpu_count() or 1
        wall, dof, _ = cpu_reference_step(a.cpu_nx, n, cores)
        out["cpu_baseline"] = {"value": dof / wall, "unit": "DOF/s", "cores": cores, "kind": "port",
                               "sample": f"{cores} strips of {a.cpu_nx}x{max(1, a.cpu_nx // cores)} cells ngl {n}, "
                                         f"oracle assemble_noslip, one process per core, {wall:.1f} s"}
    if world >= 4 and a.nx == 1024 and not a.no_configs4:
        # the largest mesh of BASELINE.json rides along as a second record inside the one JSON line
        del prob
        torch.cuda.empty_cache()
        try:
            out["configs4"] = largest_mesh_record(a, be, torch, world, barrier, peak)
        except Exception as e:      # e.g. not enough memory on this box: say so, keep the configs[3] line
            out["configs4"] = {"unavailable": f"{type(e).__name__}: {e}"[:300]}
    if world == 1 and a.nx == 1024 and not a.no_percell:
        # the per-cell path (BASELINE configs[1]/[2]) as a second record inside the one JSON line
        prob = None
        torch.cuda.empty_cache()
        try:
            out["percell"] = percell_record(a, torch)
        except Exception as e:
            out["percell"] = {"unavailable": f"{type(e).__name__}: {e}"[:300]}
    if rank == 0:
        print(json.dumps(out), flush=True)
    if world > 1:
        barrier()
        sys.stdout.flush()
        os._exit(0)      # skip destructors: NCCL communicators captured in CUDA graphs are not torn down rank by rank


def time_loop_phase(a, prob, be, torch, peak, barrier, world):
    """The per-time-step half of the metric through the case class (Cavity): operators, the two-stage solveKLE
    (free-slip solve, wall vorticity, no-slip solve: src/cases/base_problem.py:321-327) to the KSP's own tolerance,
    the RHS chain Curl (1/rho) DivSrT (2 mu SrT v - rho v(x)v) (base_problem.py:212-232), and the same solve driven from
    HOST buffers (vorticity in, velocity out).  One RK 5(4) step = 7 evalRHS = 14 solves + 7 chains."""
    import torch.distributed as dist
    fem = prob.fem
    ev = lambda: torch.cuda.Event(enable_timing=True)
    N = prob.dom.getTotalNodes()
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    fem.buildOperators()
    fem.collectCornerNodes()
    fem._bcDevice = None
    fem.setUpSolver()
    torch.cuda.synchronize()
    t_ops = (time.perf_counter() - t0) * 1e3
    dc = fem.comm.dist
    g = torch.Generator(device="cuda").manual_seed(1234)
    # a smooth synthetic vorticity field of the cavity's magnitude (the time loop solves for fields like this one, not
    # for white noise): omega = 4 sin(pi x) sin(pi y) + lid shear layer
    xy = fem.dom.fullCoordVec.data.view(-1, 2)
    w = 4.0 * torch.sin(torch.pi * xy[:, 0]) * torch.sin(torch.pi * xy[:, 1]) + 20.0 * torch.exp(-40.0 * (1.0 - xy[:, 1]))
    fem.vort.data.copy_(w)
    recs = []
    for sv in (fem.solverFS, fem.solver):
        inner = sv._ksp.solve

        def timed(b, x, _inner=inner, _sv=sv):
            s, e = ev(), ev()
            s.record()
            r = _inner(b, x)
            e.record()
            torch.cuda.synchronize()
            recs.append({"its": int(r.its), "reason": int(r.reason), "seconds": s.elapsed_time(e) * 1e-3})
            return r
        sv._ksp.solve = timed

    def maxr(x):
        if world > 1:
            t = torch.tensor([x], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t[0])
        return x

    # ---- solveKLE: device resident ----
    if a.pc != "jacobi":
        fem.solveKLE(0.0, fem.vort)           # warm-up (graph capture, hierarchy set-up) when a solve is cheap
    del recs[:]
    barrier()
    s, e = ev(), ev()
    s.record()
    fem.solveKLE(0.0, fem.vort)
    e.record()
    barrier()
    t_kle = maxr(s.elapsed_time(e) * 1e-3)
    fs, ns = recs[0], recs[1]
    # ---- RHS chain: the three operator products of evalRHS with their fused epilogues ----
    op = fem.operator.spmv

    def chain():
        op["SrT"].csr.mult(fem.vel.data, fem._Aux1.data, mode=3, p0=2.0 * fem.mu, p1=fem.rho, aux=fem.vel.data)
        fem._halo(fem._Aux1, fem.dim_s)
        op["DivSrT"].csr.mult(fem._Aux1.data, fem._rhs.data, mode=2, p0=1 / fem.rho)
        fem._halo(fem._rhs, fem.dim)
        op["Curl"].csr.mult(fem._rhs.data, f.data)

    f = fem.vort.duplicate()
    for _ in range(3):
        chain()
    barrier()
    s, e = ev(), ev()
    nrep = 10
    s.record()
    for _ in range(nrep):
        chain()
    e.record()
    barrier()
    t_chain = maxr(s.elapsed_time(e) / nrep)
    # algorithmic bytes: each product's own accounting (element operator: x, connectivity, element results, y;
    # CSR stream: 12 B per stored entry of the zero-squeezed operator + vectors) + v read again by the fused v(x)v
    chain_bytes = sum(op[k].csr.bytes_spmv() for k in ("SrT", "DivSrT", "Curl")) + 8 * 2 * N
    csr_bytes = sum(op[k].csr.bytes_spmv(templates=False) for k in ("SrT", "DivSrT", "Curl")) + 8 * 2 * N
    # ---- the same solve end to end from host buffers: vorticity host -> device, solveKLE, velocity device -> host
    w_h = fem.vort.data.cpu().pin_memory()
    v_h = torch.empty(2 * N, dtype=torch.float64).pin_memory()
    barrier()
    s, e = ev(), ev()
    s.record()
    fem.vort.data.copy_(w_h, non_blocking=True)
    fem.solveKLE(0.0, fem.vort)
    v_h.copy_(fem.vel.data, non_blocking=True)
    e.record()
    barrier()
    t_e2e = maxr(s.elapsed_time(e) * 1e-3)
    assert bool(torch.isfinite(v_h).all())
    n_rhs = 7
    per_step = n_rhs * (t_kle + t_chain * 1e-3)
    out = {
        "ksp_type": "cg", "pc": a.pc, "rtol": fem.solver.rtol, "operators_setup_ms": t_ops,
        "solve_kle": {"seconds": t_kle, "free_slip_solve": fs, "no_slip_solve": ns,
                      "what": "Cavity.solveKLE: 6 SpMV + KSP on K+Kfs + Curl + KSP on K, zero initial guess, "
                              "KSPConvergedDefault at the stated rtol"},
        "rhs_chain": {"ms": t_chain, "bytes": chain_bytes, "achieved": chain_bytes / (t_chain * 1e-3) / 1e9,
                      "peak": peak, "unit": "GB/s", "frac": chain_bytes / (t_chain * 1e-3) / 1e9 / peak,
                      "csr_stream_bytes": csr_bytes,
                      "what": "SrT (+ v(x)v, 2 mu / -rho epilogue), DivSrT (1/rho), Curl on the zero-squeezed "
                              "operators; interior rows through the element operator, bytes = that path's own "
                              "accounting; csr_stream_bytes = what the CSR stream of the same operators would move"},
        "seconds_per_time_step": per_step,
        "time_step": f"RK 5(4) Bogacki-Shampine, FSAL: {n_rhs} evalRHS per accepted step = {2 * n_rhs} KSP solves + "
                     f"{n_rhs} RHS chains (measured solveKLE + chain, times {n_rhs})",
        "e2e_solve": {"seconds": t_e2e, "h2d_bytes": 8 * N * world, "d2h_bytes": 16 * N * world,
                      "what": "pinned host vorticity -> device, Cavity.solveKLE, velocity -> pinned host"},
    }
    return out



# ------------------------------------------------------------------------------------------------------------------
# per-cell workload (BASELINE configs[1]/[2]): FreeSlip.buildKLEMats integrates EVERY cell (base_problem.py:503-506)
# ------------------------------------------------------------------------------------------------------------------
def elem_flops(n):
    """FP64 operations of one sum-factorised element integration (csrc/elem.cu), K and Rw: step 1 builds 16 + 8 T
    arrays (2 flop per multiply-add over the q2 index, plus the table product), step 2 contracts q1 for the n^4 (i, j)
    pairs.  nqF = n (full rule), nqG = n - 1 (reduced)."""
    nqF, nqG, nn = n, n - 1, n * n
    step1 = 3 * (4 * nqF * nn * nqF + 12 * nqG * nn * nqG) + 3 * (4 * nqF * nn * nqF + 4 * nqG * nn * nqG)
    step2_k = nn * nn * (nqF * 12 + nqG * 36)
    step2_r = nn * nn * (nqF * 10 + nqG * 10)
    return step1 + step2_k + step2_r


def run_percell(a, emit=True):
    """One step = integrate K_e, Rw_e of every cell from its corner coordinates + assemble K, Krhs, Rw (values on the
    prebuilt pattern) through the reference's case class (CustomFuncCase -> FreeSlip.buildKLEMats)."""
    import torch
    import yaml
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device (B200); there is no CPU fallback.")
    torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
    from pynama_b200 import backend as be
    from pynama_b200.cases.custom_func import CustomFuncCase
    from pynama_b200.common.options import Options
    Options.clear()
    for k, v in (("ksp_type", "cg"), ("pc_type", "jacobi")):
        Options.setValue(k, v)
    if a.fused is not None:
        Options.setValue("mat_fused", a.fused)
    with open(os.path.join(ROOT, "pynama_b200", "cases", "taylor-green.yaml")) as f:
        cfg = yaml.load(f, Loader=yaml.Loader)
    nx, n = a.nx, a.ngl
    fem = CustomFuncCase(cfg, case="taylor-green", ngl=n, nelem=[nx, nx])
    fem.setUpGeneral()
    fem.setUpBoundaryConditions()
    fem.setUpEmptyMats()
    fem.buildKLEMats()
    torch.cuda.synchronize()
    ev = lambda: torch.cuda.Event(enable_timing=True)
    names = ("elem", "K", "Krhs", "Rw")

    def step(events=None):
        after = None
        if events is not None:
            events[0].record()
            after = {nm: events[i + 1].record for i, nm in enumerate(names)}
        fem.mat.assembleFromMesh(fem.dom, fem.elemType, perCell=True, after=after)

    for _ in range(a.warmup):
        step()
    torch.cuda.synchronize()
    sampler = ClockSampler(0)
    sampler.start()
    time.sleep(0.25)
    l0 = be.launch_count()
    evs = [[ev() for _ in range(len(names) + 1)] for _ in range(a.steps)]
    t0, t1 = ev(), ev()
    t0.record()
    for i in range(a.steps):
        step(evs[i])
    t1.record()
    torch.cuda.synchronize()
    launches = be.launch_count() - l0
    clocks = sampler.stop()
    ms_step = t0.elapsed_time(t1) / a.steps
    phase = {nm: sum(e[i].elapsed_time(e[i + 1]) for e in evs) / a.steps for i, nm in enumerate(names)}
    peak, peak_src = measured_peak_hbm()
    K, Rw, Krhs = fem.mat.K.csr, fem.mat.Rw.csr, fem.mat.Krhs.csr
    ne, nn = fem.dom.ncell, n * n
    dof = K.shape[0]
    fused = bool(getattr(fem.mat, "fusedAssembly", False))
    elem_bytes = 64 * ne + (0 if fused else 8 * ne * (4 * nn * nn + 2 * nn * nn))
    k_bytes = 8 * K.nnz + 8 * (K.shape[0] + 1) + 64 * ne
    fl = elem_flops(n) * ne
    t_k = phase["K"] + (phase["elem"] if fused else 0.0)
    out = {
        "metric": METRIC, "value": dof / (ms_step * 1e-3), "unit": "DOF/s", "n_gpus": 1, "steps": a.steps,
        "warmup": a.warmup, "ms_per_step": ms_step, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"per-cell: Taylor-Green {nx}x{nx} elements ngl {n} (BASELINE configs[1]/[2]): "
                               f"getElemKLEMatrices for EVERY cell + K, Krhs, Rw values on the prebuilt pattern "
                               f"(FreeSlip.buildKLEMats, base_problem.py:499-552)",
                   "dof_rows": dof, "cells": ne, "nnz": {"K": K.nnz, "Krhs": Krhs.nnz, "Rw": Rw.nnz},
                   "fused_integrate_assemble": fused,
                   "l2": "outputs exceed the 126 MB L2; no flush needed"},
        "phases_ms": phase,
        "integration": {"kernel": "elem_affine_coef_kernel: 17 closed-form coefficients per cell, entries evaluated inside "
                                  "the assembly kernels (flop / tflops below do not apply)", "ms": phase["elem"],
                        "bytes": elem_bytes + 144 * ne} if fused else
                       {"kernel": "elem_kle_kernel (K_e, Rw_e of every cell)", "ms": phase["elem"], "flop": fl,
                        "tflops": fl / (phase["elem"] * 1e-3) / 1e12 if phase["elem"] > 0 else None,
                        "bytes": elem_bytes, "gbs": elem_bytes / (phase["elem"] * 1e-3) / 1e9 if phase["elem"] > 0 else None,
                        "hbm_frac": elem_bytes / (phase["elem"] * 1e-3) / 1e9 / peak if phase["elem"] > 0 else None,
                        "flop_per_byte": fl / elem_bytes,
                        "fp64_peak_tflops": 37.0, "fp64_frac": fl / (phase["elem"] * 1e-3) / 1e12 / 37.0 if phase["elem"] > 0 else None},
        "roofline": {"bound": "hbm", "kernel": "K from corner coordinates: " + ("fused integrate+assemble" if fused else
                                                                             "asm_interior_kernel<2,2> on materialised K_e"),
                     "achieved": k_bytes / (t_k * 1e-3) / 1e9, "peak": peak, "unit": "GB/s", "peak_source": peak_src,
                     "frac": k_bytes / (t_k * 1e-3) / 1e9 / peak, "ms_per_launch": t_k, "bytes_per_launch": k_bytes,
                     "traffic": None},
        "gpu_launches": int(launches), "clocks": clocks,
    }
    if emit:
        print(json.dumps(out), flush=True)
    return out


def percell_record(a, torch):
    """BASELINE configs[2] at the order of the headline case riding along in the default line: the per-cell path
    (FreeSlip.buildKLEMats, every cell integrated) at ~16 M velocity dofs, compressed to its timings and roofline."""
    import copy
    b = copy.copy(a)
    b.nx, b.steps, b.warmup, b.fused = round(2828 / (a.ngl - 1)), 5, 3, None
    r = run_percell(b, emit=False)
    return {"workload": r["config"]["workload"], "dof_rows": r["config"]["dof_rows"], "cells": r["config"]["cells"],
            "nnz_K": r["config"]["nnz"]["K"], "fused_integrate_assemble": r["config"]["fused_integrate_assemble"],
            "value": r["value"], "unit": r["unit"], "ms_per_step": r["ms_per_step"], "steps": r["steps"],
            "phases_ms": r["phases_ms"], "roofline": r["roofline"], "gpu_launches": r["gpu_launches"]}


def parity_checks(prob, be, torch, world, nx, ngl, dof_total):
    """Size-independent invariants evaluated on every rank in the measured run (tests/test_fullsize_gpu.py holds the
    full set): the owned rows of all ranks add up to the serial system, and a constant velocity field is in the
    null space of K - Krhs on every owned row ((K - Krhs) t = 0: rigid translations carry no strain), which
    fails if any element contribution, boundary class or ghost/owner decision is wrong."""
    import torch.distributed as dist
    K, Krhs = prob.mats["K"], prob.mats["Krhs"]
    N = prob.dom.getTotalNodes()
    res = 0.0
    for comp in (0, 1):
        t = be.zeros(2 * N)
        t[comp::2] = 1.0
        y, y2 = be.empty(2 * N), be.empty(2 * N)
        K.mult(t, y, planned=False)
        Krhs.mult(t, y2, planned=False)
        scale = float(y.abs().max())
        res = max(res, float((y - y2).abs().max()) / max(scale, 1e-300))
    serial_dofs = 2 * (nx * (ngl - 1) + 1) ** 2
    nnz_owned = torch.tensor([float(K.nnz), res], dtype=torch.float64, device="cuda")
    if world > 1:
        mx = nnz_owned.clone()
        dist.all_reduce(nnz_owned, op=dist.ReduceOp.SUM)
        dist.all_reduce(mx, op=dist.ReduceOp.MAX)
        res = float(mx[1])
    return {"owned_dofs_sum_equals_serial": int(dof_total) == serial_dofs, "serial_dofs": serial_dofs,
            "nnz_K_sum_over_ranks": int(nnz_owned[0]), "nullspace_residual_max_over_ranks": res,
            "nullspace_ok": bool(res < 1e-12),
            "peer_memory": bool(getattr(prob.fem.comm.dist, "peer_memory", False)) if world > 1 else None}


def largest_mesh_record(a, be, torch, world, barrier, peak):
    """BASELINE configs[4] -- 4096 x 4096 elements, ngl 4 (302 M velocity dofs, K = 181 GB: needs >= 4 GPUs) -- measured
    in the same run as configs[3] when the job has >= 4 ranks: assembly step and CG iteration, same method."""
    import torch.distributed as dist
    nx, n = 4096, 4
    prob = CavityAssembly(nx, n)
    for _ in range(2):
        prob.step()
    barrier()
    ev = lambda: torch.cuda.Event(enable_timing=True)
    steps = 3
    t0, t1 = ev(), ev()
    t0.record()
    for _ in range(steps):
        prob.step()
    t1.record()
    barrier()
    tt = torch.tensor([t0.elapsed_time(t1) / steps, float(prob.dof_owned)], dtype=torch.float64, device="cuda")
    tmax = tt.clone()
    dist.all_reduce(tmax, op=dist.ReduceOp.MAX)
    dist.all_reduce(tt, op=dist.ReduceOp.SUM)
    ms, dof = float(tmax[0]), float(tt[1])
    ksp = ksp_phase(a, prob, be, torch, peak, barrier)
    t = torch.tensor([ksp["ms_per_solve"]], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    checks = parity_checks(prob, be, torch, world, nx, n, dof)
    out = {"workload": f"cavity {nx}x{nx} elements ngl {n} (BASELINE configs[4]), {world} strips",
           "value": dof / (ms * 1e-3), "unit": "DOF/s", "ms_per_step": ms, "dof_rows": int(dof),
           "ksp": {"its": ksp["its"], "ms_per_iteration": float(t[0]) / max(ksp["its"], 1), "operator": ksp["operator"],
                   "csr_stream_ms_per_iteration": ksp["csr_stream"]["ms_per_iteration"]},
           "parity_checks": checks}
    del prob
    torch.cuda.empty_cache()
    return out


def load_traffic(kernel_substr):
    """DRAM bytes per launch of a kernel from the committed `ncu --set full` summary (profiles/traffic.json), or None."""
    p = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        for k, v in json.load(open(p)).items():
            if kernel_substr in k:
                return v
    except Exception:
        pass
    return None


def ksp_phase(a, prob, be, torch, peak, barrier):
    """CG + Jacobi (KSPCG semantics) on the assembled K with a fixed iteration budget (-ksp_max_it), rhs =
    Rw*vort + Krhs*vel_bc for the lid velocity and a synthetic vorticity field; CUDA events around the solve.  On N
    ranks the solve is the distributed one (halo exchange of the search direction + 3 all-reduced scalars per
    iteration); times are rank-local here and maxed over ranks by the caller."""
    K, Rw, Krhs = prob.mats["K"], prob.mats["Rw"], prob.mats["Krhs"]
    N = prob.dom.getTotalNodes()
    dc = prob.fem.comm.dist
    g = torch.Generator(device="cuda").manual_seed(1234)
    vort = torch.randn(N, dtype=torch.float64, device="cuda", generator=g)
    if dc is not None:
        dc.exchange(vort, 1)
    vel = be.zeros(2 * N)
    lid = prob.dom.borderNodesArray("up")
    if len(lid):
        be.vec_scatter_const(vel, be.to_device(lid) * 2, 1.0)
    rhs = be.empty(2 * N)
    Rw.mult(vort, rhs)
    Krhs.mult(vel, rhs, mode=1)
    its = a.ksp_its
    x = be.empty(2 * N)
    y = be.empty(2 * N)
    ev = lambda: torch.cuda.Event(enable_timing=True)

    def time_cg():
        ksp = be.Ksp(K, "cg", "jacobi", rtol=1e-300, atol=1e-300, max_it=its, check_every=its)
        if dc is not None:
            dc.attach(ksp, 2)
        ksp.solve(rhs, x)                       # warm-up: graph capture + one full budget
        ksp.solve(rhs, x)
        barrier()
        ms = []
        for _ in range(3):
            barrier()
            s, e = ev(), ev()
            s.record()
            r = ksp.solve(rhs, x)
            e.record()
            torch.cuda.synchronize()
            ms.append(s.elapsed_time(e))
        del ksp
        return min(ms), r

    def time_mult():
        for _ in range(3):
            K.mult(x, y)
        s, e = ev(), ev()
        s.record()
        nrep = 10
        for _ in range(nrep):
            K.mult(x, y)
        e.record()
        torch.cuda.synchronize()
        return s.elapsed_time(e) / nrep

    def leg(kernel, traffic_key):
        """one operator variant: CG with the fixed budget + the bare product y = K x, against the HBM roofline"""
        ms_solve, r = time_cg()
        ms_mult = time_mult()
        nbytes = K.bytes_spmv()
        it_bytes = nbytes + 8 * K.shape[0] * 10      # + 10 vector passes per PCG iteration (csrc/ksp.cu; SURVEY 8d: 12)
        ms_it = ms_solve / max(r.its, 1)
        return r, {"kernel": kernel, "ms_per_solve": ms_solve, "ms_per_iteration": ms_it,
                   "iteration_bytes": it_bytes, "iteration_gbs": it_bytes / (ms_it * 1e-3) / 1e9,
                   "iteration_frac_of_peak": it_bytes / (ms_it * 1e-3) / 1e9 / peak,
                   "spmv": {"ms": ms_mult, "bytes": nbytes, "achieved": nbytes / (ms_mult * 1e-3) / 1e9, "peak": peak,
                            "unit": "GB/s", "frac": nbytes / (ms_mult * 1e-3) / 1e9 / peak,
                            "traffic": load_traffic(traffic_key)}}

    # (1) the assembled CSR streamed from HBM: the north_star's "vectorised, coalesced CSR SpMV"
    saved = K._tmat_asm
    K.disable_templates()
    r, csr_leg = leg("cg_spmv_dot_kernel / spmv_stream_kernel (CSR stream, 12 B per stored entry)", "cg_spmv_dot_kernel")
    # (2) the element operator the case classes use on uniform meshes (csrc/tmat.cuh): same operator, matrix-free
    eop_leg = None
    if saved is not None or K.enable_templates():
        if saved is not None:
            K._tmat_asm = saved
        r, eop_leg = leg("eop_apply_mma_kernel + cg_tmat_interior_kernel (element operator: element products with the "
                         "element matrix in shared memory + per-node gather)", "eop_product")
    best = eop_leg or csr_leg
    out = {"type": "cg", "pc": "jacobi", "its": int(r.its), "reason": int(r.reason),
           "ms_per_solve": best["ms_per_solve"], "ms_per_iteration": best["ms_per_iteration"],
           "rows": K.shape[0], "nnz": K.nnz, "operator": "element operator" if eop_leg else "csr stream",
           "iteration_gbs": csr_leg["iteration_gbs"], "iteration_frac_of_peak": csr_leg["iteration_frac_of_peak"],
           "spmv": csr_leg["spmv"], "csr_stream": csr_leg, "element_operator": eop_leg,
           "note": f"fixed budget of {its} iterations (-ksp_max_it {its}); rows/nnz/GB/s are rank 0's strip; a converged "
                   f"Jacobi-CG solve at this size needs O(10^4) iterations (SURVEY 7.2).  ms_per_solve / "
                   f"ms_per_iteration are the default path of the case classes (`operator`); spmv / iteration_gbs "
                   f"are the CSR-stream leg (HBM-bound kernel the roofline applies to)"}
    if dc is not None:
        out["halo_bytes_per_iteration"] = dc.halo_bytes(2)
    if a.ksp_converge > 0:
        # time to solution of the reference's solve (KSPConvergedDefault, zero initial guess) with the default operator
        ksp = be.Ksp(K, "cg", "jacobi", rtol=a.ksp_converge, max_it=1000000, check_every=200)
        if dc is not None:
            dc.attach(ksp, 2)
        barrier()
        s, e = ev(), ev()
        s.record()
        r = ksp.solve(rhs, x)
        e.record()
        torch.cuda.synchronize()
        out["converged"] = {"rtol": a.ksp_converge, "its": int(r.its), "reason": int(r.reason),
                            "seconds": s.elapsed_time(e) * 1e-3, "operator": out["operator"]}
        del ksp
    return out


def e2e_phase(a, prob, be, torch, barrier):
    """The same assembly step from HOST buffers through the public API: pinned cell-0 corner coordinates and the
    per-dof boundary classes go host->device, the matrices are integrated + assembled, diag(K) comes back."""
    import numpy as np
    cc_h = torch.from_numpy(prob.dom.templateCellCorners()).pin_memory()
    host_c, host_r = prob.mat.hostDofClasses(prob.dom)
    cls_h = torch.from_numpy(host_c).pin_memory()
    rcls_h = torch.from_numpy(host_r).pin_memory() if prob.dom.partition is not None else None
    diag_h = [torch.empty(prob.nrows, dtype=torch.float64).pin_memory() for _ in range(2)]
    ev = lambda: torch.cuda.Event(enable_timing=True)
    main, side, up = torch.cuda.current_stream(), torch.cuda.Stream(), torch.cuda.Stream()
    hold = [None, None]         # inputs + boundary maps of the last two steps (released on `up`, see below)
    assembled = [None, None]    # event: the assembly that read them has finished

    def one(i):
        # Three streams, two buffers each way.  `up`: this step's inputs host -> device and their boundary maps (the
        # one scalar the map construction reads back synchronises only this stream); `main`: integration + assembly +
        # diagonal; `side`: the result device -> host.  Step i+1's upload overlaps step i's assembly, step i's download
        # overlaps step i+1's assembly; every step still pays its own H2D, kernels and D2H.
        with torch.cuda.stream(up):
            if assembled[i & 1] is not None:
                up.wait_event(assembled[i & 1])     # step i-2 no longer reads the buffers released just below
            cc = cc_h.to("cuda", non_blocking=True)
            cls = cls_h.to("cuda", non_blocking=True)
            rcls = rcls_h.to("cuda", non_blocking=True) if rcls_h is not None else None
            maps = prob.mat.prepareBoundaryMaps(prob.dom, cls, rcls)
            hold[i & 1] = (cc, cls, rcls, maps)
            ready = torch.cuda.Event()
            ready.record(up)
        main.wait_event(ready)
        for t in (cc, cls, rcls):
            if t is not None:
                t.record_stream(main)
        mats = prob.step(classes=cls, row_classes=rcls, corners=cc)
        d = mats["K"].diagonal()
        done = torch.cuda.Event()
        done.record(main)
        assembled[i & 1] = done
        side.wait_event(done)
        with torch.cuda.stream(side):
            diag_h[i & 1].copy_(d, non_blocking=True)
        d.record_stream(side)

    for i in range(2):
        one(i)
    main.wait_stream(side)
    barrier()
    steps = max(3, min(a.steps, 20))
    s, e = ev(), ev()
    s.record()
    for i in range(steps):
        one(i)
    main.wait_stream(side)
    e.record()
    torch.cuda.synchronize()
    assert float(diag_h[0][0]) != 0.0 and float(diag_h[1][0]) != 0.0
    h2d = cc_h.numel() * 8 + cls_h.numel() + (rcls_h.numel() if rcls_h is not None else 0)
    return {"ms": s.elapsed_time(e) / steps, "h2d": h2d, "d2h": diag_h[0].numel() * 8,
            "what": "pinned host corner coords + int8 dof classes -> device; integrate + assemble 8 matrices; "
                    "diag(K) -> pinned host; uploads (+ boundary maps) and downloads on their own streams, two buffers "
                    "each way: they overlap the neighbouring steps' assembly"}


if __name__ == "__main__":
    args = parse_args()
    if args.impl == "reference":
        run_reference(args)
    elif args.workload == "percell":
        run_percell(args)
    else:
        run_native(args)
