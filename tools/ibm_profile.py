"""Timings of the immersed-boundary coupling on one GPU: what ImmersedBoundaryDynamic pays every time
step -- rebuilding H, S, A where the body is (rebuildMatrix, src/cases/immersed_boundary.py:416-433) and the velocity
correction (computeVelocityCorrection, :405-414) -- next to the oracle's Python loops on a sample of the markers.

    python tools/ibm_profile.py [nelem] [ngl] > profiles/r02_ibm_profile.json
"""
import json
import os
import sys
import time

import numpy as np
import torch
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import kle_oracle as o  # noqa: E402
from pynama_b200 import backend as be  # noqa: E402
from pynama_b200.cases.immersed_boundary import ImmersedBoundaryDynamic  # noqa: E402
from pynama_b200.common.options import Options  # noqa: E402

nelem = int(sys.argv[1]) if len(sys.argv) > 1 else 400
ngl = int(sys.argv[2]) if len(sys.argv) > 2 else 3
Options.clear()
for k, v in dict(ksp_type="cg", pc_type="jacobi", ksp_rtol=1e-10).items():
    Options.setValue(k, v)
with open(os.path.join(ROOT, "pynama_b200", "cases", "ibm-dynamic.yaml")) as f:
    cfg = yaml.load(f, Loader=yaml.Loader)
cfg["domain"]["box-mesh"]["nelem"] = [nelem, nelem]
cfg["domain"]["ngl"] = ngl
fem = ImmersedBoundaryDynamic(cfg, case="ibm-dynamic")
fem.setUpGeneral()
fem.setUpBoundaryConditions()
fem.setUpEmptyMats()
fem.buildOperators()
fem.setUpBodies()
cells = fem.getAffectedCells(10)
fem.collectedNodes, fem.maxNodesPerLag = fem.collectNodes(cells)
fem.totalEulerNodes = fem.dom.getTotalNodes()
fem.buildIBMMatrix()
from pynama_b200.petsc_like import Vec  # noqa: E402
fem.vel = Vec(2 * fem.totalEulerNodes)
fem.vel.set(fem.U_ref)
fem.vel_correction = fem.vel.duplicate()
fem.virtualFlux, fem.ibm_rhs = fem.S.createVecRight(), fem.S.createVecRight()


def timed(fn, reps=5):
    fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(reps):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / reps


lagxy = be.to_device(fem.body.coordinatesArray().reshape(-1))
nodexy = fem.dom.deviceNodeCoords().view(-1)
dl, h = fem.body.getElementLong(), fem.h
t_collect = timed(lambda: be.ibm_collect(lagxy, fem._ibmCandidates, nodexy, dl, h, mode=1, kind="four"))
p, nodes, dirac = fem.collectedNodes
t_gram = timed(lambda: be.ibm_gram(p, nodes, dirac, h ** 2, dl * h))
t_rebuild = timed(fem.rebuildMatrix)
t_corr = timed(lambda: ImmersedBoundaryDynamic.__mro__[1].computeVelocityCorrection(fem))
nlag = fem.body.getTotalNodes()
# the oracle's loops on a sample of the markers, scaled to all of them
sample = min(nlag, 4)
host_xy, cand = fem.dom.deviceNodeCoords().cpu().numpy().reshape(-1, 2), fem._ibmCandidates.cpu().numpy()
t0 = time.perf_counter()
o.ibm_collect(fem.body.coordinatesArray()[:sample], cand, host_xy, dl, h, mode=1, kind="four")
t_cpu = (time.perf_counter() - t0) / sample * nlag
print(json.dumps({
    "workload": f"ibm-dynamic: circle r = 0.5 on a {nelem}x{nelem} box mesh ngl {ngl} (h = {h:.4g})",
    "markers": nlag, "euler_nodes": fem.totalEulerNodes, "candidate_nodes": int(fem._ibmCandidates.numel()),
    "close_nodes_per_marker_max": fem.maxNodesPerLag, "nnz": {"H": fem.H.csr.nnz, "A": fem.A.csr.nnz},
    "ms": {"ibm_collect (count + fill)": t_collect, "ibm_gram (count + fill)": t_gram,
           "rebuildMatrix (collect + H, S, A + solver set-up)": t_rebuild,
           "computeVelocityCorrection (H v, A q = -rhs, v += S q)": t_corr},
    "ksp_A": {"its": fem.ksp.getIterationNumber(), "reason": fem.ksp.getConvergedReason()},
    "cpu_oracle_collect_s": {"value": t_cpu, "what": f"oracle ibm_collect on {sample} markers x {len(cand)} candidates, "
                                                     f"scaled to {nlag} markers (one core)"},
}))
