"""Development aid: wall time of each segment of bench.py's end-to-end step (synchronised after every segment)."""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
from bench import CavityAssembly
from pynama_b200 import backend as be

nx = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
prob = CavityAssembly(nx, 5)
prob.step()
cc_h = torch.from_numpy(np.asarray(prob.dom.getCellCornersCoords(0)).reshape(1, 4, 2)).pin_memory()
host_c, _ = prob.mat.hostDofClasses(prob.dom)
cls_h = torch.from_numpy(host_c).pin_memory()
diag_h = torch.empty(prob.nrows, dtype=torch.float64).pin_memory()
topo = prob.dom.topology()


def seg(name, fn, acc):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = fn()
    torch.cuda.synchronize()
    acc[name] = acc.get(name, 0.0) + (time.perf_counter() - t0) * 1e3
    return out


for rep in range(3):
    acc = {}
    cls = seg("h2d classes", lambda: cls_h.to("cuda", non_blocking=True), acc)
    cc = seg("h2d corners", lambda: cc_h.to("cuda", non_blocking=True), acc)
    seg("bcmap 2x2", lambda: be.BcMap(topo, cls, 2, cls, 2), acc)
    seg("bcmap 2x1", lambda: be.BcMap(topo, cls, 2), acc)
    mats = seg("step (incl. both bcmaps again)", lambda: prob.step(classes=cls, corners=cc), acc)
    d = seg("diagonal", lambda: mats["K"].diagonal(), acc)
    seg("d2h diag", lambda: diag_h.copy_(d, non_blocking=True), acc)
    print({k: round(v, 3) for k, v in acc.items()}, flush=True)
