"""Development aid: where does a values-only K assembly spend its time (device vs host)?"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bench import CavityAssembly
from pynama_b200 import backend as be

nx = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
prob = CavityAssembly(nx, 5)
m, fem = prob.mat, prob.fem
print("nsig", fem.dom.topology().nsig, "slots", fem.dom.topology().sig_slots, flush=True)
ev = lambda: torch.cuda.Event(enable_timing=True)
for tmpl in (True, False, True):
    be.set_templates(tmpl)
    for names in (("K",), ("Rw",), ("K", "Krhs", "Kfs", "Krhsfs", "Rw", "Rwfs", "Rd", "Rdfs")):
        for _ in range(2):
            m.assembleFromMesh(fem.dom, fem.elemType, names=names, classes=m.dofClasses)
        torch.cuda.synchronize()
        s, e = ev(), ev()
        t0 = time.perf_counter(); s.record()
        for _ in range(5):
            m.assembleFromMesh(fem.dom, fem.elemType, names=names, classes=m.dofClasses)
        e.record(); t1 = time.perf_counter()
        torch.cuda.synchronize(); t2 = time.perf_counter()
        nbytes = sum(8 * getattr(m, n).csr.nnz for n in names)
        ms = s.elapsed_time(e) / 5
        print(f"templates={tmpl} {names[0]}{'+' if len(names)>1 else ''}: device {ms:.3f} ms ({nbytes/ms/1e6:.0f} GB/s), "
              f"host enqueue {(t1-t0)/5*1e3:.3f} ms, wall {(t2-t0)/5*1e3:.3f} ms", flush=True)
