"""y = K x through the CSR stream at a given mesh size (for ncu captures of the SpMV kernels).

    python tools/spmv_probe.py --nx 512 [--reps 5]
"""
import argparse
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402

from bench import CavityAssembly  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--nx", type=int, default=512)
ap.add_argument("--ngl", type=int, default=5)
ap.add_argument("--reps", type=int, default=5)
a = ap.parse_args()
from pynama_b200 import backend as be  # noqa: E402
prob = CavityAssembly(a.nx, a.ngl)
K = prob.mats["K"]
K.disable_templates()
x = torch.randn(K.shape[1], dtype=torch.float64, device="cuda")
y = be.empty(K.shape[0])
for _ in range(2):
    K.mult(x, y)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.reps):
    K.mult(x, y)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.reps
nb = K.bytes_spmv(templates=False)
print(f"nx {a.nx} nnz {K.nnz} ms {ms:.3f} GB/s {nb / ms / 1e6:.0f}")
