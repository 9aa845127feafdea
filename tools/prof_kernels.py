"""Short driver for ncu captures: K values-only assembly + planned SpMV + a few CG iterations at nx x nx, ngl n."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bench import CavityAssembly
from pynama_b200 import backend as be

nx = int(sys.argv[1]) if len(sys.argv) > 1 else 512
n = int(sys.argv[2]) if len(sys.argv) > 2 else 5
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
prob = CavityAssembly(nx, nx, n, [0.0, 0.0], [1.0, 1.0])
prob.step()
for _ in range(reps):
    prob.step()
K = prob.mats["K"]
x = torch.randn(K.shape[1], dtype=torch.float64, device="cuda")
y = be.empty(K.shape[0])
for _ in range(reps + 2):
    K.mult(x, y)
ksp = be.Ksp(K, "cg", "jacobi", rtol=1e-300, atol=1e-300, max_it=4, check_every=4)
sol = be.empty(K.shape[0])
ksp.solve(y, sol)
torch.cuda.synchronize()
print("ok", K.nnz)
