"""Scratch timing of the device path at large sizes (not the bench; development aid)."""
import sys, os, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np, torch
from pynama_b200 import backend as be
from pynama_b200.elements.spectral import Spectral

def timed(fn, reps=3):
    torch.cuda.synchronize()
    ts = []
    for _ in range(reps):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); r = fn(); b.record(); torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return min(ts), r

def run(nx, n, per_cell=False):
    s = Spectral(n, 2)
    m = n - 2
    nnodes = nx * nx * m * m + (nx + 1) ** 2 + 2 * nx * (nx + 1) * m
    t0 = time.time()
    e2n = be.box_cell_nodes(nx, nx, n)
    topo = be.Topology(e2n, nnodes)
    torch.cuda.synchronize()
    t_topo = time.time() - t0
    # boundary classes: compute with torch from coordinates
    xy = be.box_node_coords(nx, nx, n, [0, 0], [1, 1], s.nodes1D, s.localOrder, nnodes)
    bc = (xy[:, 0] == 0) | (xy[:, 0] == 1) | (xy[:, 1] == 0) | (xy[:, 1] == 1)
    cls = (bc.to(torch.int8) * 2).repeat_interleave(2).contiguous()
    cc = be.box_cell_corners(nx, nx, [0, 0], [1, 1], 0, nx * nx if per_cell else 1)
    t_el, (Ke, Rwe, _) = timed(lambda: s.getElemKLEMatricesDevice(cc, want_rd=False), 2)
    t0 = time.time()
    mp = be.BcMap(topo, cls, 2, cls, 2)
    torch.cuda.synchronize()
    t_map = time.time() - t0
    t0 = time.time()
    K = be.assemble(topo, 2, 2, Ke, cls, cls, [(0, 0)], diag_classes=0b100, diag_add=(0, 0, 1.0, 0), bcmap=mp)
    torch.cuda.synchronize()
    t_first = time.time() - t0
    t_fill, _ = timed(lambda: be.assemble(topo, 2, 2, Ke, cls, cls, [(0, 0)], diag_classes=0b100,
                                          diag_add=(0, 0, 1.0, 0), pattern=K, bcmap=mp))
    x, y = torch.randn(2 * nnodes, dtype=torch.float64, device="cuda"), be.empty(2 * nnodes)
    t_spmv, _ = timed(lambda: K.mult(x, y), 5)
    gb = K.bytes_spmv() / 1e9
    ksp = be.Ksp(K, "cg", "jacobi", rtol=1e-30, max_it=64, check_every=64)
    sol = be.empty(2 * nnodes)
    t_cg, r = timed(lambda: ksp.solve(y, sol), 2)
    print(f"nx={nx} n={n} nodes={nnodes} adj_nnz={topo.adj_nnz} Knnz={K.nnz} | topo {t_topo*1e3:.1f} ms bcmap {t_map*1e3:.1f} ms ({mp.n_boundary}) | elem({cc.shape[0]}) {t_el:.2f} ms | "
          f"asm first {t_first*1e3:.1f} ms, values-only {t_fill:.2f} ms = {8*K.nnz/t_fill/1e6:.0f} GB/s(vals) {2*nnodes/t_fill/1e6:.2f} GDOF/s | "
          f"spmv {t_spmv:.3f} ms = {gb/t_spmv*1e3:.0f} GB/s | cg {r.its} its {t_cg/ max(r.its,1):.3f} ms/it", flush=True)

if __name__ == "__main__":
    import sys as _s
    cases = [(64, 5, True), (512, 5, False), (1024, 5, False), (707, 5, True), (1414, 3, False), (354, 9, False)]
    for nx, n, pc in cases:
        run(nx, n, pc)
