"""Geometric p/h-multigrid hierarchy for the velocity-from-vorticity operator (`-pc_type mg`).

The reference solves `K v = rhs` with a factor-once sparse LU (src/solver/ksp_solver.py:13-16); the device path solves
it with CG, whose iteration count with Jacobi grows like 1/h.  This module builds the levels of a V-cycle that the
library applies as CG's preconditioner (csrc/mg.cu):

    level 0      the case's own space: nx x ny elements, ngl nodes per side, the case's K (or K + Kfs)
    p-levels     the same mesh at lower orders, ngl -> (ngl-1)//2 + 1 -> ... -> 2
    h-levels     bilinear elements on meshes coarsened by two per direction, down to <= `coarse_cells` elements

Every coarse operator is REDISCRETISED: a coarse case object of the same class (same boundary conditions, same material)
is set up on the coarse mesh through the very same calls as the fine one (setUpGeneral, setUpBoundaryConditions,
setUpEmptyMats, assembleFromMesh), i.e. the same integration / assembly / element-operator kernels run on every level.
Interpolation P evaluates the coarse tensor-product Lagrange basis at the fine nodes' coordinates (box meshes on the
same domain, nested or not), restricted to the unknowns of the solve; R = P^T.  Prescribed dofs (identity rows of K)
are outside the cycle: the library preconditions them with Jacobi.
"""
import ctypes as C

import numpy as np
import torch

from .. import backend as be
from .._lib import call
from ..elements.spectral import localNodeOrder
from ..elements.utilities import lobattoPoints


def coarsening_sequence(nx, ny, ngl, coarse_cells=1024):
    """[(nx, ny, ngl), ...] from the fine level down."""
    seq = [(nx, ny, ngl)]
    n = ngl
    while n > 2:
        n = (n - 1) // 2 + 1
        seq.append((nx, ny, n))
    while nx % 2 == 0 and ny % 2 == 0 and (nx // 2) * (ny // 2) >= 4 and nx * ny > coarse_cells:
        nx, ny = nx // 2, ny // 2
        seq.append((nx, ny, 2))
    while nx * ny > coarse_cells:          # odd counts: non-nested coarsening (interpolation by coordinates copes)
        nx, ny = max(2, (nx + 1) // 2), max(2, (ny + 1) // 2)
        seq.append((nx, ny, 2))
    return seq


def _lagrange(nodes, pts):
    """h[q, a] = l_a(pts[q]) on device (float64)."""
    n = len(nodes)
    h = torch.ones(pts.numel(), n, dtype=torch.float64, device=pts.device)
    for a in range(n):
        for m in range(n):
            if m != a:
                h[:, a] *= (pts - nodes[m]) / (nodes[a] - nodes[m])
    return h


def node_interpolation(dom_c, dom_f):
    """(rows, cols, vals): coarse basis of `dom_c` at the nodes of `dom_f` (device COO on nodes, zero weights dropped)."""
    xy = dom_f.deviceNodeCoords()
    nf = xy.shape[0]
    n = dom_c.getNGL()
    lo, up = dom_c.lower, dom_c.upper
    hx, hy = (up[0] - lo[0]) / dom_c.nx, (up[1] - lo[1]) / dom_c.ny_global
    fx = torch.clamp(torch.floor((xy[:, 0] - lo[0]) / hx - 1e-9).to(torch.int64), 0, dom_c.nx - 1)
    # dom_c may be a strip (its cell rows start at global row fy0): cells are addressed in its local numbering
    fy0 = int(getattr(dom_c, "fy0", 0))
    fy = torch.clamp(torch.floor((xy[:, 1] - lo[1]) / hy - 1e-9).to(torch.int64) - fy0, 0, dom_c.ny - 1)
    # element map of the reference: x = x0 + (1 - r) hx / 2  (SURVEY 8a'), r ascending with the 1-D node index
    r = 1.0 - 2.0 * (xy[:, 0] - (lo[0] + hx * fx)) / hx
    s = 1.0 - 2.0 * (xy[:, 1] - (lo[1] + hy * (fy + fy0))) / hy
    nodes1d, _ = lobattoPoints(n)
    hr, hs = _lagrange(nodes1d, r), _lagrange(nodes1d, s)
    order = torch.as_tensor(localNodeOrder(n), device=xy.device)
    H = (hr[:, :, None] * hs[:, None, :]).reshape(nf, n * n)[:, order]        # columns in element-local order
    cn = dom_c.deviceCellNodes()[fy * dom_c.nx + fx].to(torch.int64)             # (nf, nn_c)
    rows = torch.arange(nf, device=xy.device, dtype=torch.int64)[:, None].expand_as(cn)
    keep = H.abs() > 1e-12
    return rows[keep], cn[keep], H[keep]


class _Level(object):
    pass


class Hierarchy(object):
    """V-cycle levels of one solve of a case.  kind "K": the no-slip / Dirichlet solve (unknowns = FREE dofs);
    kind "KFS": the free-slip stage K + Kfs (unknowns = FREE and TANGENT dofs)."""

    def __init__(self, fem, kind, csr, degree=3, ratio=30.0, coarse_cells=1024, power_its=30, partition_p_levels=True):
        if getattr(fem.dom, "nx", None) is None:
            raise Exception("-pc_type mg needs a box mesh (coarse meshes are generated, not read)")
        self.kind, self.degree, self.ratio = kind, int(degree), float(ratio)
        self.levels = []
        # Partitioned mesh: level 0 is this rank's strip (halo exchange before each fine product, restricted residual
        # summed over the ranks); the coarse levels are REPLICATED -- every rank builds them on the whole mesh and
        # runs the same coarse cycle (a 2-D coarse hierarchy is a few per cent of the fine level's work and memory).
        self.dist = fem.comm.dist if fem.dom.partition is not None else None
        seq = coarsening_sequence(fem.dom.nx, fem.dom.ny_global, fem.dom.getNGL(), coarse_cells)
        self.sequence = seq
        for i, (nx, ny, ngl) in enumerate(seq):
            L = _Level()
            # the fine mesh at lower orders keeps the strip partition of level 0 (same cell rows per rank, its own halo
            # plan); the h-coarsened levels are replicated
            same_mesh = (nx, ny) == (fem.dom.nx, fem.dom.ny_global)
            if i == 0:
                self._fine = L
                L.fem, L.csr, L.dist = fem, csr, self.dist
            else:
                part = self.dist is not None and same_mesh and partition_p_levels and i + 1 < len(seq)   # coarsest: dense, replicated
                L.fem = self._coarse_case(fem, nx, ny, ngl, serial=not part)
                L.csr = self._coarse_operator(L.fem)
                L.dist = L.fem.comm.dist if part else None
            cls = L.fem.mat.dofClasses
            L.bc = (cls == 0) if kind == "K" else (cls < 2)       # not prescribed (ghost entries included)
            L.free = L.bc
            if L.dist is not None:                     # unknowns of this rank: owned rows only
                L.free = L.bc & L.dist.owned_rows(fem.dim).bool()
            L.n = int(cls.numel())
            diag = L.csr.diagonal()
            L.dinv = torch.where(L.free, 1.0 / diag, torch.zeros_like(diag)).contiguous()
            L.lmax = self._lambda_max(L, power_its)
            self.levels.append(L)
        for Lf, Lc in zip(self.levels[:-1], self.levels[1:]):
            self._transfer(Lf, Lc)
        self._coarse_inverse(self.levels[-1])
        self._h = C.c_void_p()
        call("pyn_mg_create", len(self.levels), self.degree, self.ratio, C.byref(self._h))
        for i, L in enumerate(self.levels):
            tm = L.csr.tmat()
            L.tm = tm
            call("pyn_mg_set_level", self._h, i, L.n, tm.handle if tm is not None else None, L.csr.plan(),
                 be.ptr(L.csr.indices), be.ptr(L.csr.data), be.ptr(L.dinv), float(L.lmax))
        for i, L in enumerate(self.levels[:-1]):
            call("pyn_mg_set_transfer", self._h, i, L.P.plan(), be.ptr(L.P.indices), be.ptr(L.P.data), L.R.plan(),
                 be.ptr(L.R.indices), be.ptr(L.R.data))
        call("pyn_mg_set_coarse_inverse", self._h, be.ptr(self.Ainv))
        for i, L in enumerate(self.levels):
            if L.dist is not None:
                call("pyn_mg_set_level_dist", self._h, i, L.dist._comm, L.dist._halo, int(fem.dim))

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                from .. import _lib
                _lib.load().pyn_mg_destroy(self._h)
                self._h = None
        except Exception:
            pass

    # -- levels ---------------------------------------------------------------------------------------------------
    @staticmethod
    def _coarse_case(fem, nx, ny, ngl, serial=True):
        import copy
        cfg = copy.deepcopy(fem.config)
        cfg["domain"]["ngl"] = ngl
        cfg["domain"]["box-mesh"]["nelem"] = [nx, ny]
        for k in ("lower", "upper"):                       # keyword overrides of the fine case become the coarse config
            if k in fem.opts:
                cfg["domain"]["box-mesh"][k] = fem.opts[k]
        opts = {k: v for k, v in fem.opts.items() if k not in ("ngl", "nelem", "chart", "lower", "upper", "dist_parent")}
        opts["case"] = fem.case
        opts["serial"] = serial        # replicated level: on N ranks every rank holds the whole coarse level
        if not serial:
            opts["dist_parent"] = fem.comm.dist       # same strips: share the communicator and the peer region
        c = type(fem)(cfg, **opts)
        c.setUpGeneral()
        c.setUpBoundaryConditions()
        c.setUpEmptyMats()
        return c

    def _coarse_operator(self, c):
        from ..matrices.mat_ns import MatNS
        if isinstance(c.mat, MatNS):
            c.mat.assembleFromMesh(c.dom, c.elemType, names=("K", "Kfs") if self.kind == "KFS" else ("K",))
            if self.kind == "KFS":
                A = (c.mat.K + c.mat.Kfs).csr
                if c.mat.K.csr._tmat_asm is not None:
                    A.enable_templates(like=c.mat.K.csr)
                return A
            return c.mat.K.csr
        if self.kind != "K":
            raise Exception("free-slip stage requested on a case without one")
        c.mat.assembleFromMesh(c.dom, c.elemType, perCell=False)      # coarse levels: uniform mesh, one element matrix
        return c.mat.K.csr

    def _lambda_max(self, L, its):
        """Upper bound of the spectrum of D^-1 A on the unknowns: power iteration (deterministic start), times 1.1."""
        g = torch.Generator(device=L.dinv.device).manual_seed(20260)
        x = torch.randn(L.n, dtype=torch.float64, device=L.dinv.device, generator=g) * L.free
        y = torch.empty_like(x)
        dist_ = L.dist                                     # partitioned level: ghosts refreshed, norms all-reduced

        def norm(v):
            s2 = float((v * v).sum())
            if dist_ is not None:
                s2 = dist_.allreduce_sum([s2])[0]
            return s2 ** 0.5

        lam = 1.0
        for _ in range(its):
            if dist_ is not None:
                dist_.exchange(x, L.fem.dim)
            L.csr.mult(x, y)
            y *= L.dinv
            ny_ = norm(y)
            lam = ny_ / norm(x * L.free)
            x = y / ny_
        return 1.1 * lam

    def _transfer(self, Lf, Lc):
        rows, cols, vals = node_interpolation(Lc.fem.dom, Lf.fem.dom)
        dim = Lf.fem.dim
        r2 = (rows[:, None] * dim + torch.arange(dim, device=rows.device)).reshape(-1)
        c2 = (cols[:, None] * dim + torch.arange(dim, device=rows.device)).reshape(-1)
        v2 = vals[:, None].expand(-1, dim).reshape(-1)
        i32 = lambda t: t.to(torch.int32).contiguous()
        if Lc.dist is not None:
            # strip -> strip: both operators GATHER over every local node of the other level, ghost entries included
            # (refreshed by a halo exchange first), and write the rank's own unknowns: globally R is still P^T
            kp = Lf.free[r2] & Lc.bc[c2]
            kr = Lc.free[c2] & Lf.bc[r2]
            Lf.P = be.coo_to_csr(i32(r2[kp]), i32(c2[kp]), v2[kp].contiguous(), (Lf.n, Lc.n), "P")
            Lf.R = be.coo_to_csr(i32(c2[kr]), i32(r2[kr]), v2[kr].contiguous(), (Lc.n, Lf.n), "R")
            return
        keep = Lf.free[r2] & Lc.free[c2]
        r2, c2, v2 = i32(r2[keep]), i32(c2[keep]), v2[keep].contiguous()
        Lf.P = be.coo_to_csr(r2, c2, v2, (Lf.n, Lc.n), "P")
        # (strip -> replicated: R of a strip has no entries in the rows of coarse dofs outside the strip; the SpMV plan
        # cuts its chunks by rows as well as by stored entries, so those runs of empty rows are spread over many CTAs;
        # the partial results of the ranks are then summed)
        Lf.R = be.coo_to_csr(c2, r2, v2, (Lc.n, Lf.n), "R")

    def _coarse_inverse(self, L):
        if L.n > 8192:
            raise Exception(f"coarsest multigrid level has {L.n} dofs: too many for the dense inverse")
        A = L.csr.to_scipy().toarray()
        free = L.free.cpu().numpy()
        inv = np.zeros_like(A)
        idx = np.flatnonzero(free)
        inv[np.ix_(idx, idx)] = np.linalg.inv(A[np.ix_(idx, idx)])
        self.Ainv = be.to_device(inv.reshape(-1))

    # -- use ------------------------------------------------------------------------------------------------------
    def attach(self, ksp):
        call("pyn_ksp_set_mg", ksp._h, self._h)
        ksp._mg = self

    def apply(self, r, z):
        """z = V(r) (testing hook)."""
        call("pyn_mg_apply", self._h, be.ptr(r), be.ptr(z), be.stream())
        return z

    def describe(self):
        return " -> ".join(f"{nx}x{ny} ngl {n}" for nx, ny, n in self.sequence) + \
            f"; Chebyshev({self.degree}) on [lmax/{self.ratio:g}, lmax], dense coarse inverse ({self.levels[-1].n} dofs)"
