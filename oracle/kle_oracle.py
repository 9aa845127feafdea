"""CPU ORACLE -- TEST INFRASTRUCTURE ONLY (never imported by the product package `pynama_b200`).

A NumPy/SciPy restatement of the Pynama FEM hot path (BASELINE.json north_star):
quadrature tables -> element KLE matrices / operators -> serial DMPlex box-mesh numbering -> global sparse
assembly (explicit zeros kept, duplicates added in ascending cell order) -> PETSc-semantics CG -> velocity
from vorticity -> vorticity right-hand side -> Bogacki-Shampine 5(4) Runge-Kutta.

Every function cites the reference file:line it follows (paths relative to /root/reference/).

Parity pinning (what this oracle has been checked against, see tests/test_oracle_*.py and
oracle/gen_golden.py):
  * element layer (quadrature, tabulation, element K/Rw/Rd, SrT/DivSrT/Curl/W): PINNED against the
    reference's own code imported unmodified in the build container (tests/golden/elem_*.npz) and against the
    closed-form points/weights of src/tests/test_element.py:181-229.
  * assembly loops, operators, right-hand-side chain, two-stage solveKLE, immersed-boundary coupling: PINNED against the
    reference's own case-layer code run UNMODIFIED in the build container on stand-ins for PETSc's Mat / Vec and for the
    mesh queries (oracle/ref_shim.py: run_reference_build_kle_mats, run_reference_operators_and_rhs,
    run_reference_cavity_solve_kle, run_reference_taylor_green_solve_kle, run_reference_cell_nodes,
    run_reference_ibm): IndicesManager.mapEntitiesToNodes (src/domain/indices.py:66-118) over DMPlex-style closures,
    CustomFuncCase + FreeSlip.solveKLE (custom_func.py:62-83,173-193), NoSlipFreeSlip.buildKLEMats and FreeSlip.buildKLEMats
    (src/cases/base_problem.py:329-454, 499-552) with MatNS / Mat.assembleAll / setIndices2One, buildOperators with
    Operators.setValues / assembleAll (mat_generator.py:157-190), evalRHS + computeVtensV (:212-252), solveKLE with
    Cavity.applyBoundaryConditions[FS] and NoSlipWalls (:321-327, cavity.py:53-81), the IBM case code
    (immersed_boundary.py:205-277, 318-338, 391-402).  Golden vectors: tests/golden/asm_ref.npz, ibm_ref.npz
    (oracle/gen_golden_asm.py, gen_golden_ibm.py): stored patterns (explicit zeros included) bit-exact, values to the
    round-off of the element matrices.  What the stand-ins assume of PETSc: serial AIJ keeps every inserted entry,
    ADD_VALUES accumulate in call order, INSERT replaces, MatMult / MatMatMult walk sorted rows.
  * numbering / KSP / TS: the arithmetic lives in PETSc (petsc4py; un-vendored, version not pinned, absent from this
    image).  Pinned only through the reference tests that survive at that boundary:
    src/tests/test_domain.py:22-78,187-201 (cell range, corner order, border/node counts, ngl=2 numbering) and
    src/tests/test_solver.py:20-37,66-87 (uniform flow 1e-12, Taylor-Green 2x2/ngl 11 < 2e-8, VtensV).
    The rest of that layer (edge/interior numbering for ngl>2 -- which the stand-ins take from this oracle's BoxMesh --,
    the CG recurrence, RK tableau/adaptor) is "PARITY UNPINNED": restated from PETSc's documented semantics.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np
import scipy.sparse as sp

ALPHA_W = 1e2  # src/elements/spectral.py:93
ALPHA_D = 1e3  # src/elements/spectral.py:94


# --------------------------------------------------------------------------------------------------------
# Q1/Q2  quadrature rules                                                  src/elements/utilities.py:43-92
# --------------------------------------------------------------------------------------------------------
def gauss_points(N):
    """Gauss-Legendre nodes/weights by Golub-Welsch, symmetrised (utilities.py:43-61)."""
    k = np.arange(1, N)
    beta = 0.5 / np.sqrt(1.0 - (2.0 * k) ** -2.0)
    T = np.diag(beta, 1) + np.diag(beta, -1)
    lam, V = np.linalg.eig(T)
    order = np.argsort(lam)
    x = lam[order]
    w = 2.0 * V[0, order] ** 2
    return (x - x[::-1]) / 2.0, (w + w[::-1]) / 2.0


def lobatto_points(N):
    """Gauss-Lobatto-Legendre nodes/weights: Newton iteration on the Legendre recursion started from the
    Chebyshev-GL points, stop when max|dx| <= 1e-15, then symmetrise, ascending (utilities.py:63-92)."""
    x = np.cos(np.linspace(0.0, math.pi, N))
    P = np.zeros((N, N))
    xold = 2.0
    while np.amax(np.abs(x - xold)) > 1e-15:
        xold = x
        P[:, 0] = 1.0
        P[:, 1] = x
        for k in range(2, N):
            P[:, k] = ((2 * k - 1) * x * P[:, k - 1] - (k - 1) * P[:, k - 2]) / k
        x = xold - (x * P[:, N - 1] - P[:, N - 2]) / (N * P[:, N - 1])
    w = 2.0 / ((N - 1) * N * np.square(P[:, N - 1]))
    return (x[::-1] - x) / 2.0, (w[::-1] + w) / 2.0


# --------------------------------------------------------------------------------------------------------
# Q3  1-D Lagrange basis and derivative                                     src/elements/element.py:17-48
# --------------------------------------------------------------------------------------------------------
def lagrange_1d(nodes, pts):
    """h[q,a] = l_a(pts[q]),  dh[q,a] = l_a'(pts[q]) for the Lagrange basis on `nodes` (element.py:17-48).
    Same product formulas as the reference: l_a = prod_{m!=a}(x-x_m)/prod_{m!=a}(x_a-x_m),
    l_a' = sum_{c!=a} prod_{m!=a,c}(x-x_m) / prod_{m!=a}(x_a-x_m)."""
    nodes = np.asarray(nodes, float)
    pts = np.asarray(pts, float)
    n = len(nodes)
    h = np.zeros((len(pts), n))
    dh = np.zeros((len(pts), n))
    for a in range(n):
        others = [m for m in range(n) if m != a]
        den = np.prod([nodes[a] - nodes[m] for m in others]) if others else 1.0
        for q, x in enumerate(pts):
            h[q, a] = np.prod([x - nodes[m] for m in others]) / den if others else 1.0
            s = 0.0
            for c in others:
                s += np.prod([x - nodes[m] for m in others if m != c]) if len(others) > 1 else 1.0
            dh[q, a] = s / den
    return h, dh


# --------------------------------------------------------------------------------------------------------
# Q4  element-local ordering of the (n x n) tensor grid                  src/elements/spectral.py:220-300
# --------------------------------------------------------------------------------------------------------
def local_order(n):
    """local index k -> tensor flat index a*n+b (a = r-index, b = s-index, both ascending in the 1-D nodes).
    Ring walk equivalent to the `Ind`/`Permlst`/`invPerm` construction of spectral.py:224-246:
    corners (+1,+1),(-1,+1),(-1,-1),(+1,-1); edge 1->2 (s=+1, r decreasing), edge 2->3 (r=-1, s decreasing),
    edge 3->4 (s=-1, r increasing), edge 4->1 (r=+1, s increasing); interior with a ascending (slow) and
    b descending (fast)."""
    if n == 1:
        return np.array([0])
    p = n - 1
    ab = [(p, p), (0, p), (0, 0), (p, 0)]
    ab += [(a, p) for a in range(p - 1, 0, -1)]
    ab += [(0, b) for b in range(p - 1, 0, -1)]
    ab += [(a, 0) for a in range(1, p)]
    ab += [(p, b) for b in range(1, p)]
    ab += [(a, b) for a in range(1, p) for b in range(p - 1, 0, -1)]
    return np.array([a * n + b for a, b in ab])


@dataclass
class Tab2D:
    """Tabulation in the reference's layout: H[q][k], Hrs[q][2,k], points (r,s,w), all in local order."""
    H: np.ndarray    # (nq, nn)
    Hrs: np.ndarray  # (nq, 2, nn)
    r: np.ndarray
    s: np.ndarray
    w: np.ndarray


def tabulate_2d(nodes1d, pts1d, wts1d):
    """computeMats2D (spectral.py:220-300): tensor products of the 1-D tables, nodes and points permuted into
    the element-local order."""
    h, dh = lagrange_1d(nodes1d, pts1d)
    n, nq = len(nodes1d), len(pts1d)
    pn, pq = local_order(n), local_order(nq)
    # tensor tables: point (q1,q2) flat q1*nq+q2 ; node (a,b) flat a*n+b
    H = np.einsum("qa,pb->qpab", h, h).reshape(nq * nq, n * n)
    Hr = np.einsum("qa,pb->qpab", dh, h).reshape(nq * nq, n * n)
    Hs = np.einsum("qa,pb->qpab", h, dh).reshape(nq * nq, n * n)
    H, Hr, Hs = H[pq][:, pn], Hr[pq][:, pn], Hs[pq][:, pn]
    r = np.repeat(pts1d, nq)[pq]
    s = np.tile(pts1d, nq)[pq]
    w = np.outer(wts1d, wts1d).reshape(-1)[pq]
    return Tab2D(H, np.stack([Hr, Hs], axis=1), r, s, w)


@dataclass
class SpectralTables:
    """Everything Spectral.__init__/setUpSpectralMats2D tabulates (spectral.py:17-62), for dim = 2."""
    ngl: int
    nodes1d: np.ndarray
    full: Tab2D = None
    red: Tab2D = None
    op: Tab2D = None
    coo_full: Tab2D = None
    coo_red: Tab2D = None
    coo_op: Tab2D = None
    rules: dict = field(default_factory=dict)  # 1-D rules: name -> (pts, wts)

    @property
    def nnode(self):
        return self.ngl ** 2


def spectral_tables(ngl):
    nodes, w_op = lobatto_points(ngl)
    pts_f, w_f = gauss_points(ngl) if ngl <= 3 else lobatto_points(ngl)  # spectral.py:41-42
    pts_g, w_g = gauss_points(ngl - 1)                                   # spectral.py:43
    cn, _ = lobatto_points(2)                                            # spectral.py:44
    t = SpectralTables(ngl, nodes)
    t.full = tabulate_2d(nodes, pts_f, w_f)
    t.red = tabulate_2d(nodes, pts_g, w_g)
    t.op = tabulate_2d(nodes, nodes, w_op)
    t.coo_full = tabulate_2d(cn, pts_f, w_f)
    t.coo_red = tabulate_2d(cn, pts_g, w_g)
    t.coo_op = tabulate_2d(cn, nodes, w_op)
    t.rules = {"full": (pts_f, w_f), "red": (pts_g, w_g), "op": (nodes, w_op)}
    return t


_TABLE_CACHE = {}


def tables(ngl):
    if ngl not in _TABLE_CACHE:
        _TABLE_CACHE[ngl] = spectral_tables(ngl)
    return _TABLE_CACHE[ngl]


def _phys_grads(tab, ctab, coords):
    """Per element and point: Hxy = inv(J) Hrs, detJ, with J = HrsCoo[q] . coords (spectral.py:118-122).
    coords: (ne, 4, 2) -> Hxy (ne, nq, 2, nn), det (ne, nq)."""
    J = np.einsum("qck,ekd->eqcd", ctab.Hrs, coords)          # rows: d/dr, d/ds ; cols: x, y
    det = J[..., 0, 0] * J[..., 1, 1] - J[..., 0, 1] * J[..., 1, 0]
    Jinv = np.empty_like(J)
    Jinv[..., 0, 0] = J[..., 1, 1] / det
    Jinv[..., 0, 1] = -J[..., 0, 1] / det
    Jinv[..., 1, 0] = -J[..., 1, 0] / det
    Jinv[..., 1, 1] = J[..., 0, 0] / det
    Hxy = np.einsum("eqcd,qdk->eqck", Jinv, tab.Hrs)
    return Hxy, det


# --------------------------------------------------------------------------------------------------------
# E1  element KLE matrices                                                src/elements/spectral.py:89-157
# --------------------------------------------------------------------------------------------------------
def elem_kle_matrices(ngl, coords):
    """Batched restatement of Spectral.getElemKLEMatrices for dim=2.
    coords: (ne, 4, 2) corner coordinates in closure order.  Returns K (ne,2nn,2nn), Rw (ne,2nn,nn),
    Rd (ne,2nn,nn); row/col index = localNode*2 + component (spectral.py:124-126)."""
    t = tables(ngl)
    coords = np.asarray(coords, float).reshape(-1, 4, 2)
    ne, nn = coords.shape[0], t.nnode
    K = np.zeros((ne, 2 * nn, 2 * nn))
    Rw = np.zeros((ne, 2 * nn, nn))
    Rd = np.zeros((ne, 2 * nn, nn))
    # full rule (spectral.py:117-133): one dense rank update per quadrature point, as in the reference
    Hxy, det = _phys_grads(t.full, t.coo_full, coords)
    for q in range(len(t.full.w)):
        wd = (t.full.w[q] * det[:, q])[:, None, None]
        gx, gy, H = Hxy[:, q, 0], Hxy[:, q, 1], t.full.H[q]
        B_gr = np.zeros((ne, 4, 2 * nn))
        B_gr[:, 0, 0::2], B_gr[:, 1, 0::2] = gx, gy      # grad u
        B_gr[:, 2, 1::2], B_gr[:, 3, 1::2] = gx, gy      # grad v
        Hvel = np.zeros((2, 2 * nn))
        Hvel[0, 0::2] = H
        Hvel[1, 1::2] = H
        Bw = np.stack([gy, -gx], axis=1)                 # curl of scalar w: (w_y, -w_x)  (:128-129)
        K += wd * np.einsum("eci,ecj->eij", B_gr, B_gr)
        Rw += wd * np.einsum("ci,ecj->eij", Hvel, Bw)
        Rd -= wd * np.einsum("ci,ecj->eij", Hvel, Hxy[:, q])
    # reduced rule, penalisations (spectral.py:137-156)
    Hxy, det = _phys_grads(t.red, t.coo_red, coords)
    for q in range(len(t.red.w)):
        wd = (t.red.w[q] * det[:, q])[:, None, None]
        gx, gy, H = Hxy[:, q, 0], Hxy[:, q, 1], t.red.H[q]
        B_div = np.zeros((ne, 2 * nn))
        B_div[:, 0::2], B_div[:, 1::2] = gx, gy
        B_curl = np.zeros((ne, 2 * nn))
        B_curl[:, 1::2], B_curl[:, 0::2] = gx, -gy       # v_x - u_y  (:147-148)
        K += wd * (ALPHA_D * np.einsum("ei,ej->eij", B_div, B_div)
                   + ALPHA_W * np.einsum("ei,ej->eij", B_curl, B_curl))
        Rw += wd * ALPHA_W * np.einsum("ei,j->eij", B_curl, H)
        flatF = np.empty((ne, 2 * nn))
        flatF[:, 0::2], flatF[:, 1::2] = gx, gy          # Hxy.flatten('F')  (:156)
        Rd += wd * ALPHA_D * np.einsum("ei,j->eij", flatF, H)
    return K, Rw, Rd


# --------------------------------------------------------------------------------------------------------
# E2  element operators                                                  src/elements/spectral.py:159-218
# --------------------------------------------------------------------------------------------------------
def elem_kle_operators(ngl, coords):
    """Batched restatement of Spectral.getElemKLEOperators (dim=2): SrT (ne,3nn,2nn), DivSrT (ne,2nn,3nn),
    Curl (ne,nn,2nn), W (ne,nn)."""
    t = tables(ngl)
    coords = np.asarray(coords, float).reshape(-1, 4, 2)
    ne, nn = coords.shape[0], t.nnode
    SrT = np.zeros((ne, 3 * nn, 2 * nn))
    Div = np.zeros((ne, 2 * nn, 3 * nn))
    Curl = np.zeros((ne, nn, 2 * nn))
    Wm = np.zeros((ne, nn, nn))
    Hxy, det = _phys_grads(t.op, t.coo_op, coords)
    for q in range(len(t.op.w)):
        wd = (t.op.w[q] * det[:, q])[:, None, None]
        gx, gy, H = Hxy[:, q, 0], Hxy[:, q, 1], t.op.H[q]
        Hsrt = np.zeros((3, 3 * nn))
        for c in range(3):
            Hsrt[c, c::3] = H
        Hdiv = np.zeros((2, 2 * nn))
        Hdiv[0, 0::2] = H
        Hdiv[1, 1::2] = H
        B_curl = np.zeros((ne, 1, 2 * nn))
        B_curl[:, 0, 1::2], B_curl[:, 0, 0::2] = gx, -gy
        # indBdiv = [[0,1],[1,2]] (spectral.py:28,195-199)
        B_div = np.zeros((ne, 2, 3 * nn))
        B_srt = np.zeros((ne, 3, 2 * nn))
        B_div[:, 0, 0::3], B_srt[:, 0, 0::2] = gx, gx    # x=0,i=0 : ind 0
        B_div[:, 1, 1::3], B_srt[:, 1, 1::2] = gx, gx    # x=0,i=1 : ind 1
        B_div[:, 0, 1::3], B_srt[:, 1, 0::2] = gy, gy    # x=1,i=0 : ind 1
        B_div[:, 1, 2::3], B_srt[:, 2, 1::2] = gy, gy    # x=1,i=1 : ind 2
        B_srt[:, 0, 1::2] = -gy                          # :201
        B_srt[:, 2, 0::2] = -gx                          # :202
        B_srt *= 0.5                                     # :207
        SrT += wd * np.einsum("ci,ecj->eij", Hsrt, B_srt)
        Div += wd * np.einsum("ci,ecj->eij", Hdiv, B_div)
        Curl += wd * np.einsum("i,ecj->eij", H, B_curl)
        Wm += wd * np.outer(H, H)[None]
    return SrT, Div, Curl, Wm.sum(2)


# --------------------------------------------------------------------------------------------------------
# D1-D5  serial DMPlex box mesh + PetscSection numbering (restated; PETSc-internal)
# --------------------------------------------------------------------------------------------------------
class BoxMesh:
    """Serial `DMPlexCreateBoxMesh(simplex=False)` + `setFemIndexing(ngl)` (src/domain/dmplex.py:8-61,
    src/domain/indices.py:22-30,66-118), restated from PETSc's documented point numbering:
    cells c=fy*nx+fx; vertices v=vy*(nx+1)+vx; x-edges e=vy*nx+ex; y-edges e=vx*ny+ey (after the x-edges);
    cell cone = [bottom, right, top, left] with top/left reversed; node ids = section offsets in point order
    (cell interiors, then vertices, then edges)."""

    def __init__(self, nelem, lower, upper, ngl):
        self.nx, self.ny = int(nelem[0]), int(nelem[1])
        self.lower = np.asarray(lower, float)
        self.upper = np.asarray(upper, float)
        self.ngl = int(ngl)
        nx, ny, n = self.nx, self.ny, self.ngl
        self.ncell = nx * ny
        self.nvert = (nx + 1) * (ny + 1)
        self.nxedge = nx * (ny + 1)
        self.nyedge = (nx + 1) * ny
        m = n - 2
        self.nnodes = self.ncell * m * m + self.nvert + (self.nxedge + self.nyedge) * m
        self.off_vert = self.ncell * m * m
        self.off_edge = self.off_vert + self.nvert

    # -- geometry -------------------------------------------------------------------------------------
    def cell_corner_coords(self, cells=None):
        """(nc,4,2) in closure order (x0,y0),(x1,y0),(x1,y1),(x0,y1) (dmplex.py:97-104, test_domain.py:26-30)."""
        cells = np.arange(self.ncell) if cells is None else np.asarray(cells)
        fx, fy = cells % self.nx, cells // self.nx
        hx = (self.upper[0] - self.lower[0]) / self.nx
        hy = (self.upper[1] - self.lower[1]) / self.ny
        # PETSc computes vertex coordinates as lower + (upper-lower)/faces * index
        x0, x1 = self.lower[0] + hx * fx, self.lower[0] + hx * (fx + 1)
        y0, y1 = self.lower[1] + hy * fy, self.lower[1] + hy * (fy + 1)
        return np.stack([np.stack([x0, y0], 1), np.stack([x1, y0], 1),
                         np.stack([x1, y1], 1), np.stack([x0, y1], 1)], 1)

    # -- numbering ------------------------------------------------------------------------------------
    def cell_nodes(self, cells=None):
        """(nc, nn) global node ids of each cell in element-local order: 4 vertices in closure order, 4 edges
        (bottom, right, top, left; reversed when the closure orientation is negative), cell interior
        (dmplex.py:197-200 -> indices.py:66-88,116-118)."""
        cells = np.arange(self.ncell) if cells is None else np.asarray(cells)
        nx, ny, m = self.nx, self.ny, self.ngl - 2
        fx, fy = cells % nx, cells // nx
        v = lambda vx, vy: self.off_vert + vy * (nx + 1) + vx
        cols = [v(fx, fy), v(fx + 1, fy), v(fx + 1, fy + 1), v(fx, fy + 1)]
        k = np.arange(m)
        xe = lambda ex, vy: self.off_edge + (vy * nx + ex)[:, None] * m
        ye = lambda vx, ey: self.off_edge + (self.nxedge + vx * ny + ey)[:, None] * m
        edges = [xe(fx, fy) + k,                 # bottom, +x
                 ye(fx + 1, fy) + k,             # right, +y
                 xe(fx, fy + 1) + k[::-1],       # top, traversed -x
                 ye(fx, fy) + k[::-1]]           # left, traversed -y
        interior = (cells * m * m)[:, None] + np.arange(m * m)
        out = np.concatenate([np.stack(cols, 1)] + edges + [interior], axis=1)
        return out.astype(np.int64)

    def node_coords(self):
        """computeFullCoordinates (dmplex.py:66-95): HCooOp[k] @ corners scattered to node ids."""
        t = tables(self.ngl)
        cc = self.cell_corner_coords()
        xy = np.einsum("kc,ecd->ekd", t.coo_op.H, cc)
        out = np.zeros((self.nnodes, 2))
        out[self.cell_nodes().reshape(-1)] = xy.reshape(-1, 2)   # later cells overwrite (INSERT_VALUES)
        return out

    def border_nodes(self, name):
        """getBorderNodes(name) (dmplex.py:160-166): nodes on the closure of the border's faces.
        "Face Sets": 1 bottom(down), 2 right, 3 top(up), 4 left (dmplex.py:38,168-171)."""
        nx, ny, m = self.nx, self.ny, self.ngl - 2
        k = np.arange(m)
        if name in ("down", "up"):
            vy = 0 if name == "down" else ny
            verts = self.off_vert + vy * (nx + 1) + np.arange(nx + 1)
            ed = self.off_edge + ((vy * nx + np.arange(nx))[:, None] * m + k).reshape(-1)
        elif name in ("left", "right"):
            vx = 0 if name == "left" else nx
            verts = self.off_vert + np.arange(ny + 1) * (nx + 1) + vx
            ed = self.off_edge + ((self.nxedge + vx * ny + np.arange(ny))[:, None] * m + k).reshape(-1)
        else:
            raise KeyError(name)
        return np.unique(np.concatenate([verts, ed])).astype(np.int64)

    def boundary_nodes(self):
        return np.unique(np.concatenate([self.border_nodes(b) for b in ("down", "right", "up", "left")]))


# --------------------------------------------------------------------------------------------------------
# 8f-2  unstructured quadrilateral mesh (gmsh): the numbering rule of pynama_b200/domain/quadmesh.py restated
# (DMPlex createFromCellList + interpolation + setFemIndexing, src/domain/dmplex.py:22-23,42-61; PETSc-internal and
# un-pinned by any reference test: "parity unpinned")
# --------------------------------------------------------------------------------------------------------
class QuadMesh:
    """points [nv,2], quads [nc,4] (0-based, any orientation).  Same surface as BoxMesh for the assembly functions."""

    def __init__(self, points, quads, ngl):
        self.points = np.asarray(points, float)
        q = np.array(quads, dtype=np.int64)
        for c in range(len(q)):                      # counter-clockwise cells
            p = self.points[q[c]]
            if (p[1, 0] - p[0, 0]) * (p[2, 1] - p[0, 1]) - (p[1, 1] - p[0, 1]) * (p[2, 0] - p[0, 0]) < 0:
                q[c] = q[c][::-1]
        self.quads, self.ngl = q, int(ngl)
        self.ncell, self.nvert = len(q), len(self.points)
        edge_id, self.edges, self.cell_edges, self.cell_rev, count = {}, [], [], [], []
        for c in range(self.ncell):                  # edges in order of first appearance: sides v0v1, v1v2, v2v3, v3v0
            ce, cr = [], []
            for i in range(4):
                a, b = int(q[c, i]), int(q[c, (i + 1) % 4])
                key = (min(a, b), max(a, b))
                if key not in edge_id:
                    edge_id[key] = len(self.edges)
                    self.edges.append((a, b))
                    count.append(0)
                e = edge_id[key]
                count[e] += 1
                ce.append(e)
                cr.append(self.edges[e][0] != a)
            self.cell_edges.append(ce)
            self.cell_rev.append(cr)
        self.edge_count = np.array(count)
        m = self.ngl - 2
        self.off_vert = self.ncell * m * m
        self.off_edge = self.off_vert + self.nvert
        self.nnodes = self.off_edge + len(self.edges) * m

    def cell_corner_coords(self, cells=None):
        cells = np.arange(self.ncell) if cells is None else np.asarray(cells)
        return self.points[self.quads[cells]]

    def cell_nodes(self, cells=None):
        cells = np.arange(self.ncell) if cells is None else np.asarray(cells)
        m = self.ngl - 2
        out = []
        for c in cells:
            row = [self.off_vert + int(v) for v in self.quads[c]]
            for i in range(4):
                ks = range(m - 1, -1, -1) if self.cell_rev[c][i] else range(m)
                row += [self.off_edge + self.cell_edges[c][i] * m + k for k in ks]
            row += [int(c) * m * m + k for k in range(m * m)]
            out.append(row)
        return np.array(out, dtype=np.int64)

    def node_coords(self):
        t = tables(self.ngl)
        xy = np.einsum("kc,ecd->ekd", t.coo_op.H, self.cell_corner_coords())
        out = np.zeros((self.nnodes, 2))
        out[self.cell_nodes().reshape(-1)] = xy.reshape(-1, 2)
        return out

    def boundary_nodes(self):
        m = self.ngl - 2
        nodes = []
        for e in np.flatnonzero(self.edge_count == 1):
            a, b = self.edges[e]
            nodes += [self.off_vert + a, self.off_vert + b] + [self.off_edge + e * m + k for k in range(m)]
        return np.unique(np.array(nodes, dtype=np.int64))


# --------------------------------------------------------------------------------------------------------
# deterministic COO -> CSR (PETSc MatSetValues(ADD_VALUES)+MatAssembly semantics, serial)
# --------------------------------------------------------------------------------------------------------
def coo_to_csr(rows, cols, vals, shape):
    """Keep every inserted (row,col) -- explicit zeros included -- sort columns ascending inside each row and add
    duplicates in insertion order (= ascending cell order of the reference loops)."""
    rows = np.asarray(rows, np.int64)
    cols = np.asarray(cols, np.int64)
    vals = np.asarray(vals, float)
    key = rows * shape[1] + cols
    order = np.argsort(key, kind="stable")
    key, vals = key[order], vals[order]
    first = np.ones(len(key), bool)
    first[1:] = key[1:] != key[:-1]
    seg = np.cumsum(first) - 1
    start = np.flatnonzero(first)
    rank = np.arange(len(key)) - start[seg]
    data = np.zeros(len(start))
    for r in range(int(rank.max()) + 1 if len(rank) else 0):   # explicit left-to-right summation order
        sel = rank == r
        data[seg[sel]] += vals[sel]
    ukey = key[start]
    urow, ucol = ukey // shape[1], ukey % shape[1]
    indptr = np.zeros(shape[0] + 1, np.int64)
    np.add.at(indptr, urow + 1, 1)
    indptr = np.cumsum(indptr)
    return sp.csr_matrix((data, ucol.astype(np.int32), indptr), shape=shape)


def _block_coo(rdofs, cdofs, E, rmask, cmask, sign=1.0):
    """COO triplets of E[e][rmask x cmask] scattered to (rdofs, cdofs) for all elements, in cell order."""
    ne, nr = rdofs.shape
    nc = cdofs.shape[1]
    R = np.broadcast_to(rdofs[:, :, None], (ne, nr, nc))
    C = np.broadcast_to(cdofs[:, None, :], (ne, nr, nc))
    M = rmask[:, :, None] & cmask[:, None, :]
    Eb = np.broadcast_to(E, (ne, nr, nc))
    return R[M], C[M], sign * Eb[M]


def _expand(nodes, ncomp):
    return (nodes[:, :, None] * ncomp + np.arange(ncomp)).reshape(nodes.shape[0], -1)


# --------------------------------------------------------------------------------------------------------
# A1  FreeSlip.buildKLEMats                                        src/cases/base_problem.py:499-552
# --------------------------------------------------------------------------------------------------------
def assemble_freeslip(mesh, per_cell=True):
    """K, Krhs, Rw for the FreeSlip (all-Dirichlet-boundary) path.  per_cell=True integrates every cell
    (base_problem.py:505-506)."""
    nodes = mesh.cell_nodes()
    cc = mesh.cell_corner_coords()
    Ke, Rwe, _ = elem_kle_matrices(mesh.ngl, cc if per_cell else cc[:1])
    N = mesh.nnodes
    bc = np.zeros(N, bool)
    bc[mesh.boundary_nodes()] = True
    vdofs, wdofs = _expand(nodes, 2), nodes
    setm = np.repeat(bc[nodes], 2, axis=1)          # dofSetFSNS: both components of boundary nodes (:516-520)
    free = ~setm
    allw = np.ones_like(nodes, bool)
    dir_dofs = np.flatnonzero(np.repeat(bc, 2))
    # K[free,free] += K_e ; explicit 0 on the diagonal of constrained dofs per cell (:540-544) ; +1 (:549)
    r, c, v = _block_coo(vdofs, vdofs, Ke, free, free)
    dz = vdofs[setm]
    K = coo_to_csr(np.concatenate([r, dz, dir_dofs]), np.concatenate([c, dz, dir_dofs]),
                   np.concatenate([v, np.zeros(len(dz)), np.ones(len(dir_dofs))]), (2 * N, 2 * N))
    r, c, v = _block_coo(vdofs, vdofs, Ke, free, setm, sign=-1.0)          # (:531-533)
    Krhs = coo_to_csr(np.concatenate([r, dir_dofs]), np.concatenate([c, dir_dofs]),
                      np.concatenate([v, np.ones(len(dir_dofs))]), (2 * N, 2 * N))
    r, c, v = _block_coo(vdofs, wdofs, Rwe, free, allw)                    # (:546-547)
    Rw = coo_to_csr(r, c, v, (2 * N, N))
    return dict(K=K, Krhs=Krhs, Rw=Rw, bc_nodes=np.flatnonzero(bc))


# --------------------------------------------------------------------------------------------------------
# A2  NoSlipFreeSlip.buildKLEMats (cavity: all four walls no-slip)   src/cases/base_problem.py:329-454
# --------------------------------------------------------------------------------------------------------
def dof_classes_noslip(mesh):
    """Per velocity dof: 0 free, 1 tangential on a no-slip wall (free in the FS solve, `dofFreeFSSetNS`),
    2 set in both solves (`dofSetFSNS`: wall-normal component; both components at corners)
    (base_problem.py:359-369, dmplex.py:245-261, common/nswalls.py:249-268)."""
    cls = np.zeros((mesh.nnodes, 2), np.int8)
    normal = {"down": 1, "up": 1, "left": 0, "right": 0}
    is_set = np.zeros((mesh.nnodes, 2), bool)
    is_tan = np.zeros((mesh.nnodes, 2), bool)
    for name, nrm in normal.items():
        nd = mesh.border_nodes(name)
        is_set[nd, nrm] = True
        is_tan[nd, 1 - nrm] = True
    is_tan &= ~is_set
    cls[is_tan] = 1
    cls[is_set] = 2
    return cls.reshape(-1)


def assemble_noslip(mesh):
    """The eight matrices of MatNS for the lid-driven cavity; element matrices of cell 0 only (:333-334)."""
    nodes = mesh.cell_nodes()
    Ke, Rwe, Rde = elem_kle_matrices(mesh.ngl, mesh.cell_corner_coords([0]))
    N = mesh.nnodes
    cls = dof_classes_noslip(mesh)
    vdofs = _expand(nodes, 2)
    c_ = cls[vdofs]
    free, tan, st = c_ == 0, c_ == 1, c_ == 2
    beset = tan | st
    allw = np.ones_like(nodes, bool)
    touched = beset.any(axis=1)                      # cells with nodeBCintersectNS (:387)
    cat = np.concatenate

    def build(parts, shape):
        r = cat([p[0] for p in parts]); c = cat([p[1] for p in parts]); v = cat([p[2] for p in parts])
        return coo_to_csr(r, c, v, shape)

    def diag(mask, val):
        d = vdofs[mask]
        return d, d, np.full(len(d), val)

    def gdiag(dofs, val):
        return dofs, dofs, np.full(len(dofs), val)

    ind2one = np.flatnonzero(cls > 0)                # indices2one = all gldof2beSet (:391)
    ind2onefs = np.flatnonzero(cls == 1)             # (:410)
    sq, rc = (2 * N, 2 * N), (2 * N, N)
    K = build([_block_coo(vdofs, vdofs, Ke, free, free), diag(beset, 0.0), gdiag(ind2one, 1.0)], sq)
    Krhs = build([_block_coo(vdofs, vdofs, Ke, free, beset, -1.0), diag(beset, 0.0), gdiag(ind2one, 1.0)], sq)
    Kfs = build([_block_coo(vdofs, vdofs, Ke, tan, free), _block_coo(vdofs, vdofs, Ke, free, tan),
                 _block_coo(vdofs, vdofs, Ke, tan, tan), gdiag(ind2onefs, -1.0)], sq)
    Rw = build([_block_coo(vdofs, nodes, Rwe, free, allw)], rc)
    Rd = build([_block_coo(vdofs, nodes, Rde, free, allw)], rc)
    Rwfs = build([_block_coo(vdofs, nodes, Rwe, tan, allw)], rc)
    Rdfs = build([_block_coo(vdofs, nodes, Rde, tan, allw)], rc)
    # Krhsfs: -K_e[tan|free, set]; explicit 0 then INSERT 1 on the diagonal of set dofs (:417-424,449-451)
    Krhsfs = build([_block_coo(vdofs, vdofs, Ke, tan, st, -1.0), _block_coo(vdofs, vdofs, Ke, free, st, -1.0),
                    gdiag(np.flatnonzero(cls == 2), 1.0)], sq)
    del touched
    return dict(K=K, Krhs=Krhs, Kfs=Kfs, Rw=Rw, Rd=Rd, Rwfs=Rwfs, Rdfs=Rdfs, Krhsfs=Krhsfs, cls=cls)


def mat_add(A, B):
    """`A + B` of petsc4py (duplicate + MatAXPY, DIFFERENT_NONZERO_PATTERN): union pattern, a_ij + 1.0*b_ij
    (base_problem.py:318)."""
    A = A.tocoo(); B = B.tocoo()
    r = np.concatenate([A.row, B.row]); c = np.concatenate([A.col, B.col]); v = np.concatenate([A.data, B.data])
    return coo_to_csr(r, c, v, A.shape)


# --------------------------------------------------------------------------------------------------------
# A3  buildOperators / Operators.assembleAll      base_problem.py:132-140, mat_generator.py:157-190
# --------------------------------------------------------------------------------------------------------
def assemble_operators(mesh, per_cell=False):
    """per_cell=False: cell 0's element operators for every cell, as the reference does (base_problem.py:133-137);
    per_cell=True: every cell's own -- what a non-uniform mesh needs."""
    nodes = mesh.cell_nodes()
    cc = mesh.cell_corner_coords() if per_cell else mesh.cell_corner_coords([0])
    SrTe, Dive, Curle, We = elem_kle_operators(mesh.ngl, cc)
    N = mesh.nnodes
    v, w, s = _expand(nodes, 2), nodes, _expand(nodes, 3)
    tv, tw, ts = np.ones_like(v, bool), np.ones_like(w, bool), np.ones_like(s, bool)
    Curl = coo_to_csr(*_block_coo(w, v, Curle, tw, tv), (N, 2 * N))
    SrT = coo_to_csr(*_block_coo(s, v, SrTe, ts, tv), (3 * N, 2 * N))
    Div = coo_to_csr(*_block_coo(v, s, Dive, tv, ts), (2 * N, 3 * N))
    W = np.zeros(N)
    for e in range(nodes.shape[0]):                 # VecSetValues(ADD) in cell order (mat_generator.py:168-170)
        W[nodes[e]] += We[e if per_cell else 0]
    rW = 1.0 / W                                    # VecReciprocal (:175,180,185)

    def lscale(M, ncomp):
        M = M.copy()
        l = np.repeat(rW, ncomp)
        M.data = M.data * np.repeat(l, np.diff(M.indptr))   # MatDiagonalScale(L): a_ij <- l_i * a_ij
        return M
    return dict(Curl=lscale(Curl, 1), SrT=lscale(SrT, 3), DivSrT=lscale(Div, 2), W=W)


# --------------------------------------------------------------------------------------------------------
# S1  KSP: CG + Jacobi with PETSc semantics (restated from PETSc KSPCG / KSPConvergedDefault docs)
# --------------------------------------------------------------------------------------------------------
def ksp_cg_jacobi(A, b, rtol=1e-5, atol=1e-50, dtol=1e4, max_it=10000, jacobi=True, x0=None, pc=None):
    """Left-preconditioned CG, zero initial guess, monitored norm ||M^-1 r||_2, default convergence test
    rnorm <= max(rtol*rnorm0, atol) (src/solver/ksp_solver.py:9-19 with -ksp_type cg -pc_type jacobi).
    x0 (-ksp_initial_guess_nonzero): r0 = b - A x0 and, as in KSPConvergedDefault, rnorm0 = ||M^-1 b|| (the initial
    residual if b = 0).  Returns x, iterations, reason (2 rtol, 3 atol, -3 its, -4 dtol), residual history."""
    n = A.shape[0]
    dinv = 1.0 / A.diagonal() if jacobi else np.ones(n)
    apply_pc = pc if pc is not None else (lambda v: dinv * v)      # pc: callable z = M^-1 r (e.g. MultigridPC)
    if x0 is None:
        x = np.zeros(n)
        r = b.copy()
    else:
        x = np.array(x0, dtype=float)
        r = b - A @ x
    z = apply_pc(r)
    dp = math.sqrt(z @ z)
    hist = [dp]
    rnorm0 = dp
    if x0 is not None:
        zb = apply_pc(b)
        snorm = math.sqrt(zb @ zb)
        rnorm0 = snorm if snorm > 0.0 else dp
    ttol = max(rtol * rnorm0, atol)
    if dp <= ttol:
        return x, 0, (3 if dp <= atol else 2), hist
    beta = z @ r
    p = np.zeros(n)
    betaold = 1.0
    its, reason = 0, -3
    for i in range(max_it):
        its = i + 1
        if beta == 0.0:
            reason = 3
            break
        p = z.copy() if i == 0 else z + (beta / betaold) * p
        w = A @ p
        dpi = p @ w
        betaold = beta
        a = beta / dpi
        x = x + a * p
        r = r - a * w
        z = apply_pc(r)
        dp = math.sqrt(z @ z)
        hist.append(dp)
        if dp <= ttol:
            reason = 3 if dp <= atol else 2
            break
        if dp >= dtol * rnorm0:
            reason = -4
            break
        beta = z @ r
    return x, its, reason, hist


# --------------------------------------------------------------------------------------------------------
# 8f-1  geometric p/h-multigrid V-cycle as CG's preconditioner (`-pc_type mg`)
#
# Not a restatement of reference code (the reference's answer to the cost of the solve is PETSc's factor-once LU,
# src/solver/ksp_solver.py:13-16): this is the CPU statement of the algorithm in pynama_b200/solver/multigrid.py +
# csrc/mg.cu -- same level sequence, same rediscretised operators (this oracle's assemblies), same interpolation, same
# Chebyshev recurrence, exact coarse solve -- so that the device cycle can be checked against it.
# --------------------------------------------------------------------------------------------------------
def mg_coarsening_sequence(nx, ny, ngl, coarse_cells=1024):
    seq = [(nx, ny, ngl)]
    n = ngl
    while n > 2:
        n = (n - 1) // 2 + 1
        seq.append((nx, ny, n))
    while nx % 2 == 0 and ny % 2 == 0 and (nx // 2) * (ny // 2) >= 4 and nx * ny > coarse_cells:
        nx, ny = nx // 2, ny // 2
        seq.append((nx, ny, 2))
    while nx * ny > coarse_cells:
        nx, ny = max(2, (nx + 1) // 2), max(2, (ny + 1) // 2)
        seq.append((nx, ny, 2))
    return seq


def mg_node_interpolation(mc, mf):
    """P [Nf x Nc]: the coarse mesh's tensor-product Lagrange basis at the fine mesh's node coordinates."""
    xy = mf.node_coords()
    hx = (mc.upper[0] - mc.lower[0]) / mc.nx
    hy = (mc.upper[1] - mc.lower[1]) / mc.ny
    fx = np.clip(np.floor((xy[:, 0] - mc.lower[0]) / hx - 1e-9).astype(int), 0, mc.nx - 1)
    fy = np.clip(np.floor((xy[:, 1] - mc.lower[1]) / hy - 1e-9).astype(int), 0, mc.ny - 1)
    r = 1.0 - 2.0 * (xy[:, 0] - (mc.lower[0] + hx * fx)) / hx      # x = x0 + (1 - r) hx / 2
    s = 1.0 - 2.0 * (xy[:, 1] - (mc.lower[1] + hy * fy)) / hy
    n = mc.ngl
    nodes, _ = lobatto_points(n)

    def lag(pts):
        h = np.ones((len(pts), n))
        for a in range(n):
            for m in range(n):
                if m != a:
                    h[:, a] *= (pts - nodes[m]) / (nodes[a] - nodes[m])
        return h
    H = np.einsum("qa,qb->qab", lag(r), lag(s)).reshape(len(r), n * n)[:, local_order(n)]
    cn = mc.cell_nodes()[fy * mc.nx + fx]
    rows = np.repeat(np.arange(mf.nnodes), n * n)
    keep = np.abs(H.reshape(-1)) > 1e-12
    return sp.csr_matrix((H.reshape(-1)[keep], (rows[keep], cn.reshape(-1)[keep])), shape=(mf.nnodes, mc.nnodes))


class MultigridPC:
    """z = V(r): Chebyshev(degree) on [lmax/ratio, lmax] of D^-1 A before and after the coarse correction, coarse levels
    rediscretised, exact coarsest solve; Jacobi on the prescribed dofs (identity rows).  kind: "K" | "KFS" for the
    cavity (assemble_noslip), "freeslip" for the all-Dirichlet cases (assemble_freeslip)."""

    def __init__(self, nx, ny, ngl, kind="K", lower=(0, 0), upper=(1, 1), degree=3, ratio=30.0, coarse_cells=1024,
                 power_its=30, fine=None, lmax=None):
        """fine: use this matrix as the level-0 operator; lmax: per-level spectral bounds to use instead of this
        class's own power iteration (its start vector differs from the device's, so the bounds agree to ~1e-3 only)."""
        import scipy.sparse.linalg as spl
        self.degree, self.ratio = degree, ratio
        self.levels = []
        for i, (hx, hy, p) in enumerate(mg_coarsening_sequence(nx, ny, ngl, coarse_cells)):
            m = BoxMesh([hx, hy], lower, upper, p)
            if kind == "freeslip":
                A = fine if (i == 0 and fine is not None) else assemble_freeslip(m, per_cell=False)["K"]
                bc = np.zeros(m.nnodes, bool)
                bc[m.boundary_nodes()] = True
                free = ~np.repeat(bc, 2)
            else:
                r = assemble_noslip(m)
                A = r["K"] if kind == "K" else mat_add(r["K"], r["Kfs"])
                free = (r["cls"] == 0) if kind == "K" else (r["cls"] < 2)
                if i == 0 and fine is not None:
                    A = fine
            A = sp.csr_matrix(A)
            L = dict(m=m, A=A, free=free, dinv=np.where(free, 1.0 / A.diagonal(), 0.0))
            x = np.random.default_rng(0).standard_normal(A.shape[0]) * free
            lam = 1.0
            for _ in range(power_its):
                y = L["dinv"] * (A @ x)
                lam = np.linalg.norm(y) / np.linalg.norm(x)
                x = y / np.linalg.norm(y)
            L["lmax"] = 1.1 * lam if lmax is None else float(lmax[i])
            if self.levels:
                f = self.levels[-1]
                P = sp.kron(mg_node_interpolation(m, f["m"]), sp.identity(2)).tocsr()
                P = (sp.diags(f["free"].astype(float)) @ P @ sp.diags(free.astype(float))).tocsr()
                f["P"], f["R"] = P, P.T.tocsr()
            self.levels.append(L)
        c = self.levels[-1]
        idx = np.flatnonzero(c["free"])
        c["lu"] = (idx, spl.splu(sp.csc_matrix(c["A"][idx][:, idx])))
        self.jac = 1.0 / self.levels[0]["A"].diagonal()

    def _smooth(self, L, b, x, r, fresh, want_resid):
        lmax = L["lmax"]
        lmin = lmax / self.ratio
        theta, delta = 0.5 * (lmax + lmin), 0.5 * (lmax - lmin)
        sigma = theta / delta
        rho = 1.0 / sigma
        A, dinv = L["A"], L["dinv"]
        d = dinv * (b if fresh else r) / theta
        x = d.copy() if fresh else x + d
        if fresh:
            r = b.copy()
        for _ in range(1, self.degree):
            r = r - A @ d
            rho_new = 1.0 / (2.0 * sigma - rho)
            d = rho_new * rho * d + (2.0 * rho_new / delta) * (dinv * r)
            x = x + d
            rho = rho_new
        if want_resid:
            r = r - A @ d
        return x, r

    def _cycle(self, l, b):
        L = self.levels[l]
        if l == len(self.levels) - 1:
            idx, lu = L["lu"]
            x = np.zeros_like(b)
            x[idx] = lu.solve(b[idx])
            return x
        x, r = self._smooth(L, b, None, None, True, True)
        e = L["P"] @ self._cycle(l + 1, L["R"] @ r)
        x = x + e
        r = r - L["A"] @ e
        x, _ = self._smooth(L, b, x, r, False, False)
        return x

    def __call__(self, r):
        free = self.levels[0]["free"]
        z = self._cycle(0, r)
        return np.where(free, z, self.jac * r)


# --------------------------------------------------------------------------------------------------------
# R1-R3  velocity from vorticity, convective tensor, vorticity RHS   base_problem.py:212-252,321-327,479-481
# --------------------------------------------------------------------------------------------------------
def taylor_green(xy, nu, t):
    """custom_func.py:173-193 (Lx=Ly=Uref=1)."""
    x_, y_ = 2 * math.pi * xy[:, 0], 2 * math.pi * xy[:, 1]
    e = math.exp(-4 * math.pi ** 2 * nu * t * 2.0)
    vel = np.stack([np.cos(x_) * np.sin(y_) * e, -np.sin(x_) * np.cos(y_) * e], 1)
    vort = -2 * math.pi * 2.0 * np.cos(x_) * np.cos(y_) * e
    return vel, vort


def vtensv(vel):
    """computeVtensV (base_problem.py:234-252): [vx^2, vx*vy, vy^2] per node."""
    vx, vy = vel[0::2], vel[1::2]
    out = np.empty(3 * len(vx))
    out[0::3], out[1::3], out[2::3] = vx ** 2, vx * vy, vy ** 2
    return out


def eval_rhs_from_vel(ops, vel, rho, mu):
    """evalRHS after solveKLE (base_problem.py:217-232): Curl (1/rho) DivSrT (2 mu SrT v - rho VtensV)."""
    aux = ops["SrT"] @ vel
    aux = aux * (2.0 * mu)
    aux = aux + (-1.0 * rho) * vtensv(vel)
    rhs = ops["DivSrT"] @ aux
    rhs = rhs * (1 / rho)
    return ops["Curl"] @ rhs


# --------------------------------------------------------------------------------------------------------
# T1  Bogacki-Shampine 5(4) FSAL tableau (PETSc TSRK5BS; SURVEY 8a'' -- verified by order conditions)
# --------------------------------------------------------------------------------------------------------
def rk5bs_tableau():
    from fractions import Fraction as F
    A = [[],
         [F(1, 6)],
         [F(2, 27), F(4, 27)],
         [F(183, 1372), F(-162, 343), F(1053, 1372)],
         [F(68, 297), F(-4, 11), F(42, 143), F(1960, 3861)],
         [F(597, 22528), F(81, 352), F(63099, 585728), F(58653, 366080), F(4617, 20480)],
         [F(174197, 959244), F(-30942, 79937), F(8152137, 19744439), F(666106, 1039181), F(-29421, 29068),
          F(482048, 414219)],
         [F(587, 8064), F(0), F(4440339, 15491840), F(24353, 124800), F(387, 44800), F(2152, 5985),
          F(7267, 94080)]]
    b = A[7] + [F(0)]
    bh = [F(2479, 34992), F(0), F(123, 416), F(612941, 3411720), F(43, 1440), F(2272, 6561),
          F(79937, 1113912), F(3293, 556956)]
    c = [sum(row, F(0)) for row in A]
    Af = np.zeros((8, 8))
    for i, row in enumerate(A):
        Af[i, :len(row)] = [float(v) for v in row]
    return Af, np.array([float(v) for v in b]), np.array([float(v) for v in bh]), np.array([float(v) for v in c])


def rk5bs_step(f, t, y, dt, k1=None):
    """One explicit step; returns y5, embedded error y5-y4, and the FSAL stage k8 = f(t+dt, y5)."""
    A, b, bh, c = rk5bs_tableau()
    k = [f(t, y) if k1 is None else k1]
    for i in range(1, 8):
        yi = y.copy()
        for j in range(i):
            if A[i, j] != 0.0:
                yi = yi + (dt * A[i, j]) * k[j]
        k.append(f(t + c[i] * dt, yi))
    y5 = y.copy()
    err = np.zeros_like(y)
    for j in range(8):
        if b[j] != 0.0:
            y5 = y5 + (dt * b[j]) * k[j]
        err = err + (dt * (b[j] - bh[j])) * k[j]
    return y5, err, k[7]


# --------------------------------------------------------------------------------------------------------
# R1  case-level flows: boundary vectors and the velocity-from-vorticity solves
# --------------------------------------------------------------------------------------------------------
def direct_solve(A, b):
    """Stand-in for the reference's default `preonly` + `lu` (src/solver/ksp_solver.py:13-16): SuperLU."""
    import scipy.sparse.linalg as spla
    return spla.spsolve(A.tocsc(), b)


def cavity_wall_vectors(mesh, moving=(("up", (1.0, 0.0)),)):
    """Cavity.applyBoundaryConditions / applyBoundaryConditionsFS (src/cases/cavity.py:53-81) as data:
    returns vel_bc[2N] (zero except the moving walls' tangential component) and (dofs, values) assigned to velFS
    after the free-slip solve, in assignment order (moving walls first, then the static walls' tangential dofs)."""
    normal = {"down": 1, "up": 1, "left": 0, "right": 0}
    vel = np.zeros(2 * mesh.nnodes)
    fs_d, fs_v = [], []
    moving = dict(moving)
    for name, v in moving.items():
        tang = 1 - normal[name]
        d = mesh.border_nodes(name) * 2 + tang
        vel[d] = v[tang]
        fs_d.append(d)
        fs_v.append(np.full(len(d), float(v[tang])))
    for name in ("left", "right", "up", "down"):          # NoSlipWalls side order (common/nswalls.py:13)
        if name in moving:
            continue
        d = mesh.border_nodes(name) * 2 + (1 - normal[name])
        fs_d.append(d)
        fs_v.append(np.zeros(len(d)))
    return vel, np.concatenate(fs_d), np.concatenate(fs_v)


def solve_kle_noslip(mats, ops, vel_bc, fs_dofs, fs_vals, vort, solve=direct_solve, solve_fs=None, KFS=None):
    """NoSlipFreeSlip.solveKLE (src/cases/base_problem.py:321-327)."""
    KFS = mat_add(mats["K"], mats["Kfs"]) if KFS is None else KFS
    solve_fs = solve if solve_fs is None else solve_fs
    rhs1 = mats["Rw"] @ vort + mats["Rwfs"] @ vort + mats["Krhsfs"] @ vel_bc
    vel_fs = solve_fs(KFS, rhs1)
    vel_fs = vel_fs.copy()
    vel_fs[fs_dofs] = fs_vals
    vort2 = ops["Curl"] @ vel_fs
    vel = solve(mats["K"], mats["Rw"] @ vort2 + mats["Krhs"] @ vel_bc)
    return vel, vel_fs, vort2


def solve_kle_freeslip(mats, vel_bc, vort, solve=direct_solve):
    """FreeSlip.solveKLE (src/cases/base_problem.py:479-481)."""
    return solve(mats["K"], mats["Rw"] @ vort + mats["Krhs"] @ vel_bc)


def rk_fixed_steps(f, t0, y0, dt, nsteps, post=None):
    """nsteps of RK5BS with a fixed step and FSAL reuse (TSRK with -ts_adapt_type none)."""
    t, y, k1 = t0, y0.copy(), None
    for _ in range(nsteps):
        y, _, k1 = rk5bs_step(f, t, y, dt, k1)
        t += dt
        if post is not None:
            y = post(t, y)
    return t, y


def wrms(err, y, ynew, atol, rtol):
    """TSErrorWeightedNorm (NORM_2): sqrt(mean((e_i / (atol + rtol*max(|y_i|,|ynew_i|)))^2))."""
    tol = atol + rtol * np.maximum(np.abs(y), np.abs(ynew))
    return math.sqrt(np.mean((err / tol) ** 2))


def rk_adaptive(f, t0, y0, t_end, dt0=0.1, atol=1e-4, rtol=1e-4, max_steps=5000, safety=0.9, reject_safety=0.5,
                clip=(0.1, 10.0)):
    """TSRK 5bs + TSAdapt basic + MATCHSTEP restated (SURVEY 8a''): returns t, y, accepted step sizes, rejections."""
    def match(tn, h):
        hmax = t_end - tn
        if tn < t_end and t_end < tn + h * 1.01:
            return hmax
        if tn < t_end and t_end < tn + h * 2.0:
            return hmax / 2
        return h
    t, y, k1, dt = t0, y0.copy(), None, dt0
    if t + dt > t_end:
        dt = t_end - t
    hs, nrej, steps = [], 0, 0
    while t < t_end and steps < max_steps:
        prev_rejected = False
        while True:
            y5, err, k8 = rk5bs_step(f, t, y, dt, k1)
            k1 = f(t, y) if k1 is None else k1
            e = wrms(err, y5, y5 - err, atol, rtol)
            accept = e <= 1.0
            s = safety * (reject_safety if (prev_rejected and not accept) else 1.0)
            fac = clip[1] if e == 0.0 else min(max(s * e ** (-1.0 / 5.0), clip[0]), clip[1])
            if accept:
                break
            nrej += 1
            prev_rejected = True
            dt = match(t, dt * fac)
        hs.append(dt)
        t, y, k1 = t + dt, y5, k8
        steps += 1
        dt = match(t, dt * fac)
    return t, y, hs, nrej


# --------------------------------------------------------------------------------------------------------
# S1b  KSP: restarted GMRES with PETSc semantics (restated from the KSPGMRES documentation)
# --------------------------------------------------------------------------------------------------------
def ksp_gmres(A, b, rtol=1e-5, atol=1e-50, dtol=1e4, max_it=10000, restart=30, jacobi=True, x0=None):
    """Left-preconditioned GMRES(restart), zero initial guess, classical Gram-Schmidt without refinement, residual
    norm from the Givens recurrence, KSPConvergedDefault; the true preconditioned residual is formed and tested at the
    top of every cycle (src/solver/ksp_solver.py:9-19 with -ksp_type gmres).  x0 (-ksp_initial_guess_nonzero): first
    residual from x0, rtol measured against ||M^-1 b|| (KSPConvergedDefault).  Returns x, its, reason, history."""
    n = A.shape[0]
    dinv = 1.0 / A.diagonal() if jacobi else np.ones(n)
    x = np.zeros(n) if x0 is None else np.array(x0, dtype=float)
    its, reason, hist, m = 0, 0, [], restart
    rnorm0 = ttol = None
    while not reason:
        r = dinv * (b - A @ x) if (its or x0 is not None) else dinv * b
        beta = math.sqrt(r @ r)
        if rnorm0 is None:
            rnorm0 = beta
            if x0 is not None:
                zb = dinv * b
                snorm = math.sqrt(zb @ zb)
                rnorm0 = snorm if snorm > 0.0 else beta
            ttol = max(rtol * rnorm0, atol)
            hist.append(beta)
        if beta <= ttol:
            reason = 3 if beta <= atol else 2
            break
        if beta >= dtol * rnorm0:
            reason = -4
            break
        if its >= max_it:
            reason = -3
            break
        V = [r / beta]
        H = np.zeros((m + 1, m))
        cs, sn, g = np.zeros(m), np.zeros(m), np.zeros(m + 1)
        g[0] = beta
        k = 0
        for j in range(m):
            w = dinv * (A @ V[j])
            h = np.array([V[i] @ w for i in range(j + 1)])
            for i in range(j + 1):
                w = w - h[i] * V[i]
            hn = math.sqrt(w @ w)
            V.append(w / hn if hn > 0 else w)
            H[:j + 1, j], H[j + 1, j] = h, hn
            for i in range(j):
                t = cs[i] * H[i, j] + sn[i] * H[i + 1, j]
                H[i + 1, j] = -sn[i] * H[i, j] + cs[i] * H[i + 1, j]
                H[i, j] = t
            den = math.hypot(H[j, j], H[j + 1, j])
            cs[j], sn[j] = H[j, j] / den, H[j + 1, j] / den
            H[j, j], H[j + 1, j] = den, 0.0
            g[j + 1], g[j] = -sn[j] * g[j], cs[j] * g[j]
            rn = abs(g[j + 1])
            its += 1
            k = j + 1
            hist.append(rn)
            if rn <= ttol:
                reason = 3 if rn <= atol else 2
            elif hn == 0.0:
                reason = 2
            elif rn >= dtol * rnorm0:
                reason = -4
            elif its >= max_it:
                reason = -3
            if reason:
                break
        y = np.linalg.solve(np.triu(H[:k, :k]), g[:k]) if k else np.zeros(0)
        for i in range(k):
            x = x + y[i] * V[i]
    return x, its, reason, hist


# ======================================================================================================================
# Immersed-boundary velocity correction.  PINNED against the reference's own code: the delta functions,
# getDiracs and marker coordinates against the UNMODIFIED src/domain/immersed_body.py, and the lists, H, S, A, the
# corrected velocity and the virtual flux against the UNMODIFIED collectNodes / computeClose / computeDirac /
# buildIBMMatrix / computeVelocityCorrection of src/cases/immersed_boundary.py run on scipy stand-ins for PETSc's Mat and
# Vec (oracle/ref_shim.run_reference_ibm) -- bit for bit; golden vectors tests/golden/ibm_ref.npz from
# oracle/gen_golden_ibm.py, live comparison where /root/reference exists (tests/test_oracle_golden.py).  What the
# stand-ins assume of PETSc: INSERT_VALUES setValues, sorted AIJ rows, MatMatMult accumulating along the left row.
# ======================================================================================================================
def ibm_phi(kind, r):
    """Discrete delta functions of src/domain/immersed_body.py:398-421: 'four' (fourGrid, the default of
    ImmersedBody.__init__, :136), 'three' (threeGrid), 'linear'.  r = |distance| / h."""
    if kind == "four":
        if r <= 1:
            return (3 - 2 * r + math.sqrt(1 + 4 * r - 4 * r ** 2)) / 8
        elif r <= 2:
            return (5 - 2 * r - math.sqrt(-7 + 12 * r - 4 * r ** 2)) / 8
        return 0.0
    if kind == "three":
        if r <= 0.5:
            return (1 + math.sqrt(-3 * r ** 2 + 1)) / 3
        elif r <= 1.5:
            return (5 - 3 * r - math.sqrt(-3 * (1 - r) ** 2 + 1)) / 6
        return 0.0
    if kind == "linear":
        return (1 - r) if r < 1 else 0.0
    raise ValueError(kind)


def ibm_dirac(kind, dist, h):
    """ImmersedBody.getDiracs (immersed_body.py:259-264): per direction multiply by phi, then divide by h."""
    d = 1.0
    for r in dist:
        d *= ibm_phi(kind, abs(r) / h)
        d /= h
    return d


def ibm_close(mode, dist, dl):
    """computeClose of the static (immersed_boundary.py:325-336, mode 0) and dynamic (:391-402, mode 1) cases."""
    ex, ey = dist[0] / dl, dist[1] / dl
    if mode == 0:
        return abs(ex) < 3 and abs(ey) < 3
    return abs(ex) < 2 and abs(ey * dl) < (1 + 2 * dl)


def ibm_circle(radius, h, center=(0.0, 0.0)):
    """Circle.generateBody (immersed_body.py:371-389) + the centre displacement of getNodeCoordinates (:240-243):
    ((npoints, 2) coordinates, dl)."""
    long_total = 2 * math.pi * radius
    points = math.ceil(long_total / h)
    start = math.pi / 1000
    angles = np.linspace(0, 2 * math.pi, points, endpoint=False)
    xy = np.stack([radius * np.cos(angles + start), radius * np.sin(angles + start)], 1)
    dl = 2 * math.pi * radius / points
    return xy + np.asarray(center, dtype=float), dl


def ibm_line(dl, length=2.0, center=(0.0, 0.0)):
    """Line.generateBody (immersed_body.py:294-307)."""
    div = math.ceil(length / dl)
    xy = np.stack([np.linspace(0, length, div), np.zeros(div)], 1)
    return xy + np.asarray(center, dtype=float), dl


def ibm_open_box(dl, length=1.0, center=(0.0, 0.0)):
    """OpenBox.generateBody (immersed_body.py:316-345): the diamond (0,L) -> (-L,0) -> (0,-L) -> (L,0)."""
    div = math.ceil(math.sqrt(2) / dl)
    pts = [(0, length), (-length, 0), (0, -length), (length, 0), (0, length)]
    xs = np.concatenate([np.linspace(pts[i][0], pts[i + 1][0], div, endpoint=False) for i in range(4)])
    ys = np.concatenate([np.linspace(pts[i][1], pts[i + 1][1], div, endpoint=False) for i in range(4)])
    return np.stack([xs, ys], 1) + np.asarray(center, dtype=float), dl


def ibm_collect(lagxy, cand, nodexy, dl, h, mode=0, kind="four"):
    """collectNodes + computeDirac (immersed_boundary.py:258-277, 318-323): per Lagrangian node the close Euler nodes
    (ascending id) and their delta values."""
    out = []
    cand = np.asarray(cand)
    for X in lagxy:
        nodes, dirac = [], []
        for e in cand:
            dist = nodexy[e] - X
            if ibm_close(mode, dist, dl):
                nodes.append(int(e))
                dirac.append(ibm_dirac(kind, dist, h))
        if not nodes:
            raise Exception("Lag Node without Euler")
        out.append((np.array(nodes, dtype=np.int64), np.array(dirac)))
    return out


def ibm_matrices(collected, n_euler_nodes, dl, h, dim=2):
    """buildIBMMatrix (immersed_boundary.py:222-251): H (2 rows per Lagrangian node, columns 2 per Euler node),
    S = H^T dl h, H scaled by h^2 afterwards, A = H S.  A is accumulated over the common Euler nodes in ascending
    order, the order of a sorted-row sparse product."""
    nlag = len(collected)
    rows, cols, vals = [], [], []
    for l, (nodes, dirac) in enumerate(collected):
        for dof in range(dim):
            rows += [l * dim + dof] * len(nodes)
            cols += list(nodes * dim + dof)
            vals += list(dirac)
    H0 = sp.csr_matrix((vals, (rows, cols)), shape=(nlag * dim, n_euler_nodes * dim))
    H0.sort_indices()
    S = sp.csr_matrix(H0.T)
    S.sort_indices()
    S.data = S.data * (dl * h)
    H = H0.copy()
    H.data = H.data * (h ** 2)
    a, b = h ** 2, dl * h
    ar, ac, av = [], [], []
    for i, (ni, di) in enumerate(collected):
        for j, (nj, dj) in enumerate(collected):
            common, pi, pj = np.intersect1d(ni, nj, assume_unique=True, return_indices=True)
            if len(common) == 0:
                continue
            s = 0.0
            for p, q in zip(pi, pj):
                s = s + (di[p] * a) * (dj[q] * b)
            for dof in range(dim):
                ar.append(i * dim + dof)
                ac.append(j * dim + dof)
                av.append(s)
    A = sp.csr_matrix((av, (ar, ac)), shape=(nlag * dim, nlag * dim))
    A.sort_indices()
    return H, S, A


def ibm_velocity_correction(H, S, A, vel, body_vel):
    """computeVelocityCorrection (immersed_boundary.py:205-211): rhs = H v - V_body; A q = -rhs; v += S q.  Returns
    (corrected velocity, virtual flux q)."""
    import scipy.sparse.linalg as spla
    rhs = H @ vel - body_vel
    q = spla.spsolve(sp.csc_matrix(A), -rhs)
    return vel + S @ q, q
