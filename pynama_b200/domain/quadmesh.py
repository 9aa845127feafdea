"""Unstructured quadrilateral meshes: a gmsh reader and the domain object built from it.

Reference: `DMPlexDom(fileName=...)` -> `PETSc.DMPlex.createFromFile` + interpolation (src/domain/dmplex.py:22-23); the
reference's case layer never reaches it (`readDomainData` raises "No Gmsh Implemented", src/cases/base_problem.py:107),
so this is the "next" row of the scope table rather than a path with reference results.

`read_msh` parses the gmsh v2 ASCII format (4-node quadrilaterals = element type 3; lines and points are ignored: the
boundary is found topologically, like `markBoundaryFaces`).  `QuadMeshDom` numbers the mesh the way DMPlex does after
`createFromCellList` + interpolation -- cells, then vertices in file order, then edges in order of first appearance
walking the cells and their sides v0v1, v1v2, v2v3, v3v0 -- and the nodes the way `setFemIndexing` does (section offsets
in point order: cell interiors, vertices, edges; src/domain/dmplex.py:42-61, src/domain/indices.py:66-118).  PETSc's
numbering of interpolated edges is PETSc-internal and un-pinned by any reference test ("parity unpinned"); the oracle
restates exactly this rule (oracle/kle_oracle.py QuadMesh).

Everything downstream is the general path the box mesh also takes: `pyn_topo_create` on the element-to-node table,
element matrices / operators integrated for EVERY cell from its four corners (straight-sided quads, bilinear map), the
deterministic gather assembly.  The uniform-mesh shortcuts (row templates, element operator) do not apply.
"""
import logging

import numpy as np

from ..elements.spectral import localNodeOrder
from ..petsc_like import Vec
from .indices import IndicesManager


def read_msh(path):
    """(points [nv,2] float64, quads [nc,4] int64 0-based, counter-clockwise as stored) from a gmsh v2 ASCII file."""
    with open(path) as f:
        lines = [ln.strip() for ln in f]
    def section(name):
        try:
            i = lines.index(f"${name}")
            j = lines.index(f"$End{name}")
        except ValueError:
            raise Exception(f"{path}: section ${name} not found (gmsh v2 ASCII expected)")
        return lines[i + 1:j]
    fmt = section("MeshFormat")[0].split()
    if not fmt[0].startswith("2") or int(fmt[1]) != 0:
        raise Exception(f"{path}: gmsh format {fmt[0]} / binary not supported (v2 ASCII expected)")
    nodes = section("Nodes")
    nv = int(nodes[0])
    ids = np.empty(nv, dtype=np.int64)
    pts = np.empty((nv, 3))
    for k, ln in enumerate(nodes[1:nv + 1]):
        t = ln.split()
        ids[k] = int(t[0])
        pts[k] = [float(t[1]), float(t[2]), float(t[3])]
    remap = {int(g): k for k, g in enumerate(ids)}
    quads = []
    el = section("Elements")
    for ln in el[1:int(el[0]) + 1]:
        t = ln.split()
        if int(t[1]) == 3:                        # 4-node quadrangle
            ntags = int(t[2])
            quads.append([remap[int(v)] for v in t[3 + ntags:3 + ntags + 4]])
    if not quads:
        raise Exception(f"{path}: no quadrilateral elements")
    quads = np.array(quads, dtype=np.int64)
    used = np.unique(quads)                       # drop geometry-only points, keep file order
    new = -np.ones(nv, dtype=np.int64)
    new[used] = np.arange(len(used))
    return pts[used, :2].copy(), new[quads]


def interpolate_edges(quads):
    """Edges in order of first appearance (cells in order, sides v0v1, v1v2, v2v3, v3v0).  Returns
    (edges [ne,2] in their first-seen direction, cell_edges [nc,4], cell_edge_reversed [nc,4] bool)."""
    nc = len(quads)
    a = quads.reshape(-1)
    b = np.roll(quads, -1, axis=1).reshape(-1)                  # side i: v_i -> v_{i+1}
    lo, hi = np.minimum(a, b), np.maximum(a, b)
    key = lo * (int(quads.max()) + 1) + hi
    uniq, first, inv = np.unique(key, return_index=True, return_inverse=True)
    order = np.argsort(first, kind="stable")                    # unique keys in order of first appearance
    rank = np.empty(len(uniq), dtype=np.int64)
    rank[order] = np.arange(len(uniq))
    eid = rank[inv]
    fa, fb = a[first[order]], b[first[order]]
    edges = np.stack([fa, fb], 1)
    reversed_ = a != edges[eid, 0]
    counts = np.bincount(eid, minlength=len(uniq))
    return edges, eid.reshape(nc, 4), reversed_.reshape(nc, 4), counts


class QuadMeshDom(object):
    """Domain of an unstructured mesh of straight-sided quadrilaterals, with the slice of the DMPlexDom surface the
    case layer uses."""

    def __init__(self, points=None, quads=None, fileName=None, comm=None, **kwargs):
        if fileName is not None:
            points, quads = read_msh(fileName)
        self.points = np.ascontiguousarray(points, dtype=np.float64)
        self.quads = np.ascontiguousarray(quads, dtype=np.int64)
        # counter-clockwise cells (positive Jacobian of the bilinear map), like DMPlex after orientation
        p = self.points[self.quads]
        area2 = ((p[:, 1, 0] - p[:, 0, 0]) * (p[:, 2, 1] - p[:, 0, 1]) - (p[:, 1, 1] - p[:, 0, 1]) * (p[:, 2, 0] - p[:, 0, 0]))
        flip = area2 < 0
        self.quads[flip] = self.quads[flip][:, ::-1]
        self.dim, self.dim_w, self.dim_s = 2, 1, 3
        self.comm, self.partition = comm, None
        self.logger = logging.getLogger("[0] Class")
        self.ncell, self.nvert = len(self.quads), len(self.points)
        self.edges, self.cell_edges, self.cell_edge_rev, self.edge_cells = interpolate_edges(self.quads)
        self.nedge = len(self.edges)
        self.cellStart, self.cellEnd = 0, self.ncell
        self.namingConvention = ["boundary"]
        self.lower = [float(self.points[:, 0].min()), float(self.points[:, 1].min())]
        self.upper = [float(self.points[:, 0].max()), float(self.points[:, 1].max())]
        self.nx = self.ny = self.ny_global = None          # not a box mesh
        self.fy0 = 0
        self._dev = {}
        self.indicesManager = None
        self.fullCoordVec = None

    # -- numbering --------------------------------------------------------------------------------------------
    def getDimension(self):
        return self.dim

    def setFemIndexing(self, ngl):
        self.ngl = int(ngl)
        m = self.ngl - 2
        self.indicesManager = IndicesManager(self.dim, ngl, self.comm)
        self.off_vert = self.ncell * m * m
        self.off_edge = self.off_vert + self.nvert
        self.nnodes = self.off_edge + self.nedge * m
        self._dev = {}

    def getNGL(self):
        return self.ngl

    def getTotalNodes(self):
        return self.nnodes

    def cellNodesArray(self):
        """(nc, nn) global node ids in element-local order: 4 vertices, the 4 sides' interior nodes along the cell's
        own traversal (reversed when the stored edge runs the other way), the cell interior."""
        m = self.ngl - 2
        k = np.arange(m)
        cols = [self.off_vert + self.quads]
        for i in range(4):
            base = self.off_edge + self.cell_edges[:, i][:, None] * m
            cols.append(np.where(self.cell_edge_rev[:, i][:, None], base + k[::-1], base + k))
        cols.append((np.arange(self.ncell) * m * m)[:, None] + np.arange(m * m))
        return np.concatenate(cols, axis=1).astype(np.int64)

    def cellCornersArray(self):
        return self.points[self.quads]                       # (nc, 4, 2), closure order = the cell's vertex order

    def getBoundingBox(self):
        return list(self.lower), list(self.upper)

    def getTotalCells(self):
        return self.ncell

    def deviceCellNodes(self):
        from .. import backend as be
        if "e2n" not in self._dev:
            self._dev["e2n"] = be.to_device(self.cellNodesArray().astype(np.int32))
        return self._dev["e2n"]

    def deviceCellCorners(self, cell0=0, ncells=None):
        from .. import backend as be
        ncells = self.ncell - cell0 if ncells is None else ncells
        return be.to_device(self.cellCornersArray()[cell0:cell0 + ncells])

    def deviceTemplateCorners(self):
        return self.deviceCellCorners(0, 1)

    def templateCellCorners(self):
        return self.cellCornersArray()[:1].copy()

    def getCellCornersCoords(self, cell):
        return self.cellCornersArray()[cell].reshape(-1).copy()

    def getGlobalNodesFromCell(self, cell, shared=True):
        return [int(x) for x in self.cellNodesArray()[cell]]

    def getVelocityIndex(self, nodes):
        return [int(n) * self.dim + d for n in nodes for d in range(self.dim)]

    def getVorticityIndex(self, nodes):
        return [int(n) for n in nodes]

    def topology(self):
        from .. import backend as be
        if "topo" not in self._dev:
            self._dev["topo"] = be.Topology(self.deviceCellNodes(), self.nnodes)
        return self._dev["topo"]

    def isUniform(self):
        return False

    # -- coordinates ------------------------------------------------------------------------------------------
    def computeFullCoordinates(self, spElem):
        """x(node) = HCooOp[k] @ corners of any cell holding the node (src/domain/dmplex.py:66-95)."""
        from .. import backend as be
        H = np.array(spElem.HCooOp)                           # (nn, 4): bilinear shape functions at the nodal points
        xy_e = np.einsum("kc,ecd->ekd", H, self.cellCornersArray())
        xy = np.zeros((self.nnodes, 2))
        xy[self.cellNodesArray().reshape(-1)] = xy_e.reshape(-1, 2)      # later cells overwrite (INSERT_VALUES)
        self._dev["xy"] = be.to_device(xy)
        self.fullCoordVec = Vec(data=self._dev["xy"].view(-1), name="NodeCoordinates")
        self.nodes = list(range(self.nnodes))

    def deviceNodeCoords(self):
        return self._dev["xy"]

    def ownedNodeMask(self):
        return np.ones(self.nnodes, dtype=bool)

    # -- boundary ---------------------------------------------------------------------------------------------
    def setLabelToBorders(self):
        pass

    def getBordersNames(self):
        return self.namingConvention

    def borderNodesArray(self, name="boundary"):
        """Nodes on the closure of the edges that belong to one cell only (markBoundaryFaces)."""
        m = self.ngl - 2
        be_ = np.flatnonzero(self.edge_cells == 1)
        verts = self.off_vert + np.unique(self.edges[be_].reshape(-1))
        ed = (self.off_edge + be_[:, None] * m + np.arange(m)).reshape(-1)
        return np.unique(np.concatenate([verts, ed])).astype(np.int64)

    def getBorderNodes(self, name="boundary"):
        return [int(x) for x in self.borderNodesArray(name)]

    def getBordersNodes(self):
        return set(self.getBorderNodes())

    def getNodesFromLabel(self, label, shared=False):
        if label == "External Boundary":
            return self.getBordersNodes()
        return set()

    def getGlobalIndicesDirichlet(self):
        return self.indicesManager.getDirichletNodes()

    def getGlobalIndicesNoSlip(self):
        return self.indicesManager.getNoSlipNodes()

    def setBoundaryCondition(self, freeSlipFaces=[], noSlipFaces=[]):
        if len(freeSlipFaces) or len(noSlipFaces):
            raise Exception("named borders exist on box meshes only: an unstructured mesh takes Dirichlet data on its "
                            "whole boundary")
        self.indicesManager.setDirichletNodes(self.getBordersNodes())

    def getMatIndices(self):
        return 0, self.nnodes, self.topology(), np.empty(0), None, None
