"""Structured box mesh with the serial DMPlex numbering of the reference, without PETSc.

Mirrors `DMPlexDom` (src/domain/dmplex.py): same constructor keywords, attributes (`cellStart`, `cellEnd`,
`fullCoordVec`, `indicesManager`, `namingConvention`) and query methods.  The point numbering restates
`DMPlexCreateBoxMesh(simplex=False)` + `PetscSection` offsets in point order:

    cells      c = fy*nx + fx                        points [0, nc)
    vertices   v = vy*(nx+1) + vx                    points [nc, nc+nv)
    x-edges    e = vy*nx + ex        (direction +x)  points [nc+nv, ...)
    y-edges    e = vx*ny + ey        (direction +y)  after the x-edges
    cell cone  = [bottom, right, top, left], top and left traversed backwards
    node ids   = cell interiors, then vertices, then edge interiors (section offsets in point order)

Everything per-cell that the reference does in Python loops (`getGlobalNodesFromCell`, `computeFullCoordinates`,
`getMatIndices`) is also available as whole-mesh device tables (`deviceCellNodes`, `deviceCellCorners`,
`deviceNodeCoords`, `topology`) built by CUDA kernels (csrc/topo.cu).
"""
import logging
from math import floor

import numpy as np

from ..petsc_like import Vec
from .indices import IndicesManager


class DMPlexDom(object):
    def __new__(cls, **kwargs):
        if 'boxMesh' not in kwargs and 'fileName' in kwargs:       # dmplex.py:22-23: createFromFile
            from .quadmesh import QuadMeshDom
            return QuadMeshDom(**kwargs)
        return super().__new__(cls)

    def __init__(self, **kwargs):
        if 'boxMesh' not in kwargs:
            raise Exception("DMPlexDom needs boxMesh={...} or fileName=<gmsh v2 ASCII file>")
        meshData = kwargs['boxMesh']
        lower = kwargs['lower'] if 'lower' in kwargs else meshData.get('lower')
        upper = kwargs['upper'] if 'upper' in kwargs else meshData.get('upper')
        faces = kwargs['nelem'] if 'nelem' in kwargs else meshData.get('nelem')
        lower = [eval(v) if isinstance(v, str) else v for v in lower]      # yaml may carry "2*pi" (dmplex.py:17-20)
        upper = [eval(v) if isinstance(v, str) else v for v in upper]
        self.dim = len(faces)
        if self.dim != 2:
            raise Exception("3-D meshes are outside the B200 hot path (SURVEY.md 8f item 3)")
        self.nx, self.ny = int(faces[0]), int(faces[1])
        self.lower = [float(lower[0]), float(lower[1])]
        self.upper = [float(upper[0]), float(upper[1])]
        self.comm = kwargs.get('comm', None)
        # multi-GPU: this object describes the rank's strip (its cell rows + one ghost row) in local numbering;
        # geometry and the physical borders stay those of the global mesh (domain/partition.py)
        self.partition = kwargs.get('partition', None)
        self.ny_global, self.fy0 = self.ny, 0
        if self.partition is not None:
            assert (self.partition.nx, self.partition.ny) == (self.nx, self.ny)
            self.ny, self.fy0 = self.partition.ny_local, self.partition.fy0
        self.logger = logging.getLogger("[0] Class")
        self.dim_w, self.dim_s = 1, 3
        self.namingConvention = ["down", "right", "up", "left"]
        self.ncell = self.nx * self.ny
        self.nvert = (self.nx + 1) * (self.ny + 1)
        self.nxedge = self.nx * (self.ny + 1)
        self.nyedge = (self.nx + 1) * self.ny
        self.cellStart, self.cellEnd = 0, self.ncell
        self._dev = {}
        self.indicesManager = None
        self.fullCoordVec = None

    def getDimension(self):
        return self.dim

    def isUniform(self):
        """Every cell is a copy of cell 0 (box mesh): one element matrix / operator set serves all cells."""
        return True

    # -- numbering --------------------------------------------------------------------------------------------
    def setFemIndexing(self, ngl):
        self.ngl = int(ngl)
        m = self.ngl - 2
        self.indicesManager = IndicesManager(self.dim, ngl, self.comm)
        self.off_vert = self.ncell * m * m
        self.off_edge = self.off_vert + self.nvert
        self.nnodes = self.off_edge + (self.nxedge + self.nyedge) * m
        self._dev = {}

    def getNGL(self):
        return self.indicesManager.getNGL()

    def getTotalNodes(self):
        return self.nnodes

    def getChart(self):
        return 0, self.ncell + self.nvert + self.nxedge + self.nyedge

    def getHeightStratum(self, h):
        if h == 0:
            return 0, self.ncell
        if h == 1:
            return self.ncell + self.nvert, self.ncell + self.nvert + self.nxedge + self.nyedge
        return self.ncell, self.ncell + self.nvert

    def _pointNodes(self, point):
        """node ids carried by a mesh point (section dof/offset, indices.py:107-114)."""
        m = self.ngl - 2
        if point < self.ncell:
            return list(range(point * m * m, (point + 1) * m * m))
        if point < self.ncell + self.nvert:
            return [self.off_vert + point - self.ncell]
        e = point - self.ncell - self.nvert
        return list(range(self.off_edge + e * m, self.off_edge + (e + 1) * m))

    def getTransitiveClosure(self, point):
        """(points, orientations) in DMPlex closure order: the point, its cone, then the vertices."""
        nc, nv, nx, ny = self.ncell, self.nvert, self.nx, self.ny
        vpt = lambda vx, vy: nc + vy * (nx + 1) + vx
        if point < nc:
            fx, fy = point % nx, point // nx
            e0 = nc + nv
            bottom, top = e0 + fy * nx + fx, e0 + (fy + 1) * nx + fx
            right, left = e0 + self.nxedge + (fx + 1) * ny + fy, e0 + self.nxedge + fx * ny + fy
            pts = [point, bottom, right, top, left, vpt(fx, fy), vpt(fx + 1, fy), vpt(fx + 1, fy + 1), vpt(fx, fy + 1)]
            return np.array(pts, dtype=np.int32), np.array([0, 0, 0, -1, -1, 0, 0, 0, 0], dtype=np.int32)
        if point < nc + nv:
            return np.array([point], dtype=np.int32), np.array([0], dtype=np.int32)
        e = point - nc - nv
        if e < self.nxedge:
            vy, ex = e // nx, e % nx
            ends = [vpt(ex, vy), vpt(ex + 1, vy)]
        else:
            e -= self.nxedge
            vx, ey = e // ny, e % ny
            ends = [vpt(vx, ey), vpt(vx, ey + 1)]
        return np.array([point] + ends, dtype=np.int32), np.array([0, 0, 0], dtype=np.int32)

    def getGlobalNodesFromCell(self, cell, shared=True):
        """Ordered global nodes of a mesh point's closure: vertices, edge interiors (reversed where the closure
        orientation is negative), cell interior (dmplex.py:197-200, indices.py:66-88)."""
        pts, orient = self.getTransitiveClosure(cell)
        if len(pts) == 9:
            order = [5, 6, 7, 8, 1, 2, 3, 4, 0]
        elif len(pts) == 3:
            order = [1, 2, 0]
        else:
            order = [0]
        nodes = []
        for i in order:
            nd = self._pointNodes(int(pts[i]))
            if orient[i]:
                nd = nd[::-1]
            nodes += nd
        return nodes

    def getGlobalNodesFromEntities(self, entities, shared=True):
        nodes = set()
        for entity in entities:
            nodes |= set(self.getGlobalNodesFromCell(int(entity), shared))
        return nodes

    def cellNodesArray(self, cells=None):
        """(ncells, nn) table of getGlobalNodesFromCell for many cells at once (host, vectorised)."""
        cells = np.arange(self.ncell) if cells is None else np.asarray(cells)
        nx, ny, m = self.nx, self.ny, self.ngl - 2
        fx, fy = cells % nx, cells // nx
        v = lambda vx, vy: self.off_vert + vy * (nx + 1) + vx
        k = np.arange(m)
        xe = lambda ex, vy: self.off_edge + (vy * nx + ex)[:, None] * m
        ye = lambda vx, ey: self.off_edge + (self.nxedge + vx * ny + ey)[:, None] * m
        cols = np.stack([v(fx, fy), v(fx + 1, fy), v(fx + 1, fy + 1), v(fx, fy + 1)], 1)
        parts = [cols, xe(fx, fy) + k, ye(fx + 1, fy) + k, xe(fx, fy + 1) + k[::-1], ye(fx, fy) + k[::-1],
                 (cells * m * m)[:, None] + np.arange(m * m)]
        return np.concatenate(parts, axis=1).astype(np.int64)

    def getVelocityIndex(self, nodes):
        return self.indicesManager.mapNodesToIndices(nodes, self.dim)

    def getVorticityIndex(self, nodes):
        return self.indicesManager.mapNodesToIndices(nodes, self.dim_w)

    def getSrtIndex(self, nodes):
        return self.indicesManager.mapNodesToIndices(nodes, self.dim_s)

    def getAllNodes(self):
        return list(range(self.nnodes))

    # -- geometry ---------------------------------------------------------------------------------------------
    def getCellCornersCoords(self, cell):
        """8 coordinates of the cell corners in closure order (x0,y0),(x1,y0),(x1,y1),(x0,y1) (dmplex.py:97-104)."""
        if cell + self.cellStart >= self.cellEnd:
            raise Exception('elem parameter must be in local numbering!')
        hx = (self.upper[0] - self.lower[0]) / self.nx
        hy = (self.upper[1] - self.lower[1]) / self.ny_global
        fx, fy = cell % self.nx, cell // self.nx + self.fy0
        x0, x1 = self.lower[0] + hx * fx, self.lower[0] + hx * (fx + 1)
        y0, y1 = self.lower[1] + hy * fy, self.lower[1] + hy * (fy + 1)
        return np.array([x0, y0, x1, y0, x1, y1, x0, y1], dtype=float)

    def computeFullCoordinates(self, spElem):
        """Coordinates of every GLL node (dmplex.py:66-95) -- one CUDA kernel instead of the cells x nodes loop."""
        from .. import backend as be
        xy = be.box_node_coords(self.nx, self.ny, self.ngl, self.lower, self.upper, spElem.nodes1D,
                                spElem.localOrder, self.nnodes, strip=(self.ny_global, self.fy0))
        self._dev["xy"] = xy
        self.fullCoordVec = Vec(data=xy.view(-1), name='NodeCoordinates')
        self.nodes = list(range(self.nnodes))

    def getNodesCoordinates(self, nodes=None, indices=None):
        dim = self.dim
        if nodes is not None:
            indices = self.indicesManager.mapNodesToIndices(nodes, dim)
            return self.fullCoordVec.getValues(indices).reshape((len(nodes), dim))
        assert indices is not None
        return self.fullCoordVec.getValues(indices).reshape((floor(len(indices) / dim), dim))

    # -- device tables ----------------------------------------------------------------------------------------
    def getBoundingBox(self):
        """(lower, upper) corners of the mesh (DMPlex.getBoundingBox, used by immersed_boundary.py:50)."""
        return list(self.lower), list(self.upper)

    def getTotalCells(self):
        return self.ncell

    def deviceCellNodes(self):
        from .. import backend as be
        if "e2n" not in self._dev:
            self._dev["e2n"] = be.box_cell_nodes(self.nx, self.ny, self.ngl)
        return self._dev["e2n"]

    def deviceCellCorners(self, cell0=0, ncells=None):
        from .. import backend as be
        ncells = self.ncell - cell0 if ncells is None else ncells
        return be.box_cell_corners(self.nx, self.ny_global, self.lower, self.upper, cell0 + self.fy0 * self.nx, ncells)

    def templateCellCorners(self):
        """Host (1,4,2) corner coordinates of GLOBAL cell 0: the one element the reference integrates when it reuses
        cell 0's matrices for every cell (base_problem.py:133,333).  Every rank of a partition uses the same cell, so
        the element matrices -- and with them the assembled rows -- are bit-identical to the serial ones."""
        hx = (self.upper[0] - self.lower[0]) / self.nx
        hy = (self.upper[1] - self.lower[1]) / self.ny_global
        x0, x1 = self.lower[0] + hx * 0, self.lower[0] + hx * 1
        y0, y1 = self.lower[1] + hy * 0, self.lower[1] + hy * 1
        return np.array([x0, y0, x1, y0, x1, y1, x0, y1], dtype=float).reshape(1, 4, 2)

    def deviceTemplateCorners(self):
        from .. import backend as be
        return be.box_cell_corners(self.nx, self.ny_global, self.lower, self.upper, 0, 1)

    def ownedNodeMask(self):
        """bool per local node: owned by this rank (all True without a partition)."""
        if self.partition is None:
            return np.ones(self.nnodes, dtype=bool)
        return self.partition.owned_mask()

    def deviceNodeCoords(self):
        return self._dev["xy"]

    def topology(self):
        from .. import backend as be
        if "topo" not in self._dev:
            self._dev["topo"] = be.Topology(self.deviceCellNodes(), self.nnodes)
        return self._dev["topo"]

    # -- borders / boundary conditions ------------------------------------------------------------------------
    def setLabelToBorders(self):
        pass                       # labels are implicit in the structured numbering (dmplex.py:113-131)

    def getBordersNames(self):
        return self.namingConvention

    def borderNodesArray(self, name):
        """nodes on the closure of a border's faces; "Face Sets" 1 down, 2 right, 3 up, 4 left (dmplex.py:160-171)."""
        nx, ny, m = self.nx, self.ny, self.ngl - 2
        k = np.arange(m)
        # a strip holds a physical border only where its local mesh (ghost row included) reaches the global one: the
        # ghost row of the last-but-one rank IS the top cell row when the last strip has a single row
        if self.partition is not None and ((name == "down" and self.fy0 > 0) or
                                           (name == "up" and self.fy0 + self.ny < self.ny_global)):
            return np.zeros(0, dtype=np.int64)          # that physical border lies outside this rank's local mesh
        if name in ("down", "up"):
            vy = 0 if name == "down" else ny
            verts = self.off_vert + vy * (nx + 1) + np.arange(nx + 1)
            ed = self.off_edge + ((vy * nx + np.arange(nx))[:, None] * m + k).reshape(-1)
        elif name in ("left", "right"):
            vx = 0 if name == "left" else nx
            verts = self.off_vert + np.arange(ny + 1) * (nx + 1) + vx
            ed = self.off_edge + ((self.nxedge + vx * ny + np.arange(ny))[:, None] * m + k).reshape(-1)
        else:
            raise Exception(f"unknown border {name}")
        return np.unique(np.concatenate([verts, ed])).astype(np.int64)

    def getBorderNodes(self, name):
        return [int(x) for x in self.borderNodesArray(name)]

    def getBordersNodes(self) -> set:
        nodes = set()
        for faceName in self.namingConvention:
            nodes |= set(self.getBorderNodes(faceName))
        return nodes

    def getNodesFromLabel(self, label, shared=False) -> set:
        if label == "External Boundary":
            return self.getBordersNodes()
        self.logger.warning(f"Label >> {label} << found")
        return set()

    def getGlobalIndicesDirichlet(self):
        return self.indicesManager.getDirichletNodes()

    def getGlobalIndicesNoSlip(self):
        return self.indicesManager.getNoSlipNodes()

    def setBoundaryCondition(self, freeSlipFaces=[], noSlipFaces=[]):
        """dmplex.py:181-195: named faces -> Dirichlet (free-slip) / no-slip node sets; none named -> every
        border node is Dirichlet."""
        if len(freeSlipFaces) or len(noSlipFaces):
            for fsFace in freeSlipFaces:
                self.indicesManager.setDirichletNodes(set(self.getBorderNodes(fsFace)))
            for nsFace in noSlipFaces:
                self.indicesManager.setNoSlipNodes(set(self.getBorderNodes(nsFace)))
        else:
            self.indicesManager.setDirichletNodes(self.getBordersNodes())

    def getBorderNodesWithNormal(self, cell, intersect):
        """Per border touched by the cell (through a face) whose nodes all lie in `intersect`: the node set and
        the index of the wall normal (dmplex.py:245-261)."""
        nodes, normals = [], []
        fx, fy = cell % self.nx, cell // self.nx
        pts, _ = self.getTransitiveClosure(cell)
        touch = {"down": (fy + self.fy0 == 0, pts[1], 1), "right": (fx == self.nx - 1, pts[2], 0),
                 "up": (fy + self.fy0 == self.ny_global - 1, pts[3], 1), "left": (fx == 0, pts[4], 0)}
        for faceName in self.namingConvention:
            on, edge, normal = touch[faceName]
            if on:
                borderNodes = self.getGlobalNodesFromEntities([edge], shared=True)
                if not (set(borderNodes) - set(intersect)):
                    normals.append(normal)
                    nodes.append(borderNodes)
        return nodes, normals

    # -- vector helpers (dmplex.py:263-296) ---------------------------------------------------------------------
    def applyFunctionVecToVec(self, nodes, f_vec, vec, dof):
        nodes = [int(x) for x in nodes]
        coords = self.getNodesCoordinates(nodes)
        inds = [node * dof + pos for node in nodes for pos in range(dof)]
        values = np.array(list(map(f_vec, coords)), dtype=float)
        vec.setValues(inds, values.reshape(-1), addv=False)
        return vec

    def applyFunctionScalarToVec(self, nodes, f_scalar, vec):
        nodes = [int(x) for x in nodes]
        coords = self.getNodesCoordinates(nodes)
        values = np.array(list(map(f_scalar, coords)), dtype=float)
        vec.setValues(nodes, values, addv=False)
        return vec

    def applyValuesToVec(self, nodes, values, vec):
        dof = len(values)
        assert dof <= self.dim
        if dof == 1:
            vec.set(values[0])
        else:
            nodes = [int(x) for x in nodes]
            vec.setValues(self.getVelocityIndex(nodes), np.array(list(values) * len(nodes), dtype=float), addv=False)
        return vec

    def getMatIndices(self):
        """Preallocation graph of the reference (dmplex.py:305-333): rows = nodes, columns = nodes sharing a cell.
        Here it is the node adjacency built on the device; returned as counts plus a lazy accessor because the
        per-row Python sets of the reference do not scale."""
        topo = self.topology()
        rowptr = np.empty(0)
        return 0, self.nnodes, topo, rowptr, None, None

    def getNodesOverline(self, line: str, val: float, invert=False):
        assert line in ['x', 'y']
        dof, orderDof = (0, 1) if line == 'x' else (1, 0)
        coords = self.fullCoordVec.getArray()
        nodes = np.where(coords[dof::self.dim] == val)[0]
        c = coords[nodes * self.dim + orderDof]
        order = np.argsort(c, kind="stable")
        return nodes[order].astype(int), c[order]

    def getVecArrayFromNodes(self, vec, nodes):
        arr = vec.getArray()
        return arr[nodes * self.dim], arr[nodes * self.dim + 1]
