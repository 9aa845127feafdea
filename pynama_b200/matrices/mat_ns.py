"""`MatNS`: the eight matrices of the two-stage free-slip / no-slip velocity solve (reference:
src/matrices/mat_ns.py:17-121, filled by NoSlipFreeSlip.buildKLEMats, src/cases/base_problem.py:329-454).

Every velocity dof belongs to one of three classes (include/pynama_b200.h, pyn_asm_desc):

    0 FREE     unknown in both solves
    1 TANGENT  tangential component on a no-slip wall: unknown in the free-slip solve, prescribed in the final one
               (`dofFreeFSSetNS`, base_problem.py:359-369)
    2 SET      prescribed in both (wall-normal component, both components at corners and on Dirichlet borders;
               `dofSetFSNS`)

and each matrix is "the entries of the element matrix whose (row class, column class) pair is in a list", plus
diagonal fix-ups.  The table below is that statement for the reference's thirteen `setValues` calls and three
post-assembly loops; the assembly kernel evaluates it for the whole mesh in one launch per matrix.
"""
import numpy as np

from .. import backend as be
from .mat_generator import FREE, SET, TANGENT, Mat, fusedWanted, rowClasses, templatesWanted

# name: (cdof, element matrix, kept (row class, column class) pairs, classes whose diagonal entry is stored,
#        value added on that diagonal per class, sign)
NS_DESCRIPTORS = {
    # K[free,free] += K_e ; explicit 0 on the diagonal of every dof to be set (:426-430) ; +1 there (:439)
    "K": (2, "K", [(FREE, FREE)], (1 << TANGENT) | (1 << SET), (0.0, 1.0, 1.0, 0.0), 1.0),
    # Krhs[free, toBeSet] -= K_e ; explicit 0 (:388-395) ; +1 (:439)
    "Krhs": (2, "K", [(FREE, TANGENT), (FREE, SET)], (1 << TANGENT) | (1 << SET), (0.0, 1.0, 1.0, 0.0), -1.0),
    # Kfs: three blocks around the tangential dofs (:396-407) ; -1 on their diagonal (:441-442)
    "Kfs": (2, "K", [(TANGENT, FREE), (FREE, TANGENT), (TANGENT, TANGENT)], 1 << TANGENT, (0.0, -1.0, 0.0, 0.0), 1.0),
    # Krhsfs[tangential|free, set] -= K_e ; explicit 0 then INSERT 1 on the diagonal of set dofs (:417-424,449-451)
    "Krhsfs": (2, "K", [(TANGENT, SET), (FREE, SET)], 1 << SET, (0.0, 0.0, 1.0, 0.0), -1.0),
    "Rw": (1, "Rw", [(FREE, FREE)], 0, (0.0,) * 4, 1.0),          # (:432-433)
    "Rwfs": (1, "Rw", [(TANGENT, FREE)], 0, (0.0,) * 4, 1.0),     # (:412-413)
    "Rd": (1, "Rd", [(FREE, FREE)], 0, (0.0,) * 4, 1.0),          # (:435-437)
    "Rdfs": (1, "Rd", [(TANGENT, FREE)], 0, (0.0,) * 4, 1.0),     # (:415-416)
}


def noSlipDofClasses(dom, noSlipNodes=None, dirichletNodes=None):
    """int8 class per velocity dof (host array).  No-slip walls: the wall-normal component is SET, the tangential one
    TANGENT, and a dof that is normal to any wall through the node is SET (corners: both) -- base_problem.py:359-369
    with the per-border normals of dmplex.py:245-261.  Dirichlet (free-slip-listed) borders: every component SET
    (:371-375).  Borders are taken whole, as the reference's `getBorderNodesWithNormal` does."""
    N, dim = dom.getTotalNodes(), dom.dim
    normal = {"down": 1, "up": 1, "left": 0, "right": 0}
    is_set = np.zeros((N, dim), dtype=bool)
    is_tan = np.zeros((N, dim), dtype=bool)
    ns = None if noSlipNodes is None else np.fromiter(noSlipNodes, dtype=np.int64, count=len(noSlipNodes))
    dr = None if dirichletNodes is None else np.fromiter(dirichletNodes, dtype=np.int64, count=len(dirichletNodes))
    for name in dom.getBordersNames():
        nd = dom.borderNodesArray(name)
        if ns is not None and len(ns) and np.isin(nd, ns).all():
            is_set[nd, normal[name]] = True
            is_tan[nd, 1 - normal[name]] = True
        if dr is not None and len(dr) and np.isin(nd, dr).all():
            is_set[nd, :] = True
    is_tan &= ~is_set
    cls = np.zeros((N, dim), dtype=np.int8)
    cls[is_tan] = TANGENT
    cls[is_set] = SET
    return cls.reshape(-1)


class MatNS(Mat):
    def __init__(self, dim, comm=None):
        super().__init__(dim, comm)

    def assembleAll(self):
        for m in (self.K, self.Rw, self.Rd, self.Krhs):
            m.assemble()

    def createEmptyKLEMats(self, rStart, rEnd, d_nnz_ind, o_nnz_ind, ind_d, ind_o, indicesDIR, indicesNS):
        self.globalIndicesDIR = set(indicesDIR)      # node sets, as in the reference (mat_ns.py:18-24)
        self.globalIndicesNS = set(indicesNS)
        locElRow = rEnd - rStart
        vel_dofs, vort_dofs = locElRow * self.dim, locElRow * self.dim_w
        for name, cols in (("K", vel_dofs), ("Rw", vort_dofs), ("Rd", locElRow), ("Krhs", vel_dofs),
                           ("Kfs", vel_dofs), ("Rwfs", vort_dofs), ("Rdfs", locElRow), ("Krhsfs", vel_dofs)):
            setattr(self, name, self.createEmptyMat(vel_dofs, cols, None, None, name))
        self.mats = [self.K, self.Rw, self.Rd, self.Krhs, self.Kfs, self.Rwfs, self.Rdfs, self.Krhsfs]

    # -- B200 path --------------------------------------------------------------------------------------------
    def hostDofClasses(self, dom):
        """(column classes, row classes) as host int8 arrays; they differ only on a partitioned mesh, where the rows
        of ghost nodes are marked NOT_OWNED."""
        host = noSlipDofClasses(dom, self.globalIndicesNS, self.globalIndicesDIR)
        return host, rowClasses(dom, host, 2)

    def prepareBoundaryMaps(self, dom, classes, rowClassesDev=None):
        """Boundary maps (flags + list of the boundary-affected nodes, csrc/assemble.cu pyn_bcmap_create) of a set of
        device dof classes; a no-op when `classes` is the object already prepared.  Runs on the CURRENT stream and
        reads one scalar back, so a caller that streams its inputs can call it on an upload stream ahead of
        `assembleFromMesh` (bench.py end-to-end phase) instead of stalling the assembly stream."""
        if getattr(self, "dofClasses", None) is classes:
            return self._maps
        if rowClassesDev is None:
            if dom.partition is not None:
                raise Exception("partitioned mesh: pass rowClassesDev together with classes")
            rowClassesDev = classes
        topo = dom.topology()
        self.dofClasses, self.rowClasses = classes, rowClassesDev
        self._maps = (be.BcMap(topo, self.rowClasses, 2, classes, 2), be.BcMap(topo, self.rowClasses, 2))
        return self._maps

    def assembleFromMesh(self, dom, elemType, names=None, classes=None, rowClassesDev=None, corners=None, after=None):
        """NoSlipFreeSlip.buildKLEMats for the whole mesh on the device (cell 0's element matrices for every cell,
        base_problem.py:333-334).  `names` restricts the matrices (default: all eight); re-assembly on an existing
        pattern rewrites values only.  `classes` / `rowClassesDev` (device int8, from `hostDofClasses`) and `corners`
        (device (1,4,2) corner coordinates of cell 0) let a caller feed the inputs from host buffers.  `after`:
        optional {name: callable}, called right after that matrix's kernels are enqueued (timing events)."""
        topo = dom.topology()
        if classes is None:
            host, hostRows = self.hostDofClasses(dom)
            classes = be.to_device(host)
            rowClassesDev = classes if dom.partition is None else be.to_device(hostRows)
        self.prepareBoundaryMaps(dom, classes, rowClassesDev)
        if corners is None:
            corners = dom.deviceTemplateCorners()
        # cell 0 of a box mesh is a rectangle: closed-form element matrices (unless -mat_fused 0)
        Ke, Rwe, Rde = elemType.getElemKLEMatricesDevice(corners, want_rd=True, affine=fusedWanted())
        E = {"K": Ke, "Rw": Rwe, "Rd": Rde}
        with be.assembly_batch(topo):     # boundary-list kernels of all matrices overlap the interior streams
            for name in (names or NS_DESCRIPTORS):
                cd, which, pairs, dcls, dadd, sign = NS_DESCRIPTORS[name]
                m = getattr(self, name)
                csr = be.assemble(topo, 2, cd, E[which], self.rowClasses, classes if cd == 2 else None, pairs,
                                  diag_classes=dcls, diag_add=dadd, sign=sign, pattern=m.csr, inplace=True, name=name,
                                  bcmap=self._maps[0] if cd == 2 else self._maps[1])
                m.csr, m.shape = csr, csr.shape
                if templatesWanted():
                    csr.enable_templates()
                if after and name in after:
                    after[name]()
        self._unit_diagonals_folded = True
