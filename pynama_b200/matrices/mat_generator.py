"""`Mat` / `Operators`: the KLE matrices and the projection operators (reference: src/matrices/mat_generator.py).

Same names and call surface as the reference (`createEmptyKLEMats`, `createEmptyMat`, `createNNZWithArray`,
`setIndices2One`, `assembleAll`, `Operators.createAll / setValues / assembleAll`, attributes `K Rw Rd Krhs`,
`Curl SrT DivSrT`), two ways of filling them:

  * the reference's way -- `mat.K.setValues(rows, cols, block, addv=True)` per cell followed by `assemble()` -- works
    (the staged triplets go through the deterministic COO->CSR kernels of the library), and is what a hand-written
    driver or a small test uses;
  * the B200 way -- `assembleFromMesh(dom, elemType, ...)` -- hands the whole mesh to the batched integration and
    the gather-assembly kernels (csrc/elem.cu, csrc/assemble.cu): no per-cell Python, no COO, bit-identical pattern
    and summation order.  The case classes (`cases.base_problem`) use this one.

PETSc preallocation has no meaning for the device assembly (the pattern is counted exactly on the GPU), so the
`d_nnz/o_nnz` arguments are accepted and ignored; the helpers that compute them are kept because the reference's tests
pin them (src/tests/test_mat.py:15-25).
"""
import logging

import numpy as np

from .. import backend as be
from ..petsc_like import Mat as DeviceMat
from ..petsc_like import Vec

# dof classes used by the assembly descriptors (include/pynama_b200.h: pyn_asm_desc)
FREE, TANGENT, SET = 0, 1, 2
NOT_OWNED = 3      # multi-GPU: row of a ghost node -- belongs to another rank, no entry is stored


def rowClasses(dom, colClasses, ncomp):
    """Row classes of a (possibly partitioned) mesh: the column classes with the rows of ghost nodes marked
    NOT_OWNED.  `colClasses` is a host int8 array (or None = all FREE); returns a host array or None."""
    if dom.partition is None:
        return colClasses
    own = np.repeat(dom.ownedNodeMask(), ncomp)
    base = np.zeros(own.size, dtype=np.int8) if colClasses is None else colClasses
    return np.where(own, base, NOT_OWNED).astype(np.int8)


def templatesWanted():
    """`-mat_templates 0` keeps every product on the CSR stream; default: row templates wherever a matrix is eligible
    (one element matrix for all cells on a repetitive mesh -- csrc/tmat.cuh)."""
    from ..common.options import Options
    return str(Options().getString("mat_templates", "1")).lower() not in ("0", "false", "no", "off")


def fusedWanted():
    """`-mat_fused 0` materialises the element matrices of a per-cell assembly; default: fused integrate-and-assemble
    whenever every element is affine."""
    from ..common.options import Options
    return str(Options().getString("mat_fused", "1")).lower() not in ("0", "false", "no", "off")


def _rank(comm):
    return getattr(comm, "rank", 0)


class Mat(object):
    def __init__(self, dim, comm=None):
        self.dim = dim
        self.logger = logging.getLogger(f"[{_rank(comm)}]:MatClass")
        self.dim_w = 1 if self.dim == 2 else 3
        self.dim_s = 3 if self.dim == 2 else 6
        self.comm = comm
        self.mats = list()

    # -- reference surface ------------------------------------------------------------------------------------
    def assembleAll(self):
        for m in self.mats:
            m.assemble()
            self.logger.debug(f"Mat {m.getName()} Assembled")

    def isParallel(self):
        return getattr(self.comm, "size", 1) > 1

    def getGlobalIndices(self, localIndices):
        return set(localIndices)            # BC node sets are computed replicated on every rank

    def createEmptyKLEMats(self, rStart, rEnd, d_nnz_ind, o_nnz_ind, ind_d, ind_o, indicesDIR):
        self.globalNodesDIR = self.getGlobalIndices(indicesDIR)
        nodes = np.array(sorted(self.globalNodesDIR), dtype=np.int64)
        # the reference turns the node set into the list of constrained velocity dofs (mat_generator.py:61-67)
        self.globalIndicesDIR = (nodes[:, None] * self.dim + np.arange(self.dim)).reshape(-1)
        locElRow = rEnd - rStart
        vel_dofs, vort_dofs = locElRow * self.dim, locElRow * self.dim_w
        self.K = self.createEmptyMat(vel_dofs, vel_dofs, None, None, "K")
        self.Rw = self.createEmptyMat(vel_dofs, vort_dofs, None, None, "Rw")
        self.Rd = self.createEmptyMat(vel_dofs, locElRow, None, None, "Rd")
        self.Krhs = self.createEmptyMat(vel_dofs, vel_dofs, None, None, "Krhs")
        self.mats = [self.K, self.Rw, self.Rd, self.Krhs]

    def createEmptyMat(self, rows, cols, d_nonzero=None, offset_nonzero=None, name=""):
        return DeviceMat((int(rows), int(cols)), name)

    def createNonZeroIndex(self, d_nnz, o_nnz, dim1, dim2):
        di_nnz = [x * dim1 for x in d_nnz for d in range(dim2)]
        oi_nnz = [x * dim1 for x in o_nnz for d in range(dim2)]
        return di_nnz, oi_nnz

    def createNNZWithArray(self, d_nnz, o_nnz, dim1, dim2):
        d_nnz = np.array(d_nnz, dtype=np.int32)
        o_nnz = np.array(o_nnz, dtype=np.int32)
        return np.repeat(d_nnz * dim1, dim2), np.repeat(o_nnz * dim1, dim2)

    def setIndices2One(self, indices2one):
        """+1 on the diagonal of K and Krhs for every constrained dof (mat_generator.py:113-118).  Matrices produced
        by `assembleFromMesh` already carry it (folded into the assembly descriptor)."""
        if getattr(self, "_unit_diagonals_folded", False):
            return
        idx = [int(i) for i in indices2one]
        for m in (self.Krhs, self.K):
            for i in idx:
                m.setValues(i, i, 1, addv=True)
            m.assemble()

    def printMatsInfo(self):
        print(" MATS INFO ")
        print("Mat   | Memory Used [B]  | NZ Unneeded")
        print("--------------------------------------")
        for m in self.mats:
            print(self.formatMatInfo(m.getName(), m.getInfo()))

    @staticmethod
    def formatMatInfo(name, info):
        return f"{name:{5}} | {info['memory']:{16}} | {info['nz_unneeded']:{10}}"

    # -- B200 path --------------------------------------------------------------------------------------------
    def _install(self, name, csr):
        m = getattr(self, name)
        m.csr, m.shape = csr, csr.shape
        csr.name = name
        if templatesWanted():
            csr.enable_templates()

    def assembleFromMesh(self, dom, elemType, perCell=True, corners=None, elementMatrices=None, after=None):
        """FreeSlip.buildKLEMats (src/cases/base_problem.py:499-552) for the whole mesh on the device: element
        matrices of every cell (perCell, :505-506) or of cell 0 only, then K (free x free, explicit zero + 1 on the
        diagonal of constrained dofs), Krhs (-K_e[free, set], + 1 on constrained diagonals), Rw (free rows).
        Rd stays empty, as in the reference.  A second call on the same mesh rewrites the values on the existing
        pattern.  `corners`: device (ne,4,2) corner coordinates (default: the box mesh's); `elementMatrices`: (Ke, Rwe)
        already integrated; `after`: {"elem" | matrix name: callable} called after that stage is enqueued (timing)."""
        topo = dom.topology()
        if getattr(self, "_asmState", None) is None or self._asmState[0] is not topo:
            N = dom.getTotalNodes()
            cls_np = np.zeros(N, dtype=np.int8)
            cls_np[np.array(sorted(self.globalNodesDIR), dtype=np.int64)] = SET
            cls_host = np.repeat(cls_np, self.dim)
            cls = be.to_device(cls_host)
            rcls = cls if dom.partition is None else be.to_device(rowClasses(dom, cls_host, self.dim))
            self._asmState = (topo, cls, rcls, be.BcMap(topo, rcls, 2, cls, 2), be.BcMap(topo, rcls, 2))
            first = True
        else:
            first = False          # values-only re-assembly on the existing pattern (same mesh, same classes)
        _, cls, rcls, m22, m21 = self._asmState
        if corners is None:
            corners = dom.deviceCellCorners(0, None) if perCell else dom.deviceTemplateCorners()
        self.fusedAssembly = False
        if elementMatrices is None:
            Ke = None
            if perCell and fusedWanted():
                # affine elements (every box mesh): 17 coefficients per element, entries evaluated inside the assembly
                # kernels -- the element matrices are never written to memory (csrc/elem_affine.cuh)
                coef, nonaffine = elemType.getAffineCoefficientsDevice(corners)
                if nonaffine == 0:
                    kern = elemType.kernels()
                    Ke, Rwe = be.AffineSource(kern, coef, "K"), be.AffineSource(kern, coef, "Rw")
                    self.fusedAssembly = True
            if Ke is None:
                Ke, Rwe, _ = elemType.getElemKLEMatricesDevice(corners, want_rd=False)
        else:
            Ke, Rwe = elementMatrices
        if after and "elem" in after:
            after["elem"]()
        one = (0.0, 0.0, 1.0, 0.0)
        for name, cd, E, ccls, pairs, dcls, dadd, sign, bmap in (
                ("K", 2, Ke, cls, [(FREE, FREE)], 1 << SET, one, 1.0, m22),
                ("Krhs", 2, Ke, cls, [(FREE, SET)], 1 << SET, one, -1.0, m22),
                ("Rw", 1, Rwe, None, [(FREE, FREE)], 0, (0.0,) * 4, 1.0, m21)):
            if first:
                self._install(name, be.assemble(topo, 2, cd, E, rcls, ccls, pairs, diag_classes=dcls, diag_add=dadd,
                                                sign=sign, bcmap=bmap))
            else:
                m = getattr(self, name)
                m.csr = be.assemble(topo, 2, cd, E, rcls, ccls, pairs, diag_classes=dcls, diag_add=dadd, sign=sign,
                                    bcmap=bmap, pattern=m.csr, inplace=True, name=name)
            if after and name in after:
                after[name]()
        self.dofClasses, self.rowClasses = cls, rcls
        self._unit_diagonals_folded = True


class Operators(Mat):
    def createAll(self, rStart, rEnd, d_nnz_ind=None, o_nnz_ind=None):
        locElRow = rEnd - rStart
        self.createCurl(d_nnz_ind, o_nnz_ind, locElRow)
        self.createDivSrt(d_nnz_ind, o_nnz_ind, locElRow)
        self.createSrT(d_nnz_ind, o_nnz_ind, locElRow)

    def create(self, dim1, dim2, d_nnz_ind, o_nnz_ind, locElRow, name=""):
        return self.createEmptyMat(locElRow * dim2, locElRow * dim1, None, None, name)

    def createCurl(self, d_nnz_ind, o_nnz_ind, locElRow):
        self.Curl = self.create(self.dim, self.dim_w, d_nnz_ind, o_nnz_ind, locElRow, "Curl")
        self.weigCurl = Vec(locElRow * self.dim_w)

    def createDivSrt(self, d_nnz_ind, o_nnz_ind, locElRow):
        self.DivSrT = self.create(self.dim_s, self.dim, d_nnz_ind, o_nnz_ind, locElRow, "DivSrT")
        self.weigDivSrT = Vec(locElRow * self.dim)

    def createSrT(self, d_nnz_ind, o_nnz_ind, locElRow):
        self.SrT = self.create(self.dim, self.dim_s, d_nnz_ind, o_nnz_ind, locElRow, "SrT")
        self.weigSrT = Vec(locElRow * self.dim_s)

    def setValues(self, localOperators, nodes):
        """Reference-shaped per-cell scatter (mat_generator.py:157-170); staged and assembled on the device."""
        locSrT, locDivSrT, locCurl, locWei = localOperators
        nodes = [int(n) for n in nodes]
        indicesVel = [n * self.dim + d for n in nodes for d in range(self.dim)]
        indicesW = [n * self.dim_w + d for n in nodes for d in range(self.dim_w)]
        indicesSrT = [n * self.dim_s + d for n in nodes for d in range(self.dim_s)]
        self.Curl.setValues(indicesW, indicesVel, locCurl, True)
        self.SrT.setValues(indicesSrT, indicesVel, locSrT, True)
        self.DivSrT.setValues(indicesVel, indicesSrT, locDivSrT, True)
        self.weigSrT.setValues(indicesSrT, np.repeat(locWei, self.dim_s), True)
        self.weigDivSrT.setValues(indicesVel, np.repeat(locWei, self.dim), True)
        self.weigCurl.setValues(indicesW, np.repeat(locWei, self.dim_w), True)

    def assembleAll(self):
        if getattr(self, "_from_mesh", False):
            return
        for m, w in ((self.SrT, self.weigSrT), (self.DivSrT, self.weigDivSrT), (self.Curl, self.weigCurl)):
            m.assemble()
            w.assemble()
            w.reciprocal()
            m.diagonalScale(L=w)
        self.weigSrT.destroy()
        self.weigCurl.destroy()
        self.weigDivSrT.destroy()

    # -- B200 path --------------------------------------------------------------------------------------------
    def assembleFromMesh(self, dom, elemType, compressed=None):
        """BaseProblem.buildOperators + Operators.assembleAll (src/cases/base_problem.py:132-140,
        mat_generator.py:157-190) for the whole mesh on the device: cell 0's element operators scattered to every
        cell, lumped weights summed in cell order, rows scaled by 1/W -- one gather-assembly launch per operator.

        The element operators are 70-87 % structural zeros which PETSc keeps as stored entries.  `self.SrT/DivSrT/Curl`
        carry that reference pattern when it fits (`compressed=False`, or automatically below ~2 M nodes);
        `self.spmv[...]` always holds zero-squeezed copies, which is what the time loop multiplies with."""
        topo = dom.topology()
        N = dom.getTotalNodes()
        # a non-uniform mesh integrates the operators of every cell; a box mesh uses cell 0's for all
        # cells like the reference (base_problem.py:133-137)
        perCell = not getattr(dom, "isUniform", lambda: True)()
        S, D, C, W = elemType.getElemKLEOperatorsDevice(dom.deviceCellCorners(0, None) if perCell
                                                        else dom.deviceTemplateCorners())
        Wg = be.assemble_weights(topo, W)
        self.weights = Wg
        rW = Wg.reciprocal()
        if compressed is None:
            compressed = N > 2_000_000
        self.spmv = {}
        for name, E, rd, cd in (("SrT", S, self.dim_s, self.dim), ("DivSrT", D, self.dim, self.dim_s),
                                ("Curl", C, self.dim_w, self.dim)):
            rc = rowClasses(dom, None, rd)              # multi-GPU: no rows for ghost nodes
            rc = None if rc is None else be.to_device(rc)
            bmap = be.BcMap(topo, rc, rd)
            full = be.assemble(topo, rd, cd, E, rclass=rc, pairs=[(FREE, FREE)], rowscale=rW, bcmap=bmap, name=name)
            small = full.compressed(name)
            if templatesWanted():
                # interior rows through the row templates of the full block (4 B per adjacency slot); the squeezed
                # CSR only serves the rows of flagged (ghost) nodes, where both matrices agree
                small.enable_templates(like=full)
            self.spmv[name] = DeviceMat.fromCSR(small, name)
            m = getattr(self, name)
            m.csr, m.shape = (small if compressed else full), full.shape
            del full
        self._from_mesh = True
