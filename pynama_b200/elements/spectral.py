"""Spectral (GLL tensor-product) element -- host tables + device integration.

Public surface of the reference's `Spectral(ngl, dim)` (src/elements/spectral.py:9-62): attributes `ngl, nnode,
nnodedge, nnodcell, elemType, H, Hrs, gps, HRed, HrsRed, gpsRed, HOp, HrsOp, gpsOp, HCoo*, HrsCoo*, gpsCoo*,
HCoo1D, indCurl, indWCurl, indBdiv` and the methods `getElemKLEMatrices(coords)`, `getElemKLEOperators(coords)`.
The tabulations are host tables; the element matrices are integrated on the GPU (csrc/elem.cu) -- one element
through the reference-shaped methods, or a whole batch through the `*Device` methods.
"""
import numpy as np

from .element import Element
from .utilities import gaussPoints, generateGaussPoints2D, lobattoPoints


def localNodeOrder(n):
    """Element-local node k -> tensor index a*n+b (a: r-index, b: s-index; 1-D nodes ascending).

    The reference builds this through an index matrix and its inverse permutation (spectral.py:224-246); the
    result is a walk around the element: the four corners (+,+),(-,+),(-,-),(+,-), then the four sides in the
    same sense (interior nodes only), then the interior column by column with s descending."""
    if n == 1:
        return np.zeros(1, dtype=np.int64)
    p = n - 1
    ring = [(p, p), (0, p), (0, 0), (p, 0)]
    ring += [(a, p) for a in range(p - 1, 0, -1)]     # side 1->2: s = +1, r decreasing
    ring += [(0, b) for b in range(p - 1, 0, -1)]     # side 2->3: r = -1, s decreasing
    ring += [(a, 0) for a in range(1, p)]             # side 3->4: s = -1, r increasing
    ring += [(p, b) for b in range(1, p)]             # side 4->1: r = +1, s increasing
    ring += [(a, b) for a in range(1, p) for b in range(p - 1, 0, -1)]
    return np.array([a * n + b for a, b in ring], dtype=np.int64)


class Spectral(Element):
    """Spectral element of `ngl` nodes per side in `dim` dimensions (only dim = 2 is on the B200 hot path)."""

    ALPHA_W = 1e2   # curl penalty (spectral.py:93)
    ALPHA_D = 1e3   # divergence penalty (spectral.py:94)

    def __init__(self, ngl, dim):
        super().__init__(dim)
        self.ngl = ngl
        self.nnode = ngl ** dim
        self.nnodedge = ngl - 2
        self.nnodcell = (ngl - 2) ** dim
        self.elemType = 'Spectral{}D({})'.format(dim, ngl)
        self._kernels = None
        if dim == 2:
            self.indWCurl = [[0, 0, 1], [1, 0, 0]]
            self.indCurl = [[0, 1, 0], [0, 0, 1]]
            self.indBdiv = [[0, 1], [1, 2]]
            self.setUpSpectralMats2D(ngl)
        elif dim == 3:
            self.nnodface = (ngl - 2) ** 2
            raise Exception("3-D spectral elements are outside the B200 hot path (SURVEY.md 8f item 3)")
        else:
            raise Exception

    # -- host tables ----------------------------------------------------------------------------------------
    def setUpSpectralMats2D(self, ngl):
        nodes1D, operWei = lobattoPoints(ngl)
        gps1D, fullWei = gaussPoints(ngl) if ngl <= 3 else lobattoPoints(ngl)   # full rule (spectral.py:41-42)
        gpsRed1D, redWei = gaussPoints(ngl - 1)                                 # reduced rule (:43)
        cnodes1D, _ = lobattoPoints(2)                                          # geometry: 2 nodes per side (:44)
        self.nodes1D = nodes1D
        self._rules1D = {"full": (gps1D, fullWei), "red": (gpsRed1D, redWei), "op": (nodes1D, operWei)}
        self.H, self.Hrs, self.gps = self.computeMats2D(nodes1D, gps1D, fullWei)
        self.HRed, self.HrsRed, self.gpsRed = self.computeMats2D(nodes1D, gpsRed1D, redWei)
        self.HOp, self.HrsOp, self.gpsOp = self.computeMats2D(nodes1D, nodes1D, operWei)
        self.HCoo, self.HrsCoo, self.gpsCoo = self.computeMats2D(cnodes1D, gps1D, fullWei)
        self.HCooRed, self.HrsCooRed, self.gpsCooRed = self.computeMats2D(cnodes1D, gpsRed1D, redWei)
        self.HCooOp, self.HrsCooOp, self.gpsCooOp = self.computeMats2D(cnodes1D, nodes1D, operWei)
        self.HCoo1D, _ = self.interpFun1D(cnodes1D, nodes1D)
        self.localOrder = localNodeOrder(ngl)

    def computeMats2D(self, nodes1D, gps1D, gps1Dwei):
        """Lists (one entry per quadrature point, element-local point order) of the shape-function row H[k] and of
        the 2 x nnode reference-gradient Hrs, plus the matching Gauss points (spectral.py:220-300)."""
        h, dh = self.interpFun1D(nodes1D, gps1D)
        n, nq = len(nodes1D), len(gps1D)
        nodeOrder, pointOrder = localNodeOrder(n), localNodeOrder(nq)
        # tensor tables, point (q1,q2) -> q1*nq+q2, node (a,b) -> a*n+b
        Hfull = np.einsum("qa,pb->qpab", h, h).reshape(nq * nq, n * n)
        Hr = np.einsum("qa,pb->qpab", dh, h).reshape(nq * nq, n * n)
        Hs = np.einsum("qa,pb->qpab", h, dh).reshape(nq * nq, n * n)
        pts = generateGaussPoints2D(gps1D, gps1Dwei)
        H = [Hfull[q, nodeOrder] for q in pointOrder]
        Hrs = [np.array([Hr[q, nodeOrder], Hs[q, nodeOrder]]) for q in pointOrder]
        gps = [pts[q] for q in pointOrder]
        return H, Hrs, gps

    # -- device kernels -------------------------------------------------------------------------------------
    def kernels(self):
        """Lazily created device tables (needs a GPU and libpynama_b200.so)."""
        if self._kernels is None:
            from ..backend import ElemKernels
            rules = {}
            for name in ("full", "red"):
                pts, wts = self._rules1D[name]
                h, dh = self.interpFun1D(self.nodes1D, pts)
                rules[name] = (h, dh, wts, pts)
            _, D = self.interpFun1D(self.nodes1D, self.nodes1D)
            rules["op"] = (self._rules1D["op"][1], self.nodes1D)
            self._kernels = ElemKernels(self.ngl, rules, D, self.localOrder, self.ALPHA_W, self.ALPHA_D)
        return self._kernels

    def getElemKLEMatricesDevice(self, coords, want_rd=True, affine=False):
        """Batched: coords device tensor (ne,4,2) -> device tensors K (ne,2nn,2nn), Rw, Rd (ne,2nn,nn).
        affine=True: the caller guarantees parallelogram cells (box meshes): closed form of csrc/elem_affine.cuh -- the
        same quadrature, ~10x cheaper than the sum-factorised kernel and without its one-CTA-per-element latency."""
        k = self.kernels()
        if affine:
            coef, _ = k.affine_coef(coords, check=False)
            return k.kle_affine(coef, want_rd)
        return k.kle(coords, want_rd)

    def getAffineCoefficientsDevice(self, coords):
        """(coef (ne,18) device tensor, number of non-affine elements) -- csrc/elem_affine.cuh."""
        return self.kernels().affine_coef(coords)

    def getElemKLEOperatorsDevice(self, coords):
        return self.kernels().ops(coords)

    # -- reference-shaped single-element calls ---------------------------------------------------------------
    def _coordsToDevice(self, coords):
        from ..backend import to_device
        coords.shape = (int(len(coords) / self.dim), self.dim)        # in place, like the reference (:92,160)
        if coords.shape[0] != 4:
            raise Exception("straight-sided quadrilateral expected: 4 corner coordinates")
        return to_device(np.asarray(coords, dtype=np.float64)).view(1, 4, 2)

    def getElemKLEMatrices(self, coords):
        """Elementary KLE matrices (K, Rw, Rd) as NumPy arrays (spectral.py:89-157)."""
        K, Rw, Rd = self.getElemKLEMatricesDevice(self._coordsToDevice(coords))
        return K[0].cpu().numpy(), Rw[0].cpu().numpy(), Rd[0].cpu().numpy()

    def getElemKLEOperators(self, coords):
        """(SrT, DivSrT, Curl, W) as NumPy arrays (spectral.py:159-218)."""
        S, D, Cu, W = self.getElemKLEOperatorsDevice(self._coordsToDevice(coords))
        return S[0].cpu().numpy(), D[0].cpu().numpy(), Cu[0].cpu().numpy(), W[0].cpu().numpy()
