"""Device-side objects of the hot path: thin Python owners of HBM buffers + calls into libpynama_b200.so.

PyTorch is used only as the allocator / stream provider (torch tensors own the device memory; raw pointers are
handed to the C ABI).  Nothing here computes on the CPU.
"""
from __future__ import annotations

import ctypes as C

import numpy as np
import torch

from . import _lib
from ._lib import AsmDesc, KspOpts, KspResult, call

F64, I64, I32, I8 = torch.float64, torch.int64, torch.int32, torch.int8


def device():
    if not torch.cuda.is_available():
        raise _lib.PynamaError("pynama_b200 needs a CUDA device (B200, sm_100a); there is no CPU fallback")
    return torch.device("cuda", torch.cuda.current_device())


def stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def ptr(t):
    return C.c_void_p(t.data_ptr()) if t is not None else None


def empty(n, dtype=F64):
    return torch.empty(int(n), dtype=dtype, device=device())


def zeros(n, dtype=F64):
    return torch.zeros(int(n), dtype=dtype, device=device())


def to_device(a, dtype=None):
    t = torch.as_tensor(np.ascontiguousarray(a))
    if dtype is not None:
        t = t.to(dtype)
    return t.to(device())


def launch_count():
    return int(_lib.load().pyn_launch_count())


def _np_ptr(a):
    return a.ctypes.data_as(C.c_void_p)


# ---------------------------------------------------------------------------------------------------------
class ElemKernels:
    """Device tables of one Spectral(ngl, 2) element + the batched integration kernels (pyn_elem_*)."""

    def __init__(self, ngl, rules, deriv_nodal, perm, alpha_w=1e2, alpha_d=1e3):
        """rules: dict name -> (h[q,a], d[q,a], w[q], xi[q]) for 'full', 'red'; plus 'op' -> (w, xi)."""
        self.ngl, self.nn = int(ngl), int(ngl) ** 2
        f64 = lambda a: np.ascontiguousarray(a, dtype=np.float64)
        hF, dF, wF, xF = map(f64, rules["full"])
        hG, dG, wG, xG = map(f64, rules["red"])
        wO, xO = map(f64, rules["op"])
        D = f64(deriv_nodal)
        perm = np.ascontiguousarray(perm, dtype=np.int32)
        self._h = C.c_void_p()
        call("pyn_elem_create", self.ngl, len(wF), _np_ptr(hF), _np_ptr(dF), _np_ptr(wF), _np_ptr(xF),
             len(wG), _np_ptr(hG), _np_ptr(dG), _np_ptr(wG), _np_ptr(xG), _np_ptr(D), _np_ptr(wO), _np_ptr(xO),
             _np_ptr(perm), float(alpha_w), float(alpha_d), C.byref(self._h))

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                _lib.load().pyn_elem_destroy(self._h)
                self._h = None
        except Exception:
            pass

    def kle(self, coords, want_rd=True):
        """coords: device tensor (ne,4,2) -> K (ne,2nn,2nn), Rw (ne,2nn,nn), Rd (ne,2nn,nn) on device."""
        coords = coords.contiguous()
        ne, nn = coords.numel() // 8, self.nn
        K = empty(ne * 4 * nn * nn).view(ne, 2 * nn, 2 * nn)
        Rw = empty(ne * 2 * nn * nn).view(ne, 2 * nn, nn)
        Rd = empty(ne * 2 * nn * nn).view(ne, 2 * nn, nn) if want_rd else None
        call("pyn_elem_kle", self._h, ptr(coords), ne, ptr(K), ptr(Rw), ptr(Rd), stream())
        return K, Rw, Rd

    def affine_coef(self, coords, check=True):
        """coords (ne,4,2) -> (coef (ne,18) on device, number of non-affine elements): the closed-form coefficients
        of csrc/elem_affine.cuh.  Synchronises (one scalar) unless check=False (caller knows the cells are affine)."""
        coords = coords.contiguous()
        ne = coords.numel() // 8
        coef = empty(ne * 18).view(ne, 18)
        bad = C.c_int64()
        call("pyn_elem_affine_coef", self._h, ptr(coords), ne, ptr(coef), C.byref(bad) if check else None, stream())
        return coef, (bad.value if check else 0)

    def kle_affine(self, coef, want_rd=True):
        """Materialised K, Rw, Rd of affine elements from their coefficients (same layout as kle())."""
        ne, nn = coef.shape[0], self.nn
        K = empty(ne * 4 * nn * nn).view(ne, 2 * nn, 2 * nn)
        Rw = empty(ne * 2 * nn * nn).view(ne, 2 * nn, nn)
        Rd = empty(ne * 2 * nn * nn).view(ne, 2 * nn, nn) if want_rd else None
        call("pyn_elem_kle_affine", self._h, ptr(coef), ne, ptr(K), ptr(Rw), ptr(Rd), stream())
        return K, Rw, Rd

    def ops(self, coords):
        coords = coords.contiguous()
        ne, nn = coords.numel() // 8, self.nn
        SrT = empty(ne * 6 * nn * nn).view(ne, 3 * nn, 2 * nn)
        Div = empty(ne * 6 * nn * nn).view(ne, 2 * nn, 3 * nn)
        Curl = empty(ne * 2 * nn * nn).view(ne, nn, 2 * nn)
        W = empty(ne * nn).view(ne, nn)
        call("pyn_elem_ops", self._h, ptr(coords), ne, ptr(SrT), ptr(Div), ptr(Curl), ptr(W), stream())
        return SrT, Div, Curl, W


# ---------------------------------------------------------------------------------------------------------
def box_cell_nodes(nx, ny, ngl, cell0=0, ncells=None):
    ncells = nx * ny - cell0 if ncells is None else ncells
    out = empty(ncells * ngl * ngl, I32).view(ncells, ngl * ngl)
    call("pyn_box_cell_nodes", nx, ny, ngl, cell0, ncells, ptr(out), stream())
    return out


def box_cell_corners(nx, ny, lower, upper, cell0=0, ncells=None):
    ncells = nx * ny - cell0 if ncells is None else ncells
    out = empty(ncells * 8).view(ncells, 4, 2)
    call("pyn_box_cell_corners", nx, ny, float(lower[0]), float(lower[1]), float(upper[0]), float(upper[1]),
         cell0, ncells, ptr(out), stream())
    return out


def box_node_coords(nx, ny, ngl, lower, upper, xi_nodal, perm, nnodes, strip=None):
    """strip = (ny_global, fy0): `ny` cell rows starting at row fy0 of the global mesh, local numbering."""
    out = empty(nnodes * 2).view(nnodes, 2)
    xi = np.ascontiguousarray(xi_nodal, dtype=np.float64)
    pm = np.ascontiguousarray(perm, dtype=np.int32)
    if strip is None or strip == (ny, 0):
        call("pyn_box_node_coords", nx, ny, ngl, float(lower[0]), float(lower[1]), float(upper[0]), float(upper[1]),
             _np_ptr(xi), _np_ptr(pm), ptr(out), stream())
    else:
        call("pyn_box_node_coords_strip", nx, int(strip[0]), int(strip[1]), ny, ngl, float(lower[0]), float(lower[1]),
             float(upper[0]), float(upper[1]), _np_ptr(xi), _np_ptr(pm), ptr(out), stream())
    return out


class Topology:
    """node->element incidence + node adjacency on device (pyn_topo_*)."""

    def __init__(self, elem2node, nnodes):
        self.elem2node = elem2node.contiguous()       # kept alive: the handle borrows it
        self.ne, self.nn = self.elem2node.shape
        self.nnodes = int(nnodes)
        self._h = C.c_void_p()
        call("pyn_topo_create", ptr(self.elem2node), self.ne, self.nn, self.nnodes, stream(), C.byref(self._h))
        nnz, val, row = C.c_int64(), C.c_int(), C.c_int()
        call("pyn_topo_info", self._h, C.byref(nnz), C.byref(val), C.byref(row))
        self.adj_nnz, self.max_valence, self.max_row = nnz.value, val.value, row.value
        ns, slots = C.c_int(), C.c_int64()
        call("pyn_topo_signatures", self._h, C.byref(ns), C.byref(slots))
        self.nsig, self.sig_slots = ns.value, slots.value      # distinct row signatures (0: templates unavailable)

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                _lib.load().pyn_topo_destroy(self._h)
                self._h = None
        except Exception:
            pass


class BcMap:
    """Boundary map of one (row classes, column classes) pair on a topology (pyn_bcmap_*): lets the assembly run its
    streamlined interior kernel on every node without boundary influence and visit only the listed nodes for the
    boundary-only matrices.  Holds the class tensors, which must not be modified afterwards."""

    def __init__(self, topo, rclass=None, rdof=0, cclass=None, cdof=0):
        self.topo, self.rclass, self.cclass = topo, rclass, cclass
        self._h = C.c_void_p()
        call("pyn_bcmap_create", topo._h, ptr(rclass), int(rdof if rclass is not None else 0), ptr(cclass),
             int(cdof if cclass is not None else 0), stream(), C.byref(self._h))
        n = C.c_int64()
        call("pyn_bcmap_info", self._h, C.byref(n))
        self.n_boundary = n.value

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                _lib.load().pyn_bcmap_destroy(self._h)
                self._h = None
        except Exception:
            pass


class assembly_batch:
    """`with assembly_batch(topo): ...` -- the values-only assemble() calls inside run their boundary-list kernels on
    the topology's side stream and are joined once at the end (pyn_asm_defer_join / pyn_asm_join)."""

    def __init__(self, topo):
        self.topo = topo

    def __enter__(self):
        call("pyn_asm_defer_join", 1)
        return self

    def __exit__(self, *exc):
        call("pyn_asm_defer_join", 0)
        call("pyn_asm_join", self.topo._h, stream())
        return False


def set_templates(enabled):
    """Enable/disable the row-template fast path of the values-only assembly (pyn_asm_set_templates)."""
    call("pyn_asm_set_templates", 1 if enabled else 0)


def pair_mask(pairs):
    m = 0
    for rc, cc in pairs:
        m |= 1 << (rc * 4 + cc)
    return m


ALL_PAIRS = pair_mask([(r, c) for r in range(4) for c in range(4)])


class AffineSource:
    """Element matrices that are never materialised: `assemble(..., E=AffineSource(kernels, coef, which))` evaluates
    the entries of K_e ("K"), Rw_e ("Rw") or Rd_e ("Rd") of affine elements from their coefficients while it writes the
    CSR rows (fused integrate-and-assemble, csrc/assemble.cu asm_interior_affine_kernel)."""
    WHICH = {"K": 0, "Rw": 1, "Rd": 2}

    def __init__(self, kernels, coef, which):
        self.kernels, self.coef, self.which = kernels, coef.contiguous(), self.WHICH[which]


class SpmvPlan:
    """Owner of a pyn_spmv_plan handle (keeps the borrowed row-pointer tensor alive)."""

    def __init__(self, nrows, indptr, nnz):
        self.indptr = indptr
        self.handle = C.c_void_p()
        call("pyn_spmv_plan_create", nrows, ptr(indptr), nnz, stream(), C.byref(self.handle))

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                _lib.load().pyn_spmv_plan_destroy(self.handle)
                self.handle = None
        except Exception:
            pass


class TMat:
    """Element ("template") operator (pyn_tmat_*): y = A x of a matrix assembled from one element matrix on a
    repetitive mesh, matrix-free (element products + per-node gather), without streaming its values.  `asm` is the assembly record (CSR._asm) whose descriptor defines the interior rows, `csr` the
    matrix whose stored rows serve the boundary-affected nodes (the same object unless e.g. csr = K + Kfs)."""

    def __init__(self, asm, csr):
        topo, desc, keep = asm
        self._keep = (topo, desc, keep, csr.indptr, csr.indices, csr.data)
        self.handle = C.c_void_p()
        call("pyn_tmat_create", topo._h, C.byref(desc), ptr(csr.indptr), ptr(csr.indices), ptr(csr.data), stream(),
             C.byref(self.handle))
        self.bytes = int(_lib.load().pyn_tmat_bytes(self.handle))

    def __del__(self):
        try:
            if getattr(self, "handle", None):
                _lib.load().pyn_tmat_destroy(self.handle)
                self.handle = None
        except Exception:
            pass


def tmat_supported(nn, rdof, cdof):
    """Does an element matrix of rdof*nn x cdof*nn fit the element operator's shared-memory budget?"""
    return bool(_lib.load().pyn_tmat_supported(int(nn), int(rdof), int(cdof)))


def templates_eligible(csr):
    """Can `csr` (as returned by assemble) be applied through row templates?"""
    a = getattr(csr, "_asm", None)
    if a is None:
        return False
    topo, desc, _ = a
    return bool(desc.E_stride == 0 and desc.bcmap and (desc.pair_mask & 1) and desc.d_E and
                not (desc.diag_classes & 1) and _lib.load().pyn_tmat_supported(topo.nn, desc.rdof, desc.cdof))


class CSR:
    """Device CSR matrix (int64 row pointers, int32 sorted column indices, FP64 values)."""

    def __init__(self, shape, indptr, indices, data, name=""):
        self.shape, self.indptr, self.indices, self.data, self.name = tuple(shape), indptr, indices, data, name
        self._plan = None
        self._diag_off = None     # [offsets of the diagonal entries] (a shared cell: see diagonal())
        self._asm = None          # (topology, descriptor, keep-alive refs) when produced by assemble()
        self._tmat = None         # template operator, built on first use when enabled
        self._tmat_asm = None     # assembly record the templates come from (own or another matrix's), None = off

    def enable_templates(self, like=None):
        """Apply this matrix through row templates from now on (products and CG).  `like`: another assembled CSR that
        equals this matrix on the interior rows and carries the descriptor (e.g. K for K + Kfs).  Returns False --
        and keeps the CSR stream -- when the matrix is not eligible (per-cell element matrices, irregular mesh)."""
        src = like if like is not None else self
        if not templates_eligible(src):
            return False
        self._tmat_asm, self._tmat = src._asm, None     # only the record is kept: `like` itself may be freed
        return True

    def disable_templates(self):
        self._tmat_asm = self._tmat = None

    def tmat(self):
        if self._tmat is None and self._tmat_asm is not None:
            self._tmat = TMat(self._tmat_asm, self)
        return self._tmat

    def plan(self):
        """SpMV plan of this sparsity pattern (pyn_spmv_plan_create), built on first use and shared between the
        matrices that share `indptr` (values-only re-assembly)."""
        if self._plan is None:
            self._plan = SpmvPlan(self.shape[0], self.indptr, self.nnz)
        return self._plan.handle

    @property
    def nnz(self):
        return int(self.indices.numel())

    # -- PETSc-Mat flavoured surface ----------------------------------------------------------------------
    def mult(self, x, y, mode=0, p0=0.0, p1=0.0, aux=None, planned=True):
        xd = x.data if hasattr(x, "data") and not isinstance(x, torch.Tensor) else x
        yd = y.data if hasattr(y, "data") and not isinstance(y, torch.Tensor) else y
        ad = aux.data if (aux is not None and not isinstance(aux, torch.Tensor)) else aux
        assert xd.numel() == self.shape[1] and yd.numel() == self.shape[0]
        tm = self.tmat() if planned else None
        if tm is not None:
            call("pyn_tmat_mult", tm.handle, ptr(xd), ptr(yd), mode, float(p0), float(p1), ptr(ad), stream())
        elif planned:
            call("pyn_spmv_planned", self.plan(), ptr(self.indices), ptr(self.data), ptr(xd), ptr(yd), mode,
                 float(p0), float(p1), ptr(ad), stream())
        else:       # plan-less CSR-vector kernel (one-off products)
            call("pyn_spmv", self.shape[0], ptr(self.indptr), ptr(self.indices), ptr(self.data), self.nnz,
                 ptr(xd), ptr(yd), mode, float(p0), float(p1), ptr(ad), stream())
        return y

    def diagonal(self):
        """diag(A) (MatGetDiagonal).  The offset of the diagonal entry inside every row is found once per sparsity
        pattern and shared by the matrices re-assembled on it; later calls are a plain gather."""
        d = empty(self.shape[0])
        if self._diag_off is None:
            self._diag_off = [None]
        if self._diag_off[0] is None:
            off = empty(self.shape[0], I32)
            call("pyn_csr_diag_offsets", self.shape[0], ptr(self.indptr), ptr(self.indices), ptr(off), stream())
            self._diag_off[0] = off
        call("pyn_csr_diagonal_at", self.shape[0], ptr(self.indptr), ptr(self._diag_off[0]), ptr(self.data), ptr(d),
             stream())
        return d

    def add(self, other, alpha=1.0, name=""):
        """self + alpha*other on the union pattern (MatAXPY, DIFFERENT_NONZERO_PATTERN)."""
        assert self.shape == other.shape
        n = self.shape[0]
        cp = empty(n + 1, I64)
        nnz = C.c_int64()
        call("pyn_csr_add_count", n, ptr(self.indptr), ptr(self.indices), ptr(other.indptr), ptr(other.indices),
             ptr(cp), C.byref(nnz), stream())
        ci, cv = empty(nnz.value, I32), empty(nnz.value)
        call("pyn_csr_add_fill", n, ptr(self.indptr), ptr(self.indices), ptr(self.data), float(alpha),
             ptr(other.indptr), ptr(other.indices), ptr(other.data), ptr(cp), ptr(ci), ptr(cv), stream())
        return CSR(self.shape, cp, ci, cv, name)

    def compressed(self, name=""):
        """Copy with stored zeros dropped (timed operator SpMVs; the reference pattern stays in `self`)."""
        n = self.shape[0]
        op = empty(n + 1, I64)
        nnz = C.c_int64()
        call("pyn_csr_compress_count", n, ptr(self.indptr), ptr(self.data), ptr(op), C.byref(nnz), stream())
        oi, ov = empty(nnz.value, I32), empty(nnz.value)
        call("pyn_csr_compress_fill", n, ptr(self.indptr), ptr(self.indices), ptr(self.data), ptr(op), ptr(oi),
             ptr(ov), stream())
        return CSR(self.shape, op, oi, ov, name or self.name)

    # -- exporters (copies to host) -------------------------------------------------------------------------
    def getValuesCSR(self):
        """(indptr, indices, data) as NumPy, like petsc4py Mat.getValuesCSR()."""
        ip = self.indptr.cpu().numpy()
        if ip[-1] < 2 ** 31:
            ip = ip.astype(np.int32)
        return ip, self.indices.cpu().numpy(), self.data.cpu().numpy()

    def to_scipy(self):
        import scipy.sparse as sp
        ip, ix, dv = self.getValuesCSR()
        return sp.csr_matrix((dv, ix, ip), shape=self.shape)

    def to_petsc(self):
        from petsc4py import PETSc  # only when the user has it
        ip, ix, dv = self.getValuesCSR()
        return PETSc.Mat().createAIJ(size=self.shape, csr=(ip, ix, dv))

    def bytes_spmv(self, templates=None):
        """algorithmic HBM bytes of one y = A x.  CSR stream: 12 nnz + 8 (rows+1) + 8 rows + 8 cols;
        element operator: see pyn_tmat_bytes (x, connectivity, element results written + read, y)."""
        tm = self.tmat() if (templates is None or templates) else None
        if tm is not None and templates is not False:
            return tm.bytes
        return 12 * self.nnz + 8 * (self.shape[0] + 1) + 8 * self.shape[0] + 8 * self.shape[1]


def assemble(topo, rdof, cdof, E, rclass=None, cclass=None, pairs=ALL_PAIRS, diag_classes=0,
             diag_add=(0.0, 0.0, 0.0, 0.0), sign=1.0, rowscale=None, pattern=None, name="", values=True, bcmap=None,
             inplace=False):
    """Deterministic assembly of one matrix (pyn_asm_count + pyn_asm_fill).

    E: device tensor (1|ne, rdof*nn, cdof*nn).  pattern: an existing CSR whose indptr/indices are reused
    (values-only re-assembly); with `inplace` its value array is overwritten and `pattern` itself is returned (no
    allocation at all -- what a re-assembly inside a time loop wants).  Returns CSR."""
    d = AsmDesc()
    d.rdof, d.cdof = rdof, cdof
    d.d_rclass = rclass.data_ptr() if rclass is not None else None
    d.d_cclass = cclass.data_ptr() if cclass is not None else None
    d.pair_mask = pairs if isinstance(pairs, int) else pair_mask(pairs)
    d.diag_classes = diag_classes
    for i in range(4):
        d.diag_add[i] = float(diag_add[i])
    d.sign = float(sign)
    if isinstance(E, AffineSource):
        assert E.coef.shape[0] == topo.ne and E.kernels.nn == topo.nn and rowscale is None
        assert (rdof, cdof) == ((2, 2) if E.which == 0 else (2, 1))
        d.d_coef, d.elem, d.which = E.coef.data_ptr(), E.kernels._h, E.which
        d.E_stride = 1
    elif E is not None:
        E = E.contiguous()
        assert E.shape[-2] == rdof * topo.nn and E.shape[-1] == cdof * topo.nn
        d.d_E = E.data_ptr()
        d.E_stride = 0 if E.shape[0] == 1 else rdof * cdof * topo.nn * topo.nn
        if E.shape[0] != 1:
            assert E.shape[0] == topo.ne
    d.d_rowscale = rowscale.data_ptr() if rowscale is not None else None
    if bcmap is not None:
        assert bcmap.topo is topo and bcmap.rclass is rclass and bcmap.cclass is cclass, "bcmap built for other classes"
        d.bcmap = bcmap._h
    nrows, ncols = topo.nnodes * rdof, topo.nnodes * cdof
    if pattern is None:
        indptr = empty(nrows + 1, I64)
        nnz = C.c_int64()
        call("pyn_asm_count", topo._h, C.byref(d), ptr(indptr), C.byref(nnz), stream())
        indices = empty(nnz.value, I32)
        fill_idx = indices
    else:
        indptr, indices, fill_idx = pattern.indptr, pattern.indices, None
    keep = (E, rclass, cclass, rowscale, bcmap)
    if inplace and pattern is not None and pattern.data is not None:
        call("pyn_asm_fill", topo._h, C.byref(d), ptr(indptr), None, ptr(pattern.data), stream())
        own = pattern._tmat_asm is pattern._asm
        pattern._asm = (topo, d, keep)
        if own:
            pattern._tmat_asm = pattern._asm
        pattern._tmat = None          # templates (if enabled) are rebuilt from the new element matrix on next use
        return pattern
    data = empty(indices.numel()) if values else None
    call("pyn_asm_fill", topo._h, C.byref(d), ptr(indptr), ptr(fill_idx), ptr(data), stream())
    out = CSR((nrows, ncols), indptr, indices, data, name)
    out._asm = (topo, d, keep)
    if pattern is not None:
        out._plan = pattern._plan
        if pattern._diag_off is None:
            pattern._diag_off = [None]
        out._diag_off = pattern._diag_off
        if pattern._tmat_asm is not None and pattern._tmat_asm is pattern._asm:
            out._tmat_asm = out._asm
    return out


def assemble_weights(topo, We):
    We = We.contiguous()
    W = empty(topo.nnodes)
    call("pyn_asm_weights", topo._h, ptr(We), 0 if We.shape[0] == 1 else topo.nn, ptr(W), stream())
    return W


def coo_to_csr(rows, cols, vals, shape, name=""):
    """Generic deterministic COO -> CSR on device (Mat.setValues(ADD) + assemble semantics)."""
    n = int(rows.numel())
    indptr = empty(shape[0] + 1, I64)
    nnz = C.c_int64()
    scratch = C.c_void_p()
    call("pyn_coo_count", n, ptr(rows), ptr(cols), shape[0], shape[1], ptr(indptr), C.byref(nnz), C.byref(scratch),
         stream())
    indices, data = empty(nnz.value, I32), empty(nnz.value)
    call("pyn_coo_fill", scratch, ptr(vals), ptr(indptr), ptr(indices), ptr(data), stream())
    return CSR(shape, indptr, indices, data, name)


# ---------------------------------------------------------------------------------------------------------
# Immersed-boundary coupling (csrc/ibm.cu; reference src/cases/immersed_boundary.py:222-277, 318-338)
IBM_DELTA = {"four": 0, "three": 1, "linear": 2}


def _two_pass(entry, head, tail, n, pieces):
    """count pass over (list, piece) pairs -> prefix sums on device -> fill pass of one of the pyn_ibm_* entry points.
    Returns (ptr [n + 1] of the lists, ids, values)."""
    counts = empty(n * pieces, I32)
    call(entry, *head, None, ptr(counts), None, *tail)
    p = zeros(n * pieces + 1, I64)
    torch.cumsum(counts, 0, out=p[1:])
    total = int(p[-1])                       # one scalar read-back: the size of the output arrays
    ids, vals = empty(total, I32), empty(total)
    call(entry, *head, ptr(p), ptr(ids), ptr(vals), *tail)
    return p[::pieces].contiguous(), ids, vals


def ibm_collect(lagxy, cand, nodexy, dl, h, mode=0, kind="four"):
    """Per Lagrangian node the close candidate Euler nodes (ascending) and their delta values:
    (ptr [nlag+1] int64, nodes int32, dirac f64)."""
    nlag, ncand = lagxy.numel() // 2, int(cand.numel())
    head = (nlag, ptr(lagxy), ncand, ptr(cand), ptr(nodexy), float(dl), float(h), int(mode), IBM_DELTA[kind])
    return _two_pass("pyn_ibm_collect", head, (stream(),), nlag, int(_lib.load().pyn_ibm_chunks(ncand, None)))


def ibm_gram(p, nodes, dirac, a, b):
    """One velocity component of A = H S: (ptr [nlag+1], partner Lagrangian nodes j, values)."""
    nlag = p.numel() - 1
    head = (nlag, ptr(p), ptr(nodes), ptr(dirac), float(a), float(b))
    return _two_pass("pyn_ibm_gram", head, (stream(),), nlag, (nlag + 31) // 32)


def dof_blocked_csr(p, ids, vals, ncols_nodes, dim, name=""):
    """CSR with `dim` rows per list and the columns ids * dim + dof: row l * dim + dof holds list l on component dof
    (H of buildIBMMatrix: setValues(lagNode * dim + dof, eulerIndices[dof::dim], dirac), immersed_boundary.py:234-235).
    Index arithmetic only (allocator-side plumbing)."""
    n = p.numel() - 1
    cnt = p[1:] - p[:-1]
    indptr = zeros(n * dim + 1, I64)
    rowlen = cnt.repeat_interleave(dim)
    torch.cumsum(rowlen, 0, out=indptr[1:])
    owner = torch.repeat_interleave(torch.arange(n, device=ids.device), cnt)        # list of every entry
    pos = torch.arange(ids.numel(), device=ids.device) - p[:-1][owner]             # position inside its list
    indices = empty(ids.numel() * dim, I32)
    data = empty(ids.numel() * dim)
    for dof in range(dim):
        at = indptr[owner * dim + dof] + pos
        indices[at] = ids * dim + dof
        data[at] = vals
    return CSR((n * dim, ncols_nodes * dim), indptr, indices, data, name)


# ---------------------------------------------------------------------------------------------------------
class Ksp:
    """Device Krylov solver handle (pyn_ksp_*)."""
    TYPES = {"cg": 0, "gmres": 1}
    PCS = {"none": 0, "jacobi": 1, "mg": 2}

    def __init__(self, A: CSR, ksp_type="cg", pc="jacobi", rtol=1e-5, atol=1e-50, dtol=1e4, max_it=10000,
                 restart=30, check_every=0, nonzero_guess=False):
        self.A = A
        o = KspOpts()
        o.type, o.pc = self.TYPES[ksp_type], self.PCS[pc]
        o.rtol, o.atol, o.dtol, o.max_it, o.restart, o.check_every = rtol, atol, dtol, max_it, restart, check_every
        o.nonzero_guess = 1 if nonzero_guess else 0
        self.opts = o
        self._h = C.c_void_p()
        call("pyn_ksp_create", A.shape[0], ptr(A.indptr), ptr(A.indices), ptr(A.data), A.nnz, C.byref(o), stream(),
             C.byref(self._h))
        self.its, self.reason, self.rnorm0, self.rnorm = 0, 0, 0.0, 0.0
        self._tm = A.tmat()                                     # element operator for w = A p when enabled on A
        if self._tm is not None:
            call("pyn_ksp_set_tmat", self._h, self._tm.handle)

    def solve(self, b, x):
        r = KspResult()
        call("pyn_ksp_solve", self._h, ptr(b), ptr(x), C.byref(r), stream())
        self.its, self.reason, self.rnorm0, self.rnorm = r.its, r.reason, r.rnorm0, r.rnorm
        return r

    def __del__(self):
        try:
            if getattr(self, "_h", None):
                _lib.load().pyn_ksp_destroy(self._h)
                self._h = None
        except Exception:
            pass


# ---------------------------------------------------------------------------------------------------------
def vec_dot(x, y):
    out = C.c_double()
    call("pyn_vec_dot", x.numel(), ptr(x), ptr(y), C.byref(out), stream())
    return out.value


def vec_norm(x, kind=2):
    out = C.c_double()
    call("pyn_vec_norm", x.numel(), ptr(x), kind, C.byref(out), stream())
    return out.value


def vec_axpy(y, a, x):
    call("pyn_vec_axpy", y.numel(), float(a), ptr(x), ptr(y), stream())


def vec_scale(x, a):
    call("pyn_vec_scale", x.numel(), float(a), ptr(x), stream())


def vec_set(x, a):
    call("pyn_vec_set", x.numel(), float(a), ptr(x), stream())


def vec_maxpy(y, y0, coefs, xs):
    m = len(coefs)
    cf = (C.c_double * m)(*[float(c) for c in coefs])
    px = (C.c_void_p * m)(*[x.data_ptr() for x in xs])
    call("pyn_vec_maxpy", y.numel(), m, cf, px, ptr(y0), ptr(y), stream())


def vec_wrms(err, y, ynew, atol, rtol):
    out = C.c_double()
    call("pyn_vec_wrms", err.numel(), ptr(err), ptr(y), ptr(ynew), float(atol), float(rtol), C.byref(out), stream())
    return out.value


def vec_scatter_set(x, idx, vals):
    call("pyn_vec_scatter_set", idx.numel(), ptr(idx), ptr(vals), ptr(x), stream())


def vec_scatter_const(x, idx, val):
    call("pyn_vec_scatter_const", idx.numel(), ptr(idx), float(val), ptr(x), stream())


def vtensv(vel, out):
    call("pyn_vtensv", vel.numel() // 2, ptr(vel), ptr(out), stream())


def taylor_green(xy, nu, t, vel=None, vort=None, idx=None):
    n = idx.numel() if idx is not None else xy.shape[0]
    call("pyn_taylor_green", n, ptr(idx), ptr(xy), float(nu), float(t), ptr(vel), ptr(vort), stream())
