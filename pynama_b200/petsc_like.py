"""Device-backed `Vec` / `Mat` with the slice of the petsc4py surface the reference's hot path uses
(base_problem.py:212-252,274-289,321-327,479-481; mat_generator.py:14-17,95-118,157-190; cavity.py:53-81;
custom_func.py:57-83; dmplex.py:263-296).  Storage is HBM (torch tensors); every arithmetic method is a kernel of
libpynama_b200.so.  Exporters: getArray()/getValuesCSR() copy to host, to_petsc() wraps them in real petsc4py
objects when petsc4py is importable."""
import numpy as np
import torch

from . import backend as be


class NormType:
    """petsc4py.PETSc.NormType values."""
    NORM_1, NORM_2, NORM_FROBENIUS, NORM_INFINITY = 0, 1, 2, 3
    N1, N2, FROBENIUS, FRB, INFINITY, INF = 0, 1, 2, 2, 3, 3


# Multi-GPU: set by the case layer (cases/base_problem.py setUpDomain) to the rank's DistContext.  Norms and dot products
# of vectors laid out on the partitioned mesh (size = local nodes x ncomp) then count owned entries only and are summed /
# maxed over the ranks, like PETSc's parallel Vec.
_DIST = None


def set_dist(ctx):
    global _DIST
    _DIST = ctx


def _owned_view(t):
    """(tensor restricted to owned entries, dist context) or (t, None) when the vector is not a partitioned field."""
    d = _DIST
    if d is None or d.world == 1:
        return t, None
    n = t.numel()
    if n == 0 or n % d.part.nnodes:
        return t, None
    return t * d.owned_rows(n // d.part.nnodes).to(t.dtype), d


class Vec(object):
    def __init__(self, n=None, data=None, name=""):
        self.data = data if data is not None else be.zeros(n)
        self._name = name

    # -- creation -------------------------------------------------------------------------------------------
    @staticmethod
    def createWithArray(arr):
        return Vec(data=be.to_device(np.asarray(arr, dtype=np.float64)))

    @staticmethod
    def createSeq(n):
        return Vec(n)

    def duplicate(self):
        return Vec(self.data.numel())

    def copy(self, out=None):
        if out is None:
            return Vec(data=self.data.clone(), name=self._name)
        out.data.copy_(self.data)
        return out

    def destroy(self):
        self.data = None

    # -- metadata -------------------------------------------------------------------------------------------
    def setName(self, name):
        self._name = name

    def getName(self):
        return self._name

    def getSize(self):
        return int(self.data.numel())

    size = property(getSize)

    def getOwnershipRange(self):
        return 0, self.getSize()

    owner_range = property(getOwnershipRange)

    # -- values ---------------------------------------------------------------------------------------------
    def set(self, value):
        be.vec_set(self.data, value)

    def getArray(self, readonly=False):
        """Host copy (device -> host)."""
        return self.data.cpu().numpy()

    array = property(getArray)

    def setArray(self, arr):
        self.data.copy_(torch.as_tensor(np.asarray(arr, dtype=np.float64)))

    def getValues(self, indices):
        idx = torch.as_tensor(np.asarray(indices, dtype=np.int64), device=self.data.device)
        return self.data[idx].cpu().numpy()

    def setValues(self, indices, values, addv=False):
        idx = np.atleast_1d(np.asarray(indices, dtype=np.int64))
        vals = np.atleast_1d(np.asarray(values, dtype=np.float64)).reshape(-1)
        if idx.size == 0:
            return
        if vals.size == 1 and idx.size > 1:
            vals = np.repeat(vals, idx.size)
        di, dv = be.to_device(idx), be.to_device(vals)
        if addv:
            self.data.index_add_(0, di, dv)      # host-driven, off the hot path (sequential adds are tiny sets)
        else:
            be.vec_scatter_set(self.data, di, dv)

    def setValue(self, index, value, addv=False):
        self.setValues([index], [value], addv)

    def assemble(self):
        pass

    def assemblyBegin(self):
        pass

    def assemblyEnd(self):
        pass

    # -- algebra --------------------------------------------------------------------------------------------
    def axpy(self, alpha, x):
        be.vec_axpy(self.data, alpha, x.data)

    def aypx(self, alpha, x):
        be.call("pyn_vec_aypx", self.data.numel(), float(alpha), be.ptr(x.data), be.ptr(self.data), be.stream())

    def scale(self, alpha):
        be.vec_scale(self.data, alpha)

    def reciprocal(self):
        self.data.reciprocal_()

    def dot(self, other):
        t, d = _owned_view(self.data)
        v = be.vec_dot(t, other.data)
        return d.allreduce_sum([v])[0] if d is not None else v

    def norm(self, norm_type=NormType.NORM_2):
        """petsc4py semantics: 0 = 1-norm, 1 and 2 (Frobenius) = 2-norm, 3 = max norm; global over the ranks."""
        t, d = _owned_view(self.data)
        if norm_type in (NormType.NORM_2, NormType.NORM_FROBENIUS):
            v = be.vec_norm(t, 2)
            return float(np.sqrt(d.allreduce_sum([v * v])[0])) if d is not None else v
        if norm_type == NormType.NORM_INFINITY:
            v = be.vec_norm(t, 3)
            return d.allreduce_max(v) if d is not None else v
        if norm_type == NormType.NORM_1:
            v = float(t.abs().sum().item())
            return d.allreduce_sum([v])[0] if d is not None else v
        raise Exception(f"unknown norm type {norm_type}")

    def max(self):
        v, i = torch.max(self.data, 0)
        return int(i), float(v)

    def __imul__(self, alpha):
        self.scale(alpha)
        return self

    def __mul__(self, other):
        if isinstance(other, Vec):
            return Vec(data=self.data * other.data)
        out = self.copy()
        out.scale(other)
        return out

    __rmul__ = __mul__

    def __add__(self, other):
        out = self.copy()
        out.axpy(1.0, other)
        return out

    def __sub__(self, other):
        out = self.copy()
        out.axpy(-1.0, other)
        return out

    def __len__(self):
        return self.getSize()

    def __array__(self, dtype=None, copy=None):
        a = self.getArray()
        return a.astype(dtype) if dtype is not None else a

    def to_petsc(self):
        from petsc4py import PETSc
        return PETSc.Vec().createWithArray(self.getArray())


class Mat(object):
    """AIJ matrix on device.  Either produced directly by the mesh-aware assembly (`fromCSR`) or filled through
    `setValues(..., addv=True)` + `assemble()` like a PETSc Mat (generic deterministic COO -> CSR on the GPU)."""

    def __init__(self, shape=None, name=""):
        self.shape = tuple(shape) if shape is not None else None
        self._name = name
        self.csr = None
        self._coo = []          # staged (rows, cols, vals) host triplets
        self._ins = []          # INSERT_VALUES entries applied after the sums

    @staticmethod
    def fromCSR(csr, name=""):
        m = Mat(csr.shape, name or csr.name)
        m.csr = csr
        return m

    # -- PETSc-style filling --------------------------------------------------------------------------------
    def setValues(self, rows, cols, values, addv=False):
        r = np.atleast_1d(np.asarray(rows, dtype=np.int64))
        c = np.atleast_1d(np.asarray(cols, dtype=np.int64))
        v = np.asarray(values, dtype=np.float64)
        if v.ndim == 0:
            v = np.full((r.size, c.size), float(v))
        v = v.reshape(r.size, c.size)
        R = np.repeat(r, c.size)
        Cc = np.tile(c, r.size)
        (self._coo if addv else self._ins).append((R, Cc, v.reshape(-1)))

    def assemble(self):
        if not self._coo and not self._ins:
            return
        parts = list(self._coo)
        if self.csr is not None:                       # later additions onto an assembled matrix
            ip, ix, dv = self.csr.getValuesCSR()
            rows = np.repeat(np.arange(self.shape[0], dtype=np.int64), np.diff(ip))
            parts.insert(0, (rows, ix.astype(np.int64), dv))
        for R, Cc, v in self._ins:                     # make sure inserted entries exist in the pattern
            parts.append((R, Cc, np.zeros_like(v)))
        R = np.concatenate([p[0] for p in parts]); Cc = np.concatenate([p[1] for p in parts])
        v = np.concatenate([p[2] for p in parts])
        self.csr = be.coo_to_csr(be.to_device(R.astype(np.int32)), be.to_device(Cc.astype(np.int32)),
                                 be.to_device(v), self.shape, self._name)
        if self._ins:
            ip = self.csr.indptr.cpu().numpy()
            ix = self.csr.indices.cpu().numpy()
            for R, Cc, v in self._ins:
                pos = np.array([ip[r] + np.searchsorted(ix[ip[r]:ip[r + 1]], c) for r, c in zip(R, Cc)], dtype=np.int64)
                be.vec_scatter_set(self.csr.data, be.to_device(pos), be.to_device(v))
        self._coo, self._ins = [], []

    # -- metadata -------------------------------------------------------------------------------------------
    def setName(self, name):
        self._name = name
        if self.csr is not None:
            self.csr.name = name

    def getName(self):
        return self._name

    def getSize(self):
        return self.shape

    def getOwnershipRange(self):
        return 0, self.shape[0]

    def setUp(self):
        pass

    def destroy(self):
        self.csr = None

    def getInfo(self):
        nz = self.csr.nnz if self.csr is not None else 0
        return {"memory": 12 * nz + 8 * (self.shape[0] + 1), "nz_used": nz, "nz_unneeded": 0, "nz_allocated": nz}

    def createVecRight(self):
        return Vec(self.shape[1])

    def createVecLeft(self):
        return Vec(self.shape[0])

    # -- algebra --------------------------------------------------------------------------------------------
    def mult(self, x, y):
        self.csr.mult(x.data, y.data)

    def multAdd(self, x, v, y):
        if y is not v:
            y.data.copy_(v.data)
        self.csr.mult(x.data, y.data, mode=1)

    def __mul__(self, x):
        y = self.createVecLeft()
        self.mult(x, y)
        return y

    def __add__(self, other):
        return Mat.fromCSR(self.csr.add(other.csr, 1.0), f"{self._name}+{other._name}")

    def getDiagonal(self):
        return Vec(data=self.csr.diagonal())

    def diagonalScale(self, L=None, R=None):
        if L is not None:
            counts = (self.csr.indptr[1:] - self.csr.indptr[:-1])
            self.csr.data.mul_(torch.repeat_interleave(L.data, counts))
        if R is not None:
            self.csr.data.mul_(R.data[self.csr.indices.long()])

    # -- exporters ------------------------------------------------------------------------------------------
    def getValuesCSR(self):
        return self.csr.getValuesCSR()

    def to_scipy(self):
        return self.csr.to_scipy()

    def to_petsc(self):
        return self.csr.to_petsc()
