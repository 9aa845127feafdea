"""Multi-GPU context of one rank: NCCL communicator + halo-exchange plan + ownership masks on the device.

`torch.distributed` (one process per GPU, launched by torchrun) is used for the rendezvous only -- it carries the
128-byte NCCL id from rank 0 to the others; every data-path collective (halo exchange of interface-node dofs,
all-reduce of the Krylov / time-step scalars) goes through the library's own communicator (csrc/dist.cu), so that CG
can issue them from inside its captured iteration graph.  In the reference these are PETSc's VecScatter and
MPI_Allreduce."""
import ctypes as C
import os

import numpy as np
import torch

from . import _lib
from . import backend as be
from ._lib import call
from .domain.partition import StripPartition


def world_info():
    """(rank, world) of the torch.distributed job, (0, 1) when there is none."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(), dist.get_world_size()
    return 0, 1


class DistContext(object):
    def __init__(self, partition: StripPartition, parent=None):
        """parent: another context of the same job whose communicator and peer region this one shares (same strips at
        another element order: the coarse levels of the multigrid hierarchy) -- only the halo plan is its own."""
        import torch.distributed as dist
        self.part = partition
        self.rank, self.world = partition.rank, partition.world
        self._parent = parent
        if parent is not None:
            assert (parent.rank, parent.world) == (self.rank, self.world)
            self._comm = parent._comm
        else:
            ident = C.create_string_buffer(128)
            if self.world > 1:
                if self.rank == 0:
                    call("pyn_comm_unique_id", ident)
                box = [bytes(ident.raw)]
                dist.broadcast_object_list(box, src=0)
                ident = C.create_string_buffer(box[0], 128)
            self._comm = C.c_void_p()
            call("pyn_comm_create", ident, self.rank, self.world, C.byref(self._comm))
        own = partition.owned_mask()
        self.n_owned = int(own.sum())
        self.owned = be.to_device(own.astype(np.uint8))
        self._owned_rows = {1: self.owned}
        # halo plan
        lists = partition.halo_lists()
        self._idx = [(be.to_device(s), be.to_device(r)) for _, s, r in lists]      # kept alive: the plan borrows them
        n = len(lists)
        nbr = (C.c_int32 * max(n, 1))(*[q for q, _, _ in lists])
        scnt = (C.c_int64 * max(n, 1))(*[len(s) for _, s, _ in lists])
        rcnt = (C.c_int64 * max(n, 1))(*[len(r) for _, _, r in lists])
        sidx = (C.c_void_p * max(n, 1))(*[t[0].data_ptr() for t in self._idx])
        ridx = (C.c_void_p * max(n, 1))(*[t[1].data_ptr() for t in self._idx])
        self._halo = C.c_void_p()
        call("pyn_halo_create", self._comm, n, nbr, scnt, sidx, rcnt, ridx, C.byref(self._halo))
        self._scal = be.empty(4)
        self._peer = None
        self.peer_memory = False
        if parent is not None:          # the parent's mailboxes were sized for its (finer) interface: large enough
            self.peer_memory = parent.peer_memory
        elif self.world > 1 and os.environ.get("PYN_P2P", "1") not in ("0", "false", "off"):
            self._attach_peer_memory(max([len(r) for _, _, r in lists] + [len(s) for _, s, _ in lists] + [1]))

    def _attach_peer_memory(self, max_nodes):
        """Map every rank's exchange region over NVLink (CUDA IPC) and route the Krylov scalars' all-reduces and the
        halo exchanges through it (csrc/dist.cu, peer-memory path).  Collective.  Falls back to NCCL -- on every rank
        alike -- when any rank cannot map its peers."""
        import torch.distributed as dist
        cap = torch.tensor([int(max_nodes) * 3], dtype=torch.int64, device=self.owned.device)
        dist.all_reduce(cap, op=dist.ReduceOp.MAX)
        ok, peer = 1, C.c_void_p()
        try:
            call("pyn_peer_create", self.rank, self.world, int(cap[0]), C.byref(peer))
            h = C.create_string_buffer(64)
            call("pyn_peer_handle", peer, h)
            mine = bytes(h.raw)
        except _lib.PynamaError:
            ok, mine = 0, b"\0" * 64
        handles = [None] * self.world
        dist.all_gather_object(handles, mine)
        if ok:
            try:
                call("pyn_peer_open", peer, C.create_string_buffer(b"".join(handles), 64 * self.world))
            except _lib.PynamaError:
                ok = 0
        flag = torch.tensor([ok], dtype=torch.int64, device=self.owned.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN)
        if int(flag[0]) == 1:
            call("pyn_comm_set_peer", self._comm, peer)
            self._peer, self.peer_memory = peer, True
        elif peer:
            _lib.load().pyn_peer_destroy(peer)
        dist.barrier()

    def owned_rows(self, ncomp):
        """uint8 per dof (node*ncomp + c): 1 when the node is owned."""
        if ncomp not in self._owned_rows:
            self._owned_rows[ncomp] = self.owned.repeat_interleave(ncomp).contiguous()
        return self._owned_rows[ncomp]

    def exchange(self, vec, ncomp):
        """Refresh the ghost entries of a local vector (tensor or Vec) with their owners' values."""
        t = vec.data if not isinstance(vec, torch.Tensor) else vec
        if self.world > 1:
            call("pyn_halo_exchange", self._halo, be.ptr(t), int(ncomp), be.stream())
        return vec

    def halo_bytes(self, ncomp):
        return int(_lib.load().pyn_halo_bytes(self._halo, int(ncomp)))

    def allreduce_sum(self, values):
        """Sum a few host scalars over the ranks (device all-reduce through the library communicator)."""
        vals = [float(v) for v in values]
        if self.world == 1:
            return vals
        self._scal[:len(vals)] = torch.tensor(vals, dtype=torch.float64)
        call("pyn_comm_allreduce_sum", self._comm, be.ptr(self._scal), len(vals), be.stream())
        return self._scal[:len(vals)].cpu().tolist()

    def allreduce_max(self, value):
        """Max of one host scalar over the ranks (rendezvous-layer collective: off the hot path)."""
        if self.world == 1:
            return float(value)
        import torch.distributed as dist
        t = torch.tensor([float(value)], dtype=torch.float64, device=self.owned.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t[0])

    def attach(self, ksp, ncomp=2):
        """Make a backend.Ksp distributed (halo exchange of the search direction + all-reduced scalars)."""
        call("pyn_ksp_set_dist", ksp._h, self._comm, self._halo, int(ncomp), be.ptr(self.owned_rows(ncomp)))
        ksp._dist = self         # keep the context alive as long as the solver

    def gather_owned(self, vec, ncomp):
        """(global dof ids, values) of the owned entries of a local vector -- for comparisons in the serial numbering."""
        t = vec.data if not isinstance(vec, torch.Tensor) else vec
        own = self.part.owned_mask()
        l2g = self.part.local_to_global()[own]
        vals = t.view(-1, ncomp)[be.to_device(np.flatnonzero(own))].cpu().numpy()
        dofs = (l2g[:, None] * ncomp + np.arange(ncomp)).reshape(-1)
        return dofs, vals.reshape(-1)

    def close(self):
        """Release the halo plan and the NCCL communicator.  Collective in spirit: call it on every rank, after the
        solvers that were attached to this context are gone, and before torch.distributed is shut down.  It is NOT
        called from __del__: tearing NCCL down from a garbage-collector callback at interpreter exit (other ranks may
        already be gone, captured graphs may still hold NCCL nodes) can block forever."""
        lib = _lib.load()
        torch.cuda.synchronize()
        if getattr(self, "_halo", None):
            lib.pyn_halo_destroy(self._halo)
            self._halo = None
        if getattr(self, "_peer", None):
            lib.pyn_comm_set_peer(self._comm, None)
            lib.pyn_peer_destroy(self._peer)
            self._peer = None
        if getattr(self, "_comm", None) and self._parent is None:
            lib.pyn_comm_destroy(self._comm)
        self._comm = None
