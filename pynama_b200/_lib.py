"""ctypes binding of libpynama_b200.so (include/pynama_b200.h).

The product path has NO CPU fallback: if the shared library is missing or a call fails, an exception is raised.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpynama_b200.so")

c_i64, c_i32, c_f64, c_vp = C.c_int64, C.c_int32, C.c_double, C.c_void_p


class PynamaError(RuntimeError):
    pass


class AsmDesc(C.Structure):
    """pyn_asm_desc"""
    _fields_ = [("rdof", c_i32), ("cdof", c_i32), ("d_rclass", c_vp), ("d_cclass", c_vp),
                ("pair_mask", C.c_uint32), ("diag_classes", C.c_uint32), ("diag_add", c_f64 * 4),
                ("sign", c_f64), ("d_E", c_vp), ("E_stride", c_i64), ("d_rowscale", c_vp), ("bcmap", c_vp),
                ("d_coef", c_vp), ("elem", c_vp), ("which", c_i32)]


class KspOpts(C.Structure):
    """pyn_ksp_opts"""
    _fields_ = [("type", c_i32), ("pc", c_i32), ("rtol", c_f64), ("atol", c_f64), ("dtol", c_f64),
                ("max_it", c_i32), ("restart", c_i32), ("check_every", c_i32), ("nonzero_guess", c_i32)]


class KspResult(C.Structure):
    """pyn_ksp_result"""
    _fields_ = [("its", c_i32), ("reason", c_i32), ("rnorm0", c_f64), ("rnorm", c_f64)]


P = C.POINTER
_SIGNATURES = {
    "pyn_last_error": (C.c_char_p, []),
    "pyn_version": (C.c_int, []),
    "pyn_launch_count": (c_i64, []),
    "pyn_device_info": (C.c_int, [P(C.c_int), P(C.c_int), P(C.c_int), P(c_i64)]),
    "pyn_elem_create": (C.c_int, [C.c_int, C.c_int, c_vp, c_vp, c_vp, c_vp, C.c_int, c_vp, c_vp, c_vp, c_vp,
                                  c_vp, c_vp, c_vp, c_vp, c_f64, c_f64, P(c_vp)]),
    "pyn_elem_destroy": (C.c_int, [c_vp]),
    "pyn_elem_kle": (C.c_int, [c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp]),
    "pyn_elem_affine_coef": (C.c_int, [c_vp, c_vp, c_i64, c_vp, P(c_i64), c_vp]),
    "pyn_elem_kle_affine": (C.c_int, [c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp]),
    "pyn_elem_ops": (C.c_int, [c_vp, c_vp, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "pyn_box_cell_nodes": (C.c_int, [C.c_int, C.c_int, C.c_int, c_i64, c_i64, c_vp, c_vp]),
    "pyn_box_cell_corners": (C.c_int, [C.c_int, C.c_int, c_f64, c_f64, c_f64, c_f64, c_i64, c_i64, c_vp, c_vp]),
    "pyn_box_node_coords": (C.c_int, [C.c_int, C.c_int, C.c_int, c_f64, c_f64, c_f64, c_f64, c_vp, c_vp, c_vp, c_vp]),
    "pyn_topo_create": (C.c_int, [c_vp, c_i64, C.c_int, c_i64, c_vp, P(c_vp)]),
    "pyn_topo_destroy": (C.c_int, [c_vp]),
    "pyn_topo_info": (C.c_int, [c_vp, P(c_i64), P(C.c_int), P(C.c_int)]),
    "pyn_topo_adjacency": (C.c_int, [c_vp, P(c_vp), P(c_vp)]),
    "pyn_bcmap_create": (C.c_int, [c_vp, c_vp, C.c_int, c_vp, C.c_int, c_vp, P(c_vp)]),
    "pyn_bcmap_destroy": (C.c_int, [c_vp]),
    "pyn_bcmap_info": (C.c_int, [c_vp, P(c_i64)]),
    "pyn_asm_count": (C.c_int, [c_vp, P(AsmDesc), c_vp, P(c_i64), c_vp]),
    "pyn_asm_fill": (C.c_int, [c_vp, P(AsmDesc), c_vp, c_vp, c_vp, c_vp]),
    "pyn_asm_set_templates": (C.c_int, [C.c_int]),
    "pyn_asm_defer_join": (C.c_int, [C.c_int]),
    "pyn_asm_join": (C.c_int, [c_vp, c_vp]),
    "pyn_topo_signatures": (C.c_int, [c_vp, P(C.c_int), P(c_i64)]),
    "pyn_asm_weights": (C.c_int, [c_vp, c_vp, c_i64, c_vp, c_vp]),
    "pyn_coo_count": (C.c_int, [c_i64, c_vp, c_vp, c_i64, c_i64, c_vp, P(c_i64), P(c_vp), c_vp]),
    "pyn_coo_fill": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "pyn_csr_add_count": (C.c_int, [c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, P(c_i64), c_vp]),
    "pyn_csr_add_fill": (C.c_int, [c_i64, c_vp, c_vp, c_vp, c_f64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "pyn_csr_diagonal": (C.c_int, [c_i64, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "pyn_csr_diag_offsets": (C.c_int, [c_i64, c_vp, c_vp, c_vp, c_vp]),
    "pyn_csr_diagonal_at": (C.c_int, [c_i64, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "pyn_csr_compress_count": (C.c_int, [c_i64, c_vp, c_vp, c_vp, P(c_i64), c_vp]),
    "pyn_csr_compress_fill": (C.c_int, [c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "pyn_spmv": (C.c_int, [c_i64, c_vp, c_vp, c_vp, c_i64, c_vp, c_vp, C.c_int, c_f64, c_f64, c_vp, c_vp]),
    "pyn_spmv_plan_create": (C.c_int, [c_i64, c_vp, c_i64, c_vp, P(c_vp)]),
    "pyn_spmv_plan_destroy": (C.c_int, [c_vp]),
    "pyn_spmv_planned": (C.c_int, [c_vp, c_vp, c_vp, c_vp, c_vp, C.c_int, c_f64, c_f64, c_vp, c_vp]),
    "pyn_ibm_collect": (C.c_int, [c_i64, c_vp, c_i64, c_vp, c_vp, c_f64, c_f64, C.c_int, C.c_int, c_vp, c_vp, c_vp, c_vp]),
    "pyn_ibm_chunks": (C.c_int, [c_i64, c_vp]),
    "pyn_ibm_gram": (C.c_int, [c_i64, c_vp, c_vp, c_vp, c_f64, c_f64, c_vp, c_vp, c_vp, c_vp]),
    "pyn_vec_set": (C.c_int, [c_i64, c_f64, c_vp, c_vp]),
    "pyn_vec_scale": (C.c_int, [c_i64, c_f64, c_vp, c_vp]),
    "pyn_vec_axpy": (C.c_int, [c_i64, c_f64, c_vp, c_vp, c_vp]),
    "pyn_vec_aypx": (C.c_int, [c_i64, c_f64, c_vp, c_vp, c_vp]),
    "pyn_vec_waxpy": (C.c_int, [c_i64, c_f64, c_vp, c_vp, c_vp, c_vp]),
    "pyn_vec_maxpy": (C.c_int, [c_i64, C.c_int, P(c_f64), P(c_vp), c_vp, c_vp, c_vp]),
    "pyn_vec_dot": (C.c_int, [c_i64, c_vp, c_vp, P(c_f64), c_vp]),
    "pyn_vec_norm": (C.c_int, [c_i64, c_vp, C.c_int, P(c_f64), c_vp]),
    "pyn_vec_dot_noalloc": (C.c_int, [c_i64, c_vp, c_vp, c_vp, c_vp]),
    "pyn_vec_scatter_set": (C.c_int, [c_i64, c_vp, c_vp, c_vp, c_vp]),
    "pyn_vec_scatter_const": (C.c_int, [c_i64, c_vp, c_f64, c_vp, c_vp]),
    "pyn_vtensv": (C.c_int, [c_i64, c_vp, c_vp, c_vp]),
    "pyn_taylor_green": (C.c_int, [c_i64, c_vp, c_vp, c_f64, c_f64, c_vp, c_vp, c_vp]),
    "pyn_vec_wrms": (C.c_int, [c_i64, c_vp, c_vp, c_vp, c_f64, c_f64, P(c_f64), c_vp]),
    "pyn_ksp_create": (C.c_int, [c_i64, c_vp, c_vp, c_vp, c_i64, P(KspOpts), c_vp, P(c_vp)]),
    "pyn_ksp_solve": (C.c_int, [c_vp, c_vp, c_vp, P(KspResult), c_vp]),
    "pyn_ksp_destroy": (C.c_int, [c_vp]),
    "pyn_ksp_stats": (C.c_int, [c_vp, P(c_i64), P(c_i64)]),
    "pyn_box_node_coords_strip": (C.c_int, [C.c_int, C.c_int, c_i64, C.c_int, C.c_int, c_f64, c_f64, c_f64, c_f64, c_vp,
                                            c_vp, c_vp, c_vp]),
    "pyn_comm_unique_id": (C.c_int, [c_vp]),
    "pyn_comm_create": (C.c_int, [c_vp, C.c_int, C.c_int, P(c_vp)]),
    "pyn_comm_destroy": (C.c_int, [c_vp]),
    "pyn_comm_info": (C.c_int, [c_vp, P(C.c_int), P(C.c_int), P(C.c_int)]),
    "pyn_comm_allreduce_sum": (C.c_int, [c_vp, c_vp, C.c_int, c_vp]),
    "pyn_peer_create": (C.c_int, [C.c_int, C.c_int, c_i64, P(c_vp)]),
    "pyn_peer_handle": (C.c_int, [c_vp, c_vp]),
    "pyn_peer_open": (C.c_int, [c_vp, c_vp]),
    "pyn_peer_destroy": (C.c_int, [c_vp]),
    "pyn_comm_set_peer": (C.c_int, [c_vp, c_vp]),
    "pyn_halo_create": (C.c_int, [c_vp, C.c_int, P(c_i32), P(c_i64), P(c_vp), P(c_i64), P(c_vp), P(c_vp)]),
    "pyn_halo_destroy": (C.c_int, [c_vp]),
    "pyn_halo_exchange": (C.c_int, [c_vp, c_vp, C.c_int, c_vp]),
    "pyn_halo_bytes": (c_i64, [c_vp, C.c_int]),
    "pyn_tmat_create": (C.c_int, [c_vp, P(AsmDesc), c_vp, c_vp, c_vp, c_vp, P(c_vp)]),
    "pyn_tmat_destroy": (C.c_int, [c_vp]),
    "pyn_tmat_mult": (C.c_int, [c_vp, c_vp, c_vp, C.c_int, c_f64, c_f64, c_vp, c_vp]),
    "pyn_tmat_bytes": (c_i64, [c_vp]),
    "pyn_tmat_supported": (C.c_int, [C.c_int, C.c_int, C.c_int]),
    "pyn_ksp_set_tmat": (C.c_int, [c_vp, c_vp]),
    "pyn_mg_create": (C.c_int, [C.c_int, C.c_int, c_f64, P(c_vp)]),
    "pyn_mg_set_level": (C.c_int, [c_vp, C.c_int, c_i64, c_vp, c_vp, c_vp, c_vp, c_vp, c_f64]),
    "pyn_mg_set_transfer": (C.c_int, [c_vp, C.c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp]),
    "pyn_mg_set_coarse_inverse": (C.c_int, [c_vp, c_vp]),
    "pyn_mg_apply": (C.c_int, [c_vp, c_vp, c_vp, c_vp]),
    "pyn_mg_destroy": (C.c_int, [c_vp]),
    "pyn_ksp_set_mg": (C.c_int, [c_vp, c_vp]),
    "pyn_mg_set_dist": (C.c_int, [c_vp, c_vp, c_vp, C.c_int]),
    "pyn_mg_set_level_dist": (C.c_int, [c_vp, C.c_int, c_vp, c_vp, C.c_int]),
    "pyn_ksp_set_dist": (C.c_int, [c_vp, c_vp, c_vp, C.c_int, c_vp]),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)

_lib = None


def load():
    """Load the shared library (once).  Raises PynamaError when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise PynamaError(f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
                          "or `make -C pynama_b200/csrc` (there is no CPU fallback)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc):
    if rc != 0:
        msg = load().pyn_last_error()
        raise PynamaError(f"libpynama_b200 error {rc}: {msg.decode() if msg else '?'}")


def call(name, *args):
    """Call an int-returning entry point and raise on error."""
    check(getattr(load(), name)(*args))
