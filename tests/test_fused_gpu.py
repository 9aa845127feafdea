"""Fused integrate-and-assemble on affine elements (csrc/elem_affine.cuh, asm_interior_affine_kernel): the per-cell path
of FreeSlip.buildKLEMats (src/cases/base_problem.py:499-552) without materialised element matrices.

  * closed-form element matrices == the general sum-factorised kernel == the oracle (1e-12), rectangles and sheared cells
  * fused assembly bit-identical to "materialise with the same closed form, then gather"
  * assembled K, Krhs, Rw vs the oracle's per-cell assembly: pattern bit-exact, values 1e-12
  * a mesh with one non-affine cell falls back to the general kernels (and still matches the oracle)"""
import os
import subprocess
import sys

import numpy as np
import pytest
import yaml

import kle_oracle as o
from conftest import ROOT, rel_err

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def be():
    import torch
    if not torch.cuda.is_available():
        pytest.skip("no CUDA device")
    from pynama_b200 import backend
    return backend


class ShearedMesh(o.BoxMesh):
    """Box numbering with corner coordinates mapped by an affine (shear + stretch) or a non-affine map."""

    def __init__(self, nelem, ngl, kind="shear"):
        super().__init__(nelem, [0.0, 0.0], [2.0, 0.7], ngl)
        self.kind = kind

    def cell_corner_coords(self, cells=None):
        cc = super().cell_corner_coords(cells).copy()
        x, y = cc[..., 0].copy(), cc[..., 1].copy()
        cc[..., 0] = 1.1 * x + 0.35 * y
        cc[..., 1] = 0.9 * y - 0.2 * x
        if self.kind == "bent":                     # one vertex moved: the cells around it are no longer parallelograms
            cc[..., 0] += 0.04 * np.exp(-200 * ((x - 1.0) ** 2 + (y - 0.35) ** 2))
        return cc


def _case(nx, ny, ngl, fused=True):
    from pynama_b200.cases.custom_func import CustomFuncCase
    from pynama_b200.common.options import Options
    Options.clear()
    for k, v in dict(ksp_type="cg", pc_type="jacobi", mat_fused=1 if fused else 0).items():
        Options.setValue(k, v)
    with open(f"{ROOT}/pynama_b200/cases/taylor-green.yaml") as f:
        cfg = yaml.load(f, Loader=yaml.Loader)
    fem = CustomFuncCase(cfg, case="taylor-green", ngl=ngl, nelem=[nx, ny], lower=[0.0, 0.0], upper=[2.0, 0.7])
    fem.setUpGeneral()
    fem.setUpBoundaryConditions()
    fem.setUpEmptyMats()
    return fem


@pytest.mark.parametrize("ngl", [2, 3, 4, 5, 7, 9])
def test_affine_element_matrices(be, ngl):
    from pynama_b200.elements.spectral import Spectral
    s = Spectral(ngl, 2)
    m = ShearedMesh([3, 2], ngl)
    rect = o.BoxMesh([3, 2], [0.0, 0.0], [2.0, 0.7], ngl)
    for mesh in (rect, m):
        cc = mesh.cell_corner_coords()
        coef, bad = s.getAffineCoefficientsDevice(be.to_device(cc))
        assert bad == 0
        Ka, Rwa, Rda = s.kernels().kle_affine(coef)
        Kg, Rwg, Rdg = s.getElemKLEMatricesDevice(be.to_device(cc))
        Ko, Rwo, Rdo = o.elem_kle_matrices(ngl, cc)
        for a, g, ref in ((Ka, Kg, Ko), (Rwa, Rwg, Rwo), (Rda, Rdg, Rdo)):
            assert rel_err(a.cpu().numpy(), ref) < 1e-12
            assert rel_err(a.cpu().numpy(), g.cpu().numpy()) < 1e-12
        K = Ka.cpu().numpy()
        assert np.array_equal(K, K.transpose(0, 2, 1))          # bitwise symmetric, like the general kernel
    _, bad = s.getAffineCoefficientsDevice(be.to_device(ShearedMesh([3, 2], ngl, "bent").cell_corner_coords()))
    assert bad > 0


@pytest.mark.parametrize("ngl,nx,ny", [(2, 9, 7), (3, 6, 5), (5, 6, 5), (8, 3, 4)])
def test_fused_assembly_vs_oracle_and_gather(be, ngl, nx, ny):
    fem = _case(nx, ny, ngl)
    m = ShearedMesh([nx, ny], ngl)
    corners = be.to_device(m.cell_corner_coords())
    fem.mat.assembleFromMesh(fem.dom, fem.elemType, perCell=True, corners=corners)
    assert fem.mat.fusedAssembly
    ref = o.assemble_freeslip(m, per_cell=True)
    for name in ("K", "Krhs", "Rw"):
        ip, ix, dv = getattr(fem.mat, name).getValuesCSR()
        np.testing.assert_array_equal(ip, ref[name].indptr)
        np.testing.assert_array_equal(ix, ref[name].indices)
        assert rel_err(dv, ref[name].data) < 1e-12, name
    first = {k: getattr(fem.mat, k).csr.data.clone() for k in ("K", "Krhs", "Rw")}
    # values-only re-assembly (the timed path of bench.py --workload percell): same bits
    fem.mat.assembleFromMesh(fem.dom, fem.elemType, perCell=True, corners=corners)
    for k, v in first.items():
        assert bool((getattr(fem.mat, k).csr.data == v).all()), k
    # gather path on element matrices materialised with the same closed form: bit-identical
    coef, bad = fem.elemType.getAffineCoefficientsDevice(corners)
    Ke, Rwe, _ = fem.elemType.kernels().kle_affine(coef, want_rd=False)
    other = _case(nx, ny, ngl)
    other.mat.assembleFromMesh(other.dom, other.elemType, perCell=True, elementMatrices=(Ke, Rwe))
    assert not other.mat.fusedAssembly
    for k, v in first.items():
        assert bool((getattr(other.mat, k).csr.data == v).all()), k


def test_non_affine_mesh_takes_the_general_kernels(be):
    nx, ny, ngl = 6, 5, 4
    fem = _case(nx, ny, ngl)
    m = ShearedMesh([nx, ny], ngl, "bent")
    fem.mat.assembleFromMesh(fem.dom, fem.elemType, perCell=True, corners=be.to_device(m.cell_corner_coords()))
    assert not fem.mat.fusedAssembly
    ref = o.assemble_freeslip(m, per_cell=True)
    for name in ("K", "Krhs", "Rw"):
        ip, ix, dv = getattr(fem.mat, name).getValuesCSR()
        np.testing.assert_array_equal(ix, ref[name].indices)
        assert rel_err(dv, ref[name].data) < 1e-12, name


def test_fused_rd_blocks(be):
    """Rd_e through the fused kernel (MatNS uses Rd; the FreeSlip path leaves it empty as the reference does)."""
    nx, ny, ngl = 5, 4, 4
    fem = _case(nx, ny, ngl)
    m = ShearedMesh([nx, ny], ngl)
    corners = be.to_device(m.cell_corner_coords())
    coef, bad = fem.elemType.getAffineCoefficientsDevice(corners)
    kern = fem.elemType.kernels()
    topo = fem.dom.topology()
    _, _, Rde = kern.kle_affine(coef)
    A = be.assemble(topo, 2, 1, be.AffineSource(kern, coef, "Rd"))
    B = be.assemble(topo, 2, 1, Rde)
    assert bool((A.data == B.data).all()) and bool((A.indices == B.indices).all())
    _, _, Rdo = o.elem_kle_matrices(ngl, m.cell_corner_coords())
    assert rel_err(Rde.cpu().numpy(), Rdo) < 1e-12


@pytest.mark.skipif(os.environ.get("PYN_AFF_NO_SIG") is not None, reason="already inside the flat-kernel run")
def test_flat_kernel_is_bit_identical_too():
    """On a structured mesh the fused assembly takes the signature-tile kernel (asm_sig_affine_kernel); the slot-flat
    kernel (asm_interior_affine_kernel: meshes without row signatures, valence > 4) is selected once per process with
    PYN_AFF_NO_SIG=1, so the parity cases above are re-run in a child process with it."""
    env = dict(os.environ, PYN_AFF_NO_SIG="1")
    r = subprocess.run([sys.executable, "-m", "pytest", __file__, "-q", "-x", "-k", "fused_assembly_vs_oracle or rd_blocks"],
                       env=env, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
