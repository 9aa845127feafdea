"""CPU tests of the multi-GPU host logic: the strip partition (ownership, local->global numbering, halo lists) and,
with two gloo ranks, a NumPy emulation of the distributed CG built on exactly those lists -- halo exchange of the search
direction before every product, all-reduced scalars -- against the serial oracle solve."""
import os
import sys

import numpy as np
import pytest

import kle_oracle as o
from conftest import ROOT
from pynama_b200.domain.dmplex import DMPlexDom
from pynama_b200.domain.partition import StripPartition, strip_rows


@pytest.mark.parametrize("nx,ny,n,world", [(4, 6, 3, 2), (3, 7, 4, 3), (5, 4, 2, 4), (2, 8, 5, 8), (6, 5, 6, 2)])
def test_strip_partition_invariants(nx, ny, n, world):
    gdom = DMPlexDom(boxMesh={"nelem": [nx, ny], "lower": [0, 0], "upper": [1, 1]})
    gdom.setFemIndexing(n)
    parts = [StripPartition(nx, ny, n, r, world) for r in range(world)]
    assert parts[0].fy0 == 0 and parts[-1].fy1 == ny and all(a.fy1 == b.fy0 for a, b in zip(parts, parts[1:]))
    owners = np.zeros(gdom.getTotalNodes(), dtype=int)
    for p in parts:
        assert p.nnodes_global == gdom.getTotalNodes()
        ldom = DMPlexDom(boxMesh={"nelem": [nx, p.ny_local], "lower": [0, 0], "upper": [1, 1]})
        ldom.setFemIndexing(n)
        assert ldom.getTotalNodes() == p.nnodes
        l2g = p.local_to_global()
        # the same cells seen locally and globally carry the same nodes in the same element-local order
        lc = ldom.cellNodesArray()
        gc = gdom.cellNodesArray(np.arange(p.ncell_local) + p.cell0)
        np.testing.assert_array_equal(l2g[lc], gc)
        assert len(np.unique(l2g)) == p.nnodes
        own = p.owned_mask()
        owners[l2g[own]] += 1
        # every owned node has all its incident cells in the local mesh (complete rows without matrix exchange)
        cells_of = {}
        for c, nodes in enumerate(gdom.cellNodesArray()):
            for v in nodes:
                cells_of.setdefault(int(v), []).append(c)
        for g in l2g[own][:: max(1, own.sum() // 50)]:
            assert all(p.cell0 <= c < p.cell0 + p.ncell_local for c in cells_of[int(g)])
    np.testing.assert_array_equal(owners, 1)        # every node owned exactly once
    # halo lists pair up: what r sends to q is what q receives from r, in the same (global id) order, and the
    # receive lists cover exactly the non-owned nodes
    lists = [{nbr: (s, r) for nbr, s, r in p.halo_lists()} for p in parts]
    for r, p in enumerate(parts):
        l2g = p.local_to_global()
        ghosts = []
        for q, (send, recv) in lists[r].items():
            qs, qr = lists[q][r]
            np.testing.assert_array_equal(l2g[send], parts[q].local_to_global()[qr])
            assert p.owned_mask()[send].all() and not p.owned_mask()[recv].any()
            ghosts.append(recv)
        got = np.sort(np.concatenate(ghosts)) if ghosts else np.zeros(0, int)
        np.testing.assert_array_equal(got, np.flatnonzero(~p.owned_mask()))
    with pytest.raises(Exception):
        strip_rows(3, 4, 0)


def _dist_cg_worker(rank, world, port, nx, ny, n, out):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    import torch
    m = o.BoxMesh([nx, ny], [0, 0], [1, 1], n)
    ref = o.assemble_freeslip(m, per_cell=False)
    K = ref["K"].tocsr()
    vel_ex, vort = o.taylor_green(m.node_coords(), 0.02, 0.0)
    vb = np.zeros(2 * m.nnodes)
    vb.reshape(-1, 2)[ref["bc_nodes"]] = vel_ex[ref["bc_nodes"]]
    b = ref["Rw"] @ vort + ref["Krhs"] @ vb
    p = StripPartition(nx, ny, n, rank, world)
    l2g, own = p.local_to_global(), p.owned_mask()
    dofs = (l2g[:, None] * 2 + np.arange(2)).reshape(-1)
    own2 = np.repeat(own, 2)
    Kl = K[dofs][:, dofs].tolil()
    Kl[np.flatnonzero(~own2)] = 0                      # rows of ghost nodes belong to another rank
    Kl = Kl.tocsr()
    bl = np.where(own2, b[dofs], 0.0)
    halo = p.halo_lists()

    def exchange(v):                                   # ghost entries <- owners (2 dofs per node)
        v2 = v.reshape(-1, 2)
        reqs, bufs = [], []
        for nbr, send, recv in halo:
            sb = torch.from_numpy(np.ascontiguousarray(v2[send]))
            rb = torch.empty((len(recv), 2), dtype=torch.float64)
            reqs += [dist.isend(sb, nbr), dist.irecv(rb, nbr)]
            bufs.append((recv, rb, sb))
        for q in reqs:
            q.wait()
        for recv, rb, _ in bufs:
            v2[recv] = rb.numpy()

    def gsum(*vals):
        t = torch.tensor(vals, dtype=torch.float64)
        dist.all_reduce(t)
        return t.tolist()

    d = Kl.diagonal()
    dinv = np.where(d != 0, 1.0 / np.where(d != 0, d, 1.0), 1.0)
    x, r = np.zeros_like(bl), bl.copy()
    z = dinv * r
    zz, beta = gsum(z @ z, z @ r)
    ttol, pvec, betaold, its = 1e-11 * np.sqrt(zz), np.zeros_like(bl), 1.0, 0
    for its in range(1, 5000):
        pvec = z.copy() if its == 1 else z + (beta / betaold) * pvec
        exchange(pvec)
        w = Kl @ pvec
        (dpi,) = gsum(pvec @ w)
        a = beta / dpi
        x += a * pvec
        r -= a * w
        z = dinv * r
        betaold = beta
        zz, beta = gsum(z @ z, z @ r)
        if np.sqrt(zz) <= ttol:
            break
    xs = o.direct_solve(K, b)
    err = np.abs(x[own2] - xs[dofs][own2]).max() / np.abs(xs).max()
    gerr = np.abs(x[~own2] - xs[dofs][~own2]).max() / np.abs(xs).max() if (~own2).any() else 0.0
    out.put((rank, its, err, gerr))
    dist.destroy_process_group()


def test_distributed_cg_emulation_gloo_world2():
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    out = ctx.Queue()
    port = 29500 + os.getpid() % 2000
    procs = [ctx.Process(target=_dist_cg_worker, args=(r, 2, port, 4, 6, 4, out)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(out.get(timeout=240) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    assert res[0][1] == res[1][1]                       # same iteration count on both ranks
    for rank, its, err, gerr in res:
        assert err < 1e-9 and gerr < 1e-9, res          # owned entries AND ghost entries of x hold the solution


@pytest.mark.parametrize("nx,ny,n,world", [(4, 4, 3, 4), (3, 5, 4, 4), (4, 6, 3, 2), (2, 8, 5, 8), (3, 9, 2, 8)])
def test_partition_dof_classes_match_serial(nx, ny, n, world):
    """Boundary classes of every rank's local mesh (ghost row included) equal the serial classes of the same nodes --
    in particular when the ghost row of the last-but-one rank is the top cell row of the global mesh (1-row strips)."""
    from pynama_b200.matrices.mat_ns import noSlipDofClasses
    box = {"nelem": [nx, ny], "lower": [0, 0], "upper": [1, 1]}
    gdom = DMPlexDom(boxMesh=box)
    gdom.setFemIndexing(n)
    gdom.setBoundaryCondition([], ["up", "down", "left", "right"])
    serial = noSlipDofClasses(gdom, gdom.getGlobalIndicesNoSlip(), gdom.getGlobalIndicesDirichlet()).reshape(-1, 2)
    for r in range(world):
        p = StripPartition(nx, ny, n, r, world)
        ldom = DMPlexDom(boxMesh=box, partition=p)
        ldom.setFemIndexing(n)
        ldom.setBoundaryCondition([], ["up", "down", "left", "right"])
        local = noSlipDofClasses(ldom, ldom.getGlobalIndicesNoSlip(), ldom.getGlobalIndicesDirichlet()).reshape(-1, 2)
        np.testing.assert_array_equal(local, serial[p.local_to_global()], err_msg=f"rank {r}")
        # the per-cell wall query agrees with the global one for the top and bottom cell rows of the local mesh
        for c in (0, ldom.ncell - 1):
            _, normals = ldom.getBorderNodesWithNormal(c, ldom.getBordersNodes())
            _, gnormals = gdom.getBorderNodesWithNormal(c + p.cell0, gdom.getBordersNodes())
            assert normals == gnormals
