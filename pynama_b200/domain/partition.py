"""Element-block partition of the box mesh over the ranks of one node (the reference's counterpart is
`DMPlex.distribute()`, src/domain/dmplex.py:29, plus PETSc's row-block ownership of Mats and Vecs).

The mesh is cut into `world` strips of whole cell rows.  Rank r

  * integrates the cells of its strip plus ONE ghost row of cells above it, numbered as a box mesh of its own
    (`nx x ny_local`, the same closed-form DMPlex numbering as the serial mesh) -- so every mesh kernel of the library
    runs unchanged on the local mesh;
  * owns the nodes of its strip except the bottom interface line (owned by the rank below -- "lowest rank owns"), in
    particular it owns its top interface line, whose matrix rows are complete thanks to the ghost row of cells: no matrix
    entry is ever exchanged (this replaces PETSc's MatStash);
  * holds as ghosts the bottom interface line (owner r-1) and the remaining nodes of the ghost cell row (owner r+1).

Everything here is host-side index arithmetic on closed forms (NumPy, O(local nodes)); the per-exchange lists are
ordered by GLOBAL node id on both sides so sender and receiver agree without a handshake.
"""
import numpy as np


def strip_rows(ny, world, rank):
    """Cell rows [fy0, fy1) of `rank`: as even as possible, the first `ny % world` ranks take one more."""
    base, rem = divmod(ny, world)
    if base == 0:
        raise Exception(f"cannot cut {ny} cell rows into {world} strips")
    fy0 = rank * base + min(rank, rem)
    return fy0, fy0 + base + (1 if rank < rem else 0)


class StripPartition(object):
    def __init__(self, nx, ny, ngl, rank, world):
        self.nx, self.ny, self.ngl, self.rank, self.world = int(nx), int(ny), int(ngl), int(rank), int(world)
        self.fy0, self.fy1 = strip_rows(self.ny, self.world, self.rank)
        self.nrows = self.fy1 - self.fy0
        self.ghost_above = 1 if self.fy1 < self.ny else 0
        self.ghost_below_line = self.fy0 > 0
        self.ny_local = self.nrows + self.ghost_above
        self.cell0 = self.fy0 * self.nx                       # first global cell of the local mesh
        self.ncell_local = self.nx * self.ny_local
        m = self.ngl - 2
        nx, nyl = self.nx, self.ny_local
        self.m = m
        self.nvert = (nx + 1) * (nyl + 1)
        self.nxedge, self.nyedge = nx * (nyl + 1), (nx + 1) * nyl
        self.off_vert = self.ncell_local * m * m
        self.off_edge = self.off_vert + self.nvert
        self.off_yedge = self.off_edge + self.nxedge * m
        self.nnodes = self.off_yedge + self.nyedge * m
        self.nnodes_global = nx * self.ny * m * m + (nx + 1) * (self.ny + 1) + (nx * (self.ny + 1) + (nx + 1) * self.ny) * m
        self._rows = None

    # -- per-node row coordinate of the carrying mesh entity ---------------------------------------------------
    def _entity_rows(self):
        """For every local node: kind (0 cell interior, 1 vertex, 2 x-edge, 3 y-edge) and the local row index of its
        entity (cell row / vertex line / x-edge line / y-edge row)."""
        if self._rows is None:
            nx, nyl, m = self.nx, self.ny_local, self.m
            kind = np.empty(self.nnodes, dtype=np.int8)
            row = np.empty(self.nnodes, dtype=np.int64)
            kind[:self.off_vert] = 0
            row[:self.off_vert] = (np.arange(self.off_vert) // (m * m)) // nx if m else 0
            kind[self.off_vert:self.off_edge] = 1
            row[self.off_vert:self.off_edge] = np.arange(self.nvert) // (nx + 1)
            kind[self.off_edge:self.off_yedge] = 2
            row[self.off_edge:self.off_yedge] = (np.arange(self.nxedge * m) // m) // nx if m else 0
            kind[self.off_yedge:] = 3
            row[self.off_yedge:] = (np.arange(self.nyedge * m) // m) % nyl if m else 0
            self._rows = (kind, row)
        return self._rows

    def owned_mask(self):
        """bool[nnodes_local]: node owned by this rank."""
        kind, row = self._entity_rows()
        line = (kind == 1) | (kind == 2)                     # entities living on a horizontal mesh line
        own = np.where(line, row <= self.nrows, row < self.nrows)
        if self.ghost_below_line:
            own &= ~(line & (row == 0))
        return own

    def local_to_global(self):
        """int64[nnodes_local]: serial (global) node id of every local node."""
        nx, ny, nyl, m = self.nx, self.ny, self.ny_local, self.m
        g = np.empty(self.nnodes, dtype=np.int64)
        g_off_vert = nx * ny * m * m
        g_off_edge = g_off_vert + (nx + 1) * (ny + 1)
        g_nxedge = nx * (ny + 1)
        mm = m * m
        i = np.arange(self.off_vert)
        g[:self.off_vert] = i + self.cell0 * mm                                   # cell interiors shift by whole cells
        g[self.off_vert:self.off_edge] = g_off_vert + np.arange(self.nvert) + self.fy0 * (nx + 1)
        i = np.arange(self.nxedge * m)
        g[self.off_edge:self.off_yedge] = g_off_edge + i + self.fy0 * nx * m
        i = np.arange(self.nyedge * m)
        if m:
            e, k = i // m, i % m
            vx, eyl = e // nyl, e % nyl
            g[self.off_yedge:] = g_off_edge + (g_nxedge + vx * ny + eyl + self.fy0) * m + k
        return g

    # -- halo lists -----------------------------------------------------------------------------------------------
    def halo_lists(self):
        """[(neighbour rank, send local ids, recv local ids)], each list ordered by global node id."""
        kind, row = self._entity_rows()
        l2g = self.local_to_global()
        line = (kind == 1) | (kind == 2)

        def by_global(mask):
            idx = np.flatnonzero(mask)
            return idx[np.argsort(l2g[idx], kind="stable")].astype(np.int64)

        out = []
        if self.ghost_below_line:                               # neighbour below: r-1
            recv = by_global(line & (row == 0))                 # my bottom line <- its top interface line
            send = by_global((line & (row == 1)) | (~line & (row == 0)))   # its ghost cell row = my first cell row
            out.append((self.rank - 1, send, recv))
        if self.ghost_above:                                    # neighbour above: r+1
            send = by_global(line & (row == self.nrows))        # my top interface line -> its bottom line
            recv = by_global((line & (row == self.nrows + 1)) | (~line & (row == self.nrows)))
            out.append((self.rank + 1, send, recv))
        return out

    def describe(self):
        return (f"rank {self.rank}/{self.world}: cell rows [{self.fy0},{self.fy1}) + {self.ghost_above} ghost row, "
                f"{self.nnodes} local nodes ({int(self.owned_mask().sum())} owned)")
