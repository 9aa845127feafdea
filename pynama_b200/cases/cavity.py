"""Lid-driven cavity (reference: src/cases/cavity.py): four no-slip walls, one or more of them moving.

The wall values are turned once into device index/value arrays; applying them each solve is one scatter kernel per
vector instead of the reference's per-call Python walk over border nodes (cavity.py:53-81)."""
import numpy as np

from .. import backend as be
from ..common.nswalls import NoSlipWalls
from .base_problem import NoSlipFreeSlip


class Cavity(NoSlipFreeSlip):
    def setUp(self):
        super().setUp()
        self.collectCornerNodes()
        self._bcDevice = None

    def collectCornerNodes(self):
        """Nodes shared by two walls (cavity.py:10-24)."""
        names = list(self.nsWalls.getWallsNames())
        border = {w: self.dom.borderNodesArray(w) for w in names}
        corner = set()
        for i, a in enumerate(names):
            for b in names[i + 1:]:
                corner |= set(np.intersect1d(border[a], border[b]).tolist())
        self.cornerDofs = [self.dim * node + dof for node in corner for dof in range(self.dim)]
        if not self.comm.rank:
            self.logger.info("Corner Nodes collected")

    def readBoundaryCondition(self, inputData):
        self.BoundaryCondition = list()
        if 'free-slip' in inputData:
            self.nsWalls = NoSlipWalls(self.lower, self.upper, exclude=inputData['free-slip'].keys())
        else:
            self.nsWalls = NoSlipWalls(self.lower, self.upper)
        if 'no-slip' in inputData:
            for wallName, wallVelocity in inputData['no-slip'].items():
                self.nsWalls.setWallVelocity(wallName, wallVelocity)

    def setUpBoundaryConditions(self):
        self.dom.setLabelToBorders()
        bc = self.config.get("boundary-conditions")
        fsFaces = bc['free-slip'].keys() if 'free-slip' in bc else list()
        nsFaces = self.nsWalls.getWallsNames()
        self.dom.setBoundaryCondition(fsFaces, list(nsFaces))
        if not self.comm.rank:
            self.logger.info("Boundary Conditions setted up")

    def computeInitialCondition(self, startTime):
        self.vort.set(0.0)

    # -- wall values ------------------------------------------------------------------------------------------
    def _wallValues(self):
        """(dofs, values) written by applyBoundaryConditions and applyBoundaryConditionsFS, in the reference's order
        of assignment (later walls overwrite earlier ones at shared corner nodes)."""
        def gather(walls, getter):
            dofs, vals = [], []
            for wallName in walls:
                nodes = self.dom.borderNodesArray(wallName)
                vel, velDofs = getter(wallName)
                vel = np.atleast_1d(np.asarray(vel, dtype=float))
                d = (nodes[:, None] * self.dim + np.asarray(velDofs, dtype=np.int64)[None, :]).reshape(-1)
                # cavity.py:61,73 pair the dofs with np.repeat(vel, len(nodes)): value k goes to entry k of the list
                v = np.repeat(vel, len(nodes))[:d.size] if len(vel) else np.zeros(0)
                dofs.append(d)
                vals.append(v)
            if not dofs:
                return np.zeros(0, np.int64), np.zeros(0)
            return np.concatenate(dofs), np.concatenate(vals)

        moving = self.nsWalls.getWallsWithVelocity()
        d1, v1 = gather(moving, self.nsWalls.getWallVelocity)
        static = self.nsWalls.getStaticWalls()
        d2, v2 = gather(static, lambda w: ([0.0] * len(self.nsWalls.getStaticDofsByName(w)),
                                           self.nsWalls.getStaticDofsByName(w)))
        dfs, vfs = np.concatenate([d1, d2]), np.concatenate([v1, v2])

        def last_wins(d, v):               # sequential setValues(INSERT): the last write to a dof survives
            if d.size == 0:
                return d, v
            _, first_rev = np.unique(d[::-1], return_index=True)
            keep = np.sort(d.size - 1 - first_rev)
            return d[keep], v[keep]

        return last_wins(d1, v1), last_wins(dfs, vfs)

    def _device_bc(self):
        if getattr(self, "_bcDevice", None) is None:
            (d1, v1), (d2, v2) = self._wallValues()
            self._bcDevice = tuple(be.to_device(a) for a in (d1.astype(np.int64), v1, d2.astype(np.int64), v2))
        return self._bcDevice

    def applyBoundaryConditions(self):
        """vel = 0 except the moving walls' tangential velocity (cavity.py:53-65)."""
        d1, v1, _, _ = self._device_bc()
        self.vel.set(0.0)
        if d1.numel():
            be.vec_scatter_set(self.vel.data, d1, v1)

    def applyBoundaryConditionsFS(self):
        """Overwrite the wall-tangential dofs of the free-slip solution with the wall values (cavity.py:67-81)."""
        _, _, d2, v2 = self._device_bc()
        if d2.numel():
            be.vec_scatter_set(self.velFS.data, d2, v2)
