"""No-slip wall bookkeeping (reference: src/common/nswalls.py:6-48,74-83,201-268).

Pure host-side logic: which sides of the box are no-slip walls, their normal direction (0: x, 1: y, 2: z), which
velocity components are held at zero ("static dofs") and which carry a prescribed wall velocity.  The reference
derives a wall's normal from a linked list of vertices and a cross product; for an axis-aligned box the normal is
a property of the side's name, which is what is used here -- same names, same numbers, same return values.
"""
import numpy as np

_SIDES_2D = ("left", "right", "up", "down")
_SIDES_3D = _SIDES_2D + ("back", "front")
_NORMAL = {"left": 0, "right": 0, "up": 1, "down": 1, "back": 2, "front": 2}


class Wall(object):
    def __init__(self, num, name, dim, endpoints):
        self.num, self.name, self.dim = num, name, dim
        self.endpoints = [np.asarray(p, dtype=float) for p in endpoints]
        self.totalVecs = len(self.endpoints) - 1
        self.normal = _NORMAL[name]
        self.staticDofs = [d for d in range(dim) if d != self.normal]
        self.velocity, self.velDofs = None, None

    def setWallName(self, name):
        self.name = name

    def getWallName(self):
        return self.name

    def getWallNum(self):
        return self.num

    def computeNormal(self):
        return self.normal

    def setWallVelocity(self, vel):
        """Tangential components with a non-zero value become prescribed-velocity dofs and leave the static list
        (nswalls.py:201-215); a velocity with no tangential part is rejected."""
        moving = [d for d in self.staticDofs if vel[d] != 0]
        if not moving:
            raise Exception("Velocity not valid")
        self.velocity = np.array([vel[d] for d in moving])
        self.velDofs = moving
        self.staticDofs[:] = [d for d in self.staticDofs if d not in moving]   # in place: callers hold the list

    def getWallVelocity(self):
        if self.velocity is not None:
            return self.velocity, self.velDofs
        return [0] * len(self.staticDofs), self.staticDofs

    def getStaticDofs(self):
        return self.staticDofs

    def __repr__(self):
        return f"Wall({self.name}, normal={self.normal}, static={self.staticDofs}, vel={self.velocity})"


class NoSlipWalls(object):
    def __init__(self, lower, upper, exclude=()):
        self.dim = len(lower)
        self.lower, self.upper = lower, upper
        self.walls = {}
        lo, up = [float(v) for v in lower], [float(v) for v in upper]
        for num, side in enumerate(_SIDES_2D if self.dim == 2 else _SIDES_3D):
            if side in exclude:
                continue
            self.walls[side] = Wall(num, side, self.dim, self._endpoints(side, lo, up))
        self.staticWalls = list(self.walls.keys())
        self.wallsWithVelocity = []
        self.normals = {name: w.computeNormal() for name, w in self.walls.items()}

    def _endpoints(self, side, lo, up):
        if side in ("left", "right"):
            x = lo[0] if side == "left" else up[0]
            pts = [[x, lo[1]], [x, up[1]]]
        elif side in ("up", "down"):
            y = up[1] if side == "up" else lo[1]
            pts = [[lo[0], y], [up[0], y]]
        else:
            z = up[2] if side == "front" else lo[2]
            return [[lo[0], z, 0.0], [up[0], z, 0.0]]
        if self.dim == 3:
            pts = [p + [0.0] for p in pts]
        return pts

    def __iter__(self):
        return iter(self.walls.values())

    def __len__(self):
        return len(self.walls)

    def __repr__(self):
        return f"Walls defined: {len(self)} "

    def getWallsNames(self):
        return self.walls.keys()

    def getWallsWithVelocity(self):
        return self.wallsWithVelocity

    def getWallBySideName(self, name):
        return self.walls[name]

    def getStaticWalls(self):
        return self.staticWalls

    def setWallVelocity(self, name, vel):
        """Silently ignores unknown sides / invalid velocities, like the reference (bare except, nswalls.py:74-83)."""
        if name not in self.walls or len(vel) != self.dim:
            return None
        try:
            self.walls[name].setWallVelocity(vel)
        except Exception:
            return None
        self.wallsWithVelocity.append(name)
        self.staticWalls.remove(name)

    def getWallVelocity(self, name):
        if name not in self.walls:
            print("side not defined")
            return None
        return self.walls[name].getWallVelocity()

    def getStaticDofsByName(self, name):
        if name not in self.walls:
            print("side not defined")
            return None
        return self.walls[name].getStaticDofs()

    def computeWallsNormals(self):
        self.normals = {name: w.computeNormal() for name, w in self.walls.items()}

    def getWalletNormalBySideName(self, name):
        if name not in self.normals:
            print("side not defined")
            return None
        return self.normals[name]
