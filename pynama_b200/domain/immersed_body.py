"""Immersed bodies: Lagrangian marker points of the immersed-boundary cases.

Reference: src/domain/immersed_body.py -- `BodiesContainer` (:8-132), `ImmersedBody` (:134-291), `Line` (:293-313),
`OpenBox` (:315-364), `Circle` (:366-396) and the discrete delta functions (:398-421).  The reference stores each body as
a 1-D DMPlex; only the point coordinates and their count are ever read back, so here a body is its coordinate array.
The marker velocities live in a device vector (what `H v - V_body` of the correction step consumes); the delta function
itself is evaluated on the device (csrc/ibm.cu) -- the host versions below serve `getDiracs` callers and the tests.
"""
import logging
from math import ceil, cos, pi, sin, sqrt

import numpy as np

from ..petsc_like import Vec


def threeGrid(r):
    """supports only three cell grids (immersed_body.py:398-405)"""
    if r <= 0.5:
        return (1 + sqrt(-3 * r ** 2 + 1)) / 3
    elif r <= 1.5:
        return (5 - 3 * r - sqrt(-3 * (1 - r) ** 2 + 1)) / 6
    return 0


def linear(r):
    """Lineal Dirac discretization (immersed_body.py:407-412)"""
    return (1 - r) if r < 1 else 0


def fourGrid(r):
    """(immersed_body.py:415-421)"""
    if r <= 1:
        return (3 - 2 * r + sqrt(1 + 4 * r - 4 * r ** 2)) / 8
    elif r <= 2:
        return (5 - 2 * r - sqrt(-7 + 12 * r - 4 * r ** 2)) / 8
    return 0


DELTA_NAMES = {fourGrid: "four", threeGrid: "three", linear: "linear"}


class ImmersedBody(object):
    def __init__(self, vel=[0, 0], center=[0, 0]):
        self.dirac = fourGrid
        self._centerDisplacement = list(center)
        self._startCenter = list(center)
        self._dl = None
        self._vel = list(vel)
        self._Uref = None
        self.logger = logging.getLogger("Body Immersed")
        self._history = {"times": [], "displ": [], "vel": []}
        self._isStatic = True
        self._coords = None
        self._velVec = None

    def setVelRef(self, vel):
        self._Uref = vel

    def getVelRef(self):
        return self._Uref

    def setIsMoving(self):
        self._isStatic = False

    def isMoving(self):
        return not self._isStatic

    def view(self):
        self.logger.info(f"Arc len: {self._dl} | Dirac Type: {self.dirac.__name__} | Vel Fluid Reference: {self._Uref} ")

    def viewState(self):
        self.logger.info(f"{self.__class__.__name__} vel: {self._vel} | Body center position {self._centerDisplacement}")

    def generateDMPlex(self, dh, dim=1):
        coords, cone, dl = self.generateBody(dh)
        self._dl = dl
        self._L = self.getLong()
        self._coords = np.asarray(coords, dtype=np.float64).reshape(-1, 2)
        self._cone = cone
        self._velVec = Vec(self.getTotalNodes() * len(self._vel), name="body-velocity")
        self.updateVelocity()

    def setCenter(self, val):
        self._centerDisplacement = list(val)

    def setElementLong(self, dl, normals):
        self._normals = normals
        self._dl = dl

    def computeForce(self, q):
        q = np.asarray(q)
        return float(q[0::2].sum()), float(q[1::2].sum())

    def getVelocity(self):
        return self._velVec

    def setVelocity(self, inds, values):
        self._velVec.setValues(inds, values)
        self._velVec.assemble()

    def getElementLong(self):
        return self._dl

    def getTotalNodes(self):
        return len(self._coords)

    def getLong(self):
        return None

    def setCaracteristicLong(self, val):
        self._L = val

    def getCaracteristicLong(self):
        return self._L

    def getCenterBody(self):
        return self._centerDisplacement

    def getNodeCoordinates(self, node):
        return self._coords[node] + np.asarray(self._centerDisplacement, dtype=np.float64)

    def coordinatesArray(self):
        """(npoints, 2): every marker's position, getNodeCoordinates for all nodes at once."""
        return self._coords + np.asarray(self._centerDisplacement, dtype=np.float64)

    def getRegion(self):
        return None

    def getDiracs(self, dist, h):
        dirac = 1
        for r in dist:
            dirac *= self.dirac(abs(r) / h)
            dirac /= h
        return dirac

    def updateBodyParameters(self, t):
        """Prescribed transverse oscillation of a moving body (immersed_body.py:266-282): A = 0.3, f = 5."""
        if self._isStatic:
            return
        velX = 0
        displX = 0 + self._startCenter[0]
        f = 5
        Te = f / self._Uref
        A = 0.3
        displY = A * sin(2 * pi * t / Te) + self._startCenter[1]
        velY = 2 * pi * A * cos(2 * pi * t / Te) / Te
        self._vel = [velX, velY]
        self._centerDisplacement = [displX, displY]
        self.updateVelocity()
        self._history["times"].append(t)
        self._history["displ"].append(self._centerDisplacement)
        self._history["vel"].append(self._vel)

    def updateVelocity(self):
        points = self.getTotalNodes()
        self._velVec.setArray(np.tile(np.asarray(self._vel, dtype=np.float64), points))

    def generateBody(self, *args):
        return None, None, None


class Line(ImmersedBody):
    def generateBody(self, dl, longitud=2):
        self._region = longitud
        div = ceil(longitud / dl)
        coords_x = np.linspace(0, longitud, div)
        coords = [[x, 0.0] for x in coords_x]
        cone = [[i, i + 1] for i in range(div - 1)]
        return coords, cone, dl

    def getLong(self):
        return 1

    def getRegion(self):
        return self._region


class OpenBox(ImmersedBody):
    def generateBody(self, dl, longitud=1):
        div = ceil(sqrt(2) / dl)
        pts = [(0, longitud), (-longitud, 0), (0, -longitud), (longitud, 0), (0, longitud)]
        xs = np.concatenate([np.linspace(pts[i][0], pts[i + 1][0], div, endpoint=False) for i in range(4)])
        ys = np.concatenate([np.linspace(pts[i][1], pts[i + 1][1], div, endpoint=False) for i in range(4)])
        coords = [[x, y] for x, y in zip(xs, ys)]
        cone = [[i, i + 1] for i in range(len(coords))]
        cone[-1][-1] = 0
        return coords, cone, dl

    def updateVelocity(self):
        """Lid-driven cavity rotated by 45 degrees: the markers of the first quadrant move (immersed_body.py:347-358)."""
        velRef = self.getVelRef()
        xy = self.coordinatesArray()
        vals = np.zeros(2 * len(xy))
        if velRef is not None:
            lid = (xy[:, 0] >= 0) & (xy[:, 1] >= 0)
            vals[0::2][lid] = velRef / sqrt(2)
            vals[1::2][lid] = -velRef / sqrt(2)
        self._velVec.setArray(vals)

    def getLong(self):
        return 1

    def getRegion(self):
        return 1


class Circle(ImmersedBody):
    def __init__(self, vel, center, radius):
        super().__init__(vel, center)
        self._radius = radius

    def generateBody(self, dh):
        r = self._radius
        longTotal = 2 * pi * r
        points = ceil(longTotal / dh)
        startAng = pi / 1000
        angles = np.linspace(0, 2 * pi, points, endpoint=False)
        x = r * np.cos(angles + startAng)
        y = r * np.sin(angles + startAng)
        coords = [[x[i], y[i]] for i in range(len(x))]
        cone = [[i, i + 1] for i in range(len(x))]
        cone[-1][-1] = 0
        dl = 2 * pi * r / len(coords)
        return coords, cone, dl

    def getLong(self):
        return self._radius * 2

    def getRegion(self):
        return self._radius + 2 * self.getElementLong()


class BodiesContainer(object):
    types = ['circle', 'line', 'box']

    def __init__(self, bodies):
        self.logger = logging.getLogger("Bodies Container")
        self.bodies = list()
        for cfgBody in bodies:
            center = cfgBody['center']
            if cfgBody['type'] == 'circle':
                body = Circle(vel=[0, 0], center=center, radius=cfgBody['radius'])
            elif cfgBody['type'] == 'line':
                body = Line(vel=[0, 0], center=center)
            elif cfgBody['type'] == 'box':
                body = OpenBox(vel=[0, 0], center=center)
            else:
                raise Exception("not defined")
            if cfgBody['vel'] == 'dynamic':
                body.setIsMoving()
            self.bodies.append(body)

    def createBodies(self, h):
        for i, body in enumerate(self.bodies):
            body.generateDMPlex(h)
            self.logger.info(f"Body number: {i}")
            body.viewState()
        self._offsets = np.concatenate([[0], np.cumsum([b.getTotalNodes() for b in self.bodies])]).astype(int)
        self._velAll = None

    def getTotalNodes(self):
        return int(sum(body.getTotalNodes() for body in self.bodies))

    def getRegion(self):
        radius = 0.5
        distanceCenters = 2
        hCeil = self.getElementLong()
        return radius + distanceCenters + hCeil * 4

    def mapGlobToLocal(self, globNode):
        numBody = int(np.searchsorted(self._offsets, globNode, side="right") - 1)
        return globNode - int(self._offsets[numBody]), numBody

    def getNodeCoordinates(self, globNode):
        locNode, numBody = self.mapGlobToLocal(globNode)
        return self.bodies[numBody].getNodeCoordinates(locNode)

    def coordinatesArray(self):
        return np.concatenate([b.coordinatesArray() for b in self.bodies], axis=0)

    def getDiracs(self, dist, h):
        return self.bodies[0].getDiracs(dist, h)

    def deltaName(self):
        return DELTA_NAMES[self.bodies[0].dirac]

    def getElementLong(self):
        return self.bodies[0].getElementLong()

    def getCenters(self):
        return [b.getCenterBody() for b in self.bodies]

    def computeForce(self, vec, dt):
        forces_x, forces_y = list(), list()
        q = vec.getArray()
        for b, body in enumerate(self.bodies):
            lo, hi = int(self._offsets[b]), int(self._offsets[b + 1])
            fx, fy = body.computeForce(q[lo * 2:hi * 2])
            forces_x.append(float(fx / dt))
            forces_y.append(float(fy / dt))
        return forces_x, forces_y

    def setVelRef(self, vel):
        for body in self.bodies:
            body.setVelRef(vel)

    def getVelRef(self, bodyNum=0):
        return self.bodies[bodyNum].getVelRef()

    def updateBodyParameters(self, t):
        for body in self.bodies:
            body.updateBodyParameters(t)

    def anyMoving(self):
        return any(b.isMoving() for b in self.bodies)

    def viewBodies(self):
        for i, body in enumerate(self.bodies):
            self.logger.info(f"Body num: {i}")
            body.viewState()

    def getVelocity(self):
        """All bodies' marker velocities as ONE vector (the reference nests the per-body vectors, :117-125)."""
        if len(self.bodies) == 1:
            return self.bodies[0].getVelocity()
        import torch
        if self._velAll is None:
            self._velAll = Vec(2 * self.getTotalNodes(), name="bodies-velocity")
        torch.cat([b.getVelocity().data for b in self.bodies], out=self._velAll.data)
        return self._velAll

    def updateVelocity(self):
        for body in self.bodies:
            body.updateVelocity()

    def bodyNumbers(self):
        return len(self.bodies)
