"""Quadrature rules and Gauss-point lists (host tables of the element layer).

Same public names and results as the reference's src/elements/utilities.py:
gaussPoints (:43-61), lobattoPoints (:63-92), generateGaussPoints2D (:15-27), generateGaussPoints3D (:29-41).
"""
from collections import namedtuple
from math import pi

import numpy as np

GaussPoint2D = namedtuple("GaussPoint2D", ["r", "s", "w"])
GaussPoint3D = namedtuple("GaussPoint3D", ["r", "s", "t", "w"])


def generateGaussPoints2D(gps1D, gpsWei):
    """Tensor list of 2-D points, first coordinate slowest (utilities.py:15-27)."""
    return [GaussPoint2D(r=gps1D[i], s=gps1D[j], w=gpsWei[i] * gpsWei[j])
            for i in range(len(gps1D)) for j in range(len(gps1D))]


def generateGaussPoints3D(gps1D, gpsWei):
    n = len(gps1D)
    return [GaussPoint3D(r=gps1D[i], s=gps1D[j], t=gps1D[k], w=gpsWei[i] * gpsWei[j] * gpsWei[k])
            for i in range(n) for j in range(n) for k in range(n)]


def gaussPoints(N):
    """Gauss-Legendre rule from the eigen-decomposition of the Jacobi matrix (Golub-Welsch), symmetrised so that
    x_i = -x_{N-1-i} exactly (utilities.py:43-61)."""
    k = np.arange(1, N)
    off = 0.5 / np.sqrt(1.0 - (2.0 * k) ** -2.0)
    lam, vec = np.linalg.eig(np.diag(off, 1) + np.diag(off, -1))
    order = np.argsort(lam)
    x, w = lam[order], 2.0 * vec[0, order] ** 2
    return (x - x[::-1]) / 2.0, (w + w[::-1]) / 2.0


def lobattoPoints(N):
    """Gauss-Lobatto-Legendre rule: Newton on the Legendre recursion from the Chebyshev-GL points until the update
    is below 1e-15, then symmetrised and ascending (utilities.py:63-92)."""
    x = np.cos(np.linspace(0.0, pi, N))
    P = np.zeros((N, N))
    prev = 2.0
    while np.max(np.abs(x - prev)) > 1e-15:
        prev = x
        P[:, 0], P[:, 1] = 1.0, x
        for k in range(2, N):
            P[:, k] = ((2 * k - 1) * x * P[:, k - 1] - (k - 1) * P[:, k - 2]) / k
        x = prev - (x * P[:, N - 1] - P[:, N - 2]) / (N * P[:, N - 1])
    w = 2.0 / ((N - 1) * N * P[:, N - 1] ** 2)
    return (x[::-1] - x) / 2.0, (w[::-1] + w) / 2.0
