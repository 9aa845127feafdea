"""Element base class: dimensions + the 1-D Lagrange tabulation (reference: src/elements/element.py:5-48)."""
import logging

import numpy as np


class Element(object):
    def __init__(self, dim, comm=None):
        self.comm = comm
        self.dim = dim
        self.dim_w = 1 if dim == 2 else 3
        self.dim_s = 3 if dim == 2 else 6
        rank = getattr(comm, "rank", 0) if comm is not None else 0
        self.logger = logging.getLogger("[{}] Class".format(rank))

    def interpFun1D(self, Nodes, evalPoi):
        """(h, dh) with h[q,a] = l_a(evalPoi[q]) and dh[q,a] = l_a'(evalPoi[q]) for the Lagrange basis on `Nodes`
        (element.py:17-48): products of (x - x_m) over the other nodes divided by products of (x_a - x_m)."""
        nodes = np.asarray(Nodes, dtype=float)
        pts = np.asarray(evalPoi, dtype=float)
        n = len(nodes)
        diff = pts[:, None, None] - nodes[None, None, :]              # [q, a, m] = x_q - x_m
        diff = np.broadcast_to(diff, (len(pts), n, n)).copy()
        eye = np.eye(n, dtype=bool)
        diff[:, eye] = 1.0                                            # skip m == a
        den = nodes[:, None] - nodes[None, :]
        den[eye] = 1.0
        den = np.prod(den, axis=1)
        h = np.prod(diff, axis=2) / den
        dh = np.zeros_like(h)
        for c in range(n):                                            # leave factor c out as well
            term = diff.copy()
            term[:, :, c] = np.where(eye[:, c], 0.0, 1.0)[None, :]    # a == c contributes nothing
            dh += np.prod(term, axis=2)
        return h, dh / den
