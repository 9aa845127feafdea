"""A skewed unstructured quadrilateral mesh of the unit square written as a gmsh v2 ASCII file (test helper)."""
import numpy as np


def skewed_square(n=8, amp=0.25, seed=0):
    """(points, quads): n x n cells, interior vertices displaced (non-affine cells), vertex ids shuffled, a third of the
    cells stored clockwise."""
    rng = np.random.default_rng(seed)
    g = np.linspace(0.0, 1.0, n + 1)
    X, Y = np.meshgrid(g, g, indexing="xy")
    P = np.stack([X.reshape(-1), Y.reshape(-1)], 1)
    h = 1.0 / n
    interior = (P[:, 0] > 0) & (P[:, 0] < 1) & (P[:, 1] > 0) & (P[:, 1] < 1)
    P[interior] += amp * h * (rng.random((int(interior.sum()), 2)) - 0.5) * 2
    vid = lambda i, j: j * (n + 1) + i
    Q = np.array([[vid(i, j), vid(i + 1, j), vid(i + 1, j + 1), vid(i, j + 1)] for j in range(n) for i in range(n)])
    perm = rng.permutation(len(P))                 # new id of old vertex v = perm[v]
    P2 = np.empty_like(P)
    P2[perm] = P
    Q2 = perm[Q]
    cw = rng.random(len(Q2)) < 0.33
    Q2[cw] = Q2[cw][:, ::-1]
    return P2, Q2


def write_msh(path, points, quads):
    """gmsh 2.2 ASCII with 1-based, non-contiguous node ids, a physical point, boundary lines and the quads."""
    nid = np.arange(len(points)) * 2 + 5            # sparse ids
    with open(path, "w") as f:
        f.write("$MeshFormat\n2.2 0 8\n$EndMeshFormat\n$Nodes\n%d\n" % (len(points) + 1))
        for k, (x, y) in enumerate(points):
            f.write("%d %.17g %.17g 0\n" % (nid[k], x, y))
        f.write("3 0.5 0.5 1\n")                    # a geometry-only point no element uses
        f.write("$EndNodes\n$Elements\n%d\n" % (len(quads) + 2))
        f.write("1 15 2 0 1 %d\n" % nid[0])
        f.write("2 1 2 7 1 %d %d\n" % (nid[quads[0][0]], nid[quads[0][1]]))
        for c, q in enumerate(quads):
            f.write("%d 3 2 1 1 %d %d %d %d\n" % (c + 3, *[nid[v] for v in q]))
        f.write("$EndElements\n")
