"""Node / dof index bookkeeping (reference: src/domain/indices.py).  The PetscSection of the reference is replaced
by the closed-form serial numbering of `domain.dmplex.DMPlexDom` (section offsets in point order)."""
import logging

import numpy as np


class IndicesManager:
    def __init__(self, dim, ngl, comm=None):
        self.logger = logging.getLogger("[{}] IndicesManager Class".format(getattr(comm, "rank", 0)))
        self.comm = comm
        self.dim = dim
        self._ngl = ngl
        self.__dirNodes = set()
        self.__nsNodes = set()
        self.reorderEntities = self.reorderEntities2D if dim == 2 else self.reorderEntities3D

    def getNGL(self):
        return self._ngl

    def getNumCompAndNumDof(self, componentsPerField, numFields):
        """dofs per mesh entity: 1 node per vertex, ngl-2 per edge, (ngl-2)^2 per cell (indices.py:22-30)."""
        numComp = [componentsPerField] * numFields
        nodesPerEntity = [1, self._ngl - 2, (self._ngl - 2) ** 2]
        if self.dim == 3:
            nodesPerEntity.append((self._ngl - 2) ** 3)
        return numComp, [componentsPerField * k for k in nodesPerEntity]

    def setDirichletNodes(self, nodes):
        self.__dirNodes |= set(int(x) for x in nodes)

    def getDirichletNodes(self):
        self.globalIndicesDIR = set(self.__dirNodes)     # replicated instead of the reference's allgather (:51-55)
        return self.__dirNodes

    def setNoSlipNodes(self, nodes):
        self.__nsNodes |= set(int(x) for x in nodes)

    def getNoSlipNodes(self):
        self.globalIndicesNS = set(self.__nsNodes)
        return self.__nsNodes

    def mapNodesToIndices(self, nodes, dof):
        """node list -> interleaved dof list node*dof + d (indices.py:90-92)."""
        return [int(x) * dof + d for x in nodes for d in range(dof)]

    @staticmethod
    def reorderEntities2D(entities):
        """closure order (cell, 4 edges, 4 vertices) -> (vertices, edges, cell) (indices.py:116-118)."""
        e = np.asarray(entities)
        return np.hstack((e[5:], e[1:5], e[0]))

    @staticmethod
    def reorderEntities3D(entities):
        e = np.asarray(entities)
        return np.hstack((e[19:], e[7:19], e[1:7], e[0]))
