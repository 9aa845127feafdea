"""Headless stand-in for the reference's HDF5/XDMF writer (src/viewer/paraviewer.py:18-73): offline visualisation
I/O is outside the hot path.  The hooks called from the time loop are kept; with a
save directory they dump plain .npz files, otherwise they do nothing."""
import os

import numpy as np


class Paraviewer:
    def __init__(self, dim=2, comm=None, saveDir=None):
        self.dim, self.comm = dim, comm
        self.saveDir = saveDir if os.environ.get("PYNAMA_SAVE") else None
        if self.saveDir:
            os.makedirs(self.saveDir, exist_ok=True)
        # one file set per rank of a partitioned run (every rank holds its strip + ghosts in local numbering)
        size = getattr(comm, "size", 1)
        self._tag = f"-r{getattr(comm, 'rank', 0):02d}" if size > 1 else ""

    def saveMesh(self, coords, name="mesh"):
        if self.saveDir:
            np.savez(os.path.join(self.saveDir, name + self._tag + ".npz"), coords=coords.getArray())

    def saveData(self, step, time, *vecs):
        if self.saveDir:
            np.savez(os.path.join(self.saveDir, f"step-{step:05d}{self._tag}.npz"), time=time,
                     **{(v.getName() or f"vec{i}"): v.getArray() for i, v in enumerate(vecs)})

    def saveVec(self, vec, timeStep=None):
        self.saveData(timeStep or 0, 0.0, vec)

    def saveStepInXML(self, *a, **k):
        pass

    def writeXmf(self, name):
        pass
