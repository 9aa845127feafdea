"""ParaView output of the time loop (reference: src/viewer/paraviewer.py:10-73): node coordinates once, the fields of
every saved step, and an XDMF index (`viewer/xml_generator.py`) that ParaView opens as a time series of point clouds.

The reference writes the heavy data through PETSc's HDF5 viewer.  Here the device vectors are copied to the host and
written as HDF5 with the same file / dataset names when h5py is importable, otherwise as raw little-endian float64
files that the XDMF index references with `Format="Binary"` (HDF5 is not available in this image).  Output is off the
hot path: it is written only when the case names a `save-dir` AND `PYNAMA_SAVE` is set (tests and benchmarks run the
case files as they are), or always with `PYNAMA_SAVE=always`.  On a partitioned mesh every rank writes its own strip
(`-rNN` suffix) in local numbering."""
import os

import numpy as np
import yaml

from .xml_generator import XmlGenerator

try:                                   # pragma: no cover - not installed in this image
    import h5py
except Exception:
    h5py = None


class Paraviewer:
    def __init__(self, dim=2, comm=None, saveDir=None):
        self.dim, self.comm = dim, comm
        mode = os.environ.get("PYNAMA_SAVE", "")
        self.saveDir = (saveDir or ".") if (mode == "always" or (mode and saveDir)) else None
        size = getattr(comm, "size", 1)
        self._tag = f"-r{getattr(comm, 'rank', 0):02d}" if size > 1 else ""
        self.h5name = "vec-data" + self._tag
        self.xmlWriter = XmlGenerator(dim, self.h5name, binary=h5py is None)
        if self.saveDir:
            os.makedirs(self.saveDir, exist_ok=True)

    # -- heavy data ---------------------------------------------------------------------------------------------
    def _write(self, h5file, fields):
        """fields: {dataset name: 1-D float64 array}"""
        if h5py is not None:
            with h5py.File(os.path.join(self.saveDir, h5file), "w") as f:
                g = f.create_group("fields")
                for name, a in fields.items():
                    g.create_dataset(name, data=a)
        else:
            stem = h5file[:-3]
            for name, a in fields.items():
                raw = "mesh.bin" if stem.startswith("mesh") else f"{stem}-{name}.bin"
                np.ascontiguousarray(a, dtype="<f8").tofile(os.path.join(self.saveDir, raw))

    def saveMesh(self, coords, name='mesh'):
        totalNodes = int(coords.getSize() / self.dim)
        self.xmlWriter.setUpDomainNodes(totalNodes=totalNodes)
        self.xmlWriter.generateXMLTemplate()
        if self.saveDir:
            coords.setName(name)
            self._write("mesh.h5", {"mesh": coords.getArray()})

    def saveData(self, step, time, *vecs):
        if not self.saveDir:
            return
        self.saveVec(vecs, step)
        self.saveStepInXML(step, time, vecs=vecs)

    def saveVec(self, vecs, step=None):
        if not self.saveDir:
            return
        if not isinstance(vecs, (tuple, list)):
            vecs = (vecs,)
        self._write(f"{self.h5name}-{int(step or 0):05d}.h5",
                    {(v.getName() or f"vec{i}"): v.getArray() for i, v in enumerate(vecs)})

    def saveStepInXML(self, step, time, vec=None, vecs=None):
        dataGrid = self.xmlWriter.generateMeshData("mesh1")
        self.xmlWriter.setTimeStamp(time, dataGrid)
        for i, v in enumerate([vec] if vec is not None else list(vecs)):
            name = v.getName() or f"vec{i}"
            if v.getSize() == self.xmlWriter.dimensions:
                self.xmlWriter.setScalarAttribute(name, step, dataGrid)
            else:
                self.xmlWriter.setVectorAttribute(name, step, dataGrid)

    def writeXmf(self, name):
        if self.saveDir:
            self.xmlWriter.writeFile(os.path.join(self.saveDir, (name or "case") + self._tag))

    def writeYaml(self, name, data):
        if self.saveDir:
            data = dict(data, dir=self.saveDir)
            with open(self.saveDir.rstrip("/") + ".yaml", "w") as outfile:
                yaml.dump(data, outfile, default_flow_style=False)
