"""XDMF index of the saved fields (reference: src/viewer/xml_generator.py:4-122) -- same document structure (Xdmf 2.0,
temporal collection of Polyvertex grids, XY geometry, scalar attributes and vector attributes joined from hyperslabs),
with the heavy data either in HDF5 files laid out like the reference's (`mesh.h5:/fields/mesh`,
`vec-data-00012.h5:/fields/<name>`, written when h5py is importable) or in raw little-endian float64 files (XDMF
`Format="Binary"`), which ParaView's Xdmf reader opens without HDF5."""
from xml.dom import minidom
from xml.etree.ElementTree import Element, SubElement, tostring


class XmlGenerator(object):
    def __init__(self, dim, h5name, binary=True):
        self.root = Element('Xdmf')
        self.root.set('Version', '2.0')
        self.dim = dim
        self.h5name = h5name
        self.binary = binary

    def setUpDomainNodes(self, totalNodes=None, nodesPerDim=None):
        if totalNodes is not None:
            self.dimensions = totalNodes
        else:
            assert len(nodesPerDim) == self.dim
            self.dimensions = 1
            for nodes in nodesPerDim:
                self.dimensions *= nodes

    def generateXMLTemplate(self):
        self.domain = SubElement(self.root, 'Domain')
        self.grid = SubElement(self.domain, 'Grid')
        self.grid.set('Name', "TimeSeries")
        self.grid.set('GridType', "Collection")
        self.grid.set('CollectionType', "Temporal")

    def _heavy(self, item, h5file, dataset, rawfile):
        item.set("NumberType", "Float")
        item.set("Precision", "8")
        if self.binary:
            item.set("Format", "Binary")
            item.set("Endian", "Little")
            item.text = rawfile
        else:
            item.set("Format", "HDF")
            item.text = f"{h5file}:/fields/{dataset}"

    def generateMeshData(self, name):
        meshGrid = SubElement(self.grid, 'Grid')
        meshGrid.set('Name', name)
        meshGrid.set('GridType', "uniform")
        topology = SubElement(meshGrid, "Topology")
        topology.set("TopologyType", "Polyvertex")
        topology.set("Dimensions", str(self.dimensions))
        geometry = SubElement(meshGrid, "Geometry")
        geometry.set("GeometryType", "XY" if self.dim == 2 else "XYZ")
        geometryData = SubElement(geometry, "DataItem")
        geometryData.set("Dimensions", str(self.dimensions * self.dim))
        self._heavy(geometryData, "mesh.h5", "mesh", "mesh.bin")
        return meshGrid

    def setTimeStamp(self, t, meshElem):
        timestamp = SubElement(meshElem, "Time")
        timestamp.set("Value", str(t))

    def setVectorAttribute(self, name, step, meshGrid):
        attr = SubElement(meshGrid, "Attribute")
        attr.set("Name", name)
        attr.set("AttributeType", "Vector")
        attr.set("Center", "Node")
        attrData = SubElement(attr, "DataItem")
        attrData.set("ItemType", "Function")
        attrData.set("Dimensions", "{} {}".format(self.dimensions, self.dim))
        attrData.set("Function", self.getJoinString(self.dim))
        for i in range(self.dim):
            self.setDataToAttribute(attrData, step, name, i)

    def setScalarAttribute(self, name, step, meshGrid):
        attr = SubElement(meshGrid, "Attribute")
        attr.set("Name", name)
        attr.set("AttributeType", "Scalar")
        attr.set("Center", "Node")
        attrData = SubElement(attr, "DataItem")
        attrData.set("Dimensions", "{}".format(self.dimensions))
        self._heavy(attrData, f"{self.h5name}-{step:05d}.h5", name, f"{self.h5name}-{step:05d}-{name}.bin")

    def setDataToAttribute(self, attrData, step, name, dof):
        dofs = ['X', 'Y', 'Z']
        hyperSlab = SubElement(attrData, "DataItem")
        hyperSlab.set("ItemType", "HyperSlab")
        hyperSlab.set("Dimensions", str(self.dimensions))
        hyperSlab.set("Name", "{}-{}".format(name, dofs[dof]))
        dimData = SubElement(hyperSlab, "DataItem")
        dimData.set("Dimensions", "3 1")
        dimData.set("Format", "XML")
        dimData.text = "{} {} {}".format(dof, self.dim, self.dimensions)      # start, stride, count
        velData = SubElement(hyperSlab, "DataItem")
        velData.set("Dimensions", str(self.dimensions * self.dim))
        self._heavy(velData, f"{self.h5name}-{step:05d}.h5", name, f"{self.h5name}-{step:05d}-{name}.bin")

    def writeFile(self, nameFile):
        rough = tostring(self.root, 'utf-8')
        with open("{}.xmf".format(nameFile), "w") as f:
            f.write(minidom.parseString(rough).toprettyxml(indent=" "))

    @staticmethod
    def formatStep(step):
        return f"{int(step):05d}"

    @staticmethod
    def getJoinString(dof):
        return "JOIN($0, $1)" if dof == 2 else "JOIN($0, $1, $2)"
