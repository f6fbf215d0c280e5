"""Mesh ingestion without dolfinx / meshio / h5py -- SURVEY 8 row f4 (host side, not on the GPU path).

Drop-in for the readers of femo_alpha/fea/utils_dolfinx.py:
  * readFEAMesh(meshFile, format)            :34-50    XDMF written by meshio / dolfinx ('Grid' with Topology+Geometry)
  * import_mesh(prefix, subdomains, dim, directory)   :88-142   msh2xdmf triplet: <prefix>_domain.xdmf,
                                                        <prefix>_boundaries.xdmf, <prefix>_association_table.ini
  * reconstructFEAMesh(filename, nodes, connectivity)  :653-668  (in femo_alpha_b200/mesh.py)
plus a Gmsh `.msh` reader (ASCII 2.2 and 4.1), the format of the wing meshes under
examples/advanced_examples/*/data_files/*.msh, which the reference converts with msh2xdmf.

XDMF heavy data: inline `Format="XML"` DataItems are parsed here; `Format="HDF"` DataItems need `h5py`, which this
image does not have -- the reader then raises with that explanation instead of guessing (the checked-in example
meshes are git-LFS pointer files anyway, which is also detected and reported).
"""
from __future__ import annotations

import configparser
import os
import xml.etree.ElementTree as ET

import numpy as np

from .mesh import ShellMesh

_XDMF_CELLS = {"quadrilateral": 4, "quad": 4, "triangle": 3, "polyline": 2, "line": 2, "polyvertex": 1, "hexahedron": 8}


def _check_not_lfs_stub(path):
    with open(path, "rb") as f:
        head = f.read(64)
    if head.startswith(b"version https://git-lfs"):
        raise IOError("%s is a git-LFS pointer file, not a mesh (fetch the LFS objects of the reference checkout)" % path)


def _data_item(item, base_dir):
    fmt = (item.get("Format") or "XML").upper()
    dims = [int(v) for v in (item.get("Dimensions") or "").split()]
    kind = (item.get("NumberType") or item.get("DataType") or "Float").lower()
    dtype = np.int64 if kind in ("int", "uint") else np.float64
    if fmt == "XML":
        a = np.array((item.text or "").split(), dtype=dtype)
    elif fmt == "HDF":
        fname, dset = (item.text or "").strip().split(":", 1)
        path = os.path.join(base_dir, fname)
        _check_not_lfs_stub(path)
        try:
            import h5py  # type: ignore
        except Exception as e:
            raise ImportError("XDMF heavy data in HDF5 (%s) needs h5py, which is not installed; re-export the mesh with "
                              "ASCII/XML data items (writeFEAMesh) or pass nodes + connectivity to reconstructFEAMesh"
                              % fname) from e
        with h5py.File(path, "r") as h:
            a = np.asarray(h[dset], dtype=dtype)
    else:
        raise ValueError("unsupported XDMF DataItem format %r" % fmt)
    return a.reshape(dims) if dims and a.size == int(np.prod(dims)) else a


def _read_xdmf_grid(path, name="Grid"):
    _check_not_lfs_stub(path)
    root = ET.parse(path).getroot()
    grids = [g for g in root.iter("Grid") if g.find("Topology") is not None]
    if not grids:
        raise ValueError("%s: no uniform Grid with a Topology" % path)
    grid = next((g for g in grids if g.get("Name") == name), grids[0])
    base = os.path.dirname(os.path.abspath(path))
    topo = grid.find("Topology")
    ttype = (topo.get("TopologyType") or topo.get("Type") or "").lower()
    conn = _data_item(topo.find("DataItem"), base).astype(np.int64)
    if ttype == "mixed":
        raise NotImplementedError("mixed-topology XDMF grids are not supported")
    npc = _XDMF_CELLS.get(ttype)
    if npc is None:
        raise ValueError("%s: unsupported TopologyType %r" % (path, ttype))
    conn = conn.reshape(-1, npc)
    geo = grid.find("Geometry")
    x = _data_item(geo.find("DataItem"), base).astype(np.float64)
    gdim = 2 if (geo.get("GeometryType") or "XYZ").upper() == "XY" else 3
    x = x.reshape(-1, gdim)
    attrs = {}
    for at in grid.findall("Attribute"):
        attrs[at.get("Name")] = (at.get("Center"), _data_item(at.find("DataItem"), base))
    return x, conn, ttype, attrs


def readFEAMesh(meshFile, format='HDF'):
    """fea/utils_dolfinx.py:34-50.  `format` 'HDF' (xdmf + h5) or 'XML' (self-contained ASCII xdmf)."""
    if format not in ('HDF', 'XML'):
        raise ValueError("Invalid mesh file type. Must be 'HDF' or 'XML'")
    x, conn, ttype, _ = _read_xdmf_grid(meshFile, "Grid")
    if conn.shape[1] == 8:                  # hexahedral solid for the named 3-D forms (fea/named_forms.py); XDMF = VTK order
        from .fea.named_forms import SolidMesh
        return SolidMesh(x, conn)
    if conn.shape[1] not in (3, 4):
        raise ValueError('Invalid cell shape--should be either triangular or quadrilateral')
    return ShellMesh(x, conn)


def writeFEAMesh(meshFile, nodes, connectivity, cell_data=None):
    """Self-contained ASCII XDMF ('XML' format of readFEAMesh), the layout meshio writes."""
    nodes = np.asarray(nodes, dtype=np.float64)
    conn = np.asarray(connectivity, dtype=np.int64)
    ttype = {8: "Hexahedron", 4: "Quadrilateral", 3: "Triangle", 2: "Polyline"}[conn.shape[1]]
    root = ET.Element("Xdmf", Version="3.0")
    dom = ET.SubElement(root, "Domain")
    grid = ET.SubElement(dom, "Grid", Name="Grid", GridType="Uniform")
    geo = ET.SubElement(grid, "Geometry", GeometryType="XYZ" if nodes.shape[1] == 3 else "XY")
    di = ET.SubElement(geo, "DataItem", DataType="Float", Dimensions="%d %d" % nodes.shape, Format="XML", Precision="8")
    di.text = "\n" + "\n".join(" ".join(repr(float(v)) for v in row) for row in nodes) + "\n"
    topo = ET.SubElement(grid, "Topology", TopologyType=ttype, NumberOfElements=str(conn.shape[0]),
                         NodesPerElement=str(conn.shape[1]))
    di = ET.SubElement(topo, "DataItem", DataType="Int", Dimensions="%d %d" % conn.shape, Format="XML", Precision="8")
    di.text = "\n" + "\n".join(" ".join(str(int(v)) for v in row) for row in conn) + "\n"
    for name, vals in (cell_data or {}).items():
        vals = np.asarray(vals)
        at = ET.SubElement(grid, "Attribute", Name=name, AttributeType="Scalar", Center="Cell")
        di = ET.SubElement(at, "DataItem", DataType="Int", Dimensions=str(vals.size), Format="XML", Precision="8")
        di.text = "\n" + "\n".join(str(int(v)) for v in vals.ravel()) + "\n"
    ET.ElementTree(root).write(meshFile, xml_declaration=True, encoding="utf-8")


class MeshTags:
    """`dolfinx.mesh.MeshTags` stand-in: `indices` (entity numbers), `values` (tags), `dim`; `find(tag)`."""

    def __init__(self, dim, indices, values):
        self.dim = dim
        self.indices = np.asarray(indices, dtype=np.int64)
        self.values = np.asarray(values, dtype=np.int64)

    def find(self, tag):
        return self.indices[self.values == tag]

    def as_dict(self):
        """{tag: cell indices}: the `mesh_tags` argument of RMShellModel (rm_shell_model.py:101-133)."""
        return {int(t): self.find(t) for t in np.unique(self.values)}


def import_mesh(prefix='mesh', subdomains=False, dim=2, directory='.'):
    """fea/utils_dolfinx.py:88-142 (msh2xdmf output).  Boundary tags are returned on mesh edges given by their two
    vertex numbers (`MeshTags.indices` has shape (n, 2)): edge numbering is internal to each back-end."""
    domain = os.path.join(directory, '{}_domain.xdmf'.format(prefix))
    boundaries = os.path.join(directory, '{}_boundaries.xdmf'.format(prefix))
    x, conn, _, attrs = _read_xdmf_grid(domain, 'Grid')
    mesh = ShellMesh(x, conn)
    bx, bconn, _, battrs = _read_xdmf_grid(boundaries, 'Grid')
    bvals = next(iter(battrs.values()))[1] if battrs else np.zeros(bconn.shape[0], dtype=np.int64)
    if bx.shape[0] != x.shape[0] or not np.allclose(bx, x[:, :bx.shape[1]]):
        # the boundary file carries its own point list: map by coordinates
        from scipy.spatial import cKDTree
        _, idx = cKDTree(x[:, :bx.shape[1]]).query(bx)
        bconn = idx[bconn]
    boundaries_mf = MeshTags(dim - 1, bconn, bvals)
    table_path = os.path.join(directory, '{}_association_table.ini'.format(prefix))
    cp = configparser.ConfigParser()
    cp.read(table_path)
    association_table = {k: int(v) for k, v in dict(cp['ASSOCIATION TABLE']).items()} if cp.has_section('ASSOCIATION TABLE') else {}
    if not subdomains:
        return mesh, boundaries_mf, association_table
    svals = next(iter(attrs.values()))[1] if attrs else np.zeros(conn.shape[0], dtype=np.int64)
    return mesh, boundaries_mf, MeshTags(dim, np.arange(conn.shape[0]), svals), association_table


# ---- Gmsh .msh (ASCII 2.2 and 4.1) ---------------------------------------------------------------------------------
_GMSH_NODES = {1: 2, 2: 3, 3: 4, 5: 8, 15: 1}       # line, triangle, quadrangle, hexahedron (VTK node order), point


def read_msh(path):
    """Gmsh ASCII mesh -> dict(nodes (n,3), quads, triangles, lines, hexes, quad_tags, triangle_tags, line_tags, hex_tags,
    physical_names {name: (dim, tag)}).  Tags are the physical tags (first tag of 2.2, entity -> physical map of 4.1)."""
    _check_not_lfs_stub(path)
    with open(path, "r") as f:
        lines = f.read().split("\n")
    sections = {}
    i = 0
    while i < len(lines):
        ln = lines[i].strip()
        if ln.startswith("$") and not ln.startswith("$End"):
            name = ln[1:]
            j = i + 1
            while j < len(lines) and lines[j].strip() != "$End" + name:
                j += 1
            sections[name] = lines[i + 1:j]
            i = j
        i += 1
    if "MeshFormat" not in sections:
        raise ValueError("%s: not a Gmsh ASCII mesh" % path)
    ver, ftype = sections["MeshFormat"][0].split()[:2]
    if int(ftype) != 0:
        raise NotImplementedError("binary .msh files are not supported")
    names = {}
    for ln in sections.get("PhysicalNames", [])[1:]:
        p = ln.split(None, 2)
        if len(p) == 3:
            names[p[2].strip().strip('"')] = (int(p[0]), int(p[1]))
    cells = {2: ([], []), 3: ([], []), 4: ([], []), 8: ([], [])}       # nodes-per-cell -> (connectivity, tag)
    if ver.startswith("2"):
        nn = int(sections["Nodes"][0])
        ids = np.empty(nn, dtype=np.int64); xyz = np.empty((nn, 3))
        for k, ln in enumerate(sections["Nodes"][1:nn + 1]):
            p = ln.split(); ids[k] = int(p[0]); xyz[k] = [float(v) for v in p[1:4]]
        for ln in sections["Elements"][1:]:
            p = ln.split()
            if len(p) < 4:
                continue
            et, ntags = int(p[1]), int(p[2])
            npc = _GMSH_NODES.get(et)
            if npc in cells:
                cells[npc][0].append([int(v) for v in p[3 + ntags:3 + ntags + npc]])
                cells[npc][1].append(int(p[3]) if ntags > 0 else 0)
    elif ver.startswith("4"):
        ent_phys = {}
        E = sections.get("Entities", [])
        if E:
            npnt, ncur, nsur, nvol = [int(v) for v in E[0].split()]
            row = 1 + npnt
            for d, cnt in ((1, ncur), (2, nsur), (3, nvol)):
                for ln in E[row:row + cnt]:
                    p = ln.split()
                    nphys = int(p[7])
                    ent_phys[(d, int(p[0]))] = int(p[8]) if nphys > 0 else 0
                row += cnt
        N = sections["Nodes"]
        nblocks, nn = int(N[0].split()[0]), int(N[0].split()[1])
        ids = np.empty(nn, dtype=np.int64); xyz = np.empty((nn, 3))
        row, k = 1, 0
        for _ in range(nblocks):
            nb = int(N[row].split()[3]); row += 1
            tags = [int(v) for v in N[row:row + nb]]; row += nb
            for t in range(nb):
                ids[k] = tags[t]; xyz[k] = [float(v) for v in N[row + t].split()[:3]]; k += 1
            row += nb
        El = sections["Elements"]
        nblocks = int(El[0].split()[0])
        row = 1
        for _ in range(nblocks):
            d, etag, et, nb = [int(v) for v in El[row].split()]; row += 1
            npc = _GMSH_NODES.get(et)
            for ln in El[row:row + nb]:
                p = ln.split()
                if npc in cells:
                    cells[npc][0].append([int(v) for v in p[1:1 + npc]])
                    cells[npc][1].append(ent_phys.get((d, etag), 0))
            row += nb
    else:
        raise NotImplementedError("Gmsh format %s" % ver)
    lookup = np.full(ids.max() + 1, -1, dtype=np.int64)
    lookup[ids] = np.arange(ids.size)
    out = dict(nodes=xyz, physical_names=names)
    for npc, key in ((8, "hex"), (4, "quad"), (3, "triangle"), (2, "line")):
        c = np.array(cells[npc][0], dtype=np.int64).reshape(-1, npc)
        out[key + ("es" if key == "hex" else "s")] = lookup[c] if c.size else c
        out[key + "_tags"] = np.array(cells[npc][1], dtype=np.int64)
    return out


def readMshShellMesh(path):
    """Quadrilateral shell mesh + {physical tag: cell indices} from a Gmsh file: the (mesh, mesh_tags) pair of
    RMShellModel(mesh, ..., mesh_tags=...) (rm_shell_model.py:101-133).  Nodes not used by any quad are dropped."""
    m = read_msh(path)
    quads = m["quads"]
    if quads.size == 0:
        if m["triangles"].size:
            raise NotImplementedError("triangular (P2xP1) shells are not implemented in femo_alpha_b200 yet")
        raise ValueError("%s holds no surface cells" % path)
    used = np.unique(quads)
    remap = np.full(m["nodes"].shape[0], -1, dtype=np.int64)
    remap[used] = np.arange(used.size)
    mesh = ShellMesh(m["nodes"][used], remap[quads])
    tags = MeshTags(2, np.arange(quads.shape[0]), m["quad_tags"])
    return mesh, tags.as_dict(), m["physical_names"]


def readMshSolidMesh(path):
    """Hexahedral solid mesh + {physical tag: cell indices} + physical names from a Gmsh file, for the named 3-D forms
    (fea/named_forms.py: build_fea).  Nodes not used by any hexahedron are dropped."""
    from .fea.named_forms import SolidMesh
    m = read_msh(path)
    hexes = m["hexes"]
    if hexes.size == 0:
        raise ValueError("%s holds no hexahedra" % path)
    used = np.unique(hexes)
    remap = np.full(m["nodes"].shape[0], -1, dtype=np.int64)
    remap[used] = np.arange(used.size)
    mesh = SolidMesh(m["nodes"][used], remap[hexes])
    tags = MeshTags(3, np.arange(hexes.shape[0]), m["hex_tags"])
    return mesh, tags.as_dict(), m["physical_names"]
