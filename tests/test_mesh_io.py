"""Mesh readers of the drop-in surface (femo_alpha/fea/utils_dolfinx.py:34-50,88-142): XDMF (ASCII),
msh2xdmf triplets and Gmsh .msh 2.2 / 4.1, round-tripped on generated files (the reference's checked-in meshes are
git-LFS pointer files)."""
import os

import numpy as np
import pytest

from femo_alpha_b200.mesh import createRectangleMesh
from femo_alpha_b200 import mesh_io as IO


def test_xdmf_ascii_round_trip(tmp_path):
    m = createRectangleMesh((0., 0.), (10., 2.), 5, 2)
    path = str(tmp_path / "plate.xdmf")
    IO.writeFEAMesh(path, m.geometry.x, m.cells)
    for fmt in ("XML", "HDF"):                     # an ASCII file is self-contained under either keyword
        r = IO.readFEAMesh(path, format=fmt)
        assert np.array_equal(r.cells, m.cells) and np.array_equal(r.geometry.x, m.geometry.x)
    with pytest.raises(ValueError):
        IO.readFEAMesh(path, format="VTK")


def test_lfs_pointer_and_hdf_are_reported(tmp_path):
    p = tmp_path / "stub.xdmf"
    p.write_text("version https://git-lfs.github.com/spec/v1\noid sha256:00\nsize 580\n")
    with pytest.raises(IOError, match="git-LFS"):
        IO.readFEAMesh(str(p))
    q = tmp_path / "heavy.xdmf"
    q.write_text('<Xdmf Version="3.0"><Domain><Grid Name="Grid"><Geometry GeometryType="XYZ">'
                 '<DataItem DataType="Float" Dimensions="4 3" Format="HDF" Precision="8">heavy.h5:/data0</DataItem></Geometry>'
                 '<Topology TopologyType="Quadrilateral" NumberOfElements="1" NodesPerElement="4">'
                 '<DataItem DataType="Int" Dimensions="1 4" Format="XML">0 1 2 3</DataItem></Topology></Grid></Domain></Xdmf>')
    (tmp_path / "heavy.h5").write_bytes(b"\x89HDF\r\n\x1a\n" + b"\0" * 64)
    with pytest.raises(ImportError, match="h5py"):
        IO.readFEAMesh(str(q))


def test_import_mesh_triplet(tmp_path):
    m = createRectangleMesh((0., 0.), (2., 1.), 2, 1)
    sub = np.array([7, 9])
    IO.writeFEAMesh(str(tmp_path / "wing_domain.xdmf"), m.geometry.x, m.cells, {"subdomains": sub})
    edges = np.array([[0, 1], [1, 3]])
    IO.writeFEAMesh(str(tmp_path / "wing_boundaries.xdmf"), m.geometry.x, edges, {"boundaries": [3, 4]})
    (tmp_path / "wing_association_table.ini").write_text("[ASSOCIATION TABLE]\nroot = 3\ntip = 4\nskin = 7\nspar = 9\n")
    mesh, bnd, subs, table = IO.import_mesh("wing", subdomains=True, directory=str(tmp_path))
    assert np.array_equal(mesh.cells, m.cells)
    assert table == dict(root=3, tip=4, skin=7, spar=9)
    assert np.array_equal(bnd.find(table["tip"]), [[1, 3]])
    assert subs.as_dict() == {7: [0], 9: [1]} or all(np.array_equal(subs.as_dict()[k], v) for k, v in {7: [0], 9: [1]}.items())
    mesh2, bnd2, table2 = IO.import_mesh("wing", directory=str(tmp_path))
    assert table2 == table


MSH22 = """$MeshFormat
2.2 0 8
$EndMeshFormat
$PhysicalNames
2
2 11 "upper skin"
2 12 "spar"
$EndPhysicalNames
$Nodes
7
1 0 0 0
2 1 0 0
3 2 0 0
4 0 1 0
5 1 1 0
6 2 1 0
9 5 5 5
$EndNodes
$Elements
4
1 1 2 5 1 1 2
2 3 2 11 1 1 2 5 4
3 3 2 12 2 2 3 6 5
4 15 2 0 1 9
$EndElements
"""

MSH41 = """$MeshFormat
4.1 0 8
$EndMeshFormat
$PhysicalNames
1
2 12 "spar"
$EndPhysicalNames
$Entities
0 0 2 0
1 0 0 0 1 1 0 1 11 0
2 1 0 0 2 1 0 1 12 0
$EndEntities
$Nodes
1 6 1 6
2 1 0 6
1
2
3
4
5
6
0 0 0
1 0 0
2 0 0
0 1 0
1 1 0
2 1 0
$EndNodes
$Elements
2 2 1 2
2 1 3 1
1 1 2 5 4
2 2 3 1
2 2 3 6 5
$EndElements
"""


@pytest.mark.parametrize("text", [MSH22, MSH41])
def test_gmsh_reader(tmp_path, text):
    p = tmp_path / "wing.msh"
    p.write_text(text)
    mesh, tags, names = IO.readMshShellMesh(str(p))
    assert mesh.geometry.x.shape == (6, 3)                      # the stray point is dropped
    assert np.array_equal(mesh.cells, [[0, 1, 4, 3], [1, 2, 5, 4]])
    assert np.array_equal(tags[11], [0]) and np.array_equal(tags[12], [1])
    assert names["spar"] == (2, 12)


def test_reference_wing_msh_is_an_lfs_pointer():
    ref = "/root/reference/examples/advanced_examples/lpc_shell_opt/data_files/left_wing_c.msh"
    if not os.path.exists(ref):
        pytest.skip("reference checkout not present")
    with pytest.raises(IOError, match="git-LFS"):
        IO.read_msh(ref)


# ---- hexahedral solids (named 3-D forms) ---------------------------------------------------------------------
def _two_hexes():
    nodes = np.array([[x, y, z] for z in (0.0, 1.0) for y in (0.0, 1.0) for x in (0.0, 1.0, 2.0)], dtype=float)
    # node id = x + 3 y + 6 z ; VTK_HEXAHEDRON order
    cells = np.array([[0, 1, 4, 3, 6, 7, 10, 9], [1, 2, 5, 4, 7, 8, 11, 10]])
    return nodes, cells


def test_xdmf_hexahedron_round_trip(tmp_path):
    from femo_alpha_b200.fea.named_forms import SolidMesh
    nodes, cells = _two_hexes()
    p = str(tmp_path / "block.xdmf")
    IO.writeFEAMesh(p, nodes, cells)
    mesh = IO.readFEAMesh(p, 'XML')
    assert isinstance(mesh, SolidMesh)
    assert np.array_equal(mesh.cells, cells) and np.allclose(mesh.geometry.x, nodes)
    assert mesh.topology.dim == 3 and mesh.topology.index_map(3).size_local == 2 and mesh.topology.index_map(0).size_local == 12
    assert mesh.ufl_cell().cellname() == "hexahedron"


def _msh22_hex():
    nodes, cells = _two_hexes()
    lines = ["$MeshFormat", "2.2 0 8", "$EndMeshFormat", "$PhysicalNames", "2", '3 7 "core"', '3 8 "shell"', "$EndPhysicalNames",
             "$Nodes", str(len(nodes) + 1)]
    lines += ["%d %r %r %r" % (i + 1, *map(float, nodes[i])) for i in range(len(nodes))] + ["40 9 9 9", "$EndNodes", "$Elements", "3"]
    lines += ["1 5 2 7 1 " + " ".join(str(v + 1) for v in cells[0]), "2 5 2 8 2 " + " ".join(str(v + 1) for v in cells[1]),
              "3 15 2 0 1 40", "$EndElements", ""]
    return "\n".join(lines)


def _msh41_hex():
    nodes, cells = _two_hexes()
    lines = ["$MeshFormat", "4.1 0 8", "$EndMeshFormat", "$PhysicalNames", "1", '3 8 "shell"', "$EndPhysicalNames",
             "$Entities", "0 0 0 2", "1 0 0 0 1 1 1 1 7 0", "2 1 0 0 2 1 1 1 8 0", "$EndEntities",
             "$Nodes", "1 12 1 12", "3 1 0 12"] + [str(i + 1) for i in range(12)]
    lines += ["%r %r %r" % tuple(map(float, nodes[i])) for i in range(12)] + ["$EndNodes", "$Elements", "2 2 1 2", "3 1 5 1",
              "1 " + " ".join(str(v + 1) for v in cells[0]), "3 2 5 1", "2 " + " ".join(str(v + 1) for v in cells[1]),
              "$EndElements", ""]
    return "\n".join(lines)


@pytest.mark.parametrize("text", [_msh22_hex(), _msh41_hex()], ids=["msh22", "msh41"])
def test_gmsh_hexahedra(tmp_path, text):
    nodes, cells = _two_hexes()
    p = tmp_path / "block.msh"
    p.write_text(text)
    mesh, tags, names = IO.readMshSolidMesh(str(p))
    assert mesh.geometry.x.shape == (12, 3) and np.allclose(mesh.geometry.x, nodes)      # the stray point is dropped
    assert np.array_equal(mesh.cells, cells)
    assert np.array_equal(tags[7], [0]) and np.array_equal(tags[8], [1])
    assert names["shell"] == (3, 8)
    with pytest.raises(ValueError):
        q = tmp_path / "wing.msh"; q.write_text(MSH22)
        IO.readMshSolidMesh(str(q))
    # the mesh feeds the oracle of the named forms: positive Jacobians, volume 2
    from oracle.hex_ref import HexOracle
    assert np.isclose(HexOracle(mesh.geometry.x, mesh.cells).vol.sum(), 2.0)
