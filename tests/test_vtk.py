"""iu_write_vtk (m_interp_unstructured.f90:930-985 through m_vtk.f90): the library's VTU writer against the
Python restatement of the reference's byte stream (oracle/vtk_oracle.py).  The reference's own
test_vtk.f90 only writes three files and checks nothing, so the format itself is the fixture."""
import os

import numpy as np
import pytest

from interpolate_unstructured_b200 import api, binda, meshgen
from oracle import vtk_oracle

GOLDEN = os.path.join(os.path.dirname(__file__), "golden")


def _mesh(kind, n):
    if kind == "tetra":
        return meshgen.kuhn_tet_mesh(n, jitter=0.1, seed=3)
    if kind == "quad":
        return meshgen.quad_mesh(n, jitter=0.2, seed=3)
    return meshgen.triangle_mesh(n, jitter=0.2, seed=3)


@pytest.mark.parametrize("kind,ct,n", [("triangle", 1, 5), ("quad", 2, 4), ("tetra", 3, 3)])
def test_host_writer_matches_reference_byte_stream(tmp_path, kind, ct, n):
    pts, cells = _mesh(kind, n)
    rng = np.random.default_rng(7)
    cd = {"cell volume": rng.normal(size=len(cells)), "b": rng.normal(size=len(cells))}
    icd = {"material": rng.integers(0, 4, size=len(cells)).astype(np.int32)}
    pd = {"phi": rng.normal(size=len(pts)), "Ex": rng.normal(size=len(pts)), "Ey": rng.normal(size=len(pts))}
    path = str(tmp_path / f"{kind}.vtu")
    api.write_vtk_arrays(path, kind, pts, cells + 1, cd, icd, pd)
    got = open(path, "rb").read()
    want = vtk_oracle.vtu_bytes(ct, pts, cells + 1, cd, icd, pd)
    assert got == want
    arrays = vtk_oracle.parse_vtu(got)
    assert np.array_equal(arrays["Points"].reshape(-1, 3), pts.astype(np.float32))
    assert np.array_equal(arrays["connectivity"].reshape(-1, cells.shape[1]), cells)
    assert np.array_equal(arrays["offsets"], cells.shape[1] * np.arange(1, len(cells) + 1))
    assert set(arrays["types"]) == {{1: 5, 2: 9, 3: 10}[ct]}
    assert np.array_equal(arrays["phi"], pd["phi"]) and np.array_equal(arrays["material"], icd["material"])


def test_host_writer_without_data_and_unwritable_path(tmp_path):
    pts, cells = _mesh("triangle", 2)
    path = str(tmp_path / "bare.vtu")
    api.write_vtk_arrays(path, "triangle", pts, cells + 1)
    assert open(path, "rb").read() == vtk_oracle.vtu_bytes(1, pts, cells + 1)
    with pytest.raises(api.IUError):
        api.write_vtk_arrays(str(tmp_path / "no_such_dir" / "x.vtu"), "triangle", pts, cells + 1)


@pytest.mark.gpu
@pytest.mark.parametrize("name,ct", [("triangle", 1), ("quad", 2), ("tetra", 3)])
def test_grid_write_vtk_from_device(tmp_path, name, ct):
    """iu_read_grid -> iu_add_point_data -> iu_write_vtk as in test_trace_field.f90:17-25: the file is rebuilt
    from the DEVICE copy (records, data tables) and must equal the reference stream of the host mesh."""
    src = os.path.join(GOLDEN, name + ".binda")
    g = api.Grid.iu_read_grid(src)
    e = {x.name: x for x in binda.read_binda(src)}
    pts = np.asarray(e["points"].data, dtype=np.float64).reshape(-1, 3)
    cells1 = np.asarray(e["cells"].data).reshape(g.n_cells, -1).astype(np.int32) + 1
    p2, c2 = g.export_mesh()
    assert np.array_equal(p2, pts) and np.array_equal(c2, cells1)
    i_new = g.iu_add_point_data("extra")
    extra = np.linspace(0.0, 1.0, g.n_points)
    g.set_point_data(i_new, extra)
    path = str(tmp_path / "out.vtu")
    g.iu_write_vtk(path)
    arrays = vtk_oracle.parse_vtu(open(path, "rb").read())
    assert np.array_equal(arrays["extra"], extra)
    pd = {g.point_data_name(i): g.get_point_data(i)[:, 0] for i in range(1, g.n_point_data + 1)}
    cd = {g.cell_data_name(i): g.get_cell_data(i)[:, 0] for i in range(1, g.n_cell_data + 1)} if g.n_cell_data else {}
    icd = {g.icell_data_name(i): g.get_icell_data(i)[:, 0] for i in range(1, g.n_icell_data + 1)} if g.n_icell_data else {}
    assert open(path, "rb").read() == vtk_oracle.vtu_bytes(ct, pts, cells1, cd, icd, pd)


@pytest.mark.gpu
def test_grid_write_vtk_keeps_points_no_cell_references(tmp_path):
    """meshes converted by meshio often carry points that no cell uses (and duplicates of merged points).
    iu_write_vtk writes ug%points unchanged (m_interp_unstructured.f90:961), so those points must come back with their
    coordinates, not as zeros; named columns given to from_arrays must be the VTU array names."""
    pts, cells = meshgen.kuhn_tet_mesh(3, jitter=0.1, seed=3)
    orphans = np.array([[7.5, -2.25, 3.125], [pts[5, 0], pts[5, 1], pts[5, 2]], [-1.0, -1.0, -1.0]])
    pts2 = np.concatenate([pts[:10], orphans[:2], pts[10:], orphans[2:]])          # orphans in the middle and at the end
    remap = np.concatenate([np.arange(10), np.arange(10, len(pts)) + 2])
    cells2 = remap[cells].astype(np.int32)
    nb = api.build_cell_neighbors(cells2, pts2, "tetra")
    rng = np.random.default_rng(3)
    pd = rng.normal(size=(len(pts2), 2))
    cd = rng.normal(size=(len(cells2), 1))
    g = api.Grid.from_arrays("tetra", pts2, cells2, nb, pd, cell_data=cd, point_names=["phi", "rho"], cell_names=["vol"])
    assert g.iu_get_point_data_index("rho") == 2 and g.iu_get_cell_data_index("vol") == 1 and g.iu_get_point_data_index("nope") == -1
    p_back, c_back = g.export_mesh()
    assert np.array_equal(p_back, pts2) and np.array_equal(c_back, cells2 + 1)
    path = str(tmp_path / "orphans.vtu")
    g.iu_write_vtk(path)
    got = open(path, "rb").read()
    assert got == vtk_oracle.vtu_bytes(3, pts2, cells2 + 1, {"vol": cd[:, 0]}, {}, {"phi": pd[:, 0], "rho": pd[:, 1]})
    arrays = vtk_oracle.parse_vtu(got)
    assert np.array_equal(arrays["Points"].reshape(-1, 3)[[10, 11, len(pts2) - 1]], orphans.astype(np.float32))
    # queries still resolve (the orphans belong to no cell and change nothing)
    r = meshgen.uniform_queries(2000, [0.1] * 3, [0.9] * 3, seed=2)
    _, ic = g.iu_interpolate_at(r, [1])
    assert (ic >= 1).all()
