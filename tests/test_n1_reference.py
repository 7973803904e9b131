"""Row N1 pinned against the reference's own Python (convert_to_binary.py): the golden .binda files under
tests/golden/ and neighbors_golden.json are OUTPUT OF THE UNMODIFIED REFERENCE CONVERTER (BindaWriter :11-115,
get_cell_neighbors :118-162), produced by tests/golden/make_fixtures.py with a stub for `meshio.read` only.
CPU tests: the repo's writer must reproduce the files byte for byte, and both neighbour builders (the C++ one inside
libiu_b200.so and the numpy one) must reproduce the reference's table on a 118k-tet jittered mesh with duplicated points."""
import hashlib
import importlib.util
import json
import os

import numpy as np
import pytest

from interpolate_unstructured_b200 import api, binda, meshgen
from oracle import oracle

HERE = os.path.dirname(os.path.abspath(__file__))
GOLD = os.path.join(HERE, "golden")
FIXTURES = ["triangle", "quad", "tetra", "tetra_cell", "tetra_two_cell"]
HAVE_REF = os.path.exists("/root/reference/convert_to_binary.py")


def _make_fixtures():
    spec = importlib.util.spec_from_file_location("make_fixtures", os.path.join(GOLD, "make_fixtures.py"))
    m = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(m)
    return m


@pytest.mark.parametrize("name", FIXTURES)
def test_writer_reproduces_reference_file_bytes(name, tmp_path):
    """entries of the reference's file -> binda.write_grid_binda -> identical bytes"""
    gold = os.path.join(GOLD, name + ".binda")
    ents = binda.read_binda(gold)
    by = {e.name: e for e in ents if e.name in ("points", "cells", "cell_neighbors")}
    pdata = {e.metadata: e.data for e in ents if e.name == "point_data"}
    cdata = {e.metadata: e.data for e in ents if e.name in ("cell_data", "icell_data")}
    out = tmp_path / (name + ".binda")
    binda.write_grid_binda(str(out), by["points"].data, by["cells"].data, by["cells"].metadata, by["cell_neighbors"].data, pdata, cdata)
    assert open(out, "rb").read() == open(gold, "rb").read()


@pytest.mark.parametrize("name", FIXTURES)
def test_neighbour_builders_reproduce_reference_fixture_tables(name):
    ents = {e.name: e for e in binda.read_binda(os.path.join(GOLD, name + ".binda")) if e.name in ("points", "cells", "cell_neighbors")}
    pts, cells, ct = ents["points"].data.astype(np.float64), ents["cells"].data, ents["cells"].metadata
    assert (api.build_cell_neighbors(cells, pts, ct) == ents["cell_neighbors"].data).all()
    assert (meshgen.cell_neighbors(cells, pts, ct) == ents["cell_neighbors"].data).all()
    assert (oracle.cell_neighbors(cells, pts, ct) == ents["cell_neighbors"].data).all()     # the oracle's restatement (bench.py's reference arm)


def test_neighbour_builders_match_reference_on_118k_tets_with_duplicate_points():
    gold = json.load(open(os.path.join(GOLD, "neighbors_golden.json")))
    pts, cells = _make_fixtures().neighbor_case()
    assert len(cells) == gold["n_cells"] and len(pts) == gold["n_points"] and gold["n_cells"] >= 100_000
    nb_cpp = api.build_cell_neighbors(cells, pts, "tetra")
    assert int((nb_cpp < 0).sum()) == gold["boundary_faces"]
    assert hashlib.sha256(np.ascontiguousarray(nb_cpp, dtype=np.int32).tobytes()).hexdigest() == gold["sha256_int32"]
    nb_np = meshgen.cell_neighbors(cells, pts, "tetra")
    assert hashlib.sha256(np.ascontiguousarray(nb_np, dtype=np.int32).tobytes()).hexdigest() == gold["sha256_int32"]
    nb_or = oracle.cell_neighbors(cells, pts, "tetra")
    assert hashlib.sha256(np.ascontiguousarray(nb_or, dtype=np.int32).tobytes()).hexdigest() == gold["sha256_int32"]


@pytest.mark.skipif(not HAVE_REF, reason="reference tree not present (GPU box)")
def test_golden_files_are_what_the_reference_converter_writes_today(tmp_path):
    """re-runs the unmodified converter (meshio stubbed) and compares with the committed files"""
    mf = _make_fixtures()
    for name in FIXTURES:
        mf.run_reference_converter(os.path.join(mf.REF, name + ".vtu"), str(tmp_path / name))
        assert open(tmp_path / (name + ".binda"), "rb").read() == open(os.path.join(GOLD, name + ".binda"), "rb").read()


@pytest.mark.skipif(not HAVE_REF, reason="reference tree not present (GPU box)")
def test_cpp_builder_equals_reference_function_live():
    mf = _make_fixtures()
    ref = mf.run_reference_converter(os.path.join(mf.REF, "tetra_cell.vtu"), "/tmp/_iu_n1_probe")
    rng = np.random.default_rng(5)
    for ct, npf, (pts, cells) in [("tetra", 3, meshgen.kuhn_tet_mesh(9, jitter=0.15, seed=3, shuffle=True)),
                                  ("quad", 2, meshgen.quad_mesh(40)), ("triangle", 2, meshgen.triangle_mesh(40))]:
        dup = rng.integers(0, len(pts), 50)                       # 50 extra copies of existing points, used by random cells
        pts2 = np.concatenate([pts, pts[dup]])
        cells2 = cells.copy()
        for j, d in enumerate(dup):
            hit = np.argwhere(cells2 == d)
            if len(hit):
                c, k = hit[rng.integers(len(hit))]
                cells2[c, k] = len(pts) + j
        want = ref["get_cell_neighbors"](cells2, pts2, npf)
        assert (api.build_cell_neighbors(cells2, pts2, ct) == want).all()
        assert (oracle.cell_neighbors(cells2, pts2, ct) == want).all()
