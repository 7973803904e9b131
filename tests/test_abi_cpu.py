"""CPU-only checks of the drop-in boundary: libiu_b200.so loads, exports every symbol that
include/iu_b200.h declares, refuses to compute without a GPU (no CPU fallback), and its host-only
utilities (face-neighbour builder, BINDA reader errors) behave like the reference tools."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from interpolate_unstructured_b200 import api, binda, meshgen

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(scope="module")
def lib():
    api.build_library()
    return api.lib()


def _declared_symbols():
    text = open(os.path.join(ROOT, "include", "iu_b200.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(iu_b200_[a-z0-9_]+)\s*\(", text)))


def test_every_declared_symbol_is_exported_and_bound(lib):
    names = _declared_symbols()
    assert len(names) >= 40
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/iu_b200.h but not exported"
        assert n in api._SIGNATURES, f"{n} has no ctypes signature in api.py"
    assert lib.iu_b200_version() >= 100


def test_no_cpu_fallback(lib, golden_dir):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present; the fail-loudly path is for GPU-less hosts")
    with pytest.raises(api.IUError) as e:
        api.Grid.iu_read_grid(os.path.join(golden_dir, "tetra.binda"))
    assert e.value.code == api.E_CUDA
    pts, cells = meshgen.kuhn_tet_mesh(2, jitter=0.0)
    nb = meshgen.cell_neighbors(cells, pts, "tetra")
    with pytest.raises(api.IUError) as e:
        api.Grid.from_arrays("tetra", pts, cells, nb)
    assert e.value.code == api.E_CUDA and "no CPU fallback" in str(e.value)


def test_binda_reader_errors(lib, tmp_path):
    h = C.c_void_p()
    bad = tmp_path / "bad.binda"
    bad.write_bytes(b"NOTBINDA" + b"\0" * 64)
    assert lib.iu_b200_grid_from_binda(str(bad).encode(), 1.0, -1, C.byref(h)) == api.E_IO
    assert b"Wrong file format" in lib.iu_b200_last_error()          # m_binda.f90:59
    assert lib.iu_b200_grid_from_binda(str(tmp_path / "missing.binda").encode(), 1.0, -1, C.byref(h)) == api.E_IO
    # a container without "cells"
    p = tmp_path / "nocells.binda"
    binda.write_binda(str(p), [("points", np.zeros((3, 3)), "")])
    assert lib.iu_b200_grid_from_binda(str(p).encode(), 1.0, -1, C.byref(h)) == api.E_IO
    # unsupported cell type (m_interp_unstructured.f90:846-847)
    p = tmp_path / "hexa.binda"
    binda.write_binda(str(p), [("points", np.zeros((8, 3)), ""), ("cells", np.zeros((1, 4), dtype=np.int32), "hexahedron"),
                               ("cell_neighbors", np.zeros((1, 4), dtype=np.int32), "")])
    assert lib.iu_b200_grid_from_binda(str(p).encode(), 1.0, -1, C.byref(h)) == api.E_IO
    assert b"Unsupported cell type" in lib.iu_b200_last_error()


def test_create_rejects_bad_arguments(lib):
    h = C.c_void_p()
    assert lib.iu_b200_grid_create(None, C.byref(h)) == api.E_INVAL
    pts = np.zeros((4, 3))
    cells = np.array([[1, 2, 3, 9]], dtype=np.int32)       # point 9 does not exist
    nb = np.zeros((1, 4), dtype=np.int32)
    d = api._Desc(api.TETRA, 4, 1, pts.ctypes.data_as(api.c_dp), cells.ctypes.data_as(api.c_ip), nb.ctypes.data_as(api.c_ip),
                  0, None, 0, None, 0, None, -1)
    assert lib.iu_b200_grid_create(C.byref(d), C.byref(h)) == api.E_INVAL
    assert b"point index" in lib.iu_b200_last_error()


@pytest.mark.parametrize("kind", ["tetra", "quad", "triangle"])
def test_cxx_neighbor_builder_matches_numpy(lib, kind):
    if kind == "tetra":
        pts, cells = meshgen.kuhn_tet_mesh(9, jitter=0.1, seed=2, shuffle=True)
    elif kind == "quad":
        pts, cells = meshgen.quad_mesh(40, seed=2)
    else:
        pts, cells = meshgen.triangle_mesh(40, seed=2)
    ref = meshgen.cell_neighbors(cells, pts, kind)
    got = api.build_cell_neighbors(cells, pts, kind)
    assert (ref == got).all()


def test_cxx_neighbor_builder_merges_coincident_points(lib):
    # two triangles sharing an edge through DUPLICATED points (convert_to_binary.py:130-136)
    pts = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0]], dtype=np.float64)
    cells = np.array([[0, 1, 2], [3, 4, 5]], dtype=np.int32)
    got = api.build_cell_neighbors(cells, pts, "triangle")
    assert (got == meshgen.cell_neighbors(cells, pts, "triangle")).all()
    assert got[0, 1] == 1 and got[1, 2] == 0 and (got == -1).sum() == 4


def test_large_neighbor_build_parallel_path(lib):
    pts, cells = meshgen.kuhn_tet_mesh(26, jitter=0.1, seed=4)     # 105k tets -> threaded bucket sort
    got = api.build_cell_neighbors(cells, pts, "tetra")
    ref = meshgen.cell_neighbors(cells, pts, "tetra")
    assert (ref == got).all()
    assert (got < 0).sum() == 6 * 2 * 26 * 26                       # boundary faces of the cube


def test_header_is_plain_c99_and_c_driver_fails_loudly_without_a_gpu(tmp_path):
    """include/iu_b200.h must be usable from C (no C++-isms); the C driver of tests/c_driver links against the library
    and, on a machine without a GPU, stops with the library's 'no CPU fallback' error instead of computing anything."""
    import subprocess
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    probe = tmp_path / "probe.c"
    probe.write_text('#include "iu_b200.h"\nint main(void) { return iu_b200_version() < 0; }\n')
    subprocess.run(["gcc", "-std=c99", "-pedantic", "-Wall", "-Werror", "-fsyntax-only", "-I", os.path.join(root, "include"), str(probe)], check=True)
    import __graft_entry__ as ge
    exe = ge.build_c_driver()
    try:
        import torch
        has_gpu = torch.cuda.is_available()
    except ImportError:
        has_gpu = False
    out = subprocess.run([exe, os.path.join(root, "tests", "golden")], capture_output=True, text=True)
    if has_gpu:
        assert out.returncode == 0 and "PASSED" in out.stdout, out.stderr
    else:
        assert out.returncode != 0 and "no CPU fallback" in out.stderr
