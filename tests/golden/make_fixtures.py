#!/usr/bin/env python3
"""Regenerates tests/golden/*.binda from the reference's VTU fixtures (/root/reference/test_data)
by running the reference's own, UNMODIFIED converter /root/reference/convert_to_binary.py.

The converter needs ``meshio`` (absent on this image) only for ``meshio.read`` (:8, :185); a stub
module of that name (written to a temp dir on sys.path) serves the mesh from the VTU parser below
with meshio's attribute layout (points, cells[0].type/.data, point_data, cell_data).  BindaWriter
(:11-115) and get_cell_neighbors (:118-162) are therefore the reference's code, and the committed
.binda files are reference output.  Also writes neighbors_golden.json: sha256 of the reference's
get_cell_neighbors on a 118k-tet jittered Kuhn mesh with duplicated points (tests/test_n1_reference.py).

Runs only in the build container (the reference tree does not exist on the GPU box).
"""
import os
import re
import struct
import sys
import xml.etree.ElementTree as ET

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(HERE, "..", ".."))
from interpolate_unstructured_b200 import binda, meshgen  # noqa: E402

REF = "/root/reference/test_data"
VTK_DT = {"Float32": np.float32, "Float64": np.float64, "Int32": np.int32, "Int64": np.int64,
          "UInt8": np.uint8, "UInt64": np.uint64, "UInt32": np.uint32}
VTK_CELL = {5: ("triangle", 3), 9: ("quad", 4), 10: ("tetra", 4)}


def read_vtu(path):
    raw = open(path, "rb").read()
    appended = None
    m = re.search(rb"<AppendedData[^>]*>\s*_", raw)
    if m:
        end = raw.rindex(b"</AppendedData>")
        appended = raw[m.end():end]
        raw = raw[:m.end()] + raw[end:]
    root = ET.fromstring(raw.decode("ascii", errors="replace"))
    hdr = np.dtype(VTK_DT[root.attrib.get("header_type", "UInt32")])

    def arr(da):
        dt = VTK_DT[da.attrib["type"]]
        if da.attrib.get("format", "ascii") == "appended":
            off = int(da.attrib["offset"])
            nbytes = int(np.frombuffer(appended, dtype=hdr, count=1, offset=off)[0])
            return np.frombuffer(appended, dtype=dt, count=nbytes // np.dtype(dt).itemsize, offset=off + hdr.itemsize)
        text = da.text or ""
        return np.array(text.split(), dtype=dt)

    piece = root.find(".//Piece")
    npts = int(piece.attrib["NumberOfPoints"])
    points = arr(piece.find("Points/DataArray")).reshape(npts, 3)
    cells = {d.attrib["Name"]: arr(d) for d in piece.findall("Cells/DataArray")}
    ctype, ppc = VTK_CELL[int(cells["types"][0])]
    conn = cells["connectivity"].reshape(-1, ppc)
    pdata = {d.attrib["Name"]: arr(d) for d in piece.findall("PointData/DataArray")}
    cdata = {d.attrib["Name"]: arr(d) for d in piece.findall("CellData/DataArray")}
    return points, conn, ctype, pdata, cdata


MESHIO_STUB = '''"""Stand-in for meshio on an image without it: read() only, served by tests/golden/make_fixtures.read_vtu."""
import importlib.util, os
_spec = importlib.util.spec_from_file_location("_iu_make_fixtures", os.environ["IU_MAKE_FIXTURES"])
_mf = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(_mf)


class CellBlock:
    def __init__(self, type, data):
        self.type, self.data = type, data


class Mesh:
    pass


def read(filename):
    points, conn, ctype, pdata, cdata = _mf.read_vtu(filename)
    m = Mesh()
    m.points = points
    m.cells = [CellBlock(ctype, conn)]
    m.point_data = dict(pdata)
    m.cell_data = {k: [v] for k, v in cdata.items()}
    return m
'''

CONVERTER = "/root/reference/convert_to_binary.py"


def run_reference_converter(infile, out_basename):
    """Executes the unmodified reference converter on `infile` (meshio stubbed) and returns its globals
    (BindaWriter, get_cell_neighbors, ...)."""
    import runpy
    import tempfile
    with tempfile.TemporaryDirectory() as td:
        with open(os.path.join(td, "meshio.py"), "w") as f:
            f.write(MESHIO_STUB)
        os.environ["IU_MAKE_FIXTURES"] = os.path.abspath(__file__)
        argv, path = sys.argv, list(sys.path)
        sys.modules.pop("meshio", None)
        sys.path.insert(0, td)
        sys.argv = [CONVERTER, infile, "-output_basename", out_basename, "-force"]
        try:
            return runpy.run_path(CONVERTER, run_name="__main__")
        finally:
            sys.argv, sys.path[:] = argv, path
            sys.modules.pop("meshio", None)


def neighbor_case():
    """118k-tet jittered Kuhn mesh whose points are all duplicated (every second cell refers to the copies)."""
    pts, cells = meshgen.kuhn_tet_mesh(27, jitter=0.15, seed=99, shuffle=True)
    npts = len(pts)
    pts2 = np.concatenate([pts, pts])
    cells2 = cells.copy()
    cells2[1::2] += npts
    return pts2, cells2


def main():
    import hashlib
    import json
    ref = None
    for name in ["triangle", "quad", "tetra", "tetra_cell", "tetra_two_cell"]:
        ref = run_reference_converter(os.path.join(REF, name + ".vtu"), os.path.join(HERE, name))
        ents = binda.read_binda(os.path.join(HERE, name + ".binda"))
        print(name, [(e.name, e.metadata, e.data.dtype.name, e.data.shape) for e in ents])
    pts, cells = neighbor_case()
    nb = ref["get_cell_neighbors"](cells, pts, 3)
    digest = hashlib.sha256(np.ascontiguousarray(nb, dtype=np.int32).tobytes()).hexdigest()
    with open(os.path.join(HERE, "neighbors_golden.json"), "w") as f:
        json.dump({"case": "meshgen.kuhn_tet_mesh(27, jitter=0.15, seed=99, shuffle=True), points duplicated, odd cells on the copies",
                   "n_cells": int(len(cells)), "n_points": int(len(pts)), "boundary_faces": int((nb < 0).sum()),
                   "sha256_int32": digest, "source": "get_cell_neighbors of /root/reference/convert_to_binary.py:118-162"}, f, indent=1)
    print("neighbors golden:", len(cells), "tets", digest)


if __name__ == "__main__":
    main()
