#!/usr/bin/env python3
"""Regenerates tests/golden/*.binda from the reference's VTU fixtures (/root/reference/test_data).

Runs only in the build container (the reference tree does not exist on the GPU box); the
resulting .binda files are committed.  It plays the role of ``convert_to_binary.py`` (which
needs meshio, absent here): same entry order, dtypes and neighbour convention
(convert_to_binary.py:118-162, :202-224).  dtypes follow the VTU arrays (Float32 stays float32,
integers become int32).
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


def main():
    for name in ["triangle", "quad", "tetra", "tetra_cell", "tetra_two_cell"]:
        points, conn, ctype, pdata, cdata = read_vtu(os.path.join(REF, name + ".vtu"))
        neigh = meshgen.cell_neighbors(conn, points, ctype)
        out = os.path.join(HERE, name + ".binda")
        binda.write_grid_binda(out, points, conn, ctype, neigh, pdata, cdata)
        print(name, ctype, points.shape, points.dtype, conn.shape, list(pdata), "boundary faces:", int((neigh < 0).sum()))


if __name__ == "__main__":
    main()
