"""TEST INFRASTRUCTURE ONLY: Python restatement of the byte stream of iu_write_vtk
(/root/reference m_interp_unstructured.f90:930-985, written through m_vtk.f90:45-221).
Used by tests/ to check the library's VTU writer byte for byte; never imported by the product.

Pinned by the format itself (header strings and block order are literal in m_vtk.f90); the
reference's test_vtk.f90 only writes files and checks nothing, so there is no golden file."""
import struct

import numpy as np


def vtu_bytes(cell_type, points, cells1, cell_data=None, icell_data=None, point_data=None):
    """points (Np,3) f64, cells1 (Nc,ppc) 1-based int32, *_data: dict name -> column."""
    vtk_type = {1: 5, 2: 9, 3: 10}[cell_type]                      # :941-950
    pts = np.asarray(points, dtype=np.float64)
    cells1 = np.asarray(cells1, dtype=np.int32)
    n_points, n_cells, ppc = len(pts), len(cells1), cells1.shape[1]
    xml, raw = [], []
    state = {"indent": 0, "offset": 0}

    def line(t):
        xml.append(" " * state["indent"] + t + "\n")

    def dat(name, is_open):                                        # vtk_dat_xml, m_vtk.f90:148-162
        if is_open:
            line(f"<{name}>")
            state["indent"] += 2
        else:
            state["indent"] -= 2
            line(f"</{name}>")

    def block(a):                                                  # "write(sunit) n_bytes, data"
        b = a.tobytes()
        raw.append(struct.pack("<i", len(b)) + b)
        state["offset"] += 4 + len(b)

    xml.append('<?xml version="1.0"?>\n')                          # vtk_ini_xml, :45-64
    xml.append('<VTKFile type="UnstructuredGrid" version="0.1" byte_order="LittleEndian">\n')
    state["indent"] = 2
    dat("UnstructuredGrid", True)
    line(f'<Piece NumberOfPoints="{n_points}" NumberOfCells="{n_cells}">')   # vtk_unstr_geo_xml, :66-100
    state["indent"] += 2
    dat("Points", True)
    line(f'<DataArray type="Float32" NumberOfComponents="3" Name="Points" format="appended" offset="{state["offset"]}"/>')
    block(pts.astype(np.float32))
    dat("Points", False)
    dat("Cells", True)                                             # vtk_unstr_con_xml, :107-146
    line(f'<DataArray type="Int32" Name="connectivity" format="appended" offset="{state["offset"]}"/>')
    block((cells1 - 1).astype(np.int32))
    line(f'<DataArray type="Int32" Name="offsets" format="appended" offset="{state["offset"]}"/>')
    block((np.arange(1, n_cells + 1, dtype=np.int32) * ppc).astype(np.int32))
    line(f'<DataArray type="Int32" Name="types" format="appended" offset="{state["offset"]}"/>')
    block(np.full(n_cells, vtk_type, dtype=np.int32))
    dat("Cells", False)
    dat("CellData", True)
    for name, col in (cell_data or {}).items():                    # vtk_var_r8_xml, :164-180
        line(f'<DataArray type="Float64" Name="{name}" NumberOfComponents="1" format="appended" offset="{state["offset"]}"/>')
        block(np.asarray(col, dtype=np.float64))
    for name, col in (icell_data or {}).items():                   # vtk_var_i4_xml, :182-198
        line(f'<DataArray type="Int32" Name="{name}" NumberOfComponents="1" format="appended" offset="{state["offset"]}"/>')
        block(np.asarray(col, dtype=np.int32))
    dat("CellData", False)
    dat("PointData", True)
    for name, col in (point_data or {}).items():
        line(f'<DataArray type="Float64" Name="{name}" NumberOfComponents="1" format="appended" offset="{state["offset"]}"/>')
        block(np.asarray(col, dtype=np.float64))
    dat("PointData", False)
    dat("Piece", False)
    dat("UnstructuredGrid", False)
    line('<AppendedData encoding="raw">')                          # vtk_end_xml, :200-221
    head = "".join(xml).encode()
    xml.clear()
    xml.append("\n")
    line("</AppendedData>")
    dat("VTKFile", False)
    return head + b"_" + b"".join(raw) + "".join(xml).encode()


def parse_vtu(data):
    """Minimal reader of the stream above: returns {array name: numpy array} (for round-trip checks)."""
    import re
    i = data.index(b'<AppendedData encoding="raw">')
    start = data.index(b"_", i) + 1
    out = {}
    for m in re.finditer(rb'<DataArray type="(\w+)"(?: NumberOfComponents="\d+")? Name="([^"]+)"[^>]*offset="(\d+)"', data[:i]):
        dt = {b"Float32": np.float32, b"Float64": np.float64, b"Int32": np.int32}[m.group(1)]
        off = start + int(m.group(3))
        nb = struct.unpack("<i", data[off:off + 4])[0]
        out[m.group(2).decode()] = np.frombuffer(data[off + 4:off + 4 + nb], dtype=dt)
    return out
