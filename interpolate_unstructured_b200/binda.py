"""BINDA ("BInary N-dimensional DAta") container: writer and reader.

On-disk format is the one produced by the reference's ``convert_to_binary.py`` (BindaWriter,
convert_to_binary.py:11-115) and consumed by ``m_binda.f90:52-194``: 8-byte "BINDA   " tag,
int64 n_entries, int64 total_header_size, then per entry a 464-byte record (128 B name,
128 B numpy dtype string, 128 B metadata, int64 ndim, 8 x int64 shape, int64 absolute payload
offset), then the payloads in C order.  Integer arrays are stored as int32
(convert_to_binary.py:48-49).

The C++ reader inside libiu_b200.so (csrc/host_mesh.cpp: read_binda_mesh) is the product's reader; this module
is the Python-side writer used to produce grid files (the role convert_to_binary.py plays for
the reference) plus a small reader for round-trip checks.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass

import numpy as np

_HDR = struct.Struct("<8sqq")
_ENT = struct.Struct("<128s128s128sq8qq")


@dataclass
class BindaEntry:
    name: str
    metadata: str
    data: np.ndarray


def write_binda(filename, entries):
    """entries: iterable of (name, ndarray, metadata)."""
    recs = []
    blobs = []
    offset = 0
    for name, data, metadata in entries:
        data = np.ascontiguousarray(data)
        if len(name) > 128 or len(metadata) > 128:
            raise ValueError("name/metadata must be at most 128 characters")
        if data.ndim > 8:
            raise ValueError("Number of dimensions cannot exceed 8.")
        if np.issubdtype(data.dtype, np.integer):
            data = data.astype(np.int32)
        blob = data.tobytes()
        recs.append((name, str(data.dtype), metadata, data.ndim, tuple(data.shape), offset))
        blobs.append(blob)
        offset += len(blob)
    total_header = _HDR.size + _ENT.size * len(recs)
    with open(filename, "wb") as f:
        f.write(_HDR.pack(b"BINDA".ljust(8), len(recs), total_header))
        for name, dtype, meta, ndim, shape, off in recs:
            f.write(_ENT.pack(name.ljust(128).encode("ascii"), dtype.ljust(128).encode("ascii"),
                              meta.ljust(128).encode("ascii"), ndim,
                              *(shape + (0,) * (8 - len(shape))), off + total_header))
        for b in blobs:
            f.write(b)


def read_binda(filename):
    """Returns a list of BindaEntry in file order."""
    with open(filename, "rb") as f:
        buf = f.read()
    tag, n_entries, _ = _HDR.unpack_from(buf, 0)
    if tag.rstrip() != b"BINDA":
        raise ValueError("Wrong file format")
    out = []
    pos = _HDR.size
    for _ in range(n_entries):
        name, dtype, meta, ndim, *rest = _ENT.unpack_from(buf, pos)
        shape, off = tuple(rest[:8][:ndim]), rest[8]
        pos += _ENT.size
        dt = np.dtype(dtype.decode("ascii").strip())
        count = int(np.prod(shape)) if ndim else 1
        arr = np.frombuffer(buf, dtype=dt, count=count, offset=off).reshape(shape)
        out.append(BindaEntry(name.decode("ascii").strip(), meta.decode("ascii").strip(), arr))
    return out


def write_grid_binda(filename, points, cells, cell_type, neighbors, point_data=None, cell_data=None):
    """Write a mesh the way convert_to_binary.py:202-224 does: points, cells (metadata = meshio
    cell type), cell_neighbors, then point_data / cell_data / icell_data entries (variable name
    in metadata, commas removed).  ``cells``/``neighbors`` are 0-based, -1 = no neighbour."""
    ents = [("points", np.asarray(points), ""),
            ("cells", np.asarray(cells), cell_type),
            ("cell_neighbors", np.asarray(neighbors), "")]
    for k, v in (point_data or {}).items():
        ents.append(("point_data", np.asarray(v), k.replace(",", "")))
    for k, v in (cell_data or {}).items():
        v = np.asarray(v)
        ents.append(("icell_data" if np.issubdtype(v.dtype, np.integer) else "cell_data", v, k.replace(",", "")))
    write_binda(filename, ents)
