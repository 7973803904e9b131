"""Host-side mirror of the reference's public interface (m_interp_unstructured.f90:77-101) on top
of the C ABI of ``libiu_b200.so`` (include/iu_b200.h), bound with ctypes.

Names, argument meaning and error behaviour follow the Fortran module; every call is batched:
``r`` is ``(n, 3)`` (C order = Fortran ``r(3, n)``), ``i_cell`` is an int32 array that is updated in
place exactly like the reference's ``intent(inout)`` argument, result arrays are ``intent(inout)`` too
(rows of not-found queries stay untouched).  Where the reference would ``error stop`` this module
raises :class:`IUError`.

Host arrays are numpy; the ``*_dev`` variants take CUDA ``torch`` tensors (torch is only the
owner of device memory and streams here) and enqueue on the current torch stream.

There is no CPU fallback: if ``libiu_b200.so`` is missing or no GPU is usable, calls raise.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# IU_B200_LIB selects an alternative build of the same CUDA library (kernel tuning experiments)
LIB_PATH = os.environ.get("IU_B200_LIB") or os.path.join(_HERE, "lib", "libiu_b200.so")
_lib = None

TRIANGLE, QUAD, TETRA = 1, 2, 3
CELL_TYPES = {"triangle": TRIANGLE, "quad": QUAD, "tetra": TETRA}
E_INVAL, E_CUDA, E_IO, E_HINT, E_NOMEM = -1, -2, -3, -4, -5
CELL_WALK_LIMIT = -2
OPT_SORT_QUERIES, OPT_MAX_WALK_STEPS, OPT_HOST_CHUNK, OPT_DEV_BATCH, OPT_TIME_KERNELS, OPT_DYNAMIC_TILES = 1, 2, 3, 4, 5, 6
OPT_NN_VOXEL_DENSITY, OPT_OVERLAP_SORT, OPT_TILES_PER_WARP = 7, 8, 9
STAT_NAMES = ["kernel_launches", "walk_limit_hits", "queries", "workspace_bytes", "mesh_bytes", "tree_depth",
              "not_found", "query_kernel_ns", "query_kernel_launches", "sort_ns", "nn_voxels", "nn_voxels_tree", "nn_voxel_entries",
              "nn_tree_queries", "nn_voxel_build_us", "devices", "replicate_us", "bounds_errors", "h2d_skipped_bytes", "tree_build_us"]

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int32)
c_vp = C.c_void_p


class IUError(RuntimeError):
    """Raised where the reference would ``error stop``; ``code`` is the IU_B200_E* value."""

    def __init__(self, code, msg):
        super().__init__(f"iu_b200 error {code}: {msg}")
        self.code = code


class _Desc(C.Structure):
    _fields_ = [("cell_type", C.c_int32), ("n_points", C.c_int32), ("n_cells", C.c_int32),
                ("points", c_dp), ("cells", c_ip), ("neighbors", c_ip),
                ("n_point_data", C.c_int32), ("point_data", c_dp),
                ("n_cell_data", C.c_int32), ("cell_data", c_dp),
                ("n_icell_data", C.c_int32), ("icell_data", c_ip),
                ("device", C.c_int32)]


# name -> (restype, argtypes); every symbol of include/iu_b200.h
_SIGNATURES = {
    "iu_b200_version": (C.c_int, []),
    "iu_b200_last_error": (C.c_char_p, []),
    "iu_b200_grid_create": (C.c_int, [C.POINTER(_Desc), C.POINTER(c_vp)]),
    "iu_b200_grid_from_binda": (C.c_int, [C.c_char_p, C.c_double, C.c_int, C.POINTER(c_vp)]),
    "iu_b200_grid_replicate": (C.c_int, [c_vp, C.c_int, C.POINTER(C.c_int)]),
    "iu_b200_grid_create_multi": (C.c_int, [C.POINTER(_Desc), C.c_int, C.POINTER(C.c_int), C.POINTER(c_vp)]),
    "iu_b200_grid_from_binda_multi": (C.c_int, [C.c_char_p, C.c_double, C.c_int, C.POINTER(C.c_int), C.POINTER(c_vp)]),
    "iu_b200_device_count": (C.c_int, [c_vp]),
    "iu_b200_grid_replica": (c_vp, [c_vp, C.c_int]),
    "iu_b200_grid_destroy": (C.c_int, [c_vp]),
    "iu_b200_grid_info": (C.c_int, [c_vp, c_ip]),
    "iu_b200_grid_bounds": (C.c_int, [c_vp, c_dp, c_dp]),
    "iu_b200_get_point_data_index": (C.c_int, [c_vp, C.c_char_p]),
    "iu_b200_get_cell_data_index": (C.c_int, [c_vp, C.c_char_p]),
    "iu_b200_get_icell_data_index": (C.c_int, [c_vp, C.c_char_p]),
    "iu_b200_set_data_name": (C.c_int, [c_vp, C.c_int, C.c_int, C.c_char_p]),
    "iu_b200_point_data_name": (C.c_char_p, [c_vp, C.c_int]),
    "iu_b200_cell_data_name": (C.c_char_p, [c_vp, C.c_int]),
    "iu_b200_icell_data_name": (C.c_char_p, [c_vp, C.c_int]),
    "iu_b200_add_point_data": (C.c_int, [c_vp, C.c_char_p, C.POINTER(C.c_int)]),
    "iu_b200_add_cell_data": (C.c_int, [c_vp, C.c_char_p, C.POINTER(C.c_int)]),
    "iu_b200_add_icell_data": (C.c_int, [c_vp, C.c_char_p, C.POINTER(C.c_int)]),
    "iu_b200_update_point_data": (C.c_int, [c_vp, C.c_int, C.c_int, c_dp]),
    "iu_b200_update_cell_data": (C.c_int, [c_vp, C.c_int, C.c_int, c_dp]),
    "iu_b200_update_icell_data": (C.c_int, [c_vp, C.c_int, C.c_int, c_ip]),
    "iu_b200_download_point_data": (C.c_int, [c_vp, C.c_int, C.c_int, c_dp]),
    "iu_b200_download_cell_data": (C.c_int, [c_vp, C.c_int, C.c_int, c_dp]),
    "iu_b200_download_icell_data": (C.c_int, [c_vp, C.c_int, C.c_int, c_ip]),
    "iu_b200_set_option": (C.c_int, [c_vp, C.c_int, C.c_int64]),
    "iu_b200_get_stats": (C.c_int, [c_vp, C.POINTER(C.c_int64)]),
    "iu_b200_reset_stats": (C.c_int, [c_vp]),
    "iu_b200_find_nearby_cell_batch": (C.c_int, [c_vp, C.c_int64, c_vp, c_vp]),
    "iu_b200_get_cell_batch": (C.c_int, [c_vp, C.c_int64, c_vp, c_vp]),
    "iu_b200_interpolate_at_batch": (C.c_int, [c_vp, C.c_int64, c_vp, C.c_int, c_ip, c_vp, c_vp]),
    "iu_b200_interpolate_scalar_at_batch": (C.c_int, [c_vp, C.c_int64, c_vp, C.c_int, c_vp, c_vp]),
    "iu_b200_interpolate_at_icell_batch": (C.c_int, [c_vp, C.c_int64, c_vp, C.c_int, c_ip, c_vp, c_vp]),
    "iu_b200_interpolate_and_cell_scalar_at_batch": (C.c_int, [c_vp, C.c_int64, c_vp, C.c_int, c_ip, C.c_int, c_vp, c_vp, c_vp]),
    "iu_b200_get_cell_scalar_at_batch": (C.c_int, [c_vp, C.c_int64, c_vp, C.c_int, c_vp, c_vp]),
    "iu_b200_get_icell_scalar_at_batch": (C.c_int, [c_vp, C.c_int64, c_vp, C.c_int, c_vp, c_vp]),
    "iu_b200_get_cell_through_neighbors_batch": (C.c_int, [c_vp, C.c_int64, C.c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, C.c_int]),
    "iu_b200_find_nearby_cell_batch_dev": (C.c_int, [c_vp, C.c_int64, c_vp, c_vp, c_vp]),
    "iu_b200_get_cell_batch_dev": (C.c_int, [c_vp, C.c_int64, c_vp, c_vp, c_vp]),
    "iu_b200_interpolate_at_batch_dev": (C.c_int, [c_vp, C.c_int64, c_vp, C.c_int, c_ip, c_vp, c_vp, c_vp]),
    "iu_b200_interpolate_at_batch_dev_noalloc": (C.c_int, [c_vp, C.c_int64, c_vp, C.c_int, c_ip, c_vp, c_vp, c_vp]),
    "iu_b200_interpolate_at_icell_batch_dev": (C.c_int, [c_vp, C.c_int64, c_vp, C.c_int, c_ip, c_vp, c_vp, c_vp]),
    "iu_b200_interpolate_and_cell_scalar_at_batch_dev": (C.c_int, [c_vp, C.c_int64, c_vp, C.c_int, c_ip, C.c_int, c_vp, c_vp, c_vp, c_vp]),
    "iu_b200_get_cell_scalar_at_batch_dev": (C.c_int, [c_vp, C.c_int64, c_vp, C.c_int, c_vp, c_vp, c_vp]),
    "iu_b200_get_icell_scalar_at_batch_dev": (C.c_int, [c_vp, C.c_int64, c_vp, C.c_int, c_vp, c_vp, c_vp]),
    "iu_b200_get_cell_through_neighbors_batch_dev": (C.c_int, [c_vp, C.c_int64, C.c_int, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, c_vp, C.c_int, c_vp]),
    "iu_b200_integrate_along_field_batch": (C.c_int, [c_vp, C.c_int64, C.c_int, C.c_int, C.c_int, c_ip, C.c_double, C.c_double, C.c_int,
                                                    C.c_double, C.c_double, C.c_int, c_vp, c_vp, c_vp, C.c_int, C.c_int, C.c_int, c_vp]),
    "iu_b200_integrate_along_field_batch_dev": (C.c_int, [c_vp, C.c_int64, C.c_int, C.c_int, C.c_int, c_ip, C.c_double, C.c_double, C.c_int,
                                                        C.c_double, C.c_double, C.c_int, c_vp, c_vp, c_vp, C.c_int, C.c_int, C.c_int, c_vp, c_vp]),
    "iu_b200_check_dev": (C.c_int, [c_vp, c_vp]),
    "iu_b200_build_cell_neighbors": (C.c_int, [C.c_int, C.c_int32, C.c_int32, c_dp, c_ip, c_ip]),
    "iu_b200_export_geometry": (C.c_int, [c_vp, c_dp, c_dp, c_dp, c_dp]),
    "iu_b200_export_mesh": (C.c_int, [c_vp, c_dp, c_ip]),
    "iu_b200_write_vtk": (C.c_int, [c_vp, C.c_char_p]),
    "iu_b200_write_vtk_arrays": (C.c_int, [C.c_char_p, C.c_int, C.c_int32, C.c_int32, c_dp, c_ip, C.c_int, c_dp, C.POINTER(C.c_char_p),
                                         C.c_int, c_ip, C.POINTER(C.c_char_p), C.c_int, c_dp, C.POINTER(C.c_char_p)]),
    "iu_b200_generate_queries_dev": (C.c_int, [c_vp, C.c_int64, C.c_int64, C.c_uint64, c_dp, c_dp, c_vp, c_vp]),
    "iu_b200_host_alloc": (c_vp, [C.c_size_t]),
    "iu_b200_host_free": (None, [c_vp]),
}


def build_library(force=False):
    """Compiles libiu_b200.so in-tree with nvcc for sm_100a (csrc/Makefile)."""
    src = os.path.join(_HERE, "csrc")
    if force:
        subprocess.run(["make", "-C", src, "clean"], check=True, capture_output=True)
    p = subprocess.run(["make", "-C", src], capture_output=True, text=True)
    if p.returncode != 0:
        raise RuntimeError("building libiu_b200.so failed:\n" + p.stdout[-4000:] + p.stderr[-4000:])
    return LIB_PATH


def lib():
    """The loaded C ABI.  Fails loudly when the CUDA library has not been built."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise IUError(E_CUDA, f"{LIB_PATH} is missing: run `python -c 'import __graft_entry__ as g; g.build()'` "
                                  "(there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(L, name)
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


def _check(rc):
    if rc != 0:
        raise IUError(rc, lib().iu_b200_last_error().decode(errors="replace"))


def _f64(a, shape_tail=None):
    a = np.ascontiguousarray(a, dtype=np.float64)
    if shape_tail is not None and a.shape[1:] != shape_tail:
        raise ValueError(f"expected shape (n,{','.join(map(str, shape_tail))}), got {a.shape}")
    return a


def _inout(a, dtype, shape, name):
    """intent(inout) arrays must be written in place: insist on the right dtype and layout."""
    if not isinstance(a, np.ndarray) or a.dtype != dtype or not a.flags.c_contiguous or a.shape != shape:
        raise ValueError(f"{name} must be a C-contiguous {np.dtype(dtype).name} numpy array of shape {shape}")
    return a


def _p(a):
    return a.ctypes.data_as(c_vp)


class _PinnedBlock:
    """Owner of one iu_b200_host_alloc block, exported to numpy through __array_interface__."""

    def __init__(self, nbytes):
        self.nbytes = max(int(nbytes), 1)
        self.ptr = lib().iu_b200_host_alloc(self.nbytes)
        if not self.ptr:
            raise IUError(E_NOMEM, "iu_b200_host_alloc failed")
        self.__array_interface__ = {"shape": (self.nbytes,), "typestr": "|u1", "data": (self.ptr, False), "version": 3}

    def __del__(self):
        try:
            if self.ptr:
                lib().iu_b200_host_free(self.ptr)
                self.ptr = None
        except Exception:
            pass


def pinned_empty(shape, dtype):
    """numpy array backed by page-locked memory (iu_b200_host_alloc); keeps the H2D/D2H legs of the
    host API at full PCIe speed.  The block is released when the last view of it is collected."""
    dtype = np.dtype(dtype)
    shape = (shape,) if np.isscalar(shape) else tuple(shape)
    count = int(np.prod(shape, dtype=np.int64))
    block = _PinnedBlock(count * dtype.itemsize)
    raw = np.asarray(block)            # base -> block
    return raw[:count * dtype.itemsize].view(dtype).reshape(shape)


def build_cell_neighbors(cells0, points, cell_type):
    """Face-neighbour table with the conventions of convert_to_binary.py:118-162 (0-based, -1 = none),
    computed by the sort-based C++ builder (no GPU needed)."""
    ct = CELL_TYPES.get(cell_type, cell_type)
    points = _f64(points, (3,))
    cells0 = np.ascontiguousarray(cells0, dtype=np.int32)
    out = np.empty_like(cells0)
    _check(lib().iu_b200_build_cell_neighbors(ct, len(points), len(cells0), points.ctypes.data_as(c_dp),
                                              cells0.ctypes.data_as(c_ip), out.ctypes.data_as(c_ip)))
    return out


def write_vtk_arrays(filename, cell_type, points, cells1, cell_data=None, icell_data=None, point_data=None):
    """The library's VTU writer on host arrays (iu_write_vtk byte stream, m_vtk.f90:45-221); no GPU needed.
    cells1 (Nc,ppc) 1-based; *_data = dict name -> column."""
    ct = {"triangle": 1, "quad": 2, "tetra": 3}.get(cell_type, cell_type)
    pts = np.ascontiguousarray(points, dtype=np.float64)
    cl = np.ascontiguousarray(cells1, dtype=np.int32)

    def pack(d, dt):
        d = d or {}
        names = (C.c_char_p * max(1, len(d)))(*[k.encode() for k in d])
        a = np.ascontiguousarray(np.stack([np.asarray(v, dtype=dt) for v in d.values()]) if d else np.zeros((0,), dtype=dt))
        return len(d), a, names

    ncd, cd, cdn = pack(cell_data, np.float64)
    nid, icd, icn = pack(icell_data, np.int32)
    npd, pd, pdn = pack(point_data, np.float64)
    _check(lib().iu_b200_write_vtk_arrays(os.fsencode(filename), ct, len(pts), len(cl), pts.ctypes.data_as(c_dp), cl.ctypes.data_as(c_ip),
                                          ncd, cd.ctypes.data_as(c_dp), cdn, nid, icd.ctypes.data_as(c_ip), icn,
                                          npd, pd.ctypes.data_as(c_dp), pdn))


def _dev_ptr(t, dtype_name, n_elems, name):
    """Pointer of a CUDA torch tensor after dtype / layout / size checks."""
    import torch
    want = {"float64": torch.float64, "int32": torch.int32}[dtype_name]
    if not t.is_cuda or t.dtype != want or not t.is_contiguous() or t.numel() != n_elems:
        raise ValueError(f"{name} must be a contiguous CUDA {dtype_name} tensor with {n_elems} elements")
    return C.c_void_p(t.data_ptr())


def _cur_stream(device):
    import torch
    return C.c_void_p(torch.cuda.current_stream(device).cuda_stream)


class Grid:
    """Device-resident ``type(iu_grid_t)`` (m_interp_unstructured.f90:24-59).

    Construct with :meth:`iu_read_grid` (a ``.binda`` file, :820-927) or :meth:`from_arrays` (the
    in-memory state iu_read_grid has before deriving geometry).  All cell / variable indices are
    1-based as in the reference."""

    def __init__(self, handle):
        self._h = C.c_void_p(handle)
        info = (C.c_int32 * 8)()
        _check(lib().iu_b200_grid_info(self._h, info))
        (self.cell_type, self.n_points, self.n_cells, self.n_points_per_cell, self.n_point_data, self.n_cell_data,
         self.n_icell_data, self.device) = (int(x) for x in info)
        rmin, rmax = np.zeros(3), np.zeros(3)
        _check(lib().iu_b200_grid_bounds(self._h, rmin.ctypes.data_as(c_dp), rmax.ctypes.data_as(c_dp)))
        self.rmin, self.rmax = rmin, rmax

    # ---- construction ------------------------------------------------------------------------
    @classmethod
    def iu_read_grid(cls, filename, coord_scale_factor=1.0, device=-1, devices=None):
        """iu_read_grid(filename, ug, coord_scale_factor) for ``*.binda`` files (:820-927).  ``devices``: list of GPUs that
        each get a replica of the mesh (the first one builds it)."""
        h = c_vp()
        if devices is not None:
            arr = (C.c_int * len(devices))(*[int(d) for d in devices])
            _check(lib().iu_b200_grid_from_binda_multi(os.fsencode(str(filename)), float(coord_scale_factor), len(devices), arr, C.byref(h)))
        else:
            _check(lib().iu_b200_grid_from_binda(os.fsencode(str(filename)), float(coord_scale_factor), int(device), C.byref(h)))
        return cls(h.value)

    @classmethod
    def from_arrays(cls, cell_type, points, cells0, neighbors0, point_data=None, cell_data=None, icell_data=None,
                    point_names=None, cell_names=None, icell_names=None, device=-1, devices=None):
        """cells0 / neighbors0 are 0-based as on disk (+1 applied here like :871-872); data tables
        are (n_rows, n_columns).  ``devices``: GPUs that each get a replica (the first one builds the mesh)."""
        if devices is not None:
            device = devices[0]
        ct = CELL_TYPES.get(cell_type, cell_type)
        points = _f64(points, (3,))
        def plus_one(a):          # 0-based -> 1-based in one pass (int32 input stays int32: index tables of 50M cells are 800 MB)
            a = np.asarray(a)
            if a.dtype == np.int32 and (a.size == 0 or a.max() < np.iinfo(np.int32).max):
                return np.ascontiguousarray(np.add(a, 1, dtype=np.int32))
            return np.ascontiguousarray(np.asarray(a, dtype=np.int64) + 1, dtype=np.int32)
        cells, neigh = plus_one(cells0), plus_one(neighbors0)
        if cells.shape != neigh.shape or cells.ndim != 2:
            raise ValueError("cells and neighbors must both be (n_cells, points_per_cell)")

        def colmajor(a, dt, rows):
            if a is None:
                return 0, None, None
            a = np.asarray(a, dtype=dt)
            if a.ndim == 1:
                a = a[:, None]
            if a.shape[0] != rows:
                raise ValueError("data table has the wrong number of rows")
            t = np.ascontiguousarray(a.T)
            return a.shape[1], t, t.ctypes.data_as(c_dp if dt == np.float64 else c_ip)

        npd, pd_keep, pd_p = colmajor(point_data, np.float64, len(points))
        ncd, cd_keep, cd_p = colmajor(cell_data, np.float64, len(cells))
        nicd, icd_keep, icd_p = colmajor(icell_data, np.int32, len(cells))
        d = _Desc(ct, len(points), len(cells), points.ctypes.data_as(c_dp), cells.ctypes.data_as(c_ip),
                  neigh.ctypes.data_as(c_ip), npd, pd_p, ncd, cd_p, nicd, icd_p, int(device))
        h = c_vp()
        _check(lib().iu_b200_grid_create(C.byref(d), C.byref(h)))
        g = cls(h.value)
        for kind, names, n in ((0, point_names, npd), (1, cell_names, ncd), (2, icell_names, nicd)):
            if names is not None:
                if len(names) != n:
                    raise ValueError("one name per data column expected")
                for i, nm in enumerate(names):
                    _check(lib().iu_b200_set_data_name(g._h, kind, i + 1, str(nm).encode()))
        if devices is not None and len(devices) > 1:
            g.replicate(devices[1:])
        return g

    # ---- one handle, several GPUs ---------------------------------------------------------------
    def replicate(self, devices=None):
        """Copies the built mesh device to device onto ``devices`` (None = every visible GPU).  Afterwards the host-buffer
        batch calls split their queries into one contiguous slab per GPU and run the slabs concurrently."""
        if devices is None:
            _check(lib().iu_b200_grid_replicate(self._h, 0, None))
        else:
            arr = (C.c_int * len(devices))(*[int(d) for d in devices])
            _check(lib().iu_b200_grid_replicate(self._h, len(devices), arr))
        return self

    @property
    def device_count(self):
        return int(lib().iu_b200_device_count(self._h))

    def replica(self, i):
        """Handle of the replica on the i-th GPU of this grid (0 = the grid itself), for the ``*_dev`` calls with tensors
        of that GPU.  Owned by this grid: do not close it."""
        h = lib().iu_b200_grid_replica(self._h, int(i))
        if not h:
            raise IndexError("replica index outside 0..device_count-1")
        if i == 0:
            return self
        g = Grid(h)
        g._borrowed = True
        return g

    def close(self):
        if getattr(self, "_borrowed", False):
            self._h = C.c_void_p(None)
            return
        if getattr(self, "_h", None) is not None and self._h.value:
            lib().iu_b200_grid_destroy(self._h)
            self._h = C.c_void_p(None)

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    # ---- bookkeeping (:106-247) --------------------------------------------------------------
    def iu_get_point_data_index(self, name):
        return lib().iu_b200_get_point_data_index(self._h, name.encode())

    def iu_get_cell_data_index(self, name):
        return lib().iu_b200_get_cell_data_index(self._h, name.encode())

    def iu_get_icell_data_index(self, name):
        return lib().iu_b200_get_icell_data_index(self._h, name.encode())

    def _add(self, fn, name, attr):
        iv = C.c_int()
        _check(fn(self._h, name.encode(), C.byref(iv)))
        setattr(self, attr, iv.value)
        return iv.value

    def iu_add_point_data(self, name):
        return self._add(lib().iu_b200_add_point_data, name, "n_point_data")

    def iu_add_cell_data(self, name):
        return self._add(lib().iu_b200_add_cell_data, name, "n_cell_data")

    def iu_add_icell_data(self, name):
        return self._add(lib().iu_b200_add_icell_data, name, "n_icell_data")

    def set_point_data(self, i_var, values):
        """``ug%point_data(:, i_var:i_var+k-1) = values`` followed by the device refresh."""
        v = np.asarray(values, dtype=np.float64)
        v = v[:, None] if v.ndim == 1 else v
        t = np.ascontiguousarray(v.T)
        _check(lib().iu_b200_update_point_data(self._h, int(i_var), t.shape[0], t.ctypes.data_as(c_dp)))

    def set_cell_data(self, i_var, values):
        v = np.asarray(values, dtype=np.float64)
        v = v[:, None] if v.ndim == 1 else v
        t = np.ascontiguousarray(v.T)
        _check(lib().iu_b200_update_cell_data(self._h, int(i_var), t.shape[0], t.ctypes.data_as(c_dp)))

    def set_icell_data(self, i_var, values):
        v = np.asarray(values, dtype=np.int32)
        v = v[:, None] if v.ndim == 1 else v
        t = np.ascontiguousarray(v.T)
        _check(lib().iu_b200_update_icell_data(self._h, int(i_var), t.shape[0], t.ctypes.data_as(c_ip)))

    def get_point_data(self, i_var, n_cols=1):
        """``ug%point_data(:, i_var:i_var+n_cols-1)`` read back from the device, shape (n_points, n_cols)."""
        t = np.empty((int(n_cols), self.n_points))
        _check(lib().iu_b200_download_point_data(self._h, int(i_var), int(n_cols), t.ctypes.data_as(c_dp)))
        return np.ascontiguousarray(t.T)

    def get_cell_data(self, i_var, n_cols=1):
        t = np.empty((int(n_cols), self.n_cells))
        _check(lib().iu_b200_download_cell_data(self._h, int(i_var), int(n_cols), t.ctypes.data_as(c_dp)))
        return np.ascontiguousarray(t.T)

    def get_icell_data(self, i_var, n_cols=1):
        t = np.empty((int(n_cols), self.n_cells), dtype=np.int32)
        _check(lib().iu_b200_download_icell_data(self._h, int(i_var), int(n_cols), t.ctypes.data_as(c_ip)))
        return np.ascontiguousarray(t.T)

    def set_option(self, option, value):
        _check(lib().iu_b200_set_option(self._h, int(option), int(value)))

    def stats(self):
        s = (C.c_int64 * len(STAT_NAMES))()
        _check(lib().iu_b200_get_stats(self._h, s))
        return dict(zip(STAT_NAMES, (int(x) for x in s)))

    def reset_stats(self):
        _check(lib().iu_b200_reset_stats(self._h))

    def export_geometry(self):
        """cell_points (Nc,ppc,3), cell_face_normals (Nc,ppc,3), cell_volume (Nc), centres (Nc,3) as
        the device derived them, in original cell order."""
        nc, ppc = self.n_cells, self.n_points_per_cell
        cp, fn, vol, ce = np.empty((nc, ppc, 3)), np.empty((nc, ppc, 3)), np.empty(nc), np.empty((nc, 3))
        _check(lib().iu_b200_export_geometry(self._h, cp.ctypes.data_as(c_dp), fn.ctypes.data_as(c_dp),
                                             vol.ctypes.data_as(c_dp), ce.ctypes.data_as(c_dp)))
        return cp, fn, vol, ce

    def _name(self, fn, i_var, prefix):
        b = fn(self._h, int(i_var))
        if b is None:
            raise IndexError(i_var)
        return b.decode() or f"{prefix}{i_var}"      # unnamed columns: the name iu_write_vtk gives them

    def point_data_name(self, i_var):
        return self._name(lib().iu_b200_point_data_name, i_var, "point_data_")

    def cell_data_name(self, i_var):
        return self._name(lib().iu_b200_cell_data_name, i_var, "cell_data_")

    def icell_data_name(self, i_var):
        return self._name(lib().iu_b200_icell_data_name, i_var, "icell_data_")

    def export_mesh(self):
        """points (Np,3) and cells (Nc,ppc) 1-based, rebuilt from the device records in original order."""
        pts = np.empty((self.n_points, 3))
        cells = np.empty((self.n_cells, self.n_points_per_cell), dtype=np.int32)
        _check(lib().iu_b200_export_mesh(self._h, pts.ctypes.data_as(c_dp), cells.ctypes.data_as(c_ip)))
        return pts, cells

    def iu_write_vtk(self, filename):
        """iu_write_vtk(ug, filename) (m_interp_unstructured.f90:930-985): the grid and its current data as .vtu."""
        _check(lib().iu_b200_write_vtk(self._h, os.fsencode(filename)))

    # ---- the hot path, host arrays -----------------------------------------------------------
    def _prep(self, r, i_cell):
        r = _f64(r, (3,))
        n = len(r)
        if i_cell is None:
            i_cell = np.zeros(n, dtype=np.int32)
        _inout(i_cell, np.int32, (n,), "i_cell")
        return r, n, i_cell

    def find_nearby_cell_kdtree(self, r):
        """find_nearby_cell_kdtree (:272-288): 1-based cell with the nearest centre."""
        r = _f64(r, (3,))
        idx = np.zeros(len(r), dtype=np.int32)
        _check(lib().iu_b200_find_nearby_cell_batch(self._h, len(r), _p(r), _p(idx)))
        return idx

    def iu_get_cell(self, r, i_cell=None):
        """iu_get_cell (:414-434); ``i_cell`` is updated in place and returned."""
        r, n, i_cell = self._prep(r, i_cell)
        _check(lib().iu_b200_get_cell_batch(self._h, n, _p(r), _p(i_cell)))
        return i_cell

    def iu_interpolate_at(self, r, i_vars, var_at_r=None, i_cell=None):
        """iu_interpolate_at (:480-495) -> (var_at_r (n, n_vars), i_cell)."""
        r, n, i_cell = self._prep(r, i_cell)
        iv = np.ascontiguousarray(i_vars, dtype=np.int32).reshape(-1)
        if var_at_r is None:
            var_at_r = np.zeros((n, len(iv)))
        _inout(var_at_r, np.float64, (n, len(iv)), "var_at_r")
        _check(lib().iu_b200_interpolate_at_batch(self._h, n, _p(r), len(iv), iv.ctypes.data_as(c_ip), _p(var_at_r), _p(i_cell)))
        return var_at_r, i_cell

    def iu_interpolate_scalar_at(self, r, i_var, var_at_r=None, i_cell=None):
        """iu_interpolate_scalar_at (:465-477) -> (var_at_r (n,), i_cell)."""
        r, n, i_cell = self._prep(r, i_cell)
        if var_at_r is None:
            var_at_r = np.zeros(n)
        _inout(var_at_r, np.float64, (n,), "var_at_r")
        _check(lib().iu_b200_interpolate_scalar_at_batch(self._h, n, _p(r), int(i_var), _p(var_at_r), _p(i_cell)))
        return var_at_r, i_cell

    def iu_interpolate_at_icell(self, r, i_vars, i_cell, var_at_r=None):
        """iu_interpolate_at_icell (:498-527); entries of ``i_cell`` below 1 are skipped."""
        r = _f64(r, (3,))
        n = len(r)
        ic = np.ascontiguousarray(i_cell, dtype=np.int32)
        iv = np.ascontiguousarray(i_vars, dtype=np.int32).reshape(-1)
        if var_at_r is None:
            var_at_r = np.zeros((n, len(iv)))
        _inout(var_at_r, np.float64, (n, len(iv)), "var_at_r")
        _check(lib().iu_b200_interpolate_at_icell_batch(self._h, n, _p(r), len(iv), iv.ctypes.data_as(c_ip), _p(var_at_r), _p(ic)))
        return var_at_r

    def iu_interpolate_and_cell_scalar_at(self, r, i_vars, i_cell_var, var_at_r=None, cell_res=None, i_cell=None):
        """iu_interpolate_at (:480-495) then iu_get_cell_scalar_at (:437-448) with the found cell as i_guess, in one pass
        -> (var_at_r (n, n_vars), cell_res (n,), i_cell)."""
        r, n, i_cell = self._prep(r, i_cell)
        iv = np.ascontiguousarray(i_vars, dtype=np.int32).reshape(-1)
        if var_at_r is None:
            var_at_r = np.zeros((n, len(iv)))
        if cell_res is None:
            cell_res = np.zeros(n)
        _inout(var_at_r, np.float64, (n, len(iv)), "var_at_r")
        _inout(cell_res, np.float64, (n,), "cell_res")
        _check(lib().iu_b200_interpolate_and_cell_scalar_at_batch(self._h, n, _p(r), len(iv), iv.ctypes.data_as(c_ip), int(i_cell_var),
                                                                  _p(var_at_r), _p(cell_res), _p(i_cell)))
        return var_at_r, cell_res, i_cell

    def iu_get_cell_scalar_at(self, r, i_var, i_guess=None, res=None):
        """iu_get_cell_scalar_at (:437-448) -> (res, i_guess)."""
        r, n, i_guess = self._prep(r, i_guess)
        if res is None:
            res = np.zeros(n)
        _inout(res, np.float64, (n,), "res")
        _check(lib().iu_b200_get_cell_scalar_at_batch(self._h, n, _p(r), int(i_var), _p(i_guess), _p(res)))
        return res, i_guess

    def iu_get_icell_scalar_at(self, r, i_var, i_guess=None, res=None):
        """iu_get_icell_scalar_at (:451-462) -> (res, i_guess)."""
        r, n, i_guess = self._prep(r, i_guess)
        if res is None:
            res = np.zeros(n, dtype=np.int32)
        _inout(res, np.int32, (n,), "res")
        _check(lib().iu_b200_get_icell_scalar_at_batch(self._h, n, _p(r), int(i_var), _p(i_guess), _p(res)))
        return res, i_guess

    def iu_get_cell_through_neighbors(self, max_steps, r0, r1, ic0, i_icell_mask=0, r_p=None, n_steps=None):
        """iu_get_cell_through_neighbors (:667-725) -> (ic1, r_p, n_steps, status)."""
        r0, r1 = _f64(r0, (3,)), _f64(r1, (3,))
        n = len(r0)
        ic0 = np.ascontiguousarray(ic0, dtype=np.int32)
        ic1 = np.zeros(n, dtype=np.int32)
        status = np.zeros(n, dtype=np.int32)
        if r_p is None:
            r_p = np.full((n, 3), np.nan)
        if n_steps is None:
            n_steps = np.full(n, -1, dtype=np.int32)
        _inout(r_p, np.float64, (n, 3), "r_p")
        _inout(n_steps, np.int32, (n,), "n_steps")
        _check(lib().iu_b200_get_cell_through_neighbors_batch(self._h, n, int(min(max_steps, 2**31 - 2)), _p(r0), _p(r1), _p(ic0),
                                                              _p(ic1), _p(r_p), _p(n_steps), _p(status), int(i_icell_mask)))
        return ic1, r_p, n_steps, status

    def iu_integrate_along_field(self, y, i_field, min_dx, max_dx, rtol, atol, reverse=False, nvar=0, integrand=0,
                                 y_field=None, axisymmetric=False, i_icell_mask=0, mask_value=0, boundary_material=None):
        """iu_integrate_along_field (:989-1217) for n field lines.  ``y`` is (n, max_steps, ndim+nvar)
        (= Fortran y(ndim+nvar, max_steps, n)) with y[:, 0, :] the start positions and initial values,
        updated in place; the sub_int callback is replaced by ``integrand`` (0 none, 1 dy = 1,
        2 dy = |field|).  Returns (y, y_field, n_steps, boundary_material)."""
        iv = np.ascontiguousarray(i_field, dtype=np.int32).reshape(-1)
        ndim = len(iv)
        if not isinstance(y, np.ndarray) or y.dtype != np.float64 or not y.flags.c_contiguous or y.ndim != 3 or y.shape[2] != ndim + nvar:
            raise ValueError("y must be a C-contiguous float64 array of shape (n, max_steps, ndim + nvar)")
        n, max_steps = y.shape[0], y.shape[1]
        if y_field is None:
            y_field = np.full((n, max_steps, ndim), np.nan)
        _inout(y_field, np.float64, (n, max_steps, ndim), "y_field")
        if boundary_material is None:
            boundary_material = np.full(n, -99, dtype=np.int32)
        _inout(boundary_material, np.int32, (n,), "boundary_material")
        n_steps = np.zeros(n, dtype=np.int32)
        _check(lib().iu_b200_integrate_along_field_batch(
            self._h, n, ndim, int(nvar), int(integrand), iv.ctypes.data_as(c_ip), float(min_dx), float(max_dx), int(max_steps),
            float(rtol), float(atol), int(bool(reverse)), _p(y), _p(y_field), _p(n_steps), int(bool(axisymmetric)),
            int(i_icell_mask), int(mask_value), _p(boundary_material)))
        return y, y_field, n_steps, boundary_material

    # ---- the hot path, device tensors (no copies, asynchronous on the current torch stream) ----
    def generate_queries_dev(self, r, first, seed, lo, hi):
        lo, hi = _f64(lo), _f64(hi)
        n = r.numel() // 3
        _check(lib().iu_b200_generate_queries_dev(self._h, n, int(first), int(seed), lo.ctypes.data_as(c_dp), hi.ctypes.data_as(c_dp),
                                                  _dev_ptr(r, "float64", 3 * n, "r"), _cur_stream(self.device)))

    def find_nearby_cell_kdtree_dev(self, r, idx):
        n = idx.numel()
        _check(lib().iu_b200_find_nearby_cell_batch_dev(self._h, n, _dev_ptr(r, "float64", 3 * n, "r"),
                                                        _dev_ptr(idx, "int32", n, "idx"), _cur_stream(self.device)))

    def iu_get_cell_dev(self, r, i_cell):
        n = i_cell.numel()
        _check(lib().iu_b200_get_cell_batch_dev(self._h, n, _dev_ptr(r, "float64", 3 * n, "r"),
                                                _dev_ptr(i_cell, "int32", n, "i_cell"), _cur_stream(self.device)))

    def iu_interpolate_at_dev(self, r, i_vars, var_at_r, i_cell):
        n = i_cell.numel()
        iv = np.ascontiguousarray(i_vars, dtype=np.int32).reshape(-1)
        _check(lib().iu_b200_interpolate_at_batch_dev(self._h, n, _dev_ptr(r, "float64", 3 * n, "r"), len(iv), iv.ctypes.data_as(c_ip),
                                                      _dev_ptr(var_at_r, "float64", n * len(iv), "var_at_r"),
                                                      _dev_ptr(i_cell, "int32", n, "i_cell"), _cur_stream(self.device)))

    NOALLOC_MAX_QUERIES = 1 << 20

    def iu_interpolate_at_dev_noalloc(self, r, i_vars, var_at_r, i_cell):
        """iu_interpolate_at_dev for at most NOALLOC_MAX_QUERIES queries with all scratch in a static device array of the
        library: no cudaMalloc / cudaFree on the path (include/iu_b200.h)."""
        n = i_cell.numel()
        iv = np.ascontiguousarray(i_vars, dtype=np.int32).reshape(-1)
        _check(lib().iu_b200_interpolate_at_batch_dev_noalloc(self._h, n, _dev_ptr(r, "float64", 3 * n, "r"), len(iv), iv.ctypes.data_as(c_ip),
                                                              _dev_ptr(var_at_r, "float64", n * len(iv), "var_at_r"),
                                                              _dev_ptr(i_cell, "int32", n, "i_cell"), _cur_stream(self.device)))

    def iu_interpolate_at_icell_dev(self, r, i_vars, var_at_r, i_cell):
        n = i_cell.numel()
        iv = np.ascontiguousarray(i_vars, dtype=np.int32).reshape(-1)
        _check(lib().iu_b200_interpolate_at_icell_batch_dev(self._h, n, _dev_ptr(r, "float64", 3 * n, "r"), len(iv),
                                                            iv.ctypes.data_as(c_ip),
                                                            _dev_ptr(var_at_r, "float64", n * len(iv), "var_at_r"),
                                                            _dev_ptr(i_cell, "int32", n, "i_cell"), _cur_stream(self.device)))

    def iu_interpolate_and_cell_scalar_at_dev(self, r, i_vars, i_cell_var, var_at_r, cell_res, i_cell):
        n = i_cell.numel()
        iv = np.ascontiguousarray(i_vars, dtype=np.int32).reshape(-1)
        _check(lib().iu_b200_interpolate_and_cell_scalar_at_batch_dev(
            self._h, n, _dev_ptr(r, "float64", 3 * n, "r"), len(iv), iv.ctypes.data_as(c_ip), int(i_cell_var),
            _dev_ptr(var_at_r, "float64", n * len(iv), "var_at_r"), _dev_ptr(cell_res, "float64", n, "cell_res"),
            _dev_ptr(i_cell, "int32", n, "i_cell"), _cur_stream(self.device)))

    def iu_get_cell_scalar_at_dev(self, r, i_var, i_guess, res):
        n = i_guess.numel()
        _check(lib().iu_b200_get_cell_scalar_at_batch_dev(self._h, n, _dev_ptr(r, "float64", 3 * n, "r"), int(i_var),
                                                          _dev_ptr(i_guess, "int32", n, "i_guess"),
                                                          _dev_ptr(res, "float64", n, "res"), _cur_stream(self.device)))

    def iu_get_icell_scalar_at_dev(self, r, i_var, i_guess, res):
        n = i_guess.numel()
        _check(lib().iu_b200_get_icell_scalar_at_batch_dev(self._h, n, _dev_ptr(r, "float64", 3 * n, "r"), int(i_var),
                                                           _dev_ptr(i_guess, "int32", n, "i_guess"),
                                                           _dev_ptr(res, "int32", n, "res"), _cur_stream(self.device)))

    def iu_get_cell_through_neighbors_dev(self, max_steps, r0, r1, ic0, ic1, r_p, n_steps, status, i_icell_mask=0):
        n = ic0.numel()
        _check(lib().iu_b200_get_cell_through_neighbors_batch_dev(
            self._h, n, int(min(max_steps, 2**31 - 2)), _dev_ptr(r0, "float64", 3 * n, "r0"), _dev_ptr(r1, "float64", 3 * n, "r1"),
            _dev_ptr(ic0, "int32", n, "ic0"), _dev_ptr(ic1, "int32", n, "ic1"), _dev_ptr(r_p, "float64", 3 * n, "r_p"),
            _dev_ptr(n_steps, "int32", n, "n_steps"), _dev_ptr(status, "int32", n, "status"), int(i_icell_mask),
            _cur_stream(self.device)))

    def iu_integrate_along_field_dev(self, y, i_field, min_dx, max_dx, rtol, atol, n_steps, reverse=False, nvar=0, integrand=0,
                                     y_field=None, axisymmetric=False, i_icell_mask=0, mask_value=0, boundary_material=None):
        """iu_integrate_along_field on device tensors: y (n, max_steps, ndim+nvar) f64 with y[:, 0] set, n_steps (n) i32."""
        iv = np.ascontiguousarray(i_field, dtype=np.int32).reshape(-1)
        ndim = len(iv)
        n, max_steps = y.shape[0], y.shape[1]
        _check(lib().iu_b200_integrate_along_field_batch_dev(
            self._h, n, ndim, int(nvar), int(integrand), iv.ctypes.data_as(c_ip), float(min_dx), float(max_dx), int(max_steps),
            float(rtol), float(atol), int(bool(reverse)), _dev_ptr(y, "float64", n * max_steps * (ndim + nvar), "y"),
            None if y_field is None else _dev_ptr(y_field, "float64", n * max_steps * ndim, "y_field"),
            _dev_ptr(n_steps, "int32", n, "n_steps"), int(bool(axisymmetric)), int(i_icell_mask), int(mask_value),
            None if boundary_material is None else _dev_ptr(boundary_material, "int32", n, "boundary_material"),
            _cur_stream(self.device)))

    def check_dev(self):
        """Synchronises the current stream; raises IUError(E_HINT) if a ``*_dev`` call since the last
        check carried ``i_cell > n_cells`` (the reference's ``error stop`` at :490)."""
        _check(lib().iu_b200_check_dev(self._h, _cur_stream(self.device)))


def iu_read_grid(filename, coord_scale_factor=1.0, device=-1):
    """Module-level spelling of the reference entry point (:820-927)."""
    return Grid.iu_read_grid(filename, coord_scale_factor, device)
