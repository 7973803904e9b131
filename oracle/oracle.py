"""ctypes binding of oracle/_build/libiu_oracle.so -- TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / ``--impl reference`` legs
may import this module; the product package never does.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libiu_oracle.so")
_lib = None

c_dp = C.POINTER(C.c_double)
c_ip = C.POINTER(C.c_int)


def build(force=False):
    src = os.path.join(_HERE, "iu_oracle.cpp")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.run(["make", "-C", _HERE], check=True, capture_output=True)
    return _SO


def lib():
    global _lib
    if _lib is None:
        build()
        L = C.CDLL(_SO)
        L.iu_oracle_last_error.restype = C.c_char_p
        L.iu_oracle_grid_create.restype = C.c_void_p
        L.iu_oracle_grid_create.argtypes = [C.c_int, C.c_int, C.c_int, c_dp, c_ip, c_ip, C.c_int, c_dp, C.c_int, c_dp, C.c_int, c_ip]
        L.iu_oracle_grid_from_binda.restype = C.c_void_p
        L.iu_oracle_grid_from_binda.argtypes = [C.c_char_p, C.c_double]
        L.iu_oracle_grid_destroy.argtypes = [C.c_void_p]
        L.iu_oracle_grid_info.argtypes = [C.c_void_p, c_ip]
        L.iu_oracle_f64_array.restype = c_dp
        L.iu_oracle_f64_array.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int64)]
        L.iu_oracle_i32_array.restype = c_ip
        L.iu_oracle_i32_array.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int64)]
        L.iu_oracle_data_index.argtypes = [C.c_void_p, C.c_int, C.c_char_p]
        L.iu_oracle_set_point_data.argtypes = [C.c_void_p, C.c_int, c_dp]
        L.iu_oracle_nearest.argtypes = [C.c_void_p, C.c_int64, c_dp, C.c_int, c_ip, c_dp, c_dp, C.c_int]
        L.iu_oracle_get_cell.argtypes = [C.c_void_p, C.c_int64, c_dp, c_ip, C.c_int, c_ip, c_dp, C.c_int]
        L.iu_oracle_interpolate_at.argtypes = [C.c_void_p, C.c_int64, c_dp, C.c_int, c_ip, c_dp, c_ip, C.c_int, C.c_int]
        L.iu_oracle_interpolate_at_icell.argtypes = [C.c_void_p, C.c_int64, c_dp, C.c_int, c_ip, c_dp, c_ip]
        L.iu_oracle_get_cell_scalar_at.argtypes = [C.c_void_p, C.c_int64, c_dp, C.c_int, c_ip, c_dp, C.c_int]
        L.iu_oracle_get_cell_scalar_at_mt.argtypes = [C.c_void_p, C.c_int64, c_dp, C.c_int, c_ip, c_dp, C.c_int, C.c_int]
        L.iu_oracle_get_icell_scalar_at.argtypes = [C.c_void_p, C.c_int64, c_dp, C.c_int, c_ip, c_ip, C.c_int]
        L.iu_oracle_walk.argtypes = [C.c_void_p, C.c_int64, C.c_int, c_dp, c_dp, c_ip, c_ip, c_dp, c_ip, c_ip, C.c_int]
        L.iu_oracle_point_is_inside_cell.argtypes = [C.c_void_p, C.c_int64, c_dp, c_ip, c_ip]
        L.iu_oracle_integrate_along_field.argtypes = [C.c_void_p, C.c_int64, C.c_int, C.c_int, C.c_int, c_ip, C.c_double, C.c_double,
                                                      C.c_int, C.c_double, C.c_double, C.c_int, c_dp, c_dp, c_ip, C.c_int, C.c_int,
                                                      C.c_int, c_ip, C.c_int, C.c_int]
        L.iu_oracle_cell_neighbors.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, c_dp, c_ip, c_ip]
        L.iu_oracle_max_threads.restype = C.c_int
        L.iu_oracle_last_trace_margin.restype = C.c_double
        L.iu_oracle_benchmark_pass.restype = C.c_double
        L.iu_oracle_benchmark_pass.argtypes = [C.c_void_p, C.c_int64, c_dp, C.c_int, c_ip, c_dp, c_ip, C.c_int]
        _lib = L
    return _lib


def _d(a):
    return a.ctypes.data_as(c_dp)


def _i(a):
    return a.ctypes.data_as(c_ip)


def _f64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _i32(a):
    return np.ascontiguousarray(a, dtype=np.int32)


def cell_neighbors(cells0, points, cell_type):
    """get_cell_neighbors (convert_to_binary.py:118-162): 0-based (Nc, ppc) neighbour table, -1 = none."""
    cells0 = _i32(cells0)
    points = _f64(points)
    nc, ppc = cells0.shape
    npf = {"triangle": 2, "quad": 2, "tetra": 3}[cell_type]
    out = np.empty((nc, ppc), dtype=np.int32)
    rc = lib().iu_oracle_cell_neighbors(len(points), nc, ppc, npf, _d(points), _i(cells0), _i(out))
    if rc != 0:
        raise RuntimeError(lib().iu_oracle_last_error().decode())
    return out


class OracleGrid:
    """Restated ``iu_grid_t`` (m_interp_unstructured.f90:24-59).  All indices 1-based."""

    def __init__(self, handle):
        if not handle:
            raise RuntimeError(lib().iu_oracle_last_error().decode())
        self.h = handle
        info = np.zeros(7, dtype=np.int32)
        lib().iu_oracle_grid_info(self.h, _i(info))
        (self.cell_type, self.n_points, self.n_cells, self.ppc, self.n_point_data, self.n_cell_data,
         self.n_icell_data) = (int(x) for x in info)

    @classmethod
    def from_binda(cls, path, coord_scale_factor=1.0):
        return cls(lib().iu_oracle_grid_from_binda(str(path).encode(), float(coord_scale_factor)))

    @classmethod
    def from_arrays(cls, cell_type, points, cells0, neighbors0, point_data=None, cell_data=None, icell_data=None):
        """cells0 / neighbors0 are 0-based (as on disk); +1 applied here as iu_read_grid does (:871-872).
        point_data (Np, nv), cell_data (Nc, nv), icell_data (Nc, nv) -- converted to column-major."""
        ct = {"triangle": 1, "quad": 2, "tetra": 3}.get(cell_type, cell_type)
        points = _f64(points)
        cells = _i32(np.asarray(cells0) + 1)
        neigh = _i32(np.asarray(neighbors0) + 1)

        def cm(a, dt):
            if a is None:
                return 0, np.zeros(1, dtype=dt)
            a = np.asarray(a, dtype=dt)
            if a.ndim == 1:
                a = a[:, None]
            return a.shape[1], np.ascontiguousarray(a.T)

        npd, pd = cm(point_data, np.float64)
        ncd, cd = cm(cell_data, np.float64)
        nicd, icd = cm(icell_data, np.int32)
        return cls(lib().iu_oracle_grid_create(ct, len(points), len(cells), _d(points), _i(cells), _i(neigh),
                                               npd, _d(pd), ncd, _d(cd), nicd, _i(icd)))

    def __del__(self):
        try:
            if self.h:
                lib().iu_oracle_grid_destroy(self.h)
                self.h = None
        except Exception:
            pass

    def _arr_f64(self, which, shape):
        n = C.c_int64()
        p = lib().iu_oracle_f64_array(self.h, which, C.byref(n))
        return np.ctypeslib.as_array(p, shape=(n.value,)).copy().reshape(shape)

    def _arr_i32(self, which, shape):
        n = C.c_int64()
        p = lib().iu_oracle_i32_array(self.h, which, C.byref(n))
        return np.ctypeslib.as_array(p, shape=(n.value,)).copy().reshape(shape)

    # derived geometry, C-order views of the Fortran arrays: (Nc, ppc, 3) etc.
    @property
    def points(self):
        return self._arr_f64(0, (self.n_points, 3))

    @property
    def cell_points(self):
        return self._arr_f64(1, (self.n_cells, self.ppc, 3))

    @property
    def cell_face_normals(self):
        return self._arr_f64(2, (self.n_cells, self.ppc, 3))

    @property
    def cell_volume(self):
        return self._arr_f64(3, (self.n_cells,))

    @property
    def cell_centers(self):
        return self._arr_f64(4, (self.n_cells, 3))

    @property
    def point_data(self):
        return self._arr_f64(5, (self.n_point_data, self.n_points)).T

    @property
    def cells(self):
        return self._arr_i32(0, (self.n_cells, self.ppc))

    @property
    def neighbors(self):
        return self._arr_i32(1, (self.n_cells, self.ppc))

    def get_point_data_index(self, name):
        return lib().iu_oracle_data_index(self.h, 0, name.encode())

    def get_cell_data_index(self, name):
        return lib().iu_oracle_data_index(self.h, 1, name.encode())

    def get_icell_data_index(self, name):
        return lib().iu_oracle_data_index(self.h, 2, name.encode())

    def set_point_data(self, col, data):
        data = _f64(data)
        assert data.shape == (self.n_points,)
        lib().iu_oracle_set_point_data(self.h, int(col), _d(data))
        self.n_point_data = max(self.n_point_data, int(col))

    def nearest(self, r, mode="brute", nthreads=0):
        r = _f64(r)
        n = len(r)
        idx = np.zeros(n, dtype=np.int32)
        best = np.zeros(n)
        second = np.zeros(n)
        lib().iu_oracle_nearest(self.h, n, _d(r), 0 if mode == "kdtree2" else 1, _i(idx), _d(best), _d(second), nthreads)
        return (idx, best, second) if mode != "kdtree2" else idx

    def get_cell(self, r, i_cell=None, nn="kdtree2", diagnostics=False, nthreads=0):
        r = _f64(r)
        n = len(r)
        ic = np.zeros(n, dtype=np.int32) if i_cell is None else _i32(i_cell).copy()
        ns = np.zeros(n, dtype=np.int32)
        mg = np.zeros((n, 2)) if diagnostics else None
        lib().iu_oracle_get_cell(self.h, n, _d(r), _i(ic), 0 if nn == "kdtree2" else 1, _i(ns),
                                 _d(mg) if diagnostics else None, nthreads)
        return (ic, ns, mg) if diagnostics else ic

    def interpolate_at(self, r, i_vars, i_cell=None, var_at_r=None, nn="kdtree2", nthreads=0):
        r = _f64(r)
        n = len(r)
        iv = _i32(i_vars)
        ic = np.zeros(n, dtype=np.int32) if i_cell is None else _i32(i_cell).copy()
        out = np.zeros((n, len(iv))) if var_at_r is None else _f64(var_at_r).copy()
        rc = lib().iu_oracle_interpolate_at(self.h, n, _d(r), len(iv), _i(iv), _d(out), _i(ic), 0 if nn == "kdtree2" else 1, nthreads)
        if rc != 0:
            raise RuntimeError(lib().iu_oracle_last_error().decode())
        return out, ic

    def interpolate_at_icell(self, r, i_vars, i_cell):
        r = _f64(r)
        iv = _i32(i_vars)
        ic = _i32(i_cell)
        out = np.zeros((len(r), len(iv)))
        lib().iu_oracle_interpolate_at_icell(self.h, len(r), _d(r), len(iv), _i(iv), _d(out), _i(ic))
        return out

    def get_cell_scalar_at(self, r, i_var, i_guess=None, res=None, nn="kdtree2", nthreads=1):
        r = _f64(r)
        n = len(r)
        ic = np.zeros(n, dtype=np.int32) if i_guess is None else _i32(i_guess).copy()
        out = np.zeros(n) if res is None else _f64(res).copy()
        lib().iu_oracle_get_cell_scalar_at_mt(self.h, n, _d(r), int(i_var), _i(ic), _d(out), 0 if nn == "kdtree2" else 1, int(nthreads))
        return out, ic

    def get_icell_scalar_at(self, r, i_var, i_guess=None, res=None, nn="kdtree2"):
        r = _f64(r)
        n = len(r)
        ic = np.zeros(n, dtype=np.int32) if i_guess is None else _i32(i_guess).copy()
        out = np.zeros(n, dtype=np.int32) if res is None else _i32(res).copy()
        lib().iu_oracle_get_icell_scalar_at(self.h, n, _d(r), int(i_var), _i(ic), _i(out), 0 if nn == "kdtree2" else 1)
        return out, ic

    def get_cell_through_neighbors(self, max_steps, r0, r1, ic0, i_icell_mask=0):
        r0, r1 = _f64(r0), _f64(r1)
        n = len(r0)
        ic0 = _i32(ic0)
        ic1 = np.zeros(n, dtype=np.int32)
        r_p = np.full((n, 3), np.nan)
        ns = np.full(n, -1, dtype=np.int32)
        st = np.zeros(n, dtype=np.int32)
        lib().iu_oracle_walk(self.h, n, int(max_steps), _d(r0), _d(r1), _i(ic0), _i(ic1), _d(r_p), _i(ns), _i(st), int(i_icell_mask))
        return ic1, r_p, ns, st

    def integrate_along_field(self, y0, i_field, min_dx, max_dx, max_steps, rtol, atol, reverse=False, nvar=0, integrand=0,
                              axisymmetric=False, i_icell_mask=0, mask_value=0, nn="brute", nthreads=0):
        """Batched iu_integrate_along_field (:989-1217) with built-in integrands.  y0: (n, ndim+nvar)
        initial position + state.  Returns y (n, max_steps, ndim+nvar), y_field (n, max_steps, ndim),
        n_steps (n), boundary_material (n; -99 where the reference leaves it unset)."""
        y0 = _f64(y0)
        iv = _i32(i_field)
        n, nt = y0.shape
        ndim = len(iv)
        assert nt == ndim + nvar
        y = np.full((n, max_steps, nt), np.nan)
        y[:, 0, :] = y0
        yf = np.full((n, max_steps, ndim), np.nan)
        ns = np.zeros(n, dtype=np.int32)
        bm = np.full(n, -99, dtype=np.int32)
        lib().iu_oracle_integrate_along_field(self.h, n, ndim, int(nvar), int(integrand), _i(iv), float(min_dx), float(max_dx),
                                              int(max_steps), float(rtol), float(atol), int(bool(reverse)), _d(y), _d(yf), _i(ns),
                                              int(bool(axisymmetric)), int(i_icell_mask), int(mask_value), _i(bm),
                                              0 if nn == "kdtree2" else 1, nthreads)
        self.last_trace_margin = float(lib().iu_oracle_last_trace_margin())   # min |err - 1| over the accept / reject decisions
        return y, yf, ns, bm

    def point_is_inside_cell(self, r, i_cell):
        r = _f64(r)
        ic = _i32(i_cell)
        out = np.zeros(len(r), dtype=np.int32)
        lib().iu_oracle_point_is_inside_cell(self.h, len(r), _d(r), _i(ic), _i(out))
        return out.astype(bool)

    def benchmark_pass(self, r, i_vars, i_cell, nthreads=0):
        """benchmark.f90:46-55: times one pass; i_cell updated in place. Returns (seconds, res)."""
        r = _f64(r)
        iv = _i32(i_vars)
        res = np.zeros((len(r), len(iv)))
        assert i_cell.dtype == np.int32 and i_cell.flags.c_contiguous
        t = lib().iu_oracle_benchmark_pass(self.h, len(r), _d(r), len(iv), _i(iv), _d(res), _i(i_cell), nthreads)
        return t, res
