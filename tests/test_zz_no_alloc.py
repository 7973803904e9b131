"""iu_b200_interpolate_at_batch_dev_noalloc: the allocation-free variant of iu_b200_interpolate_at_batch_dev (scratch in a
static __device__ array of the library; no counterpart in the reference).

Collected last on purpose (file name): a failure here must not hide other tests.
Through the C ABI (ctypes), two calls in a row on the same stream with different batches (the second smaller, so stale
scratch contents of the first would show), compared with the existing entry point: cell indices and values bit-identical
(the contractual tolerance of the other GPU tests is 1e-12 relative)."""
import ctypes as C

import numpy as np
import pytest

from interpolate_unstructured_b200 import api, meshgen

pytestmark = pytest.mark.gpu
RTOL = 1e-12


def _free_bytes():
    import torch
    return torch.cuda.mem_get_info(0)[0]


def test_no_alloc_entry_point_matches_the_existing_one_twice_in_a_row():
    import torch
    dev = torch.device("cuda", 0)
    pts, cells = meshgen.kuhn_tet_mesh(14, jitter=0.1, seed=21)
    nb = api.build_cell_neighbors(cells, pts, "tetra")
    pd = meshgen.smooth_fields(pts, 4)
    g = api.Grid.from_arrays("tetra", pts, cells, nb, pd, device=0)
    L = api.lib()
    iv = np.array([1, 2, 3, 4], dtype=np.int32)
    ivp = iv.ctypes.data_as(C.POINTER(C.c_int32))
    lo, hi = pts.min(0), pts.max(0)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)

    def both(nq, seed, hinted):
        r = meshgen.uniform_queries(nq, lo - 0.03 * (hi - lo), hi + 0.03 * (hi - lo), seed=seed)     # some queries outside the mesh
        hint = np.zeros(nq, dtype=np.int32)
        if hinted:
            rng = np.random.default_rng(seed)
            hint = np.where(np.arange(nq) % 3 == 0, rng.integers(1, len(cells) + 1, size=nq), 0).astype(np.int32)
        r_t = torch.from_numpy(r).to(dev)
        out = []
        for fn in (L.iu_b200_interpolate_at_batch_dev, L.iu_b200_interpolate_at_batch_dev_noalloc):
            ic_t = torch.from_numpy(hint.copy()).to(dev)
            val_t = torch.full((nq, 4), -7.25, dtype=torch.float64, device=dev)
            torch.cuda.synchronize()
            free0 = _free_bytes()
            rc = fn(g._h, nq, C.c_void_p(r_t.data_ptr()), 4, ivp, C.c_void_p(val_t.data_ptr()), C.c_void_p(ic_t.data_ptr()), stream)
            assert rc == 0, L.iu_b200_last_error().decode()
            torch.cuda.synchronize()
            out.append((val_t.cpu().numpy(), ic_t.cpu().numpy(), free0 - _free_bytes()))
        g.check_dev()
        return out

    for call, (nq, seed, hinted) in enumerate(((400_000, 5, False), (150_001, 6, True))):
        (val_a, ic_a, _), (val_b, ic_b, grew) = both(nq, seed, hinted)
        assert (ic_b >= 1).any() and (ic_b < 1).any()
        assert np.array_equal(ic_a, ic_b), (call, int((ic_a != ic_b).sum()))
        assert (np.abs(val_a - val_b) <= RTOL * np.maximum(1.0, np.abs(val_a))).all(), call
        assert np.array_equal(val_a, val_b), call                     # in fact bit-identical, untouched rows included
        if call == 1:                                                 # first call: lazy loading of the kernel image may take memory
            assert grew <= 0, f"device memory in use grew by {grew} bytes across an allocation-free call"

    # sizes beyond the stated maximum are rejected with a clear error, nothing is launched
    n_big = api.Grid.NOALLOC_MAX_QUERIES + 1
    assert L.iu_b200_interpolate_at_batch_dev_noalloc(g._h, n_big, C.c_void_p(1), 4, ivp, C.c_void_p(1), C.c_void_p(1), stream) == api.E_INVAL
    assert b"IU_B200_NOALLOC_MAX_QUERIES" in L.iu_b200_last_error()
    # the Python mirror of the entry point
    r_t = torch.from_numpy(meshgen.uniform_queries(1000, [0.1] * 3, [0.9] * 3, seed=9)).to(dev)
    ic_t = torch.zeros(1000, dtype=torch.int32, device=dev)
    val_t = torch.zeros((1000, 4), dtype=torch.float64, device=dev)
    g.iu_interpolate_at_dev_noalloc(r_t, [1, 2, 3, 4], val_t, ic_t)
    g.check_dev()
    assert int((ic_t >= 1).sum()) == 1000
    assert float((val_t[:, 0] - (r_t.sum(1) + 1)).abs().max()) < 1e-12
    g.close()
