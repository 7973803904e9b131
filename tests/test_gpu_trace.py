"""GPU parity of the batched field-line tracer (iu_b200_integrate_along_field_batch) against the
oracle's restatement of iu_integrate_along_field (m_interp_unstructured.f90:989-1217).

Tolerance: step counts, boundary materials and untouched columns must be identical; positions and
integrated variables agree to 1e-10 (absolute, unit-size domains).  They are not required to be
bit-identical because the step-size controller evaluates (1/err)**(1/3.0) (:1187) through pow(),
whose last bit differs between glibc (oracle / gfortran's libm) and the CUDA math library."""
import os

import numpy as np
import pytest

from interpolate_unstructured_b200 import api, binda, meshgen
from oracle.oracle import OracleGrid

pytestmark = pytest.mark.gpu
TOL = 1e-10


def _compare(gy, gf, gns, gbm, oy, of, ons, obm, max_steps):
    assert (gns == ons).all()
    assert (gbm == obm).all()
    for i in range(len(gns)):
        m = min(int(gns[i]), max_steps)
        assert np.abs(gy[i, :m] - oy[i, :m]).max() <= TOL
        if gns[i] == 1 and np.isnan(of[i, 0]).all():          # invalid seed: y_field(:, 1) is never written (:1057-1070)
            assert np.isnan(gf[i, 0]).all()
            continue
        assert np.abs(gf[i, :m] - of[i, :m]).max() <= TOL * max(1.0, np.abs(of[i, :m]).max())
    untouched = np.arange(max_steps)[None, :] >= np.minimum(gns, max_steps)[:, None]
    assert np.isnan(gy[untouched]).all() and np.isnan(gf[untouched]).all()


def test_trace_field_reference_criterion(golden_dir):
    """test_trace_field.f90:16-53 through the C ABI: rotation field on triangle.binda from (1.5, 0)."""
    ug = api.iu_read_grid(os.path.join(golden_dir, "triangle.binda"))
    i_vx, i_vy = ug.iu_add_point_data("vx"), ug.iu_add_point_data("vy")
    pts = {e.name: e.data for e in binda.read_binda(os.path.join(golden_dir, "triangle.binda")) if e.name == "points"}["points"].astype(np.float64)
    ug.set_point_data(i_vx, -pts[:, 1])
    ug.set_point_data(i_vy, pts[:, 0])
    max_steps = 100
    y = np.full((1, max_steps, 3), np.nan)
    y[0, 0] = [1.5, 0.0, -0.75 * np.pi]
    y, yf, ns, bm = ug.iu_integrate_along_field(y, [i_vx, i_vy], 1e-5, 1e-1, 1e-3, 1e-3, nvar=1, integrand=1)
    assert 1 < ns[0] <= max_steps                           # "Boundary not reached" otherwise
    end = y[0, ns[0] - 1]
    assert abs(end[0]) < 2e-2 and abs(end[1] - 1.5) < 2e-2 and abs(end[2]) < 2e-2 and bm[0] == -1
    ent = {e.name: e.data for e in binda.read_binda(os.path.join(golden_dir, "triangle.binda")) if e.name != "point_data"}
    o = OracleGrid.from_arrays("triangle", pts, ent["cells"], ent["cell_neighbors"], np.stack([np.zeros(len(pts)), np.zeros(len(pts)), -pts[:, 1], pts[:, 0]], axis=1))
    oy, of, ons, obm = o.integrate_along_field(np.array([[1.5, 0.0, -0.75 * np.pi]]), [3, 4], 1e-5, 1e-1, max_steps, 1e-3, 1e-3, nvar=1, integrand=1)
    _compare(y, yf, ns, bm, oy, of, ons, obm, max_steps)


@pytest.mark.parametrize("kind,n,ndim", [("quad", 40, 2), ("triangle", 40, 2), ("tetra", 10, 3)])
@pytest.mark.parametrize("variant", ["arc", "magnitude_mask", "reverse_axisymmetric_short"])
def test_trace_parity(kind, n, ndim, variant):
    pts, cells = getattr(meshgen, {"quad": "quad_mesh", "triangle": "triangle_mesh", "tetra": "kuhn_tet_mesh"}[kind])(n, jitter=0.1, seed=3)
    nb = api.build_cell_neighbors(cells, pts, kind)
    c = pts - 0.5
    # swirl + drift: lines spiral outwards and leave through the boundary; third component for tets
    field = np.stack([-c[:, 1] + 0.3 * c[:, 0], c[:, 0] + 0.3 * c[:, 1], 0.2 + 0.5 * c[:, 2] if ndim == 3 else np.zeros(len(pts))], axis=1)
    cen = pts[cells].mean(1)
    material = (np.hypot(cen[:, 0] - 0.5, cen[:, 1] - 0.5) > 0.33).astype(np.int32) * 7      # ring of another material
    g = api.Grid.from_arrays(kind, pts, cells, nb, field, icell_data=material)
    o = OracleGrid.from_arrays(kind, pts, cells, nb, field, icell_data=material)
    rng = np.random.default_rng(11)
    nl = 3000
    seeds = 0.5 + (rng.random((nl, 3)) - 0.5) * np.array([0.5, 0.5, 0.5 if ndim == 3 else 0.0])
    seeds[:20] += 2.0                                                                          # seeds outside the mesh
    i_field = [1, 2, 3][:ndim]
    kw = dict(arc=dict(nvar=1, integrand=1, max_steps=120),
              magnitude_mask=dict(nvar=2, integrand=2, i_icell_mask=1, mask_value=0, max_steps=80),
              reverse_axisymmetric_short=dict(nvar=0, integrand=0, reverse=True, axisymmetric=True, max_steps=12))[variant]
    max_steps = kw.pop("max_steps")
    nvar = kw["nvar"]
    y0 = np.concatenate([seeds[:, :ndim], rng.random((nl, nvar))], axis=1)
    y = np.full((nl, max_steps, ndim + nvar), np.nan)
    y[:, 0] = y0
    gy, gf, gns, gbm = g.iu_integrate_along_field(y, i_field, 1e-4, 2e-2, 1e-4, 1e-6, **kw)
    oy, of, ons, obm = o.integrate_along_field(y0, i_field, 1e-4, 2e-2, max_steps, 1e-4, 1e-6, **kw)
    _compare(gy, gf, gns, gbm, oy, of, ons, obm, max_steps)
    # margin report: the closest any accept / reject decision (err <= 1, :1165) of this batch came to flipping.  The curves
    # agree to 1e-10 and err is a ratio of O(1) quantities, so a margin orders of magnitude above that rules out a
    # decision that went the other way on the device (n_steps equality above is then not a coincidence).
    assert o.last_trace_margin > 1e-7, f"an accept/reject decision came within {o.last_trace_margin:.1e} of flipping"
    print(f"trace {kind} {variant}: closest |err - 1| over all decisions = {o.last_trace_margin:.2e}")
    # lines traced in Morton order of their seeds (forced on for this small batch) give the same bits as in input order
    g.set_option(api.OPT_SORT_QUERIES, 1)
    y2 = np.full((nl, max_steps, ndim + nvar), np.nan)
    y2[:, 0] = y0
    sy, sf, sns, sbm = g.iu_integrate_along_field(y2, i_field, 1e-4, 2e-2, 1e-4, 1e-6, **kw)
    g.set_option(api.OPT_SORT_QUERIES, 0)
    y3 = np.full((nl, max_steps, ndim + nvar), np.nan)
    y3[:, 0] = y0
    uy, uf, uns, ubm = g.iu_integrate_along_field(y3, i_field, 1e-4, 2e-2, 1e-4, 1e-6, **kw)
    assert np.array_equal(sy, uy, equal_nan=True) and np.array_equal(sf, uf, equal_nan=True)
    assert np.array_equal(sns, uns) and np.array_equal(sbm, ubm) and np.array_equal(sy, gy, equal_nan=True)
    assert (gns == 1).sum() >= 20
    assert (gbm[:20] == -1).all()
    if variant == "magnitude_mask":
        assert (gbm == 7).sum() > nl // 2                                                                # stopped at the material boundary


def test_trace_argument_errors():
    pts, cells = meshgen.quad_mesh(8, jitter=0.0)
    nb = api.build_cell_neighbors(cells, pts, "quad")
    g = api.Grid.from_arrays("quad", pts, cells, nb, np.ones((len(pts), 2)))
    y = np.zeros((2, 5, 2))
    with pytest.raises(api.IUError, match="max_dx < min_dx"):
        g.iu_integrate_along_field(y, [1, 2], 1e-1, 1e-2, 1e-3, 1e-3)
    with pytest.raises(api.IUError, match="i_field"):
        g.iu_integrate_along_field(y, [1, 3], 1e-3, 1e-2, 1e-3, 1e-3)
    with pytest.raises(api.IUError, match="integrand"):
        g.iu_integrate_along_field(np.zeros((2, 5, 3)), [1, 2], 1e-3, 1e-2, 1e-3, 1e-3, nvar=1, integrand=0)
