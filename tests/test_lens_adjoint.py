"""Lens adjoint, CPU part: the hand-derived reverse sweep over the interfaces (lrt_lens_adj.cuh, run here through
the library's host entry lrt_sensor_trace_vjp_host — same LRT_HD code as the CUDA adjoint kernel) against the
oracle's forward-mode dual-number Jacobian in float64 (oracle/oracle_spot.cpp orc_trace_jac). The reference has
no derivative of its lens code, so there is no golden vector to pin against; see DESIGN.md ("adjoint oracle")."""
import ctypes as C

import numpy as np
import pytest

from oracle_lib import *  # noqa

ld = load_lens_designs()
COLS = {0: [1, 3, 13, 14], 1: [1, 13, 14], 2: [1], 3: [1, 3, 4, 6, 7, 8, 13, 14]}


def jac_oracle(rows, from_world, od, sel):
    L = oracle()
    n_sel = len(sel)
    out6 = np.zeros(6); jac = np.zeros((6, 6 + n_sel))
    ok = L.orc_trace_jac(C.c_int(len(rows)), fp(np.ascontiguousarray(rows, f32)), C.c_int(int(from_world)), fp(np.ascontiguousarray(od, f32)),
                         C.c_int(n_sel), np.ascontiguousarray(sel, np.int32).ctypes.data_as(IP),
                         out6.ctypes.data_as(C.POINTER(C.c_double)), jac.ctypes.data_as(C.POINTER(C.c_double)))
    return bool(ok), out6, jac


@pytest.mark.parametrize("name,from_world", [("doublet", True), ("doublet", False), ("gauss", True), ("asph", True), ("asph", False),
                                             ("exp2a", True), ("fisheye", False), ("tessar", True)])
def test_trace_vjp_matches_dual_number_jacobian(mi, name, from_world):
    if name not in ld.all_designs():
        pytest.skip(name)
    s = mi.RealisticLensCamera(mi.Film(128, 128), name)
    rows = s.interfaces()
    n = len(rows)
    rng = np.random.default_rng(7)
    # paraxial-ish rays that make it through long stacks
    k = 64
    od = np.zeros((k, 6), f32)
    if from_world:
        od[:, 0:2] = (rng.random((k, 2)) - .5) * .3 * rows[-1, 2]; od[:, 2] = rows[-1, 1] + 0.05
        d = np.zeros((k, 3)); d[:, 0:2] = (rng.random((k, 2)) - .5) * .04; d[:, 2] = -1
    else:
        od[:, 0:2] = (rng.random((k, 2)) - .5) * .1 * rows[0, 2]; od[:, 2] = 0
        d = np.zeros((k, 3)); d[:, 0:2] = (rng.random((k, 2)) - .5) * .1; d[:, 2] = 1
    od[:, 3:] = (d / np.linalg.norm(d, axis=1, keepdims=True)).astype(f32)
    sel = [(i, c) for i in range(n) for c in COLS[int(rows[i, 0])]][:56]
    bar = rng.standard_normal((k, 6)).astype(f32)
    checked = 0
    for r in range(k):
        ok, out6, jac = jac_oracle(rows, from_world, od[r], sel)
        in_bar, grad, act = s.trace_vjp(od[r, :3], od[r, 3:], bar[r, :3], bar[r, 3:], from_world=from_world)
        assert act[0] == ok
        if not ok:
            continue
        checked += 1
        want = bar[r].astype(np.float64) @ jac              # [6 + n_sel]
        got_p = np.array([grad[i, c] for i, c in sel], np.float64)
        scale = np.abs(want[6:]).max() + 1e-12
        assert np.abs(got_p - want[6:]).max() <= 2e-3 * scale, (name, r, got_p, want[6:])
        # incoming-ray adjoint: origin exactly; direction on the tangent space of the unit vector
        dd = od[r, 3:].astype(np.float64)
        proj = np.eye(3) - np.outer(dd, dd)
        so = np.abs(want[:6]).max() + 1e-12
        assert np.abs(in_bar[0, :3] - want[:3]).max() <= 2e-3 * so
        assert np.abs(proj @ in_bar[0, 3:] - proj @ want[3:6]).max() <= 2e-3 * so
    assert checked >= 3


# ------------------------------------------------------------------------------------------------ spot path
def spot_case(name="doublet", angle=1.5, n_pix=24, full=8192, spp=2, props=None, distance=1000.0):
    """A field source at `angle` degrees and a crop window of n_pix^2 pixels of a full^2 film centred on the chief ray."""
    L = oracle()
    design = ld.design(name, props or {})
    big = FilmSpec(full, full, None, 35.0)
    h = orc_camera(design, big)
    rows = cam_interfaces(L, "orc", h)
    src = np.array([distance * np.tan(np.radians(angle)), 0.0, rows[-1, 1] + distance], f32)
    o = np.zeros(11, f32)
    L.orc_cam_sample_direction(h, C.c_int(1), fp(src), fp(np.array([.5, .5], f32)), fp(o))
    assert o[9] > 0, "chief ray blocked"
    cx, cy = int(o[0]) - n_pix // 2, int(o[1]) - n_pix // 2
    L.orc_cam_destroy(h)
    film = FilmSpec(full, full, [n_pix, n_pix, cx, cy], 35.0)
    h = orc_camera(design, film)
    f2s = np.zeros(16, f32); L.orc_cam_film_to_sample(h, fp(f2s))
    s2f = np.zeros(16, f32); L.orc_cam_sample_to_film(h, fp(s2f))
    norm = 1.0 / (abs(float(s2f[0])) * abs(float(s2f[5])))
    return dict(design=design, film=film, h=h, rows=rows, src=src, f2s=f2s, norm=norm, spp=spp, name=name, props=props or {})


@pytest.mark.parametrize("filter_kind,tol", [(2, 2e-4), (1, 2e-2)])
def test_checker_gradient_agrees_with_finite_differences(filter_kind, tol):
    """The dual-number checker against central differences of its own value (float64): pins the derivative that the
    CUDA adjoint is compared with on the GPU."""
    c = spot_case("doublet", 1.5, n_pix=24, spp=2)
    rng = np.random.default_rng(3)
    G = rng.standard_normal((24, 24, 4)).astype(f32)
    rows = c["rows"]
    sel = [(i, col) for i in range(len(rows)) for col in COLS[int(rows[i, 0])]]
    args = (rows, c["f2s"], c["norm"], 1e-2, 1e4, c["film"], c["src"], [1.0, 0.8, 0.6], 5, c["spp"], G, sel)
    # gaussian: smooth, tight tolerance; tent: the estimate is piecewise linear in the splat position, so a finite
    # step that carries samples across pixel centres sees the kinks (loose tolerance)
    kw = dict(filter_kind=filter_kind)
    phi, grad = orc_spot_grad(*args, **kw)
    assert np.isfinite(grad).all() and np.abs(grad).max() > 0
    for s, (i, col) in enumerate(sel):
        hstep = 1e-7 * max(abs(float(rows[i, col])), 1e-2)
        d = np.zeros(len(sel)); d[s] = hstep
        pp, _ = orc_spot_grad(*args, delta=d, **kw)
        pm, _ = orc_spot_grad(*args, delta=-d, **kw)
        fd = (pp - pm) / (2 * hstep)
        assert abs(fd - grad[s]) <= tol * max(abs(grad[s]), 1e-2 * np.abs(grad).max()), (i, col, fd, grad[s])
