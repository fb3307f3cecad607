"""Mesh-lens parity mode, first slice: the notebooks' meshed surfaces (icosphere caps with vertex z from the
sag function, smooth vertex normals, black baffles), traced with the scene kernels, against the ANALYTIC interface table the
rest of the package traces — the comparison the notebook itself makes (doublet_full_optim.ipynb cell 7 reports a meshing error
of 3.8e-4 mm for its subdivision level). Plus the device BVH refit that stands in for params.update() of vertex positions."""
import numpy as np
import pytest

from oracle_lib import *  # noqa

pytestmark = pytest.mark.gpu


def doublet_rows(mi):
    L = mi.lens_design
    n1, n2 = L.cauchy_ior(1.45608, 0.021087), L.cauchy_ior(1.56895, 0.026305)
    return L.asphere_table([24.0, -24.0, -168.0107], [3.0, 2.2529, 46.7471], [8, 8, 8], [n1, n2, 1.000277])


def analytic_positions(mi, rows, ap, d, origin_offset):
    from mitsuba3_camera_b200.lens_designs import InterfaceList
    cam = mi.RealisticLensCamera(mi.Film(256, 256), "custom", prescription=InterfaceList(rows, film_z=0.0, fine_focus=False))
    o = (ap - d * origin_offset).astype(f32)
    po, pd, act = cam.trace(o, d.astype(f32), from_world=True)
    t = -po[:, 2] / pd[:, 2]
    return (po + t[:, None] * pd)[:, :2], act


def spot_deviation(mi, rows, subdiv, angle, n=20000):
    lens = mi.mesh_lens.MeshLens(mi, rows, subdiv=subdiv)
    src = mi.lens_design.field_source(angle, 1.0e3, front_z=float(rows[-1, 1]))
    off = 2.0 * float(rows[-1, 2])
    pf, alive, ap, d = lens.spot_positions(src, n, seed=1, origin_offset=off)
    pf, alive, ap, d = pf.cpu().numpy(), alive.cpu().numpy(), ap.cpu().numpy(), d.cpu().numpy()
    want, act = analytic_positions(mi, rows, ap, d, off)
    both = alive & act
    inner = np.linalg.norm(ap[:, :2], axis=1) < 0.95 * float(rows[-1, 2])
    return lens, pf, want, both, (alive[inner] == act[inner]).mean()


@pytest.mark.parametrize("angle", [0.0, 3.0])
def test_meshed_doublet_spot_matches_analytic_table(mi, angle):
    """Landing points on the film, ray by ray, meshed surfaces vs analytic table. The deviation is the meshing error of the
    notebooks' representation (flat facets + angle-weighted vertex normals, geometry.py / mesh.cpp:321-375): second order in
    the edge length — measured ~9e-3 mm rms at subdivision 5 and ~4x less per extra level (the notebook quotes 3.8e-4 mm for
    its own, finer meshes) — while centroid and rms radius of the spot, which the design loss uses, agree far better."""
    rows = doublet_rows(mi)
    dev_rms = {}
    for subdiv in (4, 5, 6):
        lens, pf, want, both, agree = spot_deviation(mi, rows, subdiv, angle)
        dev = np.linalg.norm(pf[both] - want[both], axis=1)
        dev_rms[subdiv] = float(np.sqrt((dev ** 2).mean()))
        spot = np.linalg.norm(want[both] - want[both].mean(0), axis=1)
        print(f"field {angle} deg, subdivision {subdiv} ({lens.scene.info()['triangles']} triangles): landing-point deviation rms "
              f"{dev_rms[subdiv] * 1e3:.3e} mm, max {dev.max() * 1e3:.3e} mm; analytic spot rms radius {np.sqrt((spot ** 2).mean()) * 1e3:.3e} mm; "
              f"alive agreement inside 0.95 R {agree:.4f}")
        assert agree > 0.998 and both.sum() > 15000
    assert dev_rms[5] < 1.5e-5 and dev_rms[6] < 0.45 * dev_rms[5] < 0.45 * 0.45 * dev_rms[4] * 1.5      # ~h^2 convergence
    c_m, c_a = pf[both].mean(0), want[both].mean(0)                                                     # subdivision 6
    assert np.linalg.norm(c_m - c_a) < 3e-7                      # centroids agree to 0.3 micron
    r_m = np.sqrt(((pf[both] - c_m) ** 2).sum(1).mean()); r_a = np.sqrt(((want[both] - c_a) ** 2).sum(1).mean())
    print(f"  centroid offset {np.linalg.norm(c_m - c_a) * 1e3:.3e} mm, rms radius mesh {r_m * 1e3:.4e} mm vs analytic {r_a * 1e3:.4e} mm")
    assert abs(r_m - r_a) < 0.1 * r_a + 5e-7


def test_refit_follows_a_design_step(mi):
    """New curvatures / vertex positions -> vertex z on the device -> lrt_scene_refit: the refitted scene traces like a scene
    built from scratch for the new table (same triangles; only the BVH boxes differ), and like the analytic table."""
    import torch
    rows = doublet_rows(mi)
    lens = mi.mesh_lens.MeshLens(mi, rows, subdiv=4)
    src = mi.lens_design.field_source(2.0, 1.0e3, front_z=float(rows[-1, 1]))
    off = 2.0 * float(rows[-1, 2])
    new = rows.copy()
    new[2, 3] *= 1.1; new[0, 3] *= 0.85; new[1, 1] += 2e-4; new[1, 4] = -0.5; new[2, 6] = 0.004      # curvatures, a vertex, a conic, a4
    lens.update(torch.tensor(new, device="cuda"))
    pf1, al1, ap, d = lens.spot_positions(src, 8000, seed=2, origin_offset=off)
    fresh = mi.mesh_lens.MeshLens(mi, rows, subdiv=4)            # same xy grid (the notebooks keep it fixed: only z moves)
    fresh.update(torch.tensor(new, device="cuda")); fresh.rebuild() # ... and a BVH built from scratch over the moved vertices
    pf2, al2, _, _ = fresh.spot_positions(src, 8000, seed=2, origin_offset=off)
    assert (al1 == al2).all()
    m = (al1 & al2)
    assert (pf1[m] - pf2[m]).abs().max().item() < 1e-9          # identical triangles: the refitted and the rebuilt tree find the same hits
    want, act = analytic_positions(mi, new, ap.cpu().numpy(), d.cpu().numpy(), off)
    both = m.cpu().numpy() & act
    dev = np.linalg.norm(pf1.cpu().numpy()[both] - want[both], axis=1)
    # subdivision 4: ~3x the subdivision-5 meshing error for the bulk of the rays. Rim vertices average their normal over half
    # a fan of faces, so rays through the outermost ring of triangles deviate most (worst rays sit at r > 0.98 R): the bound is
    # on the inner 90 % of the aperture and on quantiles
    inner = np.linalg.norm(ap.cpu().numpy()[both, :2], axis=1) < 0.9 * float(rows[-1, 2])
    print(f"refit: deviation median {np.median(dev) * 1e3:.3e} mm; inside 0.9 R: 95th percentile {np.quantile(dev[inner], 0.95) * 1e3:.3e} mm, "
          f"max {dev[inner].max() * 1e3:.3e} mm; whole aperture max {dev.max() * 1e3:.3e} mm")
    assert np.median(dev) < 3e-5 and np.quantile(dev[inner], 0.95) < 1.5e-4
    # and the un-refitted tree would have been wrong: the old boxes do not contain the moved surfaces everywhere
    old = mi.mesh_lens.MeshLens(mi, rows, subdiv=4)
    pf0, al0, _, _ = old.spot_positions(src, 8000, seed=2, origin_offset=off)
    assert (pf0[m] - pf1[m]).abs().max().item() > 1e-5


# ------------------------------------------------------------------------------------------------ the adjoint (lrt_mesh_adj.cu)
def _golden():
    import os
    return dict(np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "ref_mesh_lens.npz")))


def _mesh_trace():
    import os
    import sys
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "oracle"))
    import mesh_trace
    return mesh_trace


@pytest.mark.parametrize("smooth", [True, False])
def test_trace_adjoint_matches_oracle_reverse_mode(mi, smooth):
    """lrt_scene_trace_refractive_backward against torch.autograd over the float64 restatement (oracle/mesh_trace.py, itself
    pinned to the reference build on this very lens: tests/test_mesh_lens_parity.py): gradients with respect to the vertex
    positions, the vertex normals (as independent inputs), the relative indices and the rays, for a random adjoint of
    (o, d, weight). Interpolated vertex normals and flat facets."""
    import torch
    MT = _mesh_trace()
    G = _golden()
    lens = mi.mesh_lens.MeshLens(mi, G["rows"], subdiv=3, smooth=smooth)
    rng = np.random.default_rng(11)
    od = np.concatenate([G["rays"], G["rays"] + rng.normal(0, 2e-3, G["rays"].shape).astype(f32) * [1, 1, 0, 1, 1, 0]], 0).astype(f32)
    od_t = torch.tensor(od, device="cuda")
    n = len(od)
    out = torch.empty((n, 8), dtype=torch.float32, device="cuda")
    capi, lib = mi.capi, mi.capi.lib
    import ctypes as C
    K = C.c_int32(lens.n_surfaces + 1)
    capi.check(lib.lrt_scene_trace_refractive(lens.scene.handle, C.c_int64(n), capi.dev_ptr(od_t), capi.dev_ptr(lens.eta), K, capi.dev_ptr(out), None))
    torch.cuda.synchronize()
    # the oracle on the same float32 inputs
    a = lens.scene.arrays
    V = torch.tensor(lens.vertices.cpu().numpy(), dtype=torch.float64, requires_grad=True)
    N = torch.tensor(a["vertex_normals"], dtype=torch.float64, requires_grad=True) if smooth else None
    eta = torch.tensor(lens.eta.cpu().numpy(), dtype=torch.float64, requires_grad=True)
    odc = torch.tensor(od, dtype=torch.float64, requires_grad=True)
    F = torch.tensor(a["indices"].astype(np.int64)); ts = torch.tensor(a["tri_shape"].astype(np.int64))
    o, d, alive, w = MT.trace(V, F, N, ts, eta, odc[:, :3], odc[:, 3:], lens.n_surfaces + 1)
    got = out.cpu().numpy()
    assert (alive.numpy() == (got[:, 6] > 0)).mean() > 0.99
    both = alive.numpy() & (got[:, 6] > 0)
    assert both.sum() > 0.8 * n
    assert np.abs(got[both, 0:3] - o.detach().numpy()[both]).max() < 1e-6 and np.abs(got[both, 3:6] - d.detach().numpy()[both]).max() < 2e-6
    assert np.abs(got[both, 7] - w.detach().numpy()[both]).max() < 1e-5
    # a random adjoint of (o, d, weight) for the rays both sides carry through (a ray on a facet edge may be classified differently)
    g_np = rng.standard_normal((n, 8)).astype(f32) * both[:, None]
    g_out = torch.tensor(g_np, device="cuda")
    g_od = torch.empty((n, 6), dtype=torch.float32, device="cuda")
    g_V = torch.zeros_like(lens.vertices); g_N = torch.zeros_like(lens.vertices); g_eta = torch.zeros_like(lens.eta)
    capi.check(lib.lrt_scene_trace_refractive_backward(lens.scene.handle, C.c_int64(n), capi.dev_ptr(od_t), capi.dev_ptr(lens.eta), K, capi.dev_ptr(g_out),
                                                       capi.dev_ptr(g_od), capi.dev_ptr(g_V), capi.dev_ptr(g_N) if smooth else None, capi.dev_ptr(g_eta), None))
    torch.cuda.synchronize()
    go = g_out.cpu().double()
    m = torch.tensor(both)
    loss = ((o * go[:, 0:3]).sum(1) + (d * go[:, 3:6]).sum(1) + w * go[:, 7])[m].sum()
    loss.backward()
    # rays alive on one side only contribute to one side: drop them from the per-ray comparison, bound them in the sums
    def close(name, x, y, tol):
        x = x.cpu().double().numpy(); y = y.numpy()
        err = np.linalg.norm(x - y) / max(np.linalg.norm(y), 1e-30)
        print(f"smooth={smooth} {name}: |oracle| {np.linalg.norm(y):.4e} rel err {err:.2e}")
        assert err < tol, (name, err)
    close("d/d rays", g_od[m.cuda()], odc.grad[m], 2e-4)
    close("d/d vertices", g_V, V.grad, 5e-4)
    close("d/d eta", g_eta, eta.grad, 5e-4)
    if smooth:
        close("d/d normals", g_N, N.grad, 5e-4)


def test_mesh_lens_gradients_match_finite_differences_of_the_reference(mi):
    """The whole differentiable route of the product — table rows -> vertices (sag) -> vertex normals -> refit -> trace ->
    landing points, backward through lrt_scene_trace_refractive_backward and torch — against central finite differences of the
    REFERENCE's own forward on the same meshed singlet (tests/golden/ref_mesh_lens.npz), variable by variable: curvatures, a
    vertex position, a conic constant, an asphere coefficient."""
    import torch
    G = _golden()
    lens = mi.mesh_lens.MeshLens(mi, G["rows"], subdiv=3)
    rows = torch.tensor(G["rows"], dtype=torch.float64, device="cuda", requires_grad=True)
    od = G["rays"]
    po, pd, alive, w = lens.trace_ad(rows, od[:, :3], od[:, 3:])
    assert alive.cpu().numpy().tolist() == G["alive"].tolist()
    t = -po[:, 2] / pd[:, 2]
    pf = (po + t[:, None] * pd)[:, :2]
    assert np.abs(pf.detach().cpu().numpy() - G["landing"]).max() < 2e-7
    a = torch.tensor(np.random.default_rng(3).standard_normal((len(od), 2)), device="cuda")
    (pf.double() * a).sum().backward()
    fd = G["fd_landing"]
    for j, (ri, ci) in enumerate(G["var_index"]):
        want = float((fd[j] * a.cpu().numpy()).sum()); got = float(rows.grad[ri, ci])
        print(f"{G['variables'][j]}: product {got:.6e}  reference FD {want:.6e}")
        assert abs(got - want) <= 1e-2 * abs(want) + 1e-6, (G["variables"][j], got, want)
    # per-ray Jacobian columns for a few rays
    for ray in (0, 17, 55, 90):
        for c in range(2):
            rows.grad = None
            po, pd, alive, w = lens.trace_ad(rows, od[:, :3], od[:, 3:])
            pfr = (po[ray] - po[ray, 2] / pd[ray, 2] * pd[ray])[c]
            pfr.backward()
            for j, (ri, ci) in enumerate(G["var_index"]):
                assert abs(float(rows.grad[ri, ci]) - fd[j, ray, c]) <= 3e-2 * np.abs(fd[j]).max(), (G["variables"][j], ray, c)


def test_mesh_lens_gradient_agrees_with_the_analytic_lens_adjoint(mi):
    """A design loss differentiated through the MESHED doublet (subdivision 5: 2 x 5 000 triangles) equals, up to the meshing
    error, the same loss differentiated through the ANALYTIC interface table (central differences of the bit-exact lens trace):
    the two lens representations of the package give the same design gradients for curvatures and vertex positions."""
    import torch
    rows_np = doublet_rows(mi)
    src = mi.lens_design.field_source(2.0, 1.0e3, front_z=float(rows_np[-1, 1]))
    off = 2.0 * float(rows_np[-1, 2])
    lens = mi.mesh_lens.MeshLens(mi, rows_np, subdiv=5)
    rows = torch.tensor(rows_np, dtype=torch.float64, device="cuda", requires_grad=True)
    pf, alive = lens.spot_positions_ad(rows, src, 20000, seed=4, origin_offset=off)
    ap, d, o = lens._spot_rays(src, 20000, 4, off)
    R = float(rows_np[-1, 2])
    inner = alive & (ap[:, :2].norm(dim=1) < 0.9 * R)
    # a loss LINEAR in the landing points (defocus-like moment + centroid): its gradient does not pick up the facet noise of the
    # landing points themselves, only the meshing error of their derivatives
    def design_loss(pf, apxy, m):
        return ((pf[m] * apxy[m]).sum(1).mean() / R + pf[m][:, 0].mean())
    loss = design_loss(pf.double(), ap[:, :2].double(), inner)
    loss.backward()
    g_mesh = rows.grad.cpu().numpy()
    # the same rays through the analytic table, differentiated by central differences of the (float32, bit-exact) lens trace
    o_np, d_np, ap_np, in_np = o.cpu().numpy(), d.cpu().numpy(), ap.cpu().numpy(), inner.cpu().numpy()

    def analytic_fd(rp, rm):
        """central difference over the rays BOTH perturbed tables transmit (a ray that one of them vignettes would turn the
        difference quotient into the derivative of the mask)"""
        (qp, actp), (qm, actm) = (analytic_positions(mi, r.astype(f32), ap_np, d_np, off) for r in (rp, rm))
        m = in_np & actp & actm
        assert m.sum() > 0.99 * in_np.sum()
        val = lambda q: ((q[m].astype(np.float64) * ap_np[m, :2]).sum(1).mean() / R + q[m][:, 0].astype(np.float64).mean())  # noqa: E731
        return val(qp) - val(qm)
    groups = {"curvature": [(2, 3, 1e-3), (1, 3, 1e-3), (0, 3, 1e-3)], "vertex z": [(2, 1, 1e-4), (1, 1, 1e-4), (0, 1, 1e-4)]}
    for name, cols in groups.items():
        pairs = []
        for ri, ci, h in cols:
            rp, rm = rows_np.astype(np.float64).copy(), rows_np.astype(np.float64).copy()
            rp[ri, ci] += h; rm[ri, ci] -= h
            rp, rm = rp.astype(f32), rm.astype(f32)
            fdv = analytic_fd(rp, rm) / (float(rp[ri, ci]) - float(rm[ri, ci]))
            pairs.append((g_mesh[ri, ci], fdv))
            print(f"{name} row {ri}: mesh adjoint {g_mesh[ri, ci]:.5e}  analytic FD {fdv:.5e}")
        scale = max(abs(f) for _, f in pairs)
        for gm, fdv in pairs:
            assert abs(gm - fdv) <= 0.01 * scale, (name, gm, fdv)


def test_mesh_design_loop_refocuses_a_perturbed_doublet(mi):
    """MeshDesignProblem: parallel ray bundles (the notebooks' `directionalarea` field sources) through the meshed doublet,
    geometric rms spot loss, Adam on the three curvatures. Starting from a doublet whose rear curvature was changed
    by 6 %, the loop brings the spot size back down; its loss agrees with the same bundles traced through the analytic table."""
    import torch
    rows = doublet_rows(mi)
    bad = rows.copy(); bad[0, 3] *= 1.06
    prob = mi.mesh_lens.MeshDesignProblem(mi, bad, [0.0, 2.5], subdiv=5, rays=1 << 14, lr=1e-4, free=[(i, 3) for i in range(3)])
    l0 = float(prob.loss(prob.rows()).detach())
    # the same bundles through the analytic table: geometric rms about the centroid, mm^2
    want = 0.0
    for o, d in prob.bundles:
        o_np, d_np = o.cpu().numpy(), d.cpu().numpy()
        from mitsuba3_camera_b200.lens_designs import InterfaceList
        cam = mi.RealisticLensCamera(mi.Film(256, 256), "custom", prescription=InterfaceList(bad, film_z=0.0, fine_focus=False))
        po, pd, act = cam.trace(o_np, d_np, from_world=True)
        t = -po[:, 2] / pd[:, 2]
        q = (po + t[:, None] * pd)[act, :2].astype(np.float64) * 1e3
        want += ((q - q.mean(0)) ** 2).sum(1).mean()
    print(f"perturbed doublet: mesh loss {l0:.5e} mm^2, analytic {want:.5e} mm^2")
    assert abs(l0 - want) <= 0.03 * want
    losses = [prob.step() for _ in range(60)]
    print("loss", losses[0], "->", losses[-1], "min", min(losses))
    assert np.isfinite(losses).all() and np.mean(losses[-10:]) < 0.25 * losses[0]
    assert prob.theta.grad is not None and float(prob.theta.grad.abs().max()) > 0
    # moving the lens between a trace and its backward is refused (the adjoint kernel replays the trace)
    r = prob.rows().detach().clone().requires_grad_(True)
    po, pd, alive, w = prob.lens.trace_ad(r, *prob.bundles[0])
    prob.lens.trace_ad(r * 1.0001, *prob.bundles[0])
    with pytest.raises(RuntimeError):
        po.sum().backward()


def test_trace_adjoint_limits(mi):
    """edge cases of lrt_scene_trace_refractive_backward: no rays, too many interactions, rays that miss everything (the
    adjoint passes straight through: d/d rays = the normalisation's Jacobian), absorbed rays (nothing)."""
    import ctypes as C
    import torch
    capi, lib = mi.capi, mi.capi.lib
    rows = doublet_rows(mi)
    lens = mi.mesh_lens.MeshLens(mi, rows, subdiv=3)
    z = torch.zeros((4, 8), dtype=torch.float32, device="cuda")
    g_od = torch.full((4, 6), 7.0, dtype=torch.float32, device="cuda")
    g_V = torch.zeros_like(lens.vertices)
    od = torch.zeros((4, 6), dtype=torch.float32, device="cuda")
    assert lib.lrt_scene_trace_refractive_backward(lens.scene.handle, C.c_int64(0), None, None, C.c_int32(4), None, None, None, None, None, None) == 0
    rc = lib.lrt_scene_trace_refractive_backward(lens.scene.handle, C.c_int64(4), capi.dev_ptr(od), capi.dev_ptr(lens.eta), C.c_int32(33), capi.dev_ptr(z),
                                                 capi.dev_ptr(g_od), capi.dev_ptr(g_V), None, None, None)
    assert rc != 0 and b"at most 32" in lib.lrt_last_error()
    R, fz = float(rows[-1, 2]), float(rows[-1, 1])
    # ray 0 flies past everything (sideways, far from the lens); ray 1 hits the front baffle; rays 2, 3 go through the glass
    od_np = np.array([[1.0, 1.0, fz + 0.5, 0.0, 2.0, 0.0], [2.0 * R, 0.0, fz + 0.02, 0.0, 0.0, -1.0],
                      [0.2 * R, 0.0, fz + 0.02, 0.0, 0.0, -1.0], [0.0, 0.3 * R, fz + 0.02, 0.0, 0.0, -3.0]], f32)
    od = torch.tensor(od_np, device="cuda")
    out = torch.empty((4, 8), dtype=torch.float32, device="cuda")
    K = C.c_int32(lens.n_surfaces + 1)
    capi.check(lib.lrt_scene_trace_refractive(lens.scene.handle, C.c_int64(4), capi.dev_ptr(od), capi.dev_ptr(lens.eta), K, capi.dev_ptr(out), None))
    assert out[:, 6].cpu().numpy().tolist() == [1.0, 0.0, 1.0, 1.0]
    g = torch.zeros((4, 8), dtype=torch.float32, device="cuda"); g[:, 0:6] = 1.0
    capi.check(lib.lrt_scene_trace_refractive_backward(lens.scene.handle, C.c_int64(4), capi.dev_ptr(od), capi.dev_ptr(lens.eta), K, capi.dev_ptr(g),
                                                       capi.dev_ptr(g_od), capi.dev_ptr(g_V), None, None, None))
    got = g_od.cpu().numpy()
    assert np.allclose(got[0], [1, 1, 1, 0.5, 0.0, 0.5], atol=1e-6)          # (I - d d^T) / |d_in| applied to (1,1,1), d = +y, |d_in| = 2
    assert (got[1] == 0).all()                                                # absorbed by the baffle
    assert np.isfinite(got).all() and np.abs(got[2:]).max() > 0
    assert float(g_V.abs().max()) > 0


def test_mesh_design_loop_with_the_notebooks_image_loss(mi):
    """loss="image": the landing points splatted with the tent filter into a spot image (what the notebooks' perspective sensors
    see on the film plane) and rms_loss of that image (losses.py:8-28) — differentiable through the filter weights, the
    splat-position adjoint the notebooks' AD propagates. Interpolating onto pixel centres preserves the centroid and adds
    f (1 - f) px^2 per axis to the second moment (1/6 on average): the image loss is the geometric one plus pitch^2 / 3. The loop
    brings the perturbed doublet back with this loss as well."""
    import torch
    rows = doublet_rows(mi)
    bad = rows.copy(); bad[0, 3] *= 1.06
    free = [(i, 3) for i in range(3)]
    geo = mi.mesh_lens.MeshDesignProblem(mi, bad, [0.0, 2.5], subdiv=5, rays=1 << 14, lr=1e-4, free=free)
    prob = mi.mesh_lens.MeshDesignProblem(mi, bad, [0.0, 2.5], subdiv=5, rays=1 << 14, lr=1e-4, free=free, loss="image", resolution=96, pitch=1.5e-5)
    lg = float(geo.loss(geo.rows()).detach()); li = float(prob.loss(prob.rows()).detach())
    want = lg + 2 * (prob.pitch * 1e3) ** 2 / 3.0            # two fields
    print(f"geometric loss {lg:.5e} mm^2, image loss {li:.5e} mm^2 (geometric + pitch^2/3 per field: {want:.5e})")
    assert abs(li - want) <= 0.03 * want
    imgs = prob.images()
    assert imgs.shape == (2, 96, 96, 3) and float(imgs[0, ..., 0].sum()) > 0.99 * (1 << 14) * 0.9       # the window holds the spot
    g_geo = torch.autograd.grad(geo.loss(geo.rows()), geo.theta)[0]; g_img = torch.autograd.grad(prob.loss(prob.rows()), prob.theta)[0]
    cos = float((g_geo * g_img).sum() / (g_geo.norm() * g_img.norm()))
    print("gradient cosine image vs geometric", cos)
    assert cos > 0.99
    losses = [prob.step() for _ in range(60)]
    print("image loss", losses[0], "->", losses[-1])
    assert np.isfinite(losses).all() and np.mean(losses[-10:]) - 2 * (prob.pitch * 1e3) ** 2 / 3.0 < 0.3 * (losses[0] - 2 * (prob.pitch * 1e3) ** 2 / 3.0)


def test_mesh_lens_spot_image_agrees_with_the_particle_tracer_on_path_b(mi):
    """Two independent routes to the same picture: (1) the particle tracer on the notebooks' pipeline (path_b_scene: emitter rays of
    the `directionalarea` source -> dielectric BSDF sampling on the meshed caps -> diffuse film plane -> perspective sensor,
    pinned to the reference's render), (2) MeshLens: the same parallel bundle through lrt_scene_trace_refractive, the landing
    points splatted (splat_tent, Fresnel weights) at the sensor's pixel pitch on the film plane. Different samples, different
    kernels: the normalised images agree to the Monte Carlo noise."""
    import torch
    import ref_scenes as rs
    res, spp = 64, 256
    d = rs.path_b_scene(lambda m: mi.Transform4f(m), spp=spp, res=res)
    img_p = mi.render(mi.load_dict(d), seed=1).cpu().numpy()[..., 0].astype(np.float64)
    rows = mi.lens_design.asphere_table([30.0, -45.0], [4.0, 52.0], [8, 8], [1.5168, 1.000277]).astype(f32)       # the caps of the assets
    lens = mi.mesh_lens.MeshLens(mi, rows, subdiv=3)
    M = d["field_source"]["to_world"].matrix.astype(np.float64)
    g = torch.Generator(device="cuda").manual_seed(3)
    n = res * res * spp
    uv = (2.0 * torch.rand((n, 2), device="cuda", generator=g, dtype=torch.float64) - 1.0)
    Mt = torch.tensor(M, device="cuda")
    o = (Mt[:3, 3][None] + uv[:, :1] * Mt[:3, 0][None] + uv[:, 1:] * Mt[:3, 1][None]).float()
    nrm = torch.nn.functional.normalize(Mt[:3, 2], dim=0).float()
    po, pd, alive, w = lens.trace(o, nrm.expand(n, 3).contiguous())
    t = -po[:, 2] / pd[:, 2]
    pf = (po + t[:, None] * pd)[:, :2].double()
    pitch = 2.0 * 0.03 * np.tan(np.radians(15.0)) / res
    pts = torch.stack([pf[:, 0], -pf[:, 1]], 1)[alive]                     # image column ~ +x, row ~ -y (the sensor looks down -z, up = +y)
    img_s = mi.mesh_lens.splat_tent(pts, w[alive].double(), res, pitch, [0.0, 0.0])[..., 0].cpu().numpy()
    a, b = img_p / img_p.sum(), img_s / img_s.sum()
    lit = (a + b) > 0.2 * max(a.max(), b.max())
    err = np.sqrt(((a - b)[lit] ** 2).mean()) / a[lit].mean()
    corr = np.corrcoef(a.ravel(), b.ravel())[0, 1]
    yy, xx = np.mgrid[0:res, 0:res]
    ca, cb = np.array([(a * xx).sum(), (a * yy).sum()]), np.array([(b * xx).sum(), (b * yy).sum()])
    print(f"path B vs mesh-lens spot image: rel-RMSE over lit pixels {err:.3e}, correlation {corr:.5f}, centroids {ca} / {cb} px")
    assert corr > 0.995 and err < 0.06 and np.abs(ca - cb).max() < 0.05
