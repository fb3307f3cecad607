"""Integrator step. CPU: known-answer checks of the oracle's path tracer. GPU: BVH traversal and full renders
(pinhole and realistic lens) through the C ABI against the oracle at matched seeds/spp.

Stated image tolerance: relative RMSE <= 2e-3 of the developed image at matched spp and seed (the sample
streams are identical, so the residual is float rounding of traversal/shading plus the rare path whose
branch flips on a rounding difference; measured values are ~1e-5..1e-4)."""
import ctypes as C

import numpy as np
import pytest

from oracle_lib import *  # noqa

ld = load_lens_designs()
IMG_TOL = 2e-3


def cornell(mi, res=64, spp=16, sensor=None):
    d = mi.cornell_box()
    d["sensor"]["film"]["width"] = d["sensor"]["film"]["height"] = res
    d["sensor"]["sampler"]["sample_count"] = spp
    if sensor is not None:
        film, smp = d["sensor"]["film"], d["sensor"]["sampler"]
        d["sensor"] = dict(sensor, film=film, sampler=smp)
    return d


def scene_arrays(mi, d):
    """The flat arrays the product would upload (built without touching the GPU)."""
    return mi.scene_data_from_dict(d)[0].arrays()


# ------------------------------------------------------------------------------------------------ CPU
def test_cornell_scene_data(mi):
    a = scene_arrays(mi, mi.cornell_box())
    assert a["indices"].shape == (6 * 2 + 2 * 12, 3) and len(a["shape_bsdf"]) == 8
    assert len(a["emitters"]) == 1 and a["shape_emitter"].tolist().count(-1) == 7
    # the light faces down (rectangle normal = to_world * +z)
    e = a["emitters"][0]
    assert np.allclose(np.cross(e[7:10], e[10:13]) / np.linalg.norm(np.cross(e[7:10], e[10:13])), [0, -1, 0], atol=1e-6)


def test_oracle_intersect_vs_bruteforce(mi):
    rng = np.random.default_rng(3)
    nt = 500
    verts = rng.random((nt * 3, 3)).astype(f32)
    verts[1::3] = verts[0::3] + (rng.random((nt, 3)).astype(f32) - 0.5) * 0.2
    verts[2::3] = verts[0::3] + (rng.random((nt, 3)).astype(f32) - 0.5) * 0.2
    a = dict(vertices=verts, indices=np.arange(nt * 3, dtype=np.int32).reshape(-1, 3), tri_shape=np.zeros(nt, np.int32),
             shape_bsdf=np.zeros(1, np.int32), shape_emitter=-np.ones(1, np.int32),
             bsdfs=np.zeros((1, 16), f32), emitters=np.zeros((0, 16), f32))
    h = orc_scene(a)
    n = 2000
    o = (rng.random((n, 3)).astype(f32) - 0.5) * 3 + 0.5
    d = rng.standard_normal((n, 3)).astype(f32); d /= np.linalg.norm(d, axis=1, keepdims=True)
    got = orc_intersect(h, o, d, np.float32(3.4e38))
    # numpy Moeller-Trumbore in float64 over all triangles
    p0, p1, p2 = (verts[i::3].astype(np.float64) for i in range(3))
    e1, e2 = p1 - p0, p2 - p0
    best = np.full(n, np.inf)
    for i in range(n):
        pv = np.cross(d[i].astype(np.float64), e2); det = (e1 * pv).sum(1)
        tv = o[i].astype(np.float64) - p0; u = (tv * pv).sum(1) / det
        qv = np.cross(tv, e1); v = (qv * d[i].astype(np.float64)).sum(1) / det; t = (e2 * qv).sum(1) / det
        ok = (u >= 0) & (v >= 0) & (u + v <= 1) & (t >= 0)
        if ok.any():
            best[i] = t[ok].min()
    hit = np.isfinite(best)
    assert ((got[:, 3] >= 0) == hit).mean() > 0.999          # grazing-edge rays may differ between f32 and f64
    both = hit & (got[:, 3] >= 0)
    assert np.allclose(got[both, 0], best[both], rtol=1e-4, atol=1e-5)


def test_oracle_emitter_seen_directly(mi):
    # a camera staring at an area light sees exactly its radiance (path.cpp:158-174 with prev_bsdf_delta = true)
    T = mi.Transform4f
    d = {"type": "scene",
         "light": {"type": "rectangle", "to_world": T.translate([0, 0, 0]).scale(10.0), "bsdf": {"type": "diffuse"},
                   "emitter": {"type": "area", "radiance": {"type": "rgb", "value": [2.0, 3.0, 5.0]}}}}
    h = orc_scene(scene_arrays(mi, d))
    tw = T.look_at(origin=[0, 0, 4], target=[0, 0, 0], up=[0, 1, 0]).matrix
    st, _ = orc_render(h, FilmSpec(8, 8), 4, fov=30.0, to_world=tw)
    img = develop(st)
    assert np.allclose(img, [2.0, 3.0, 5.0], rtol=1e-5)


def test_oracle_cornell_statistics(mi):
    """Energy sanity of the CPU path tracer: the Cornell box is brighter under the light than in the corners,
    red/green walls bleed, and two seeds agree within Monte Carlo noise."""
    a = scene_arrays(mi, mi.cornell_box())
    h = orc_scene(a)
    tw = mi.cornell_box()["sensor"]["to_world"].matrix
    imgs = [develop(orc_render(h, FilmSpec(32, 32), 64, seed=s, fov=39.3077, to_world=tw, near=1e-3, far=100.0)[0]) for s in (0, 1)]
    m = imgs[0]
    assert 0.1 < m.mean() < 1.5
    assert m[:, :4, 0].mean() > 1.5 * m[:, :4, 1].mean()       # left wall is red
    assert m[:, -4:, 1].mean() > 1.5 * m[:, -4:, 0].mean()     # right wall is green
    assert rel_rmse(imgs[0], imgs[1]) < 0.35


# ------------------------------------------------------------------------------------------------ GPU
@pytest.mark.gpu
@pytest.mark.parametrize("nt", [1, 3, 200, 20000])
def test_gpu_bvh_vs_oracle(mi, nt):
    import torch
    rng = np.random.default_rng(nt)
    verts = rng.random((nt * 3, 3)).astype(f32)
    s = 0.5 / max(nt, 1) ** (1 / 3)
    verts[1::3] = verts[0::3] + (rng.random((nt, 3)).astype(f32) - 0.5) * s * 2
    verts[2::3] = verts[0::3] + (rng.random((nt, 3)).astype(f32) - 0.5) * s * 2
    faces = np.arange(nt * 3, dtype=np.int32).reshape(-1, 3)
    d = {"type": "scene", "soup": {"type": "mesh", "vertices": verts, "faces": faces, "bsdf": {"type": "diffuse"}}}
    sc = mi.load_dict(d)
    h = orc_scene(sc.arrays)
    n = 100000
    o = (rng.random((n, 3)).astype(f32) - 0.5) * 2.5 + 0.5
    dd = (0.5 + (rng.random((n, 3)) - 0.5) * 0.8 - o).astype(f32); dd /= np.linalg.norm(dd, axis=1, keepdims=True)
    maxt = np.where(rng.random(n) < 0.3, rng.random(n) * 2, 3.4e38).astype(f32)
    want = orc_intersect(h, o, dd, maxt)
    t, u, v, prim = (x.cpu().numpy() for x in sc.ray_intersect(o, dd, maxt))
    hit_w, hit_g = want[:, 3] >= 0, prim >= 0
    assert (hit_w == hit_g).mean() > 0.9995
    both = hit_w & hit_g
    assert np.allclose(t[both], want[both, 0], rtol=2e-5, atol=1e-6)
    same = prim[both] == want[both, 3].astype(np.int64)
    assert same.mean() > 0.999
    # barycentrics of small far-away triangles amplify float rounding by distance/size: compare loosely
    uv_tol = 2e-4 * max(1.0, nt ** (1 / 3))
    assert np.allclose(u[both][same], want[both, 1][same], atol=uv_tol) and np.allclose(v[both][same], want[both, 2][same], atol=uv_tol)
    # any-hit (ray_test) agrees with closest-hit on occlusion
    _, _, _, prim_any = (x.cpu().numpy() for x in sc.ray_intersect(o, dd, maxt, any_hit=True))
    assert ((prim_any >= 0) == hit_g).all()
    info = sc.info()
    assert info["triangles"] == nt


@pytest.mark.gpu
def test_gpu_cornell_box_vs_oracle(mi):
    """BASELINE config 1 shape (Cornell box, path, max_depth as given) at reduced size, pinhole sensor."""
    d = cornell(mi, res=64, spp=16)
    d["integrator"]["max_depth"] = 6
    sc = mi.load_dict(d)
    img = mi.render(sc, seed=0).cpu().numpy()
    st, _ = orc_render(orc_scene(sc.arrays), FilmSpec(64, 64), 16, seed=0, max_depth=6, fov=39.3077,
                       to_world=d["sensor"]["to_world"].matrix, near=1e-3, far=100.0)
    want = develop(st)
    err = rel_rmse(img, want)
    assert err < IMG_TOL, err
    assert sc.stats()["samples"] == 64 * 64 * 16


@pytest.mark.gpu
@pytest.mark.parametrize("design", ["doublet", "gauss", "exp1a"])
@pytest.mark.parametrize("rfilter", ["gaussian", "tent", "box"])
def test_gpu_cornell_through_realistic_lens(mi, design, rfilter):
    T = mi.Transform4f
    tw = T.look_at(origin=[0, 0, 3.9], target=[0, 0, 0], up=[0, 1, 0])
    sensor = {"type": "realisticlens", "lens_design": design, "to_world": tw, "object_distance": 3.9, "near_clip": 1e-3, "far_clip": 100.0}
    d = cornell(mi, res=48, spp=8, sensor=sensor)
    d["sensor"]["film"]["rfilter"] = {"type": rfilter}
    d["sensor"]["film"]["diagonal"] = 30.0
    d["integrator"]["max_depth"] = 5
    sc = mi.load_dict(d)
    img = mi.render(sc, seed=3).cpu().numpy()
    cam = orc_camera(ld.design(design, {"object_distance": 3.9}), FilmSpec(48, 48, None, 30.0), to_world=tw.matrix, near=1e-3, far=100.0,
                     object_distance=3.9)
    st, miss = orc_render(orc_scene(sc.arrays), FilmSpec(48, 48, None, 30.0), 8, seed=3, max_depth=5, cam_h=cam,
                          filter_kind={"box": 0, "tent": 1, "gaussian": 2}[rfilter])
    want = develop(st)
    assert sc.stats()["lens_misses"] == miss, "lens ray-miss count differs from the oracle"
    err = rel_rmse(img, want)
    assert want.mean() > 1e-3 and err < IMG_TOL, err


@pytest.mark.gpu
def test_gpu_cornell_exit_pupil_mode(mi):
    """sample_exit_pupil = true with an open shutter: LUT-guided aperture sampling, weight shutter * area / z^2, and the
    extra time draw of render_sample (integrator.cpp:442-444) -- image and miss count vs the oracle in the same mode."""
    T = mi.Transform4f
    tw = T.look_at(origin=[0, 0, 3.9], target=[0, 0, 0], up=[0, 1, 0])
    sensor = {"type": "realisticlens", "lens_design": "gauss", "to_world": tw, "object_distance": 3.9, "near_clip": 1e-3, "far_clip": 100.0,
              "sample_exit_pupil": True, "shutter_open": 0.0, "shutter_close": 2.0}
    d = cornell(mi, res=48, spp=8, sensor=sensor)
    d["sensor"]["film"]["diagonal"] = 30.0
    d["integrator"]["max_depth"] = 5
    sc = mi.load_dict(d)
    img = mi.render(sc, seed=5).cpu().numpy()
    cam = orc_camera(ld.design("gauss", {"object_distance": 3.9}), FilmSpec(48, 48, None, 30.0), to_world=tw.matrix, near=1e-3, far=100.0,
                     object_distance=3.9)
    lut = cam_exit_pupil(oracle(), "orc", cam, shutter_open_time=2.0)
    assert (bits(sc.sensors()[0].exit_pupil_bounds()) == bits(lut)).all()
    st, miss = orc_render(orc_scene(sc.arrays), FilmSpec(48, 48, None, 30.0), 8, seed=5, max_depth=5, cam_h=cam, filter_kind=2)
    want = develop(st)
    assert sc.stats()["lens_misses"] == miss, "lens ray-miss count differs from the oracle"
    err = rel_rmse(img, want)
    assert want.mean() > 1e-6 and err < IMG_TOL, err


@pytest.mark.gpu
def test_gpu_sharding_and_chunking_invariance(mi):
    """Samples keep their PCG32 stream under any tile/chunk/shard split: images agree to float summation order."""
    import torch
    d = cornell(mi, res=40, spp=12)
    sc = mi.load_dict(d)
    full = mi.render(sc, seed=1, develop=False)
    parts = torch.zeros_like(full)
    for r in range(3):
        mi.render(sc, seed=1, develop=False, shard=(r, 3), storage=parts)
    assert torch.allclose(full, parts, rtol=1e-5, atol=1e-6)
    assert sc.stats()["samples"] <= 40 * 40 * 12   # the last call rendered one shard only


@pytest.mark.gpu
def test_gpu_conductor_and_dielectric(mi):
    T = mi.Transform4f
    d = cornell(mi, res=48, spp=16)
    d["small-box"]["bsdf"] = {"type": "conductor", "eta": [0.2, 0.92, 1.1], "k": [3.9, 2.45, 2.14]}
    d["large-box"]["bsdf"] = {"type": "dielectric", "int_ior": 1.5, "ext_ior": 1.0}
    d["integrator"]["max_depth"] = 7
    sc = mi.load_dict(d)
    img = mi.render(sc, seed=2).cpu().numpy()
    st, _ = orc_render(orc_scene(sc.arrays), FilmSpec(48, 48), 16, seed=2, max_depth=7, fov=39.3077,
                       to_world=d["sensor"]["to_world"].matrix, near=1e-3, far=100.0)
    err = rel_rmse(img, develop(st))
    assert err < 5 * IMG_TOL, err    # delta lobes amplify rounding differences of the reflected direction


@pytest.mark.gpu
def test_gpu_render_backward_vs_oracle_finite_differences(mi):
    """lrt_render_backward (reflectance / specular reflectance / radiance gradients) against central finite differences of
    the ORACLE's render in float64 accumulators. max_depth = 4 < rr_depth keeps Russian roulette out of the way (prb detaches
    it; a finite difference of the primal would not). Tolerance: cosine >= 0.999, per entry 3 % of the largest."""
    import torch
    d = cornell(mi, res=32, spp=16)
    d["integrator"]["max_depth"] = 4
    d["small-box"]["bsdf"] = {"type": "conductor", "material": "Cu", "specular_reflectance": {"type": "rgb", "value": [0.9, 0.8, 0.7]}}
    sc = mi.load_dict(d)
    tw = d["sensor"]["to_world"].matrix
    rng = np.random.default_rng(1)
    G = np.zeros((32, 32, 4), f32); G[..., :3] = rng.standard_normal((32, 32, 3))

    def phi(arrays):
        st, _ = orc_render(orc_scene(arrays), FilmSpec(32, 32), 16, seed=5, max_depth=4, fov=39.3077, to_world=tw, near=1e-3, far=100.0)
        return float((st * G).sum())

    gb, ge = mi.render_backward(sc, torch.from_numpy(G).cuda(), seed=5, spp=16)
    gb, ge = gb.cpu().numpy(), ge.cpu().numpy()
    got, want = [], []
    base = {k: v.copy() for k, v in sc.arrays.items()}
    for b in range(len(base["bsdfs"])):
        if not np.any(base["tri_shape"][:, None] == np.nonzero(base["shape_bsdf"] == b)[0][None, :]):
            continue
        for c in range(3):
            h = 0.02
            ap = {k: v.copy() for k, v in base.items()}; am = {k: v.copy() for k, v in base.items()}
            ap["bsdfs"][b, 2 + c] += h; am["bsdfs"][b, 2 + c] -= h
            want.append((phi(ap) - phi(am)) / (2 * h)); got.append(gb[b, 2 + c])
    for e in range(len(base["emitters"])):
        for c in range(3):
            h = 0.5
            ap = {k: v.copy() for k, v in base.items()}; am = {k: v.copy() for k, v in base.items()}
            ap["emitters"][e, c] += h; am["emitters"][e, c] -= h
            want.append((phi(ap) - phi(am)) / (2 * h)); got.append(ge[e, c])
    got, want = np.array(got, np.float64), np.array(want, np.float64)
    cos = float(got @ want / (np.linalg.norm(got) * np.linalg.norm(want)))
    assert cos >= 0.999, (cos, got, want)
    assert np.abs(got - want).max() <= 3e-2 * np.abs(want).max(), (got, want)
    # the traverse() view: keys, update() round trip, gradients() split
    params = mi.traverse(sc)
    assert "white.reflectance.value" in params and "light.emitter.radiance.value" in params
    g = params.gradients(torch.from_numpy(gb), torch.from_numpy(ge))
    assert g["red.reflectance.value"].shape == (3,)
    img0 = mi.render(sc, seed=1, spp=4).cpu().numpy()
    params["light.emitter.radiance.value"] = params["light.emitter.radiance.value"] * 2
    params.update()
    img1 = mi.render(sc, seed=1, spp=4).cpu().numpy()
    assert rel_rmse(img1, 2 * img0) < 1e-5


# ------------------------------------------------------------------------------------------------ edge cases
@pytest.mark.gpu
@pytest.mark.parametrize("size,crop,spp", [((37, 19), None, 5), ((70, 33), (41, 17, 13, 9), 3), ((33, 1), None, 16), ((1, 45), None, 2)])
def test_gpu_ragged_films_and_crops_vs_oracle(mi, size, crop, spp):
    """Film sizes that are not multiples of the 32-pixel row tile, crop windows with odd offsets, odd sample counts
    (spp < 4 also switches ImageBlock::put to its non-coalesced branch, integrator.cpp:290)."""
    w, h = size
    d = cornell(mi, res=w, spp=spp)
    d["sensor"]["film"]["height"] = h
    fs = FilmSpec(w, h, list(crop) if crop else None)
    if crop:
        d["sensor"]["film"].update(crop_width=crop[0], crop_height=crop[1], crop_offset_x=crop[2], crop_offset_y=crop[3])
    d["integrator"]["max_depth"] = 4
    sc = mi.load_dict(d)
    img = mi.render(sc, seed=9).cpu().numpy()
    cw, ch = (crop[0], crop[1]) if crop else (w, h)
    assert img.shape == (ch, cw, 3)
    # the Cornell sensor's fov is given along the smaller axis (fov_axis = "smaller"); the oracle takes the x fov
    st, _ = orc_render(orc_scene(sc.arrays), fs, spp, seed=9, max_depth=4, fov=float(sc.sensors()[0].x_fov),
                       to_world=d["sensor"]["to_world"].matrix, near=1e-3, far=100.0)
    want = develop(st)
    assert sc.stats()["samples"] == cw * ch * spp
    assert rel_rmse(img, want) < IMG_TOL


@pytest.mark.gpu
def test_gpu_empty_scene_and_empty_batches(mi):
    """No triangles at all: every path misses, RGB stays 0 and the weight channel still receives the filter weights;
    zero-length batches of the unit entry points are accepted."""
    import torch
    d = {"type": "scene", "integrator": {"type": "path", "max_depth": 4},
         "sensor": {"type": "realisticlens", "lens_design": "doublet", "film": {"type": "hdrfilm", "width": 40, "height": 24},
                    "sampler": {"type": "independent", "sample_count": 4}}}
    sc = mi.load_dict(d)
    st = mi.render(sc, seed=1, develop=False).cpu().numpy()
    assert np.all(st[..., :3] == 0) and st[..., 3].sum() > 0
    stats = sc.stats()
    assert stats["samples"] == 40 * 24 * 4 and stats["triangles"] == 0
    s = sc.sensors()[0]
    o, dd, maxt, w = s.sample_ray(0.0, 0.0, np.zeros((0, 2), f32), np.zeros((0, 2), f32))
    assert len(o) == 0 and len(w) == 0
    out = torch.zeros((0, 4), device="cuda")
    mi.capi.check(mi.capi.lib.lrt_scene_intersect(sc.handle, C.c_int32(0), C.c_int64(0), mi.capi.dev_ptr(out), mi.capi.dev_ptr(out), None))


@pytest.mark.gpu
@pytest.mark.parametrize("rf,kind,param", [({"type": "gaussian", "stddev": 0.8}, 2, 0.8), ({"type": "tent", "radius": 3.0}, 1, 3.0)])
def test_gpu_wide_reconstruction_filters(mi, rf, kind, param):
    """Filters wider than the defaults (gaussian stddev 0.8 -> radius 3.2, tent radius 3: 7 taps per axis, the build's
    maximum is 8): the private film tiles grow with the footprint and the image still matches the oracle. Anything
    wider is refused with an error, not rendered wrongly."""
    d0 = cornell(mi, res=32, spp=4)
    d0["sensor"]["film"]["rfilter"] = {"type": "gaussian", "stddev": 1.25}
    with pytest.raises(mi.capi.LensrtError):
        mi.render(mi.load_dict(d0), seed=0)
    d = cornell(mi, res=56, spp=8)
    d["sensor"]["film"]["rfilter"] = rf
    d["integrator"]["max_depth"] = 4
    sc = mi.load_dict(d)
    img = mi.render(sc, seed=4).cpu().numpy()
    st, _ = orc_render(orc_scene(sc.arrays), FilmSpec(56, 56), 8, seed=4, max_depth=4, fov=39.3077,
                       to_world=d["sensor"]["to_world"].matrix, near=1e-3, far=100.0, filter_kind=kind, filter_param=param)
    assert rel_rmse(img, develop(st)) < IMG_TOL
