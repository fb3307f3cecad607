"""GPU particle tracer (lrt_ptrace, csrc/lrt_ptrace.cu) on the XML scenes of tests/golden/assets/xml seen through a realistic lens.

  * default mode (one PCG32 stream per sample) vs the PINNED oracle's JIT-convention render (oracle/oracle_ptracer_impl.h, which
    equals the reference's own images to 7e-8 in the scalar convention, tests/test_ptracer_parity.py) at matched seed / spp:
    rel-RMSE <= 2e-3 [measured 2e-5 .. 3e-4] — every splat is a single sample landing in a few pixels, so a path that flips a branch under the product's
    fused arithmetic moves whole splats; the image mean must agree to 2e-3;
  * rng_mode 1 (the scalar AdjointIntegrator's ranges, n_threads from the golden file) vs the reference's images DIRECTLY: one
    stream feeds 1/32 of all samples, so a single flipped branch rewrites the rest of its range — stated bar: >= 95 % of the
    pixels within 1e-4 relative [measured 97.8 .. 99.9 %] and the mean within 1 %;
  * conservation between shards: two half renders (shard 0/2 + 1/2) sum to the full render."""
import os

import numpy as np
import pytest

from oracle_lib import *  # noqa
from test_ptracer_parity import XML, CASES, case_key, oracle_ptracer, P  # noqa: F401  (P is a fixture)


@pytest.mark.gpu
@pytest.mark.parametrize("name,kw", CASES)
def test_gpu_ptracer_default_mode_vs_pinned_oracle(mi, name, kw):
    d = mi.xml_to_dict(path=os.path.join(XML, name + ".xml"), **kw)
    sc = mi.load_dict(d)
    assert isinstance(sc.integrator(), mi.PTracerIntegrator)
    img = mi.render(sc, seed=9).cpu().numpy()
    want = oracle_ptracer(mi, d, 9, 8, scalar=False)
    err = rel_rmse(img, want)
    st = sc.stats()
    print(f"{name} {kw}: rel-RMSE vs oracle {err:.3e}; light paths {st['samples']}, connections {st['lens_misses']}, rays {st['rays']}")
    assert img.shape == want.shape and np.isfinite(img).all()
    assert err <= 2e-3, err
    assert abs(img.mean() / want.mean() - 1) <= 2e-3


@pytest.mark.gpu
@pytest.mark.parametrize("name,kw", CASES)
def test_gpu_ptracer_scalar_convention_vs_reference_images(mi, P, name, kw):
    sc = mi.load_file(os.path.join(XML, name + ".xml"), **kw)
    img = sc.integrator().render(sc, seed=0, rng_mode=1, n_threads=int(P["n_threads"])).cpu().numpy()
    ref = P[case_key(name, kw)]
    close = np.abs(img - ref).max(-1) <= 1e-4 * np.maximum(ref.max(-1), 1e-3)
    print(f"{name} {kw}: rel-RMSE vs the reference's image {rel_rmse(img, ref):.3e}, pixels within 1e-4: {close.mean():.4f}")
    assert close.mean() >= 0.95
    assert abs(img.mean() / ref.mean() - 1) <= 1e-2


@pytest.mark.gpu
def test_gpu_ptracer_shards_sum_to_the_whole(mi):
    sc = mi.load_file(os.path.join(XML, "lens_room.xml"))
    it = sc.integrator()
    full = it.render(sc, seed=4)
    halves = it.render(sc, seed=4, shard=(0, 2)) + it.render(sc, seed=4, shard=(1, 2))
    assert rel_rmse(halves.cpu().numpy(), full.cpu().numpy()) <= 1e-5


@pytest.mark.gpu
def test_gpu_ptracer_launches_of_one_scene_may_overlap(mi):
    """two renders of ONE scene issued on two streams (their kernels may run at the same time): each has its own work counter
    (a slot of the scene's per-launch ring), so each image equals the one rendered alone"""
    import torch
    sc = mi.load_dict(triangle_scene(mi))          # (no spheres: the queue kernel and its chunk counter)
    it = sc.integrator()
    alone = [it.render(sc, seed=s).cpu().numpy() for s in (7, 8)]
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    out = [None, None]
    torch.cuda.synchronize()
    for rep in range(3):
        for k, s in enumerate((7, 8)):
            with torch.cuda.stream(streams[k]):
                out[k] = it.render(sc, seed=s)
        torch.cuda.synchronize()
        for k in range(2):
            assert rel_rmse(out[k].cpu().numpy(), alone[k]) <= 1e-5, (rep, k)


def triangle_scene(mi, lens="gauss", spp=16, res=96):
    """a triangle-only scene: the Cornell box with a copper box, a glass box and
    two emitters (tests/ref_scenes.py cbox_materials: conductor, dielectric, twosided, roulette active) behind a realistic lens"""
    import ref_scenes as rs
    d = dict(rs.scene_dicts(lambda m: mi.Transform4f(m))["cbox_materials"][0])
    cam = d["sensor"]["to_world"]
    d["integrator"] = {"type": "ptracer", "max_depth": 10, "rr_depth": 3}
    d["sensor"] = {"type": "realisticlens", "lens_design": lens, "object_distance": 3.9, "to_world": cam, "near_clip": 1e-3, "far_clip": 100.0,
                   "sampler": {"type": "independent", "sample_count": spp},
                   "film": {"type": "hdrfilm", "width": res, "height": res, "rfilter": {"type": "tent"}}}
    return d


@pytest.mark.gpu
@pytest.mark.parametrize("lens", ["gauss", "doublet"])
def test_gpu_ptracer_queue_kernel_vs_pinned_oracle(mi, lens):
    """the default forward kernel of triangle scenes (k_ptrace_queue) against the pinned oracle, JIT convention, same seed"""
    d = triangle_scene(mi, lens)
    sc = mi.load_dict(d)
    img = mi.render(sc, seed=5).cpu().numpy()
    want = oracle_ptracer(mi, d, 5, 8, scalar=False)
    err = rel_rmse(img, want)
    st = sc.stats()
    print(f"triangle scene, {lens}: rel-RMSE vs oracle {err:.3e}; light paths {st['samples']}, connections {st['lens_misses']}, rays {st['rays']}")
    assert np.isfinite(img).all() and img.mean() > 0
    assert err <= 2e-3, err
    assert abs(img.mean() / want.mean() - 1) <= 2e-3


@pytest.mark.gpu
def test_gpu_ptracer_queue_kernel_equals_per_thread_kernel(mi):
    """the queue kernel (persistent warps, per-lane path state machines) draws the same samples per path as the
    one-thread-per-path kernel it replaced (still behind LRT_PTRACE_SIMPLE / the adjoint / the scalar convention):
    same image up to the order of the float atomics, same path / ray / connection counts"""
    import subprocess
    import sys
    import tempfile
    code = ("import os, sys, numpy as np; sys.path.insert(0, %r); sys.path.insert(0, %r); import lensrt as mi\n"
            "from test_gpu_ptracer import triangle_scene\n"
            "sc = mi.load_dict(triangle_scene(mi)); img = mi.render(sc, seed=3).cpu().numpy()\n"
            "st = sc.stats(); G = np.random.default_rng(1).standard_normal(img.shape).astype(np.float32)\n"
            "rows = sc.integrator().render_backward_lens(sc, G, seed=3).cpu().numpy(); st2 = sc.stats()\n"
            "np.savez(sys.argv[1], img=img, rows=rows, st=np.array([st['samples'], st['rays'], st['lens_misses']]),\n"
            "         st2=np.array([st2['samples'], st2['rays'], st2['lens_misses']]))\n"
            % (ROOT, os.path.join(ROOT, "tests")))
    res = []
    with tempfile.TemporaryDirectory() as tmp:
        for k, simple in enumerate(("0", "1")):
            fn = os.path.join(tmp, f"r{k}.npz")
            env = dict(os.environ, LRT_PTRACE_SIMPLE=simple)
            subprocess.run([sys.executable, "-c", code, fn], check=True, env=env, timeout=300)
            res.append(dict(np.load(fn)))
    assert (res[0]["st"] == res[1]["st"]).all(), (res[0]["st"], res[1]["st"])
    assert res[0]["st"][0] > 0 and rel_rmse(res[0]["img"], res[1]["img"]) <= 1e-5
    # ... and the lens adjoint (k_ptrace_queue<true> vs k_ptrace<true>): same connections, same gradient of the interface table
    assert (res[0]["st2"] == res[1]["st2"]).all(), (res[0]["st2"], res[1]["st2"])
    g0, g1 = res[0]["rows"].astype(np.float64), res[1]["rows"].astype(np.float64)
    assert np.abs(g1).max() > 0 and np.abs(g0 - g1).max() <= 1e-4 * np.abs(g1).max(), (np.abs(g0 - g1).max(), np.abs(g1).max())


@pytest.mark.gpu
def test_gpu_ptracer_through_a_perspective_sensor(mi):
    """the particle tracer connects to a pinhole as well (PerspectiveCamera::sample_direction, perspective.cpp:283-386): the room
    scene (spheres, mirror, glass) through its own perspective sensor, default mode against the pinned oracle"""
    d = mi.xml_to_dict(path=os.path.join(XML, "room.xml"), integrator="ptracer")
    sc = mi.load_dict(d)
    img = mi.render(sc, seed=2).cpu().numpy()
    want = oracle_ptracer(mi, d, 2, 8, scalar=False)
    err = rel_rmse(img, want)
    print(f"room.xml, perspective sensor, ptracer: rel-RMSE vs oracle {err:.3e}, mean {img.mean():.5f}")
    assert np.isfinite(img).all() and img.mean() > 0 and err <= 2e-3


@pytest.mark.gpu
def test_gpu_ptracer_lens_adjoint_equals_spot_adjoint_for_a_point_like_emitter(mi):
    """lrt_ptrace_backward: d<G, image>/d(interface table). With max_depth = 1 (directly visible emitters only) and a tiny
    rectangle light, the particle tracer IS the spot path of a point source of intensity radiance x area x cos(theta_emitter) at
    the light's position — whose adjoint is pinned (float64 dual numbers + finite differences, tests/test_gpu_spot.py). Both are
    Monte Carlo estimates over different streams (2^20 samples each): stated tolerance cosine >= 0.999, per entry 3 % of the largest.
    The forward images must agree the same way (rel-RMSE <= 2 %)."""
    import torch
    T = mi.Transform4f
    P = np.array([0.12, -0.08, 2.0], f32); half = 2e-4
    radiance = np.array([3.0, 2.0, 1.0], f32)
    d = {"type": "scene", "integrator": {"type": "ptracer", "max_depth": 1},
         "sensor": {"type": "realisticlens", "lens_design": "gauss", "object_distance": 2.0,
                    "film": {"type": "hdrfilm", "width": 32, "height": 32, "rfilter": {"type": "tent"}},
                    "sampler": {"type": "independent", "sample_count": 1024}},
         "light": {"type": "rectangle", "to_world": T.translate(P).rotate([1, 0, 0], 180).scale(half),
                   "emitter": {"type": "area", "radiance": {"type": "rgb", "value": radiance.tolist()}}}}
    sc = mi.load_dict(d)
    it, s = sc.integrator(), sc.sensors()[0]
    H, W = s.film.crop_h, s.film.crop_w
    yy, xx = np.meshgrid(np.arange(H), np.arange(W), indexing="ij")
    G = np.stack([1 + .5 * np.cos(2 * np.pi * xx / W), 1 + .5 * np.sin(2 * np.pi * yy / H), 1 + .5 * np.cos(2 * np.pi * (xx + yy) / W)], -1).astype(f32)
    Gt = torch.as_tensor(G, device="cuda")
    front_z = float(s.interfaces()[-1, 1])
    rel = P - np.array([0, 0, front_z], f32)
    intensity = radiance * (2 * half) ** 2 * (rel[2] / np.linalg.norm(rel))
    img_pt = it.render(sc, seed=5).cpu().numpy()
    img_sp = mi.spot_render(s, P, intensity=tuple(intensity), seed=7, spp=1024)[..., :3].cpu().numpy()
    assert img_pt.sum() > 0 and rel_rmse(img_pt, img_sp) <= 2e-2, rel_rmse(img_pt, img_sp)
    g_pt = it.render_backward_lens(sc, Gt, seed=5).cpu().numpy()
    gs = torch.zeros((H, W, 4), device="cuda"); gs[..., :3] = Gt
    g_sp = mi.spot_render_backward(s, P, gs, intensity=tuple(intensity), seed=7, spp=1024).cpu().numpy()
    rows = s.interfaces()
    scale = np.ones(16, f32); scale[1] = 1e-4; scale[3] = 1e-3; scale[13] = scale[14] = 1e-3    # first-order loss changes for typical steps
    a, b = (g_pt * scale)[:, [1, 3, 13, 14]].ravel(), (g_sp * scale)[:, [1, 3, 13, 14]].ravel()
    cos = float(a @ b / (np.linalg.norm(a) * np.linalg.norm(b)))
    print("ptracer adjoint", a); print("spot adjoint   ", b); print("cosine", cos, "forward rel-RMSE", rel_rmse(img_pt, img_sp))
    assert np.abs(b).max() > 0 and len(rows) == 11
    assert cos >= 0.999
    assert np.abs(a - b).max() <= 0.03 * np.abs(b).max()


# ------------------------------------------------------------------------------------------------ the notebooks' "Path B"
@pytest.mark.gpu
def test_gpu_path_b_mesh_lens_directional_source_perspective_sensor(mi):
    """The notebooks' own pipeline in miniature (tests/ref_scenes.py path_b_scene; design_problem.py:46-120,355-428): a MESH lens
    (dielectric caps with smooth normals), a `directionalarea` field source, a diffuse film plane, a PERSPECTIVE sensor looking
    at it, the particle tracer. Default mode against the pinned oracle (same seed); the scalar convention against the
    reference's own renders of the same scene dictionary."""
    import ref_scenes as rs
    R = np.load(os.path.join(ROOT, "tests", "golden", "ref_pathb.npz"))
    d = rs.path_b_scene(lambda m: mi.Transform4f(m))
    sc = mi.load_dict(d)
    assert isinstance(sc.integrator(), mi.PTracerIntegrator) and isinstance(sc.sensors()[0], mi.PerspectiveCamera)
    img = mi.render(sc, seed=5).cpu().numpy()
    want = oracle_ptracer(mi, d, 5, 8, scalar=False)
    err = rel_rmse(img, want)
    st = sc.stats()
    print(f"path B, default mode: rel-RMSE vs oracle {err:.3e}; light paths {st['samples']}, connections {st['lens_misses']}, rays {st['rays']}")
    assert err <= 2e-3 and abs(img.mean() / want.mean() - 1) <= 2e-3
    for seed in (0, 3):
        img = sc.integrator().render(sc, seed=seed, rng_mode=1, n_threads=int(R["n_threads"])).cpu().numpy()
        ref = R[f"image_seed{seed}"]
        close = np.abs(img - ref).max(-1) <= 1e-4 * np.maximum(ref.max(-1), 1e-3)
        print(f"path B, scalar convention, seed {seed}: rel-RMSE vs the reference's image {rel_rmse(img, ref):.3e}, pixels within 1e-4: {close.mean():.4f}")
        assert close.mean() >= 0.95 and abs(img.mean() / ref.mean() - 1) <= 1e-2
    # `directionalarea` emitters exist for the particle tracer only
    d2 = dict(d); d2["integrator"] = {"type": "path", "max_depth": 4}
    with pytest.raises(mi.capi.LensrtError, match="particle tracer"):
        mi.render(mi.load_dict(d2), spp=1)
    with pytest.raises(ValueError, match="realistic lens"):
        sc.integrator().render_backward_lens(sc, np.zeros((64, 64, 3), f32))


@pytest.mark.gpu
def test_gpu_disk_field_source(mi):
    """`disk` (src/shapes/disk.cpp) as the notebooks use it — the carrier of a `directionalarea` field source
    (design_problem.py:355-371): Scene::ray_intersect on an elliptic, tilted disk bit-exact against the reference's own vectors,
    the particle tracer against the pinned oracle (default mode) and the reference's renders (scalar convention); an `area` light on a disk through
    the path tracer against the oracle."""
    import ref_scenes as rs
    R = np.load(os.path.join(ROOT, "tests", "golden", "ref_disk.npz"))
    d = rs.path_b_scene(lambda m: mi.Transform4f(m), source="disk")
    sc = mi.load_dict(d)
    o, dd, ref = R["isect_o"], R["isect_d"], R["isect"]
    t, _, _, prim = sc.ray_intersect(o, dd)
    t = t.cpu().numpy(); prim = prim.cpu().numpy()
    hit = np.isfinite(ref[:, 0])
    assert (np.isfinite(t) == hit).all()
    dk = hit & (ref[:, 21] == sc.data.shape_names.index("field_source"))
    assert dk.sum() > 500 and (prim[dk] == -2).all() and (prim[hit & ~dk] >= 0).all()
    bits = lambda x: np.ascontiguousarray(x, f32).view(np.uint32)   # noqa: E731
    assert (bits(t[dk]) == bits(ref[dk, 0])).all()
    img = mi.render(sc, seed=5).cpu().numpy()
    want = oracle_ptracer(mi, d, 5, 8, scalar=False)
    err = rel_rmse(img, want)
    print(f"disk source, default mode: rel-RMSE vs oracle {err:.3e}, mean {img.mean():.5f} / {want.mean():.5f}")
    assert err <= 2e-3 and abs(img.mean() / want.mean() - 1) <= 2e-3
    for seed in (0, 3):
        img = sc.integrator().render(sc, seed=seed, rng_mode=1, n_threads=int(R["n_threads"])).cpu().numpy()
        refi = R[f"image_seed{seed}"]
        close = np.abs(img - refi).max(-1) <= 1e-4 * np.maximum(refi.max(-1), 1e-3)
        print(f"disk source, scalar convention, seed {seed}: rel-RMSE vs the reference's image {rel_rmse(img, refi):.3e}, within 1e-4: {close.mean():.4f}")
        assert close.mean() >= 0.95 and abs(img.mean() / refi.mean() - 1) <= 1e-2
    # an `area` light on a disk in the path tracer (Disk::sample_position under Shape::sample_direction): product vs the oracle,
    # whose emitter sampling is pinned to the reference's vectors (test_disk_area_light_oracle_vs_reference)
    from test_gpu_ref_parity import oracle_jit_render
    d3 = rs.disk_light_scene(lambda m: mi.Transform4f(m))
    sc3 = mi.load_dict(d3)
    img = mi.render(sc3, seed=2).cpu().numpy()
    want, _ = oracle_jit_render(mi, sc3, d3, 16, 2)
    print(f"disk lamp, path: rel-RMSE vs oracle {rel_rmse(img, want):.3e}, mean {img.mean():.5f} / {want.mean():.5f}")
    assert rel_rmse(img, want) <= 2e-3 and abs(img.mean() / want.mean() - 1) <= 2e-3
