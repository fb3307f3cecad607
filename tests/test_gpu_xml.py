"""GPU side of the XML scenes (tests/golden/assets/xml): `mi.load_file`, analytic spheres, the constant environment emitter and
the procedural textures in both render kernels.

  * rng_mode = 1 vs the images the reference's scalar_rgb build rendered from the same XML files (tests/golden/ref_xml.npz),
    sample for sample: rel-RMSE <= 2e-3 and >= 98 % of the pixels within 1e-4 relative (the product's traversal and shading use
    fused multiply-adds; a sample that lands on a texture edge or a silhouette may take the other branch);
  * default mode vs the pinned oracle's JIT-convention render at matched seed / spp: rel-RMSE <= 2e-3, lens misses bit-equal;
  * Scene.ray_intersect over triangles + spheres vs the reference's ray_intersect / ray_test vectors;
  * the adjoint with an environment emitter: the image is linear in the emitters' radiances."""
import os

import numpy as np
import pytest

from oracle_lib import *  # noqa
from test_gpu_ref_parity import oracle_jit_render, IMG_TOL
from test_xml_parity import XML, xml_cases, case_key, G  # noqa: F401  (G is a fixture)


def rel_rmse(a, b):
    return float(np.sqrt(((a.astype(np.float64) - b) ** 2).mean()) / np.sqrt((b.astype(np.float64) ** 2).mean()))


@pytest.mark.gpu
@pytest.mark.parametrize("name,kw", xml_cases())
def test_gpu_xml_scalar_convention_vs_reference_images(mi, G, name, kw):
    sc = mi.load_file(os.path.join(XML, name + ".xml"), **kw)
    img = mi.render(sc, seed=0, rng_mode=1).cpu().numpy()
    ref = G[case_key(name, kw)]
    err = rel_rmse(img, ref)
    close = np.abs(img - ref).max(-1) <= 1e-4 * np.maximum(ref.max(-1), 1e-3)
    print(f"{name} {kw}: rel-RMSE vs the reference's scalar_rgb image {err:.3e}, pixels within 1e-4: {close.mean():.4f}")
    assert err <= 2e-3, err
    assert close.mean() >= 0.98


@pytest.mark.gpu
@pytest.mark.parametrize("name,kw", xml_cases())
def test_gpu_xml_default_mode_vs_pinned_oracle(mi, G, name, kw):
    d = mi.xml_to_dict(path=os.path.join(XML, name + ".xml"), **kw)
    sc = mi.load_dict(d)
    spp = sc.sensors()[0].sampler.sample_count()
    img = mi.render(sc, seed=5).cpu().numpy()
    want, miss = oracle_jit_render(mi, sc, d, spp, 5)
    err = rel_rmse(img, want)
    print(f"{name} {kw}: rel-RMSE vs oracle {err:.3e}")
    assert sc.stats()["lens_misses"] == miss
    assert err < 5 * IMG_TOL, err
    assert abs(img.mean() / G[case_key(name, kw)].mean() - 1) < 0.05       # same integrand as the reference's render


@pytest.mark.gpu
def test_gpu_ray_intersect_with_spheres_vs_reference(mi, G):
    import torch
    sc = mi.load_file(os.path.join(XML, "charts.xml"))
    for nm in ("isect", "inside"):
        o, d, ref = G[f"charts/{nm}_o"], G[f"charts/{nm}_d"], G[f"charts/{nm}"]
        t, _, _, prim = sc.ray_intersect(o, d)
        t = t.cpu().numpy(); prim = prim.cpu().numpy()
        hit = np.isfinite(ref[:, 0])
        assert (np.isfinite(t) == hit).all()
        sph = hit & (ref[:, 21] == sc.data.shape_names.index("probe"))
        assert (prim[sph] == -2).all() and (prim[hit & ~sph] >= 0).all()
        assert (bits(t[sph]) == bits(ref[sph, 0])).all()                 # the double-precision quadratic, rounded once
        assert np.abs(t[hit] - ref[hit, 0]).max() <= 2e-6 * ref[hit, 0].max()
    o, d, maxt = G["charts/isect_o"], G["charts/isect_d"], G["charts/test_maxt"]
    t, _, _, _ = sc.ray_intersect(o, d, maxt=maxt, any_hit=True)
    occ = np.isfinite(t.cpu().numpy()).astype(np.int32)
    assert (occ != G["charts/test_occluded"]).mean() <= 1e-3


@pytest.mark.gpu
def test_gpu_adjoint_with_environment_is_linear_in_radiance(mi):
    """Every path's value is a sum of terms each linear in one emitter's radiance, so <dL/dI, I> = sum_e <dI/dradiance_e, radiance_e>
    for any cotangent: checks the environment's two gradient sites (escaped rays, emitter sampling) next to a mesh light's."""
    import torch
    d = mi.xml_to_dict(path=os.path.join(XML, "studio.xml"), res=48, spp=32)
    d["_arg_0"]["type"] = "prb"
    sc = mi.load_dict(d)
    s = sc.sensors()[0]
    st = mi.render(sc, seed=3, develop=False)
    gs = torch.zeros_like(st); gs[..., :3] = torch.rand((s.film.crop_h, s.film.crop_w, 3), device="cuda", generator=torch.Generator("cuda").manual_seed(1))
    gb, ge = mi.scene.render_backward(sc, gs, seed=3)
    rad = torch.as_tensor(sc.arrays["emitters"][:, :3], device="cuda")
    lhs = float((gs[..., :3].double() * st[..., :3].double()).sum())
    rhs = float((ge[:, :3].double() * rad.double()).sum())
    assert ge[:, :3].abs().sum(1).min() > 0                                   # both emitters receive a gradient
    assert abs(lhs - rhs) <= 2e-3 * abs(lhs), (lhs, rhs)
