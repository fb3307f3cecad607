"""`prb` through the Python surface: mi.load_dict({'type': 'prb'}), mi.render(scene, params, seed_grad=, spp_grad=) returning a
tensor with a grad_fn (_RenderOp, src/python/python/util.py:356-407), develop's adjoint, dielectric lobe gradients, Russian
roulette active and detached, mi.ad.Adam. Checker: central finite differences of the PINNED oracle's render (float64
accumulators) of the differential phase's own samples, with the roulette frozen to the unperturbed scene (orc_set_rr_scene)
exactly as prb detaches it. Stated tolerance: cosine >= 0.999, per entry 3 % of the largest."""
import ctypes as C

import numpy as np
import pytest

from oracle_lib import *  # noqa
from test_ref_parity import x_fov


def box_scene(mi, max_depth, rr_depth, res=32, spp=16):
    d = mi.cornell_box()
    d["sensor"]["film"]["width"] = d["sensor"]["film"]["height"] = res
    d["sensor"]["sampler"]["sample_count"] = spp
    d["integrator"] = {"type": "prb", "max_depth": max_depth, "rr_depth": rr_depth}
    d["small-box"]["bsdf"] = {"type": "conductor", "material": "Cu", "specular_reflectance": {"type": "rgb", "value": [0.9, 0.8, 0.7]}}
    d["large-box"]["to_world"] = mi.Transform4f.translate([-0.33, -0.38, -0.28]).rotate([0, 1, 0], 18.25).scale([0.3, 0.61, 0.3])
    d["large-box"]["bsdf"] = {"type": "dielectric", "int_ior": 1.5, "ext_ior": 1.0, "specular_reflectance": {"type": "rgb", "value": [0.9, 0.85, 0.8]},
                              "specular_transmittance": {"type": "rgb", "value": [0.7, 0.8, 0.9]}}
    return d


def fd_gradients(mi, sc, d, G, seed, spp, max_depth, rr_depth, keys, params):
    """d<G, develop(render)>/d(param) by central differences of the oracle, roulette detached."""
    tw = d["sensor"]["to_world"].matrix
    res = d["sensor"]["film"]["width"]
    base = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in sc.arrays.items()}
    h0 = orc_scene(base)
    oracle().orc_set_rr_scene(h0)
    try:
        def phi(arrays):
            st, _ = orc_render(orc_scene(arrays), FilmSpec(res, res), spp, seed=seed, max_depth=max_depth, rr_depth=rr_depth, fov=39.3077,
                               to_world=tw, near=1e-3, far=100.0)
            return float((develop(st).astype(np.float64) * G).sum())
        out = {}
        for key in keys:
            kind, i = params._where[key]
            tab, c0, h = {"bsdf": ("bsdfs", 2, 0.02), "bsdf_t": ("bsdfs", 11, 0.02), "emitter": ("emitters", 0, 0.5)}[kind]
            g = np.zeros(3)
            for c in range(3):
                ap = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in base.items()}
                am = {k: (v.copy() if isinstance(v, np.ndarray) else v) for k, v in base.items()}
                ap[tab][i, c0 + c] += h; am[tab][i, c0 + c] -= h
                g[c] = (phi(ap) - phi(am)) / (2 * h)
            out[key] = g
        return out
    finally:
        oracle().orc_set_rr_scene(None)


@pytest.mark.gpu
@pytest.mark.parametrize("max_depth,rr_depth", [(4, 5), (8, 3)])
def test_gpu_prb_render_autograd_vs_oracle_fd(mi, max_depth, rr_depth):
    import torch
    d = box_scene(mi, max_depth, rr_depth)
    sc = mi.load_dict(d)
    assert isinstance(sc.integrator(), mi.PRBIntegrator)
    params = mi.traverse(sc)
    keys = ["red.reflectance.value", "white.reflectance.value", "small-box.bsdf.specular_reflectance.value",
            "large-box.bsdf.specular_reflectance.value", "large-box.bsdf.specular_transmittance.value", "light.emitter.radiance.value"]
    for k in keys:
        assert k in params, (k, list(params))
        params[k].requires_grad_(True)
    G = np.random.default_rng(1).standard_normal((32, 32, 3))
    img = mi.render(sc, params, seed=3, seed_grad=11, spp=16, spp_grad=16)
    assert img.grad_fn is not None and img.shape == (32, 32, 3)
    # the primal image is the plain render with the primal seed
    assert torch.allclose(img.detach(), mi.render(sc, seed=3, spp=16), rtol=1e-5, atol=1e-6)
    (img * torch.from_numpy(G).float().cuda()).sum().backward()
    want = fd_gradients(mi, sc, d, G, 11, 16, max_depth, rr_depth, keys, params)
    got = np.concatenate([params[k].grad.cpu().numpy().astype(np.float64) for k in keys])
    ref = np.concatenate([want[k] for k in keys])
    cos = float(got @ ref / (np.linalg.norm(got) * np.linalg.norm(ref)))
    print(f"max_depth {max_depth} rr_depth {rr_depth}: cosine {cos:.6f}, max rel dev {np.abs(got - ref).max() / np.abs(ref).max():.3e}")
    assert cos >= 0.999, (cos, got, ref)
    assert np.abs(got - ref).max() <= 3e-2 * np.abs(ref).max(), (got, ref)
    # every parameter kind received a gradient
    for k in keys:
        assert np.abs(params[k].grad.cpu().numpy()).max() > 0, k
    assert sc.stats()["stack_clamps"] == 0


@pytest.mark.gpu
def test_gpu_prb_limits_and_errors(mi):
    import torch
    d = box_scene(mi, 18, 5)
    sc = mi.load_dict(d)
    params = mi.traverse(sc)
    params["red.reflectance.value"].requires_grad_(True)
    img = mi.render(sc, params, seed=1, spp=4)
    with pytest.raises(mi.capi.LensrtError):       # 17 recorded bounces do not fit: refused, not truncated
        img.sum().backward()
    with pytest.raises(Exception):                 # equal primal / differential seeds (util.py:521-523)
        mi.render(sc, params, seed=5, seed_grad=5, spp=4)
    # the `path` plugin has no reverse mode: mi.render falls back to a plain image only when nothing requires a gradient
    d2 = box_scene(mi, 4, 5); d2["integrator"]["type"] = "path"
    sc2 = mi.load_dict(d2); p2 = mi.traverse(sc2)
    assert mi.render(sc2, p2, seed=1, spp=4).grad_fn is None
    p2["red.reflectance.value"].requires_grad_(True)
    with pytest.raises(Exception):
        mi.render(sc2, p2, seed=1, spp=4).sum().backward()


@pytest.mark.gpu
def test_gpu_adam_recovers_wall_colour(mi):
    """The reference's introductory inverse-rendering loop (tutorials: optimise the red wall's reflectance towards a target
    image) with mi.ad.Adam + params.update(opt)."""
    import torch
    d = mi.cornell_box()
    d["sensor"]["film"]["width"] = d["sensor"]["film"]["height"] = 48
    d["integrator"] = {"type": "prb", "max_depth": 5}
    sc = mi.load_dict(d)
    params = mi.traverse(sc)
    key = "red.reflectance.value"
    target_value = params[key].clone()
    target = mi.render(sc, seed=100, spp=256)
    opt = mi.ad.Adam(lr=0.05)
    opt[key] = torch.tensor([0.2, 0.5, 0.3])
    params.update(opt)
    def clean_loss():
        with torch.no_grad():
            return float(((mi.render(sc, seed=999, spp=256) - target) ** 2).mean())
    loss0 = clean_loss()
    for it in range(40):
        img = mi.render(sc, params, seed=it, spp=16)
        loss = ((img - target) ** 2).mean()
        loss.backward()
        opt.step()
        with torch.no_grad():
            opt[key].clamp_(0.0, 1.0)
        params.update(opt)
    loss1 = clean_loss()
    err = (opt[key].detach() - target_value).abs().max().item()
    print("loss", loss0, "->", loss1, "recovered", opt[key].detach().tolist(), "err", err)
    assert loss1 < 0.3 * loss0 and err < 0.08
