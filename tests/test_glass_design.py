"""Glass variables and wavelengths in the lens-design loop (what the notebooks differentiate — the `dispersive`
BSDF's (int_ior_d, int_V_d), src/bsdfs/dispersive_dielectric.cpp:236-239,425-431 — and the chromatic side of their designs).
CPU: the glass model itself. GPU: LensDesignProblem(glasses=..., wavelengths=...) — the table at a wavelength, the gradient
with respect to (n_d, V_d) through the IOR columns of the spot adjoint, the polychromatic loss, the graph-captured iteration."""
import numpy as np
import pytest

from oracle_lib import *  # noqa


def test_glass_ior_is_the_dispersive_bsdfs_cauchy_fit(mi):
    """eval_ior(n_d, V_d, w): n(0.5893) = n_d, n_F - n_C = (n_d - 1) / V_d at the model's own F and C lines, and the float32
    restatement the scene loader uses for `dispersive` BSDFs (pinned by the reference's renders) agrees."""
    import torch
    L = mi.lens_design
    for n_d, V_d in ((1.5168, 64.17), (1.62, 36.4), (1.7, 30.0)):
        assert abs(L.glass_ior(n_d, V_d, 0.5893) - n_d) < 1e-12
        assert abs((L.glass_ior(n_d, V_d, 0.48613) - L.glass_ior(n_d, V_d, 0.65627)) - (n_d - 1) / V_d) < 1e-12
        r = mi.bsdf_row({"type": "dispersive", "int_ior": n_d, "int_V_d": V_d})
        assert abs(float(r[14]) - L.glass_ior(n_d, V_d, 0.5893) / L.glass_ior(1.0, 0.0, 0.5893)) < 2e-7
    assert L.glass_ior(1.000277, 0.0, 0.45) == 1.000277                     # V_d = 0: no dispersion (the reference's air)
    n = torch.tensor([1.5168, 1.62], dtype=torch.float64, requires_grad=True)
    V = torch.tensor([64.17, 36.4], dtype=torch.float64, requires_grad=True)
    assert torch.autograd.gradcheck(lambda a, b: L.glass_ior(a, b, 0.45), (n, V))


def doublet(mi):
    """the notebooks' cemented doublet (doublet_full_optim.ipynb cell 6 == realisticlens.cpp:1283-1320): its two Cauchy glasses
    (A, B) restated as (n_d, V_d) of the `dispersive` model — the same curve: n_d = A + B / 0.5893^2, V_d = (n_d - 1) / (B k)"""
    L = mi.lens_design
    k = 1.0 / 0.48613 ** 2 - 1.0 / 0.65627 ** 2
    as_glass = lambda A, B: (A + B / 0.5893 ** 2, (A + B / 0.5893 ** 2 - 1.0) / (B * k))   # noqa: E731
    crown, flint = as_glass(1.45608, 0.021087), as_glass(1.56895, 0.026305)
    glasses = {2: crown, 1: flint}                        # medium 2: between rows 1|2 (front element), medium 1: rows 0|1 (rear element)
    rows = L.asphere_table([24.0, -24.0, -168.0107], [3.0, 2.2529, 46.7471], [8, 8, 8], [crown[0], flint[0], 1.000277])
    return rows, glasses


@pytest.mark.gpu
def test_gpu_glass_table_loss_and_gradient(mi):
    import torch
    L = mi.lens_design
    rows, glasses = doublet(mi)
    kw = dict(resolution=64, full_res=2048, spp=64, spp_grad=64, lr=2e-4)
    plain = L.LensDesignProblem(mi, rows, [0.0, 2.5], **kw)
    prob = L.LensDesignProblem(mi, rows, [0.0, 2.5], glasses=glasses, free_glass=True, **kw)
    # the table at 589.3 nm is the plain table; at other wavelengths only the glass media's IOR columns move
    r0 = prob.rows().detach().cpu().numpy()
    assert np.abs(r0 - rows).max() < 2e-7
    rb = prob.rows(450.0).detach().cpu().numpy()
    moved = np.argwhere(np.abs(rb - r0) > 1e-6)
    assert sorted(map(tuple, moved)) == [(0, 14), (1, 13), (1, 14), (2, 13)]
    assert rb[1, 13] > r0[1, 13] and rb[2, 13] > r0[2, 13]              # normal dispersion: higher index in the blue
    # same seeds, same table -> the same loss as the problem without glass variables
    l0 = plain.loss(plain.rows(), 3); l1 = prob.loss(prob.rows(), 3)
    assert abs(float(l0) - float(l1)) <= 1e-5 * abs(float(l0))
    # gradient with respect to (n_d, V_d) = the IOR columns' gradient of the adjoint, chained through glass_ior
    w = 486.1
    prob.opt.zero_grad(set_to_none=True)
    prob.loss(prob.rows(w), 5).backward()
    got = prob.glass.grad.detach().cpu().double().numpy().copy()
    leaf = prob.rows(w).detach().clone().requires_grad_(True)
    prob.loss(leaf, 5).backward()
    g = leaf.grad.detach().cpu().double()
    gl = prob.glass.detach().cpu().double().requires_grad_(True)
    ior = L.glass_ior(gl[:, 0], gl[:, 1], 1e-3 * w)                     # media 1, 2
    (ior[0] * (g[1, 13] + g[0, 14]) + ior[1] * (g[2, 13] + g[1, 14])).backward()     # prob.glass rows: media 1, 2 (sorted)
    want = gl.grad.numpy()
    assert np.abs(want).max() > 0
    assert np.abs(got - want).max() <= 1e-4 * np.abs(want).max(), (got, want)
    # (the IOR columns' gradient itself is checked against float64 dual numbers and finite differences at kernel level:
    #  tests/test_lens_adjoint.py, tests/test_gpu_spot.py — columns 13, 14)


@pytest.mark.gpu
def test_gpu_polychromatic_design_iterations(mi):
    """three wavelengths (F, d, C lines), glass variables free: the loss is the mean of the three spot losses, blue and red
    spots differ (the chromatic aberration the loss sees), iterations run eagerly and as a captured CUDA graph."""
    import torch
    L = mi.lens_design
    rows, glasses = doublet(mi)
    prob = L.LensDesignProblem(mi, rows, [0.0, 2.5], resolution=64, full_res=2048, spp=64, spp_grad=32, lr=2e-4, glasses=glasses,
                               wavelengths=[486.1, 589.3, 656.3], free_glass=True, lr_glass=1e-4)
    parts = [float(prob.loss(prob.rows(w), 0, stream=j)) for j, w in enumerate(prob.wavelengths)]
    total = float(prob.loss_all_wavelengths(0))
    assert abs(total - np.mean(parts)) <= 1e-5 * abs(total)
    assert abs(parts[0] - parts[2]) > 1e-3 * max(parts)                   # the spots are not the same in blue and in red
    g0 = prob.glass.detach().clone()
    losses = [prob.step() for _ in range(12)]
    assert np.isfinite(losses).all()
    assert (prob.glass.detach() - g0).abs().max() > 0                    # the glass variables moved
    assert np.mean(losses[-4:]) < np.mean(losses[:4])
    prob.capture()
    for _ in range(3):
        l = prob.step_graph()
    torch.cuda.synchronize()
    assert np.isfinite(float(l))
    with pytest.raises(ValueError):
        L.LensDesignProblem(mi, rows, [0.0], resolution=32, full_res=2048, spp=8, spp_grad=8, wavelengths=[450.0, 650.0])
