"""Mesh-lens parity mode, CPU side: the float64 restatement oracle/mesh_trace.py of the differentiable chain
the notebooks' mesh lenses rely on — Moeller–Trumbore, interpolated vertex normals, refraction — against the REFERENCE build's
own outputs (tests/golden/ref_mesh_lens.npz, written by tests/golden/gen_ref_mesh_lens.py): every interaction of 96 rays
through a meshed singlet from scene.ray_intersect + bsdf.sample, the vertex normals the reference's Mesh computed, and central
finite differences of the landing points with respect to five lens variables (each perturbed scene traced by the reference).
The oracle's reverse-mode derivative (torch.autograd over its own expressions) is what csrc/lrt_mesh_adj.cu is checked against
on the GPU (tests/test_gpu_mesh_lens.py)."""
import os
import sys

import numpy as np
import pytest
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.join(os.path.dirname(HERE), "oracle"))
import mesh_trace as MT  # noqa: E402

f64 = torch.float64


@pytest.fixture(scope="module")
def G():
    return dict(np.load(os.path.join(HERE, "golden", "ref_mesh_lens.npz")))


def oracle_scene(G, rows):
    """(V, F, tri_shape, eta, slices): both caps as one vertex / face list; vertex z = sag(rows) in float64 (differentiable)."""
    Vs, Fs, shape, nv, slices = [], [], [], 0, []
    for i in range(2):
        xy = torch.tensor(G[f"xy{i}"], dtype=f64)
        Vs.append(torch.cat([xy, MT.sag(rows[i], xy[:, 0], xy[:, 1])[:, None]], 1))
        F = torch.tensor(G[f"faces{i}"].astype(np.int64))
        Fs.append(F + nv); shape += [i] * len(F); slices.append((nv, nv + len(xy))); nv += len(xy)
    eta = torch.stack([rows[i, 13] / rows[i, 14] for i in range(2)])
    return torch.cat(Vs), torch.cat(Fs), torch.tensor(shape), eta, slices


def landing(o, d):
    t = -o[:, 2] / d[:, 2]
    return (o + t[:, None] * d)[:, :2]


def test_vertex_normals_match_the_reference_mesh(G):
    rows = torch.tensor(G["rows"], dtype=f64)
    V, F, _, _, slices = oracle_scene(G, rows)
    V = V.float().double()                               # the reference sees float32 positions
    N = MT.vertex_normals(V, F)
    for i, (a, b) in enumerate(slices):
        want = G[f"normals_lens{i}"]
        err = np.abs(N[a:b].numpy() - want).max()
        assert err < 2e-6, (i, err)


def test_interactions_match_the_reference(G):
    """t, p, shading normal, refracted direction and Fresnel reflectance of every interaction, on the reference's own
    incoming rays (float32 reference, float64 oracle: agreement to float32 rounding)."""
    rows = torch.tensor(G["rows"], dtype=f64)
    V, F, tri_shape, eta, _ = oracle_scene(G, rows)
    V = V.float().double()
    N = MT.vertex_normals(V, F)
    rec = G["records"]
    n_checked = 0
    for k in range(rec.shape[1]):
        m = np.isfinite(rec[:, k, 0])
        if not m.any():
            continue
        r = torch.tensor(rec[m, k])
        it = MT.interaction(V, F, N, tri_shape, eta, r[:, 0:3], r[:, 3:6])
        assert it["valid"].all()
        assert (it["shape"].numpy() == rec[m, k, 17].astype(int)).all()
        assert np.abs(it["t"].numpy() - rec[m, k, 6]).max() < 2e-7
        assert np.abs(it["p"].numpy() - rec[m, k, 7:10]).max() < 2e-8
        assert np.abs(it["n"].numpy() - rec[m, k, 10:13]).max() < 1e-6
        assert np.abs(it["d_out"].numpy() - rec[m, k, 13:16]).max() < 1e-6
        assert np.abs(it["r"].numpy() - rec[m, k, 16]).max() < 1e-6
        n_checked += int(m.sum())
    assert n_checked == 2 * len(rec)
    # the whole chain: landing points
    od = torch.tensor(G["rays"], dtype=f64)
    o, d, alive, w = MT.trace(V, F, N, tri_shape, eta, od[:, :3], od[:, 3:], 4)
    assert alive.numpy().tolist() == G["alive"].tolist()
    assert np.abs(landing(o, d).numpy() - G["landing"]).max() < 5e-8        # metres (float32 reference)


def test_reverse_mode_matches_finite_differences_of_the_reference(G):
    """d(landing point) / d(lens variable), ray by ray: torch.autograd over the oracle (sag -> vertices -> vertex normals ->
    Moeller–Trumbore -> interpolated normal -> refraction, twice) vs central differences of the reference's own forward.
    Tolerance: the reference is float32 (vertex rounding alone moves a landing point by ~1e-7 m, the steps move it by
    1e-5 .. 6e-5 m) and the difference quotient straddles triangle edges."""
    rows0 = torch.tensor(G["rows"], dtype=f64)
    od = torch.tensor(G["rays"], dtype=f64)
    a = torch.tensor(np.random.default_rng(3).standard_normal((len(od), 2)))
    rows = rows0.clone().requires_grad_(True)
    V, F, tri_shape, eta, _ = oracle_scene(G, rows)
    N = MT.vertex_normals(V, F)
    o, d, alive, w = MT.trace(V, F, N, tri_shape, eta, od[:, :3], od[:, 3:], 4)
    pf = landing(o, d)
    # per-ray Jacobian rows through two scalar losses would cost 192 backward passes: check (1) a random projection over all
    # rays, (2) the per-ray derivative of a handful of rays
    (pf * a).sum().backward(retain_graph=True)
    g = rows.grad.clone()
    fd = G["fd_landing"]
    for j, (ri, ci) in enumerate(G["var_index"]):
        want = float((fd[j] * a.numpy()).sum())
        got = float(g[ri, ci])
        assert abs(got - want) <= 1e-2 * abs(want) + 1e-6, (G["variables"][j], got, want)
    for ray in (0, 17, 55, 90):
        for c in range(2):
            rows.grad = None
            pf[ray, c].backward(retain_graph=True)
            for j, (ri, ci) in enumerate(G["var_index"]):
                want, got = fd[j, ray, c], float(rows.grad[ri, ci])
                scale = np.abs(fd[j]).max()
                assert abs(got - want) <= 3e-2 * scale, (G["variables"][j], ray, c, got, want)
