"""TEST INFRASTRUCTURE — not part of the product (only tests/ may import this).

CPU restatement, in float64 torch tensors, of what the reference does to a ray that meets a refractive triangle mesh — the
differentiable chain the notebooks' mesh lenses rely on:

  * Moeller–Trumbore (include/mitsuba/render/mesh.h:432-452), brute force over all triangles, closest hit;
  * the shading normal interpolated from the vertex normals (src/render/mesh.cpp:1448-1500), the vertex normals themselves
    angle-weighted face normals (Mesh::recompute_vertex_normals, mesh.cpp:321-375);
  * fresnel / refract (include/mitsuba/render/fresnel.h:35-84,311-315) as the transmission lobe of `dielectric` /
    `dispersive` applies them (src/bsdfs/dielectric.cpp:233-300);
  * the sag functions that tie the vertex z to the lens variables (notebooks/scripts/lens_design/lens.py:67-85,210-237).

Everything is an ordinary torch expression, so torch.autograd differentiates it with respect to the vertex positions, the
vertex normals, the relative indices and the rays: the checker for csrc/lrt_mesh_adj.cu. Forward pinned to the reference build
by tests/golden/ref_mesh_lens.npz (per-interaction t, p, shading normal, refracted direction from the reference's own
scene.ray_intersect + bsdf.sample); gradients pinned by central finite differences of the REFERENCE's forward (same file) —
the reference's AD variants cannot run here (no LLVM), so gradient parity with its AD itself stays unpinned."""
import numpy as np
import torch

RAY_EPSILON = 8.94069671630859375e-05      # math::RayEpsilon (include/mitsuba/core/math.h:18-22)
f64 = torch.float64


def unit_angle(a, b):
    """math.h unit_angle: numerically robust angle between unit vectors."""
    dp = (a * b).sum(-1)
    return torch.where(dp < 0, np.pi - 2 * torch.asin((0.5 * (b + a).norm(dim=-1)).clamp(-1, 1)),
                       2 * torch.asin((0.5 * (b - a).norm(dim=-1)).clamp(-1, 1)))


def vertex_normals(V, F):
    """mesh.cpp:321-375: face normals weighted by the face angle at each vertex, normalised."""
    p = [V[F[:, k]] for k in range(3)]
    n = torch.linalg.cross(p[1] - p[0], p[2] - p[0])
    n = n / n.norm(dim=1, keepdim=True).clamp_min(1e-300)
    out = torch.zeros_like(V)
    for k in range(3):
        d0 = p[(k + 1) % 3] - p[k]; d1 = p[(k + 2) % 3] - p[k]
        d0 = d0 / d0.norm(dim=1, keepdim=True); d1 = d1 / d1.norm(dim=1, keepdim=True)
        out = out.index_add(0, F[:, k], n * unit_angle(d0, d1)[:, None])
    return out / out.norm(dim=1, keepdim=True).clamp_min(1e-300)


def moeller_trumbore(o, d, p0, p1, p2):
    """mesh.h:432-452 for rays [n,3] against triangles [m,3]: (t, u, v, hit) each [n,m]."""
    e1, e2 = p1 - p0, p2 - p0
    pvec = torch.linalg.cross(d[:, None, :], e2[None])
    det = (e1[None] * pvec).sum(-1)
    inv = 1.0 / det
    tvec = o[:, None, :] - p0[None]
    u = (tvec * pvec).sum(-1) * inv
    qvec = torch.linalg.cross(tvec, e1[None].expand_as(tvec))
    v = (d[:, None, :] * qvec).sum(-1) * inv
    t = (e2[None] * qvec).sum(-1) * inv
    hit = (u >= 0) & (v >= 0) & (u + v <= 1) & (t >= 0) & (det != 0)
    return t, u, v, hit


def fresnel(ci, eta):
    """fresnel.h:35-84: (r, cos_theta_t, eta_it, eta_ti)."""
    outside = ci >= 0
    rcp = 1.0 / eta
    eta_it = torch.where(outside, eta, rcp); eta_ti = torch.where(outside, rcp, eta)
    ct2 = 1.0 - (1.0 - ci * ci) * eta_ti * eta_ti
    cia, cta = ci.abs(), ct2.clamp_min(1e-30).sqrt()
    a_s = (cia - eta_it * cta) / (cia + eta_it * cta)
    a_p = (cta - eta_it * cia) / (cta + eta_it * cia)
    r = 0.5 * (a_s * a_s + a_p * a_p)
    r = torch.where(eta == 1.0, torch.zeros_like(r), torch.where(cia == 0, torch.ones_like(r), r))
    r = torch.where(ct2 <= 0, torch.ones_like(r), r)
    return r, torch.where(outside, -cta, cta), eta_it, eta_ti


def interaction(V, F, N, tri_shape, eta, o, d):
    """One surface: closest hit of every ray, shading normal, refracted direction.
    -> dict(valid, tri, t, u, v, p, n, d_out, r, shape); rays without a hit have valid = False."""
    p0, p1, p2 = V[F[:, 0]], V[F[:, 1]], V[F[:, 2]]
    with torch.no_grad():                                   # which triangle: a discrete event (detached in the reference too)
        t, u, v, hit = moeller_trumbore(o, d, p0, p1, p2)
        t = torch.where(hit, t, torch.full_like(t, float("inf")))
        tri = t.argmin(dim=1)
        valid = torch.isfinite(t.gather(1, tri[:, None])[:, 0])
    q0, q1, q2 = p0[tri], p1[tri], p2[tri]
    e1, e2 = q1 - q0, q2 - q0
    pvec = torch.linalg.cross(d, e2)
    inv = 1.0 / (e1 * pvec).sum(-1)
    tvec = o - q0
    u = (tvec * pvec).sum(-1) * inv
    qvec = torch.linalg.cross(tvec, e1)
    v = (d * qvec).sum(-1) * inv
    t = (e2 * qvec).sum(-1) * inv
    p = o + t[:, None] * d
    h = torch.linalg.cross(e1, e2)
    if N is not None:                                       # shapes without vertex normals carry zero rows: face normal there
        f = F[tri]
        hs = N[f[:, 0]] * (1 - u - v)[:, None] + N[f[:, 1]] * u[:, None] + N[f[:, 2]] * v[:, None]
        has = (N[f].detach().abs().sum(dim=(1, 2)) > 0)
        h = torch.where(has[:, None], hs, h)
    n = h / h.norm(dim=1, keepdim=True)
    shape = tri_shape[tri]
    et = eta[shape]
    ci = -(d * n).sum(-1)
    r, ct, eta_it, eta_ti = fresnel(ci, torch.where(et > 0, et, torch.ones_like(et)))
    g = d * eta_ti[:, None] + n * (eta_ti * ci + ct)[:, None]
    d_out = g / g.norm(dim=1, keepdim=True)
    return dict(valid=valid, tri=tri, t=t, u=u, v=v, p=p, n=n, d_out=d_out, r=r, shape=shape, absorb=~(et > 0))


def trace(V, F, N, tri_shape, eta, o, d, max_interactions):
    """Rays through refractive meshes: (o, d, alive, weight) after at most max_interactions surfaces — the function
    lrt_scene_trace_refractive computes. Shapes with eta = 0 absorb; total internal reflection kills."""
    d = d / d.norm(dim=1, keepdim=True)
    n = o.shape[0]
    alive = torch.ones(n, dtype=torch.bool); going = torch.ones(n, dtype=torch.bool)
    weight = torch.ones(n, dtype=f64)
    for _ in range(max_interactions):
        it = interaction(V, F, N, tri_shape, eta, o, d)
        hit = going & it["valid"]
        going = hit
        dead = hit & (it["absorb"] | (it["r"] >= 1.0))
        alive = alive & ~dead
        ok = hit & ~dead
        going = going & ~dead
        if not ok.any():
            break
        c = RAY_EPSILON * 0.01 * (1.0 + it["p"].detach().abs().max(dim=1).values)
        weight = torch.where(ok, weight * (1.0 - it["r"]), weight)
        d_new = torch.where(ok[:, None], it["d_out"], d)
        o = torch.where(ok[:, None], it["p"] + d_new * c[:, None], o)
        d = d_new
    return o, d, alive, torch.where(alive, weight, torch.zeros_like(weight))


def sag(row, x, y):
    """Vertex z of an interface-table row (rows16, metres) at (x, y): sphere (kind 0), plane / stop (1, 2), even asphere in the
    unitless radius r / R_e (kind 3; realisticlens.cpp:680-704 == lens.py:210-237). `row` is a float64 tensor (differentiable)."""
    kind = int(row[0])
    r2 = x * x + y * y
    if kind == 0:
        R = row[3]
        return row[1] + R - torch.sign(R) * torch.sqrt(R * R - r2)
    if kind in (1, 2):
        return row[1] + 0 * r2
    Re = row[2]
    q = r2 / (Re * Re)
    z = 0 * q
    for i in range(6, -1, -1):
        z = z * q + row[6 + i]
    z = z * q * q
    z = z + row[3] * q / (1 + torch.sqrt((1 - (1 + row[4]) * row[3] * row[3] * q).clamp_min(0)))
    return row[1] - Re * z
