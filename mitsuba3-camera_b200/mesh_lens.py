"""Mesh-lens parity mode: the notebooks' representation of a lens — every refracting surface a triangle mesh
whose vertex z follows the sag function — rendered with the B200 scene kernels.

Restates notebooks/scripts/lens_design/geometry.py (get_hemisphere_ico :44-70, get_spherical_cap :121-160,
create_surface_geometry :296-390) and the sag functions of lens.py (:67-85 conic, :210-237 even asphere) on top of the interface
table this package uses everywhere else (rows16, rear -> front, metres). One optimisation step of the notebooks
(lens.py:509-565: new sag -> params['lensN.vertex_positions'] -> params.update()) becomes: new vertex z on the device ->
lrt_scene_refit (BVH refit kernels) -> lrt_scene_trace_refractive."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi as capi

f32 = np.float32


def icosphere(n):
    """gpytoolbox.icosphere(n): icosahedron subdivided n times (each triangle into four), vertices pushed to the unit sphere."""
    t = (1 + 5 ** 0.5) / 2
    v = [[-1, t, 0], [1, t, 0], [-1, -t, 0], [1, -t, 0], [0, -1, t], [0, 1, t], [0, -1, -t], [0, 1, -t], [t, 0, -1], [t, 0, 1],
         [-t, 0, -1], [-t, 0, 1]]
    f = [[0, 11, 5], [0, 5, 1], [0, 1, 7], [0, 7, 10], [0, 10, 11], [1, 5, 9], [5, 11, 4], [11, 10, 2], [10, 7, 6], [7, 1, 8],
         [3, 9, 4], [3, 4, 2], [3, 2, 6], [3, 6, 8], [3, 8, 9], [4, 9, 5], [2, 4, 11], [6, 2, 10], [8, 6, 7], [9, 8, 1]]
    v = [np.asarray(p, np.float64) / np.linalg.norm(p) for p in v]
    for _ in range(n):
        cache, nf = {}, []

        def mid(a, b):
            k = (min(a, b), max(a, b))
            if k not in cache:
                p = v[a] + v[b]; v.append(p / np.linalg.norm(p)); cache[k] = len(v) - 1
            return cache[k]
        for a, b, c in f:
            ab, bc, ca = mid(a, b), mid(b, c), mid(c, a)
            nf += [[a, ab, ca], [b, bc, ab], [c, ca, bc], [ab, bc, ca]]
        f = nf
    return np.asarray(v, np.float64), np.asarray(f, np.int64)


def hemisphere_ico(n):
    """geometry.py:44-70: the +z half of the icosphere, open at z = 0 (faces with a vertex below -edge/2 are dropped)."""
    V, F = icosphere(n)
    avg_edge = np.sqrt(4.0 * np.pi / len(F))
    bad = V[:, 2] < -avg_edge / 2.0
    F = F[~bad[F].any(axis=1)]
    used, inv = np.unique(F.reshape(-1), return_inverse=True)
    return V[used], inv.reshape(-1, 3)


def spherical_cap(n, elem_radius, abs_curvature):
    """geometry.py:121-160: the hemisphere's polar angles rescaled so that the cap subtends the surface's own aperture angle."""
    V, F = hemisphere_ico(n)
    target = np.arcsin(elem_radius * max(abs_curvature, 1e-3))
    vlen = np.linalg.norm(V, axis=1)
    ang = np.arccos(np.clip(V[:, 2] / vlen, -1, 1)) * (target / (0.5 * np.pi))
    rxy = np.linalg.norm(V[:, :2], axis=1)
    out = V.copy()
    out[:, 2] = vlen * np.cos(ang)
    with np.errstate(invalid="ignore", divide="ignore"):
        out[:, :2] *= (vlen * np.sin(ang) / rxy)[:, None]
    out[np.isnan(out)] = 0.0
    out[:, 2] -= np.cos(target)
    out *= elem_radius / np.sin(target)
    return out, F


def sag(row, x, y, xp=np):
    """Vertex z of an interface-table row at (x, y), metres, camera space (film at z = 0, +z towards the scene): sphere
    (kind 0: z_vertex + R - sign(R) sqrt(R^2 - r^2)), plane / stop, or the even asphere in unitless radius r / R_e
    (kind 3; realisticlens.cpp:680-704 == lens.py:210-237). `xp` = numpy or torch."""
    kind = int(row[0])
    r2 = x * x + y * y
    if kind == 0:
        R = row[3]
        return row[1] + R - xp.sign(R) * xp.sqrt(R * R - r2) if xp is not np else row[1] + R - np.sign(R) * np.sqrt(R * R - r2)
    if kind in (1, 2):
        return row[1] + 0 * r2
    Re = row[2]
    r2_ = r2 / (Re * Re)
    z = 0 * r2_
    for i in range(6, -1, -1):
        z = z * r2_ + row[6 + i]
    z = z * r2_ * r2_
    z = z + row[3] * r2_ / (1 + xp.sqrt(xp.clip(1 - (1 + row[4]) * row[3] * row[3] * r2_, 0, None) if xp is np
                                        else xp.clamp(1 - (1 + row[4]) * row[3] * row[3] * r2_, min=0)))
    return row[1] - Re * z


def curvature(row):
    """|second-order curvature| of a table row (1 / m)."""
    kind = int(row[0])
    if kind == 0:
        return abs(1.0 / float(row[3]))
    if kind == 3:
        return abs(float(row[3]) / float(row[2]))
    return 0.0


def surface_meshes(rows, subdiv=5, baffle_scale=3.0):
    """create_surface_geometry (geometry.py:296-390) for every refracting row of the table: (caps, baffles), each a list of
    (vertices [nv,3] float64, faces [nf,3]) — the cap's xy grid is fixed, its z is sag(row, x, y); normals face +z."""
    caps, baffles = [], []
    for row in rows:
        if int(row[0]) == 2:          # the aperture stop has no glass: an annular baffle only
            caps.append(None)
            V, F = hemisphere_ico(2)
            rim = _boundary_loop(F)
            el = V[rim] * np.array([row[2], row[2], 0.0]) / np.maximum(np.linalg.norm(V[rim, :2], axis=1, keepdims=True), 1e-30)
            el[:, 2] = row[1]
            baffles.append(_annulus(el, baffle_scale * float(row[2])))
            continue
        V, F = spherical_cap(subdiv, float(row[2]), curvature(row))
        V[:, 2] = sag(np.asarray(row, np.float64), V[:, 0], V[:, 1])
        caps.append((V, F))
        rim = _boundary_loop(F)
        baffles.append(_annulus(V[rim], baffle_scale * float(row[2])))
    return caps, baffles


def _boundary_loop(F):
    """Ordered vertex loop of the open boundary (gpytoolbox.boundary_loops(F)[0])."""
    edges = {}
    for a, b, c in F:
        for u, v in ((a, b), (b, c), (c, a)):
            edges[(u, v)] = edges.get((u, v), 0) + 1
    nxt = {u: v for (u, v) in edges if (v, u) not in edges}
    start = next(iter(nxt)); loop = [start]
    while nxt[loop[-1]] != start:
        loop.append(nxt[loop[-1]])
    return np.asarray(loop)


def _annulus(rim_pts, outer):
    """Flat-ish ring from the cap's rim out to the square |x|, |y| <= outer (geometry.py:355-388, both orientations)."""
    el = np.asarray(rim_pts, np.float64)
    edge = el.copy()
    edge[:, :2] *= outer / np.max(np.abs(el[:, :2]), axis=1, keepdims=True)
    n = len(el)
    F = []
    for i in range(n):
        j = (i + 1) % n
        F += [[j, i, n + j], [n + i, n + j, i]]
    return np.vstack([edge, el]), np.asarray(F, np.int64)


class MeshLens:
    """The meshed lens as a lensrt scene: one shape per refracting surface (relative index across it) + absorbing baffles.

    rows: interface table (rear -> front). The ray direction convention of the trace: light travels from the scene (+z) to
    the film (z = 0), so it enters surface i from the medium ior_right (world side) into ior_left (film side); cap normals
    face +z, i.e. the 'interior' (behind the normal) is the film-side medium: eta_i = ior_left / ior_right."""

    def __init__(self, mi, rows, subdiv=5, smooth=True):
        import torch
        from .scene import SceneData, Scene, PathIntegrator, vertex_normals
        self.mi, self.rows0 = mi, np.ascontiguousarray(rows, f32).copy()
        caps, baffles = surface_meshes(self.rows0, subdiv)
        data = SceneData()
        glass = data.add_bsdf("glass", np.zeros(16, f32)); black = data.add_bsdf("black", np.zeros(16, f32))
        eta, self.cap_slices, self.cap_xy, self.baffle_slices, nv = [], [], [], [], 0
        for i, cap in enumerate(caps):
            if cap is not None:
                V, F = cap
                normals = vertex_normals(V.astype(f32), F) if smooth else None      # from the stored (float32) positions, as Mesh does
                data.add_shape(f"lens{i}", V.astype(f32), F.astype(np.int32), glass, normals=normals)
                eta.append(float(self.rows0[i, 13]) / float(self.rows0[i, 14]))
                self.cap_slices.append((i, nv, nv + len(V))); self.cap_xy.append(V[:, :2].copy()); nv += len(V)
            V, F = baffles[i]
            data.add_shape(f"baffle{i}", V.astype(f32), F.astype(np.int32), black)
            # the ring's 2 x n vertices (outer edge, then the cap's rim) follow the rim's sag when the surface moves
            self.baffle_slices.append((i, nv, len(V) // 2, V[len(V) // 2:, :2].copy()))
            eta.append(0.0); nv += len(V)
        self.scene = Scene(data, [], PathIntegrator())
        self.faces = self.scene.arrays["indices"]
        self.vertices = torch.tensor(self.scene.arrays["vertices"], device="cuda")
        self.eta = torch.tensor(eta, dtype=torch.float32, device="cuda")
        self.n_surfaces = len(caps)
        self._xy = [torch.tensor(xy, dtype=torch.float64, device="cuda") for xy in self.cap_xy]
        self._faces_t = torch.tensor(self.faces.astype(np.int64), device="cuda")
        self._rim_xy = [torch.tensor(xy, dtype=torch.float64, device="cuda") for (_, _, _, xy) in self.baffle_slices]
        self.smooth = smooth
        self._normals = None

    # ---- one design step: new table -> vertex z from the sag function -> refit
    def update(self, rows):
        """rows: torch CUDA tensor [n,16] (or array): moves every cap's vertices onto the new surfaces and refits the BVH."""
        import torch
        # float64 sag, rounded once to float32 — like the host path that builds a scene from scratch: a float32 sag would add
        # ~4e-9 m of noise per vertex, i.e. ~1e-5 rad of facet-normal noise at this mesh density (a micron on the film)
        rows = torch.as_tensor(rows, device="cuda").double()
        V = self.vertices
        for (i, a, b), xy in zip(self.cap_slices, self._xy):
            V[a:b, 2] = sag(rows[i], xy[:, 0], xy[:, 1], xp=torch).float()
        for (i, a, n, _), xy in zip(self.baffle_slices, self._rim_xy):        # baffles stay sewn to their surface's rim
            z = sag(rows[i], xy[:, 0], xy[:, 1], xp=torch).float()
            V[a:a + n, 2] = z; V[a + n:a + 2 * n, 2] = z
        normals = self._vertex_normals(V) if self.smooth else None
        capi.check(capi.lib.lrt_scene_refit(self.scene.handle, capi.dev_ptr(V), capi.dev_ptr(normals) if normals is not None else None,
                                            capi.stream_ptr()))
        self._normals = normals

    def rebuild(self):
        """Full rebuild of the acceleration structure from the current vertices (host round trip): what the reference does on
        every params.update(); here only needed when a deformation is large enough for the refitted tree to degrade."""
        a = self.scene.arrays
        a["vertices"] = np.ascontiguousarray(self.vertices.detach().cpu().numpy(), f32)
        if getattr(self, "_normals", None) is not None:
            a["vertex_normals"] = np.ascontiguousarray(self._normals.detach().cpu().numpy(), f32)
        self.scene._upload()

    def _vertex_normals(self, V):
        """Mesh::recompute_vertex_normals (src/render/mesh.cpp:321-375: face normals weighted by the angle at the vertex) on the device."""
        import torch
        F = self._faces_t
        V = V.double()          # edge vectors of a float32 mesh carry ~1e-5 relative noise: accumulate the normals in float64
        p = [V[F[:, k]] for k in range(3)]
        n = torch.linalg.cross(p[1] - p[0], p[2] - p[0])
        n = n / n.norm(dim=1, keepdim=True).clamp_min(1e-30)
        out = torch.zeros_like(V)
        for k in range(3):
            d0 = torch.nn.functional.normalize(p[(k + 1) % 3] - p[k], dim=1); d1 = torch.nn.functional.normalize(p[(k + 2) % 3] - p[k], dim=1)
            ang = torch.acos((d0 * d1).sum(1).clamp(-1, 1))
            out.index_add_(0, F[:, k], n * ang[:, None])
        return torch.nn.functional.normalize(out, dim=1).float().contiguous()

    # ---- the differentiable route: table rows -> vertices -> vertex normals -> trace, with the adjoint kernel behind autograd
    def vertices_of(self, rows):
        """Vertex positions [nv,3] (float32, CUDA) as a differentiable function of the table rows (torch tensor [n,16]): what
        update() writes, as an expression — the sag in float64, rounded once."""
        import torch
        rows = torch.as_tensor(rows, device="cuda").double()
        z = self.vertices[:, 2].detach().double().clone()
        for (i, a, b), xy in zip(self.cap_slices, self._xy):
            z[a:b] = sag(rows[i], xy[:, 0], xy[:, 1], xp=torch)
        for (i, a, n, _), xy in zip(self.baffle_slices, self._rim_xy):
            zr = sag(rows[i], xy[:, 0], xy[:, 1], xp=torch)
            z[a:a + n] = zr; z[a + n:a + 2 * n] = zr
        return torch.cat([self.vertices[:, :2].detach().double(), z[:, None]], 1).float()

    def eta_of(self, rows):
        """Relative index of every shape (caps: ior_left / ior_right of their row; baffles: 0 = absorber), differentiable."""
        import torch
        rows = torch.as_tensor(rows, device="cuda").double()
        eta = self.eta.detach().double().clone()
        k = 0
        for i in range(self.n_surfaces):
            if any(ci == i for (ci, _, _) in self.cap_slices):
                eta[k] = rows[i, 13] / rows[i, 14]; k += 1
            k += 1                                                    # the surface's baffle
        return eta.float()

    def trace_ad(self, rows, o, d):
        """trace() as a differentiable function of the table rows (and of the rays): moves the vertices onto the surfaces of
        `rows`, refits the BVH, traces; backward runs lrt_scene_trace_refractive_backward and torch chains its vertex / normal /
        index gradients through the vertex normals (mesh.cpp:321-375) and the sag function down to `rows`.
        The scene must not be moved again between this call and backward(): the adjoint kernel replays the trace."""
        return self.trace_prepared(self.prepare(rows), o, d)

    def prepare(self, rows):
        """(vertices, vertex normals | None, relative indices) of the table `rows`, differentiable: computed once per design
        iteration and shared by every bundle traced through the lens in that iteration (trace_prepared)."""
        V = self.vertices_of(rows)
        return V, (self._vertex_normals(V) if self.smooth else None), self.eta_of(rows)

    def trace_prepared(self, prepared, o, d):
        import torch
        V, N, eta = prepared
        od = torch.cat([torch.as_tensor(o, dtype=torch.float32, device="cuda").reshape(-1, 3),
                        torch.as_tensor(d, dtype=torch.float32, device="cuda").reshape(-1, 3)], 1)
        out = _RefractiveTrace.apply(self, V, N, eta, od)
        return out[:, 0:3], out[:, 3:6], out[:, 6] > 0, out[:, 7]

    def spot_positions_ad(self, rows, src_p, n, seed=0, origin_offset=0.0):
        """spot_positions() with the landing points differentiable in `rows`: (pf [n,2], alive)."""
        ap, d, o = self._spot_rays(src_p, n, seed, origin_offset)
        po, pd, alive, w = self.trace_ad(rows, o, d)
        t = -po[:, 2] / pd[:, 2]
        return (po + t[:, None] * pd)[:, :2], alive & (pd[:, 2] < 0)

    def _spot_rays(self, src_p, n, seed, origin_offset):
        import torch
        g = torch.Generator(device="cuda").manual_seed(seed)
        u = torch.rand((n, 2), device="cuda", generator=g)
        r, phi = float(self.rows0[-1, 2]) * torch.sqrt(u[:, 0]), 2 * np.pi * u[:, 1]
        ap = torch.stack([r * torch.cos(phi), r * torch.sin(phi), torch.full_like(r, float(self.rows0[-1, 1]))], 1)
        src = torch.tensor(np.asarray(src_p, f32), device="cuda")
        d = torch.nn.functional.normalize(ap - src, dim=1)
        o = ap - d * float(origin_offset) if origin_offset > 0 else src.expand_as(ap)
        return ap, d, o

    def trace(self, o, d):
        """Rays (camera space, travelling towards -z) through the meshed surfaces: (o, d, alive, weight) torch tensors."""
        import torch
        od = torch.cat([torch.as_tensor(o, dtype=torch.float32, device="cuda").reshape(-1, 3),
                        torch.as_tensor(d, dtype=torch.float32, device="cuda").reshape(-1, 3)], 1).contiguous()
        out = torch.empty((od.shape[0], 8), dtype=torch.float32, device="cuda")
        capi.check(capi.lib.lrt_scene_trace_refractive(self.scene.handle, C.c_int64(od.shape[0]), capi.dev_ptr(od), capi.dev_ptr(self.eta),
                                                       C.c_int32(self.n_surfaces + 1), capi.dev_ptr(out), capi.stream_ptr()))
        return out[:, 0:3], out[:, 3:6], out[:, 6] > 0, out[:, 7]

    def spot_positions(self, src_p, n, seed=0, origin_offset=0.0):
        """Film-plane (z = 0) landing points of n rays from the point source `src_p` through uniformly sampled points of the front
        element's disk: the geometric spot diagram (what the notebooks' prb_basic geo_tracer draws, design_problem.py:399-428)."""
        ap, d, o = self._spot_rays(src_p, n, seed, origin_offset)
        po, pd, alive, w = self.trace(o, d)
        t = -po[:, 2] / pd[:, 2]
        pf = po + t[:, None] * pd
        return pf[:, :2], alive & (pd[:, 2] < 0), ap, d


def _make_trace_op():
    import torch

    class RefractiveTrace(torch.autograd.Function):
        """lrt_scene_refit + lrt_scene_trace_refractive forward, lrt_scene_trace_refractive_backward backward: the analogue, for the
        mesh-lens trace, of what Dr.Jit's reverse mode does for the notebooks (design_problem.py:495-584)."""

        @staticmethod
        def forward(ctx, lens, V, N, eta, od):
            V = V.detach().contiguous(); od = od.detach().contiguous(); eta = eta.detach().contiguous()
            Nc = N.detach().contiguous() if N is not None else None
            if lens.vertices.data_ptr() != V.data_ptr():      # (several bundles of one iteration share one set of vertices: one refit)
                capi.check(capi.lib.lrt_scene_refit(lens.scene.handle, capi.dev_ptr(V), capi.dev_ptr(Nc) if Nc is not None else None, capi.stream_ptr()))
                lens.vertices, lens._normals = V, Nc
            lens.eta = eta
            out = torch.empty((od.shape[0], 8), dtype=torch.float32, device="cuda")
            capi.check(capi.lib.lrt_scene_trace_refractive(lens.scene.handle, C.c_int64(od.shape[0]), capi.dev_ptr(od), capi.dev_ptr(eta),
                                                           C.c_int32(lens.n_surfaces + 1), capi.dev_ptr(out), capi.stream_ptr()))
            ctx.lens, ctx.has_normals = lens, Nc is not None
            ctx.save_for_backward(V, eta, od)
            ctx.mark_non_differentiable()
            return out

        @staticmethod
        def backward(ctx, grad_out):
            V, eta, od = ctx.saved_tensors
            lens = ctx.lens
            if lens.vertices.data_ptr() != V.data_ptr():
                raise RuntimeError("the mesh lens was moved between trace_ad() and backward(): the adjoint kernel replays the trace")
            grad_out = grad_out.contiguous().float()
            g_od = torch.empty_like(od)
            g_V = torch.zeros_like(V); g_eta = torch.zeros_like(eta)
            g_N = torch.zeros_like(V) if ctx.has_normals else None
            capi.check(capi.lib.lrt_scene_trace_refractive_backward(
                lens.scene.handle, C.c_int64(od.shape[0]), capi.dev_ptr(od), capi.dev_ptr(eta), C.c_int32(lens.n_surfaces + 1),
                capi.dev_ptr(grad_out), capi.dev_ptr(g_od), capi.dev_ptr(g_V), capi.dev_ptr(g_N) if g_N is not None else None,
                capi.dev_ptr(g_eta), capi.stream_ptr()))
            return None, g_V, g_N, g_eta, g_od

    return RefractiveTrace


class _LazyOp:
    """torch is imported on first use (the package imports without it for host-only work)."""
    _op = None

    def apply(self, *a):
        if _LazyOp._op is None:
            _LazyOp._op = _make_trace_op()
        return _LazyOp._op.apply(*a)


_RefractiveTrace = _LazyOp()


def splat_tent(points, weights, res, pitch, centre):
    """The film-plane spot IMAGE of landing points (what the notebooks' perspective sensors see on their diffuse film plane, up
    to a constant): `points` [n,2] (metres, film plane), `weights` [n] -> image [res,res,3] of pixels of size `pitch` centred on
    `centre`, splatted with the tent filter (radius 1: ImageBlock::put, src/render/imageblock.cpp:264-379, continuous weights,
    footprint normalised to 1 like the particle tracer's film). Plain torch: differentiable in the points (the derivative of the
    filter weights is the splat-position adjoint the notebooks' AD propagates) and in the weights."""
    import torch
    pos = (points - torch.as_tensor(centre, dtype=points.dtype, device=points.device)) / pitch + 0.5 * res      # pixel coordinates
    base = torch.floor(pos - 0.5)                                   # the 2 x 2 pixel centres within the tent's radius
    img = torch.zeros((res * res,), dtype=points.dtype, device=points.device)
    fx = pos[:, 0] - (base[:, 0] + 0.5); fy = pos[:, 1] - (base[:, 1] + 0.5)        # in [0, 1): distance to the lower centre
    for dy in (0, 1):
        wy = (1.0 - fy) if dy == 0 else fy
        for dx in (0, 1):
            wx = (1.0 - fx) if dx == 0 else fx
            ix = base[:, 0].long() + dx; iy = base[:, 1].long() + dy
            ok = (ix >= 0) & (ix < res) & (iy >= 0) & (iy < res)
            img = img.index_add(0, (iy * res + ix)[ok], (weights * wx * wy)[ok])
    return img.reshape(res, res, 1).expand(res, res, 3)


class MeshDesignProblem:
    """The notebooks' design loop on the MESH representation of the lens (design_problem.py:355-584 over lens.py's mesh lenses):
    per field a bundle of PARALLEL rays filling the front element — what the notebooks' `directionalarea` disk field sources
    emit (design_problem.py:46-120), which only a mesh lens can receive: a delta direction cannot be connected to the analytic
    lens — traced through the meshed surfaces; loss = sum over fields of the geometric rms spot radius squared about the
    (detached) centroid, mm^2 (rms_loss, losses.py:8-28, on the landing points instead of a splatted image); Adam over table
    entries. One step = sag -> vertices -> vertex normals -> BVH refit -> trace -> loss -> adjoint kernel -> torch chain."""

    def __init__(self, mi, rows0, field_angles, subdiv=5, rays=1 << 16, free=None, lr=3e-5, seed=0, fill=0.95, loss="geometric",
                 resolution=64, pitch=None):
        """loss="geometric": rms radius of the landing points. loss="image": the notebooks' loss itself — rms_loss
        (losses.py:8-28) of the spot IMAGE, the landing points splatted with the tent filter into `resolution`^2 pixels of size
        `pitch` (default: 1/400 of the front semi-diameter; it should resolve the spot) centred on each field's initial centroid; mm^2 either way."""
        import torch
        self.torch, self.mi = torch, mi
        self.loss_kind, self.resolution = loss, int(resolution)
        self.pitch = float(pitch) if pitch else float(np.asarray(rows0, f32)[-1, 2]) / 400.0
        self._centres = None
        self.rows0 = np.ascontiguousarray(rows0, f32).copy()
        self.lens = MeshLens(mi, self.rows0, subdiv=subdiv)
        n = len(self.rows0)
        if free is None:      # every curvature and vertex position of the refracting rows (+ conic and polynomial terms of aspheres)
            free = [(i, c) for i in range(n) for c in ([1, 3] + ([4] + list(range(6, 13)) if self.rows0[i, 0] == 3 else []))
                    if self.rows0[i, 0] != 2]
        self.free = list(free)
        self._rows0_t = torch.tensor(self.rows0, dtype=torch.float64, device="cuda")
        self._idx = (torch.tensor([i for i, _ in self.free], device="cuda"), torch.tensor([c for _, c in self.free], device="cuda"))
        self.theta = torch.tensor([self.rows0[i, c] for i, c in self.free], dtype=torch.float64, device="cuda", requires_grad=True)
        self.opt = torch.optim.Adam([self.theta], lr=lr)
        front_z, R = float(self.rows0[-1, 1]), float(self.rows0[-1, 2])
        g = torch.Generator(device="cuda").manual_seed(seed)
        self.bundles = []
        for a in field_angles:
            a = np.radians(a)
            d = np.array([-np.sin(a), 0.0, -np.cos(a)])
            u = np.array([np.cos(a), 0.0, -np.sin(a)]); v = np.array([0.0, 1.0, 0.0])          # the disk's plane, normal = d
            centre = np.array([0.0, 0.0, front_z]) - d * (2.0 * R)
            s = torch.rand((rays, 2), device="cuda", generator=g)
            r, phi = fill * R * torch.sqrt(s[:, 0]), 2 * np.pi * s[:, 1]
            o = (torch.tensor(centre, device="cuda")[None] + (r * torch.cos(phi))[:, None] * torch.tensor(u, device="cuda")[None]
                 + (r * torch.sin(phi))[:, None] * torch.tensor(v, device="cuda")[None]).float()
            self.bundles.append((o.contiguous(), torch.tensor(d, dtype=torch.float32, device="cuda").expand(rays, 3).contiguous()))
        self.iteration, self.losses = 0, []

    def rows(self):
        return self._rows0_t.index_put(self._idx, self.theta)

    def spots(self, rows):
        """per field (landing points [n,2], alive mask) on the film plane z = 0, differentiable in `rows`"""
        out = []
        prepared = self.lens.prepare(rows)
        for o, d in self.bundles:
            po, pd, alive, w = self.lens.trace_prepared(prepared, o, d)
            t = -po[:, 2] / pd[:, 2]
            out.append(((po + t[:, None] * pd)[:, :2], alive & (pd[:, 2] < 0)))
        return out

    def loss(self, rows):
        # one trace_ad per field; each keeps its own autograd node, and every node replays the SAME scene state (the table
        # does not move inside an iteration), so backward() through all of them is consistent
        total = 0.0
        spots = self.spots(rows)
        if self.loss_kind == "image":
            from .lens_design import rms_loss
            if self._centres is None:        # the windows stay where the first iteration's spots are (the notebooks' sensors do not move)
                self._centres = [pf[alive].detach().double().mean(0) for pf, alive in spots]
            for (pf, alive), c in zip(spots, self._centres):
                img = splat_tent(pf[alive].double(), self.torch.ones(int(alive.sum()), dtype=self.torch.float64, device=pf.device),
                                 self.resolution, self.pitch, c)
                total = total + rms_loss(img) * (self.pitch * 1e3) ** 2
            return total
        for pf, alive in spots:
            q = pf[alive].double() * 1e3                     # mm
            total = total + ((q - q.mean(0).detach()) ** 2).sum(1).mean()
        return total

    def images(self):
        """the spot images of the current table (loss="image" windows), [fields, res, res, 3]"""
        torch = self.torch
        with torch.no_grad():
            spots = self.spots(self.rows())
            cs = self._centres or [pf[alive].double().mean(0) for pf, alive in spots]
            return torch.stack([splat_tent(pf[alive].double(), torch.ones(int(alive.sum()), dtype=torch.float64, device=pf.device),
                                           self.resolution, self.pitch, c) for (pf, alive), c in zip(spots, cs)])

    def step(self):
        self.opt.zero_grad(set_to_none=True)
        l = self.loss(self.rows())
        l.backward()
        self.opt.step()
        self.iteration += 1
        self.losses.append(float(l.detach()))
        return self.losses[-1]
