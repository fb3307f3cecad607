"""Lens-design loop on the analytic lens table: spot images, their adjoint, the notebooks' loss, an optimiser.

Stands in for the notebooks' DesignProblem.optimize() (notebooks/scripts/lens_design/design_problem.py:495-584):
there, every iteration re-meshes the lens surfaces from their sag functions (lens.py:67-85,210-237), renders one
`ptracer` image per field source and back-propagates the RMS spot loss (losses.py:8-28) through Dr.Jit's tape.
Here the lens is the realisticlens interface table itself; one iteration is, per field source, one
lrt_spot_render (primal), the loss and its image gradient in torch, and one lrt_spot_render_backward that returns
d(loss)/d(table). Design variables map to the table through ordinary torch ops, so torch.autograd finishes the
chain rule (curvatures, thicknesses -> vertex positions, Cauchy / (n_d, V_d) glass models -> refractive indices)."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi as capi
from .film import FilmDesc, ReconstructionFilter

f32 = np.float32


class SpotParams(C.Structure):
    _fields_ = [("seed", C.c_uint32), ("spp", C.c_int32), ("src_p", C.c_float * 3), ("intensity", C.c_float * 3),
                ("shard_index", C.c_int32), ("shard_count", C.c_int32), ("origin_offset", C.c_float), ("fast", C.c_int32),
                ("seed_dev", C.c_void_p), ("seed_stride", C.c_uint32), ("seed_xor", C.c_uint32)]


def field_source(angle_deg, distance=1.0e3, azimuth_deg=0.0, front_z=0.0):
    """A distant point emitter at field angle `angle_deg` (camera space): the stand-in for the notebooks'
    directional field sources (design_problem.py:46-120). Returns src_p[3]."""
    a, phi = np.radians(angle_deg), np.radians(azimuth_deg)
    return np.array([distance * np.tan(a) * np.cos(phi), distance * np.tan(a) * np.sin(phi), front_z + distance], f32)


def _film_desc(film):
    rf = ReconstructionFilter(film.rfilter, getattr(film, "rfilter_param", 0.0))
    return FilmDesc(film.crop_w, film.crop_h, film.crop_x, film.crop_y, 0, 4, rf.kind, rf.param, 0, 1)


class DeviceSeed:
    """seed = (base + counter * stride) ^ xor with `counter` a uint32 CUDA tensor (for loops replayed from a CUDA graph)."""

    def __init__(self, counter, base=0, stride=1, xor=0):
        self.counter, self.base, self.stride, self.xor = counter, int(base), int(stride), int(xor)

    def __xor__(self, mask):
        return DeviceSeed(self.counter, self.base, self.stride, self.xor ^ int(mask))


def _params(seed, spp, src_p, intensity, shard, origin_offset=0.0, fast=False):
    p = SpotParams()
    if isinstance(seed, DeviceSeed):
        p.seed, p.seed_dev, p.seed_stride, p.seed_xor = seed.base & 0xffffffff, seed.counter.data_ptr(), seed.stride & 0xffffffff, seed.xor & 0xffffffff
    else:
        p.seed = int(seed) & 0xffffffff
    p.spp = int(spp)
    for k in range(3):
        p.src_p[k] = float(src_p[k]); p.intensity[k] = float(intensity[k])
    p.shard_index, p.shard_count = int(shard[0]), int(shard[1])
    p.origin_offset = float(origin_offset)
    p.fast = int(bool(fast))
    return p


def spot_render(sensor, src_p, intensity=(1.0, 1.0, 1.0), seed=0, spp=None, shard=(0, 1), storage=None, stats=None,
                origin_offset=0.0, fast=False):
    """One light-traced spot image: storage [H,W,4] (R,G,B,W accumulators; the image is [..., :3])."""
    import torch
    film = sensor.film
    spp = int(spp if spp else sensor.sampler.sample_count())
    if storage is None:
        storage = torch.zeros((film.crop_h, film.crop_w, 4), dtype=torch.float32, device="cuda")
    fd, p = _film_desc(film), _params(seed, spp, src_p, intensity, shard, origin_offset, fast)
    capi.check(capi.lib.lrt_spot_render(sensor.handle, C.byref(fd), C.byref(p), capi.dev_ptr(storage),
                                        capi.dev_ptr(stats) if stats is not None else None, capi.stream_ptr()))
    return storage


def spot_render_backward(sensor, src_p, grad_storage, intensity=(1.0, 1.0, 1.0), seed=0, spp=None, shard=(0, 1), grad_rows=None,
                         stats=None, origin_offset=0.0, fast=False):
    """Adjoint of spot_render: grad_storage [H,W,4] -> d/d(interface table) [n_if,16] (torch CUDA tensor, accumulated)."""
    import torch
    film = sensor.film
    spp = int(spp if spp else sensor.sampler.sample_count())
    if grad_rows is None:
        grad_rows = torch.zeros((sensor.num_interfaces(), 16), dtype=torch.float32, device="cuda")
    g = grad_storage.contiguous()
    fd, p = _film_desc(film), _params(seed, spp, src_p, intensity, shard, origin_offset, fast)
    capi.check(capi.lib.lrt_spot_render_backward(sensor.handle, C.byref(fd), C.byref(p), capi.dev_ptr(g), capi.dev_ptr(grad_rows),
                                                 capi.dev_ptr(stats) if stats is not None else None, capi.stream_ptr()))
    return grad_rows


def rms_loss(image):
    """losses.py:8-28 rms_loss: squared RMS radius (pixels^2) of the spot in `image` [H,W,3] (torch)."""
    import torch
    scaled = image / image.detach().mean()
    I = scaled.sum(-1) / 3.0
    H, W = I.shape
    ii = torch.arange(H, device=I.device, dtype=I.dtype)[:, None]
    jj = torch.arange(W, device=I.device, dtype=I.dtype)[None, :]
    inv = 1.0 / I.sum()
    ibar = ((ii * I).sum() * inv).detach()
    jbar = ((jj * I).sum() * inv).detach()
    return (I * ((ii - ibar) ** 2 + (jj - jbar) ** 2)).sum() * inv


def _set_rows(sensor, rows):
    """The interface table of a design iteration: device tensors stay on the device (lrt_sensor_set_interfaces_dev)."""
    if rows.is_cuda:
        sensor.set_interfaces_dev(rows)
    else:
        sensor.set_interfaces(rows.numpy())


def _all_reduce_image():
    import torch
    import torch.distributed as dist

    class SumOverRanks(torch.autograd.Function):
        """image = sum over ranks of the partial spot images (each rank traces its share of the sample indices). Every rank
        then evaluates the SAME loss on the SAME image, so d(loss)/d(partial image of rank r) = d(loss)/d(image): the
        backward is the identity, and the table gradients (each rank: its own samples) are summed afterwards."""

        @staticmethod
        def forward(ctx, img):
            out = img.contiguous().clone()
            dist.all_reduce(out)
            return out

        @staticmethod
        def backward(ctx, g):
            return g

    return SumOverRanks


_SumOverRanks = None


def sum_over_ranks(img):
    """All-reduce (sum) of a partial image with the autograd rule above; the identity for a single process."""
    global _SumOverRanks
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return img
    if _SumOverRanks is None:
        _SumOverRanks = _all_reduce_image()
    return _SumOverRanks.apply(img)


def _spot_function():
    import torch

    class SpotRender(torch.autograd.Function):
        """The analogue of _RenderOp (src/python/python/util.py:356-407) for the spot path: rows16 (CPU or CUDA
        tensor [n_if,16]) -> spot image [H,W,3]; backward replays the samples with seed_grad / spp_grad."""

        @staticmethod
        def forward(ctx, rows, sensor, src_p, intensity, seed, spp, seed_grad, spp_grad, shard, origin_offset, fast):
            ctx.rows = rows.detach()
            _set_rows(sensor, ctx.rows)
            st = spot_render(sensor, src_p, intensity, seed, spp, shard, origin_offset=origin_offset, fast=fast)
            ctx.args = (sensor, src_p, intensity, seed_grad, spp_grad, shard, rows.device, origin_offset, fast)
            return st[..., :3]

        @staticmethod
        def backward(ctx, grad_img):
            sensor, src_p, intensity, seed_grad, spp_grad, shard, dev, origin_offset, fast = ctx.args
            _set_rows(sensor, ctx.rows)
            g = torch.zeros(grad_img.shape[:2] + (4,), dtype=torch.float32, device=grad_img.device)
            g[..., :3] = grad_img
            gr = spot_render_backward(sensor, src_p, g, intensity, seed_grad, spp_grad, shard, origin_offset=origin_offset, fast=fast)
            return (gr.to(dev),) + (None,) * 10

    return SpotRender


_SpotRender = None


def spot_image(rows, sensor, src_p, intensity=(1.0, 1.0, 1.0), seed=0, spp=None, seed_grad=None, spp_grad=None, shard=(0, 1),
               origin_offset=0.0, fast=False):
    """Differentiable spot image (torch autograd): d(image)/d(rows) comes from lrt_spot_render_backward.
    seed_grad defaults to seed (same samples in the adjoint: the exact derivative of the primal estimate)."""
    global _SpotRender
    if _SpotRender is None:
        _SpotRender = _spot_function()
    spp = int(spp if spp else sensor.sampler.sample_count())
    keep = lambda v: v if isinstance(v, DeviceSeed) else int(v)     # noqa: E731
    return _SpotRender.apply(rows, sensor, src_p, tuple(intensity), keep(seed), spp, keep(seed if seed_grad is None else seed_grad),
                             int(spp_grad if spp_grad else spp), shard, float(origin_offset), bool(fast))


# ------------------------------------------------------------------------------------------------ design problem
def asphere_table(curv_mm, thickness_mm, semi_diameter_mm, glass_iors, air=1.000277):
    """Even-asphere interface table (rear -> front rows16, metres) from a front-to-rear prescription in millimetres, the
    way AsphericalLens's constructor scales it (realisticlens.cpp:557-572: unitless radius r/R_e, c = R_e / curv).
    glass_iors[i] = index of the medium behind surface i (front to rear); the last one is the image space (air)."""
    n = len(curv_mm)
    rows, z = [], 0.0
    media = [air] + list(glass_iors)             # media[i] in front of surface i, media[i+1] behind it
    for i in range(n - 1, -1, -1):
        z += float(thickness_mm[i])
        r = np.zeros(16, f32)
        r[0], r[1], r[2] = 3, 1e-3 * z, 1e-3 * float(semi_diameter_mm[i])
        r[3] = float(semi_diameter_mm[i]) / float(curv_mm[i]) if np.isfinite(curv_mm[i]) and curv_mm[i] != 0 else 0.0
        r[5] = float(semi_diameter_mm[i])
        r[13], r[14] = media[i + 1], media[i]
        rows.append(r)
    return np.stack(rows)


def glass_ior(n_d, V_d, wavelength_um=0.5893):
    """DispersiveDielectric::eval_ior (src/bsdfs/dispersive_dielectric.cpp:425-431): the Cauchy fit through (n_d, V_d) the
    notebooks' lens materials are parameterised by — B = (n_d - 1) / (V_d (1/0.48613^2 - 1/0.65627^2)) (0 if V_d = 0),
    A = n_d - B / 0.5893^2, n(w) = A + B / w^2. Works on floats, numpy arrays and torch tensors (differentiable)."""
    k = 1.0 / (0.48613 * 0.48613) - 1.0 / (0.65627 * 0.65627)
    if hasattr(V_d, "clamp_min"):
        import torch
        B = torch.where(V_d > 5.9604644775390625e-08, (n_d - 1.0) / (V_d.clamp_min(1e-30) * k), torch.zeros_like(V_d))
    else:
        B = np.where(np.asarray(V_d) > 5.9604644775390625e-08, (n_d - 1.0) / (np.maximum(V_d, 1e-30) * k), 0.0)
    A = n_d - B / (0.5893 * 0.5893)
    return A + B / (wavelength_um * wavelength_um)


def cauchy_ior(A, B, wavelength_um=0.5893):
    return A + B / (wavelength_um * wavelength_um)     # realisticlens.cpp:169-179


def paraxial_trace(z, kappa, n_left, n_right, upto=0):
    """Marginal paraxial ray (height 1, parallel to the axis) from the front surface down to surface `upto` (rear = 0):
    returns (y, u) just behind that surface. z, kappa (= curvature seen by a ray travelling towards the film,
    -1/R of the table), n_left/n_right: torch 1-D tensors, rear -> front. Stand-in for the notebooks' ABCD matrices
    (notebooks/scripts/lens_design/lens.py:850-960)."""
    y, u = 1.0, 0.0
    n = len(z)
    for i in range(n - 1, upto - 1, -1):
        u = (n_right[i] * u - y * (n_left[i] - n_right[i]) * kappa[i]) / n_left[i]
        if i > upto:
            y = y + u * (z[i] - z[i - 1])
    return y, u


def efl_bfl(z, kappa, n_left, n_right):
    """Effective focal length and back focal distance (measured from the rear vertex) of the table."""
    y, u = paraxial_trace(z, kappa, n_left, n_right, 0)
    return -1.0 / u, -y / u


def efl_constrained_rear(z, kappa, n_left, n_right, f):
    """Rear-surface vertex position and curvature that give the system the focal length `f` and put its focus on the film
    (z = 0): the role of LensSystem.get_rear_surface_params (lens.py:970-997). Differentiable in every input."""
    y1, u1 = paraxial_trace(z, kappa, n_left, n_right, 1)      # behind surface 1, in the medium n_right[0]
    z0 = (y1 + u1 * z[1]) / (u1 + 1.0 / f)
    y0 = y1 + u1 * (z[1] - z0)
    k0 = (n_right[0] * u1 + n_left[0] / f) / (y0 * (n_left[0] - n_right[0]))
    return z0, k0


class LensDesignProblem:
    """The notebooks' multi-spot design loop (design_problem.py:355-584) on the analytic table.

    variables: dict name -> (row index, column) of free table entries (all start from `rows0`); every field source gets
    its own sensor whose crop window of `resolution`^2 pixels (pixel pitch = film diagonal / full_res) is centred on
    the chief ray. loss = sum over fields of rms_loss (pixel^2) * pitch^2. One `step()` = primal renders at `spp`,
    image-space loss gradient, adjoint renders at `spp_grad` (fresh seed), Adam update."""

    def __init__(self, mi, rows0, field_angles, resolution=128, full_res=4096, diagonal=35.0, spp=256, spp_grad=128,
                 free=None, lr=3e-5, source_distance=1.0e3, rfilter="tent", efl_target=None, shard=(0, 1), film_z=0.0, device="cuda", fast=True,
                 glasses=None, wavelengths=None, free_glass=False, lr_glass=None):
        import torch
        from .lens_designs import InterfaceList
        self.mi, self.torch = mi, torch
        self.rows0 = np.ascontiguousarray(rows0, f32).copy()
        self.n = len(self.rows0)
        self.spp, self.spp_grad, self.shard, self.fast = int(spp), int(spp_grad), shard, bool(fast)
        self.origin_offset = 2.0 * float(self.rows0[-1, 2])      # rays start two front semi-diameters before the front plane
        self.pitch = 1e-3 * diagonal / np.sqrt(2.0) / full_res          # square film: side = diagonal / sqrt(2)
        self.efl_target = efl_target
        if free is None:   # every curvature, vertex position and (for aspheres) conic + polynomial coefficient
            free = [(i, c) for i in range(self.n) for c in ([1, 3] + ([4] + list(range(6, 13)) if self.rows0[i, 0] == 3 else []))
                    if self.rows0[i, 0] != 2]
        if efl_target is not None:
            free = [(i, c) for (i, c) in free if not (i == 0 and c in (1, 3))]   # the rear vertex / curvature are constrained
        self.free = list(free)
        # the design variables, the table they map to and the optimiser state all live on the device: an iteration has no
        # host <-> device round trip (the table reaches the sensors through lrt_sensor_set_interfaces_dev)
        self.device = torch.device(device)
        self.theta = torch.tensor([self.rows0[i, c] for i, c in self.free], dtype=torch.float32, device=self.device, requires_grad=True)
        self.opt = torch.optim.Adam([self.theta], lr=lr, capturable=self.device.type == "cuda")
        self._rows0_t = torch.tensor(self.rows0, device=self.device)
        self._idx = (torch.tensor([i for i, _ in self.free], device=self.device), torch.tensor([c for _, c in self.free], device=self.device))
        self._sph = torch.tensor(self.rows0[:, 0] == 0, device=self.device); self._asp = torch.tensor(self.rows0[:, 0] == 3, device=self.device)
        self._rear_idx = (torch.tensor([0, 0], device=self.device), torch.tensor([1, 3], device=self.device))
        # glass variables and wavelengths (the notebooks' lens materials: `dispersive` BSDFs parameterised by (n_d, V_d),
        # lens.py:343-420; chromatic designs evaluate the spots at several wavelengths). glasses: {medium: (n_d, V_d)} with
        # medium m = the space on the film side of table row m (row m's ior_left, row m-1's ior_right; m = n: object space).
        # The IOR columns of those media then follow glass_ior(n_d, V_d, wavelength) for every wavelength of `wavelengths`
        # (nm; default: the rgb variants' single 589.3) and the loss is the mean over the wavelengths. free_glass=True makes
        # (n_d, V_d) design variables (their gradients arrive through the IOR columns of the adjoint).
        self.wavelengths = [589.3] if wavelengths is None else [float(w) for w in wavelengths]
        if len(self.wavelengths) * len(field_angles) > 64:
            raise ValueError("fields x wavelengths must not exceed 64 (one sample stream each per iteration)")
        self.glass_media = sorted(glasses) if glasses else []
        if self.glass_media:
            g = torch.tensor([[glasses[m][0], glasses[m][1]] for m in self.glass_media], dtype=torch.float32, device=self.device)
            self.glass = g.clone().requires_grad_(bool(free_glass))
            if free_glass:
                self.opt.add_param_group({"params": [self.glass], "lr": lr if lr_glass is None else lr_glass})
            il = [(m, 13) for m in self.glass_media if m < self.n]; ir = [(m - 1, 14) for m in self.glass_media if m >= 1]
            self._glass_idx = (torch.tensor([i for i, _ in il + ir], device=self.device), torch.tensor([c for _, c in il + ir], device=self.device))
            self._glass_src = torch.tensor([self.glass_media.index(m) for m in self.glass_media if m < self.n] +
                                           [self.glass_media.index(m) for m in self.glass_media if m >= 1], device=self.device)
        elif wavelengths is not None and len(self.wavelengths) > 1:
            raise ValueError("several wavelengths need `glasses`: the table itself holds one IOR per medium")
        self.sensors, self.sources = [], []
        probe = mi.RealisticLensCamera(mi.Film(full_res, full_res, diagonal=diagonal, rfilter=rfilter), "custom",
                                       prescription=InterfaceList(self.rows0, film_z=film_z, fine_focus=False))
        s2f = probe.info()["sample_to_film"]
        front_z, front_r = float(self.rows0[-1, 1]), float(self.rows0[-1, 2])
        for a in field_angles:
            src = field_source(a, source_distance, front_z=front_z)
            d = np.array([0, 0, front_z], f32) - src
            d = (d / np.linalg.norm(d)).astype(f32)
            o0 = (np.array([0, 0, front_z], f32) - d * f32(self.origin_offset)).astype(f32)   # same ray, started near the lens
            o, dd, act = probe.trace(o0[None], d[None], from_world=True)
            if not act[0]:
                raise ValueError(f"chief ray of the {a} degree field is blocked")
            t = -o[0, 2] / dd[0, 2]
            pf = o[0] + t * dd[0]
            px = (pf[0] - s2f[0, 3]) / s2f[0, 0] * full_res
            py = (pf[1] - s2f[1, 3]) / s2f[1, 1] * full_res
            cx, cy = int(np.floor(px)) - resolution // 2, int(np.floor(py)) - resolution // 2
            film = mi.Film(full_res, full_res, crop=(resolution, resolution, cx, cy), diagonal=diagonal, rfilter=rfilter)
            s = mi.RealisticLensCamera(film, "custom", prescription=InterfaceList(self.rows0, film_z=film_z, fine_focus=False))
            self.sensors.append(s); self.sources.append(src)
        self.iteration, self.losses = 0, []

    def rows(self, wavelength_nm=None):
        """theta -> interface table (torch, differentiable); applies the EFL constraint when one is set. With glass variables:
        the table at `wavelength_nm` (default: the first of self.wavelengths)."""
        torch = self.torch
        rows = self._rows0_t.index_put(self._idx, self.theta)
        if self.glass_media:
            w = 1e-3 * (self.wavelengths[0] if wavelength_nm is None else float(wavelength_nm))
            ior = glass_ior(self.glass[:, 0], self.glass[:, 1], w)
            rows = rows.index_put(self._glass_idx, ior[self._glass_src])
        if self.efl_target is not None:
            Re = rows[:, 2]
            safe_R = torch.where(self._sph, rows[:, 3], torch.ones_like(Re))
            kappa = torch.where(self._sph, -1.0 / safe_R, torch.where(self._asp, rows[:, 3] / Re, torch.zeros_like(Re)))
            z0, k0 = efl_constrained_rear(rows[:, 1], kappa, rows[:, 13], rows[:, 14], self.efl_target)
            col3 = k0 * Re[0] if self.rows0[0, 0] == 3 else -1.0 / k0
            rows = rows.index_put(self._rear_idx, torch.stack([z0, col3]))
        return rows

    def loss_all_wavelengths(self, seed):
        """mean over self.wavelengths of loss(rows(wavelength)): the polychromatic spot loss"""
        if len(self.wavelengths) == 1:
            return self.loss(self.rows(), seed)
        total = 0.0
        for j, w in enumerate(self.wavelengths):
            total = total + self.loss(self.rows(w), seed, stream=j)
        return total / len(self.wavelengths)

    def loss(self, rows, seed, stream=0):
        """seed: the iteration number (int), or a uint32 CUDA tensor holding it (graph replay). stream: decorrelates the
        samples of the wavelengths of one iteration."""
        total = 0.0
        for k, (s, src) in enumerate(zip(self.sensors, self.sources)):
            k = k + len(self.sensors) * stream          # (fields x wavelengths <= 64 sample streams per iteration)
            sd = DeviceSeed(seed, base=k, stride=64) if hasattr(seed, "data_ptr") else seed * 64 + k
            img = spot_image(rows, s, src, seed=sd, spp=self.spp, seed_grad=sd ^ 0x5bd1e995,
                             spp_grad=self.spp_grad, shard=self.shard, origin_offset=self.origin_offset, fast=self.fast)
            # sample-index sharding: the (non-linear) loss is evaluated on the image summed over ranks, never on a partial one
            img = sum_over_ranks(img) if self.shard[1] > 1 else img
            total = total + rms_loss(img) * (self.pitch * 1e3) ** 2       # mm^2, as evaluate_rms_loss (design_problem.py:268-282)
        return total

    def step(self, all_reduce=None, record_loss=True):
        """One optimisation iteration. Sharded runs (shard[1] > 1): every rank renders its share of the sample indices, the spot
        accumulators are all-reduced BEFORE the loss (loss()), and the table gradients — each rank's own samples against the
        common image gradient — are summed here: the result equals the single-process gradient up to float summation order.
        record_loss=False keeps the loss on the device (no synchronisation in the iteration)."""
        from .parallel import all_reduce_sum
        self.opt.zero_grad(set_to_none=True)
        l = self.loss_all_wavelengths(self.iteration)
        l.backward()
        if self.shard[1] > 1:
            (all_reduce or all_reduce_sum)(self.theta.grad)
            if self.glass_media and self.glass.grad is not None:
                (all_reduce or all_reduce_sum)(self.glass.grad)
        self.opt.step()
        self.iteration += 1
        self.last_loss = l.detach()
        if record_loss:
            self.losses.append(float(self.last_loss))
            return self.losses[-1]
        return self.last_loss

    # ---- the whole iteration as one CUDA graph (ms-scale iterations are launch / host-latency bound)
    def capture(self, warmup=3):
        """Captures rows() -> per-field primal renders -> loss -> adjoint renders -> (all-reduces) -> Adam step into a CUDA graph.
        The iteration counter is a device scalar the spot kernels read their seeds from, so every replay draws fresh samples.
        Afterwards step_graph() replays it; self.last_loss holds the loss of the last replay (device tensor)."""
        torch = self.torch
        self._it_dev = torch.full((1,), self.iteration, dtype=torch.int32, device=self.device)
        side = torch.cuda.Stream()
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):
                self._graph_body()
        torch.cuda.current_stream().wait_stream(side)
        self._graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self._graph):
            self._graph_body()
        return self

    def _graph_body(self):
        from .parallel import all_reduce_sum
        if self.theta.grad is not None:         # (static gradient tensor: created by the first warm-up iteration, zeroed afterwards)
            self.theta.grad.zero_()
        if self.glass_media and self.glass.grad is not None:
            self.glass.grad.zero_()
        l = self.loss_all_wavelengths(self._it_dev)
        l.backward()
        if self.shard[1] > 1:
            all_reduce_sum(self.theta.grad)
            if self.glass_media and self.glass.grad is not None:
                all_reduce_sum(self.glass.grad)
        self.opt.step()
        self._it_dev += 1
        self.last_loss = l.detach()

    def step_graph(self):
        self._graph.replay()
        self.iteration += 1
        return self.last_loss
