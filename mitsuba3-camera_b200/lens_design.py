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
                ("shard_index", C.c_int32), ("shard_count", C.c_int32)]


def field_source(angle_deg, distance=1.0e3, azimuth_deg=0.0, front_z=0.0):
    """A distant point emitter at field angle `angle_deg` (camera space): the stand-in for the notebooks'
    directional field sources (design_problem.py:46-120). Returns src_p[3]."""
    a, phi = np.radians(angle_deg), np.radians(azimuth_deg)
    return np.array([distance * np.tan(a) * np.cos(phi), distance * np.tan(a) * np.sin(phi), front_z + distance], f32)


def _film_desc(film):
    rf = ReconstructionFilter(film.rfilter, getattr(film, "rfilter_param", 0.0))
    return FilmDesc(film.crop_w, film.crop_h, film.crop_x, film.crop_y, 0, 4, rf.kind, rf.param, 0, 1)


def _params(seed, spp, src_p, intensity, shard):
    p = SpotParams()
    p.seed, p.spp = int(seed) & 0xffffffff, int(spp)
    for k in range(3):
        p.src_p[k] = float(src_p[k]); p.intensity[k] = float(intensity[k])
    p.shard_index, p.shard_count = int(shard[0]), int(shard[1])
    return p


def spot_render(sensor, src_p, intensity=(1.0, 1.0, 1.0), seed=0, spp=None, shard=(0, 1), storage=None, stats=None):
    """One light-traced spot image: storage [H,W,4] (R,G,B,W accumulators; the image is [..., :3])."""
    import torch
    film = sensor.film
    spp = int(spp if spp else sensor.sampler.sample_count())
    if storage is None:
        storage = torch.zeros((film.crop_h, film.crop_w, 4), dtype=torch.float32, device="cuda")
    fd, p = _film_desc(film), _params(seed, spp, src_p, intensity, shard)
    capi.check(capi.lib.lrt_spot_render(sensor.handle, C.byref(fd), C.byref(p), capi.dev_ptr(storage),
                                        capi.dev_ptr(stats) if stats is not None else None, capi.stream_ptr()))
    return storage


def spot_render_backward(sensor, src_p, grad_storage, intensity=(1.0, 1.0, 1.0), seed=0, spp=None, shard=(0, 1), grad_rows=None,
                         stats=None):
    """Adjoint of spot_render: grad_storage [H,W,4] -> d/d(interface table) [n_if,16] (torch CUDA tensor, accumulated)."""
    import torch
    film = sensor.film
    spp = int(spp if spp else sensor.sampler.sample_count())
    if grad_rows is None:
        grad_rows = torch.zeros((sensor.num_interfaces(), 16), dtype=torch.float32, device="cuda")
    g = grad_storage.contiguous()
    fd, p = _film_desc(film), _params(seed, spp, src_p, intensity, shard)
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


def _spot_function():
    import torch

    class SpotRender(torch.autograd.Function):
        """The analogue of _RenderOp (src/python/python/util.py:356-407) for the spot path: rows16 (CPU or CUDA
        tensor [n_if,16]) -> spot image [H,W,3]; backward replays the samples with seed_grad / spp_grad."""

        @staticmethod
        def forward(ctx, rows, sensor, src_p, intensity, seed, spp, seed_grad, spp_grad, shard):
            sensor.set_interfaces(rows.detach().cpu().numpy())
            st = spot_render(sensor, src_p, intensity, seed, spp, shard)
            ctx.args = (sensor, src_p, intensity, seed_grad, spp_grad, shard, rows.device)
            ctx.rows_np = rows.detach().cpu().numpy().copy()
            return st[..., :3]

        @staticmethod
        def backward(ctx, grad_img):
            sensor, src_p, intensity, seed_grad, spp_grad, shard, dev = ctx.args
            sensor.set_interfaces(ctx.rows_np)
            g = torch.zeros(grad_img.shape[:2] + (4,), dtype=torch.float32, device=grad_img.device)
            g[..., :3] = grad_img
            gr = spot_render_backward(sensor, src_p, g, intensity, seed_grad, spp_grad, shard)
            return (gr.to(dev),) + (None,) * 8

    return SpotRender


_SpotRender = None


def spot_image(rows, sensor, src_p, intensity=(1.0, 1.0, 1.0), seed=0, spp=None, seed_grad=None, spp_grad=None, shard=(0, 1)):
    """Differentiable spot image (torch autograd): d(image)/d(rows) comes from lrt_spot_render_backward.
    seed_grad defaults to seed (same samples in the adjoint: the exact derivative of the primal estimate)."""
    global _SpotRender
    if _SpotRender is None:
        _SpotRender = _spot_function()
    spp = int(spp if spp else sensor.sampler.sample_count())
    return _SpotRender.apply(rows, sensor, src_p, tuple(intensity), int(seed), spp, int(seed if seed_grad is None else seed_grad),
                             int(spp_grad if spp_grad else spp), shard)
