"""Film side of the hot path: ImageBlock (src/render/imageblock.cpp, include/mitsuba/render/imageblock.h)
and HDRFilm's prepare/develop (src/films/hdrfilm.cpp:245-418) over torch CUDA tensors and liblensrt."""
from __future__ import annotations

import ctypes as C
import math

import numpy as np

from . import _capi as capi

FILTERS = {"box": 0, "tent": 1, "gaussian": 2}


class FilmDesc(C.Structure):
    _fields_ = [("size_x", C.c_int32), ("size_y", C.c_int32), ("off_x", C.c_int32), ("off_y", C.c_int32),
                ("border", C.c_int32), ("channels", C.c_int32), ("filter", C.c_int32), ("filter_param", C.c_float),
                ("coalesce", C.c_int32), ("normalize", C.c_int32)]


class ReconstructionFilter:
    """`gaussian` (stddev, default .5 -> radius 2), `tent` (radius, default 1), `box` plugins."""

    def __init__(self, name="gaussian", param=0.0):
        self.name, self.kind, self.param = name, FILTERS[name], float(param)
        r = C.c_float()
        capi.check(capi.lib.lrt_filter_eval_host(C.c_int32(self.kind), capi.c_f(self.param), C.c_int32(0), None, None, C.byref(r)))
        self._radius = r.value

    def radius(self):
        return self._radius

    def border_size(self):   # include/mitsuba/core/rfilter.h: ceil(radius - 1/2 - 2*RayEpsilon)
        return int(math.ceil(self._radius - 0.5 - 2 * 8.94069671630859375e-05))

    def is_box_filter(self):
        return self.kind == 0

    def eval(self, x):
        x = np.ascontiguousarray(np.atleast_1d(np.asarray(x, np.float32)))
        out = np.zeros_like(x)
        capi.check(capi.lib.lrt_filter_eval_host(C.c_int32(self.kind), capi.c_f(self.param), C.c_int32(x.size),
                                                 capi.fp(x.ravel()), capi.fp(out.ravel()), None))
        return out


class ImageBlock:
    """ImageBlock(size, offset, channel_count, rfilter, border, normalize, coalesce) — imageblock.h:52-72."""

    def __init__(self, size, offset=(0, 0), channel_count=4, rfilter: ReconstructionFilter | None = None,
                 border=None, normalize=False, coalesce=True, device="cuda"):
        import torch
        self.rfilter = rfilter if rfilter is not None else ReconstructionFilter("box")
        self._size = (int(size[0]), int(size[1]))
        self._offset = (int(offset[0]), int(offset[1]))
        self.channels = int(channel_count)
        if border is None:
            border = not self.rfilter.is_box_filter()
        self._border = self.rfilter.border_size() if border else 0
        self.normalize, self.coalesce = bool(normalize), bool(coalesce)
        h, w = self._size[1] + 2 * self._border, self._size[0] + 2 * self._border
        self._tensor = torch.zeros((h, w, self.channels), dtype=torch.float32, device=device)

    def desc(self) -> FilmDesc:
        return FilmDesc(self._size[0], self._size[1], self._offset[0], self._offset[1], self._border, self.channels,
                        self.rfilter.kind, self.rfilter.param, int(self.coalesce), int(self.normalize))

    def size(self):
        return self._size

    def offset(self):
        return self._offset

    def set_offset(self, o):
        self._offset = (int(o[0]), int(o[1]))

    def border_size(self):
        return self._border

    def set_coalesce(self, c):
        self.coalesce = bool(c)

    def channel_count(self):
        return self.channels

    def tensor(self):
        return self._tensor

    def clear(self):
        self._tensor.zero_()

    def put(self, pos, values, active=None):
        """ImageBlock::put(pos, values): pos [n,2] film pixel coordinates, values [n,channels] (or a list of
        per-channel arrays, as the reference's Python binding takes)."""
        import torch
        dev = self._tensor.device
        pos = torch.as_tensor(pos, dtype=torch.float32, device=dev).reshape(-1, 2).contiguous()
        if isinstance(values, (list, tuple)):
            values = torch.stack([torch.as_tensor(v, dtype=torch.float32, device=dev).reshape(-1).expand(pos.shape[0])
                                  for v in values], dim=1)
        values = torch.as_tensor(values, dtype=torch.float32, device=dev).reshape(-1, self.channels).contiguous()
        if active is not None:
            m = torch.as_tensor(active, device=dev).reshape(-1)
            pos, values = pos[m].contiguous(), values[m].contiguous()
        d = self.desc()
        capi.check(capi.lib.lrt_film_put(C.byref(d), C.c_int64(pos.shape[0]), capi.dev_ptr(pos), capi.dev_ptr(values),
                                         capi.dev_ptr(self._tensor), capi.stream_ptr()))


def develop_rgb(storage):
    """HDRFilm::develop (pixel_format rgb): [H,W,4] weighted storage -> [H,W,3]."""
    import torch
    h, w, c = storage.shape
    assert c == 4
    out = torch.empty((h, w, 3), dtype=torch.float32, device=storage.device)
    capi.check(capi.lib.lrt_film_develop_rgb(C.c_int64(h * w), capi.dev_ptr(storage.contiguous()), capi.dev_ptr(out), capi.stream_ptr()))
    return out
