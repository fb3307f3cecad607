"""Sensors of the hot path: the fork's `realisticlens` (src/sensors/realisticlens.cpp:824-2466) and
`perspective` (src/sensors/perspective.cpp), as thin Python objects over lrt_sensor handles.

Method names and argument order follow the reference's Sensor interface
(include/mitsuba/render/endpoint.h sample_ray(time, sample1, sample2, sample3)) so that unit tests read
like the reference's sensor tests."""
from __future__ import annotations

import ctypes as C

import numpy as np

from . import _capi as capi
from . import lens_designs as ld

f32 = np.float32


class Film:
    """Geometry part of `hdrfilm` (src/films/hdrfilm.cpp, src/render/film.cpp:30-60): size, crop window,
    the fork's physical `diagonal` (mm), reconstruction filter."""

    def __init__(self, width=768, height=576, crop=None, diagonal=35.0, rfilter="gaussian", rfilter_radius=None,
                 pixel_format="rgb", sample_border=False):
        self.width, self.height = int(width), int(height)
        if crop is None:
            crop = (self.width, self.height, 0, 0)
        self.crop_w, self.crop_h, self.crop_x, self.crop_y = (int(c) for c in crop)
        self.diagonal = float(diagonal)
        self.rfilter = rfilter
        default_r = {"gaussian": 2.0, "tent": 1.0, "box": 0.5}
        self.rfilter_radius = float(rfilter_radius if rfilter_radius is not None else default_r[rfilter])
        self.pixel_format = pixel_format
        self.sample_border = sample_border

    def geom(self) -> capi.FilmGeom:
        return capi.FilmGeom(self.width, self.height, self.crop_w, self.crop_h, self.crop_x, self.crop_y, self.diagonal)

    def size(self):
        return (self.width, self.height)

    def crop_size(self):
        return (self.crop_w, self.crop_h)

    def crop_offset(self):
        return (self.crop_x, self.crop_y)


class Sensor:
    handle = None

    def __del__(self):
        if getattr(self, "handle", None):
            if capi is not None and capi.lib is not None:
                capi.lib.lrt_sensor_destroy(self.handle)
            self.handle = None

    def needs_aperture_sample(self) -> bool:
        return False

    def shutter_open(self):
        return 0.0

    def shutter_open_time(self):
        return 0.0

    # ---- Sensor::sample_ray for arrays of samples; returns (o[n,3], d[n,3], maxt[n], weight[n,3])
    def sample_ray(self, time, sample1, sample2, sample3, active=True):
        import torch
        pos = torch.as_tensor(sample2, dtype=torch.float32).reshape(-1, 2)
        n = pos.shape[0]
        ap = torch.full((n, 2), 0.5) if sample3 is None else torch.as_tensor(sample3, dtype=torch.float32).reshape(-1, 2)
        if pos.is_cuda:
            ap = ap.to(pos.device)
            out = torch.empty((n, 10), dtype=torch.float32, device=pos.device)
            capi.check(capi.lib.lrt_sensor_sample_ray(self.handle, C.c_int64(n), capi.dev_ptr(pos.contiguous()),
                                                      capi.dev_ptr(ap.contiguous()), capi.dev_ptr(out), capi.stream_ptr()))
        else:
            out_np = np.empty((n, 10), f32)
            capi.check(capi.lib.lrt_sensor_sample_ray_host(self.handle, C.c_int64(n),
                                                           capi.fp(np.ascontiguousarray(pos.numpy())),
                                                           capi.fp(np.ascontiguousarray(ap.numpy())), capi.fp(out_np)))
            out = torch.from_numpy(out_np)
        return out[:, 0:3], out[:, 3:6], out[:, 6], out[:, 7:10]


def _to_world_array(to_world):
    m = np.eye(4, dtype=f32) if to_world is None else np.asarray(getattr(to_world, "matrix", to_world), dtype=f32).reshape(4, 4)
    return np.ascontiguousarray(m)


class RealisticLensCamera(Sensor):
    """`realisticlens` plugin. Properties as in realisticlens.cpp:844-851,874-878: lens_design,
    object_distance, lens_focal_length, lens_diameter, stop_ratio, ap_size, ap_offset (+ to_world,
    near_clip, far_clip of ProjectiveCamera, src/render/sensor.cpp:127-141).
    New (the reference hard-codes its lenses): lens_design="custom" with `prescription=` a
    lens_designs.Prescription / InterfaceList."""

    def __init__(self, film: Film, lens_design="singlet", to_world=None, near_clip=1e-2, far_clip=1e4,
                 prescription=None, **props):
        self.film = film
        self.lens_design = lens_design
        self.props = dict(props)
        self.near_clip, self.far_clip = float(near_clip), float(far_clip)
        self.to_world = _to_world_array(to_world)
        # the inverse the scene description carries next to the matrix (Transform4f), else inverted numerically
        self.to_camera = (np.ascontiguousarray(to_world.inverse_matrix, f32) if hasattr(to_world, "inverse_matrix")
                          else np.linalg.inv(self.to_world.astype(np.float64)).astype(f32))
        self.object_distance = float(props.get("object_distance", 6.0))
        # ProjectiveCamera / Sensor: shutter_open, shutter_close (src/render/sensor.cpp:27-33); the pupil-mode weight
        # is scaled by shutter_open_time() = close - open, which is 0 unless the scene sets them (as in the reference)
        self._shutter_open = float(props.get("shutter_open", 0.0))
        self._shutter_open_time = float(props.get("shutter_close", self._shutter_open)) - self._shutter_open
        self.sample_exit_pupil = bool(props.get("sample_exit_pupil", False))
        self.spectral = bool(props.get("spectral", False))
        d = prescription if prescription is not None else ld.design(lens_design, props)
        self.design = d
        h = C.c_void_p()
        g = film.geom()
        if hasattr(d, "rows"):
            rows = np.ascontiguousarray(d.rows, f32)
            rc = capi.lib.lrt_sensor_create_interfaces(
                C.c_int32(len(rows)), capi.fp(rows), capi.c_f(d.film_z), C.c_int32(int(d.fine_focus)),
                capi.c_f(self.object_distance), C.byref(g), capi.fp(self.to_world), capi.c_f(self.near_clip),
                capi.c_f(self.far_clip), C.byref(h))
        else:
            # spectral=True: the table a `*_spectral*` variant of the fork builds (nominal IORs at 0.001f * 589.3f um)
            create = capi.lib.lrt_sensor_create_prescription_spectral if self.spectral else capi.lib.lrt_sensor_create_prescription
            rc = create(
                C.c_int32(len(d.kinds)), capi.ip(d.kinds), capi.fp(d.thickness), capi.fp(d.radius),
                capi.fp(d.semi_diameter), capi.fp(d.kappa), capi.fp(np.ascontiguousarray(d.ai)),
                capi.fp(np.ascontiguousarray(d.materials)), capi.c_f(d.scale), capi.c_f(d.stop_ratio),
                C.c_int32(int(d.fine_focus)), capi.c_f(self.object_distance), C.byref(g), capi.fp(self.to_world),
                capi.c_f(self.near_clip), capi.c_f(self.far_clip), C.byref(h))
        capi.check(rc)
        self.handle = h
        if hasattr(d, "rows") and getattr(d, "materials", None) is not None:
            self.set_materials(d.materials)
        # realisticlens.cpp:913-921: build the exit-pupil LUT (a GPU kernel here). An open shutter is registered either way:
        # render_sample draws a time sample whenever shutter_open_time() > 0 (integrator.cpp:442-444)
        if self.sample_exit_pupil or self._shutter_open_time != 0.0:
            fn = capi.lib.lrt_sensor_set_exit_pupil_spectral if self.spectral else capi.lib.lrt_sensor_set_exit_pupil
            capi.check(fn(self.handle, C.c_int32(int(bool(self.sample_exit_pupil))),
                                                          capi.c_f(self._shutter_open_time), None))

    def needs_aperture_sample(self) -> bool:   # m_needs_sample_3 = true (:840)
        return True

    def shutter_open(self):
        return self._shutter_open

    def shutter_open_time(self):
        return self._shutter_open_time

    def set_exit_pupil(self, enable=True, shutter_open_time=None):
        """Switch `sample_exit_pupil` on an existing sensor (builds the LUT when enabling)."""
        if shutter_open_time is not None:
            self._shutter_open_time = float(shutter_open_time)
        fn = capi.lib.lrt_sensor_set_exit_pupil_spectral if self.spectral else capi.lib.lrt_sensor_set_exit_pupil
        capi.check(fn(self.handle, C.c_int32(int(enable)), capi.c_f(self._shutter_open_time), None))
        self.sample_exit_pupil = bool(enable)

    def exit_pupil_bounds(self) -> np.ndarray:
        """The 64 x (min.x, min.y, max.x, max.y) LUT of compute_exit_pupil_bounds (realisticlens.cpp:1769-1809), metres."""
        out = np.zeros((64, 4), f32)
        capi.check(capi.lib.lrt_sensor_get_exit_pupil(self.handle, capi.fp(out)))
        return out

    def num_interfaces(self) -> int:
        return capi.lib.lrt_sensor_num_interfaces(self.handle)

    def interfaces(self) -> np.ndarray:
        """Interface table after autofocus, rear -> front, rows of 16 floats (include/lensrt.h)."""
        rows = np.zeros((self.num_interfaces(), 16), f32)
        capi.check(capi.lib.lrt_sensor_get_interfaces(self.handle, capi.fp(rows)))
        return rows

    def set_interfaces(self, rows):
        rows = np.ascontiguousarray(rows, f32)
        capi.check(capi.lib.lrt_sensor_set_interfaces(self.handle, C.c_int32(len(rows)), capi.fp(rows)))

    def set_interfaces_dev(self, rows):
        """parameters_changed() with the table on the device (torch CUDA tensor [n,16] float32): no host round trip."""
        assert rows.is_cuda and rows.dtype.is_floating_point and rows.shape[-1] == 16
        rows = rows.detach().contiguous().float()
        capi.check(capi.lib.lrt_sensor_set_interfaces_dev(self.handle, C.c_int32(rows.shape[0]), capi.dev_ptr(rows), capi.stream_ptr()))
        self._rows_dev = rows        # keeps the tensor alive until the kernel has read it

    def set_to_world(self, to_world):
        self.to_world = _to_world_array(to_world)
        self.to_camera = (np.ascontiguousarray(to_world.inverse_matrix, f32) if hasattr(to_world, "inverse_matrix")
                          else np.linalg.inv(self.to_world.astype(np.float64)).astype(f32))
        capi.check(capi.lib.lrt_sensor_set_to_world(self.handle, capi.fp(self.to_world)))

    def info(self):
        m = np.zeros(16, f32)
        fd, fz = C.c_float(), C.c_float()
        capi.check(capi.lib.lrt_sensor_get_info(self.handle, capi.fp(m), C.byref(fd), C.byref(fz)))
        return {"sample_to_film": m.reshape(4, 4), "focus_delta": fd.value, "film_z": fz.value}

    # ---- dispersion (the fork's spectral variants: every ray carries a wavelength, realisticlens.cpp:169-183)
    def set_materials(self, mat_lr):
        """Cauchy / Sellmeier rows of the two media of every interface, [n_if, 2, 7] (left = film side, right), for sensors
        built from ready-made interfaces; sensors built from a prescription carry them already."""
        m = np.ascontiguousarray(mat_lr, f32).reshape(-1, 14)
        capi.check(capi.lib.lrt_sensor_set_materials(self.handle, C.c_int32(len(m)), capi.fp(m)))

    def materials(self):
        m = np.zeros((self.num_interfaces(), 2, 7), f32)
        capi.check(capi.lib.lrt_sensor_get_materials(self.handle, capi.fp(m.reshape(-1))))
        return m

    def sample_ray_wavelength(self, sample2, sample3, wavelength_nm):
        """Sensor::sample_ray of a spectral variant for [n,2] film / aperture samples and one wavelength (nm) per ray (the first
        entry of the reference's wavelength vector decides the IORs): (o, d, maxt, weight) torch CUDA tensors; the weight is
        the geometric one (multiply by sample_wavelengths()' own)."""
        import torch
        pos = torch.as_tensor(sample2, dtype=torch.float32, device="cuda").reshape(-1, 2).contiguous()
        ap = torch.as_tensor(sample3, dtype=torch.float32, device="cuda").reshape(-1, 2).contiguous()
        wl = torch.as_tensor(wavelength_nm, dtype=torch.float32, device="cuda").reshape(-1).expand(pos.shape[0]).contiguous()
        out = torch.empty((pos.shape[0], 10), dtype=torch.float32, device="cuda")
        capi.check(capi.lib.lrt_sensor_sample_ray_wl(self.handle, C.c_int64(pos.shape[0]), capi.dev_ptr(pos), capi.dev_ptr(ap), capi.dev_ptr(wl),
                                                     capi.dev_ptr(out), capi.stream_ptr()))
        return out[:, 0:3], out[:, 3:6], out[:, 6], out[:, 7]

    def trace_wavelength(self, origins, directions, wavelength_nm, from_world=False, device=True):
        """trace_ray_from_film / trace_ray_from_world with per-ray wavelengths (nm), camera space: (o, d, active).
        device=True runs the kernel (torch CUDA tensors in and out); device=False the host twin (numpy)."""
        if not device:
            od = np.ascontiguousarray(np.concatenate([np.asarray(origins, f32).reshape(-1, 3), np.asarray(directions, f32).reshape(-1, 3)], axis=1))
            wl = np.ascontiguousarray(np.broadcast_to(np.asarray(wavelength_nm, f32).reshape(-1), (len(od),)))
            out = np.zeros((len(od), 7), f32)
            capi.check(capi.lib.lrt_sensor_trace_wl_host(self.handle, C.c_int32(int(from_world)), C.c_int64(len(od)), capi.fp(od), capi.fp(wl), capi.fp(out)))
            return out[:, 0:3], out[:, 3:6], out[:, 6] > 0
        import torch
        od = torch.cat([torch.as_tensor(origins, dtype=torch.float32, device="cuda").reshape(-1, 3),
                        torch.as_tensor(directions, dtype=torch.float32, device="cuda").reshape(-1, 3)], 1).contiguous()
        wl = torch.as_tensor(wavelength_nm, dtype=torch.float32, device="cuda").reshape(-1).expand(od.shape[0]).contiguous()
        out = torch.empty((od.shape[0], 7), dtype=torch.float32, device="cuda")
        capi.check(capi.lib.lrt_sensor_trace_wl(self.handle, C.c_int32(int(from_world)), C.c_int64(od.shape[0]), capi.dev_ptr(od), capi.dev_ptr(wl),
                                                capi.dev_ptr(out), capi.stream_ptr()))
        return out[:, 0:3], out[:, 3:6], out[:, 6] > 0

    def trace(self, origins, directions, from_world=False):
        """trace_ray_from_film / trace_ray_from_world in camera space (host; chief-ray queries)."""
        od = np.ascontiguousarray(np.concatenate([np.asarray(origins, f32).reshape(-1, 3),
                                                  np.asarray(directions, f32).reshape(-1, 3)], axis=1))
        out = np.zeros((len(od), 7), f32)
        capi.check(capi.lib.lrt_sensor_trace_host(self.handle, C.c_int32(int(from_world)), C.c_int64(len(od)),
                                                  capi.fp(od), capi.fp(out)))
        return out[:, 0:3], out[:, 3:6], out[:, 6] > 0


    def sample_direction(self, it_p, sample, active=True):
        """Sensor::sample_direction (realisticlens.cpp:2301-2394) for [n,3] reference points (world space) and
        [n,2] aperture samples -> dict(uv, p, d, dist, pdf, weight); p/d are returned in world space."""
        import torch
        p = np.asarray(it_p.cpu() if hasattr(it_p, "cpu") else it_p, f32).reshape(-1, 3)
        s2 = np.ascontiguousarray(np.asarray(sample.cpu() if hasattr(sample, "cpu") else sample, f32).reshape(-1, 2))
        tw = self.to_world.astype(np.float64)
        inv = np.linalg.inv(tw)
        pc = np.ascontiguousarray((p.astype(np.float64) @ inv[:3, :3].T + inv[:3, 3]).astype(f32))
        out = np.zeros((len(pc), 11), f32)
        capi.check(capi.lib.lrt_sensor_sample_direction_host(self.handle, C.c_int64(len(pc)), capi.fp(pc), capi.fp(s2), capi.fp(out)))
        R = tw[:3, :3].astype(f32)
        valid = out[:, 9] > 0
        return {"uv": out[:, 0:2], "p": np.where(valid[:, None], out[:, 2:5] @ R.T + tw[:3, 3].astype(f32), 0), "d": out[:, 5:8] @ R.T,
                "dist": out[:, 8], "pdf": out[:, 9], "weight": out[:, 10]}

    def trace_vjp(self, origins, directions, bar_o, bar_d, from_world=False):
        """Vector-Jacobian product of one lens trace (host): returns (adjoint of the incoming ray [n,6], gradient
        w.r.t. the interface table [n_if,16] summed over rays, active mask)."""
        od = np.ascontiguousarray(np.concatenate([np.asarray(origins, f32).reshape(-1, 3),
                                                  np.asarray(directions, f32).reshape(-1, 3)], axis=1))
        bar = np.ascontiguousarray(np.concatenate([np.asarray(bar_o, f32).reshape(-1, 3),
                                                   np.asarray(bar_d, f32).reshape(-1, 3)], axis=1))
        out = np.zeros((len(od), 6), f32)
        grad = np.zeros((self.num_interfaces(), 16), f32)
        act = np.zeros(len(od), np.int32)
        capi.check(capi.lib.lrt_sensor_trace_vjp_host(self.handle, C.c_int32(int(from_world)), C.c_int64(len(od)), capi.fp(od),
                                                      capi.fp(bar), capi.fp(out), capi.fp(grad), capi.ip(act)))
        return out, grad, act > 0


class PerspectiveCamera(Sensor):
    """`perspective` plugin (src/sensors/perspective.cpp): fov (x axis, degrees), to_world, clips."""

    def __init__(self, film: Film, fov=45.0, to_world=None, near_clip=1e-2, far_clip=1e4, fov_axis="x"):
        self.film = film
        self.near_clip, self.far_clip = float(near_clip), float(far_clip)
        self.to_world = _to_world_array(to_world)
        # the inverse the scene description carries next to the matrix (the particle tracer's sample_direction uses it)
        self.to_camera = (np.ascontiguousarray(to_world.inverse_matrix, f32) if hasattr(to_world, "inverse_matrix")
                          else np.linalg.inv(self.to_world.astype(np.float64)).astype(f32))
        aspect = film.width / film.height
        fov = float(fov)
        if fov_axis == "smaller":
            fov_axis = "y" if aspect > 1 else "x"
        elif fov_axis == "larger":
            fov_axis = "x" if aspect > 1 else "y"
        if fov_axis == "y":   # src/render/sensor.cpp:187-190
            fov = np.degrees(2.0 * np.arctan(np.tan(0.5 * np.radians(fov)) * aspect))
        elif fov_axis == "diagonal":
            diag = 2.0 * np.tan(0.5 * np.radians(fov))
            width = diag / np.sqrt(1.0 + 1.0 / (aspect * aspect))
            fov = np.degrees(2.0 * np.arctan(width * 0.5))
        self.x_fov = fov
        h = C.c_void_p()
        g = film.geom()
        capi.check(capi.lib.lrt_sensor_create_perspective(capi.c_f(fov), C.byref(g), capi.fp(self.to_world),
                                                          capi.c_f(self.near_clip), capi.c_f(self.far_clip), C.byref(h)))
        self.handle = h

    def set_to_world(self, to_world):
        self.to_world = _to_world_array(to_world)
        self.to_camera = (np.ascontiguousarray(to_world.inverse_matrix, f32) if hasattr(to_world, "inverse_matrix")
                          else np.linalg.inv(self.to_world.astype(np.float64)).astype(f32))
        capi.check(capi.lib.lrt_sensor_set_to_world(self.handle, capi.fp(self.to_world)))
