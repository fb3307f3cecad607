"""lrt_render_rgb_noalloc (include/lensrt.h; csrc/lrt_noalloc.cu): render + develop with the film accumulator in a fixed device
array inside the library — an entry point for callers that cannot call cudaMalloc. Called through the C ABI, twice in a row,
and compared with the existing entry points (lrt_render into a caller-owned accumulator + lrt_film_develop_rgb): the second call
must not see the first call's film. Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest


def rel_rmse(a, b):
    a = np.asarray(a, np.float64); b = np.asarray(b, np.float64)
    return float(np.sqrt(np.mean((a - b) ** 2)) / max(np.sqrt(np.mean(b ** 2)), 1e-12))


def small_box(mi, res=64, spp=16):
    d = mi.cornell_box()
    d["sensor"]["film"]["width"] = d["sensor"]["film"]["height"] = res
    d["sensor"]["sampler"]["sample_count"] = spp
    d["integrator"]["max_depth"] = 5
    return mi.load_dict(d)


def reference_image(mi, sc, rp, fd, npix):
    """the existing entry points: lrt_render into a caller-owned accumulator, then lrt_film_develop_rgb"""
    import torch
    capi = mi.capi
    s = sc.sensors()[0]
    acc = torch.zeros((npix, 4), dtype=torch.float32, device="cuda")
    rgb = torch.empty((npix, 3), dtype=torch.float32, device="cuda")
    capi.check(capi.lib.lrt_render(sc.handle, s.handle, C.byref(fd), C.byref(rp), capi.dev_ptr(acc), None))
    capi.check(capi.lib.lrt_film_develop_rgb(C.c_int64(npix), capi.dev_ptr(acc), capi.dev_ptr(rgb), None))
    torch.cuda.synchronize()
    return rgb.cpu().numpy()


@pytest.mark.gpu
def test_gpu_render_rgb_noalloc_matches_render_twice_in_a_row(mi):
    import torch
    capi = mi.capi
    lib = capi.lib
    lib.lrt_render_rgb_noalloc_max_pixels.restype = C.c_int64
    assert lib.lrt_render_rgb_noalloc_max_pixels() == 1 << 18
    sc = small_box(mi)
    s = sc.sensors()[0]
    fd = mi.scene.film_desc(s.film)
    npix = s.film.crop_w * s.film.crop_h
    out = torch.empty((npix, 3), dtype=torch.float32, device="cuda")      # the caller's output buffer, allocated at set-up
    # two calls in a row with DIFFERENT seeds: the second must start from a cleared accumulator
    for seed in (11, 12):
        rp = mi.RenderParams(seed, 16, 5, 5, 0, 1, 0, 0, 0, 0, 0)
        want = reference_image(mi, sc, rp, fd, npix)                       # (also the sensor's first use: it uploads its table)
        out.fill_(-1.0)
        capi.check(lib.lrt_render_rgb_noalloc(sc.handle, s.handle, C.byref(fd), C.byref(rp), capi.dev_ptr(out), None))
        torch.cuda.synchronize()
        got = out.cpu().numpy()
        assert np.isfinite(got).all() and got.min() >= 0.0
        # same samples, same kernel; only the order of the float atomics differs (the tolerance of tests/test_render.py)
        err = rel_rmse(got, want)
        assert err <= 1e-5, (seed, err)
        assert np.abs(got - want).max() <= 1e-4 * max(1.0, float(want.max()))
    # the same seed again gives the same image again (nothing leaks from call to call)
    capi.check(lib.lrt_render_rgb_noalloc(sc.handle, s.handle, C.byref(fd), C.byref(rp), capi.dev_ptr(out), None))
    torch.cuda.synchronize()
    assert rel_rmse(out.cpu().numpy(), got) <= 1e-5


@pytest.mark.gpu
def test_gpu_render_rgb_noalloc_limits(mi):
    import torch
    capi = mi.capi
    lib = capi.lib
    sc = small_box(mi, res=600, spp=1)        # 360 000 pixels > 2^18
    s = sc.sensors()[0]
    fd = mi.scene.film_desc(s.film)
    rp = mi.RenderParams(0, 1, 2, 5, 0, 1, 0, 0, 0, 0, 0)
    out = torch.empty((600 * 600, 3), dtype=torch.float32, device="cuda")
    acc = torch.zeros((600 * 600, 4), dtype=torch.float32, device="cuda")
    capi.check(lib.lrt_render(sc.handle, s.handle, C.byref(fd), C.byref(rp), capi.dev_ptr(acc), None))
    rc = lib.lrt_render_rgb_noalloc(sc.handle, s.handle, C.byref(fd), C.byref(rp), capi.dev_ptr(out), None)
    assert rc != 0 and b"at most 262144" in lib.lrt_last_error()
    # a sensor that was never used has no device table yet: refused with an instruction, not allocated behind the caller's back
    sc2 = small_box(mi, res=32, spp=1)
    s2 = sc2.sensors()[0]
    fd2 = mi.scene.film_desc(s2.film)
    rc = lib.lrt_render_rgb_noalloc(sc2.handle, s2.handle, C.byref(fd2), C.byref(rp), capi.dev_ptr(out), None)
    assert rc != 0 and b"has not been uploaded" in lib.lrt_last_error()
    rp1 = mi.RenderParams(0, 1, 2, 5, 0, 1, 0, 1, 0, 0, 0)
    rc = lib.lrt_render_rgb_noalloc(sc2.handle, s2.handle, C.byref(fd2), C.byref(rp1), capi.dev_ptr(out), None)
    assert rc != 0
