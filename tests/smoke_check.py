"""__graft_entry__.smoke(): one small invocation of the hot path on cuda:0, checked against the oracle."""
import numpy as np

from oracle_lib import *  # noqa


def run():
    import lensrt as mi
    ld = load_lens_designs()
    T = mi.Transform4f
    tw = T.look_at(origin=[0, 0, 3.9], target=[0, 0, 0], up=[0, 1, 0])
    d = mi.cornell_box()
    d["integrator"]["max_depth"] = 6
    d["sensor"] = {"type": "realisticlens", "lens_design": "doublet", "to_world": tw, "object_distance": 3.9, "near_clip": 1e-3,
                   "far_clip": 100.0, "sampler": {"type": "independent", "sample_count": 8},
                   "film": {"type": "hdrfilm", "width": 64, "height": 64, "diagonal": 30.0, "rfilter": {"type": "gaussian"}}}
    sc = mi.load_dict(d)
    img = mi.render(sc, seed=0).cpu().numpy()
    film = FilmSpec(64, 64, None, 30.0)
    cam = orc_camera(ld.design("doublet", {"object_distance": 3.9}), film, to_world=tw.matrix, near=1e-3, far=100.0, object_distance=3.9)
    st, miss = orc_render(orc_scene(sc.arrays), film, 8, seed=0, max_depth=6, cam_h=cam)
    want = develop(st)
    err = rel_rmse(img, want)
    stats = sc.stats()
    assert stats["lens_misses"] == miss, (stats, miss)
    assert err < 2e-3, err
    print(f"smoke ok: 64x64x8 Cornell box through the doublet lens, rel-RMSE vs oracle {err:.2e}, lens misses {miss} (bit-equal), "
          f"kernels launched {mi.capi.lib.lrt_launch_count()}")
