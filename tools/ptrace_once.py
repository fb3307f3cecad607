"""bench.py's particle-tracer block on its own (headline scene and lens, 1024x576 film, 4 spp): the command tools/ncu_ptrace.sh
profiles. Prints M light paths/s of 5 renders."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

import lensrt as mi  # noqa: E402

d = mi.workloads.heightfield_scene(n_side=708, width=3840, height=2160, spp=2, lens_design="gauss", max_depth=6)
scene = mi.load_dict(d)
s0 = scene.sensors()[0]
film = mi.Film(1024, 576, diagonal=s0.film.diagonal, rfilter="tent")
sens = mi.RealisticLensCamera(film, s0.lens_design, to_world=mi.Transform4f(s0.to_world), near_clip=s0.near_clip, far_clip=s0.far_clip, **s0.props)
sens.sampler = mi.IndependentSampler(4, 0)
it = mi.PTracerIntegrator(6, 5, False)
for i in range(3):
    it.render(scene, sensor=sens, seed=i, spp=4)
torch.cuda.synchronize()
e = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
e[0].record()
for i in range(5):
    it.render(scene, sensor=sens, seed=10 + i, spp=4)
e[1].record(); torch.cuda.synchronize()
ms = e[0].elapsed_time(e[1]) / 5
print(f"ptracer: {1024 * 576 * 4 / ms / 1e3:.1f} M light paths/s ({ms:.2f} ms per render)", scene.stats())
