"""One forward + one adjoint spot render at the C3 shape (profiling target for ncu; also prints CUDA-event timings)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
import lensrt as mi
import bench
spp = int(sys.argv[1]) if len(sys.argv) > 1 else 64
rows = bench.triplet_table(mi)
prob = mi.lens_design.LensDesignProblem(mi, rows, [3.75], resolution=512, full_res=4096, spp=spp, spp_grad=spp, efl_target=None)
s, src = prob.sensors[0], prob.sources[0]
G = torch.randn((512, 512, 4), device="cuda")
for it in range(3):
    e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
    stats = torch.zeros(2, dtype=torch.int64, device="cuda")
    e[0].record()
    st = mi.spot_render(s, src, seed=it, spp=spp, origin_offset=prob.origin_offset, stats=stats)
    e[1].record()
    gr = mi.spot_render_backward(s, src, G, seed=it, spp=spp, origin_offset=prob.origin_offset)
    e[2].record(); torch.cuda.synchronize()
    n = 512 * 512 * spp
    print(f"forward {e[0].elapsed_time(e[1]):.2f} ms ({n / e[0].elapsed_time(e[1]) / 1e6:.2f} Grays/s)  adjoint {e[1].elapsed_time(e[2]):.2f} ms "
          f"({n / e[1].elapsed_time(e[2]) / 1e6:.2f} Grays/s) valid {stats.cpu().numpy()}")
