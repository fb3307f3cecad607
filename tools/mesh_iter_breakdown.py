import sys, torch, numpy as np
sys.path.insert(0, '/root/repo'); import lensrt as mi
sys.path.insert(0, '/root/repo'); import bench
rows = bench.doublet_table(mi)
lens = mi.mesh_lens.MeshLens(mi, rows, subdiv=6)
src = mi.lens_design.field_source(2.0, 1.0e3, front_z=float(rows[-1, 1])); off = 2.0 * float(rows[-1, 2])
ap, d, o = lens._spot_rays(src, 1 << 20, 1, off)
rt = torch.tensor(rows, dtype=torch.float64, device="cuda", requires_grad=True)
def T(fn, reps=5):
    for _ in range(2): fn()
    torch.cuda.synchronize(); e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps): r = fn()
    e1.record(); torch.cuda.synchronize(); return e0.elapsed_time(e1) / reps
print("vertices_of          %.2f ms" % T(lambda: lens.vertices_of(rt)))
V = lens.vertices_of(rt)
print("vertex normals fwd   %.2f ms" % T(lambda: lens._vertex_normals(V)))
def fb_normals():
    Vd = V.detach().requires_grad_(True); N = lens._vertex_normals(Vd); N.sum().backward()
print("vertex normals f+b   %.2f ms" % T(fb_normals))
print("prepare              %.2f ms" % T(lambda: lens.prepare(rt)))
def full():
    rt.grad = None
    po, pd, alive, w = lens.trace_ad(rt, o, d)
    t = -po[:, 2] / pd[:, 2]; pf = (po + t[:, None] * pd)[:, :2][alive]
    ((pf - pf.mean(0)) ** 2).sum(1).mean().backward()
print("full iteration       %.2f ms" % T(full))
def fwd_only():
    with torch.no_grad():
        po, pd, alive, w = lens.trace_ad(rt, o, d)
print("fwd (prepare+refit+trace) %.2f ms" % T(fwd_only))
print("trace only           %.2f ms" % T(lambda: lens.trace(o, d)))
from torch.profiler import profile, ProfilerActivity
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    for _ in range(3): full()
    torch.cuda.synchronize()
rowsp = sorted(prof.key_averages(), key=lambda e: -e.device_time_total)[:8]
for e in rowsp:
    print(f"{e.key[:70]:70s} calls {e.count:4d}  cuda total {e.device_time_total / 3e3:.3f} ms/iter")
import time
t0 = time.perf_counter()
for _ in range(5): full()
torch.cuda.synchronize(); print("wall per iteration %.2f ms" % ((time.perf_counter() - t0) / 5 * 1e3))
