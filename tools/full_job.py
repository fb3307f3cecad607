"""The headline configuration rendered ONCE in full: C5 = 3840x2160, 4096 spp (3.4e10 samples > 2^32: eight passes of 512 spp,
src/render/integrator.cpp:247-265), through the public API (mi.render via parallel.render_sharded), on however many ranks
torchrun started. Prints one JSON line with the wall clock (scene build excluded) and writes the developed image's statistics.

  python tools/full_job.py                       (1 GPU, ~41 s)
  python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29533 tools/full_job.py"""
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

spp = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
import lensrt as mi  # noqa: E402

d = mi.workloads.heightfield_scene(n_side=708, width=3840, height=2160, spp=spp, lens_design="gauss", max_depth=6)
scene = mi.load_dict(d)
passes = mi.scene.render_passes(3840 * 2160, spp, 0)
mi.parallel.render_sharded(mi, scene, seed=1, spp=4)          # warm-up (NCCL communicator, kernel load)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
t0 = time.perf_counter()
img = mi.parallel.render_sharded(mi, scene, seed=0, spp=spp)
torch.cuda.synchronize()
if world > 1:
    dist.barrier()
sec = time.perf_counter() - t0
if rank == 0:
    st = scene.stats()
    print(json.dumps({"what": "C5 in full through mi.render / render_sharded", "n_gpus": world, "spp": spp, "passes": [p[1] for p in passes],
                      "seconds": sec, "Mpaths_per_s": 3840 * 2160 * spp / sec / 1e6, "image_mean": float(img.mean()),
                      "image_finite": bool(torch.isfinite(img).all()), "stack_clamps_last_pass_rank0": int(st["stack_clamps"])}))
if world > 1:
    dist.destroy_process_group()
