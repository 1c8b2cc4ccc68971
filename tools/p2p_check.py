#!/usr/bin/env python
"""Direct tile exchange over NVLink peer memory (dist.py mode "tiles_p2p") against the NCCL gather (mode "tiles"):
the root's frame must be byte-identical, then per-frame times of both.  Run on a multi-GPU box:
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tools/p2p_check.py
Writes gpurun_out/p2p_check.json on rank 0.  Test/measurement infrastructure, not product."""
import json, os, sys
ROOT = os.path.abspath(os.path.join(os.path.dirname(__file__), ".."))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import torch.distributed as dist
from pathtracer_b200 import dist as D
from pathtracer_b200 import host, render

rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
torch.cuda.set_device(local)
if world > 1:
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
stream = torch.cuda.Stream()
torch.cuda.set_stream(stream)
out = {"world": world, "cases": []}
for w, (W, H, pp) in [(5, (640, 360, 4)), (1, (1280, 720, 8)), (5, (1920, 1080, 16)), (2, (1280, 720, 16))]:
    sc = host.HostWorld(w).scene(W, H)
    frames, times = {}, {}
    for mode in ("tiles", "tiles_p2p"):
        core = render.RenderCore(local, stream=stream.cuda_stream)
        core.upload_scene(sc)
        fr = D.ShardedFrame(core, W, H, mode)
        for i in range(3):
            res = fr.render(pp, seed=7)
        torch.cuda.synchronize()
        if rank == 0:
            frames[mode] = res.cpu().numpy().copy()
        if world > 1:
            dist.barrier()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for i in range(5):
            fr.render(pp, seed=20 + i)
        e1.record()
        torch.cuda.synchronize()
        t = torch.tensor([e0.elapsed_time(e1) / 5], device="cuda")
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        times[mode] = float(t.item())
        fr.close()
        core.close()
    if rank == 0:
        same = bool(np.array_equal(frames["tiles"], frames["tiles_p2p"]))
        out["cases"].append({"world_scene": w, "size": [W, H], "pp": pp, "identical": same, "ms_per_frame": times,
                             "differing_pixels": int((frames["tiles"] != frames["tiles_p2p"]).sum())})
        print(json.dumps(out["cases"][-1]), flush=True)
if rank == 0:
    os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "gpurun_out", "p2p_check.json"), "w"), indent=1)
if world > 1:
    dist.barrier()
    dist.destroy_process_group()
