"""Developer tool (GPU box, torchrun): wall-clock of every cupt_group call of bench.py's N-GPU e2e step, per rank.
   python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29531 tools/group_breakdown.py"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch, torch.distributed as dist
import bench
from cudapathtracer_b200 import api

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
sc, s, cam = bench.build_scene()
bvh = s.build_bvh()
uid = torch.zeros(128, dtype=torch.uint8, device="cuda")
if rank == 0:
    uid.copy_(torch.frombuffer(bytearray(api.group_unique_id()), dtype=torch.uint8))
dist.broadcast(uid, src=0)
g = api.Group(device=local, rank=rank, nranks=world, unique_id=bytes(uid.cpu().numpy().tobytes()))
g.set_options(max_depth=bench.DEPTH, seed=bench.SEED)
img = torch.empty((bench.W * bench.H, 3), dtype=torch.float32, pin_memory=True).numpy()
root = rank == 0
for it in range(4):
    dist.barrier(); torch.cuda.synchronize()
    t = [time.perf_counter()]
    g.set_scene(s if root else None); t.append(time.perf_counter())
    g.set_camera(cam); g.set_bvh(bvh if root else None); t.append(time.perf_counter())
    g.reset_accum(); g.render(0, bench.SPP * world); t.append(time.perf_counter())
    if root:
        g.read_accum(False, out=img)
    else:
        g.sync()
    t.append(time.perf_counter())
    dist.barrier(); torch.cuda.synchronize(); t.append(time.perf_counter())
    print(f"rank {rank} it {it}: " + " ".join(f"{n}={1e3 * (b - a):.2f}" for n, a, b in zip(("set_scene", "set_bvh", "render_call", "read_or_sync", "barrier"), t, t[1:])) + f" total={1e3 * (t[-1] - t[0]):.2f}", flush=True)
g.close()
dist.destroy_process_group()
