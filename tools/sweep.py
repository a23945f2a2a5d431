"""BASELINE.json configs[4]: env-count sweep on one rank (N per GPU from the list in NS) for walk_imitation and
flight_imitation, random actions, auto-reset.  One JSON line per (task, N).  Under torchrun every rank runs the same N
(weak scaling) and rank 0 prints the aggregate, timed as max over ranks."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from flybody_b200 import fly_envs

rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
local = int(os.environ.get("LOCAL_RANK", 0))
if world > 1:
    import torch.distributed as dist
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
NS = [int(x) for x in os.environ.get("NS", "1024,4096,16384,65536,262144").split(",")]
for task in os.environ.get("TASKS", "walk_imitation,flight_imitation").split(","):
    for N in NS:
        if task == "walk_imitation":
            env = fly_envs.walk_imitation("synthetic:64:600:0", num_envs=N, device=local, seed=1, env_offset=rank * N)
        else:
            env = fly_envs.flight_imitation("synthetic", "synthetic:32:3006:0", num_envs=N, device=local, seed=1, env_offset=rank * N)
        dev = env.flat_observation.device
        spec = env.action_spec()
        lo = torch.as_tensor(spec.minimum, device=dev); hi = torch.as_tensor(spec.maximum, device=dev)
        gen = torch.Generator(device=dev); gen.manual_seed(rank)
        scale = 0.3 if task == "flight_imitation" else 1.0
        steps = max(3, min(20, int(2_000_000 // N)))
        acts = [scale * (lo + torch.rand((N, spec.shape[0]), device=dev, generator=gen) * (hi - lo)) for _ in range(4)]
        env.reset()
        for k in range(3):
            env.step(acts[k % 4])
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(); e0.record()
        for k in range(steps):
            env.step(acts[k % 4])
        e1.record(); torch.cuda.synchronize()
        ms = torch.tensor([e0.elapsed_time(e1) / steps], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(json.dumps({"task": task, "envs_per_gpu": N, "n_gpus": world, "ms_per_step": float(ms), "steps": steps,
                              "env_steps_per_s": N * world / float(ms) * 1e3}), flush=True)
        env.close(); del env
        torch.cuda.empty_cache()
if world > 1:
    dist.destroy_process_group()
