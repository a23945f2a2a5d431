"""Times one control step of a task at the library's default configuration (FLYB_LIB_PATH picks the build).
   env: N (16384), TASK (walk_imitation), STEPS (8), WARM (8).  Prints one line; not the bench."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from flybody_b200 import fly_envs
from flybody_b200.engine import LIB_PATH

N = int(os.environ.get("N", 16384)); task = os.environ.get("TASK", "walk_imitation")
steps = int(os.environ.get("STEPS", 8)); warm = int(os.environ.get("WARM", 8))
make = {"walk_imitation": lambda: fly_envs.walk_imitation("synthetic:32:600:0", num_envs=N, seed=1),
        "walk_on_ball": lambda: fly_envs.walk_on_ball(num_envs=N, seed=1),
        "template": lambda: fly_envs.template_task(num_envs=N, seed=1),
        "flight_imitation": lambda: fly_envs.flight_imitation("synthetic", "synthetic:32:3006:0", num_envs=N, seed=1)}[task]
env = make()
spec = env.action_spec()
lo = torch.as_tensor(spec.minimum, device=env.device); hi = torch.as_tensor(spec.maximum, device=env.device)
gen = torch.Generator(device=env.device); gen.manual_seed(0)
scale = 0.3 if task == "flight_imitation" else 1.0
acts = [scale * (lo + torch.rand((N, spec.shape[0]), device=env.device, generator=gen) * (hi - lo)) for _ in range(steps + warm)]
env.reset()
for k in range(warm):
    out = env.step(acts[k])
ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
torch.cuda.synchronize()
ev[0].record()
for k in range(steps):
    out = env.step(acts[warm + k])
    ev[k + 1].record()
torch.cuda.synchronize()
ms = [ev[k].elapsed_time(ev[k + 1]) for k in range(steps)]
obs = torch.cat([v.reshape(N, -1) for v in out.observation.values()], dim=1)
print(f"{os.path.basename(LIB_PATH)} {task} N={N} ms/step med {np.median(ms):.2f} min {min(ms):.2f} "
      f"-> {N / np.median(ms) * 1e3:.0f} env-steps/s  obs_sum {float(obs.double().nan_to_num().sum()):.6f} nan {int(torch.isnan(obs).any())}", flush=True)
