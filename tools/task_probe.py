"""Step-time of every task at the default launch configuration (informational; bench.py measures the north-star workload)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from flybody_b200 import fly_envs

N = int(os.environ.get("N", 16384))
for task, make in (("walk_imitation", lambda: fly_envs.walk_imitation("synthetic:32:600:0", num_envs=N, seed=1)),
                   ("walk_on_ball", lambda: fly_envs.walk_on_ball(num_envs=N, seed=1)),
                   ("template", lambda: fly_envs.template_task(num_envs=N, seed=1)),
                   ("flight_imitation", lambda: fly_envs.flight_imitation("synthetic", "synthetic:32:3006:0", num_envs=N, seed=1))):
    env = make()
    spec = env.action_spec()
    lo = torch.as_tensor(spec.minimum, device=env.device); hi = torch.as_tensor(spec.maximum, device=env.device)
    gen = torch.Generator(device=env.device); gen.manual_seed(0)
    env.reset()
    ts, last, rew = [], 0, 0.0
    for k in range(12):
        a = lo + torch.rand((N, spec.shape[0]), device=env.device, generator=gen) * (hi - lo)
        if task == "flight_imitation":
            a = 0.3 * a
        torch.cuda.synchronize(); t0 = time.time()
        out = env.step(a)
        torch.cuda.synchronize(); ts.append(time.time() - t0)
        last += int((out.step_type == 2).sum()); rew += float(out.reward.mean())
    t = float(np.median(ts[4:]))
    print(f"{task:18s} N {N}: {t*1e3:8.2f} ms/step -> {N/t:10.0f} env-steps/s ({N*int(env.M['n_substeps'])/t/1e6:6.2f} M physics steps/s); "
          f"episodes ended {last}, mean reward {rew/12:.3f}", flush=True)
    del env
