"""BASELINE.json configs[3]: the actor side of the DMPO loop with GPU-batched environments (flybody_b200/actors.py).
walk_imitation, N envs, the reference's policy shape (512 x 4 LayerNorm-MLP -> diagonal Gaussian,
config/train_config_fly.yaml:6), n_step 50 / discount 0.95 transitions assembled on the device.  Prints env-steps/s of
(a) env.step with random actions, (b) + policy inference, (c) + n-step adder (the whole actor loop).  Not the bench."""
import json, os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from flybody_b200 import actors as A, fly_envs

N = int(os.environ.get("N", 16384)); steps = int(os.environ.get("STEPS", 20)); warm = int(os.environ.get("WARM", 5))
env = fly_envs.walk_imitation("synthetic:64:600:0", num_envs=N, seed=1)
dev = env.flat_observation.device
spec = env.action_spec()
O, Adim = env.flat_observation.shape[1], spec.shape[0]
torch.manual_seed(0)
policy = A.LayerNormMLPPolicy(O, Adim, layer_sizes=(512, 512, 512, 512)).to(dev)
count = {"n": 0}
adder = A.NStepTransitionAdder(N, O, Adim, n_step=50, discount=0.95, device=dev, sink=lambda b: count.__setitem__("n", count["n"] + int(b["reward"].shape[0])))
gen = torch.Generator(device=dev); gen.manual_seed(0)
lo = torch.as_tensor(spec.minimum, device=dev); hi = torch.as_tensor(spec.maximum, device=dev)


def timed(fn, k):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize(); e0.record()
    for _ in range(k):
        fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / k


ts = [env.reset()]
def env_only():
    ts[0] = env.step(lo + torch.rand((N, Adim), device=dev, generator=gen) * (hi - lo))
def env_policy():
    ts[0] = env.step(policy.sample(ts[0].observation, gen).clamp(lo, hi))
actor = A.BatchedFeedForwardActor(policy, adder, generator=gen)
loop = A.BatchedEnvironmentLoop(env, actor)
out = {"num_envs": N, "policy": "LayerNormMLP 512x4 -> diag Gaussian, fp32", "n_step": 50, "discount": 0.95}
for name, fn in (("env_random_actions", env_only), ("env_plus_policy", env_policy)):
    timed(fn, warm)
    ms = timed(fn, steps)
    out[name] = {"ms_per_step": ms, "env_steps_per_s": N / ms * 1e3}
loop.run(warm + 50)   # fill the n-step windows so every step emits N transitions
t0 = count["n"]
ms = timed(lambda: loop.run(1), steps)
out["actor_loop"] = {"ms_per_step": ms, "env_steps_per_s": N / ms * 1e3, "transitions_per_step": (count["n"] - t0) / steps}
print(json.dumps(out), flush=True)
