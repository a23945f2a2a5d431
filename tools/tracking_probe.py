"""walk_imitation throughput with reference-tracking actions instead of U(ctrlrange) noise.
Every env plays the same synthetic clip; the action of control step k is the clip's joint angles of step k + 1 written to the
position servos that drive those joints (other actuators mid-range), plus per-env Gaussian noise of 2 % of the control range.
Episodes then last and the feet stay on the floor, which is the contact pattern training sees.  Not the bench."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from flybody_b200 import clips as C, fly_envs

N = int(os.environ.get("N", 16384)); steps = int(os.environ.get("STEPS", 60))
env = fly_envs.walk_imitation("synthetic:1:600:0", num_envs=N, seed=1)
M = env.M
rows, off, ln = C.synthetic_table(M, 1, 600, 0)
dev = env.flat_observation.device
spec = env.action_spec()
lo = torch.as_tensor(spec.minimum, device=dev); hi = torch.as_tensor(spec.maximum, device=dev)
A = spec.shape[0]
mocap = {int(j): i for i, j in enumerate(np.asarray(M["mocap_joint"]))}
col = np.full(A, -1)                       # action index -> column of the clip row holding its joint's reference angle
for a in range(int(M["nu"])):
    ai, j = int(M["ctrl_action"][a]), int(M["act_trnid"][a])
    if ai >= 0 and int(M["act_trntype"][a]) == 0 and j in mocap:
        col[ai] = 7 + mocap[j]
base = (0.5 * (lo + hi)).cpu().numpy()
gen = torch.Generator(device=dev); gen.manual_seed(0)
env.reset()
ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
ended, rew = 0, 0.0
torch.cuda.synchronize(); ev[0].record()
for k in range(steps):
    a = base.copy()
    a[col >= 0] = rows[min(k + 1, ln[0] - 1), col[col >= 0]]
    act = torch.as_tensor(a, device=dev)[None, :] + 0.02 * (hi - lo) * torch.randn((N, A), device=dev, generator=gen)
    out = env.step(torch.minimum(torch.maximum(act, lo), hi))
    ev[k + 1].record()
    ended += int((out.step_type == 2).sum()); rew += float(out.reward.mean())
torch.cuda.synchronize()
ms = np.array([ev[k].elapsed_time(ev[k + 1]) for k in range(steps)])   # includes this script's per-step host work
print(json.dumps({"task": "walk_imitation", "actions": "reference tracking + 2 % noise", "num_envs": N, "steps": steps,
                  "tracked_actions": int((col >= 0).sum()), "ms_per_step_median": float(np.median(ms[5:])),
                  "env_steps_per_s": float(N / np.median(ms[5:]) * 1e3), "episodes_ended": ended,
                  "mean_reward_per_step": rew / steps}), flush=True)
