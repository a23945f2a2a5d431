"""Small fixed workload for ncu: reset + a few control steps of walk_imitation (N, LEVEL, WARPS, STEPS from env)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from flybody_b200 import clips as C
from flybody_b200.compile_fly import load_model
from flybody_b200.engine import Engine
from flybody_b200.model_pack import asset_path

N = int(os.environ.get("N", 4144)); lvl = os.environ.get("LEVEL"); warps = os.environ.get("WARPS")   # default: the library's configuration
steps = int(os.environ.get("STEPS", 4)); task = os.environ.get("TASK", "walk_imitation")
M = load_model(asset_path(task))
eng = Engine(M, N, 0, seed=1)
if task == "walk_imitation":
    eng.load_clips(*C.synthetic_table(M, 32, 600, 0))
if lvl is not None and warps is not None:
    eng.configure(int(lvl), int(warps), int(os.environ.get("BPS", 0)))
dev = torch.device("cuda:0")
obs = torch.zeros((N, eng.obs_dim), device=dev); r = torch.zeros(N, device=dev); g = torch.zeros(N, device=dev)
st = torch.zeros(N, dtype=torch.int32, device=dev)
amin = torch.as_tensor(np.asarray(M["action_min"], np.float32), device=dev); amax = torch.as_tensor(np.asarray(M["action_max"], np.float32), device=dev)
gen = torch.Generator(device=dev); gen.manual_seed(0)
eng.reset_ptr(0, obs.data_ptr(), st.data_ptr(), 0)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
acts = [amin + torch.rand((N, eng.action_dim), device=dev, generator=gen) * (amax - amin) for _ in range(steps)]
torch.cuda.synchronize()
ev[0].record()
for k in range(steps):
    eng.step_ptr(acts[k].data_ptr(), obs.data_ptr(), r.data_ptr(), g.data_ptr(), st.data_ptr(), 0)
    ev[k + 1].record()
torch.cuda.synchronize()
print("ms per step", [round(ev[k].elapsed_time(ev[k + 1]), 3) for k in range(steps)], "grid", eng.info(11))
