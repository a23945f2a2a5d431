"""Times one control step of a task at the library's default configuration (FLYB_LIB_PATH picks the build).
   env: N (16384), TASK (walk_imitation), STEPS (8), WARM (4).  Prints one line; not the bench."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
from flybody_b200 import clips as C
from flybody_b200.compile_fly import load_model
from flybody_b200.engine import Engine, LIB_PATH
from flybody_b200.model_pack import asset_path

N = int(os.environ.get("N", 16384)); task = os.environ.get("TASK", "walk_imitation")
steps = int(os.environ.get("STEPS", 8)); warm = int(os.environ.get("WARM", 4))
M = load_model(asset_path(task))
eng = Engine(M, N, 0, seed=1)
if task == "walk_imitation":
    eng.load_clips(*C.synthetic_table(M, 32, 600, 0))
dev = torch.device("cuda:0")
obs = torch.zeros((N, eng.obs_dim), device=dev); r = torch.zeros(N, device=dev); g = torch.zeros(N, device=dev)
st = torch.zeros(N, dtype=torch.int32, device=dev)
amin = torch.as_tensor(np.asarray(M["action_min"], np.float32), device=dev); amax = torch.as_tensor(np.asarray(M["action_max"], np.float32), device=dev)
gen = torch.Generator(device=dev); gen.manual_seed(0)
eng.reset_ptr(0, obs.data_ptr(), st.data_ptr(), 0)
acts = [amin + torch.rand((N, eng.action_dim), device=dev, generator=gen) * (amax - amin) for _ in range(steps + warm)]
for k in range(warm):
    eng.step_ptr(acts[k].data_ptr(), obs.data_ptr(), r.data_ptr(), g.data_ptr(), st.data_ptr(), 0)
ev = [torch.cuda.Event(enable_timing=True) for _ in range(steps + 1)]
torch.cuda.synchronize()
ev[0].record()
for k in range(steps):
    eng.step_ptr(acts[warm + k].data_ptr(), obs.data_ptr(), r.data_ptr(), g.data_ptr(), st.data_ptr(), 0)
    ev[k + 1].record()
torch.cuda.synchronize()
ms = [ev[k].elapsed_time(ev[k + 1]) for k in range(steps)]
print(f"{os.path.basename(LIB_PATH)} {task} N={N} grid {eng.info(11)} ms/step med {np.median(ms):.2f} min {min(ms):.2f} "
      f"-> {N / np.median(ms) * 1e3:.0f} env-steps/s  obs_sum {float(obs.double().sum()):.6f} nan {int(torch.isnan(obs).any())}", flush=True)
