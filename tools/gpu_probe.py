"""First-contact GPU script: smoke, then step-time sweep over launch configurations (not the bench)."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import __graft_entry__ as G
from flybody_b200 import clips as C
from flybody_b200.compile_fly import load_model
from flybody_b200.engine import Engine, FLAG_TRACE
from flybody_b200.model_pack import asset_path

G.smoke()
N = int(os.environ.get("N", 16384))
M = load_model(asset_path("walk_imitation"))
tab = C.synthetic_table(M, 32, 600, 0)
eng = Engine(M, N, 0, seed=1, flags=FLAG_TRACE)
eng.load_clips(*tab)
dev = torch.device("cuda:0")
obs = torch.zeros((N, eng.obs_dim), device=dev); r = torch.zeros(N, device=dev); g = torch.zeros(N, device=dev)
st = torch.zeros(N, dtype=torch.int32, device=dev)
amin = torch.as_tensor(np.asarray(M["action_min"], np.float32), device=dev); amax = torch.as_tensor(np.asarray(M["action_max"], np.float32), device=dev)
gen = torch.Generator(device=dev); gen.manual_seed(0)
for (lvl, warps) in [(1, 16), (1, 14), (1, 12), (2, 8)]:
    try:
        eng.configure(lvl, warps, 0)
        eng.reset_ptr(0, obs.data_ptr(), st.data_ptr(), 0)
        torch.cuda.synchronize()
        ts = []
        for k in range(6):
            a = amin + torch.rand((N, eng.action_dim), device=dev, generator=gen) * (amax - amin)
            torch.cuda.synchronize(); t0 = time.time()
            eng.step_ptr(a.data_ptr(), obs.data_ptr(), r.data_ptr(), g.data_ptr(), st.data_ptr(), 0)
            torch.cuda.synchronize(); ts.append(time.time() - t0)
        tr = eng.get_trace()
        print(f"level {lvl} warps {warps}: grid {eng.info(11)} smem/block {eng.info(9)} ms/step {[round(t*1e3,2) for t in ts]} "
              f"-> {N/min(ts[2:]):.0f} env-steps/s ; ncon mean {tr['ncon'].mean():.1f} max {tr['ncon'].max()} nefc mean {tr['nefc'].mean():.1f} "
              f"max {tr['nefc'].max()} niter mean {tr['niter'].mean():.2f} overflow {np.bitwise_or.reduce(tr['overflow'])} "
              f"last {int((st==2).sum())} nan {int(torch.isnan(obs).any())}", flush=True)
    except Exception as e:
        print("config", lvl, warps, "failed:", e, flush=True)
