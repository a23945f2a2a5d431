"""vision_flight at N envs: device time of the step kernel and of the eye renderer, separately (CUDA events)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from flybody_b200 import terrain as T, wbpg as W
from flybody_b200.compile_fly import load_model
from flybody_b200.engine import Engine
from flybody_b200.model_pack import asset_path

N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
kind = sys.argv[2] if len(sys.argv) > 2 else "trench"
dev = torch.device("cuda:0")
M = load_model(asset_path("vision_flight"))
eng = Engine(M, N, 0, seed=1)
eng.load_wbpg(W.build_tables(W.synthetic_base_pattern()))
noterrain = bool(os.environ.get("NOTERRAIN"))     # experiment: what the height-field narrow phase costs (no pool: its pairs return at once)
if not noterrain:
    eng.load_terrain(T.pack_pool(T.terrain_table(kind, 16, seed=0)))
A, O = eng.action_dim, eng.obs_dim
obs = torch.zeros((N, O), device=dev); rew = torch.zeros(N, device=dev); disc = torch.zeros(N, device=dev)
st = torch.zeros(N, dtype=torch.int32, device=dev)
eyes = torch.zeros((N, 2, 32, 32, 3), dtype=torch.uint8, device=dev)
amin = torch.as_tensor(np.asarray(M["action_min"], np.float32), device=dev); amax = torch.as_tensor(np.asarray(M["action_max"], np.float32), device=dev)
gen = torch.Generator(device=dev); gen.manual_seed(0)
pool = [0.3 * (amin + torch.rand((N, A), device=dev, generator=gen) * (amax - amin)) for _ in range(8)]
s = torch.cuda.current_stream(dev).cuda_stream
eng.reset_ptr(0, obs.data_ptr(), st.data_ptr(), s)
ev = lambda: torch.cuda.Event(enable_timing=True)
ts, tr = [], []
for k in range(30):
    a, b, c = ev(), ev(), ev()
    a.record(); eng.step_ptr(pool[k % 8].data_ptr(), obs.data_ptr(), rew.data_ptr(), disc.data_ptr(), st.data_ptr(), s)
    b.record()
    if not noterrain:
        eng.render_ptr(eyes.data_ptr(), stream=s)
    c.record()
    torch.cuda.synchronize()
    if k >= 10:
        ts.append(a.elapsed_time(b)); tr.append(b.elapsed_time(c))
print(f"N {N} {kind}: step {np.mean(ts):.2f} ms, render {np.mean(tr):.2f} ms, stats {eng.episode_stats()}")
print("image mean", float(eyes.float().mean()), "nonzero share", float((eyes > 0).float().mean()))
