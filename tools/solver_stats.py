"""Distribution of the constraint solver's work over the envs of a batch (trace of the last physics step of a control step):
contacts, constraint rows, Newton iterations.  usage: python tools/solver_stats.py [task] [N] [steps]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from flybody_b200 import clips as C
from flybody_b200.compile_fly import load_model
from flybody_b200.engine import Engine, FLAG_TRACE
from flybody_b200.model_pack import asset_path

task = sys.argv[1] if len(sys.argv) > 1 else "walk_on_ball"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 4096
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 30
M = load_model(asset_path(task))
eng = Engine(M, N, 0, seed=1, flags=FLAG_TRACE)
if task == "walk_imitation":
    eng.load_clips(*C.pack_clips(M, [C.synth_walk_clip(M, T=600, seed=s) for s in range(8)]))
eng.reset_host()
rng = np.random.RandomState(0)
lo, hi = np.asarray(M["action_min"], np.float32), np.asarray(M["action_max"], np.float32)
for k in range(steps):
    eng.step_host((lo + rng.rand(N, len(lo)).astype(np.float32) * (hi - lo)))
    if k >= steps - 3:
        t = eng.get_trace()
        q = lambda x: "mean %.1f  p50 %d  p90 %d  p99 %d  max %d" % (x.mean(), *np.percentile(x, [50, 90, 99]).astype(int), x.max())
        print(f"{task} step {k}: ncon {q(t['ncon'])} | nefc {q(t['nefc'])} | niter {q(t['niter'])}")
        blocks = t["niter"][: N // 28 * 28].reshape(-1, 28)
        print("   per block of 28 envs: mean of max niter %.2f vs mean niter %.2f; nefc: mean of max %.1f vs mean %.1f" % (
            blocks.max(1).mean(), t["niter"].mean(), t["nefc"][: N // 28 * 28].reshape(-1, 28).max(1).mean(), t["nefc"].mean()))
