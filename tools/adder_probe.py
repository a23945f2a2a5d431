"""Times NStepTransitionAdder.add at N envs (default 16 384 x 741 observations, 59 actions, n_step 50): the one-kernel path
(flyb_adder_step) against the same arithmetic as torch indexing operations."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from flybody_b200 import actors as A
from flybody_b200.fly_envs import TimeStep

N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
O, Ad, n = 741, 59, 50
dev = torch.device("cuda:0")
g = torch.Generator(device=dev); g.manual_seed(0)
def ts(kind):
    st = torch.full((N,), kind, dtype=torch.int32, device=dev)
    if kind == 1:
        u = torch.rand(N, device=dev, generator=g)
        st[u < 0.02] = 2
        st[u > 0.98] = 0
    return TimeStep(st, torch.rand(N, device=dev, generator=g), torch.ones(N, device=dev), {"walker/obs": torch.randn((N, O), device=dev, generator=g)})
def run(use_kernel):
    adder = A.NStepTransitionAdder(N, O, Ad, n, 0.95, device=dev, sink=lambda b: None)
    if not use_kernel:
        A.NStepTransitionAdder._force_torch = True
    adder.add_first(ts(0))
    steps = [ts(1) for _ in range(8)]
    acts = [torch.randn((N, Ad), device=dev, generator=g) for _ in range(8)]
    for k in range(60):
        adder.add(acts[k % 8], steps[k % 8])
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(41)]
    t0 = time.perf_counter()
    ev[0].record()
    tot = 0.0
    for k in range(40):
        if FLUSH:
            flush.fill_(float(k))
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        adder.add(acts[k % 8], steps[k % 8])
        b.record(); ev[k + 1].record()
        torch.cuda.synchronize()
        tot += a.elapsed_time(b)
    print("   events around add only: %.3f ms" % (tot / 40))
    torch.cuda.synchronize()
    wall = (time.perf_counter() - t0) / 40 * 1e3
    ms = [ev[k].elapsed_time(ev[k + 1]) for k in range(40)]
    A.NStepTransitionAdder._force_torch = False
    return float(np.median(ms)), wall
FLUSH = len(sys.argv) > 2
flush = torch.empty(256 * 1024 * 1024 // 4, device=dev)
print("kernel  : %.3f ms per add (events), %.3f ms wall" % run(True))
print("torch   : %.3f ms per add (events), %.3f ms wall" % run(False))
