"""The fused policy kernel alone at N rows (default 16 384 x 741 -> 59): CUDA-event time of flyb_policy_forward split into the
observation cast and the fused kernel (two launches), TFLOP/s against the measured bf16 peak, and the fp32 torch module beside it."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from flybody_b200 import actors as A

N = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
obs_dim, act_dim = 741, 59
dev = torch.device("cuda:0")
torch.manual_seed(0)
pol = A.LayerNormMLPPolicy(obs_dim, act_dim).to(dev)
fused = A.FusedPolicy(pol, max_rows=N)
obs = torch.randn((N, obs_dim), device=dev)
flush = torch.empty(256 * 1024 * 1024 // 4, device=dev)
ev = lambda: torch.cuda.Event(enable_timing=True)
def timed(fn, n=20, warm=5):
    for _ in range(warm):
        fn()
    ts = []
    for k in range(n):
        flush.fill_(float(k))
        a, b = ev(), ev()
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    return float(np.median(ts))
with torch.no_grad():
    t_fused = timed(lambda: fused(obs))
    t_torch = timed(lambda: pol(obs))
    torch.backends.cuda.matmul.allow_tf32 = True
    t_tf32 = timed(lambda: pol(obs))
flop = 2.0 * N * (obs_dim * 512 + 3 * 512 * 512 + 512 * 2 * act_dim)
try:   # a -DFLY_POLICY_PROF build (tools/build_variant.py) carries clock64 stamps of two CTAs
    from flybody_b200 import engine as _E
    _E.lib().flyb_policy_prof_dump()
except AttributeError:
    pass
peaks = json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json")))
print(json.dumps({"rows": N, "fused_ms": t_fused, "fused_tflops": flop / t_fused / 1e9, "frac_of_measured_bf16_peak": flop / t_fused / 1e9 / peaks["bf16_tflops"],
                  "torch_fp32_ms": t_torch, "torch_tf32_ms": t_tf32, "gflop": flop / 1e9,
                  "note": "fused_ms = observation cast (fp32 -> bf16, padded) + fly_policy_kernel, L2 flushed between calls"}))
