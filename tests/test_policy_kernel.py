"""The fused policy kernel (csrc/flyb_policy.cu: tcgen05.mma / TMEM / TMA, through the C ABI `flyb_policy_*`) against the fp32
torch module it takes its parameters from (flybody_b200/actors.py: LayerNormMLPPolicy = the network of
vnl_ray/agents/network_factory.py:73-101).  Run with -m gpu.

Tolerances: the kernel multiplies bf16 operands (8 mantissa bits) and accumulates in fp32; activations are rounded to bf16 between
layers.  Against a torch evaluation that rounds the same tensors to bf16 at the same places the outputs agree to 5e-3 of their
scale (measured on B200: 0.9e-3 .. 1.9e-3 - accumulation order and tanh / exp implementations flip the bf16 rounding of single
activations); against the plain fp32 module to 3e-2 (measured 3.6e-3 .. 5.3e-3: bf16 rounding of weights and activations
through five layers)."""
import math

import numpy as np
import pytest
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.timeout(300)]


def _bf16_reference(pol, obs):
    """The module's forward with operands rounded to bf16 where the kernel rounds them (fp32 accumulation)."""
    r = lambda t: t.to(torch.bfloat16).to(torch.float32)   # noqa: E731
    hp = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    try:
        x = r(obs) @ r(pol.first.weight).T + pol.first.bias
        x = torch.tanh(torch.nn.functional.layer_norm(x, (512,), pol.norm.weight, pol.norm.bias, pol.norm.eps))
        for lin in pol.rest:
            x = torch.nn.functional.elu(r(x) @ r(lin.weight).T + lin.bias)
        mean = r(x) @ r(pol.mean.weight).T + pol.mean.bias
        scale = torch.nn.functional.softplus(r(x) @ r(pol.scale.weight).T + pol.scale.bias) * (pol.init_scale / math.log(2.0)) + pol.min_scale
    finally:
        torch.backends.cuda.matmul.allow_tf32 = hp
    return mean, scale


@pytest.mark.parametrize("obs_dim,act_dim,rows", [(741, 59, 1000), (289, 59, 128), (104, 12, 37), (64, 64, 300)])
def test_fused_policy_matches_the_torch_module(obs_dim, act_dim, rows):
    from flybody_b200 import actors as A
    torch.manual_seed(obs_dim + rows)
    dev = torch.device("cuda:0")
    pol = A.LayerNormMLPPolicy(obs_dim, act_dim).to(dev)
    with torch.no_grad():   # heads as large as a trained policy's, non-trivial LayerNorm affine and biases
        for lin in (pol.mean, pol.scale):
            lin.weight.normal_(std=1.0 / math.sqrt(512)); lin.bias.normal_(std=0.3)
        pol.norm.weight.uniform_(0.5, 1.5); pol.norm.bias.normal_(std=0.2)
        for lin in [pol.first, *pol.rest]:
            lin.bias.normal_(std=0.2)
    fused = A.FusedPolicy(pol, max_rows=1024)
    obs = torch.randn((rows, obs_dim), device=dev) * torch.linspace(0.1, 3.0, obs_dim, device=dev)
    with torch.no_grad():
        mean, scale = fused(obs)
        m32, s32 = pol(obs)
        mb, sb = _bf16_reference(pol, obs)
    torch.cuda.synchronize()
    assert mean.shape == (rows, act_dim) and torch.isfinite(mean).all() and torch.isfinite(scale).all()
    ms, ss = float(m32.abs().max()), float(s32.abs().max())
    e_bf = max(float((mean - mb).abs().max()) / ms, float((scale - sb).abs().max()) / ss)
    e_32 = max(float((mean - m32).abs().max()) / ms, float((scale - s32).abs().max()) / ss)
    print(f"obs {obs_dim} act {act_dim} rows {rows}: vs bf16-rounded torch {e_bf:.2e}, vs fp32 torch {e_32:.2e}")
    assert e_bf < 5e-3, e_bf
    assert e_32 < 3e-2, e_32
    # the sampled action is mean + scale * eps with the generator's noise; a second call reproduces it
    g = torch.Generator(device=dev); g.manual_seed(7)
    a1 = fused.sample(obs, g)
    g.manual_seed(7)
    eps = torch.randn((rows, act_dim), device=dev, generator=g)
    assert float((a1 - (mean + scale * eps)).abs().max()) < 1e-5 * max(1.0, float(a1.abs().max()))
    # parameters changed by a learner reach the kernel through refresh()
    with torch.no_grad():
        pol.mean.bias += 1.0
    fused.refresh()
    mean2, _ = fused(obs)
    assert float((mean2 - mean - 1.0).abs().max()) < 1e-5
    assert fused.launches >= 8
    with pytest.raises(RuntimeError, match="more rows"):
        fused(torch.zeros((2048, obs_dim), device=dev))
    fused.close()


def test_fused_policy_head_options():
    """tanh_mean and fixed_scale of the head (network_factory.py:90-98 passes them through to MultivariateNormalDiagHead)."""
    from flybody_b200 import actors as A
    torch.manual_seed(3)
    dev = torch.device("cuda:0")
    obs = torch.randn((200, 104), device=dev)
    for kw in (dict(tanh_mean=True), dict(fixed_scale=True, init_scale=0.3), dict(min_scale=0.05, init_scale=1.2)):
        pol = A.LayerNormMLPPolicy(104, 12, **kw).to(dev)
        with torch.no_grad():
            for lin in (pol.mean, pol.scale):
                lin.weight.normal_(std=2.0 / math.sqrt(512)); lin.bias.normal_(std=0.5)
        fused = A.FusedPolicy(pol, max_rows=256)
        with torch.no_grad():
            mean, scale = fused(obs)
            m32, s32 = pol(obs)
        assert float((mean - m32).abs().max()) < 3e-2 * max(1.0, float(m32.abs().max())), kw
        assert float((scale - s32).abs().max()) < 3e-2 * max(1.0, float(s32.abs().max())), kw
        if kw.get("tanh_mean"):
            assert float(mean.abs().max()) <= 1.0
        if kw.get("fixed_scale"):
            assert float((scale - 0.3).abs().max()) == 0.0
        g = torch.Generator(device=dev); g.manual_seed(1)
        a = fused.sample(obs, g)
        g.manual_seed(1)
        eps = torch.randn((200, 12), device=dev, generator=g)
        assert float((a - (mean + scale * eps)).abs().max()) < 1e-5 * max(1.0, float(a.abs().max())), kw
        fused.close()


def test_actor_loop_with_the_fused_policy():
    """config 4 (BASELINE.json configs[3]): the batched environment loop with the kernel as the actor's policy."""
    from flybody_b200 import actors as A, fly_envs
    dev = torch.device("cuda:0")
    env = fly_envs.walk_on_ball(num_envs=256, seed=0)
    torch.manual_seed(0)
    obs_dim = sum(int(np.prod(v.shape)) for v in env.observation_spec().values())
    act_dim = env.action_spec().shape[0]
    pol = A.LayerNormMLPPolicy(obs_dim, act_dim).to(dev)
    fused = A.FusedPolicy(pol, max_rows=256)
    got = []
    adder = A.NStepTransitionAdder(256, obs_dim, act_dim, n_step=5, discount=0.99, device=dev, sink=lambda b: got.append(int(b["valid"].sum())))
    g = torch.Generator(device=dev); g.manual_seed(1)
    loop = A.BatchedEnvironmentLoop(env, A.BatchedFeedForwardActor(fused, adder, generator=g))
    steps, reward = loop.run(6)
    assert steps == 6 * 256 and np.isfinite(reward) and sum(got) >= 5 * 256
    env.close(); fused.close()


def test_fused_policy_is_bit_reproducible_at_the_bench_size():
    """16 384 rows = 64 CTA pairs, 300 calls on the same observations with an L2 flush in between now and then: every call must give
    the first call's bits.  The kernel has no atomics in its arithmetic, so any difference would be a synchronisation hole in the
    TMA / MMA / epilogue hand-offs (stale activations, a stage reused too early, an accumulator read before its commit)."""
    from flybody_b200 import actors as A
    dev = torch.device("cuda:0")
    torch.manual_seed(3)
    pol = A.LayerNormMLPPolicy(741, 59).to(dev)
    with torch.no_grad():   # heads away from zero so that all five layers matter
        pol.mean.weight.normal_(std=0.05); pol.scale.weight.normal_(std=0.05)
    fused = A.FusedPolicy(pol, max_rows=16384)
    obs = torch.randn((16384, 741), device=dev)
    flush = torch.empty(192 * 1024 * 1024 // 4, device=dev)
    mean0, scale0 = [t.clone() for t in fused(obs)]
    assert torch.isfinite(mean0).all() and torch.isfinite(scale0).all()
    for k in range(300):
        if k % 7 == 0:
            flush.fill_(float(k))
        mean, scale = fused(obs)
        assert torch.equal(mean, mean0) and torch.equal(scale, scale0), f"call {k} differs"
    fused.close()
