"""The batched actor (flybody_b200/actors.py): the n-step adder against a per-env list implementation of acme's
NStepTransitionAdder sequence (growing horizons at the start of an episode, the sliding full horizon, shrinking horizons at its
end), the policy head, action delay and the environment loop - on CPU tensors, and (-m gpu) the same checks on CUDA tensors
plus the whole actor loop over the real batched environment."""
import collections

import numpy as np
import pytest
import torch

from flybody_b200 import actors as A

TimeStep = collections.namedtuple("TimeStep", ["step_type", "reward", "discount", "observation"])


class FakeEnv:
    """Auto-resetting batched env with seeded episode lengths, rewards and discounts (no physics)."""

    def __init__(self, N, O, seed):
        self.N, self.O = N, O
        self.rng = np.random.RandomState(seed)
        self.left = np.zeros(N, np.int64)
        self.ended = np.ones(N, bool)

    def _obs(self):
        return {"walker/b": torch.as_tensor(self.rng.normal(size=(self.N, self.O - 2)), dtype=torch.float32),
                "walker/a": torch.as_tensor(self.rng.normal(size=(self.N, 2)), dtype=torch.float32)}

    def reset(self):
        self.left = self.rng.randint(1, 12, self.N)
        self.ended[:] = False
        return TimeStep(torch.zeros(self.N, dtype=torch.int32), None, None, self._obs())

    def step(self, action):
        st = np.ones(self.N, np.int32)
        for i in range(self.N):
            if self.ended[i]:
                st[i] = 0; self.left[i] = self.rng.randint(1, 12); self.ended[i] = False
            else:
                self.left[i] -= 1
                if self.left[i] == 0:
                    st[i] = 2; self.ended[i] = True
        r = self.rng.uniform(0, 2, self.N).astype(np.float32)
        d = np.where((st == 2) & (self.rng.uniform(size=self.N) < 0.5), 0.0, 1.0).astype(np.float32)
        return TimeStep(torch.as_tensor(st), torch.as_tensor(r), torch.as_tensor(d), self._obs())


def _reference_transitions(steps, n, gamma, finished=True):
    """One episode as [(o, a, r, d, o_next), ...] -> the transitions acme's NStepTransitionAdder writes, in its order
    (adders/reverb/transition.py): after every step t the transition from max(0, t - n) to t; at the end of the episode the
    shrinking ones from max(0, T - n) + 1 .. T - 1 to T."""
    T = len(steps)
    spans = [(max(0, t - n), t) for t in range(1, T + 1)]
    if finished:
        spans += [(s, T) for s in range(max(0, T - n) + 1, T)]
    out = []
    for s, e in spans:
        R, run = 0.0, 1.0
        for k in range(s, e):
            R += run * steps[k][2]
            run *= gamma * steps[k][3]
        out.append((s, steps[s][0], steps[s][1], R, run / gamma, steps[e - 1][4]))
    return out


def _check_adder(n_step, device):
    N, O, Adim, gamma = 9, 6, 4, 0.95
    env = FakeEnv(N, O, seed=n_step)
    got = collections.defaultdict(list)
    dev = torch.device(device)

    def to(ts):
        return TimeStep(ts.step_type.to(dev), None if ts.reward is None else ts.reward.to(dev),
                        None if ts.discount is None else ts.discount.to(dev), {k: v.to(dev) for k, v in ts.observation.items()})

    def sink(b):
        b = {k: v.cpu() for k, v in b.items()}
        for i in range(len(b["env"])):
            if bool(b["valid"][i]):
                got[int(b["env"][i])].append(tuple(np.asarray(b[k][i]) for k in ("start", "observation", "action", "reward", "discount", "next_observation")))

    adder = A.NStepTransitionAdder(N, O, Adim, n_step, gamma, device=device, sink=sink)
    ts = env.reset()
    adder.add_first(to(ts))
    episodes = [[] for _ in range(N)]       # finished + running episodes per env
    cur = [[] for _ in range(N)]
    prev = A.batch_concat(ts.observation).numpy().copy()
    rng = np.random.RandomState(0)
    for t in range(60):
        a = torch.as_tensor(rng.normal(size=(N, Adim)), dtype=torch.float32)
        ts = env.step(a)
        adder.add(a.to(dev), to(ts))
        o = A.batch_concat(ts.observation).numpy()
        for i in range(N):
            if int(ts.step_type[i]) == 0:
                cur[i] = []
            else:
                cur[i].append((prev[i].copy(), a[i].numpy().copy(), float(ts.reward[i]), float(ts.discount[i]), o[i].copy()))
                if int(ts.step_type[i]) == 2:
                    episodes[i].append(cur[i]); cur[i] = []
        prev = o.copy()
    total = 0
    for i in range(N):
        want = []
        for ep in episodes[i]:
            want += _reference_transitions(ep, n_step, gamma)
        want += _reference_transitions(cur[i], n_step, gamma, finished=False)
        have = got[i]
        assert len(have) == len(want), (i, len(have), len(want))
        for h, w in zip(have, want):
            assert int(h[0]) == w[0]
            np.testing.assert_allclose(h[1], w[1], atol=0); np.testing.assert_allclose(h[2], w[2], atol=0)
            np.testing.assert_allclose(h[3], w[3], rtol=1e-5, atol=1e-6); np.testing.assert_allclose(h[4], w[4], rtol=1e-5, atol=1e-7)
            np.testing.assert_allclose(h[5], w[5], atol=0)
        total += len(want)
    assert adder.num_written == total and total > 200


@pytest.mark.parametrize("n_step", [1, 3, 5])
def test_nstep_adder_matches_per_env_reference(n_step):
    _check_adder(n_step, "cpu")


@pytest.mark.gpu
@pytest.mark.parametrize("n_step", [1, 5, 50])
def test_nstep_adder_matches_per_env_reference_on_cuda(n_step):
    _check_adder(n_step, "cuda")


@pytest.mark.gpu
def test_actor_loop_over_the_batched_fly_env():
    """BASELINE.json configs[3]: policy (741 -> 512 x 4 -> 2 x 59), env and n-step adder on one GPU, 256 walk_imitation envs."""
    from flybody_b200 import fly_envs
    env = fly_envs.walk_imitation("synthetic:4:200:0", num_envs=256, seed=2)
    O, Adim = env.flat_observation.shape[1], env.action_spec().shape[0]
    torch.manual_seed(0)
    pi = A.LayerNormMLPPolicy(O, Adim).to(env.device)
    got = {"n": 0, "bad": 0}

    def sink(b):
        v = b["valid"]
        got["n"] += int(v.sum())
        got["bad"] += int((~torch.isfinite(b["reward"][v])).sum()) + int((~torch.isfinite(b["next_observation"][v])).sum())
    adder = A.NStepTransitionAdder(256, O, Adim, 5, 0.95, device=env.device, sink=sink)
    loop = A.BatchedEnvironmentLoop(env, A.BatchedFeedForwardActor(pi, adder, generator=torch.Generator(device=env.device).manual_seed(1)))
    n, rsum = loop.run(40)
    assert n == 40 * 256 and np.isfinite(rsum) and rsum > 0
    assert got["n"] == adder.num_written >= 40 * 256 - 256 and got["bad"] == 0
    env.close()


def test_policy_head_and_batch_concat_order():
    torch.manual_seed(0)
    obs = {"walker/b": torch.randn(5, 3), "walker/a": torch.randn(5, 2, 2)}
    flat = A.batch_concat(obs)
    assert flat.shape == (5, 7) and torch.equal(flat[:, :4], obs["walker/a"].reshape(5, 4))
    pi = A.LayerNormMLPPolicy(7, 4, layer_sizes=(32, 16))
    mean, scale = pi(obs)
    assert mean.shape == scale.shape == (5, 4)
    # near-zero head at init: scale ~ init_scale (0.7), mean ~ 0 (acme MultivariateNormalDiagHead with init_scale)
    assert float((scale - 0.7).abs().max()) < 0.05 and float(mean.abs().max()) < 0.05
    g = torch.Generator().manual_seed(1)
    a1 = pi.sample(obs, g)
    g = torch.Generator().manual_seed(1)
    assert torch.equal(a1, pi.sample(obs, g))


def test_action_delay_and_environment_loop():
    torch.manual_seed(0)
    N, O, Adim = 4, 6, 3
    env = FakeEnv(N, O, seed=3)
    pi = A.LayerNormMLPPolicy(O, Adim, layer_sizes=(16,))
    seen = []
    adder = A.NStepTransitionAdder(N, O, Adim, 2, 0.9, device="cpu", sink=lambda b: seen.append(int(b["valid"].sum())))
    actor = A.BatchedFeedForwardActor(pi, adder, action_delay=2, generator=torch.Generator().manual_seed(0))
    obs = env.reset().observation
    assert float(actor.select_action(obs).abs().max()) == 0 and float(actor.select_action(obs).abs().max()) == 0
    assert float(actor.select_action(obs).abs().max()) > 0          # third call returns the first queued action
    loop = A.BatchedEnvironmentLoop(FakeEnv(N, O, seed=4), A.BatchedFeedForwardActor(pi, adder))
    n, rsum = loop.run(25)
    assert n == 25 * N and rsum > 0 and sum(seen) == adder.num_written > 0
