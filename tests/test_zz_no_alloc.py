"""flyb_step_noalloc (include/flyb.h): the step with its scratch in the library's static device array must give what
flyb_step gives.  Collected last on purpose.  Run with -m gpu."""
import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _pair(M, N, walk_clip, seed):
    from flybody_b200.engine import Engine
    engs = [Engine(M, N, 0, seed=seed) for _ in range(2)]
    for e in engs:
        e.load_clips(*walk_clip)
        e.set_clip_ids(np.zeros(N, np.int32))
    return engs


def test_noalloc_step_equals_step_twice_in_a_row(models, walk_clip):
    """Two engines in the same state; one is stepped through flyb_step, the other through flyb_step_noalloc, two calls in a
    row with per-env random actions.  Results are independent of the launch configuration and of which warp slot steps an
    env (tests/test_gpu_parity.py holds the engine to that bit for bit), so the tolerance here is the same: equality."""
    import torch
    M = models["walk_imitation"]
    N = 40
    ref, noa = _pair(M, N, walk_clip, seed=3)
    dev = torch.device("cuda:0")
    O, A = ref.obs_dim, ref.action_dim

    def bufs():
        return (torch.zeros(N, O, device=dev), torch.zeros(N, device=dev), torch.zeros(N, device=dev),
                torch.zeros(N, dtype=torch.int32, device=dev))
    out_r, out_n = bufs(), bufs()
    for e, o in ((ref, out_r), (noa, out_n)):   # reset launches the step kernel once: its module is loaded from here on
        e.reset_ptr(None, o[0].data_ptr(), o[3].data_ptr())
    torch.cuda.synchronize()
    assert torch.equal(out_r[0], out_n[0])
    rng = np.random.RandomState(11)
    lo, hi = np.asarray(M["action_min"], np.float32), np.asarray(M["action_max"], np.float32)
    acts = [torch.from_numpy(rng.uniform(lo, hi, (N, A)).astype(np.float32)).to(dev) for _ in range(2)]
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for k, a in enumerate(acts):
        noa.step_noalloc_ptr(a.data_ptr(), out_n[0].data_ptr(), out_n[1].data_ptr(), out_n[2].data_ptr(), out_n[3].data_ptr())
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0, "flyb_step_noalloc must not allocate device memory"
    for a in acts:
        ref.step_ptr(a.data_ptr(), out_r[0].data_ptr(), out_r[1].data_ptr(), out_r[2].data_ptr(), out_r[3].data_ptr())
    torch.cuda.synchronize()
    for r, n, name in zip(out_r, out_n, ("obs", "reward", "discount", "step_type")):
        assert torch.equal(r, n), name
    sr, sn = ref.get_state(), noa.get_state()
    for k in ("qpos", "qvel", "act", "warm", "ints"):
        assert np.array_equal(sr[k], sn[k]), k
    assert float(out_n[0].abs().max()) > 0 and bool((out_n[3] == 1).any())   # it really stepped
    # and step by step: a third call on each, compared right after it
    a = torch.from_numpy(rng.uniform(lo, hi, (N, A)).astype(np.float32)).to(dev)
    noa.step_noalloc_ptr(a.data_ptr(), out_n[0].data_ptr(), out_n[1].data_ptr(), out_n[2].data_ptr(), out_n[3].data_ptr())
    ref.step_ptr(a.data_ptr(), out_r[0].data_ptr(), out_r[1].data_ptr(), out_r[2].data_ptr(), out_r[3].data_ptr())
    torch.cuda.synchronize()
    for r, n, name in zip(out_r, out_n, ("obs", "reward", "discount", "step_type")):
        assert torch.equal(r, n), name
    ref.close(); noa.close()


def test_noalloc_rejects_batches_beyond_its_maximum(models):
    import torch
    from flybody_b200.engine import Engine
    M = models["walk_on_ball"]
    N = 1025   # FLYB_NOALLOC_MAX_ENVS + 1
    eng = Engine(M, N, 0)
    dev = torch.device("cuda:0")
    obs, r, g = torch.zeros(N, eng.obs_dim, device=dev), torch.zeros(N, device=dev), torch.zeros(N, device=dev)
    st, a = torch.zeros(N, dtype=torch.int32, device=dev), torch.zeros(N, eng.action_dim, device=dev)
    with pytest.raises(RuntimeError, match="FLYB_NOALLOC_MAX_ENVS"):
        eng.step_noalloc_ptr(a.data_ptr(), obs.data_ptr(), r.data_ptr(), g.data_ptr(), st.data_ptr())
    eng.close()
