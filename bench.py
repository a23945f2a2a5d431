"""bench.py - fly env-steps/s of walk_imitation (BASELINE.json configs[1]: synthetic reference clips, 16384 envs per B200).

  python bench.py --gpus N --steps K --warmup W            own arm (CUDA engine through the C ABI)
  python bench.py --impl reference --gpus N ...            the reference's CPU algorithm on the host cores

A "step" is one control step (10 physics substeps) of every env of the shard.  `value` is timed with CUDA events
on the launching stream with actions already resident in HBM; `e2e` goes through the host-pointer C-ABI call
(`flyb_step_host`: pinned H2D of the actions, launch, observations / reward / discount / step_type written to the pinned host
buffers - by the kernel itself, row by row as envs finish - inside the timed region).  Envs shard across ranks with no data-path collective (weak scaling: 16384 envs per GPU); NCCL is
used for the barrier, the max-over-ranks of the timing and the episode-statistics all-reduce.

The reference's own implementation of the path (MuJoCo C via dm_control) is not installable here (no network, not
vendored): `--impl reference` and `cpu_baseline` time the float64 CPU restatement in oracle/ (kind "port")."""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402

ENVS_PER_GPU = 16384
# compulsory fp32 bytes per env-step; vision_flight (8f rank 4, not in the survey's table): state R+W 2 x (43 + 42 +
# 42 + 8) x 4 = 1 080, action 48, observations 64 x 4 = 256, two 32 x 32 x 3 uint8 eye images 6 144, reward / discount / type 12
ALGO_BYTES = {"walk_imitation": 9984, "walk_on_ball": 4436, "flight_imitation": 1924, "vision_flight": 7540}
METRIC = "fly env-steps/s (walk_imitation, 16384 envs per GPU, synthetic clips, random actions)"




def workload(clips: int) -> str:
    """config.workload of both arms (the reference arm times a bounded sample of exactly this workload)."""
    return ("walk_imitation (BASELINE.json configs[1]): 16384 envs per GPU, 10 physics substeps per control step, "
            f"{clips} synthetic reference clips x 600 steps, U(ctrlrange) random actions, auto-reset")


_REAL_STDOUT = None


def _emit(obj):
    """The one JSON line, on the process's original stdout."""
    sys.stdout.flush()
    line = (json.dumps(obj) + "\n").encode()
    os.write(_REAL_STDOUT if _REAL_STDOUT is not None else 1, line)


def _peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        return float(json.load(open(p))["hbm_gbs"]), "measured"
    return 6650.0, "fallback"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md recipe)."""

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
             "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-lms", "100", "-i", str(index)],
                                         stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.rows.append([x.strip() for x in line.split(",")])

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm = [float(r[0]) for r in self.rows if r and r[0].replace(".", "").isdigit()]
        mx = [float(r[1]) for r in self.rows if len(r) > 1 and r[1].replace(".", "").isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({names[i] for r in self.rows if len(r) >= 6 for i in range(4) if r[2 + i].lower().startswith("active")})
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None, "reasons": reasons,
                "samples": len(sm)}


def _model_and_clips(n_clips, T):
    from flybody_b200 import clips as C
    from flybody_b200.compile_fly import load_model
    from flybody_b200.model_pack import asset_path
    M = load_model(asset_path("walk_imitation"))
    return M, C.synthetic_table(M, n_clips, T, 0, cache=False)   # every rank builds its own copy (about a minute for 1024 clips)


# ------------------------------------------------------------------------------------------------ CPU arms
def _cpu_worker(args):
    """`steps` control steps of one oracle env (auto-reset on LAST); returns elapsed seconds."""
    wid, steps, rows_path = args
    from flybody_b200.compile_fly import load_model
    from flybody_b200.model_pack import asset_path
    from oracle.oracle import OracleEnv
    M = load_model(asset_path("walk_imitation"))
    d = np.load(rows_path)
    rows, off, ln = d["rows"], d["offsets"], d["lengths"]
    c = wid % len(off)
    r = rows[off[c]:off[c] + ln[c]].astype(np.float64)
    nj, ns = int(M["n_mocap_joints"]), int(M["n_mocap_sites"])
    o = OracleEnv(M)
    o.set_clip(r[:, :7 + nj], r[:, 7 + nj:13 + 2 * nj], r[:, 13 + 2 * nj:13 + 2 * nj + 3 * ns], r[:, 13 + 2 * nj + 3 * ns:])
    o.env_reset()
    rng = np.random.RandomState(wid)
    acts = rng.uniform(M["action_min"], M["action_max"], (steps, int(M["action_dim"])))
    t0 = time.perf_counter()
    for k in range(steps):
        o.env_step(acts[k])
    return time.perf_counter() - t0


def cpu_throughput(cores: int, steps_per_core: int, table):
    """env-steps/s of the float64 port on `cores` processes (one env each, same workload)."""
    import multiprocessing as mp
    import tempfile
    from oracle import oracle as O
    O.build()
    path = os.path.join(tempfile.gettempdir(), f"flyb_bench_clips_{os.getpid()}.npz")
    np.savez(path, rows=table[0][:table[2][0] * min(len(table[1]), 8)], offsets=table[1][:8], lengths=table[2][:8])
    if cores == 1:
        el = [_cpu_worker((0, steps_per_core, path))]
    else:
        with mp.get_context("spawn").Pool(cores) as pool:
            el = pool.map(_cpu_worker, [(i, steps_per_core, path) for i in range(cores)])
    os.remove(path)
    dt = max(el)   # stepping only: process start-up and model load are outside the timed region
    return cores * steps_per_core / dt, dt


def run_reference(args):
    """The reference arm: the CPU algorithm on all host cores, each step a bounded sample of the workload."""
    rank = int(os.environ.get("RANK", 0))
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    M, table = _model_and_clips(8, 600)
    per_core = 20   # control steps per core per "step" (about 0.3 s of CPU work per core)
    for _ in range(args.warmup):
        cpu_throughput(cores, 5, table)
    vals, ms = [], []
    for _ in range(args.steps):
        v, dt = cpu_throughput(cores, per_core, table)
        vals.append(v); ms.append(dt * 1e3)
    v = float(np.mean(vals))
    sample = (f"{cores} oracle envs (one per host core, clips 0..7 of the same table) x {per_core} control steps per step, same model "
              "and action distribution; float64 CPU port of the reference algorithm (MuJoCo / dm_control not installable here)")
    _emit({
        "impl": "reference", "metric": METRIC, "value": v, "unit": "env-steps/s", "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": float(np.mean(ms)), "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic",
        "config": {"workload": workload(args.clips), "sample": sample},
        "cpu_baseline": {"value": v, "unit": "env-steps/s", "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": v, "unit": "env-steps/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    })


def _time_steps(torch, eng, dev, acts, steps, warm, flush, out, eyes=None):
    """`steps` control steps timed with CUDA events on the launching stream after `warm` untimed ones, L2 flushed in between;
    returns ms per step (list)."""
    obs, rew, disc, st = out
    stream = torch.cuda.current_stream(dev).cuda_stream
    for k in range(warm):
        eng.step_ptr(acts(k).data_ptr(), obs.data_ptr(), rew.data_ptr(), disc.data_ptr(), st.data_ptr(), stream)
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    torch.cuda.synchronize(dev)
    for k in range(steps):
        flush.fill_(float(k))
        a = acts(warm + k)
        ev[k][0].record()
        eng.step_ptr(a.data_ptr(), obs.data_ptr(), rew.data_ptr(), disc.data_ptr(), st.data_ptr(), stream)
        if eyes is not None:
            eng.render_ptr(eyes.data_ptr(), stream=stream)
        ev[k][1].record()
    torch.cuda.synchronize(dev)
    return [a.elapsed_time(b) for a, b in ev]


def _other_task(torch, task, N, dev, local, flush, peak, steps=10, warm=5):
    """One more task at the same batch size, device-timed like the headline (random actions; flight at 0.3 x ctrlrange so that
    episodes last, as tools/task_probe.py does)."""
    from flybody_b200 import clips as C, wbpg as W
    from flybody_b200.compile_fly import load_model
    from flybody_b200.engine import Engine
    from flybody_b200.model_pack import asset_path
    M = load_model(asset_path(task))
    eng = Engine(M, N, local, seed=1)
    scale = 1.0
    eyes = None
    if task == "flight_imitation":
        eng.load_clips(*C.synthetic_flight_table(M, 64, 3006, 0))
        eng.load_wbpg(W.build_tables(W.synthetic_base_pattern()))
        scale = 0.3
    if task == "vision_flight":   # 16 trench terrains; the eye images are rendered inside the timed region (flyb_render)
        from flybody_b200 import terrain as T
        eng.load_wbpg(W.build_tables(W.synthetic_base_pattern()))
        eng.load_terrain(T.pack_pool(T.terrain_table("trench", 16, seed=0)))
        eyes = torch.zeros((N, 2, 32, 32, 3), dtype=torch.uint8, device=dev)
        scale = 0.3
    A, O = eng.action_dim, eng.obs_dim
    out = (torch.zeros((N, O), device=dev), torch.zeros(N, device=dev), torch.zeros(N, device=dev),
           torch.zeros(N, dtype=torch.int32, device=dev))
    amin = torch.as_tensor(np.asarray(M["action_min"], np.float32), device=dev)
    amax = torch.as_tensor(np.asarray(M["action_max"], np.float32), device=dev)
    gen = torch.Generator(device=dev); gen.manual_seed(99)
    pool = [scale * (amin + torch.rand((N, A), device=dev, generator=gen) * (amax - amin)) for _ in range(8)]
    eng.reset_ptr(0, out[0].data_ptr(), out[3].data_ptr(), torch.cuda.current_stream(dev).cuda_stream)
    ms = _time_steps(torch, eng, dev, lambda k: pool[k % 8], steps, warm, flush, out, eyes)
    stats = eng.episode_stats()
    eng.close()
    t = float(np.mean(ms)) * 1e-3
    ach = ALGO_BYTES[task] * N / t / 1e9
    return {"value": N / t, "unit": "env-steps/s", "ms_per_step": t * 1e3, "envs": N, "n_substeps": int(M["n_substeps"]),
            "algorithmic_bytes_per_env_step": ALGO_BYTES[task], "achieved_gbs": ach, "frac": ach / peak,
            "actions": "U(ctrlrange)" if scale == 1.0 else f"{scale} x U(ctrlrange)", "overflow_steps": stats["overflow_steps"],
            "reset_share": stats["resets"] / max(1.0, stats["resets"] + stats["steps"])}


def _actor_loop(torch, eng, N, dev, flush, out, amin, amax, env_only_ms, steps=10, warm=5):
    """config 4 (BASELINE.json configs[3]): the actor side of the DMPO loop on the GPU the environments live on - fused policy
    kernel (flyb_policy_forward: sample an action from the observations) -> clamp -> step kernel -> n-step adder (n_step 50,
    discount 0.99: train_dmpo_ray.py:236), device-timed per iteration like the headline; the policy alone is timed beside it."""
    import collections
    from flybody_b200 import actors as A
    obs, rew, disc, st = out
    TS = collections.namedtuple("TimeStep", ["step_type", "reward", "discount", "observation"])
    stream = torch.cuda.current_stream(dev).cuda_stream
    torch.manual_seed(0)
    pol = A.LayerNormMLPPolicy(obs.shape[1], amin.shape[0]).to(dev)
    fused = A.FusedPolicy(pol, max_rows=N)
    adder = A.NStepTransitionAdder(N, obs.shape[1], amin.shape[0], n_step=50, discount=0.99, device=dev, sink=None)
    gen = torch.Generator(device=dev); gen.manual_seed(5)
    adder.add_first(TS(st, None, None, obs), torch.ones(N, dtype=torch.bool, device=dev))
    l0 = fused.launches

    inner = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    adder_end = [torch.cuda.Event(enable_timing=True) for _ in range(steps)]

    def one(k=-1):
        a = torch.minimum(torch.maximum(fused.sample(obs, gen), amin), amax)
        if k >= 0:
            inner[k][0].record()
        eng.step_ptr(a.data_ptr(), obs.data_ptr(), rew.data_ptr(), disc.data_ptr(), st.data_ptr(), stream)
        if k >= 0:
            inner[k][1].record()
        adder.add(a, TS(st, rew, disc, obs))
        if k >= 0:
            adder_end[k].record()
    for _ in range(warm):
        one()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    torch.cuda.synchronize(dev)
    for k in range(steps):
        flush.fill_(float(k))
        ev[k][0].record(); one(k); ev[k][1].record()
    torch.cuda.synchronize(dev)
    ms = float(np.mean([a.elapsed_time(b) for a, b in ev]))
    env_ms = float(np.mean([a.elapsed_time(b) for a, b in inner]))   # the step kernel inside the loop, under the policy's actions
    adder_ms = float(np.mean([inner[k][1].elapsed_time(adder_end[k]) for k in range(steps)]))
    pe = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(steps)]
    for k in range(steps):
        flush.fill_(float(k))
        pe[k][0].record(); fused.sample(obs, gen); pe[k][1].record()
    torch.cuda.synchronize(dev)
    pms = float(np.mean([a.elapsed_time(b) for a, b in pe]))
    flop = 2.0 * N * (obs.shape[1] * 512 + 3 * 512 * 512 + 512 * 2 * amin.shape[0])
    res = {"value": N / (ms * 1e-3), "unit": "env-steps/s", "ms_per_step": ms, "envs": N,
           "loop": "fused policy kernel (sample) -> clamp -> step kernel -> n-step adder (n_step 50)",
           "step_kernel_ms_in_loop": env_ms, "ratio_to_env_only": env_ms / ms,
           "note": "ratio_to_env_only = the step kernel's own time inside this loop / the loop's time per step (the policy's N(0, 0.7) "
                   "actions keep the flies on their feet: more contacts per step than under U(ctrlrange), headline_ms = %.2f)" % env_only_ms,
           "adder_ms": adder_ms, "policy_ms": pms, "policy_tflops": flop / (pms * 1e-3) / 1e12,
           "policy_kernel": "fly_policy_kernel (tcgen05.mma kind::f16 bf16 -> fp32 in TMEM, TMA weight tiles; incl. the fp32 -> bf16 cast of "
                            "the observations and torch.randn for the noise)",
           "policy_launches_per_step": (fused.launches - l0) / float(warm + 2 * steps), "transitions_written": adder.num_written}
    fused.close()
    return res


def run_gpu(args):
    import torch
    import torch.distributed as dist
    from flybody_b200.engine import (Engine, INFO_LAUNCHES, INFO_GRID, INFO_SMEM_BYTES, INFO_SPEC_KERNEL, INFO_WARPS_PER_BLOCK,
                                     fma_peak_tflops)
    from flybody_b200.sharding import shard_range, reduce_stats
    rank, world = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1))
    local = int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    N = args.envs
    lo, hi = shard_range(N * world, rank, world)
    M, table = _model_and_clips(args.clips, 600)
    eng = Engine(M, N, local, seed=args.seed, env_offset=lo)
    eng.load_clips(*table)
    A, O = eng.action_dim, eng.obs_dim
    obs = torch.zeros((N, O), device=dev); rew = torch.zeros(N, device=dev); disc = torch.zeros(N, device=dev)
    st = torch.zeros(N, dtype=torch.int32, device=dev)
    amin = torch.as_tensor(np.asarray(M["action_min"], np.float32), device=dev)
    amax = torch.as_tensor(np.asarray(M["action_max"], np.float32), device=dev)
    gen = torch.Generator(device=dev); gen.manual_seed(1234 + rank)
    pool = [amin + torch.rand((N, A), device=dev, generator=gen) * (amax - amin) for _ in range(8)]
    flush = torch.empty(256 * 1024 * 1024 // 4, device=dev)   # > 126 MB L2
    stream = torch.cuda.current_stream(dev).cuda_stream

    def barrier():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)

    def gmax(x):
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    eng.reset_ptr(0, obs.data_ptr(), st.data_ptr(), stream)
    for k in range(args.warmup):
        eng.step_ptr(pool[k % 8].data_ptr(), obs.data_ptr(), rew.data_ptr(), disc.data_ptr(), st.data_ptr(), stream)
    # ---- device-resident timing: CUDA events on the launching stream, L2 flushed between timed steps
    eng.episode_stats()   # counters from here on: the timed steps and the end-to-end leg
    l0 = eng.info(INFO_LAUNCHES)
    clocks = ClockSampler(local) if rank == 0 else None
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(args.steps)]
    barrier()
    for k in range(args.steps):
        flush.fill_(float(k))
        ev[k][0].record()
        eng.step_ptr(pool[(k + args.warmup) % 8].data_ptr(), obs.data_ptr(), rew.data_ptr(), disc.data_ptr(), st.data_ptr(), stream)
        ev[k][1].record()
    barrier()
    kernel_ms = [a.elapsed_time(b) for a, b in ev]
    total_ms = gmax(sum(kernel_ms))
    launches = eng.info(INFO_LAUNCHES) - l0
    value = N * world * args.steps / (total_ms * 1e-3)
    finite = bool(torch.isfinite(obs).all().item())
    # ---- end to end through the host-pointer C-ABI call
    # inputs and outputs in page-locked host memory (the contract's "pinned host memory"): flyb_step_host then copies the actions
    # from the caller's buffer and the kernel writes the outputs into the caller's buffers directly; with pageable numpy arrays it stages through its own pinned buffers (one more
    # host copy each way, ~6 ms per step at this size)
    h_pin = [torch.empty((N, A), dtype=torch.float32).pin_memory() for _ in pool]
    for hp, p in zip(h_pin, pool):
        hp.copy_(p)
    h_act = [hp.numpy() for hp in h_pin]
    o_pin = (torch.zeros((N, O)).pin_memory(), torch.zeros(N).pin_memory(), torch.zeros(N).pin_memory(),
             torch.zeros(N, dtype=torch.int32).pin_memory())
    out = tuple(t.numpy() for t in o_pin)
    for k in range(2):
        eng.step_host(h_act[k], out)
    barrier()
    t0 = time.perf_counter()
    for k in range(args.steps):
        eng.step_host(h_act[k % 8], out)
    torch.cuda.synchronize(dev)
    e2e_s = gmax(time.perf_counter() - t0)
    clk = clocks.stop() if clocks else None
    e2e = N * world * args.steps / e2e_s
    stats = reduce_stats(eng.episode_stats(), dev)
    extras = {}
    if world == 1 and not args.no_extras:
        # ---- the other workloads of SURVEY.md 8(d), same process, device-timed like the headline (N = 1 run only)
        peak0, _ = _peaks()
        # config 2 (ii): reference-tracking actions - every env is sent the joint angles of its own clip's next row on the
        # position servos of the mocap joints, other actuators mid-range, + 2 % of the control range of Gaussian noise
        rows_t = torch.as_tensor(table[0], device=dev)
        off_t, len_t = torch.as_tensor(table[1], device=dev).long(), torch.as_tensor(table[2], device=dev).long()
        mocap = {int(j): i for i, j in enumerate(np.asarray(M["mocap_joint"]))}
        col = np.full(A, -1)
        for a in range(int(M["nu"])):
            ai, j = int(M["ctrl_action"][a]), int(M["act_trnid"][a])
            if ai >= 0 and int(M["act_trntype"][a]) == 0 and j in mocap:
                col[ai] = 7 + mocap[j]
        dst = torch.as_tensor(np.nonzero(col >= 0)[0], device=dev)
        src = torch.as_tensor(col[col >= 0], device=dev)
        mid = 0.5 * (amin + amax)
        mask = torch.ones(N, dtype=torch.uint8, device=dev)
        eng.reset_ptr(mask.data_ptr(), obs.data_ptr(), st.data_ptr(), stream)

        def tracking(k):
            ints = torch.as_tensor(eng.get_state()["ints"], device=dev).long()
            row = off_t[ints[:, 3]] + torch.minimum(ints[:, 0] + 1, len_t[ints[:, 3]] - 1)
            a = mid.repeat(N, 1)
            a[:, dst] = rows_t[row][:, src]
            a = a + 0.02 * (amax - amin) * torch.randn((N, A), device=dev, generator=gen)
            tracking.keep = torch.minimum(torch.maximum(a, amin), amax).contiguous()
            return tracking.keep
        eng.episode_stats()
        ms_t = _time_steps(torch, eng, dev, tracking, 10, 20, flush, (obs, rew, disc, st))
        ts = eng.episode_stats()
        tt = float(np.mean(ms_t)) * 1e-3
        extras["tracking_actions"] = {"value": N / tt, "unit": "env-steps/s", "ms_per_step": tt * 1e3, "envs": N,
                                      "actions": "clip joint angles of the next row on the position servos + 2 % noise (SURVEY 8d config 2 ii)",
                                      "mean_reward": float(rew.mean().item()), "episodes_ended": ts["episodes"],
                                      "overflow_steps": ts["overflow_steps"],
                                      "frac": ALGO_BYTES["walk_imitation"] * N / tt / 1e9 / peak0}
        del rows_t
        extras["actor_loop"] = _actor_loop(torch, eng, N, dev, flush, (obs, rew, disc, st), amin, amax, total_ms / args.steps)
        for task in ("flight_imitation", "walk_on_ball", "vision_flight"):
            extras[task] = _other_task(torch, task, N, dev, local, flush, peak0)
        extras["fma_peak_tflops"] = fma_peak_tflops(local)
    if rank == 0:
        peak, how = _peaks()
        per_launch_s = float(np.mean(kernel_ms)) * 1e-3
        achieved = ALGO_BYTES["walk_imitation"] * N / per_launch_s / 1e9
        traffic = None
        tp = os.path.join(ROOT, "profiles", "traffic.json")
        if os.path.exists(tp):
            traffic = json.load(open(tp)).get("dram_bytes_per_launch")
        cpu_steps = args.cpu_steps
        cores = os.cpu_count() or 1
        cpu_v, cpu_dt = cpu_throughput(cores, cpu_steps, table)
        fp32 = None
        fp = os.path.join(ROOT, "profiles", "r2_fp32.json")
        if os.path.exists(fp) and "fma_peak_tflops" in extras:
            flop = float(json.load(open(fp))["fp32_flop_per_env_step"])
            ach_tf = flop * N / per_launch_s / 1e12
            fp32 = {"flop_per_env_step": flop, "flop_source": "profiles/r2_fp32.json (ncu smsp__sass_thread_inst_executed_op_f{add,mul,fma} of the kept capture)",
                    "achieved_tflops": ach_tf, "peak_tflops": extras["fma_peak_tflops"],
                    "peak_source": "FMA micro-benchmark measured in this run (flyb_fma_peak)", "frac": ach_tf / extras["fma_peak_tflops"]}
        timed_steps = max(1.0, stats.get("steps", 0.0) + stats.get("resets", 0.0))
        _emit({
            "metric": METRIC, "value": value, "unit": "env-steps/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": total_ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic",
            "config": {"workload": workload(args.clips), "clip_table_gb": table[0].nbytes / 1e9,
                       "envs_per_gpu": N, "global_envs": N * world, "l2": "256 MB buffer written between timed steps",
                       "grid": eng.info(INFO_GRID), "warps_per_block": eng.info(INFO_WARPS_PER_BLOCK),
                       "smem_bytes_per_block": eng.info(INFO_SMEM_BYTES), "outputs_finite": finite,
                       "overflow_envs": stats.get("overflow_steps"),
                       "reset_share_of_env_steps": stats.get("resets", 0.0) / timed_steps,
                       "step_kernel": ("compiled for this task model (index %d of csrc/fly_dims_gen.h)" % eng.info(INFO_SPEC_KERNEL))
                       if eng.info(INFO_SPEC_KERNEL) >= 0 else "generic"},
            "clocks": clk, "gpu_launches": int(launches),
            "e2e": {"value": e2e, "unit": "env-steps/s", "h2d_bytes_per_step": N * A * 4, "d2h_bytes_per_step": N * (O + 3) * 4},
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak, "traffic": traffic,
                         "peak_source": how, "kernel": "fly_step_kernel", "algorithmic_bytes_per_env_step": ALGO_BYTES["walk_imitation"],
                         "note": "HBM is the contract roofline; the kernel is FP32-issue/latency bound (DESIGN.md)", "fp32": fp32},
            "cpu_baseline": {"value": cpu_v, "unit": "env-steps/s", "cores": cores, "kind": "port",
                             "sample": f"{cores} oracle envs (one per host core) x {cpu_steps} control steps, float64 CPU restatement "
                                       f"(oracle/), same model, clips and action distribution, {cpu_dt:.1f} s"},
            "episode_stats": stats,
            **{k: v for k, v in extras.items() if k != "fma_peak_tflops"},
        })
    eng.close()
    if world > 1:
        dist.destroy_process_group()


def main():
    # stdout carries ONE JSON line.  Libraries write there too (NCCL prints a version banner to fd 1 at communicator
    # creation on the GPU boxes), so fd 1 points at stderr while the benchmark runs and _emit() writes to the real one.
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="flyb", choices=["flyb", "reference"])
    ap.add_argument("--envs", type=int, default=ENVS_PER_GPU, help="envs per GPU")
    ap.add_argument("--clips", type=int, default=1024, help="synthetic reference clips x 600 steps (SURVEY 8d config 2: 1024 = 1.05 GB in HBM)")
    ap.add_argument("--no-extras", action="store_true", help="skip the tracking-action / flight / ball arms of the N = 1 run")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--cpu-steps", type=int, default=400, help="control steps per host core of the cpu_baseline leg")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "flyb" else args.warmup
    if args.impl == "reference":
        run_reference(args)
    else:
        run_gpu(args)


if __name__ == "__main__":
    main()
