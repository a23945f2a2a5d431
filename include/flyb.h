/* flyb.h - C ABI of the B200 batched fruit-fly environment engine (libflyb.so).
 *
 * What it replaces.  The reference has no FFI of its own for this path: `env.step(action)` /
 * `env.reset()` of the objects returned by vnl_ray/fly_envs.py:29-157 reach the MuJoCo C library through
 * dm_control (`composer.Environment.step` -> `mjcf.Physics.step` -> `mj_step2`/`mj_step1`) one environment
 * and one Python call at a time.  This library is the boundary a maintainer would bind instead (ctypes stub
 * in INTEGRATION.md): N environments are stepped by one call, state stays resident in HBM.
 *
 *   flyb_create      <- fly_envs.walk_on_ball / walk_imitation / flight_imitation constructors
 *                       (fly_envs.py:29,75,125) + composer.Environment.__init__ (fly_envs.py:67-72,117-122,152-157)
 *   flyb_load_clips  <- HDF5WalkingTrajectoryLoader / HDF5FlightTrajectoryLoader (tasks/trajectory_loaders.py:10-223)
 *   flyb_reset       <- composer.Environment.reset (task hooks walk_imitation.py:91-136)
 *   flyb_step        <- composer.Environment.step (walk_imitation.py:138-191, tasks/base.py:190-217,
 *                       fruitfly/fruitfly.py:480-492,532-669, tasks/rewards.py:9-111)
 *   flyb_step_host   <- the same call as the reference makes it: numpy (host) action in, host TimeStep out
 *   flyb_get_state / flyb_set_state <- physics.data.qpos/qvel/act/qacc_warmstart access (tasks/base.py:216,
 *                       walk_imitation.py:119-123)
 *
 * Conventions.  All functions return 0 on success, non-zero on error (message: flyb_last_error).  Nothing
 * throws across the ABI.  Physics divergence is not an error: it ends the episode (step_type LAST, discount
 * 0), as vnl_ray/tasks/base.py:214-217 and dm_control's PhysicsError handling do.  The caller owns every I/O
 * buffer; the library owns the model constants, clip table and per-env state.  One handle per GPU and
 * process; calls on one handle must be serialised by the caller; work is enqueued on the given CUDA stream
 * (cudaStream_t passed as void*, NULL = default stream) without hidden synchronisation.  The *_host calls run on a
 * stream of the library's own, first wait (on the device) for the handle's last device-pointer launch, and synchronise
 * before returning.  There is no CPU fallback: without a CUDA device
 * flyb_create fails. */
#ifndef FLYB_H_
#define FLYB_H_
#include <stdint.h>
#include "flyb_model.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct FlybEngine* flyb_handle;

enum { FLYB_STEP_FIRST = 0, FLYB_STEP_MID = 1, FLYB_STEP_LAST = 2 };
/* flags for flyb_create (tests switch stages off to isolate them) */
enum { FLYB_FLAG_NOCONTACT = 1, FLYB_FLAG_NOLIMIT = 2, FLYB_FLAG_NOFLUID = 4, FLYB_FLAG_NONOSLIP = 8,
       FLYB_FLAG_NOACTUATION = 16, FLYB_FLAG_NOWARMSTART = 32, FLYB_FLAG_TRACE = 1 << 16,
       FLYB_FLAG_LANES16 = 1 << 17 /* two envs per warp, 16 lanes each (csrc/fly_math.cuh: FLY_NL); default: one env per warp */ };
/* flyb_info selectors */
enum { FLYB_INFO_NUM_ENVS = 0, FLYB_INFO_OBS_DIM, FLYB_INFO_ACTION_DIM, FLYB_INFO_NQ, FLYB_INFO_NV, FLYB_INFO_NA,
       FLYB_INFO_TRACE_INTS, FLYB_INFO_TRACE_REALS, FLYB_INFO_LAUNCHES, FLYB_INFO_SMEM_BYTES, FLYB_INFO_BULK_BYTES,
       FLYB_INFO_GRID, FLYB_INFO_MAXCON, FLYB_INFO_MAXEFC, FLYB_INFO_STATE_INTS, FLYB_INFO_WARPS_PER_BLOCK,
       FLYB_INFO_SMEM_LEVEL,
       FLYB_INFO_SPEC_KERNEL /* after a launch: index of the model-specialised step kernel in use, -1 = generic */,
       FLYB_INFO_LANES /* lanes of a warp that step one env: 32 (one env per warp), 16 or 8 (two / four envs per warp) */ };

/* Compiles `model` for the device and allocates state for num_envs environments on CUDA device `device`.
 * env_offset is the global index of this shard's first env (multi-GPU: rank * num_envs); it only enters the
 * counter-based RNG that picks clips, so results do not depend on how envs are sharded. */
int flyb_create(const FlybModel* model, int num_envs, int device, uint64_t seed, int env_offset, int flags,
                flyb_handle* out);
void flyb_destroy(flyb_handle h);
const char* flyb_last_error(flyb_handle h);  /* h may be NULL: error of the last failed flyb_create */
int64_t flyb_info(flyb_handle h, int what);

/* Tuning knobs: smem_level 0..6 (how many workspace tiers live in shared memory, csrc/fly_ws.cuh; default 0: the whole
 * workspace in per-warp global scratch behind a full-size L1, 28 warps per SM), warps per block (only the combinations built into the library: KERNEL_UNITS in
 * flybody_b200/engine.py) and blocks per SM (0 = as many as fit).  Takes effect at the next launch. */
int flyb_configure(flyb_handle h, int smem_level, int warps_per_block, int blocks_per_sm);
/* The default configuration runs a step kernel compiled for the model's array sizes when the model is one of the shipped task
 * models (workspace offsets are immediates: 25 % faster); generic != 0 forces the any-model kernel.  The two agree to float32
 * rounding, not bit for bit (the compiler contracts multiply-adds differently in the two instantiations). */
int flyb_use_generic_kernel(flyb_handle h, int generic);

/* Reference clips (host pointers): rows of clip_width floats, [qpos 7+nj | qvel 6+nj | root2site 3ns | joint_quat 4nj]
 * (walk_imitation) or [qpos 7 | qvel 6] (flight_imitation); clip c occupies rows offsets[c] .. offsets[c]+lengths[c]-1. */
int flyb_load_clips(flyb_handle h, const float* rows, const int32_t* offsets, const int32_t* lengths, int n_clips);
/* Wing-beat pattern generator tables for flight_imitation (host pointers, float64): what the reference's
 * WingBeatPatternGenerator.__init__ builds (vnl_ray/tasks/pattern_generators.py:18-119; host builder: flybody_b200/wbpg.py):
 * beat_freqs [n_freqs], offsets [n_freqs+1] (first row of each frequency), traj [rows, 6], phase [rows]; rate is the
 * per-control-step low-pass factor exp(-dt_ctrl / ctrl_filter) (0 = no filter).  reset/step of the generator
 * (pattern_generators.py:121-191) then run inside the kernel, per env. */
int flyb_load_wbpg(flyb_handle h, const double* beat_freqs, const int32_t* offsets, const double* traj, const double* phase,
                   int n_freqs, double base_freq, double rel_freq_range, double rate, double dt_ctrl);
/* Initial wing-beat phase in [0,1) per env for the following resets (host pointer, num_envs doubles) - the value the
 * reference draws with random_state.uniform() (flight_imitation.py:137).  NULL = counter-based hash of (seed, env, episode). */
int flyb_set_wbpg_phase(flyb_handle h, const double* phase);
/* Per-env clip choice for each env's NEXT episode only (host pointer, num_envs entries; -1 = counter-based random), as
 * task.set_next_trajectory_index (walk_imitation.py:86-89).  NULL = all random.  Ids beyond the loaded table are an error. */
int flyb_set_clip_ids(flyb_handle h, const int32_t* ids);

/* Device-pointer API.  mask: NULL (all envs) or num_envs bytes, non-zero = reset that env.
 * Outputs of envs not reset are left untouched.  obs [N, obs_dim] float32, step_type [N] int32. */
int flyb_reset(flyb_handle h, const uint8_t* mask, float* obs, int32_t* step_type, void* stream);
/* One control step (n_substeps physics steps) of every env.  An env whose previous step returned LAST is
 * reset instead (step_type FIRST, reward 0, discount 1): dm_env auto-reset semantics, per env. */
int flyb_step(flyb_handle h, const float* action, float* obs, float* reward, float* discount, int32_t* step_type,
              void* stream);

/* flyb_step for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that
 * forbid allocation).  No counterpart in the reference.  Same arguments,
 * results and stream semantics as flyb_step (device pointers, caller-owned inputs and outputs), but the step kernel's
 * scratch memory - the per-warp workspace slots and the work counter, which flyb_step keeps in buffers that flyb_create
 * cudaMalloc'ed - lives in a file-scope `static __device__` array of FLYB_NOALLOC_SCRATCH_BYTES inside the library (every
 * GPU that loads the library holds it until exit; its device address is resolved with cudaGetSymbolAddress at flyb_create).
 * The call itself allocates nothing: no cudaMalloc, cudaFree or allocating library routine.
 * Limits: handles with more than FLYB_NOALLOC_MAX_ENVS environments, or whose one-block workspace does not fit the array,
 * are rejected with an error (flyb_last_error).  The array holds few workspace slots, so fewer blocks run than under
 * flyb_step: this entry point is for small batches.
 * Sharing: every call on a device uses the same array, which keeps the previous call's contents; the work counter is
 * cleared by each call (the workspace needs no clearing: every stage writes what it reads).  Calls that overlap on the
 * GPU - two streams, or two handles on one device - race on the array; serialise them. */
#define FLYB_NOALLOC_SCRATCH_BYTES (16u << 20)
#define FLYB_NOALLOC_MAX_ENVS 1024
int flyb_step_noalloc(flyb_handle h, const float* action, float* obs, float* reward, float* discount, int32_t* step_type,
                      void* stream);

/* Host-pointer API (what a numpy caller uses): copies through pinned staging buffers, synchronises. */
int flyb_reset_host(flyb_handle h, const uint8_t* mask, float* obs, int32_t* step_type);
/* Host-pointer variants: page-locked caller buffers are read and written directly, pageable ones are staged through the library's
 * own page-locked buffers.  flyb_step_host copies the actions in, launches, and the kernel itself writes every env's finished
 * observation row, reward, discount and step type into the (device-visible) host buffers while the other envs are still
 * stepping: nothing is left to copy when the launch ends.  FLYB_HOST_DIRECT=0 in the environment restores device -> host copies
 * after the launch. */
int flyb_step_host(flyb_handle h, const float* action, float* obs, float* reward, float* discount, int32_t* step_type);

/* State access (host pointers, any may be NULL).  qpos [N,nq] qvel [N,nv] act [N,na] qacc_warmstart [N,nv]
 * ints [N, FLYB_INFO_STATE_INTS]: step, episode_steps, reset_next, clip, episode, forced_clip, episode_len, - */
int flyb_get_state(flyb_handle h, float* qpos, float* qvel, float* act, float* warm, int32_t* ints);
int flyb_set_state(flyb_handle h, const float* qpos, const float* qvel, const float* act, const float* warm,
                   const int32_t* ints);
/* Trace of the last launch (needs FLYB_FLAG_TRACE; host pointers, may be NULL).
 * ints [N, TRACE_INTS]: ncon, nefc, nlimit, solver iterations, overflow bits, candidates (all of the control step's last full
 * physics step), ncon0, -, contact pair ids [MAXCON] of that step, contact pair ids [MAXCON] of the state the control step
 * started from (ncon0 of them; = the reference's contact list after the previous control step).
 * reals [N, TRACE_REALS]: contact dist [MAXCON] | qacc [nv] | sensordata | efc_force [MAXEFC]. */
int flyb_get_trace(flyb_handle h, int32_t* ints, float* reals);
/* Statistics summed over this shard's envs since the last call (host, 8 doubles): [0] sum of returns, [1] sum of lengths,
 * [2] fatal terminations, [3] count of the episodes that finished; [4] control steps (or reset passes) in which a capacity was
 * exceeded and contacts or constraint rows were dropped (FLYB_INFO_MAXCON / MAXEFC; MuJoCo would raise a warning counter:
 * must be 0 for the physics to be the reference's); [5] reset passes; [6] control steps; [7] unused. */
int flyb_episode_stats(flyb_handle h, double* out8);

/* vision_guided_flight (vnl_ray/fly_envs.py:160-206, tasks/vision_flight.py, tasks/arenas/hills.py): the terrain pool of a
 * handle whose model has a height-field geom.  heights [n_terrains, hfield_nrow, hfield_ncol] float32 (host); for trench
 * terrains trench_meta [n_terrains, 3] = points of the centre line (0: none), x of its first and of its last point, and
 * trench_y [n_terrains, trench_stride] its y values (the arena's `trench_specs`, hills.py:377-379); both NULL for bumps.
 * The reference regenerates its terrain every episode on the host; here every (env, episode) picks one of the pool by a
 * counter-based hash, the way reference clips are picked.  Collision with the terrain runs inside flyb_step (MuJoCo's
 * mjc_ConvexHField prism walk, at most 16 contacts per geom); an active contact with the terrain or the ground plane ends
 * the episode (vision_flight.py:214-233). */
int flyb_load_terrain(flyb_handle h, const float* heights, int n_terrains, const double* trench_meta, const float* trench_y,
                      int trench_stride);
/* The `right_eye` / `left_eye` observables (vnl_ray/fruitfly/fruitfly.py:676-708) of every env after the last flyb_reset /
 * flyb_step, ray-cast on the device: out [num_envs, 2 (right, left), height, width, 3] uint8 (DEVICE pointer), enqueued on
 * `stream` (same stream as the step, or ordered after it).  colors8 as flyb_render_eyes.  The reference renders through
 * MuJoCo's OpenGL renderer (unavailable on this box); what the ray caster draws is stated in csrc/fly_eyes.cuh. */
int flyb_render(flyb_handle h, int width, int height, float fovy_deg, const float* colors8, uint8_t* out, void* stream);

/* Height field against convex geoms, stand-alone (host pointers; the narrow phase flyb_step runs for the terrain of vision_guided_flight, vnl_ray/fly_envs.py:160-206,
 * tasks/arenas/hills.py:160-166, 186-221: the `terrain` hfield geom that MuJoCo collides with the fly through mjc_ConvexHField).
 * data [nrow, ncol] float32 heights (x elevation scale), pose12 = position (3) and row-major orientation (9) of the hfield geom,
 * size4 = x half extent, y half extent, elevation scale, base thickness.  Query q: types[q] (FLYB_GEOM_SPHERE / CAPSULE /
 * ELLIPSOID / CYLINDER), geoms16[q] = world position (3), orientation (9), size (3), margin.  out [nquery, 1 + 7 maxcon]:
 * number of contacts, then dist, position (3), normal (3, field -> geom) of each.  Error text: flyb_last_error(NULL). */
int flyb_hfield_contacts(int device, const float* data, int nrow, int ncol, const double* pose12, const double* size4, int nquery,
                         const int32_t* types, const double* geoms16, int maxcon, double* out);

/* The fly's eye cameras as a ray caster over the terrain, stand-alone (host pointers, every cell walked: the reference path flyb_render is tested against; the `right_eye` / `left_eye` observables,
 * vnl_ray/fruitfly/fruitfly.py:676-708 with tasks/vision_flight.py:25-26: 32 x 32 pixels, fovy 150; the reference renders them
 * through MuJoCo's OpenGL renderer).  data [nterrain, nrow, ncol] heights over [-half_extent, half_extent]^2, the hfield geom
 * z_offset (-0.01) below the ground plane; camera i looks at terrain terrain_id[i] from cams12[i] = position (3) and row-major
 * orientation (9; columns = the camera's x (right), y (up), z (backward) axes in the world).  colors8 = terrain rgb, background
 * rgb, headlight ambient and diffuse.  out [ncam, height, width, 3] uint8.  What is and is not rendered: csrc/fly_eyes.cuh. */
int flyb_render_eyes(int device, const float* data, int nterrain, int nrow, int ncol, float half_extent, float z_offset, int ncam,
                     const int32_t* terrain_id, const float* cams12, int width, int height, float fovy_deg, const float* colors8,
                     uint8_t* out);

/* The actor's policy network, batched (replaces the per-actor TensorFlow policy call of
 * vnl_ray/agents/actors.py:59-89 on the network of agents/network_factory.py:73-101: batch_concat -> acme LayerNormMLP((512, 512,
 * 512, 512), activate_final=True) -> MultivariateNormalDiagHead).  One fused sm_100a kernel per call: bf16 operands, fp32
 * accumulation in tensor memory (tcgen05.mma), TMA-fed weight tiles, activations resident in shared memory
 * (csrc/flyb_policy.cu).  All pointers are DEVICE pointers to fp32 (torch parameters / observation tensors); weights are
 * [out, in] row-major as torch.nn.Linear stores them, wm / ws are the mean and scale heads [act_dim, 512].
 * forward: obs [n_rows, obs_dim]; eps (nullable) standard-normal noise [n_rows, act_dim]; outputs (each nullable) [n_rows,
 * act_dim]: mean (tanh'ed when tanh_mean), scale = softplus(.) * init_scale / ln 2 + min_scale, action = mean + scale * eps.
 * Work is enqueued on `stream`, no synchronisation.  Limits: hidden width 512, 4 layers, act_dim <= 64, obs_dim <= 4096,
 * n_rows <= max_rows of the handle.  Errors: non-zero return, text from flyb_policy_last_error. */
typedef void* flyb_policy_handle;
int flyb_policy_create(int device, int obs_dim, int act_dim, int max_rows, flyb_policy_handle* out);
void flyb_policy_destroy(flyb_policy_handle h);
const char* flyb_policy_last_error(flyb_policy_handle h);
long long flyb_policy_launches(flyb_policy_handle h);   /* kernels launched through this handle so far */
int flyb_policy_set_weights(flyb_policy_handle h, const float* w1, const float* b1, const float* ln_g, const float* ln_b,
                            const float* w2, const float* b2, const float* w3, const float* b3, const float* w4, const float* b4,
                            const float* wm, const float* bm, const float* ws, const float* bs, float ln_eps, float init_scale,
                            float min_scale, int tanh_mean, void* stream);
int flyb_policy_forward(flyb_policy_handle h, const float* obs, int n_rows, const float* eps, float* mean, float* scale,
                        float* action, void* stream);

/* One step of the actor's n-step transition adder for N environments (replaces the per-environment calls of acme's
 * NStepTransitionAdder.add that vnl_ray/agents/actors.py:59-89 makes; adder built in vnl_ray/agents/ray_distributed_dmpo.py:268-300).
 * All pointers are DEVICE buffers owned by the caller (flybody_b200/actors.py: NStepTransitionAdder); the function keeps no state.
 *   ring_obs [N, n_step + 1, obs_dim], ring_act [N, n_step, action_dim], ring_rew / ring_disc [N, n_step], count [N] int64
 *   (transitions added to the running episode), started [N] bytes: the rings, updated in place;
 *   action [N, action_dim], next_obs [N, obs_dim], reward / step_discount [N], step_type [N] int32 (0 FIRST, 1 MID, 2 LAST): the step.
 *   An env whose step_type is FIRST starts a new episode (its action is ignored, as the env ignored it).
 *   out_*: the transition acme writes after every step, from s = max(0, t - n_step) over h = min(t, n_step) steps:
 *   out_obs = o_s, out_act = a_s, out_return = sum_k (prod_{j<k} discount d_{s+j}) r_{s+k}, out_discount = discount^(h-1) prod d,
 *   out_valid [N] bytes (the env is inside an episode), out_start / out_horizon [N] int64; the transition ends at next_obs.
 *   written [1] int64 += number of valid transitions; any_ended [1] int32 = 1 when a valid env's step_type is LAST (the caller then
 *   flushes the shrinking-horizon transitions of those envs).  Enqueued on `stream`; returns non-zero on bad arguments. */
int flyb_adder_step(int device, int num_envs, int obs_dim, int action_dim, int n_step, float discount, float* ring_obs,
                    float* ring_act, float* ring_rew, float* ring_disc, int64_t* count, uint8_t* started, const float* action,
                    const float* next_obs, const float* reward, const float* step_discount, const int32_t* step_type,
                    float* out_obs, float* out_act, float* out_return, float* out_discount, uint8_t* out_valid, int64_t* out_start,
                    int64_t* out_horizon, int64_t* written, int32_t* any_ended, void* stream);

/* Measurement utility (no reference counterpart): FP32 fused-multiply-add throughput of the device in TFLOP/s, measured with
 * a register-resident FMA micro-benchmark - the denominator of the FP32-pipe fractions bench.py reports. */
int flyb_fma_peak(int device, double* tflops);

#ifdef __cplusplus
}
#endif
#endif
