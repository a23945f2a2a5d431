/* flyb_actor.cu - the per-step work of the actor's n-step transition adder as ONE kernel.
 *
 * The reference's actors hand every step to acme's NStepTransitionAdder (vnl_ray/agents/ray_distributed_dmpo.py:268-300 builds it
 * with n_step / discount of the config; vnl_ray/agents/actors.py:59-89 calls add_first / add), one environment and one Python call
 * at a time.  flybody_b200/actors.py keeps N environments' ring buffers on the device; as torch indexing operations one `add` is
 * ~40 launches that move the 48 MB observation batch four times (1.2 ms of a 50 ms control step at 16 384 envs).  Here a warp per
 * environment does the whole step: start a new episode where the env was auto-reset, append (a_t, r_t, d_t, o_t+1) to the rings,
 * and assemble the transition acme writes after every step - from s = max(0, t - n) over h = min(t, n) steps:
 *   (o_s, a_s, R = sum_{k<h} (prod_{j<k} discount d_{s+j}) r_{s+k}, D = discount^(h-1) prod_{k<h} d_{s+k}, o_t+1)
 * as fixed-shape [N, ...] outputs with a `valid` mask.  The shrinking-horizon flush at an episode's end stays with the caller (rare);
 * `any_ended` tells it when.  No state of its own: every buffer is the caller's. */
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/flyb.h"

namespace {

constexpr int kFirst = 0, kLast = 2;   /* dm_env.StepType */

__global__ void __launch_bounds__(256) fly_adder_step_kernel(
    int N, int O, int A, int n, float gamma, float* __restrict__ ring_obs, float* __restrict__ ring_act, float* __restrict__ ring_rew,
    float* __restrict__ ring_disc, long long* __restrict__ count, unsigned char* __restrict__ started, const float* __restrict__ action,
    const float* __restrict__ next_obs, const float* __restrict__ reward, const float* __restrict__ discount,
    const int* __restrict__ step_type, float* __restrict__ out_obs, float* __restrict__ out_act, float* __restrict__ out_R,
    float* __restrict__ out_D, unsigned char* __restrict__ out_valid, long long* __restrict__ out_start, long long* __restrict__ out_hor,
    unsigned long long* __restrict__ written, int* __restrict__ any_ended) {
  const int lane = threadIdx.x & 31;
  const int e = (int)((blockIdx.x * (size_t)blockDim.x + threadIdx.x) >> 5);
  if (e >= N) return;
  const int st = step_type[e];
  const bool first = st == kFirst;
  const bool was_started = started[e] != 0;
  const bool live = was_started && !first;
  long long c = first ? 0 : count[e];
  const float* src = next_obs + (size_t)e * O;
  float* ro = ring_obs + (size_t)e * (n + 1) * O;
  float* ra = ring_act + (size_t)e * n * A;
  const int slot = (int)(c % n), oslot = (int)((c + 1) % (n + 1));
  /* unconditional appends: an env that is not live has count 0 (just (re)started, or never started), so they land in slots 0 / 1,
     which its first live step overwrites before anything reads them */
  for (int i = lane; i < O; i += 32) {
    const float x = src[i];
    ro[(size_t)oslot * O + i] = x;
    if (first) ro[i] = x;   /* the new episode's o_0 */
  }
  for (int i = lane; i < A; i += 32) ra[(size_t)slot * A + i] = action[(size_t)e * A + i];
  if (lane == 0) {
    ring_rew[(size_t)e * n + slot] = reward[e];
    ring_disc[(size_t)e * n + slot] = discount[e];
  }
  __syncwarp();
  if (live) c += 1;
  const long long start = c - n > 0 ? c - n : 0;
  int hor = (int)(c < n ? c : n);
  if (hor < 1) hor = 1;
  /* n-step return and discount over the horizon (every lane alike: 2 n cached loads) */
  float R = 0.f, run = 1.f;
  int idx = (int)(start % n);
  for (int k = 0; k < hor; k++) {
    R += ring_rew[(size_t)e * n + idx] * run;
    run *= gamma * ring_disc[(size_t)e * n + idx];
    if (++idx == n) idx = 0;
  }
  const float* os = ro + (size_t)(start % (n + 1)) * O;
  for (int i = lane; i < O; i += 32) out_obs[(size_t)e * O + i] = os[i];
  const float* as = ra + (size_t)(start % n) * A;
  for (int i = lane; i < A; i += 32) out_act[(size_t)e * A + i] = as[i];
  if (lane == 0) {
    count[e] = c;
    if (first) started[e] = 1;
    out_R[e] = R;
    out_D[e] = run / gamma;
    out_valid[e] = live ? 1 : 0;
    out_start[e] = start;
    out_hor[e] = (long long)hor;
    if (live) {
      atomicAdd(written, 1ull);
      if (st == kLast) atomicOr(any_ended, 1);
    }
  }
}

}  // namespace

extern "C" int flyb_adder_step(int device, int num_envs, int obs_dim, int action_dim, int n_step, float discount, float* ring_obs,
                               float* ring_act, float* ring_rew, float* ring_disc, int64_t* count, uint8_t* started,
                               const float* action, const float* next_obs, const float* reward, const float* step_discount,
                               const int32_t* step_type, float* out_obs, float* out_act, float* out_return, float* out_discount,
                               uint8_t* out_valid, int64_t* out_start, int64_t* out_horizon, int64_t* written, int32_t* any_ended,
                               void* stream) {
  if (num_envs < 1 || obs_dim < 1 || action_dim < 1 || n_step < 1 || !(discount > 0.f)) return 1;
  if (!ring_obs || !ring_act || !ring_rew || !ring_disc || !count || !started || !action || !next_obs || !reward || !step_discount ||
      !step_type || !out_obs || !out_act || !out_return || !out_discount || !out_valid || !out_start || !out_horizon || !written ||
      !any_ended)
    return 1;
  if (cudaSetDevice(device) != cudaSuccess) return 1;
  cudaStream_t st = (cudaStream_t)stream;
  if (cudaMemsetAsync(any_ended, 0, sizeof(int32_t), st) != cudaSuccess) return 1;
  const unsigned blocks = (unsigned)(((size_t)num_envs * 32 + 255) / 256);
  fly_adder_step_kernel<<<blocks, 256, 0, st>>>(num_envs, obs_dim, action_dim, n_step, discount, ring_obs, ring_act, ring_rew, ring_disc,
                                                (long long*)count, started, action, next_obs, reward, step_discount, step_type, out_obs,
                                                out_act, out_return, out_discount, out_valid, (long long*)out_start,
                                                (long long*)out_horizon, (unsigned long long*)written, any_ended);
  return cudaGetLastError() == cudaSuccess ? 0 : 1;
}
