/* fly_kernel.cuh - the step kernel template and its argument block.
 *
 * Every instantiation (warps per block, shared-memory level, model specialisation) is compiled in a translation unit of
 * its own (flyb_kern.cu, one nvcc process each, in parallel - flybody_b200/engine.py: build_library) and handed to the
 * host code in flyb_cuda.cu as a function pointer: the fully inlined kernel takes ~45 s to compile and there are nine. */
#ifndef FLY_KERNEL_CUH_
#define FLY_KERNEL_CUH_
#include <cuda_runtime.h>
#include "fly_env.cuh"
#include "fly_dims_gen.h"

/* sizes of the shipped task models: kernels with a compile-time workspace layout (fly_ws.cuh: ws_layout_c) */
template <int S> struct SpecDims;
#define X(i, ...) template <> struct SpecDims<i> { static constexpr WsDims value = {__VA_ARGS__}; };
FLY_SPEC_MODELS(X)
#undef X

struct KArgs {
  FlyClips clips;
  FlyWbpg wb;
  FlyIO<float> io;
  char* bulk;
  size_t bulk_stride, fast_stride;
  int* counter;
  const int* perm;   /* nullable: position in the launch order -> env id */
  int num_envs, level;
};

/* SPEC >= 0: the model is entry SPEC of fly_dims_gen.h and array offsets are compile-time constants; SPEC < 0: any model,
 * offsets from the kernel parameter `off` */
/* NL (= FLY_NL of the translation unit) is a template parameter only so that units compiled for different lane counts do not
 * define the same symbol: their host stubs would be merged at link time and a launch could run the other unit's code */
template <int NL, int WARPS, int LEVEL, int SPEC>
__global__ void __launch_bounds__(WARPS * 32, 1) fly_step_kernel(const __grid_constant__ DModel<float> m,
                                                                const __grid_constant__ KArgs a,
                                                                const __grid_constant__ WsOff off) {
  static_assert(NL == FLY_NL, "lanes per env are fixed per translation unit");
  extern __shared__ __align__(16) char smem[];
#ifdef FLY_SYNC_GROUP
  constexpr int GW = FLY_SYNC_GROUP;   /* warps of one lockstep group: fetches its own envs, runs on its own barrier */
#else
  constexpr int GW = WARPS;
#endif
  static_assert(WARPS % GW == 0, "lockstep group must divide the block");
  __shared__ int s_base[WARPS / GW];
  const int warp = threadIdx.x >> 5;
  const int eiw = warp * FLY_SUBS + FLY_SUB();            /* this env group's index in the block: FLY_SUBS envs per warp */
  const int slot = blockIdx.x * (WARPS * FLY_SUBS) + eiw;
  const int grp = warp / GW, gwarp = warp % GW;
  /* every thread keeps its warp's array pointers as (arena base + constant offset): registers / immediates, with the
     address space of each array known to the compiler */
  Ws<float> w;
  if constexpr (SPEC >= 0) {
    constexpr WsOff OFF = ws_layout_c<float>(SpecDims<SPEC>::value, LEVEL, FLY_NL);
    ws_attach<float, LEVEL>(w, OFF, smem + eiw * a.fast_stride, a.bulk + (size_t)slot * a.bulk_stride);
  } else {
    ws_attach<float, LEVEL>(w, off, smem + eiw * a.fast_stride, a.bulk + (size_t)slot * a.bulk_stride);
  }
  for (;;) {  /* the group takes GW * FLY_SUBS consecutive envs at a time and its warps step them stage by stage together */
    FLY_BLOCK_SYNC();
    if (gwarp == 0 && (threadIdx.x & 31) == 0) s_base[grp] = atomicAdd(a.counter, GW * FLY_SUBS);
    FLY_BLOCK_SYNC();
    const int base = s_base[grp];
    const int pos = base + gwarp * FLY_SUBS + FLY_SUB();
    if (base >= a.num_envs) break;
    fly_env_run<float>(m, w, a.clips, a.wb, a.io, pos < a.num_envs ? (a.perm ? a.perm[pos] : pos) : -1);
  }
}


/* Which (lanes per env, warps per block, shared-memory level, model) instantiations exist is the build list of
 * flybody_b200/engine.py (KERNEL_UNITS); each unit registers its kernel here when the library is loaded. */
void fly_register_kernel(int lanes, int warps, int level, int spec, const void* fn);
#endif
