/* flyb_kern.cu - one instantiation of the step kernel per translation unit (see fly_kernel.cuh).
 * Compiled with -DFLY_TU_WARPS=.. -DFLY_TU_LEVEL=.. -DFLY_TU_SPEC=.. -DFLY_TU_NAME=flyb_kernel_<W>_<L>_<g|sN>. */
#include "fly_kernel.cuh"
extern "C" const void* FLY_TU_NAME() { return (const void*)fly_step_kernel<FLY_TU_WARPS, FLY_TU_LEVEL, FLY_TU_SPEC>; }
