/* flyb_kern.cu - one instantiation of the step kernel per translation unit (see fly_kernel.cuh).
 * Compiled with -DFLY_NL=.. (lanes per env) -DFLY_TU_WARPS=.. -DFLY_TU_LEVEL=.. -DFLY_TU_SPEC=.. (model index, -1 = any model). */
#include "fly_dims_gen.h"
#if FLY_TU_SPEC >= 0 && !defined(FLY_NO_CONST_DIMS)
/* this unit's model: its array sizes as constants (fly_model.h: FLY_DIM) */
template <int S> struct FlyDimsOf;
#define X(i, a, b, c, d, e, f, g, h, k, l, p, q, r, s, t, u, v, hf)                                                    \
  template <> struct FlyDimsOf<i> {                                                                       \
    static constexpr int nq = a, nv = b, na = c, nu = d, nbody = e, njnt = f, ngeom = g, nsite = h, nM = k, \
                         nsensordata = l, nfluidbox = p, nfluidgeom = q, npair = r, nsensor = s, ntrunk = t,   \
                         gsteps = u, nlevel = v, nhfield = hf;                                            \
  };
FLY_SPEC_MODELS(X)
#undef X
#define FLY_DIM(m, f) (FlyDimsOf<FLY_TU_SPEC>::f)
#endif
#include "fly_kernel.cuh"
/* the unit announces its kernel to the host code (flyb_cuda.cu: fly_register_kernel) when the library is loaded: which
 * instantiations exist is decided by the build list alone (flybody_b200/engine.py: KERNEL_UNITS) */
namespace {
struct FlyKernelRegistrar {
  FlyKernelRegistrar() {
    fly_register_kernel(FLY_NL, FLY_TU_WARPS, FLY_TU_LEVEL, FLY_TU_SPEC, (const void*)fly_step_kernel<FLY_NL, FLY_TU_WARPS, FLY_TU_LEVEL, FLY_TU_SPEC>);
  }
} fly_kernel_registrar;
}
