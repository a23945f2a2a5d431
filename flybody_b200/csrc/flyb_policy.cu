/* flyb_policy.cu - the actor's policy network as ONE fused sm_100a kernel (BASELINE.json configs[3]).
 *
 * The network of vnl_ray/agents/network_factory.py:73-101 (acme LayerNormMLP((512, 512, 512, 512), activate_final=True) ->
 * MultivariateNormalDiagHead):  x = tanh(LayerNorm(W1 obs + b1)); x = elu(Wl x + bl), l = 2..4; mean = Wm x + bm;
 * scale = softplus(Ws x + bs) * init_scale / ln 2 + min_scale; action = mean + scale * eps.
 * For 16 384 envs that is 2 x 16 384 x (741 x 512 + 3 x 512^2 + 512 x 118) = 40 GFLOP per control step: the one GEMM-shaped piece of
 * the actor side, so it runs on the 5th-generation tensor cores:
 *
 *   - a CTA per 128 observation rows, CTAs in PAIRS (cluster of 2, tcgen05.mma.cta_group::2, M = 256): the leader issues the MMAs of
 *     both; each CTA's tensor core takes its own 128 rows and its own half of the weight tile.  18 warps: warps 0-15 epilogue (a
 *     thread per row = TMEM lane and quarter of the columns), warp 16 TMA producer, warp 17 MMA issuer (one elected thread each);
 *   - the whole 128 x 512 fp32 layer output lives in TMEM (512 columns), written by tcgen05.mma.kind::f16 (bf16 x bf16 -> fp32,
 *     N = 256 per instruction), read back by tcgen05.ld.32x32b for bias + LayerNorm + tanh / ELU;
 *   - activations never leave the SM: the epilogue writes them as bf16 into shared memory in the canonical K-major
 *     SWIZZLE_128B layout, where the next layer's MMAs read them as the A operand (128 KB: 8 K-chunks of 128 rows x 128 B);
 *   - MMA and epilogue overlap by N halves: the epilogue of half 0 runs under the MMAs of half 1, the next layer's half 0 starts
 *     on K chunks 0-3 under the epilogue of half 1;
 *   - weights (bf16, [out, in] row-major = K-major, L2-resident: 2.3 MB) and the first layer's input tiles stream through a
 *     6-stage TMA pipeline (cp.async.bulk.tensor.2d.cta_group::2, SWIZZLE_128B, complete_tx on the leader's mbarriers), freed by
 *     tcgen05.commit multicast to both CTAs.
 *
 * Every spin on an mbarrier is bounded (trap after ~2 s) so that a protocol bug cannot hang the device. */
#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>

#include <new>
#include <string>

#include "../../include/flyb.h"

namespace {

constexpr int PH = 512;            /* hidden width of every layer */
constexpr int PKC = 64;            /* bf16 elements per K chunk = one 128-byte swizzle row */
constexpr int PM = 128;            /* rows per CTA */
constexpr int PNH = 256;           /* N per MMA instruction */
constexpr int PSTAGES = 6;
constexpr int PA_BYTES = PM * PKC * 2;      /* 16 KB: an A tile (128 rows x 64 bf16) */
constexpr int PCLUSTER = 2;                 /* a CTA pair (cta_group::2): the MMAs span both CTAs' 128 rows, each CTA holds half of a weight tile */
constexpr int PB_BYTES = PNH / PCLUSTER * PKC * 2;   /* 16 KB: this CTA's half of a B tile (128 of its 256 weight rows x 64 bf16) */
constexpr int PSTAGE_BYTES = PB_BYTES;       /* a stage holds that half; the first layer's A tile of the stage lies in the (then unused) activation buffer */
constexpr int PACT_BYTES = PM * PH * 2;     /* 128 KB */
static_assert(PSTAGES <= 7, "layer 0 stages its A tiles in K chunks 0 .. PSTAGES - 1 of the activation buffer, chunk 7 holds the LayerNorm sums");
constexpr int PHEAD_N = 128;                /* head output columns: mean at 0.., scale at 64.. */
constexpr int PSMEM_BYTES = PACT_BYTES + PSTAGES * PSTAGE_BYTES + 1024 /* barriers */ + 1024 /* alignment slack */;
constexpr int PEPI_WARPS = 16;              /* epilogue warps: 4 per quarter of the TMEM lanes, each takes 128 of the 512 columns */
constexpr int PEPI_THREADS = PEPI_WARPS * 32;
constexpr int PTHREADS = PEPI_THREADS + 64; /* + TMA producer warp + MMA issuer warp */

struct PolicyParams {
  const float* b1; const float* ln_g; const float* ln_b;
  const float* b[3];
  const float* bm; const float* bs;
  const float* eps;                /* nullable [N, A] */
  float* mean; float* scale; float* action;   /* each nullable, [N, A] */
  int n_rows, act_dim, nk1;        /* nk1: K chunks of the first layer (obs_dim padded to 64) */
  const uint8_t* xb; uint32_t x_row_bytes;       /* the bf16 observations, for the L2 warm-up */
  const uint8_t* w_ptr[5]; uint32_t w_bytes[5];  /* the bf16 weight matrices, likewise */
  float ln_eps, scale_mul, min_scale;
  int tanh_mean;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint32_t bar, uint32_t parity) {
  uint32_t done;
  asm volatile(
      "{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(done) : "r"(bar), "r"(parity) : "memory");
  return done != 0;
}
/* (the clock is only read once the first probe has failed: the MMA issuer's waits mostly succeed at once, and its instruction
   count per MMA is what paces the tensor core) */
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try(bar, parity)) return;
  const long long t0 = clock64();
  while (!mbar_try(bar, parity))
    if (clock64() - t0 > 4000000000LL) __trap();   /* ~2 s: a protocol bug must not hang the GPU */
}
/* a 2-D tile load issued by either CTA of a pair, completing on a barrier that may lie in the peer CTA (the leader's: the pair's MMAs are issued there) */
__device__ __forceinline__ void tma_load_2d_pair(uint32_t dst, const CUtensorMap* map, int x, int y, uint32_t bar_cluster) {
  asm volatile("cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
               ::"r"(dst), "l"(map), "r"(x), "r"(y), "r"(bar_cluster) : "memory");
}
/* shared::cluster address of the same shared-memory offset in CTA `rank` of the cluster */
__device__ __forceinline__ uint32_t mapa_u32(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
/* arrival on a barrier of the leader CTA.  Default semantics, as in CUTLASS' ClusterBarrier::arrive(cta_id): the writes it publishes
   are the thread's own shared-memory stores, already handed to the async proxy by fence.proxy.async; `.release.cluster` puts a
   MEMBAR.ALL.GPU in front of every arrival (9 % of the kernel's stall samples) */
__device__ __forceinline__ void mbar_arrive_cluster(uint32_t bar_cluster) {
  asm volatile("mbarrier.arrive.shared::cluster.b64 _, [%0];" ::"r"(bar_cluster) : "memory");
}
/* shared-memory matrix descriptor, K-major, SWIZZLE_128B: start >> 4 | LBO (unused) | SBO = 1024 B (8 rows x 128 B) | version 1 */
/* (given the low word, (address & 0x3FFFF) >> 4: 32 bytes further along the swizzled row is + 2) */
__device__ __forceinline__ uint64_t umma_desc_lo(uint32_t lo) {
  return (uint64_t)lo | ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}
/* instruction descriptor of kind::f16: D fp32 (bit 4), A / B bf16 (bits 7, 10), both K-major, N >> 3 at bit 17, M >> 4 at bit 24 */
__device__ __forceinline__ constexpr uint32_t umma_idesc(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
/* issued by the leader CTA of a pair: D (both CTAs' TMEM, 128 lanes each) += A (each CTA's own 128 rows) x B (N split over the CTAs) */
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::2.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
/* arrives on the barrier at this offset in both CTAs of the pair once the MMAs issued so far are complete */
__device__ __forceinline__ void umma_commit(uint32_t bar) {
  asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar), "h"((uint16_t)3) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release;\n\tbarrier.cluster.wait.acquire;" ::: "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, float* v) {
  uint32_t r[32];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]), "=r"(r[17]), "=r"(r[18]),
        "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]),
        "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 32; i++) v[i] = __uint_as_float(r[i]);
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
        "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr) : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int i = 0; i < 16; i++) v[i] = __uint_as_float(r[i]);
}
/* activations are rounded to bf16 (8 mantissa bits) right after: the MUFU approximations (tanh.approx, ex2.approx: ~2^-11 relative)
   are below that rounding */
__device__ __forceinline__ float fast_tanh(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
/* ELU without the range handling of __expf (12 instructions and a predicate shuffle per value in the first version): ex2 of a
   large positive argument is +inf and is not selected */
__device__ __forceinline__ float fast_elu(float x) {
  float t;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(t) : "f"(x * 1.4426950408889634f));
  return x > 0.f ? x : t - 1.0f;
}
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&h);
}
/* 32 activations of a row as the four 16-byte groups of the A-operand buffer */
__device__ __forceinline__ void pack_act32(const float* y, uint4* q) {
#pragma unroll
  for (int g = 0; g < 4; g++) {
    q[g].x = pack_bf16(y[8 * g + 0], y[8 * g + 1]); q[g].y = pack_bf16(y[8 * g + 2], y[8 * g + 3]);
    q[g].z = pack_bf16(y[8 * g + 4], y[8 * g + 5]); q[g].w = pack_bf16(y[8 * g + 6], y[8 * g + 7]);
  }
}
/* ... of row r, columns [32 c, 32 c + 32): K chunk c / 2, 16-byte groups swizzled by row (act_s: shared-space address of the buffer) */
__device__ __forceinline__ void store_packed32(uint32_t act_s, int r, int c, const uint4* q) {
  const uint32_t row = act_s + (c >> 1) * PA_BYTES + r * 128;
  const int g0 = (c & 1) * 4;
#pragma unroll
  for (int g = 0; g < 4; g++)
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row + (((g0 + g) ^ (r & 7)) << 4)), "r"(q[g].x), "r"(q[g].y),
                 "r"(q[g].z), "r"(q[g].w) : "memory");
}

/* 16 activations of row r, columns [64 kc + 16 cg, + 16): 16-byte groups 2 cg, 2 cg + 1 of K chunk kc */
__device__ __forceinline__ void store_act16(uint32_t act_s, int r, int kc, int cg, const float* y) {
  const uint32_t row = act_s + kc * PA_BYTES + r * 128;
#pragma unroll
  for (int g = 0; g < 2; g++)
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(row + (((2 * cg + g) ^ (r & 7)) << 4)),
                 "r"(pack_bf16(y[8 * g + 0], y[8 * g + 1])), "r"(pack_bf16(y[8 * g + 2], y[8 * g + 3])),
                 "r"(pack_bf16(y[8 * g + 4], y[8 * g + 5])), "r"(pack_bf16(y[8 * g + 6], y[8 * g + 7])) : "memory");
}
__device__ __forceinline__ void sts_f32(uint32_t addr, float x) { asm volatile("st.shared.f32 [%0], %1;" ::"r"(addr), "f"(x) : "memory"); }
__device__ __forceinline__ float lds_f32(uint32_t addr) {
  float x;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(x) : "r"(addr) : "memory");
  return x;
}

#ifdef FLY_POLICY_PROF   /* kernel experiments: clock64 stamps of two CTAs, read by flyb_policy_prof_dump */
__device__ long long g_pprof[2][6][6];   /* [5][.]: stamps inside the head epilogue */
#define PPROF(slot, l, val) do { if (blockIdx.x == 0 || blockIdx.x == 100) g_pprof[blockIdx.x ? 1 : 0][l][slot] = (val); } while (0)
#else
#define PPROF(slot, l, val) do { } while (0)
#endif

__global__ void __cluster_dims__(PCLUSTER, 1, 1) __launch_bounds__(PTHREADS, 1)
fly_policy_kernel(const __grid_constant__ CUtensorMap map_x, const __grid_constant__ CUtensorMap map_w1,
                  const __grid_constant__ CUtensorMap map_w2, const __grid_constant__ CUtensorMap map_w3,
                  const __grid_constant__ CUtensorMap map_w4, const __grid_constant__ CUtensorMap map_wh,
                  const __grid_constant__ PolicyParams p) {
  extern __shared__ uint8_t smem_raw[];
  if (threadIdx.x == PEPI_THREADS + 32) PPROF(4, 5, clock64());
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint8_t* act = smem;                                   /* 8 chunks x 16 KB */
  uint8_t* stage = smem + PACT_BYTES;                    /* PSTAGES x 16 KB: this CTA's halves of the B tiles */
  uint64_t* bars = reinterpret_cast<uint64_t*>(stage + PSTAGES * PSTAGE_BYTES);
  /* bars: full[PSTAGES], empty[PSTAGES], acc[2] (an N half of the layer is complete), ka (the second half's MMAs are past K chunk 3),
     epi[5] (the epilogue has drained half 0 from TMEM and written K chunks 0-3 of the next layer's input | ... K chunk 4 | 5 | 6 | 7 of half 1) ;
     then the TMEM base address */
  const uint32_t bar_full = smem_u32(bars), bar_empty = smem_u32(bars + PSTAGES);
  const uint32_t bar_acc = smem_u32(bars + 2 * PSTAGES), bar_ka = smem_u32(bars + 2 * PSTAGES + 2);
  const uint32_t bar_epi = smem_u32(bars + 2 * PSTAGES + 3);
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * PSTAGES + 8);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row0 = blockIdx.x * PM;

  if (threadIdx.x == 0) {
    for (int s = 0; s < PSTAGES; s++) { mbar_init(bar_full + 8 * s, 1); mbar_init(bar_empty + 8 * s, 1); }
    mbar_init(bar_acc, 1); mbar_init(bar_acc + 8, 1);
    mbar_init(bar_ka, 1);
    for (int i = 0; i < 5; i++) mbar_init(bar_epi + 8 * i, PCLUSTER * PEPI_WARPS);   /* (the leader's: a lane per epilogue warp of both CTAs) */
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == PEPI_WARPS) {   /* the allocating warp also frees: all 512 columns (one CTA per SM: nothing to contend with) */
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(512) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = *tmem_slot;
  /* A CTA pair works as one: the leader (rank 0) issues tcgen05.mma.cta_group::2 with M = 256 - each CTA's tensor core takes its own
     128 rows (its A tiles, its TMEM) and HALF of the weight tile from its own shared memory, the other half comes from the peer.
     Per MMA a CTA reads 4 + 4 KB of shared memory instead of 4 + 8, and the TMA ring writes half as much: the single-CTA kernel
     ran its MMAs at 210 cycles each against a floor of 128 with shared memory at ~100 of its 128 B / cycle.  Both CTAs' loads
     complete on the LEADER's full barriers; the leader's commits arrive on the empty / acc / ka barriers of both; the epilogue
     warps of both arrive on the leader's epi barriers. */
  uint32_t cta_rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(cta_rank));
  const bool leader = cta_rank == 0;
  cluster_sync_all();   /* the barriers of both CTAs are initialised before a peer signals them */

  /* The item sequence both pipeline ends walk: layer 0: 2 N halves x nk1 chunks (A + B tiles); layers 1-3: 2 x 8 (B only, A =
     the activation buffer); layer 4 = the head: 8 chunks of a 128-row B tile.
     MMA and epilogue overlap by halves.  TMEM holds one layer's accumulators (2 halves of 256 columns) and the activation buffer one
     layer's input, so: the epilogue of half 0 runs while the MMAs of half 1 are issued, and stores into K chunks 0-3 once those
     MMAs are past them (ka); the next layer's half 0 starts on K chunks 0-3 as soon as the epilogue of half 0 has written them and
     drained its TMEM columns (epi[0]) - i.e. during the epilogue of half 1 - and goes on to chunks 4-7 after epi[1]. */
  if (warp == PEPI_WARPS && lane == 0) {
    /* ------------------------------------------------------------------------------------------------ TMA producer */
    const CUtensorMap* wmap[5] = {&map_w1, &map_w2, &map_w3, &map_w4, &map_wh};
    uint32_t s = 0, ph = 1;   /* stage; parity of the empty phase to wait for (a fresh barrier passes parity 1) */
    for (int l = 0; l < 5; l++) {
      const int nk = l == 0 ? p.nk1 : PH / PKC, nhalf = l == 4 ? 1 : 2;
      for (int nh = 0; nh < nhalf; nh++)
        for (int kc = 0; kc < nk; kc++) {
          mbar_wait(bar_empty + 8 * s, ph);
          const uint32_t a_dst = smem_u32(act + s * PA_BYTES), b_dst = smem_u32(stage + s * PSTAGE_BYTES);
          const int share = (l == 4 ? PHEAD_N : PNH) / PCLUSTER;   /* weight rows of the tile this CTA holds */
          const uint32_t bytes = PCLUSTER * ((l == 0 ? PA_BYTES : 0) + share * PKC * 2);   /* of both CTAs */
          const uint32_t full = mapa_u32(bar_full + 8 * s, 0);
          if (leader) mbar_expect_tx(bar_full + 8 * s, bytes);   /* (the peer's bytes may land before this: the transaction count runs negative until then) */
          if (l == 0) tma_load_2d_pair(a_dst, &map_x, kc * PKC, row0, full);
          tma_load_2d_pair(b_dst, wmap[l], kc * PKC, nh * PNH + cta_rank * share, full);
          if (++s == PSTAGES) { s = 0; ph ^= 1; }
        }
    }
  } else if (warp == PEPI_WARPS + 1 && lane == 0 && leader) {
    /* ------------------------------------------------------------------------------- MMA issuer (of the pair: leader only) */
    uint32_t s = 0, ph = 0;   /* stage and its phase parity */
    const uint32_t act_lo = (smem_u32(act) & 0x3FFFF) >> 4, stage_lo = (smem_u32(stage) & 0x3FFFF) >> 4;
    for (int l = 0; l < 5; l++) {
      const int nk = l == 0 ? p.nk1 : PH / PKC, nhalf = l == 4 ? 1 : 2;
      const int kka = nk - 1 < 3 ? nk - 1 : 3;
      const uint32_t idesc = l == 4 ? umma_idesc(PCLUSTER * PM, PHEAD_N) : umma_idesc(PCLUSTER * PM, PNH);
#ifdef FLY_POLICY_PROF
      long long wfull = 0;
#endif
      PPROF(0, l, clock64());
      for (int nh = 0; nh < nhalf; nh++) {
        const uint32_t d_tmem = tmem + nh * PNH;
        for (int kc = 0; kc < nk; kc++) {
          if (l > 0 && nh == 0 && (kc == 0 || kc >= 4)) {   /* K chunks 0-3 (chunk kc) of this layer's input are written; TMEM half 0 is
                                                              drained (half 1: when chunk 7 is, before the MMAs of half 1) */
            mbar_wait(bar_epi + (kc ? 8 * (kc - 3) : 0), (l - 1) & 1);
            asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          }
#ifdef FLY_POLICY_PROF
          const long long tw = clock64();
#endif
          mbar_wait(bar_full + 8 * s, ph);
#ifdef FLY_POLICY_PROF
          if (kc == 0 && nh == 0) PPROF(1, l, tw);
          wfull += clock64() - tw;
#endif
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          const uint32_t a_lo = act_lo + (l == 0 ? s : (uint32_t)kc) * (PA_BYTES >> 4);
          const uint32_t b_lo = stage_lo + s * (PSTAGE_BYTES >> 4);
#pragma unroll
          for (int k = 0; k < PKC / 16; k++)   /* UMMA K = 16 bf16 = 32 bytes along the swizzled row */
            umma_bf16(d_tmem, umma_desc_lo(a_lo + 2 * k), umma_desc_lo(b_lo + 2 * k), idesc, k ? 1u : (uint32_t)kc);
          umma_commit(bar_empty + 8 * s);     /* releases the stage in both CTAs once these MMAs have read it */
          if (nh == 1 && kc == kka) umma_commit(bar_ka);
          if (++s == PSTAGES) { s = 0; ph ^= 1; }
        }
        umma_commit(bar_acc + 8 * nh);        /* this half of the layer's accumulator is complete */
      }
      PPROF(2, l, wfull); PPROF(3, l, clock64());
    }
  } else if (warp < PEPI_WARPS) {
    /* ---------------------------------------------------------------------------------------------------- epilogue */
    const int q = warp & 3, cg = warp >> 2;                /* quarter of the TMEM lanes this warp may read; its column group */
    const int r = q * 32 + lane;                           /* row of the tile = TMEM lane */
    const uint32_t trow = tmem + ((uint32_t)(q * 32) << 16);
    const uint32_t act_s = smem_u32(act);
    const uint32_t epi_leader = mapa_u32(bar_epi, 0);
    /* per N half the thread takes the two 32-column chunks 8 h + 2 cg + {0, 1} of its row, i.e. K chunk 4 h + cg of the next layer */
    const uint32_t red = act_s + 7 * PA_BYTES;                   /* LayerNorm partial sums [4 column groups][128 rows][2]: K chunk 7 is
                                                                    neither a staging slot of layer 0's A tiles nor written before the sums are read */
    float v[32];
    {   /* L2 warm-up while the first tiles are on their way: the TMA ring holds six tiles, so a tile that misses L2 costs a DRAM
           round trip per six items.  These warps are idle until the first half of layer 0 is done: they ask for the CTA's own
           observation rows and for its 1 / gridDim.x of every weight matrix, a line per thread. */
      const uint8_t* xp = p.xb + (size_t)row0 * p.x_row_bytes;
      const uint32_t xbytes = PM * p.x_row_bytes;
      for (uint32_t o = threadIdx.x * 128u; o < xbytes; o += PEPI_THREADS * 128u)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(xp + o));
#pragma unroll
      for (int w = 0; w < 5; w++) {
        const uint32_t share = ((p.w_bytes[w] + gridDim.x - 1) / gridDim.x + 127u) & ~127u;
        const uint32_t lo = blockIdx.x * share, hi = lo + share < p.w_bytes[w] ? lo + share : p.w_bytes[w];
        for (uint32_t o = lo + threadIdx.x * 128u; o < hi; o += PEPI_THREADS * 128u)
          asm volatile("prefetch.global.L2 [%0];" ::"l"(p.w_ptr[w] + o));
      }
    }
    for (int l = 0; l < 4; l++) {
      if (l == 0) {         /* bias, LayerNorm over the 512 features of the row, tanh */
        float sum = 0, sq = 0;
        for (int h = 0; h < 2; h++) {
          mbar_wait(bar_acc + 8 * h, 0);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          if (threadIdx.x == 0 && h == 0) PPROF(4, l, clock64());
          for (int j = 0; j < 2; j++) {
            const int c = 8 * h + 2 * cg + j;
            tmem_ld32(trow + 32 * c, v);
#pragma unroll
            for (int j4 = 0; j4 < 8; j4++) {
              const float4 b = __ldg(reinterpret_cast<const float4*>(p.b1 + 32 * c) + j4);
              const float x0 = v[4 * j4] + b.x, x1 = v[4 * j4 + 1] + b.y, x2 = v[4 * j4 + 2] + b.z, x3 = v[4 * j4 + 3] + b.w;
              sum += (x0 + x1) + (x2 + x3);
              sq += (x0 * x0 + x1 * x1) + (x2 * x2 + x3 * x3);
            }
          }
        }
        sts_f32(red + (cg * PM + r) * 8, sum); sts_f32(red + (cg * PM + r) * 8 + 4, sq);
        asm volatile("bar.sync 1, %0;" ::"n"(PEPI_THREADS) : "memory");
        sum = 0; sq = 0;
#pragma unroll
        for (int g = 0; g < 4; g++) { sum += lds_f32(red + (g * PM + r) * 8); sq += lds_f32(red + (g * PM + r) * 8 + 4); }
        asm volatile("bar.sync 1, %0;" ::"n"(PEPI_THREADS) : "memory");   /* everyone has read the sums: act may be written */
        const float mu = sum * (1.0f / PH);
        const float var = fmaxf(sq * (1.0f / PH) - mu * mu, 0.f);
        const float rstd = rsqrtf(var + p.ln_eps);
        mbar_wait(bar_ka, 0);   /* (complete: both halves are; waited on to keep the phase in step) */
        for (int j = 0; j < 2; j++) {   /* half 0: K chunk cg of the next layer */
          const int c = 2 * cg + j;
          tmem_ld32(trow + 32 * c, v);
#pragma unroll
          for (int j4 = 0; j4 < 8; j4++) {
            const float4 b = __ldg(reinterpret_cast<const float4*>(p.b1 + 32 * c) + j4);
            const float4 g = __ldg(reinterpret_cast<const float4*>(p.ln_g + 32 * c) + j4);
            const float4 e = __ldg(reinterpret_cast<const float4*>(p.ln_b + 32 * c) + j4);
            v[4 * j4 + 0] = fast_tanh((v[4 * j4 + 0] + b.x - mu) * rstd * g.x + e.x);
            v[4 * j4 + 1] = fast_tanh((v[4 * j4 + 1] + b.y - mu) * rstd * g.y + e.y);
            v[4 * j4 + 2] = fast_tanh((v[4 * j4 + 2] + b.z - mu) * rstd * g.z + e.z);
            v[4 * j4 + 3] = fast_tanh((v[4 * j4 + 3] + b.w - mu) * rstd * g.w + e.w);
          }
          uint4 qq[4];
          pack_act32(v, qq);
          store_packed32(act_s, r, c, qq);
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   /* generic-proxy stores -> visible to the tensor core */
        asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(epi_leader);
        for (int kc = 4; kc < 8; kc++) {   /* half 1, all warps on one K chunk at a time: the next layer's MMAs follow chunk by chunk */
          const int c0 = 64 * kc + 16 * cg;
          tmem_ld16(trow + c0, v);
#pragma unroll
          for (int j4 = 0; j4 < 4; j4++) {
            const float4 b = __ldg(reinterpret_cast<const float4*>(p.b1 + c0) + j4);
            const float4 g = __ldg(reinterpret_cast<const float4*>(p.ln_g + c0) + j4);
            const float4 e = __ldg(reinterpret_cast<const float4*>(p.ln_b + c0) + j4);
            v[4 * j4 + 0] = fast_tanh((v[4 * j4 + 0] + b.x - mu) * rstd * g.x + e.x);
            v[4 * j4 + 1] = fast_tanh((v[4 * j4 + 1] + b.y - mu) * rstd * g.y + e.y);
            v[4 * j4 + 2] = fast_tanh((v[4 * j4 + 2] + b.z - mu) * rstd * g.z + e.z);
            v[4 * j4 + 3] = fast_tanh((v[4 * j4 + 3] + b.w - mu) * rstd * g.w + e.w);
          }
          store_act16(act_s, r, kc, cg, v);
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(epi_leader + 8 * (kc - 3));
        }
      } else {              /* bias, ELU */
        const float* bl = p.b[l - 1];
        {   /* half 0 (under the MMAs of half 1): the thread's 64 columns = K chunk cg of the next layer, held until ka */
          mbar_wait(bar_acc, l & 1);
          asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
          if (threadIdx.x == 0) PPROF(4, l, clock64());
          uint4 qq[2][4];
#pragma unroll
          for (int j = 0; j < 2; j++) {
            const int c = 2 * cg + j;
            tmem_ld32(trow + 32 * c, v);
#pragma unroll
            for (int j4 = 0; j4 < 8; j4++) {
              const float4 b = __ldg(reinterpret_cast<const float4*>(bl + 32 * c) + j4);
              v[4 * j4 + 0] = fast_elu(v[4 * j4 + 0] + b.x); v[4 * j4 + 1] = fast_elu(v[4 * j4 + 1] + b.y);
              v[4 * j4 + 2] = fast_elu(v[4 * j4 + 2] + b.z); v[4 * j4 + 3] = fast_elu(v[4 * j4 + 3] + b.w);
            }
            pack_act32(v, qq[j]);
          }
          mbar_wait(bar_ka, l & 1);   /* the MMAs of half 1 have read K chunks 0-3: they may be overwritten */
#pragma unroll
          for (int j = 0; j < 2; j++) store_packed32(act_s, r, 2 * cg + j, qq[j]);
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(epi_leader);
        }
        mbar_wait(bar_acc + 8, l & 1);   /* all of this layer's MMAs are complete */
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        for (int kc = 4; kc < 8; kc++) {   /* half 1 is on the critical path: all warps on one K chunk at a time, the next layer's
                                              MMAs follow chunk by chunk */
          const int c0 = 64 * kc + 16 * cg;
          tmem_ld16(trow + c0, v);
#pragma unroll
          for (int j4 = 0; j4 < 4; j4++) {
            const float4 b = __ldg(reinterpret_cast<const float4*>(bl + c0) + j4);
            v[4 * j4 + 0] = fast_elu(v[4 * j4 + 0] + b.x); v[4 * j4 + 1] = fast_elu(v[4 * j4 + 1] + b.y);
            v[4 * j4 + 2] = fast_elu(v[4 * j4 + 2] + b.z); v[4 * j4 + 3] = fast_elu(v[4 * j4 + 3] + b.w);
          }
          store_act16(act_s, r, kc, cg, v);
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
          __syncwarp();
          if (lane == 0) mbar_arrive_cluster(epi_leader + 8 * (kc - 3));
        }
      }
      if (threadIdx.x == 0) PPROF(5, l, clock64());
    }
    {   /* heads: mean (TMEM columns 0..63) | scale (64..127).  Every warp takes 32 columns of one of them; the values go through shared
           memory (the activation buffer is free: all MMAs are complete) so that the three outputs - a contiguous block of
           128 x act_dim floats each - leave as coalesced stores */
      const int rows_here = p.n_rows - row0 < PM ? p.n_rows - row0 : PM;
      const int valid = rows_here * p.act_dim;
      const size_t o0 = (size_t)row0 * p.act_dim;
      constexpr int kPer = PM * 64 / PEPI_THREADS;   /* output elements per thread at most */
      float ev[kPer];                                /* the noise of this thread's output elements, fetched under the head's MMAs */
#pragma unroll
      for (int k = 0; k < kPer; k++) {
        const int i = threadIdx.x + k * PEPI_THREADS;
        ev[k] = (p.eps && i < valid) ? __ldg(p.eps + o0 + i) : 0.f;
      }
      const bool is_scale = cg >= 2;
      const int c = cg & 1;
      float bv[32];                                  /* this thread's 32 biases (the vectors are zero-padded to 64) */
#pragma unroll
      for (int j4 = 0; j4 < 8; j4++) {
        const float4 b = __ldg(reinterpret_cast<const float4*>((is_scale ? p.bs : p.bm) + 32 * c) + j4);
        bv[4 * j4] = b.x; bv[4 * j4 + 1] = b.y; bv[4 * j4 + 2] = b.z; bv[4 * j4 + 3] = b.w;
      }
      if (threadIdx.x == 0) PPROF(0, 5, clock64());
      mbar_wait(bar_acc, 0);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      if (threadIdx.x == 0) PPROF(4, 4, clock64());
      const uint32_t sm_mean = act_s, sm_scale = act_s + PM * 64 * 4;
      tmem_ld32(trow + (is_scale ? 64 : 0) + 32 * c, v);
      if (threadIdx.x == 0) PPROF(1, 5, clock64());
      if (is_scale) {
#pragma unroll
        for (int j = 0; j < 32; j++) {   /* softplus(x) * init_scale / ln 2 + min_scale */
          const float x = v[j] + bv[j];
          v[j] = (fmaxf(x, 0.f) + __logf(1.0f + __expf(-fabsf(x)))) * p.scale_mul + p.min_scale;
        }
      } else if (p.tanh_mean) {
#pragma unroll
        for (int j = 0; j < 32; j++) v[j] = tanhf(v[j] + bv[j]);
      } else {
#pragma unroll
        for (int j = 0; j < 32; j++) v[j] += bv[j];
      }
      const uint32_t dst = (is_scale ? sm_scale : sm_mean) + (uint32_t)(r * p.act_dim + 32 * c) * 4u;
#pragma unroll
      for (int j = 0; j < 32; j++)
        if (32 * c + j < p.act_dim) sts_f32(dst + 4 * j, v[j]);
      if (threadIdx.x == 0) PPROF(2, 5, clock64());
      asm volatile("bar.sync 1, %0;" ::"n"(PEPI_THREADS) : "memory");
      if (threadIdx.x == 0) PPROF(3, 5, clock64());
#pragma unroll
      for (int k = 0; k < kPer; k++) {
        const int i = threadIdx.x + k * PEPI_THREADS;
        if (i < valid) {
          const float mean = lds_f32(sm_mean + 4 * i), scale = lds_f32(sm_scale + 4 * i);
          if (p.mean) p.mean[o0 + i] = mean;
          if (p.scale) p.scale[o0 + i] = scale;
          if (p.action) p.action[o0 + i] = mean + scale * ev[k];
        }
      }
      if (threadIdx.x == 0) PPROF(5, 4, clock64());
    }
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  cluster_sync_all();   /* no CTA leaves while a peer may still signal its barriers */
  if (warp == PEPI_WARPS) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512) : "memory");
  }
  if (threadIdx.x == PEPI_THREADS + 32) PPROF(5, 5, clock64());
}

/* fp32 [rows, cols] -> bf16 [rows_pad, cols_pad], zero padded (weights once, observations every call): a block per PCAST_ROWS rows, a
 * thread per pair of columns - coalesced 4-byte loads (rows of 741 floats start unaligned) and stores, no integer division, and the
 * loads of all the block's rows in flight before the first store (a row per block kept 2.3 TB/s of reads in flight: 21 us for 16 384
 * observation rows) */
constexpr int PCAST_ROWS = 8;
__global__ void __launch_bounds__(256) fly_policy_cast(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, int rows, int cols,
                                                        int rows_pad, int cols_pad, int dst_row0) {
  for (int r0 = blockIdx.x * PCAST_ROWS; r0 < rows_pad; r0 += gridDim.x * PCAST_ROWS) {
    for (int c2 = threadIdx.x; 2 * c2 < cols_pad; c2 += blockDim.x) {
      const int c = 2 * c2;
      float a[PCAST_ROWS], b[PCAST_ROWS];
#pragma unroll
      for (int k = 0; k < PCAST_ROWS; k++) {
        const int r = r0 + k;
        const float* s = src + (size_t)r * cols;
        a[k] = (r < rows && c < cols) ? __ldg(s + c) : 0.f;
        b[k] = (r < rows && c + 1 < cols) ? __ldg(s + c + 1) : 0.f;
      }
#pragma unroll
      for (int k = 0; k < PCAST_ROWS; k++) {
        const int r = r0 + k;
        if (r < rows_pad)
          reinterpret_cast<__nv_bfloat162*>(dst + (size_t)(dst_row0 + r) * cols_pad)[c2] = __floats2bfloat162_rn(a[k], b[k]);
      }
    }
  }
}

struct Policy {
  int device = 0, obs_dim = 0, act_dim = 0, k1pad = 0, max_rows = 0;
  __nv_bfloat16 *w1 = nullptr, *w2 = nullptr, *w3 = nullptr, *w4 = nullptr, *wh = nullptr, *xb = nullptr;
  float* vec = nullptr;            /* b1 | ln_g | ln_b | b2 | b3 | b4 | bm (64) | bs (64) */
  CUtensorMap m_w1, m_w2, m_w3, m_w4, m_wh, m_x;
  int x_rows = -1;
  float ln_eps = 1e-5f, init_scale = 0.7f, min_scale = 1e-6f;
  int tanh_mean = 0;
  long long launches = 0;
  std::string err;
};

typedef CUresult (*EncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                             const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                             CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeFn encode_fn() {
  static EncodeFn fn = nullptr;
  if (!fn) {
    void* ptr = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = (EncodeFn)ptr;
  }
  return fn;
}

/* bf16 [rows, cols] row-major, box = 64 columns x box_rows, 128-byte swizzle */
bool make_map(CUtensorMap* map, void* base, int rows, int cols, int box_rows) {
  EncodeFn fn = encode_fn();
  if (!fn) return false;
  cuuint64_t gdim[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t gstride[1] = {(cuuint64_t)cols * 2};
  cuuint32_t box[2] = {PKC, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  return fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

int pfail(Policy* P, const char* what, cudaError_t e = cudaSuccess) {
  P->err = std::string(what) + (e != cudaSuccess ? std::string(": ") + cudaGetErrorString(e) : std::string());
  return 1;
}

#define PCK(call)                                       \
  do {                                                  \
    cudaError_t e_ = (call);                            \
    if (e_ != cudaSuccess) return pfail(P, #call, e_);  \
  } while (0)

}  // namespace

extern "C" {

int flyb_policy_create(int device, int obs_dim, int act_dim, int max_rows, flyb_policy_handle* out) {
  if (!out) return 1;
  *out = nullptr;
  if (obs_dim < 1 || obs_dim > 4096 || act_dim < 1 || act_dim > 64 || max_rows < 1) return 1;
  Policy* P = new (std::nothrow) Policy();
  if (!P) return 1;
  *out = P;
  P->device = device; P->obs_dim = obs_dim; P->act_dim = act_dim;
  P->k1pad = (obs_dim + PKC - 1) / PKC * PKC;
  P->max_rows = (max_rows + PM * PCLUSTER - 1) / (PM * PCLUSTER) * (PM * PCLUSTER);
  PCK(cudaSetDevice(device));
  int cc_major = 0;
  PCK(cudaDeviceGetAttribute(&cc_major, cudaDevAttrComputeCapabilityMajor, device));
  if (cc_major != 10) return pfail(P, "flyb_policy needs an sm_100 device (tcgen05 / TMEM)");
  if (!encode_fn()) return pfail(P, "cuTensorMapEncodeTiled is not available from this driver");
  PCK(cudaMalloc(&P->w1, (size_t)PH * P->k1pad * 2));
  PCK(cudaMalloc(&P->w2, (size_t)PH * PH * 2));
  PCK(cudaMalloc(&P->w3, (size_t)PH * PH * 2));
  PCK(cudaMalloc(&P->w4, (size_t)PH * PH * 2));
  PCK(cudaMalloc(&P->wh, (size_t)PHEAD_N * PH * 2));
  PCK(cudaMalloc(&P->xb, (size_t)P->max_rows * P->k1pad * 2));
  PCK(cudaMalloc(&P->vec, (size_t)(6 * PH + 128) * sizeof(float)));
  PCK(cudaMemset(P->vec, 0, (size_t)(6 * PH + 128) * sizeof(float)));
  PCK(cudaMemset(P->wh, 0, (size_t)PHEAD_N * PH * 2));
  if (!make_map(&P->m_w1, P->w1, PH, P->k1pad, PNH / PCLUSTER) || !make_map(&P->m_w2, P->w2, PH, PH, PNH / PCLUSTER) ||
      !make_map(&P->m_w3, P->w3, PH, PH, PNH / PCLUSTER) || !make_map(&P->m_w4, P->w4, PH, PH, PNH / PCLUSTER) ||
      !make_map(&P->m_wh, P->wh, PHEAD_N, PH, PHEAD_N / PCLUSTER))
    return pfail(P, "cuTensorMapEncodeTiled failed for a weight matrix");
  PCK(cudaFuncSetAttribute(fly_policy_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, PSMEM_BYTES));
  return 0;
}

void flyb_policy_destroy(flyb_policy_handle h) {
  Policy* P = (Policy*)h;
  if (!P) return;
  cudaSetDevice(P->device);
  cudaFree(P->w1); cudaFree(P->w2); cudaFree(P->w3); cudaFree(P->w4); cudaFree(P->wh); cudaFree(P->xb); cudaFree(P->vec);
  delete P;
}

const char* flyb_policy_last_error(flyb_policy_handle h) { return h ? ((Policy*)h)->err.c_str() : "null handle"; }

long long flyb_policy_launches(flyb_policy_handle h) { return h ? ((Policy*)h)->launches : 0; }

int flyb_policy_set_weights(flyb_policy_handle h, const float* w1, const float* b1, const float* ln_g, const float* ln_b,
                            const float* w2, const float* b2, const float* w3, const float* b3, const float* w4, const float* b4,
                            const float* wm, const float* bm, const float* ws, const float* bs, float ln_eps, float init_scale,
                            float min_scale, int tanh_mean, void* stream) {
  Policy* P = (Policy*)h;
  if (!P) return 1;
  if (!w1 || !b1 || !ln_g || !ln_b || !w2 || !b2 || !w3 || !b3 || !w4 || !b4 || !wm || !bm || !ws || !bs)
    return pfail(P, "flyb_policy_set_weights: null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  PCK(cudaSetDevice(P->device));
  fly_policy_cast<<<256, 256, 0, st>>>(w1, P->w1, PH, P->obs_dim, PH, P->k1pad, 0);
  fly_policy_cast<<<256, 256, 0, st>>>(w2, P->w2, PH, PH, PH, PH, 0);
  fly_policy_cast<<<256, 256, 0, st>>>(w3, P->w3, PH, PH, PH, PH, 0);
  fly_policy_cast<<<256, 256, 0, st>>>(w4, P->w4, PH, PH, PH, PH, 0);
  fly_policy_cast<<<64, 256, 0, st>>>(wm, P->wh, P->act_dim, PH, 64, PH, 0);    /* head rows 0..63: mean, 64..127: scale */
  fly_policy_cast<<<64, 256, 0, st>>>(ws, P->wh, P->act_dim, PH, 64, PH, 64);
  const float* src[6] = {b1, ln_g, ln_b, b2, b3, b4};
  for (int i = 0; i < 6; i++) PCK(cudaMemcpyAsync(P->vec + i * PH, src[i], PH * sizeof(float), cudaMemcpyDeviceToDevice, st));
  PCK(cudaMemcpyAsync(P->vec + 6 * PH, bm, P->act_dim * sizeof(float), cudaMemcpyDeviceToDevice, st));
  PCK(cudaMemcpyAsync(P->vec + 6 * PH + 64, bs, P->act_dim * sizeof(float), cudaMemcpyDeviceToDevice, st));
  PCK(cudaGetLastError());
  P->ln_eps = ln_eps; P->init_scale = init_scale; P->min_scale = min_scale; P->tanh_mean = tanh_mean;
  P->launches += 6;
  return 0;
}

int flyb_policy_forward(flyb_policy_handle h, const float* obs, int n_rows, const float* eps, float* mean, float* scale,
                        float* action, void* stream) {
  Policy* P = (Policy*)h;
  if (!P) return 1;
  if (!obs || n_rows < 1) return pfail(P, "flyb_policy_forward: no observations");
  if (n_rows > P->max_rows) return pfail(P, "flyb_policy_forward: more rows than the handle was created for");
  cudaStream_t st = (cudaStream_t)stream;
  PCK(cudaSetDevice(P->device));
  const int rows_pad = (n_rows + PM * PCLUSTER - 1) / (PM * PCLUSTER) * (PM * PCLUSTER);   /* whole clusters; rows beyond n_rows are zero */
  if (P->x_rows != rows_pad) {
    if (!make_map(&P->m_x, P->xb, rows_pad, P->k1pad, PM)) return pfail(P, "cuTensorMapEncodeTiled failed for the observations");
    P->x_rows = rows_pad;
  }
  fly_policy_cast<<<(rows_pad + PCAST_ROWS - 1) / PCAST_ROWS, 256, 0, st>>>(obs, P->xb, n_rows, P->obs_dim, rows_pad, P->k1pad, 0);
  PolicyParams pp;
  pp.b1 = P->vec; pp.ln_g = P->vec + PH; pp.ln_b = P->vec + 2 * PH;
  pp.b[0] = P->vec + 3 * PH; pp.b[1] = P->vec + 4 * PH; pp.b[2] = P->vec + 5 * PH;
  pp.bm = P->vec + 6 * PH; pp.bs = P->vec + 6 * PH + 64;
  pp.eps = eps; pp.mean = mean; pp.scale = scale; pp.action = action;
  pp.n_rows = n_rows; pp.act_dim = P->act_dim; pp.nk1 = P->k1pad / PKC;
  pp.ln_eps = P->ln_eps; pp.scale_mul = P->init_scale / 0.6931471805599453f; pp.min_scale = P->min_scale;
  pp.tanh_mean = P->tanh_mean;
  pp.xb = (const uint8_t*)P->xb; pp.x_row_bytes = (uint32_t)P->k1pad * 2;
  const __nv_bfloat16* wp[5] = {P->w1, P->w2, P->w3, P->w4, P->wh};
  const uint32_t wb[5] = {(uint32_t)PH * P->k1pad * 2, PH * PH * 2, PH * PH * 2, PH * PH * 2, PHEAD_N * PH * 2};
  for (int i = 0; i < 5; i++) { pp.w_ptr[i] = (const uint8_t*)wp[i]; pp.w_bytes[i] = wb[i]; }
  fly_policy_kernel<<<rows_pad / PM, PTHREADS, PSMEM_BYTES, st>>>(P->m_x, P->m_w1, P->m_w2, P->m_w3, P->m_w4, P->m_wh, pp);
  PCK(cudaGetLastError());
  P->launches += 2;
  return 0;
}

#ifdef FLY_POLICY_PROF
void flyb_policy_prof_dump(void) {
  long long h[2][6][6];
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(h, g_pprof, sizeof(h));
  for (int c = 0; c < 2; c++) {
    const long long t0 = h[c][0][0];
    for (int l = 0; l < 5; l++)
      printf("cta %d layer %d: mma start %lld, first full-wait at %lld, waited on full %lld, issue end %lld | epilogue: acc at %lld, done %lld\n",
             c ? 100 : 0, l, h[c][l][0] - t0, h[c][l][1] - t0, h[c][l][2], h[c][l][3] - t0, h[c][l][4] - t0, h[c][l][5] - t0);
    printf("cta %d head epilogue: noise loads issued %lld, TMEM read %lld, staged %lld, past the barrier %lld\n", c ? 100 : 0,
           h[c][5][0] - t0, h[c][5][1] - t0, h[c][5][2] - t0, h[c][5][3] - t0);
    printf("cta %d kernel entry %lld, exit %lld\n", c ? 100 : 0, h[c][5][4] - t0, h[c][5][5] - t0);
  }
}
#endif

}  // extern "C"
