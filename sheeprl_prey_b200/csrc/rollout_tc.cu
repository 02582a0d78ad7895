// Imagination rollout as ONE persistent kernel (bf16 tensor-core mode).
//
// Reference: the loop dreamer_v3.py:219-257 -- RSSM.imagination (agent.py:439-455: RecurrentModel -> LayerNormGRUCell -> transition
// MLP -> unimix -> one-hot sample) followed by Actor.forward (agent.py:682-739) on every imagined state, H times, starting from every
// posterior state (N = T*B rows).  Discrete actions: forward only, nothing is kept for a backward pass through the dynamics; the
// actor's activations ARE kept (the policy loss reuses them instead of a second actor pass, engine.py).
//
// The N rows are independent, so the kernel splits them into blocks of 128 rows (the M of tcgen05.mma) and gives each block to one
// thread-block cluster of CL CTAs; CTA c owns 1/CL of the output features of every layer.  A rollout step is a chain of 7 dense
// layers; each is run as a "stage":
//     producer warps: TMA (SWIZZLE_128B boxes) of the 128-row activation tile [128 x 64 k] (warp 0) and of this CTA's weight rows
//                     [n_own x 64 k] (warp 18) into one mbarrier ring;
//     MMA warp      : tcgen05.mma M128 x N n_own x K16, fp32 accumulator in TMEM; the four MMAs of a k-block and its commits are
//                     issued under one elect.sync (an election per instruction cost a quarter of the issue loop);
//     epilogue warps: 16 warps, thread = (row, quarter of the CTA's columns); ONE TMEM -> register pass per stage; bias / one-hot
//                     action gather; LayerNorm statistics over the CTA's columns (partials of the four quarters through shared
//                     memory), exchanged with the peers by st.async on the receiver's mbarrier; normalise, SiLU / GRU gates /
//                     unimix + inverse-CDF sampling; results to global memory (fp32 for the losses, bf16 as the next operand).
//                     (A first version with one thread per row and 4 epilogue warps was latency-starved: 57-75 k clk per stage.)
// The layer outputs of all CTAs of a cluster form the next layer's input: they meet in L2-resident global buffers (the trajectory
// and actor-activation buffers that the behaviour phase needs anyway) and are re-read by TMA after a cluster barrier; generic-proxy
// stores are ordered before the async-proxy reads with fence.proxy.async.
// The 128-row activation tile of a k-block is the same for all CL CTAs of a cluster: each CTA loads RM / CL of its rows and multicasts
// them into every CTA's ring slot (cp.async.bulk.tensor ... .multicast::cluster), so the tile crosses L2 -> SM once per cluster
// instead of CL times; ring slots are therefore recycled cluster-wide (every MMA warp commits to every CTA's empty barrier).  TMA
// multicast loads here read only this GPU's memory, since on B200 a TMA multicast load from another GPU's memory hung the GPU: the
// host entry point refuses buffers that are not allocations of the launching device.
#include "common.cuh"
#include "dv3_b200.h"
#include <cuda.h>
#include <cuda_bf16.h>
#include <stdio.h>
#include <stdlib.h>

int dv3_internal_make_map_2d(CUtensorMap* tm, const void* base, unsigned long long rows, unsigned long long cols, unsigned long long ld_elems,
                             unsigned box_rows, unsigned box_cols);      // gemm_tma.cu

namespace {

constexpr int RM = 128;                 // rows per block = M of the MMA
constexpr int KB = 64;                  // k per ring stage (one 128 B swizzle row of bf16)
constexpr int NSTG = 4;                 // (5 stages fit and were measured: no change, the stages are not bound by the load round trip)
constexpr int A_BYTES = RM * KB * 2;    // 16 KB
constexpr int NMAX = 192;               // most output features per CTA and stage
constexpr int B_BYTES = NMAX * KB * 2;  // 24 KB
constexpr int NEPI = 512;               // 16 epilogue warps: 4 per TMEM lane quadrant, each taking a quarter of the stage's columns
constexpr int NTHR = 96 + NEPI;         // warp 0 activation producer, warp 1 MMA, warps 2-17 epilogue, warp 18 weight producer
constexpr int PB_WARP = 18;             // (a TMA issue costs ~160 clk: one warp feeding both operands bounded the short-K stages)
constexpr int MAXL = DV3_ROLLOUT_MAX_LAYERS;
constexpr int MAXCL = 16;

struct RollK {
  CUtensorMap m_traj, m_feat, m_thid, m_act[MAXL], m_wz, m_wgru, m_wt0, m_wt1, m_wa[MAXL], m_wh;
  Dv3RolloutArgs a;
  int CL, tmem_cols;
  long long* dbg;
};

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n.reg .b32 rx;\n.reg .pred px;\nelect.sync rx|px, %1;\n@px mov.s32 %0, 1;\n}\n" : "+r"(pred) : "r"(0xFFFFFFFFu));
  return pred;
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n.reg .pred P1;\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\nselp.u32 %0, 1, 0, P1;\n}\n"
               : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
// Bounded, call-free wait (a protocol error must trap, not hang; no call so that the single-role warps stay in uniform registers)
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try(bar, parity)) {
    if (++spins > (1u << 24)) __trap();
  }
}
__device__ __forceinline__ void tma_load_2d(uint32_t smem_dst, const CUtensorMap* tmap, uint64_t* bar, int c_inner, int c_outer) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_dst),
               "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c_inner), "r"(c_outer) : "memory");
}
__device__ __forceinline__ void tma_load_2d_mc(uint32_t smem_dst, const CUtensorMap* tmap, uint64_t* bar, int c_inner, int c_outer, uint16_t mask) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.multicast::cluster [%0], [%1, {%3, %4}], [%2], %5;"
               ::"r"(smem_dst), "l"(reinterpret_cast<uint64_t>(tmap)), "r"(smem_u32(bar)), "r"(c_inner), "r"(c_outer), "h"(mask) : "memory");
}
__device__ __forceinline__ uint64_t make_desc_sw128(uint32_t smem_addr) {   // K-major SWIZZLE_128B: SBO 1024, version 1, layout 2
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__device__ __forceinline__ uint32_t mapa(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_async_f2(uint32_t dst_cluster_addr, float x, float y, uint32_t bar_cluster_addr) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];" ::"r"(dst_cluster_addr), "f"(x), "f"(y),
               "r"(bar_cluster_addr) : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ float fsigmoid(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }
__device__ __forceinline__ float ftanh(float x) { const float e = __expf(-2.f * fabsf(x)); return copysignf(__fdividef(1.f - e, 1.f + e), x); }

// One dense layer of the chain for this CTA: which activation tile(s) to stream, which weight rows, how many output columns.
struct Stage {
  const CUtensorMap* a0; int a0_col, a0_row, K0;      // first source of the 128-row operand: columns [a0_col, a0_col + K0) of rows a0_row..
  const CUtensorMap* a1; int a1_col, a1_row, K1;      // optional second source (concatenated along K)
  const CUtensorMap* b; int nb, b_row[3], b_rows;     // nb boxes of b_rows weight rows each (GRU: the three gates)
  int n_pad;                                          // N of the MMA (multiple of 16, >= nb * b_rows)
  int npass, pass_row_step, pass_tmem_step;           // the stage's N split into npass sweeps over K (pass p: weight rows + p * pass_row_step,
                                                      // accumulator columns + p * pass_tmem_step); 1 except for a GRU with D / CL > 64
};
struct Pipe {            // ring position, advanced identically by the producer and the MMA warp
  uint32_t slot, phase;
};

__global__ void __launch_bounds__(NTHR, 1) rollout_tc_kernel(const __grid_constant__ RollK k) {
  const Dv3RolloutArgs& a = k.a;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* ring_a = smem;
  uint8_t* ring_b = smem + NSTG * A_BYTES;
  float2* lnp = reinterpret_cast<float2*>(smem + NSTG * (A_BYTES + B_BYTES));       // [MAXCL][RM] (mean, M2) pushed by the peers
  float2* red = lnp + MAXCL * RM;                                                   // [4][RM] partials of the four column quarters
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(red + 4 * RM);
  uint64_t* empty_bar = full_bar + NSTG;
  uint64_t* acc_bar = empty_bar + NSTG;
  uint64_t* stat_bar = acc_bar + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(stat_bar + 1);

  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int CL = k.CL;
  uint32_t crank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
  const int c = (int)crank;
  const int blk = (int)blockIdx.x / CL;                     // row block
  const int N = a.N, H = a.H, S = a.S, G = a.G, Kc = a.Kc, D = a.D, U = a.U, Hh = a.Hh, A = a.A, lat = a.lat, L = a.L, Ua = a.Ua;
  const int uper = U / CL, dper = D / CL, hper = Hh / CL, gper = G / CL, sper = gper * Kc, aper = Ua / CL;
  const int u0 = c * uper, d0 = c * dper, h0 = c * hper, s0 = c * sper, au0 = c * aper;
  const int dpp = dper < 64 ? dper : 64;                    // GRU units per pass (accumulator: [pass][gate][dpp] columns)
  const int row_blk = blk * RM;                             // first row of the block inside a step's N rows

  for (int i = tid; i < NSTG * (A_BYTES + B_BYTES) / 16; i += NTHR) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0u, 0u, 0u, 0u);
  if (tid == 0) {
    for (int s = 0; s < NSTG; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], (uint32_t)k.CL); }   // a slot is free when ALL CTAs' MMAs left it
    mbar_init(acc_bar, 1);
    mbar_init(stat_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(k.tmem_cols));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_slot;
  cluster_sync_all();

  int dbg_i = 0;
#define RDBG() do { if (k.dbg && blockIdx.x == 0 && tid == 64 && dbg_on && dbg_i < 60) k.dbg[dbg_i++] = clock64(); } while (0)
  bool dbg_on = false;
  Pipe pp = {0u, 0u};          // producer / MMA ring position (each role keeps its own copy in step)
  uint32_t acc_par = 0, stat_par = 0;

  // ---- the three roles of one stage -------------------------------------------------------------------------------------------------
  const int a_slice_row = c * (RM / CL);
  const uint32_t a_slice_off = (uint32_t)a_slice_row * KB * 2;
  const uint16_t mc_mask = (uint16_t)((1u << CL) - 1u);
  // two producer warps share the ring: warp 0 arms the slot's barrier and loads (multicasts) the activation rows, warp PB_WARP loads
  // the weight boxes; both wait for the slot on the same empty barrier and advance the same ring position
  auto produce = [&](const Stage& st, bool weights) {
    const int nk0 = st.K0 / KB, nk = nk0 + st.K1 / KB;
    const uint32_t bytes = A_BYTES + (uint32_t)st.nb * st.b_rows * KB * 2;
#pragma unroll 1
    for (int ps = 0; ps < st.npass; ++ps) {
      const int roff = ps * st.pass_row_step;
#pragma unroll 1
      for (int kb = 0; kb < nk; ++kb) {
        mbar_wait(&empty_bar[pp.slot], pp.phase ^ 1u);
        if (elect_one()) {
          if (!weights) {
            mbar_expect_tx(&full_bar[pp.slot], bytes);
            const uint32_t da = smem_u32(ring_a + pp.slot * A_BYTES);
            // own RM / CL rows of the activation tile, multicast into the same slot of every CTA of the cluster (this GPU's memory only)
            if (kb < nk0) tma_load_2d_mc(da + a_slice_off, st.a0, &full_bar[pp.slot], st.a0_col + kb * KB, st.a0_row + a_slice_row, mc_mask);
            else tma_load_2d_mc(da + a_slice_off, st.a1, &full_bar[pp.slot], st.a1_col + (kb - nk0) * KB, st.a1_row + a_slice_row, mc_mask);
          } else {
            const uint32_t db = smem_u32(ring_b + pp.slot * B_BYTES);
            for (int j = 0; j < st.nb; ++j) tma_load_2d(db + (uint32_t)j * st.b_rows * KB * 2, st.b, &full_bar[pp.slot], kb * KB, st.b_row[j] + roff);
          }
        }
        __syncwarp();
        if (++pp.slot == NSTG) { pp.slot = 0; pp.phase ^= 1u; }
      }
    }
  };
  auto issue = [&](const Stage& st) {
    const int nk = st.K0 / KB + st.K1 / KB;
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(st.n_pad >> 3) << 17) | ((uint32_t)(RM >> 4) << 24);
    const uint64_t ad0 = make_desc_sw128(smem_u32(ring_a)), bd0 = make_desc_sw128(smem_u32(ring_b));
#pragma unroll 1
    for (int ps = 0; ps < st.npass; ++ps) {
      const uint32_t dcol = tmem_base + (uint32_t)(ps * st.pass_tmem_step);
      const bool last_pass = (ps == st.npass - 1);
#pragma unroll 1
      for (int kb = 0; kb < nk; ++kb) {
        mbar_wait(&full_bar[pp.slot], pp.phase);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        const uint64_t ad = ad0 + (uint64_t)((pp.slot * A_BYTES) >> 4), bd = bd0 + (uint64_t)((pp.slot * B_BYTES) >> 4);
        if (elect_one()) {              // one election per k-block: the four MMAs and the commits are issued by the same lane
#pragma unroll
          for (int kk = 0; kk < KB / 16; ++kk) {
            const uint32_t accum = (kb > 0 || kk > 0) ? 1u : 0u;
            asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(dcol),
                         "l"(ad + (uint64_t)(kk * 2)), "l"(bd + (uint64_t)(kk * 2)), "r"(idesc), "r"(accum) : "memory");
          }
          asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(&empty_bar[pp.slot])),
                       "h"(mc_mask) : "memory");
          if (last_pass && kb == nk - 1)
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(acc_bar)) : "memory");
        }
        __syncwarp();
        if (++pp.slot == NSTG) { pp.slot = 0; pp.phase ^= 1u; }
      }
    }
  };
  // ---- epilogue: thread = (row rl, column quarter cq).  Hardware warp w reaches TMEM lanes 32 (w % 4) ..: the four warps of one cq
  // cover the four lane quadrants.
  const int ewarp = warp - 2, cq = ewarp >> 2, rl = 32 * (warp & 3) + lane;
  const uint32_t t_lane = tmem_base + ((uint32_t)(32 * (warp & 3)) << 16);
  auto epi_bar = [&]() { asm volatile("bar.sync 1, %0;" ::"n"(NEPI) : "memory"); };
  auto ld8 = [&](int col, float* v) {
    uint32_t r[8];
    asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(t_lane + (uint32_t)col));
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[j]);
  };
  // four 8-column loads in flight, one wait: x[0..32) = accumulator columns [col, col + 32) of this thread's row (chunks at or past
  // `ncols` columns are skipped and left untouched).  The registers pass through the wait statement so that no use is scheduled above it.
  auto ld32 = [&](int col, int ncols, float* v) {
    uint32_t r[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) r[j] = 0u;
#pragma unroll
    for (int c0 = 0; c0 < 32; c0 += 8)
      if (c0 < ncols)
        asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
                     : "=r"(r[c0]), "=r"(r[c0 + 1]), "=r"(r[c0 + 2]), "=r"(r[c0 + 3]), "=r"(r[c0 + 4]), "=r"(r[c0 + 5]), "=r"(r[c0 + 6]), "=r"(r[c0 + 7])
                     : "r"(t_lane + (uint32_t)(col + c0)));
#pragma unroll
    for (int c0 = 0; c0 < 32; c0 += 8)
      asm volatile("tcgen05.wait::ld.sync.aligned;"
                   : "+r"(r[c0]), "+r"(r[c0 + 1]), "+r"(r[c0 + 2]), "+r"(r[c0 + 3]), "+r"(r[c0 + 4]), "+r"(r[c0 + 5]), "+r"(r[c0 + 6]), "+r"(r[c0 + 7])
                   :: "memory");
#pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
  };
  // three 8-column loads (the GRU's three gates) in flight, one wait
  auto ld8x3 = [&](int c0, int c1, int c2, float* v0, float* v1, float* v2) {
    uint32_t r[24];
    const int cols[3] = {c0, c1, c2};
#pragma unroll
    for (int g = 0; g < 3; ++g)
      asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
                   : "=r"(r[8 * g]), "=r"(r[8 * g + 1]), "=r"(r[8 * g + 2]), "=r"(r[8 * g + 3]), "=r"(r[8 * g + 4]), "=r"(r[8 * g + 5]),
                     "=r"(r[8 * g + 6]), "=r"(r[8 * g + 7])
                   : "r"(t_lane + (uint32_t)cols[g]));
#pragma unroll
    for (int g = 0; g < 3; ++g)
      asm volatile("tcgen05.wait::ld.sync.aligned;"
                   : "+r"(r[8 * g]), "+r"(r[8 * g + 1]), "+r"(r[8 * g + 2]), "+r"(r[8 * g + 3]), "+r"(r[8 * g + 4]), "+r"(r[8 * g + 5]),
                     "+r"(r[8 * g + 6]), "+r"(r[8 * g + 7])
                   :: "memory");
#pragma unroll
    for (int j = 0; j < 8; ++j) { v0[j] = __uint_as_float(r[j]); v1[j] = __uint_as_float(r[8 + j]); v2[j] = __uint_as_float(r[16 + j]); }
  };
  auto st8_f32 = [](float* dst, const float* v) {
    *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(dst + 4) = make_float4(v[4], v[5], v[6], v[7]);
  };
  auto st8_bf16 = [](__nv_bfloat16* dst, const float* v) {
    uint32_t w[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) { __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]); w[j] = *reinterpret_cast<uint32_t*>(&h); }
    *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
  };
  // LayerNorm statistics of row rl over columns spread over the 4 quarters of this CTA and the CL CTAs of the cluster: quarter partials
  // (sum, sum of squares) meet in shared memory; the cq == 0 thread of the row sends the CTA's (mean, M2) to every peer's lnp[c][rl]
  // (st.async, completes on the peer's barrier); then every thread merges the CL partials (Chan, equal counts).
  auto ln_exchange = [&](float s, float q, int n, float eps, float& mu, float& rs) {
    red[cq * RM + rl] = make_float2(s, q);
    if (tid == 64) mbar_expect_tx(stat_bar, (uint32_t)CL * RM * 8);
    epi_bar();
    if (cq == 0) {
      float ss = 0.f, qq = 0.f;
#pragma unroll
      for (int x = 0; x < 4; ++x) { const float2 p = red[x * RM + rl]; ss += p.x; qq += p.y; }
      const float mu_c = ss / (float)n, m2_c = fmaxf(qq - ss * mu_c, 0.f);
      const uint32_t my = smem_u32(&lnp[c * RM + rl]), bar = smem_u32(stat_bar);
      for (int p = 0; p < CL; ++p) st_async_f2(mapa(my, (uint32_t)p), mu_c, m2_c, mapa(bar, (uint32_t)p));
    }
    mbar_wait(stat_bar, stat_par);
    stat_par ^= 1u;
    float sm = 0.f;
    for (int p = 0; p < CL; ++p) sm += lnp[p * RM + rl].x;
    mu = sm / (float)CL;
    float m2 = 0.f;
    for (int p = 0; p < CL; ++p) { const float2 pc = lnp[p * RM + rl]; const float d = pc.x - mu; m2 += pc.y + (float)n * d * d; }
    rs = rsqrtf(m2 / (float)(n * CL) + eps);
  };
  // dense layer + LayerNorm + SiLU on the thread's n / 4 columns, streamed from TMEM twice in 8-column chunks (statistics, then
  // normalisation: keeping the columns in registers across the exchange spilled at 576 threads per CTA):
  // acc (+ bias) (+ gathered action rows) -> [pre] -> act (fp32) / act16 (bf16) at columns col0 + [0, n)
  auto chunk = [&](int tcol, int gcol, const float* bias, const float* add_tab, int add_ld, const int* add_idx, float* x) {
    ld8(tcol, x);
    if (bias) {
      const float4 b0 = __ldg(reinterpret_cast<const float4*>(bias + gcol)), b1 = __ldg(reinterpret_cast<const float4*>(bias + gcol + 4));
      x[0] += b0.x; x[1] += b0.y; x[2] += b0.z; x[3] += b0.w; x[4] += b1.x; x[5] += b1.y; x[6] += b1.z; x[7] += b1.w;
    }
    if (add_tab) {
      for (int h = 0; h < a.nh; ++h) {
        const float* tr = add_tab + (size_t)(a.head_off[h] + add_idx[h]) * add_ld + gcol;
        const float4 t0 = __ldg(reinterpret_cast<const float4*>(tr)), t1 = __ldg(reinterpret_cast<const float4*>(tr + 4));
        x[0] += t0.x; x[1] += t0.y; x[2] += t0.z; x[3] += t0.w; x[4] += t1.x; x[5] += t1.y; x[6] += t1.z; x[7] += t1.w;
      }
    }
  };
  // ---- staged stores.  With thread = row every global store of the epilogue would touch 32 different cache lines (16 B used of each
  // 32 B sector): measured, the LSU then bounds the stage (8-22 k clk per epilogue).  The TMA ring is idle between the stage's last
  // MMA (acc_bar) and the cluster barrier that ends the stage, so results are parked there as fp32 [RM][n + 4] tiles (pitch = 16 mod
  // 64 B: conflict-free 16 B accesses for thread = row) and written out row-contiguously by all 16 warps; the proxy fence of
  // end_stage orders these generic-proxy writes before the next stage's TMA writes to the same bytes.
  float* stg0 = reinterpret_cast<float*>(ring_a);
  float* stg1 = reinterpret_cast<float*>(ring_a + 72 * 1024);
  const int etid = tid - 64;
  auto sts8 = [](float* dst, const float* v) {
    *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(dst + 4) = make_float4(v[4], v[5], v[6], v[7]);
  };
  auto lds8 = [](const float* src, float* v) {
    const float4 a0 = *reinterpret_cast<const float4*>(src), a1 = *reinterpret_cast<const float4*>(src + 4);
    v[0] = a0.x; v[1] = a0.y; v[2] = a0.z; v[3] = a0.w; v[4] = a1.x; v[5] = a1.y; v[6] = a1.z; v[7] = a1.w;
  };
  // tile [RM][n] (pitch floats) -> rows grow0.. of a global fp32 matrix (and / or its bf16 mirror) at column col0; all epilogue threads
  auto write_tile = [&](const float* tile, int pitch, int n, float* g32, __nv_bfloat16* g16, size_t grow0, int ld, int col0) {
    const int n4 = n >> 2, total = RM * n4;
    for (int i = etid; i < total; i += NEPI) {
      const int r = i / n4, c4 = i - r * n4;
      const float4 v = *reinterpret_cast<const float4*>(tile + r * pitch + 4 * c4);
      const size_t o = (grow0 + r) * (size_t)ld + col0 + 4 * c4;
      if (g32) *reinterpret_cast<float4*>(g32 + o) = v;
      if (g16) {
        __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
        uint2 pk;
        pk.x = *reinterpret_cast<uint32_t*>(&lo);
        pk.y = *reinterpret_cast<uint32_t*>(&hi);
        *reinterpret_cast<uint2*>(g16 + o) = pk;
      }
    }
  };
  auto ln_silu_epilogue = [&](int n, const float* bias, const float* gamma, const float* beta, float eps, const float* add_tab, int add_ld,
                              const int* add_idx, float* pre, float* act, __nv_bfloat16* act16, int ld_out, size_t grow, int col0, float* gmean,
                              float* grstd) {
    const int nq = (((n + 3) >> 2) + 7) & ~7, cb = cq * nq;   // this thread's columns [cb, min(cb + nq, n)) of the CTA's n (n % 8 == 0)
    const int pitch = n + 4;
    float s = 0.f, q = 0.f;
#pragma unroll 1
    for (int c0 = 0; c0 < nq && cb + c0 < n; c0 += 8) {       // one TMEM pass: statistics; the pre-LayerNorm values wait in stg1
      float x[8];
      chunk(cb + c0, col0 + cb + c0, bias, add_tab, add_ld, add_idx, x);
#pragma unroll
      for (int j = 0; j < 8; ++j) { s += x[j]; q = fmaf(x[j], x[j], q); }
      sts8(stg1 + rl * pitch + cb + c0, x);
    }
    float mu, rs;
    ln_exchange(s, q, n, eps, mu, rs);
    if (gmean && c == 0 && cq == 0) { gmean[grow] = mu; grstd[grow] = rs; }
#pragma unroll 1
    for (int c0 = 0; c0 < nq && cb + c0 < n; c0 += 8) {
      float x[8];
      const int gcol = col0 + cb + c0;
      lds8(stg1 + rl * pitch + cb + c0, x);                   // (own values)
      const float4 g0 = __ldg(reinterpret_cast<const float4*>(gamma + gcol)), g1 = __ldg(reinterpret_cast<const float4*>(gamma + gcol + 4));
      const float4 e0 = __ldg(reinterpret_cast<const float4*>(beta + gcol)), e1 = __ldg(reinterpret_cast<const float4*>(beta + gcol + 4));
      const float gg[8] = {g0.x, g0.y, g0.z, g0.w, g1.x, g1.y, g1.z, g1.w}, bb[8] = {e0.x, e0.y, e0.z, e0.w, e1.x, e1.y, e1.z, e1.w};
#pragma unroll
      for (int j = 0; j < 8; ++j) { const float t = (x[j] - mu) * rs * gg[j] + bb[j]; x[j] = t * fsigmoid(t); }
      sts8(stg0 + rl * pitch + cb + c0, x);
    }
    epi_bar();
    const size_t grow0 = grow - (size_t)rl;
    if (pre) write_tile(stg1, pitch, n, pre, nullptr, grow0, ld_out, col0);
    write_tile(stg0, pitch, n, act, act16, grow0, ld_out, col0);
  };
  auto wait_acc = [&]() {
    RDBG();
    mbar_wait(acc_bar, acc_par);
    acc_par ^= 1u;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    RDBG();
  };
  auto end_stage = [&]() {
    // this CTA's global stores (generic proxy) must be visible to the peers' TMA loads (async proxy) of the next stage
    RDBG();
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    if (warp >= 2 && warp != PB_WARP) {          // only the epilogue warps wrote anything (global results, staging tiles in the ring)
      asm volatile("fence.proxy.async.global;" ::: "memory");
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    }
    cluster_sync_all();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  };

  __nv_bfloat16* traj16 = reinterpret_cast<__nv_bfloat16*>(a.traj16);
  __nv_bfloat16* feat16 = reinterpret_cast<__nv_bfloat16*>(a.feat16);
  __nv_bfloat16* thid16 = reinterpret_cast<__nv_bfloat16*>(a.thid16);

  // the actor on the rows of imagination step i (agent.py:682-739): L x [Linear, LN 1e-3, SiLU], head (+ bias), unimix, sample
  auto policy = [&](int i) {
    const size_t grow = (size_t)i * N + row_blk + rl;        // this thread's row in the [M1, .] buffers
    const int arow = i * N + row_blk;
    for (int l = 0; l < L; ++l) {
      Stage st;
      if (l == 0) { st.a0 = &k.m_traj; st.a0_col = 0; st.a0_row = arow; st.K0 = lat; }
      else { st.a0 = &k.m_act[l - 1]; st.a0_col = 0; st.a0_row = arow; st.K0 = Ua; }
      st.a1 = nullptr; st.a1_col = st.a1_row = st.K1 = 0;
      st.b = &k.m_wa[l]; st.nb = 1; st.b_row[0] = au0; st.b_rows = aper; st.n_pad = aper; st.npass = 1; st.pass_row_step = st.pass_tmem_step = 0;
      if (warp == 0 || warp == PB_WARP) produce(st, warp == PB_WARP);
      else if (warp == 1) issue(st);
      else {
        wait_acc();
        ln_silu_epilogue(aper, nullptr, a.ga[l], a.ba[l], 1e-3f, nullptr, 0, nullptr, a.apre[l], a.aact[l], reinterpret_cast<__nv_bfloat16*>(a.aact16[l]), Ua,
                         grow, au0, a.amean[l], a.arstd[l]);
      }
      end_stage();
    }
    {   // head: all action logits of a row live in CTA 0 of the cluster; the four column quarters are 32 logits wide (the last one shorter)
      Stage st;
      st.a0 = &k.m_act[L - 1]; st.a0_col = 0; st.a0_row = arow; st.K0 = Ua;
      st.a1 = nullptr; st.a1_col = st.a1_row = st.K1 = 0;
      st.b = &k.m_wh; st.nb = 1; st.b_row[0] = 0; st.b_rows = A; st.n_pad = (A + 15) & ~15; st.npass = 1; st.pass_row_step = st.pass_tmem_step = 0;
      // every CTA runs the head's loads and MMAs (the ring positions and the cluster-wide slot barriers stay in lockstep); only CTA 0
      // has an epilogue: all action logits of a row live there
      if (warp == 0 || warp == PB_WARP) produce(st, warp == PB_WARP);
      else if (warp == 1) issue(st);
      else {
        const float u_head = (c != 0) ? 0.f : ((i == 0) ? a.u_act0[(size_t)(row_blk + rl)] : a.u_act[(size_t)(i - 1) * N + row_blk + rl]);
        wait_acc();
        if (c == 0) {
          // this thread: logits [32 cq, 32 cq + 32) of the row, in registers (one TMEM pass, one exponential per logit); the four
          // quarters meet in shared memory for the row maximum, the softmax denominator, the total unimix mass and the draw
          const int dim = a.head_dim[0], cb = 32 * cq, npad = st.n_pad;
          float x[32];
          ld32(cb, npad - cb, x);
          float mx = -INFINITY;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const int col = cb + j;
            x[j] = (col < dim) ? x[j] + __ldg(a.bh + col) : -INFINITY;
            mx = fmaxf(mx, x[j]);
          }
          const int hpitch = npad + 4;
          if (cb < npad) {
#pragma unroll
            for (int j = 0; j < 32; j += 8)
              if (cb + j < npad) sts8(stg0 + rl * hpitch + cb + j, x + j);
          }
          red[cq * RM + rl].x = mx;
          epi_bar();
          mx = fmaxf(fmaxf(red[rl].x, red[RM + rl].x), fmaxf(red[2 * RM + rl].x, red[3 * RM + rl].x));
          float se = 0.f;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            x[j] = __expf(x[j] - mx);           // exp(-inf) = 0 for the columns past the head
            se += x[j];
          }
          red[cq * RM + rl].y = se;
          epi_bar();
          se = red[rl].y + red[RM + rl].y + red[2 * RM + rl].y + red[3 * RM + rl].y;
          const float mixk = a.unimix / (float)dim, inv = __fdividef(1.f, se);
          float mass = 0.f;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            const float p = x[j] * inv;
            x[j] = (cb + j < dim) ? ((a.unimix > 0.f) ? fminf(fmaxf((1.f - a.unimix) * p + mixk, DV3_EPS32), 1.f - DV3_EPS32) : p) : 0.f;
            mass += x[j];
          }
          epi_bar();                                   // (everyone has read red[].x / .y)
          red[cq * RM + rl].x = mass;
          epi_bar();
          float before = 0.f, tot = 0.f;
#pragma unroll
          for (int qq = 0; qq < 4; ++qq) { const float mq = red[qq * RM + rl].x; tot += mq; if (qq < cq) before += mq; }
          const float thr = u_head * tot;                   // idx = #{k : cumsum_k(p) / tot <= u}
          float cs = before;
          int cnt = 0;
#pragma unroll
          for (int j = 0; j < 32; ++j) { cs += x[j]; cnt += (cb + j < dim && cs <= thr) ? 1 : 0; }
          red[cq * RM + rl].y = __int_as_float(cnt);
          epi_bar();
          if (cq == 0) {
            int idx = __float_as_int(red[rl].y) + __float_as_int(red[RM + rl].y) + __float_as_int(red[2 * RM + rl].y) + __float_as_int(red[3 * RM + rl].y);
            idx = min(idx, dim - 1);
            a.act_idx[grow] = a.f_act ? a.f_act[grow] : idx;
          }
          // the block's logits are one contiguous piece of the [M1][A] output: coalesced copy from the staged tile
          for (int e = etid; e < RM * A; e += NEPI) {
            const int r = e / A, col = e - r * A;
            a.aout[((size_t)arow) * A + e] = stg0[r * hpitch + col];
          }
        }
      }
      end_stage();
    }
  };

  policy(0);
#pragma unroll 1
  for (int i = 1; i <= H; ++i) {
    dbg_on = (i == 2);
    const int prow = (i - 1) * N + row_blk, crow = i * N + row_blk;     // block's first row in the previous / current imagined step
    const size_t gp = (size_t)prow + rl, gc = (size_t)crow + rl;
    const int lrow = row_blk + rl;                                      // row inside the per-step scratch buffers
    // ---- S1: feat = SiLU(LN(Wmlp . [z_{i-1}, a_{i-1}])): dense on the one-hot z (bf16, exact), action columns as a row gather
    {
      Stage st;
      st.a0 = &k.m_traj; st.a0_col = 0; st.a0_row = prow; st.K0 = S;
      st.a1 = nullptr; st.a1_col = st.a1_row = st.K1 = 0;
      st.b = &k.m_wz; st.nb = 1; st.b_row[0] = u0; st.b_rows = uper; st.n_pad = uper; st.npass = 1; st.pass_row_step = st.pass_tmem_step = 0;
      if (warp == 0 || warp == PB_WARP) produce(st, warp == PB_WARP);
      else if (warp == 1) issue(st);
      else {
        int aidx[1];
        aidx[0] = a.act_idx[gp];                          // (loaded in the shadow of the MMAs)
        wait_acc();
        ln_silu_epilogue(uper, nullptr, a.g_mlp, a.b_mlp, 1e-3f, a.W_mlpaT, a.ld_mlpaT, aidx, nullptr, nullptr, feat16, U, (size_t)lrow, u0, nullptr, nullptr);
      }
      end_stage();
    }
    // ---- S2: GRU: gates = LN(Wgru . [h_{i-1}, feat]) -> h_i   (own units d0.. of the three gates; this thread: dper / 4 of them)
    {
      Stage st;
      st.a0 = &k.m_traj; st.a0_col = S; st.a0_row = prow; st.K0 = D;
      st.a1 = &k.m_feat; st.a1_col = 0; st.a1_row = row_blk; st.K1 = U;
      st.b = &k.m_wgru; st.nb = 3; st.b_row[0] = d0; st.b_row[1] = D + d0; st.b_row[2] = 2 * D + d0; st.b_rows = dpp; st.n_pad = 3 * dpp;
      st.npass = dper / dpp; st.pass_row_step = dpp; st.pass_tmem_step = 3 * dpp;
      if (warp == 0 || warp == PB_WARP) produce(st, warp == PB_WARP);
      else if (warp == 1) issue(st);
      else {
        wait_acc();
        const int dq = dper >> 2, db = cq * dq;            // this thread: units [db, db + dq) of the CTA's dper, 8 at a time
        const int hpitch = dper + 4;
        auto gate_col = [&](int ul, int g) { const int ps = ul / dpp; return ps * 3 * dpp + g * dpp + (ul - ps * dpp); };
        {   // h_{i-1}[:, d0 .. d0 + dper) of the block's rows -> stg1, row-contiguous cp.async in the shadow of the statistics pass
          const int n4 = dper >> 2, total = RM * n4;
          for (int e = etid; e < total; e += NEPI) {
            const int r = e / n4, c4 = e - r * n4;
            const float* src = a.traj + ((size_t)prow + r) * lat + S + d0 + 4 * c4;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(stg1 + r * hpitch + 4 * c4)), "l"(src) : "memory");
          }
        }
        float s = 0.f, q = 0.f;
#pragma unroll 1
        for (int c0 = 0; c0 < dq; c0 += 8) {
          float vr[8], vc[8], vu[8];
          ld8x3(gate_col(db + c0, 0), gate_col(db + c0, 1), gate_col(db + c0, 2), vr, vc, vu);
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            s += vr[j] + vc[j] + vu[j];
            q = fmaf(vr[j], vr[j], fmaf(vc[j], vc[j], fmaf(vu[j], vu[j], q)));
          }
        }
        asm volatile("cp.async.wait_all;" ::: "memory");   // (own copies; the barrier inside ln_exchange publishes everyone's)
        float mu, rs;
        ln_exchange(s, q, 3 * dper, 1e-5f, mu, rs);
#pragma unroll 1
        for (int c0 = 0; c0 < dq; c0 += 8) {
          float vr[8], vc[8], vu[8], hv[8], hn[8];
          ld8x3(gate_col(db + c0, 0), gate_col(db + c0, 1), gate_col(db + c0, 2), vr, vc, vu);
          lds8(stg1 + rl * hpitch + db + c0, hv);
          const int d = d0 + db + c0;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const float r = fsigmoid((vr[j] - mu) * rs * __ldg(a.g_gru + d + j) + __ldg(a.b_gru + d + j));
            const float cp = (vc[j] - mu) * rs * __ldg(a.g_gru + D + d + j) + __ldg(a.b_gru + D + d + j);
            const float cd = ftanh(r * cp);
            const float u = fsigmoid((vu[j] - mu) * rs * __ldg(a.g_gru + 2 * D + d + j) + __ldg(a.b_gru + 2 * D + d + j) - 1.f);
            hn[j] = u * cd + (1.f - u) * hv[j];
          }
          sts8(stg0 + rl * hpitch + db + c0, hn);
        }
        epi_bar();
        write_tile(stg0, hpitch, dper, a.traj, traj16, (size_t)crow, lat, S + d0);
      }
      end_stage();
    }
    // ---- S3: prior hidden = SiLU(LN(Wt0 . h_i + b))
    {
      Stage st;
      st.a0 = &k.m_traj; st.a0_col = S; st.a0_row = crow; st.K0 = D;
      st.a1 = nullptr; st.a1_col = st.a1_row = st.K1 = 0;
      st.b = &k.m_wt0; st.nb = 1; st.b_row[0] = h0; st.b_rows = hper; st.n_pad = hper; st.npass = 1; st.pass_row_step = st.pass_tmem_step = 0;
      if (warp == 0 || warp == PB_WARP) produce(st, warp == PB_WARP);
      else if (warp == 1) issue(st);
      else {
        wait_acc();
        ln_silu_epilogue(hper, a.bt0, a.g_t, a.b_t, 1e-5f, nullptr, 0, nullptr, nullptr, nullptr, thid16, Hh, (size_t)lrow, h0, nullptr, nullptr);
      }
      end_stage();
    }
    // ---- S4: prior logits of the own categoricals -> unimix -> sample z_i (one-hot into the trajectory); this thread: groups cq, cq + 4, ..
    {
      Stage st;
      st.a0 = &k.m_thid; st.a0_col = 0; st.a0_row = row_blk; st.K0 = Hh;
      st.a1 = nullptr; st.a1_col = st.a1_row = st.K1 = 0;
      st.b = &k.m_wt1; st.nb = 1; st.b_row[0] = s0; st.b_rows = sper; st.n_pad = sper; st.npass = 1; st.pass_row_step = st.pass_tmem_step = 0;
      if (warp == 0 || warp == PB_WARP) produce(st, warp == PB_WARP);
      else if (warp == 1) issue(st);
      else {
        // the first group's uniform is fetched in the shadow of the MMAs (one uncoalesced L2 round trip per thread)
        const float uz_first = (!a.f_z && cq < gper) ? a.u_z[((size_t)(i - 1) * N + row_blk + rl) * G + c * gper + cq] : 0.f;
        wait_acc();
#pragma unroll 1
        for (int gl = cq; gl < gper; gl += 4) {
          const int g = c * gper + gl;
          const size_t zi = ((size_t)(i - 1) * N + row_blk + rl) * G + g;
          int idx;
          if (a.f_z) {
            idx = a.f_z[zi];
          } else {
            // the group's Kc <= 32 logits in registers: one TMEM pass, one exponential per class
            float x[32];
            ld32(gl * Kc, Kc, x);
            const float* bp = a.bt1 + s0 + gl * Kc;
            float mx = -INFINITY;
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              if (j < Kc) {
                const float4 b4 = __ldg(reinterpret_cast<const float4*>(bp + j));
                x[j] += b4.x; x[j + 1] += b4.y; x[j + 2] += b4.z; x[j + 3] += b4.w;
                mx = fmaxf(fmaxf(mx, fmaxf(x[j], x[j + 1])), fmaxf(x[j + 2], x[j + 3]));
              }
            }
            float se = 0.f;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              x[j] = (j < Kc) ? __expf(x[j] - mx) : 0.f;
              se += x[j];
            }
            const float mixk = a.unimix / (float)Kc, inv = __fdividef(1.f, se);
            // unimix probabilities sum to 1 up to round-off unless the clamp bites: the draw is against their actual total
            float tot = 0.f;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              const float p = x[j] * inv;
              x[j] = (j < Kc) ? ((a.unimix > 0.f) ? fminf(fmaxf((1.f - a.unimix) * p + mixk, DV3_EPS32), 1.f - DV3_EPS32) : p) : 0.f;
              tot += x[j];
            }
            const float thr = ((gl == cq) ? uz_first : a.u_z[zi]) * tot;         // idx = #{k : cumsum_k(p) / tot <= u}
            float cs = 0.f;
            idx = 0;
#pragma unroll
            for (int j = 0; j < 32; ++j) {
              cs += x[j];
              idx += (j < Kc && cs <= thr) ? 1 : 0;
            }
            idx = min(idx, Kc - 1);
          }
          reinterpret_cast<int*>(stg0)[rl * gper + gl] = idx;
        }
        epi_bar();
        {   // write-out, row-contiguous: the sampled indices and their one-hot rows (fp32 trajectory + bf16 mirror)
          const int* zs = reinterpret_cast<const int*>(stg0);
          for (int e = etid; e < RM * gper; e += NEPI) {
            const int r = e / gper, gl = e - r * gper;
            a.z_idx[((size_t)(i - 1) * N + row_blk + r) * G + c * gper + gl] = zs[e];
          }
          const int n4 = sper >> 2, total = RM * n4;
          for (int e = etid; e < total; e += NEPI) {
            const int r = e / n4, col = (e - r * n4) << 2;
            const int gl = col / Kc, k0 = col - gl * Kc, idx = zs[r * gper + gl];
            const float4 v = make_float4(k0 == idx ? 1.f : 0.f, k0 + 1 == idx ? 1.f : 0.f, k0 + 2 == idx ? 1.f : 0.f, k0 + 3 == idx ? 1.f : 0.f);
            const size_t o = ((size_t)crow + r) * lat + s0 + col;
            *reinterpret_cast<float4*>(a.traj + o) = v;
            __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
            uint2 pk;
            pk.x = *reinterpret_cast<uint32_t*>(&lo);
            pk.y = *reinterpret_cast<uint32_t*>(&hi);
            *reinterpret_cast<uint2*>(traj16 + o) = pk;
          }
        }
      }
      end_stage();
    }
    policy(i);
  }

  __syncthreads();
  if (warp == 1) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(k.tmem_cols));
  }
}

constexpr size_t SMEM_BYTES = NSTG * (A_BYTES + B_BYTES) + MAXCL * RM * 8 + 4 * RM * 8 + (2 * NSTG + 2) * 8 + 16 + 1024;

int pick_cluster(const Dv3RolloutArgs* a) {
  for (int CL = 8; CL <= 16; CL <<= 1) {
    if (a->U % CL || a->D % CL || a->Hh % CL || a->G % CL || a->Ua % CL) continue;
    const int uper = a->U / CL, dper = a->D / CL, hper = a->Hh / CL, sper = (a->G / CL) * a->Kc, aper = a->Ua / CL;
    // MMA N: multiples of 16; a thread owns a quarter of the CTA's columns in 8-column TMEM chunks (GRU: dper / 4 units per thread);
    // a GRU with more than 64 own units runs as passes of 64 (3 x 64 accumulator columns each)
    if ((uper & 15) || (hper & 15) || (aper & 15) || (dper & 31) || (sper & 15) || (a->Kc & 7)) continue;
    if (uper > NMAX || hper > NMAX || aper > NMAX || sper > NMAX) continue;
    if (dper > 64 && ((dper & 63) || dper > 128)) continue;
    return CL;
  }
  return 0;
}

}  // namespace

extern "C" int dv3_rollout_tc(const Dv3RolloutArgs* a, void* stream_) {
  if (!a) return DV3_ERR_ARG;
  if (a->N <= 0 || a->N % RM || a->H < 1 || a->L < 1 || a->L > MAXL || a->nh < 1 || a->nh > 4 || a->Kc > 32 || (a->Kc & 3)) return DV3_ERR_UNSUPPORTED;
  if (a->S != a->G * a->Kc || a->lat != a->S + a->D || (a->S % KB) || (a->D % KB) || (a->U % KB) || (a->Hh % KB) || (a->Ua % KB) || (a->lat % KB))
    return DV3_ERR_UNSUPPORTED;
  if (a->A > 128 || a->nh != 1) return DV3_ERR_UNSUPPORTED;           // one action head of <= 128 logits (they live in one CTA, 32 per column quarter)
  if ((a->ld_mlpaT & 3)) return DV3_ERR_UNSUPPORTED;
  const int CL = pick_cluster(a);
  if (!CL) return DV3_ERR_UNSUPPORTED;
  const int M1 = (a->H + 1) * a->N;
  {   // the multicast loads must read allocations of the launching device only (first and last byte of every activation buffer)
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess) return DV3_ERR_UNSUPPORTED;
    auto local = [&](const void* p, size_t bytes) {
      for (const char* q : {static_cast<const char*>(p), static_cast<const char*>(p) + bytes - 1}) {
        cudaPointerAttributes at;
        if (cudaPointerGetAttributes(&at, q) != cudaSuccess) { cudaGetLastError(); return false; }
        if (at.type != cudaMemoryTypeDevice || at.device != dev) return false;
      }
      return true;
    };
    bool ok = local(a->traj16, (size_t)M1 * a->lat * 2) && local(a->feat16, (size_t)a->N * a->U * 2) && local(a->thid16, (size_t)a->N * a->Hh * 2);
    for (int l = 0; l < a->L && ok; ++l) ok = local(a->aact16[l], (size_t)M1 * a->Ua * 2);
    if (!ok) return DV3_ERR_UNSUPPORTED;
  }
  RollK* k = new RollK;
  k->a = *a;
  k->CL = CL;
  {
    const int dper_ = a->D / CL, need = dper_ > 64 ? 3 * dper_ : 256;
    k->tmem_cols = need > 256 ? 512 : 256;
  }
  k->dbg = getenv("DV3_SCAN_DEBUG") ? reinterpret_cast<long long*>(a->dbg) : nullptr;
  int rc = 0;
  auto mk = [&](CUtensorMap* tm, const void* base, int rows, int cols, int ld, int box_rows) {
    if (!rc) rc = dv3_internal_make_map_2d(tm, base, (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld, (unsigned)box_rows, KB);
  };
  const int uper = a->U / CL, dper = a->D / CL, hper = a->Hh / CL, sper = (a->G / CL) * a->Kc, aper = a->Ua / CL;
  // activation maps: one box = the RM / CL rows a CTA multicasts to its cluster
  mk(&k->m_traj, a->traj16, M1, a->lat, a->lat, RM / CL);
  mk(&k->m_feat, a->feat16, a->N, a->U, a->U, RM / CL);
  mk(&k->m_thid, a->thid16, a->N, a->Hh, a->Hh, RM / CL);
  for (int l = 0; l < a->L; ++l) mk(&k->m_act[l], a->aact16[l], M1, a->Ua, a->Ua, RM / CL);
  mk(&k->m_wz, a->Wz16, a->U, a->S, a->ld_wz, uper);
  mk(&k->m_wgru, a->Wgru16, 3 * a->D, a->D + a->U, a->ld_wgru, dper < 64 ? dper : 64);      // one gate of one pass per box
  mk(&k->m_wt0, a->Wt016, a->Hh, a->D, a->ld_wt0, hper);
  mk(&k->m_wt1, a->Wt116, a->S, a->Hh, a->ld_wt1, sper);
  for (int l = 0; l < a->L; ++l) mk(&k->m_wa[l], a->Wa16[l], a->Ua, l == 0 ? a->lat : a->Ua, a->ld_wa[l], aper);
  mk(&k->m_wh, a->Wh16, a->A, a->Ua, a->ld_wh, a->A);
  if (rc) { delete k; return rc == DV3_ERR_UNSUPPORTED ? rc : DV3_ERR_UNSUPPORTED; }
  auto kern = rollout_tc_kernel;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SMEM_BYTES);
  if (e == cudaSuccess && CL > 8) e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
  if (e != cudaSuccess) { cudaGetLastError(); delete k; return DV3_ERR_UNSUPPORTED; }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((a->N / RM) * CL, 1, 1);
  cfg.blockDim = dim3(NTHR, 1, 1);
  cfg.dynamicSmemBytes = SMEM_BYTES;
  cfg.stream = (cudaStream_t)stream_;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int nmax = 0;
  e = cudaOccupancyMaxActiveClusters(&nmax, kern, &cfg);
  if (e != cudaSuccess || nmax < 1) { cudaGetLastError(); delete k; return DV3_ERR_UNSUPPORTED; }
  if (getenv("DV3_SCAN_DEBUG")) fprintf(stderr, "[dv3 rollout] CL=%d blocks=%d smem %zu KB, max active clusters %d\n", CL, a->N / RM, SMEM_BYTES / 1024, nmax);
  e = cudaLaunchKernelEx(&cfg, kern, *k);
  delete k;
  return e == cudaSuccess ? DV3_OK : (int)e;
}
