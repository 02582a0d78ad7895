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
//     producer warp : TMA (SWIZZLE_128B boxes) of the 128-row activation tile [128 x 64 k] and of this CTA's weight rows [n_own x 64 k]
//                     into an mbarrier ring;
//     MMA warp      : tcgen05.mma M128 x N n_own x K16, fp32 accumulator in TMEM;
//     epilogue warps: thread = row; TMEM -> registers; bias / one-hot action gather; LayerNorm statistics over the CTA's columns,
//                     exchanged with the peers by st.async on the receiver's mbarrier; normalise, SiLU / GRU gates / unimix +
//                     inverse-CDF sampling; results to global memory (fp32 for the losses, bf16 as the next layer's operand).
// The layer outputs of all CTAs of a cluster form the next layer's input: they meet in L2-resident global buffers (the trajectory
// and actor-activation buffers that the behaviour phase needs anyway) and are re-read by TMA after a cluster barrier; generic-proxy
// stores are ordered before the async-proxy reads with fence.proxy.async.  All TMA loads read this GPU's memory only.
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
constexpr int NSTG = 4;
constexpr int A_BYTES = RM * KB * 2;    // 16 KB
constexpr int NMAX = 192;               // most output features per CTA and stage
constexpr int B_BYTES = NMAX * KB * 2;  // 24 KB
constexpr int NTHR = 192;               // warp 0 producer, warp 1 MMA, warps 2-5 epilogue (thread = row)
constexpr int MAXL = DV3_ROLLOUT_MAX_LAYERS;
constexpr int MAXCL = 16;

struct RollK {
  CUtensorMap m_traj, m_feat, m_thid, m_act[MAXL], m_wz, m_wgru, m_wt0, m_wt1, m_wa[MAXL], m_wh;
  Dv3RolloutArgs a;
  int CL;
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
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
                 "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
}
__device__ __forceinline__ float fsigmoid(float x) { return __fdividef(1.f, 1.f + __expf(-x)); }
__device__ __forceinline__ float ftanh(float x) { const float e = __expf(-2.f * fabsf(x)); return copysignf(__fdividef(1.f - e, 1.f + e), x); }
__device__ __forceinline__ void store16_f32(float* dst, const float* v) {
#pragma unroll
  for (int j = 0; j < 16; j += 4) *reinterpret_cast<float4*>(dst + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
}
__device__ __forceinline__ void store16_bf16(__nv_bfloat16* dst, const float* v) {
  uint32_t w[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) { __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]); w[j] = *reinterpret_cast<uint32_t*>(&h); }
  *reinterpret_cast<uint4*>(dst) = make_uint4(w[0], w[1], w[2], w[3]);
  *reinterpret_cast<uint4*>(dst + 8) = make_uint4(w[4], w[5], w[6], w[7]);
}

// One dense layer of the chain for this CTA: which activation tile(s) to stream, which weight rows, how many output columns.
struct Stage {
  const CUtensorMap* a0; int a0_col, a0_row, K0;      // first source of the 128-row operand: columns [a0_col, a0_col + K0) of rows a0_row..
  const CUtensorMap* a1; int a1_col, a1_row, K1;      // optional second source (concatenated along K)
  const CUtensorMap* b; int nb, b_row[3], b_rows;     // nb boxes of b_rows weight rows each (GRU: the three gates)
  int n_pad;                                          // N of the MMA (multiple of 16, >= nb * b_rows)
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
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(lnp + MAXCL * RM);
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
  const int row_blk = blk * RM;                             // first row of the block inside a step's N rows

  for (int i = tid; i < NSTG * (A_BYTES + B_BYTES) / 16; i += NTHR) reinterpret_cast<uint4*>(smem)[i] = make_uint4(0u, 0u, 0u, 0u);
  if (tid == 0) {
    for (int s = 0; s < NSTG; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    mbar_init(acc_bar, 1);
    mbar_init(stat_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(256));
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
  auto produce = [&](const Stage& st) {
    const int nk0 = st.K0 / KB, nk = nk0 + st.K1 / KB;
    const uint32_t bytes = A_BYTES + (uint32_t)st.nb * st.b_rows * KB * 2;
#pragma unroll 1
    for (int kb = 0; kb < nk; ++kb) {
      mbar_wait(&empty_bar[pp.slot], pp.phase ^ 1u);
      if (elect_one()) {
        mbar_expect_tx(&full_bar[pp.slot], bytes);
        const uint32_t da = smem_u32(ring_a + pp.slot * A_BYTES), db = smem_u32(ring_b + pp.slot * B_BYTES);
        if (kb < nk0) tma_load_2d(da, st.a0, &full_bar[pp.slot], st.a0_col + kb * KB, st.a0_row);
        else tma_load_2d(da, st.a1, &full_bar[pp.slot], st.a1_col + (kb - nk0) * KB, st.a1_row);
        for (int j = 0; j < st.nb; ++j) tma_load_2d(db + (uint32_t)j * st.b_rows * KB * 2, st.b, &full_bar[pp.slot], kb * KB, st.b_row[j]);
      }
      __syncwarp();
      if (++pp.slot == NSTG) { pp.slot = 0; pp.phase ^= 1u; }
    }
  };
  auto issue = [&](const Stage& st) {
    const int nk = st.K0 / KB + st.K1 / KB;
    const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(st.n_pad >> 3) << 17) | ((uint32_t)(RM >> 4) << 24);
    const uint64_t ad0 = make_desc_sw128(smem_u32(ring_a)), bd0 = make_desc_sw128(smem_u32(ring_b));
#pragma unroll 1
    for (int kb = 0; kb < nk; ++kb) {
      mbar_wait(&full_bar[pp.slot], pp.phase);
      asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
      const uint64_t ad = ad0 + (uint64_t)((pp.slot * A_BYTES) >> 4), bd = bd0 + (uint64_t)((pp.slot * B_BYTES) >> 4);
#pragma unroll
      for (int kk = 0; kk < KB / 16; ++kk) {
        const uint32_t accum = (kb > 0 || kk > 0) ? 1u : 0u;
        if (elect_one())
          asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_base),
                       "l"(ad + (uint64_t)(kk * 2)), "l"(bd + (uint64_t)(kk * 2)), "r"(idesc), "r"(accum) : "memory");
      }
      if (elect_one()) {
        asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&empty_bar[pp.slot])) : "memory");
        if (kb == nk - 1) asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(acc_bar)) : "memory");
      }
      __syncwarp();
      if (++pp.slot == NSTG) { pp.slot = 0; pp.phase ^= 1u; }
    }
  };
  // epilogue helpers (thread = row rl of the block)
  const int ew = warp & 3, rl = 32 * ew + lane;
  const uint32_t t_lane = tmem_base + ((uint32_t)(32 * ew) << 16);
  // LayerNorm statistics of a row whose columns are spread over the cluster: (mean, M2) of the own n columns go to every peer's
  // lnp[c][rl] (st.async, completes on the peer's barrier), then the CL partials are merged (Chan, equal counts).
  auto ln_exchange = [&](float s, float q, int n, float eps, float& mu, float& rs) {
    const float mu_c = s / (float)n, m2_c = fmaxf(q - s * mu_c, 0.f);
    if (tid == 64) mbar_expect_tx(stat_bar, (uint32_t)CL * RM * 8);
    const uint32_t my = smem_u32(&lnp[c * RM + rl]), bar = smem_u32(stat_bar);
    for (int p = 0; p < CL; ++p) st_async_f2(mapa(my, (uint32_t)p), mu_c, m2_c, mapa(bar, (uint32_t)p));
    mbar_wait(stat_bar, stat_par);
    stat_par ^= 1u;
    float sm = 0.f;
    for (int p = 0; p < CL; ++p) sm += lnp[p * RM + rl].x;
    mu = sm / (float)CL;
    float m2 = 0.f;
    for (int p = 0; p < CL; ++p) { const float2 pc = lnp[p * RM + rl]; const float d = pc.x - mu; m2 += pc.y + (float)n * d * d; }
    rs = rsqrtf(m2 / (float)(n * CL) + eps);
  };
  // dense layer + LayerNorm + SiLU: acc (+ bias) (+ gathered action rows) -> [pre] -> act (fp32) / act16 (bf16) at columns col0 + [0, n)
  auto ln_silu_epilogue = [&](int n, const float* bias, const float* gamma, const float* beta, float eps, const float* add_tab, int add_ld,
                              const int* add_idx, float* pre, float* act, __nv_bfloat16* act16, int ld_out, size_t grow, int col0, float* gmean,
                              float* grstd) {
    float s = 0.f, q = 0.f;
    float v[16];
#pragma unroll 1
    for (int c0 = 0; c0 < n; c0 += 16) {
      tmem_ld16(t_lane + (uint32_t)c0, v);
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        float x = v[j] + (bias ? __ldg(bias + col0 + c0 + j) : 0.f);
        if (add_tab) for (int h = 0; h < a.nh; ++h) x += __ldg(add_tab + (size_t)(a.head_off[h] + add_idx[h]) * add_ld + col0 + c0 + j);
        if (c0 + j < n) { s += x; q = fmaf(x, x, q); }
        v[j] = x;
      }
      if (pre) store16_f32(pre + grow * ld_out + col0 + c0, v);
    }
    float mu, rs;
    ln_exchange(s, q, n, eps, mu, rs);
    if (gmean && c == 0) { gmean[grow] = mu; grstd[grow] = rs; }
#pragma unroll 1
    for (int c0 = 0; c0 < n; c0 += 16) {
      tmem_ld16(t_lane + (uint32_t)c0, v);
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        float x = v[j] + (bias ? __ldg(bias + col0 + c0 + j) : 0.f);
        if (add_tab) for (int h = 0; h < a.nh; ++h) x += __ldg(add_tab + (size_t)(a.head_off[h] + add_idx[h]) * add_ld + col0 + c0 + j);
        x = (x - mu) * rs * __ldg(gamma + col0 + c0 + j) + __ldg(beta + col0 + c0 + j);
        v[j] = x * fsigmoid(x);
      }
      if (act) store16_f32(act + grow * ld_out + col0 + c0, v);
      store16_bf16(act16 + grow * ld_out + col0 + c0, v);
    }
  };
  // unimix + inverse-CDF draw of one categorical whose Kn raw logits sit in TMEM columns [tc0, tc0 + Kn) (+ bias): the sampling
  // probabilities are the unimix probabilities themselves, renormalised (bf16-mode contract, see dv3_unimix_sample)
  auto sample_cat = [&](int tc0, int Kn, const float* bias, float u, int forced, float* raw_out) -> int {
    float v[16];
    float mx = -INFINITY;
    for (int c0 = 0; c0 < Kn; c0 += 16) {
      tmem_ld16(t_lane + (uint32_t)(tc0 + c0), v);
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (c0 + j < Kn) { const float x = v[j] + __ldg(bias + c0 + j); mx = fmaxf(mx, x); if (raw_out) raw_out[c0 + j] = x; }
    }
    if (forced >= 0) return forced;
    float se = 0.f;
    for (int c0 = 0; c0 < Kn; c0 += 16) {
      tmem_ld16(t_lane + (uint32_t)(tc0 + c0), v);
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (c0 + j < Kn) se += __expf(v[j] + __ldg(bias + c0 + j) - mx);
    }
    const float mixk = a.unimix / (float)Kn;
    float tot = 0.f;
    for (int c0 = 0; c0 < Kn; c0 += 16) {
      tmem_ld16(t_lane + (uint32_t)(tc0 + c0), v);
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (c0 + j < Kn) {
          const float p = __fdividef(__expf(v[j] + __ldg(bias + c0 + j) - mx), se);
          tot += (a.unimix > 0.f) ? fminf(fmaxf((1.f - a.unimix) * p + mixk, DV3_EPS32), 1.f - DV3_EPS32) : p;
        }
    }
    const float thr = u * tot;      // idx = #{k : cumsum_k(p) / tot <= u}
    float cs = 0.f;
    int idx = 0;
    for (int c0 = 0; c0 < Kn; c0 += 16) {
      tmem_ld16(t_lane + (uint32_t)(tc0 + c0), v);
#pragma unroll
      for (int j = 0; j < 16; ++j)
        if (c0 + j < Kn) {
          const float p = __fdividef(__expf(v[j] + __ldg(bias + c0 + j) - mx), se);
          cs += (a.unimix > 0.f) ? fminf(fmaxf((1.f - a.unimix) * p + mixk, DV3_EPS32), 1.f - DV3_EPS32) : p;
          idx += (cs <= thr) ? 1 : 0;
        }
    }
    return min(idx, Kn - 1);
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
    asm volatile("fence.proxy.async;" ::: "memory");
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
      st.b = &k.m_wa[l]; st.nb = 1; st.b_row[0] = au0; st.b_rows = aper; st.n_pad = (aper + 15) & ~15;
      if (warp == 0) produce(st);
      else if (warp == 1) issue(st);
      else {
        wait_acc();
        ln_silu_epilogue(aper, nullptr, a.ga[l], a.ba[l], 1e-3f, nullptr, 0, nullptr, a.apre[l], a.aact[l], reinterpret_cast<__nv_bfloat16*>(a.aact16[l]), Ua,
                         grow, au0, a.amean[l], a.arstd[l]);
      }
      end_stage();
    }
    {   // head(s): all action logits of a row live in CTA 0 of the cluster
      Stage st;
      st.a0 = &k.m_act[L - 1]; st.a0_col = 0; st.a0_row = arow; st.K0 = Ua;
      st.a1 = nullptr; st.a1_col = st.a1_row = st.K1 = 0;
      st.b = &k.m_wh; st.nb = 1; st.b_row[0] = 0; st.b_rows = A; st.n_pad = (A + 15) & ~15;
      if (c == 0) {
        if (warp == 0) produce(st);
        else if (warp == 1) issue(st);
        else {
          wait_acc();
          int col = 0;
          for (int h = 0; h < a.nh; ++h) {
            const int dim = a.head_dim[h];
            const float u = (i == 0) ? a.u_act0[(size_t)(row_blk + rl) * a.nh + h] : a.u_act[((size_t)(i - 1) * N + row_blk + rl) * a.nh + h];
            const int forced = a.f_act ? a.f_act[grow * a.nh + h] : -1;
            const int idx = sample_cat(col, dim, a.bh + col, u, forced, a.aout + grow * A + col);
            a.act_idx[grow * a.nh + h] = idx;
            if (a.acts) for (int j = 0; j < dim; ++j) a.acts[grow * A + col + j] = (j == idx) ? 1.f : 0.f;
            col += dim;
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
      st.b = &k.m_wz; st.nb = 1; st.b_row[0] = u0; st.b_rows = uper; st.n_pad = (uper + 15) & ~15;
      if (warp == 0) produce(st);
      else if (warp == 1) issue(st);
      else {
        wait_acc();
        int aidx[4];
        for (int h = 0; h < a.nh; ++h) aidx[h] = a.act_idx[gp * a.nh + h];
        ln_silu_epilogue(uper, nullptr, a.g_mlp, a.b_mlp, 1e-3f, a.W_mlpaT, a.ld_mlpaT, aidx, nullptr, nullptr, feat16, U, (size_t)lrow, u0, nullptr, nullptr);
      }
      end_stage();
    }
    // ---- S2: GRU: gates = LN(Wgru . [h_{i-1}, feat]) -> h_i   (own units d0.. of the three gates)
    {
      Stage st;
      st.a0 = &k.m_traj; st.a0_col = S; st.a0_row = prow; st.K0 = D;
      st.a1 = &k.m_feat; st.a1_col = 0; st.a1_row = row_blk; st.K1 = U;
      st.b = &k.m_wgru; st.nb = 3; st.b_row[0] = d0; st.b_row[1] = D + d0; st.b_row[2] = 2 * D + d0; st.b_rows = dper; st.n_pad = (3 * dper + 15) & ~15;
      if (warp == 0) produce(st);
      else if (warp == 1) issue(st);
      else {
        wait_acc();
        float s = 0.f, q = 0.f, v[16];
#pragma unroll 1
        for (int c0 = 0; c0 < 3 * dper; c0 += 16) {
          tmem_ld16(t_lane + (uint32_t)c0, v);
#pragma unroll
          for (int j = 0; j < 16; ++j) { s += v[j]; q = fmaf(v[j], v[j], q); }
        }
        float mu, rs;
        ln_exchange(s, q, 3 * dper, 1e-5f, mu, rs);
        float vr[16], vc[16], vu[16], hn[16];
#pragma unroll 1
        for (int c0 = 0; c0 < dper; c0 += 16) {
          tmem_ld16(t_lane + (uint32_t)c0, vr);
          tmem_ld16(t_lane + (uint32_t)(dper + c0), vc);
          tmem_ld16(t_lane + (uint32_t)(2 * dper + c0), vu);
          const float* hp = a.traj + gp * lat + S + d0 + c0;
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int d = d0 + c0 + j;
            const float r = fsigmoid((vr[j] - mu) * rs * __ldg(a.g_gru + d) + __ldg(a.b_gru + d));
            const float cp = (vc[j] - mu) * rs * __ldg(a.g_gru + D + d) + __ldg(a.b_gru + D + d);
            const float cd = ftanh(r * cp);
            const float u = fsigmoid((vu[j] - mu) * rs * __ldg(a.g_gru + 2 * D + d) + __ldg(a.b_gru + 2 * D + d) - 1.f);
            hn[j] = u * cd + (1.f - u) * hp[j];
          }
          store16_f32(a.traj + gc * lat + S + d0 + c0, hn);
          store16_bf16(traj16 + gc * lat + S + d0 + c0, hn);
        }
      }
      end_stage();
    }
    // ---- S3: prior hidden = SiLU(LN(Wt0 . h_i + b))
    {
      Stage st;
      st.a0 = &k.m_traj; st.a0_col = S; st.a0_row = crow; st.K0 = D;
      st.a1 = nullptr; st.a1_col = st.a1_row = st.K1 = 0;
      st.b = &k.m_wt0; st.nb = 1; st.b_row[0] = h0; st.b_rows = hper; st.n_pad = (hper + 15) & ~15;
      if (warp == 0) produce(st);
      else if (warp == 1) issue(st);
      else {
        wait_acc();
        ln_silu_epilogue(hper, a.bt0, a.g_t, a.b_t, 1e-5f, nullptr, 0, nullptr, nullptr, nullptr, thid16, Hh, (size_t)lrow, h0, nullptr, nullptr);
      }
      end_stage();
    }
    // ---- S4: prior logits of the own categoricals -> unimix -> sample z_i (one-hot into the trajectory)
    {
      Stage st;
      st.a0 = &k.m_thid; st.a0_col = 0; st.a0_row = row_blk; st.K0 = Hh;
      st.a1 = nullptr; st.a1_col = st.a1_row = st.K1 = 0;
      st.b = &k.m_wt1; st.nb = 1; st.b_row[0] = s0; st.b_rows = sper; st.n_pad = (sper + 15) & ~15;
      if (warp == 0) produce(st);
      else if (warp == 1) issue(st);
      else {
        wait_acc();
#pragma unroll 1
        for (int gl = 0; gl < gper; ++gl) {
          const int g = c * gper + gl;
          const size_t zi = ((size_t)(i - 1) * N + row_blk + rl) * G + g;
          const int idx = sample_cat(gl * Kc, Kc, a.bt1 + s0 + gl * Kc, a.u_z[zi], a.f_z ? a.f_z[zi] : -1, nullptr);
          a.z_idx[zi] = idx;
          float* zo = a.traj + gc * lat + s0 + gl * Kc;
          __nv_bfloat16* zo16 = traj16 + gc * lat + s0 + gl * Kc;
          for (int j = 0; j < Kc; j += 4) {
            *reinterpret_cast<float4*>(zo + j) = make_float4(j == idx ? 1.f : 0.f, j + 1 == idx ? 1.f : 0.f, j + 2 == idx ? 1.f : 0.f, j + 3 == idx ? 1.f : 0.f);
            __nv_bfloat162 h0_ = __floats2bfloat162_rn(j == idx ? 1.f : 0.f, j + 1 == idx ? 1.f : 0.f), h1_ = __floats2bfloat162_rn(j + 2 == idx ? 1.f : 0.f, j + 3 == idx ? 1.f : 0.f);
            *reinterpret_cast<uint2*>(zo16 + j) = make_uint2(*reinterpret_cast<uint32_t*>(&h0_), *reinterpret_cast<uint32_t*>(&h1_));
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
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(256));
  }
}

constexpr size_t SMEM_BYTES = NSTG * (A_BYTES + B_BYTES) + MAXCL * RM * 8 + (2 * NSTG + 2) * 8 + 16 + 1024;

int pick_cluster(const Dv3RolloutArgs* a) {
  for (int CL = 8; CL <= 16; CL <<= 1) {
    if (a->U % CL || a->D % CL || a->Hh % CL || a->G % CL || a->Ua % CL) continue;
    const int uper = a->U / CL, dper = a->D / CL, hper = a->Hh / CL, sper = (a->G / CL) * a->Kc, aper = a->Ua / CL;
    // 16-column TMEM chunks and 16-element vector stores per thread: every per-CTA slice must be a multiple of 16
    if ((uper & 15) || (dper & 15) || (hper & 15) || (sper & 15) || (aper & 15)) continue;
    if (3 * dper > NMAX || uper > NMAX || hper > NMAX || sper > NMAX || aper > NMAX) continue;
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
  if (a->A > NMAX || a->nh != 1) return DV3_ERR_UNSUPPORTED;          // one action head (its logits live in one CTA); several heads: per-op path
  if ((a->ld_mlpaT & 3)) return DV3_ERR_UNSUPPORTED;
  const int CL = pick_cluster(a);
  if (!CL) return DV3_ERR_UNSUPPORTED;
  const int M1 = (a->H + 1) * a->N;
  RollK* k = new RollK;
  k->a = *a;
  k->CL = CL;
  k->dbg = getenv("DV3_SCAN_DEBUG") ? reinterpret_cast<long long*>(a->dbg) : nullptr;
  int rc = 0;
  auto mk = [&](CUtensorMap* tm, const void* base, int rows, int cols, int ld, int box_rows) {
    if (!rc) rc = dv3_internal_make_map_2d(tm, base, (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld, (unsigned)box_rows, KB);
  };
  const int uper = a->U / CL, dper = a->D / CL, hper = a->Hh / CL, sper = (a->G / CL) * a->Kc, aper = a->Ua / CL;
  mk(&k->m_traj, a->traj16, M1, a->lat, a->lat, RM);
  mk(&k->m_feat, a->feat16, a->N, a->U, a->U, RM);
  mk(&k->m_thid, a->thid16, a->N, a->Hh, a->Hh, RM);
  for (int l = 0; l < a->L; ++l) mk(&k->m_act[l], a->aact16[l], M1, a->Ua, a->Ua, RM);
  mk(&k->m_wz, a->Wz16, a->U, a->S, a->ld_wz, uper);
  mk(&k->m_wgru, a->Wgru16, 3 * a->D, a->D + a->U, a->ld_wgru, dper);
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
