// RSSM observe scan (forward) with tcgen05 tensor cores: the whole loop dreamer_v3.py:122-129 over RSSM.dynamic (agent.py:350-388)
// -- RecurrentModel (agent.py:296-309), LayerNormGRUCell (models/models.py:388-395), _representation (agent.py:404-420), unimix +
// straight-through one-hot sampling -- as ONE kernel for all T steps.  Machinery: scan_tc.cuh.
//
// Per step (own = this CTA's 1/CL slice of every feature dimension), contractions in the order the MMA warp issues them:
//   gh(t)  gates  += Wgru[own rows, :D]  . h_{t-1}      issued as soon as h_{t-1} has arrived (overlaps the posterior of step t-1)
//   P1     feat    = SiLU(LN(gather of Wmlp[:, :S]^T rows by the one-hot z_{t-1} + hoisted action projection))      (epilogue only)
//   gf(t)  gates  += Wgru[own rows, D:]  . feat         -> LN over 3D (stats exchange) -> GRU gates -> h_t  (exchange)
//   r0(t)  hid     = Wr0[own rows, :D]   . h_t + (e_t . Wr0[:, D:]^T + b)  -> LN (stats exchange) -> SiLU   (exchange)
//   r1(t)  logits  = Wr1[own rows]       . hid + b      -> unimix, inverse-CDF sample of the own categoricals (index exchange)
// Everything the hand-written backward needs is written to the same fp32 buffers as by the per-op path.
#include "scan_tc.cuh"
#include "dv3_b200.h"
#include <stdio.h>
#include <stdlib.h>

namespace {
using namespace tcs;

enum { SEG_GH = 0, SEG_GF = 1, SEG_R0 = 2, SEG_R1 = 3 };
enum { X_FEAT = 0, X_H = 1, X_HID = 2 };            // activation exchanges (consumed by the MMA warp)
enum { ST_FEAT = 0, ST_GRU = 1, ST_HID = 2 };        // LayerNorm-partial exchanges (consumed by the epilogue)

struct FwdK {
  Dv3ScanFwdArgs a;
  Plan plan;
  const uint8_t* stream;        // packed weights: [CL][step_bytes]
  const float* WmlpzT;          // [S, U] fp32 = Wmlp[:, :S]^T
  const float* aW;              // [T*B, U] fp32 = masked action . Wmlp[:, S:]^T
  long long* dbg;
  int CL, nslots;
  int tab_smem;                  // this CTA's columns of Wmlp[:, :S]^T live in shared memory as bf16 (one-hot gather without L2 latency)
};

#define SDBG(p) do { if (k.dbg && c == 0 && blockIdx.x == 0 && et == 0 && t == 2) k.dbg[p] = clock64(); } while (0)

__global__ void __launch_bounds__(NT, 1) rssm_observe_fwd_tc_kernel(const __grid_constant__ FwdK k) {
  const Dv3ScanFwdArgs& a = k.a;
  const Plan& pl = k.plan;
  extern __shared__ __align__(128) uint8_t smraw[];
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int CL = k.CL;
  uint32_t crank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
  const int c = (int)crank;
  const int b0 = ((int)blockIdx.x / CL) * NB;                 // first batch row of this cluster
  const int Bv = min(NB, a.B - b0);                           // valid batch rows
  const int B = a.B, G = a.G, Kc = a.Kc, D = a.D, U = a.U, Hh = a.Hh;
  const int S = G * Kc;
  const int uper = U / CL, dper = D / CL, hper = Hh / CL, gper = G / CL, sper = gper * Kc;
  const int u0 = c * uper, d0 = c * dper, h0 = c * hper, s0 = c * sper;
  const int npairs = NB * gper;

  // ---- shared memory carve-up
  uint8_t* sp = smraw;
  auto take = [&](size_t bytes) { uint8_t* p = sp; sp += (bytes + 127) & ~(size_t)127; return p; };
  uint8_t* ring_p = take((size_t)k.nslots * pl.slot_bytes + 2048);          // + slack: M = 128 reads past a packed tile's rows
  uint8_t* xh = take((size_t)(D >> 3) * BROW);                              // activation operands [k8][NB rows][8 k] bf16
  // feat_t and the posterior hidden share one buffer: feat_{t+1} cannot arrive before every CTA has sampled z_t (after r1(t) read the
  // hidden), the hidden of step t not before every CTA's r0(t) has completed (issued after gf(t), which read feat_t)
  uint8_t* xf = take((size_t)(max(U, Hh) >> 3) * BROW);
  uint8_t* xhid = xf;
  const int fmax = max(max(uper, dper), hper);
  uint8_t* stage = take((size_t)(fmax >> 3) * BROW);                         // own slice in operand layout (source of the pushes)
  const int rows_max = max(max(3 * dper, uper), max(hper, sper));
  float* outb = reinterpret_cast<float*>(take((size_t)rows_max * OB * 4));   // own pre-activations [row][OB]
  float* outp = reinterpret_cast<float*>(take((size_t)sper * OB * 4));       // own posterior raw logits  } kept until the next step
  float* outl = reinterpret_cast<float*>(take((size_t)sper * OB * 4));       // own unimix logits         } stores them to global
  float* hown = reinterpret_cast<float*>(take((size_t)NB * dper * 4));       // own fp32 slice of h  [m][dl] (double-buffered: the saved
  float* hnew = reinterpret_cast<float*>(take((size_t)NB * dper * 4));       // h_in of a step is stored to global after its gates ran)
  float* lnp = reinterpret_cast<float*>(take((size_t)3 * CL * NB * 8));      // [3 exchanges][CL][NB] (mean, M2)
  int* idxf = reinterpret_cast<int*>(take((size_t)G * NB * 4));              // sampled z indices of all groups [g][m] (peers write)
  int* idxo = reinterpret_cast<int*>(take((size_t)npairs * 4));              // own sampled indices [m][gl]
  int* s_zreset = reinterpret_cast<int*>(take((size_t)G * 4));
  uint32_t* s_rmask = reinterpret_cast<uint32_t*>(take((size_t)a.T * 4));    // bit m: batch row m is reset at step t (or padding)
  float* p_mlp = reinterpret_cast<float*>(take((size_t)2 * uper * 4));
  float* p_gru = reinterpret_cast<float*>(take((size_t)6 * dper * 4));
  float* p_r = reinterpret_cast<float*>(take((size_t)2 * hper * 4));
  float* p_br1 = reinterpret_cast<float*>(take((size_t)sper * 4));
  float* s_aw = reinterpret_cast<float*>(take((size_t)NB * uper * 4));       // cp.async landing zones (next step's inputs)
  float* s_ew = reinterpret_cast<float*>(take((size_t)NB * hper * 4));
  float* s_uu = reinterpret_cast<float*>(take((size_t)npairs * 4));
  __nv_bfloat16* tab = k.tab_smem ? reinterpret_cast<__nv_bfloat16*>(take((size_t)S * uper * 2)) : nullptr;   // [S][uper]
  uint64_t* bars = reinterpret_cast<uint64_t*>(take((size_t)(2 * k.nslots + 16) * 8));
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + k.nslots;
  uint64_t* acc_bar = bars + 2 * k.nslots;        // [3]: gates, r0, r1 accumulators complete
  uint64_t* act_bar = acc_bar + 3;                // [3]: feat, h, hid slices of all peers landed
  uint64_t* stat_bar = act_bar + 3;               // [3]
  uint64_t* idx_bar = stat_bar + 3;               // [1]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(idx_bar + 2);

  // ---- one-time set-up (all threads)
  for (int i = tid; i < ((D + max(U, Hh)) >> 3) * BROW / 16; i += NT) reinterpret_cast<uint4*>(xh)[i] = make_uint4(0u, 0u, 0u, 0u);   // xh | xf contiguous
  for (int i = tid; i < (int)((k.nslots * pl.slot_bytes + 2048) / 16); i += NT) reinterpret_cast<uint4*>(ring_p)[i] = make_uint4(0u, 0u, 0u, 0u);
  for (int i = tid; i < (fmax >> 3) * BROW / 16; i += NT) reinterpret_cast<uint4*>(stage)[i] = make_uint4(0u, 0u, 0u, 0u);
  for (int i = tid; i < NB * dper; i += NT) { hown[i] = 0.f; hnew[i] = 0.f; }
  for (int i = tid; i < G * NB; i += NT) idxf[i] = 0;
  for (int i = tid; i < NB * uper; i += NT) s_aw[i] = 0.f;
  for (int i = tid; i < NB * hper; i += NT) s_ew[i] = 0.f;
  for (int i = tid; i < npairs; i += NT) s_uu[i] = 0.f;
  for (int i = tid; i < G; i += NT) s_zreset[i] = __ldg(a.zreset_idx + i);
  for (int t = tid; t < a.T; t += NT) {
    uint32_t msk = 0;
    for (int m = 0; m < NB; ++m) if (m >= Bv || t == 0 || __ldg(a.first + (size_t)t * B + b0 + m) != 0.f) msk |= 1u << m;
    s_rmask[t] = msk;
  }
  for (int i = tid; i < uper; i += NT) { p_mlp[i] = __ldg(a.g_mlp + u0 + i); p_mlp[uper + i] = __ldg(a.b_mlp + u0 + i); }
  for (int i = tid; i < 3 * dper; i += NT) {
    const int gate = i / dper, dl = i - gate * dper;
    p_gru[i] = __ldg(a.g_gru + gate * D + d0 + dl);
    p_gru[3 * dper + i] = __ldg(a.b_gru + gate * D + d0 + dl);
  }
  for (int i = tid; i < hper; i += NT) { p_r[i] = __ldg(a.g_r + h0 + i); p_r[hper + i] = __ldg(a.b_r + h0 + i); }
  for (int i = tid; i < sper; i += NT) p_br1[i] = __ldg(a.br1 + s0 + i);
  if (tab) {
    for (int i = tid; i < S * (uper >> 1); i += NT) {
      const int srow = i / (uper >> 1), ul = (i - srow * (uper >> 1)) << 1;
      const float2 w = *reinterpret_cast<const float2*>(k.WmlpzT + (size_t)srow * U + u0 + ul);
      *reinterpret_cast<__nv_bfloat162*>(tab + srow * uper + ul) = __floats2bfloat162_rn(w.x, w.y);
    }
  }
  if (tid == 0) {
    for (int s = 0; s < k.nslots; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int i = 0; i < 10; ++i) mbar_init(&acc_bar[i], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)), "r"(128));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  fence_proxy_async();
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  cluster_sync_all();          // every CTA's barriers and buffers exist before any peer writes into them

  Ring rg;
  rg.base = smem_u32(ring_p); rg.slot_bytes = pl.slot_bytes; rg.nslots = k.nslots; rg.full = full_bar; rg.empty = empty_bar;
  const uint8_t* stream_cta = k.stream + (size_t)c * pl.step_bytes;

  if (warp == 0) {
    // ===================================================== producer =====================================================
    RingPos pos = {0u, 0u};
    produce_seg(pl, SEG_GH, stream_cta, rg, pos);
#pragma unroll 1
    for (int t = 0; t < a.T; ++t) {
      produce_seg(pl, SEG_GF, stream_cta, rg, pos);
      produce_seg(pl, SEG_R0, stream_cta, rg, pos);
      if (t + 1 < a.T) produce_seg(pl, SEG_GH, stream_cta, rg, pos);
      produce_seg(pl, SEG_R1, stream_cta, rg, pos);
    }
  } else if (warp == 1) {
    // ===================================================== MMA issuer =====================================================
    RingPos pos = {0u, 0u};
    const uint32_t xh_a = smem_u32(xh), xf_a = smem_u32(xf), xhid_a = smem_u32(xhid);
    const uint32_t feat_bytes = (uint32_t)CL * (uper >> 3) * BROW, h_bytes = (uint32_t)CL * (dper >> 3) * BROW,
                   hid_bytes = (uint32_t)CL * (hper >> 3) * BROW;
    mma_seg(pl, SEG_GH, xh_a, tmem_base, rg, pos);                       // h_{-1} = 0
#pragma unroll 1
    for (int t = 0; t < a.T; ++t) {
      const uint32_t par = (uint32_t)t & 1u;
#define MDBG(p) do { if (k.dbg && c == 0 && blockIdx.x == 0 && lane == 0 && t == 2) k.dbg[p] = clock64(); } while (0)
      MDBG(40);
      if (elect_one()) mbar_expect_tx(&act_bar[X_FEAT], feat_bytes);
      __syncwarp();
      mbar_wait(&act_bar[X_FEAT], par);
      MDBG(41);
      mma_seg(pl, SEG_GF, xf_a, tmem_base, rg, pos);
      if (elect_one()) umma_commit(&acc_bar[0]);
      __syncwarp();
      MDBG(42);
      if (elect_one()) mbar_expect_tx(&act_bar[X_H], h_bytes);
      __syncwarp();
      mbar_wait(&act_bar[X_H], par);
      MDBG(43);
      mma_seg(pl, SEG_R0, xh_a, tmem_base, rg, pos);
      if (elect_one()) umma_commit(&acc_bar[1]);
      __syncwarp();
      MDBG(44);
      if (t + 1 < a.T) {
        // reset rows of step t+1 see h = 0 (agent.py:379-381); the posterior of step t above needed the unmasked h_t
        const uint32_t rm = s_rmask[t + 1] & ((1u << Bv) - 1u);
        if (rm) {
          mbar_wait(&acc_bar[1], par);                                   // r0(t) has finished reading xh
#pragma unroll 1
          for (int m = 0; m < NB; ++m)
            if ((rm >> m) & 1u)
              for (int k8 = lane; k8 < (D >> 3); k8 += 32) *reinterpret_cast<uint4*>(xh + b_off(m, k8 * 8)) = make_uint4(0u, 0u, 0u, 0u);
          fence_proxy_async();
          __syncwarp();
        }
        mma_seg(pl, SEG_GH, xh_a, tmem_base, rg, pos);
      }
      MDBG(45);
      if (elect_one()) mbar_expect_tx(&act_bar[X_HID], hid_bytes);
      __syncwarp();
      mbar_wait(&act_bar[X_HID], par);
      MDBG(46);
      mma_seg(pl, SEG_R1, xhid_a, tmem_base, rg, pos);
      if (elect_one()) umma_commit(&acc_bar[2]);
      __syncwarp();
      MDBG(47);
    }
  } else {
    // ===================================================== epilogue =====================================================
    const int et = tid - 64, ew = et >> 5;
    const uint32_t stage_a = smem_u32(stage);
    const int col_g = pl.seg[SEG_GH].tmem_col, col_r0 = pl.seg[SEG_R0].tmem_col, col_r1 = pl.seg[SEG_R1].tmem_col;
    float* lnp0 = lnp;
    float* lnp1 = lnp + (size_t)CL * NB * 2;
    float* lnp2 = lnp + (size_t)2 * CL * NB * 2;

    auto fetch_aw = [&](int t) {
      for (int e = et; e < Bv * uper; e += NT_EPI) { const int m = e / uper, ul = e - m * uper; cp_async4(s_aw + m * uper + ul, k.aW + ((size_t)t * B + b0 + m) * U + u0 + ul); }
    };
    auto fetch_ew = [&](int t) {
      for (int e = et; e < Bv * hper; e += NT_EPI) { const int m = e / hper, hl = e - m * hper; cp_async4(s_ew + m * hper + hl, a.eW + ((size_t)t * B + b0 + m) * Hh + h0 + hl); }
    };
    auto fetch_uu = [&](int t) {
      for (int e = et; e < Bv * gper; e += NT_EPI) {
        const int m = e / gper, gl = e - m * gper;
        const size_t src = ((size_t)t * B + b0 + m) * G + c * gper + gl;
        if (a.forced_idx) cp_async4(s_uu + m * gper + gl, a.forced_idx + src);
        else cp_async4(s_uu + m * gper + gl, a.u_post + src);
      }
    };
    // global stores of the posterior sampled in the step whose first row is prow0 (outp / outl / idxo live until the next P4)
    auto flush_posterior = [&](size_t prow0) {
#pragma unroll 1
      for (int m = ew; m < Bv; m += EW) {
        for (int sl = lane; sl < sper; sl += 32) {
          const int gl = (int)__umulhi((uint32_t)sl, 0xFFFFFFFFu / (uint32_t)Kc + 1u);
          a.post_raw[(prow0 + m) * S + s0 + sl] = outp[sl * OB + m];
          a.post_logits[(prow0 + m) * S + s0 + sl] = outl[sl * OB + m];
          a.latent[(prow0 + m) * a.ld_lat + s0 + sl] = (idxo[m * gper + gl] == (sl - gl * Kc)) ? 1.f : 0.f;
        }
        if (lane < gper) a.post_idx[(prow0 + m) * G + c * gper + lane] = idxo[m * gper + lane];
      }
    };
    fetch_aw(0);
    fetch_ew(0);
    fetch_uu(0);

    // Thread <-> data mapping of the element-wise work: epilogue warp ew owns batch rows 2 ew and 2 ew + 1, one half-warp each
    // (hl = half, l16 = lane in the half); the 16 lanes stride over the features of the row.  The LayerNorm partials, their
    // cluster-wide combination and the normalisation of a row then live in the registers of one half-warp: no shared-memory round
    // trip and no block barrier between them.  Global stores use one warp per row (128 B segments).
    const int hl = lane >> 4, l16 = lane & 15, mrow = 2 * ew + hl;
    const uint32_t fd_kc = 0xFFFFFFFFu / (uint32_t)Kc + 1u;      // n / Kc = __umulhi(n, fd_kc) for n * Kc < 2^32

    // mean / rstd of batch row mrow over a vector spread over the cluster: own values vals[j * OB + mrow], j < n.
    // Partials (mean, M2) go to lnp[c][m] of every peer (st.async, completes on the peer's barrier); after the wait the CL partials
    // are merged with Chan's formula (equal counts).  Returns the statistics in every lane of the half-warp.
    auto row_stats = [&](const float* vals, int n, float* lnp_x, uint64_t* bar, uint32_t parity, float eps, float* gmean, float* grstd,
                         float& mean_out, float& rstd_out, long long* dbg4) {
      float sx = 0.f;
      for (int j = l16; j < n; j += 16) sx += vals[j * OB + mrow];
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) sx += __shfl_xor_sync(0xffffffffu, sx, o);
      const float mu_c = sx / (float)n;
      float q = 0.f;
      for (int j = l16; j < n; j += 16) { const float d = vals[j * OB + mrow] - mu_c; q = fmaf(d, d, q); }
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
      if (dbg4) dbg4[0] = clock64();
      if (l16 < CL) st_async_f2(mapa(smem_u32(lnp_x + (c * NB + mrow) * 2), (uint32_t)l16), mu_c, q, mapa(smem_u32(bar), (uint32_t)l16));
      if (dbg4) dbg4[1] = clock64();
      mbar_wait(bar, parity);
      if (dbg4) dbg4[2] = clock64();
      const float2 pc = (l16 < CL) ? reinterpret_cast<const float2*>(lnp_x)[l16 * NB + mrow] : make_float2(0.f, 0.f);
      float sm = pc.x;
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) sm += __shfl_xor_sync(0xffffffffu, sm, o);
      const float mu = sm / (float)CL;
      const float dm = pc.x - mu;
      float m2 = (l16 < CL) ? (pc.y + (float)n * dm * dm) : 0.f;
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) m2 += __shfl_xor_sync(0xffffffffu, m2, o);
      const float rs = rsqrtf(m2 / (float)(n * CL) + eps);
      if (c == 0 && l16 == 0 && mrow < Bv) { gmean[mrow] = mu; grstd[mrow] = rs; }
      mean_out = mu; rstd_out = rs;
    };
    // bf16-mode activations: exp / divide through the fast paths (error ~1e-6, far below the bf16 rounding of the contractions)
    auto fsigmoid = [](float x) { return __fdividef(1.f, 1.f + __expf(-x)); };
    auto ftanh = [](float x) { const float e = __expf(-2.f * fabsf(x)); return copysignf(__fdividef(1.f - e, 1.f + e), x); };

#pragma unroll 1
    for (int t = 0; t < a.T; ++t) {
      const uint32_t par = (uint32_t)t & 1u;
      const size_t row0 = (size_t)t * B + b0;
      const uint32_t rmask = s_rmask[t];
      const bool rst = (rmask >> mrow) & 1u;                     // reset rows (agent.py:379-384): z = mode(transition(0)), h = 0
      SDBG(0);
      // ================= P0: arm this step's receive barriers; wait for z_{t-1} =================
      if (et == 0) {
        mbar_expect_tx(&stat_bar[ST_FEAT], (uint32_t)CL * NB * 8);
        mbar_expect_tx(&stat_bar[ST_GRU], (uint32_t)CL * NB * 8);
        mbar_expect_tx(&stat_bar[ST_HID], (uint32_t)CL * NB * 8);
      }
      if (t > 0) mbar_wait(&idx_bar[0], par ^ 1u);                         // z_{t-1} of every group has arrived
      if (et == 0) mbar_expect_tx(&idx_bar[0], (uint32_t)G * NB * 4);      // (armed only once the previous phase has completed)
      cp_async_wait_all();
      epi_bar();                                                           // s_aw / s_ew / s_uu landed (other threads' cp.async)
      SDBG(1);
#define ZIDX(g, m) ((((rmask) >> (m)) & 1u) ? s_zreset[(g)] : idxf[(g) * NB + (m)])
      // ================= P1: feat = SiLU(LN(Wmlp . [z, a])) -- one-hot gather + hoisted action projection =================
      if (tab) {
#pragma unroll 1
        for (int ul = 2 * l16; ul < uper; ul += 32) {             // two adjacent features per lane, G independent 4 B loads
          float a0 = s_aw[mrow * uper + ul], a1 = s_aw[mrow * uper + ul + 1];
          const __nv_bfloat16* tcol = tab + ul;
#pragma unroll 8
          for (int g = 0; g < G; ++g) {
            const int zi = rst ? s_zreset[g] : idxf[g * NB + mrow];
            const float2 w = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(tcol + (g * Kc + zi) * uper));
            a0 += w.x; a1 += w.y;
          }
          outb[ul * OB + mrow] = a0;
          outb[(ul + 1) * OB + mrow] = a1;
        }
      } else {
#pragma unroll 1
        for (int ul = l16; ul < uper; ul += 16) {
          float acc = s_aw[mrow * uper + ul];
          const float* wcol = k.WmlpzT + u0 + ul;
#pragma unroll 1
          for (int g0 = 0; g0 < G; g0 += 16) {                    // 16 independent row gathers in flight
            float v[16];
#pragma unroll
            for (int q = 0; q < 16; ++q) v[q] = (g0 + q < G) ? __ldg(wcol + (size_t)((g0 + q) * Kc + (rst ? s_zreset[g0 + q] : idxf[(g0 + q) * NB + mrow])) * U) : 0.f;
#pragma unroll
            for (int q = 0; q < 16; ++q) acc += v[q];
          }
          outb[ul * OB + mrow] = acc;
        }
      }
      __syncwarp();
      SDBG(2);
      float mu, rs;
      row_stats(outb, uper, lnp0, &stat_bar[ST_FEAT], par, 1e-3f, a.feat_mean + row0, a.feat_rstd + row0, mu, rs, nullptr);
      SDBG(3);
#pragma unroll 1
      for (int ul = 2 * l16; ul < uper; ul += 32) {              // -> one 4 B store into the operand layout
        float f0 = (outb[ul * OB + mrow] - mu) * rs * p_mlp[ul] + p_mlp[uper + ul];
        float f1 = (outb[(ul + 1) * OB + mrow] - mu) * rs * p_mlp[ul + 1] + p_mlp[uper + ul + 1];
        f0 *= fsigmoid(f0); f1 *= fsigmoid(f1);
        if (mrow < Bv) *reinterpret_cast<float2*>(a.x_gru + (row0 + mrow) * a.ld_xgru + D + u0 + ul) = make_float2(f0, f1);
        else { f0 = 0.f; f1 = 0.f; }
        *reinterpret_cast<__nv_bfloat162*>(stage + b_off(mrow, ul)) = __floats2bfloat162_rn(f0, f1);
      }
      fence_proxy_async();
      epi_bar();
      SDBG(4);
      push_act(stage_a, (uint32_t)(uper >> 3) * BROW, smem_u32(xf), &act_bar[X_FEAT], c, CL, ew);
      SDBG(5);
      // while the slices travel and the MMA warp contracts: saved inputs of the step (one-hot z_in of the own categoricals, h_in of
      // the own units), the pre-LN feature, the posterior of the PREVIOUS step, next step's action projection
#pragma unroll 1
      for (int m = ew; m < Bv; m += EW) {
        const bool rm = (rmask >> m) & 1u;
        for (int q = lane; q < (sper >> 2); q += 32) {
          const int sc_ = s0 + 4 * q, g = (int)__umulhi((uint32_t)sc_, fd_kc), kk = sc_ - g * Kc, ch = rm ? s_zreset[g] : idxf[g * NB + m];
          *reinterpret_cast<float4*>(a.x_mlp + (row0 + m) * a.ld_xmlp + sc_) =
              make_float4(ch == kk ? 1.f : 0.f, ch == kk + 1 ? 1.f : 0.f, ch == kk + 2 ? 1.f : 0.f, ch == kk + 3 ? 1.f : 0.f);
        }
        for (int dl = lane; dl < dper; dl += 32) a.x_gru[(row0 + m) * a.ld_xgru + d0 + dl] = rm ? 0.f : hown[m * dper + dl];
        for (int ul = lane; ul < uper; ul += 32) a.feat_pre[(row0 + m) * U + u0 + ul] = outb[ul * OB + m];
      }
      if (t > 0) flush_posterior(row0 - B);
      if (t + 1 < a.T) fetch_aw(t + 1);
      SDBG(6);

      // ================= P2: GRU gates: own rows {d, D+d, 2D+d} of Wgru . [h, feat] =================
      mbar_wait(&acc_bar[0], par);
      tc_fence_after();
      SDBG(7);
      tmem_to_outb(tmem_base, col_g, 3 * dper, pl.seg[SEG_GH].m64, outb, nullptr, 0, ew, lane);          // outb row = gate * dper + dl
      epi_bar();
      SDBG(8);
      row_stats(outb, 3 * dper, lnp1, &stat_bar[ST_GRU], par, 1e-5f, a.gru_mean + row0, a.gru_rstd + row0, mu, rs,
                (k.dbg && c == 0 && blockIdx.x == 0 && et == 0 && t == 2) ? k.dbg + 32 : nullptr);
      SDBG(9);
#pragma unroll 1
      for (int dl = 2 * l16; dl < dper; dl += 32) {
        float hn[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int d = dl + j;
          const float r = fsigmoid((outb[d * OB + mrow] - mu) * rs * p_gru[d] + p_gru[3 * dper + d]);
          const float cp = (outb[(dper + d) * OB + mrow] - mu) * rs * p_gru[dper + d] + p_gru[4 * dper + d];
          const float cd = ftanh(r * cp);
          const float u = fsigmoid((outb[(2 * dper + d) * OB + mrow] - mu) * rs * p_gru[2 * dper + d] + p_gru[5 * dper + d] - 1.f);
          hn[j] = u * cd + (1.f - u) * (rst ? 0.f : hown[mrow * dper + d]);
          if (mrow >= Bv) hn[j] = 0.f;
        }
        if (mrow < Bv) *reinterpret_cast<float2*>(a.x_rep + (row0 + mrow) * a.ld_xrep + d0 + dl) = make_float2(hn[0], hn[1]);
        *reinterpret_cast<__nv_bfloat162*>(stage + b_off(mrow, dl)) = __floats2bfloat162_rn(hn[0], hn[1]);
        hnew[mrow * dper + dl] = hn[0];
        hnew[mrow * dper + dl + 1] = hn[1];
      }
      fence_proxy_async();
      epi_bar();
      SDBG(10);
      push_act(stage_a, (uint32_t)(dper >> 3) * BROW, smem_u32(xh), &act_bar[X_H], c, CL, ew);
      SDBG(11);
#pragma unroll 1
      for (int m = ew; m < Bv; m += EW)                         // saved pre-LN gates (own rows of the three gates)
        for (int j = lane; j < 3 * dper; j += 32) {
          const int gate = (j >= 2 * dper) ? 2 : (j >= dper ? 1 : 0);
          a.gates_pre[(row0 + m) * 3 * D + gate * D + d0 + (j - gate * dper)] = outb[j * OB + m];
        }
      { float* tmp = hown; hown = hnew; hnew = tmp; }          // (the x_gru store of this step, above, read the old buffer before the gates)
      SDBG(12);

      // ================= P3: posterior hidden: own rows of Wr0[:, :D] . h' + (e_t . Wr0[:, D:]^T + b) =================
      mbar_wait(&acc_bar[1], par);
      tc_fence_after();
      SDBG(13);
      tmem_to_outb(tmem_base, col_r0, hper, pl.seg[SEG_R0].m64, outb, s_ew, hper, ew, lane);
      epi_bar();
      SDBG(14);
      row_stats(outb, hper, lnp2, &stat_bar[ST_HID], par, 1e-5f, a.hidr_mean + row0, a.hidr_rstd + row0, mu, rs, nullptr);
      SDBG(15);
#pragma unroll 1
      for (int hl_ = 2 * l16; hl_ < hper; hl_ += 32) {
        float f0 = (outb[hl_ * OB + mrow] - mu) * rs * p_r[hl_] + p_r[hper + hl_];
        float f1 = (outb[(hl_ + 1) * OB + mrow] - mu) * rs * p_r[hl_ + 1] + p_r[hper + hl_ + 1];
        f0 *= fsigmoid(f0); f1 *= fsigmoid(f1);
        if (mrow < Bv) *reinterpret_cast<float2*>(a.hidr + (row0 + mrow) * Hh + h0 + hl_) = make_float2(f0, f1);
        else { f0 = 0.f; f1 = 0.f; }
        *reinterpret_cast<__nv_bfloat162*>(stage + b_off(mrow, hl_)) = __floats2bfloat162_rn(f0, f1);
      }
      fence_proxy_async();
      epi_bar();
      SDBG(16);
      push_act(stage_a, (uint32_t)(hper >> 3) * BROW, smem_u32(xhid), &act_bar[X_HID], c, CL, ew);
      SDBG(17);
#pragma unroll 1
      for (int m = ew; m < Bv; m += EW)
        for (int j = lane; j < hper; j += 32) a.hidr_pre[(row0 + m) * Hh + h0 + j] = outb[j * OB + m];
      if (t + 1 < a.T) fetch_ew(t + 1);
      SDBG(18);

      // ================= P4: posterior logits of the own categoricals, unimix, inverse-CDF sample =================
      mbar_wait(&acc_bar[2], par);
      tc_fence_after();
      SDBG(19);
      tmem_to_outb(tmem_base, col_r1, sper, pl.seg[SEG_R1].m64, outp, nullptr, 0, ew, lane);
      epi_bar();
      SDBG(20);
      // One warp per (batch row, own categorical), lane = class (Kc <= 32); the warp's pairs two at a time so that their shuffle
      // chains overlap.  Sampling probabilities: the unimix probabilities themselves (they sum to 1 up to fp32 round-off; the
      // reference renormalises log(p) twice, a difference far below the bf16 error of the logits) and a warp prefix sum instead of
      // the sequential cumsum of the fp32 kernels.
      const float um = a.unimix, um_k = um / (float)Kc;
#pragma unroll 1
      for (int pb = ew; pb < npairs; pb += 2 * EW) {
        float raw[2], e1[2], lu[2], prob[2], mx[2], se[2], cs[2];
        int pm[2], pg[2];
        const bool on = lane < Kc;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int p = min(pb + j * EW, npairs - 1);
          pg[j] = p % gper; pm[j] = p / gper;
          raw[j] = on ? (outp[(size_t)(pg[j] * Kc + lane) * OB + pm[j]] + p_br1[pg[j] * Kc + lane]) : -INFINITY;
          mx[j] = raw[j];
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
          for (int j = 0; j < 2; ++j) mx[j] = fmaxf(mx[j], __shfl_xor_sync(0xffffffffu, mx[j], o));
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) { e1[j] = on ? __expf(raw[j] - mx[j]) : 0.f; se[j] = e1[j]; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
          for (int j = 0; j < 2; ++j) se[j] += __shfl_xor_sync(0xffffffffu, se[j], o);
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          lu[j] = raw[j];
          prob[j] = __fdividef(e1[j], se[j]);
          if (um > 0.f) {
            const float pr = fminf(fmaxf((1.f - um) * prob[j] + um_k, DV3_EPS32), 1.f - DV3_EPS32);
            lu[j] = on ? __logf(pr) : -INFINITY;
            prob[j] = on ? pr : 0.f;
          }
          cs[j] = prob[j];
        }
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
#pragma unroll
          for (int j = 0; j < 2; ++j) { const float v = __shfl_up_sync(0xffffffffu, cs[j], o); if (lane >= o) cs[j] += v; }
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int p = pb + j * EW;
          int choice;
          if (a.forced_idx) {
            choice = __float_as_int(s_uu[min(p, npairs - 1)]);
          } else {                                           // idx = #{k : cumsum_k(p) <= u}
            choice = __popc(__ballot_sync(0xffffffffu, on && cs[j] <= s_uu[min(p, npairs - 1)]));
            if (choice > Kc - 1) choice = Kc - 1;
          }
          if (p < npairs) {
            if (on) {
              outp[(size_t)(pg[j] * Kc + lane) * OB + pm[j]] = raw[j];
              outl[(size_t)(pg[j] * Kc + lane) * OB + pm[j]] = lu[j];
            }
            if (lane == 0) idxo[p] = choice;
            if (lane < CL)
              st_async_b32(mapa(smem_u32(idxf + (c * gper + pg[j]) * NB + pm[j]), (uint32_t)lane), (uint32_t)choice, mapa(smem_u32(&idx_bar[0]), (uint32_t)lane));
          }
        }
      }
      SDBG(21);
      if (t + 1 < a.T) fetch_uu(t + 1);
      epi_bar();                                               // idxo / outp / outl complete before the next step reads / flushes them
      SDBG(22);
    }
    flush_posterior((size_t)(a.T - 1) * B + b0);
    tc_fence_before();
  }

  // ---- teardown: nobody may still be writing into a peer that has left
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(128));
  }
}

// ---- host ------------------------------------------------------------------------------------------------------------------------
bool build_plan(int D, int U, int Hh, int G, int Kc, int CL, int slot_max, Plan* pl) {
  if (CL < 1 || D % CL || U % CL || Hh % CL || G % CL) return false;
  const int uper = U / CL, dper = D / CL, hper = Hh / CL, gper = G / CL, sper = gper * Kc;
  if ((uper & 7) || (dper & 7) || (hper & 7) || (sper & 3) || Kc > 32 || (D & 15) || (U & 15) || (Hh & 15)) return false;
  if (hper > 128 || sper > 128 || 3 * dper > 128 * 6) return false;
  memset(pl, 0, sizeof(Plan));
  uint32_t off16 = 0;
  const int gt = (3 * dper + 127) / 128;                      // gate tiles: TMEM columns [0, gt * NB)
  RowBlk gb[3] = {{0, dper}, {D, dper}, {2 * D, dper}};
  RowBlk hb[1] = {{0, hper}};
  RowBlk sb[1] = {{0, sper}};
  if (plan_add_seg(*pl, slot_max, 3 * dper, D, 0, 1, 0, 3, gb, 0, off16) != SEG_GH) return false;
  if (plan_add_seg(*pl, slot_max, 3 * dper, U, 0, 0, 0, 3, gb, D, off16) != SEG_GF) return false;
  if (plan_add_seg(*pl, slot_max, hper, D, gt * NB, 1, 0, 1, hb, 0, off16) != SEG_R0) return false;
  if (plan_add_seg(*pl, slot_max, sper, Hh, gt * NB + NB, 1, 0, 1, sb, 0, off16) != SEG_R1) return false;
  pl->seg[SEG_GF].m64 = pl->seg[SEG_GH].m64;
  if (gt * NB + 2 * NB > 128) return false;
  pl->step_bytes = off16 * 16;
  return true;
}

constexpr int T_MAX = 1024;      // s_rmask is sized for the launch's T; the plan selection reserves room for this many steps

size_t fixed_smem(int D, int U, int Hh, int G, int Kc, int CL, int T) {
  const int uper = U / CL, dper = D / CL, hper = Hh / CL, gper = G / CL, sper = gper * Kc;
  auto r = [](size_t b) { return (b + 127) & ~(size_t)127; };
  auto mx = [](int x, int y) { return x > y ? x : y; };
  const int fmax = mx(mx(uper, dper), hper), rows_max = mx(mx(3 * dper, uper), mx(hper, sper)), npairs = NB * gper;
  return r((size_t)(D >> 3) * BROW) + r((size_t)(mx(U, Hh) >> 3) * BROW) + r((size_t)(fmax >> 3) * BROW) +
         r((size_t)rows_max * OB * 4) + 2 * r((size_t)sper * OB * 4) + 2 * r((size_t)NB * dper * 4) + r((size_t)3 * CL * NB * 8) + r(2 * NB * 4) +
         r((size_t)G * NB * 4) + r((size_t)npairs * 4) + r((size_t)G * 4) + r((size_t)T * 4) + r((size_t)2 * uper * 4) +
         r((size_t)6 * dper * 4) + r((size_t)2 * hper * 4) + r((size_t)sper * 4) + r((size_t)NB * uper * 4) + r((size_t)NB * hper * 4) +
         r((size_t)npairs * 4) + 2048 + 128 /* ring slack */ + 128 /* alignment of the dynamic segment */;
}

struct Choice { int CL, nslots, tab_smem; size_t smem; };
// The one place that decides cluster size, ring slot size / depth and whether the gather table is resident: the stream size query,
// the packing kernel and the launch must all see the same plan.  Preference: 16 CTAs over 8; the largest slots that still leave a
// ring of >= 4 of them (then >= 3); the resident table if at least 64 KB of ring remain next to it.
bool select_plan(int D, int U, int Hh, int G, int Kc, int max_smem, Plan* pl, Choice* ch) {
  const bool table_ok = !getenv("DV3_SCAN_NO_TABLE");
  for (int CL = 16; CL >= 8; CL >>= 1) {
    const size_t tab_bytes = ((size_t)G * Kc * (U / CL) * 2 + 127) & ~(size_t)127;
    for (int tab = table_ok ? 1 : 0; tab >= 0; --tab) {
      for (int want = 4; want >= (tab ? 4 : 2); --want) {
        for (int slot_max = 32768; slot_max >= 4096; slot_max >>= 1) {
          if (!build_plan(D, U, Hh, G, Kc, CL, slot_max, pl)) continue;
          const size_t fixed = fixed_smem(D, U, Hh, G, Kc, CL, T_MAX);
          const long long room = (long long)max_smem - (long long)fixed - 512 - (tab ? (long long)tab_bytes : 0);
          if (room < (tab ? 49152 : 0)) break;
          int nslots = (int)(room / (long long)pl->slot_bytes);
          if (nslots < want) continue;
          if (nslots > 12) nslots = 12;
          ch->CL = CL; ch->nslots = nslots; ch->tab_smem = tab;
          ch->smem = fixed + (tab ? tab_bytes : 0) + (size_t)nslots * pl->slot_bytes + (((size_t)(2 * nslots + 16) * 8 + 127) & ~(size_t)127);
          return true;
        }
      }
    }
  }
  return false;
}
int device_max_smem() {
  int dev = 0, max_smem = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  return max_smem;
}

int launch_tc(FwdK& k, int nclusters, size_t smem_bytes, cudaStream_t stream) {
  auto kern = rssm_observe_fwd_tc_kernel;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
  if (e != cudaSuccess) { cudaGetLastError(); return DV3_ERR_UNSUPPORTED; }
  if (k.CL > 8) {
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess) { cudaGetLastError(); return DV3_ERR_UNSUPPORTED; }
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(k.CL * nclusters, 1, 1);
  cfg.blockDim = dim3(NT, 1, 1);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = k.CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int nmax = 0;
  e = cudaOccupancyMaxActiveClusters(&nmax, kern, &cfg);
  if (e != cudaSuccess || nmax < 1) { cudaGetLastError(); return DV3_ERR_UNSUPPORTED; }   // clusters are independent: they need not be co-resident
  e = cudaLaunchKernelEx(&cfg, kern, k);
  if (e != cudaSuccess) return (int)e;
  return DV3_OK;
}

}  // namespace

// Plan of the forward scan's weight stream for a shape (the same function sizes the buffer, drives the packing kernel and the scan).
extern "C" int dv3_scan_tc_fwd_config(int D, int U, int Hh, int G, int Kc, int* cluster_size, size_t* stream_bytes) {
  Plan pl;
  Choice ch;
  if (!select_plan(D, U, Hh, G, Kc, device_max_smem(), &pl, &ch)) return DV3_ERR_UNSUPPORTED;
  if (cluster_size) *cluster_size = ch.CL;
  if (stream_bytes) *stream_bytes = (size_t)ch.CL * pl.step_bytes;
  return DV3_OK;
}

// ---- packing: fp32 weights -> per-CTA, per-step bf16 stream in the MMA's shared-memory image -----------------------------------
namespace {
struct PackSrc { const float* W; long long sr, sc; };
struct PackK { Plan plan; PackSrc src[MAX_SEG]; uint8_t* dst; int CL; };
// one block per (cta, chunk); unit = 8 consecutive k of one row (16 B in the stream)
__global__ void __launch_bounds__(256) scan_tc_pack_kernel(const __grid_constant__ PackK p) {
  const int cta = blockIdx.x / p.plan.nchunks, ci = blockIdx.x - cta * p.plan.nchunks;
  const Chunk ch = p.plan.c[ci];
  const Seg& sg = p.plan.seg[ch.seg];
  const PackSrc src = p.src[ch.seg];
  const int rp = ch.lbo16, nunits = (int)ch.bytes16;
  uint4* dst = reinterpret_cast<uint4*>(p.dst + (size_t)cta * p.plan.step_bytes) + ch.off16;
  for (int u = threadIdx.x; u < nunits; u += blockDim.x) {
    const int k8 = u / rp, r = u - k8 * rp;
    const int j = ch.tile * sg.tile_rows + r;                         // local row of the segment
    uint4 out = make_uint4(0u, 0u, 0u, 0u);
    if (j < sg.rows) {
      int start = 0, grow = -1;
      for (int b = 0; b < sg.nblk; ++b) {
        if (j < start + sg.blk[b].size) { grow = sg.blk[b].base + cta * sg.blk[b].size + (j - start); break; }
        start += sg.blk[b].size;
      }
      const int kk = ch.k0 + k8 * 8;
      const float* s = src.W + (size_t)grow * src.sr + (size_t)(sg.col0 + kk) * src.sc;
      float v[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) v[q] = (kk + q < sg.K) ? __ldg(s + (size_t)q * src.sc) : 0.f;
      __nv_bfloat162 h0 = __floats2bfloat162_rn(v[0], v[1]), h1 = __floats2bfloat162_rn(v[2], v[3]), h2 = __floats2bfloat162_rn(v[4], v[5]),
                     h3 = __floats2bfloat162_rn(v[6], v[7]);
      out.x = *reinterpret_cast<uint32_t*>(&h0); out.y = *reinterpret_cast<uint32_t*>(&h1);
      out.z = *reinterpret_cast<uint32_t*>(&h2); out.w = *reinterpret_cast<uint32_t*>(&h3);
    }
    // stream order inside a chunk: [k8][row group][8 rows] x 16 B  ==  unit index u
    dst[u] = out;
  }
}
}  // namespace

extern "C" int dv3_scan_tc_fwd_pack(int D, int U, int Hh, int G, int Kc, const float* Wgru, int ld_wgru, const float* Wr0, int ld_wr0,
                                    const float* Wr1, int ld_wr1, void* dst, void* stream) {
  if (!Wgru || !Wr0 || !Wr1 || !dst) return DV3_ERR_ARG;
  PackK p;
  Choice ch;
  if (!select_plan(D, U, Hh, G, Kc, device_max_smem(), &p.plan, &ch)) return DV3_ERR_UNSUPPORTED;
  const int CL = ch.CL;
  p.CL = CL;
  p.dst = reinterpret_cast<uint8_t*>(dst);
  p.src[SEG_GH] = {Wgru, ld_wgru, 1};
  p.src[SEG_GF] = {Wgru, ld_wgru, 1};
  p.src[SEG_R0] = {Wr0, ld_wr0, 1};
  p.src[SEG_R1] = {Wr1, ld_wr1, 1};
  scan_tc_pack_kernel<<<CL * p.plan.nchunks, 256, 0, (cudaStream_t)stream>>>(p);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}

extern "C" int dv3_rssm_observe_fwd_tc(const Dv3ScanFwdArgs* h, const Dv3ScanTcExtra* x, void* stream_) {
  if (!h || !x) return DV3_ERR_ARG;
  if (h->B < 1 || h->T < 1 || h->G < 1 || h->Kc < 1 || h->Kc > 32 || (h->Kc & 3) || (h->ld_xmlp & 3) || ((uintptr_t)h->x_mlp & 15) ||
      (h->ld_xgru & 1) || (h->ld_xrep & 1) || ((uintptr_t)h->x_gru & 7) || ((uintptr_t)h->x_rep & 7) || ((uintptr_t)h->hidr & 7) || (h->D & 1))
    return DV3_ERR_UNSUPPORTED;
  if (!x->stream || !x->WmlpzT || !x->aW || ((uintptr_t)x->stream & 15)) return DV3_ERR_UNSUPPORTED;
  if (h->T > T_MAX) return DV3_ERR_UNSUPPORTED;
  const int max_smem = device_max_smem();
  FwdK* k = new FwdK;
  Choice ch;
  int rc = DV3_ERR_UNSUPPORTED;
  if (select_plan(h->D, h->U, h->Hh, h->G, h->Kc, max_smem, &k->plan, &ch)) {
    k->a = *h;
    k->stream = reinterpret_cast<const uint8_t*>(x->stream);
    k->WmlpzT = x->WmlpzT;
    k->aW = x->aW;
    k->CL = ch.CL;
    k->nslots = ch.nslots;
    k->tab_smem = (ch.tab_smem && !((uintptr_t)x->WmlpzT & 7) && !(h->U & 1)) ? 1 : 0;
    k->dbg = (getenv("DV3_SCAN_DEBUG") && h->scratch && h->scratch_bytes >= 64 * sizeof(long long)) ? reinterpret_cast<long long*>(h->scratch) : nullptr;
    const int nclusters = (h->B + NB - 1) / NB;
    if (getenv("DV3_SCAN_DEBUG"))
      fprintf(stderr, "[dv3 tc scan] CL=%d clusters=%d slots=%d x %u B, table %d, step stream %u B/CTA, %d chunks, smem %zu KB (max %d KB)\n", ch.CL,
              nclusters, ch.nslots, k->plan.slot_bytes, k->tab_smem, k->plan.step_bytes, k->plan.nchunks, ch.smem / 1024, max_smem / 1024);
    rc = launch_tc(*k, nclusters, ch.smem, (cudaStream_t)stream_);
  }
  delete k;
  return rc;
}
