// Reverse scan of the RSSM observe loop (backward through time of RSSM.dynamic) with tcgen05 tensor cores.  Machinery: scan_tc.cuh.
//
// Reference: what loss.backward() does for the T sequential steps of dreamer_v3.py:122-129 -- the hand-written reverse of
// RSSM.dynamic (agent.py:350-388): straight-through one-hot + unimix categorical (agent.py:404-420, distribution.py), representation
// MLP, LayerNormGRUCell (models/models.py:388-395), RecurrentModel.mlp (agent.py:296-309).  Only dL/dh_t and dL/dz_t propagate
// sequentially; all parameter gradients are K = T*B contractions done after the scan from the tensors saved here.
//
// Per step t = T-1 .. 0 (own = this CTA's 1/CL slice of the feature dimension; every contraction multiplies this CTA's rows of a
// TRANSPOSED weight matrix, streamed through the bulk-copy ring, with a gradient vector that all CTAs assembled by pushing their
// slices into each other's operand buffers):
//   B1  dz_t (own categoricals) = head gradient + carry  ->  d post_raw  (unimix + straight-through Jacobian; the softmax
//       quantities come from dv3_unimix_bwd_prep, computed for all T*B rows before the scan)                      -> exchange
//   B2  d hid  = Wr1^T[own Hh rows]       . d post_raw;   LayerNorm+SiLU backward (row sums over the cluster) -> d hidr_pre  -> exchange
//   B3  dh_t  += Wr0[:, :D]^T[own D rows] . d hidr_pre;   GRU gate backward + LayerNorm backward over 3D      -> d gates_pre -> exchange
//   B4  [dh_{t-1} part | d feat] = Wgru^T[own rows] . d gates_pre;  LayerNorm+SiLU backward                   -> d feat_pre  -> exchange
//   B5  dz_{t-1} = Wmlp[:, :S]^T[own rows] . d feat_pre;  carry = (1 - first_t) * (dz_{t-1}, dh_{t-1} part + dh_t (1 - u))
// Saved-activation inputs of step t-1 are fetched with cp.async while step t runs.
#include "scan_tc.cuh"
#include "dv3_b200.h"
#include <stdio.h>
#include <stdlib.h>

namespace {
using namespace tcs;

enum { SEG_R1T = 0, SEG_R0T = 1, SEG_GRUT = 2, SEG_MLPT = 3 };
enum { X_PR = 0, X_HP = 1, X_GP = 2, X_FP = 3 };      // gradient-vector exchanges (consumed by the MMA warp)
enum { SM_HID = 0, SM_GRU = 1, SM_FEAT = 2 };          // LayerNorm-backward row-sum exchanges (consumed by the epilogue)

struct BwdK {
  Dv3ScanBwdArgs a;
  Plan plan;
  const uint8_t* stream;        // packed transposed weights: [CL][step_bytes]
  int CL, nslots;
  int nb;                       // batch rows per cluster (N of the MMA): 16, or 8 when the [nb, 3D] gradient operand would not fit
};

#define BDBG(p) do { if (a.dbg && c == 0 && blockIdx.x == 0 && et == 0 && t == a.T - 3) a.dbg[p] = clock64(); } while (0)

__global__ void __launch_bounds__(NT, 1) rssm_observe_bwd_tc_kernel(const __grid_constant__ BwdK k) {
  const Dv3ScanBwdArgs& a = k.a;
  const Plan& pl = k.plan;
  extern __shared__ __align__(128) uint8_t smraw[];
  const int tid = threadIdx.x, lane = tid & 31;
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  const int CL = k.CL;
  uint32_t crank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(crank));
  const int c = (int)crank;
  const int nb = k.nb, ob = nb + 1, brow = nb * 16;          // batch tile of this launch (the namespace constants NB / OB / BROW are the 16-row case)
  const int b0 = ((int)blockIdx.x / CL) * nb;
  const int Bv = min(nb, a.B - b0);
  const int B = a.B, G = a.G, Kc = a.Kc, D = a.D, U = a.U, Hh = a.Hh;
  const int S = G * Kc, D3 = 3 * D, lat = a.ld_dlat;
  const int uper = U / CL, dper = D / CL, hper = Hh / CL, gper = G / CL, sper = gper * Kc;
  const int u0 = c * uper, d0 = c * dper, h0 = c * hper, s0 = c * sper;
  typedef __nv_bfloat16 bf;
  bf* m_pr = reinterpret_cast<bf*>(a.dpost_raw_bf16);
  bf* m_hp = reinterpret_cast<bf*>(a.dhidr_pre_bf16);
  bf* m_gp = reinterpret_cast<bf*>(a.dgates_pre_bf16);
  bf* m_fp = reinterpret_cast<bf*>(a.dfeat_pre_bf16);

  // ---- shared memory carve-up
  uint8_t* sp = smraw;
  auto take = [&](size_t bytes) { uint8_t* p = sp; sp += (bytes + 127) & ~(size_t)127; return p; };
  auto takef = [&](size_t n) { return reinterpret_cast<float*>(take(n * 4)); };
  uint8_t* ring_p = take((size_t)k.nslots * pl.slot_bytes + 2048);
  // operand buffers (gradient vectors of all CL slices, bf16, MMA layout).  d post_raw / d gates_pre share one buffer and
  // d hidr_pre / d feat_pre the other: a slice of the later vector can only arrive after every CTA's contraction with the earlier
  // one has completed (its result is needed, cluster-wide, before the later vector exists).
  uint8_t* xa = take((size_t)(max(S, D3) >> 3) * brow);
  uint8_t* xb = take((size_t)(max(Hh, U) >> 3) * brow);
  const int fmax = max(max(sper, hper), max(3 * dper, uper));
  uint8_t* stage = take((size_t)(fmax >> 3) * brow);
  const int rows_max = max(max(hper, 3 * dper), max(dper + uper, sper));
  float* outb = takef((size_t)rows_max * ob);          // accumulator [row][m]
  float* wg = takef((size_t)rows_max * ob);            // gg = g * gamma  [row][m]
  float* wx = takef((size_t)rows_max * ob);            // xhat            [row][m]
  float* lnp = takef((size_t)3 * CL * nb * 2);
  float* dzc = takef((size_t)nb * sper);               // carried dL/dz (own categoricals) [m][sl]
  float* dhc = takef((size_t)nb * dper);               // carried dL/dh (own units) [m][dl]
  float* dhd = takef((size_t)nb * dper);               // dh_t * (1 - u): direct path to h_{t-1}
  float* dhp = takef((size_t)nb * dper);               // h_{t-1} part of Wgru^T . d gates_pre
  float* f_ps = takef((size_t)nb * sper);              // cp.async landing zones (saved forward tensors of the next step to process)
  float* f_cf = takef((size_t)nb * sper);
  float* f_pn = takef((size_t)nb * sper);
  float* f_dl = takef((size_t)nb * sper);
  float* f_dz = takef((size_t)nb * sper);
  float* f_hp = takef((size_t)nb * hper);
  float* f_dh = takef((size_t)nb * dper);
  float* f_hi = takef((size_t)nb * dper);
  float* f_gp = takef((size_t)3 * nb * dper);          // [gate][m][dl]
  float* f_fp = takef((size_t)nb * uper);
  float* f_st = takef(6 * nb);                         // hidr mean|rstd, gru mean|rstd, feat mean|rstd
  uint32_t* s_rmask = reinterpret_cast<uint32_t*>(take((size_t)a.T * 4));
  float* p_r = takef(2 * hper);
  float* p_gru = takef(6 * dper);
  float* p_mlp = takef(2 * uper);
  uint64_t* bars = reinterpret_cast<uint64_t*>(take((size_t)(2 * k.nslots + 16) * 8));
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + k.nslots;
  uint64_t* acc_bar = bars + 2 * k.nslots;        // [4]
  uint64_t* act_bar = acc_bar + 4;                // [4]
  uint64_t* sum_bar = act_bar + 4;                // [3]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(sum_bar + 4);

  // ---- one-time set-up
  for (int i = tid; i < ((max(S, D3) + max(Hh, U)) >> 3) * brow / 16; i += NT) reinterpret_cast<uint4*>(xa)[i] = make_uint4(0u, 0u, 0u, 0u);   // xa | xb contiguous
  for (int i = tid; i < (int)((k.nslots * pl.slot_bytes + 2048) / 16); i += NT) reinterpret_cast<uint4*>(ring_p)[i] = make_uint4(0u, 0u, 0u, 0u);
  for (int i = tid; i < (fmax >> 3) * brow / 16; i += NT) reinterpret_cast<uint4*>(stage)[i] = make_uint4(0u, 0u, 0u, 0u);
  for (int i = tid; i < nb * sper; i += NT) { dzc[i] = 0.f; f_ps[i] = f_cf[i] = f_pn[i] = f_dl[i] = f_dz[i] = 0.f; }
  for (int i = tid; i < nb * dper; i += NT) { dhc[i] = dhd[i] = dhp[i] = 0.f; f_dh[i] = f_hi[i] = 0.f; }
  for (int i = tid; i < 3 * nb * dper; i += NT) f_gp[i] = 0.f;
  for (int i = tid; i < nb * hper; i += NT) f_hp[i] = 0.f;
  for (int i = tid; i < nb * uper; i += NT) f_fp[i] = 0.f;
  for (int i = tid; i < 6 * nb; i += NT) f_st[i] = (i / nb) & 1 ? 1.f : 0.f;
  for (int t = tid; t < a.T; t += NT) {
    uint32_t msk = 0;
    for (int m = 0; m < nb; ++m) if (m >= Bv || t == 0 || __ldg(a.first + (size_t)t * B + b0 + m) != 0.f) msk |= 1u << m;
    s_rmask[t] = msk;
  }
  for (int i = tid; i < hper; i += NT) { p_r[i] = __ldg(a.g_r + h0 + i); p_r[hper + i] = __ldg(a.b_r + h0 + i); }
  for (int i = tid; i < 3 * dper; i += NT) {
    const int gate = i / dper, dl = i - gate * dper;
    p_gru[i] = __ldg(a.g_gru + gate * D + d0 + dl);
    p_gru[3 * dper + i] = __ldg(a.b_gru + gate * D + d0 + dl);
  }
  for (int i = tid; i < uper; i += NT) { p_mlp[i] = __ldg(a.g_mlp + u0 + i); p_mlp[uper + i] = __ldg(a.b_mlp + u0 + i); }
  if (tid == 0) {
    for (int s = 0; s < k.nslots; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int i = 0; i < 11; ++i) mbar_init(&acc_bar[i], 1);
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
  cluster_sync_all();

  Ring rg;
  rg.base = smem_u32(ring_p); rg.slot_bytes = pl.slot_bytes; rg.nslots = k.nslots; rg.full = full_bar; rg.empty = empty_bar;
  const uint8_t* stream_cta = k.stream + (size_t)c * pl.step_bytes;

  if (warp == 0) {
    // ===================================================== producer =====================================================
    RingPos pos = {0u, 0u};
#pragma unroll 1
    for (int t = 0; t < a.T; ++t)
      for (int si = 0; si < 4; ++si) produce_seg(pl, si, stream_cta, rg, pos);
  } else if (warp == 1) {
    // ===================================================== MMA issuer =====================================================
    RingPos pos = {0u, 0u};
    const uint32_t xa_a = smem_u32(xa), xb_a = smem_u32(xb);
    const uint32_t nbytes[4] = {(uint32_t)CL * (sper >> 3) * brow, (uint32_t)CL * (hper >> 3) * brow, (uint32_t)CL * 3 * (dper >> 3) * brow,
                                (uint32_t)CL * (uper >> 3) * brow};
#pragma unroll 1
    for (int it = 0; it < a.T; ++it) {
      const uint32_t par = (uint32_t)it & 1u;
#pragma unroll
      for (int si = 0; si < 4; ++si) {
        if (elect_one()) mbar_expect_tx(&act_bar[si], nbytes[si]);
        __syncwarp();
        mbar_wait(&act_bar[si], par);
        mma_seg(pl, si, (si == SEG_R1T || si == SEG_GRUT) ? xa_a : xb_a, tmem_base, rg, pos, nb);
        if (elect_one()) umma_commit(&acc_bar[si]);
        __syncwarp();
      }
    }
  } else {
    // ===================================================== epilogue =====================================================
    const int et = tid - 64, ew = et >> 5;
    // nb = 16: one half-warp per batch row (see rssm_scan_tc.cu); nb = 8: warps 0-3 take two rows each, warps 4-7 only keep the barriers
    const int hl = lane >> 4, l16 = lane & 15, mrow_raw = 2 * ew + hl;
    const bool has_row = mrow_raw < nb;
    const int mrow = has_row ? mrow_raw : 0;        // (row-less warps skip the per-row work: they only keep the barriers and do their share of the pushes)
    const uint32_t stage_a = smem_u32(stage);
    float* lnp0 = lnp;
    float* lnp1 = lnp + (size_t)CL * nb * 2;
    float* lnp2 = lnp + (size_t)2 * CL * nb * 2;

    // asynchronous fetch of rows m < Bv of a global [.., ld] slice (cols floats, 16 B aligned) into smem [m][cols]
    auto fetch_rows = [&](float* dst, const float* src, int ld, int cols) {
      const int cpr = cols >> 2;
#pragma unroll 1
      for (int m = ew; m < Bv; m += EW)
        for (int ch = lane; ch < cpr; ch += 32) cp_async16(dst + m * cols + ch * 4, src + (size_t)m * ld + ch * 4);
    };
    auto fetch_b1 = [&](int t) {
      const size_t r = (size_t)t * B + b0;
      fetch_rows(f_ps, a.ps + r * S + s0, S, sper);
      fetch_rows(f_cf, a.coef + r * S + s0, S, sper);
      fetch_rows(f_pn, a.pn + r * S + s0, S, sper);
      fetch_rows(f_dl, a.dpost_logits + r * S + s0, S, sper);
      fetch_rows(f_dz, a.dlatent + r * lat + s0, lat, sper);
    };
    auto fetch_b2 = [&](int t) {
      const size_t r = (size_t)t * B + b0;
      fetch_rows(f_hp, a.hidr_pre + r * Hh + h0, Hh, hper);
      if (et < Bv) { cp_async4(f_st + et, a.hidr_mean + r + et); cp_async4(f_st + nb + et, a.hidr_rstd + r + et); }
    };
    auto fetch_b3 = [&](int t) {
      const size_t r = (size_t)t * B + b0;
      fetch_rows(f_dh, a.dlatent + r * lat + S + d0, lat, dper);
      fetch_rows(f_hi, a.x_gru + r * a.ld_xgru + d0, a.ld_xgru, dper);
      for (int gate = 0; gate < 3; ++gate) fetch_rows(f_gp + gate * nb * dper, a.gates_pre + r * D3 + gate * D + d0, D3, dper);
      if (et < Bv) { cp_async4(f_st + 2 * nb + et, a.gru_mean + r + et); cp_async4(f_st + 3 * nb + et, a.gru_rstd + r + et); }
    };
    auto fetch_b4 = [&](int t) {
      const size_t r = (size_t)t * B + b0;
      fetch_rows(f_fp, a.feat_pre + r * U + u0, U, uper);
      if (et < Bv) { cp_async4(f_st + 4 * nb + et, a.feat_mean + r + et); cp_async4(f_st + 5 * nb + et, a.feat_rstd + r + et); }
    };
    // row sums over a vector spread over the cluster (LayerNorm backward): s1 = sum gg, s2 = sum gg * xh over this CTA's n values of
    // batch row mrow -> every peer (st.async on its barrier) -> after the wait the CL partials are added.  Returns s1 / n_total, s2 / n_total.
    auto row_sums = [&](const float* gg, const float* xh, int n, int n_total, float* lnp_x, uint64_t* bar, uint32_t parity, float& m1, float& m2) {
      float s1 = 0.f, s2 = 0.f;
      for (int j = l16; j < n; j += 16) { const float g = gg[j * ob + mrow]; s1 += g; s2 = fmaf(g, xh[j * ob + mrow], s2); }
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) { s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
      if (l16 < CL && has_row) st_async_f2(mapa(smem_u32(lnp_x + (c * nb + mrow) * 2), (uint32_t)l16), s1, s2, mapa(smem_u32(bar), (uint32_t)l16));
      mbar_wait(bar, parity);
      const float2 pc = (l16 < CL) ? reinterpret_cast<const float2*>(lnp_x)[l16 * nb + mrow] : make_float2(0.f, 0.f);
      s1 = pc.x; s2 = pc.y;
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) { s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
      m1 = s1 / (float)n_total; m2 = s2 / (float)n_total;
    };
    auto fsigmoid = [](float x) { return __fdividef(1.f, 1.f + __expf(-x)); };
    auto ftanh = [](float x) { const float e = __expf(-2.f * fabsf(x)); return copysignf(__fdividef(1.f - e, 1.f + e), x); };

    fetch_b1(a.T - 1);
    fetch_b2(a.T - 1);
    fetch_b3(a.T - 1);
    fetch_b4(a.T - 1);

#pragma unroll 1
    for (int t = a.T - 1; t >= 0; --t) {
      const uint32_t par = (uint32_t)(a.T - 1 - t) & 1u;
      const size_t row0 = (size_t)t * B + b0;
      const bool rst = (s_rmask[t] >> mrow) & 1u;
      const bool live = has_row && mrow < Bv;
      BDBG(0);
      if (et == 0) {
        mbar_expect_tx(&sum_bar[SM_HID], (uint32_t)CL * nb * 8);
        mbar_expect_tx(&sum_bar[SM_GRU], (uint32_t)CL * nb * 8);
        mbar_expect_tx(&sum_bar[SM_FEAT], (uint32_t)CL * nb * 8);
      }
      cp_async_wait_all();
      epi_bar();                                                 // this step's saved tensors landed (other threads' cp.async)
      // ================= B1: dz_t -> d post_raw (own categoricals of batch row mrow; two classes per lane) =================
#pragma unroll 1
      for (int gl = 0; gl < (has_row ? gper : 0); ++gl) {     // (row-less warps only keep the barriers and the pushes)
        float o[2];
        float dz[2], pk[2], dot = 0.f;
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int cls = l16 + 16 * j, e = mrow * sper + gl * Kc + cls;
          const bool on = cls < Kc;
          dz[j] = on ? (f_dz[e] + dzc[e]) : 0.f;
          pk[j] = on ? f_pn[e] : 0.f;
          dot = fmaf(pk[j], dz[j], dot);
        }
#pragma unroll
        for (int of = 8; of > 0; of >>= 1) dot += __shfl_xor_sync(0xffffffffu, dot, of);
        float dot2 = 0.f, ps[2], gps[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int cls = l16 + 16 * j, e = mrow * sper + gl * Kc + cls;
          const bool on = cls < Kc;
          o[j] = (on ? f_dl[e] : 0.f) + pk[j] * (dz[j] - dot);
          ps[j] = on ? f_ps[e] : 0.f;
          gps[j] = (on ? f_cf[e] : 0.f) * o[j];
          dot2 = fmaf(ps[j], gps[j], dot2);
        }
        if (a.unimix > 0.f) {
#pragma unroll
          for (int of = 8; of > 0; of >>= 1) dot2 += __shfl_xor_sync(0xffffffffu, dot2, of);
#pragma unroll
          for (int j = 0; j < 2; ++j) o[j] = ps[j] * (gps[j] - dot2);
        }
#pragma unroll
        for (int j = 0; j < 2; ++j) {
          const int cls = l16 + 16 * j, sl = gl * Kc + cls;
          if (cls < Kc) {
            const float ov = live ? o[j] : 0.f;
            if (live) {
              a.dpost_raw[(row0 + mrow) * S + s0 + sl] = ov;
              if (m_pr) m_pr[(row0 + mrow) * S + s0 + sl] = __float2bfloat16_rn(ov);
            }
            *reinterpret_cast<bf*>(stage + b_off(mrow, sl, brow)) = __float2bfloat16_rn(ov);
          }
        }
      }
      fence_proxy_async();
      epi_bar();
      push_act(stage_a, (uint32_t)(sper >> 3) * brow, smem_u32(xa), &act_bar[X_PR], c, CL, ew);
      BDBG(1);
      if (t > 0) fetch_b1(t - 1);

      // ================= B2: d hid = Wr1^T . d post_raw (own rows), LayerNorm + SiLU backward =================
      mbar_wait(&acc_bar[SEG_R1T], par);
      tc_fence_after();
      BDBG(2);
      tmem_to_outb(tmem_base, pl.seg[SEG_R1T].tmem_col, hper, pl.seg[SEG_R1T].m64, outb, nullptr, 0, ew, lane, nb);
      epi_bar();
      if (has_row) {
        const float mean = f_st[mrow], rstd = f_st[nb + mrow];
        for (int j = l16; j < hper; j += 16) {
          const float xh = (f_hp[mrow * hper + j] - mean) * rstd;
          const float ln = xh * p_r[j] + p_r[hper + j];
          const float sg = fsigmoid(ln);
          const float g = outb[j * ob + mrow] * sg * (1.f + ln * (1.f - sg));
          if (live) a.dln_r[(row0 + mrow) * Hh + h0 + j] = g;
          wg[j * ob + mrow] = g * p_r[j];
          wx[j * ob + mrow] = xh;
        }
        __syncwarp();
        float m1, m2;
        row_sums(wg, wx, hper, Hh, lnp0, &sum_bar[SM_HID], par, m1, m2);
        BDBG(3);
        for (int j = 2 * l16; j < hper; j += 32) {
          float o0 = rstd * (wg[j * ob + mrow] - m1 - wx[j * ob + mrow] * m2);
          float o1 = rstd * (wg[(j + 1) * ob + mrow] - m1 - wx[(j + 1) * ob + mrow] * m2);
          if (!live) { o0 = 0.f; o1 = 0.f; }
          const __nv_bfloat162 o16 = __floats2bfloat162_rn(o0, o1);
          if (live) {
            *reinterpret_cast<float2*>(a.dhidr_pre + (row0 + mrow) * Hh + h0 + j) = make_float2(o0, o1);
            if (m_hp) *reinterpret_cast<__nv_bfloat162*>(m_hp + (row0 + mrow) * Hh + h0 + j) = o16;
          }
          *reinterpret_cast<__nv_bfloat162*>(stage + b_off(mrow, j, brow)) = o16;
        }
      }
      fence_proxy_async();
      epi_bar();
      push_act(stage_a, (uint32_t)(hper >> 3) * brow, smem_u32(xb), &act_bar[X_HP], c, CL, ew);
      BDBG(4);
      if (t > 0) fetch_b2(t - 1);

      // ================= B3: dh_t += Wr0[:, :D]^T . d hidr_pre (own units), GRU gate backward =================
      mbar_wait(&acc_bar[SEG_R0T], par);
      tc_fence_after();
      BDBG(5);
      tmem_to_outb(tmem_base, pl.seg[SEG_R0T].tmem_col, dper, pl.seg[SEG_R0T].m64, outb, nullptr, 0, ew, lane, nb);
      epi_bar();
      if (has_row) {
        const float mu = f_st[2 * nb + mrow], rs = f_st[3 * nb + mrow];
        for (int dl = l16; dl < dper; dl += 16) {
          const int e = mrow * dper + dl;
          const float dho = outb[dl * ob + mrow] + f_dh[e] + dhc[e];
          const float xr = (f_gp[e] - mu) * rs, xc = (f_gp[nb * dper + e] - mu) * rs, xu = (f_gp[2 * nb * dper + e] - mu) * rs;
          const float r = fsigmoid(xr * p_gru[dl] + p_gru[3 * dper + dl]);
          const float cp = xc * p_gru[dper + dl] + p_gru[4 * dper + dl];
          const float cc = ftanh(r * cp);
          const float u = fsigmoid(xu * p_gru[2 * dper + dl] + p_gru[5 * dper + dl] - 1.f);
          const float du = dho * (cc - f_hi[e]), dc = dho * u;
          dhd[e] = dho * (1.f - u);
          const float dpc = dc * (1.f - cc * cc);
          const float g_r = dpc * cp * r * (1.f - r), g_c = dpc * r, g_u = du * u * (1.f - u);
          if (live) {
            float* dl_ = a.dln_gru + (row0 + mrow) * D3 + d0 + dl;
            dl_[0] = g_r; dl_[D] = g_c; dl_[2 * D] = g_u;
          }
          wx[dl * ob + mrow] = xr; wx[(dper + dl) * ob + mrow] = xc; wx[(2 * dper + dl) * ob + mrow] = xu;
          wg[dl * ob + mrow] = g_r * p_gru[dl];
          wg[(dper + dl) * ob + mrow] = g_c * p_gru[dper + dl];
          wg[(2 * dper + dl) * ob + mrow] = g_u * p_gru[2 * dper + dl];
        }
        __syncwarp();
        float m1, m2;
        row_sums(wg, wx, 3 * dper, D3, lnp1, &sum_bar[SM_GRU], par, m1, m2);
        BDBG(6);
        for (int gate = 0; gate < 3; ++gate)
          for (int dl = 2 * l16; dl < dper; dl += 32) {
            const int j = gate * dper + dl;
            float o0 = rs * (wg[j * ob + mrow] - m1 - wx[j * ob + mrow] * m2);
            float o1 = rs * (wg[(j + 1) * ob + mrow] - m1 - wx[(j + 1) * ob + mrow] * m2);
            if (!live) { o0 = 0.f; o1 = 0.f; }
            const __nv_bfloat162 o16 = __floats2bfloat162_rn(o0, o1);
            if (live) {
              *reinterpret_cast<float2*>(a.dgates_pre + (row0 + mrow) * D3 + gate * D + d0 + dl) = make_float2(o0, o1);
              if (m_gp) *reinterpret_cast<__nv_bfloat162*>(m_gp + (row0 + mrow) * D3 + gate * D + d0 + dl) = o16;
            }
            *reinterpret_cast<__nv_bfloat162*>(stage + (size_t)gate * (dper >> 3) * brow + b_off(mrow, dl, brow)) = o16;
          }
      }
      fence_proxy_async();
      epi_bar();
      // the own slice sits at three places of the [gate][D] operand: one bulk copy per gate and peer
#pragma unroll 1
      for (int p = ew; p < 3 * CL; p += EW) {
        const int peer = p % CL, gate = p / CL;
        const uint32_t bytes = (uint32_t)(dper >> 3) * brow;
        if (elect_one())
          bulk_s2peer(mapa(smem_u32(xa) + (uint32_t)(gate * (D >> 3)) * brow + (uint32_t)c * bytes, (uint32_t)peer), stage_a + (uint32_t)gate * bytes, bytes,
                      mapa(smem_u32(&act_bar[X_GP]), (uint32_t)peer));
        __syncwarp();
      }
      BDBG(7);
      if (t > 0) fetch_b3(t - 1);

      // ================= B4: [dh_{t-1} part | d feat] = Wgru^T . d gates_pre (own rows), LayerNorm + SiLU backward =================
      mbar_wait(&acc_bar[SEG_GRUT], par);
      tc_fence_after();
      BDBG(8);
      tmem_to_outb(tmem_base, pl.seg[SEG_GRUT].tmem_col, dper + uper, pl.seg[SEG_GRUT].m64, outb, nullptr, 0, ew, lane, nb);
      epi_bar();
      if (has_row) {
        for (int dl = l16; dl < dper; dl += 16) dhp[mrow * dper + dl] = outb[dl * ob + mrow];
        const float mean = f_st[4 * nb + mrow], rstd = f_st[5 * nb + mrow];
        for (int j = l16; j < uper; j += 16) {
          const float xh = (f_fp[mrow * uper + j] - mean) * rstd;
          const float ln = xh * p_mlp[j] + p_mlp[uper + j];
          const float sg = fsigmoid(ln);
          const float g = outb[(dper + j) * ob + mrow] * sg * (1.f + ln * (1.f - sg));
          if (live) a.dln_feat[(row0 + mrow) * U + u0 + j] = g;
          wg[j * ob + mrow] = g * p_mlp[j];
          wx[j * ob + mrow] = xh;
        }
        __syncwarp();
        float m1, m2;
        row_sums(wg, wx, uper, U, lnp2, &sum_bar[SM_FEAT], par, m1, m2);
        BDBG(9);
        for (int j = 2 * l16; j < uper; j += 32) {
          float o0 = rstd * (wg[j * ob + mrow] - m1 - wx[j * ob + mrow] * m2);
          float o1 = rstd * (wg[(j + 1) * ob + mrow] - m1 - wx[(j + 1) * ob + mrow] * m2);
          if (!live) { o0 = 0.f; o1 = 0.f; }
          const __nv_bfloat162 o16 = __floats2bfloat162_rn(o0, o1);
          if (live) {
            *reinterpret_cast<float2*>(a.dfeat_pre + (row0 + mrow) * U + u0 + j) = make_float2(o0, o1);
            if (m_fp) *reinterpret_cast<__nv_bfloat162*>(m_fp + (row0 + mrow) * U + u0 + j) = o16;
          }
          *reinterpret_cast<__nv_bfloat162*>(stage + b_off(mrow, j, brow)) = o16;
        }
      }
      fence_proxy_async();
      epi_bar();
      push_act(stage_a, (uint32_t)(uper >> 3) * brow, smem_u32(xb), &act_bar[X_FP], c, CL, ew);
      BDBG(10);
      if (t > 0) fetch_b4(t - 1);

      // ================= B5: dz_{t-1} = Wmlp[:, :S]^T . d feat_pre (own categoricals); carry through the reset mask =================
      mbar_wait(&acc_bar[SEG_MLPT], par);
      tc_fence_after();
      BDBG(11);
      tmem_to_outb(tmem_base, pl.seg[SEG_MLPT].tmem_col, sper, pl.seg[SEG_MLPT].m64, outb, nullptr, 0, ew, lane, nb);
      epi_bar();
      if (has_row) {
        for (int sl = l16; sl < sper; sl += 16) dzc[mrow * sper + sl] = rst ? 0.f : outb[sl * ob + mrow];
        for (int dl = l16; dl < dper; dl += 16) { const int e = mrow * dper + dl; dhc[e] = rst ? 0.f : (dhp[e] + dhd[e]); }
      }
      BDBG(12);
    }
    tc_fence_before();
  }

  // ---- teardown
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(128));
  }
}

// ---- host ------------------------------------------------------------------------------------------------------------------------
bool build_plan_bwd(int D, int U, int Hh, int G, int Kc, int CL, int slot_max, int nb, Plan* pl) {
  if (CL < 1 || D % CL || U % CL || Hh % CL || G % CL) return false;
  const int uper = U / CL, dper = D / CL, hper = Hh / CL, gper = G / CL, sper = gper * Kc, S = G * Kc;
  if ((uper & 7) || (dper & 7) || (hper & 7) || (sper & 7) || Kc > 32 || (D & 15) || (U & 15) || (Hh & 15) || (S & 15)) return false;
  if (hper > 128 || sper > 128 || dper > 128 || dper + uper > 128 * 3) return false;
  memset(pl, 0, sizeof(Plan));
  uint32_t off16 = 0;
  const int tr = nb >= 16 ? 128 : 64;                         // N = 8 is only legal with M = 64 tiles
  const int gt = (dper + uper + tr - 1) / tr;
  RowBlk hb[1] = {{0, hper}};
  RowBlk db[1] = {{0, dper}};
  RowBlk gb[2] = {{0, dper}, {D, uper}};
  RowBlk sb[1] = {{0, sper}};
  const int t1 = (hper + tr - 1) / tr, t2 = (dper + tr - 1) / tr, t4 = (sper + tr - 1) / tr;
  if (plan_add_seg(*pl, slot_max, hper, S, 0, 1, 0, 1, hb, 0, off16, tr) != SEG_R1T) return false;
  if (plan_add_seg(*pl, slot_max, dper, Hh, t1 * NB, 1, 0, 1, db, 0, off16, tr) != SEG_R0T) return false;
  if (plan_add_seg(*pl, slot_max, dper + uper, 3 * D, (t1 + t2) * NB, 1, 0, 2, gb, 0, off16, tr) != SEG_GRUT) return false;
  if (plan_add_seg(*pl, slot_max, sper, U, (t1 + t2 + gt) * NB, 1, 0, 1, sb, 0, off16, tr) != SEG_MLPT) return false;
  if ((t1 + t2 + gt + t4) * NB > 128) return false;
  pl->step_bytes = off16 * 16;
  return true;
}

constexpr int T_MAX_B = 1024;

size_t fixed_smem_bwd(int D, int U, int Hh, int G, int Kc, int CL, int T, int nb) {
  const int NB = nb, OB = nb + 1, BROW = nb * 16;      // shadow the 16-row namespace constants
  const int uper = U / CL, dper = D / CL, hper = Hh / CL, gper = G / CL, sper = gper * Kc, S = G * Kc;
  auto r = [](size_t b) { return (b + 127) & ~(size_t)127; };
  auto mx = [](int x, int y) { return x > y ? x : y; };
  const int fmax = mx(mx(sper, hper), mx(3 * dper, uper)), rows_max = mx(mx(hper, 3 * dper), mx(dper + uper, sper));
  return r((size_t)(mx(S, 3 * D) >> 3) * BROW) + r((size_t)(mx(Hh, U) >> 3) * BROW) + r((size_t)(fmax >> 3) * BROW) + 3 * r((size_t)rows_max * OB * 4) +
         r((size_t)3 * CL * NB * 8) + 6 * r((size_t)NB * sper * 4) + 5 * r((size_t)NB * dper * 4) + r((size_t)NB * hper * 4) + r((size_t)3 * NB * dper * 4) +
         r((size_t)NB * uper * 4) + r(6 * NB * 4) + r((size_t)T * 4) + r((size_t)2 * hper * 4) + r((size_t)6 * dper * 4) + r((size_t)2 * uper * 4) +
         2048 + 128 + 128;
}

struct ChoiceB { int CL, nslots, nb; size_t smem; };
// Preference: 16 batch rows per cluster over 8 (half the clusters, same per-cluster time); 16 CTAs over 8; a ring of >= 4 slots of >= 8 KB.
bool select_plan_bwd(int D, int U, int Hh, int G, int Kc, int max_smem, Plan* pl, ChoiceB* ch) {
  for (int nb = 16; nb >= 8; nb >>= 1)
  for (int CL = 16; CL >= 8; CL >>= 1) {
    for (int want = 4; want >= 3; --want) {
      for (int slot_max = 32768; slot_max >= 8192; slot_max >>= 1) {
        if (!build_plan_bwd(D, U, Hh, G, Kc, CL, slot_max, nb, pl)) continue;
        const size_t fixed = fixed_smem_bwd(D, U, Hh, G, Kc, CL, T_MAX_B, nb);
        const long long room = (long long)max_smem - (long long)fixed - 512;
        int nslots = (int)(room / (long long)pl->slot_bytes);
        if (nslots < want) continue;
        if (nslots > 12) nslots = 12;
        ch->CL = CL; ch->nslots = nslots; ch->nb = nb;
        ch->smem = fixed + (size_t)nslots * pl->slot_bytes + (((size_t)(2 * nslots + 16) * 8 + 127) & ~(size_t)127);
        return true;
      }
    }
  }
  return false;
}
int device_max_smem_b() {
  int dev = 0, max_smem = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 0;
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  return max_smem;
}

struct PackSrcB { const float* W; long long sr, sc; };
struct PackKB { Plan plan; PackSrcB src[MAX_SEG]; uint8_t* dst; int CL; };
// one block per (cta, chunk); unit = 8 consecutive k of one row (16 B in the stream); see rssm_scan_tc.cu
__global__ void __launch_bounds__(256) scan_tc_pack_bwd_kernel(const __grid_constant__ PackKB p) {
  const int cta = blockIdx.x / p.plan.nchunks, ci = blockIdx.x - cta * p.plan.nchunks;
  const Chunk ch = p.plan.c[ci];
  const Seg& sg = p.plan.seg[ch.seg];
  const PackSrcB src = p.src[ch.seg];
  const int rp = ch.lbo16, nunits = (int)ch.bytes16;
  uint4* dst = reinterpret_cast<uint4*>(p.dst + (size_t)cta * p.plan.step_bytes) + ch.off16;
  for (int u = threadIdx.x; u < nunits; u += blockDim.x) {
    const int k8 = u / rp, r = u - k8 * rp;
    const int j = ch.tile * sg.tile_rows + r;
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
    dst[u] = out;
  }
}

}  // namespace

extern "C" int dv3_scan_tc_bwd_config(int D, int U, int Hh, int G, int Kc, int* cluster_size, size_t* stream_bytes) {
  Plan pl;
  ChoiceB ch;
  if (!select_plan_bwd(D, U, Hh, G, Kc, device_max_smem_b(), &pl, &ch)) return DV3_ERR_UNSUPPORTED;
  if (cluster_size) *cluster_size = ch.CL;
  if (stream_bytes) *stream_bytes = (size_t)ch.CL * pl.step_bytes;
  return DV3_OK;
}

// Transposed weights straight from the fp32 parameters (strides swapped; no transposed copy is materialised).
extern "C" int dv3_scan_tc_bwd_pack(int D, int U, int Hh, int G, int Kc, const float* Wr1, int ld_wr1, const float* Wr0, int ld_wr0,
                                    const float* Wgru, int ld_wgru, const float* Wmlp, int ld_wmlp, void* dst, void* stream) {
  if (!Wr1 || !Wr0 || !Wgru || !Wmlp || !dst) return DV3_ERR_ARG;
  PackKB p;
  ChoiceB ch;
  if (!select_plan_bwd(D, U, Hh, G, Kc, device_max_smem_b(), &p.plan, &ch)) return DV3_ERR_UNSUPPORTED;
  p.CL = ch.CL;
  p.dst = reinterpret_cast<uint8_t*>(dst);
  p.src[SEG_R1T] = {Wr1, 1, ld_wr1};       // row h of Wr1^T = column h of Wr1 [S, Hh];  k = s
  p.src[SEG_R0T] = {Wr0, 1, ld_wr0};       // row d of Wr0[:, :D]^T;                     k = hh
  p.src[SEG_GRUT] = {Wgru, 1, ld_wgru};    // row (d | D + u) of Wgru^T [D+U, 3D];       k = gate row
  p.src[SEG_MLPT] = {Wmlp, 1, ld_wmlp};    // row s of Wmlp[:, :S]^T;                    k = u
  scan_tc_pack_bwd_kernel<<<ch.CL * p.plan.nchunks, 256, 0, (cudaStream_t)stream>>>(p);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}

extern "C" int dv3_rssm_observe_bwd_tc(const Dv3ScanBwdArgs* h, const void* stream_pk, void* stream_) {
  if (!h || !stream_pk) return DV3_ERR_ARG;
  if (h->B < 1 || h->T < 1 || h->T > T_MAX_B || h->G < 1 || h->Kc < 1 || h->Kc > 32) return DV3_ERR_UNSUPPORTED;
  const int S = h->G * h->Kc;
  if ((S & 15) || (h->Hh & 15) || (h->D & 15) || (h->U & 15) || (h->ld_dlat & 3) || (h->ld_xgru & 3) ||
      (((uintptr_t)stream_pk | (uintptr_t)h->ps | (uintptr_t)h->coef | (uintptr_t)h->pn | (uintptr_t)h->dpost_logits | (uintptr_t)h->dlatent |
        (uintptr_t)h->hidr_pre | (uintptr_t)h->gates_pre | (uintptr_t)h->x_gru | (uintptr_t)h->feat_pre | (uintptr_t)h->dpost_raw_bf16 |
        (uintptr_t)h->dhidr_pre_bf16 | (uintptr_t)h->dgates_pre_bf16 | (uintptr_t)h->dfeat_pre_bf16 | (uintptr_t)h->dhidr_pre |
        (uintptr_t)h->dgates_pre | (uintptr_t)h->dfeat_pre) & 15))
    return DV3_ERR_UNSUPPORTED;
  const int max_smem = device_max_smem_b();
  BwdK* k = new BwdK;
  ChoiceB ch;
  int rc = DV3_ERR_UNSUPPORTED;
  if (select_plan_bwd(h->D, h->U, h->Hh, h->G, h->Kc, max_smem, &k->plan, &ch)) {
    k->a = *h;
    if (!getenv("DV3_SCAN_DEBUG")) k->a.dbg = nullptr;
    k->stream = reinterpret_cast<const uint8_t*>(stream_pk);
    k->CL = ch.CL;
    k->nslots = ch.nslots;
    k->nb = ch.nb;
    const int nclusters = (h->B + ch.nb - 1) / ch.nb;
    if (getenv("DV3_SCAN_DEBUG"))
      fprintf(stderr, "[dv3 tc reverse scan] rows/cluster=%d CL=%d clusters=%d slots=%d x %u B, step stream %u B/CTA, %d chunks, smem %zu KB (max %d KB)\n", ch.nb, ch.CL, nclusters,
              ch.nslots, k->plan.slot_bytes, k->plan.step_bytes, k->plan.nchunks, ch.smem / 1024, max_smem / 1024);
    auto kern = rssm_observe_bwd_tc_kernel;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ch.smem);
    if (e == cudaSuccess && ch.CL > 8) e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess) { cudaGetLastError(); delete k; return DV3_ERR_UNSUPPORTED; }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(ch.CL * nclusters, 1, 1);
    cfg.blockDim = dim3(NT, 1, 1);
    cfg.dynamicSmemBytes = ch.smem;
    cfg.stream = (cudaStream_t)stream_;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = ch.CL;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    int nmax = 0;
    e = cudaOccupancyMaxActiveClusters(&nmax, kern, &cfg);
    if (e != cudaSuccess || nmax < 1) { cudaGetLastError(); delete k; return DV3_ERR_UNSUPPORTED; }
    e = cudaLaunchKernelEx(&cfg, kern, *k);
    rc = (e == cudaSuccess) ? DV3_OK : (int)e;
  }
  delete k;
  return rc;
}
