// Reverse scan of the RSSM observe loop (backward through time of RSSM.dynamic) on ONE thread-block cluster.
//
// Reference: what loss.backward() does for the T sequential steps of dreamer_v3.py:122-129 -- the hand-written reverse of
// RSSM.dynamic (agent.py:350-388): straight-through one-hot + unimix categorical (agent.py:404-420, distribution.py), representation
// MLP, LayerNormGRUCell (models/models.py:388-395), RecurrentModel.mlp (agent.py:296-309).  Only dL/dh_t and dL/dz_t propagate
// sequentially; all parameter gradients are K = T*B contractions done after the scan from the tensors saved here.
//
// Per step t = T-1 .. 0 (own = this CTA's 1/CL slice of the feature dimension; every contraction streams bf16 rows of a TRANSPOSED
// weight matrix from L2 into mma.sync fragments, see scan_cluster.cuh):
//   B1  dz_t (own categoricals) = head gradient + carry  ->  d post_raw  (unimix + straight-through Jacobian; the softmax
//       quantities come from dv3_unimix_bwd_prep, computed for all T*B rows before the scan)              -> all-gather (bf16)
//   B2  d hid  = d post_raw . Wr1        (own Hh columns);  LayerNorm+SiLU backward (row sums over the cluster) -> d hidr_pre
//   B3  dh_t  += d hidr_pre . Wr0[:, :D] (own D columns);   GRU gate backward + LayerNorm backward over 3D      -> d gates_pre
//   B4  [dh_{t-1} part | d feat] = d gates_pre . Wgru (own columns);  LayerNorm+SiLU backward                    -> d feat_pre
//   B5  dz_{t-1} = d feat_pre . Wmlp[:, :S] (own categoricals);  carry = (1 - first_t) * (dz_{t-1}, dh_{t-1} part + dh_t (1 - u))
// Seven hardware cluster barriers per step; saved-activation inputs of step t-1 are fetched with cp.async while step t runs.
#define DV3_SCAN_CT 256   // threads per CTA (8 warps: 1.84 ms; 16 warps at the 128-register cap: 1.93 ms)
#include "scan_cluster.cuh"
#include "dv3_b200.h"
#include <stdio.h>
#include <stdlib.h>

namespace {
using namespace sc;

struct KArgs {
  Dv3ScanBwdArgs a;
  int CL;
};

// per-row sums over a vector spread over the cluster: s1 = sum gg, s2 = sum gg * xh (LayerNorm backward); pushed to every peer
__device__ __noinline__ void sum2_push(const float* gg, const float* xh, int n, float* lnp, int c, int CL) {
  cg::cluster_group cluster = cg::this_cluster();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll 1
  for (int m = warp; m < MB; m += CW) {
    float s1 = 0.f, s2 = 0.f;
    for (int j = lane; j < n; j += 32) { const float g = gg[j * OB + m]; s1 += g; s2 = fmaf(g, xh[j * OB + m], s2); }
    s1 = warp_sum(s1);
    s2 = warp_sum(s2);
    if (lane < CL) reinterpret_cast<float2*>(cluster.map_shared_rank(lnp, lane))[c * MB + m] = make_float2(s1, s2);
  }
}
__device__ __noinline__ void sum2_combine(const float* lnp, int CL, int n_total, float* s_stat) {   // after the cluster barrier
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll 1
  for (int m = warp; m < MB; m += CW) {
    const float2 pc = (lane < CL) ? reinterpret_cast<const float2*>(lnp)[lane * MB + m] : make_float2(0.f, 0.f);
    const float s1 = warp_sum(pc.x), s2 = warp_sum(pc.y);
    if (lane == 0) { s_stat[m] = s1 / (float)n_total; s_stat[MB + m] = s2 / (float)n_total; }
  }
  __syncthreads();
}
// LayerNorm(+SiLU) backward of the own slice, first half: dact[col][m] (in place -> gg = g * gamma), pre[m][col] (saved pre-LN
// activation), statistics mean / rstd[m], prm = [gamma(cols) | beta(cols)]; writes xh to xhb[col][m] and g (= dL/d LN-output) to gsave.
__device__ __noinline__ void ln_bwd_a(float* dact, const float* pre, int cols, const float* mean, const float* rstd, const float* prm,
                                      float* xhb, float* gsave) {
  const int lg = 31 - __clz(cols);
#pragma unroll 1
  for (int e = threadIdx.x; e < MB * cols; e += CT) {
    const int m = e >> lg, col = e & (cols - 1);
    const float xh = (pre[e] - mean[m]) * rstd[m];
    const float ln = xh * prm[col] + prm[cols + col];
    const float sg = sigmoidf_(ln);
    const float g = dact[col * OB + m] * sg * (1.f + ln * (1.f - sg));
    gsave[e] = g;
    dact[col * OB + m] = g * prm[col];
    xhb[col * OB + m] = xh;
  }
}
// second half: dpre = rstd * (gg - m1 - xh * m2) -> dsave[m][col] (fp32, stored later) and stage[m * sld + col] (bf16, all-gathered)
__device__ __noinline__ void ln_bwd_b(const float* gg, const float* xhb, int cols, const float* rstd, const float* s_stat, float* dsave,
                                      __nv_bfloat16* stage, int sld) {
  const int lg = 31 - __clz(cols);
#pragma unroll 1
  for (int e = threadIdx.x; e < MB * cols; e += CT) {
    const int m = e >> lg, col = e & (cols - 1);
    const float o = rstd[m] * (gg[col * OB + m] - s_stat[m] - xhb[col * OB + m] * s_stat[MB + m]);
    dsave[e] = o;
    stage[m * sld + col] = __float2bfloat16_rn(o);
  }
}
// rows m < B of a [MB][cols] fp32 buffer -> global (row stride ld) and, optionally, the bf16 staging copy -> the global bf16 mirror
__device__ __noinline__ void store_saved(float* dst, __nv_bfloat16* dst16, int ld, const float* src, const __nv_bfloat16* stage, int sld, int cols,
                                         int B) {
  const int lg = 31 - __clz(cols);
#pragma unroll 1
  for (int e = threadIdx.x; e < B * cols; e += CT) {
    const int m = e >> lg, col = e & (cols - 1);
    dst[(size_t)m * ld + col] = src[e];
  }
  if (dst16) {
    const int cpr = cols >> 3, lgc = 31 - __clz(cpr);
#pragma unroll 1
    for (int i = threadIdx.x; i < B * cpr; i += CT) {
      const int m = i >> lgc, ch = i & (cpr - 1);
      *reinterpret_cast<uint4*>(dst16 + (size_t)m * ld + ch * 8) = *reinterpret_cast<const uint4*>(stage + m * sld + ch * 8);
    }
  }
}
// asynchronous fetch of rows m < B of a global [.., ld] slice (cols floats, 16 B aligned) into smem [m][cols]
__device__ __noinline__ void fetch_rows(float* dst, const float* src, int ld, int cols, int B) {
  const int cpr = cols >> 2, lgc = 31 - __clz(cpr);
#pragma unroll 1
  for (int i = threadIdx.x; i < B * cpr; i += CT) {
    const int m = i >> lgc, ch = i & (cpr - 1);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(dst + m * cols + ch * 4)),
                 "l"(src + (size_t)m * ld + ch * 4)
                 : "memory");
  }
}

#define BDBG(p) do { if (a.dbg && c == 0 && tid == 0 && t == a.T - 2) a.dbg[p] = clock64(); } while (0)

__global__ void __launch_bounds__(CT, 1) rssm_observe_bwd_cluster_kernel(const __grid_constant__ KArgs k) {
  cg::cluster_group cluster = cg::this_cluster();
  const Dv3ScanBwdArgs& a = k.a;
  extern __shared__ __align__(16) uint8_t smraw[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int c = (int)cluster.block_rank(), CL = k.CL;
  const int B = a.B, G = a.G, Kc = a.Kc, D = a.D, U = a.U, Hh = a.Hh;
  const int S = G * Kc, D3 = 3 * D, lat = a.ld_dlat;
  const int uper = U / CL, dper = D / CL, hper = Hh / CL, gper = G / CL, sper = gper * Kc;   // powers of two, multiples of 16
  const int u0 = c * uper, d0 = c * dper, h0 = c * hper, s0 = c * sper;
  const int lg_d = 31 - __clz(dper), lg_gp = 31 - __clz(gper), lg_s = 31 - __clz(sper);
  const int ld_pr = S + 32, ld_hp = Hh + 32, ld_gp = D3 + 32, ld_fp = U + 32;
  const int npairs = B * gper;
  const int sld = max(max(sper, hper), max(3 * dper, uper));
  typedef __nv_bfloat16 bf;
  const uint4* Wr1T = reinterpret_cast<const uint4*>(a.Wr1T_pk);      // transposed weights packed by dv3_pack_mma_a
  const uint4* Wr0hT = reinterpret_cast<const uint4*>(a.Wr0hT_pk);
  const uint4* WgruT = reinterpret_cast<const uint4*>(a.WgruT_pk);
  const uint4* WmlpzT = reinterpret_cast<const uint4*>(a.WmlpzT_pk);
  bf* m_pr = reinterpret_cast<bf*>(a.dpost_raw_bf16);
  bf* m_hp = reinterpret_cast<bf*>(a.dhidr_pre_bf16);
  bf* m_gp = reinterpret_cast<bf*>(a.dgates_pre_bf16);
  bf* m_fp = reinterpret_cast<bf*>(a.dfeat_pre_bf16);

  uint8_t* sp = smraw;
  auto take = [&](size_t bytes) { uint8_t* p = sp; sp += (bytes + 15) & ~(size_t)15; return p; };
  auto takef = [&](size_t n) { return reinterpret_cast<float*>(take(n * 4)); };
  __nv_bfloat16* x_pr = reinterpret_cast<__nv_bfloat16*>(take((size_t)MB * ld_pr * 2));   // d post_raw   [MB][S]   (MMA operands)
  __nv_bfloat16* x_hp = reinterpret_cast<__nv_bfloat16*>(take((size_t)MB * ld_hp * 2));   // d hidr_pre   [MB][Hh]
  __nv_bfloat16* x_gp = reinterpret_cast<__nv_bfloat16*>(take((size_t)MB * ld_gp * 2));   // d gates_pre  [MB][3D]
  __nv_bfloat16* x_fp = reinterpret_cast<__nv_bfloat16*>(take((size_t)MB * ld_fp * 2));   // d feat_pre   [MB][U]
  __nv_bfloat16* stage = reinterpret_cast<__nv_bfloat16*>(take((size_t)MB * sld * 2));
  const int rows_max = max(max(hper, 3 * dper), max(dper + uper, sper));
  float* outb = takef((size_t)rows_max * OB);
  float* xhb = takef((size_t)rows_max * OB);
  float* part = takef((size_t)max(max(part_kb(hper >> 4), part_kb(dper >> 4)), max(part_kb((dper + uper) >> 4), part_kb(sper >> 4))) * 256);
  float* lnp = takef((size_t)CL * MB * 2);
  float* s_stat = takef(2 * MB);
  int* s_first = reinterpret_cast<int*>(take((size_t)a.T * MB * 4));
  float* p_r = takef(2 * hper);
  float* p_gru = takef(6 * dper);
  float* p_mlp = takef(2 * uper);
  // Own-slice buffers, grouped by size so that each one is `group base + index * size` (cheap to rematerialise: keeping ~25 separate
  // pointers live across the time loop is what pushes the 16-warp build over its 128-register cap).
  float* gS = takef((size_t)7 * MB * sper);
  float* gD = takef((size_t)14 * MB * dper);
  float* gH = takef((size_t)3 * MB * hper);
  float* gU = takef((size_t)3 * MB * uper);
#define dzc (gS)                               /* carried dL/dz (own categoricals) */
#define sv_pr (gS + 1 * MB * sper)             /* saved gradients whose global store is deferred by one barrier: sv_* */
#define f_ps (gS + 2 * MB * sper)              /* cp.async landing zones (saved forward tensors of the next step to process): f_* */
#define f_cf (gS + 3 * MB * sper)
#define f_pn (gS + 4 * MB * sper)
#define f_dl (gS + 5 * MB * sper)
#define f_dz (gS + 6 * MB * sper)
#define dhc (gD)                               /* carried dL/dh (own units) */
#define dhd (gD + 1 * MB * dper)               /* dh_t * (1 - u): direct path to h_{t-1} */
#define dhp (gD + 2 * MB * dper)               /* h_{t-1} part of d gates_pre . Wgru */
#define f_dh (gD + 3 * MB * dper)
#define f_hi (gD + 4 * MB * dper)
#define sv_gp (gD + 5 * MB * dper)             /* [gate][m][dl] */
#define sv_lg (gD + 8 * MB * dper)
#define f_gp (gD + 11 * MB * dper)
#define sv_hp (gH)
#define sv_lr (gH + 1 * MB * hper)
#define f_hp (gH + 2 * MB * hper)
#define sv_fp (gU)
#define sv_lf (gU + 1 * MB * uper)
#define f_fp (gU + 2 * MB * uper)
  float* f_st = takef(6 * MB);                            // hidr mean|rstd, gru mean|rstd, feat mean|rstd

  for (int i = tid; i < MB * ld_pr / 2; i += CT) reinterpret_cast<uint32_t*>(x_pr)[i] = 0u;
  for (int i = tid; i < MB * ld_hp / 2; i += CT) reinterpret_cast<uint32_t*>(x_hp)[i] = 0u;
  for (int i = tid; i < MB * ld_gp / 2; i += CT) reinterpret_cast<uint32_t*>(x_gp)[i] = 0u;
  for (int i = tid; i < MB * ld_fp / 2; i += CT) reinterpret_cast<uint32_t*>(x_fp)[i] = 0u;
  for (int i = tid; i < MB * sper; i += CT) { dzc[i] = 0.f; f_ps[i] = f_cf[i] = f_pn[i] = f_dl[i] = f_dz[i] = 0.f; }
  for (int i = tid; i < MB * dper; i += CT) { dhc[i] = 0.f; f_dh[i] = f_hi[i] = 0.f; }
  for (int i = tid; i < MB * 3 * dper; i += CT) f_gp[i] = 0.f;
  for (int i = tid; i < MB * hper; i += CT) f_hp[i] = 0.f;
  for (int i = tid; i < MB * uper; i += CT) f_fp[i] = 0.f;
  for (int i = tid; i < 6 * MB; i += CT) f_st[i] = (i / MB) & 1 ? 1.f : 0.f;
  for (int i = tid; i < a.T * MB; i += CT) {
    const int t = i / MB, m = i - t * MB;
    s_first[i] = (m >= B || t == 0 || __ldg(a.first + t * B + m) != 0.f) ? 1 : 0;
  }
  for (int i = tid; i < hper; i += CT) { p_r[i] = __ldg(a.g_r + h0 + i); p_r[hper + i] = __ldg(a.b_r + h0 + i); }
  for (int i = tid; i < 3 * dper; i += CT) {
    const int gate = i / dper, dl = i - gate * dper;
    p_gru[i] = __ldg(a.g_gru + gate * D + d0 + dl);
    p_gru[3 * dper + i] = __ldg(a.b_gru + gate * D + d0 + dl);
  }
  for (int i = tid; i < uper; i += CT) { p_mlp[i] = __ldg(a.g_mlp + u0 + i); p_mlp[uper + i] = __ldg(a.b_mlp + u0 + i); }
  __syncthreads();

  auto fetch_b1 = [&](int t) {
    const size_t r = (size_t)t * B;
    fetch_rows(f_ps, a.ps + r * S + s0, S, sper, B);
    fetch_rows(f_cf, a.coef + r * S + s0, S, sper, B);
    fetch_rows(f_pn, a.pn + r * S + s0, S, sper, B);
    fetch_rows(f_dl, a.dpost_logits + r * S + s0, S, sper, B);
    fetch_rows(f_dz, a.dlatent + r * lat + s0, lat, sper, B);
  };
  auto fetch_b2 = [&](int t) {
    const size_t r = (size_t)t * B;
    fetch_rows(f_hp, a.hidr_pre + r * Hh + h0, Hh, hper, B);
    if (tid < B) { cp_async4(f_st + tid, a.hidr_mean + r + tid); cp_async4(f_st + MB + tid, a.hidr_rstd + r + tid); }
  };
  auto fetch_b3 = [&](int t) {
    const size_t r = (size_t)t * B;
    fetch_rows(f_dh, a.dlatent + r * lat + S + d0, lat, dper, B);
    fetch_rows(f_hi, a.x_gru + r * a.ld_xgru + d0, a.ld_xgru, dper, B);
    for (int gate = 0; gate < 3; ++gate) fetch_rows(f_gp + gate * MB * dper, a.gates_pre + r * D3 + gate * D + d0, D3, dper, B);
    if (tid < B) { cp_async4(f_st + 2 * MB + tid, a.gru_mean + r + tid); cp_async4(f_st + 3 * MB + tid, a.gru_rstd + r + tid); }
  };
  auto fetch_b4 = [&](int t) {
    const size_t r = (size_t)t * B;
    fetch_rows(f_fp, a.feat_pre + r * U + u0, U, uper, B);
    if (tid < B) { cp_async4(f_st + 4 * MB + tid, a.feat_mean + r + tid); cp_async4(f_st + 5 * MB + tid, a.feat_rstd + r + tid); }
  };
  fetch_b1(a.T - 1);
  fetch_b2(a.T - 1);
  fetch_b3(a.T - 1);
  fetch_b4(a.T - 1);
  cluster.sync();          // every CTA's buffers exist before any peer pushes into them

#pragma unroll 1
  for (int t = a.T - 1; t >= 0; --t) {
    const size_t row0 = (size_t)t * B;
    BDBG(0);
    // ================= B1: dz_t -> d post_raw (own categoricals; lane = class) =================
    cp_async_wait_all();
    __syncthreads();
#pragma unroll 1
    for (int p = warp; p < npairs; p += CW) {
      const int m = p >> lg_gp, gl = p & (gper - 1), e = m * sper + gl * Kc + lane;
      const bool on = lane < Kc;
      const float dz = on ? (f_dz[e] + dzc[e]) : 0.f;
      const float pk = on ? f_pn[e] : 0.f;
      const float dot = warp_sum(pk * dz);
      float o = (on ? f_dl[e] : 0.f) + pk * (dz - dot);
      if (a.unimix > 0.f) {
        const float ps = on ? f_ps[e] : 0.f;
        const float gps = (on ? f_cf[e] : 0.f) * o;
        const float dot2 = warp_sum(ps * gps);
        o = ps * (gps - dot2);
      }
      if (on) { sv_pr[e] = o; stage[m * sld + gl * Kc + lane] = __float2bfloat16_rn(o); }
    }
    __syncthreads();
    allgather(x_pr, ld_pr, s0, sper, stage, sld, CL);
    BDBG(1);
    cluster_arrive();                                                                 // #1 (stores / prefetches hide in the barrier)
    if (t > 0) fetch_b1(t - 1);
    store_saved(a.dpost_raw + row0 * S + s0, m_pr ? m_pr + row0 * S + s0 : nullptr, S, sv_pr, stage, sld, sper, B);
    cluster_wait();
    BDBG(2);

    // ================= B2: d hid = d post_raw . Wr1 (own columns), LayerNorm + SiLU backward =================
    contract(Wr1T, hper >> 4, h0, hper >> 4, 0, S, x_pr, ld_pr, part, outb);
    ln_bwd_a(outb, f_hp, hper, f_st, f_st + MB, p_r, xhb, sv_lr);
    __syncthreads();
    sum2_push(outb, xhb, hper, lnp, c, CL);
    BDBG(3);
    cluster.sync();                                                                   // #2
    BDBG(4);
    sum2_combine(lnp, CL, Hh, s_stat);
    ln_bwd_b(outb, xhb, hper, f_st + MB, s_stat, sv_hp, stage, sld);
    __syncthreads();
    allgather(x_hp, ld_hp, h0, hper, stage, sld, CL);
    BDBG(5);
    cluster_arrive();                                                                 // #3
    store_saved(a.dhidr_pre + row0 * Hh + h0, m_hp ? m_hp + row0 * Hh + h0 : nullptr, Hh, sv_hp, stage, sld, hper, B);
    store_saved(a.dln_r + row0 * Hh + h0, nullptr, Hh, sv_lr, nullptr, 0, hper, B);
    if (t > 0) fetch_b2(t - 1);
    cluster_wait();
    BDBG(6);

    // ================= B3: dh_t += d hidr_pre . Wr0[:, :D] (own units), GRU gate backward =================
    contract(Wr0hT, dper >> 4, d0, dper >> 4, 0, Hh, x_hp, ld_hp, part, outb);
#pragma unroll 1
    for (int e = tid; e < MB * dper; e += CT) {
      const int m = e >> lg_d, dl = e & (dper - 1);
      const float mu = f_st[2 * MB + m], rs = f_st[3 * MB + m];
      const float dho = outb[dl * OB + m] + f_dh[e] + dhc[e];
      const float xr = (f_gp[e] - mu) * rs, xc = (f_gp[MB * dper + e] - mu) * rs, xu = (f_gp[2 * MB * dper + e] - mu) * rs;
      const float r = sigmoidf_(xr * p_gru[dl] + p_gru[3 * dper + dl]);
      const float cp = xc * p_gru[dper + dl] + p_gru[4 * dper + dl];
      const float cc = tanhf(r * cp);
      const float u = sigmoidf_(xu * p_gru[2 * dper + dl] + p_gru[5 * dper + dl] - 1.f);
      const float du = dho * (cc - f_hi[e]), dc = dho * u;
      dhd[e] = dho * (1.f - u);
      const float dpc = dc * (1.f - cc * cc);
      const float g_r = dpc * cp * r * (1.f - r), g_c = dpc * r, g_u = du * u * (1.f - u);
      sv_lg[e] = g_r; sv_lg[MB * dper + e] = g_c; sv_lg[2 * MB * dper + e] = g_u;
      xhb[dl * OB + m] = xr; xhb[(dper + dl) * OB + m] = xc; xhb[(2 * dper + dl) * OB + m] = xu;
      outb[dl * OB + m] = g_r * p_gru[dl];                  // outb row dl was read above by this same thread only
      outb[(dper + dl) * OB + m] = g_c * p_gru[dper + dl];
      outb[(2 * dper + dl) * OB + m] = g_u * p_gru[2 * dper + dl];
    }
    __syncthreads();
    sum2_push(outb, xhb, 3 * dper, lnp, c, CL);
    BDBG(7);
    cluster.sync();                                                                   // #4
    BDBG(8);
    sum2_combine(lnp, CL, D3, s_stat);
    for (int gate = 0; gate < 3; ++gate)
      ln_bwd_b(outb + gate * dper * OB, xhb + gate * dper * OB, dper, f_st + 3 * MB, s_stat, sv_gp + gate * MB * dper, stage + gate * dper, sld);
    __syncthreads();
    for (int gate = 0; gate < 3; ++gate) allgather(x_gp, ld_gp, gate * D + d0, dper, stage + gate * dper, sld, CL);
    BDBG(9);
    cluster_arrive();                                                                 // #5
    for (int gate = 0; gate < 3; ++gate) {
      store_saved(a.dgates_pre + row0 * D3 + gate * D + d0, m_gp ? m_gp + row0 * D3 + gate * D + d0 : nullptr, D3,
                  sv_gp + gate * MB * dper, stage + gate * dper, sld, dper, B);
      store_saved(a.dln_gru + row0 * D3 + gate * D + d0, nullptr, D3, sv_lg + gate * MB * dper, nullptr, 0, dper, B);
    }
    if (t > 0) fetch_b3(t - 1);
    cluster_wait();
    BDBG(10);

    // ================= B4: [dh_{t-1} part | d feat] = d gates_pre . Wgru (own columns), LayerNorm + SiLU backward =================
    contract(WgruT, (dper + uper) >> 4, d0, dper >> 4, D + u0 - d0, D3, x_gp, ld_gp, part, outb);   // dper == uper
#pragma unroll 1
    for (int e = tid; e < MB * dper; e += CT) dhp[e] = outb[(e & (dper - 1)) * OB + (e >> lg_d)];
    ln_bwd_a(outb + dper * OB, f_fp, uper, f_st + 4 * MB, f_st + 5 * MB, p_mlp, xhb, sv_lf);
    __syncthreads();
    sum2_push(outb + dper * OB, xhb, uper, lnp, c, CL);
    BDBG(11);
    cluster.sync();                                                                   // #6
    BDBG(12);
    sum2_combine(lnp, CL, U, s_stat);
    ln_bwd_b(outb + dper * OB, xhb, uper, f_st + 5 * MB, s_stat, sv_fp, stage, sld);
    __syncthreads();
    allgather(x_fp, ld_fp, u0, uper, stage, sld, CL);
    BDBG(13);
    cluster_arrive();                                                                 // #7
    store_saved(a.dfeat_pre + row0 * U + u0, m_fp ? m_fp + row0 * U + u0 : nullptr, U, sv_fp, stage, sld, uper, B);
    store_saved(a.dln_feat + row0 * U + u0, nullptr, U, sv_lf, nullptr, 0, uper, B);
    if (t > 0) fetch_b4(t - 1);
    cluster_wait();
    BDBG(14);

    // ================= B5: dz_{t-1} = d feat_pre . Wmlp[:, :S] (own categoricals); carry through the reset mask =================
    contract(WmlpzT, sper >> 4, s0, sper >> 4, 0, U, x_fp, ld_fp, part, outb);
    const int* rs_t = s_first + t * MB;
#pragma unroll 1
    for (int e = tid; e < MB * sper; e += CT) { const int m = e >> lg_s; dzc[e] = rs_t[m] ? 0.f : outb[(e & (sper - 1)) * OB + m]; }
#pragma unroll 1
    for (int e = tid; e < MB * dper; e += CT) dhc[e] = rs_t[e >> lg_d] ? 0.f : (dhp[e] + dhd[e]);
    BDBG(15);
  }
}

#undef dzc
#undef sv_pr
#undef f_ps
#undef f_cf
#undef f_pn
#undef f_dl
#undef f_dz
#undef dhc
#undef dhd
#undef dhp
#undef f_dh
#undef f_hi
#undef sv_gp
#undef sv_lg
#undef f_gp
#undef sv_hp
#undef sv_lr
#undef f_hp
#undef sv_fp
#undef sv_lf
#undef f_fp

bool shape_ok(const Dv3ScanBwdArgs* h, int CL, size_t* smem_bytes, int max_smem) {
  if (h->U % (16 * CL) || h->D % (16 * CL) || h->Hh % (16 * CL) || h->G % CL || ((h->G / CL) * h->Kc) % 16) return false;
  const int S = h->G * h->Kc, D3 = 3 * h->D;
  const int uper = h->U / CL, dper = h->D / CL, hper = h->Hh / CL, gper = h->G / CL, sper = gper * h->Kc;
  auto pow2 = [](int v) { return v > 0 && (v & (v - 1)) == 0; };
  if (!pow2(uper) || !pow2(dper) || !pow2(hper) || !pow2(gper) || !pow2(h->Kc) || h->Kc > 32 || dper != uper) return false;
  if (((dper + uper) >> 4) > MAXT || (sper >> 4) > MAXT) return false;
  auto r16 = [](size_t b) { return (b + 15) & ~(size_t)15; };
  auto mx = [](int x, int y) { return x > y ? x : y; };
  const int sld = mx(mx(sper, hper), mx(3 * dper, uper)), rows_max = mx(mx(hper, 3 * dper), mx(dper + uper, sper));
  const int pk = mx(mx(part_kb(hper >> 4), part_kb(dper >> 4)), mx(part_kb((dper + uper) >> 4), part_kb(sper >> 4)));
  size_t b = r16((size_t)MB * (S + 32) * 2) + r16((size_t)MB * (h->Hh + 32) * 2) + r16((size_t)MB * (D3 + 32) * 2) + r16((size_t)MB * (h->U + 32) * 2) +
             r16((size_t)MB * sld * 2) + 2 * r16((size_t)rows_max * OB * 4) + r16((size_t)pk * 1024) + r16((size_t)CL * MB * 8) + r16(2 * MB * 4) +
             r16((size_t)h->T * MB * 4) + r16((size_t)2 * hper * 4) + r16((size_t)6 * dper * 4) + r16((size_t)2 * uper * 4) +
             r16((size_t)MB * sper * 4) + 3 * r16((size_t)MB * dper * 4) +                                          // dzc, dhc, dhd, dhp (dhc counted here)
             r16((size_t)MB * dper * 4) +
             r16((size_t)MB * sper * 4) + 2 * r16((size_t)MB * hper * 4) + 2 * r16((size_t)MB * 3 * dper * 4) + 2 * r16((size_t)MB * uper * 4) +   // sv_*
             5 * r16((size_t)MB * sper * 4) + 2 * r16((size_t)MB * dper * 4) + r16((size_t)MB * hper * 4) + r16((size_t)MB * 3 * dper * 4) +
             r16((size_t)MB * uper * 4) + r16(6 * MB * 4);
  *smem_bytes = b + 64;
  return (long long)*smem_bytes <= (long long)max_smem;
}

int launch_bwd(KArgs& k, int CL, size_t smem_bytes, cudaStream_t stream, bool probe_only) {
  auto kern = rssm_observe_bwd_cluster_kernel;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
  if (e != cudaSuccess) { cudaGetLastError(); return DV3_ERR_UNSUPPORTED; }
  if (CL > 8) {
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 1);
    if (e != cudaSuccess) { cudaGetLastError(); return DV3_ERR_UNSUPPORTED; }
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(CL, 1, 1);
  cfg.blockDim = dim3(CT, 1, 1);
  cfg.dynamicSmemBytes = smem_bytes;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = CL;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  int nclusters = 0;
  e = cudaOccupancyMaxActiveClusters(&nclusters, kern, &cfg);
  if (e != cudaSuccess || nclusters < 1) { cudaGetLastError(); return DV3_ERR_UNSUPPORTED; }
  if (probe_only) return DV3_OK;
  k.CL = CL;
  e = cudaLaunchKernelEx(&cfg, kern, k);
  if (e != cudaSuccess) return (int)e;
  return DV3_OK;
}

}  // namespace

extern "C" int dv3_rssm_observe_bwd_cluster(const Dv3ScanBwdArgs* h, void* stream_) {
  if (!h) return DV3_ERR_ARG;
  if (h->B < 1 || h->B > MB || h->T < 1 || h->G < 1 || h->Kc < 1) return DV3_ERR_UNSUPPORTED;
  const int S = h->G * h->Kc;
  if (!h->Wr1T_pk || !h->Wr0hT_pk || !h->WgruT_pk || !h->WmlpzT_pk) return DV3_ERR_UNSUPPORTED;
  if ((S & 31) || (h->Hh & 31) || (h->D & 31) || (h->U & 31) || (h->ld_dlat & 3) || (h->ld_xgru & 3) ||
      (((uintptr_t)h->Wr1T_pk | (uintptr_t)h->Wr0hT_pk | (uintptr_t)h->WgruT_pk | (uintptr_t)h->WmlpzT_pk | (uintptr_t)h->ps | (uintptr_t)h->coef |
        (uintptr_t)h->pn | (uintptr_t)h->dpost_logits | (uintptr_t)h->dlatent | (uintptr_t)h->hidr_pre | (uintptr_t)h->gates_pre | (uintptr_t)h->x_gru |
        (uintptr_t)h->feat_pre | (uintptr_t)h->dpost_raw_bf16 | (uintptr_t)h->dhidr_pre_bf16 | (uintptr_t)h->dgates_pre_bf16 |
        (uintptr_t)h->dfeat_pre_bf16) & 15))
    return DV3_ERR_UNSUPPORTED;
  int dev = 0, max_smem = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  KArgs k;
  k.a = *h;
  k.CL = 0;
  const char* force = getenv("DV3_SCAN_CLUSTER");
  const int want = force ? atoi(force) : 16;
  size_t smem = 0;
  for (int CL = 16; CL >= 8; CL >>= 1) {
    if (CL > want || !shape_ok(h, CL, &smem, max_smem)) continue;
    if (launch_bwd(k, CL, smem, (cudaStream_t)stream_, true) != DV3_OK) continue;
    return launch_bwd(k, CL, smem, (cudaStream_t)stream_, false);
  }
  return DV3_ERR_UNSUPPORTED;
}
