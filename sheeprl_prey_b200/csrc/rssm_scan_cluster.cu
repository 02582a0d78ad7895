// RSSM observe scan (forward) on ONE thread-block cluster: distributed shared memory instead of grid-wide barriers.
//
// Reference: the Python loop sheeprl/algos/dreamer_v3/dreamer_v3.py:122-129 over RSSM.dynamic (agent.py:350-388):
// RecurrentModel (agent.py:296-309) -> LayerNormGRUCell (models/models.py:388-395) -> _representation (agent.py:404-420).
//
// Why a second kernel: the cooperative kernel (rssm_scan.cu) spreads every contraction over all 148 SMs and pays ~6 grid-wide
// barriers + L2 round trips per time step (~48 us/step measured, 35 % of warp samples spinning in the barrier).  The recurrence
// only has B <= 16 independent rows, so the arithmetic is tiny; what matters is the latency of the dependent chain.  Here a
// single cluster of CL CTAs (16, or 8 if the larger size cannot be scheduled) runs the whole scan:
//   * each CTA owns 1/CL of the output features of every in-loop contraction and streams its bf16 weight rows from L2 straight
//     into tensor-core fragments (mma.sync m16n8k16, weights = A operand, the 16 batch rows = two n8 tiles): the per-step weight
//     traffic (~0.3 MB per CTA at config S) is the only bulk data movement and it does not depend on the recurrence;
//   * activations are exchanged by pushing each CTA's bf16 slice into every peer's shared memory (DSMEM, 16 B stores) and the
//     LayerNorm statistics by pushing per-CTA (mean, M2) partials (Chan's parallel variance); ordering comes from the hardware
//     cluster barrier (barrier.cluster, ~0.2 us) -- 7 per step instead of 6 grid barriers at ~1.1 us + L2 latency each;
//   * the one-hot stochastic state never becomes a dense operand: feat_pre = sum_g WmlpT[g*Kc + idx_g, :] (32 coalesced row
//     gathers per batch row) + the hoisted action projection aW_t.
// Requires the bf16 weight mirrors (bf16 precision mode); fp32 mode keeps the cooperative kernel.
// Everything the hand-written backward needs is written to the same buffers, in fp32, as by the per-op path.
#define DV3_SCAN_CT 512   // threads per CTA (16 warps: measured 1.79 ms vs 1.94 ms with 8)
#include "scan_cluster.cuh"
#include "dv3_b200.h"
#include <stdio.h>
#include <stdlib.h>

namespace {
using namespace sc;

struct KArgs {
  Dv3ScanFwdArgs a;
  const uint4* Wgru_pk;         // in-loop weights packed in MMA-fragment order (dv3_pack_mma_a)
  const uint4* Wr0_pk;
  const uint4* Wr1_pk;
  const float* WmlpzT;          // [S, U] fp32 = Wmlp[:, :S]^T
  const float* aW;              // [T*B, U] fp32 = masked action . Wmlp[:, S:]^T
  long long* dbg;
  int CL;
  int w_resident;               // Wr0 / Wr1 tiles of the CTA fit in shared memory next to the working set
};

#define SDBG(p) do { if (k.dbg && c == 0 && tid == 0 && t == 1) k.dbg[p] = clock64(); } while (0)

__global__ void __launch_bounds__(CT, 1) rssm_observe_fwd_cluster_kernel(const __grid_constant__ KArgs k) {
  cg::cluster_group cluster = cg::this_cluster();
  const Dv3ScanFwdArgs& a = k.a;
  extern __shared__ __align__(16) uint8_t smraw[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int c = (int)cluster.block_rank(), CL = k.CL;
  const int B = a.B, G = a.G, Kc = a.Kc, D = a.D, U = a.U, Hh = a.Hh;
  const int S = G * Kc, KG = D + U;
  const int uper = U / CL, dper = D / CL, hper = Hh / CL, gper = G / CL, sper = gper * Kc;   // all multiples of 16 (host check)
  const int u0 = c * uper, d0 = c * dper, h0 = c * hper, s0 = c * sper;
  const int ldxg = KG + 32, ldh = Hh + 32;                 // +32 bf16: the two 8-row halves of a fragment load hit different banks
  const int tg = dper >> 4, nt_g = 3 * tg, nt_h = hper >> 4, nt_s = sper >> 4;
  const int npairs = B * gper;
  // per-CTA slice sizes are powers of two (host check): index arithmetic by shifts keeps the loop body small
  const int lg_u = 31 - __clz(uper), lg_d = 31 - __clz(dper), lg_h = 31 - __clz(hper), lg_gp = 31 - __clz(gper), lg_s = 31 - __clz(sper),
            lg_kc = 31 - __clz(Kc), lg_G = 31 - __clz(G);

  // ---- shared memory carve-up (16 B aligned regions)
  uint8_t* sp = smraw;
  auto take = [&](size_t bytes) { uint8_t* p = sp; sp += (bytes + 15) & ~(size_t)15; return p; };
  __nv_bfloat16* xg = reinterpret_cast<__nv_bfloat16*>(take((size_t)MB * ldxg * 2));   // [MB][h | feat] bf16 (MMA operand)
  __nv_bfloat16* hid = reinterpret_cast<__nv_bfloat16*>(take((size_t)MB * ldh * 2));   // [MB][Hh] bf16
  const int rows_max = max(max(3 * dper, uper), max(hper, sper));
  float* outb = reinterpret_cast<float*>(take((size_t)rows_max * OB * 4));             // own pre-activations [row][OB]
  float* outl = reinterpret_cast<float*>(take((size_t)sper * OB * 4));                 // own unimix logits (kept until flushed)
  float* part = reinterpret_cast<float*>(take((size_t)max(max(part_kb(nt_g), part_kb(nt_h)), part_kb(nt_s)) * 1024));
  float* hown = reinterpret_cast<float*>(take((size_t)MB * dper * 4));                 // own fp32 slice of h
  const int sld = max(max(uper, dper), hper);
  __nv_bfloat16* stage = reinterpret_cast<__nv_bfloat16*>(take((size_t)MB * sld * 2));
  float* dfr = reinterpret_cast<float*>(take((size_t)MB * sld * 4));                   // fp32 activations whose global store is deferred
  float* lnp = reinterpret_cast<float*>(take((size_t)CL * MB * 2 * 4));                // [CL][MB](mean, M2) pushed by the peers
  float* s_stat = reinterpret_cast<float*>(take(2 * MB * 4));
  int* idxf = reinterpret_cast<int*>(take((size_t)MB * G * 4));                        // z indices [MB][G] (all groups)
  int* idxo = reinterpret_cast<int*>(take((size_t)MB * gper * 4));                     // own sampled indices (kept until flushed)
  int* s_zreset = reinterpret_cast<int*>(take((size_t)G * 4));
  int* s_first = reinterpret_cast<int*>(take((size_t)a.T * MB * 4));                   // reset flag of every (t, m)
  float* p_mlp = reinterpret_cast<float*>(take((size_t)2 * uper * 4));                 // own LayerNorm weight | bias slices
  float* p_gru = reinterpret_cast<float*>(take((size_t)2 * 3 * dper * 4));
  float* p_r = reinterpret_cast<float*>(take((size_t)2 * hper * 4));
  float* p_br1 = reinterpret_cast<float*>(take((size_t)sper * 4));
  float* s_aw = reinterpret_cast<float*>(take((size_t)MB * uper * 4));                 // cp.async landing zones (next step's inputs)
  float* s_ew = reinterpret_cast<float*>(take((size_t)MB * hper * 4));
  float* s_uu = reinterpret_cast<float*>(take((size_t)MB * gper * 4));                 // uniforms (or forced indices, as int bits)
  // resident weights (k.w_resident): this CTA's packed tiles of Wr0[:, :D] and Wr1 stay in shared memory for all T steps
  uint4* w_r0_s = k.w_resident ? reinterpret_cast<uint4*>(take((size_t)nt_h * (D >> 5) * 1024)) : nullptr;
  uint4* w_r1_s = k.w_resident ? reinterpret_cast<uint4*>(take((size_t)nt_s * (Hh >> 5) * 1024)) : nullptr;
  if (k.w_resident) {
    const uint4* g0p = k.Wr0_pk + (size_t)(h0 >> 4) * (D >> 5) * 64;
    for (int i = tid; i < nt_h * (D >> 5) * 64; i += CT) w_r0_s[i] = __ldg(g0p + i);
    const uint4* g1p = k.Wr1_pk + (size_t)(s0 >> 4) * (Hh >> 5) * 64;
    for (int i = tid; i < nt_s * (Hh >> 5) * 64; i += CT) w_r1_s[i] = __ldg(g1p + i);
  }

  for (int i = tid; i < MB * ldxg / 2; i += CT) reinterpret_cast<uint32_t*>(xg)[i] = 0u;
  for (int i = tid; i < MB * ldh / 2; i += CT) reinterpret_cast<uint32_t*>(hid)[i] = 0u;
  for (int i = tid; i < MB * dper; i += CT) hown[i] = 0.f;
  for (int i = tid; i < MB * G; i += CT) idxf[i] = 0;
  for (int i = tid; i < MB * uper; i += CT) s_aw[i] = 0.f;
  for (int i = tid; i < MB * hper; i += CT) s_ew[i] = 0.f;
  for (int i = tid; i < G; i += CT) s_zreset[i] = __ldg(a.zreset_idx + i);
  for (int i = tid; i < a.T * MB; i += CT) {
    const int t = i / MB, m = i - t * MB;
    s_first[i] = (m >= B || t == 0 || __ldg(a.first + t * B + m) != 0.f) ? 1 : 0;
  }
  for (int i = tid; i < uper; i += CT) { p_mlp[i] = __ldg(a.g_mlp + u0 + i); p_mlp[uper + i] = __ldg(a.b_mlp + u0 + i); }
  for (int i = tid; i < 3 * dper; i += CT) {
    const int gate = i / dper, dl = i - gate * dper;
    p_gru[i] = __ldg(a.g_gru + gate * D + d0 + dl);
    p_gru[3 * dper + i] = __ldg(a.b_gru + gate * D + d0 + dl);
  }
  for (int i = tid; i < hper; i += CT) { p_r[i] = __ldg(a.g_r + h0 + i); p_r[hper + i] = __ldg(a.b_r + h0 + i); }
  for (int i = tid; i < sper; i += CT) p_br1[i] = __ldg(a.br1 + s0 + i);
  __syncthreads();

  // inputs of step t that do not depend on the recurrence, fetched asynchronously one step ahead (cp.async, 4 B each)
  auto fetch_aw = [&](int t) {
    for (int e = tid; e < B * uper; e += CT) { const int m = e >> lg_u, ul = e & (uper - 1); cp_async4(s_aw + e, k.aW + (size_t)(t * B + m) * U + u0 + ul); }
  };
  auto fetch_ew = [&](int t) {
    for (int e = tid; e < B * hper; e += CT) { const int m = e >> lg_h, hl = e & (hper - 1); cp_async4(s_ew + e, a.eW + (size_t)(t * B + m) * Hh + h0 + hl); }
  };
  auto fetch_uu = [&](int t) {
    for (int e = tid; e < npairs; e += CT) {
      const int m = e >> lg_gp, g = c * gper + (e & (gper - 1));
      if (a.forced_idx) cp_async4(s_uu + e, a.forced_idx + (size_t)(t * B + m) * G + g);
      else cp_async4(s_uu + e, a.u_post + (size_t)(t * B + m) * G + g);
    }
  };
  // global stores of the posterior of the step that just ended (issued after the barrier, see store_rows)
  auto flush_posterior = [&](int row0) {
    store_cols(a.post_raw + (size_t)row0 * S + s0, S, outb, sper, B);
    store_cols(a.post_logits + (size_t)row0 * S + s0, S, outl, sper, B);
#pragma unroll 1
    for (int e = tid; e < B * sper; e += CT) {
      const int m = e >> lg_s, sl = e & (sper - 1), gl = sl >> lg_kc;
      a.latent[(size_t)(row0 + m) * a.ld_lat + s0 + sl] = (idxo[m * gper + gl] == (sl & (Kc - 1))) ? 1.f : 0.f;
    }
    for (int e = tid; e < npairs; e += CT) { const int m = e >> lg_gp, gl = e & (gper - 1); a.post_idx[(size_t)(row0 + m) * G + c * gper + gl] = idxo[e]; }
  };
  fetch_aw(0);
  fetch_ew(0);
  fetch_uu(0);
  cluster.sync();          // every CTA's buffers exist before any peer pushes into them

#pragma unroll 1
  for (int t = 0; t < a.T; ++t) {
    const int row0 = t * B;
    SDBG(0);
    // ================= P0: reset masking (agent.py:379-384), saved inputs of the step =================
    const int* rs_t = s_first + t * MB;
    cp_async_wait_all();
    __syncthreads();                                          // flush_posterior has read idxo / outb / outl; s_aw / s_ew / s_uu landed
#pragma unroll 1
    for (int i = tid; i < MB * G; i += CT) { if (rs_t[i >> lg_G]) idxf[i] = s_zreset[i & (G - 1)]; }
#pragma unroll 1
    for (int m = 0; m < MB; ++m)
      if (rs_t[m]) for (int ch = tid; ch < (D >> 3); ch += CT) *reinterpret_cast<uint4*>(xg + m * ldxg + ch * 8) = make_uint4(0u, 0u, 0u, 0u);
#pragma unroll 1
    for (int e = tid; e < MB * dper; e += CT) if (rs_t[e >> lg_d]) hown[e] = 0.f;
    __syncthreads();
#pragma unroll 1
    for (int i = tid; i < B * (sper >> 2); i += CT) {        // one-hot z_in, 4 classes per thread (Kc % 4 == 0)
      const int m = i >> (lg_s - 2), sl = (i & ((sper >> 2) - 1)) << 2, s = s0 + sl, g = s >> lg_kc, kk = s & (Kc - 1), ch = idxf[m * G + g];
      *reinterpret_cast<float4*>(a.x_mlp + (size_t)(row0 + m) * a.ld_xmlp + s) =
          make_float4(ch == kk ? 1.f : 0.f, ch == kk + 1 ? 1.f : 0.f, ch == kk + 2 ? 1.f : 0.f, ch == kk + 3 ? 1.f : 0.f);
    }
    store_rows(a.x_gru + (size_t)row0 * a.ld_xgru + d0, a.ld_xgru, hown, dper, dper, B);
    SDBG(1);

    // ================= P1: feat = SiLU(LN(Wmlp . [z, a])) -- one-hot gather + hoisted action projection =================
#pragma unroll 1
    for (int e = tid; e < MB * uper; e += CT) {
      const int m = e >> lg_u, ul = e & (uper - 1);
      float acc = s_aw[e];
      const int* ix = idxf + m * G;
      const float* wcol = k.WmlpzT + u0 + ul;
#pragma unroll 1
      for (int g0 = 0; g0 < G; g0 += 16) {                    // 16 independent row gathers in flight
        float v[16];
#pragma unroll
        for (int q = 0; q < 16; ++q) v[q] = (g0 + q < G) ? __ldg(wcol + (size_t)((g0 + q) * Kc + ix[g0 + q]) * U) : 0.f;
#pragma unroll
        for (int q = 0; q < 16; ++q) acc += v[q];
      }
      outb[ul * OB + m] = acc;
    }
    __syncthreads();
    ln_push(outb, uper, lnp, c, CL);
    SDBG(2);
    SDBG(11);
    cluster_arrive();                                                                 // #1
    store_cols(a.feat_pre + (size_t)row0 * U + u0, U, outb, uper, B);
    cluster_wait();
    SDBG(12);
    ln_combine(lnp, CL, uper, 1e-3f, s_stat, a.feat_mean + row0, a.feat_rstd + row0, c == 0, B);
    ln_silu_slice(outb, uper, s_stat, p_mlp, dfr, stage, sld);
    __syncthreads();
    allgather(xg, ldxg, D + u0, uper, stage, sld, CL);
    SDBG(3);
    SDBG(13);
    cluster_arrive();                                                                 // #2
    if (t + 1 < a.T) fetch_aw(t + 1);
    store_rows(a.x_gru + (size_t)row0 * a.ld_xgru + D + u0, a.ld_xgru, dfr, sld, uper, B);
    cluster_wait();
    SDBG(14);

    // ================= P2: GRU gates: own rows {d, D+d, 2D+d} of Wgru . [h, feat] =================
    contract(k.Wgru_pk, nt_g, d0, tg, D, KG, xg, ldxg, part, outb);      // outb row = gate * dper + dl
    SDBG(4);
    ln_push(outb, 3 * dper, lnp, c, CL);
    SDBG(15);
    cluster_arrive();                                                                 // #3
#pragma unroll 1
    for (int e = tid; e < B * dper; e += CT) {                // saved pre-LN gates (own rows of the three gates)
      const int m = e >> lg_d, dl = e & (dper - 1);
      float* gp = a.gates_pre + (size_t)(row0 + m) * 3 * D + d0 + dl;
      gp[0] = outb[dl * OB + m]; gp[D] = outb[(dper + dl) * OB + m]; gp[2 * D] = outb[(2 * dper + dl) * OB + m];
    }
    cluster_wait();
    SDBG(16);
    ln_combine(lnp, CL, 3 * dper, 1e-5f, s_stat, a.gru_mean + row0, a.gru_rstd + row0, c == 0, B);
#pragma unroll 1
    for (int e = tid; e < MB * dper; e += CT) {
      const int m = e >> lg_d, dl = e & (dper - 1);
      const float mu = s_stat[m], rs = s_stat[MB + m];
      const float g0v = outb[dl * OB + m], g1v = outb[(dper + dl) * OB + m], g2v = outb[(2 * dper + dl) * OB + m];
      const float r = sigmoidf_((g0v - mu) * rs * p_gru[dl] + p_gru[3 * dper + dl]);
      const float cp = (g1v - mu) * rs * p_gru[dper + dl] + p_gru[4 * dper + dl];
      const float cd = tanhf(r * cp);
      const float u = sigmoidf_((g2v - mu) * rs * p_gru[2 * dper + dl] + p_gru[5 * dper + dl] - 1.f);
      const float hn = u * cd + (1.f - u) * hown[e];
      hown[e] = hn;
      stage[m * sld + dl] = __float2bfloat16_rn(hn);
    }
    __syncthreads();
    allgather(xg, ldxg, d0, dper, stage, sld, CL);
    SDBG(5);
    SDBG(17);
    cluster_arrive();                                                                 // #4
    store_rows(a.x_rep + (size_t)row0 * a.ld_xrep + d0, a.ld_xrep, hown, dper, dper, B);
    cluster_wait();
    SDBG(18);

    // ================= P3: posterior hidden: own rows of Wr0[:, :D] . h' + (e_t . Wr0[:, D:]^T + b) =================
    if (k.w_resident) contract_t<true>(w_r0_s, nt_h, 0, nt_h, 0, D, xg, ldxg, part, outb);
    else contract(k.Wr0_pk, nt_h, h0, nt_h, 0, D, xg, ldxg, part, outb);
#pragma unroll 1
    for (int e = tid; e < B * hper; e += CT) { const int m = e >> lg_h, hl = e & (hper - 1); outb[hl * OB + m] += s_ew[e]; }
    __syncthreads();
    ln_push(outb, hper, lnp, c, CL);
    SDBG(6);
    SDBG(19);
    cluster_arrive();                                                                 // #5
    if (t + 1 < a.T) fetch_ew(t + 1);
    store_cols(a.hidr_pre + (size_t)row0 * Hh + h0, Hh, outb, hper, B);
    cluster_wait();
    SDBG(20);
    ln_combine(lnp, CL, hper, 1e-5f, s_stat, a.hidr_mean + row0, a.hidr_rstd + row0, c == 0, B);
    ln_silu_slice(outb, hper, s_stat, p_r, dfr, stage, sld);
    __syncthreads();
    allgather(hid, ldh, h0, hper, stage, sld, CL);
    SDBG(7);
    SDBG(21);
    cluster_arrive();                                                                 // #6
    store_rows(a.hidr + (size_t)row0 * Hh + h0, Hh, dfr, sld, hper, B);
    cluster_wait();
    SDBG(22);

    // ================= P4: posterior logits of the own categoricals, unimix, inverse-CDF sample =================
    if (k.w_resident) contract_t<true>(w_r1_s, nt_s, 0, nt_s, 0, Hh, hid, ldh, part, outb);
    else contract(k.Wr1_pk, nt_s, s0, nt_s, 0, Hh, hid, ldh, part, outb);
    SDBG(8);
#pragma unroll 1
    for (int p = warp; p < npairs; p += CW) {                // one warp per (batch row, own categorical); lane = class (Kc <= 32)
      const int m = p >> lg_gp, gl = p & (gper - 1), g = c * gper + gl;
      const bool on = lane < Kc;
      const float raw = on ? (outb[(size_t)(gl * Kc + lane) * OB + m] + p_br1[gl * Kc + lane]) : -INFINITY;
      const float mx = warp_max(raw);
      const float e1 = on ? expf(raw - mx) : 0.f;
      const float se = warp_sum(e1);
      float lu = raw;
      if (a.unimix > 0.f) {
        const float pr = (1.f - a.unimix) * (e1 / se) + a.unimix / (float)Kc;
        lu = on ? logf(fminf(fmaxf(pr, DV3_EPS32), 1.f - DV3_EPS32)) : -INFINITY;
      }
      // Sampling probabilities: Categorical(logits = lu).probs = softmax(lu) = pm / sum(pm) up to fp32 round-off (the reference
      // renormalises log(pm) twice); with bf16 contractions upstream that round-off is far below the logits' own error, so the
      // draw uses pm directly and a warp prefix sum instead of the sequential cumsum of the fp32 kernels.
      float prob = (a.unimix > 0.f) ? (on ? fminf(fmaxf((1.f - a.unimix) * (e1 / se) + a.unimix / (float)Kc, DV3_EPS32), 1.f - DV3_EPS32) : 0.f)
                                    : (e1 / se);
      prob = prob / warp_sum(prob);
      int choice;
      if (a.forced_idx) {
        choice = __float_as_int(s_uu[p]);
      } else {                                             // idx = #{k : cumsum_k(p) <= u}
        float cs = prob;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const float v = __shfl_up_sync(0xffffffffu, cs, o); if (lane >= o) cs += v; }
        choice = __popc(__ballot_sync(0xffffffffu, on && cs <= s_uu[p]));
        if (choice > Kc - 1) choice = Kc - 1;
      }
      if (on) {
        outb[(size_t)(gl * Kc + lane) * OB + m] = raw;
        outl[(size_t)(gl * Kc + lane) * OB + m] = lu;
      }
      if (lane == 0) idxo[p] = choice;
      if (lane < CL) cluster.map_shared_rank(idxf, lane)[m * G + g] = choice;
    }
    __syncthreads();
    SDBG(9);
    SDBG(23);
    cluster_arrive();                                                                 // #7
    if (t + 1 < a.T) fetch_uu(t + 1);
    flush_posterior(row0);
    cluster_wait();
    SDBG(24);
    SDBG(10);
  }
}

int launch_cluster(KArgs& k, int CL, size_t smem_bytes, cudaStream_t stream, bool probe_only) {
  auto kern = rssm_observe_fwd_cluster_kernel;
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

bool shape_ok(const Dv3ScanFwdArgs* h, int CL, size_t* smem_bytes, int max_smem) {
  if (h->U % (16 * CL) || h->D % (16 * CL) || h->Hh % (16 * CL) || h->G % CL || ((h->G / CL) * h->Kc) % 16) return false;
  const int KG = h->D + h->U, uper = h->U / CL, dper = h->D / CL, hper = h->Hh / CL, gper = h->G / CL, sper = gper * h->Kc;
  const int nt_g = 3 * (dper >> 4), nt_h = hper >> 4, nt_s = sper >> 4;
  if (nt_g > MAXT || nt_s > MAXT) return false;
  auto pow2 = [](int v) { return v > 0 && (v & (v - 1)) == 0; };
  if (!pow2(uper) || !pow2(dper) || !pow2(hper) || !pow2(gper) || !pow2(h->Kc) || !pow2(h->G)) return false;   // index math by shifts
  auto r16 = [](size_t b) { return (b + 15) & ~(size_t)15; };
  int rows_max = 3 * dper;
  if (uper > rows_max) rows_max = uper;
  if (hper > rows_max) rows_max = hper;
  if (sper > rows_max) rows_max = sper;
  int stg = uper > dper ? uper : dper;
  if (hper > stg) stg = hper;
  int pk = part_kb(nt_g);
  if (part_kb(nt_h) > pk) pk = part_kb(nt_h);
  if (part_kb(nt_s) > pk) pk = part_kb(nt_s);
  size_t b = r16((size_t)MB * (KG + 32) * 2) + r16((size_t)MB * (h->Hh + 32) * 2) + r16((size_t)rows_max * OB * 4) + r16((size_t)sper * OB * 4) +
             r16((size_t)pk * 1024) + r16((size_t)MB * dper * 4) + r16((size_t)MB * stg * 2) + r16((size_t)MB * stg * 4) +
             r16((size_t)CL * MB * 2 * 4) + r16(2 * MB * 4) + r16((size_t)MB * h->G * 4) + r16((size_t)MB * gper * 4) + r16((size_t)h->G * 4) +
             r16((size_t)h->T * MB * 4) + r16((size_t)2 * uper * 4) + r16((size_t)6 * dper * 4) + r16((size_t)2 * hper * 4) + r16((size_t)sper * 4) +
             r16((size_t)MB * uper * 4) + r16((size_t)MB * hper * 4) + r16((size_t)MB * gper * 4);
  *smem_bytes = b + 16;
  return (long long)*smem_bytes <= (long long)max_smem;
}
// extra shared memory for the resident Wr0 / Wr1 tiles of one CTA
size_t resident_bytes(const Dv3ScanFwdArgs* h, int CL) {
  const int hper = h->Hh / CL, sper = (h->G / CL) * h->Kc;
  return (size_t)(hper >> 4) * (h->D >> 5) * 1024 + (size_t)(sper >> 4) * (h->Hh >> 5) * 1024 + 32;
}

}  // namespace

extern "C" int dv3_rssm_observe_fwd_cluster(const Dv3ScanFwdArgs* h, const Dv3ScanClusterExtra* x, void* stream_) {
  if (!h || !x) return DV3_ERR_ARG;
  if (h->B < 1 || h->B > MB || h->T < 1 || h->G < 1 || h->Kc < 1 || h->Kc > 32 || (h->Kc & 3) || (h->ld_xmlp & 3) ||
      ((uintptr_t)h->x_mlp & 15))
    return DV3_ERR_UNSUPPORTED;
  if (!x->Wgru_pk || !x->Wr0_pk || !x->Wr1_pk || !x->WmlpzT || !x->aW) return DV3_ERR_UNSUPPORTED;
  const int KG = h->D + h->U;
  if ((KG & 31) || (h->D & 31) || (h->Hh & 31) || ((uintptr_t)x->Wgru_pk & 15) || ((uintptr_t)x->Wr0_pk & 15) || ((uintptr_t)x->Wr1_pk & 15))
    return DV3_ERR_UNSUPPORTED;   // 16 B fragment loads, 32-wide k chunks
  int dev = 0, max_smem = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  KArgs k;
  k.a = *h;
  k.Wgru_pk = reinterpret_cast<const uint4*>(x->Wgru_pk);
  k.Wr0_pk = reinterpret_cast<const uint4*>(x->Wr0_pk);
  k.Wr1_pk = reinterpret_cast<const uint4*>(x->Wr1_pk);
  k.WmlpzT = x->WmlpzT;
  k.aW = x->aW;
  k.dbg = (getenv("DV3_SCAN_DEBUG") && h->scratch && h->scratch_bytes >= 64 * sizeof(long long)) ? reinterpret_cast<long long*>(h->scratch) : nullptr;
  // cluster size: 16 (non-portable) when the device schedules it and the shape divides, else 8; $DV3_SCAN_CLUSTER caps it
  const char* force = getenv("DV3_SCAN_CLUSTER");
  const int want = force ? atoi(force) : 16;
  size_t smem = 0;
  for (int CL = 16; CL >= 8; CL >>= 1) {
    if (CL > want || !shape_ok(h, CL, &smem, max_smem)) continue;
    k.w_resident = 0;
    if (!getenv("DV3_SCAN_NO_RESIDENT") && smem + resident_bytes(h, CL) <= (size_t)max_smem) { k.w_resident = 1; smem += resident_bytes(h, CL); }
    if (getenv("DV3_SCAN_DEBUG")) fprintf(stderr, "[dv3 cluster scan] CL=%d smem=%zu KB resident=%d (max %d KB)\n", CL, smem / 1024, k.w_resident, max_smem / 1024);
    if (launch_cluster(k, CL, smem, (cudaStream_t)stream_, true) != DV3_OK) continue;
    return launch_cluster(k, CL, smem, (cudaStream_t)stream_, false);
  }
  return DV3_ERR_UNSUPPORTED;
}

// ---- weights in MMA-fragment order -----------------------------------------------------------------------------------------------
// dst[((tile * (K/32) + chunk) * 32 + lane) * 16 + j] (bf16) with lane = 4 gid + tig, j = 8 mma + 2 reg + half:
//   row = 16 tile + gid + 8 (reg & 1),   k = 32 chunk + 8 tig + 4 mma + 2 (reg >> 1) + half
// = the a0..a3 registers of the two mma.sync.m16n8k16 of a 32-wide k chunk (k order inside the chunk permuted as in sc::contract).
// src element (row, k) is src[row * stride_r + k * stride_k] (fp32), so the same kernel packs W and W^T.
namespace {
__global__ void __launch_bounds__(256) pack_mma_a_kernel(int R, int K, const float* __restrict__ src, long long stride_r, long long stride_k,
                                                         __nv_bfloat16* __restrict__ dst) {
  const size_t n = (size_t)R * K;
  const int nch = K >> 5;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int j = (int)(i & 15), lane = (int)((i >> 4) & 31);
    const size_t tc = i >> 9;
    const int chunk = (int)(tc % nch), tile = (int)(tc / nch);
    const int gid = lane >> 2, tig = lane & 3, mma = j >> 3, reg = (j >> 1) & 3, half = j & 1;
    const int row = 16 * tile + gid + 8 * (reg & 1), kk = 32 * chunk + 8 * tig + 4 * mma + 2 * (reg >> 1) + half;
    dst[i] = __float2bfloat16_rn(src[(size_t)row * stride_r + (size_t)kk * stride_k]);
  }
}
}  // namespace

extern "C" int dv3_pack_mma_a(int R, int K, const float* src, long long stride_r, long long stride_k, void* dst_bf16, void* stream) {
  if (R <= 0 || K <= 0 || !src || !dst_bf16) return DV3_ERR_ARG;
  if ((R & 15) || (K & 31)) return DV3_ERR_UNSUPPORTED;
  size_t g = ((size_t)R * K + 255) / 256;
  if (g > 148 * 16) g = 148 * 16;
  pack_mma_a_kernel<<<(int)g, 256, 0, (cudaStream_t)stream>>>(R, K, src, stride_r, stride_k, reinterpret_cast<__nv_bfloat16*>(dst_bf16));
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
