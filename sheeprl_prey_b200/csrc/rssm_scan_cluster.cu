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
#include "common.cuh"
#include "dv3_b200.h"
#include <cooperative_groups.h>
#include <cuda_bf16.h>
#include <stdio.h>
#include <stdlib.h>
namespace cg = cooperative_groups;

namespace {

constexpr int MB = 16;          // batch rows (B <= 16) = the two n8 tiles of the MMA
constexpr int CT = 512;         // threads per CTA
constexpr int CW = CT / 32;     // warps per CTA (== MB: warp m reduces batch row m)
constexpr int RP = 3;           // weight row tiles (16 rows each) per contraction pass
constexpr int OB = MB + 1;      // row stride of the own-output buffer (odd: conflict-free by row and by column)

struct KArgs {
  Dv3ScanFwdArgs a;
  const __nv_bfloat16* Wgru16; int ld_wgru16;
  const __nv_bfloat16* Wr0_16; int ld_wr0_16;
  const __nv_bfloat16* Wr1_16; int ld_wr1_16;
  const float* WmlpzT;          // [S, U] fp32 = Wmlp[:, :S]^T
  const float* aW;              // [T*B, U] fp32 = masked action . Wmlp[:, S:]^T
  long long* dbg;
};

__device__ __forceinline__ void mma_bf16(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint4 ldg_nc16(const __nv_bfloat16* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

// One pass of a contraction: `ntiles` (<= RP) tiles of 16 weight rows times X[MB][K] (bf16, shared memory) -> per-warp partial
// sums part[warp][tile][16 rows][MB].  K is split over the warps in 32-wide chunks.  Within a chunk thread (gid, tig) loads the 8
// consecutive k of its two weight rows and its two batch rows (16 B each); the contraction index is permuted consistently
// between the A and B fragments (sums are order-independent), so no shuffles / ldmatrix are needed.
__device__ __forceinline__ void gemv_pass(const __nv_bfloat16* __restrict__ W, int ldw, const int (&rbase)[RP], const int (&rvalid)[RP],
                                          int ntiles, int K, const __nv_bfloat16* X, int ldx, float* part) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, gid = lane >> 2, tig = lane & 3;
  float acc[RP][2][4];
  const __nv_bfloat16* w0[RP];
  const __nv_bfloat16* w1[RP];
#pragma unroll
  for (int i = 0; i < RP; ++i) {
#pragma unroll
    for (int j = 0; j < 2; ++j)
#pragma unroll
      for (int q = 0; q < 4; ++q) acc[i][j][q] = 0.f;
    const int ii = i < ntiles ? i : ntiles - 1;
    w0[i] = W + (size_t)(rbase[ii] + min(gid, rvalid[ii] - 1)) * ldw + tig * 8;
    w1[i] = W + (size_t)(rbase[ii] + min(gid + 8, rvalid[ii] - 1)) * ldw + tig * 8;
  }
  const __nv_bfloat16* xa_p = X + gid * ldx + tig * 8;
  const __nv_bfloat16* xb_p = X + (gid + 8) * ldx + tig * 8;
  const int nch = K >> 5;
#pragma unroll 2
  for (int ch = warp; ch < nch; ch += CW) {
    const int k0 = ch << 5;
    uint4 wa[RP], wb[RP];
#pragma unroll
    for (int i = 0; i < RP; ++i) { wa[i] = ldg_nc16(w0[i] + k0); wb[i] = ldg_nc16(w1[i] + k0); }
    const uint4 xa = *reinterpret_cast<const uint4*>(xa_p + k0);
    const uint4 xb = *reinterpret_cast<const uint4*>(xb_p + k0);
#pragma unroll
    for (int i = 0; i < RP; ++i) {
      mma_bf16(acc[i][0], wa[i].x, wb[i].x, wa[i].y, wb[i].y, xa.x, xa.y);
      mma_bf16(acc[i][0], wa[i].z, wb[i].z, wa[i].w, wb[i].w, xa.z, xa.w);
      mma_bf16(acc[i][1], wa[i].x, wb[i].x, wa[i].y, wb[i].y, xb.x, xb.y);
      mma_bf16(acc[i][1], wa[i].z, wb[i].z, wa[i].w, wb[i].w, xb.z, xb.w);
    }
  }
#pragma unroll
  for (int i = 0; i < RP; ++i) {
    if (i < ntiles) {
      float* p = part + (size_t)((warp * RP + i) * 16) * MB;
      *reinterpret_cast<float2*>(p + gid * MB + 2 * tig) = make_float2(acc[i][0][0], acc[i][0][1]);
      *reinterpret_cast<float2*>(p + (gid + 8) * MB + 2 * tig) = make_float2(acc[i][0][2], acc[i][0][3]);
      *reinterpret_cast<float2*>(p + gid * MB + 8 + 2 * tig) = make_float2(acc[i][1][0], acc[i][1][1]);
      *reinterpret_cast<float2*>(p + (gid + 8) * MB + 8 + 2 * tig) = make_float2(acc[i][1][2], acc[i][1][3]);
    }
  }
}

// outb[(tile0 + ti) * 16 + r][m] = sum over warps of the partials (fixed order -> deterministic)
__device__ __forceinline__ void reduce_pass(const float* part, int ntiles, int tile0, float* outb) {
  for (int i = threadIdx.x; i < ntiles * 256; i += CT) {
    const int ti = i >> 8, e = i & 255;
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < CW; ++w) s += part[(size_t)(w * RP + ti) * 256 + e];
    outb[(size_t)((tile0 + ti) * 16 + (e >> 4)) * OB + (e & 15)] = s;
  }
}

// Full contraction of `nblk` row blocks (block b = rows blk_base(b) + [0, rows_per_blk), padded to whole 16-row tiles) into
// outb[(b * tpb + tile) * 16 + r][m].
template <typename F>
__device__ __forceinline__ void cta_contract(const __nv_bfloat16* W, int ldw, int nblk, int rows_per_blk, F blk_base, int K,
                                             const __nv_bfloat16* X, int ldx, float* part, float* outb) {
  const int tpb = (rows_per_blk + 15) >> 4, nt = nblk * tpb;
  for (int t0 = 0; t0 < nt; t0 += RP) {
    int rbase[RP], rvalid[RP];
    const int n = min(RP, nt - t0);
#pragma unroll
    for (int i = 0; i < RP; ++i) {
      const int ti = min(t0 + i, nt - 1), b = ti / tpb, sub = ti - b * tpb;
      rbase[i] = blk_base(b) + sub * 16;
      rvalid[i] = min(16, rows_per_blk - sub * 16);
    }
    gemv_pass(W, ldw, rbase, rvalid, n, K, X, ldx, part);
    __syncthreads();
    reduce_pass(part, n, t0, outb);
    __syncthreads();
  }
}

#define SDBG(p) do { if (k.dbg && c == 0 && tid == 0 && t == 1) k.dbg[p] = clock64(); } while (0)

template <int CL>
__global__ void __launch_bounds__(CT, 1) rssm_observe_fwd_cluster_kernel(KArgs k) {
  cg::cluster_group cluster = cg::this_cluster();
  const Dv3ScanFwdArgs& a = k.a;
  extern __shared__ __align__(16) uint8_t smraw[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int c = (int)cluster.block_rank();
  const int B = a.B, G = a.G, Kc = a.Kc, D = a.D, U = a.U, Hh = a.Hh;
  const int S = G * Kc, KG = D + U;
  const int uper = U / CL, dper = D / CL, hper = Hh / CL, gper = G / CL, sper = gper * Kc;
  const int u0 = c * uper, d0 = c * dper, h0 = c * hper, s0 = c * sper;
  const int dpad = ((dper + 15) >> 4) << 4;
  const int ldxg = KG + 32, ldh = Hh + 32;                 // +32 bf16: the two 8-row halves of a fragment load hit different banks

  // ---- shared memory carve-up (16 B aligned regions)
  uint8_t* sp = smraw;
  auto take = [&](size_t bytes) { uint8_t* p = sp; sp += (bytes + 15) & ~(size_t)15; return p; };
  __nv_bfloat16* xg = reinterpret_cast<__nv_bfloat16*>(take((size_t)MB * ldxg * 2));   // [MB][h | feat] bf16 (MMA operand)
  __nv_bfloat16* hid = reinterpret_cast<__nv_bfloat16*>(take((size_t)MB * ldh * 2));   // [MB][Hh] bf16
  const int rows_max = max(max(3 * dpad, ((uper + 15) >> 4) << 4), max(((hper + 15) >> 4) << 4, ((sper + 15) >> 4) << 4));
  float* outb = reinterpret_cast<float*>(take((size_t)rows_max * OB * 4));             // own outputs [row][OB]
  float* part = reinterpret_cast<float*>(take((size_t)CW * RP * 256 * 4));             // per-warp partials of one pass
  float* hown = reinterpret_cast<float*>(take((size_t)MB * dper * 4));                 // own fp32 slice of h
  const int stg_cols = max(max(uper, dper), hper);
  __nv_bfloat16* stage = reinterpret_cast<__nv_bfloat16*>(take((size_t)MB * stg_cols * 2));
  float* lnp = reinterpret_cast<float*>(take((size_t)CL * MB * 2 * 4));                // [CL][MB](mean, M2) pushed by the peers
  float* s_stat = reinterpret_cast<float*>(take(2 * MB * 4));
  int* idxf = reinterpret_cast<int*>(take((size_t)MB * G * 4));                        // z indices [MB][G] (all groups)
  int* s_reset = reinterpret_cast<int*>(take(MB * 4));

  for (int i = tid; i < MB * ldxg / 2; i += CT) reinterpret_cast<uint32_t*>(xg)[i] = 0u;
  for (int i = tid; i < MB * ldh / 2; i += CT) reinterpret_cast<uint32_t*>(hid)[i] = 0u;
  for (int i = tid; i < MB * dper; i += CT) hown[i] = 0.f;
  for (int i = tid; i < MB * G; i += CT) idxf[i] = 0;
  __syncthreads();
  cluster.sync();          // every CTA's buffers exist before any peer pushes into them

  // LayerNorm statistics of a vector spread over the cluster: val(j, m), j < n own values of batch row m.
  auto ln_push = [&](auto val, int n) {
    const int m = warp;                                   // CW == MB
    float s = 0.f;
    for (int j = lane; j < n; j += 32) s += val(j, m);
    const float mu = warp_sum(s) / (float)n;
    float q = 0.f;
    for (int j = lane; j < n; j += 32) { const float dlt = val(j, m) - mu; q = fmaf(dlt, dlt, q); }
    q = warp_sum(q);
    if (lane < CL) {
      float2* dst = reinterpret_cast<float2*>(cluster.map_shared_rank(lnp, lane)) + (c * MB + m);
      *dst = make_float2(mu, q);
    }
  };
  auto ln_combine = [&](int n_per_cta, float eps, float* gmean, float* grstd, int row0) {   // after cluster.sync()
    if (tid < MB) {
      float mu = 0.f;
#pragma unroll
      for (int cc = 0; cc < CL; ++cc) mu += lnp[(cc * MB + tid) * 2];
      mu /= (float)CL;
      float m2 = 0.f;
#pragma unroll
      for (int cc = 0; cc < CL; ++cc) { const float dm = lnp[(cc * MB + tid) * 2] - mu; m2 += lnp[(cc * MB + tid) * 2 + 1] + (float)n_per_cta * dm * dm; }
      const float rs = rsqrtf(m2 / (float)(n_per_cta * CL) + eps);
      s_stat[tid] = mu;
      s_stat[MB + tid] = rs;
      if (c == 0 && tid < B) { gmean[row0 + tid] = mu; grstd[row0 + tid] = rs; }
    }
    __syncthreads();
  };
  // push the own bf16 slice stage[MB][ncols] into column col0 of `buf` (row stride ld) of every CTA of the cluster
  auto allgather = [&](__nv_bfloat16* buf, int ld, int col0, int ncols) {
    const int cpr = ncols >> 3, per_peer = MB * cpr;
    for (int i = tid; i < CL * per_peer; i += CT) {
      const int peer = i / per_peer, rem = i - peer * per_peer, m = rem / cpr, ch = rem - m * cpr;
      __nv_bfloat16* dst = cluster.map_shared_rank(buf, peer) + m * ld + col0 + ch * 8;
      *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(stage + m * stg_cols + ch * 8);
    }
  };

  for (int t = 0; t < a.T; ++t) {
    const int row0 = t * B;
    SDBG(0);
    // ================= P0: reset masking (agent.py:379-384), saved inputs of the step =================
    if (tid < MB) s_reset[tid] = (tid >= B || t == 0 || __ldg(a.first + row0 + tid) != 0.f) ? 1 : 0;
    __syncthreads();
    for (int i = tid; i < MB * G; i += CT) { const int m = i / G; if (s_reset[m]) idxf[i] = __ldg(a.zreset_idx + (i - m * G)); }
    for (int i = tid; i < MB * (D >> 3); i += CT) {
      const int m = i / (D >> 3), ch = i - m * (D >> 3);
      if (s_reset[m]) *reinterpret_cast<uint4*>(xg + m * ldxg + ch * 8) = make_uint4(0u, 0u, 0u, 0u);
    }
    for (int i = tid; i < MB * dper; i += CT) if (s_reset[i / dper]) hown[i] = 0.f;
    __syncthreads();
    for (int i = tid; i < B * sper; i += CT) {
      const int m = i / sper, sl = i - m * sper, s = s0 + sl, g = s / Kc;
      a.x_mlp[(size_t)(row0 + m) * a.ld_xmlp + s] = (idxf[m * G + g] == s - g * Kc) ? 1.f : 0.f;
    }
    for (int i = tid; i < B * dper; i += CT) { const int m = i / dper, dl = i - m * dper; a.x_gru[(size_t)(row0 + m) * a.ld_xgru + d0 + dl] = hown[m * dper + dl]; }
    SDBG(1);

    // ================= P1: feat = SiLU(LN(Wmlp . [z, a])) -- one-hot gather + hoisted action projection =================
    for (int i = tid; i < MB * uper; i += CT) {
      const int m = i / uper, ul = i - m * uper;
      float acc = (m < B) ? __ldg(k.aW + (size_t)(row0 + m) * U + u0 + ul) : 0.f;
      const int* ix = idxf + m * G;
#pragma unroll 8
      for (int g = 0; g < G; ++g) acc += __ldg(k.WmlpzT + (size_t)(g * Kc + ix[g]) * U + u0 + ul);
      outb[ul * OB + m] = acc;
      if (m < B) a.feat_pre[(size_t)(row0 + m) * U + u0 + ul] = acc;
    }
    __syncthreads();
    ln_push([&](int j, int m) { return outb[j * OB + m]; }, uper);
    SDBG(2);
    cluster.sync();                                                                   // #1
    ln_combine(uper, 1e-3f, a.feat_mean, a.feat_rstd, row0);
    for (int i = tid; i < MB * uper; i += CT) {
      const int m = i / uper, ul = i - m * uper, u = u0 + ul;
      const float x = (outb[ul * OB + m] - s_stat[m]) * s_stat[MB + m] * __ldg(a.g_mlp + u) + __ldg(a.b_mlp + u);
      const float f = siluf_(x);
      if (m < B) a.x_gru[(size_t)(row0 + m) * a.ld_xgru + D + u] = f;
      stage[m * stg_cols + ul] = __float2bfloat16_rn(f);
    }
    __syncthreads();
    allgather(xg, ldxg, D + u0, uper);
    SDBG(3);
    cluster.sync();                                                                   // #2

    // ================= P2: GRU gates: own rows {d, D+d, 2D+d} of Wgru . [h, feat] =================
    cta_contract(k.Wgru16, k.ld_wgru16, 3, dper, [&](int gate) { return gate * D + d0; }, KG, xg, ldxg, part, outb);
    SDBG(4);
    for (int i = tid; i < 3 * dper * B; i += CT) {
      const int rl = i / B, m = i - rl * B, gate = rl / dper, dl = rl - gate * dper;
      a.gates_pre[(size_t)(row0 + m) * 3 * D + gate * D + d0 + dl] = outb[(gate * dpad + dl) * OB + m];
    }
    ln_push([&](int j, int m) { const int gate = j / dper, dl = j - gate * dper; return outb[(gate * dpad + dl) * OB + m]; }, 3 * dper);
    cluster.sync();                                                                   // #3
    ln_combine(3 * dper, 1e-5f, a.gru_mean, a.gru_rstd, row0);
    for (int i = tid; i < MB * dper; i += CT) {
      const int m = i / dper, dl = i - m * dper, d = d0 + dl;
      const float mu = s_stat[m], rs = s_stat[MB + m];
      const float r = sigmoidf_((outb[(0 * dpad + dl) * OB + m] - mu) * rs * __ldg(a.g_gru + d) + __ldg(a.b_gru + d));
      const float cp = (outb[(1 * dpad + dl) * OB + m] - mu) * rs * __ldg(a.g_gru + D + d) + __ldg(a.b_gru + D + d);
      const float cd = tanhf(r * cp);
      const float u = sigmoidf_((outb[(2 * dpad + dl) * OB + m] - mu) * rs * __ldg(a.g_gru + 2 * D + d) + __ldg(a.b_gru + 2 * D + d) - 1.f);
      const float hn = u * cd + (1.f - u) * hown[m * dper + dl];
      hown[m * dper + dl] = hn;
      if (m < B) a.x_rep[(size_t)(row0 + m) * a.ld_xrep + d] = hn;
      stage[m * stg_cols + dl] = __float2bfloat16_rn(hn);
    }
    __syncthreads();
    allgather(xg, ldxg, d0, dper);
    SDBG(5);
    cluster.sync();                                                                   // #4

    // ================= P3: posterior hidden: own rows of Wr0[:, :D] . h' + (e_t . Wr0[:, D:]^T + b) =================
    cta_contract(k.Wr0_16, k.ld_wr0_16, 1, hper, [&](int) { return h0; }, D, xg, ldxg, part, outb);
    for (int i = tid; i < MB * hper; i += CT) {
      const int m = i / hper, hl = i - m * hper;
      float v = outb[hl * OB + m];
      if (m < B) { v += __ldg(a.eW + (size_t)(row0 + m) * Hh + h0 + hl); a.hidr_pre[(size_t)(row0 + m) * Hh + h0 + hl] = v; }
      outb[hl * OB + m] = v;
    }
    __syncthreads();
    ln_push([&](int j, int m) { return outb[j * OB + m]; }, hper);
    SDBG(6);
    cluster.sync();                                                                   // #5
    ln_combine(hper, 1e-5f, a.hidr_mean, a.hidr_rstd, row0);
    for (int i = tid; i < MB * hper; i += CT) {
      const int m = i / hper, hl = i - m * hper, j = h0 + hl;
      const float x = (outb[hl * OB + m] - s_stat[m]) * s_stat[MB + m] * __ldg(a.g_r + j) + __ldg(a.b_r + j);
      const float f = siluf_(x);
      if (m < B) a.hidr[(size_t)(row0 + m) * Hh + j] = f;
      stage[m * stg_cols + hl] = __float2bfloat16_rn(f);
    }
    __syncthreads();
    allgather(hid, ldh, h0, hper);
    SDBG(7);
    cluster.sync();                                                                   // #6

    // ================= P4: posterior logits of the own categoricals, unimix, inverse-CDF sample =================
    cta_contract(k.Wr1_16, k.ld_wr1_16, 1, sper, [&](int) { return s0; }, Hh, hid, ldh, part, outb);
    SDBG(8);
    for (int p = warp; p < B * gper; p += CW) {          // one warp per (batch row, own categorical)
      const int m = p / gper, gl = p - m * gper, g = c * gper + gl;
      const size_t orow = (size_t)(row0 + m);
      float* rawp = outb + (size_t)(gl * Kc) * OB + m;   // class k at rawp[k * OB]
      float mx = -INFINITY;
      for (int kk = lane; kk < Kc; kk += 32) {
        const float raw = rawp[kk * OB] + __ldg(a.br1 + g * Kc + kk);
        rawp[kk * OB] = raw;
        a.post_raw[orow * S + g * Kc + kk] = raw;
        mx = fmaxf(mx, raw);
      }
      mx = warp_max(mx);
      float se = 0.f;
      for (int kk = lane; kk < Kc; kk += 32) se += expf(rawp[kk * OB] - mx);
      se = warp_sum(se);
      float mx2 = -INFINITY;
      for (int kk = lane; kk < Kc; kk += 32) {
        const float raw = rawp[kk * OB];
        float lu = raw;
        if (a.unimix > 0.f) {
          const float pr = (1.f - a.unimix) * (expf(raw - mx) / se) + a.unimix / (float)Kc;
          lu = logf(fminf(fmaxf(pr, DV3_EPS32), 1.f - DV3_EPS32));
        }
        a.post_logits[orow * S + g * Kc + kk] = lu;
        rawp[kk * OB] = lu;
        mx2 = fmaxf(mx2, lu);
      }
      mx2 = warp_max(mx2);
      float se2 = 0.f;
      for (int kk = lane; kk < Kc; kk += 32) se2 += expf(rawp[kk * OB] - mx2);
      se2 = warp_sum(se2);
      const float lse = mx2 + logf(se2);
      float mx3 = -INFINITY;                              // probs = softmax(lu - lse), as torch's Categorical(logits=lu).probs
      for (int kk = lane; kk < Kc; kk += 32) mx3 = fmaxf(mx3, rawp[kk * OB] - lse);
      mx3 = warp_max(mx3);
      float se3 = 0.f;
      for (int kk = lane; kk < Kc; kk += 32) se3 += expf(rawp[kk * OB] - lse - mx3);
      se3 = warp_sum(se3);
      __syncwarp();
      for (int kk = lane; kk < Kc; kk += 32) rawp[kk * OB] = expf(rawp[kk * OB] - lse - mx3) / se3;
      __syncwarp();
      int choice = 0;
      if (a.forced_idx) {
        choice = __ldg(a.forced_idx + orow * G + g);
      } else if (lane == 0) {                             // idx = #{k : cumsum_k(p) <= u}, sequential fp32 cumsum (the sampling contract)
        const float uu = __ldg(a.u_post + orow * G + g);
        float cs = 0.f;
#pragma unroll 8
        for (int kk = 0; kk < Kc; ++kk) { cs += rawp[kk * OB]; choice += (cs <= uu) ? 1 : 0; }
        if (choice > Kc - 1) choice = Kc - 1;
      }
      choice = __shfl_sync(0xffffffffu, choice, 0);
      if (lane == 0) a.post_idx[orow * G + g] = choice;
      for (int kk = lane; kk < Kc; kk += 32) a.latent[orow * a.ld_lat + g * Kc + kk] = (kk == choice) ? 1.f : 0.f;
      if (lane < CL) cluster.map_shared_rank(idxf, lane)[m * G + g] = choice;
    }
    SDBG(9);
    cluster.sync();                                                                   // #7
    SDBG(10);
  }
}

template <int CL>
int launch_cluster(const KArgs& k, size_t smem_bytes, cudaStream_t stream, bool probe_only) {
  auto kern = rssm_observe_fwd_cluster_kernel<CL>;
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
  e = cudaLaunchKernelEx(&cfg, kern, k);
  if (e != cudaSuccess) return (int)e;
  return DV3_OK;
}

template <int CL>
bool shape_ok(const Dv3ScanFwdArgs* h, size_t* smem_bytes, int max_smem) {
  if (h->U % (8 * CL) || h->D % (8 * CL) || h->Hh % (8 * CL) || h->G % CL) return false;
  const int KG = h->D + h->U, uper = h->U / CL, dper = h->D / CL, hper = h->Hh / CL, sper = h->G / CL * h->Kc;
  auto r16 = [](size_t b) { return (b + 15) & ~(size_t)15; };
  auto pad16 = [](int n) { return ((n + 15) >> 4) << 4; };
  int rows_max = 3 * pad16(dper);
  if (pad16(uper) > rows_max) rows_max = pad16(uper);
  if (pad16(hper) > rows_max) rows_max = pad16(hper);
  if (pad16(sper) > rows_max) rows_max = pad16(sper);
  int stg = uper > dper ? uper : dper;
  if (hper > stg) stg = hper;
  size_t b = r16((size_t)MB * (KG + 32) * 2) + r16((size_t)MB * (h->Hh + 32) * 2) + r16((size_t)rows_max * OB * 4) +
             r16((size_t)CW * RP * 256 * 4) + r16((size_t)MB * dper * 4) + r16((size_t)MB * stg * 2) + r16((size_t)CL * MB * 2 * 4) +
             r16(2 * MB * 4) + r16((size_t)MB * h->G * 4) + r16(MB * 4);
  *smem_bytes = b + 16;
  return (long long)*smem_bytes <= (long long)max_smem;
}

}  // namespace

extern "C" int dv3_rssm_observe_fwd_cluster(const Dv3ScanFwdArgs* h, const Dv3ScanClusterExtra* x, void* stream_) {
  if (!h || !x) return DV3_ERR_ARG;
  if (h->B < 1 || h->B > MB || h->T < 1 || h->G < 1 || h->Kc < 1) return DV3_ERR_UNSUPPORTED;
  if (!x->Wgru_bf16 || !x->Wr0_bf16 || !x->Wr1_bf16 || !x->WmlpzT || !x->aW) return DV3_ERR_UNSUPPORTED;
  const int KG = h->D + h->U;
  if ((KG & 31) || (h->D & 31) || (h->Hh & 31) || (x->ld_wgru & 7) || (x->ld_wr0 & 7) || (x->ld_wr1 & 7) ||
      ((uintptr_t)x->Wgru_bf16 & 15) || ((uintptr_t)x->Wr0_bf16 & 15) || ((uintptr_t)x->Wr1_bf16 & 15))
    return DV3_ERR_UNSUPPORTED;   // 16 B fragment loads, 32-wide k chunks
  int dev = 0, max_smem = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  KArgs k;
  k.a = *h;
  k.Wgru16 = reinterpret_cast<const __nv_bfloat16*>(x->Wgru_bf16); k.ld_wgru16 = x->ld_wgru;
  k.Wr0_16 = reinterpret_cast<const __nv_bfloat16*>(x->Wr0_bf16); k.ld_wr0_16 = x->ld_wr0;
  k.Wr1_16 = reinterpret_cast<const __nv_bfloat16*>(x->Wr1_bf16); k.ld_wr1_16 = x->ld_wr1;
  k.WmlpzT = x->WmlpzT;
  k.aW = x->aW;
  k.dbg = (getenv("DV3_SCAN_DEBUG") && h->scratch && h->scratch_bytes >= 64 * sizeof(long long)) ? reinterpret_cast<long long*>(h->scratch) : nullptr;
  static int cl_choice = -1;       // 16, 8, or 0 (= cluster launch unavailable)
  const char* force = getenv("DV3_SCAN_CLUSTER");
  size_t smem = 0;
  if (cl_choice < 0 || force) {
    cl_choice = 0;
    const int want = force ? atoi(force) : 16;
    if (want >= 16 && shape_ok<16>(h, &smem, max_smem) && launch_cluster<16>(k, smem, (cudaStream_t)stream_, true) == DV3_OK) cl_choice = 16;
    else if (want >= 8 && shape_ok<8>(h, &smem, max_smem) && launch_cluster<8>(k, smem, (cudaStream_t)stream_, true) == DV3_OK) cl_choice = 8;
  }
  if (cl_choice == 16 && shape_ok<16>(h, &smem, max_smem)) return launch_cluster<16>(k, smem, (cudaStream_t)stream_, false);
  if (cl_choice >= 8 && shape_ok<8>(h, &smem, max_smem)) return launch_cluster<8>(k, smem, (cudaStream_t)stream_, false);
  return DV3_ERR_UNSUPPORTED;
}
