// Persistent RSSM observe scan (forward): the T sequential steps of RSSM.dynamic run inside ONE cooperative kernel.
//
// Reference: the Python loop sheeprl/algos/dreamer_v3/dreamer_v3.py:122-129 over RSSM.dynamic (agent.py:350-388):
// RecurrentModel (agent.py:296-309: Linear(no bias)->LN(1e-3)->SiLU, then LayerNormGRUCell models/models.py:388-395),
// _representation (agent.py:404-420: Linear->LN(1e-5)->SiLU->Linear, unimix, straight-through one-hot sample).
// (The prior/_transition head does not feed the recurrence and is evaluated for all T*B rows after the scan.)
//
// Design: grid = one CTA per SM (cooperative launch); each CTA owns a fixed slice of the output rows of every in-loop
// weight matrix and keeps that slice in shared memory for the whole scan (XS/S/M sizes: the in-loop weights fit in the
// 148 x 227 KB of aggregate shared memory; slices that do not fit are streamed from L2).  Per step there are six
// grid-wide barriers (one per dependent contraction / normalisation), and the activation vectors (B <= 16 rows) are
// exchanged through L2.  All LayerNorms, the GRU gates, unimix, and the inverse-CDF sampling are fused into the stages:
//   P1  feat_pre slice  = gather-sum of W_mlp columns selected by the one-hot z  +  W_mlp[:, S:] . a_t
//   P2  feat = SiLU(LN(feat_pre)) (every CTA, redundantly);  own GRU gate rows {d, D+d, 2D+d} = W_gru . [h, feat];
//       per-CTA (mean, M2) of the own gate values (Chan's parallel variance)
//   P3  LayerNorm statistics of the 3D gate vector from the per-CTA partials; gates -> h'[own d]
//   P4  hidr_pre slice  = W_r0[:, :D] . h' + (e_t . W_r0[:, D:]^T + b)   (the embedding part is hoisted out of the scan)
//   P5a hid = SiLU(LN(hidr_pre)) (every CTA, redundantly); own posterior logit rows = W_r1 rows . hid + b  (balanced over all CTAs)
//   P5b CTAs 0..G-1 own one categorical each: unimix; sample z_t by inverse CDF of the injected uniform; write unimix logits,
//       index, one-hot.   (six grid-wide barriers per step in total)
// This is the fp32-precision kernel; in bf16 precision the step uses the cluster kernel of rssm_scan_cluster.cu instead.
// Everything the hand-written backward needs (pre-LN values, LN statistics, inputs of every contraction) is written to the
// same buffers the per-op path fills.
#include "common.cuh"
#include <cooperative_groups.h>
#include <stdlib.h>
#include <stdio.h>
namespace cg = cooperative_groups;

namespace {

constexpr int MB = 16;             // batch rows handled (B <= 16)
constexpr int ST = 256;            // threads per CTA
constexpr int NW = ST / 32;

struct ScanFwdArgs {
  int T, B, G, Kc, D, U, Hh, A, E;
  float unimix;
  // weights (fp32, row-major [out, in])
  const float* Wmlp; int ld_wmlp;      // [U, S+A]
  const float* g_mlp; const float* b_mlp;
  const float* Wgru; int ld_wgru;      // [3D, D+U]
  const float* g_gru; const float* b_gru;
  const float* Wr0; int ld_wr0;        // [Hh, D+E] (only the first D columns are used here)
  const float* g_r; const float* b_r;
  const float* Wr1; int ld_wr1;        // [S, Hh]
  const float* br1;
  // inputs
  const float* eW;                     // [T*B, Hh] = e_t . W_r0[:, D:]^T + b_r0
  const float* first;                  // [T*B]
  const float* u_post;                 // [T*B, G]
  const int* forced_idx;               // [T*B, G] or null
  const int* zreset_idx;               // [G]
  // outputs / saved activations
  float* x_mlp; int ld_xmlp;           // [T*B, S+A]: z part written here, action part is input
  float* feat_pre; float* feat_mean; float* feat_rstd;        // [T*B, U], [T*B]
  float* x_gru; int ld_xgru;           // [T*B, D+U] = [h_in, feat]
  float* gates_pre; float* gru_mean; float* gru_rstd;         // [T*B, 3D]
  float* x_rep; int ld_xrep;           // [T*B, D+E]: h' written to the first D columns
  float* hidr_pre; float* hidr; float* hidr_mean; float* hidr_rstd;   // [T*B, Hh]
  float* post_raw; float* post_logits; int* post_idx;         // [T*B, S], [T*B, S], [T*B, G]
  float* latent; int ld_lat;           // [T*B, S+D]: one-hot z written to the first S columns
  float* lnpart;                       // [gridDim.x][MB][2] scratch (per-CTA mean, M2 of the own gate values)
  long long* dbg;                      // optional: clock64 stamps of CTA 0 during step t == 1
  // shared-memory plan (floats): which weight slices are resident
  int smem_wmlp, smem_wgru, smem_wr0, smem_wr1;
};

__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }

// v[i] (i < 32) per lane -> lane l returns the warp-wide sum of v[l]  (31 shuffles instead of 32 x 5)
__device__ __forceinline__ float warp_transpose_reduce32(float (&v)[32], int lane) {
#pragma unroll
  for (int s = 16; s >= 1; s >>= 1) {
    const bool up = (lane & s) != 0;
#pragma unroll
    for (int j = 0; j < s; ++j) {
      const float send = up ? v[j] : v[j + s];
      const float keep = up ? v[j + s] : v[j];
      v[j] = keep + __shfl_xor_sync(0xffffffffu, send, s);
    }
  }
  return v[0];
}

// rows [0, R) of a weight slice (row stride ldw, 16 B aligned rows, K % 4 == 0) times act[MB][K] (shared memory):
// one warp per pair of rows, lanes over k (float4), 32 accumulators per lane, transposed reduction; out(r, m, value).
template <typename F>
__device__ __forceinline__ void cta_gemv(const float* W, int ldw, int R, int K, const float* act, int ldact, F out) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int r0 = warp * 2; r0 < R; r0 += NW * 2) {
    const bool two = (r0 + 1 < R);
    const float* w0 = W + (size_t)r0 * ldw;
    const float* w1 = W + (size_t)(two ? r0 + 1 : r0) * ldw;
    float acc[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) acc[i] = 0.f;
#pragma unroll 2
    for (int k = lane * 4; k < K; k += 128) {
      const float4 wa = ld4(w0 + k), wb = ld4(w1 + k);
#pragma unroll
      for (int m = 0; m < MB; ++m) {
        const float4 x = ld4(act + m * ldact + k);
        acc[m] = fmaf(x.x, wa.x, acc[m]); acc[m] = fmaf(x.y, wa.y, acc[m]);
        acc[m] = fmaf(x.z, wa.z, acc[m]); acc[m] = fmaf(x.w, wa.w, acc[m]);
        acc[MB + m] = fmaf(x.x, wb.x, acc[MB + m]); acc[MB + m] = fmaf(x.y, wb.y, acc[MB + m]);
        acc[MB + m] = fmaf(x.z, wb.z, acc[MB + m]); acc[MB + m] = fmaf(x.w, wb.w, acc[MB + m]);
      }
    }
    const float v = warp_transpose_reduce32(acc, lane);
    const int r = r0 + (lane >> 4);
    if (r < R) out(r, lane & 15, v);
  }
}

// smem dst[m][0..cols) (row stride ldd) <- global src[m][0..cols) (row stride lds) for m < rows, zero rows up to MB.
// cols % 4 == 0 and 16 B aligned rows; every thread keeps 8 independent 16 B L2 loads in flight (ld.global.cg: the data
// was written by other CTAs during this kernel, L1 must be bypassed).
__device__ __forceinline__ void load_rows_cg(float* dst, int ldd, const float* src, int lds, int rows, int cols) {
  const int c4 = cols >> 2, n4 = MB * c4;
  for (int base = 0; base < n4; base += ST * 8) {
    float4 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int i = base + threadIdx.x + j * ST;
      v[j] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (i < n4) {
        const int m = i / c4, k = (i - m * c4) << 2;
        if (m < rows) v[j] = __ldcg(reinterpret_cast<const float4*>(src + (size_t)m * lds + k));
      }
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      const int i = base + threadIdx.x + j * ST;
      if (i < n4) {
        const int m = i / c4, k = (i - m * c4) << 2;
        *reinterpret_cast<float4*>(dst + m * ldd + k) = v[j];
      }
    }
  }
}

// Row-wise LayerNorm + SiLU of buf[MB][U] in shared memory, in place: 16 threads per row, 4-way unrolled passes.
__device__ __forceinline__ void smem_ln_silu(float* buf, int ld, int U, const float* __restrict__ gamma,
                                             const float* __restrict__ beta, float eps, float* s_mean, float* s_rstd) {
  const int m = threadIdx.x >> 4, sub = threadIdx.x & 15;
  float* row = buf + m * ld;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  int i = sub;
  for (; i + 48 < U; i += 64) { s0 += row[i]; s1 += row[i + 16]; s2 += row[i + 32]; s3 += row[i + 48]; }
  for (; i < U; i += 16) s0 += row[i];
  float s = (s0 + s1) + (s2 + s3);
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mu = s / (float)U;
  s0 = s1 = s2 = s3 = 0.f;
  i = sub;
  for (; i + 48 < U; i += 64) {
    const float a0 = row[i] - mu, a1 = row[i + 16] - mu, a2 = row[i + 32] - mu, a3 = row[i + 48] - mu;
    s0 = fmaf(a0, a0, s0); s1 = fmaf(a1, a1, s1); s2 = fmaf(a2, a2, s2); s3 = fmaf(a3, a3, s3);
  }
  for (; i < U; i += 16) { const float a0 = row[i] - mu; s0 = fmaf(a0, a0, s0); }
  float v = (s0 + s1) + (s2 + s3);
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const float rs = rsqrtf(v / (float)U + eps);
  // SiLU through ex2.approx / rcp.approx: this loop runs redundantly in every CTA on all MB x U values, so it is
  // instruction-issue bound; the approximations are ~1e-7 absolute, far inside the fp32-mode parity tolerance.
#pragma unroll 4
  for (i = sub; i < U; i += 16) {
    const float x = (row[i] - mu) * rs * __ldg(gamma + i) + __ldg(beta + i);
    row[i] = __fdividef(x, 1.f + __expf(-x));
  }
  if (sub == 0) { s_mean[m] = mu; s_rstd[m] = rs; }
}

#define DBG(p) do { if (a.dbg && c == 0 && tid == 0 && t == 1) a.dbg[p] = clock64(); } while (0)

__global__ void __launch_bounds__(ST, 1) rssm_observe_fwd_kernel(ScanFwdArgs a) {
  cg::grid_group grid = cg::this_grid();
  extern __shared__ __align__(16) float smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int c = blockIdx.x, NC = gridDim.x;
  const int B = a.B, G = a.G, Kc = a.Kc, D = a.D, U = a.U, Hh = a.Hh, A = a.A;
  const int S = G * Kc, KA = S + A, KG = D + U;
  const int KAp = (KA + 3) & ~3, Ap = A + 4;

  // ---- static partition of the output rows of every in-loop contraction
  const int ua = (U + NC - 1) / NC, ua0 = min(c * ua, U), nA = min(ua0 + ua, U) - ua0;      // feat rows
  const int dd = (D + NC - 1) / NC, d0 = min(c * dd, D), nD = min(d0 + dd, D) - d0;         // recurrent units (3 gate rows each)
  const int hr = (Hh + NC - 1) / NC, hr0 = min(c * hr, Hh), nH = min(hr0 + hr, Hh) - hr0;   // posterior hidden rows
  const int sr = (S + NC - 1) / NC, sr0 = min(c * sr, S), nS = min(sr0 + sr, S) - sr0;      // posterior logit rows
  const bool owner = c < G;                                                                  // samples categorical c

  // ---- shared memory carve-up (floats; every region 16 B aligned)
  float* sp = smem;
  auto take = [&](int n) { float* p = sp; sp += (n + 3) & ~3; return p; };
  float* act = take(MB * KG);              // [MB][D+U]: h (first D) and feat / hid (from D on)
  float* acta = take(MB * Ap);             // [MB][A] masked previous action
  float* gown = take(3 * max(dd, 1) * MB); // own gate pre-activations [3*nD][MB]
  float* s_stat = take(4 * MB);            // mean / rstd scratch
  float* s_log = take(MB * Kc);            // owner: logits / probs [MB][Kc]
  int* s_idx = reinterpret_cast<int*>(take(MB * G));   // z indices of the previous step [MB][G]
  int* s_reset = reinterpret_cast<int*>(take(MB));
  float* lnp = take(NC * MB * 2);          // staging of every CTA's LN partials
  float* w_mlp_s = a.smem_wmlp ? take(max(ua, 1) * KAp) : nullptr;
  float* w_gru_s = a.smem_wgru ? take(3 * max(dd, 1) * KG) : nullptr;
  float* w_r0_s = a.smem_wr0 ? take(max(hr, 1) * D) : nullptr;
  float* w_r1_s = a.smem_wr1 ? take(max(sr, 1) * Hh) : nullptr;

  // ---- preload the resident weight slices
  if (w_mlp_s)
    for (int i = tid; i < nA * KAp; i += ST) { const int r = i / KAp, k = i % KAp; w_mlp_s[i] = (k < KA) ? __ldg(a.Wmlp + (size_t)(ua0 + r) * a.ld_wmlp + k) : 0.f; }
  if (w_gru_s)
    for (int i = tid; i < 3 * nD * KG; i += ST) {
      const int rl = i / KG, k = i % KG, gate = rl / nD, dl = rl % nD;
      w_gru_s[i] = __ldg(a.Wgru + (size_t)(gate * D + d0 + dl) * a.ld_wgru + k);
    }
  if (w_r0_s)
    for (int i = tid; i < nH * D; i += ST) { const int r = i / D, k = i % D; w_r0_s[i] = __ldg(a.Wr0 + (size_t)(hr0 + r) * a.ld_wr0 + k); }
  if (w_r1_s)
    for (int i = tid; i < nS * Hh; i += ST) { const int r = i / Hh, k = i % Hh; w_r1_s[i] = __ldg(a.Wr1 + (size_t)(sr0 + r) * a.ld_wr1 + k); }
  for (int i = tid; i < MB * KG; i += ST) act[i] = 0.f;
  for (int i = tid; i < MB * Ap; i += ST) acta[i] = 0.f;
  __syncthreads();

  for (int t = 0; t < a.T; ++t) {
    const int row0 = t * B;
    DBG(0);
    // ================= P0: reset masking, previous z indices, previous action =================
    {
      int pv[2] = {0, 0};
#pragma unroll
      for (int j = 0; j < 2; ++j) {            // G * MB <= 2 * ST is checked by the host wrapper
        const int i = tid + j * ST;
        if (i < MB * G && t > 0) { const int m = i / G, g = i - m * G; if (m < B) pv[j] = __ldcg(a.post_idx + (size_t)(row0 - B + m) * G + g); }
      }
      float av[8];
      const int nact = B * A;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int i = tid + j * ST;
        av[j] = 0.f;
        if (i < nact) { const int m = i / A, k = i - m * A; av[j] = __ldg(a.x_mlp + (size_t)(row0 + m) * a.ld_xmlp + S + k); }
      }
      if (tid < MB) s_reset[tid] = (tid >= B || t == 0 || __ldg(a.first + row0 + tid) != 0.f) ? 1 : 0;
      __syncthreads();
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int i = tid + j * ST;
        if (i < MB * G) { const int m = i / G, g = i - m * G; s_idx[i] = s_reset[m] ? __ldg(a.zreset_idx + g) : pv[j]; }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) { const int i = tid + j * ST; if (i < nact) { const int m = i / A, k = i - m * A; acta[m * Ap + k] = av[j]; } }
      for (int i = tid + 8 * ST; i < nact; i += ST) { const int m = i / A, k = i - m * A; acta[m * Ap + k] = __ldg(a.x_mlp + (size_t)(row0 + m) * a.ld_xmlp + S + k); }
      for (int m = 0; m < MB; ++m)
        if (s_reset[m]) for (int k = tid; k < D; k += ST) act[m * KG + k] = 0.f;   // else act[:, :D] holds h'_{t-1} (loaded in P4)
    }
    __syncthreads();
    DBG(11);
    // saved inputs of the step: dense one-hot z_in and h_in (column ranges spread over the CTAs)
    for (int i = c * ST + tid; i < B * S; i += NC * ST) {
      const int m = i / S, s = i - m * S, g = s / Kc;
      a.x_mlp[(size_t)(row0 + m) * a.ld_xmlp + s] = (s_idx[m * G + g] == (s - g * Kc)) ? 1.f : 0.f;
    }
    for (int i = c * ST + tid; i < B * D; i += NC * ST) { const int m = i / D, k = i - m * D; a.x_gru[(size_t)(row0 + m) * a.ld_xgru + k] = act[m * KG + k]; }
    DBG(12);

    // ================= P1: feat_pre slice (one-hot gather + action part), one warp per pair of rows ================
    for (int r0 = warp * 2; r0 < nA; r0 += NW * 2) {
      const bool two = r0 + 1 < nA;
      const float* wr0 = w_mlp_s ? (w_mlp_s + r0 * KAp) : (a.Wmlp + (size_t)(ua0 + r0) * a.ld_wmlp);
      const float* wr1 = two ? (w_mlp_s ? (w_mlp_s + (r0 + 1) * KAp) : (a.Wmlp + (size_t)(ua0 + r0 + 1) * a.ld_wmlp)) : wr0;
      float acc[32];
#pragma unroll
      for (int i = 0; i < 32; ++i) acc[i] = 0.f;
      for (int g = lane; g < G; g += 32) {
#pragma unroll
        for (int m = 0; m < MB; ++m) {
          const int col = g * Kc + s_idx[m * G + g];
          acc[m] += wr0[col];
          acc[MB + m] += wr1[col];
        }
      }
      for (int k = lane; k < A; k += 32) {
        const float wa = wr0[S + k], wb = wr1[S + k];
#pragma unroll
        for (int m = 0; m < MB; ++m) { const float x = acta[m * Ap + k]; acc[m] = fmaf(wa, x, acc[m]); acc[MB + m] = fmaf(wb, x, acc[MB + m]); }
      }
      const float v = warp_transpose_reduce32(acc, lane);
      const int r = r0 + (lane >> 4), m = lane & 15;
      if (r < nA && m < B) a.feat_pre[(size_t)(row0 + m) * U + ua0 + r] = v;
    }
    DBG(1);
    grid.sync();

    // ================= P2: feat = SiLU(LN(feat_pre)); own GRU gate rows =================
    load_rows_cg(act + D, KG, a.feat_pre + (size_t)row0 * U, U, B, U);
    __syncthreads();
    DBG(2);
    smem_ln_silu(act + D, KG, U, a.g_mlp, a.b_mlp, 1e-3f, s_stat, s_stat + MB);
    __syncthreads();
    DBG(13);
    if (c == 0 && tid < B) { a.feat_mean[row0 + tid] = s_stat[tid]; a.feat_rstd[row0 + tid] = s_stat[MB + tid]; }
    for (int i = c * ST + tid; i < B * U; i += NC * ST) { const int m = i / U, k = i - m * U; a.x_gru[(size_t)(row0 + m) * a.ld_xgru + D + k] = act[m * KG + D + k]; }
    DBG(14);
    if (w_gru_s) {
      cta_gemv(w_gru_s, KG, 3 * nD, KG, act, KG, [&](int rl, int m, float v) { gown[rl * MB + m] = v; });
    } else {
      for (int gate = 0; gate < 3; ++gate)
        cta_gemv(a.Wgru + (size_t)(gate * D + d0) * a.ld_wgru, a.ld_wgru, nD, KG, act, KG,
                 [&](int dl, int m, float v) { gown[(gate * nD + dl) * MB + m] = v; });
    }
    __syncthreads();
    DBG(15);
    for (int i = tid; i < 3 * nD * B; i += ST) {
      const int rl = i / B, m = i - rl * B, gate = rl / nD, dl = rl - gate * nD;
      a.gates_pre[(size_t)(row0 + m) * 3 * D + gate * D + d0 + dl] = gown[rl * MB + m];
    }
    if (tid < MB) {   // Chan partials (mean, M2) of the own 3*nD values of row m
      const int n = 3 * nD;
      float mu = 0.f, m2 = 0.f;
      if (n > 0) {
        for (int r = 0; r < n; ++r) mu += gown[r * MB + tid];
        mu /= (float)n;
        for (int r = 0; r < n; ++r) { const float dlt = gown[r * MB + tid] - mu; m2 = fmaf(dlt, dlt, m2); }
      }
      a.lnpart[((size_t)c * MB + tid) * 2 + 0] = mu;
      a.lnpart[((size_t)c * MB + tid) * 2 + 1] = m2;
    }
    DBG(3);
    grid.sync();

    // ================= P3: LN statistics of the gate vector, gates, h'[own d] =================
    {   // all partials (NC x MB x 2 floats) into shared memory with every load in flight at once
      const int n4 = NC * MB * 2 / 4;
      for (int base = 0; base < n4; base += ST * 8) {
        float4 v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) { const int i = base + tid + j * ST; if (i < n4) v[j] = __ldcg(reinterpret_cast<const float4*>(a.lnpart) + i); }
#pragma unroll
        for (int j = 0; j < 8; ++j) { const int i = base + tid + j * ST; if (i < n4) reinterpret_cast<float4*>(lnp)[i] = v[j]; }
      }
    }
    __syncthreads();
    DBG(4);
    {   // 16 threads per row: mean = sum n_c mean_c / 3D ; M2 = sum (M2_c + n_c (mean_c - mean)^2)
      const int m = tid >> 4, sub = tid & 15;
      float sw = 0.f;
      for (int cc = sub; cc < NC; cc += 16) {
        const int lo = min(cc * dd, D);
        sw += (float)(3 * (min(lo + dd, D) - lo)) * lnp[(cc * MB + m) * 2];
      }
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) sw += __shfl_xor_sync(0xffffffffu, sw, o);
      const float mean = sw / (float)(3 * D);
      float sm2 = 0.f;
      for (int cc = sub; cc < NC; cc += 16) {
        const int lo = min(cc * dd, D);
        const float dm = lnp[(cc * MB + m) * 2] - mean;
        sm2 += lnp[(cc * MB + m) * 2 + 1] + (float)(3 * (min(lo + dd, D) - lo)) * dm * dm;
      }
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) sm2 += __shfl_xor_sync(0xffffffffu, sm2, o);
      if (sub == 0) { s_stat[m] = mean; s_stat[MB + m] = rsqrtf(sm2 / (float)(3 * D) + 1e-5f); }
    }
    __syncthreads();
    if (c == 0 && tid < B) { a.gru_mean[row0 + tid] = s_stat[tid]; a.gru_rstd[row0 + tid] = s_stat[MB + tid]; }
    for (int i = tid; i < nD * B; i += ST) {
      const int dl = i / B, m = i - dl * B, d = d0 + dl;
      const float mu = s_stat[m], rs = s_stat[MB + m];
      const float r = sigmoidf_((gown[(0 * nD + dl) * MB + m] - mu) * rs * __ldg(a.g_gru + d) + __ldg(a.b_gru + d));
      const float cp = (gown[(1 * nD + dl) * MB + m] - mu) * rs * __ldg(a.g_gru + D + d) + __ldg(a.b_gru + D + d);
      const float cd = tanhf(r * cp);
      const float u = sigmoidf_((gown[(2 * nD + dl) * MB + m] - mu) * rs * __ldg(a.g_gru + 2 * D + d) + __ldg(a.b_gru + 2 * D + d) - 1.f);
      const float h = act[m * KG + d];
      a.x_rep[(size_t)(row0 + m) * a.ld_xrep + d] = u * cd + (1.f - u) * h;
    }
    DBG(5);
    grid.sync();

    // ================= P4: h' (all rows) -> own posterior-hidden rows =================
    load_rows_cg(act, KG, a.x_rep + (size_t)row0 * a.ld_xrep, a.ld_xrep, B, D);
    __syncthreads();
    DBG(6);
    {
      const float* W = w_r0_s ? w_r0_s : (a.Wr0 + (size_t)hr0 * a.ld_wr0);
      const int ldw = w_r0_s ? D : a.ld_wr0;
      cta_gemv(W, ldw, nH, D, act, KG, [&](int r, int m, float v) {
        if (m < B) a.hidr_pre[(size_t)(row0 + m) * Hh + hr0 + r] = v + __ldg(a.eW + (size_t)(row0 + m) * Hh + hr0 + r);
      });
    }
    DBG(7);
    grid.sync();
    DBG(8);

    // ================= P5a: hid = SiLU(LN(hidr_pre)) (every CTA); own posterior logit rows =================
    {
      float* hid = act + D;   // feat is no longer needed in this step
      load_rows_cg(hid, KG, a.hidr_pre + (size_t)row0 * Hh, Hh, B, Hh);
      __syncthreads();
      DBG(16);
      smem_ln_silu(hid, KG, Hh, a.g_r, a.b_r, 1e-5f, s_stat + 2 * MB, s_stat + 3 * MB);
      __syncthreads();
      DBG(17);
      if (c == 0 && tid < B) { a.hidr_mean[row0 + tid] = s_stat[2 * MB + tid]; a.hidr_rstd[row0 + tid] = s_stat[3 * MB + tid]; }
      for (int i = c * ST + tid; i < B * Hh; i += NC * ST) { const int m = i / Hh, k = i - m * Hh; a.hidr[(size_t)(row0 + m) * Hh + k] = hid[m * KG + k]; }
      const float* W = w_r1_s ? w_r1_s : (a.Wr1 + (size_t)sr0 * a.ld_wr1);
      const int ldw = w_r1_s ? Hh : a.ld_wr1;
      cta_gemv(W, ldw, nS, Hh, hid, KG, [&](int r, int m, float v) {
        if (m < B) a.post_raw[(size_t)(row0 + m) * S + sr0 + r] = v + __ldg(a.br1 + sr0 + r);
      });
    }
    DBG(18);
    grid.sync();

    // ================= P5b: CTA c < G samples categorical c of every batch row =================
    if (owner) {
      for (int i = tid; i < B * Kc; i += ST) { const int m = i / Kc, k = i - m * Kc; s_log[m * Kc + k] = __ldcg(a.post_raw + (size_t)(row0 + m) * S + c * Kc + k); }
      __syncthreads();
      // one warp per batch row: softmax -> unimix -> log -> normalise -> inverse CDF
      for (int m = warp; m < B; m += NW) {
        const size_t orow = (size_t)(row0 + m);
        float mx = -INFINITY;
        for (int k = lane; k < Kc; k += 32) mx = fmaxf(mx, s_log[m * Kc + k]);
        mx = warp_max(mx);
        float se = 0.f;
        for (int k = lane; k < Kc; k += 32) se += expf(s_log[m * Kc + k] - mx);
        se = warp_sum(se);
        float mx2 = -INFINITY;
        for (int k = lane; k < Kc; k += 32) {
          const float raw = s_log[m * Kc + k];
          float lu = raw;
          if (a.unimix > 0.f) {
            const float p = (1.f - a.unimix) * (expf(raw - mx) / se) + a.unimix / (float)Kc;
            lu = logf(fminf(fmaxf(p, DV3_EPS32), 1.f - DV3_EPS32));
          }
          a.post_logits[orow * S + c * Kc + k] = lu;
          s_log[m * Kc + k] = lu;
          mx2 = fmaxf(mx2, lu);
        }
        mx2 = warp_max(mx2);
        float se2 = 0.f;
        for (int k = lane; k < Kc; k += 32) se2 += expf(s_log[m * Kc + k] - mx2);
        se2 = warp_sum(se2);
        const float lse = mx2 + logf(se2);
        // probs = softmax(lu - lse) exactly as torch's Categorical(logits=lu).probs
        float mx3 = -INFINITY;
        for (int k = lane; k < Kc; k += 32) mx3 = fmaxf(mx3, s_log[m * Kc + k] - lse);
        mx3 = warp_max(mx3);
        float se3 = 0.f;
        for (int k = lane; k < Kc; k += 32) se3 += expf(s_log[m * Kc + k] - lse - mx3);
        se3 = warp_sum(se3);
        __syncwarp();
        for (int k = lane; k < Kc; k += 32) s_log[m * Kc + k] = expf(s_log[m * Kc + k] - lse - mx3) / se3;
        __syncwarp();
        int choice = 0;
        if (a.forced_idx) {
          choice = __ldg(a.forced_idx + orow * G + c);
        } else if (lane == 0) {
          const float uu = __ldg(a.u_post + orow * G + c);
          float cs = 0.f;
#pragma unroll 8
          for (int k = 0; k < Kc; ++k) { cs += s_log[m * Kc + k]; choice += (cs <= uu) ? 1 : 0; }
          if (choice > Kc - 1) choice = Kc - 1;
        }
        choice = __shfl_sync(0xffffffffu, choice, 0);
        if (lane == 0) a.post_idx[orow * G + c] = choice;
        for (int k = lane; k < Kc; k += 32) a.latent[orow * a.ld_lat + c * Kc + k] = (k == choice) ? 1.f : 0.f;
      }
    }
    DBG(9);
    grid.sync();
    DBG(10);
  }
}

}  // namespace

extern "C" {

// Plain-C argument block of dv3_rssm_observe_fwd (all pointers are device pointers; see the field comments above).
struct Dv3ScanFwdArgs {
  int T, B, G, Kc, D, U, Hh, A, E;
  float unimix;
  const float* Wmlp; int ld_wmlp; const float* g_mlp; const float* b_mlp;
  const float* Wgru; int ld_wgru; const float* g_gru; const float* b_gru;
  const float* Wr0; int ld_wr0; const float* g_r; const float* b_r;
  const float* Wr1; int ld_wr1; const float* br1;
  const float* eW; const float* first; const float* u_post; const int* forced_idx; const int* zreset_idx;
  float* x_mlp; int ld_xmlp;
  float* feat_pre; float* feat_mean; float* feat_rstd;
  float* x_gru; int ld_xgru;
  float* gates_pre; float* gru_mean; float* gru_rstd;
  float* x_rep; int ld_xrep;
  float* hidr_pre; float* hidr; float* hidr_mean; float* hidr_rstd;
  float* post_raw; float* post_logits; int* post_idx;
  float* latent; int ld_lat;
  float* scratch; size_t scratch_bytes;
};

size_t dv3_rssm_observe_fwd_scratch_bytes(void) { return (size_t)1024 * MB * 2 * sizeof(float); }

int dv3_rssm_observe_fwd(const Dv3ScanFwdArgs* h, void* stream_) {
  if (!h) return DV3_ERR_ARG;
  if (h->B < 1 || h->B > MB || h->T < 1 || h->G < 1 || h->Kc < 1 || h->Hh > h->U) return DV3_ERR_UNSUPPORTED;
  const int S = h->G * h->Kc, KG = h->D + h->U;
  if ((KG & 3) || (h->D & 3) || (h->Hh & 3) || (h->ld_wgru & 3) || (h->ld_wr0 & 3) || (h->ld_wr1 & 3) ||
      ((uintptr_t)h->Wgru & 15) || ((uintptr_t)h->Wr0 & 15) || ((uintptr_t)h->Wr1 & 15))
    return DV3_ERR_UNSUPPORTED;   // the float4 contraction path needs 16 B aligned weight rows
  int dev = 0, sms = 0, max_smem = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  cudaDeviceGetAttribute(&max_smem, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev);
  int NC = sms;
  if (NC < h->G || h->G * MB > 2 * ST || ((NC * MB * 2) & 3)) return DV3_ERR_UNSUPPORTED;   // one CTA per categorical in P5b
  if (NC > 1024) NC = 1024;
  if (h->scratch_bytes < (size_t)NC * MB * 2 * sizeof(float)) return DV3_ERR_SCRATCH;

  ScanFwdArgs a;
  a.T = h->T; a.B = h->B; a.G = h->G; a.Kc = h->Kc; a.D = h->D; a.U = h->U; a.Hh = h->Hh; a.A = h->A; a.E = h->E;
  a.unimix = h->unimix;
  a.Wmlp = h->Wmlp; a.ld_wmlp = h->ld_wmlp; a.g_mlp = h->g_mlp; a.b_mlp = h->b_mlp;
  a.Wgru = h->Wgru; a.ld_wgru = h->ld_wgru; a.g_gru = h->g_gru; a.b_gru = h->b_gru;
  a.Wr0 = h->Wr0; a.ld_wr0 = h->ld_wr0; a.g_r = h->g_r; a.b_r = h->b_r;
  a.Wr1 = h->Wr1; a.ld_wr1 = h->ld_wr1; a.br1 = h->br1;
  a.eW = h->eW; a.first = h->first; a.u_post = h->u_post; a.forced_idx = h->forced_idx; a.zreset_idx = h->zreset_idx;
  a.x_mlp = h->x_mlp; a.ld_xmlp = h->ld_xmlp;
  a.feat_pre = h->feat_pre; a.feat_mean = h->feat_mean; a.feat_rstd = h->feat_rstd;
  a.x_gru = h->x_gru; a.ld_xgru = h->ld_xgru;
  a.gates_pre = h->gates_pre; a.gru_mean = h->gru_mean; a.gru_rstd = h->gru_rstd;
  a.x_rep = h->x_rep; a.ld_xrep = h->ld_xrep;
  a.hidr_pre = h->hidr_pre; a.hidr = h->hidr; a.hidr_mean = h->hidr_mean; a.hidr_rstd = h->hidr_rstd;
  a.post_raw = h->post_raw; a.post_logits = h->post_logits; a.post_idx = h->post_idx;
  a.latent = h->latent; a.ld_lat = h->ld_lat;
  a.lnpart = h->scratch;
  a.dbg = (h->scratch_bytes >= (size_t)(NC * MB * 2 + 64) * sizeof(float) && getenv("DV3_SCAN_DEBUG"))
              ? reinterpret_cast<long long*>(h->scratch + NC * MB * 2 + ((NC * MB * 2) & 1)) : nullptr;

  // shared-memory plan: fixed buffers first, then the weight slices in order of benefit while they fit
  auto r4 = [](long long n) { return (n + 3) & ~3LL; };
  const int KA = S + h->A, KAp = (KA + 3) & ~3;
  const int ua = (h->U + NC - 1) / NC, dd = (h->D + NC - 1) / NC, hr = (h->Hh + NC - 1) / NC;
  long long used = r4((long long)MB * KG) + r4((long long)MB * (h->A + 4)) + r4(3LL * (dd > 1 ? dd : 1) * MB) + r4(4 * MB) +
                   r4((long long)MB * h->Kc) + r4((long long)MB * h->G) + r4(MB) + r4((long long)NC * MB * 2);
  const long long budget = (long long)max_smem / 4 - 64;
  const long long n_gru = r4(3LL * (dd > 1 ? dd : 1) * KG), n_mlp = r4((long long)(ua > 1 ? ua : 1) * KAp),
                  n_r0 = r4((long long)(hr > 1 ? hr : 1) * h->D), n_r1 = r4((long long)(((S + NC - 1) / NC) > 1 ? ((S + NC - 1) / NC) : 1) * h->Hh);
  a.smem_wgru = a.smem_wmlp = a.smem_wr0 = a.smem_wr1 = 0;
  if (used > budget) return DV3_ERR_UNSUPPORTED;
  const long long fixed_used = used;
  if (used + n_gru <= budget) { a.smem_wgru = 1; used += n_gru; }
  if (used + n_mlp <= budget) { a.smem_wmlp = 1; used += n_mlp; }
  if (used + n_r0 <= budget) { a.smem_wr0 = 1; used += n_r0; }
  if (used + n_r1 <= budget) { a.smem_wr1 = 1; used += n_r1; }
  if (!a.smem_wmlp && ((h->ld_wmlp & 3) || ((uintptr_t)h->Wmlp & 15))) {
    // gather path reads scalars, so unaligned rows are fine; nothing to check
  }
  const size_t smem_bytes = (size_t)used * 4 + 64;
  if (getenv("DV3_SCAN_DEBUG"))
    fprintf(stderr, "[dv3 scan] NC=%d fixed=%lld KB budget=%lld KB resident: gru=%d mlp=%d r0=%d r1=%d total=%zu KB\n", NC,
            fixed_used * 4 / 1024, budget * 4 / 1024, a.smem_wgru, a.smem_wmlp, a.smem_wr0, a.smem_wr1, smem_bytes / 1024);
  e = cudaFuncSetAttribute(rssm_observe_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_bytes);
  if (e != cudaSuccess) return (int)e;
  int occ = 0;
  e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, rssm_observe_fwd_kernel, ST, smem_bytes);
  if (e != cudaSuccess) return (int)e;
  if (occ < 1) return DV3_ERR_UNSUPPORTED;
  void* params[] = {(void*)&a};
  e = cudaLaunchCooperativeKernel((void*)rssm_observe_fwd_kernel, dim3(NC), dim3(ST), params, smem_bytes, (cudaStream_t)stream_);
  if (e != cudaSuccess) return (int)e;
  return DV3_OK;
}

}  // extern "C"
