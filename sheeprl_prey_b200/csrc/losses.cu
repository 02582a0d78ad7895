// Categorical / loss kernels of the DreamerV3 step (forward value and hand-written gradient in one pass).
//   unimix + straight-through one-hot sampling : sheeprl/algos/dreamer_v3/agent.py:390-402, :741-747;
//                                                sheeprl/utils/distribution.py:355-360, :390-393
//   KL balance with free nats                  : sheeprl/algos/dreamer_v3/loss.py:69-98 (+ torch _kl_categorical_categorical)
//   two-hot symlog log-prob / mean             : sheeprl/utils/distribution.py:245-270
//   symlog MSE                                 : sheeprl/utils/distribution.py:177-193
//   Bernoulli continue loss                    : sheeprl/algos/dreamer_v3/dreamer_v3.py:150-155, loss.py:100-101
//   lambda-returns + discount                  : sheeprl/algos/dreamer_v3/utils.py:44-55, dreamer_v3.py:262-280
//   Moments (percentile EMA)                   : sheeprl/algos/dreamer_v3/utils.py:34-41
//   REINFORCE + entropy actor objective        : sheeprl/algos/dreamer_v3/dreamer_v3.py:292-317
// One warp owns one categorical group / one row; reductions are warp shuffles.
#include "common.cuh"
#include <stddef.h>
#include <cuda_bf16.h>

namespace {

constexpr int KR = 8;            // classes per lane kept in registers -> K <= 256
constexpr int WARPS_PER_BLOCK = 4;

template <int KRT>
struct Group {  // per-lane slice of one K-class group (K <= 32*KRT)
  float v[KRT];
};

template <int KRT, typename F>
static __device__ __forceinline__ void for_k(int K, int lane, F f) {
#pragma unroll
  for (int j = 0; j < KRT; ++j) {
    const int k = lane + 32 * j;
    if (k < K) f(j, k);
  }
}
template <int KRT>
static __device__ __forceinline__ float group_max(const Group<KRT>& g, int K, int lane) {
  float m = -INFINITY;
  for_k<KRT>(K, lane, [&](int j, int) { m = fmaxf(m, g.v[j]); });
  return warp_max(m);
}
template <int KRT>
static __device__ __forceinline__ float group_sum(const Group<KRT>& g, int K, int lane) {
  float s = 0.f;
  for_k<KRT>(K, lane, [&](int j, int) { s += g.v[j]; });
  return warp_sum(s);
}
// ps = softmax(raw)
template <int KRT>
static __device__ __forceinline__ void softmax_group(const Group<KRT>& raw, Group<KRT>& ps, int K, int lane) {
  const float m = group_max(raw, K, lane);
  for_k<KRT>(K, lane, [&](int j, int) { ps.v[j] = expf(raw.v[j] - m); });
  const float s = group_sum(ps, K, lane);
  for_k<KRT>(K, lane, [&](int j, int) { ps.v[j] = ps.v[j] / s; });
}
// unimix: pm = (1-mix) ps + mix/K ; lu = log(clamp(pm, eps, 1-eps)); inside = clamp inactive
template <int KRT>
static __device__ __forceinline__ void unimix_group(const Group<KRT>& ps, Group<KRT>& pm, Group<KRT>& lu, float mix, int K, int lane) {
  for_k<KRT>(K, lane, [&](int j, int) {
    const float p = (1.f - mix) * ps.v[j] + mix / (float)K;
    pm.v[j] = p;
    lu.v[j] = logf(fminf(fmaxf(p, DV3_EPS32), 1.f - DV3_EPS32));
  });
}
// ln = l - logsumexp(l) ; p = softmax(ln)   (torch Categorical(logits=l))
template <int KRT>
static __device__ __forceinline__ void normalise_group(const Group<KRT>& l, Group<KRT>& ln, Group<KRT>& p, int K, int lane) {
  const float m = group_max(l, K, lane);
  Group<KRT> e;
  for_k<KRT>(K, lane, [&](int j, int) { e.v[j] = expf(l.v[j] - m); });
  const float lse = m + logf(group_sum(e, K, lane));
  for_k<KRT>(K, lane, [&](int j, int) { ln.v[j] = l.v[j] - lse; });
  const float m2 = group_max(ln, K, lane);
  for_k<KRT>(K, lane, [&](int j, int) { p.v[j] = expf(ln.v[j] - m2); });
  const float s2 = group_sum(p, K, lane);
  for_k<KRT>(K, lane, [&](int j, int) { p.v[j] = p.v[j] / s2; });
}
template <int KRT>
static __device__ __forceinline__ void load_group(Group<KRT>& g, const float* p, int K, int lane) {
  for_k<KRT>(K, lane, [&](int j, int k) { g.v[j] = p[k]; });
}

// ---------------------------------------------------------------- unimix forward
template <int KRT>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
unimix_fwd_kernel(int R, int groups, int K, float mix, const float* __restrict__ raw, int ldr, float* __restrict__ out, int ldo) {
  const int lane = threadIdx.x & 31;
  const int gid = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
  if (gid >= R * groups) return;
  const int row = gid / groups, g = gid % groups;
  Group<KRT> r, ps, pm, lu;
  load_group(r, raw + (size_t)row * ldr + g * K, K, lane);
  softmax_group(r, ps, K, lane);
  if (mix > 0.f) unimix_group(ps, pm, lu, mix, K, lane);
  else lu = r;
  float* o = out + (size_t)row * ldo + g * K;
  for_k<KRT>(K, lane, [&](int j, int k) { o[k] = lu.v[j]; });
}

// ---------------------------------------------------------------- sampling (inverse CDF on injected uniforms)
template <int KRT>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
sample_onehot_kernel(int R, int groups, int K, const float* __restrict__ logits, int ldl, const float* __restrict__ u, int ldu,
                     const int* __restrict__ forced, int ldf, int mode, int* __restrict__ idx, int ldi,
                     float* __restrict__ onehot, int ldo, __nv_bfloat16* __restrict__ onehot16) {
  __shared__ float sp[WARPS_PER_BLOCK][KR * 32];
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int gid = blockIdx.x * WARPS_PER_BLOCK + w;
  if (gid >= R * groups) return;
  const int row = gid / groups, g = gid % groups;
  int choice;
  if (forced) {
    choice = forced[(size_t)row * ldf + g];
  } else {
    Group<KRT> l, ln, p;
    load_group(l, logits + (size_t)row * ldl + g * K, K, lane);
    normalise_group(l, ln, p, K, lane);
    for_k<KRT>(K, lane, [&](int j, int k) { sp[w][k] = p.v[j]; });
    __syncwarp();
    choice = 0;
    if (lane == 0) {
      if (mode) {  // argmax, first maximal index
        float best = sp[w][0];
        for (int k = 1; k < K; ++k) if (sp[w][k] > best) { best = sp[w][k]; choice = k; }
      } else {     // idx = #{k : c_k <= u}, c = sequential fp32 cumsum
        const float uu = u[(size_t)row * ldu + g];
        float c = 0.f;
        for (int k = 0; k < K; ++k) { c += sp[w][k]; choice += (c <= uu) ? 1 : 0; }
        if (choice > K - 1) choice = K - 1;
      }
    }
    choice = __shfl_sync(0xffffffffu, choice, 0);
  }
  if (lane == 0) idx[(size_t)row * ldi + g] = choice;
  float* o = onehot + (size_t)row * ldo + g * K;
  for_k<KRT>(K, lane, [&](int, int k) { o[k] = (k == choice) ? 1.f : 0.f; });
  if (onehot16) {
    __nv_bfloat16* o16 = onehot16 + (size_t)row * ldo + g * K;
    for_k<KRT>(K, lane, [&](int, int k) { o16[k] = __float2bfloat16_rn((k == choice) ? 1.f : 0.f); });
  }
}

// ---------------------------------------------------------------- unimix + sample in one pass (imagination rollout)
// = unimix_fwd followed by sample_onehot(mode 0) on its output, same arithmetic; the unimix logits are written only if asked for.
// fast != 0 (bf16 precision mode): the draw uses a warp prefix sum instead of the sequential fp32 cumsum of the sampling contract.
template <int KRT>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
unimix_sample_kernel(int R, int groups, int K, float mix, const float* __restrict__ raw, int ldr, const float* __restrict__ u, int ldu,
                     const int* __restrict__ forced, int ldf, int* __restrict__ idx, int ldi, float* __restrict__ onehot, int ldo,
                     __nv_bfloat16* __restrict__ onehot16, float* __restrict__ logits, int ldl, int fast) {
  __shared__ float sp[WARPS_PER_BLOCK][KR * 32];
  pdl_sync();
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int gid = blockIdx.x * WARPS_PER_BLOCK + w;
  if (gid >= R * groups) return;
  const int row = gid / groups, g = gid % groups;
  Group<KRT> r, ps, pm, lu;
  int choice;
  if (forced && !logits) {
    choice = forced[(size_t)row * ldf + g];
  } else {
    load_group(r, raw + (size_t)row * ldr + g * K, K, lane);
    softmax_group(r, ps, K, lane);
    if (mix > 0.f) unimix_group(ps, pm, lu, mix, K, lane);
    else lu = r;
    if (logits) { float* lo = logits + (size_t)row * ldl + g * K; for_k<KRT>(K, lane, [&](int j, int k) { lo[k] = lu.v[j]; }); }
    if (forced) {
      choice = forced[(size_t)row * ldf + g];
    } else {
      Group<KRT> ln, p;
      const float uu = u[(size_t)row * ldu + g];
      if (fast && KRT == 1 && mix > 0.f) {
        // bf16 precision mode: Categorical(logits = log pm).probs = pm / sum(pm) up to fp32 round-off (the reference renormalises
        // log(pm) twice) -- far below the bf16 error of the logits -- and a warp prefix sum replaces the sequential cumsum
        float pc = (lane < K) ? fminf(fmaxf(pm.v[0], DV3_EPS32), 1.f - DV3_EPS32) : 0.f;
        float cs = pc / warp_sum(pc);
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const float v = __shfl_up_sync(0xffffffffu, cs, o); if (lane >= o) cs += v; }
        choice = __popc(__ballot_sync(0xffffffffu, lane < K && cs <= uu));
      } else if (fast && KRT == 1) {
        normalise_group(lu, ln, p, K, lane);
        float cs = (lane < K) ? p.v[0] : 0.f;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const float v = __shfl_up_sync(0xffffffffu, cs, o); if (lane >= o) cs += v; }
        choice = __popc(__ballot_sync(0xffffffffu, lane < K && cs <= uu));
      } else {
        normalise_group(lu, ln, p, K, lane);
        for_k<KRT>(K, lane, [&](int j, int k) { sp[w][k] = p.v[j]; });
        __syncwarp();
        choice = 0;
        if (lane == 0) {
          float c = 0.f;
          for (int k = 0; k < K; ++k) { c += sp[w][k]; choice += (c <= uu) ? 1 : 0; }
        }
        choice = __shfl_sync(0xffffffffu, choice, 0);
      }
      if (choice > K - 1) choice = K - 1;
    }
  }
  if (lane == 0) idx[(size_t)row * ldi + g] = choice;
  float* o = onehot + (size_t)row * ldo + g * K;
  for_k<KRT>(K, lane, [&](int, int k) { o[k] = (k == choice) ? 1.f : 0.f; });
  if (onehot16) {
    __nv_bfloat16* o16 = onehot16 + (size_t)row * ldo + g * K;
    for_k<KRT>(K, lane, [&](int, int k) { o16[k] = __float2bfloat16_rn((k == choice) ? 1.f : 0.f); });
  }
}

// ---------------------------------------------------------------- unimix backward (+ straight-through)
template <int KRT>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
unimix_bwd_kernel(int R, int groups, int K, float mix, const float* __restrict__ raw, int ldr,
                  const float* __restrict__ dlogits, int lddl, const float* __restrict__ dz, int lddz,
                  float* __restrict__ draw, int lddr) {
  const int lane = threadIdx.x & 31;
  const int gid = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
  if (gid >= R * groups) return;
  const int row = gid / groups, g = gid % groups;
  Group<KRT> r, ps, pm, lu, gl;
  load_group(r, raw + (size_t)row * ldr + g * K, K, lane);
  softmax_group(r, ps, K, lane);
  if (mix > 0.f) unimix_group(ps, pm, lu, mix, K, lane);
  else { lu = r; pm = ps; }
  for_k<KRT>(K, lane, [&](int j, int k) { gl.v[j] = dlogits ? dlogits[(size_t)row * lddl + g * K + k] : 0.f; });
  if (dz) {
    Group<KRT> ln, p, d;
    normalise_group(lu, ln, p, K, lane);
    load_group(d, dz + (size_t)row * lddz + g * K, K, lane);
    float dot = 0.f;
    for_k<KRT>(K, lane, [&](int j, int) { dot = fmaf(p.v[j], d.v[j], dot); });
    dot = warp_sum(dot);
    for_k<KRT>(K, lane, [&](int j, int) { gl.v[j] += p.v[j] * (d.v[j] - dot); });
  }
  float* o = draw + (size_t)row * lddr + g * K;
  if (mix > 0.f) {
    Group<KRT> gps;
    float dot = 0.f;
    for_k<KRT>(K, lane, [&](int j, int) {
      const bool inside = (pm.v[j] >= DV3_EPS32) && (pm.v[j] <= 1.f - DV3_EPS32);
      gps.v[j] = inside ? (1.f - mix) * (gl.v[j] / pm.v[j]) : 0.f;
      dot = fmaf(ps.v[j], gps.v[j], dot);
    });
    dot = warp_sum(dot);
    for_k<KRT>(K, lane, [&](int j, int k) { o[k] = ps.v[j] * (gps.v[j] - dot); });
  } else {
    for_k<KRT>(K, lane, [&](int j, int k) { o[k] = gl.v[j]; });
  }
}

// step-independent factors of unimix_bwd for every (row, group): ps = softmax(raw), pn = probs of the unimix logits, coef
template <int KRT>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
unimix_bwd_prep_kernel(int R, int groups, int K, float mix, const float* __restrict__ raw, int ldr, float* __restrict__ ps_o,
                       float* __restrict__ coef_o, float* __restrict__ pn_o) {
  const int lane = threadIdx.x & 31;
  const int gid = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
  if (gid >= R * groups) return;
  const int row = gid / groups, g = gid % groups;
  Group<KRT> r, ps, pm, lu, ln, p;
  load_group(r, raw + (size_t)row * ldr + g * K, K, lane);
  softmax_group(r, ps, K, lane);
  if (mix > 0.f) unimix_group(ps, pm, lu, mix, K, lane);
  else { lu = r; pm = ps; }
  normalise_group(lu, ln, p, K, lane);
  const size_t o = (size_t)row * groups * K + g * K;
  for_k<KRT>(K, lane, [&](int j, int k) {
    const bool inside = (pm.v[j] >= DV3_EPS32) && (pm.v[j] <= 1.f - DV3_EPS32);
    ps_o[o + k] = ps.v[j];
    pn_o[o + k] = p.v[j];
    coef_o[o + k] = (mix > 0.f) ? (inside ? (1.f - mix) / pm.v[j] : 0.f) : 1.f;
  });
}

// ---------------------------------------------------------------- KL balance (block = one row, warp = one group)
template <int KRT>
__global__ void __launch_bounds__(1024)
kl_balance_kernel(int groups, int K, const float* __restrict__ post, const float* __restrict__ prior, float free_nats,
                  float w_dyn, float w_rep, float scale, float* __restrict__ rows_out, float* __restrict__ dpost,
                  float* __restrict__ dprior) {
  __shared__ float s_kl[32], s_ep[32], s_eq[32];
  const int lane = threadIdx.x & 31, g = threadIdx.x >> 5, row = blockIdx.x;
  const size_t off = ((size_t)row * groups + g) * K;
  Group<KRT> lp, lq, p, q, lpn, lqn;
  load_group(lp, post + off, K, lane);
  load_group(lq, prior + off, K, lane);
  normalise_group(lp, lpn, p, K, lane);
  normalise_group(lq, lqn, q, K, lane);
  float t = 0.f, ep = 0.f, eq = 0.f;
  for_k<KRT>(K, lane, [&](int j, int) {
    float tt = p.v[j] * (lpn.v[j] - lqn.v[j]);
    if (q.v[j] == 0.f) tt = INFINITY;
    if (p.v[j] == 0.f) tt = 0.f;
    t += tt;
    ep -= fmaxf(lpn.v[j], -3.4028234663852886e38f) * p.v[j];
    eq -= fmaxf(lqn.v[j], -3.4028234663852886e38f) * q.v[j];
  });
  const float klg = warp_sum(t);
  ep = warp_sum(ep);
  eq = warp_sum(eq);
  if (lane == 0) { s_kl[g] = klg; s_ep[g] = ep; s_eq[g] = eq; }
  __syncthreads();
  float kl = 0.f, sep = 0.f, seq = 0.f;
  for (int i = 0; i < groups; ++i) { kl += s_kl[i]; sep += s_ep[i]; seq += s_eq[i]; }  // fixed order, every thread
  if (threadIdx.x == 0) {
    rows_out[(size_t)row * 4 + 0] = kl;
    rows_out[(size_t)row * 4 + 1] = (w_dyn + w_rep) * fmaxf(kl, free_nats);
    rows_out[(size_t)row * 4 + 2] = sep;
    rows_out[(size_t)row * 4 + 3] = seq;
  }
  const float active = (kl > free_nats) ? 1.f : ((kl == free_nats) ? 0.5f : 0.f);
  for_k<KRT>(K, lane, [&](int j, int k) {
    dprior[off + k] = scale * w_dyn * active * (q.v[j] - p.v[j]);
    dpost[off + k] = scale * w_rep * active * p.v[j] * ((lpn.v[j] - lqn.v[j]) - klg);
  });
}

// ---------------------------------------------------------------- two-hot
static __device__ __forceinline__ float bin_value(int i, int K) {
  // torch.linspace(-20, 20, K): start + i*step below the midpoint, end - (K-1-i)*step above it
  const float step = 40.f / (float)(K - 1);
  return (i < K / 2) ? (-20.f + step * (float)i) : (20.f - step * (float)(K - 1 - i));
}
// two-hot target of symlog(x): indices (below, above) and weights
static __device__ __forceinline__ void twohot_target(float x, int K, int lane, int& below, int& above, float& wb, float& wa) {
  const float y = symlogf_(x);
  int le = 0, gt = 0;
  for (int k = lane; k < K; k += 32) {
    const float b = bin_value(k, K);
    le += (b <= y) ? 1 : 0;
    gt += (b > y) ? 1 : 0;
  }
  le = warp_sum_i(le);
  gt = warp_sum_i(gt);
  below = min(max(le - 1, 0), K - 1);
  above = min(max(K - gt, 0), K - 1);
  float db, da;
  if (below == above) { db = 1.f; da = 1.f; }
  else { db = fabsf(bin_value(below, K) - y); da = fabsf(bin_value(above, K) - y); }
  const float tot = db + da;
  wb = da / tot;
  wa = db / tot;
}

template <int KRT>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
twohot_loss_kernel(int M, int K, const float* __restrict__ logits, int ldl, const float* __restrict__ x,
                   const float* __restrict__ w, float scale, float* __restrict__ loss_rows, float* __restrict__ dlogits,
                   int ldd, int accumulate) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
  if (row >= M) return;
  Group<KRT> l, e;
  load_group(l, logits + (size_t)row * ldl, K, lane);
  const float m = group_max(l, K, lane);
  for_k<KRT>(K, lane, [&](int j, int) { e.v[j] = expf(l.v[j] - m); });
  const float s = group_sum(e, K, lane);
  const float lse = m + logf(s);
  int below, above;
  float wb, wa;
  twohot_target(x[row], K, lane, below, above, wb, wa);
  float part = 0.f;
  for_k<KRT>(K, lane, [&](int j, int k) {
    float tgt = 0.f;
    if (k == below) tgt += wb;
    if (k == above) tgt += wa;
    part += tgt * (l.v[j] - lse);
  });
  const float nll = -warp_sum(part);
  const float ww = w ? w[row] : 1.f;
  if (lane == 0) loss_rows[row] = (accumulate ? loss_rows[row] : 0.f) + ww * nll;
  float* d = dlogits + (size_t)row * ldd;
  for_k<KRT>(K, lane, [&](int j, int k) {
    float tgt = 0.f;
    if (k == below) tgt += wb;
    if (k == above) tgt += wa;
    const float gval = scale * ww * (expf(l.v[j] - lse) * (wb + wa) - tgt);
    d[k] = (accumulate ? d[k] : 0.f) + gval;
  });
}

template <int KRT>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
twohot_mean_kernel(int M, int K, const float* __restrict__ logits, int ldl, float* __restrict__ out) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
  if (row >= M) return;
  Group<KRT> l, ps;
  load_group(l, logits + (size_t)row * ldl, K, lane);
  softmax_group(l, ps, K, lane);
  float part = 0.f;
  for_k<KRT>(K, lane, [&](int j, int k) { part = fmaf(ps.v[j], bin_value(k, K), part); });
  const float v = warp_sum(part);
  if (lane == 0) out[row] = symexpf_(v);
}

// ---------------------------------------------------------------- symlog-MSE / Bernoulli
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
symlog_mse_kernel(int M, int Dd, const float* __restrict__ mode, int ldm, const float* __restrict__ target, int ldt,
                  float scale, float* __restrict__ loss_rows, float* __restrict__ dmode, int ldd) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
  if (row >= M) return;
  float s = 0.f;
  for (int i = lane; i < Dd; i += 32) {
    const float diff = mode[(size_t)row * ldm + i] - target[(size_t)row * ldt + i];
    const float d = diff * diff;
    const bool keep = d >= 1e-8f;
    s += keep ? d : 0.f;
    dmode[(size_t)row * ldd + i] = keep ? scale * 2.f * diff : 0.f;
  }
  s = warp_sum(s);
  if (lane == 0) loss_rows[row] = s;
}
__global__ void bernoulli_loss_kernel(int M, const float* __restrict__ logit, const float* __restrict__ dones, float scale,
                                      float* __restrict__ loss_rows, float* __restrict__ dlogit) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= M) return;
  const float l = logit[i], t = 1.f - dones[i];
  loss_rows[i] = fmaxf(l, 0.f) - l * t + log1pf(expf(-fabsf(l)));
  dlogit[i] = scale * (sigmoidf_(l) - t);
}

// out[slot] (+)= scale * sum_m x[m*ldx + col]   (single block; deterministic order)
__global__ void __launch_bounds__(1024)
reduce_mean_kernel(int M, const float* __restrict__ x, int ldx, int col, float* __restrict__ out, int slot, float scale,
                   int accumulate) {
  __shared__ float red[64];
  float s = 0.f;
  for (int m = threadIdx.x; m < M; m += blockDim.x) s += x[(size_t)m * ldx + col];
  s = block_sum(s, red);
  if (threadIdx.x == 0) out[slot] = (accumulate ? out[slot] : 0.f) + scale * s;
}

// ---------------------------------------------------------------- lambda-returns
__global__ void lambda_returns_kernel(int H, int N, const float* __restrict__ values, const float* __restrict__ rewards,
                                      const float* __restrict__ cont_logit, const float* __restrict__ dones, float gamma,
                                      float lmbda, float* __restrict__ lam, float* __restrict__ discount) {
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= N) return;
  auto cont = [&](int t) -> float {
    if (t == 0) return 1.f - dones[n];
    const float p = sigmoidf_(cont_logit[(size_t)t * N + n]);
    return (p == 0.5f) ? NAN : ((p > 0.5f) ? 1.f : 0.f);   // Bernoulli.mode
  };
  float nxt = values[(size_t)H * N + n];
  for (int t = H - 1; t >= 0; --t) {  // output index t <-> imagined step t+1
    const float cg = cont(t + 1) * gamma;
    const float interm = rewards[(size_t)(t + 1) * N + n] + cg * values[(size_t)(t + 1) * N + n] * (1.f - lmbda);
    nxt = interm + cg * lmbda * nxt;
    lam[(size_t)t * N + n] = nxt;
  }
  float prod = 1.f;
  for (int t = 0; t <= H; ++t) {
    prod *= cont(t) * gamma;
    discount[(size_t)t * N + n] = prod / gamma;
  }
}

// ---------------------------------------------------------------- Moments: exact quantiles by radix select (one block)
static __device__ __forceinline__ unsigned f2key(float f) {
  const unsigned u = __float_as_uint(f);
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
static __device__ __forceinline__ float key2f(unsigned k) {
  const unsigned u = (k & 0x80000000u) ? (k & 0x7fffffffu) : ~k;
  return __uint_as_float(u);
}
// k-th smallest (0-based) of x[0..n): MSB-first 8-bit radix select; all threads of the block participate.  The histogram update is
// aggregated per warp (match.any on the bin, one atomic per distinct bin): lambda-values share their exponent, so in the first
// passes nearly every element hits the same bin and plain shared-memory atomics serialise (with the values of 8 ranks gathered, the
// one-block kernel took several hundred microseconds on the critical path of the data-parallel step).
static __device__ unsigned select_kth_key(const float* __restrict__ x, int n, int kth, unsigned* hist, unsigned* bcast) {
  unsigned prefix = 0, mask = 0;
  int rank = kth;
  const int lane = threadIdx.x & 31;
  const int n_pad = (n + 31) & ~31;                          // whole warps iterate together (match.any needs the full mask)
  for (int shift = 24; shift >= 0; shift -= 8) {
    for (int i = threadIdx.x; i < 256; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    for (int i = threadIdx.x; i < n_pad; i += blockDim.x) {
      unsigned bin = 0xffffffffu;                            // "no contribution"
      if (i < n) {
        const unsigned key = f2key(x[i]);
        if ((key & mask) == prefix) bin = (key >> shift) & 255u;
      }
      const unsigned peers = __match_any_sync(0xffffffffu, bin);
      if (bin != 0xffffffffu && lane == __ffs(peers) - 1) atomicAdd(&hist[bin], (unsigned)__popc(peers));
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned acc = 0;
      int b = 0;
      for (; b < 256; ++b) {
        if (acc + hist[b] > (unsigned)rank) break;
        acc += hist[b];
      }
      bcast[0] = (unsigned)b;
      bcast[1] = acc;
    }
    __syncthreads();
    prefix |= bcast[0] << shift;
    mask |= 255u << shift;
    rank -= (int)bcast[1];
    __syncthreads();
  }
  return prefix;
}
// the next order statistic after the kth (whose key is `akey`): the same value if at least kth + 2 elements are <= it, else the
// smallest larger element -- one counting pass instead of a second radix select
static __device__ unsigned next_after_key(const float* __restrict__ x, int n, int kth, unsigned akey, unsigned* hist, unsigned* bcast) {
  if (threadIdx.x == 0) { hist[0] = 0u; hist[1] = 0xffffffffu; }
  __syncthreads();
  unsigned cnt = 0, mn = 0xffffffffu;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const unsigned key = f2key(x[i]);
    if (key <= akey) ++cnt;
    else mn = min(mn, key);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
    mn = min(mn, __shfl_xor_sync(0xffffffffu, mn, o));
  }
  if ((threadIdx.x & 31) == 0) { atomicAdd(&hist[0], cnt); atomicMin(&hist[1], mn); }
  __syncthreads();
  const unsigned res = (hist[0] >= (unsigned)(kth + 2)) ? akey : hist[1];
  __syncthreads();
  return res;
}
static __device__ float quantile_linear(const float* x, int n, float q, unsigned* hist, unsigned* bcast) {
  const float rank = q * (float)(n - 1);       // fp32, like torch.quantile
  const float fl = floorf(rank), ce = ceilf(rank);
  const float w = rank - fl;
  const unsigned ak = select_kth_key(x, n, (int)fl, hist, bcast);
  const float a = key2f(ak);
  const float b = ((int)ce == (int)fl) ? a : key2f(next_after_key(x, n, (int)fl, ak, hist, bcast));
  return (w < 0.5f) ? (a + w * (b - a)) : (b - (b - a) * (1.f - w));  // at::lerp
}
__global__ void __launch_bounds__(1024)
moments_kernel(int n, const float* __restrict__ x, float* __restrict__ state, float* __restrict__ out, float decay,
               float max_, float plow, float phigh) {
  __shared__ unsigned hist[256];
  __shared__ unsigned bcast[2];
  const float lo = quantile_linear(x, n, plow, hist, bcast);
  const float hi = quantile_linear(x, n, phigh, hist, bcast);
  if (threadIdx.x == 0) {
    const float nl = decay * state[0] + (1.f - decay) * lo;
    const float nh = decay * state[1] + (1.f - decay) * hi;
    state[0] = nl;
    state[1] = nh;
    out[0] = nl;
    out[1] = fmaxf(1.f / max_, nh - nl);
  }
}

// ---- Moments, multi-block (n >= 8192: every data-parallel step gathers the lambda-values of all ranks) -----------------------------
// The one-block kernel above is bound by its dependent global loads (one per thread and iteration, 10 passes): 66 us at 15 360 values,
// 474 us at the 122 880 values of 8 ranks -- on the critical path of the step.  Here the radix select runs as three grid-wide
// histogram passes (11 + 11 + 10 key bits; both quantiles share the launches, each block keeps its slice of x in registers), one
// counting pass for the upper neighbours of the two order statistics, and a one-block finish; the pass-to-pass decisions (which bin
// holds the rank) are recomputed by every block from the global histograms, so there is no grid barrier.  State lives in a fixed
// device array of the library (80 KB): calls on one GPU must not overlap; the finishing kernel leaves it zeroed for the next call (it loads as zeros, so the first call needs no set-up).
constexpr int MQ_BINS = 2048, MQ_THREADS = 256, MQ_PER_THREAD = 16, MQ_MAX_BLOCKS = 148;
struct MomScratch {
  unsigned hist[3][2][MQ_BINS];        // [pass][quantile][bin]  (pass 0 is shared: only [0][0] is used)
  unsigned cnt[2];                     // #{key <= a_q}
  unsigned mn[2];                      // max{~key : key > a_q}  (0 = no such element; so the zero-initialised array needs no priming)
};
static __device__ MomScratch g_mom;

// bin of `hist` (nbins) that holds 0-based rank `rank`; returns the bin, `below` = elements in lower bins.  All threads of the block.
static __device__ unsigned mq_find_bin(const unsigned* __restrict__ hist, int nbins, unsigned rank, unsigned* below, unsigned* sh) {
  // sh: [MQ_THREADS + 2] words of shared memory.  Thread t owns `per` consecutive bins; a block-wide prefix sum of the per-thread
  // totals names the one thread whose bins straddle the rank, and only that thread walks its bins.
  const int per = nbins / MQ_THREADS;                       // 8 or 4
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  unsigned v[8];
  unsigned local = 0;
#pragma unroll
  for (int j = 0; j < 8; ++j) { v[j] = (j < per) ? hist[threadIdx.x * per + j] : 0u; local += v[j]; }
  unsigned incl = local;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) { const unsigned t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
  if (lane == 31) sh[warp] = incl;
  __syncthreads();
  unsigned woff = 0;
  for (int w2 = 0; w2 < warp; ++w2) woff += sh[w2];
  const unsigned excl = woff + incl - local;
  if (rank >= excl && rank < excl + local) {                // exactly one thread (the total exceeds the rank by construction)
    unsigned acc = excl;
    int b = 0;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (j < per - 1 && acc + v[j] <= rank && b == j) { acc += v[j]; b = j + 1; }
    }
    sh[MQ_THREADS] = (unsigned)(threadIdx.x * per + b);
    sh[MQ_THREADS + 1] = acc;
  }
  __syncthreads();
  const unsigned bsel = sh[MQ_THREADS];
  *below = sh[MQ_THREADS + 1];
  __syncthreads();
  return bsel;
}
// resolves, for quantile q with 0-based rank `rank`, the key prefix after `npass` passes (bits 31..21, 20..10, 9..0)
static __device__ unsigned mq_prefix(int q, unsigned rank, int npass, unsigned* sh) {
  unsigned prefix = 0, below;
  if (npass >= 1) { const unsigned b = mq_find_bin(g_mom.hist[0][0], MQ_BINS, rank, &below, sh); prefix = b << 21; rank -= below; }
  if (npass >= 2) { const unsigned b = mq_find_bin(g_mom.hist[1][q], MQ_BINS, rank, &below, sh); prefix |= b << 10; rank -= below; }
  if (npass >= 3) { const unsigned b = mq_find_bin(g_mom.hist[2][q], 1024, rank, &below, sh); prefix |= b; }
  return prefix;
}
// PASS 0..2: histogram pass; PASS 3: counting pass for the upper neighbours
template <int PASS>
__global__ void __launch_bounds__(MQ_THREADS) moments_pass_kernel(int n, const float* __restrict__ x, unsigned rank_lo, unsigned rank_hi) {
  __shared__ unsigned sh[MQ_THREADS + 2];
  __shared__ unsigned lh[(PASS == 1 || PASS == 2) ? 2 * MQ_BINS : (PASS == 0 ? MQ_BINS : 4)];
  // this block's slice of x, loaded before anything that depends on earlier passes: all loads are in flight together
  unsigned key[MQ_PER_THREAD];
  const int per_block = MQ_THREADS * MQ_PER_THREAD;
  for (int base = blockIdx.x * per_block; base < n; base += gridDim.x * per_block) {
#pragma unroll
    for (int j = 0; j < MQ_PER_THREAD; ++j) {
      const int i = base + j * MQ_THREADS + threadIdx.x;
      key[j] = (i < n) ? f2key(x[i]) : 0u;
    }
    unsigned p_lo = 0, p_hi = 0;
    if (PASS > 0) { p_lo = mq_prefix(0, rank_lo, PASS, sh); p_hi = mq_prefix(1, rank_hi, PASS, sh); }
    if (PASS < 3) {
      const int nb = (PASS == 0) ? MQ_BINS : 2 * MQ_BINS;
      for (int i = threadIdx.x; i < nb; i += MQ_THREADS) lh[i] = 0;
      __syncthreads();
#pragma unroll
      for (int j = 0; j < MQ_PER_THREAD; ++j) {
        const int i = base + j * MQ_THREADS + threadIdx.x;
        if (i < n) {
          const unsigned kk = key[j];
          if (PASS == 0) atomicAdd(&lh[kk >> 21], 1u);
          if (PASS == 1) {
            if ((kk >> 21) == (p_lo >> 21)) atomicAdd(&lh[(kk >> 10) & 2047u], 1u);
            if ((kk >> 21) == (p_hi >> 21)) atomicAdd(&lh[MQ_BINS + ((kk >> 10) & 2047u)], 1u);
          }
          if (PASS == 2) {
            if ((kk >> 10) == (p_lo >> 10)) atomicAdd(&lh[kk & 1023u], 1u);
            if ((kk >> 10) == (p_hi >> 10)) atomicAdd(&lh[MQ_BINS + (kk & 1023u)], 1u);
          }
        }
      }
      __syncthreads();
      for (int i = threadIdx.x; i < nb; i += MQ_THREADS) {
        const unsigned v = lh[i];
        if (v) atomicAdd(PASS == 0 ? &g_mom.hist[0][0][i] : &g_mom.hist[PASS][i / MQ_BINS][i % MQ_BINS], v);
      }
      __syncthreads();
    } else {
      unsigned c0 = 0, c1 = 0, m0 = 0u, m1 = 0u;
#pragma unroll
      for (int j = 0; j < MQ_PER_THREAD; ++j) {
        const int i = base + j * MQ_THREADS + threadIdx.x;
        if (i < n) {
          const unsigned kk = key[j];
          if (kk <= p_lo) ++c0; else m0 = max(m0, ~kk);
          if (kk <= p_hi) ++c1; else m1 = max(m1, ~kk);
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        c0 += __shfl_xor_sync(0xffffffffu, c0, o); c1 += __shfl_xor_sync(0xffffffffu, c1, o);
        m0 = max(m0, __shfl_xor_sync(0xffffffffu, m0, o)); m1 = max(m1, __shfl_xor_sync(0xffffffffu, m1, o));
      }
      if ((threadIdx.x & 31) == 0) {
        atomicAdd(&g_mom.cnt[0], c0); atomicAdd(&g_mom.cnt[1], c1);
        atomicMax(&g_mom.mn[0], m0); atomicMax(&g_mom.mn[1], m1);
      }
    }
  }
}
__global__ void __launch_bounds__(MQ_THREADS) moments_finish_kernel(int n, unsigned rank_lo, unsigned rank_hi, float w_lo, float w_hi,
                                                                    float* __restrict__ state, float* __restrict__ out, float decay, float max_) {
  __shared__ unsigned sh[MQ_THREADS + 2];
  const unsigned a_lo = mq_prefix(0, rank_lo, 3, sh), a_hi = mq_prefix(1, rank_hi, 3, sh);
  if (threadIdx.x == 0) {
    const unsigned b_lo = (g_mom.cnt[0] >= rank_lo + 2u || g_mom.mn[0] == 0u) ? a_lo : ~g_mom.mn[0];
    const unsigned b_hi = (g_mom.cnt[1] >= rank_hi + 2u || g_mom.mn[1] == 0u) ? a_hi : ~g_mom.mn[1];
    auto lerp = [](float a, float b, float w) { return (w < 0.5f) ? (a + w * (b - a)) : (b - (b - a) * (1.f - w)); };   // at::lerp
    const float lo = (w_lo == 0.f) ? key2f(a_lo) : lerp(key2f(a_lo), key2f(b_lo), w_lo);
    const float hi = (w_hi == 0.f) ? key2f(a_hi) : lerp(key2f(a_hi), key2f(b_hi), w_hi);
    const float nl = decay * state[0] + (1.f - decay) * lo;
    const float nh = decay * state[1] + (1.f - decay) * hi;
    state[0] = nl;
    state[1] = nh;
    out[0] = nl;
    out[1] = fmaxf(1.f / max_, nh - nl);
  }
  __syncthreads();
  // leave the scratch zeroed (min slots at +inf) for the next call
  unsigned* h = &g_mom.hist[0][0][0];
  for (int i = threadIdx.x; i < 3 * 2 * MQ_BINS; i += MQ_THREADS) h[i] = 0u;
  if (threadIdx.x < 2) { g_mom.cnt[threadIdx.x] = 0u; g_mom.mn[threadIdx.x] = 0u; }
}

// ---------------------------------------------------------------- actor objective (warp = one imagined state)
struct HeadDims { int n; int k[8]; };

template <int KRT>
__global__ void __launch_bounds__(WARPS_PER_BLOCK * 32)
actor_loss_kernel(int M, HeadDims hd, const float* __restrict__ raw, int ldr, const int* __restrict__ act_idx, int lda,
                  const float* __restrict__ lam, const float* __restrict__ values, const float* __restrict__ discount,
                  const float* __restrict__ mom, float mix, float ent_coef, float scale, float* __restrict__ loss_rows,
                  float* __restrict__ draw, int ldd) {
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * WARPS_PER_BLOCK + (threadIdx.x >> 5);
  if (row >= M) return;
  const float off = mom[0], inv = mom[1];
  const float adv = (lam[row] - off) / inv - (values[row] - off) / inv;
  const float disc = discount[row];
  float total = 0.f;
  int col = 0;
  for (int h = 0; h < hd.n; ++h) {
    const int K = hd.k[h];
    Group<KRT> r, ps, pm, lu, ln, p;
    load_group(r, raw + (size_t)row * ldr + col, K, lane);
    softmax_group(r, ps, K, lane);
    if (mix > 0.f) unimix_group(ps, pm, lu, mix, K, lane);
    else { lu = r; pm = ps; }
    normalise_group(lu, ln, p, K, lane);
    const int a = act_idx[(size_t)row * lda + h];
    float lp = 0.f, en = 0.f;
    for_k<KRT>(K, lane, [&](int j, int k) {
      if (k == a) lp = ln.v[j];
      en -= ln.v[j] * p.v[j];
    });
    lp = warp_sum(lp);
    en = warp_sum(en);
    total += lp * adv + ent_coef * en;
    Group<KRT> glu;
    for_k<KRT>(K, lane, [&](int j, int k) {
      const float oh = (k == a) ? 1.f : 0.f;
      glu.v[j] = -(scale * disc) * (adv * (oh - p.v[j]) + ent_coef * (-p.v[j] * (ln.v[j] + en)));
    });
    float* o = draw + (size_t)row * ldd + col;
    if (mix > 0.f) {
      Group<KRT> gps;
      float dot = 0.f;
      for_k<KRT>(K, lane, [&](int j, int) {
        const bool inside = (pm.v[j] >= DV3_EPS32) && (pm.v[j] <= 1.f - DV3_EPS32);
        gps.v[j] = inside ? (1.f - mix) * (glu.v[j] / pm.v[j]) : 0.f;
        dot = fmaf(ps.v[j], gps.v[j], dot);
      });
      dot = warp_sum(dot);
      for_k<KRT>(K, lane, [&](int j, int k) { o[k] = ps.v[j] * (gps.v[j] - dot); });
    } else {
      for_k<KRT>(K, lane, [&](int j, int k) { o[k] = glu.v[j]; });
    }
    col += K;
  }
  if (lane == 0) loss_rows[row] = -disc * total;
}

// ---------------------------------------------------------------- counter-based uniforms
static __device__ __forceinline__ unsigned long long splitmix64(unsigned long long z) {
  z += 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  return z ^ (z >> 31);
}
__global__ void uniform_fill_kernel(size_t n, float* __restrict__ out, unsigned long long seed,
                                    const long long* __restrict__ counter) {
  const unsigned long long base = splitmix64(seed ^ splitmix64((unsigned long long)counter[0]));
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const unsigned long long r = splitmix64(base + i);
    out[i] = (float)(r >> 40) * (1.0f / 16777216.0f);   // 24 random bits -> [0,1)
  }
}
__global__ void bump_counter_kernel(long long* counter) { counter[0] += 1; }

}  // namespace

#define DV3_DISPATCH_KR(K, ...)                      \
  do {                                               \
    if ((K) <= 32) { constexpr int KRT = 1; __VA_ARGS__; }        \
    else if ((K) <= 128) { constexpr int KRT = 4; __VA_ARGS__; }  \
    else { constexpr int KRT = 8; __VA_ARGS__; }                  \
  } while (0)

extern "C" {

static inline int groups_grid(long long n) { return (int)((n + WARPS_PER_BLOCK - 1) / WARPS_PER_BLOCK); }

int dv3_unimix_fwd(int R, int groups, int K, float mix, const float* raw, int ldr, float* logits, int ldo, void* stream) {
  if (R <= 0 || groups <= 0 || K <= 0) return DV3_ERR_ARG;
  if (K > KR * 32) return DV3_ERR_UNSUPPORTED;
  DV3_DISPATCH_KR(K, unimix_fwd_kernel<KRT><<<groups_grid((long long)R * groups), WARPS_PER_BLOCK * 32, 0, (cudaStream_t)stream>>>(R, groups, K, mix, raw, ldr, logits, ldo));
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_sample_onehot(int R, int groups, int K, const float* logits, int ldl, const float* u, int ldu, const int* forced,
                      int ldf, int mode, int* idx, int ldi, float* onehot, int ldo, void* onehot_bf16, void* stream) {
  if (R <= 0 || groups <= 0 || K <= 0) return DV3_ERR_ARG;
  if (K > KR * 32) return DV3_ERR_UNSUPPORTED;
  if (!forced && !mode && !u) return DV3_ERR_ARG;
  DV3_DISPATCH_KR(K, sample_onehot_kernel<KRT><<<groups_grid((long long)R * groups), WARPS_PER_BLOCK * 32, 0, (cudaStream_t)stream>>>(
      R, groups, K, logits, ldl, u, ldu, forced, ldf, mode, idx, ldi, onehot, ldo, reinterpret_cast<__nv_bfloat16*>(onehot_bf16)));
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_unimix_sample(int R, int groups, int K, float mix, const float* raw, int ldr, const float* u, int ldu, const int* forced, int ldf,
                      int* idx, int ldi, float* onehot, int ldo, void* onehot_bf16, float* logits, int ldl, int fast, void* stream) {
  if (R <= 0 || groups <= 0 || K <= 0) return DV3_ERR_ARG;
  if (K > KR * 32) return DV3_ERR_UNSUPPORTED;
  if (!forced && !u) return DV3_ERR_ARG;
  DV3_DISPATCH_KR(K, dv3_launch_pdl(unimix_sample_kernel<KRT>, dim3(groups_grid((long long)R * groups)), dim3(WARPS_PER_BLOCK * 32), 0,
                                    (cudaStream_t)stream, R, groups, K, mix, raw, ldr, u, ldu, forced, ldf, idx, ldi, onehot, ldo,
                                    reinterpret_cast<__nv_bfloat16*>(onehot_bf16), logits, ldl, fast));
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_unimix_bwd(int R, int groups, int K, float mix, const float* raw, int ldr, const float* dlogits, int lddl,
                   const float* dz, int lddz, float* draw, int lddr, void* stream) {
  if (R <= 0 || groups <= 0 || K <= 0) return DV3_ERR_ARG;
  if (K > KR * 32) return DV3_ERR_UNSUPPORTED;
  DV3_DISPATCH_KR(K, unimix_bwd_kernel<KRT><<<groups_grid((long long)R * groups), WARPS_PER_BLOCK * 32, 0, (cudaStream_t)stream>>>(
      R, groups, K, mix, raw, ldr, dlogits, lddl, dz, lddz, draw, lddr));
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_unimix_bwd_prep(int R, int groups, int K, float mix, const float* raw, int ldr, float* ps, float* coef, float* pn, void* stream) {
  if (R <= 0 || groups <= 0 || K <= 0) return DV3_ERR_ARG;
  if (K > KR * 32) return DV3_ERR_UNSUPPORTED;
  DV3_DISPATCH_KR(K, unimix_bwd_prep_kernel<KRT><<<groups_grid((long long)R * groups), WARPS_PER_BLOCK * 32, 0, (cudaStream_t)stream>>>(
      R, groups, K, mix, raw, ldr, ps, coef, pn));
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_kl_balance(int R, int groups, int K, const float* post, const float* prior, float free_nats, float w_dyn,
                   float w_rep, float scale, float* rows_out, float* dpost, float* dprior, void* stream) {
  if (R <= 0 || groups <= 0 || groups > 32 || K <= 0) return DV3_ERR_ARG;
  if (K > KR * 32) return DV3_ERR_UNSUPPORTED;
  DV3_DISPATCH_KR(K, kl_balance_kernel<KRT><<<R, groups * 32, 0, (cudaStream_t)stream>>>(groups, K, post, prior, free_nats, w_dyn, w_rep, scale,
                                                                rows_out, dpost, dprior));
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_twohot_loss(int M, int K, const float* logits, int ldl, const float* x, const float* w, float scale,
                    float* loss_rows, float* dlogits, int ldd, int accumulate, void* stream) {
  if (M <= 0 || K <= 1) return DV3_ERR_ARG;
  if (K > KR * 32) return DV3_ERR_UNSUPPORTED;
  DV3_DISPATCH_KR(K, twohot_loss_kernel<KRT><<<groups_grid(M), WARPS_PER_BLOCK * 32, 0, (cudaStream_t)stream>>>(M, K, logits, ldl, x, w, scale,
                                                                                       loss_rows, dlogits, ldd, accumulate));
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_twohot_mean(int M, int K, const float* logits, int ldl, float* out, void* stream) {
  if (M <= 0 || K <= 1) return DV3_ERR_ARG;
  if (K > KR * 32) return DV3_ERR_UNSUPPORTED;
  DV3_DISPATCH_KR(K, twohot_mean_kernel<KRT><<<groups_grid(M), WARPS_PER_BLOCK * 32, 0, (cudaStream_t)stream>>>(M, K, logits, ldl, out));
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_symlog_mse(int M, int Dd, const float* mode, int ldm, const float* target, int ldt, float scale, float* loss_rows,
                   float* dmode, int ldd, void* stream) {
  if (M <= 0 || Dd <= 0) return DV3_ERR_ARG;
  symlog_mse_kernel<<<groups_grid(M), WARPS_PER_BLOCK * 32, 0, (cudaStream_t)stream>>>(M, Dd, mode, ldm, target, ldt, scale,
                                                                                      loss_rows, dmode, ldd);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_bernoulli_loss(int M, const float* logit, const float* dones, float scale, float* loss_rows, float* dlogit, void* stream) {
  if (M <= 0) return DV3_ERR_ARG;
  bernoulli_loss_kernel<<<(M + 255) / 256, 256, 0, (cudaStream_t)stream>>>(M, logit, dones, scale, loss_rows, dlogit);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_reduce_mean(int M, const float* x, int ldx, int col, float* out, int slot, float scale, int accumulate, void* stream) {
  if (M <= 0) return DV3_ERR_ARG;
  reduce_mean_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(M, x, ldx, col, out, slot, scale, accumulate);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_lambda_returns(int H, int N, const float* values, const float* rewards, const float* cont_logit, const float* dones,
                       float gamma, float lmbda, float* lam, float* discount, void* stream) {
  if (H <= 0 || N <= 0) return DV3_ERR_ARG;
  lambda_returns_kernel<<<(N + 127) / 128, 128, 0, (cudaStream_t)stream>>>(H, N, values, rewards, cont_logit, dones, gamma,
                                                                          lmbda, lam, discount);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_moments(int n, const float* x, float* state, float* out, float decay, float max_, float plow, float phigh, void* stream) {
  if (n <= 0) return DV3_ERR_ARG;
  if (n >= 8192) {
    // ranks and interpolation weights as torch.quantile computes them (fp32)
    const float r_lo = plow * (float)(n - 1), r_hi = phigh * (float)(n - 1);
    const float f_lo = floorf(r_lo), f_hi = floorf(r_hi);
    const float w_lo = (ceilf(r_lo) == f_lo) ? 0.f : r_lo - f_lo, w_hi = (ceilf(r_hi) == f_hi) ? 0.f : r_hi - f_hi;
    const unsigned k_lo = (unsigned)f_lo, k_hi = (unsigned)f_hi;
    cudaStream_t st = (cudaStream_t)stream;
    int blocks = (n + MQ_THREADS * MQ_PER_THREAD - 1) / (MQ_THREADS * MQ_PER_THREAD);
    if (blocks > MQ_MAX_BLOCKS) blocks = MQ_MAX_BLOCKS;
    moments_pass_kernel<0><<<blocks, MQ_THREADS, 0, st>>>(n, x, k_lo, k_hi);
    moments_pass_kernel<1><<<blocks, MQ_THREADS, 0, st>>>(n, x, k_lo, k_hi);
    moments_pass_kernel<2><<<blocks, MQ_THREADS, 0, st>>>(n, x, k_lo, k_hi);
    moments_pass_kernel<3><<<blocks, MQ_THREADS, 0, st>>>(n, x, k_lo, k_hi);
    moments_finish_kernel<<<1, MQ_THREADS, 0, st>>>(n, k_lo, k_hi, w_lo, w_hi, state, out, decay, max_);
    DV3_LAUNCH_CHECK();
    return DV3_OK;
  }
  moments_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(n, x, state, out, decay, max_, plow, phigh);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_actor_loss(int M, int n_heads, const int* head_dims, const float* raw, int ldr, const int* act_idx, int lda,
                   const float* lam, const float* values, const float* discount, const float* mom, float mix, float ent_coef,
                   float scale, float* loss_rows, float* draw, int ldd, void* stream) {
  if (M <= 0 || n_heads <= 0 || n_heads > 8) return DV3_ERR_ARG;
  HeadDims hd;
  hd.n = n_heads;
  int kmax = 0;
  for (int i = 0; i < n_heads; ++i) {
    if (head_dims[i] > KR * 32 || head_dims[i] <= 0) return DV3_ERR_UNSUPPORTED;
    hd.k[i] = head_dims[i];
    if (head_dims[i] > kmax) kmax = head_dims[i];
  }
  DV3_DISPATCH_KR(kmax, actor_loss_kernel<KRT><<<groups_grid(M), WARPS_PER_BLOCK * 32, 0, (cudaStream_t)stream>>>(
      M, hd, raw, ldr, act_idx, lda, lam, values, discount, mom, mix, ent_coef, scale, loss_rows, draw, ldd));
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_uniform_fill(size_t n, float* out, unsigned long long seed, long long* counter, void* stream) {
  if (n == 0) return DV3_OK;
  size_t g = (n + 255) / 256;
  if (g > 1184) g = 1184;
  uniform_fill_kernel<<<(int)g, 256, 0, (cudaStream_t)stream>>>(n, out, seed, counter);
  DV3_LAUNCH_CHECK();
  bump_counter_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(counter);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}

}  // extern "C"
