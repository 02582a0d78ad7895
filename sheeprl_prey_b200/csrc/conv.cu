// Data-movement kernels that turn the k4-s2-p1 convolutions / transposed convolutions of the DreamerV3 pixel encoder and
// decoder into dense contractions (which then run on dv3_gemm_*):
//   Conv2d(k4,s2,p1)          forward = im2col -> GEMM(NT, W[Cout][16*Cin])      backward-data = GEMM(NN) -> col2im
//   ConvTranspose2d(k4,s2,p1) forward = GEMM(NN, W[Cin][16*Cout]) -> col2im       backward-data = im2col -> GEMM(NT)
// Reference: CNNEncoder sheeprl/algos/dreamer_v3/agent.py:28-80 (Conv2d k4 s2 p1, no bias, LayerNormChannelLast eps 1e-3,
// SiLU), CNNDecoder agent.py:136-206 (Linear -> Unflatten(C,4,4) -> 3x[ConvTranspose2d, LN-CL, SiLU] -> ConvTranspose2d(+bias)
// -> +0.5), cnn_forward / LayerNormChannelLast sheeprl/utils/model.py:164-235.
// Activations are kept channel-last (NHWC) so that the channel LayerNorm is a row LayerNorm and every gather below moves
// contiguous channel runs; the 16 kernel taps are the slow index of the contraction dimension (k = tap*C + c), and the
// torch-layout weights ([Cout][Cin][4][4] / [Cin][Cout][4][4]) are re-ordered once per step with dv3_permute_last2.
#include "common.cuh"
#include <cuda_bf16.h>

namespace {

// col[(n*Ho + oh)*Wo + ow][tap*C + c] = x[n][2*oh-1+kh][2*ow-1+kw][c]  (0 outside), tap = kh*4 + kw, Ho = H/2, Wo = W/2
// One block per output row (n, oh): 32-bit index arithmetic with a single division per element; `col` (fp32) may be null when only the
// bf16 copy is consumed (bf16 precision: the contraction reads col16 and nothing else reads col).
__global__ void __launch_bounds__(256) im2col_k4s2p1_kernel(int H, int W, int C, const float* __restrict__ x,
                                                           float* __restrict__ col, __nv_bfloat16* __restrict__ col16) {
  const int Ho = H >> 1, Wo = W >> 1;
  const int n = blockIdx.x / Ho, oh = blockIdx.x - n * Ho;
  const int per_row = Wo * 16 * C;
  const size_t base = (size_t)blockIdx.x * per_row;
  for (int j = threadIdx.x; j < per_row; j += blockDim.x) {
    const int t = j / C, c = j - t * C;
    const int tap = t & 15, ow = t >> 4;
    const int ih = 2 * oh - 1 + (tap >> 2), iw = 2 * ow - 1 + (tap & 3);
    float v = 0.f;
    if (ih >= 0 && ih < H && iw >= 0 && iw < W) v = __ldg(x + (((size_t)n * H + ih) * W + iw) * C + c);
    if (col) col[base + j] = v;
    if (col16) col16[base + j] = __float2bfloat16_rn(v);
  }
}
// Same for C % 4 == 0: one float4 of channels per thread (16 B loads, 16 B fp32 / 8 B bf16 stores, all coalesced along tap*C + c)
__global__ void __launch_bounds__(256) im2col_k4s2p1_vec4_kernel(int H, int W, int C4, const float4* __restrict__ x,
                                                                float4* __restrict__ col, uint2* __restrict__ col16) {
  const int Ho = H >> 1, Wo = W >> 1;
  const int n = blockIdx.x / Ho, oh = blockIdx.x - n * Ho;
  const int per_row = Wo * 16 * C4;
  const size_t base = (size_t)blockIdx.x * per_row;
#pragma unroll 4
  for (int j = threadIdx.x; j < per_row; j += blockDim.x) {
    const int t = j / C4, c = j - t * C4;
    const int tap = t & 15, ow = t >> 4;
    const int ih = 2 * oh - 1 + (tap >> 2), iw = 2 * ow - 1 + (tap & 3);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ih >= 0 && ih < H && iw >= 0 && iw < W) v = __ldg(x + (((size_t)n * H + ih) * W + iw) * C4 + c);
    if (col) col[base + j] = v;
    if (col16) {
      __nv_bfloat162 lo = __floats2bfloat162_rn(v.x, v.y), hi = __floats2bfloat162_rn(v.z, v.w);
      uint2 pk;
      pk.x = *reinterpret_cast<uint32_t*>(&lo);
      pk.y = *reinterpret_cast<uint32_t*>(&hi);
      col16[base + j] = pk;
    }
  }
}

// y[n][oh][ow][c] = sum over the (<= 4) taps that reach it of col[(n*Hi + ih)*Wi + iw][tap*C + c]  (+ bias[c] + add)
// oh = 2*ih - 1 + kh  =>  kh = oh + 1 - 2*ih, two valid ih per oh (same for w).  Output is [N][2Hi][2Wi][C].
// One block per output row (n, oh); VEC = 4 reads / writes float4 runs of channels (C % 4 == 0), VEC = 1 is the scalar form.
template <int VEC>
__global__ void __launch_bounds__(256) col2im_k4s2p1_kernel(int Hi, int Wi, int C, const float* __restrict__ col,
                                                           const float* __restrict__ bias, float add, float* __restrict__ y) {
  const int Ho = Hi * 2, Wo = Wi * 2, Cv = C / VEC;
  const int n = blockIdx.x / Ho, oh = blockIdx.x - n * Ho;
  const int per_row = Wo * Cv;
  const size_t base = (size_t)blockIdx.x * per_row * VEC;
  for (int j = threadIdx.x; j < per_row; j += blockDim.x) {
    const int ow = j / Cv, c = (j - ow * Cv) * VEC;
    float s[VEC];
#pragma unroll
    for (int e = 0; e < VEC; ++e) s[e] = add + (bias ? __ldg(bias + c + e) : 0.f);
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      const int ih = ((oh + 1) >> 1) - a, kh = oh + 1 - 2 * ih;     // kh in {0..3} by construction when ih is valid
      if (ih < 0 || ih >= Hi || kh < 0 || kh > 3) continue;
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const int iw = ((ow + 1) >> 1) - b, kw = ow + 1 - 2 * iw;
        if (iw < 0 || iw >= Wi || kw < 0 || kw > 3) continue;
        const float* src = col + (((size_t)n * Hi + ih) * Wi + iw) * (16 * (size_t)C) + (kh * 4 + kw) * C + c;
        if (VEC == 4) {
          const float4 t = __ldg(reinterpret_cast<const float4*>(src));
          s[0] += t.x; s[1 % VEC] += t.y; s[2 % VEC] += t.z; s[3 % VEC] += t.w;
        } else {
          s[0] += __ldg(src);
        }
      }
    }
    if (VEC == 4) *reinterpret_cast<float4*>(y + base + (size_t)j * 4) = make_float4(s[0], s[1 % VEC], s[2 % VEC], s[3 % VEC]);
    else y[base + j] = s[0];
  }
}

// dst[a][c][b] = src[a][b][c]   (weights [Cout][Cin][16] <-> [Cout][16][Cin]; features [N][C][16] <-> [N][16][C])
__global__ void __launch_bounds__(256) permute_last2_kernel(int A, int B, int Cc, const float* __restrict__ src, size_t lds,
                                                           float* __restrict__ dst, size_t ldd) {
  const size_t total = (size_t)A * B * Cc;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int b = (int)(i % B);
    size_t r = i / B;
    const int c = (int)(r % Cc);
    const int a = (int)(r / Cc);
    dst[(size_t)a * ldd + (size_t)c * B + b] = __ldg(src + (size_t)a * lds + (size_t)b * Cc + c);
  }
}

// uint8 NCHW [N][C][HW] -> float NHWC [N][HW][C] scaled by `scale` (the /255 of dreamer_v3.py:102)
// (several image keys are concatenated on channels: this key's C channels land at channel offset c_off of Ctot)
__global__ void __launch_bounds__(256) u8_nchw_to_nhwc_kernel(int N, int C, int HW, const unsigned char* __restrict__ src,
                                                             float scale, float* __restrict__ dst, int Ctot, int c_off) {
  const size_t total = (size_t)N * HW * C;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % C);
    size_t r = i / C;
    const int p = (int)(r % HW);
    const int n = (int)(r / HW);
    dst[((size_t)n * HW + p) * Ctot + c_off + c] = (float)src[((size_t)n * C + c) * HW + p] * scale;
  }
}

// ---- warp-per-row LayerNorm (+SiLU) for narrow rows (channel-last LN over C = 24..1024 on N*H*W pixel rows) -------
constexpr int LW_WARPS = 8;
__global__ void __launch_bounds__(LW_WARPS * 32)
ln_act_fwd_warp_kernel(int M, int U, const float* __restrict__ x, int ldx, const float* __restrict__ gamma,
                       const float* __restrict__ beta, float eps, int act, float* __restrict__ y, int ldy,
                       float* __restrict__ mean, float* __restrict__ rstd, __nv_bfloat16* __restrict__ y16) {
  const int lane = threadIdx.x & 31;
  const size_t row = (size_t)blockIdx.x * LW_WARPS + (threadIdx.x >> 5);
  if (row >= (size_t)M) return;
  const float* xr = x + row * ldx;
  float s = 0.f;
  for (int i = lane; i < U; i += 32) s += xr[i];
  const float mu = warp_sum(s) / (float)U;
  float v = 0.f;
  for (int i = lane; i < U; i += 32) { const float d = xr[i] - mu; v = fmaf(d, d, v); }
  const float rs = rsqrtf(warp_sum(v) / (float)U + eps);
  float* yr = y + row * ldy;
  for (int i = lane; i < U; i += 32) {
    const float ln = (xr[i] - mu) * rs * __ldg(gamma + i) + __ldg(beta + i);
    const float o = act ? siluf_(ln) : ln;
    yr[i] = o;
    if (y16) y16[row * ldy + i] = __float2bfloat16_rn(o);
  }
  if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
}

// Same, row kept in registers: one 16 B load per lane and 128 columns, a single pass over global memory (U % 4 == 0, U <= 128 * NV,
// 16 B aligned rows).  The imagination rollout calls this 60+ times per step on 1024 x 512 rows, where the three dependent
// passes of the generic kernel cost three L2 round trips.
template <int NV>
__global__ void __launch_bounds__(LW_WARPS * 32)
ln_act_fwd_warp_vec_kernel(int M, int U, const float* __restrict__ x, int ldx, const float* __restrict__ gamma,
                           const float* __restrict__ beta, float eps, int act, float* __restrict__ y, int ldy,
                           float* __restrict__ mean, float* __restrict__ rstd, __nv_bfloat16* __restrict__ y16,
                           const float* __restrict__ add_table, int ld_table, const int* __restrict__ add_idx, int ld_idx, int n_heads,
                           int4 head_off) {
  pdl_sync();
  const int lane = threadIdx.x & 31;
  const size_t row = (size_t)blockIdx.x * LW_WARPS + (threadIdx.x >> 5);
  if (row >= (size_t)M) return;
  const float4* xr = reinterpret_cast<const float4*>(x + row * ldx);
  const int n4 = U >> 2;
  float4 v[NV];
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int i = lane + 32 * j;
    v[j] = (i < n4) ? xr[i] : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  // optional: x += sum_h table[head_off[h] + idx[row, h], :]  (a one-hot input block of the Linear in front, as a row gather)
  for (int h = 0; h < n_heads; ++h) {
    const int off = (h == 0) ? head_off.x : ((h == 1) ? head_off.y : ((h == 2) ? head_off.z : head_off.w));
    const float4* tr = reinterpret_cast<const float4*>(add_table + (size_t)(off + add_idx[row * ld_idx + h]) * ld_table);
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int i = lane + 32 * j;
      if (i < n4) { const float4 t = __ldg(tr + i); v[j].x += t.x; v[j].y += t.y; v[j].z += t.z; v[j].w += t.w; }
    }
  }
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < NV; ++j) s += (v[j].x + v[j].y) + (v[j].z + v[j].w);
  const float mu = warp_sum(s) / (float)U;
  float q = 0.f;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    if (lane + 32 * j < n4) {
      const float a = v[j].x - mu, b = v[j].y - mu, c = v[j].z - mu, d = v[j].w - mu;
      q = fmaf(a, a, q); q = fmaf(b, b, q); q = fmaf(c, c, q); q = fmaf(d, d, q);
    }
  }
  const float rs = rsqrtf(warp_sum(q) / (float)U + eps);
  float4* yr = reinterpret_cast<float4*>(y + row * ldy);
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int i = lane + 32 * j;
    if (i < n4) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(gamma) + i), b = __ldg(reinterpret_cast<const float4*>(beta) + i);
      float4 o;
      o.x = (v[j].x - mu) * rs * g.x + b.x; o.y = (v[j].y - mu) * rs * g.y + b.y;
      o.z = (v[j].z - mu) * rs * g.z + b.z; o.w = (v[j].w - mu) * rs * g.w + b.w;
      if (act) { o.x = siluf_(o.x); o.y = siluf_(o.y); o.z = siluf_(o.z); o.w = siluf_(o.w); }
      yr[i] = o;
      if (y16) {
        __nv_bfloat162 lo = __floats2bfloat162_rn(o.x, o.y), hi = __floats2bfloat162_rn(o.z, o.w);
        uint2 pk;
        pk.x = *reinterpret_cast<uint32_t*>(&lo);
        pk.y = *reinterpret_cast<uint32_t*>(&hi);
        *reinterpret_cast<uint2*>(y16 + row * ldy + 4 * i) = pk;
      }
    }
  }
  if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
}

__global__ void __launch_bounds__(LW_WARPS * 32)
ln_act_bwd_warp_kernel(int M, int U, const float* __restrict__ dy, int lddy, const float* __restrict__ x, int ldx,
                       const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ mean,
                       const float* __restrict__ rstd, int act, float* __restrict__ dx, int lddx, float* __restrict__ dln,
                       int lddln, __nv_bfloat16* __restrict__ dx16) {
  pdl_sync();
  const int lane = threadIdx.x & 31;
  const size_t row = (size_t)blockIdx.x * LW_WARPS + (threadIdx.x >> 5);
  if (row >= (size_t)M) return;
  const float* xr = x + row * ldx;
  const float* dyr = dy + row * lddy;
  const float mu = mean[row], rs = rstd[row];
  float s1 = 0.f, s2 = 0.f;
  for (int i = lane; i < U; i += 32) {
    const float xh = (xr[i] - mu) * rs;
    float g = dyr[i];
    if (act) {
      const float ln = xh * __ldg(gamma + i) + __ldg(beta + i);
      const float sg = sigmoidf_(ln);
      g *= sg * (1.f + ln * (1.f - sg));
    }
    if (dln) dln[row * lddln + i] = g;
    const float gg = g * __ldg(gamma + i);
    s1 += gg;
    s2 = fmaf(gg, xh, s2);
  }
  s1 = warp_sum(s1) / (float)U;
  s2 = warp_sum(s2) / (float)U;
  for (int i = lane; i < U; i += 32) {
    const float xh = (xr[i] - mu) * rs;
    float g = dyr[i];
    if (act) {
      const float ln = xh * __ldg(gamma + i) + __ldg(beta + i);
      const float sg = sigmoidf_(ln);
      g *= sg * (1.f + ln * (1.f - sg));
    }
    const float o = rs * (g * __ldg(gamma + i) - s1 - xh * s2);
    dx[row * lddx + i] = o;
    if (dx16) dx16[row * lddx + i] = __float2bfloat16_rn(o);
  }
}

// LayerNorm(+SiLU) backward with the parameter gradients fused in: one pass over dy / x produces dx (+ its bf16 mirror) AND the
// column sums dgamma[u] = sum_m g xhat, dbeta[u] = sum_m g, dbias[u] = sum_m dx (the bias gradient of the Linear in front of the
// LayerNorm), instead of writing g to memory and re-reading it together with x in two more kernels.  Warp per row, rows in
// registers (U % 4 == 0, U <= 128 * NV), grid-stride over rows with per-lane column accumulators; per-block partials go to
// scratch[block][3][U] and are summed in block order by the second kernel (deterministic).
template <int NV>
__global__ void __launch_bounds__(LW_WARPS * 32)
ln_act_bwd_colsum_kernel(int M, int U, const float* __restrict__ dy, int lddy, const float* __restrict__ x, int ldx,
                         const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ mean,
                         const float* __restrict__ rstd, int act, float* __restrict__ dx, int lddx, __nv_bfloat16* __restrict__ dx16,
                         float* __restrict__ partial) {
  extern __shared__ float red[];                     // [LW_WARPS][U]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n4 = U >> 2;
  float4 ag[NV], ab[NV], ax[NV], gm[NV], bt[NV];
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int i = lane + 32 * j;
    ag[j] = ab[j] = ax[j] = make_float4(0.f, 0.f, 0.f, 0.f);
    gm[j] = (i < n4) ? __ldg(reinterpret_cast<const float4*>(gamma) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
    bt[j] = (i < n4) ? __ldg(reinterpret_cast<const float4*>(beta) + i) : make_float4(0.f, 0.f, 0.f, 0.f);
  }
  for (size_t row = (size_t)blockIdx.x * LW_WARPS + warp; row < (size_t)M; row += (size_t)gridDim.x * LW_WARPS) {
    const float4* xr = reinterpret_cast<const float4*>(x + row * ldx);
    const float4* dr = reinterpret_cast<const float4*>(dy + row * lddy);
    const float mu = mean[row], rs = rstd[row];
    float4 xh[NV], gg[NV];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int i = lane + 32 * j;
      if (i < n4) {
        const float4 xv = xr[i], dv = dr[i];
        float4 h, g;
        h.x = (xv.x - mu) * rs; h.y = (xv.y - mu) * rs; h.z = (xv.z - mu) * rs; h.w = (xv.w - mu) * rs;
        g = dv;
        if (act) {
          float ln, sg;
          ln = h.x * gm[j].x + bt[j].x; sg = sigmoidf_(ln); g.x *= sg * (1.f + ln * (1.f - sg));
          ln = h.y * gm[j].y + bt[j].y; sg = sigmoidf_(ln); g.y *= sg * (1.f + ln * (1.f - sg));
          ln = h.z * gm[j].z + bt[j].z; sg = sigmoidf_(ln); g.z *= sg * (1.f + ln * (1.f - sg));
          ln = h.w * gm[j].w + bt[j].w; sg = sigmoidf_(ln); g.w *= sg * (1.f + ln * (1.f - sg));
        }
        ab[j].x += g.x; ab[j].y += g.y; ab[j].z += g.z; ab[j].w += g.w;
        ag[j].x = fmaf(g.x, h.x, ag[j].x); ag[j].y = fmaf(g.y, h.y, ag[j].y); ag[j].z = fmaf(g.z, h.z, ag[j].z); ag[j].w = fmaf(g.w, h.w, ag[j].w);
        g.x *= gm[j].x; g.y *= gm[j].y; g.z *= gm[j].z; g.w *= gm[j].w;
        s1 += (g.x + g.y) + (g.z + g.w);
        s2 = fmaf(g.x, h.x, s2); s2 = fmaf(g.y, h.y, s2); s2 = fmaf(g.z, h.z, s2); s2 = fmaf(g.w, h.w, s2);
        xh[j] = h;
        gg[j] = g;
      }
    }
    s1 = warp_sum(s1) / (float)U;
    s2 = warp_sum(s2) / (float)U;
    float4* oxr = reinterpret_cast<float4*>(dx + row * lddx);
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int i = lane + 32 * j;
      if (i < n4) {
        float4 o;
        o.x = rs * (gg[j].x - s1 - xh[j].x * s2); o.y = rs * (gg[j].y - s1 - xh[j].y * s2);
        o.z = rs * (gg[j].z - s1 - xh[j].z * s2); o.w = rs * (gg[j].w - s1 - xh[j].w * s2);
        oxr[i] = o;
        ax[j].x += o.x; ax[j].y += o.y; ax[j].z += o.z; ax[j].w += o.w;
        if (dx16) {
          __nv_bfloat162 lo = __floats2bfloat162_rn(o.x, o.y), hi = __floats2bfloat162_rn(o.z, o.w);
          uint2 pk;
          pk.x = *reinterpret_cast<uint32_t*>(&lo);
          pk.y = *reinterpret_cast<uint32_t*>(&hi);
          *reinterpret_cast<uint2*>(dx16 + row * lddx + 4 * i) = pk;
        }
      }
    }
  }
  // block partials: warps -> shared -> one [3][U] record per block
#pragma unroll 1
  for (int k = 0; k < 3; ++k) {
    __syncthreads();
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int i = lane + 32 * j;
      if (i < n4) reinterpret_cast<float4*>(red + warp * U)[i] = (k == 0) ? ag[j] : ((k == 1) ? ab[j] : ax[j]);
    }
    __syncthreads();
    for (int u = threadIdx.x; u < U; u += LW_WARPS * 32) {
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < LW_WARPS; ++w) s += red[w * U + u];
      partial[((size_t)blockIdx.x * 3 + k) * U + u] = s;
    }
  }
}
// ---- narrow rows (U <= 128: the channel LayerNorm of the convolution stages, 24..128 channels on up to N*32*32 pixel rows) ----------
// LPR lanes (8, 16 or 32) hold one row, one float4 per lane, so a warp works on 32 / LPR rows at once instead of leaving 3/4 of its
// lanes idle; the row reductions are xor-shuffles inside the LPR-lane group.
template <int LPR>
__device__ __forceinline__ float group_sum(float v) {
#pragma unroll
  for (int o = LPR / 2; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
template <int LPR>
__global__ void __launch_bounds__(LW_WARPS * 32)
ln_act_fwd_sub_kernel(int M, int U, const float* __restrict__ x, int ldx, const float* __restrict__ gamma, const float* __restrict__ beta,
                      float eps, int act, float* __restrict__ y, int ldy, float* __restrict__ mean, float* __restrict__ rstd,
                      __nv_bfloat16* __restrict__ y16) {
  constexpr int RPW = 32 / LPR;
  pdl_sync();
  const int lane = threadIdx.x & 31, sub = lane % LPR, grp = lane / LPR;
  const size_t row = ((size_t)blockIdx.x * LW_WARPS + (threadIdx.x >> 5)) * RPW + grp;
  const bool ok = row < (size_t)M && sub < (U >> 2);
  float4 v = make_float4(0.f, 0.f, 0.f, 0.f), g = v, b = v;
  if (ok) {
    v = reinterpret_cast<const float4*>(x + row * ldx)[sub];
    g = __ldg(reinterpret_cast<const float4*>(gamma) + sub);
    b = __ldg(reinterpret_cast<const float4*>(beta) + sub);
  }
  const float mu = group_sum<LPR>((v.x + v.y) + (v.z + v.w)) / (float)U;
  float q = 0.f;
  if (ok) {
    const float a0 = v.x - mu, a1 = v.y - mu, a2 = v.z - mu, a3 = v.w - mu;
    q = fmaf(a0, a0, fmaf(a1, a1, fmaf(a2, a2, a3 * a3)));
  }
  const float rs = rsqrtf(group_sum<LPR>(q) / (float)U + eps);
  if (!ok) return;
  float4 o;
  o.x = (v.x - mu) * rs * g.x + b.x; o.y = (v.y - mu) * rs * g.y + b.y;
  o.z = (v.z - mu) * rs * g.z + b.z; o.w = (v.w - mu) * rs * g.w + b.w;
  if (act) { o.x = siluf_(o.x); o.y = siluf_(o.y); o.z = siluf_(o.z); o.w = siluf_(o.w); }
  reinterpret_cast<float4*>(y + row * ldy)[sub] = o;
  if (y16) {
    __nv_bfloat162 lo = __floats2bfloat162_rn(o.x, o.y), hi = __floats2bfloat162_rn(o.z, o.w);
    uint2 pk;
    pk.x = *reinterpret_cast<uint32_t*>(&lo);
    pk.y = *reinterpret_cast<uint32_t*>(&hi);
    *reinterpret_cast<uint2*>(y16 + row * ldy + 4 * sub) = pk;
  }
  if (sub == 0) { mean[row] = mu; rstd[row] = rs; }
}
// backward + column sums (see ln_act_bwd_colsum_kernel): every lane keeps the three partial sums of its own four columns
template <int LPR>
__global__ void __launch_bounds__(LW_WARPS * 32)
ln_act_bwd_colsum_sub_kernel(int M, int U, const float* __restrict__ dy, int lddy, const float* __restrict__ x, int ldx,
                             const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ mean,
                             const float* __restrict__ rstd, int act, float* __restrict__ dx, int lddx, __nv_bfloat16* __restrict__ dx16,
                             float* __restrict__ partial) {
  constexpr int RPW = 32 / LPR, NG = LW_WARPS * RPW;
  extern __shared__ float red[];                     // [NG][U]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, sub = lane % LPR, grp = lane / LPR;
  const bool col_ok = sub < (U >> 2);
  float4 ag = make_float4(0.f, 0.f, 0.f, 0.f), ab = ag, ax = ag, gm = ag, bt = ag;
  if (col_ok) {
    gm = __ldg(reinterpret_cast<const float4*>(gamma) + sub);
    bt = __ldg(reinterpret_cast<const float4*>(beta) + sub);
  }
  const size_t stride = (size_t)gridDim.x * NG;
  const size_t rows_pad = ((size_t)M + stride - 1) / stride * stride;       // every lane runs the same number of iterations (shuffles)
  for (size_t row = (size_t)blockIdx.x * NG + warp * RPW + grp; row < rows_pad; row += stride) {
    const bool ok = col_ok && row < (size_t)M;
    float4 h = make_float4(0.f, 0.f, 0.f, 0.f), g = h;
    float rs = 0.f;
    if (ok) {
      const float4 xv = reinterpret_cast<const float4*>(x + row * ldx)[sub];
      g = reinterpret_cast<const float4*>(dy + row * lddy)[sub];
      const float mu = mean[row];
      rs = rstd[row];
      h.x = (xv.x - mu) * rs; h.y = (xv.y - mu) * rs; h.z = (xv.z - mu) * rs; h.w = (xv.w - mu) * rs;
      if (act) {
        float ln, sg;
        ln = h.x * gm.x + bt.x; sg = sigmoidf_(ln); g.x *= sg * (1.f + ln * (1.f - sg));
        ln = h.y * gm.y + bt.y; sg = sigmoidf_(ln); g.y *= sg * (1.f + ln * (1.f - sg));
        ln = h.z * gm.z + bt.z; sg = sigmoidf_(ln); g.z *= sg * (1.f + ln * (1.f - sg));
        ln = h.w * gm.w + bt.w; sg = sigmoidf_(ln); g.w *= sg * (1.f + ln * (1.f - sg));
      }
      ab.x += g.x; ab.y += g.y; ab.z += g.z; ab.w += g.w;
      ag.x = fmaf(g.x, h.x, ag.x); ag.y = fmaf(g.y, h.y, ag.y); ag.z = fmaf(g.z, h.z, ag.z); ag.w = fmaf(g.w, h.w, ag.w);
      g.x *= gm.x; g.y *= gm.y; g.z *= gm.z; g.w *= gm.w;
    }
    const float s1 = group_sum<LPR>((g.x + g.y) + (g.z + g.w)) / (float)U;
    const float s2 = group_sum<LPR>(fmaf(g.x, h.x, fmaf(g.y, h.y, fmaf(g.z, h.z, g.w * h.w)))) / (float)U;
    if (ok) {
      float4 o;
      o.x = rs * (g.x - s1 - h.x * s2); o.y = rs * (g.y - s1 - h.y * s2);
      o.z = rs * (g.z - s1 - h.z * s2); o.w = rs * (g.w - s1 - h.w * s2);
      reinterpret_cast<float4*>(dx + row * lddx)[sub] = o;
      ax.x += o.x; ax.y += o.y; ax.z += o.z; ax.w += o.w;
      if (dx16) {
        __nv_bfloat162 lo = __floats2bfloat162_rn(o.x, o.y), hi = __floats2bfloat162_rn(o.z, o.w);
        uint2 pk;
        pk.x = *reinterpret_cast<uint32_t*>(&lo);
        pk.y = *reinterpret_cast<uint32_t*>(&hi);
        *reinterpret_cast<uint2*>(dx16 + row * lddx + 4 * sub) = pk;
      }
    }
  }
#pragma unroll 1
  for (int k = 0; k < 3; ++k) {
    __syncthreads();
    if (col_ok) reinterpret_cast<float4*>(red + (warp * RPW + grp) * U)[sub] = (k == 0) ? ag : ((k == 1) ? ab : ax);
    __syncthreads();
    for (int u = threadIdx.x; u < U; u += LW_WARPS * 32) {
      float s = 0.f;
#pragma unroll
      for (int w = 0; w < NG; ++w) s += red[w * U + u];
      partial[((size_t)blockIdx.x * 3 + k) * U + u] = s;
    }
  }
}

// 32 columns x 8 row lanes per block: each lane sums every 8th block partial, then a fixed-order reduction in shared memory
__global__ void __launch_bounds__(256) colsum3_stage2_kernel(int U, int nblocks, const float* __restrict__ partial, float* __restrict__ o0,
                                                             float* __restrict__ o1, float* __restrict__ o2) {
  __shared__ float red[8][33];
  const int c = threadIdx.x & 31, rl = threadIdx.x >> 5;
  const int u = blockIdx.x * 32 + c, k = blockIdx.y;
  float* o = (k == 0) ? o0 : ((k == 1) ? o1 : o2);
  if (!o) return;
  float s = 0.f;
  if (u < U)
    for (int b = rl; b < nblocks; b += 8) s += partial[((size_t)b * 3 + k) * U + u];
  red[rl][c] = s;
  __syncthreads();
  if (rl == 0 && u < U) {
    float t = 0.f;
#pragma unroll
    for (int r = 0; r < 8; ++r) t += red[r][c];
    o[u] = t;
  }
}

// MSEDistribution.log_prob (sheeprl/utils/distribution.py:212-221): loss[m] = sum_d (mode - x)^2, dmode = scale*2*(mode-x)
__global__ void __launch_bounds__(256) mse_rows_kernel(int M, int Dd, const float* __restrict__ mode, const float* __restrict__ target,
                                                      float scale, float* __restrict__ loss_rows, float* __restrict__ dmode) {
  __shared__ float red[64];
  const int row = blockIdx.x;
  const float* a = mode + (size_t)row * Dd;
  const float* b = target + (size_t)row * Dd;
  float* d = dmode + (size_t)row * Dd;
  float s = 0.f;
  for (int i = threadIdx.x; i < Dd; i += blockDim.x) {
    const float diff = a[i] - b[i];
    s = fmaf(diff, diff, s);
    d[i] = scale * 2.f * diff;
  }
  s = block_sum(s, red);
  if (threadIdx.x == 0) loss_rows[row] = s;
}

static inline int grid_for(size_t n) {
  size_t g = (n + 255) / 256;
  if (g > 148 * 32) g = 148 * 32;
  if (g < 1) g = 1;
  return (int)g;
}

}  // namespace

extern "C" {

int dv3_im2col_k4s2p1(int N, int H, int W, int C, const float* x, float* col, void* col_bf16, void* stream) {
  if (N <= 0 || H < 2 || W < 2 || (H & 1) || (W & 1) || C <= 0 || (!col && !col_bf16)) return DV3_ERR_ARG;
  if ((size_t)N * (H / 2) > 0x7fffffffu || (size_t)(W / 2) * 16 * C > 0x7fffffffu) return DV3_ERR_TOO_LARGE;
  const int blocks = N * (H / 2);
  if ((C & 3) == 0 && (((uintptr_t)x | (uintptr_t)col) & 15) == 0 && ((uintptr_t)col_bf16 & 7) == 0)
    im2col_k4s2p1_vec4_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(H, W, C / 4, reinterpret_cast<const float4*>(x),
                                                                        reinterpret_cast<float4*>(col), reinterpret_cast<uint2*>(col_bf16));
  else
    im2col_k4s2p1_kernel<<<blocks, 256, 0, (cudaStream_t)stream>>>(H, W, C, x, col, reinterpret_cast<__nv_bfloat16*>(col_bf16));
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_col2im_k4s2p1(int N, int Hi, int Wi, int C, const float* col, const float* bias, float add, float* y, void* stream) {
  if (N <= 0 || Hi <= 0 || Wi <= 0 || C <= 0) return DV3_ERR_ARG;
  if ((size_t)N * Hi * 2 > 0x7fffffffu || (size_t)Wi * 2 * C > 0x7fffffffu) return DV3_ERR_TOO_LARGE;
  const int blocks = N * Hi * 2;
  if ((C & 3) == 0 && (((uintptr_t)col | (uintptr_t)y) & 15) == 0)
    col2im_k4s2p1_kernel<4><<<blocks, 256, 0, (cudaStream_t)stream>>>(Hi, Wi, C, col, bias, add, y);
  else
    col2im_k4s2p1_kernel<1><<<blocks, 256, 0, (cudaStream_t)stream>>>(Hi, Wi, C, col, bias, add, y);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_permute_last2(int A, int B, int C, const float* src, long long lds, float* dst, long long ldd, void* stream) {
  if (A <= 0 || B <= 0 || C <= 0) return DV3_ERR_ARG;
  permute_last2_kernel<<<grid_for((size_t)A * B * C), 256, 0, (cudaStream_t)stream>>>(A, B, C, src, (size_t)lds, dst, (size_t)ldd);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_u8_nchw_to_nhwc(int N, int C, int HW, const unsigned char* src, float scale, float* dst, int Ctot, int c_off, void* stream) {
  if (N <= 0 || C <= 0 || HW <= 0 || c_off < 0 || c_off + C > Ctot) return DV3_ERR_ARG;
  u8_nchw_to_nhwc_kernel<<<grid_for((size_t)N * C * HW), 256, 0, (cudaStream_t)stream>>>(N, C, HW, src, scale, dst, Ctot, c_off);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_ln_act_fwd_rows(int M, int U, const float* x, int ldx, const float* gamma, const float* beta, float eps, int act,
                        float* y, int ldy, float* mean, float* rstd, void* y_bf16, void* stream) {
  if (M <= 0 || U <= 0) return DV3_ERR_ARG;
  const bool vec = (U & 3) == 0 && U <= 1024 && (ldx & 3) == 0 && (ldy & 3) == 0 &&
                   ((((uintptr_t)x | (uintptr_t)y | (uintptr_t)gamma | (uintptr_t)beta) & 15) == 0) && (((uintptr_t)y_bf16 & 7) == 0);
  const dim3 grid((M + LW_WARPS - 1) / LW_WARPS);
  if (vec && U <= 128) {
    const int lpr = U <= 32 ? 8 : (U <= 64 ? 16 : 32), rpb = LW_WARPS * (32 / lpr);
    const dim3 g2((M + rpb - 1) / rpb);
    auto* y16 = reinterpret_cast<__nv_bfloat16*>(y_bf16);
    if (lpr == 8) dv3_launch_pdl(ln_act_fwd_sub_kernel<8>, g2, dim3(LW_WARPS * 32), 0, (cudaStream_t)stream, M, U, x, ldx, gamma, beta, eps, act, y, ldy, mean, rstd, y16);
    else if (lpr == 16) dv3_launch_pdl(ln_act_fwd_sub_kernel<16>, g2, dim3(LW_WARPS * 32), 0, (cudaStream_t)stream, M, U, x, ldx, gamma, beta, eps, act, y, ldy, mean, rstd, y16);
    else dv3_launch_pdl(ln_act_fwd_sub_kernel<32>, g2, dim3(LW_WARPS * 32), 0, (cudaStream_t)stream, M, U, x, ldx, gamma, beta, eps, act, y, ldy, mean, rstd, y16);
  } else if (vec && U <= 512)
    dv3_launch_pdl(ln_act_fwd_warp_vec_kernel<4>, grid, dim3(LW_WARPS * 32), 0, (cudaStream_t)stream, M, U, x, ldx, gamma, beta, eps, act, y, ldy,
                   mean, rstd, reinterpret_cast<__nv_bfloat16*>(y_bf16), (const float*)nullptr, 0, (const int*)nullptr, 0, 0, make_int4(0, 0, 0, 0));
  else if (vec)
    dv3_launch_pdl(ln_act_fwd_warp_vec_kernel<8>, grid, dim3(LW_WARPS * 32), 0, (cudaStream_t)stream, M, U, x, ldx, gamma, beta, eps, act, y, ldy,
                   mean, rstd, reinterpret_cast<__nv_bfloat16*>(y_bf16), (const float*)nullptr, 0, (const int*)nullptr, 0, 0, make_int4(0, 0, 0, 0));
  else
    ln_act_fwd_warp_kernel<<<grid, LW_WARPS * 32, 0, (cudaStream_t)stream>>>(
        M, U, x, ldx, gamma, beta, eps, act, y, ldy, mean, rstd, reinterpret_cast<__nv_bfloat16*>(y_bf16));
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_ln_act_fwd_rows_gather(int M, int U, const float* x, int ldx, const float* gamma, const float* beta, float eps, int act, float* y, int ldy,
                               float* mean, float* rstd, void* y_bf16, const float* add_table, int ld_table, const int* add_idx, int ld_idx,
                               int n_heads, const int* head_off_host, void* stream) {
  if (M <= 0 || U <= 0 || !add_table || !add_idx || n_heads < 1 || !head_off_host) return DV3_ERR_ARG;
  const bool vec = (U & 3) == 0 && U <= 1024 && (ldx & 3) == 0 && (ldy & 3) == 0 && (ld_table & 3) == 0 && n_heads <= 4 &&
                   ((((uintptr_t)x | (uintptr_t)y | (uintptr_t)gamma | (uintptr_t)beta | (uintptr_t)add_table) & 15) == 0) &&
                   (((uintptr_t)y_bf16 & 7) == 0);
  if (!vec) return DV3_ERR_UNSUPPORTED;
  int4 ho = make_int4(head_off_host[0], n_heads > 1 ? head_off_host[1] : 0, n_heads > 2 ? head_off_host[2] : 0, n_heads > 3 ? head_off_host[3] : 0);
  const dim3 grid((M + LW_WARPS - 1) / LW_WARPS);
  if (U <= 512)
    dv3_launch_pdl(ln_act_fwd_warp_vec_kernel<4>, grid, dim3(LW_WARPS * 32), 0, (cudaStream_t)stream, M, U, x, ldx, gamma, beta, eps, act, y, ldy,
                   mean, rstd, reinterpret_cast<__nv_bfloat16*>(y_bf16), add_table, ld_table, add_idx, ld_idx, n_heads, ho);
  else
    dv3_launch_pdl(ln_act_fwd_warp_vec_kernel<8>, grid, dim3(LW_WARPS * 32), 0, (cudaStream_t)stream, M, U, x, ldx, gamma, beta, eps, act, y, ldy,
                   mean, rstd, reinterpret_cast<__nv_bfloat16*>(y_bf16), add_table, ld_table, add_idx, ld_idx, n_heads, ho);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_ln_act_bwd_rows(int M, int U, const float* dy, int lddy, const float* x, int ldx, const float* gamma, const float* beta,
                        const float* mean, const float* rstd, int act, float* dx, int lddx, float* dln, int lddln, void* dx_bf16,
                        void* stream) {
  if (M <= 0 || U <= 0) return DV3_ERR_ARG;
  if ((const void*)dx == (const void*)dy) return DV3_ERR_ARG;
  dv3_launch_pdl(ln_act_bwd_warp_kernel, dim3((M + LW_WARPS - 1) / LW_WARPS), dim3(LW_WARPS * 32), 0, (cudaStream_t)stream, M, U, dy, lddy, x, ldx,
                 gamma, beta, mean, rstd, act, dx, lddx, dln, lddln, reinterpret_cast<__nv_bfloat16*>(dx_bf16));
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_ln_act_bwd_fused(int M, int U, const float* dy, int lddy, const float* x, int ldx, const float* gamma, const float* beta,
                         const float* mean, const float* rstd, int act, float* dx, int lddx, void* dx_bf16, float* dgamma, float* dbeta,
                         float* dbias, float* scratch, size_t scratch_bytes, void* stream) {
  if (M <= 0 || U <= 0 || !scratch) return DV3_ERR_ARG;
  if ((const void*)dx == (const void*)dy) return DV3_ERR_ARG;
  if ((U & 3) || U > 1024 || (ldx & 3) || (lddy & 3) || (lddx & 3) ||
      ((((uintptr_t)x | (uintptr_t)dy | (uintptr_t)dx | (uintptr_t)gamma | (uintptr_t)beta) & 15) != 0) || (((uintptr_t)dx_bf16 & 7) != 0))
    return DV3_ERR_UNSUPPORTED;
  if (U <= 128) {
    const int lpr = U <= 32 ? 8 : (U <= 64 ? 16 : 32), ng = LW_WARPS * (32 / lpr);
    int nb = (M + ng - 1) / ng;
    if (nb > 148 * 6) nb = 148 * 6;
    if (scratch_bytes < (size_t)nb * 3 * U * sizeof(float)) return DV3_ERR_SCRATCH;
    const size_t sm = (size_t)ng * U * sizeof(float);
    auto* d16 = reinterpret_cast<__nv_bfloat16*>(dx_bf16);
    if (lpr == 8) ln_act_bwd_colsum_sub_kernel<8><<<nb, LW_WARPS * 32, sm, (cudaStream_t)stream>>>(M, U, dy, lddy, x, ldx, gamma, beta, mean, rstd, act, dx, lddx, d16, scratch);
    else if (lpr == 16) ln_act_bwd_colsum_sub_kernel<16><<<nb, LW_WARPS * 32, sm, (cudaStream_t)stream>>>(M, U, dy, lddy, x, ldx, gamma, beta, mean, rstd, act, dx, lddx, d16, scratch);
    else ln_act_bwd_colsum_sub_kernel<32><<<nb, LW_WARPS * 32, sm, (cudaStream_t)stream>>>(M, U, dy, lddy, x, ldx, gamma, beta, mean, rstd, act, dx, lddx, d16, scratch);
    DV3_LAUNCH_CHECK();
    colsum3_stage2_kernel<<<dim3((U + 31) / 32, 3), 256, 0, (cudaStream_t)stream>>>(U, nb, scratch, dgamma, dbeta, dbias);
    DV3_LAUNCH_CHECK();
    return DV3_OK;
  }
  int nblocks = (M + LW_WARPS - 1) / LW_WARPS;
  if (nblocks > 296) nblocks = 296;
  if (scratch_bytes < (size_t)nblocks * 3 * U * sizeof(float)) return DV3_ERR_SCRATCH;
  const size_t smem = (size_t)LW_WARPS * U * sizeof(float);
  if (U <= 256)
    ln_act_bwd_colsum_kernel<2><<<nblocks, LW_WARPS * 32, smem, (cudaStream_t)stream>>>(M, U, dy, lddy, x, ldx, gamma, beta, mean, rstd, act, dx, lddx,
                                                                                      reinterpret_cast<__nv_bfloat16*>(dx_bf16), scratch);
  else if (U <= 512)
    ln_act_bwd_colsum_kernel<4><<<nblocks, LW_WARPS * 32, smem, (cudaStream_t)stream>>>(M, U, dy, lddy, x, ldx, gamma, beta, mean, rstd, act, dx, lddx,
                                                                                      reinterpret_cast<__nv_bfloat16*>(dx_bf16), scratch);
  else
    ln_act_bwd_colsum_kernel<8><<<nblocks, LW_WARPS * 32, smem, (cudaStream_t)stream>>>(M, U, dy, lddy, x, ldx, gamma, beta, mean, rstd, act, dx, lddx,
                                                                                      reinterpret_cast<__nv_bfloat16*>(dx_bf16), scratch);
  DV3_LAUNCH_CHECK();
  colsum3_stage2_kernel<<<dim3((U + 31) / 32, 3), 256, 0, (cudaStream_t)stream>>>(U, nblocks, scratch, dgamma, dbeta, dbias);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_mse_rows(int M, int D, const float* mode, const float* target, float scale, float* loss_rows, float* dmode, void* stream) {
  if (M <= 0 || D <= 0) return DV3_ERR_ARG;
  mse_rows_kernel<<<M, 256, 0, (cudaStream_t)stream>>>(M, D, mode, target, scale, loss_rows, dmode);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}

}  // extern "C"
