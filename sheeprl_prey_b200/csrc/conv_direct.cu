// Direct kernels for the two narrow-channel convolutions of the pixel path: the first encoder layer (image channels -> C_out) and the
// last decoder layer (C_in -> image channels).
// Reference: CNNEncoder sheeprl/algos/dreamer_v3/agent.py:28-80 (first Conv2d k4 s2 p1 + LayerNormChannelLast + SiLU), CNNDecoder
// agent.py:136-206 (last ConvTranspose2d k4 s2 p1 + bias, then + 0.5).
// As GEMMs these two layers have K = N = 16 * 3 = 48: one k-block per 128-row tile, 8 192 CTAs that each live for a few microseconds
// (0.30 ms per launch at 16 TFLOP/s, measured), plus an im2col / col2im pass over a 16-tap expansion of the image.  Here every output
// pixel is one thread: the 48 gathered inputs (or the 2 x 2 contributing taps) sit in registers / shared memory, the 48 x 48 weights in
// shared memory (broadcast reads), results leave through a shared-memory tile so that global stores are row-contiguous.
//   conv_narrow_fwd   y[(n,oh,ow)][o] = sum_k col(x)[.][k] w[o][k]  (+ channel LayerNorm + SiLU)     k = (kh*4+kw)*C + c
//   conv_narrow_wgrad dw[a][k]        = sum_rows g[row][a] col(x)[row][k]                            (two-stage, deterministic)
//   deconv_narrow_fwd y[n][oh][ow][c] = bias[c] + add + sum_{taps, ci} x[(n,ih,iw)][ci] w[ci][(kh*4+kw)*C + c]
// The data gradient of the last decoder layer is conv_narrow_fwd on the image-shaped gradient (same arithmetic, no LayerNorm).
#include "common.cuh"
#include <cuda_bf16.h>

namespace {

constexpr int CN_THREADS = 128;

// x[n][2oh-1+kh][2ow-1+kw][c] (0 outside) for the thread's output pixel -> v[16*C]
template <int C>
__device__ __forceinline__ void gather_patch(const float* __restrict__ x, int H, int W, int n, int oh, int ow, float* v) {
#pragma unroll
  for (int kh = 0; kh < 4; ++kh) {
    const int ih = 2 * oh - 1 + kh;
#pragma unroll
    for (int kw = 0; kw < 4; ++kw) {
      const int iw = 2 * ow - 1 + kw;
      const bool ok = ih >= 0 && ih < H && iw >= 0 && iw < W;
      const float* p = x + (((size_t)n * H + ih) * W + iw) * C;
#pragma unroll
      for (int c = 0; c < C; ++c) v[(kh * 4 + kw) * C + c] = ok ? __ldg(p + c) : 0.f;
    }
  }
}

template <int C>
__global__ void __launch_bounds__(CN_THREADS)
conv_narrow_fwd_kernel(int N, int H, int W, int O, const float* __restrict__ x, const float* __restrict__ w, int ldw, float* __restrict__ y,
                       const float* __restrict__ gamma, const float* __restrict__ beta, float eps, float* __restrict__ act,
                       __nv_bfloat16* __restrict__ act16, float* __restrict__ mean, float* __restrict__ rstd) {
  constexpr int K = 16 * C;
  extern __shared__ float sm[];
  float* ws = sm;                        // [O][K]
  float* tile = sm + O * K;              // [CN_THREADS][O + 1]
  const int Ho = H >> 1, Wo = W >> 1, rows = N * Ho * Wo, pitch = O + 1;
  const int tid = threadIdx.x, row0 = blockIdx.x * CN_THREADS, row = row0 + tid;
  for (int i = tid; i < O * K; i += CN_THREADS) ws[i] = __ldg(w + (size_t)(i / K) * ldw + (i % K));
  float v[K];
  {
    const int r = row < rows ? row : rows - 1;
    const int n = r / (Ho * Wo), rem = r - n * Ho * Wo, oh = rem / Wo, ow = rem - oh * Wo;
    gather_patch<C>(x, H, W, n, oh, ow, v);
  }
  __syncthreads();
  float s1 = 0.f, s2 = 0.f;
  for (int o = 0; o < O; ++o) {
    const float4* wr = reinterpret_cast<const float4*>(ws + o * K);
    float a0 = 0.f, a1 = 0.f;
#pragma unroll
    for (int k4 = 0; k4 < K / 4; ++k4) {
      const float4 q = wr[k4];
      a0 = fmaf(v[4 * k4], q.x, a0); a1 = fmaf(v[4 * k4 + 1], q.y, a1);
      a0 = fmaf(v[4 * k4 + 2], q.z, a0); a1 = fmaf(v[4 * k4 + 3], q.w, a1);
    }
    const float s = a0 + a1;
    tile[tid * pitch + o] = s;
    s1 += s;
  }
  __syncthreads();
  // row-contiguous write-out of the convolution output
  for (int i = tid; i < CN_THREADS * O; i += CN_THREADS) {
    const int r = i / O, c = i - r * O;
    if (row0 + r < rows) y[(size_t)(row0 + r) * O + c] = tile[r * pitch + c];
  }
  if (gamma == nullptr) return;
  // channel LayerNorm (+ SiLU) of the thread's own pixel, in place in the tile
  const float mu = s1 / (float)O;
  for (int o = 0; o < O; ++o) { const float d = tile[tid * pitch + o] - mu; s2 = fmaf(d, d, s2); }
  const float rs = rsqrtf(s2 / (float)O + eps);
  __syncthreads();                       // (the write-out above has read every row)
  for (int o = 0; o < O; ++o) {
    const float t = (tile[tid * pitch + o] - mu) * rs * __ldg(gamma + o) + __ldg(beta + o);
    tile[tid * pitch + o] = siluf_(t);
  }
  if (row < rows) { mean[row] = mu; rstd[row] = rs; }
  __syncthreads();
  for (int i = tid; i < CN_THREADS * O; i += CN_THREADS) {
    const int r = i / O, c = i - r * O;
    if (row0 + r < rows) {
      const float t = tile[r * pitch + c];
      const size_t o = (size_t)(row0 + r) * O + c;
      act[o] = t;
      if (act16) act16[o] = __float2bfloat16_rn(t);
    }
  }
}

// ---- weight gradient.  Block = 8 warps; a sub-tile of 64 rows (g [64][A], gathered col [64][K]) is staged in shared memory, warp w
// takes rows w, w + 8, ..; lane (ta, tk) = (lane / 4, lane % 4) owns the TA x 4C block dw[ta*TA .., tk*4C ..] in registers.
template <int C, int TA>
__global__ void __launch_bounds__(256)
conv_narrow_wgrad_kernel(int N, int H, int W, int A, const float* __restrict__ x, const float* __restrict__ g, int rows_per_block,
                         float* __restrict__ partial) {
  constexpr int K = 16 * C, TK = 4 * C, R = 64, AP = 8 * TA;
  __shared__ __align__(16) float gs[R][AP];
  __shared__ __align__(16) float cs[R][K];
  __shared__ float accs[AP * K];
  const int Ho = H >> 1, Wo = W >> 1, rows = N * Ho * Wo;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, ta = lane >> 2, tk = lane & 3;
  const int r_begin = blockIdx.x * rows_per_block, r_end = min(rows, r_begin + rows_per_block);
  float acc[TA][TK];
#pragma unroll
  for (int i = 0; i < TA; ++i)
#pragma unroll
    for (int j = 0; j < TK; ++j) acc[i][j] = 0.f;
  for (int r0 = r_begin; r0 < r_end; r0 += R) {
    __syncthreads();
    for (int i = tid; i < R * AP; i += 256) {
      const int r = i / AP, a = i - r * AP;
      gs[r][a] = (r0 + r < r_end && a < A) ? g[(size_t)(r0 + r) * A + a] : 0.f;
    }
    for (int i = tid; i < R * 16; i += 256) {          // one (row, tap) per item: C contiguous channels
      const int r = i >> 4, tap = i & 15;
      const int rr = min(r0 + r, rows - 1);
      const int n = rr / (Ho * Wo), rem = rr - n * Ho * Wo, oh = rem / Wo, ow = rem - oh * Wo;
      const int ih = 2 * oh - 1 + (tap >> 2), iw = 2 * ow - 1 + (tap & 3);
      const bool ok = r0 + r < r_end && ih >= 0 && ih < H && iw >= 0 && iw < W;
      const float* p = x + (((size_t)n * H + ih) * W + iw) * C;
#pragma unroll
      for (int c = 0; c < C; ++c) cs[r][tap * C + c] = ok ? __ldg(p + c) : 0.f;
    }
    __syncthreads();
#pragma unroll 2
    for (int r = warp; r < R; r += 8) {
      float gv[TA], cv[TK];
#pragma unroll
      for (int i = 0; i < TA; ++i) gv[i] = gs[r][ta * TA + i];
#pragma unroll
      for (int j = 0; j < TK; j += 4) {
        const float4 q = *reinterpret_cast<const float4*>(&cs[r][tk * TK + j]);
        cv[j] = q.x; cv[j + 1] = q.y; cv[j + 2] = q.z; cv[j + 3] = q.w;
      }
#pragma unroll
      for (int i = 0; i < TA; ++i)
#pragma unroll
        for (int j = 0; j < TK; ++j) acc[i][j] = fmaf(gv[i], cv[j], acc[i][j]);
    }
  }
  // warps add their blocks in warp order (fixed order: deterministic), then one [A][K] record per block
  for (int i = tid; i < AP * K; i += 256) accs[i] = 0.f;
  for (int wv = 0; wv < 8; ++wv) {
    __syncthreads();
    if (warp == wv) {
#pragma unroll
      for (int i = 0; i < TA; ++i)
#pragma unroll
        for (int j = 0; j < TK; ++j) accs[(ta * TA + i) * K + tk * TK + j] += acc[i][j];
    }
  }
  __syncthreads();
  for (int i = tid; i < A * K; i += 256) partial[(size_t)blockIdx.x * A * K + i] = accs[i];
}
__global__ void __launch_bounds__(256) conv_narrow_wgrad_stage2(int AK, int K, int nblocks, const float* __restrict__ partial,
                                                               float* __restrict__ dw, int lddw) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= AK) return;
  float s = 0.f;
  for (int b = 0; b < nblocks; ++b) s += partial[(size_t)b * AK + i];
  dw[(size_t)(i / K) * lddw + (i % K)] = s;
}

// ---- transposed convolution to a narrow-channel image.  Block = (n, j): output rows 2j and 2j + 1, which read input rows j-1, j, j+1.
template <int C>
__global__ void __launch_bounds__(128)
deconv_narrow_fwd_kernel(int Hi, int Wi, int Ci, const float* __restrict__ x, const float* __restrict__ w, int ldw,
                         const float* __restrict__ bias, float add, float* __restrict__ y) {
  extern __shared__ float sm[];
  const int xp = Ci + 4;                          // pixel pitch in xs (spreads consecutive pixels over the banks)
  float* xs = sm;                                 // [3][Wi][xp]
  float* ws = sm + 3 * Wi * xp;                   // [16][C][Ci]
  const int n = blockIdx.x / Hi, j = blockIdx.x - n * Hi;
  const int Wo = 2 * Wi, Ho = 2 * Hi, tid = threadIdx.x;
  const int c4 = Ci >> 2;
  for (int i = tid; i < 3 * Wi * c4; i += blockDim.x) {
    const int q = i % c4, px = (i / c4) % Wi, rl = i / (c4 * Wi);
    const int ih = j - 1 + rl;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (ih >= 0 && ih < Hi) v = __ldg(reinterpret_cast<const float4*>(x + (((size_t)n * Hi + ih) * Wi + px) * Ci) + q);
    *reinterpret_cast<float4*>(xs + (rl * Wi + px) * xp + 4 * q) = v;
  }
  for (int i = tid; i < 16 * C * Ci; i += blockDim.x) {
    const int ci = i % Ci, c = (i / Ci) % C, tap = i / (Ci * C);
    ws[i] = __ldg(w + (size_t)ci * ldw + tap * C + c);
  }
  __syncthreads();
  for (int t = tid; t < 2 * Wo; t += blockDim.x) {
    const int rr = t / Wo, ow = t - rr * Wo, oh = 2 * j + rr;
    float acc[C];
#pragma unroll
    for (int c = 0; c < C; ++c) acc[c] = add + (bias ? __ldg(bias + c) : 0.f);
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      const int ih = ((oh + 1) >> 1) - a, kh = oh + 1 - 2 * ih;            // (as in col2im_k4s2p1_kernel)
      if (ih < 0 || ih >= Hi) continue;
#pragma unroll
      for (int b = 0; b < 2; ++b) {
        const int iw = ((ow + 1) >> 1) - b, kw = ow + 1 - 2 * iw;
        if (iw < 0 || iw >= Wi) continue;
        const float4* xv = reinterpret_cast<const float4*>(xs + ((ih - (j - 1)) * Wi + iw) * xp);
        const float* wt = ws + (kh * 4 + kw) * C * Ci;
        for (int q = 0; q < c4; ++q) {
          const float4 xx = xv[q];
#pragma unroll
          for (int c = 0; c < C; ++c) {
            const float4 ww = *reinterpret_cast<const float4*>(wt + c * Ci + 4 * q);
            acc[c] = fmaf(xx.x, ww.x, fmaf(xx.y, ww.y, fmaf(xx.z, ww.z, fmaf(xx.w, ww.w, acc[c]))));
          }
        }
      }
    }
    float* yo = y + (((size_t)n * Ho + oh) * Wo + ow) * C;
#pragma unroll
    for (int c = 0; c < C; ++c) yo[c] = acc[c];
  }
}

}  // namespace

extern "C" {

int dv3_conv_k4s2p1_narrow_fwd(int N, int H, int W, int C, int O, const float* x, const float* w, int ldw, float* y, const float* gamma,
                               const float* beta, float eps, float* act, void* act_bf16, float* mean, float* rstd, void* stream) {
  if (N <= 0 || H < 2 || W < 2 || (H & 1) || (W & 1) || !x || !w || !y) return DV3_ERR_ARG;
  if (gamma && (!beta || !act || !mean || !rstd)) return DV3_ERR_ARG;
  if (C < 1 || C > 4 || C == 2 || O < 1 || O > 96 || ((uintptr_t)w & 3)) return DV3_ERR_UNSUPPORTED;
  const size_t rows = (size_t)N * (H / 2) * (W / 2);
  if (rows > 0x7fffffffu) return DV3_ERR_TOO_LARGE;
  const int K = 16 * C;
  const size_t smem = ((size_t)O * K + (size_t)CN_THREADS * (O + 1)) * sizeof(float);
  const int blocks = (int)((rows + CN_THREADS - 1) / CN_THREADS);
  auto* a16 = reinterpret_cast<__nv_bfloat16*>(act_bf16);
  cudaError_t e = cudaSuccess;
#define DV3_CNF(CC)                                                                                                                   \
  {                                                                                                                                   \
    if (smem > 48 * 1024) e = cudaFuncSetAttribute(conv_narrow_fwd_kernel<CC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    if (e == cudaSuccess)                                                                                                             \
      conv_narrow_fwd_kernel<CC><<<blocks, CN_THREADS, smem, (cudaStream_t)stream>>>(N, H, W, O, x, w, ldw, y, gamma, beta, eps, act, a16, mean, rstd); \
  }
  if (C == 1) DV3_CNF(1) else if (C == 3) DV3_CNF(3) else DV3_CNF(4)
#undef DV3_CNF
  if (e != cudaSuccess) return (int)e;
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}

size_t dv3_conv_k4s2p1_narrow_wgrad_scratch_bytes(int N, int H, int W, int C, int A) {
  const size_t rows = (size_t)N * (H / 2) * (W / 2);
  size_t nb = (rows + 511) / 512;
  if (nb > 592) nb = 592;
  return nb * (size_t)A * 16 * C * sizeof(float);
}
int dv3_conv_k4s2p1_narrow_wgrad(int N, int H, int W, int C, int A, const float* x, const float* g, float* dw, int lddw, float* scratch,
                                 size_t scratch_bytes, void* stream) {
  if (N <= 0 || H < 2 || W < 2 || (H & 1) || (W & 1) || !x || !g || !dw || !scratch) return DV3_ERR_ARG;
  if (C < 1 || C > 4 || C == 2 || A < 1 || A > 64 || (C == 4 && A > 48)) return DV3_ERR_UNSUPPORTED;
  const size_t rows = (size_t)N * (H / 2) * (W / 2);
  if (rows > 0x7fffffffu) return DV3_ERR_TOO_LARGE;
  if (scratch_bytes < dv3_conv_k4s2p1_narrow_wgrad_scratch_bytes(N, H, W, C, A)) return DV3_ERR_SCRATCH;
  int nb = (int)((rows + 511) / 512);
  if (nb > 592) nb = 592;
  int rpb = (int)((rows + nb - 1) / nb);
  rpb = (rpb + 63) / 64 * 64;
  nb = (int)((rows + rpb - 1) / rpb);
  const int K = 16 * C;
  cudaStream_t st = (cudaStream_t)stream;
#define DV3_CNW(CC, TT) conv_narrow_wgrad_kernel<CC, TT><<<nb, 256, 0, st>>>(N, H, W, A, x, g, rpb, scratch)
  const int TA = A <= 32 ? 4 : (A <= 48 ? 6 : 8);
  if (C == 1) { if (TA == 4) DV3_CNW(1, 4); else if (TA == 6) DV3_CNW(1, 6); else DV3_CNW(1, 8); }
  else if (C == 3) { if (TA == 4) DV3_CNW(3, 4); else if (TA == 6) DV3_CNW(3, 6); else DV3_CNW(3, 8); }
  else { if (TA == 4) DV3_CNW(4, 4); else DV3_CNW(4, 6); }
#undef DV3_CNW
  DV3_LAUNCH_CHECK();
  conv_narrow_wgrad_stage2<<<(A * K + 255) / 256, 256, 0, st>>>(A * K, K, nb, scratch, dw, lddw);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}

int dv3_deconv_k4s2p1_narrow_fwd(int N, int Hi, int Wi, int Ci, int C, const float* x, const float* w, int ldw, const float* bias, float add,
                                 float* y, void* stream) {
  if (N <= 0 || Hi <= 0 || Wi <= 0 || !x || !w || !y) return DV3_ERR_ARG;
  if (C < 1 || C > 4 || C == 2 || Ci < 4 || (Ci & 3) || ((uintptr_t)x & 15)) return DV3_ERR_UNSUPPORTED;
  if ((size_t)N * Hi > 0x7fffffffu) return DV3_ERR_TOO_LARGE;
  const size_t smem = ((size_t)3 * Wi * (Ci + 4) + (size_t)16 * C * Ci) * sizeof(float);
  if (smem > 200 * 1024) return DV3_ERR_UNSUPPORTED;
  cudaError_t e = cudaSuccess;
#define DV3_DNF(CC)                                                                                                                     \
  {                                                                                                                                     \
    if (smem > 48 * 1024) e = cudaFuncSetAttribute(deconv_narrow_fwd_kernel<CC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
    if (e == cudaSuccess) deconv_narrow_fwd_kernel<CC><<<N * Hi, 128, smem, (cudaStream_t)stream>>>(Hi, Wi, Ci, x, w, ldw, bias, add, y); \
  }
  if (C == 1) DV3_DNF(1) else if (C == 3) DV3_DNF(3) else DV3_DNF(4)
#undef DV3_DNF
  if (e != cudaSuccess) return (int)e;
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}

}  // extern "C"
