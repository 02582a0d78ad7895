// Row-wise fused kernels: LayerNorm(+SiLU) forward/backward, LayerNorm-GRU gates forward/backward,
// column reductions for LN-parameter / bias gradients, symlog, strided copy / transpose, reset masking.
// Reference arithmetic: nn.LayerNorm + nn.SiLU inside sheeprl/models/models.py:83-99 (miniblock order
// sheeprl/utils/model.py:71-87); LayerNormGRUCell sheeprl/models/models.py:388-395; symlog
// sheeprl/utils/utils.py:120-121; reset masking sheeprl/algos/dreamer_v3/agent.py:379-384.
#include "common.cuh"
#include <cuda_bf16.h>

namespace {

constexpr int ROW_THREADS = 256;

// ---------------------------------------------------------------- LayerNorm (+SiLU) forward
__global__ void __launch_bounds__(ROW_THREADS)
ln_act_fwd_kernel(int U, const float* __restrict__ x, int ldx, const float* __restrict__ gamma,
                  const float* __restrict__ beta, float eps, int act, float* __restrict__ y, int ldy,
                  float* __restrict__ mean, float* __restrict__ rstd, __nv_bfloat16* __restrict__ y16) {
  __shared__ float red[64];
  const int row = blockIdx.x;
  const float* xr = x + (size_t)row * ldx;
  float s = 0.f;
  for (int i = threadIdx.x; i < U; i += blockDim.x) s += xr[i];
  const float mu = block_sum(s, red) / (float)U;
  float v = 0.f;
  for (int i = threadIdx.x; i < U; i += blockDim.x) { const float d = xr[i] - mu; v = fmaf(d, d, v); }
  const float var = block_sum(v, red) / (float)U;
  const float rs = rsqrtf(var + eps);
  float* yr = y + (size_t)row * ldy;
  for (int i = threadIdx.x; i < U; i += blockDim.x) {
    float ln = (xr[i] - mu) * rs * gamma[i] + beta[i];
    const float o = act ? siluf_(ln) : ln;
    yr[i] = o;
    if (y16) y16[(size_t)row * ldy + i] = __float2bfloat16_rn(o);      // bf16 mirror for the next contraction
  }
  if (threadIdx.x == 0) { mean[row] = mu; rstd[row] = rs; }
}

// ---------------------------------------------------------------- LayerNorm (+SiLU) backward (rows)
__global__ void __launch_bounds__(ROW_THREADS)
ln_act_bwd_kernel(int U, const float* __restrict__ dy, int lddy, const float* __restrict__ x, int ldx,
                  const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ mean,
                  const float* __restrict__ rstd, int act, float* __restrict__ dx, int lddx,
                  float* __restrict__ dln, int lddln, __nv_bfloat16* __restrict__ dx16) {
  __shared__ float red[64];
  const int row = blockIdx.x;
  const float* xr = x + (size_t)row * ldx;
  const float* dyr = dy + (size_t)row * lddy;
  const float mu = mean[row], rs = rstd[row];
  float s1 = 0.f, s2 = 0.f;
  for (int i = threadIdx.x; i < U; i += blockDim.x) {
    const float xh = (xr[i] - mu) * rs;
    float g = dyr[i];
    if (act) {
      const float ln = xh * gamma[i] + beta[i];
      const float sg = sigmoidf_(ln);
      g *= sg * (1.f + ln * (1.f - sg));
    }
    if (dln) dln[(size_t)row * lddln + i] = g;
    const float gg = g * gamma[i];
    s1 += gg;
    s2 = fmaf(gg, xh, s2);
  }
  block_sum2(s1, s2, red);
  const float m1 = s1 / (float)U, m2 = s2 / (float)U;
  __syncthreads();
  for (int i = threadIdx.x; i < U; i += blockDim.x) {
    const float xh = (xr[i] - mu) * rs;
    float g = dyr[i];
    if (act) {
      const float ln = xh * gamma[i] + beta[i];
      const float sg = sigmoidf_(ln);
      g *= sg * (1.f + ln * (1.f - sg));
    }
    const float gg = g * gamma[i];
    const float o = rs * (gg - m1 - xh * m2);
    dx[(size_t)row * lddx + i] = o;
    if (dx16) dx16[(size_t)row * lddx + i] = __float2bfloat16_rn(o);
  }
}

// ---------------------------------------------------------------- column reductions (two stages, scratch)
// stage 1: partial[chunk][0][u] = sum_{m in chunk} d[m,u]*xhat[m,u]; partial[chunk][1][u] = sum d[m,u]
constexpr int CR_COLS = 32, CR_ROWLANES = 8;
__global__ void __launch_bounds__(CR_COLS * CR_ROWLANES)
col_reduce_stage1(int M, int U, const float* __restrict__ d, int ldd, const float* __restrict__ x, int ldx,
                  const float* __restrict__ mean, const float* __restrict__ rstd, int rows_per_chunk,
                  float* __restrict__ partial) {
  __shared__ float sa[CR_ROWLANES][CR_COLS + 1], sb[CR_ROWLANES][CR_COLS + 1];
  pdl_sync();
  const int c = threadIdx.x % CR_COLS, rl = threadIdx.x / CR_COLS;
  const int u = blockIdx.x * CR_COLS + c;
  const int r0 = blockIdx.y * rows_per_chunk, r1 = min(M, r0 + rows_per_chunk);
  float a = 0.f, b = 0.f;
  if (u < U) {
    for (int m = r0 + rl; m < r1; m += CR_ROWLANES) {
      const float g = d[(size_t)m * ldd + u];
      b += g;
      if (x) a = fmaf(g, (x[(size_t)m * ldx + u] - mean[m]) * rstd[m], a);
    }
  }
  sa[rl][c] = a; sb[rl][c] = b;
  __syncthreads();
  if (rl == 0 && u < U) {
#pragma unroll
    for (int j = 1; j < CR_ROWLANES; ++j) { a += sa[j][c]; b += sb[j][c]; }
    partial[((size_t)blockIdx.y * 2 + 0) * U + u] = a;
    partial[((size_t)blockIdx.y * 2 + 1) * U + u] = b;
  }
}
// Narrow matrices (U <= 8: the image-channel bias gradient of the last transposed convolution, N*64*64 rows x 3 columns): one row per
// thread with the U partial sums in registers -- the 32-column tiling above would leave 29 of 32 lanes idle.
constexpr int CRN_MAX_U = 8;
__global__ void __launch_bounds__(256)
col_reduce_narrow_stage1(int M, int U, const float* __restrict__ d, int ldd, const float* __restrict__ x, int ldx,
                         const float* __restrict__ mean, const float* __restrict__ rstd, int rows_per_chunk, float* __restrict__ partial) {
  __shared__ float sa[8][CRN_MAX_U], sb[8][CRN_MAX_U];
  pdl_sync();
  const int r0 = blockIdx.x * rows_per_chunk, r1 = min(M, r0 + rows_per_chunk);
  float a[CRN_MAX_U], b[CRN_MAX_U];
#pragma unroll
  for (int u = 0; u < CRN_MAX_U; ++u) a[u] = b[u] = 0.f;
  for (int m = r0 + threadIdx.x; m < r1; m += blockDim.x) {
    const float mu = x ? mean[m] : 0.f, rs = x ? rstd[m] : 0.f;
#pragma unroll
    for (int u = 0; u < CRN_MAX_U; ++u) {
      if (u < U) {
        const float g = d[(size_t)m * ldd + u];
        b[u] += g;
        if (x) a[u] = fmaf(g, (x[(size_t)m * ldx + u] - mu) * rs, a[u]);
      }
    }
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int u = 0; u < CRN_MAX_U; ++u) {
    const float ta = warp_sum(a[u]), tb = warp_sum(b[u]);
    if (lane == 0) { sa[warp][u] = ta; sb[warp][u] = tb; }
  }
  __syncthreads();
  if (threadIdx.x < U) {
    float ta = 0.f, tb = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) { ta += sa[w][threadIdx.x]; tb += sb[w][threadIdx.x]; }
    partial[((size_t)blockIdx.x * 2 + 0) * U + threadIdx.x] = ta;
    partial[((size_t)blockIdx.x * 2 + 1) * U + threadIdx.x] = tb;
  }
}
__global__ void __launch_bounds__(256)
col_reduce_narrow_stage2(int U, int nchunks, const float* __restrict__ partial, float* __restrict__ out_a, float* __restrict__ out_b) {
  __shared__ float red[64];
  pdl_sync();
  for (int k = 0; k < 2 * U; ++k) {
    float s = 0.f;
    for (int ch = threadIdx.x; ch < nchunks; ch += blockDim.x) s += partial[(size_t)ch * 2 * U + k];
    s = block_sum(s, red);
    float* o = (k < U) ? out_a : out_b;
    if (threadIdx.x == 0 && o) o[k < U ? k : k - U] = s;
    __syncthreads();
  }
}
__global__ void col_reduce_stage2(int U, int nchunks, const float* __restrict__ partial, float* __restrict__ out_a,
                                  float* __restrict__ out_b) {
  pdl_sync();
  const int u = blockIdx.x * blockDim.x + threadIdx.x;
  if (u >= U) return;
  float a = 0.f, b = 0.f;
  for (int ch = 0; ch < nchunks; ++ch) {
    a += partial[((size_t)ch * 2 + 0) * U + u];
    b += partial[((size_t)ch * 2 + 1) * U + u];
  }
  if (out_a) out_a[u] = a;
  if (out_b) out_b[u] = b;
}

// ---------------------------------------------------------------- GRU gates
__global__ void __launch_bounds__(ROW_THREADS)
gru_gate_fwd_kernel(int D, const float* __restrict__ gp, int ldg, const float* __restrict__ lnw,
                    const float* __restrict__ lnb, const float* __restrict__ h_in, int ldhi, float* __restrict__ h_out,
                    int ldho, float* __restrict__ mean, float* __restrict__ rstd, __nv_bfloat16* __restrict__ h16) {
  __shared__ float red[64];
  const int row = blockIdx.x, U3 = 3 * D;
  const float* g = gp + (size_t)row * ldg;
  float s = 0.f;
  for (int i = threadIdx.x; i < U3; i += blockDim.x) s += g[i];
  const float mu = block_sum(s, red) / (float)U3;
  float v = 0.f;
  for (int i = threadIdx.x; i < U3; i += blockDim.x) { const float d = g[i] - mu; v = fmaf(d, d, v); }
  const float rs = rsqrtf(block_sum(v, red) / (float)U3 + 1e-5f);
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    const float r = sigmoidf_((g[d] - mu) * rs * lnw[d] + lnb[d]);
    const float cp = (g[D + d] - mu) * rs * lnw[D + d] + lnb[D + d];
    const float c = tanhf(r * cp);
    const float u = sigmoidf_((g[2 * D + d] - mu) * rs * lnw[2 * D + d] + lnb[2 * D + d] - 1.f);
    const float h = h_in[(size_t)row * ldhi + d];
    const float o = u * c + (1.f - u) * h;
    h_out[(size_t)row * ldho + d] = o;
    if (h16) h16[(size_t)row * ldho + d] = __float2bfloat16_rn(o);
  }
  if (threadIdx.x == 0) { mean[row] = mu; rstd[row] = rs; }
}

// Warp-per-row variant for many rows (imagination rollout: 1024 rows per call): the 3D gate pre-activations of a row are held in
// registers (D <= 32 * GD), one pass over global memory, warp-shuffle reductions only.
template <int GD>
__global__ void __launch_bounds__(256)
gru_gate_fwd_warp_kernel(int R, int D, const float* __restrict__ gp, int ldg, const float* __restrict__ lnw,
                         const float* __restrict__ lnb, const float* __restrict__ h_in, int ldhi, float* __restrict__ h_out,
                         int ldho, float* __restrict__ mean, float* __restrict__ rstd, __nv_bfloat16* __restrict__ h16) {
  pdl_sync();
  const int lane = threadIdx.x & 31;
  const size_t row = (size_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= (size_t)R) return;
  const float* g = gp + row * ldg;
  float v[3][GD];
  float s = 0.f;
#pragma unroll
  for (int j = 0; j < GD; ++j) {
    const int d = lane + 32 * j;
#pragma unroll
    for (int q = 0; q < 3; ++q) { v[q][j] = (d < D) ? g[q * D + d] : 0.f; s += v[q][j]; }
  }
  const float mu = warp_sum(s) / (float)(3 * D);
  float qq = 0.f;
#pragma unroll
  for (int j = 0; j < GD; ++j)
    if (lane + 32 * j < D) {
#pragma unroll
      for (int q = 0; q < 3; ++q) { const float dl = v[q][j] - mu; qq = fmaf(dl, dl, qq); }
    }
  const float rs = rsqrtf(warp_sum(qq) / (float)(3 * D) + 1e-5f);
#pragma unroll
  for (int j = 0; j < GD; ++j) {
    const int d = lane + 32 * j;
    if (d < D) {
      const float r = sigmoidf_((v[0][j] - mu) * rs * __ldg(lnw + d) + __ldg(lnb + d));
      const float cp = (v[1][j] - mu) * rs * __ldg(lnw + D + d) + __ldg(lnb + D + d);
      const float c = tanhf(r * cp);
      const float u = sigmoidf_((v[2][j] - mu) * rs * __ldg(lnw + 2 * D + d) + __ldg(lnb + 2 * D + d) - 1.f);
      const float o = u * c + (1.f - u) * h_in[row * ldhi + d];
      h_out[row * ldho + d] = o;
      if (h16) h16[row * ldho + d] = __float2bfloat16_rn(o);
    }
  }
  if (lane == 0) { mean[row] = mu; rstd[row] = rs; }
}

__global__ void __launch_bounds__(ROW_THREADS)
gru_gate_bwd_kernel(int D, const float* __restrict__ dh_out, int lddho, const float* __restrict__ gp, int ldg,
                    const float* __restrict__ lnw, const float* __restrict__ lnb, const float* __restrict__ mean,
                    const float* __restrict__ rstd, const float* __restrict__ h_in, int ldhi,
                    float* __restrict__ dgp, int lddg, float* __restrict__ dln, int lddln, float* __restrict__ dh_in,
                    int lddhi) {
  __shared__ float red[64];
  const int row = blockIdx.x, U3 = 3 * D;
  const float* g = gp + (size_t)row * ldg;
  const float mu = mean[row], rs = rstd[row];
  float* dl = dln + (size_t)row * lddln;
  float s1 = 0.f, s2 = 0.f;
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    const float xr = (g[d] - mu) * rs, xc = (g[D + d] - mu) * rs, xu = (g[2 * D + d] - mu) * rs;
    const float r = sigmoidf_(xr * lnw[d] + lnb[d]);
    const float cp = xc * lnw[D + d] + lnb[D + d];
    const float c = tanhf(r * cp);
    const float u = sigmoidf_(xu * lnw[2 * D + d] + lnb[2 * D + d] - 1.f);
    const float h = h_in[(size_t)row * ldhi + d];
    const float dho = dh_out[(size_t)row * lddho + d];
    const float du = dho * (c - h), dc = dho * u;
    dh_in[(size_t)row * lddhi + d] = dho * (1.f - u);
    const float dpc = dc * (1.f - c * c);
    const float g_r = dpc * cp * r * (1.f - r);
    const float g_c = dpc * r;
    const float g_u = du * u * (1.f - u);
    dl[d] = g_r; dl[D + d] = g_c; dl[2 * D + d] = g_u;
    const float a = g_r * lnw[d], b = g_c * lnw[D + d], e = g_u * lnw[2 * D + d];
    s1 += a + b + e;
    s2 += a * xr + b * xc + e * xu;
  }
  block_sum2(s1, s2, red);
  const float m1 = s1 / (float)U3, m2 = s2 / (float)U3;
  __syncthreads();
  float* o = dgp + (size_t)row * lddg;
  for (int d = threadIdx.x; d < D; d += blockDim.x) {  // same thread wrote dl[d], dl[D+d], dl[2D+d]
#pragma unroll
    for (int j = 0; j < 3; ++j) {
      const int i = j * D + d;
      const float xh = (g[i] - mu) * rs;
      o[i] = rs * (dl[i] * lnw[i] - m1 - xh * m2);
    }
  }
}

// ---------------------------------------------------------------- elementwise / data movement
__global__ void symlog_kernel(size_t n, const float* __restrict__ x, float* __restrict__ y) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    y[i] = symlogf_(x[i]);
}
__global__ void copy2d_kernel(int R, int C, const float* __restrict__ s, int lds, float* __restrict__ d, int ldd) {
  pdl_sync();
  const size_t n = (size_t)R * C;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int r = (int)(i / C), c = (int)(i % C);
    d[(size_t)r * ldd + c] = s[(size_t)r * lds + c];
  }
}
__global__ void transpose_kernel(int R, int C, const float* __restrict__ s, int lds, float* __restrict__ d, int ldd) {
  __shared__ float tile[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int r = r0 + j, c = c0 + threadIdx.x;
    tile[j][threadIdx.x] = (r < R && c < C) ? s[(size_t)r * lds + c] : 0.f;
  }
  __syncthreads();
  for (int j = threadIdx.y; j < 32; j += blockDim.y) {
    const int c = c0 + j, r = r0 + threadIdx.x;  // output row = c, col = r
    if (c < C && r < R) d[(size_t)c * ldd + r] = tile[threadIdx.x][j];
  }
}
// reset masking: rows with first==1 get h=0, z=zreset; others copy (prev pointers null => all rows reset)
__global__ void scan_prep_kernel(int B, int D, int S, const float* __restrict__ first, const float* __restrict__ h_prev,
                                 int ldhp, const float* __restrict__ z_prev, int ldzp, const float* __restrict__ zreset,
                                 float* __restrict__ h_in, int ldhi, float* __restrict__ z_in, int ldzi) {
  const int b = blockIdx.x;
  const bool reset = (h_prev == nullptr) || (first[b] != 0.f);
  for (int i = threadIdx.x; i < D; i += blockDim.x) h_in[(size_t)b * ldhi + i] = reset ? 0.f : h_prev[(size_t)b * ldhp + i];
  for (int i = threadIdx.x; i < S; i += blockDim.x) z_in[(size_t)b * ldzi + i] = reset ? zreset[i] : z_prev[(size_t)b * ldzp + i];
}
__global__ void scan_carry_kernel(int B, int D, int S, const float* __restrict__ first, const float* __restrict__ dz_prev,
                                  int lddz, const float* __restrict__ dh_a, int lda, const float* __restrict__ dh_b, int ldb,
                                  float* __restrict__ dz_dst, int ldzd, float* __restrict__ dh_dst, int ldhd) {
  const int b = blockIdx.x;
  if (first[b] != 0.f) return;
  for (int i = threadIdx.x; i < S; i += blockDim.x) dz_dst[(size_t)b * ldzd + i] += dz_prev[(size_t)b * lddz + i];
  for (int i = threadIdx.x; i < D; i += blockDim.x)
    dh_dst[(size_t)b * ldhd + i] += dh_a[(size_t)b * lda + i] + dh_b[(size_t)b * ldb + i];
}
// out[t,b,:] = (1-first[t,b]) * actions[t-1,b,:], zeros at t == 0
__global__ void mask_shift_actions_kernel(int T, int B, int A, const float* __restrict__ actions,
                                          const float* __restrict__ first, float* __restrict__ out, int ldo) {
  const int tb = blockIdx.x, t = tb / B;
  const bool zero = (t == 0) || (first[tb] != 0.f);
  for (int i = threadIdx.x; i < A; i += blockDim.x)
    out[(size_t)tb * ldo + i] = zero ? 0.f : actions[(size_t)(tb - B) * A + i];
}
__global__ void ema_kernel(size_t n, float* __restrict__ tgt, const float* __restrict__ src, float tau) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    tgt[i] = tau * src[i] + (1.f - tau) * tgt[i];
}
__global__ void fill_kernel(size_t n, float* __restrict__ p, float v) {
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = v;
}

static inline int grid_for(size_t n, int threads) {
  size_t g = (n + threads - 1) / threads;
  if (g > 148 * 16) g = 148 * 16;
  if (g < 1) g = 1;
  return (int)g;
}

}  // namespace

extern "C" {

int dv3_ln_act_fwd(int M, int U, const float* x, int ldx, const float* gamma, const float* beta, float eps, int act,
                   float* y, int ldy, float* mean, float* rstd, void* y_bf16, void* stream) {
  if (M <= 0 || U <= 0) return DV3_ERR_ARG;
  ln_act_fwd_kernel<<<M, ROW_THREADS, 0, (cudaStream_t)stream>>>(U, x, ldx, gamma, beta, eps, act, y, ldy, mean, rstd,
                                                                reinterpret_cast<__nv_bfloat16*>(y_bf16));
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}

int dv3_ln_act_bwd(int M, int U, const float* dy, int lddy, const float* x, int ldx, const float* gamma,
                   const float* beta, const float* mean, const float* rstd, int act, float* dx, int lddx, float* dln,
                   int lddln, void* dx_bf16, void* stream) {
  if (M <= 0 || U <= 0) return DV3_ERR_ARG;
  if ((const void*)dx == (const void*)dy) return DV3_ERR_ARG;  // two passes read dy after dx is written
  ln_act_bwd_kernel<<<M, ROW_THREADS, 0, (cudaStream_t)stream>>>(U, dy, lddy, x, ldx, gamma, beta, mean, rstd, act, dx,
                                                                lddx, dln, lddln, reinterpret_cast<__nv_bfloat16*>(dx_bf16));
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}

// out_a[u] = sum_m d[m,u]*xhat[m,u] (x may be null -> skipped), out_b[u] = sum_m d[m,u].
// scratch: >= 2*U*nchunks floats, nchunks = min(64, ceil(M/64)).
size_t dv3_col_reduce_scratch_bytes(int M, int U) {
  int nchunks = (M + 63) / 64;
  if (nchunks > 64) nchunks = 64;
  return (size_t)nchunks * 2 * U * sizeof(float);
}
int dv3_col_reduce(int M, int U, const float* d, int ldd, const float* x, int ldx, const float* mean, const float* rstd,
                   float* out_a, float* out_b, float* scratch, size_t scratch_bytes, void* stream) {
  if (M <= 0 || U <= 0 || !d || !scratch) return DV3_ERR_ARG;
  if (scratch_bytes < dv3_col_reduce_scratch_bytes(M, U)) return DV3_ERR_SCRATCH;
  if (U <= CRN_MAX_U && M >= 65536) {
    int nch = 148 * 4;
    if (scratch_bytes < (size_t)nch * 2 * U * sizeof(float)) nch = 64;
    const int rpc = (M + nch - 1) / nch;
    dv3_launch_pdl(col_reduce_narrow_stage1, dim3(nch), dim3(256), 0, (cudaStream_t)stream, M, U, d, ldd, x, ldx, mean, rstd, rpc, scratch);
    DV3_LAUNCH_CHECK();
    dv3_launch_pdl(col_reduce_narrow_stage2, dim3(1), dim3(256), 0, (cudaStream_t)stream, U, nch, (const float*)scratch, out_a, out_b);
    DV3_LAUNCH_CHECK();
    return DV3_OK;
  }
  int nchunks = (M + 63) / 64;
  if (nchunks > 64) nchunks = 64;
  const int rows_per_chunk = (M + nchunks - 1) / nchunks;
  dim3 g1((U + CR_COLS - 1) / CR_COLS, nchunks);
  dv3_launch_pdl(col_reduce_stage1, g1, dim3(CR_COLS * CR_ROWLANES), 0, (cudaStream_t)stream, M, U, d, ldd, x, ldx, mean, rstd,
                                                                          rows_per_chunk, scratch);
  DV3_LAUNCH_CHECK();
  dv3_launch_pdl(col_reduce_stage2, dim3((U + 127) / 128), dim3(128), 0, (cudaStream_t)stream, U, nchunks, (const float*)scratch, out_a, out_b);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}

// ---- dv3_col_reduce for callers that must not allocate ------------------------------------------------------------
// Some hosts of this library cannot hand us a workspace: a plugin that is only given its input and output buffers, a
// real-time loop in which cudaMalloc's implicit device-wide synchronisation is not acceptable, or an embedding runtime that
// owns every byte of device memory through its own pool and forbids side allocations.  For them the two-stage reduction
// keeps its partial sums in the fixed array below instead of a caller-provided scratch buffer.
//   * the array lives in every GPU that loads this library until the process exits (2 MiB per device);
//   * it bounds the problem: U (columns) <= DV3_COL_REDUCE_NOALLOC_MAX_U, any M; larger U is refused with DV3_ERR_TOO_LARGE;
//   * nothing in it has to start at zero: stage 1 overwrites every partial that stage 2 reads, so leftovers of the previous
//     call are harmless -- but all calls on a device share the one array, so two calls whose kernels can overlap on the GPU
//     (different streams without an ordering between them) race and must be serialised by the caller;
//   * the first call on each device looks the array's device address up (cudaGetSymbolAddress, with that device current) and
//     caches it; make that first call before starting any stream capture.
// No cudaMalloc / cudaFree / allocating library routine is reached from this entry point.
#define DV3_COL_REDUCE_NOALLOC_MAX_U 4096
#define DV3_COL_REDUCE_MAX_CHUNKS 64
static __device__ __align__(256) float g_col_reduce_partials[DV3_COL_REDUCE_MAX_CHUNKS * 2 * DV3_COL_REDUCE_NOALLOC_MAX_U];
static float* g_col_reduce_partials_addr[64] = {nullptr};   // per device ordinal: DEVICE address of the array above

int dv3_col_reduce_noalloc(int M, int U, const float* d, int ldd, const float* x, int ldx, const float* mean,
                           const float* rstd, float* out_a, float* out_b, void* stream) {
  if (M <= 0 || U <= 0 || !d) return DV3_ERR_ARG;
  if (U > DV3_COL_REDUCE_NOALLOC_MAX_U) return DV3_ERR_TOO_LARGE;
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return (int)e;
  if (dev < 0 || dev >= 64) return DV3_ERR_UNSUPPORTED;
  float* partials = g_col_reduce_partials_addr[dev];
  if (partials == nullptr) {
    // In host code `g_col_reduce_partials` is only a handle for the symbol, not an address a kernel may dereference.
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_col_reduce_partials);
    if (e != cudaSuccess) return (int)e;
    partials = static_cast<float*>(p);
    g_col_reduce_partials_addr[dev] = partials;
  }
  int nchunks = (M + 63) / 64;
  if (nchunks > DV3_COL_REDUCE_MAX_CHUNKS) nchunks = DV3_COL_REDUCE_MAX_CHUNKS;
  const int rows_per_chunk = (M + nchunks - 1) / nchunks;
  dim3 g1((U + CR_COLS - 1) / CR_COLS, nchunks);
  dv3_launch_pdl(col_reduce_stage1, g1, dim3(CR_COLS * CR_ROWLANES), 0, (cudaStream_t)stream, M, U, d, ldd, x, ldx, mean, rstd, rows_per_chunk, partials);
  DV3_LAUNCH_CHECK();
  dv3_launch_pdl(col_reduce_stage2, dim3((U + 127) / 128), dim3(128), 0, (cudaStream_t)stream, U, nchunks, (const float*)partials, out_a, out_b);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}

int dv3_gru_gate_fwd(int R, int D, const float* gates_pre, int ldg, const float* lnw, const float* lnb,
                     const float* h_in, int ldhi, float* h_out, int ldho, float* mean, float* rstd, void* h_out_bf16,
                     void* stream) {
  if (R <= 0 || D <= 0) return DV3_ERR_ARG;
  // (measured: the warp-per-row variant is slower at 1 024 rows -- 8 heavy warps per SM, 26 us vs 11 us -- so it is only taken when
  //  there are enough rows to fill the machine with warps)
  if (R >= 8192 && D <= 512)
    dv3_launch_pdl(gru_gate_fwd_warp_kernel<16>, dim3((R + 7) / 8), dim3(256), 0, (cudaStream_t)stream, R, D, gates_pre, ldg, lnw, lnb, h_in, ldhi,
                   h_out, ldho, mean, rstd, reinterpret_cast<__nv_bfloat16*>(h_out_bf16));
  else
    gru_gate_fwd_kernel<<<R, ROW_THREADS, 0, (cudaStream_t)stream>>>(D, gates_pre, ldg, lnw, lnb, h_in, ldhi, h_out, ldho,
                                                                    mean, rstd, reinterpret_cast<__nv_bfloat16*>(h_out_bf16));
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}

int dv3_gru_gate_bwd(int R, int D, const float* dh_out, int lddho, const float* gates_pre, int ldg, const float* lnw,
                     const float* lnb, const float* mean, const float* rstd, const float* h_in, int ldhi,
                     float* dgates_pre, int lddg, float* dln, int lddln, float* dh_in, int lddhi, void* stream) {
  if (R <= 0 || D <= 0) return DV3_ERR_ARG;
  gru_gate_bwd_kernel<<<R, ROW_THREADS, 0, (cudaStream_t)stream>>>(D, dh_out, lddho, gates_pre, ldg, lnw, lnb, mean, rstd,
                                                                  h_in, ldhi, dgates_pre, lddg, dln, lddln, dh_in, lddhi);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}

int dv3_symlog(size_t n, const float* x, float* y, void* stream) {
  if (n == 0) return DV3_OK;
  symlog_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(n, x, y);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_copy2d(int R, int C, const float* s, int lds, float* d, int ldd, void* stream) {
  if (R <= 0 || C <= 0) return DV3_ERR_ARG;
  dv3_launch_pdl(copy2d_kernel, dim3(grid_for((size_t)R * C, 256)), dim3(256), 0, (cudaStream_t)stream, R, C, s, lds, d, ldd);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_transpose(int R, int C, const float* s, int lds, float* d, int ldd, void* stream) {
  if (R <= 0 || C <= 0) return DV3_ERR_ARG;
  dim3 grid((C + 31) / 32, (R + 31) / 32), block(32, 8);
  transpose_kernel<<<grid, block, 0, (cudaStream_t)stream>>>(R, C, s, lds, d, ldd);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_scan_prep(int B, int D, int S, const float* first, const float* h_prev, int ldhp, const float* z_prev, int ldzp,
                  const float* zreset, float* h_in, int ldhi, float* z_in, int ldzi, void* stream) {
  if (B <= 0) return DV3_ERR_ARG;
  scan_prep_kernel<<<B, 256, 0, (cudaStream_t)stream>>>(B, D, S, first, h_prev, ldhp, z_prev, ldzp, zreset, h_in, ldhi, z_in, ldzi);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_scan_carry(int B, int D, int S, const float* first, const float* dz_prev, int lddz, const float* dh_a, int lda,
                   const float* dh_b, int ldb, float* dz_dst, int ldzd, float* dh_dst, int ldhd, void* stream) {
  if (B <= 0) return DV3_ERR_ARG;
  scan_carry_kernel<<<B, 256, 0, (cudaStream_t)stream>>>(B, D, S, first, dz_prev, lddz, dh_a, lda, dh_b, ldb, dz_dst, ldzd, dh_dst, ldhd);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_mask_shift_actions(int T, int B, int A, const float* actions, const float* first, float* out, int ldo, void* stream) {
  if (T <= 0 || B <= 0 || A <= 0) return DV3_ERR_ARG;
  mask_shift_actions_kernel<<<T * B, 128, 0, (cudaStream_t)stream>>>(T, B, A, actions, first, out, ldo);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_ema(size_t n, float* target, const float* src, float tau, void* stream) {
  if (n == 0) return DV3_OK;
  ema_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(n, target, src, tau);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
int dv3_fill(size_t n, float* p, float v, void* stream) {
  if (n == 0) return DV3_OK;
  fill_kernel<<<grid_for(n, 256), 256, 0, (cudaStream_t)stream>>>(n, p, v);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}

}  // extern "C"
