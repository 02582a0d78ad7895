// fp32 CUDA-core contractions for the strict-parity ("fp32") mode of libdv3b200.
//   * dv3_gemm_f32 tiled kernel: 128x128x16 tiles, 8x8 register micro-tile, register-prefetched global loads.
//     Handles NT / NN / TN operand layouts with arbitrary leading dimensions and ragged edges.
//   * skinny NT kernel (M <= 32): the per-step contractions of the RSSM observe scan, where the batch is the
//     M dimension (B = 16) and the problem is streaming the weight matrix once.
// Replaces, for the reference, every torch `addmm/mm` on the DreamerV3 train path
// (sheeprl/models/models.py:115-118 MLP.forward, :389 LayerNormGRUCell linear, and their autograd backward).
#include "common.cuh"

namespace {

constexpr int BM = 128, BN = 128, BK = 16, TM = 8, TN_ = 8, NT_THREADS = 256;

enum { MODE_NT = 0, MODE_NN = 1, MODE_TN = 2 };

template <int MODE>
__global__ void __launch_bounds__(NT_THREADS)
gemm_tiled_kernel(int M, int N, int K, const float* __restrict__ A, int lda, const float* __restrict__ B, int ldb,
                  float* __restrict__ C, int ldc, const float* __restrict__ bias, int accumulate) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.y * BM, n0 = blockIdx.x * BN;
  const int tx = tid & 15, ty = tid >> 4;  // 16 x 16 threads

  // element mappings for the global->shared staging (8 A + 8 B elements per thread per k-tile)
  constexpr bool A_KCONTIG = (MODE != MODE_TN);  // A(m,k) contiguous in k for NT/NN
  constexpr bool B_KCONTIG = (MODE == MODE_NT);  // B(k,n) contiguous in k for NT
  float ra[8], rb[8];

  auto load_tiles = [&](int k0) {
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int lin = tid + e * NT_THREADS;
      int m, k;
      if (A_KCONTIG) { m = lin / BK; k = lin % BK; } else { k = lin / BM; m = lin % BM; }
      const int gm = m0 + m, gk = k0 + k;
      float v = 0.f;
      if (gm < M && gk < K) v = A_KCONTIG ? __ldg(A + (size_t)gm * lda + gk) : __ldg(A + (size_t)gk * lda + gm);
      ra[e] = v;
      int n, kb;
      if (B_KCONTIG) { n = lin / BK; kb = lin % BK; } else { kb = lin / BN; n = lin % BN; }
      const int gn = n0 + n, gkb = k0 + kb;
      float w = 0.f;
      if (gn < N && gkb < K) w = B_KCONTIG ? __ldg(B + (size_t)gn * ldb + gkb) : __ldg(B + (size_t)gkb * ldb + gn);
      rb[e] = w;
    }
  };
  auto store_tiles = [&]() {
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      const int lin = tid + e * NT_THREADS;
      int m, k;
      if (A_KCONTIG) { m = lin / BK; k = lin % BK; } else { k = lin / BM; m = lin % BM; }
      As[k][m] = ra[e];
      int n, kb;
      if (B_KCONTIG) { n = lin / BK; kb = lin % BK; } else { kb = lin / BN; n = lin % BN; }
      Bs[kb][n] = rb[e];
    }
  };

  float acc[TM][TN_];
#pragma unroll
  for (int i = 0; i < TM; ++i)
#pragma unroll
    for (int j = 0; j < TN_; ++j) acc[i][j] = 0.f;

  load_tiles(0);
  for (int k0 = 0; k0 < K; k0 += BK) {
    store_tiles();
    __syncthreads();
    if (k0 + BK < K) load_tiles(k0 + BK);
#pragma unroll
    for (int k = 0; k < BK; ++k) {
      float a[TM], b[TN_];
      const float4 a0 = *reinterpret_cast<const float4*>(&As[k][ty * 4]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[k][ty * 4 + 64]);
      const float4 b0 = *reinterpret_cast<const float4*>(&Bs[k][tx * 4]);
      const float4 b1 = *reinterpret_cast<const float4*>(&Bs[k][tx * 4 + 64]);
      a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
      b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
#pragma unroll
      for (int i = 0; i < TM; ++i)
#pragma unroll
        for (int j = 0; j < TN_; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }

#pragma unroll
  for (int i = 0; i < TM; ++i) {
    const int gm = m0 + ty * 4 + (i & 3) + (i >> 2) * 64;
    if (gm >= M) continue;
#pragma unroll
    for (int j = 0; j < TN_; ++j) {
      const int gn = n0 + tx * 4 + (j & 3) + (j >> 2) * 64;
      if (gn >= N) continue;
      float v = acc[i][j];
      if (bias) v += __ldg(bias + gn);
      float* c = C + (size_t)gm * ldc + gn;
      if (accumulate) v += *c;
      *c = v;
    }
  }
}

// ---- skinny NT: C[M<=32, N] = A[M,K] * B[N,K]^T ------------------------------------------------------
constexpr int SK_WARPS = 8, SK_NPW = 2, SK_KT = 256;

template <int MT>
__global__ void __launch_bounds__(SK_WARPS * 32)
gemm_skinny_nt_kernel(int M, int N, int K, const float* __restrict__ A, int lda, const float* __restrict__ B, int ldb,
                      float* __restrict__ C, int ldc, const float* __restrict__ bias, int accumulate) {
  __shared__ float As[MT][SK_KT];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nbase = (blockIdx.x * SK_WARPS + warp) * SK_NPW;
  float acc[SK_NPW][MT];
#pragma unroll
  for (int c = 0; c < SK_NPW; ++c)
#pragma unroll
    for (int m = 0; m < MT; ++m) acc[c][m] = 0.f;

  for (int k0 = 0; k0 < K; k0 += SK_KT) {
    __syncthreads();
    for (int i = threadIdx.x; i < MT * SK_KT; i += blockDim.x) {
      const int m = i / SK_KT, kk = i % SK_KT;
      float v = 0.f;
      if (m < M && k0 + kk < K) v = A[(size_t)m * lda + k0 + kk];
      As[m][kk] = v;
    }
    __syncthreads();
    const int kmax = min(SK_KT, K - k0);
    for (int kk = lane; kk < kmax; kk += 32) {
      float bv[SK_NPW];
#pragma unroll
      for (int c = 0; c < SK_NPW; ++c) {
        const int n = nbase + c;
        bv[c] = (n < N) ? __ldg(B + (size_t)n * ldb + k0 + kk) : 0.f;
      }
#pragma unroll
      for (int m = 0; m < MT; ++m) {
        const float a = As[m][kk];
#pragma unroll
        for (int c = 0; c < SK_NPW; ++c) acc[c][m] = fmaf(a, bv[c], acc[c][m]);
      }
    }
  }
#pragma unroll
  for (int c = 0; c < SK_NPW; ++c) {
    const int n = nbase + c;
#pragma unroll
    for (int m = 0; m < MT; ++m) {
      float v = warp_sum(acc[c][m]);
      if (lane == 0 && n < N && m < M) {
        if (bias) v += bias[n];
        float* cp = C + (size_t)m * ldc + n;
        if (accumulate) v += *cp;
        *cp = v;
      }
    }
  }
}


// ---- skinny NT, streaming version: no shared memory, no barriers --------------------------------------
// One warp owns SKV_NPW output columns; each lane streams V-wide vectors of the weight rows (coalesced, several loads in
// flight) and reads the <= 16 activation rows through L1 (they are shared by every warp of the SM).  The per-step
// contractions of the RSSM scan have M = B = 16: the problem is to read the weight matrix once at L2/HBM speed.
constexpr int SKV_WARPS = 8, SKV_NPW = 2;

template <int V> struct VecT;
template <> struct VecT<4> { using T = float4; };
template <> struct VecT<2> { using T = float2; };
template <> struct VecT<1> { using T = float; };
template <int V> __device__ __forceinline__ void vec_load(const float* p, float (&o)[V]) {
  typename VecT<V>::T v = __ldg(reinterpret_cast<const typename VecT<V>::T*>(p));
  const float* f = reinterpret_cast<const float*>(&v);
#pragma unroll
  for (int i = 0; i < V; ++i) o[i] = f[i];
}

template <int MT, int V>
__global__ void __launch_bounds__(SKV_WARPS * 32)
gemm_skinny_stream_kernel(int M, int N, int K, const float* __restrict__ A, int lda, const float* __restrict__ B, int ldb,
                          float* __restrict__ C, int ldc, const float* __restrict__ bias, int accumulate) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nbase = (blockIdx.x * SKV_WARPS + warp) * SKV_NPW;
  if (nbase >= N) return;
  float acc[SKV_NPW][MT];
#pragma unroll
  for (int c = 0; c < SKV_NPW; ++c)
#pragma unroll
    for (int m = 0; m < MT; ++m) acc[c][m] = 0.f;
  const int KV = (K / V) * V;
  const float* brow[SKV_NPW];
#pragma unroll
  for (int c = 0; c < SKV_NPW; ++c) brow[c] = B + (size_t)min(nbase + c, N - 1) * ldb;
#pragma unroll 2
  for (int k = lane * V; k < KV; k += 32 * V) {
    float w[SKV_NPW][V];
#pragma unroll
    for (int c = 0; c < SKV_NPW; ++c) vec_load<V>(brow[c] + k, w[c]);
#pragma unroll
    for (int m = 0; m < MT; ++m) {
      float a[V];
      if (m < M) vec_load<V>(A + (size_t)m * lda + k, a);
      else {
#pragma unroll
        for (int i = 0; i < V; ++i) a[i] = 0.f;
      }
#pragma unroll
      for (int c = 0; c < SKV_NPW; ++c)
#pragma unroll
        for (int i = 0; i < V; ++i) acc[c][m] = fmaf(a[i], w[c][i], acc[c][m]);
    }
  }
  for (int k = KV + lane; k < K; k += 32) {   // scalar tail
#pragma unroll
    for (int m = 0; m < MT; ++m) {
      const float a = (m < M) ? __ldg(A + (size_t)m * lda + k) : 0.f;
#pragma unroll
      for (int c = 0; c < SKV_NPW; ++c) acc[c][m] = fmaf(a, __ldg(brow[c] + k), acc[c][m]);
    }
  }
#pragma unroll
  for (int c = 0; c < SKV_NPW; ++c) {
    const int n = nbase + c;
#pragma unroll
    for (int m = 0; m < MT; ++m) {
      float v = warp_sum(acc[c][m]);
      if (lane == 0 && n < N && m < M) {
        if (bias) v += bias[n];
        float* cp = C + (size_t)m * ldc + n;
        if (accumulate) v += *cp;
        *cp = v;
      }
    }
  }
}


// ---- skinny NT, split-K inside the block: maximum memory-level parallelism -------------------------------------------
// Block = 8 warps = 8 output columns; warp w streams the k-slice [w*K/8, (w+1)*K/8) of all 8 weight rows, so every lane has
// 8 independent 16-byte loads in flight per 128-wide chunk and the whole 8 x K weight panel is fetched in about one memory
// round trip.  Per-lane partial sums (8 columns x 16 rows) are reduced with four 31-shuffle transposed reductions, then across
// the 8 warps through shared memory (fixed order -> deterministic).
constexpr int SK3_COLS = 8, SK3_WARPS = 8;

__device__ __forceinline__ float sk3_transpose_reduce32(float (&v)[32], int lane) {
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

__global__ void __launch_bounds__(SK3_WARPS * 32)
gemm_skinny_splitk_kernel(int M, int N, int K, const float* __restrict__ A, int lda, const float* __restrict__ B, int ldb,
                          float* __restrict__ C, int ldc, const float* __restrict__ bias, int accumulate) {
  __shared__ float part[SK3_WARPS][SK3_COLS * 16];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n0 = blockIdx.x * SK3_COLS;
  // k-slice of this warp, in units of 128-element chunks
  const int chunks = (K + 127) / 128;
  const int cpw = (chunks + SK3_WARPS - 1) / SK3_WARPS;
  const int kbeg = warp * cpw * 128, kend = min(K, kbeg + cpw * 128);
  float acc[SK3_COLS][16];
#pragma unroll
  for (int c = 0; c < SK3_COLS; ++c)
#pragma unroll
    for (int m = 0; m < 16; ++m) acc[c][m] = 0.f;
  for (int k = kbeg + lane * 4; k < kend; k += 128) {      // K % 4 == 0 and 16 B aligned rows (checked by the launcher)
    float4 w[SK3_COLS];
#pragma unroll
    for (int c = 0; c < SK3_COLS; ++c) {
      const int n = min(n0 + c, N - 1);
      w[c] = __ldg(reinterpret_cast<const float4*>(B + (size_t)n * ldb + k));
    }
#pragma unroll
    for (int m = 0; m < 16; ++m) {
      float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
      if (m < M) a = __ldg(reinterpret_cast<const float4*>(A + (size_t)m * lda + k));
#pragma unroll
      for (int c = 0; c < SK3_COLS; ++c) {
        acc[c][m] = fmaf(a.x, w[c].x, acc[c][m]);
        acc[c][m] = fmaf(a.y, w[c].y, acc[c][m]);
        acc[c][m] = fmaf(a.z, w[c].z, acc[c][m]);
        acc[c][m] = fmaf(a.w, w[c].w, acc[c][m]);
      }
    }
  }
  // lane-reduce: 4 groups of (2 columns x 16 rows) = 32 values each
#pragma unroll
  for (int g = 0; g < SK3_COLS / 2; ++g) {
    float v[32];
#pragma unroll
    for (int m = 0; m < 16; ++m) { v[m] = acc[2 * g][m]; v[16 + m] = acc[2 * g + 1][m]; }
    const float r = sk3_transpose_reduce32(v, lane);          // lane l: column 2g + (l>>4), row l&15
    part[warp][(2 * g + (lane >> 4)) * 16 + (lane & 15)] = r;
  }
  __syncthreads();
  if (threadIdx.x < SK3_COLS * 16) {
    const int c = threadIdx.x >> 4, m = threadIdx.x & 15, n = n0 + c;
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < SK3_WARPS; ++w) s += part[w][threadIdx.x];
    if (n < N && m < M) {
      if (bias) s += bias[n];
      float* cp = C + (size_t)m * ldc + n;
      if (accumulate) s += *cp;
      *cp = s;
    }
  }
}


// ---- narrow outputs ---------------------------------------------------------------------------------------------------
// NT with N <= 32 (scalar / few-column heads: continue logit, low-dimensional observation decoder, small action spaces): one warp per row of A,
// memory-bound on A (read once, coalesced 16 B per lane); the N weight rows stay in L1.
constexpr int NARROW = 16;
constexpr int NT_NARROW = 32;   // NT: up to 32 output columns per row (one accumulator each)
__global__ void __launch_bounds__(256)
gemm_nt_narrow_kernel(int M, int N, int K, const float* __restrict__ A, int lda, const float* __restrict__ B, int ldb,
                      float* __restrict__ C, int ldc, const float* __restrict__ bias, int accumulate, int vec) {
  const int lane = threadIdx.x & 31;
  const size_t row = (size_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (row >= (size_t)M) return;
  const float* a = A + row * lda;
  float acc[NT_NARROW];
#pragma unroll
  for (int n = 0; n < NT_NARROW; ++n) acc[n] = 0.f;
  if (vec) {
    for (int k = lane * 4; k < K; k += 128) {
      const float4 av = __ldg(reinterpret_cast<const float4*>(a + k));
#pragma unroll
      for (int n = 0; n < NT_NARROW; ++n) {
        if (n < N) {
          const float4 w = __ldg(reinterpret_cast<const float4*>(B + (size_t)n * ldb + k));
          acc[n] = fmaf(av.x, w.x, fmaf(av.y, w.y, fmaf(av.z, w.z, fmaf(av.w, w.w, acc[n]))));
        }
      }
    }
  } else {
    for (int k = lane; k < K; k += 32) {
      const float av = __ldg(a + k);
#pragma unroll
      for (int n = 0; n < NT_NARROW; ++n)
        if (n < N) acc[n] = fmaf(av, __ldg(B + (size_t)n * ldb + k), acc[n]);
    }
  }
#pragma unroll
  for (int n = 0; n < NT_NARROW; ++n) {
    if (n < N) {
      float v = warp_sum(acc[n]);
      if (lane == 0) {
        if (bias) v += bias[n];
        float* cp = C + row * ldc + n;
        *cp = accumulate ? (*cp + v) : v;
      }
    }
  }
}

// TN with one narrow side (<= 16): C = sum_k P[k, :]^T Q[k, :] with P [K, p <= 16] and Q [K, q] (bias / few-column weight
// gradients over K = T*B(*H) rows).  Threads own q columns (coalesced reads of Q), the P rows of a 64-deep chunk are staged in
// shared memory and broadcast; K is split over blockIdx.y and reduced with fp32 atomics into a zeroed (or accumulated) C.
__global__ void __launch_bounds__(128)
gemm_tn_narrow_kernel(int P, int Q, int K, const float* __restrict__ Pm, int ldp, const float* __restrict__ Qm, int ldq,
                      float* __restrict__ C, int ldc, int transposed_out, int k_per_block) {
  __shared__ float sp[64][NARROW];
  const int q = blockIdx.x * 128 + threadIdx.x;
  const int k0 = blockIdx.y * k_per_block, k1 = min(K, k0 + k_per_block);
  float acc[NARROW];
#pragma unroll
  for (int p = 0; p < NARROW; ++p) acc[p] = 0.f;
  for (int kc = k0; kc < k1; kc += 64) {
    __syncthreads();
    for (int i = threadIdx.x; i < 64 * NARROW; i += 128) {
      const int kk = i >> 4, p = i & 15;
      sp[kk][p] = (kc + kk < k1 && p < P) ? __ldg(Pm + (size_t)(kc + kk) * ldp + p) : 0.f;
    }
    __syncthreads();
    if (q < Q) {
      const int kn = min(64, k1 - kc);
#pragma unroll 8
      for (int kk = 0; kk < kn; ++kk) {
        const float b = __ldg(Qm + (size_t)(kc + kk) * ldq + q);
#pragma unroll
        for (int p = 0; p < NARROW; ++p) acc[p] = fmaf(sp[kk][p], b, acc[p]);
      }
    }
  }
  if (q < Q) {
    for (int p = 0; p < P; ++p) atomicAdd(transposed_out ? (C + (size_t)q * ldc + p) : (C + (size_t)p * ldc + q), acc[p]);
  }
}

// C = (accumulate ? C : 0) + bias: the atomics of gemm_tn_narrow_kernel then add the product
__global__ void init2d_simt_kernel(int M, int N, float* __restrict__ C, int ldc, const float* __restrict__ bias, int accumulate) {
  const size_t n = (size_t)M * N;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    float* cp = C + (i / N) * ldc + (i % N);
    *cp = (accumulate ? *cp : 0.f) + (bias ? bias[i % N] : 0.f);
  }
}

template <int MT>
static int launch_skinny_stream(int M, int N, int K, const float* A, int lda, const float* B, int ldb, float* C, int ldc,
                                const float* bias, int accumulate, cudaStream_t stream) {
  dim3 grid((N + SKV_WARPS * SKV_NPW - 1) / (SKV_WARPS * SKV_NPW));
  const uintptr_t al = (uintptr_t)A | (uintptr_t)B | ((uintptr_t)lda * 4) | ((uintptr_t)ldb * 4);
  if ((al & 15) == 0)
    gemm_skinny_stream_kernel<MT, 4><<<grid, SKV_WARPS * 32, 0, stream>>>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate);
  else if ((al & 7) == 0)
    gemm_skinny_stream_kernel<MT, 2><<<grid, SKV_WARPS * 32, 0, stream>>>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate);
  else
    gemm_skinny_stream_kernel<MT, 1><<<grid, SKV_WARPS * 32, 0, stream>>>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate);
  cudaError_t e = cudaGetLastError();
  return e == cudaSuccess ? DV3_OK : (int)e;
}

}  // namespace

extern "C" int dv3_gemm_f32(int mode, int M, int N, int K, const float* A, int lda, const float* B, int ldb, float* C,
                            int ldc, const float* bias, int accumulate, void* stream_) {
  if (M <= 0 || N <= 0 || K <= 0 || !A || !B || !C) return DV3_ERR_ARG;
  cudaStream_t stream = (cudaStream_t)stream_;
  if (mode == MODE_NT && M <= 16 && K >= 256 && (K & 3) == 0 && (lda & 3) == 0 && (ldb & 3) == 0 &&
      (((uintptr_t)A | (uintptr_t)B) & 15) == 0) {
    gemm_skinny_splitk_kernel<<<(N + SK3_COLS - 1) / SK3_COLS, SK3_WARPS * 32, 0, stream>>>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate);
    DV3_LAUNCH_CHECK();
    return DV3_OK;
  }
  if (mode == MODE_NT && N <= NT_NARROW && M > 32) {
    const int vec = (K & 3) == 0 && (lda & 3) == 0 && (ldb & 3) == 0 && (((uintptr_t)A | (uintptr_t)B) & 15) == 0;
    gemm_nt_narrow_kernel<<<(M + 7) / 8, 256, 0, stream>>>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate, vec);
    DV3_LAUNCH_CHECK();
    return DV3_OK;
  }
  if (mode == MODE_TN && (M <= NARROW || N <= NARROW) && K >= 64) {
    if (!accumulate || bias) {
      size_t n = (size_t)M * N;
      init2d_simt_kernel<<<(int)((n + 255) / 256 > 1184 ? 1184 : (n + 255) / 256), 256, 0, stream>>>(M, N, C, ldc, bias, accumulate);
      DV3_LAUNCH_CHECK();
    }
    const bool a_narrow = M <= N;                       // the narrow operand is staged, the wide one is streamed
    const int P = a_narrow ? M : N, Q = a_narrow ? N : M;
    const int qb = (Q + 127) / 128;
    int ks = (296 + qb - 1) / qb;                       // ~2 blocks per SM
    int kpb = (K + ks - 1) / ks;
    kpb = (kpb + 63) / 64 * 64;
    ks = (K + kpb - 1) / kpb;
    gemm_tn_narrow_kernel<<<dim3(qb, ks), 128, 0, stream>>>(P, Q, K, a_narrow ? A : B, a_narrow ? lda : ldb, a_narrow ? B : A,
                                                            a_narrow ? ldb : lda, C, ldc, a_narrow ? 0 : 1, kpb);
    DV3_LAUNCH_CHECK();
    return DV3_OK;
  }
  if (mode == MODE_NT && M <= 16) {
    if (M <= 8) return launch_skinny_stream<8>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate, stream);
    return launch_skinny_stream<16>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate, stream);
  }
  if (mode == MODE_NT && M <= 32) {
    dim3 grid((N + SK_WARPS * SK_NPW - 1) / (SK_WARPS * SK_NPW));
    if (M <= 8)
      gemm_skinny_nt_kernel<8><<<grid, SK_WARPS * 32, 0, stream>>>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate);
    else if (M <= 16)
      gemm_skinny_nt_kernel<16><<<grid, SK_WARPS * 32, 0, stream>>>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate);
    else
      gemm_skinny_nt_kernel<32><<<grid, SK_WARPS * 32, 0, stream>>>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate);
    DV3_LAUNCH_CHECK();
    return DV3_OK;
  }
  dim3 grid((N + BN - 1) / BN, (M + BM - 1) / BM);
  if (grid.y > 65535) return DV3_ERR_UNSUPPORTED;
  switch (mode) {
    case MODE_NT: gemm_tiled_kernel<MODE_NT><<<grid, NT_THREADS, 0, stream>>>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate); break;
    case MODE_NN: gemm_tiled_kernel<MODE_NN><<<grid, NT_THREADS, 0, stream>>>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate); break;
    case MODE_TN: gemm_tiled_kernel<MODE_TN><<<grid, NT_THREADS, 0, stream>>>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate); break;
    default: return DV3_ERR_ARG;
  }
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}
