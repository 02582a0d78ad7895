// Shared pieces of the two cluster scan kernels (rssm_scan_cluster.cu forward, rssm_scan_bwd_cluster.cu backward):
// one thread-block cluster of CL CTAs, bf16 weight rows streamed from L2 into mma.sync fragments, activations and LayerNorm
// partials exchanged through distributed shared memory, hardware cluster barriers.
//
// The kernels run B <= 16 rows per time step on 16 SMs, so they are latency bound -- and, measured with ncu on the first version
// (14.5 k SASS instructions, 233 KB): instruction-FETCH bound (sm__icc_request_hit_rate 81 %).  The time-loop body must fit the
// 32 KB L1.5 instruction cache, hence: helpers are __noinline__ and shared by all phases, loops are not unrolled except the
// MMA inner body, CL is a run-time value (one kernel instance).
#pragma once
#include "common.cuh"
#include <cooperative_groups.h>
#include <cuda_bf16.h>
namespace cg = cooperative_groups;

namespace sc {

constexpr int MB = 16;          // batch rows (B <= 16) = the two n8 tiles of the MMA
#ifndef DV3_SCAN_CT
#define DV3_SCAN_CT 512
#endif
constexpr int CT = DV3_SCAN_CT; // threads per CTA
constexpr int CW = CT / 32;     // warps per CTA
constexpr int OB = MB + 1;      // row stride of the own-output buffers (odd: conflict-free by row and by column)
constexpr int MAXT = 12;        // at most 12 row tiles (192 rows) per contraction and CTA

__host__ __device__ inline int ksplit_of(int ntiles) {    // CW / gcd(ntiles, CW): (tile, k-part) units divide evenly over the warps
  int g = 1;
  while (g < CW && (ntiles & g) == 0) g <<= 1;
  return CW / g;
}
__host__ __device__ inline int part_kb(int ntiles) { return ntiles * ksplit_of(ntiles); }   // partial-sum slots (1 KB each)

__device__ __forceinline__ void mma_bf16(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// Cluster barrier in two halves: between the arrive (release: own DSMEM pushes are ordered before it) and the wait (acquire) a CTA
// does work that needs nothing from its peers -- the global stores of activations it already holds, cp.async prefetches -- so that
// this work hides in the barrier's latency (~1.4 k clk here) instead of following it.
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }

// Contraction of `ntiles` tiles of 16 weight rows with X[MB][K] (bf16, shared memory, row stride ldx):
//   outb[(ti * 16 + r) * OB + m] = sum_k W[tile_row(ti) + r, k] * X[m, k],  tile_row(ti) = row_base + (ti / tpb) * blk_stride + (ti % tpb) * 16
// W is PACKED in MMA-fragment order by dv3_pack_mma_a (see rssm_scan_cluster.cu): for every (16-row tile, 32-wide k chunk) the 32
// lanes' A fragments of the two m16n8k16 MMAs of the chunk are 32 contiguous bytes each -> two coalesced LDG.128 per lane and chunk,
// no register shuffling, immediate-offset addressing in the unrolled loop.  The k order inside a chunk is permuted (thread tig owns
// k = 8 tig .. 8 tig + 7); the B fragments read the same 8 consecutive k of the two batch rows from shared memory (LDS.128), so
// the permutation is consistent and the sum is unchanged.
// Work unit = (tile, k-part): ksplit = CW / gcd(ntiles, CW) so that the units divide evenly over the CW warps.  Partial sums go
// through part[tile][kq][16][MB] and are added in a fixed order (deterministic).
// WSMEM: the packed tiles were copied into shared memory once (weights resident for the whole scan): plain loads instead of LDG.
template <bool WSMEM>
static __device__ __noinline__ void contract_t(const uint4* __restrict__ Wp, int ntiles, int row_base, int tpb, int blk_stride, int K,
                                               const __nv_bfloat16* X, int ldx, float* part, float* outb) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, gid = lane >> 2, tig = lane & 3;
  const int ksplit = ksplit_of(ntiles), ks_shift = 31 - __clz(ksplit);
  const int upw = (ntiles * ksplit) / CW, nch = K >> 5, lg_tpb = 31 - __clz(tpb);
  const __nv_bfloat16* xa_p = X + gid * ldx + tig * 8;
  const __nv_bfloat16* xb_p = X + (gid + 8) * ldx + tig * 8;
#pragma unroll 1
  for (int j = 0; j < upw; ++j) {
    const int u = warp * upw + j, tile = u >> ks_shift, kq = u & (ksplit - 1);
    const int ch0 = (kq * nch) >> ks_shift, ch1 = ((kq + 1) * nch) >> ks_shift;
    const int blk = tile >> lg_tpb, trow = row_base + blk * blk_stride + (tile & (tpb - 1)) * 16;   // tpb is a power of two
    const uint4* wp = Wp + ((size_t)((trow >> 4) * nch + ch0) * 32 + lane) * 2;
    const __nv_bfloat16* xa = xa_p + (ch0 << 5);
    const __nv_bfloat16* xb = xb_p + (ch0 << 5);
    float acc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll 4
    for (int ch = ch0; ch < ch1; ++ch) {
      const uint4 w1 = WSMEM ? wp[0] : __ldg(wp), w2 = WSMEM ? wp[1] : __ldg(wp + 1);
      const uint4 va = *reinterpret_cast<const uint4*>(xa);
      const uint4 vb = *reinterpret_cast<const uint4*>(xb);
      mma_bf16(acc[0], w1.x, w1.y, w1.z, w1.w, va.x, va.y);
      mma_bf16(acc[0], w2.x, w2.y, w2.z, w2.w, va.z, va.w);
      mma_bf16(acc[1], w1.x, w1.y, w1.z, w1.w, vb.x, vb.y);
      mma_bf16(acc[1], w2.x, w2.y, w2.z, w2.w, vb.z, vb.w);
      wp += 64;
      xa += 32;
      xb += 32;
    }
    float* p = part + (size_t)u * 256;
    *reinterpret_cast<float2*>(p + gid * MB + 2 * tig) = make_float2(acc[0][0], acc[0][1]);
    *reinterpret_cast<float2*>(p + (gid + 8) * MB + 2 * tig) = make_float2(acc[0][2], acc[0][3]);
    *reinterpret_cast<float2*>(p + gid * MB + 8 + 2 * tig) = make_float2(acc[1][0], acc[1][1]);
    *reinterpret_cast<float2*>(p + (gid + 8) * MB + 8 + 2 * tig) = make_float2(acc[1][2], acc[1][3]);
  }
  __syncthreads();
#pragma unroll 1
  for (int i = threadIdx.x; i < ntiles * 256; i += CT) {
    const int tile = i >> 8, e = i & 255;
    const float* p = part + (size_t)(tile << ks_shift) * 256 + e;
    float s = p[0];
#pragma unroll 1
    for (int q = 1; q < ksplit; ++q) s += p[q * 256];
    outb[(size_t)(tile * 16 + (e >> 4)) * OB + (e & 15)] = s;
  }
  __syncthreads();
}

static __device__ __forceinline__ void contract(const uint4* Wp, int ntiles, int row_base, int tpb, int blk_stride, int K,
                                                const __nv_bfloat16* X, int ldx, float* part, float* outb) {
  contract_t<false>(Wp, ntiles, row_base, tpb, blk_stride, K, X, ldx, part, outb);
}

// LayerNorm statistics of a vector spread over the cluster.  Own values of batch row m: vals[j * OB + m], j < n.  Every CTA pushes
// its (mean, M2) into lnp[c][m] of every peer; after the next cluster barrier ln_combine() merges them (Chan, equal counts).
static __device__ __noinline__ void ln_push(const float* vals, int n, float* lnp, int c, int CL) {
  cg::cluster_group cluster = cg::this_cluster();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll 1
  for (int m = warp; m < MB; m += CW) {
    float s = 0.f;
    for (int j = lane; j < n; j += 32) s += vals[j * OB + m];
    const float mu = warp_sum(s) / (float)n;
    float q = 0.f;
    for (int j = lane; j < n; j += 32) { const float dlt = vals[j * OB + m] - mu; q = fmaf(dlt, dlt, q); }
    q = warp_sum(q);
    if (lane < CL) reinterpret_cast<float2*>(cluster.map_shared_rank(lnp, lane))[c * MB + m] = make_float2(mu, q);
  }
}
static __device__ __noinline__ void ln_combine(const float* lnp, int CL, int n_per_cta, float eps, float* s_stat, float* gmean, float* grstd,
                                        bool write_global, int B) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#pragma unroll 1
  for (int m = warp; m < MB; m += CW) {
    const float2 pc = (lane < CL) ? reinterpret_cast<const float2*>(lnp)[lane * MB + m] : make_float2(0.f, 0.f);
    const float mu = warp_sum(pc.x) / (float)CL;
    const float dm = pc.x - mu;
    const float m2 = warp_sum((lane < CL) ? (pc.y + (float)n_per_cta * dm * dm) : 0.f);
    const float rs = rsqrtf(m2 / (float)(n_per_cta * CL) + eps);
    if (lane == 0) {
      s_stat[m] = mu;
      s_stat[MB + m] = rs;
      if (write_global && m < B) { gmean[m] = mu; grstd[m] = rs; }
    }
  }
  __syncthreads();
}
// push the own bf16 slice stage[MB][ncols] (row stride sld) into column col0 of `buf` (row stride ld) of every CTA of the cluster
static __device__ __noinline__ void allgather(__nv_bfloat16* buf, int ld, int col0, int ncols, const __nv_bfloat16* stage, int sld, int CL) {
  cg::cluster_group cluster = cg::this_cluster();
  const int cpr = ncols >> 3, per_peer = MB * cpr, lg_pp = 31 - __clz(per_peer), lg_cpr = 31 - __clz(cpr);   // powers of two
#pragma unroll 1
  for (int i = threadIdx.x; i < CL * per_peer; i += CT) {
    const int peer = i >> lg_pp, rem = i & (per_peer - 1), m = rem >> lg_cpr, ch = rem & (cpr - 1);
    __nv_bfloat16* dst = cluster.map_shared_rank(buf, peer) + m * ld + col0 + ch * 8;
    *reinterpret_cast<uint4*>(dst) = *reinterpret_cast<const uint4*>(stage + m * sld + ch * 8);
  }
}
// global dst[m * ld + col] = src[m * sld + col] for m < B, col < cols (shared -> global; the deferred stores of saved activations:
// a cluster barrier's release fence waits for every earlier store of the thread, so they are issued right AFTER a barrier)
static __device__ __noinline__ void store_rows(float* dst, int ld, const float* src, int sld, int cols, int B) {
  const int lg = 31 - __clz(cols);                        // cols is a power of two (host check)
#pragma unroll 1
  for (int e = threadIdx.x; e < B * cols; e += CT) {
    const int m = e >> lg, col = e & (cols - 1);
    dst[(size_t)m * ld + col] = src[m * sld + col];
  }
}
// same, source laid out [col][OB] (the contraction output buffer)
static __device__ __noinline__ void store_cols(float* dst, int ld, const float* src, int cols, int B) {
  const int lg = 31 - __clz(cols);
#pragma unroll 1
  for (int e = threadIdx.x; e < B * cols; e += CT) {
    const int m = e >> lg, col = e & (cols - 1);
    dst[(size_t)m * ld + col] = src[col * OB + m];
  }
}
// y = SiLU(LN(pre)) of the own slice: pre[col * OB + m], statistics in s_stat, parameters prm = [weight(cols) | bias(cols)];
// writes the fp32 value to dfr[m * sld + col] (stored to global one barrier later) and bf16 to stage[m * sld + col] (all-gathered).
static __device__ __noinline__ void ln_silu_slice(const float* pre, int cols, const float* s_stat, const float* prm, float* dfr, __nv_bfloat16* stage,
                                           int sld) {
  const int lg = 31 - __clz(cols);
#pragma unroll 1
  for (int e = threadIdx.x; e < MB * cols; e += CT) {
    const int m = e >> lg, col = e & (cols - 1);
    const float x = (pre[col * OB + m] - s_stat[m]) * s_stat[MB + m] * prm[col] + prm[cols + col];
    const float f = siluf_(x);
    dfr[m * sld + col] = f;
    stage[m * sld + col] = __float2bfloat16_rn(f);
  }
}

}  // namespace sc
