// Stand-alone B200 probe for the mechanisms the tcgen05 scan kernels are built on (not part of the library):
//   T1  tcgen05.mma kind::f16 with SWIZZLE_NONE K-major shared-memory descriptors, M = 128, N = 16 / 32: which of LBO / SBO is the
//       stride between core matrices along K, packed tiles with fewer than 128 valid rows, a padded (non power of two) LBO
//   T2  where the rows of an M = 64 accumulator land in TMEM
//   T3  tcgen05.mma issue rate at N = 16 / 32 with A in shared memory; cp.async.bulk global -> shared ring throughput
//   T4  16-CTA cluster all-to-all: cp.async.bulk shared::cta -> shared::cluster with complete_tx on the peer's mbarrier
//       (correctness + latency), against st.shared::cluster + barrier.cluster
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o probe_umma tools/probe_umma.cu
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cooperative_groups.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <math.h>
#include <vector>
namespace cg = cooperative_groups;

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("CUDA error %s at %s:%d\n", cudaGetErrorString(e_), __FILE__, __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n.reg .pred P1;\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\nselp.u32 %0, 1, 0, P1;\n}\n"
               : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  for (uint32_t it = 0; it < (1u << 24); ++it) if (mbar_try(bar, parity)) return;
  printf("mbar_wait timeout block %d thread %d\n", blockIdx.x, threadIdx.x);
  __trap();
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(smem_dst)),
               "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ uint32_t mapa(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void bulk_s2peer(uint32_t dst_cluster_addr, const void* smem_src, uint32_t bytes, uint32_t bar_cluster_addr) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_cluster_addr),
               "r"(smem_u32(smem_src)), "r"(bytes), "r"(bar_cluster_addr) : "memory");
}
// SWIZZLE_NONE descriptor: start address, leading byte offset, stride byte offset (all >> 4), version 1 (sm_100), layout type 0
__device__ __forceinline__ uint64_t desc_ns(uint32_t addr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
__device__ __forceinline__ void umma(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
  asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d), "l"(adesc),
               "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ uint32_t make_idesc(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, float* v) {
  uint32_t r[16];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]), "=r"(r[9]),
                 "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
               : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
  for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
}

// ------------------------------------------------------------------------------------------------------------------------ T1 / T2
struct T1Args {
  const uint8_t* a_img; int a_bytes; const uint8_t* b_img; int b_bytes;
  uint32_t lbo_a, sbo_a, lbo_b, sbo_b, a_kstep, b_kstep;
  int nk, M, N, prezero;
  float* dout;   // [128][32]
};
__global__ void __launch_bounds__(128) t1_kernel(T1Args p) {
  extern __shared__ __align__(1024) uint8_t sm[];
  uint8_t* sa = sm;
  uint8_t* sb = sm + ((p.a_bytes + 1023) & ~1023);
  uint8_t* sz = sb + ((p.b_bytes + 1023) & ~1023);     // 8 KB of zeros (A operand of the pre-zeroing MMA)
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = tid >> 5;
  for (int i = tid; i < p.a_bytes / 4; i += 128) reinterpret_cast<uint32_t*>(sa)[i] = reinterpret_cast<const uint32_t*>(p.a_img)[i];
  for (int i = tid; i < p.b_bytes / 4; i += 128) reinterpret_cast<uint32_t*>(sb)[i] = reinterpret_cast<const uint32_t*>(p.b_img)[i];
  for (int i = tid; i < 8192 / 4; i += 128) reinterpret_cast<uint32_t*>(sz)[i] = 0u;
  if (tid == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(32));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_slot;
  if (tid == 0) {
    if (p.prezero) {   // D[128 lanes][N] = 0 through an M = 128 MMA on a zero A tile (so untouched lanes read back as 0)
      umma(tmem, desc_ns(smem_u32(sz), 128, 256), desc_ns(smem_u32(sb), p.lbo_b, p.sbo_b), make_idesc(128, p.N), 0u);
    }
    const uint32_t idesc = make_idesc(p.M, p.N);
    for (int k = 0; k < p.nk; ++k)
      umma(tmem, desc_ns(smem_u32(sa) + k * p.a_kstep, p.lbo_a, p.sbo_a), desc_ns(smem_u32(sb) + k * p.b_kstep, p.lbo_b, p.sbo_b), idesc,
           (k > 0) ? 1u : 0u);
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  float v[16];
  for (int c0 = 0; c0 < p.N; c0 += 16) {
    tmem_ld16(tmem + ((uint32_t)(warp * 32) << 16) + c0, v);
    for (int j = 0; j < 16; ++j) p.dout[tid * 32 + c0 + j] = v[j];
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(32));
}

// host: build the shared-memory image of a K-major operand: element (row, k) at (k / 8) * kstride + (row / 8) * rstride + (row % 8) * 16 + (k % 8) * 2
static void build_img(std::vector<uint8_t>& img, const std::vector<float>& X, int rows, int K, int kstride, int rstride, int total) {
  img.assign(total, 0);
  for (int r = 0; r < rows; ++r)
    for (int k = 0; k < K; ++k) {
      __nv_bfloat16 b = __float2bfloat16(X[(size_t)r * K + k]);
      size_t off = (size_t)(k / 8) * kstride + (size_t)(r / 8) * rstride + (r % 8) * 16 + (k % 8) * 2;
      if (off + 2 > img.size()) { printf("build_img overflow\n"); exit(1); }
      memcpy(&img[off], &b, 2);
    }
}
static float bf(float x) { return __bfloat162float(__float2bfloat16(x)); }

static double run_t1(const char* name, int rowsA, int M, int N, int K, int kstrideA, int kstrideB, bool swap_hyp, int check_rows) {
  std::vector<float> A((size_t)rowsA * K), B((size_t)N * K);
  srand(1234);
  for (auto& x : A) x = bf((rand() % 2001 - 1000) / 1000.f);
  for (auto& x : B) x = bf((rand() % 2001 - 1000) / 1000.f);
  std::vector<uint8_t> ia, ib;
  const int a_total = (K / 8) * kstrideA + 16 * 128 + 4096;     // slack: M = 128 reads past a packed tile
  const int b_total = (K / 8) * kstrideB + 4096;
  build_img(ia, A, rowsA, K, kstrideA, 128, a_total);
  build_img(ib, B, N, K, kstrideB, 128, b_total);
  uint8_t *da, *db; float* dd;
  CK(cudaMalloc(&da, a_total)); CK(cudaMalloc(&db, b_total)); CK(cudaMalloc(&dd, 128 * 32 * 4));
  CK(cudaMemcpy(da, ia.data(), a_total, cudaMemcpyHostToDevice));
  CK(cudaMemcpy(db, ib.data(), b_total, cudaMemcpyHostToDevice));
  CK(cudaMemset(dd, 0, 128 * 32 * 4));
  T1Args p;
  p.a_img = da; p.a_bytes = a_total; p.b_img = db; p.b_bytes = b_total;
  // hypothesis 1: LBO = stride between core matrices along K, SBO = stride between 8-row groups;  hypothesis 2: swapped
  p.lbo_a = swap_hyp ? 128 : kstrideA; p.sbo_a = swap_hyp ? kstrideA : 128;
  p.lbo_b = swap_hyp ? 128 : kstrideB; p.sbo_b = swap_hyp ? kstrideB : 128;
  p.a_kstep = 2 * kstrideA; p.b_kstep = 2 * kstrideB;    // one MMA = K 16 = two 8-wide k chunks
  p.nk = K / 16; p.M = M; p.N = N; p.prezero = (M == 64) ? 1 : 0; p.dout = dd;
  size_t smem = ((a_total + 1023) & ~1023) + ((b_total + 1023) & ~1023) + 8192 + 1024;
  CK(cudaFuncSetAttribute(t1_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  t1_kernel<<<1, 128, smem>>>(p);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("[T1 %s] kernel failed: %s\n", name, cudaGetErrorString(e)); exit(2); }
  std::vector<float> D(128 * 32);
  CK(cudaMemcpy(D.data(), dd, 128 * 32 * 4, cudaMemcpyDeviceToHost));
  double maxerr = 0;
  if (M == 128) {
    for (int r = 0; r < check_rows; ++r)
      for (int n = 0; n < N; ++n) {
        double ref = 0;
        for (int k = 0; k < K; ++k) ref += (double)A[(size_t)r * K + k] * B[(size_t)n * K + k];
        maxerr = fmax(maxerr, fabs(ref - D[r * 32 + n]));
      }
    printf("[T1 %s] M=%d N=%d K=%d rowsA=%d kstrideA=%d kstrideB=%d hyp=%s  max|err| = %.3e\n", name, M, N, K, rowsA, kstrideA, kstrideB,
           swap_hyp ? "LBO=rowgroup,SBO=k" : "LBO=k,SBO=rowgroup", maxerr);
  } else {
    // M = 64: find where each row landed
    printf("[T2 %s] M=64 accumulator placement (hyp=%s): ", name, swap_hyp ? "swapped" : "LBO=k");
    int found = 0;
    for (int r = 0; r < 64; ++r) {
      double ref0 = 0;
      for (int k = 0; k < K; ++k) ref0 += (double)A[(size_t)r * K + k] * B[k];
      int where = -1;
      for (int lane = 0; lane < 128; ++lane) {
        bool ok = true;
        for (int n = 0; n < N && ok; ++n) {
          double ref = 0;
          for (int k = 0; k < K; ++k) ref += (double)A[(size_t)r * K + k] * B[(size_t)n * K + k];
          if (fabs(ref - D[lane * 32 + n]) > 1e-2) ok = false;
        }
        if (ok) { where = lane; break; }
      }
      if (where >= 0) ++found;
      if (r < 4 || r == 15 || r == 16 || r == 31 || r == 32 || r == 47 || r == 48 || r == 63) printf("row%d->lane%d ", r, where);
      (void)ref0;
    }
    int nz = 0;
    for (int lane = 0; lane < 128; ++lane) if (D[lane * 32] != 0.f) ++nz;
    printf("| rows located %d/64, nonzero lanes %d\n", found, nz);
    maxerr = (found == 64) ? 0 : 1;
  }
  cudaFree(da); cudaFree(db); cudaFree(dd);
  return maxerr;
}

// ------------------------------------------------------------------------------------------------------------------------ T3
// (a) MMA issue rate: nmma MMAs of M=128, N, K=16 on shared-memory operands, one commit; cycles per MMA.
//   VAR 0: issued from a divergent branch (tid == 0), descriptors rebuilt per MMA
//   VAR 1: the whole warp runs the loop on warp-uniform values, elect.sync around the MMA, descriptor = base + constant
//   NACC independent accumulators used round-robin (dependent-accumulate latency vs issue rate)
__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n.reg .b32 rx;\n.reg .pred px;\nelect.sync rx|px, %1;\n@px mov.s32 %0, 1;\n}\n" : "+r"(pred) : "r"(0xFFFFFFFFu));
  return pred;
}
template <int VAR, int NACC>
__global__ void __launch_bounds__(128) t3a_kernel(int N, int nmma, int a_rows_bytes, long long* out, int Mrows = 128) {
  extern __shared__ __align__(1024) uint8_t sm[];
  __shared__ uint64_t bar;
  __shared__ uint32_t tmem_slot;
  const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  for (int i = tid; i < (160 * 1024) / 4; i += 128) reinterpret_cast<uint32_t*>(sm)[i] = 0x3c003c00u;
  if (tid == 0) { mbar_init(&bar, 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_slot)), "r"(512));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem = tmem_slot;
  const uint32_t idesc = make_idesc(Mrows, N);
  const uint32_t a0 = smem_u32(sm), b0 = smem_u32(sm) + 128 * 1024;
  if (VAR == 0) {
    if (tid == 0) {
      for (int rep = 0; rep < 3; ++rep) {
        const long long t0 = clock64();
        for (int k = 0; k < nmma; ++k)   // A: k-chunk stride a_rows_bytes (rows/8*128), B: k-chunk stride N*16
          umma(tmem + (k % NACC) * N, desc_ns(a0 + (k % 32) * 2 * a_rows_bytes, a_rows_bytes, 128), desc_ns(b0 + (k % 8) * 2 * (N * 16), N * 16, 128),
               idesc, k >= NACC);
        const long long t1 = clock64();
        umma_commit(&bar);
        mbar_wait(&bar, rep & 1);
        const long long t2 = clock64();
        out[rep * 2] = t1 - t0; out[rep * 2 + 1] = t2 - t0;
      }
    }
  } else {
    if (warp == 0) {
      const uint64_t ad0 = desc_ns(a0, a_rows_bytes, 128), bd0 = desc_ns(b0, N * 16, 128);
      const uint32_t astep = (uint32_t)(2 * a_rows_bytes) >> 4, bstep = (uint32_t)(2 * N * 16) >> 4;
      for (int rep = 0; rep < 3; ++rep) {
        const long long t0 = clock64();
#pragma unroll 8
        for (int k = 0; k < nmma; ++k) {
          const uint64_t ad = ad0 + (uint64_t)((k & 31) * astep), bd = bd0 + (uint64_t)((k & 7) * bstep);
          if (elect_one()) umma(tmem + (k % NACC) * N, ad, bd, idesc, k >= NACC);
        }
        const long long t1 = clock64();
        if (elect_one()) umma_commit(&bar);
        __syncwarp();
        mbar_wait(&bar, rep & 1);
        const long long t2 = clock64();
        if (tid == 0) { out[rep * 2] = t1 - t0; out[rep * 2 + 1] = t2 - t0; }
      }
    }
  }
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}
// (b) bulk-copy ring throughput: one thread streams nchunks chunks of `bytes` through an NS-stage ring, a second thread consumes
__global__ void __launch_bounds__(64) t3b_kernel(const uint8_t* src, size_t src_bytes, int bytes, int nchunks, int NS, long long* out) {
  extern __shared__ __align__(1024) uint8_t sm[];
  __shared__ uint64_t full[16], empty[16];
  const int tid = threadIdx.x;
  if (tid == 0) {
    for (int s = 0; s < NS; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  const uint8_t* base = src + (size_t)blockIdx.x * (4u << 20);
  const uint32_t nslot = (2u << 20) / (uint32_t)bytes;      // chunk slots inside this CTA's 4 MB window (wraps: L2 resident)
  long long t0 = clock64();
  const int warp = __shfl_sync(0xffffffffu, tid >> 5, 0);
  if (warp == 0) {            // producer: warp-uniform loop, one elected lane issues; addresses advance incrementally
    uint32_t slot = 0, s = 0, ph = 1;
    for (int i = 0; i < nchunks; ++i) {
      mbar_wait(&empty[s], ph);
      if (elect_one()) {
        mbar_expect_tx(&full[s], bytes);
        bulk_g2s(sm + (size_t)s * bytes, base + (size_t)slot * bytes, bytes, &full[s]);
      }
      __syncwarp();
      if (++slot == nslot) slot = 0;
      if (++s == (uint32_t)NS) { s = 0; ph ^= 1; }
    }
  } else {
    uint32_t s = 0, ph = 0;
    for (int i = 0; i < nchunks; ++i) {
      mbar_wait(&full[s], ph);
      if (elect_one()) mbar_arrive(&empty[s]);
      __syncwarp();
      if (++s == (uint32_t)NS) { s = 0; ph ^= 1; }
    }
    if (tid == 32) out[blockIdx.x] = clock64() - t0;
  }
}

// ------------------------------------------------------------------------------------------------------------------------ T4
constexpr int CLS = 16;
// mode 0: bulk copy smem -> peer smem, complete_tx on the peer's mbarrier.  mode 1: 16 B st.shared::cluster + barrier.cluster.
__global__ void __launch_bounds__(256) t4_kernel(int mode, int bytes, int iters, long long* out_clk, int* out_err) {
  cg::cluster_group cluster = cg::this_cluster();
  extern __shared__ __align__(1024) uint8_t sm[];
  uint8_t* stage = sm;                                   // own slice [bytes]
  uint8_t* recv = sm + 4096;                             // [2][CLS][bytes]
  __shared__ uint64_t bars[2];
  const int tid = threadIdx.x, c = (int)cluster.block_rank();
  if (tid == 0) { mbar_init(&bars[0], 1); mbar_init(&bars[1], 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
  __syncthreads();
  cluster.sync();
  int errs = 0;
  long long t_sum = 0;
  for (int it = 0; it < iters; ++it) {
    const int pb = it & 1;
    uint8_t* rb = recv + (size_t)pb * CLS * bytes;
    for (int i = tid; i < bytes / 4; i += 256) reinterpret_cast<uint32_t*>(stage)[i] = (uint32_t)(it * 1000003 + c * 4099 + i);
    long long t0 = 0;
    if (mode == 0) {
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncthreads();
      t0 = clock64();
      if (tid == 0) mbar_expect_tx(&bars[pb], CLS * bytes);
      if (tid < CLS) bulk_s2peer(mapa(smem_u32(rb + (size_t)c * bytes), tid), stage, bytes, mapa(smem_u32(&bars[pb]), tid));
      mbar_wait(&bars[pb], (it >> 1) & 1);
    } else {
      __syncthreads();
      t0 = clock64();
      const int per = bytes / 16;
      for (int i = tid; i < CLS * per; i += 256) {
        const int peer = i / per, ch = i - peer * per;
        uint4* dst = reinterpret_cast<uint4*>(cluster.map_shared_rank(rb + (size_t)c * bytes, peer)) + ch;
        *dst = reinterpret_cast<const uint4*>(stage)[ch];
      }
      asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
      asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
    }
    const long long t1 = clock64();
    if (it >= 4) t_sum += t1 - t0;
    // verify: word 0 and the last word of every peer's slice
    if (tid < CLS) {
      const uint32_t w0 = reinterpret_cast<const uint32_t*>(rb + (size_t)tid * bytes)[0];
      const uint32_t wl = reinterpret_cast<const uint32_t*>(rb + (size_t)tid * bytes)[bytes / 4 - 1];
      if (w0 != (uint32_t)(it * 1000003 + tid * 4099) || wl != (uint32_t)(it * 1000003 + tid * 4099 + bytes / 4 - 1)) ++errs;
    }
    __syncthreads();      // stage is rewritten next iteration: the copies that read it have completed at every peer only once every
                          // peer's barrier completed; the next iteration's data cannot arrive before that (see DESIGN: all-to-all ordering)
    if (mode == 0) {      // the bulk copy reads `stage` asynchronously: make sure our own copies have landed everywhere before rewriting
      cluster.sync();     // (probe only: conservative)
    }
  }
  if (tid == 0) out_clk[c] = t_sum / (iters - 4);
  if (tid < CLS) atomicAdd(out_err, errs);
  cluster.sync();
}

static void run_t4(int mode, int bytes) {
  long long* dclk; int* derr;
  CK(cudaMalloc(&dclk, CLS * 8)); CK(cudaMalloc(&derr, 4));
  CK(cudaMemset(derr, 0, 4));
  size_t smem = 4096 + 2 * CLS * bytes + 1024;
  CK(cudaFuncSetAttribute(t4_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  CK(cudaFuncSetAttribute(t4_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(CLS); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = smem;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension; at[0].val.clusterDim.x = CLS; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  CK(cudaLaunchKernelEx(&cfg, t4_kernel, mode, bytes, 200, dclk, derr));
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { printf("[T4 mode %d bytes %d] kernel failed: %s\n", mode, bytes, cudaGetErrorString(e)); exit(3); }
  long long clk[CLS]; int err;
  CK(cudaMemcpy(clk, dclk, CLS * 8, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(&err, derr, 4, cudaMemcpyDeviceToHost));
  long long mx = 0, mn = 1ll << 60;
  for (int i = 0; i < CLS; ++i) { mx = clk[i] > mx ? clk[i] : mx; mn = clk[i] < mn ? clk[i] : mn; }
  printf("[T4 %s] %d B per peer (x16 peers): exchange %lld..%lld clk per all-to-all, errors %d\n",
         mode == 0 ? "bulk smem->peer + mbarrier complete_tx" : "st.shared::cluster 16B + barrier.cluster", bytes, mn, mx, err);
  cudaFree(dclk); cudaFree(derr);
}

int main(int argc, char** argv) {
  int dev = 0;
  CK(cudaSetDevice(dev));
  cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, dev));
  printf("device %s sm_%d%d, %d SMs, clock %d kHz\n", prop.name, prop.major, prop.minor, prop.multiProcessorCount, prop.clockRate);
  const char* only = argc > 1 ? argv[1] : "all";
  auto want = [&](const char* t) { return !strcmp(only, "all") || !strcmp(only, t); };

  if (want("t1")) {
    // full tile: 128 rows, K = 64, k-chunk stride 2048 (16 row groups x 128 B); B: N = 16, stride 256
    run_t1("full-h1", 128, 128, 16, 64, 2048, 256, false, 128);
    // packed 96-row tile (k-chunk stride 1536) read by an M = 128 MMA; padded B stride 272
    run_t1("packed96-h1", 96, 128, 16, 128, 1536, 272, false, 96);
    run_t1("packed32-h1", 32, 128, 16, 256, 512, 272, false, 32);
    run_t1("n32-h1", 128, 128, 32, 64, 2048, 512, false, 128);
    run_t1("n32-pad-h1", 40, 128, 32, 128, 640, 528, false, 40);
  }
  if (want("t2")) {
    run_t1("m64-h1", 64, 64, 16, 64, 1024, 256, false, 64);
  }
  if (want("t3")) {
    long long* dout; CK(cudaMalloc(&dout, 256 * 8));
    auto run3a = [&](auto kern, const char* name, int N, int rows_bytes) {
      CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      kern<<<1, 128, 200 * 1024>>>(N, 64, rows_bytes, dout, 128);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("[T3a] failed: %s\n", cudaGetErrorString(e)); exit(4); }
      long long h[6]; CK(cudaMemcpy(h, dout, 48, cudaMemcpyDeviceToHost));
      printf("[T3a %s] 64 MMAs M=128 N=%d K=16, A k-chunk stride %d B: issue %lld clk, issue+complete %lld clk (%.1f clk/MMA)\n", name, N, rows_bytes, h[4],
             h[5], h[5] / 64.0);
    };
    for (int N : {16, 32, 128, 256}) {
      run3a(t3a_kernel<0, 1>, "divergent,1acc", N, 2048);
      run3a(t3a_kernel<1, 1>, "uniform,1acc", N, 2048);
      if (N <= 128) run3a(t3a_kernel<1, 2>, "uniform,2acc", N, 2048);
      if (N <= 64) run3a(t3a_kernel<1, 4>, "uniform,4acc", N, 2048);
      if (N <= 64) run3a(t3a_kernel<0, 4>, "divergent,4acc", N, 2048);
    }
    run3a(t3a_kernel<1, 4>, "uniform,4acc,packed32", 16, 512);
    // M = 64 tiles (A k-chunk = 8 row groups x 128 B): does the per-instruction cost follow the A-operand bytes?
    auto run3m = [&](const char* name, int M, int N, int rows_bytes) {
      auto kern = t3a_kernel<1, 1>;
      CK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
      kern<<<1, 128, 200 * 1024>>>(N, 64, rows_bytes, dout, M);
      cudaError_t e = cudaDeviceSynchronize();
      if (e != cudaSuccess) { printf("[T3m] failed: %s\n", cudaGetErrorString(e)); exit(4); }
      long long h[6]; CK(cudaMemcpy(h, dout, 48, cudaMemcpyDeviceToHost));
      printf("[T3m %s] 64 MMAs M=%d N=%d K=16, A k-chunk stride %d B: issue %lld clk, issue+complete %lld clk (%.1f clk/MMA)\n", name, M, N, rows_bytes,
             h[4], h[5], h[5] / 64.0);
    };
    for (int N : {8, 16, 32, 64}) run3m("uniform,1acc", 64, N, 1024);
    for (int N : {16, 64}) run3m("uniform,1acc", 128, N, 2048);
    size_t src_bytes = 148ull * (4u << 20);
    uint8_t* src; CK(cudaMalloc(&src, src_bytes)); CK(cudaMemset(src, 1, src_bytes));
    CK(cudaFuncSetAttribute(t3b_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    for (int grid : {1, 16, 148}) {
      for (int bytes : {4096, 12288, 16384}) {
        const int NS = 192 * 1024 / bytes > 12 ? 12 : 192 * 1024 / bytes;
        for (int rep = 0; rep < 2; ++rep) {   // second repetition: source resident in L2
          t3b_kernel<<<grid, 64, 200 * 1024>>>(src, src_bytes, bytes, 512, NS, dout);
          cudaError_t e = cudaDeviceSynchronize();
          if (e != cudaSuccess) { printf("[T3b] failed: %s\n", cudaGetErrorString(e)); exit(5); }
          long long h[148]; CK(cudaMemcpy(h, dout, grid * 8, cudaMemcpyDeviceToHost));
          long long mx = 0;
          for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
          if (rep == 1)
            printf("[T3b] grid %d: 512 chunks of %d B through %d stages: %lld clk -> %.1f B/clk per SM\n", grid, bytes, NS, mx, 512.0 * bytes / mx);
        }
      }
    }
    cudaFree(src); cudaFree(dout);
  }
  if (want("t4")) {
    for (int bytes : {128, 1024, 2048}) { run_t4(0, bytes); run_t4(1, bytes); }
  }
  printf("probe done\n");
  return 0;
}
