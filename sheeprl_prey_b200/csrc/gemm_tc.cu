// bf16 tensor-core contraction on the 5th-generation tensor cores (tcgen05) with fp32 accumulation in TMEM.
//
//   C[M,N] (+)= op(A) * op(B) (+ bias)      NT / NN / TN, fp32 operands in HBM, arbitrary leading dimensions, ragged edges.
//
// Design (sm_100a only):
//   * one CTA per 128 x 128 output tile, 384 threads, warp-specialised:
//       warps 0-3  producers A, warps 4-7 producers B : read the fp32 operand tiles from global memory (coalesced for either
//                              majorness), round to bf16 in registers and write them into shared memory directly in the canonical
//                              SWIZZLE_128B UMMA layout -- K-major, or MN-major for operands whose contiguous dimension is M / N
//                              (no transposition) -- so the fp32->bf16 conversion is fused into the load path;
//       warp  8    issuer    : allocates TMEM, one elected lane issues tcgen05.mma (M128 x N128 x K16, kind::f16) per
//                              16-wide k slice, and tcgen05.commit's the stage's "empty" mbarrier;
//       warps 8-11 epilogue  : tcgen05.ld the 128x128 fp32 accumulator (one TMEM lane = one output row), transpose 32 x 32
//                              chunks through shared memory, add bias / accumulate / split-K atomics, coalesced stores.
//   * 4-stage shared-memory ring (32 KB / stage) with full / empty mbarriers; generic-proxy writes are made visible to
//     the tensor core's async proxy with fence.proxy.async before the "full" arrive.
// Used for operands that have no bf16 mirror (gemm_tma.cu takes the mirrored ones).  Replaces torch addmm/mm (cuBLAS) on the large-M contractions of the step (a3, a5, a8, a9, a12, a13 and
// the batched weight-gradient contractions of a7).  Numerics: bf16 operands, fp32 accumulate -> the north-star's
// "2e-2 for bf16 tensor-core GEMMs" mode.
#include "common.cuh"
#include <cuda_bf16.h>
#include <stdlib.h>

namespace {

constexpr int TBM = 128, TBN = 128, TBK = 64, STAGES = 4;
constexpr int NUM_PRODUCERS = 256, NUM_THREADS = 384;
constexpr int STAGE_A_BYTES = TBM * TBK * 2, STAGE_B_BYTES = TBN * TBK * 2;
constexpr int TMEM_COLS = 128;
enum { MODE_NT = 0, MODE_NN = 1, MODE_TN = 2 };

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n"
      ".reg .pred P1;\n"
      "WAIT_LOOP:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 P1, [%0], %1;\n"
      "@P1 bra DONE;\n"
      "bra WAIT_LOOP;\n"
      "DONE:\n"
      "}\n" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// K-major SWIZZLE_128B shared-memory matrix descriptor (cute::UMMA::SmemDescriptor): 8-row groups 1024 B apart.
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);   // start address, 16 B units
  d |= (uint64_t)1 << 16;                        // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;              // stride byte offset: next 8-row group
  d |= (uint64_t)1 << 46;                        // descriptor version (Blackwell)
  d |= (uint64_t)2 << 61;                        // SWIZZLE_128B
  return d;
}
// MN-major SWIZZLE_128B descriptor: the tile is stored as [64-element MN groups][8-deep k groups][8 k rows x 128 B];
// LBO = byte stride between 64-element MN groups, SBO = byte stride between 8-deep k groups (cute canonical layout
// ((T,8,m),(8,k)) : ((1,T,LBO),(8T,SBO)) for Major-MN, T = 8 bf16).
__device__ __forceinline__ uint64_t make_smem_desc_mn(uint32_t smem_addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// Instruction descriptor (cute::UMMA::InstrDescriptor): D=F32, A=B=BF16, both K-major, N>>3 at [17,23), M>>4 at [24,29)
__device__ __forceinline__ uint32_t make_instr_desc(int m, int n, bool a_mn_major, bool b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((a_mn_major ? 1u : 0u) << 15) | ((b_mn_major ? 1u : 0u) << 16) |
         ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// Loads one [128 rows][64 k] operand tile (fp32, either majorness) and stores it as bf16 in the UMMA layout.
//   kcontig: element (r, k) at src[(r0 + r) * ld + k0 + k]     else: src[(k0 + k) * ld + r0 + r]
constexpr int TILE_THREADS = 128;   // producer threads per operand tile (A and B tiles are staged concurrently)

template <bool KCONTIG>
__device__ __forceinline__ void load_tile(const float* __restrict__ src, int ld, int r0, int k0, int R, int K, bool vec_ok,
                                          uint8_t* smem_tile, int p) {
  constexpr int PER = (TBM * TBK / 8) / TILE_THREADS;
  float v[PER][8];
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    const int item = p + TILE_THREADS * i;
    if (KCONTIG) {
      const int row = item >> 3, chunk = item & 7;
      const int gr = r0 + row, gk = k0 + chunk * 8;
      const float* ptr = src + (size_t)gr * ld + gk;
      if (gr < R && vec_ok && gk + 8 <= K) {
        const float4 a = __ldg(reinterpret_cast<const float4*>(ptr));
        const float4 b = __ldg(reinterpret_cast<const float4*>(ptr) + 1);
        v[i][0] = a.x; v[i][1] = a.y; v[i][2] = a.z; v[i][3] = a.w;
        v[i][4] = b.x; v[i][5] = b.y; v[i][6] = b.z; v[i][7] = b.w;
      } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) v[i][j] = (gr < R && gk + j < K) ? __ldg(ptr + j) : 0.f;
      }
    } else {
      const int row = item & 127, chunk = item >> 7;
      const int gr = r0 + row, gk = k0 + chunk * 8;
#pragma unroll
      for (int j = 0; j < 8; ++j) v[i][j] = (gr < R && gk + j < K) ? __ldg(src + (size_t)(gk + j) * ld + gr) : 0.f;
    }
  }
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    const int item = p + TILE_THREADS * i;
    const int row = KCONTIG ? (item >> 3) : (item & 127);
    const int chunk = KCONTIG ? (item & 7) : (item >> 7);
    const uint32_t off = (uint32_t)(row >> 3) * 1024u + (uint32_t)(row & 7) * 128u + (uint32_t)((chunk ^ (row & 7)) << 4);
    uint4 q;
    q.x = pack_bf16(v[i][0], v[i][1]);
    q.y = pack_bf16(v[i][2], v[i][3]);
    q.z = pack_bf16(v[i][4], v[i][5]);
    q.w = pack_bf16(v[i][6], v[i][7]);
    *reinterpret_cast<uint4*>(smem_tile + off) = q;
  }
}

// MN-contiguous source kept MN-major in shared memory (no transposition): element (r, k) at src[(k0 + k) * ld + r0 + r].
// Each item is 8 consecutive r (two float4 loads) at one k -> one 16 B store.  Tile = 2 MN groups x 8 k groups of 1 KB atoms.
__device__ __forceinline__ void load_tile_mn(const float* __restrict__ src, int ld, int r0, int k0, int R, int K, bool vec_ok,
                                             uint8_t* smem_tile, int p) {
  constexpr int ITEMS = (TBM / 8) * TBK;              // 16 chunks x 64 k = 1024
  constexpr int PER = ITEMS / TILE_THREADS;
  float v[PER][8];
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    const int item = p + TILE_THREADS * i;
    const int chunk = item & 15, k = item >> 4;       // 16 consecutive threads cover 128 contiguous r
    const int gr = r0 + chunk * 8, gk = k0 + k;
    const float* ptr = src + (size_t)gk * ld + gr;
    if (gk < K && vec_ok && gr + 8 <= R) {
      const float4 a = __ldg(reinterpret_cast<const float4*>(ptr));
      const float4 b = __ldg(reinterpret_cast<const float4*>(ptr) + 1);
      v[i][0] = a.x; v[i][1] = a.y; v[i][2] = a.z; v[i][3] = a.w;
      v[i][4] = b.x; v[i][5] = b.y; v[i][6] = b.z; v[i][7] = b.w;
    } else {
#pragma unroll
      for (int j = 0; j < 8; ++j) v[i][j] = (gk < K && gr + j < R) ? __ldg(ptr + j) : 0.f;
    }
  }
#pragma unroll
  for (int i = 0; i < PER; ++i) {
    const int item = p + TILE_THREADS * i;
    const int chunk = item & 15, k = item >> 4;
    const uint32_t off = (uint32_t)(chunk >> 3) * 8192u + (uint32_t)(k >> 3) * 1024u + (uint32_t)(k & 7) * 128u +
                         (uint32_t)(((chunk & 7) ^ (k & 7)) << 4);
    uint4 q;
    q.x = pack_bf16(v[i][0], v[i][1]);
    q.y = pack_bf16(v[i][2], v[i][3]);
    q.z = pack_bf16(v[i][4], v[i][5]);
    q.w = pack_bf16(v[i][6], v[i][7]);
    *reinterpret_cast<uint4*>(smem_tile + off) = q;
  }
}

__global__ void zero2d_kernel(int M, int N, float* __restrict__ C, int ldc) {
  const size_t n = (size_t)M * N;
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    C[(i / N) * ldc + (i % N)] = 0.f;
}

__device__ unsigned long long g_tc_stamps[16];     // diagnostics: %globaltimer stamps of CTA (0,0,0) (tools/gemm_bench.py)
__device__ __forceinline__ unsigned long long gtimer() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define TCSTAMP(i) do { if (blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (threadIdx.x & 31) == 0) g_tc_stamps[i] = gtimer(); } while (0)

template <int MODE, bool MNMAJ>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tc_kernel(int M, int N, int K, const float* __restrict__ A, int lda, const float* __restrict__ B, int ldb,
               float* __restrict__ C, int ldc, const float* __restrict__ bias, int accumulate, int kb_per_split) {
  extern __shared__ uint8_t smem_raw[];
  // SWIZZLE_128B needs 1024 B alignment.  Plain pointer arithmetic on the __shared__ array (no integer round trip) keeps the
  // address space visible to the compiler, so the staging stores / epilogue loads compile to STS / LDS instead of generic ST / LD.
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* smem_a = smem;
  uint8_t* smem_b = smem + STAGES * STAGE_A_BYTES;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + STAGES * (STAGE_A_BYTES + STAGE_B_BYTES));
  uint64_t* empty_bar = full_bar + STAGES;
  uint64_t* tmem_full_bar = empty_bar + STAGES;
  uint32_t* tmem_ptr_smem = reinterpret_cast<uint32_t*>(tmem_full_bar + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.y * TBM, n0 = blockIdx.x * TBN;
  if (threadIdx.x == 0) TCSTAMP(0);
  const int total_kb = (K + TBK - 1) / TBK;
  const int kb_begin = blockIdx.z * kb_per_split;                 // split-K: this CTA owns k-blocks [kb_begin, kb_begin + num_kb)
  const int num_kb = min(kb_per_split, total_kb - kb_begin);
  const bool split = gridDim.z > 1;                                // partial tiles are combined with red.global.add

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], NUM_PRODUCERS); mbar_init(&empty_bar[s], 1); }
    mbar_init(tmem_full_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == NUM_PRODUCERS / 32) {   // TMEM allocation is warp-collective; the same warp frees it at the end
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_ptr_smem)), "r"(TMEM_COLS));
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tmem_base = *tmem_ptr_smem;
  if (threadIdx.x == 0) TCSTAMP(1);

  constexpr bool A_KCONTIG = (MODE != MODE_TN);
  constexpr bool B_KCONTIG = (MODE == MODE_NT);
  constexpr bool A_MN = MNMAJ && !A_KCONTIG;      // operand stays MN-major in shared memory
  constexpr bool B_MN = MNMAJ && !B_KCONTIG;
  if (warp < NUM_PRODUCERS / 32) {
    // ===================== producers =====================
    const int p = threadIdx.x & (TILE_THREADS - 1);
    const bool is_b = threadIdx.x >= TILE_THREADS;    // warps 0-3 stage the A tile, warps 4-7 the B tile, concurrently
    const bool a_vec = ((lda & 3) == 0) && ((((uintptr_t)A) & 15) == 0);
    const bool b_vec = ((ldb & 3) == 0) && ((((uintptr_t)B) & 15) == 0);
    for (int kb = 0; kb < num_kb; ++kb) {
      const int s = kb % STAGES;
      const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
      mbar_wait(&empty_bar[s], ph ^ 1u);
      const int k0 = (kb_begin + kb) * TBK;
      if (!is_b) {
        if (A_MN) load_tile_mn(A, lda, m0, k0, M, K, a_vec, smem_a + s * STAGE_A_BYTES, p);
        else load_tile<A_KCONTIG>(A, lda, m0, k0, M, K, a_vec && A_KCONTIG, smem_a + s * STAGE_A_BYTES, p);
      } else {
        if (B_MN) load_tile_mn(B, ldb, n0, k0, N, K, b_vec, smem_b + s * STAGE_B_BYTES, p);
        else load_tile<B_KCONTIG>(B, ldb, n0, k0, N, K, b_vec && B_KCONTIG, smem_b + s * STAGE_B_BYTES, p);
      }
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy stores -> visible to the async (UMMA) proxy
      mbar_arrive(&full_bar[s]);
      if (threadIdx.x == 0 && kb == 0) TCSTAMP(2);
      if (threadIdx.x == 0 && kb == num_kb - 1) TCSTAMP(3);
    }
  } else {
    if (warp == NUM_PRODUCERS / 32) {
      // ===================== MMA issuer =====================
      const uint32_t idesc = make_instr_desc(TBM, TBN, A_MN, B_MN);
      for (int kb = 0; kb < num_kb; ++kb) {
        const int s = kb % STAGES;
        const uint32_t ph = (uint32_t)(kb / STAGES) & 1u;
        mbar_wait(&full_bar[s], ph);
        asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
        if (lane == 0) {
          const uint32_t a_addr = smem_u32(smem_a + s * STAGE_A_BYTES);
          const uint32_t b_addr = smem_u32(smem_b + s * STAGE_B_BYTES);
#pragma unroll
          for (int k = 0; k < TBK / 16; ++k) {
            // K-major: 16 bf16 = 32 B along K inside the swizzle atom.  MN-major: 16 k = two 1 KB k-groups.
            const uint64_t adesc = A_MN ? make_smem_desc_mn(a_addr + k * 2048, 8192, 1024) : make_smem_desc(a_addr + k * 32);
            const uint64_t bdesc = B_MN ? make_smem_desc_mn(b_addr + k * 2048, 8192, 1024) : make_smem_desc(b_addr + k * 32);
            const uint32_t accum = (kb > 0 || k > 0) ? 1u : 0u;
            asm volatile(
                "{\n"
                ".reg .pred p;\n"
                "setp.ne.b32 p, %4, 0;\n"
                "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
                "}\n" ::"r"(tmem_base),
                "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accum)
                : "memory");
          }
          // frees the smem stage once the MMAs above have read it (implies tcgen05.fence::before_thread_sync)
          asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(&empty_bar[s])) : "memory");
          if (kb == num_kb - 1)
            asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(tmem_full_bar)) : "memory");
        }
        __syncwarp();
      }
    }
    // ===================== epilogue (4 warps; warp e reads TMEM lanes [32e, 32e+32) = output rows) =====================
    // The accumulator arrives as one row per thread.  Each 32x32 chunk is transposed through shared memory (the operand ring
    // is idle once the last MMA has committed) so that global stores / read-modify-writes / split-K reductions are issued with
    // the 32 lanes of a warp on 32 consecutive columns of one row: full 128-byte transactions instead of 32 scattered sectors.
    mbar_wait(tmem_full_bar, 0);
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const int ew = warp & 3;
    if (ew == 0) TCSTAMP(4);
    float* stg = reinterpret_cast<float*>(smem) + ew * (32 * 33);
#pragma unroll 1
    for (int c0 = 0; c0 < TBN; c0 += 32) {
      if (n0 + c0 >= N) break;
      uint32_t r[32];
      const uint32_t taddr = tmem_base + ((uint32_t)(ew * 32) << 16) + (uint32_t)c0;
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
          "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
          "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];\n"
          : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
            "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
            "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
            "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
          : "r"(taddr));
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
      if (ew == 0 && c0 == 0) TCSTAMP(7);
#pragma unroll
      for (int j = 0; j < 32; ++j) stg[lane * 33 + j] = __uint_as_float(r[j]);      // bank (lane + j) % 32: conflict-free
      __syncwarp();
      if (ew == 0 && c0 == 0) TCSTAMP(8);
      const int gn = n0 + c0 + lane;
      const bool col_ok = gn < N;
      const float bv = (bias && col_ok && (!split || blockIdx.z == 0)) ? __ldg(bias + gn) : 0.f;
      const int rows_here = col_ok ? min(32, M - (m0 + ew * 32)) : 0;
      float* dst0 = C + (size_t)(m0 + ew * 32) * ldc + gn;
      float o[32];
#pragma unroll
      for (int rr = 0; rr < 32; ++rr) o[rr] = stg[rr * 33 + lane] + bv;            // all 32 reads in flight, conflict-free
      // Full 32 x 32 chunk (the common case): branch-free unrolled stores.  Ragged edges take the predicated path.
      const bool full = __all_sync(0xffffffffu, rows_here == 32);
      const size_t ld = (size_t)ldc;
      if (full) {
        if (split) {
#pragma unroll
          for (int rr = 0; rr < 32; ++rr) atomicAdd(dst0 + rr * ld, o[rr]);
        } else if (accumulate) {
          float old[32];
#pragma unroll
          for (int rr = 0; rr < 32; ++rr) old[rr] = dst0[rr * ld];             // independent loads first
#pragma unroll
          for (int rr = 0; rr < 32; ++rr) dst0[rr * ld] = old[rr] + o[rr];
        } else {
#pragma unroll
          for (int rr = 0; rr < 32; ++rr) dst0[rr * ld] = o[rr];
        }
      } else {
#pragma unroll 1
        for (int rr = 0; rr < rows_here; ++rr) {
          float* d = dst0 + rr * ld;
          if (split) atomicAdd(d, o[rr]);
          else *d = accumulate ? (*d + o[rr]) : o[rr];
        }
      }
      __syncwarp();
      if (ew == 0 && c0 == 0) TCSTAMP(9);
    }
    if (ew == 0) TCSTAMP(5);
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  }
  __syncthreads();
  if (threadIdx.x == 0) TCSTAMP(6);
  if (warp == NUM_PRODUCERS / 32) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS));
  }
}

constexpr int SMEM_BYTES = STAGES * (STAGE_A_BYTES + STAGE_B_BYTES) + (2 * STAGES + 1) * 8 + 16 + 1024;

template <int MODE, bool MNMAJ>
int launch_tc(int M, int N, int K, const float* A, int lda, const float* B, int ldb, float* C, int ldc, const float* bias,
              int accumulate, cudaStream_t stream) {
  static unsigned long long configured = 0;      // per device: the attribute does not carry over to other GPUs of the process
  if (dv3_first_call_on_device(&configured)) {
    cudaError_t e = cudaFuncSetAttribute(gemm_tc_kernel<MODE, MNMAJ>, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_BYTES);
    if (e != cudaSuccess) return (int)e;
  }
  dim3 grid((N + TBN - 1) / TBN, (M + TBM - 1) / TBM, 1);
  if (grid.y > 65535) return DV3_ERR_UNSUPPORTED;
  // split-K when the output has too few tiles to fill the 148 SMs (weight-gradient contractions: K = T*B*H tokens)
  const int total_kb = (K + TBK - 1) / TBK;
  const int tiles = (int)(grid.x * grid.y);
  int splits = 1;
  if (tiles < 120 && total_kb >= 32) {   // only long-K problems: a split needs a zero-fill launch and atomic epilogues
    splits = (296 + tiles - 1) / tiles;
    if (splits > total_kb / 4) splits = total_kb / 4;
    if (splits > 64) splits = 64;
    if (splits < 1) splits = 1;
  }
  int kb_per_split = (total_kb + splits - 1) / splits;
  splits = (total_kb + kb_per_split - 1) / kb_per_split;
  grid.z = splits;
  if (splits > 1 && !accumulate) {
    size_t n = (size_t)M * N;
    int zb = (int)((n + 255) / 256);
    if (zb > 1184) zb = 1184;
    zero2d_kernel<<<zb, 256, 0, stream>>>(M, N, C, ldc);
    DV3_LAUNCH_CHECK();
  }
  gemm_tc_kernel<MODE, MNMAJ><<<grid, NUM_THREADS, SMEM_BYTES, stream>>>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate, kb_per_split);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}

}  // namespace

extern "C" int dv3_gemm_tc_debug_stamps(unsigned long long* host_out) {
  return (int)cudaMemcpyFromSymbol(host_out, g_tc_stamps, sizeof(unsigned long long) * 16);
}

static int g_tc_mn_major = -1;   // operands whose contiguous dimension is M/N: keep them MN-major in smem (1) or transpose (0)

extern "C" int dv3_gemm_bf16_tc(int mode, int M, int N, int K, const float* A, int lda, const float* B, int ldb, float* C,
                                int ldc, const float* bias, int accumulate, void* stream_) {
  if (M <= 0 || N <= 0 || K <= 0 || !A || !B || !C) return DV3_ERR_ARG;
  cudaStream_t stream = (cudaStream_t)stream_;
  if (g_tc_mn_major < 0) {
    const char* e = getenv("DV3_TC_MNMAJOR");
    g_tc_mn_major = (e && e[0] == '0') ? 0 : 1;
  }
  switch (mode) {
    case MODE_NT: return launch_tc<MODE_NT, false>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate, stream);
    case MODE_NN:
      return g_tc_mn_major ? launch_tc<MODE_NN, true>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate, stream)
                           : launch_tc<MODE_NN, false>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate, stream);
    case MODE_TN:
      return g_tc_mn_major ? launch_tc<MODE_TN, true>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate, stream)
                           : launch_tc<MODE_TN, false>(M, N, K, A, lda, B, ldb, C, ldc, bias, accumulate, stream);
    default: return DV3_ERR_ARG;
  }
}
