// Shared machinery of the two tcgen05 scan kernels (rssm_scan_tc.cu forward, rssm_scan_bwd_tc.cu reverse).
//
// One thread-block cluster of CL CTAs runs all T steps of the recurrence for NB = 16 batch rows (B > 16: one cluster per 16 rows).
// Every in-loop contraction is "swap-AB": this CTA's rows of a weight matrix are the M = 128 operand of tcgen05.mma (kind::f16, bf16
// in, fp32 accumulate in TMEM), the 16 batch rows are N = 16.
//   * warp 0 (producer): streams this CTA's weight rows, pre-packed per CTA and per step in consumption order in the shared-memory
//     image the MMA reads (SWIZZLE_NONE K-major core matrices), with cp.async.bulk into an mbarrier ring.  The weights do not depend
//     on the recurrence, so the ring runs ahead of the dependent chain by its whole depth.
//   * warp 1 (MMA): waits for the activations of a phase (mbarrier armed with the bytes the peers will deliver), issues the MMAs of
//     the phase chunk by chunk (warp-uniform code, elect.sync: 44 clk per instruction at N = 16, profiles/r02a), commits the ring
//     slot back to the producer and the accumulator to the epilogue (tcgen05.commit).
//   * warps 2-9 (epilogue): TMEM -> shared memory, LayerNorm statistics, gates / SiLU / sampling, global stores of everything the
//     hand-written backward needs, and the exchange with the peers.
// Exchange between CTAs needs no cluster barrier: LayerNorm partials and sampled indices travel as st.async (remote store that
// completes transaction bytes on the RECEIVER's mbarrier), activation slices as cp.async.bulk shared::cta -> shared::cluster
// straight into the peers' MMA operand buffers, again completing on the receiver's mbarrier.  Every exchange point has its own
// mbarrier and buffer, used once per step; a CTA cannot run a whole step ahead of a peer because every phase needs data from all
// peers, which is what makes the reuse of buffers and barrier phases safe ("tcgen05 scans").
// All bulk copies and TMA-class loads here read this GPU's memory only.
#pragma once
#include "common.cuh"
#include <cuda_bf16.h>

namespace tcs {

constexpr int NB = 16;                 // batch rows per cluster = N of the MMA
constexpr int OB = NB + 1;             // row stride of the [feature][batch] staging buffers (odd: conflict-free both ways)
constexpr int EW = 8;                  // epilogue warps
constexpr int NT_EPI = EW * 32;
constexpr int NT = 64 + NT_EPI;        // warp 0 producer, warp 1 MMA issuer, warps 2..9 epilogue
constexpr int MAX_CHUNKS = 200;
constexpr int MAX_SEG = 6;
constexpr int BROW = NB * 16;          // bytes of one 8-wide k chunk of an activation operand (NB rows x 16 B)

// One bulk copy of the weight stream = `nmma` MMAs of one 128-row tile.
struct Chunk {
  uint32_t off16;      // offset inside this CTA's per-step stream, 16 B units
  uint16_t bytes16;    // size, 16 B units
  uint16_t nmma;       // K of the chunk / 16
  uint16_t lbo16;      // rows of the tile rounded up to 8 (= byte stride between 8-wide k chunks / 16)
  uint16_t b_k8;       // first 8-wide k chunk of the activation operand this chunk multiplies
  uint16_t k0;         // first k of the chunk inside the segment
  uint8_t tile;        // 128-row tile of the segment
  uint8_t seg;
};
struct RowBlk { int base, size; };     // local rows [start, start + size) -> weight rows base + cta * size + (j - start)
struct Seg {
  int rows, ntiles, K, chunk_begin, chunk_end, tmem_col, fresh, m64;   // m64: tiles of <= 64 rows, issued as M = 64 MMAs
  int tile_rows;                       // 128, or 64 (an 8-row batch tile needs M = 64: M = 128 requires N % 16 == 0)
  int nblk; RowBlk blk[3]; int col0;   // packing: which rows / columns of the source matrix
};
struct Plan {
  int nseg, nchunks;
  uint32_t step_bytes;                 // per CTA and step
  uint32_t slot_bytes;                 // largest chunk
  Seg seg[MAX_SEG];
  Chunk c[MAX_CHUNKS];
};

#ifdef __CUDACC__
// ---- PTX wrappers -----------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ uint32_t elect_one() {
  uint32_t pred = 0;
  asm volatile("{\n.reg .b32 rx;\n.reg .pred px;\nelect.sync rx|px, %1;\n@px mov.s32 %0, 1;\n}\n" : "+r"(pred) : "r"(0xFFFFFFFFu));
  return pred;
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ bool mbar_try(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile("{\n.reg .pred P1;\nmbarrier.try_wait.parity.shared::cta.b64 P1, [%1], %2;\nselp.u32 %0, 1, 0, P1;\n}\n"
               : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
  return ok != 0;
}
// Bounded wait: a protocol error must surface as a trapped kernel (an error code on the host), never as a hung GPU.
static __device__ __noinline__ void mbar_wait_slow(uint64_t* bar, uint32_t parity) {
  const long long t0 = clock64();
  while (!mbar_try(bar, parity)) {
    if (clock64() - t0 > 4000000000ll) __trap();
  }
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  if (mbar_try(bar, parity)) return;
  mbar_wait_slow(bar, parity);
}
// the same wait without a function call: used by the single-role warps, whose loops must stay free of calls so that ptxas keeps their
// descriptor arithmetic in uniform registers (a call inside the loop pushed it into vector registers + R2UR: ~90 clk per MMA issued)
__device__ __forceinline__ void mbar_wait_inline(uint64_t* bar, uint32_t parity) {
  uint32_t spins = 0;
  while (!mbar_try(bar, parity)) {
    if (++spins > (1u << 24)) __trap();
  }
}
__device__ __forceinline__ uint32_t mapa(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void bulk_g2s(uint32_t smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_dst), "l"(gsrc), "r"(bytes),
               "r"(smem_u32(bar)) : "memory");
}
// this CTA's shared memory -> a peer's shared memory; completes `bytes` on the PEER's mbarrier
__device__ __forceinline__ void bulk_s2peer(uint32_t dst_cluster_addr, uint32_t src_cta_addr, uint32_t bytes, uint32_t bar_cluster_addr) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_cluster_addr),
               "r"(src_cta_addr), "r"(bytes), "r"(bar_cluster_addr) : "memory");
}
__device__ __forceinline__ void st_async_f2(uint32_t dst_cluster_addr, float x, float y, uint32_t bar_cluster_addr) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.v2.f32 [%0], {%1, %2}, [%3];" ::"r"(dst_cluster_addr), "f"(x), "f"(y),
               "r"(bar_cluster_addr) : "memory");
}
__device__ __forceinline__ void st_async_b32(uint32_t dst_cluster_addr, uint32_t v, uint32_t bar_cluster_addr) {
  asm volatile("st.async.weak.shared::cluster.mbarrier::complete_tx::bytes.b32 [%0], %1, [%2];" ::"r"(dst_cluster_addr), "r"(v), "r"(bar_cluster_addr)
               : "memory");
}
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void cluster_sync_all() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 1, %0;" ::"n"(NT_EPI) : "memory"); }
__device__ __forceinline__ void cp_async4(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// SWIZZLE_NONE K-major shared-memory descriptor: element (row, k) of an operand lives at
//   (k / 8) * lbo + (row / 8) * 128 + (row % 8) * 16 + (k % 8) * 2     (lbo = stride between 8-wide k chunks, sbo = 128)
__device__ __forceinline__ uint64_t desc_ns(uint32_t addr, uint32_t lbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((addr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)(128 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
// instruction descriptor: fp32 accumulate, bf16 x bf16, both K-major, N = NB, M = 128 or 64
__device__ __forceinline__ uint32_t idesc_n16(int M, int nb = NB) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(nb >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void umma_n16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
  asm volatile("{\n.reg .pred p;\nsetp.ne.b32 p, %4, 0;\ntcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n}\n" ::"r"(tmem_d), "l"(adesc),
               "l"(bdesc), "r"(idesc), "r"(accum) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld8(uint32_t taddr, float* v) {
  uint32_t r[8];
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0, %1, %2, %3, %4, %5, %6, %7}, [%8];\n"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]) : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
  for (int j = 0; j < 8; ++j) v[j] = __uint_as_float(r[j]);
}

// ---- the three warp roles' building blocks ----------------------------------------------------------------------------------------
struct Ring {
  uint32_t base;            // shared address of slot 0
  uint32_t slot_bytes;
  int nslots;
  uint64_t* full;
  uint64_t* empty;
};
struct RingPos { uint32_t slot, phase; };
__device__ __forceinline__ void ring_advance(RingPos& p, int nslots) { if (++p.slot == (uint32_t)nslots) { p.slot = 0; p.phase ^= 1u; } }

// producer warp: all chunks of one segment (whole warp runs this; one elected lane issues)
__device__ __forceinline__ void produce_seg(const Plan& pl, int si, const uint8_t* stream_cta, const Ring& rg, RingPos& pos) {
  const int cb = pl.seg[si].chunk_begin, ce = pl.seg[si].chunk_end;
#pragma unroll 1
  for (int ci = cb; ci < ce; ++ci) {
    const uint32_t off = pl.c[ci].off16, bytes = (uint32_t)pl.c[ci].bytes16 << 4;
    mbar_wait_inline(&rg.empty[pos.slot], pos.phase ^ 1u);
    if (elect_one()) {
      mbar_expect_tx(&rg.full[pos.slot], bytes);
      bulk_g2s(rg.base + pos.slot * rg.slot_bytes, stream_cta + ((size_t)off << 4), bytes, &rg.full[pos.slot]);
    }
    __syncwarp();
    ring_advance(pos, rg.nslots);
  }
}
// MMA warp: all chunks of one segment against the activation operand at shared address `bbuf`
// nb: batch rows per cluster (N of the MMA): 16, or 8 for shapes whose operand buffers do not fit otherwise (reverse scan of large D)
__device__ __forceinline__ void mma_seg(const Plan& pl, int si, uint32_t bbuf, uint32_t tmem_base, const Ring& rg, RingPos& pos, int nb = NB) {
  const int cb = pl.seg[si].chunk_begin, ce = pl.seg[si].chunk_end;
  const uint32_t col = (uint32_t)pl.seg[si].tmem_col, fresh = (uint32_t)pl.seg[si].fresh;
  const uint32_t idesc = idesc_n16(pl.seg[si].m64 ? 64 : 128, nb);
  const uint32_t brow = (uint32_t)nb * 16;
#pragma unroll 1
  for (int ci = cb; ci < ce; ++ci) {
    const uint32_t nmma = pl.c[ci].nmma, lbo16 = pl.c[ci].lbo16, bk8 = pl.c[ci].b_k8, tile = pl.c[ci].tile, k0 = pl.c[ci].k0;
    mbar_wait_inline(&rg.full[pos.slot], pos.phase);
    tc_fence_after();
    const uint64_t ad0 = desc_ns(rg.base + pos.slot * rg.slot_bytes, lbo16 << 4);
    const uint64_t bd0 = desc_ns(bbuf + bk8 * brow, brow);
    const uint32_t dcol = tmem_base + col + tile * NB;
    const uint32_t astep = lbo16 * 2, bstep = (2 * brow) >> 4;       // one MMA = K 16 = two 8-wide k chunks
    const uint32_t acc0 = (fresh && k0 == 0) ? 0u : 1u;
    if (elect_one()) {                 // one election per chunk: its MMAs and the slot's commit are issued by the same lane
#pragma unroll 4
      for (uint32_t j = 0; j < nmma; ++j) umma_n16(dcol, ad0 + (uint64_t)(j * astep), bd0 + (uint64_t)(j * bstep), idesc, j == 0 ? acc0 : 1u);
      umma_commit(&rg.empty[pos.slot]);
    }
    __syncwarp();
    ring_advance(pos, rg.nslots);
  }
}
// epilogue warps: accumulator of a segment (rows x NB) -> outb[row * OB + m].  ew = epilogue warp 0..7 (hardware warp ew + 2: TMEM
// lane quadrant (ew + 2) % 4, column half ew / 4).
// An M = 64 accumulator keeps row r in TMEM lane 32 (r / 16) + r % 16 (profiles/r02a_probe_umma.md).
// `add` (optional): add[m * add_ld + row] is added on the way (a hoisted, batched part of the same pre-activation).
// nb = 8: only the first column half exists; ob = row stride of outb (nb + 1).
__device__ __forceinline__ void tmem_to_outb(uint32_t tmem_base, int col, int rows, int m64, float* outb, const float* add, int add_ld, int ew, int lane,
                                             int nb = NB) {
  const int q = (ew + 2) & 3, half = ew >> 2, ob = nb + 1;
  if (8 * half >= nb) { tc_fence_before(); return; }
  if (m64) {                                   // tiles of 64 rows: row 64 t + 16 q + lane (lane < 16) sits in TMEM lane 32 q + lane
#pragma unroll 1
    for (int t0 = 16 * q; t0 < rows; t0 += 64) {
      float v[8];
      tmem_ld8(tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)(col + (t0 >> 6) * NB + 8 * half), v);
      const int r = t0 + lane;
      if (lane < 16 && r < rows) {
#pragma unroll
        for (int j = 0; j < 8; ++j) outb[r * ob + 8 * half + j] = v[j] + (add ? add[(8 * half + j) * add_ld + r] : 0.f);
      }
    }
    tc_fence_before();
    return;
  }
#pragma unroll 1
  for (int r0 = 32 * q; r0 < rows; r0 += 128) {
    float v[8];
    tmem_ld8(tmem_base + ((uint32_t)(32 * q) << 16) + (uint32_t)(col + (r0 >> 7) * NB + 8 * half), v);
    const int r = r0 + lane;
    if (r < rows) {
#pragma unroll
      for (int j = 0; j < 8; ++j) outb[r * ob + 8 * half + j] = v[j] + (add ? add[(8 * half + j) * add_ld + r] : 0.f);
    }
  }
  tc_fence_before();
}

// LayerNorm statistics of a vector spread over the cluster: own values vals[j * OB + m], j < n.  (mean, M2) of every batch row goes
// to lnp[c][m] of every peer as one st.async per lane; the receiver's barrier `bar` expects CL * NB * 8 bytes.
__device__ __forceinline__ void stats_push(const float* vals, int n, float* lnp, uint64_t* bar, int c, int CL, int ew, int lane) {
  const int hl = lane >> 4, l16 = lane & 15, m = 2 * ew + hl;
  float s = 0.f;
  for (int j = l16; j < n; j += 16) s += vals[j * OB + m];
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mu = s / (float)n;
  float q = 0.f;
  for (int j = l16; j < n; j += 16) { const float d = vals[j * OB + m] - mu; q = fmaf(d, d, q); }
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  if (l16 < CL) st_async_f2(mapa(smem_u32(lnp + (c * NB + m) * 2), l16), mu, q, mapa(smem_u32(bar), l16));
}
// after the barrier: Chan's combination of the CL partials (equal counts) -> s_stat[m] = mean, s_stat[NB + m] = rstd
__device__ __forceinline__ void stats_combine(const float* lnp, int CL, int n_per_cta, float eps, float* s_stat, float* gmean, float* grstd,
                                              int n_store, int ew, int lane) {
  const int hl = lane >> 4, l16 = lane & 15, m = 2 * ew + hl;
  const float2 pc = (l16 < CL) ? reinterpret_cast<const float2*>(lnp)[l16 * NB + m] : make_float2(0.f, 0.f);
  float s = pc.x;
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  const float mu = s / (float)CL;
  const float dm = pc.x - mu;
  float q = (l16 < CL) ? (pc.y + (float)n_per_cta * dm * dm) : 0.f;
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) q += __shfl_xor_sync(0xffffffffu, q, o);
  const float rs = rsqrtf(q / (float)(n_per_cta * CL) + eps);
  if (l16 == 0) {
    s_stat[m] = mu;
    s_stat[NB + m] = rs;
    if (m < n_store) { gmean[m] = mu; grstd[m] = rs; }
  }
}
// plain sums (LayerNorm backward): s1 = sum a, s2 = sum a * b over the own n values of every batch row
__device__ __forceinline__ void sum2_push(const float* va, const float* vb, int n, float* lnp, uint64_t* bar, int c, int CL, int ew, int lane) {
  const int hl = lane >> 4, l16 = lane & 15, m = 2 * ew + hl;
  float s1 = 0.f, s2 = 0.f;
  for (int j = l16; j < n; j += 16) { const float g = va[j * OB + m]; s1 += g; s2 = fmaf(g, vb[j * OB + m], s2); }
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) { s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
  if (l16 < CL) st_async_f2(mapa(smem_u32(lnp + (c * NB + m) * 2), l16), s1, s2, mapa(smem_u32(bar), l16));
}
__device__ __forceinline__ void sum2_combine(const float* lnp, int CL, int n_total, float* s_stat, int ew, int lane) {
  const int hl = lane >> 4, l16 = lane & 15, m = 2 * ew + hl;
  const float2 pc = (l16 < CL) ? reinterpret_cast<const float2*>(lnp)[l16 * NB + m] : make_float2(0.f, 0.f);
  float s1 = pc.x, s2 = pc.y;
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) { s1 += __shfl_xor_sync(0xffffffffu, s1, o); s2 += __shfl_xor_sync(0xffffffffu, s2, o); }
  if (l16 == 0) { s_stat[m] = s1 / (float)n_total; s_stat[NB + m] = s2 / (float)n_total; }
}
// push this CTA's activation slice (already in operand layout in `stage`, `bytes` long) into slot c of every peer's operand buffer
__device__ __forceinline__ void push_act(uint32_t stage_addr, uint32_t bytes, uint32_t buf_addr, uint64_t* bar, int c, int CL, int ew) {   // (bytes already reflect the batch tile)
#pragma unroll 1
  for (int p = ew; p < CL; p += EW) {
    if (elect_one()) bulk_s2peer(mapa(buf_addr + (uint32_t)c * bytes, (uint32_t)p), stage_addr, bytes, mapa(smem_u32(bar), (uint32_t)p));
    __syncwarp();
  }
}
// offset (bytes) of element (m, k) inside an activation operand / slice that starts at k = 0
__device__ __forceinline__ uint32_t b_off(int m, int k, int brow = BROW) { return (uint32_t)((k >> 3) * brow + (m >> 3) * 128 + (m & 7) * 16 + (k & 7) * 2); }
#endif  // __CUDACC__

// ---- host: plan of the weight stream ----------------------------------------------------------------------------------------------
inline int plan_add_seg(Plan& pl, int slot_max_bytes, int rows, int K, int tmem_col, int fresh, int b_k8_base, int nblk, const RowBlk* blk, int col0,
                        uint32_t& off16, int tile_rows = 128) {
  if (pl.nseg >= MAX_SEG || rows <= 0 || (K & 15)) return -1;
  Seg& s = pl.seg[pl.nseg];
  s.rows = rows; s.K = K; s.tile_rows = tile_rows; s.ntiles = (rows + tile_rows - 1) / tile_rows; s.tmem_col = tmem_col; s.fresh = fresh;
  s.m64 = (rows <= 64 || tile_rows == 64) ? 1 : 0;
  s.nblk = nblk; for (int i = 0; i < nblk; ++i) s.blk[i] = blk[i];
  s.col0 = col0;
  s.chunk_begin = pl.nchunks;
  // chunk depth: as deep as a slot allows for the widest tile (fewer, larger copies: ~160 clk of producer time per copy)
  const int rows_pad_max = ((rows < tile_rows ? rows : tile_rows) + 7) & ~7;
  int KC = 256;
  while (KC > 16 && rows_pad_max * KC * 2 > slot_max_bytes) KC >>= 1;
  for (int k0 = 0; k0 < K; k0 += KC) {
    const int kc = (K - k0 < KC) ? (K - k0) : KC;
    for (int t = 0; t < s.ntiles; ++t) {
      if (pl.nchunks >= MAX_CHUNKS) return -1;
      const int trows = (rows - t * tile_rows < tile_rows) ? rows - t * tile_rows : tile_rows;
      const int rp = (trows + 7) & ~7;
      Chunk& c = pl.c[pl.nchunks++];
      c.off16 = off16; c.bytes16 = (uint16_t)(rp * kc * 2 / 16); c.nmma = (uint16_t)(kc / 16); c.lbo16 = (uint16_t)rp;
      c.b_k8 = (uint16_t)(b_k8_base + k0 / 8); c.k0 = (uint16_t)k0; c.tile = (uint8_t)t; c.seg = (uint8_t)pl.nseg;
      off16 += c.bytes16;
      const uint32_t bytes = (uint32_t)c.bytes16 * 16;
      if (bytes > pl.slot_bytes) pl.slot_bytes = bytes;
    }
  }
  s.chunk_end = pl.nchunks;
  return pl.nseg++;
}

}  // namespace tcs
