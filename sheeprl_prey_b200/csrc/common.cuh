// Shared device helpers for libdv3b200 (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <math.h>

#define DV3_OK 0
#define DV3_ERR_ARG (-1)
#define DV3_ERR_UNSUPPORTED (-2)
#define DV3_ERR_SCRATCH (-3)
#define DV3_ERR_TOO_LARGE (-4)

#define DV3_LAUNCH_CHECK()                                    \
  do {                                                        \
    cudaError_t e__ = cudaGetLastError();                     \
    if (e__ != cudaSuccess) return (int)e__;                  \
  } while (0)

// ---- programmatic dependent launch (PDL) ---------------------------------------------------------------------------------
// A kernel launched through dv3_launch_pdl may be scheduled while its predecessor on the stream is still draining; it must call
// pdl_sync() before its first access to global memory (griddepcontrol.wait = predecessor complete and flushed).  The prologue
// before that point (index math, barrier / TMEM set-up, descriptor prefetch) overlaps the predecessor's tail and the launch
// latency; inside the step's CUDA graph the attribute becomes a programmatic dependency edge.  Off unless $DV3_PDL=1 (no gain measured).
// In a kernel launched the ordinary way both instructions are no-ops.
static __device__ __forceinline__ void pdl_sync() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
#ifndef __CUDA_ARCH__
#include <stdlib.h>
static inline int dv3_pdl_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("DV3_PDL"); v = (e && e[0] == '1') ? 1 : 0; }   // opt-in: measured 120.6 (on) vs 122.2 (off) steps/s
  return v;
}
#endif
// One-time per-DEVICE set-up (cudaFuncSetAttribute applies to the current device only): true the first time it is called with a given
// device current.  `mask` is a function-local static, one bit per device ordinal.
static inline bool dv3_first_call_on_device(unsigned long long* mask) {
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return true;
  const unsigned long long bit = 1ull << dev;
  if (*mask & bit) return false;
  *mask |= bit;
  return true;
}
template <typename... KA, typename... A>
static inline cudaError_t dv3_launch_pdl(void (*kern)(KA...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, A... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
#ifndef __CUDA_ARCH__
  at[0].val.programmaticStreamSerializationAllowed = dv3_pdl_enabled();
#endif
  cfg.attrs = at;
  cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, KA(args)...);
}

static __device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
static __device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
static __device__ __forceinline__ int warp_sum_i(int v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// Block-wide sum of one float; `red` is >= 32 floats of shared memory. All threads get the result.
static __device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();  // protect `red` from a previous use
  if (lane == 0) red[w] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : 0.f;
  r = warp_sum(r);
  return r;
}
// Two sums at once.
static __device__ __forceinline__ void block_sum2(float& a, float& b, float* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  a = warp_sum(a);
  b = warp_sum(b);
  __syncthreads();
  if (lane == 0) { red[w] = a; red[32 + w] = b; }
  __syncthreads();
  float ra = (lane < nw) ? red[lane] : 0.f;
  float rb = (lane < nw) ? red[32 + lane] : 0.f;
  a = warp_sum(ra);
  b = warp_sum(rb);
}

static __device__ __forceinline__ float sigmoidf_(float x) { return 1.f / (1.f + expf(-x)); }
static __device__ __forceinline__ float siluf_(float x) { return x * sigmoidf_(x); }
static __device__ __forceinline__ float symlogf_(float x) {
  float s = (x > 0.f) ? 1.f : ((x < 0.f) ? -1.f : 0.f);
  return s * logf(1.f + fabsf(x));
}
static __device__ __forceinline__ float symexpf_(float x) {
  float s = (x > 0.f) ? 1.f : ((x < 0.f) ? -1.f : 0.f);
  return s * (expf(fabsf(x)) - 1.f);
}
#define DV3_EPS32 1.1920928955078125e-07f
