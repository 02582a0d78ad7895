// Device-resident replay: ring-buffer write and the sequence-gather kernel.
// Reference: ReplayBuffer.add sheeprl/data/buffers.py:99-151 (ring write, wrap-around) and
// SequentialReplayBuffer.sample/_get_samples :238-339 + AsyncReplayBuffer.sample :646-690, which the reference
// executes as CPU fancy-indexing followed by a host->device copy (dreamer_v3.py:703-709).
// Index selection (which env, which start) stays on the host with the reference's exact RNG call order; this file
// is pure byte movement and is bit-exact by construction.
// Layout: per key one allocation [n_envs][size][row_bytes]; output [n_samples][T][B][row_bytes] (time-major).
#include "common.cuh"

namespace {

constexpr int MAX_KEYS = 8;
struct GatherKeys {
  int n;
  const unsigned char* src[MAX_KEYS];
  unsigned char* dst[MAX_KEYS];
  long long row_bytes[MAX_KEYS];
};

// one warp per (key, sequence j, step t) row copy
template <typename V>
static __device__ __forceinline__ void copy_row(const unsigned char* s, unsigned char* d, long long bytes, int lane) {
  const long long nv = bytes / (long long)sizeof(V);
  const V* sv = reinterpret_cast<const V*>(s);
  V* dv = reinterpret_cast<V*>(d);
  for (long long i = lane; i < nv; i += 32) dv[i] = sv[i];
}

__global__ void __launch_bounds__(256)
seq_gather_kernel(GatherKeys keys, int size, int T, int Bt, int n_samples, const int* __restrict__ env_id,
                  const long long* __restrict__ start, const int* __restrict__ col, const int* __restrict__ samp) {
  // sequence j: env_id[j], start[j], output column col[j] (batch index), sample index samp[j]
  const int lane = threadIdx.x & 31;
  const long long wid = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const long long nseq = (long long)Bt * n_samples;
  const long long rows = nseq * T;
  const int key = blockIdx.y;
  if (wid >= rows) return;
  const int j = (int)(wid / T), t = (int)(wid % T);
  const long long rb = keys.row_bytes[key];
  const long long pos = (start[j] + t) % size;
  const unsigned char* s = keys.src[key] + ((long long)env_id[j] * size + pos) * rb;
  unsigned char* d = keys.dst[key] + (((long long)samp[j] * T + t) * Bt + col[j]) * rb;
  const uintptr_t al = (uintptr_t)s | (uintptr_t)d | (uintptr_t)rb;
  if ((al & 15) == 0) copy_row<uint4>(s, d, rb, lane);
  else if ((al & 7) == 0) copy_row<uint2>(s, d, rb, lane);
  else if ((al & 3) == 0) copy_row<unsigned int>(s, d, rb, lane);
  else copy_row<unsigned char>(s, d, rb, lane);
}

// rows l = 0..L-1 of src ([L][row_bytes]) -> ring[env][(pos + l) % size]
__global__ void __launch_bounds__(256)
buffer_add_kernel(const unsigned char* __restrict__ src, unsigned char* __restrict__ ring, int env, long long pos, int L,
                  int size, long long rb) {
  const int lane = threadIdx.x & 31;
  const long long l = (long long)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (l >= L) return;
  const unsigned char* s = src + l * rb;
  unsigned char* d = ring + ((long long)env * size + (pos + l) % size) * rb;
  const uintptr_t al = (uintptr_t)s | (uintptr_t)d | (uintptr_t)rb;
  if ((al & 15) == 0) copy_row<uint4>(s, d, rb, lane);
  else if ((al & 3) == 0) copy_row<unsigned int>(s, d, rb, lane);
  else copy_row<unsigned char>(s, d, rb, lane);
}

}  // namespace

extern "C" {

int dv3_seq_gather(int n_keys, const void* const* src, void* const* dst, const long long* row_bytes, int size, int T, int B,
                   int n_samples, const int* env_id, const long long* start, const int* col, const int* samp, void* stream) {
  if (n_keys <= 0 || n_keys > MAX_KEYS || size <= 0 || T <= 0 || B <= 0 || n_samples <= 0) return DV3_ERR_ARG;
  GatherKeys k;
  k.n = n_keys;
  for (int i = 0; i < n_keys; ++i) {
    if (!src[i] || !dst[i] || row_bytes[i] <= 0) return DV3_ERR_ARG;
    k.src[i] = (const unsigned char*)src[i];
    k.dst[i] = (unsigned char*)dst[i];
    k.row_bytes[i] = row_bytes[i];
  }
  const long long rows = (long long)B * n_samples * T;
  dim3 grid((unsigned)((rows + 7) / 8), n_keys);
  seq_gather_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(k, size, T, B, n_samples, env_id, start, col, samp);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}

int dv3_buffer_add(const void* src, void* ring, int env, long long pos, int L, int size, long long row_bytes, void* stream) {
  if (!src || !ring || L <= 0 || size <= 0 || row_bytes <= 0 || pos < 0 || env < 0) return DV3_ERR_ARG;
  buffer_add_kernel<<<(L + 7) / 8, 256, 0, (cudaStream_t)stream>>>((const unsigned char*)src, (unsigned char*)ring, env, pos, L, size, row_bytes);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}

int dv3_version(void) { return 100; }

}  // extern "C"
