// Fused global-norm gradient clipping + Adam on a flat fp32 parameter arena, and the target-critic EMA.
// Reference: fabric.clip_gradients == torch.nn.utils.clip_grad_norm_ (L2; coef = max_norm/(norm+1e-6) clamped to 1)
// followed by torch.optim.Adam.step (betas .9/.999, no weight decay, bias-corrected), at
// sheeprl/algos/dreamer_v3/dreamer_v3.py:179-187, :318-323, :339-344.
// HBM-bound: reads p, g, m, v and writes p, m, v (28 B/param) + one 4 B/param pass for the norm.
#include "common.cuh"

namespace {

constexpr int NORM_BLOCKS_MAX = 1024;

__global__ void __launch_bounds__(256) sumsq_partial_kernel(size_t n, const float* __restrict__ g, float* __restrict__ partial) {
  __shared__ float red[64];
  float s = 0.f;
  const size_t n4 = n / 4;
  const float4* g4 = reinterpret_cast<const float4*>(g);
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
    const float4 v = g4[i];
    s = fmaf(v.x, v.x, s); s = fmaf(v.y, v.y, s); s = fmaf(v.z, v.z, s); s = fmaf(v.w, v.w, s);
  }
  for (size_t i = n4 * 4 + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    s = fmaf(g[i], g[i], s);
  s = block_sum(s, red);
  if (threadIdx.x == 0) partial[blockIdx.x] = s;
}

// ctl[0] = clip coef, ctl[1] = lr / (1 - beta1^t), ctl[2] = sqrt(1 - beta2^t); step += 1; norm_out[slot] = norm
__global__ void __launch_bounds__(256) adam_prepare_kernel(int nblocks, const float* __restrict__ partial, float* __restrict__ step,
                                                           double lr, double beta1, double beta2, float max_norm,
                                                           float* __restrict__ norm_out, int slot, float* __restrict__ ctl) {
  __shared__ double red[256];
  double s = 0.0;
  for (int i = threadIdx.x; i < nblocks; i += blockDim.x) s += (double)partial[i];
  red[threadIdx.x] = s;
  __syncthreads();
  for (int o = 128; o > 0; o >>= 1) {
    if (threadIdx.x < o) red[threadIdx.x] += red[threadIdx.x + o];
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    const float norm = (float)sqrt(red[0]);
    if (norm_out) norm_out[slot] = norm;
    float coef = max_norm / (norm + 1e-6f);
    ctl[0] = coef > 1.f ? 1.f : coef;
    const float t = step[0] + 1.f;
    step[0] = t;
    const double bc1 = 1.0 - pow(beta1, (double)t);
    const double bc2 = 1.0 - pow(beta2, (double)t);
    ctl[1] = (float)(lr / bc1);
    ctl[2] = (float)sqrt(bc2);
  }
}

__global__ void __launch_bounds__(256) adam_apply_kernel(size_t n, float* __restrict__ p, const float* __restrict__ g,
                                                         float* __restrict__ m, float* __restrict__ v, float omb1, float beta2,
                                                         float omb2, float eps, const float* __restrict__ ctl) {
  const float coef = ctl[0], step_size = ctl[1], bc2s = ctl[2];
  for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const float gg = g[i] * coef;
    const float mm = m[i] + (gg - m[i]) * omb1;                   // exp_avg.lerp_(grad, 1-beta1)
    const float vv = v[i] * beta2 + omb2 * gg * gg;               // exp_avg_sq.mul_(b2).addcmul_(g, g, 1-b2)
    m[i] = mm;
    v[i] = vv;
    const float denom = sqrtf(vv) / bc2s + eps;
    p[i] = p[i] - step_size * (mm / denom);
  }
}

}  // namespace

extern "C" {

// scratch: >= (NORM_BLOCKS_MAX + 4) floats
size_t dv3_adam_clip_scratch_bytes(void) { return (NORM_BLOCKS_MAX + 4) * sizeof(float); }

// lr / betas / eps are doubles because torch derives 1-beta and the bias corrections in Python double precision
// (1 - 0.999 rounded to fp32 differs from fp32(1) - fp32(0.999) by 1.3e-5 relative).
static int adam_clip_impl(size_t n, float* p, const float* g, float* m, float* v, float* step, double lr, double beta1,
                          double beta2, double eps, double max_norm, float* norm_out, int slot, float* scratch,
                          cudaStream_t stream) {
  int nb = (int)((n / 4 + 255) / 256);
  if (nb < 1) nb = 1;
  if (nb > NORM_BLOCKS_MAX) nb = NORM_BLOCKS_MAX;
  float* ctl = scratch + NORM_BLOCKS_MAX;
  sumsq_partial_kernel<<<nb, 256, 0, stream>>>(n, g, scratch);
  DV3_LAUNCH_CHECK();
  adam_prepare_kernel<<<1, 256, 0, stream>>>(nb, scratch, step, lr, beta1, beta2, (float)max_norm, norm_out, slot, ctl);
  DV3_LAUNCH_CHECK();
  size_t gb = (n + 255) / 256;
  if (gb > 148 * 8) gb = 148 * 8;
  adam_apply_kernel<<<(int)gb, 256, 0, stream>>>(n, p, g, m, v, (float)(1.0 - beta1), (float)beta2, (float)(1.0 - beta2), (float)eps, ctl);
  DV3_LAUNCH_CHECK();
  return DV3_OK;
}

int dv3_adam_clip(size_t n, float* p, const float* g, float* m, float* v, float* step, double lr, double beta1, double beta2,
                  double eps, double max_norm, float* norm_out, int slot, float* scratch, size_t scratch_bytes, void* stream) {
  if (n == 0 || !p || !g || !m || !v || !step || !scratch) return DV3_ERR_ARG;
  if (scratch_bytes < dv3_adam_clip_scratch_bytes()) return DV3_ERR_SCRATCH;
  if (((uintptr_t)g & 15) != 0) return DV3_ERR_ARG;
  return adam_clip_impl(n, p, g, m, v, step, lr, beta1, beta2, eps, max_norm, norm_out, slot, scratch, (cudaStream_t)stream);
}

}  // extern "C"
