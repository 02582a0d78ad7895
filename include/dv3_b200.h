/* libdv3b200 -- C ABI of the B200 (sm_100a) kernels behind the DreamerV3 training step.
 *
 * The reference (hanshuo-shuo/sheeprl_prey, SheepRL 0.4.4) is 100 % Python and has no FFI for this path; its
 * boundary is the Python-level `train()` / `build_models()` / replay-buffer API.  This
 * header is the C-ABI a maintainer binds from Python with ctypes (INTEGRATION.md shows the stub); each entry
 * point cites the reference lines whose arithmetic it replaces (paths relative to the reference root).
 *
 * Conventions
 *   - every function returns 0 on success, a negative DV3_ERR_* code for bad arguments, or a positive
 *     cudaError_t if the launch failed; nothing throws, nothing allocates, nothing synchronises;
 *   - all pointers are DEVICE pointers owned by the caller (float = fp32, int = int32, long long = int64);
 *   - matrices are row-major; `ld*` is the row stride in elements (the last dimension is contiguous);
 *   - `stream` is a cudaStream_t passed as void*; kernels are enqueued on it and the call returns immediately;
 *   - functions taking `scratch` use it as device workspace (size from the matching *_scratch_bytes());
 *   - no hidden global state: calls on different streams with different buffers are independent.
 */
#ifndef DV3_B200_H
#define DV3_B200_H
#include <stddef.h>

#ifdef __cplusplus
extern "C" {
#endif

#define DV3_OK 0
#define DV3_ERR_ARG (-1)
#define DV3_ERR_UNSUPPORTED (-2)
#define DV3_ERR_SCRATCH (-3)
#define DV3_ERR_TOO_LARGE (-4)

#define DV3_GEMM_NT 0 /* C[M,N] = A[M,K] * B[N,K]^T   (nn.Linear forward)            */
#define DV3_GEMM_NN 1 /* C[M,N] = A[M,K] * B[K,N]     (input gradient dX = dY * W)    */
#define DV3_GEMM_TN 2 /* C[M,N] = A[K,M]^T * B[K,N]   (weight gradient dW = dY^T * X) */

int dv3_version(void);

/* ---- replay (sheeprl/data/buffers.py) ------------------------------------------------------------------- */
/* ReplayBuffer.add, buffers.py:99-151: rows l=0..L-1 of `src` ([L][row_bytes]) go to ring[env][(pos+l) % size]. */
int dv3_buffer_add(const void* src, void* ring, int env, long long pos, int L, int size, long long row_bytes, void* stream);
/* SequentialReplayBuffer.sample/_get_samples buffers.py:238-339 + AsyncReplayBuffer.sample :646-690 (the byte
 * movement; index selection stays on the host with the reference's RNG call order).  For every key k:
 * dst[k][samp[j]][t][col[j]][:] = src[k][env_id[j]][(start[j]+t) % size][:], j < B*n_samples, t < T. Bit-exact. */
int dv3_seq_gather(int n_keys, const void* const* src, void* const* dst, const long long* row_bytes, int size, int T, int B,
                   int n_samples, const int* env_id, const long long* start, const int* col, const int* samp, void* stream);

/* ---- dense contractions: every torch addmm/mm of the step (models.py:115-118, :389 and their backward) ---- */
int dv3_gemm_f32(int mode, int M, int N, int K, const float* A, int lda, const float* B, int ldb, float* C, int ldc,
                 const float* bias, int accumulate, void* stream);

/* Same contract on the tcgen05 tensor cores: operands are rounded to bf16 while being staged into shared memory,
 * products accumulate in fp32 in TMEM (north_star: "2e-2 for bf16 tensor-core GEMMs"). */
int dv3_gemm_bf16_tc(int mode, int M, int N, int K, const float* A, int lda, const float* B, int ldb, float* C, int ldc,
                     const float* bias, int accumulate, void* stream);

/* bf16-resident variant (same modes): A and B are bf16 in HBM (leading dimensions multiples of 8, 16 B aligned bases, else
 * DV3_ERR_UNSUPPORTED); operand tiles are fetched by TMA (cp.async.bulk.tensor, SWIZZLE_128B) straight into the UMMA layout
 * (K-major or MN-major, no transposition), C is fp32; split-K with fp32 atomics when the output has few tiles.
 * Producers below take an optional `*_bf16` pointer: a bf16 mirror of their fp32 output with the SAME leading dimension,
 * written in the same pass, so the next contraction reads half the bytes (NULL = no mirror). */
int dv3_gemm_bf16_tma(int mode, int M, int N, int K, const void* A_bf16, int lda, const void* B_bf16, int ldb, float* C, int ldc,
                      const float* bias, int accumulate, void* stream);
/* Linear + LayerNorm + SiLU (the reference's miniblock, utils/model.py:71-87) in ONE kernel: y = act(LN(A . B^T + bias)).  The
 * N / 128 column tiles of a 128-row block run as a thread-block cluster and exchange per-row (mean, M2) partials through distributed
 * shared memory; pre (optional) receives the pre-LN values the backward needs, y_bf16 (optional) the bf16 mirror of y.
 * N % 128 == 0 and N <= 1024, bf16 K-major operands as for dv3_gemm_bf16_tma, else DV3_ERR_UNSUPPORTED. */
int dv3_gemm_ln_bf16_tma(int M, int N, int K, const void* A_bf16, int lda, const void* B_bf16, int ldb, const float* bias, const float* gamma,
                         const float* beta, float eps, int act, float* pre, int ldp, float* y, int ldy, void* y_bf16, float* mean, float* rstd,
                         void* stream);
int dv3_f32_to_bf16(int R, int C, const float* src, int lds, void* dst_bf16, int ldd, void* stream);

/* ---- LayerNorm(+SiLU) blocks: miniblock order Linear -> LayerNorm -> SiLU, utils/model.py:71-87 ---------- */
int dv3_ln_act_fwd(int M, int U, const float* x, int ldx, const float* gamma, const float* beta, float eps, int act,
                   float* y, int ldy, float* mean, float* rstd, void* y_bf16, void* stream);
int dv3_ln_act_bwd(int M, int U, const float* dy, int lddy, const float* x, int ldx, const float* gamma, const float* beta,
                   const float* mean, const float* rstd, int act, float* dx, int lddx, float* dln, int lddln, void* dx_bf16,
                   void* stream);
/* out_a[u] = sum_m d[m,u]*xhat[m,u] (LN weight grad; skipped when x == NULL), out_b[u] = sum_m d[m,u] (LN bias /
 * Linear bias grad).  Two-stage reduction through `scratch`. */
size_t dv3_col_reduce_scratch_bytes(int M, int U);
int dv3_col_reduce(int M, int U, const float* d, int ldd, const float* x, int ldx, const float* mean, const float* rstd,
                   float* out_a, float* out_b, float* scratch, size_t scratch_bytes, void* stream);

/* No counterpart in the reference.  Same reduction as dv3_col_reduce for hosts that cannot give it a workspace (a plugin
 * handed only inputs and outputs, a loop that cannot afford cudaMalloc's device-wide synchronisation, a runtime that owns
 * all device memory): the partial sums live in a fixed 2 MiB __device__ array inside the library, so nothing is allocated.
 * Limit: U <= 4096 columns (any M); larger U returns DV3_ERR_TOO_LARGE.  The first call on a device resolves the array's
 * address -- issue it before any stream capture.  All calls on a device share the array: calls that can overlap on the
 * GPU race, serialise them. */
int dv3_col_reduce_noalloc(int M, int U, const float* d, int ldd, const float* x, int ldx, const float* mean, const float* rstd,
                           float* out_a, float* out_b, void* stream);

/* ---- pixel encoder / decoder: CNNEncoder agent.py:28-80, CNNDecoder agent.py:136-206, LayerNormChannelLast
 * utils/model.py:225-235.  Activations are channel-last; k = tap*C + c with tap = kh*4 + kw; the contractions run on
 * dv3_gemm_*: Conv2d(k4,s2,p1) = im2col + GEMM, ConvTranspose2d(k4,s2,p1) = GEMM + col2im (and vice versa backward). */
int dv3_im2col_k4s2p1(int N, int H, int W, int C, const float* x, float* col, void* col_bf16, void* stream);
int dv3_col2im_k4s2p1(int N, int Hi, int Wi, int C, const float* col, const float* bias, float add, float* y, void* stream);
int dv3_permute_last2(int A, int B, int C, const float* src, long long lds, float* dst, long long ldd, void* stream);
/* The two narrow-channel layers (image channels C = 1, 3 or 4) without im2col / col2im and without a GEMM (csrc/conv_direct.cu):
 *  - first encoder layer (agent.py:28-80): y[(n,oh,ow)][o] = sum_k col(x)[.][k] * w[o][k], w = [O][16*C] tap-major (ldw), O <= 96;
 *    with gamma != NULL also act = SiLU(LayerNormChannelLast(y)) (+ bf16 copy), mean, rstd.  The same call with the image-shaped
 *    output gradient as x and the re-ordered transposed-convolution weight [C_in][16*C] as w is the data gradient of the last
 *    decoder layer (agent.py:136-206).
 *  - weight gradient of either layer: dw[a][k] = sum_rows g[row][a] * col(x)[row][k], A <= 64 (two-stage, deterministic).
 *  - last decoder layer: y[n][oh][ow][c] = bias[c] + add + sum_{taps,ci} x[(n,ih,iw)][ci] * w[ci][(kh*4+kw)*C + c], w = [Ci][16*C]. */
int dv3_conv_k4s2p1_narrow_fwd(int N, int H, int W, int C, int O, const float* x, const float* w, int ldw, float* y, const float* gamma,
                               const float* beta, float eps, float* act, void* act_bf16, float* mean, float* rstd, void* stream);
size_t dv3_conv_k4s2p1_narrow_wgrad_scratch_bytes(int N, int H, int W, int C, int A);
int dv3_conv_k4s2p1_narrow_wgrad(int N, int H, int W, int C, int A, const float* x, const float* g, float* dw, int lddw, float* scratch,
                                 size_t scratch_bytes, void* stream);
int dv3_deconv_k4s2p1_narrow_fwd(int N, int Hi, int Wi, int Ci, int C, const float* x, const float* w, int ldw, const float* bias, float add,
                                 float* y, void* stream);
int dv3_u8_nchw_to_nhwc(int N, int C, int HW, const unsigned char* src, float scale, float* dst, int Ctot, int c_off, void* stream);
/* warp-per-row variants of dv3_ln_act_fwd / _bwd for narrow rows (channel LayerNorm over N*H*W pixel rows) */
int dv3_ln_act_fwd_rows(int M, int U, const float* x, int ldx, const float* gamma, const float* beta, float eps, int act,
                        float* y, int ldy, float* mean, float* rstd, void* y_bf16, void* stream);
/* dv3_ln_act_fwd_rows on x + sum_h add_table[head_off[h] + add_idx[m, h], :]: the one-hot (discrete action) input block of the Linear
 * in front of the LayerNorm as a row gather of its transposed weights instead of a K = A contraction (rollout, agent.py:296-309).
 * n_heads <= 4, head_off is a HOST array; DV3_ERR_UNSUPPORTED for shapes outside the vector kernel (U % 4, U > 1024, alignment). */
int dv3_ln_act_fwd_rows_gather(int M, int U, const float* x, int ldx, const float* gamma, const float* beta, float eps, int act, float* y, int ldy,
                               float* mean, float* rstd, void* y_bf16, const float* add_table, int ld_table, const int* add_idx, int ld_idx,
                               int n_heads, const int* head_off_host, void* stream);
int dv3_ln_act_bwd_rows(int M, int U, const float* dy, int lddy, const float* x, int ldx, const float* gamma, const float* beta,
                        const float* mean, const float* rstd, int act, float* dx, int lddx, float* dln, int lddln, void* dx_bf16,
                        void* stream);
/* dv3_ln_act_bwd + the two column reductions that follow it in a dense layer's backward, in one pass: dx (+ bf16 mirror),
 * dgamma[u] = sum_m g xhat, dbeta[u] = sum_m g, dbias[u] = sum_m dx (NULL = not wanted).  U % 4 == 0, U <= 1024, 16 B aligned
 * rows, else DV3_ERR_UNSUPPORTED; scratch >= 296 * 3 * U floats. */
int dv3_ln_act_bwd_fused(int M, int U, const float* dy, int lddy, const float* x, int ldx, const float* gamma, const float* beta,
                         const float* mean, const float* rstd, int act, float* dx, int lddx, void* dx_bf16, float* dgamma, float* dbeta,
                         float* dbias, float* scratch, size_t scratch_bytes, void* stream);
/* MSEDistribution.log_prob, utils/distribution.py:212-221: loss[m] = sum_d (mode-x)^2, dmode = scale*2*(mode-x) */
int dv3_mse_rows(int M, int D, const float* mode, const float* target, float scale, float* loss_rows, float* dmode, void* stream);

/* ---- LayerNormGRUCell gates, models.py:388-395 ------------------------------------------------------------ */
int dv3_gru_gate_fwd(int R, int D, const float* gates_pre, int ldg, const float* lnw, const float* lnb, const float* h_in,
                     int ldhi, float* h_out, int ldho, float* mean, float* rstd, void* h_out_bf16, void* stream);
int dv3_gru_gate_bwd(int R, int D, const float* dh_out, int lddho, const float* gates_pre, int ldg, const float* lnw,
                     const float* lnb, const float* mean, const float* rstd, const float* h_in, int ldhi, float* dgates_pre,
                     int lddg, float* dln, int lddln, float* dh_in, int lddhi, void* stream);

/* ---- RSSM.dynamic plumbing, agent.py:379-384 and dreamer_v3.py:108 ---------------------------------------- */
int dv3_scan_prep(int B, int D, int S, const float* first, const float* h_prev, int ldhp, const float* z_prev, int ldzp,
                  const float* zreset, float* h_in, int ldhi, float* z_in, int ldzi, void* stream);
int dv3_scan_carry(int B, int D, int S, const float* first, const float* dz_prev, int lddz, const float* dh_a, int lda,
                   const float* dh_b, int ldb, float* dz_dst, int ldzd, float* dh_dst, int ldhd, void* stream);
int dv3_mask_shift_actions(int T, int B, int A, const float* actions, const float* first, float* out, int ldo, void* stream);

/* ---- persistent RSSM observe scan: the whole loop dreamer_v3.py:122-129 over RSSM.dynamic (agent.py:350-388) in one
 * cooperative kernel (posterior path; the prior head is batched outside).  B <= 16.  Returns DV3_ERR_UNSUPPORTED when the
 * shape does not fit the kernel's plan (caller then uses the per-step entry points above). */
typedef struct Dv3ScanFwdArgs {
  int T, B, G, Kc, D, U, Hh, A, E;
  float unimix;
  const float* Wmlp; int ld_wmlp; const float* g_mlp; const float* b_mlp;     /* RecurrentModel.mlp: [U, G*Kc+A] + LN(1e-3) */
  const float* Wgru; int ld_wgru; const float* g_gru; const float* b_gru;     /* LayerNormGRUCell: [3D, D+U] + LN(1e-5)     */
  const float* Wr0; int ld_wr0; const float* g_r; const float* b_r;           /* representation layer 0: [Hh, D+E] + LN      */
  const float* Wr1; int ld_wr1; const float* br1;                             /* representation head: [G*Kc, Hh] + bias      */
  const float* eW;            /* [T*B, Hh] = e_t . Wr0[:, D:]^T + b_r0 (hoisted out of the scan)                            */
  const float* first;         /* [T*B] is_first, row 0 forced to 1                                                           */
  const float* u_post;        /* [T*B, G] uniforms of the posterior draws                                                    */
  const int* forced_idx;      /* [T*B, G] teacher-forced draws, or NULL                                                      */
  const int* zreset_idx;      /* [G] mode(unimix(transition(0)))                                                             */
  float* x_mlp; int ld_xmlp;  /* [T*B, G*Kc+A]: one-hot z_in written, masked previous action read                           */
  float* feat_pre; float* feat_mean; float* feat_rstd;
  float* x_gru; int ld_xgru;  /* [T*B, D+U] = [h_in, feat]                                                                   */
  float* gates_pre; float* gru_mean; float* gru_rstd;
  float* x_rep; int ld_xrep;  /* [T*B, D+E]: h_t written to the first D columns                                              */
  float* hidr_pre; float* hidr; float* hidr_mean; float* hidr_rstd;
  float* post_raw; float* post_logits; int* post_idx;
  float* latent; int ld_lat;  /* [T*B, G*Kc+D]: one-hot z_t written to the first G*Kc columns                                */
  float* scratch; size_t scratch_bytes;
} Dv3ScanFwdArgs;
size_t dv3_rssm_observe_fwd_scratch_bytes(void);
int dv3_rssm_observe_fwd(const Dv3ScanFwdArgs* args, void* stream);
/* Same scan on ONE thread-block cluster (16 CTAs, else 8): bf16 weight rows streamed from L2 into mma.sync fragments, activations
 * and LayerNorm partials exchanged through distributed shared memory, hardware cluster barriers instead of grid-wide ones.
 * Needs the three in-loop weight matrices packed by dv3_pack_mma_a, Wmlp[:, :G*Kc]^T in fp32 (one-hot z = row gather) and the hoisted
 * action projection aW[t] = a_t . Wmlp[:, G*Kc:]^T.  DV3_ERR_UNSUPPORTED when the shape / device cannot run it. */
typedef struct Dv3ScanClusterExtra {
  const void* Wgru_pk;                  /* dv3_pack_mma_a(3D, D+U, Wgru)        */
  const void* Wr0_pk;                   /* dv3_pack_mma_a(Hh, D, Wr0[:, :D])    */
  const void* Wr1_pk;                   /* dv3_pack_mma_a(G*Kc, Hh, Wr1)        */
  const float* WmlpzT;                  /* [G*Kc, U] fp32 */
  const float* aW;                      /* [T*B, U] fp32 */
} Dv3ScanClusterExtra;
int dv3_rssm_observe_fwd_cluster(const Dv3ScanFwdArgs* args, const Dv3ScanClusterExtra* extra, void* stream);
/* The same scan with tcgen05 tensor cores (rssm_scan_tc.cu): per 16 batch rows one thread-block cluster (16 CTAs, else 8) whose
 * CTAs each own 1/CL of every feature dimension.  Weight rows are the M = 128 operand of tcgen05.mma (batch rows = N = 16, fp32
 * accumulate in TMEM), streamed by cp.async.bulk through an mbarrier ring from a per-CTA, per-step bf16 stream that
 * dv3_scan_tc_fwd_pack lays out in consumption order; activations, LayerNorm partials and sampled indices cross the cluster as
 * bulk copies / st.async that complete on the receiver's mbarrier (no cluster barriers).  Any B (one cluster per 16 rows), per-CTA
 * slices need not be powers of two.  dv3_scan_tc_fwd_config: cluster size and stream size (bytes) for a shape, DV3_ERR_UNSUPPORTED
 * when the shape does not divide.  Replaces the same reference loop as dv3_rssm_observe_fwd (dreamer_v3.py:122-129). */
typedef struct Dv3ScanTcExtra {
  const void* stream;                   /* dv3_scan_tc_fwd_pack output */
  const float* WmlpzT;                  /* [G*Kc, U] fp32 = Wmlp[:, :G*Kc]^T */
  const float* aW;                      /* [T*B, U] fp32 = masked action . Wmlp[:, G*Kc:]^T */
} Dv3ScanTcExtra;
int dv3_scan_tc_fwd_config(int D, int U, int Hh, int G, int Kc, int* cluster_size, size_t* stream_bytes);
int dv3_scan_tc_fwd_pack(int D, int U, int Hh, int G, int Kc, const float* Wgru, int ld_wgru, const float* Wr0, int ld_wr0,
                         const float* Wr1, int ld_wr1, void* dst, void* stream);
int dv3_rssm_observe_fwd_tc(const Dv3ScanFwdArgs* args, const Dv3ScanTcExtra* extra, void* stream);
/* Weight matrix -> bf16 in mma.sync.m16n8k16 A-fragment order (the operand layout of the two cluster scan kernels): element
 * (row, k) is src[row * stride_r + k * stride_k], R % 16 == 0, K % 32 == 0; dst holds R*K bf16. */
int dv3_pack_mma_a(int R, int K, const float* src, long long stride_r, long long stride_k, void* dst_bf16, void* stream);

/* Reverse scan (backward through time of the same loop) on one thread-block cluster: what autograd does for RSSM.dynamic inside
 * rec_loss.backward() (dreamer_v3.py:177).  Inputs: the tensors saved by the forward scan, dL/d post_logits (KL) and dL/d latent
 * (heads) for all T*B rows, the softmax quantities of dv3_unimix_bwd_prep, bf16 TRANSPOSED weights.  Outputs: the pre-activation
 * gradients from which the caller forms every parameter gradient with K = T*B contractions (and, optionally, their bf16 mirrors).
 * DV3_ERR_UNSUPPORTED when the shape / device cannot run it (the caller then uses the per-step entry points). */
typedef struct Dv3ScanBwdArgs {
  int T, B, G, Kc, D, U, Hh;
  float unimix;
  const float* first;                                   /* [T*B] */
  const float* ps; const float* coef; const float* pn;  /* [T*B, G*Kc] from dv3_unimix_bwd_prep */
  const float* dpost_logits;                            /* [T*B, G*Kc] */
  const float* dlatent; int ld_dlat;                    /* [T*B, G*Kc + D]: head gradients wrt (z_t, h_t) */
  const float* hidr_pre; const float* hidr_mean; const float* hidr_rstd;
  const float* gates_pre; const float* gru_mean; const float* gru_rstd;
  const float* x_gru; int ld_xgru;                      /* [T*B, D+U]: h_in in the first D columns */
  const float* feat_pre; const float* feat_mean; const float* feat_rstd;
  const float* g_r; const float* b_r; const float* g_gru; const float* b_gru; const float* g_mlp; const float* b_mlp;
  const void* Wr1T_pk;                                  /* dv3_pack_mma_a of Wr1^T [Hh, G*Kc], Wr0[:, :D]^T [D, Hh],       */
  const void* Wr0hT_pk;                                 /* Wgru^T [D+U, 3D], Wmlp[:, :G*Kc]^T [G*Kc, U] (strides swapped:  */
  const void* WgruT_pk;                                 /* no transposed copy is materialised)                             */
  const void* WmlpzT_pk;
  float* dpost_raw; float* dhidr_pre; float* dln_r; float* dgates_pre; float* dln_gru; float* dfeat_pre; float* dln_feat;
  void* dpost_raw_bf16; void* dhidr_pre_bf16; void* dgates_pre_bf16; void* dfeat_pre_bf16;   /* optional mirrors (NULL = none) */
  long long* dbg;                                       /* optional: 16 clock64 stamps of CTA 0 at step T-2 */
} Dv3ScanBwdArgs;
int dv3_rssm_observe_bwd_cluster(const Dv3ScanBwdArgs* args, void* stream);
/* The same reverse scan with tcgen05 tensor cores (rssm_scan_bwd_tc.cu; machinery as dv3_rssm_observe_fwd_tc): the four transposed
 * in-loop weight matrices come as one per-CTA, per-step bf16 stream (dv3_scan_tc_bwd_pack reads the fp32 parameters with swapped
 * strides); the *_pk fields of the argument block are ignored.  Any B (one cluster per 16 rows); DV3_ERR_UNSUPPORTED when the shape
 * does not fit (e.g. the [16, 3D] gradient operand of very large D does not fit next to a weight ring in shared memory). */
int dv3_scan_tc_bwd_config(int D, int U, int Hh, int G, int Kc, int* cluster_size, size_t* stream_bytes);
int dv3_scan_tc_bwd_pack(int D, int U, int Hh, int G, int Kc, const float* Wr1, int ld_wr1, const float* Wr0, int ld_wr0,
                         const float* Wgru, int ld_wgru, const float* Wmlp, int ld_wmlp, void* dst, void* stream);
int dv3_rssm_observe_bwd_tc(const Dv3ScanBwdArgs* args, const void* weight_stream, void* stream);
/* ps = softmax(raw), pn = Categorical(logits = unimix(raw)).probs, coef = (1-mix)/pm inside the clamp else 0 (mix = 0: ps = pn,
 * coef = 1): the step-independent factors of the unimix + straight-through Jacobian (see dv3_unimix_bwd). */
int dv3_unimix_bwd_prep(int R, int groups, int K, float mix, const float* raw, int ldr, float* ps, float* coef, float* pn, void* stream);

/* ---- categorical: unimix agent.py:390-402/:741-747; ST one-hot sample distribution.py:355-360,:390-393 ---- */
int dv3_unimix_fwd(int R, int groups, int K, float mix, const float* raw, int ldr, float* logits, int ldo, void* stream);
int dv3_sample_onehot(int R, int groups, int K, const float* logits, int ldl, const float* u, int ldu, const int* forced,
                      int ldf, int mode, int* idx, int ldi, float* onehot, int ldo, void* onehot_bf16, void* stream);
/* dv3_unimix_fwd + dv3_sample_onehot(mode 0) in one launch (imagination rollout, dreamer_v3.py:232-257): `logits` (optional)
 * receives the unimix logits; fast != 0 draws with a warp prefix sum instead of the sequential fp32 cumsum (bf16 precision mode). */
int dv3_unimix_sample(int R, int groups, int K, float mix, const float* raw, int ldr, const float* u, int ldu, const int* forced, int ldf,
                      int* idx, int ldi, float* onehot, int ldo, void* onehot_bf16, float* logits, int ldl, int fast, void* stream);
int dv3_unimix_bwd(int R, int groups, int K, float mix, const float* raw, int ldr, const float* dlogits, int lddl,
                   const float* dz, int lddz, float* draw, int lddr, void* stream);

/* ---- world-model losses, loss.py:65-110 with distribution.py:177-193 (symlog MSE), :253-270 (two-hot) ---- */
int dv3_kl_balance(int R, int groups, int K, const float* post, const float* prior, float free_nats, float w_dyn, float w_rep,
                   float scale, float* rows_out, float* dpost, float* dprior, void* stream);
int dv3_twohot_loss(int M, int K, const float* logits, int ldl, const float* x, const float* w, float scale, float* loss_rows,
                    float* dlogits, int ldd, int accumulate, void* stream);
int dv3_twohot_mean(int M, int K, const float* logits, int ldl, float* out, void* stream); /* distribution.py:245-247 */
int dv3_symlog(size_t n, const float* x, float* y, void* stream);                           /* utils/utils.py:120-121 */
int dv3_symlog_mse(int M, int D, const float* mode, int ldm, const float* target, int ldt, float scale, float* loss_rows,
                   float* dmode, int ldd, void* stream);
int dv3_bernoulli_loss(int M, const float* logit, const float* dones, float scale, float* loss_rows, float* dlogit, void* stream);
int dv3_reduce_mean(int M, const float* x, int ldx, int col, float* out, int slot, float scale, int accumulate, void* stream);

/* ---- behaviour learning --------------------------------------------------------------------------------- */
/* compute_lambda_values algos/dreamer_v3/utils.py:44-55 + continues/discount dreamer_v3.py:262-280 */
int dv3_lambda_returns(int H, int N, const float* values, const float* rewards, const float* cont_logit, const float* dones,
                       float gamma, float lmbda, float* lam, float* discount, void* stream);
/* Moments.forward algos/dreamer_v3/utils.py:34-41 (exact torch.quantile 'linear' by radix select) */
int dv3_moments(int n, const float* x, float* state, float* out, float decay, float max_, float plow, float phigh, void* stream);
/* actor objective dreamer_v3.py:292-317 (discrete): value per row + gradient wrt the raw head logits */
int dv3_actor_loss(int M, int n_heads, const int* head_dims_host, const float* raw, int ldr, const int* act_idx, int lda,
                   const float* lam, const float* values, const float* discount, const float* mom, float mix, float ent_coef,
                   float scale, float* loss_rows, float* draw, int ldd, void* stream);

/* ---- imagination rollout as one persistent kernel (rollout_tc.cu): RSSM.imagination (agent.py:439-455) + Actor.forward
 * (agent.py:682-739) for H steps from N start states, dreamer_v3.py:219-257, bf16 tensor-core mode.  Rows are split into 128-row
 * blocks, each run by a thread-block cluster whose CTAs own 1/CL of every layer's output features; every layer is a TMA-fed
 * tcgen05 GEMM with its LayerNorm / GRU / unimix-sampling epilogue, layer outputs meet in the global trajectory / activation buffers.
 * All operands with a 16 suffix are bf16 copies with the layout of their fp32 originals.  Inputs: rows [0, N) of traj / traj16 hold
 * the start states [z one-hot | h].  DV3_ERR_UNSUPPORTED when the shape is outside the plan (N % 128, several action heads,
 * per-CTA slices that are not multiples of 16, ...): the caller then composes the rollout from the per-op entry points. */
#define DV3_ROLLOUT_MAX_LAYERS 6
typedef struct Dv3RolloutArgs {
  int N, H, G, Kc, S, D, U, Hh, A, lat, L, Ua, nh;
  int head_dim[4], head_off[4];
  float unimix;
  /* weights (bf16, K contiguous) and their row strides in elements */
  const void* Wz16; int ld_wz;            /* Wmlp[:, :S]        [U, S]     */
  const void* Wgru16; int ld_wgru;        /* rnn.linear.weight  [3D, D+U]  */
  const void* Wt016; int ld_wt0;          /* transition layer 0 [Hh, D]    */
  const void* Wt116; int ld_wt1;          /* transition head    [S, Hh]    */
  const void* Wa16[DV3_ROLLOUT_MAX_LAYERS]; int ld_wa[DV3_ROLLOUT_MAX_LAYERS];   /* actor layers [Ua, lat] / [Ua, Ua] */
  const void* Wh16; int ld_wh;            /* actor head         [A, Ua]    */
  /* fp32 parameters */
  const float* W_mlpaT; int ld_mlpaT;     /* Wmlp[:, S:]^T [A, U]: one-hot actions enter as a row gather */
  const float* g_mlp; const float* b_mlp; const float* g_gru; const float* b_gru;
  const float* bt0; const float* g_t; const float* b_t; const float* bt1;
  const float* ga[DV3_ROLLOUT_MAX_LAYERS]; const float* ba[DV3_ROLLOUT_MAX_LAYERS]; const float* bh;
  /* sampling noise / teacher forcing */
  const float* u_z; const float* u_act0; const float* u_act;      /* [H, N, G], [N, nh], [H, N, nh] */
  const int* f_z; const int* f_act;                               /* forced draws [H, N, G] / [(H+1) N, nh], or NULL */
  /* state and outputs */
  float* traj; void* traj16;              /* [(H+1) N, lat] */
  void* feat16; void* thid16;             /* scratch [N, U], [N, Hh] bf16 */
  float* apre[DV3_ROLLOUT_MAX_LAYERS]; float* aact[DV3_ROLLOUT_MAX_LAYERS]; void* aact16[DV3_ROLLOUT_MAX_LAYERS];   /* [(H+1) N, Ua] */
  float* amean[DV3_ROLLOUT_MAX_LAYERS]; float* arstd[DV3_ROLLOUT_MAX_LAYERS];                                       /* [(H+1) N]     */
  float* aout; int* act_idx; float* acts; int* z_idx;             /* [(H+1) N, A], [(H+1) N, nh], [(H+1) N, A] or NULL, [H, N, G] */
  long long* dbg;                         /* optional: 60 clock64 stamps of one epilogue thread during step 2 ($DV3_SCAN_DEBUG) */
} Dv3RolloutArgs;
int dv3_rollout_tc(const Dv3RolloutArgs* args, void* stream);

/* ---- optimizer: clip_grad_norm_ + torch.optim.Adam.step, dreamer_v3.py:179-187,:318-323,:339-344 ---------- */
size_t dv3_adam_clip_scratch_bytes(void);
int dv3_adam_clip(size_t n, float* p, const float* g, float* m, float* v, float* step, double lr, double beta1, double beta2,
                  double eps, double max_norm, float* norm_out, int slot, float* scratch, size_t scratch_bytes, void* stream);
int dv3_ema(size_t n, float* target, const float* src, float tau, void* stream); /* dreamer_v3.py:713-716 */

/* ---- utilities ------------------------------------------------------------------------------------------- */
int dv3_copy2d(int R, int C, const float* s, int lds, float* d, int ldd, void* stream);
int dv3_transpose(int R, int C, const float* s, int lds, float* d, int ldd, void* stream);
int dv3_fill(size_t n, float* p, float v, void* stream);
int dv3_uniform_fill(size_t n, float* out, unsigned long long seed, long long* counter, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* DV3_B200_H */
