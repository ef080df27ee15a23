/*
 * b200_transducer.h — C ABI of libb200_transducer.so
 *
 * Drop-in boundary for the transducer hot path of WeNet (reference tree:
 * Vindicator645/wenet-celoss).  Every entry point takes raw DEVICE pointers
 * owned by the caller, plain sizes and a CUDA stream handle; the library
 * allocates nothing, keeps no global state except a thread-local error string,
 * enqueues its kernels on the given stream and returns immediately.
 * Return value: 0 = ok, non-zero = error (text via b200_last_error()).
 * Nothing throws across this boundary.
 *
 * The reference has no FFI of its own for this path: the arithmetic lives in
 * torchaudio / ATen and is reached through Python call sites.  Each function
 * below names the reference call site (file:line under /root/reference) whose
 * work it replaces.  The Python wrappers in wenet_celoss_b200/ bind these via
 * ctypes.
 */
#ifndef B200_TRANSDUCER_H_
#define B200_TRANSDUCER_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* cudaStream_t without pulling in cuda_runtime.h */
typedef void* b200_stream_t;

/* element types of a logits / gradient buffer */
enum b200_dtype { B200_F32 = 0, B200_F16 = 1, B200_BF16 = 2 };

enum b200_status {
  B200_OK = 0,
  B200_ERR_ARG = 1,      /* bad shape / null pointer / unsupported dtype */
  B200_ERR_CUDA = 2,     /* a CUDA runtime call or launch failed */
  B200_ERR_WORKSPACE = 3 /* workspace too small */
};

/* ---- library ---------------------------------------------------------- */
int b200_version(void);
const char* b200_last_error(void);
/* compute capability the kernels were compiled for (100 => sm_100a) */
int b200_compiled_arch(void);
/* number of kernel launches this library has enqueued in this process */
uint64_t b200_launch_count(void);

/* ---- RNN-T loss over materialised logits ------------------------------
 * replaces torchaudio.functional.rnnt_loss as called at
 *   wenet/transducer/transducer.py:142-147 (training, reduction="mean")
 *   wenet/transducer/transducer.py:296-301 (rescoring, reduction="none")
 * logits  [B,T,U1,V] contiguous, dtype f32/f16/bf16   (U1 = max target len + 1)
 * targets [B,U1-1] int32 zero padded; logit_len/target_len [B] int32.
 * lse, lp_blank, lp_emit, alpha, beta: [B,T,U1] f32; costs [B] f32.
 */

/* pass 1: per lattice cell, log-sum-exp over V and the two log-probs the
 * lattice needs (blank, and the target symbol y[b,u] for u < U_b). Cells with
 * t >= T_b or u > U_b are not read. */
int b200_rnnt_logprobs(const void* logits, int dtype, const int32_t* targets,
                       const int32_t* logit_len, const int32_t* target_len,
                       int B, int T, int U1, int V, int blank,
                       float* lse, float* lp_blank, float* lp_emit,
                       b200_stream_t stream);

/* pass 2: alpha / beta lattices by anti-diagonal wavefront; costs[b] = -beta[b,0,0].
 * alpha/beta are stored RESCALED: true alpha[b,t,u] = alpha[b,t,u] + diag_off[b][0][t+u]
 * (beta: diag_off[b][1][t+u]); diag_off is [B][2][T+U1] fp64 and slot
 * diag_off[b][0][T+U1-1] holds the log-likelihood of utterance b in fp64. */
int b200_rnnt_lattice(const float* lp_blank, const float* lp_emit,
                      const int32_t* logit_len, const int32_t* target_len,
                      int B, int T, int U1,
                      float* alpha, float* beta, float* costs, double* diag_off,
                      b200_stream_t stream);

/* pass 2b: per-cell gradient coefficients {c0, cb, ce, scale} (float4 per cell):
 *   grad[v] = scale * ( exp(x[v]+c0) - [v==blank] exp(x[v]+cb) - [v==y_u] exp(x[v]+ce) )
 * with scale = dy[b]; padded cells get c0=cb=ce=-inf (gradient exactly 0). */
int b200_rnnt_grad_coef(const float* lse, const float* alpha, const float* beta,
                        const double* diag_off, const float* dy,
                        const int32_t* targets,
                        const int32_t* logit_len, const int32_t* target_len,
                        int B, int T, int U1, int blank,
                        float* coef /* [B,T,U1,4] */, b200_stream_t stream);

/* pass 3: gradient w.r.t. logits, same dtype/layout as logits; softmax is
 * recomputed from logits + coef; clamp <= 0 disables clamping. */
int b200_rnnt_grad(const void* logits, int dtype, const int32_t* targets,
                   const float* coef, int B, int T, int U1, int V, int blank,
                   float clamp, void* grad, b200_stream_t stream);

/* convenience: passes 1+2 back to back */
int b200_rnnt_loss_fwd(const void* logits, int dtype, const int32_t* targets,
                       const int32_t* logit_len, const int32_t* target_len,
                       int B, int T, int U1, int V, int blank,
                       float* costs, float* lse, float* lp_blank, float* lp_emit,
                       float* alpha, float* beta, double* diag_off, b200_stream_t stream);

/* ---- CTC loss over logits ----------------------------------------------
 * replaces, behind CTC.forward (wenet/transformer/ctc.py:46-64), the
 * log_softmax (ctc.py:60) + torch.nn.CTCLoss(reduction='sum') (ctc.py:44,61) pair.
 * logits [B,T,V] contiguous, batch-major (the module's own layout: no [T,B,V] copy),
 * dtype f32/f16/bf16.  ys_pad [B,Lmax] int64 padded with -1 (as the reference passes
 * it), hlens / ys_lens [B] int32.  Smax = 2*Lmax+1 (<= 1024).
 * lse [B,T]; lp_ext, alpha, beta [B,T,Smax] f32; nll [B] f32; leader [B,Smax] int32.
 */

/* pass 1: lse per frame + log-probs of the blank-interleaved label sequence */
int b200_ctc_logprobs(const void* logits, int dtype, const int64_t* ys_pad,
                      const int32_t* hlens, const int32_t* ys_lens,
                      int B, int T, int V, int Lmax, int blank,
                      float* lse, float* lp_ext, b200_stream_t stream);

/* pass 2: alpha / beta trellis over T, nll[b] (inf when no alignment exists:
 * zero_infinity=False), and the duplicate-label leader table used by pass 3.
 * alpha/beta are stored RESCALED: true alpha_t(s) = alpha[b,t,s] + frame_off[b][0][t]
 * (beta: frame_off[b][1][t]); frame_off is [B][2][T+1] fp64, frame_off[b][0][T] = log-likelihood. */
int b200_ctc_lattice(const float* lp_ext, const int64_t* ys_pad,
                     const int32_t* hlens, const int32_t* ys_lens,
                     int B, int T, int Lmax, int blank,
                     float* alpha, float* beta, float* nll, int32_t* leader,
                     double* frame_off, b200_stream_t stream);

/* pass 3: grad[b,t,c] = scale[b] * (softmax[t,c] - posterior of class c at t), 0 for t >= hlens[b] */
int b200_ctc_grad(const void* logits, int dtype, const float* lse, const float* lp_ext,
                  const float* alpha, const float* beta, const double* frame_off,
                  const float* scale, const int64_t* ys_pad,
                  const int32_t* hlens, const int32_t* ys_lens, const int32_t* leader,
                  int B, int T, int V, int Lmax, int blank,
                  void* grad, b200_stream_t stream);

/* ---- Fused joint + RNN-T loss (logits never materialised) ----------------
 * replaces wenet/transducer/transducer.py:132-147: TransducerJoint.forward
 * (wenet/transducer/joint.py:58-69: broadcast add, tanh, ffn_out) fused with the
 * log-softmax statistics / logits-gradient of torchaudio.functional.rnnt_loss.
 * e [B*T, J], p [B*U1, J] fp32 = outputs of enc_ffn / pred_ffn (joint.py:55-56, library
 * GEMMs on the caller's side); w_bf16 [V, J] bf16 = ffn_out.weight; bias_pad [>= V] fp32 =
 * ffn_out.bias (only the first V entries are read).  J multiple of 64, <= 512.
 * Tensor-core kernel (tcgen05 / TMEM / TMA); max_ctas <= 0 means one CTA per SM.
 */

/* bound[0] (device) = max_v (|W_v|_1 + |b_v|) >= |logit| of any cell, since |tanh| <= 1 */
int b200_joint_logit_bound(const void* w_bf16, const float* bias, int V, int J, float* bound,
                           b200_stream_t stream);

/* forward: per lattice cell lse over V and the blank / target log-probs [B,T,U1].
 * logit_bound: NULL, or the device scalar of b200_joint_logit_bound for THESE weights; when
 * it is small enough for 2^(bound*log2 e) to stay inside fp32 the log-sum-exp runs without a
 * running maximum (same result, fewer instructions per logit). */
int b200_joint_fwd(const float* e, const float* p, const void* w_bf16, const float* bias_pad,
                   const int32_t* targets, const int32_t* logit_len, const int32_t* target_len,
                   int B, int T, int U1, int V, int J, int blank,
                   float* lse, float* lp_blank, float* lp_emit, const float* logit_bound,
                   int max_ctas, b200_stream_t stream);

/* training forward: as b200_joint_fwd, and the A-tile builders also store H = tanh(e + p)
 * (bf16, [B*T*U1, ldh], ldh >= J, ldh % 8 == 0) through TMA, so that the backward pass
 * (b200_joint_grad_chunk_h, b200_joint_dw, b200_joint_dz_reduce) does not rebuild it.  H NULL =
 * b200_joint_fwd. */
int b200_joint_fwd_h(const float* e, const float* p, const void* w_bf16, const float* bias_pad,
                     const int32_t* targets, const int32_t* logit_len, const int32_t* target_len,
                     int B, int T, int U1, int V, int J, int blank,
                     float* lse, float* lp_blank, float* lp_emit, const float* logit_bound,
                     void* H, int ldh, int max_ctas, b200_stream_t stream);

/* training forward of the NO-RECOMPUTE backward: as b200_joint_fwd_h (H [B*T*U1, ldh] bf16 is written
 * first, by a kernel of its own, and read back through TMA), and the epilogue also stores
 * Pt [B*T*U1, ldg] bf16 = exp(logit) (ldg >= V, ldg % 16 == 0; columns >= V are written as zero).
 * Requires the logit bound (b200_joint_logit_bound) and bound * log2(e) < 100; callers should stay
 * below 60 so that rowscale * Pt keeps its relative precision.  The backward pass is then
 * b200_rnnt_grad_coef -> b200_joint_rowscale -> b200_joint_dw_db + b200_joint_dh_scaled: the logits
 * GEMM of joint.py:68-69 runs once, not twice. */
int b200_joint_fwd_p(const float* e, const float* p, const void* w_bf16, const float* bias_pad,
                     const int32_t* targets, const int32_t* logit_len, const int32_t* target_len,
                     int B, int T, int U1, int V, int J, int blank,
                     float* lse, float* lp_blank, float* lp_emit, const float* logit_bound,
                     void* H, int ldh, void* Pt, int ldg, int max_ctas, b200_stream_t stream);

/* b200_joint_fwd_p split in its two launches: H = tanh(e + p) alone, and the tensor-core kernel on a ready H */
int b200_joint_h(const float* e, const float* p, int B, int T, int U1, int J, void* H, int ldh,
                 b200_stream_t stream);
int b200_joint_fwd_p_h(const float* e, const float* p, const void* w_bf16, const float* bias_pad,
                       const int32_t* targets, const int32_t* logit_len, const int32_t* target_len,
                       int B, int T, int U1, int V, int J, int blank,
                       float* lse, float* lp_blank, float* lp_emit, const float* logit_bound,
                       const void* H, int ldh, void* Pt, int ldg, int max_ctas, b200_stream_t stream);

/* d loss / d logits = rowscale[c] * Pt'[c, v]: writes rowscale [B*T*U1] fp32 (= scale * exp(c0) of the
 * coefficients of b200_rnnt_grad_coef), overwrites the blank and target entries of every row of Pt with
 * exp(x) - exp(x + cb - c0) / exp(x) - exp(x + ce - c0) (x from lse + lp_blank / lp_emit), and writes
 * Hs [B*T*U1, ldh] bf16 = rowscale * H (the second operand of the weight-gradient GEMM). */
int b200_joint_rowscale(const float* coef, const float* lse, const float* lp_blank,
                        const float* lp_emit, const int32_t* targets, int B, int T, int U1, int V,
                        int J, int blank, void* Pt, int ldg, const void* H, int ldh, void* Hs,
                        float* rowscale, b200_stream_t stream);

/* backward, one chunk of cells [cell0, cell0+ncells) (cell = (b*T+t)*U1+u): recompute the
 * logits tile and write G = d loss / d logits (bf16, [ncells, ldg], ldg % 16 == 0) from the
 * coefficients of b200_rnnt_grad_coef.  H [ncells, ldh] bf16 (ldh >= J, ldh % 8 == 0) is a
 * required scratch: columns [0,J) receive tanh(e+p) (the kernel's A operand, also the
 * activations the dW GEMM needs); columns >= J are left untouched.
 * db_part: NULL, or fp32 [16][ceil(V/128)*128] (zeroed by the caller before the first chunk):
 * receives partial column sums of G, i.e. d b_out = sum over the 16 rows, first V entries. */
int b200_joint_grad_chunk(const float* e, const float* p, const void* w_bf16,
                          const float* bias_pad, const int32_t* targets, const float* coef,
                          int B, int T, int U1, int V, int J, int blank,
                          long long cell0, long long ncells,
                          void* G, int ldg, void* H, int ldh, float* db_part,
                          int max_ctas, b200_stream_t stream);

/* the same with H READY: H points at the chunk's first row of the scratch b200_joint_fwd_h
 * filled (row cell0 of the whole-lattice H); no tanh is recomputed. */
int b200_joint_grad_chunk_h(const float* e, const float* p, const void* w_bf16,
                            const float* bias_pad, const int32_t* targets, const float* coef,
                            int B, int T, int U1, int V, int J, int blank,
                            long long cell0, long long ncells,
                            void* G, int ldg, const void* H, int ldh, float* db_part,
                            int max_ctas, b200_stream_t stream);

/* backward tail: dZ = dH * (1 - H^2), d e[b,t,:] = sum_u dZ, d p[b,u,:] += sum_t dZ
 * (autograd through joint.py:58-68) for nb whole utterances: dH [nb*T*U1, J] fp32,
 * H [nb*T*U1, ldh] bf16, de [nb,T,J] (written), dp [nb,U1,J] (accumulated: zero it first). */
int b200_joint_dz_reduce(const float* dH, const void* H, int ldh, int nb, int T, int U1, int J,
                         float* de, float* dp, b200_stream_t stream);

/* weight gradient of ffn_out for one chunk: dW [V, J] fp32 += G^T H (autograd of
 * joint.py:68-69), G [n, ldg] bf16 and H [n, ldh] bf16 as written by b200_joint_grad_chunk.
 * 2-CTA tcgen05 kernel (cta_group::2, M = 256, both operands MN-major through TMA); J must
 * be a multiple of 128 in [128, 512]; dW must be zeroed before the first chunk. */
int b200_joint_dw(const void* G, int ldg, const void* H, int ldh, long long n, int V, int J,
                  float* dW, int max_ctas, b200_stream_t stream);

/* b200_joint_dw with the bias gradient: db [V] fp32 += sum over cells of rowscale[c] * G[c, v], computed
 * by the kernel's otherwise idle epilogue warps from the G boxes already staged in shared memory
 * (G = Pt' of b200_joint_rowscale, H = Hs).  db must be zeroed before the first call. */
int b200_joint_dw_db(const void* G, int ldg, const void* H, int ldh, long long n, int V, int J,
                     float* dW, const float* rowscale, float* db, int max_ctas, b200_stream_t stream);

/* b200_joint_dh with a per-cell factor: dH = rowscale[c] * (G W_out)[c, :] (G = Pt' of
 * b200_joint_rowscale; rowscale [nb*T*U1] fp32). */
int b200_joint_dh_scaled(const void* G, int ldg, const void* w_bf16, const void* H, int ldh,
                         const float* rowscale, int nb, int T, int U1, int V, int J,
                         const int32_t* logit_len, const int32_t* target_len, float* de, float* dp,
                         int max_ctas, b200_stream_t stream);

/* hidden-side gradient of the fused joint for one chunk of nb whole utterances, in ONE kernel:
 * dH = G W_out (autograd of ffn_out, joint.py:68-69), dZ = dH * (1 - H^2) (tanh, joint.py:63-64),
 * de [nb,T,J] += sum_u dZ, dp [nb,U1,J] += sum_t dZ (broadcast add, joint.py:58-60) - both outputs are
 * ACCUMULATED with fp32 atomics: zero them first.  G [nb*T*U1, ldg] bf16, H [nb*T*U1, ldh] bf16 as
 * written by b200_joint_grad_chunk / b200_joint_fwd_h, w_bf16 [V, J].  2-CTA tcgen05 kernel computing
 * the product transposed (TMEM lane = j, columns = cells); J must be 256 or 512.  logit_len / target_len
 * (optional, [nb]) let it skip tiles beyond an utterance's lengths, where G is exactly zero. */
int b200_joint_dh(const void* G, int ldg, const void* w_bf16, const void* H, int ldh, int nb, int T,
                  int U1, int V, int J, const int32_t* logit_len, const int32_t* target_len,
                  float* de, float* dp, int max_ctas, b200_stream_t stream);

/* ---- fp32-accurate (split-bf16) tensor-core path -----------------------------------
 * For the configurations the reference runs in fp32 (BASELINE configs[1]: joint.py:58-69 +
 * torchaudio rnnt_loss, rtol 1e-4; and the ctc_lo projection + log_softmax + CTCLoss of
 * wenet/transformer/ctc.py:57-61).  Every GEMM operand is a PAIR of bf16 arrays,
 * X = X_hi + X_lo with X_hi = bf16(X), X_lo = bf16(X - X_hi), and each product is issued as
 * hi.hi + lo.hi + hi.lo on tcgen05 with fp32 accumulation (~2^-17 relative operand error; the
 * few logits the lattices read are recomputed as exact fp32 dot products).  The logits never
 * reach HBM.  *_lo == NULL everywhere = plain bf16 operands (one product). */

/* x [rows, cols] fp32 (pitch ldx) -> hi, lo [rows, cols] bf16 (pitch ldo) */
int b200_split_f32(const float* x, long long rows, int cols, long long ldx, void* hi, void* lo,
                   long long ldo, b200_stream_t stream);

/* x[r,c] = sum_k A[r,k] B[c,k] + bias[c]; A [rows, K] (pitch lda), B [N, K] (pitch ldb), both
 * bf16 pairs, K % 8 == 0; bias_pad fp32 with ceil(N/256)*256 entries.
 *  _lse: lse[r] = log sum_c exp x[r,c]                        (log_softmax denominators)
 *  _exp: G[r,c] = scale_r exp(x[r,c] + c0_r), coef [rows] float4 {c0, -, -, scale}, written as
 *        the split pair g_hi / g_lo [rows, N] bf16 (pitch ldg): the softmax part of the loss
 *        gradient; the sparse posterior terms follow through b200_rnnt_fixup / b200_ctc_fixup
 *  _lsm: out[r,c] = x[r,c] - lse[r] in fp32 (pitch ldo)       (CTC.log_softmax, ctc.py:66-74) */
int b200_split_gemm_lse(const void* a_hi, const void* a_lo, long long rows, int K, int lda,
                        const void* b_hi, const void* b_lo, int N, int ldb, const float* bias_pad,
                        float* lse, int max_ctas, b200_stream_t stream);
int b200_split_gemm_exp(const void* a_hi, const void* a_lo, long long rows, int K, int lda,
                        const void* b_hi, const void* b_lo, int N, int ldb, const float* bias_pad,
                        const float* coef, void* g_hi, void* g_lo, long long ldg, int max_ctas,
                        b200_stream_t stream);
/* b200_split_gemm_exp with a sparse term subtracted inside the epilogue: G = scale exp(x + c0) - sub, the
 * whole CTC gradient w.r.t. the ctc_lo logits (autograd of log_softmax + CTCLoss, ctc.py:60-61) in one pass
 * instead of b200_split_gemm_exp + b200_ctc_fixup (a read-modify-write of the label entries of G).
 * b200_ctc_post writes the term: rows come in groups of group_rows = T (utterances);
 *   words [B, ceil(V/32), 2] uint32: {bitmap over the classes: blank + the utterance's distinct valid labels,
 *                                     number of set bits in the words before this one}
 *   val   [B, Lmax + 1, T] fp32: val[b, rank(c), t] = scale_b * posterior(class c at frame t), rank(c) = set bits
 *                                below c (entries of ranks the utterance does not use are never read).
 * alpha / beta / frame_off / leader as b200_ctc_lattice wrote them; rows must be B * T. */
int b200_ctc_post(const float* lp_ext, const float* alpha, const float* beta, const double* frame_off,
                  const float* scale, const int64_t* ys_pad, const int32_t* hlens,
                  const int32_t* ys_lens, const int32_t* leader, int B, int T, int V, int Lmax,
                  int blank, uint32_t* words, float* val, b200_stream_t stream);
int b200_split_gemm_exp_sub(const void* a_hi, const void* a_lo, long long rows, int K, int lda,
                            const void* b_hi, const void* b_lo, int N, int ldb, const float* bias_pad,
                            const float* coef, const uint32_t* sub_words, const float* sub_val,
                            int group_rows, int nranks, void* g_hi, void* g_lo, long long ldg,
                            int max_ctas, b200_stream_t stream);
int b200_split_gemm_lsm(const void* a_hi, const void* a_lo, long long rows, int K, int lda,
                        const void* b_hi, const void* b_lo, int N, int ldb, const float* bias_pad,
                        const float* lse, float* out, long long ldo, int max_ctas,
                        b200_stream_t stream);

/* out [rows, J] fp32 = G W: G [rows, V] bf16 pair (pitch ldg), W [V, J] bf16 pair, J 256 or 512
 * (d hs of the ctc_lo projection: autograd of ctc.py:57) */
int b200_split_gemm_nn(const void* G_hi, const void* G_lo, int ldg, const void* w_hi,
                       const void* w_lo, long long rows, int V, int J, float* out, int max_ctas,
                       b200_stream_t stream);

/* H = tanh(e[b,t,:] + p[b,u,:]) in fp32 (joint.py:58-64) as a split pair [B*T*U1, ldh] */
int b200_joint_h_split(const float* e, const float* p, int B, int T, int U1, int J, void* H_hi,
                       void* H_lo, int ldh, b200_stream_t stream);

/* the two logits per lattice cell the RNN-T lattice reads (blank, target), as fp32 dot products
 * with W [V, J] fp32: lp_blank / lp_emit = logit - lse [B,T,U1] (cells beyond the lengths are
 * not written), xbe [B,T,U1,2] = the raw logits for b200_rnnt_fixup */
int b200_joint_pick(const float* e, const float* p, const float* W, const float* bias,
                    const int32_t* targets, const int32_t* logit_len, const int32_t* target_len,
                    int B, int T, int U1, int V, int J, int blank, const float* lse,
                    float* lp_blank, float* lp_emit, float* xbe, b200_stream_t stream);

/* lp_ext [B,T,2*Lmax+1] of b200_ctc_logprobs from hs [B,T,E] and ctc_lo (W [V,E], bias) directly:
 * fp32 dot products with the gathered rows of the blank and the utterance's labels, minus lse */
int b200_ctc_lp_ext(const float* hs, const float* W, const float* bias, const int64_t* ys_pad,
                    const int32_t* hlens, const int32_t* ys_lens, int B, int T, int E, int V,
                    int Lmax, int blank, const float* lse, float* lp_ext, b200_stream_t stream);

/* G[cell, blank] -= scale exp(x_blank + cb), G[cell, y] -= scale exp(x_y + ce) with coef [cell]
 * = {c0, cb, ce, scale} of b200_joint_grad_coef (the sparse terms of the rnnt_loss gradient) */
int b200_rnnt_fixup(const float* coef, const float* xbe, const int32_t* targets, int B, int T,
                    int U1, int V, int blank, void* G_hi, void* G_lo, long long ldg,
                    b200_stream_t stream);

/* G[b,t,c] -= scale[b] * posterior(c at t) for the blank and every distinct label (the sparse
 * terms of pass 3 of the CTC loss); alpha / beta / frame_off / leader as b200_ctc_lattice wrote */
int b200_ctc_fixup(const float* lp_ext, const float* alpha, const float* beta,
                   const double* frame_off, const float* scale, const int64_t* ys_pad,
                   const int32_t* hlens, const int32_t* ys_lens, const int32_t* leader, int B,
                   int T, int V, int Lmax, int blank, void* G_hi, void* G_lo, long long ldg,
                   b200_stream_t stream);

/* out[c] += sum_r (G_hi + G_lo)[r, c] (the bias gradient); out must be zeroed first */
int b200_colsum_split(const void* G_hi, const void* G_lo, long long rows, int N, long long ldg,
                      float* out, b200_stream_t stream);

/* ---- no-recompute backward in fp32 mode (cf. b200_joint_fwd_p / b200_joint_rowscale for bf16) ----
 * b200_split_gemm_explse: lse [rows] AND Pt = exp(x) as a split pair (p_hi, p_lo [rows, ldp]) from ONE pass
 * over split operands; logit_bound = the device scalar of b200_joint_logit_bound: with
 * bound * log2(e) >= 60 the kernel only computes lse (running maximum) and stores nothing.
 * b200_joint_rowscale_split: rowscale [cells], the blank / target entries of the split Pt replaced by
 * exp(x) - exp(x + c_b - c_0) (x = the exact fp32 logits xbe of b200_joint_pick), Hs = rowscale * H split again.
 * b200_colsum_split_w: out[c] += sum_r w[r] (G_hi + G_lo)[r, c].   b200_joint_dh_split_scaled: b200_joint_dh_split
 * with dH = rowscale[cell] * (G W). */
int b200_split_gemm_explse(const void* a_hi, const void* a_lo, long long rows, int K, int lda,
                           const void* b_hi, const void* b_lo, int N, int ldb, const float* bias_pad,
                           const float* logit_bound, float* lse, void* p_hi, void* p_lo, long long ldp,
                           int max_ctas, b200_stream_t stream);
int b200_joint_rowscale_split(const float* coef, const float* xbe, const int32_t* targets, int B, int T,
                              int U1, int V, int J, int blank, void* P_hi, void* P_lo, long long ldg,
                              const void* H_hi, const void* H_lo, int ldh, void* Hs_hi, void* Hs_lo,
                              float* rowscale, b200_stream_t stream);
int b200_colsum_split_w(const void* G_hi, const void* G_lo, long long rows, int N, long long ldg,
                        const float* w, float* out, b200_stream_t stream);
int b200_joint_dh_split_scaled(const void* G_hi, const void* G_lo, int ldg, const void* w_hi,
                               const void* w_lo, const void* H_hi, const void* H_lo, int ldh,
                               const float* rowscale, int nb, int T, int U1, int V, int J,
                               const int32_t* logit_len, const int32_t* target_len, float* de, float* dp,
                               int max_ctas, b200_stream_t stream);

/* b200_joint_dw / b200_joint_dh on split operands (three products each) */
int b200_joint_dw_split(const void* G_hi, const void* G_lo, int ldg, const void* H_hi,
                        const void* H_lo, int ldh, long long n, int V, int J, float* dW,
                        int max_ctas, b200_stream_t stream);
/* b200_joint_dw_split with the bias gradient (autograd of ffn_out.bias, joint.py:69, and of ctc_lo.bias,
 * ctc.py:57): db [V] += sum_r w[r] (G_hi + G_lo)[r, v], w = rowscale [n] or NULL for plain column sums;
 * taken from the operand tiles in shared memory (no second pass over G).  db is accumulated: zero it first. */
int b200_joint_dw_split_db(const void* G_hi, const void* G_lo, int ldg, const void* H_hi,
                           const void* H_lo, int ldh, long long n, int V, int J, float* dW,
                           const float* rowscale, float* db, int max_ctas, b200_stream_t stream);
int b200_joint_dh_split(const void* G_hi, const void* G_lo, int ldg, const void* w_hi,
                        const void* w_lo, const void* H_hi, const void* H_lo, int ldh, int nb,
                        int T, int U1, int V, int J, const int32_t* logit_len,
                        const int32_t* target_len, float* de, float* dp, int max_ctas,
                        b200_stream_t stream);

/* ---- Batched RNN-T greedy search ------------------------------------------
 * replaces the Python loop of wenet/transducer/search/greedy_search.py:34-176 (core
 * algorithm: "greedy_search copy.py":6-63): per lattice step predictor.forward_step
 * (predictor.py:180-200), joint (joint.py:45-70), log_softmax/argmax, and the host-side
 * token / frame bookkeeping - for a whole batch, without host round trips.
 * enc_proj [B,T,J] fp32 = enc_ffn(encoder_out); lens [B]; embed [V,Emb];
 * lstm = HOST array of 4*L device pointers {w_ih, w_hh, b_ih, b_hh} per layer (PyTorch
 * nn.LSTM layout, gate order i,f,g,o); w_c [J,H], b_c [J] = pred_ffn o projection;
 * w_out [V,J], b_out [V] = ffn_out.  tokens [B,max_tokens] / n_tok [B] receive the
 * hypotheses.  Blocks until every utterance has consumed its frames (the only entry
 * point that synchronises the stream: termination is data dependent). */
size_t b200_greedy_workspace_bytes(int B, int V, int H, int L, int J);
int b200_rnnt_greedy(const float* enc_proj, const int32_t* lens, int B, int T,
                     const float* embed, const float* const* lstm,
                     const float* w_c, const float* b_c, const float* w_out, const float* b_out,
                     int V, int Emb, int H, int L, int J, int blank, int n_steps,
                     int32_t* tokens, int32_t* n_tok, int max_tokens,
                     void* workspace, size_t ws_bytes, int* rounds_out, b200_stream_t stream);

/* greedy search with the fork's hot-word predictor bias applied at every predictor step
 * (wenet/transducer/search/greedy_search.py:92-144 with context_filter_state='off':
 * ContextBias.forward_predictor_bias, wenet/transformer/context_bias.py:375-381, over the EMPTY
 * context).  All pointers are device pointers to fp32 row-major weights of the reference modules:
 * predictor.projection [P,H]; predictor_bias.linear_q [P,P]; k / v = linear_k / linear_v applied
 * to the Nc context vectors [Nc,P]; linear_out [P,P]; predictor_bias_bias_norm; predictor_bias_combine
 * [P,2P]; predictor_bias_out_norm; joint.pred_ffn [J,P]. */
typedef struct {
  const float *proj_w, *proj_b, *wq, *bq, *k, *v, *wo, *bo, *ln1_g, *ln1_b, *wc, *bc, *ln2_g, *ln2_b,
      *wf, *bf;
  int P, Nc, heads;
  float ln_eps;
} b200_pred_bias_t;
int b200_rnnt_greedy_bias(const float* enc_proj, const int32_t* lens, int B, int T,
                          const float* embed, const float* const* lstm, const float* w_out,
                          const float* b_out, int V, int Emb, int H, int L, int J, int blank,
                          int n_steps, const b200_pred_bias_t* bias, int32_t* tokens,
                          int32_t* n_tok, int max_tokens, void* workspace, size_t ws_bytes,
                          int* rounds_out, b200_stream_t stream);

/* greedy search with the fork's hot-word GATE and "go-back" - context_filter_state='on', the
 * default of Transducer.greedy_search (wenet/transducer/transducer.py:525) and of
 * wenet/bin/recognize.py:139-141; loop at wenet/transducer/search/greedy_search.py:88-165.
 * Per predictor call a gate decides 0 / 1, or a go-back (:116-131) that drops the last decision
 * and token, rewinds the frame index and re-runs the previous predictor call with decisions
 * forced to 1.  Decision 1 joins the encoder stream biased with the given phrase list
 * (`enc_proj`), decision 0 the one biased with the EMPTY list (gate->enc_proj_empty); the
 * predictor stream is biased with bias->k / bias->v on decision 1, gate->k0 / v0 on decision 0.
 *  loss_mode 'pred' (basic_greedy_search, :34-176): gate = argmax ContextBias.forward_hw_pred
 *    (wenet/transformer/context_bias.py:382-387) of the predictor output biased with the given
 *    list (= k0 / v0; bias->k / v hold the EMPTY list); gate_frames = NULL, hw head weights
 *    given: hw_bias.linear_q / linear_out, hk / hv = its linear_k / linear_v over the given
 *    phrase vectors [Nc_hw,P], hw_bias_norm, hw_output_layer [n_labels,P] (unified_hw_odim == P).
 *  loss_mode 'both' (basic_greedy_search_both, :297-430): gate = forward_hw_pred_both
 *    (context_bias.py:388-394), whose attention has a single key (the frame's encoder bias) and
 *    therefore depends on the frame only: gate_frames [B,T] int32 holds it; bias->k / v = given
 *    list, k0 / v0 = EMPTY list (:399-400); hw head pointers unused.
 * K / V tensors are linear_k / linear_v of the attention applied to the phrase vectors [Nc,P].
 * Outputs as b200_rnnt_greedy plus decisions [B,max_dec] / n_dec [B] = the reference's
 * `result` list. */
typedef struct {
  const float* enc_proj_empty;
  const float *k0, *v0;
  const float *hq_w, *hq_b, *hk, *hv, *ho_w, *ho_b, *hln_g, *hln_b, *hout_w, *hout_b;
  const int32_t* gate_frames;
  int Nc0, Nc_hw, hw_heads, n_labels;
} b200_gate_t;
size_t b200_greedy_gated_workspace_bytes(int B, int V, int H, int L, int J);
int b200_rnnt_greedy_gated(const float* enc_proj, const int32_t* lens, int B, int T,
                           const float* embed, const float* const* lstm, const float* w_out,
                           const float* b_out, int V, int Emb, int H, int L, int J, int blank,
                           int n_steps, const b200_pred_bias_t* bias, const b200_gate_t* gate,
                           int32_t* tokens, int32_t* n_tok, int max_tokens, int32_t* decisions,
                           int32_t* n_dec, int max_dec, void* workspace, size_t ws_bytes,
                           int* rounds_out, b200_stream_t stream);

/* ---- Batched RNN-T prefix beam search ---------------------------------------
 * replaces PrefixBeamSearch.prefix_beam_search (wenet/transducer/search/
 * prefix_beam_search.py:42-148) for a batch: per frame predictor step + joint + log_softmax
 * on every live hypothesis, CTC shallow fusion (:99-101), top-k (:104), N*N expansion
 * (:108-127), prefix merge with fp64 log_add (:130-142), stable sort + prune (:145-146).
 * ctc_lp [B,T,V] fp32 = CTC.log_softmax(encoder_out) (ctc.py:66-74); other operands as for
 * b200_rnnt_greedy.  beam <= 16.  Outputs: out_hyp [B,beam,T+2] (each hypothesis starts
 * with the seed blank, padded with -1), out_len [B,beam], out_score [B,beam] fp64,
 * out_n [B] = hypotheses in the final beam.  Fully asynchronous on the stream. */
size_t b200_beam_workspace_bytes(int B, int T, int V, int H, int L, int J, int beam);
int b200_rnnt_beam(const float* enc_proj, const float* ctc_lp, const int32_t* lens, int B, int T,
                   const float* embed, const float* const* lstm,
                   const float* w_c, const float* b_c, const float* w_out, const float* b_out,
                   int V, int Emb, int H, int L, int J, int blank,
                   int beam, float ctc_weight, float transducer_weight,
                   int32_t* out_hyp, int32_t* out_len, double* out_score, int32_t* out_n,
                   void* workspace, size_t ws_bytes, b200_stream_t stream);

/* ---- Batched CTC prefix beam search ------------------------------------------
 * replaces the Python loop of ASRModel._ctc_prefix_beam_search
 * (wenet/transformer/asr_model.py:326-409; C++ twin runtime/core/decoder/
 * ctc_prefix_beam_search.cc:107-200), reached from transducer_attention_rescoring
 * (wenet/transducer/transducer.py:447-456).  logp [B,T,V] fp32 = CTC.log_softmax
 * (ctc.py:66-74); per frame: top-k (k = beam), blank / repeat / extend updates of
 * (pb, pnb) in fp64 with log_add (common.py:268-276), stable sort by log_add(pb,pnb),
 * keep beam.  One CTA per utterance, prefixes in a trie inside the workspace.
 * Outputs: out_tokens [B,beam,T] int32 (first out_len entries valid), out_len [B,beam]
 * (-1 = no hypothesis), out_score [B,beam] fp64, out_n [B]. */
size_t b200_ctc_prefix_beam_workspace_bytes(int B, int T, int beam);
int b200_ctc_prefix_beam(const float* logp, const int32_t* lens, int B, int T, int V,
                         int beam, int blank, int32_t* out_tokens, int32_t* out_len,
                         double* out_score, int32_t* out_n, void* workspace,
                         size_t ws_bytes, b200_stream_t stream);

/* ---- Batched CTC forced alignment -----------------------------------------------
 * replaces forced_align (wenet/utils/ctc_util.py:27-83, driven by wenet/bin/alignment.py):
 * Viterbi over the blank-interleaved labels, fp32 max/add only (bit-identical alignments,
 * the reference's negative-index wrap of state 0 included).  logp [B,T,V] fp32 log-probs,
 * ys_pad [B,Lmax] int64, hlens / ys_lens [B] (ys_lens >= 1); alignment [B,T] int32 = token
 * per frame, -1 beyond hlens.  workspace: back-pointers [B,T,2*Lmax+1] int16. */
size_t b200_ctc_forced_align_workspace_bytes(int B, int T, int Lmax);
int b200_ctc_forced_align(const float* logp, const int64_t* ys_pad, const int32_t* hlens,
                          const int32_t* ys_lens, int B, int T, int V, int Lmax, int blank,
                          int32_t* alignment, void* workspace, size_t ws_bytes,
                          b200_stream_t stream);

/* b200_ctc_forced_align without a caller workspace and without any allocation.  No counterpart in the
 * reference: added for callers that cannot call cudaMalloc (CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  The back-pointers live in a static
 * device array of b200_ctc_forced_align_static_max_bytes() = 32 MiB that every GPU loading the library
 * holds until exit; B*T*(2*Lmax+1)*2 bytes beyond that are rejected (B200_ERR_ARG).  All calls on a
 * device share the array: calls that overlap on the GPU race - serialise them or use the workspace form.
 * Nothing in the array has to start at zero (every back-pointer read was written by the same launch).
 * Same inputs, outputs and results as b200_ctc_forced_align. */
size_t b200_ctc_forced_align_static_max_bytes(void);
int b200_ctc_forced_align_static(const float* logp, const int64_t* ys_pad, const int32_t* hlens,
                                 const int32_t* ys_lens, int B, int T, int V, int Lmax, int blank,
                                 int32_t* alignment, b200_stream_t stream);

/* ---- Hot-word cross-entropy loss ------------------------------------------------
 * replaces self.hw_criterion = nn.CrossEntropyLoss() (wenet/transducer/transducer.py:66) as
 * applied at :170-200: logits [B, C, U] fp32 (class dimension 1), labels [B, U] int64, mean over
 * the positions with label != ignore_index.  loss / count: device scalars.  bwd: grad [B, C, U] =
 * dy / count * (softmax - onehot), zero at ignored positions. */
int b200_ce_loss_fwd(const float* logits, const int64_t* labels, int B, int C, int U,
                     long long ignore_index, float* loss, float* count, b200_stream_t stream);
int b200_ce_loss_bwd(const float* logits, const int64_t* labels, int B, int C, int U,
                     long long ignore_index, const float* dy, const float* count, float* grad,
                     b200_stream_t stream);

/* ---- n-best rescoring --------------------------------------------
 * The score-combine step of Transducer.transducer_attention_rescoring (transducer.py:459-513) over the
 * attention decoder's logits (_cal_attn_score, transducer.py:304-330): per hypothesis i
 *   attn_i = sum_j log_softmax(dec[i, j])[w_ij] + log_softmax(dec[i, len_i])[eos]   (and the right-to-left
 *   analogue on rdec, rows len_i - 1 - j, mixed with reverse_weight when it is > 0),
 *   scores[i] = attn_i * attn_weight + beam_score[i] * ctc_weight + td_score[i] * transducer_weight,
 *   best = first index of the maximum.
 * dec / rdec [N, L1, V] fp32 (L1 >= Lmax + 1), hyps_pad [N, Lmax] int64 (entries beyond hyps_lens ignored),
 * beam_score [N] fp64, td_score [N] fp32 (= -rnnt_loss per hypothesis), scores [N] fp64,
 * attn_scores optional [N] fp64 (the attention part alone), best [1] int32. */
int b200_attn_rescore(const float* dec, const float* rdec, const int64_t* hyps_pad,
                      const int32_t* hyps_lens, int N, int L1, int Lmax, int V, int eos,
                      double reverse_weight, double attn_weight, const double* beam_score,
                      double ctc_weight, const float* td_score, double transducer_weight,
                      double* scores, double* attn_scores, int32_t* best, b200_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_TRANSDUCER_H_ */
