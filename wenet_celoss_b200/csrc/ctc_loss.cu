// CTC loss over logits [B,T,V]: log-sum-exp pass, alpha/beta trellis, gradient pass.
//
// Replaces, behind CTC.forward (wenet/transformer/ctc.py:46-64):
//   ys_hat.log_softmax(2) (ctc.py:60) + torch.nn.CTCLoss(reduction='sum') (ctc.py:44,61)
// i.e. ATen's ctc_loss_log_alpha / log_beta / collect kernels.
// The ctc_lo projection (ctc.py:57) stays a library GEMM in the Python wrapper.
// Arithmetic restated from SURVEY.md §8(a) row a8 (Graves 2006 eq. 6-16), blank = 0 in
// the reference (passed explicitly here), zero_infinity=False:
//   l' = blank-interleaved labels, S = 2L+1
//   alpha_t(s) = lae(alpha_{t-1}(s), alpha_{t-1}(s-1), [l'_s != blank, l'_s != l'_{s-2}] alpha_{t-1}(s-2)) + lp[t,l'_s]
//   nll = -lae(alpha_{T-1}(S-1), alpha_{T-1}(S-2))
//   d nll / d x[t,c] = softmax[t,c] - sum_{s: l'_s = c} exp(alpha_t(s) + beta_t(s) + nll - lp[t,c])
//
// HBM layout: logits/grad [B,T,V] in the module's natural batch-major layout (no [T,B,V]
// transpose copy); lse [B,T]; lp_ext / alpha / beta [B,T,Smax] f32 with Smax = 2*Lmax+1;
// leader [B,Smax] int32 (first state carrying the same label, for duplicate labels);
// frame_off [B][2][T+1] fp64: alpha/beta are stored rescaled, alpha_t(s) = alpha_hat + off[b][0][t],
// beta likewise with off[b][1][t]; off[b][0][T] = log-likelihood (fp64).
// Roofline: passes 1 and 3 stream 3*B*T*V*s bytes (HBM-bound); the trellis is a T-step chain.
#include "common.cuh"
#include "row_ops.cuh"

namespace b200 {

__device__ __forceinline__ int ext_label(const long long* __restrict__ ys, int s, int blank) {
  return (s & 1) ? (int)ys[s >> 1] : blank;
}

// ---------------------------------------------------------------------------
// pass 1: one warp per (b,t) row: lse + gather of the S extended-label log-probs
// ---------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kRowWarps * 32, 4)
ctc_logprobs_kernel(const T* __restrict__ logits, const long long* __restrict__ ys_pad,
                    const int32_t* __restrict__ hlens, const int32_t* __restrict__ ys_lens, int B,
                    int Tm, int V, int Lmax, int blank, float* __restrict__ lse,
                    float* __restrict__ lp_ext) {
  const long long rows = (long long)B * Tm;
  const long long rowi = (long long)blockIdx.x * kRowWarps + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (rowi >= rows) return;
  const int b = int(rowi / Tm), t = int(rowi - (long long)b * Tm);
  if (t >= hlens[b]) return;
  const T* row = logits + rowi * V;
  const int S = 2 * ys_lens[b] + 1;
  const int Smax = 2 * Lmax + 1;
  const long long* ys = ys_pad + (long long)b * Lmax;
  // The label gather goes FIRST: issued before the row is streamed, its 64-byte DRAM fetches are
  // reused by the stream a few microseconds later (L2 hits).  Gathering after the stream re-fetched
  // ~100 x 64 B per row from DRAM (ncu: 1.73 GB read for 1.28 GB of logits).
  constexpr int kPre = 8;  // label states per lane held in registers: S <= 512 (odd states only)
  float xg[kPre];
#pragma unroll
  for (int i = 0; i < kPre; ++i) {
    const int s = 1 + 2 * (lane + 32 * i);
    xg[i] = -INFINITY;
    if (s < S) {
      const int c = (int)ys[s >> 1];
      if (c >= 0 && c < V) xg[i] = Elem<T>::to_float(row[c]);
    }
  }
  const float xb = Elem<T>::to_float(row[blank]);
  const float l = warp_row_lse<T>(row, V, lane);
  if (lane == 0) lse[rowi] = l;
  float* out = lp_ext + rowi * Smax;
  const float lpb = xb - l;
  for (int s = 2 * lane; s < S; s += 64) out[s] = lpb;  // blank states
#pragma unroll
  for (int i = 0; i < kPre; ++i) {
    const int s = 1 + 2 * (lane + 32 * i);
    if (s < S) out[s] = xg[i] - l;
  }
  for (int s = 1 + 2 * (lane + 32 * kPre); s < S; s += 64) {  // longer label sequences
    const int c = (int)ys[s >> 1];
    out[s] = (c >= 0 && c < V) ? Elem<T>::to_float(row[c]) - l : -INFINITY;
  }
}

// ---------------------------------------------------------------------------
// pass 2: trellis.  grid (B,2): y==0 alpha (+ leader table, nll), y==1 beta.
// One thread per extended state; previous frame in shared memory (ping-pong).
// ---------------------------------------------------------------------------
constexpr int kCtcPf = 4;

__device__ __forceinline__ float lae3(float a, float b, float c) {
  float m = fmaxf(a, fmaxf(b, c));
  if (m == -INFINITY) return -INFINITY;
  return m + logf(expf(a - m) + expf(b - m) + expf(c - m));
}

__global__ void ctc_lattice_kernel(const float* __restrict__ lp_ext,
                                   const long long* __restrict__ ys_pad,
                                   const int32_t* __restrict__ hlens,
                                   const int32_t* __restrict__ ys_lens, int Tm, int Lmax, int blank,
                                   float* __restrict__ alpha, float* __restrict__ beta,
                                   float* __restrict__ nll, int32_t* __restrict__ leader,
                                   double* __restrict__ frame_off) {
  extern __shared__ float sm[];  // [2][Smax+4] floats, [2][32] lagged-max scratch, then [Smax] ints
  const int b = blockIdx.x;
  const int s = threadIdx.x;
  const int Smax = 2 * Lmax + 1;
  const int Tb = hlens[b], L = ys_lens[b];
  const int S = 2 * L + 1;
  const int stride = Smax + 4;
  float* buf0 = sm + 2;  // indices -2 .. Smax+1 addressable
  float* buf1 = sm + stride + 2;
  float* red = sm + 2 * stride;
  int* lab = reinterpret_cast<int*>(sm + 2 * stride + 64);
  double* offs = frame_off + ((long long)b * 2 + blockIdx.y) * (Tm + 1);
  double run = 0.0;  // running offset: alpha_t(s) = alpha_hat_t(s) + off_a[t] (same for beta)
  const long long* ys = ys_pad + (long long)b * Lmax;
  for (int i = s; i < 2 * stride; i += blockDim.x) sm[i] = -INFINITY;
  for (int i = s; i < Smax; i += blockDim.x) lab[i] = (i < S) ? ext_label(ys, i, blank) : -1;
  __syncthreads();
  const bool mine = s < S;
  const int my = mine ? lab[s] : -1;
  const long long off = (long long)b * Tm * Smax;
  const float* lp = lp_ext + off;

  if (blockIdx.y == 0) {
    // leader[s] = smallest state index with the same label (blank states -> 0)
    if (mine) {
      int ld = s;
      if ((s & 1) == 0) {
        ld = 0;
      } else {
        for (int j = 1; j < s; j += 2)
          if (lab[j] == my) {
            ld = j;
            break;
          }
      }
      leader[(long long)b * Smax + s] = ld;
    }
    if (Tb <= 0) {
      if (s == 0) {
        nll[b] = (L == 0) ? 0.f : INFINITY;
        offs[Tm] = (L == 0) ? 0.0 : (double)-INFINITY;
      }
      return;
    }
    float* al = alpha + off;
    const bool skip_ok = mine && s >= 2 && my != blank && my != lab[s - 2];
    float* prev = buf0;
    float* cur = buf1;
    {
      float v = (mine && s <= 1) ? lp[s] : -INFINITY;
      if (mine) {
        prev[s] = v;
        al[s] = v;
      }
      lagged_max_write(v, red, 0);
    }
    int it = kLag;
    if (s == 0) offs[0] = 0.0;
    __syncthreads();
    float r[kCtcPf];
#pragma unroll
    for (int k = 0; k < kCtcPf; ++k) r[k] = (mine && 1 + k < Tb) ? lp[(long long)(1 + k) * Smax + s] : 0.f;
    for (int t0 = 1; t0 < Tb; t0 += kCtcPf) {
      float nr[kCtcPf];
#pragma unroll
      for (int k = 0; k < kCtcPf; ++k) {
        int t = t0 + kCtcPf + k;
        nr[k] = (mine && t < Tb) ? lp[(long long)t * Smax + s] : 0.f;
      }
#pragma unroll
      for (int k = 0; k < kCtcPf; ++k) {
        int t = t0 + k;
        if (t < Tb) {  // block-uniform
          float v = -INFINITY;
          if (mine) v = lae3(prev[s], prev[s - 1], skip_ok ? prev[s - 2] : -INFINITY) + r[k];
          {  // pull the previous frame's maximum into the fp64 offset (no extra barrier)
            // every kLag-th step: subtract the maximum published kLag steps ago
            if ((it & (kLag - 1)) == 0) {
              const float M = lagged_max_read(red, ((it / kLag) - 1) & 1);
              v -= M;
              run += (double)M;
              lagged_max_write(v, red, (it / kLag) & 1);
            }
            ++it;
          }
          if (mine) {
            cur[s] = v;
            al[(long long)t * Smax + s] = v;
          }
          if (s == 0) offs[t] = run;
          __syncthreads();
          float* tmp = prev;
          prev = cur;
          cur = tmp;
        }
      }
#pragma unroll
      for (int k = 0; k < kCtcPf; ++k) r[k] = nr[k];
    }
    if (s == 0) {
      double ll = (double)log_add_exp(prev[S - 1], S > 1 ? prev[S - 2] : -INFINITY) + run;
      nll[b] = (float)(-ll);
      offs[Tm] = ll;  // [b][0][T] = log-likelihood in fp64
    }
  } else {
    if (Tb <= 0) return;
    float* be = beta + off;
    const bool skip_ok = mine && s + 2 < S && lab[s + 2] != blank && lab[s + 2] != my;
    float* prev = buf0;
    float* cur = buf1;
    {
      float v = (mine && s >= S - 2) ? lp[(long long)(Tb - 1) * Smax + s] : -INFINITY;
      if (mine) {
        prev[s] = v;
        be[(long long)(Tb - 1) * Smax + s] = v;
      }
      lagged_max_write(v, red, 0);
    }
    int it = kLag;
    if (s == 0) offs[Tb - 1] = 0.0;
    __syncthreads();
    float r[kCtcPf];
#pragma unroll
    for (int k = 0; k < kCtcPf; ++k) {
      int t = Tb - 2 - k;
      r[k] = (mine && t >= 0) ? lp[(long long)t * Smax + s] : 0.f;
    }
    for (int t0 = Tb - 2; t0 >= 0; t0 -= kCtcPf) {
      float nr[kCtcPf];
#pragma unroll
      for (int k = 0; k < kCtcPf; ++k) {
        int t = t0 - kCtcPf - k;
        nr[k] = (mine && t >= 0) ? lp[(long long)t * Smax + s] : 0.f;
      }
#pragma unroll
      for (int k = 0; k < kCtcPf; ++k) {
        int t = t0 - k;
        if (t >= 0) {  // block-uniform
          float v = -INFINITY;
          if (mine) v = lae3(prev[s], prev[s + 1], skip_ok ? prev[s + 2] : -INFINITY) + r[k];
          {
            // every kLag-th step: subtract the maximum published kLag steps ago
            if ((it & (kLag - 1)) == 0) {
              const float M = lagged_max_read(red, ((it / kLag) - 1) & 1);
              v -= M;
              run += (double)M;
              lagged_max_write(v, red, (it / kLag) & 1);
            }
            ++it;
          }
          if (mine) {
            cur[s] = v;
            be[(long long)t * Smax + s] = v;
          }
          if (s == 0) offs[t] = run;
          __syncthreads();
          float* tmp = prev;
          prev = cur;
          cur = tmp;
        }
      }
#pragma unroll
      for (int k = 0; k < kCtcPf; ++k) r[k] = nr[k];
    }
  }
}

// ---------------------------------------------------------------------------
// pass 3: gradient w.r.t. logits, one warp per (b,t) row
//   grad[c] = scale_b * ( exp(x[c]-lse) - occ[c] ),  0 for t >= T_b
// ---------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kRowWarps * 32)
ctc_grad_kernel(const T* __restrict__ logits, const float* __restrict__ lse,
                const float* __restrict__ lp_ext, const float* __restrict__ alpha,
                const float* __restrict__ beta, const double* __restrict__ frame_off,
                const float* __restrict__ scale, const long long* __restrict__ ys_pad,
                const int32_t* __restrict__ hlens, const int32_t* __restrict__ ys_lens,
                const int32_t* __restrict__ leader, int B, int Tm, int V, int Lmax, int blank,
                T* __restrict__ grad) {
  extern __shared__ float occ_all[];  // [kRowWarps][Smax]
  const long long rows = (long long)B * Tm;
  const long long rowi = (long long)blockIdx.x * kRowWarps + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (rowi >= rows) return;
  const int Smax = 2 * Lmax + 1;
  float* occ = occ_all + (threadIdx.x >> 5) * Smax;
  const int b = int(rowi / Tm), t = int(rowi - (long long)b * Tm);
  const T* row = logits + rowi * V;
  T* grow = grad + rowi * V;
  const float sc = scale[b];
  const bool dead = (t >= hlens[b]) || (sc == 0.f);
  if (dead) {
    warp_row_softmax_store<T>(row, grow, V, lane, true, 0.f, 0.f, -1.f);
    return;
  }
  const int S = 2 * ys_lens[b] + 1;
  const float l = lse[rowi];
  // alpha_hat + beta_hat + (off_a[t] + off_b[t] - ll): the bracket is O(1) because
  // sum_s alpha_t(s) beta_t(s) / y_t(l'_s) = exp(ll) for every t
  const double* oa = frame_off + (long long)b * 2 * (Tm + 1);
  const float nl = (float)(oa[t] + oa[Tm + 1 + t] - oa[Tm]);
  const float* al = alpha + rowi * Smax;
  const float* be = beta + rowi * Smax;
  const float* lp = lp_ext + rowi * Smax;
  const int32_t* ld = leader + (long long)b * Smax;
  // logits at the label classes, gathered BEFORE the row is streamed (see ctc_logprobs_kernel)
  const long long* ys = ys_pad + (long long)b * Lmax;
  constexpr int kPre = 8;
  float xg[kPre];
#pragma unroll
  for (int i = 0; i < kPre; ++i) {
    const int s = 1 + 2 * (lane + 32 * i);
    xg[i] = 0.f;
    if (s < S && ld[s] == s) {
      const int c = (int)ys[s >> 1];
      if (c >= 0 && c < V) xg[i] = Elem<T>::to_float(row[c]);
    }
  }
  const float xblank = Elem<T>::to_float(row[blank]);
  // the first kOcc * 32 states: all four loads of every state issued before the first use (a
  // rolled loop with the shared-memory atomic in its body paid one memory round trip per 32 states)
  constexpr int kOcc = 8;
  float pa[kOcc], pb[kOcc], pl[kOcc];
  int pd[kOcc];
#pragma unroll
  for (int i = 0; i < kOcc; ++i) {
    const int s = lane + 32 * i;
    pa[i] = pb[i] = pl[i] = 0.f;
    pd[i] = 0;
    if (s < S) {
      pa[i] = al[s];
      pb[i] = be[s];
      pl[i] = lp[s];
      pd[i] = ld[s];
    }
  }
  for (int s = lane; s < S; s += 32) occ[s] = 0.f;
  __syncwarp();
  float occ_blank = 0.f;
#pragma unroll
  for (int i = 0; i < kOcc; ++i) {
    const int s = lane + 32 * i;
    if (s >= S) continue;
    // alpha*beta counts the emission at t twice: take lp out once
    const float post = expf(pa[i] + pb[i] + nl - pl[i]);
    if (s & 1)
      atomicAdd(&occ[pd[i]], post);
    else
      occ_blank += post;
  }
  for (int s = lane + 32 * kOcc; s < S; s += 32) {  // longer label sequences
    float post = expf(al[s] + be[s] + nl - lp[s]);
    if (s & 1)
      atomicAdd(&occ[ld[s]], post);
    else
      occ_blank += post;
  }
  occ_blank = warp_sum(occ_blank);
  warp_row_softmax_store<T>(row, grow, V, lane, false, -l * kLog2e, sc, -1.f);
  __syncwarp();  // streaming stores and occ[] atomics are ordered before the fix-ups
  if (lane == 0) grow[blank] = Elem<T>::from_float(sc * (expf(xblank - l) - occ_blank));
#pragma unroll
  for (int i = 0; i < kPre; ++i) {
    const int s = 1 + 2 * (lane + 32 * i);
    if (s >= S || ld[s] != s) continue;  // only the first state of each distinct label writes
    const int c = (int)ys[s >> 1];
    if (c < 0 || c >= V || c == blank) continue;
    grow[c] = Elem<T>::from_float(sc * (expf(xg[i] - l) - occ[s]));
  }
  for (int s = 1 + 2 * (lane + 32 * kPre); s < S; s += 64) {  // longer label sequences
    if (ld[s] != s) continue;
    const int c = (int)ys[s >> 1];
    if (c < 0 || c >= V || c == blank) continue;
    const float x = Elem<T>::to_float(row[c]);
    grow[c] = Elem<T>::from_float(sc * (expf(x - l) - occ[s]));
  }
}

}  // namespace b200

// ===========================================================================
// C ABI
// ===========================================================================
using namespace b200;

static int ctc_check(const char* fn, int B, int T, int V, int Lmax, int blank) {
  B200_CHECK_ARG(B > 0 && T > 0 && V > 0 && Lmax >= 0, "%s: bad shape B=%d T=%d V=%d Lmax=%d", fn, B,
                 T, V, Lmax);
  B200_CHECK_ARG(blank >= 0 && blank < V, "%s: blank %d out of range", fn, blank);
  B200_CHECK_ARG(2 * Lmax + 1 <= 1024, "%s: 2*Lmax+1 = %d > 1024 states not supported", fn,
                 2 * Lmax + 1);
  return 0;
}

extern "C" int b200_ctc_logprobs(const void* logits, int dtype, const int64_t* ys_pad,
                                 const int32_t* hlens, const int32_t* ys_lens, int B, int T, int V,
                                 int Lmax, int blank, float* lse, float* lp_ext,
                                 b200_stream_t stream) {
  if (int rc = ctc_check("b200_ctc_logprobs", B, T, V, Lmax, blank)) return rc;
  B200_CHECK_ARG(logits && hlens && ys_lens && lse && lp_ext && (Lmax == 0 || ys_pad),
                 "b200_ctc_logprobs: null pointer");
  const long long rows = (long long)B * T;
  const unsigned grid = (unsigned)((rows + kRowWarps - 1) / kRowWarps);
  cudaStream_t st = (cudaStream_t)stream;
  const long long* ys = reinterpret_cast<const long long*>(ys_pad);
  switch (dtype) {
    case B200_F32:
      ctc_logprobs_kernel<float><<<grid, kRowWarps * 32, 0, st>>>(
          (const float*)logits, ys, hlens, ys_lens, B, T, V, Lmax, blank, lse, lp_ext);
      break;
    case B200_F16:
      ctc_logprobs_kernel<__half><<<grid, kRowWarps * 32, 0, st>>>(
          (const __half*)logits, ys, hlens, ys_lens, B, T, V, Lmax, blank, lse, lp_ext);
      break;
    case B200_BF16:
      ctc_logprobs_kernel<__nv_bfloat16><<<grid, kRowWarps * 32, 0, st>>>(
          (const __nv_bfloat16*)logits, ys, hlens, ys_lens, B, T, V, Lmax, blank, lse, lp_ext);
      break;
    default:
      B200_CHECK_ARG(false, "b200_ctc_logprobs: unsupported dtype %d", dtype);
  }
  B200_CHECK_LAUNCH("ctc_logprobs_kernel");
  return B200_OK;
}

extern "C" int b200_ctc_lattice(const float* lp_ext, const int64_t* ys_pad, const int32_t* hlens,
                                const int32_t* ys_lens, int B, int T, int Lmax, int blank,
                                float* alpha, float* beta, float* nll, int32_t* leader,
                                double* frame_off, b200_stream_t stream) {
  if (int rc = ctc_check("b200_ctc_lattice", B, T, blank + 1, Lmax, blank)) return rc;
  B200_CHECK_ARG(lp_ext && hlens && ys_lens && alpha && beta && nll && leader && frame_off &&
                     (Lmax == 0 || ys_pad),
                 "b200_ctc_lattice: null pointer");
  const int Smax = 2 * Lmax + 1;
  const int threads = ((Smax + 31) / 32) * 32;
  const size_t smem = (2 * (size_t)(Smax + 4) + 64) * sizeof(float) + (size_t)Smax * sizeof(int);
  ctc_lattice_kernel<<<dim3(B, 2), threads, smem, (cudaStream_t)stream>>>(
      lp_ext, reinterpret_cast<const long long*>(ys_pad), hlens, ys_lens, T, Lmax, blank, alpha,
      beta, nll, leader, frame_off);
  B200_CHECK_LAUNCH("ctc_lattice_kernel");
  return B200_OK;
}

extern "C" int b200_ctc_grad(const void* logits, int dtype, const float* lse, const float* lp_ext,
                             const float* alpha, const float* beta, const double* frame_off,
                             const float* scale, const int64_t* ys_pad, const int32_t* hlens,
                             const int32_t* ys_lens, const int32_t* leader, int B, int T, int V,
                             int Lmax, int blank, void* grad, b200_stream_t stream) {
  if (int rc = ctc_check("b200_ctc_grad", B, T, V, Lmax, blank)) return rc;
  B200_CHECK_ARG(logits && lse && lp_ext && alpha && beta && frame_off && scale && hlens && ys_lens &&
                     leader && grad && (Lmax == 0 || ys_pad),
                 "b200_ctc_grad: null pointer");
  B200_CHECK_ARG(((reinterpret_cast<uintptr_t>(logits) ^ reinterpret_cast<uintptr_t>(grad)) & 15) == 0,
                 "b200_ctc_grad: logits and grad must share 16-byte alignment phase");
  const long long rows = (long long)B * T;
  const unsigned grid = (unsigned)((rows + kRowWarps - 1) / kRowWarps);
  const int Smax = 2 * Lmax + 1;
  const size_t smem = (size_t)kRowWarps * Smax * sizeof(float);
  cudaStream_t st = (cudaStream_t)stream;
  const long long* ys = reinterpret_cast<const long long*>(ys_pad);
  switch (dtype) {
    case B200_F32:
      ctc_grad_kernel<float><<<grid, kRowWarps * 32, smem, st>>>(
          (const float*)logits, lse, lp_ext, alpha, beta, frame_off, scale, ys, hlens, ys_lens, leader, B,
          T, V, Lmax, blank, (float*)grad);
      break;
    case B200_F16:
      ctc_grad_kernel<__half><<<grid, kRowWarps * 32, smem, st>>>(
          (const __half*)logits, lse, lp_ext, alpha, beta, frame_off, scale, ys, hlens, ys_lens, leader, B,
          T, V, Lmax, blank, (__half*)grad);
      break;
    case B200_BF16:
      ctc_grad_kernel<__nv_bfloat16><<<grid, kRowWarps * 32, smem, st>>>(
          (const __nv_bfloat16*)logits, lse, lp_ext, alpha, beta, frame_off, scale, ys, hlens, ys_lens,
          leader, B, T, V, Lmax, blank, (__nv_bfloat16*)grad);
      break;
    default:
      B200_CHECK_ARG(false, "b200_ctc_grad: unsupported dtype %d", dtype);
  }
  B200_CHECK_LAUNCH("ctc_grad_kernel");
  return B200_OK;
}
