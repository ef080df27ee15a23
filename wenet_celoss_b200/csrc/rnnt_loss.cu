// RNN-T loss over materialised logits: three HBM-streaming passes + the lattice DP.
//
// Replaces torchaudio.functional.rnnt_loss at the reference call sites
//   wenet/transducer/transducer.py:142-147 and :296-301
// Arithmetic restated from SURVEY.md §8(a) row a6 (Graves 2012, §2.4-2.5):
//   lp[t,u,v] = x[t,u,v] - lse[t,u]
//   alpha[t,u] = lae(alpha[t-1,u] + lp[t-1,u,blank], alpha[t,u-1] + lp[t,u-1,y[u-1]])
//   beta [t,u] = lae(beta [t+1,u] + lp[t,u,blank],   beta [t,u+1] + lp[t,u,y[u]])
//   cost = -beta[0,0]
//
// Data layout in HBM: logits/grad [B,T,U1,V] row = one lattice cell (V contiguous);
// lse / lp_blank / lp_emit / alpha / beta [B,T,U1] f32; coef [B,T,U1] float4.
//
// Roofline: pass 1 reads N*s bytes, pass 3 reads N*s and writes N*s (N = B*T*U1*V,
// s = sizeof(logit)) -> HBM-bound; the lattice pass is a T+U-1 step dependency chain.
#include "common.cuh"
#include "row_ops.cuh"

namespace b200 {


// ---------------------------------------------------------------------------
// pass 1: one warp per lattice cell; online (max, sum-exp) over the V logits.
// ---------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(kRowWarps * 32, 4)
rnnt_logprobs_kernel(const T* __restrict__ logits, const int32_t* __restrict__ targets,
                     const int32_t* __restrict__ logit_len, const int32_t* __restrict__ target_len,
                     int B, int Tm, int U1, int V, int blank, float* __restrict__ lse,
                     float* __restrict__ lp_blank, float* __restrict__ lp_emit) {
  const long long cells = (long long)B * Tm * U1;
  const long long cell = (long long)blockIdx.x * kRowWarps + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (cell >= cells) return;
  const int b = int(cell / ((long long)Tm * U1));
  const int r = int(cell - (long long)b * Tm * U1);
  const int t = r / U1, u = r - t * U1;
  const int Tb = logit_len[b], Ub = target_len[b];
  if (t >= Tb || u > Ub) return;  // padded cell: never read

  const T* row = logits + cell * V;
  const float l = warp_row_lse<T>(row, V, lane);
  if (lane == 0) {
    lse[cell] = l;
    lp_blank[cell] = Elem<T>::to_float(row[blank]) - l;
  } else if (lane == 1) {
    if (u < Ub) {
      int y = targets[(long long)b * (U1 - 1) + u];
      lp_emit[cell] = Elem<T>::to_float(row[y]) - l;
    }
  }
}

// ---------------------------------------------------------------------------
// pass 2: anti-diagonal wavefront.  grid (B, 2): y==0 alpha, y==1 beta.
// One thread per u; the previous diagonal lives in shared memory (ping-pong),
// one __syncthreads per diagonal; log-prob loads run kPf diagonals ahead.
// The diagonal step is the serial chain of the loss (T + U dependent steps): values are kept in log2
// units so that the step is LDS -> 2 FADD -> FMNMX x2 -> ex2 -> lg2 -> STS -> barrier (2 MUFU, no
// expf / logf range code on the chain); alpha / beta leave in natural-log units.
//
// Rescaling: every path crosses each anti-diagonal exactly once, so a per-diagonal
// offset can be pulled out of alpha and beta without changing any posterior.  Every kPf = 8
// diagonals the block subtracts the maximum it published 8 diagonals earlier (one redux.sync per
// warp into shared memory, so no extra barrier) and adds it to a running fp64 offset:  alpha(t,u) = alpha_hat(t,u) + off_a[t+u]  (same for beta).
// The stored fp32 values stay O(10) in magnitude, which removes the cost-sized
// cancellation alpha+beta-cost that limits plain fp32 log-space to ~cost*2^-23.
// diag_off layout: [B][2][T+U1] fp64; [b][0][d] alpha offsets, [b][1][d] beta offsets,
// [b][0][T+U1-1] = log-likelihood of utterance b.
// ---------------------------------------------------------------------------
constexpr int kPf = 8;  // diagonals of log-probs in registers ahead of the chain == diagonals between rescalings

__global__ void rnnt_lattice_kernel(const float* __restrict__ lp_blank,
                                    const float* __restrict__ lp_emit,
                                    const int32_t* __restrict__ logit_len,
                                    const int32_t* __restrict__ target_len, int Tm, int U1,
                                    float* __restrict__ alpha, float* __restrict__ beta,
                                    float* __restrict__ costs, double* __restrict__ diag_off) {
  extern __shared__ float diag[];  // [2][U1 + 2] + [2][32] (lagged-max scratch)
  const int b = blockIdx.x;
  const int u = threadIdx.x;
  const int Tb = logit_len[b], Ub = target_len[b];
  const long long off = (long long)b * Tm * U1;
  const float* lpb = lp_blank + off;
  const float* lpe = lp_emit + off;
  const int stride = U1 + 2;
  const int nd = Tm + U1;  // slots per offset row
  float* buf0 = diag + 1;  // index -1 .. U1 addressable
  float* buf1 = diag + stride + 1;
  float* red = diag + 2 * stride;
  double* offs = diag_off + ((long long)b * 2 + blockIdx.y) * nd;
  if (Tb <= 0 || Ub < 0 || Ub >= U1 || Tb > Tm) {
    if (u == 0 && blockIdx.y == 1) {
      costs[b] = 0.f;
      diag_off[(long long)b * 2 * nd + nd - 1] = 0.0;
    }
    return;
  }
  const int D = Tb + Ub;  // number of anti-diagonals (d = 0 .. D-1)
  for (int i = u; i < stride; i += blockDim.x) {
    diag[i] = -INFINITY;
    diag[stride + i] = -INFINITY;
  }
  __syncthreads();
  const bool mine = (u <= Ub);
  double run = 0.0;  // running offset (log2 units inside the loop), tracked identically by every thread
  constexpr double kLn2d = 0.6931471805599453094;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;

  if (blockIdx.y == 0) {
    // ---------------- alpha: d ascending ----------------
    float* al = alpha + off;
    if (u == 0) {
      buf0[0] = 0.f;
      al[0] = 0.f;
      offs[0] = 0.0;
    }
    {
      const float wm = warp_max_redux(u == 0 ? 0.f : -INFINITY);
      if (lane == 0) red[warp] = wm;
    }
    __syncthreads();
    float* prev = buf0;
    float* cur = buf1;
    float rb[kPf], re[kPf];
    auto fetch = [&](int d, float& vb, float& ve) {  // log2 units
      vb = -INFINITY;
      ve = -INFINITY;
      int t = d - u;
      if (mine && d < D && t >= 0 && t < Tb) {
        if (t > 0) vb = lpb[(long long)(t - 1) * U1 + u] * kLog2e;
        if (u > 0) ve = lpe[(long long)t * U1 + u - 1] * kLog2e;
      }
    };
#pragma unroll
    for (int k = 0; k < kPf; ++k) fetch(1 + k, rb[k], re[k]);
    int blk = 0;
    for (int d0 = 1; d0 < D; d0 += kPf, ++blk) {
      float nb[kPf], ne[kPf];
#pragma unroll
      for (int k = 0; k < kPf; ++k) fetch(d0 + kPf + k, nb[k], ne[k]);
      const float M = block_max_read(red, blk & 1, nw);  // published one block (kPf diagonals) ago
#pragma unroll
      for (int k = 0; k < kPf; ++k) {
        int d = d0 + k;
        if (d < D) {  // block-uniform
          int t = d - u;
          const bool live = mine && t >= 0 && t < Tb;
          float v = live ? lae2_log2(prev[u] + rb[k], prev[u - 1] + re[k]) : -INFINITY;
          if (k == 0) {  // first diagonal of a block: subtract the maximum published kPf diagonals ago
            v -= M;
            run += (double)M;
            const float wm = warp_max_redux(v);
            if (lane == 0) red[((blk & 1) ^ 1) * 32 + warp] = wm;
          }
          if (mine) cur[u] = v;
          if (live) al[(long long)t * U1 + u] = v * kLn2;
          if (u == 0) offs[d] = run * kLn2d;
          __syncthreads();
          float* tmp = prev;
          prev = cur;
          cur = tmp;
        }
      }
#pragma unroll
      for (int k = 0; k < kPf; ++k) {
        rb[k] = nb[k];
        re[k] = ne[k];
      }
    }
  } else {
    // ---------------- beta: d descending ----------------
    float* be = beta + off;
    float* prev = buf0;
    float* cur = buf1;
    if (u == Ub) {
      float v = lpb[(long long)(Tb - 1) * U1 + Ub];
      buf0[Ub] = v * kLog2e;
      be[(long long)(Tb - 1) * U1 + Ub] = v;
      offs[D - 1] = 0.0;
      if (D == 1) {
        costs[b] = -v;
        diag_off[(long long)b * 2 * nd + nd - 1] = (double)v;
      }
    }
    {
      const float wm = warp_max_redux(u == Ub ? lpb[(long long)(Tb - 1) * U1 + Ub] * kLog2e : -INFINITY);
      if (lane == 0) red[warp] = wm;
    }
    __syncthreads();
    float rb[kPf], re[kPf];
    auto fetch = [&](int d, float& vb, float& ve) {  // log2 units
      vb = -INFINITY;
      ve = -INFINITY;
      int t = d - u;
      if (mine && d >= 0 && t >= 0 && t < Tb) {
        if (t + 1 < Tb) vb = lpb[(long long)t * U1 + u] * kLog2e;
        if (u < Ub) ve = lpe[(long long)t * U1 + u] * kLog2e;
      }
    };
#pragma unroll
    for (int k = 0; k < kPf; ++k) fetch(D - 2 - k, rb[k], re[k]);
    int blk = 0;
    for (int d0 = D - 2; d0 >= 0; d0 -= kPf, ++blk) {
      float nb[kPf], ne[kPf];
#pragma unroll
      for (int k = 0; k < kPf; ++k) fetch(d0 - kPf - k, nb[k], ne[k]);
      const float M = block_max_read(red, blk & 1, nw);
#pragma unroll
      for (int k = 0; k < kPf; ++k) {
        int d = d0 - k;
        if (d >= 0) {  // block-uniform
          int t = d - u;
          const bool live = mine && t >= 0 && t < Tb;
          float v = live ? lae2_log2(prev[u] + rb[k], prev[u + 1] + re[k]) : -INFINITY;
          if (k == 0) {
            v -= M;
            run += (double)M;
            const float wm = warp_max_redux(v);
            if (lane == 0) red[((blk & 1) ^ 1) * 32 + warp] = wm;
          }
          if (mine) cur[u] = v;
          if (live) {
            be[(long long)t * U1 + u] = v * kLn2;
            if (d == 0) {
              double ll = ((double)v + run) * kLn2d;
              costs[b] = (float)(-ll);
              diag_off[(long long)b * 2 * nd + nd - 1] = ll;
            }
          }
          if (u == 0) offs[d] = run * kLn2d;
          __syncthreads();
          float* tmp = prev;
          prev = cur;
          cur = tmp;
        }
      }
#pragma unroll
      for (int k = 0; k < kPf; ++k) {
        rb[k] = nb[k];
        re[k] = ne[k];
      }
    }
  }
}

// ---------------------------------------------------------------------------
// pass 2b: gradient coefficients per cell (tiny, one thread per cell)
// ---------------------------------------------------------------------------
__global__ void rnnt_grad_coef_kernel(const float* __restrict__ lse, const float* __restrict__ alpha,
                                      const float* __restrict__ beta,
                                      const double* __restrict__ diag_off,
                                      const float* __restrict__ dy,
                                      const int32_t* __restrict__ targets,
                                      const int32_t* __restrict__ logit_len,
                                      const int32_t* __restrict__ target_len, int B, int Tm, int U1,
                                      int blank, float4* __restrict__ coef) {
  const long long cells = (long long)B * Tm * U1;
  const long long cell = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (cell >= cells) return;
  const int b = int(cell / ((long long)Tm * U1));
  const int r = int(cell - (long long)b * Tm * U1);
  const int t = r / U1, u = r - t * U1;
  const int Tb = logit_len[b], Ub = target_len[b];
  float4 c = make_float4(-INFINITY, -INFINITY, -INFINITY, 0.f);
  if (t < Tb && u <= Ub) {
    const int nd = Tm + U1;
    const double* oa = diag_off + (long long)b * 2 * nd;
    const double* ob = oa + nd;
    const double ll = oa[nd - 1];
    const int d = t + u;
    // gradient terms are exp(x - lse + alpha + beta_* - ll); offsets combine in fp64,
    // the per-diagonal sums are O(1) because sum over a diagonal of alpha*beta = exp(ll)
    const float base = alpha[cell] - lse[cell];
    const float k0 = (float)(oa[d] + ob[d] - ll);
    c.x = base + beta[cell] + k0;
    const bool last_t = (t == Tb - 1);
    if (last_t && u == Ub)
      c.y = base + (float)(oa[d] - ll);
    else if (!last_t)
      c.y = base + beta[cell + U1] + (float)(oa[d] + ob[d + 1] - ll);
    if (u < Ub) {
      int y = targets[(long long)b * (U1 - 1) + u];
      // if/else-if order of the reference: a target equal to blank only gets the
      // emit term when no blank term applied (t == Tb-1, u < Ub).
      if (!(y == blank && !last_t)) c.z = base + beta[cell + 1] + (float)(oa[d] + ob[d + 1] - ll);
    }
    c.w = dy[b];
  }
  coef[cell] = c;
}

// ---------------------------------------------------------------------------
// pass 3: gradient w.r.t. logits; softmax recomputed from logits + coef
// ---------------------------------------------------------------------------
__device__ __forceinline__ float clampf(float g, float c) {
  return c > 0.f ? fminf(fmaxf(g, -c), c) : g;
}

template <typename T>
__global__ void __launch_bounds__(kRowWarps * 32)
rnnt_grad_kernel(const T* __restrict__ logits, const int32_t* __restrict__ targets,
                 const float4* __restrict__ coef, int B, int Tm, int U1, int V, int blank,
                 float clamp, T* __restrict__ grad) {
  const long long cells = (long long)B * Tm * U1;
  const long long cell = (long long)blockIdx.x * kRowWarps + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (cell >= cells) return;
  const float4 c = coef[cell];
  const T* row = logits + cell * V;
  T* grow = grad + cell * V;
  // grad has the same element offset as logits -> same 16-byte phase when the two
  // base pointers share alignment (checked on the host).
  const bool dead = (c.x == -INFINITY) || (c.w == 0.f);
  const float scale = c.w;
  warp_row_softmax_store<T>(row, grow, V, lane, dead, c.x * kLog2e, scale, clamp);
  if (dead) return;
  __syncwarp();  // order the streaming stores before the two fix-ups below
  if (lane == 0) {
    if (c.y != -INFINITY) {
      float x = Elem<T>::to_float(row[blank]);
      grow[blank] = Elem<T>::from_float(clampf(scale * (expf(x + c.x) - expf(x + c.y)), clamp));
    }
  } else if (lane == 1) {
    if (c.z != -INFINITY) {
      const int b = int(cell / ((long long)Tm * U1));
      const int r = int(cell - (long long)b * Tm * U1);
      const int u = r % U1;
      int y = targets[(long long)b * (U1 - 1) + u];
      float x = Elem<T>::to_float(row[y]);
      float sub = expf(x + c.z);
      // y == blank with both terms live cannot happen (see coef kernel)
      grow[y] = Elem<T>::from_float(clampf(scale * (expf(x + c.x) - sub), clamp));
    }
  }
}

}  // namespace b200

// ===========================================================================
// C ABI
// ===========================================================================
using namespace b200;

static int check_shape(const char* fn, int B, int T, int U1, int V) {
  B200_CHECK_ARG(B > 0 && T > 0 && U1 > 0 && V > 0, "%s: bad shape B=%d T=%d U1=%d V=%d", fn, B, T,
                 U1, V);
  return 0;
}

extern "C" int b200_rnnt_logprobs(const void* logits, int dtype, const int32_t* targets,
                                  const int32_t* logit_len, const int32_t* target_len, int B, int T,
                                  int U1, int V, int blank, float* lse, float* lp_blank,
                                  float* lp_emit, b200_stream_t stream) {
  if (int rc = check_shape("b200_rnnt_logprobs", B, T, U1, V)) return rc;
  B200_CHECK_ARG(logits && logit_len && target_len && lse && lp_blank && lp_emit,
                 "b200_rnnt_logprobs: null pointer");
  B200_CHECK_ARG(U1 == 1 || targets, "b200_rnnt_logprobs: null targets");
  B200_CHECK_ARG(blank >= 0 && blank < V, "b200_rnnt_logprobs: blank %d out of range", blank);
  const long long cells = (long long)B * T * U1;
  const unsigned grid = (unsigned)((cells + kRowWarps - 1) / kRowWarps);
  cudaStream_t st = (cudaStream_t)stream;
  switch (dtype) {
    case B200_F32:
      rnnt_logprobs_kernel<float><<<grid, kRowWarps * 32, 0, st>>>(
          (const float*)logits, targets, logit_len, target_len, B, T, U1, V, blank, lse, lp_blank,
          lp_emit);
      break;
    case B200_F16:
      rnnt_logprobs_kernel<__half><<<grid, kRowWarps * 32, 0, st>>>(
          (const __half*)logits, targets, logit_len, target_len, B, T, U1, V, blank, lse, lp_blank,
          lp_emit);
      break;
    case B200_BF16:
      rnnt_logprobs_kernel<__nv_bfloat16><<<grid, kRowWarps * 32, 0, st>>>(
          (const __nv_bfloat16*)logits, targets, logit_len, target_len, B, T, U1, V, blank, lse,
          lp_blank, lp_emit);
      break;
    default:
      B200_CHECK_ARG(false, "b200_rnnt_logprobs: unsupported dtype %d", dtype);
  }
  B200_CHECK_LAUNCH("rnnt_logprobs_kernel");
  return B200_OK;
}

extern "C" int b200_rnnt_lattice(const float* lp_blank, const float* lp_emit,
                                 const int32_t* logit_len, const int32_t* target_len, int B, int T,
                                 int U1, float* alpha, float* beta, float* costs, double* diag_off,
                                 b200_stream_t stream) {
  if (int rc = check_shape("b200_rnnt_lattice", B, T, U1, 1)) return rc;
  B200_CHECK_ARG(lp_blank && lp_emit && logit_len && target_len && alpha && beta && costs &&
                     diag_off,
                 "b200_rnnt_lattice: null pointer");
  B200_CHECK_ARG(U1 <= 1024, "b200_rnnt_lattice: U+1 = %d > 1024 not supported", U1);
  const int threads = ((U1 + 31) / 32) * 32;
  const size_t smem = (2 * (size_t)(U1 + 2) + 64) * sizeof(float);
  rnnt_lattice_kernel<<<dim3(B, 2), threads, smem, (cudaStream_t)stream>>>(
      lp_blank, lp_emit, logit_len, target_len, T, U1, alpha, beta, costs, diag_off);
  B200_CHECK_LAUNCH("rnnt_lattice_kernel");
  return B200_OK;
}

extern "C" int b200_rnnt_grad_coef(const float* lse, const float* alpha, const float* beta,
                                   const double* diag_off, const float* dy, const int32_t* targets,
                                   const int32_t* logit_len, const int32_t* target_len, int B, int T,
                                   int U1, int blank, float* coef, b200_stream_t stream) {
  if (int rc = check_shape("b200_rnnt_grad_coef", B, T, U1, 1)) return rc;
  B200_CHECK_ARG(lse && alpha && beta && diag_off && dy && logit_len && target_len && coef,
                 "b200_rnnt_grad_coef: null pointer");
  B200_CHECK_ARG(U1 == 1 || targets, "b200_rnnt_grad_coef: null targets");
  B200_CHECK_ARG((reinterpret_cast<uintptr_t>(coef) & 15) == 0,
                 "b200_rnnt_grad_coef: coef must be 16-byte aligned");
  const long long cells = (long long)B * T * U1;
  const unsigned grid = (unsigned)((cells + 255) / 256);
  rnnt_grad_coef_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
      lse, alpha, beta, diag_off, dy, targets, logit_len, target_len, B, T, U1, blank,
      reinterpret_cast<float4*>(coef));
  B200_CHECK_LAUNCH("rnnt_grad_coef_kernel");
  return B200_OK;
}

extern "C" int b200_rnnt_grad(const void* logits, int dtype, const int32_t* targets,
                              const float* coef, int B, int T, int U1, int V, int blank,
                              float clamp, void* grad, b200_stream_t stream) {
  if (int rc = check_shape("b200_rnnt_grad", B, T, U1, V)) return rc;
  B200_CHECK_ARG(logits && coef && grad, "b200_rnnt_grad: null pointer");
  B200_CHECK_ARG(U1 == 1 || targets, "b200_rnnt_grad: null targets");
  B200_CHECK_ARG(blank >= 0 && blank < V, "b200_rnnt_grad: blank %d out of range", blank);
  B200_CHECK_ARG(((reinterpret_cast<uintptr_t>(logits) ^ reinterpret_cast<uintptr_t>(grad)) & 15) == 0,
                 "b200_rnnt_grad: logits and grad must share 16-byte alignment phase");
  const long long cells = (long long)B * T * U1;
  const unsigned grid = (unsigned)((cells + kRowWarps - 1) / kRowWarps);
  cudaStream_t st = (cudaStream_t)stream;
  const float4* cf = reinterpret_cast<const float4*>(coef);
  switch (dtype) {
    case B200_F32:
      rnnt_grad_kernel<float><<<grid, kRowWarps * 32, 0, st>>>(
          (const float*)logits, targets, cf, B, T, U1, V, blank, clamp, (float*)grad);
      break;
    case B200_F16:
      rnnt_grad_kernel<__half><<<grid, kRowWarps * 32, 0, st>>>(
          (const __half*)logits, targets, cf, B, T, U1, V, blank, clamp, (__half*)grad);
      break;
    case B200_BF16:
      rnnt_grad_kernel<__nv_bfloat16><<<grid, kRowWarps * 32, 0, st>>>(
          (const __nv_bfloat16*)logits, targets, cf, B, T, U1, V, blank, clamp,
          (__nv_bfloat16*)grad);
      break;
    default:
      B200_CHECK_ARG(false, "b200_rnnt_grad: unsupported dtype %d", dtype);
  }
  B200_CHECK_LAUNCH("rnnt_grad_kernel");
  return B200_OK;
}

extern "C" int b200_rnnt_loss_fwd(const void* logits, int dtype, const int32_t* targets,
                                  const int32_t* logit_len, const int32_t* target_len, int B, int T,
                                  int U1, int V, int blank, float* costs, float* lse,
                                  float* lp_blank, float* lp_emit, float* alpha, float* beta,
                                  double* diag_off, b200_stream_t stream) {
  int rc = b200_rnnt_logprobs(logits, dtype, targets, logit_len, target_len, B, T, U1, V, blank, lse,
                              lp_blank, lp_emit, stream);
  if (rc) return rc;
  return b200_rnnt_lattice(lp_blank, lp_emit, logit_len, target_len, B, T, U1, alpha, beta, costs,
                           diag_off, stream);
}
