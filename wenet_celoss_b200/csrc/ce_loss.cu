// Hot-word cross-entropy loss ("celoss"): nn.CrossEntropyLoss() over hw_output [B, C, U]
// with labels [B, U], as Transducer.forward applies it to the hot-word classifier head
// (wenet/transducer/transducer.py:66 hw_criterion, :170-200 the three loss modes, :262
// loss += hw_weight * hw_loss).  Class dimension = 1 (the reference permutes to [B, C, U] at
// :175,:185,:194), mean over the positions whose label != ignore_index (PyTorch default -100).
//   loss = (1/n) sum_{b,u} ( lse_c x[b,c,u] - x[b,y,u] )
//   d loss / d x[b,c,u] = dy/n * ( softmax_c - [c == y] )
// The tensor is tiny (B*U positions x num_labels = 2 classes): one CTA, fixed reduction order
// (deterministic), everything stays on the device and on the caller's stream.
#include "common.cuh"

namespace b200 {
namespace ce {

constexpr int kThreads = 1024;

__global__ void __launch_bounds__(kThreads)
ce_fwd_kernel(const float* __restrict__ x, const long long* __restrict__ y, int B, int C, int U,
              long long ignore_index, float* __restrict__ loss, float* __restrict__ count) {
  __shared__ double s_sum[kThreads / 32];
  __shared__ int s_cnt[kThreads / 32];
  const long long n = (long long)B * U;
  double sum = 0.0;
  int cnt = 0;
  for (long long i = threadIdx.x; i < n; i += kThreads) {
    const long long lab = y[i];
    if (lab == ignore_index || lab < 0 || lab >= C) continue;
    const int b = (int)(i / U), u = (int)(i - (long long)b * U);
    const float* row = x + (size_t)b * C * U + u;  // class stride U
    float m = -INFINITY;
    for (int c = 0; c < C; ++c) m = fmaxf(m, row[(size_t)c * U]);
    float s = 0.f;
    for (int c = 0; c < C; ++c) s += expf(row[(size_t)c * U] - m);
    sum += (double)((m + logf(s)) - row[(size_t)lab * U]);
    ++cnt;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    sum += __shfl_xor_sync(0xffffffffu, sum, o);
    cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
  }
  if (lane == 0) {
    s_sum[warp] = sum;
    s_cnt[warp] = cnt;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double t = 0.0;
    int k = 0;
    for (int w = 0; w < kThreads / 32; ++w) {
      t += s_sum[w];
      k += s_cnt[w];
    }
    *count = (float)k;
    *loss = k > 0 ? (float)(t / k) : NAN;  // torch: mean over zero elements is nan
  }
}

__global__ void ce_bwd_kernel(const float* __restrict__ x, const long long* __restrict__ y, int B, int C,
                              int U, long long ignore_index, const float* __restrict__ dy,
                              const float* __restrict__ count, float* __restrict__ grad) {
  const long long n = (long long)B * U;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int b = (int)(i / U), u = (int)(i - (long long)b * U);
  const float* row = x + (size_t)b * C * U + u;
  float* grow = grad + (size_t)b * C * U + u;
  const long long lab = y[i];
  if (lab == ignore_index || lab < 0 || lab >= C) {
    for (int c = 0; c < C; ++c) grow[(size_t)c * U] = 0.f;
    return;
  }
  const float scale = dy[0] / count[0];
  float m = -INFINITY;
  for (int c = 0; c < C; ++c) m = fmaxf(m, row[(size_t)c * U]);
  float s = 0.f;
  for (int c = 0; c < C; ++c) s += expf(row[(size_t)c * U] - m);
  const float inv = 1.f / s;
  for (int c = 0; c < C; ++c) {
    const float p = expf(row[(size_t)c * U] - m) * inv;
    grow[(size_t)c * U] = scale * (p - (c == lab ? 1.f : 0.f));
  }
}

}  // namespace ce
}  // namespace b200

using namespace b200;

extern "C" int b200_ce_loss_fwd(const float* logits, const int64_t* labels, int B, int C, int U,
                                long long ignore_index, float* loss, float* count,
                                b200_stream_t stream) {
  B200_CHECK_ARG(logits && labels && loss && count, "b200_ce_loss_fwd: null pointer");
  B200_CHECK_ARG(B > 0 && C > 0 && U > 0, "b200_ce_loss_fwd: bad shape B=%d C=%d U=%d", B, C, U);
  ce::ce_fwd_kernel<<<1, ce::kThreads, 0, (cudaStream_t)stream>>>(
      logits, reinterpret_cast<const long long*>(labels), B, C, U, ignore_index, loss, count);
  B200_CHECK_LAUNCH("ce_fwd_kernel");
  return B200_OK;
}

extern "C" int b200_ce_loss_bwd(const float* logits, const int64_t* labels, int B, int C, int U,
                                long long ignore_index, const float* dy, const float* count,
                                float* grad, b200_stream_t stream) {
  B200_CHECK_ARG(logits && labels && dy && count && grad, "b200_ce_loss_bwd: null pointer");
  B200_CHECK_ARG(B > 0 && C > 0 && U > 0, "b200_ce_loss_bwd: bad shape B=%d C=%d U=%d", B, C, U);
  const long long n = (long long)B * U;
  ce::ce_bwd_kernel<<<(unsigned)((n + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      logits, reinterpret_cast<const long long*>(labels), B, C, U, ignore_index, dy, count, grad);
  B200_CHECK_LAUNCH("ce_bwd_kernel");
  return B200_OK;
}
