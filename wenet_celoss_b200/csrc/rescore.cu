// n-best rescoring: the score-combine step of Transducer.transducer_attention_rescoring
// (wenet/transducer/transducer.py:459-513) on the attention decoder's outputs (_cal_attn_score,
// transducer.py:304-330), without the reference's .cpu().numpy() round trip and Python loops.
//
//   dec  [N, L1, V] fp32  attention-decoder logits of the N hypotheses (L1 >= max_len + 1 rows each)
//   rdec [N, L1, V] fp32  right-to-left decoder logits (optional; used when reverse_weight > 0)
//   attn_i  = sum_{j < len_i} log_softmax(dec[i, j])[w_ij]           + log_softmax(dec[i, len_i])[eos]
//   rattn_i = sum_{j < len_i} log_softmax(rdec[i, len_i - 1 - j])[w_ij] + log_softmax(rdec[i, len_i])[eos]
//   score_i = (attn_i (1 - rw) + rattn_i rw) attn_weight + beam_score_i ctc_weight + td_score_i transducer_weight
//   best    = first i with the largest score (the reference's strict '>')
// One CTA per hypothesis; a row's log-sum-exp is a block reduction over V (online maximum), the
// per-hypothesis sums run in fp64 like the reference's Python floats.  HBM-bound: reads dec (and rdec)
// once, N * (len + 1) * V * 4 bytes.
#include "common.cuh"

namespace b200 {
namespace rs {

constexpr int kThreads = 256;

// log-sum-exp of one row of V floats by the whole block; every thread returns the same value
__device__ float block_lse(const float* __restrict__ row, int V, float* red) {
  float m = -INFINITY, s = 0.f;
  for (int v = threadIdx.x; v < V; v += kThreads) {
    const float x = __ldg(row + v);
    if (x > m) {
      s = s * __expf(m - x) + 1.f;
      m = x;
    } else if (x != -INFINITY) {
      s += __expf(x - m);
    }
  }
  // combine (m, s) pairs: warp, then block
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, m, o), s2 = __shfl_xor_sync(0xffffffffu, s, o);
    const float mm = fmaxf(m, m2);
    s = (m == -INFINITY ? 0.f : s * __expf(m - mm)) + (m2 == -INFINITY ? 0.f : s2 * __expf(m2 - mm));
    m = mm;
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  __syncthreads();  // red reuse across rows
  if (lane == 0) {
    red[2 * warp] = m;
    red[2 * warp + 1] = s;
  }
  __syncthreads();
  float M = -INFINITY, S = 0.f;
#pragma unroll
  for (int w = 0; w < kThreads / 32; ++w) {
    const float m2 = red[2 * w], s2 = red[2 * w + 1];
    const float mm = fmaxf(M, m2);
    S = (M == -INFINITY ? 0.f : S * __expf(M - mm)) + (m2 == -INFINITY ? 0.f : s2 * __expf(m2 - mm));
    M = mm;
  }
  return M + logf(S);
}

__global__ void __launch_bounds__(kThreads)
attn_rescore_kernel(const float* __restrict__ dec, const float* __restrict__ rdec,
                    const long long* __restrict__ hyps_pad, const int32_t* __restrict__ hyps_lens, int L1,
                    int Lmax, int V, int eos, double reverse_weight, double attn_weight,
                    const double* __restrict__ beam_score, double ctc_weight, const float* __restrict__ td_score,
                    double transducer_weight, double* __restrict__ scores, double* __restrict__ attn_out) {
  __shared__ float red[2 * (kThreads / 32)];
  const int i = blockIdx.x;
  const int len = hyps_lens[i];
  const long long* hyp = hyps_pad + (long long)i * Lmax;
  double attn = 0.0, rattn = 0.0;
  for (int j = 0; j <= len; ++j) {
    const int w = j < len ? (int)hyp[j] : eos;
    const float* row = dec + ((long long)i * L1 + j) * V;
    const float lse = block_lse(row, V, red);
    attn += (double)(__ldg(row + w) - lse);  // the reference adds fp32 log-probs into a Python float
  }
  if (rdec != nullptr && reverse_weight > 0.0) {
    for (int j = 0; j <= len; ++j) {
      const int w = j < len ? (int)hyp[j] : eos;
      const int r = j < len ? len - j - 1 : len;
      const float* row = rdec + ((long long)i * L1 + r) * V;
      const float lse = block_lse(row, V, red);
      rattn += (double)(__ldg(row + w) - lse);
    }
    attn = attn * (1.0 - reverse_weight) + rattn * reverse_weight;
  }
  if (threadIdx.x == 0) {
    if (attn_out) attn_out[i] = attn;
    scores[i] = attn * attn_weight + beam_score[i] * ctc_weight + (double)td_score[i] * transducer_weight;
  }
}

__global__ void first_argmax_kernel(const double* __restrict__ scores, int N, int32_t* __restrict__ best) {
  if (threadIdx.x == 0 && blockIdx.x == 0) {
    int bi = 0;
    double bs = -INFINITY;
    for (int i = 0; i < N; ++i)
      if (scores[i] > bs) {
        bs = scores[i];
        bi = i;
      }
    *best = bi;
  }
}

}  // namespace rs
}  // namespace b200

using namespace b200;

extern "C" int b200_attn_rescore(const float* dec, const float* rdec, const int64_t* hyps_pad,
                                 const int32_t* hyps_lens, int N, int L1, int Lmax, int V, int eos,
                                 double reverse_weight, double attn_weight, const double* beam_score,
                                 double ctc_weight, const float* td_score, double transducer_weight,
                                 double* scores, double* attn_scores, int32_t* best, b200_stream_t stream) {
  B200_CHECK_ARG(dec && hyps_lens && beam_score && td_score && scores && best && (Lmax == 0 || hyps_pad),
                 "b200_attn_rescore: null pointer");
  B200_CHECK_ARG(N > 0 && L1 > 0 && Lmax >= 0 && L1 >= Lmax + 1 && V > 0,
                 "b200_attn_rescore: bad shape N=%d L1=%d Lmax=%d V=%d (L1 >= Lmax + 1)", N, L1, Lmax, V);
  B200_CHECK_ARG(eos >= 0 && eos < V, "b200_attn_rescore: eos %d out of range", eos);
  B200_CHECK_ARG(!(reverse_weight > 0.0) || rdec, "b200_attn_rescore: reverse_weight > 0 needs rdec");
  rs::attn_rescore_kernel<<<N, rs::kThreads, 0, (cudaStream_t)stream>>>(
      dec, rdec, reinterpret_cast<const long long*>(hyps_pad), hyps_lens, L1, Lmax, V, eos, reverse_weight,
      attn_weight, beam_score, ctc_weight, td_score, transducer_weight, scores, attn_scores);
  B200_CHECK_LAUNCH("attn_rescore_kernel");
  rs::first_argmax_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(scores, N, best);
  B200_CHECK_LAUNCH("first_argmax_kernel");
  return B200_OK;
}
