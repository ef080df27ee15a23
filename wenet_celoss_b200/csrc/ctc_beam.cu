// Batched CTC prefix beam search, one CTA per utterance, all bookkeeping on the device.
//
// Replaces the Python loop of ASRModel._ctc_prefix_beam_search
//   (wenet/transformer/asr_model.py:326-409; the C++ twin is
//    runtime/core/decoder/ctc_prefix_beam_search.cc:107-200, golden test
//    runtime/core/test/ctc_prefix_beam_search_test.cc:30-73), which
//    transducer_attention_rescoring(beam_search_type='ctc') reaches at
//    wenet/transducer/transducer.py:447-456.
// Per frame t of utterance b (log-probs logp = ctc_probs[b,t,:], fp32):
//   1. top_k_logp, top_k_index = logp.topk(beam)                       (asr_model.py:377)
//   2. for s in top_k_index, for (prefix,(pb,pnb)) in cur_hyps:        (:378-401)
//        s == blank : next[prefix].pb   = log_add(pb', pb+ps, pnb+ps)
//        s == last  : next[prefix].pnb  = log_add(pnb', pnb+ps); next[prefix+s].pnb = log_add(pnb'', pb+ps)
//        else       : next[prefix+s].pnb = log_add(pnb'', pb+ps, pnb+ps)
//      ps = logp[s].item(): fp32 widened to a Python double; all sums and log_add in double
//      (wenet/utils/common.py:268-276)
//   3. stable sort of next (dict insertion order) by log_add(pb,pnb) descending, keep beam (:404-407)
//
// Prefixes live in a per-utterance trie in global memory (node = parent, token); a prefix IS its
// node id, so "prefix_j + s == prefix_j'" is parent[node_j'] == node_j && tok[node_j'] == s: no
// token-list copies, no hashes.  Two entries of `next` can only coincide when one is "prefix_j'
// kept" and the other "prefix_j + s" with prefix_j' == prefix_j + s (current prefixes are
// distinct), and then s == last(prefix_j'), so both contributions happen for the same top-k
// symbol and their order is the beam order of j and j'.
#include "common.cuh"

namespace b200 {
namespace cb {

constexpr int kMaxBeam = 16;
constexpr int kThreads = 256;      // top-k kernel
constexpr int kBeamThreads = 288;  // >= kMaxBeam * (kMaxBeam + 1) = 272 entries
constexpr int kMaxEnt = kMaxBeam * (kMaxBeam + 1);
static_assert(kMaxEnt <= 288, "one thread per entry");

__device__ __forceinline__ bool better(float v, int i, float w, int k) {
  return v > w || (v == w && i < k);
}
__device__ __forceinline__ void argmax_pair(float& bv, int& bi, float ov, int oi) {
  if (better(ov, oi, bv, bi)) {
    bv = ov;
    bi = oi;
  }
}
// log_add over up to three terms in the reference's order (common.py:268-276)
// A term of -inf contributes exp(-inf) = +0.0 to the sum, and x + log(1.0 + 0.0) == x: the shortcuts below
// return bit-for-bit what the full expression returns (most entries of a frame have such a term, and the
// fp64 exp / log chain of the entry phase is what a frame waits for).
__device__ __forceinline__ double log_add2(double a, double b) {
  if (a == -INFINITY) return b;
  if (b == -INFINITY) return a;
  const double mx = fmax(a, b);
  return mx + log(exp(a - mx) + exp(b - mx));
}
__device__ __forceinline__ double log_add3(double a, double b, double c) {
  if (a == -INFINITY) return log_add2(b, c);
  if (b == -INFINITY) return log_add2(a, c);
  if (c == -INFINITY) return log_add2(a, b);
  const double mx = fmax(a, fmax(b, c));
  return mx + log(exp(a - mx) + exp(b - mx) + exp(c - mx));
}

struct Params {
  const float* logp;    // [B,T,V] log-probs (CTC.log_softmax, ctc.py:66-74)
  const int32_t* lens;  // [B]
  int B, T, V, beam, blank;
  int32_t* parent;      // [B][T*beam+1] trie: node 0 = empty prefix
  int32_t* tok;         // [B][T*beam+1]
  float* topv;          // [B][T][beam] per-frame top-k log-probs (sorted)
  int32_t* topi;        // [B][T][beam]
  int32_t* out_node;    // [B][beam] final node of each hypothesis, best first
  double* out_score;    // [B][beam] log_add(pb, pnb)
  int32_t* out_n;       // [B]
};

// ---- kernel 1: top-k of every frame, all B*T rows in parallel (one CTA per row).  The frames are
// independent; only the prefix bookkeeping below is a T-step chain.
__global__ void __launch_bounds__(kThreads) ctc_topk_kernel(const Params P) {
  extern __shared__ float sx[];  // [V]
  __shared__ float s_red[2][kThreads / 32];
  __shared__ int s_idx[2][kThreads / 32];
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  constexpr int NW = kThreads / 32;
  const int row = blockIdx.x;
  const int b = row / P.T, t = row - b * P.T;
  if (t >= P.lens[b]) return;
  const int V = P.V, K = P.beam;
  // sorted descending, lower index wins ties
  const float* x = P.logp + (size_t)row * V;
  float bv = -INFINITY;
  int bi = 0x7fffffff;
#pragma unroll 4
  for (int v = tid; v < V; v += kThreads) {
    const float xv = x[v];
    sx[v] = xv;
    argmax_pair(bv, bi, xv, v);
  }
  for (int k = 0; k < K; ++k) {
    float wv = bv;
    int wi = bi;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1)
      argmax_pair(wv, wi, __shfl_xor_sync(0xffffffffu, wv, o), __shfl_xor_sync(0xffffffffu, wi, o));
    if (lane == 0) {
      s_red[k & 1][warp] = wv;
      s_idx[k & 1][warp] = wi;
    }
    __syncthreads();
    wv = s_red[k & 1][0];
    wi = s_idx[k & 1][0];
#pragma unroll
    for (int w = 1; w < NW; ++w) argmax_pair(wv, wi, s_red[k & 1][w], s_idx[k & 1][w]);
    if (tid == 0) {
      P.topv[(size_t)row * K + k] = wv;
      P.topi[(size_t)row * K + k] = wi;
    }
    if (wi == bi && wi != 0x7fffffff) {
      sx[wi] = __int_as_float(0x7fc00000);  // NaN never compares "better"
      bv = -INFINITY;
      bi = 0x7fffffff;
      for (int v = tid; v < V; v += kThreads) argmax_pair(bv, bi, sx[v], v);
    }
  }
}

// ---- kernel 2: one CTA per utterance walks the frames
__global__ void __launch_bounds__(kBeamThreads) ctc_prefix_beam_kernel(const Params P) {
  __shared__ float s_topv[kMaxBeam];
  __shared__ int s_topi[kMaxBeam];
  // current beam
  __shared__ int c_node[kMaxBeam], c_last[kMaxBeam], c_par[kMaxBeam];
  __shared__ double c_pb[kMaxBeam], c_pnb[kMaxBeam];
  __shared__ int s_n;
  // entries of `next`: e = j*(K+1) + slot, slot 0 = prefix_j kept, slot 1+k = prefix_j + topk[k]
  __shared__ double e_pb[kMaxEnt], e_pnb[kMaxEnt], e_key[kMaxEnt];
  __shared__ int e_order[kMaxEnt];  // insertion order in the reference's dict, -1 = absent
  __shared__ int sel[kMaxBeam];

  const int tid = threadIdx.x;
  const int b = blockIdx.x;
  const int K = P.beam;
  const int Tb = min(P.lens[b], P.T);
  const size_t pool = (size_t)P.T * K + 1;
  int32_t* parent = P.parent + (size_t)b * pool;
  int32_t* tokens = P.tok + (size_t)b * pool;
  const float* topv = P.topv + (size_t)b * P.T * K;
  const int32_t* topi = P.topi + (size_t)b * P.T * K;
  int n_nodes = 1;  // tracked by warp 0 (uniform across its lanes)
  if (tid == 0) {
    parent[0] = -1;
    tokens[0] = -1;
    c_node[0] = 0;
    c_last[0] = -1;
    c_par[0] = -1;
    c_pb[0] = 0.0;
    c_pnb[0] = -INFINITY;
    s_n = 1;
  }
  if (tid < K && Tb > 0) {
    s_topv[tid] = topv[tid];
    s_topi[tid] = topi[tid];
  }
  __syncthreads();

  for (int t = 0; t < Tb; ++t) {
    // next frame's top-k: in flight while this frame is processed
    float nv = 0.f;
    int ni = 0;
    if (tid < K && t + 1 < Tb) {
      nv = topv[(size_t)(t + 1) * K + tid];
      ni = topi[(size_t)(t + 1) * K + tid];
    }
    const int N = s_n;
    const int nent = N * (K + 1);
    // ---- 2. entries of `next`
    if (tid < nent) {
      const int e = tid;
      const int j = e / (K + 1), slot = e - j * (K + 1);
      double pb = -INFINITY, pnb = -INFINITY;
      int order = -1;
      if (slot == 0) {
        // prefix_j kept: blank (-> pb), s == last_j (-> pnb), and "prefix_j' + s" folded in
        int kb = -1, kl = -1;
        for (int k = 0; k < K; ++k) {
          if (s_topi[k] == P.blank) kb = k;
          if (s_topi[k] == c_last[j]) kl = k;  // c_last = -1 for the empty prefix: never a class
        }
        int jp = -1;  // j' with prefix_j == prefix_j' + last_j
        for (int q = 0; q < N; ++q)
          if (c_node[q] == c_par[j]) jp = q;
        if (kb >= 0) {
          const double ps = (double)s_topv[kb];
          pb = log_add3(-INFINITY, c_pb[j] + ps, c_pnb[j] + ps);
          order = (kb * N + j) * 2;
        }
        if (kl >= 0 && P.blank != c_last[j]) {
          const double ps = (double)s_topv[kl];
          const double own = c_pnb[j] + ps;
          if (jp >= 0) {
            const bool rep = c_last[jp] == c_last[j];  // s == last(prefix_j') there: pb only
            if (j < jp) {
              pnb = log_add2(-INFINITY, own);
              pnb = rep ? log_add2(pnb, c_pb[jp] + ps) : log_add3(pnb, c_pb[jp] + ps, c_pnb[jp] + ps);
            } else {
              pnb = rep ? log_add2(-INFINITY, c_pb[jp] + ps)
                        : log_add3(-INFINITY, c_pb[jp] + ps, c_pnb[jp] + ps);
              pnb = log_add2(pnb, own);
            }
            const int o = (kl * N + (j < jp ? j : jp)) * 2 + (j < jp ? 0 : 1);
            order = order < 0 ? o : min(order, o);
          } else {
            pnb = log_add2(-INFINITY, own);
            const int o = (kl * N + j) * 2;
            order = order < 0 ? o : min(order, o);
          }
        }
      } else {
        const int k = slot - 1;
        const int s = s_topi[k];
        if (s != P.blank) {
          // folded into a kept prefix when prefix_j + s is already in the beam
          bool folded = false;
          for (int q = 0; q < N; ++q)
            if (c_par[q] == c_node[j] && c_last[q] == s) folded = true;
          if (!folded) {
            const double ps = (double)s_topv[k];
            pnb = (s == c_last[j]) ? log_add2(-INFINITY, c_pb[j] + ps)
                                   : log_add3(-INFINITY, c_pb[j] + ps, c_pnb[j] + ps);
            order = (k * N + j) * 2 + 1;
          }
        }
      }
      e_pb[e] = pb;
      e_pnb[e] = pnb;
      e_order[e] = order;
      e_key[e] = log_add2(pb, pnb);
    }
    const int cnt = __syncthreads_count(tid < nent && e_order[tid] >= 0);
    // ---- 3. stable sort by key (descending), ties by insertion order; keep K
    if (tid < nent && e_order[tid] >= 0) {
      const double v = e_key[tid];
      const int o = e_order[tid];
      int rank = 0;
      for (int d = 0; d < nent; ++d)
        if (e_order[d] >= 0 && (e_key[d] > v || (e_key[d] == v && e_order[d] < o))) ++rank;
      if (rank < K) sel[rank] = tid;
    }
    __syncthreads();
    // ---- 4. new beam: lane r of warp 0 builds hypothesis r, new trie nodes by ballot
    if (tid < 32) {
      const int Nn = cnt < K ? cnt : K;
      int my_node = -1, my_last = -1, my_par = -1;
      double my_pb = 0.0, my_pnb = 0.0;
      bool ext = false;
      if (tid < Nn) {
        const int e = sel[tid];
        const int j = e / (K + 1), slot = e - j * (K + 1);
        my_pb = e_pb[e];
        my_pnb = e_pnb[e];
        if (slot == 0) {
          my_node = c_node[j];
          my_last = c_last[j];
          my_par = c_par[j];
        } else {
          ext = true;
          my_last = s_topi[slot - 1];
          my_par = c_node[j];
        }
      }
      const unsigned m = __ballot_sync(0xffffffffu, ext);
      if (ext) {
        my_node = n_nodes + __popc(m & ((1u << tid) - 1u));
        parent[my_node] = my_par;
        tokens[my_node] = my_last;
      }
      n_nodes += __popc(m);
      __syncwarp();  // every lane has read the old beam
      if (tid < Nn) {
        c_node[tid] = my_node;
        c_last[tid] = my_last;
        c_par[tid] = my_par;
        c_pb[tid] = my_pb;
        c_pnb[tid] = my_pnb;
      }
      if (tid == 0) s_n = Nn;
      if (tid < K) {
        s_topv[tid] = nv;
        s_topi[tid] = ni;
      }
    }
    __syncthreads();
  }
  const int N = s_n;
  if (tid < K) {
    P.out_node[(size_t)b * K + tid] = tid < N ? c_node[tid] : -1;
    P.out_score[(size_t)b * K + tid] = tid < N ? log_add2(c_pb[tid], c_pnb[tid]) : -INFINITY;
  }
  if (tid == 0) P.out_n[b] = N;
}

}  // namespace cb
}  // namespace b200

namespace b200 {
namespace cb {

// token lists of the final hypotheses: walk the trie back from each final node
__global__ void ctc_beam_backtrack_kernel(const Params P, int32_t* out_tokens, int32_t* out_len) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= P.B * P.beam) return;
  const int b = i / P.beam;
  const size_t pool = (size_t)P.T * P.beam + 1;
  const int32_t* parent = P.parent + (size_t)b * pool;
  const int32_t* tok = P.tok + (size_t)b * pool;
  int node = P.out_node[i];
  if (node < 0) {
    out_len[i] = -1;
    return;
  }
  int len = 0;
  for (int n = node; n > 0; n = parent[n]) ++len;
  out_len[i] = len;
  int32_t* dst = out_tokens + (size_t)i * P.T;
  for (int n = node, p = len - 1; n > 0; n = parent[n], --p) dst[p] = tok[n];
}

}  // namespace cb
}  // namespace b200

using namespace b200;

extern "C" size_t b200_ctc_prefix_beam_workspace_bytes(int B, int T, int beam) {
  const size_t pool = (size_t)T * beam + 1;
  return 2 * (size_t)B * pool * sizeof(int32_t) + (size_t)B * beam * sizeof(int32_t) +
         2 * (size_t)B * T * beam * sizeof(int32_t);
}

extern "C" int b200_ctc_prefix_beam(const float* logp, const int32_t* lens, int B, int T, int V,
                                    int beam, int blank, int32_t* out_tokens, int32_t* out_len,
                                    double* out_score, int32_t* out_n, void* workspace,
                                    size_t ws_bytes, b200_stream_t stream) {
  B200_CHECK_ARG(logp && lens && out_tokens && out_len && out_score && out_n && workspace,
                 "b200_ctc_prefix_beam: null pointer");
  B200_CHECK_ARG(B > 0 && T > 0 && V > 0, "b200_ctc_prefix_beam: bad shape B=%d T=%d V=%d", B, T, V);
  B200_CHECK_ARG(beam >= 1 && beam <= cb::kMaxBeam && beam <= V,
                 "b200_ctc_prefix_beam: beam %d must be in [1, %d] and <= V", beam, cb::kMaxBeam);
  B200_CHECK_ARG(blank >= 0 && blank < V, "b200_ctc_prefix_beam: blank %d out of range", blank);
  B200_CHECK_ARG((size_t)V * sizeof(float) <= 200 * 1024,
                 "b200_ctc_prefix_beam: V=%d does not fit the shared-memory score buffer", V);
  B200_CHECK_ARG(ws_bytes >= b200_ctc_prefix_beam_workspace_bytes(B, T, beam),
                 "b200_ctc_prefix_beam: workspace too small (%zu < %zu bytes)", ws_bytes,
                 b200_ctc_prefix_beam_workspace_bytes(B, T, beam));
  B200_CHECK_ARG(((uintptr_t)workspace & 7) == 0 && ((uintptr_t)out_score & 7) == 0,
                 "b200_ctc_prefix_beam: workspace / out_score must be 8-byte aligned");
  const size_t pool = (size_t)T * beam + 1;
  cb::Params P{};
  P.logp = logp;
  P.lens = lens;
  P.B = B;
  P.T = T;
  P.V = V;
  P.beam = beam;
  P.blank = blank;
  P.parent = reinterpret_cast<int32_t*>(workspace);
  P.tok = P.parent + (size_t)B * pool;
  P.out_node = P.tok + (size_t)B * pool;
  P.topi = P.out_node + (size_t)B * beam;
  P.topv = reinterpret_cast<float*>(P.topi + (size_t)B * T * beam);
  P.out_score = out_score;
  P.out_n = out_n;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t smem = (size_t)V * sizeof(float);
  if (smem > 48 * 1024)
    B200_CHECK_CUDA(cudaFuncSetAttribute(cb::ctc_topk_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  cb::ctc_topk_kernel<<<B * T, cb::kThreads, smem, st>>>(P);
  B200_CHECK_LAUNCH("ctc_topk_kernel");
  cb::ctc_prefix_beam_kernel<<<B, cb::kBeamThreads, 0, st>>>(P);
  B200_CHECK_LAUNCH("ctc_prefix_beam_kernel");
  const int n = B * beam;
  cb::ctc_beam_backtrack_kernel<<<(n + 127) / 128, 128, 0, st>>>(P, out_tokens, out_len);
  B200_CHECK_LAUNCH("ctc_beam_backtrack_kernel");
  return B200_OK;
}
