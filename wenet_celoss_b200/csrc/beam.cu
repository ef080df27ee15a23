// Batched RNN-T prefix beam search, hypothesis bookkeeping on the device.
//
// Replaces PrefixBeamSearch.prefix_beam_search
//   (wenet/transducer/search/prefix_beam_search.py:42-148; one utterance, Python lists,
//    N*N .item() syncs per frame) for a whole batch.  Per frame i and utterance b with
//    N live hypotheses ("rows", row = b*beam + j):
//   lstm / pred     predictor.forward_step from the row's cached state (prefix_beam_search.py:31-40,
//                   predictor.py:180-200) and pred_ffn of its output (joint.py:57) - only for the rows
//                   whose hypothesis was EXTENDED by the previous frame ("dirty" rows).  The reference
//                   repeats the step for every hypothesis on every frame, but a hypothesis kept through
//                   a blank has the same last token and the same cache, i.e. the same step: its result
//                   (new state, projected output) is carried with the row instead (bit-identical).
//   z               z = tanh(enc_ffn(enc)[i] + pred_ffn(pred)) for every live row      (joint.py:55-68)
//   logits          ffn_out(z) for all rows x V                                 (joint.py:69)
//   topk            log_softmax, shallow fusion log(tw*exp(lp) + cw*exp(ctc_lp[i])) in fp32
//                   (:99-101) and a warp-level top-k (k = beam, sorted descending, :104)
//   update          candidate scores fp32(score) + top_lp -> fp64 (:105, :115), blank keeps the
//                   hypothesis and its cache, a symbol appends and takes the new cache
//                   (:108-127); equal hypotheses merge in encounter order with a fp64
//                   log-add (:130-142, wenet/utils/common.py:268-276); stable sort by score,
//                   keep beam (:145-146).
// Two candidates can only be equal when one is "row j1 + blank" and the other "row j2 + s"
// with hyp[j1] == hyp[j2] + [s] (live rows hold distinct hypotheses), so the merge is an
// N x N test (length, last token, 64-bit prefix hash, then an exact compare).
#include "common.cuh"

namespace b200 {
namespace bs {

constexpr int kMaxLayers = 4;
constexpr int kMaxBeam = 16;

struct Model {
  const float* embed;
  const float* w_ih[kMaxLayers];
  const float* w_hh[kMaxLayers];
  const float* b_ih[kMaxLayers];
  const float* b_hh[kMaxLayers];
  const float* w_c;
  const float* b_c;
  const float* w_out;
  const float* b_out;
  int V, Emb, H, L, J, blank;
};

struct State {
  const float* enc_proj;  // [B,T,J]
  const float* ctc_lp;    // [B,T,V] log-softmax of the CTC head
  const int32_t* lens;    // [B]
  int B, T, beam, max_len, n_rows_max;
  float tw, cw;
  // hypotheses, ping-pong over frames: set s = frame & 1 is read, s^1 written
  int32_t* hyp;        // [2][B*beam][max_len]
  int32_t* hyp_len;    // [2][B*beam]
  unsigned long long* hash;       // [2][B*beam]
  unsigned long long* prev_hash;  // [2][B*beam]  hash of the hypothesis without its last token
  double* score;       // [2][B*beam]
  int32_t* n_beam;     // [2][B]
  float* cache_h;      // [2][L][B*beam][H]  state BEFORE consuming the row's last token
  float* cache_c;
  float* new_h;        // [2][L][B*beam][H]  state AFTER consuming it (computed once, then carried)
  float* new_c;
  float* pproj;        // [2][B*beam][J]     pred_ffn(new_h of the last layer) + bias, carried likewise
  float* z;            // [B*beam][J]
  float* logits;       // [B*beam][V]
  float* topv;         // [B*beam][beam]
  int32_t* topi;       // [B*beam][beam]
  int32_t* row_list;   // [2][B*beam]  live rows
  int32_t* dirty_list; // [2][B*beam]  live rows whose predictor step has not been computed yet
  int32_t* n_rows;     // [4]: live rows of set 0 / 1, dirty rows of set 0 / 1
};

__device__ __forceinline__ unsigned long long hash_push(unsigned long long h, int tok) {
  h ^= (unsigned long long)(unsigned)tok + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2);
  return h * 0xff51afd7ed558ccdull;
}
__device__ __forceinline__ float sigm(float x) { return 1.f / (1.f + expf(-x)); }

// Programmatic dependent launch (beam_host.inc launches the frame kernels with
// cudaLaunchAttributeProgrammaticStreamSerialization): FIRST statement of every frame kernel.  The next kernel
// of the stream is set up while this one runs and starts as soon as this grid's CTAs have exited (no explicit
// trigger: with griddepcontrol.launch_dependents at the top the waiting CTAs of up to six later kernels sat on
// the SMs' shared memory and registers and the decode got SLOWER, 656 vs 694 utt/s); the wait blocks until
// the grid launched before this one has completed and its writes are visible, so nothing of a kernel's global
// memory traffic moves ahead of its predecessor.  Measured: 694 -> 755 utt/s at cfg4 (launch gaps 18 -> 3 us
// per frame).  B200_BEAM_NO_PDL=1 launches without the attribute.
__device__ __forceinline__ void pdl_enter() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

__global__ void init_kernel(State S, int blank) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b == 0) {
    S.n_rows[0] = 0;
    S.n_rows[1] = 0;
    S.n_rows[2] = 0;
    S.n_rows[3] = 0;
  }
  if (b >= S.B) return;
  const int row = b * S.beam;
  S.hyp[(size_t)row * S.max_len] = blank;  // every hypothesis starts with the seed blank
  S.hyp_len[row] = 1;
  S.prev_hash[row] = 0ull;
  S.hash[row] = hash_push(0ull, blank);
  S.score[row] = 0.0;
  S.n_beam[b] = 1;
  S.n_beam[S.B + b] = 0;
}
__global__ void init_rows_kernel(State S) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= S.B) return;
  if (S.lens[b] > 0) {
    S.row_list[atomicAdd(&S.n_rows[0], 1)] = b * S.beam;
    S.dirty_list[atomicAdd(&S.n_rows[2], 1)] = b * S.beam;  // the seed hypothesis has no predictor step yet
  }
}

#include "beam_gemm.inc"
#include "beam_select.inc"

}  // namespace bs
}  // namespace b200

#include "beam_host.inc"
