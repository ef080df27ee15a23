// Batched CTC forced alignment (Viterbi over the blank-interleaved label sequence).
//
// Replaces forced_align (wenet/utils/ctc_util.py:27-83; a Python double loop over T x S with
// tensor construction per cell, driven by wenet/bin/alignment.py) for a whole batch: one CTA per
// utterance, one thread per extended state, previous frame in shared memory (ping-pong),
// back-pointers [T,S] int16 in the workspace, backtrack by one thread.
// The arithmetic is fp32 max / add only, so alignments are bit-identical to the reference,
// including its quirks, which are kept on purpose (drop-in):
//   * state 0 takes candidates [alpha[t-1,0], alpha[t-1,-1]]: Python's negative index wraps to the
//     LAST state (ctc_util.py:54-57 with s = 0), back-pointer -1 likewise addresses the last state;
//   * torch.argmax takes the first maximal candidate (ties, all -inf -> the state itself);
//   * the final state is argmax([alpha[T-1,S-1], alpha[T-1,S-2]]) (:69-74).
#include "common.cuh"

namespace b200 {
namespace al {

struct Params {
  const float* logp;       // [B,T,V] log-probs
  const long long* ys;     // [B,Lmax] labels (padding ignored beyond ys_lens)
  const int32_t* hlens;    // [B] frames
  const int32_t* ys_lens;  // [B] labels (>= 1)
  int B, T, V, Lmax, blank;
  int16_t* path;           // [B][T][Smax] back-pointers
  int32_t* out;            // [B][T] aligned token per frame (-1 beyond hlens)
};

__global__ void ctc_forced_align_kernel(const Params P) {
  extern __shared__ float sm[];  // [2][Smax] alpha, then [Smax] int labels
  const int b = blockIdx.x, s = threadIdx.x;
  const int Smax = 2 * P.Lmax + 1;
  const int L = P.ys_lens[b], S = 2 * L + 1;
  const int Tb = min(P.hlens[b], P.T);
  float* alpha = sm;
  int* lab = reinterpret_cast<int*>(sm + 2 * Smax);
  const long long* ys = P.ys + (size_t)b * P.Lmax;
  int32_t* out = P.out + (size_t)b * P.T;
  for (int t = threadIdx.x; t < P.T; t += blockDim.x) out[t] = -1;
  if (Tb <= 0 || L <= 0) return;
  int my = P.blank;
  if (s < S) {
    my = (s & 1) ? (int)ys[s >> 1] : P.blank;
    lab[s] = my;
  }
  const bool ok = my >= 0 && my < P.V;
  __syncthreads();
  const bool two = s < S && (my == P.blank || s < 2 || my == lab[s - 2 < 0 ? 0 : s - 2]);
  const float* lp = P.logp + (size_t)b * P.T * P.V;
  int16_t* path = P.path + (size_t)b * P.T * Smax;
  if (s < S) alpha[s] = (s < 2 && ok) ? lp[my] : -INFINITY;  // (:48-50)
  __syncthreads();
  int cur = 0;
  float nxt_lp = (s < S && ok && Tb > 1) ? lp[(size_t)1 * P.V + my] : -INFINITY;
  for (int t = 1; t < Tb; ++t) {
    const float* a = alpha + cur * Smax;
    const float e = nxt_lp;
    if (s < S && ok && t + 1 < Tb) nxt_lp = lp[(size_t)(t + 1) * P.V + my];  // prefetch
    if (s < S) {
      const int s1 = s == 0 ? S - 1 : s - 1;  // python index -1
      float best = a[s];
      int from = s;
      if (a[s1] > best) {
        best = a[s1];
        from = s - 1;  // -1 for s == 0, as stored by the reference
      }
      if (!two && a[s - 2] > best) {
        best = a[s - 2];
        from = s - 2;
      }
      alpha[(cur ^ 1) * Smax + s] = best + e;
      path[(size_t)t * Smax + s] = (int16_t)from;
    }
    cur ^= 1;
    __syncthreads();
  }
  if (s == 0) {
    const float* a = alpha + cur * Smax;
    // argmax([alpha[-1, S-1], alpha[-1, S-2]]): first maximum
    int st = (a[S - 2] > a[S - 1]) ? S - 2 : S - 1;
    for (int t = Tb - 1; t >= 0; --t) {
      const int idx = st < 0 ? st + S : st;  // negative states index from the end
      out[t] = lab[idx];
      if (t > 0) st = path[(size_t)t * Smax + idx];
    }
  }
}

}  // namespace al
}  // namespace b200

using namespace b200;

extern "C" size_t b200_ctc_forced_align_workspace_bytes(int B, int T, int Lmax) {
  return (size_t)B * T * (2 * Lmax + 1) * sizeof(int16_t);
}

extern "C" int b200_ctc_forced_align(const float* logp, const int64_t* ys_pad, const int32_t* hlens,
                                     const int32_t* ys_lens, int B, int T, int V, int Lmax, int blank,
                                     int32_t* alignment, void* workspace, size_t ws_bytes,
                                     b200_stream_t stream) {
  B200_CHECK_ARG(logp && ys_pad && hlens && ys_lens && alignment && workspace,
                 "b200_ctc_forced_align: null pointer");
  B200_CHECK_ARG(B > 0 && T > 0 && V > 0 && Lmax > 0,
                 "b200_ctc_forced_align: bad shape B=%d T=%d V=%d Lmax=%d", B, T, V, Lmax);
  B200_CHECK_ARG(blank >= 0 && blank < V, "b200_ctc_forced_align: blank %d out of range", blank);
  const int Smax = 2 * Lmax + 1;
  B200_CHECK_ARG(Smax <= 1024, "b200_ctc_forced_align: 2*Lmax+1 = %d > 1024 states not supported", Smax);
  B200_CHECK_ARG(ws_bytes >= b200_ctc_forced_align_workspace_bytes(B, T, Lmax),
                 "b200_ctc_forced_align: workspace too small");
  al::Params P{};
  P.logp = logp;
  P.ys = reinterpret_cast<const long long*>(ys_pad);
  P.hlens = hlens;
  P.ys_lens = ys_lens;
  P.B = B;
  P.T = T;
  P.V = V;
  P.Lmax = Lmax;
  P.blank = blank;
  P.path = reinterpret_cast<int16_t*>(workspace);
  P.out = alignment;
  const int threads = (Smax + 31) / 32 * 32;
  const size_t smem = (size_t)3 * Smax * sizeof(float);
  al::ctc_forced_align_kernel<<<B, threads, smem, (cudaStream_t)stream>>>(P);
  B200_CHECK_LAUNCH("ctc_forced_align_kernel");
  return B200_OK;
}

// ---------------------------------------------------------------------------------------------
// Allocation-free variant (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: CUDA graph capture, tight serving loops,
// harnesses that forbid it).  The back-pointer scratch [B,T,2*Lmax+1] int16 lives in a file-scope
// device array of fixed size; every GPU that loads the library holds these bytes until exit.
//   * limit: B * T * (2*Lmax+1) * 2 bytes <= B200_CTC_FORCED_ALIGN_STATIC_MAX_BYTES (32 MiB;
//     the cfg5 shape B=64, T=1000, Lmax=100 needs 25.7 MB); larger problems are rejected;
//   * every call on a device shares the one array: calls that overlap on the GPU (two streams,
//     two host threads) race on it - serialise them, or use b200_ctc_forced_align with an own
//     workspace.  The array keeps the previous call's contents; nothing has to start at zero:
//     the backtrack reads only back-pointers path[t, s], 1 <= t < T_b, s < S_b, and the forward
//     sweep of the same launch has written every one of them;
//   * in host code the array's NAME is a host-side stand-in, not a device address: the address is
//     taken with cudaGetSymbolAddress, once per device, with that device current.
// ---------------------------------------------------------------------------------------------
namespace b200 {
namespace al {
constexpr size_t kStaticScratchBytes = (size_t)32 << 20;
static __device__ __align__(256) unsigned char g_align_scratch[kStaticScratchBytes];

constexpr int kMaxDevices = 64;
static void* g_scratch_ptr[kMaxDevices];  // device address of g_align_scratch per device, nullptr = not yet asked
}  // namespace al
}  // namespace b200

extern "C" size_t b200_ctc_forced_align_static_max_bytes(void) { return al::kStaticScratchBytes; }

extern "C" int b200_ctc_forced_align_static(const float* logp, const int64_t* ys_pad, const int32_t* hlens,
                                            const int32_t* ys_lens, int B, int T, int V, int Lmax, int blank,
                                            int32_t* alignment, b200_stream_t stream) {
  B200_CHECK_ARG(logp && ys_pad && hlens && ys_lens && alignment, "b200_ctc_forced_align_static: null pointer");
  B200_CHECK_ARG(B > 0 && T > 0 && V > 0 && Lmax > 0,
                 "b200_ctc_forced_align_static: bad shape B=%d T=%d V=%d Lmax=%d", B, T, V, Lmax);
  B200_CHECK_ARG(blank >= 0 && blank < V, "b200_ctc_forced_align_static: blank %d out of range", blank);
  const int Smax = 2 * Lmax + 1;
  B200_CHECK_ARG(Smax <= 1024, "b200_ctc_forced_align_static: 2*Lmax+1 = %d > 1024 states not supported", Smax);
  const size_t need = b200_ctc_forced_align_workspace_bytes(B, T, Lmax);
  B200_CHECK_ARG(need <= al::kStaticScratchBytes,
                 "b200_ctc_forced_align_static: B*T*(2*Lmax+1)*2 = %zu bytes of back-pointers exceed the static "
                 "scratch of %zu bytes (use b200_ctc_forced_align with a caller workspace)",
                 need, al::kStaticScratchBytes);
  int dev = 0;
  B200_CHECK_CUDA(cudaGetDevice(&dev));
  B200_CHECK_ARG(dev >= 0 && dev < al::kMaxDevices, "b200_ctc_forced_align_static: device %d not supported", dev);
  // set-up, once per device, with that device current (a benign race: every thread would store the same value)
  void* scratch = __atomic_load_n(&al::g_scratch_ptr[dev], __ATOMIC_ACQUIRE);
  if (scratch == nullptr) {
    B200_CHECK_CUDA(cudaGetSymbolAddress(&scratch, al::g_align_scratch));
    B200_CHECK_ARG(scratch != nullptr, "b200_ctc_forced_align_static: no device address for the scratch array");
    __atomic_store_n(&al::g_scratch_ptr[dev], scratch, __ATOMIC_RELEASE);
  }
  al::Params P{};
  P.logp = logp;
  P.ys = reinterpret_cast<const long long*>(ys_pad);
  P.hlens = hlens;
  P.ys_lens = ys_lens;
  P.B = B;
  P.T = T;
  P.V = V;
  P.Lmax = Lmax;
  P.blank = blank;
  P.path = reinterpret_cast<int16_t*>(scratch);
  P.out = alignment;
  const int threads = (Smax + 31) / 32 * 32;
  const size_t smem = (size_t)3 * Smax * sizeof(float);
  al::ctc_forced_align_kernel<<<B, threads, smem, (cudaStream_t)stream>>>(P);
  B200_CHECK_LAUNCH("ctc_forced_align_kernel");
  return B200_OK;
}
