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
//
// The frame step is the serial chain of the whole CTC loss (T dependent frames), so nothing but the
// recurrence itself sits on it - no block barrier, no shared-memory round trip:
//   * a lane owns TWO adjacent extended states (alpha: 2g, 2g+1; beta: 2g-1, 2g), kept in registers.
//     With that pairing both updates need exactly ONE value from the neighbouring lane (its odd state):
//     one __shfl per frame; the skip transition of the odd state reads that same value;
//   * values live in log2 units (ex2.approx / lg2.approx directly); log-add with the maximum factored
//     out: 5 MUFU per lane and frame;
//   * 64 states per warp, ceil((S+1)/64) warps per direction.  Warps are NOT synchronised per frame: the
//     one value that crosses a warp boundary travels through a 32-slot shared-memory ring of
//     {value, frame} words (one 8-byte volatile store / load), so a warp simply runs a frame or two
//     behind its upstream neighbour (a software pipeline over the state axis);
//   * rescaling (alpha / beta are stored relative to a running fp64 offset per frame, so that the stored
//     fp32 numbers stay O(10)): every 8 frames each warp publishes its maximum (one redux.sync); two
//     blocks later all warps read all published maxima - by then present without waiting - and derive the
//     SAME per-frame correction c from level + drift (deterministic arithmetic on identical inputs);
//     c is folded into the emission term off the chain.  Waiting for the publications also bounds the
//     lead of any warp to < 24 frames, which is what makes the 32-slot ring safe.
// cfg5 (T = 1000, S <= 201): 330 us (one thread per state, block barrier + expf / logf per frame)
// -> see profiles/README.md.
// ---------------------------------------------------------------------------
constexpr int kCtcPf = 8;     // frames of lp_ext in registers ahead of the chain = frames per rescale block
constexpr int kCtcRing = 32;  // slots of the warp-boundary ring

__device__ __forceinline__ void st_tagged(unsigned long long* p, float v, int tag) {
  const unsigned long long w = ((unsigned long long)(unsigned)tag << 32) | (unsigned)__float_as_int(v);
  *reinterpret_cast<volatile unsigned long long*>(p) = w;
}
__device__ __forceinline__ bool ld_tagged(const unsigned long long* p, int tag, float* v) {
  const unsigned long long w = *reinterpret_cast<const volatile unsigned long long*>(p);
  *v = __int_as_float((int)(unsigned)w);
  return (int)(w >> 32) == tag;
}

// One direction of the trellis for one warp.  DIR = +1: alpha (frames ascending, value from lane - 1);
// DIR = -1: beta (frames descending, value from lane + 1).  Returns the final fp64 offset (log2 units);
// va / vb leave holding the last frame's even / odd state (log2 units, relative to the offset).
//
// One warp per scheduler and a serial chain: every instruction of the frame body is paid at its full
// issue latency (ncu on the first version of this kernel: 113 instructions per frame, 4.8 cycles per
// instruction, 1.1 of them branch resolution).  Hence: states outside [0, S) carry an emission of -inf and
// need no select; full blocks of 8 frames run without a bound check; the neighbour-warp test is hoisted
// (HAS_UP); offsets are written per block by lanes 0..7; addresses are 32-bit row offsets.
template <int DIR, bool HAS_UP>
__device__ __forceinline__ double ctc_trellis(const float* __restrict__ lp, float* __restrict__ out,
                                              double* __restrict__ offs, const int* lab,
                                              unsigned long long* ring, unsigned long long* pub, int S,
                                              int Smax, int Tb, int blank, float& va, float& vb) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const int g = threadIdx.x;
  const int e = 2 * g, o = DIR > 0 ? 2 * g + 1 : 2 * g - 1;
  const bool mine_e = e < S, mine_o = o >= 0 && o < S;
  bool skip;
  if (DIR > 0) skip = mine_o && o >= 3 && lab[o] != blank && lab[o] != lab[o - 2];
  else skip = mine_o && o + 2 < S && lab[o + 2] != blank && lab[o + 2] != lab[o];
  const bool has_down = warp + DIR >= 0 && warp + DIR < nw;
  const int lane_in = DIR > 0 ? 0 : 31, lane_out = DIR > 0 ? 31 : 0;
  const bool is_in = lane == lane_in, is_out = lane == lane_out && has_down;
  const unsigned long long* ring_in = ring + (HAS_UP ? warp - DIR : 0) * kCtcRing;
  unsigned long long* ring_out = ring + warp * kCtcRing;
  const int step = DIR * Smax;                      // row offset from one processed frame to the next
  const int row0 = (DIR > 0 ? 0 : Tb - 1) * Smax;   // row offset of processing index 0
  const int last_row = (DIR > 0 ? Tb - 1 : 0) * Smax;
  const float* lpe = lp + e;
  const float* lpo = lp + o;
  float* oute = out + e;
  float* outo = out + o;
  // ---- frame 0
  {
    const bool se = DIR > 0 ? e == 0 : e == S - 1;
    const bool so = DIR > 0 ? o == 1 : o == S - 2;
    va = (mine_e && se) ? lpe[row0] * kLog2e : -INFINITY;
    vb = (mine_o && so) ? lpo[row0] * kLog2e : -INFINITY;
    if (mine_e) oute[row0] = va * kLn2;
    if (mine_o) outo[row0] = vb * kLn2;
    const float wm = warp_max_redux(fmaxf(va, vb));
    if (lane == 0) st_tagged(pub + warp * 4, wm, 0);
    if (is_out) st_tagged(ring_out, vb, 0);
    if (g == 0) offs[DIR > 0 ? 0 : Tb - 1] = 0.0;
  }
  // emissions of the next block, natural log; states outside [0, S): -inf (they then stay -inf by themselves)
  float xe[kCtcPf], xo[kCtcPf];
  auto load_block = [&](int i0, float* de, float* dn) {
#pragma unroll
    for (int k = 0; k < kCtcPf; ++k) {
      const int r = (i0 + k < Tb) ? row0 + (i0 + k) * step : last_row;
      de[k] = mine_e ? lpe[r] : -INFINITY;
      dn[k] = mine_o ? lpo[r] : -INFINITY;
    }
  };
  load_block(1, xe, xo);
  // rescale controller state (identical in every warp): O = offset so far, (x_last, i_last) = the last
  // published maximum in absolute log2 units and the frame it was measured at
  double O = 0.0, O_pub[4] = {0.0, 0.0, 0.0, 0.0};
  int i_pub[4] = {0, 0, 0, 0};
  double x_last = 0.0, vel = 0.0;
  int i_last = -1, p_last = -1;
  int blk = 0;
  for (int i0 = 1; i0 < Tb; i0 += kCtcPf, ++blk) {
    float nxe[kCtcPf], nxo[kCtcPf];
    load_block(i0 + kCtcPf, nxe, nxo);
    // ---- per-frame correction of this block from the maxima published two blocks ago
    const int pn = blk > 0 ? blk - 1 : 0;
    float M;
    for (;;) {
      float v = -INFINITY;
      bool ok = true;
      if (lane < nw) ok = ld_tagged(pub + lane * 4 + (pn & 3), pn, &v);
      if (__all_sync(0xffffffffu, ok)) {
        M = warp_max_redux(lane < nw ? v : -INFINITY);
        break;
      }
    }
    if (pn != p_last) {
      const double x_a = (M == -INFINITY) ? O_pub[pn & 3] : (double)M + O_pub[pn & 3];
      const int i_a = i_pub[pn & 3];
      if (p_last >= 0 && i_a > i_last) vel = (x_a - x_last) / (double)(i_a - i_last);
      if (p_last < 0) vel = x_a;  // first guess of the drift per frame: what frame 0 itself contributed
      x_last = x_a;
      i_last = i_a;
      p_last = pn;
    }
    const int i_end = i0 + kCtcPf - 1;
    const float c = (float)((x_last + vel * (double)(i_end - i_last) - O) * (1.0 / kCtcPf));
    // offsets of this block's frames (k + 1 corrections of c after O): lanes 0..7 of warp 0
    if (warp == 0 && lane < kCtcPf && i0 + lane < Tb)
      offs[DIR > 0 ? i0 + lane : Tb - 1 - i0 - lane] = (O + (double)c * (double)(lane + 1)) * 0.6931471805599453094;
    float pe[kCtcPf], po[kCtcPf];  // emissions in log2 units minus the correction
#pragma unroll
    for (int k = 0; k < kCtcPf; ++k) {
      pe[k] = fmaf(xe[k], kLog2e, -c);
      po[k] = fmaf(xo[k], kLog2e, -c);
    }
    const int rowb = row0 + i0 * step;
    auto frame_step = [&](int k) {
      const int i = i0 + k;
      // the one value from the neighbouring lane / warp: the odd state of frame i - 1
      float pO = DIR > 0 ? __shfl_up_sync(0xffffffffu, vb, 1) : __shfl_down_sync(0xffffffffu, vb, 1);
      if (HAS_UP) {
        float bv;
        while (!ld_tagged(ring_in + ((i - 1) & (kCtcRing - 1)), i - 1, &bv)) {
        }
        pO = is_in ? bv : pO;
      } else {
        pO = is_in ? -INFINITY : pO;
      }
      const float na = lae2_log2(va, pO) + pe[k];
      const float nb = lae3_log2(vb, va, skip ? pO : -INFINITY) + po[k];
      va = na;
      vb = nb;
      if (is_out) st_tagged(ring_out + (i & (kCtcRing - 1)), vb, i);
      if (mine_e) oute[rowb + k * step] = va * kLn2;
      if (mine_o) outo[rowb + k * step] = vb * kLn2;
    };
    auto publish = [&]() {  // publication blk + 1: after this block's first frame
      const float wm = warp_max_redux(fmaxf(va, vb));
      if (lane == 0) st_tagged(pub + warp * 4 + ((blk + 1) & 3), wm, blk + 1);
      O_pub[(blk + 1) & 3] = O + (double)c;
      i_pub[(blk + 1) & 3] = i0;
    };
    if (i0 + kCtcPf <= Tb) {  // full block: no bound checks
      frame_step(0);
      publish();
#pragma unroll
      for (int k = 1; k < kCtcPf; ++k) frame_step(k);
      O += (double)c * (double)kCtcPf;
    } else {
      frame_step(0);
      publish();
      int n = 1;
#pragma unroll
      for (int k = 1; k < kCtcPf; ++k)
        if (i0 + k < Tb) {  // warp-uniform
          frame_step(k);
          ++n;
        }
      O += (double)c * (double)n;
    }
#pragma unroll
    for (int k = 0; k < kCtcPf; ++k) {
      xe[k] = nxe[k];
      xo[k] = nxo[k];
    }
  }
  return O;
}

__global__ void ctc_lattice_kernel(const float* __restrict__ lp_ext,
                                   const long long* __restrict__ ys_pad,
                                   const int32_t* __restrict__ hlens,
                                   const int32_t* __restrict__ ys_lens, int Tm, int Lmax, int blank,
                                   float* __restrict__ alpha, float* __restrict__ beta,
                                   float* __restrict__ nll, int32_t* __restrict__ leader,
                                   double* __restrict__ frame_off) {
  extern __shared__ unsigned long long sm64[];  // ring [nw][32] | pub [nw][4] | lab [Smax] | fin [2]
  const int b = blockIdx.x;
  const int tid = threadIdx.x, nw = blockDim.x >> 5;
  const int Smax = 2 * Lmax + 1;
  const int Tb = hlens[b], L = ys_lens[b];
  const int S = 2 * L + 1;
  unsigned long long* ring = sm64;
  unsigned long long* pub = sm64 + nw * kCtcRing;
  int* lab = reinterpret_cast<int*>(pub + nw * 4);
  float* fin = reinterpret_cast<float*>(lab + Smax + 1);
  double* offs = frame_off + ((long long)b * 2 + blockIdx.y) * (Tm + 1);
  const long long* ys = ys_pad + (long long)b * Lmax;
  for (int i = tid; i < nw * (kCtcRing + 4); i += blockDim.x) sm64[i] = 0xffffffff00000000ull;  // tag -1
  for (int i = tid; i < Smax; i += blockDim.x) lab[i] = (i < S) ? ext_label(ys, i, blank) : -1;
  __syncthreads();
  const long long off = (long long)b * Tm * Smax;
  const float* lp = lp_ext + off;

  if (blockIdx.y == 0) {
    // leader[s] = smallest state index with the same label (blank states -> 0)
    for (int s = tid; s < S; s += blockDim.x) {
      int ld = s;
      if ((s & 1) == 0) {
        ld = 0;
      } else {
        for (int j = 1; j < s; j += 2)
          if (lab[j] == lab[s]) {
            ld = j;
            break;
          }
      }
      leader[(long long)b * Smax + s] = ld;
    }
    if (Tb <= 0) {
      if (tid == 0) {
        nll[b] = (L == 0) ? 0.f : INFINITY;
        offs[Tm] = (L == 0) ? 0.0 : (double)-INFINITY;
      }
      return;
    }
    float va, vb;
    const double O = (tid >> 5) > 0
                         ? ctc_trellis<1, true>(lp, alpha + off, offs, lab, ring, pub, S, Smax, Tb, blank, va, vb)
                         : ctc_trellis<1, false>(lp, alpha + off, offs, lab, ring, pub, S, Smax, Tb, blank, va, vb);
    // final frame: states S-1 (even, lane L) and S-2 (odd, lane L-1)
    if (2 * tid == S - 1) fin[0] = va;
    if (2 * tid + 1 == S - 2) fin[1] = vb;
    if (S == 1 && tid == 0) fin[1] = -INFINITY;
    __syncthreads();
    if (tid == 0) {
      const double ll = ((double)lae2_log2(fin[0], fin[1]) + O) * 0.6931471805599453094;
      nll[b] = (float)(-ll);
      offs[Tm] = ll;  // [b][0][T] = log-likelihood in fp64
    }
  } else {
    if (Tb <= 0) return;
    float va, vb;
    if ((tid >> 5) + 1 < nw) ctc_trellis<-1, true>(lp, beta + off, offs, lab, ring, pub, S, Smax, Tb, blank, va, vb);
    else ctc_trellis<-1, false>(lp, beta + off, offs, lab, ring, pub, S, Smax, Tb, blank, va, vb);
  }
}

// ---------------------------------------------------------------------------
// pass 3: gradient w.r.t. logits
//   grad[c] = scale_b * ( exp(x[c]-lse) - occ[c] ),  0 for t >= T_b
// Persistent CTAs of kRowWarps STREAMING warps + one PATCH warp.  A work unit ("group") is
// kRowWarps consecutive frames of one utterance; CTA i takes groups i, i + grid, ...
//   * streaming warp w streams row w of the group: scale * softmax, 16-byte loads and stores - the
//     2*B*T*V*s bytes that bound the pass - and needs nothing from memory to start: lse, scale and the
//     live flag of its row were left in shared memory by the patch warp;
//   * the patch warp works ONE GROUP AHEAD (two shared-memory buffers, named barriers full / empty
//     per buffer): it forms the state posteriors of the group's rows,
//       occ[leader(s)] += exp(alpha_t(s) + beta_t(s) - lp_t(s) + off),
//     and leaves, per row, the final entries scale * (exp(lp) - occ) of the utterance's distinct label
//     classes (and blank) IN CLASS ORDER, next to a bitmap over the classes and its running popcount:
//     the entry of class c sits at rank(c) = number of label classes below c.  lp_ext[s] IS x[l'_s] - lse
//     (pass 1 stored exactly that difference), so nothing is gathered from the logits row;
//   * the streaming warp substitutes these entries INSIDE its stream (one bitmap word per 16-byte
//     segment; the rare hit costs a popcount): every byte of the gradient is written once, by a full
//     512-byte warp store.
// Why: the previous kernels wrote the <= S patched entries AFTER the row's stream.  6.4 M scattered
// 4-byte writes into lines that had already left L2 are 6.4 M read-modify-writes of a DRAM burst each:
// 156 us of a 597 us kernel (measured by switching the scatter off), i.e. as much as 1 GB of streaming.
// A round trip to memory takes 2-4 us while 32 warps per SM stream (queueing), so nothing latency-bound
// may sit in front of a row's stream: the patch warp makes one round trip per group (alpha / beta /
// lp_ext of the group's rows are one contiguous piece of each array: cp.async, no registers; per-row
// scalars by lanes 0..7 next to them; leader / label tables per utterance).
// Shared memory is kept small on purpose (28 KB per CTA at cfg5): the loads in flight live in the L1
// part of the same 256 KB, and a 52 KB variant (class -> state table of 2 V bytes per buffer) that left
// 18 KB of L1 per SM streamed at 569 us where 36 KB per CTA gave 530.
// History at cfg5 (2.56 GB): 537 us = 0.73 of the HBM peak (one warp per row did occupancies, stream,
// scatter); 597 us with a patch warp but the scatter kept; 441 us without the scatter.
// ---------------------------------------------------------------------------
constexpr int kPatchWarps = 4;                          // each takes kRowWarps / kPatchWarps rows of a group
constexpr int kPatchRows = kRowWarps / kPatchWarps;
constexpr int kGradThreads = (kRowWarps + kPatchWarps) * 32;
constexpr int kGradInfo = 32;  // floats of per-group row info per buffer: c0l2[8] | scale[8]

__device__ __forceinline__ void cp_async_f(float* dst, const float* src, bool vec) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(dst);
  if (vec)
    asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" ::"r"(d), "l"(src) : "memory");
  else
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"(d), "l"(src) : "memory");
}
// n floats global -> shared by one warp; dst and src have the same position within their 16-byte line
// (the caller offsets dst by the phase of src), so all but the ragged ends go as 16-byte copies
__device__ __forceinline__ void cp_async_span(float* dst, const float* src, int n, int lane) {
  int head = (int)((16 - ((uintptr_t)src & 15)) & 15) >> 2;
  if (head > n) head = n;
  const int nvec = (n - head) >> 2;
  if (lane < head) cp_async_f(dst + lane, src + lane, false);
  for (int i = lane; i < nvec; i += 32) cp_async_f(dst + head + 4 * i, src + head + 4 * i, true);
  const int done = head + 4 * nvec;
  if (lane < n - done) cp_async_f(dst + done + lane, src + done + lane, false);
}
__device__ __forceinline__ void named_bar_sync(int id, int n) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(n) : "memory");
}
__device__ __forceinline__ void named_bar_arrive(int id, int n) {
  asm volatile("bar.arrive %0, %1;" ::"r"(id), "r"(n) : "memory");
}

// shared-memory layout (32-bit words): per patch warp its staging alpha | beta | lp_ext (the alpha part
// doubles as the occupancy accumulator), then per buffer: values [kRowWarps][Lmax + 1],
// row info, class bitmap [nbm], running popcount [nbm] (uint16)
struct GradLayout {
  int nst, nval, nbm, nbuf;
  size_t bytes;
};
__host__ __device__ inline GradLayout ctc_grad_layout(int Lmax, int V) {
  GradLayout L;
  L.nst = (kPatchRows * (2 * Lmax + 1) + 4 + 3) & ~3;  // per patch warp and array, + the 16-byte phase of the source
  L.nval = (kRowWarps * (Lmax + 1) + 3) & ~3;
  L.nbm = ((V + 31) / 32 + 3) & ~3;
  L.nbuf = L.nval + kGradInfo + L.nbm + L.nbm / 2;
  L.bytes = ((size_t)3 * kPatchWarps * L.nst + 2 * (size_t)L.nbuf) * sizeof(float);
  return L;
}

// shared-memory reads by 32-bit shared-space address (the pointers below are selected per buffer; as
// generic pointers every read re-derived the shared window: a dozen instructions per patched entry)
__device__ __forceinline__ uint32_t lds_u32(uint32_t a) {
  uint32_t v;
  asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ uint32_t lds_u16(uint32_t a) {
  uint16_t v;
  asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a));
  return v;
}
__device__ __forceinline__ float lds_f32(uint32_t a) {
  float v;
  asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(a));
  return v;
}

// the patched entry of class e, if e is a label class of the utterance: rank = set bits below e
// (bm: class bitmap, pre: running popcount per word (uint16), val: entries in class order)
__device__ __forceinline__ void ctc_patch_elem(float& f, int e, uint32_t bm, uint32_t pre, uint32_t val) {
  const uint32_t word = lds_u32(bm + 4u * (e >> 5));
  const uint32_t sh = e & 31;
  if ((word >> sh) & 1u) f = lds_f32(val + 4u * (lds_u16(pre + 2u * (e >> 5)) + __popc(word & ((1u << sh) - 1u))));
}

// One row of the gradient by one warp: out[e] = scale * exp(x[e] - lse), entries of label classes
// replaced from val while they pass through registers.
template <typename T>
__device__ __forceinline__ void ctc_grad_row(const T* __restrict__ row, T* __restrict__ out, int V, int lane,
                                             float c0l2, float scale, uint32_t bm, uint32_t pre, uint32_t val) {
  constexpr int K = Elem<T>::kPerVec;
  RowSegs<T> rs(row, V);
  uint4* oseg0 = reinterpret_cast<uint4*>(reinterpret_cast<uintptr_t>(out) & ~uintptr_t(15));
  uint4 v[kUnroll], nv[kUnroll];
  warp_row_stream_begin<T>(rs, lane, nv);
  const bool aligned = rs.head == 0;  // a segment's K classes are K adjacent bits of ONE bitmap word
  for (int base = 0; base < rs.nseg; base += 32 * kUnroll) {
#pragma unroll
    for (int j = 0; j < kUnroll; ++j) v[j] = nv[j];
#pragma unroll
    for (int j = 0; j < kUnroll; ++j) {  // software pipeline: next block's loads in flight
      int idx = base + 32 * kUnroll + j * 32 + lane;
      if (idx < rs.nseg) nv[j] = ldg_stream16(rs.seg0 + idx);
    }
#pragma unroll
    for (int j = 0; j < kUnroll; ++j) {
      int idx = base + j * 32 + lane;
      if (idx >= rs.nseg) continue;
      float f[K];
      Elem<T>::unpack(v[j], f);
#pragma unroll
      for (int k = 0; k < K; ++k) f[k] = scale * ex2_approx(fmaf(f[k], kLog2e, c0l2));
      const int e0 = idx * K - rs.head;
      if (e0 >= 0 && e0 + K <= V) {
        if (aligned) {
          const uint32_t word = lds_u32(bm + 4u * (e0 >> 5));
          const uint32_t sh = e0 & 31;
          const uint32_t bits = (word >> sh) & ((1u << K) - 1u);
          if (bits) {  // ~L of V/K segments (but some lane of the warp nearly every time)
            uint32_t a = val + 4u * (lds_u16(pre + 2u * (e0 >> 5)) + __popc(word & ((1u << sh) - 1u)));
#pragma unroll
            for (int k = 0; k < K; ++k)
              if ((bits >> k) & 1u) {
                f[k] = lds_f32(a);
                a += 4u;
              }
          }
        } else {
#pragma unroll
          for (int k = 0; k < K; ++k) ctc_patch_elem(f[k], e0 + k, bm, pre, val);
        }
        stg_stream16(oseg0 + idx, Elem<T>::pack(f));
      } else {  // edge segments shared with the neighbouring rows: element stores
#pragma unroll
        for (int k = 0; k < K; ++k) {
          int e = e0 + k;
          if (e >= 0 && e < V) {
            ctc_patch_elem(f[k], e, bm, pre, val);
            out[e] = Elem<T>::from_float(f[k]);
          }
        }
      }
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(kGradThreads, 3)
ctc_grad_kernel(const T* __restrict__ logits, const float* __restrict__ lse,
                const float* __restrict__ lp_ext, const float* __restrict__ alpha,
                const float* __restrict__ beta, const double* __restrict__ frame_off,
                const float* __restrict__ scale, const long long* __restrict__ ys_pad,
                const int32_t* __restrict__ hlens, const int32_t* __restrict__ ys_lens,
                const int32_t* __restrict__ leader, int B, int Tm, int V, int Lmax, int blank,
                T* __restrict__ grad) {
  extern __shared__ __align__(16) float smem_f[];
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int Smax = 2 * Lmax + 1;
  const GradLayout L = ctc_grad_layout(Lmax, V);
  const int gpu = (Tm + kRowWarps - 1) / kRowWarps;  // groups per utterance
  const int ngroups = B * gpu;
  float* bufs = smem_f + 3 * kPatchWarps * L.nst;
  // named barriers: full[buf] = 1 + buf (patch warps arrive, streaming warps wait),
  //                 empty[buf] = 3 + buf (streaming warps arrive, patch warps wait),
  //                 5: among the patch warps (class bitmap of the group built by the first of them)

  if (warp >= kRowWarps) {  // ---- patch warps: one group ahead of the streaming warps ----
    const int pw = warp - kRowWarps;  // takes rows pw * kPatchRows .. of every group
    float* st_al = smem_f + 3 * pw * L.nst;
    float* st_be = st_al + L.nst;
    float* st_lp = st_be + L.nst;
    constexpr int kOcc = 8;  // states s = lane + 32 i, i < kOcc, keep leader / class / rank in registers
    int pd[kOcc], pc[kOcc], pr[kOcc];
    int cur_b = -1;
    bool dups = false;  // the utterance repeats a label (among its first 32 kOcc states)
    int k = 0;
    for (int g = blockIdx.x; g < ngroups; g += gridDim.x, ++k) {
      const int buf = k & 1;
      float* val_all = bufs + (size_t)buf * L.nbuf;
      float* info = val_all + L.nval;
      uint32_t* bm = reinterpret_cast<uint32_t*>(info + kGradInfo);
      uint16_t* pre = reinterpret_cast<uint16_t*>(bm + L.nbm);
      const int b = g / gpu;
      const int t0 = (g - b * gpu) * kRowWarps;
      const int nrows = Tm - t0 < kRowWarps ? Tm - t0 : kRowWarps;
      const long long row0 = (long long)b * Tm + t0;
      // (1) alpha / beta / lp_ext of this warp's rows: one contiguous piece of each array
      const int w0 = pw * kPatchRows;
      const int myrows = nrows - w0 < kPatchRows ? (nrows - w0 > 0 ? nrows - w0 : 0) : kPatchRows;
      const float* g_al = alpha + (row0 + w0) * Smax;
      const float* g_be = beta + (row0 + w0) * Smax;
      const float* g_lp = lp_ext + (row0 + w0) * Smax;
      const int ph_al = (int)((uintptr_t)g_al & 15) >> 2, ph_be = (int)((uintptr_t)g_be & 15) >> 2,
                ph_lp = (int)((uintptr_t)g_lp & 15) >> 2;
      cp_async_span(st_al + ph_al, g_al, myrows * Smax, lane);
      cp_async_span(st_be + ph_be, g_be, myrows * Smax, lane);
      cp_async_span(st_lp + ph_lp, g_lp, myrows * Smax, lane);
      // (2) per-row scalars (lane i: this warp's row i) and the utterance's leader / class per state
      const float sc = scale[b];
      const int hl = hlens[b];
      const int S = 2 * ys_lens[b] + 1;
      float my_nl = 0.f, my_c0 = 0.f;
      if (lane < myrows) {
        const int t = t0 + w0 + lane;
        // alpha_hat + beta_hat + (off_a[t] + off_b[t] - ll): the bracket is O(1) because
        // sum_s alpha_t(s) beta_t(s) / y_t(l'_s) = exp(ll) for every t
        const double* oa = frame_off + (long long)b * 2 * (Tm + 1);
        my_nl = (float)(oa[t] + oa[Tm + 1 + t] - oa[Tm]);
        my_c0 = -lse[row0 + w0 + lane] * kLog2e;
      }
      const int32_t* ld = leader + (long long)b * Smax;
      const long long* ys = ys_pad + (long long)b * Lmax;
      if (b != cur_b) {
        cur_b = b;
#pragma unroll
        for (int i = 0; i < kOcc; ++i) {
          const int s = lane + 32 * i;
          pd[i] = 0;
          pc[i] = -1;
          if (s < S) {
            pd[i] = ld[s];
            if (s & 1) pc[i] = (int)ys[s >> 1];
          }
        }
#pragma unroll
        bool d = false;
#pragma unroll
        for (int i = 0; i < kOcc; ++i) {  // only the first state of each distinct label patches
          const int s = lane + 32 * i;
          d |= s < S && (s & 1) && pd[i] != s;
          if (pd[i] != s || pc[i] >= V || pc[i] == blank) pc[i] = -1;
        }
        dups = __any_sync(0xffffffffu, d);
      }
      // (3) this buffer's previous group (k - 2) has been streamed
      if (k >= 2) named_bar_sync(3 + buf, kGradThreads);
      // class bitmap of utterance b and its running popcount (first patch warp), the rank of every
      // patching state (all of them)
      if (pw == 0) {
        for (int i = lane; i < L.nbm; i += 32) bm[i] = 0u;
        __syncwarp();
#pragma unroll
        for (int i = 0; i < kOcc; ++i)
          if (pc[i] >= 0) atomicOr(&bm[pc[i] >> 5], 1u << (pc[i] & 31));
        for (int s = lane + 32 * kOcc; s < S; s += 32) {  // longer label sequences
          if (!(s & 1) || ld[s] != s) continue;
          const int c = (int)ys[s >> 1];
          if (c >= 0 && c < V && c != blank) atomicOr(&bm[c >> 5], 1u << (c & 31));
        }
        if (lane == 0) atomicOr(&bm[blank >> 5], 1u << (blank & 31));
        __syncwarp();
        {
          const int chunk = (L.nbm + 31) / 32;
          int cnt = 0;
          for (int j = 0; j < chunk; ++j) {
            const int w = lane * chunk + j;
            if (w < L.nbm) cnt += __popc(bm[w]);
          }
          int incl = cnt;
#pragma unroll
          for (int d = 1; d < 32; d <<= 1) {
            const int n = __shfl_up_sync(0xffffffffu, incl, d);
            if (lane >= d) incl += n;
          }
          int run = incl - cnt;
          for (int j = 0; j < chunk; ++j) {
            const int w = lane * chunk + j;
            if (w < L.nbm) {
              pre[w] = (uint16_t)run;
              run += __popc(bm[w]);
            }
          }
        }
        __syncwarp();
      }
      named_bar_sync(5, kPatchWarps * 32);
      auto rank_of = [&](int c) { return (int)pre[c >> 5] + __popc(bm[c >> 5] & ((1u << (c & 31)) - 1u)); };
#pragma unroll
      for (int i = 0; i < kOcc; ++i) pr[i] = pc[i] >= 0 ? rank_of(pc[i]) : -1;
      const int rank_blank = rank_of(blank);
      asm volatile("cp.async.wait_all;" ::: "memory");
      __syncwarp();
      // (4) this warp's rows of the group
      for (int r = 0; r < myrows; ++r) {
        const int w = w0 + r;
        const bool live = t0 + w < hl && sc != 0.f;
        const float nl = __shfl_sync(0xffffffffu, my_nl, r);
        const float c0 = __shfl_sync(0xffffffffu, my_c0, r);
        if (lane == 0) {
          info[w] = c0;
          info[kRowWarps + w] = live ? sc : 0.f;  // 0: the streaming warp writes a row of zeros
        }
        if (!live) continue;
        float* al = st_al + ph_al + r * Smax;  // becomes the occupancy accumulator (repeated labels / long sequences)
        float* be = st_be + ph_be + r * Smax;  // becomes the state posterior (long sequences)
        const float* lp = st_lp + ph_lp + r * Smax;
        float* val = val_all + (size_t)w * (Lmax + 1);
        // alpha*beta counts the emission at t twice: take lp out once.  The first state of a label
        // (leader) owns its occupancy slot and stores into it; the other states of a repeated label add
        // to it afterwards - none in most utterances (dups, a vote per utterance), so that the common row
        // is: three shared-memory reads, one exp per state, registers from there on.
        float post[kOcc], pl[kOcc];
        float occ_blank = 0.f;
#pragma unroll
        for (int i = 0; i < kOcc; ++i) {
          const int s = lane + 32 * i;
          post[i] = pl[i] = 0.f;
          if (s < S) {
            pl[i] = lp[s];
            post[i] = expf(al[s] + be[s] + nl - pl[i]);
            if (!(s & 1)) occ_blank += post[i];
          }
        }
        for (int s = lane + 32 * kOcc; s < S; s += 32) {  // longer label sequences: through shared memory
          const float p = expf(al[s] + be[s] + nl - lp[s]);
          be[s] = p;
          if (!(s & 1)) occ_blank += p;
        }
        if (dups || S > 32 * kOcc) {
          __syncwarp();
#pragma unroll
          for (int i = 0; i < kOcc; ++i) {
            const int s = lane + 32 * i;
            if (s < S && (s & 1) && pd[i] == s) al[s] = post[i];
          }
          for (int s = lane + 32 * kOcc; s < S; s += 32)
            if ((s & 1) && ld[s] == s) al[s] = be[s];
          __syncwarp();
#pragma unroll
          for (int i = 0; i < kOcc; ++i) {
            const int s = lane + 32 * i;
            if (s < S && (s & 1) && pd[i] != s) atomicAdd(&al[pd[i]], post[i]);
          }
          for (int s = lane + 32 * kOcc; s < S; s += 32)
            if ((s & 1) && ld[s] != s) atomicAdd(&al[ld[s]], be[s]);
          __syncwarp();
#pragma unroll
          for (int i = 0; i < kOcc; ++i) {
            const int s = lane + 32 * i;
            if (s < S && (s & 1) && pd[i] == s) post[i] = al[s];
          }
        }
        occ_blank = warp_sum(occ_blank);
        // occupancy -> gradient entry at the rank of its class
#pragma unroll
        for (int i = 0; i < kOcc; ++i)
          if (pr[i] >= 0) val[pr[i]] = sc * (expf(pl[i]) - post[i]);
        for (int s = lane + 32 * kOcc; s < S; s += 32) {
          if (!(s & 1) || ld[s] != s) continue;
          const int c = (int)ys[s >> 1];
          if (c >= 0 && c < V && c != blank) val[rank_of(c)] = sc * (expf(lp[s]) - al[s]);
        }
        if (lane == 0) val[rank_blank] = sc * (expf(lp[0]) - occ_blank);
      }
      __syncwarp();
      __threadfence_block();
      named_bar_arrive(1 + buf, kGradThreads);
    }
    return;
  }

  // ---- streaming warps ----
  int k = 0;
  for (int g = blockIdx.x; g < ngroups; g += gridDim.x, ++k) {
    const int buf = k & 1;
    const float* val_all = bufs + (size_t)buf * L.nbuf;
    const float* info = val_all + L.nval;
    const uint32_t* bm = reinterpret_cast<const uint32_t*>(info + kGradInfo);
    const uint16_t* pre = reinterpret_cast<const uint16_t*>(bm + L.nbm);
    named_bar_sync(1 + buf, kGradThreads);
    const int b = g / gpu;
    const int t = (g - b * gpu) * kRowWarps + warp;
    if (t < Tm) {
      const long long rowi = (long long)b * Tm + t;
      const T* row = logits + rowi * V;
      T* grow = grad + rowi * V;
      const float sc = info[kRowWarps + warp];
      if (sc == 0.f)
        warp_row_softmax_store<T>(row, grow, V, lane, true, 0.f, 0.f, -1.f);
      else
        ctc_grad_row<T>(row, grow, V, lane, info[warp], sc, (uint32_t)__cvta_generic_to_shared(bm),
                        (uint32_t)__cvta_generic_to_shared(pre),
                        (uint32_t)__cvta_generic_to_shared(val_all + (size_t)warp * (Lmax + 1)));
    }
    if ((long long)g + 2ll * gridDim.x < ngroups) named_bar_arrive(3 + buf, kGradThreads);
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
  const int nw = (Smax + 1 + 63) / 64;  // 64 states per warp (two per lane); beta's pairing starts at state -1
  const size_t smem = (size_t)nw * (kCtcRing + 4) * 8 + ((size_t)Smax + 1) * sizeof(int) + 2 * sizeof(float);
  ctc_lattice_kernel<<<dim3(B, 2), nw * 32, smem, (cudaStream_t)stream>>>(
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
  cudaStream_t st = (cudaStream_t)stream;
  const long long* ys = reinterpret_cast<const long long*>(ys_pad);
  int dev = 0, sms = 0, smem_max = 0;
  B200_CHECK_CUDA(cudaGetDevice(&dev));
  B200_CHECK_CUDA(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev));
  B200_CHECK_CUDA(cudaDeviceGetAttribute(&smem_max, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  const GradLayout L = ctc_grad_layout(Lmax, V);  // 28 KB at cfg5; 1 CTA per SM from about 2 Lmax + 1 = 700 states
  B200_CHECK_ARG(L.bytes <= (size_t)smem_max, "b200_ctc_grad: %zu bytes of shared memory needed, %d available",
                 L.bytes, smem_max);
  const long long ngroups = (long long)B * ((T + kRowWarps - 1) / kRowWarps);
  B200_CHECK_ARG(ngroups < (1ll << 31), "b200_ctc_grad: B*T too large");
#define B200_CTC_GRAD_LAUNCH(TYPE)                                                                           \
  do {                                                                                                       \
    auto* fn = ctc_grad_kernel<TYPE>;                                                                        \
    B200_CHECK_CUDA(cudaFuncSetAttribute(fn, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)L.bytes));    \
    int per_sm = 0;                                                                                          \
    B200_CHECK_CUDA(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, kGradThreads, L.bytes));      \
    if (per_sm < 1) per_sm = 1;                                                                              \
    const long long grid = ngroups < (long long)sms * per_sm ? ngroups : (long long)sms * per_sm;            \
    fn<<<(unsigned)grid, kGradThreads, L.bytes, st>>>((const TYPE*)logits, lse, lp_ext, alpha, beta,         \
                                                      frame_off, scale, ys, hlens, ys_lens, leader, B, T, V, \
                                                      Lmax, blank, (TYPE*)grad);                             \
  } while (0)
  switch (dtype) {
    case B200_F32:
      B200_CTC_GRAD_LAUNCH(float);
      break;
    case B200_F16:
      B200_CTC_GRAD_LAUNCH(__half);
      break;
    case B200_BF16:
      B200_CTC_GRAD_LAUNCH(__nv_bfloat16);
      break;
    default:
      B200_CHECK_ARG(false, "b200_ctc_grad: unsupported dtype %d", dtype);
  }
#undef B200_CTC_GRAD_LAUNCH
  B200_CHECK_LAUNCH("ctc_grad_kernel");
  return B200_OK;
}
