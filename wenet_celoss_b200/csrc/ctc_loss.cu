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
