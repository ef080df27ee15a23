// Small kernels around the fp32-accurate (split-bf16) tensor-core path of csrc/split_gemm.cu:
//   split_f32_kernel        x (fp32) -> x_hi = bf16(x), x_lo = bf16(x - x_hi)
//   joint_h_split_kernel    H = tanh(e[b,t,:] + p[b,u,:]) in fp32 (joint.py:58-64) -> H_hi, H_lo
//   joint_pick_kernel       the two logits per lattice cell the RNN-T lattice needs (blank, target), as exact
//                           fp32 dot products -> lp_blank, lp_emit = logit - lse, and the raw logits for the fix-up
//   ctc_lp_ext_kernel       log-probs of the blank-interleaved label states, exact fp32 dot products of
//                           hs[b,t,:] with the gathered ctc_lo rows (ctc.py:57-60) minus lse
//   rnnt_fixup_kernel       G[cell, blank] -= scale exp(x_b + cb), G[cell, y] -= scale exp(x_y + ce) on the
//                           split G written by split_gemm MODE_EXP (the sparse terms of the rnnt_loss gradient)
//   ctc_fixup_kernel        G[b,t,c] -= scale_b * posterior(c at t) for the blank and every distinct label
//   colsum_split_kernel     d bias[c] = sum_r w[r] (G_hi + G_lo)[r, c]   (w optional)
//   joint_patch_split_kernel / hs_split_kernel   the no-recompute backward in fp32 mode (cf. joint_nr.cu):
//                           rowscale[c] = scale exp(c0), the blank / target entries of the split Pt = exp(logit)
//                           replaced by exp(x) - exp(x + c_b - c_0) (exact fp32 logits from joint_pick_kernel),
//                           Hs = rowscale * (H_hi + H_lo) split again
// All of them are small next to the GEMMs (gathered dot products: 2 or L+1 classes per row out of V).
#include "common.cuh"

namespace b200 {
namespace sa {

__device__ __forceinline__ void split1(float x, __nv_bfloat16& hi, __nv_bfloat16& lo) {
  hi = __float2bfloat16_rn(x);
  lo = __float2bfloat16_rn(x - __bfloat162float(hi));
}

__global__ void __launch_bounds__(256)
split_f32_kernel(const float* __restrict__ x, long long rows, int cols, long long ldx, __nv_bfloat16* __restrict__ hi,
                 __nv_bfloat16* __restrict__ lo, long long ldo) {
  const long long total = rows * cols;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / cols;
    const int c = (int)(i - r * cols);
    __nv_bfloat16 h, l;
    split1(x[r * ldx + c], h, l);
    hi[r * ldo + c] = h;
    lo[r * ldo + c] = l;
  }
}

// H[cell, 0:J] for all cells of B utterances; 8 elements per thread (16-byte stores)
__global__ void __launch_bounds__(256)
joint_h_split_kernel(const float* __restrict__ e, const float* __restrict__ p, int T, int U1, int J, long long ncells,
                     __nv_bfloat16* __restrict__ Hhi, __nv_bfloat16* __restrict__ Hlo, int ldh) {
  const int g8 = J >> 3;
  const long long total = ncells * g8;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long cell = i / g8;
    const int k0 = (int)(i - cell * g8) << 3;
    const long long bt = cell / U1;
    const int u = (int)(cell - bt * U1);
    const long long b = bt / T;
    const float4* ep = reinterpret_cast<const float4*>(e + bt * J + k0);
    const float4* pp = reinterpret_cast<const float4*>(p + (b * U1 + u) * J + k0);
    const float4 e0 = __ldg(ep), e1 = __ldg(ep + 1), p0 = __ldg(pp), p1 = __ldg(pp + 1);
    const float v[8] = {tanhf(e0.x + p0.x), tanhf(e0.y + p0.y), tanhf(e0.z + p0.z), tanhf(e0.w + p0.w),
                        tanhf(e1.x + p1.x), tanhf(e1.y + p1.y), tanhf(e1.z + p1.z), tanhf(e1.w + p1.w)};
    __align__(16) __nv_bfloat16 h[8], l[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) split1(v[k], h[k], l[k]);
    *reinterpret_cast<uint4*>(Hhi + cell * ldh + k0) = *reinterpret_cast<const uint4*>(h);
    *reinterpret_cast<uint4*>(Hlo + cell * ldh + k0) = *reinterpret_cast<const uint4*>(l);
  }
}

// one warp per lattice cell: x_blank, x_target as fp32 dot products of tanh(e + p) with two rows of W_out
__global__ void __launch_bounds__(256)
joint_pick_kernel(const float* __restrict__ e, const float* __restrict__ p, const float* __restrict__ W,
                  const float* __restrict__ bias, const int32_t* __restrict__ targets,
                  const int32_t* __restrict__ logit_len, const int32_t* __restrict__ target_len, int T, int U1, int J,
                  int blank, long long ncells, const float* __restrict__ lse, float* __restrict__ lp_blank,
                  float* __restrict__ lp_emit, float2* __restrict__ xbe) {
  const long long cell = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (cell >= ncells) return;
  const long long bt = cell / U1;
  const int u = (int)(cell - bt * U1);
  const int b = (int)(bt / T), t = (int)(bt - (long long)b * T);
  if (t >= logit_len[b] || u > target_len[b]) return;
  const int y = u < target_len[b] ? targets[(long long)b * (U1 - 1) + u] : -1;
  const float* er = e + bt * J;
  const float* pr = p + ((long long)b * U1 + u) * J;
  const float* wb = W + (long long)blank * J;
  const float* wy = W + (long long)(y >= 0 ? y : blank) * J;
  float sb = 0.f, sy = 0.f;
  for (int k = lane * 4; k < J; k += 128) {
    const float4 ev = __ldg(reinterpret_cast<const float4*>(er + k));
    const float4 pv = __ldg(reinterpret_cast<const float4*>(pr + k));
    const float4 a = __ldg(reinterpret_cast<const float4*>(wb + k));
    const float4 c = __ldg(reinterpret_cast<const float4*>(wy + k));
    const float h0 = tanhf(ev.x + pv.x), h1 = tanhf(ev.y + pv.y), h2 = tanhf(ev.z + pv.z), h3 = tanhf(ev.w + pv.w);
    sb += h0 * a.x + h1 * a.y + h2 * a.z + h3 * a.w;
    sy += h0 * c.x + h1 * c.y + h2 * c.z + h3 * c.w;
  }
  sb = warp_sum(sb);
  sy = warp_sum(sy);
  if (lane == 0) {
    const float l = lse[cell];
    const float xb = sb + bias[blank];
    const float xy = y >= 0 ? sy + bias[y] : 0.f;
    lp_blank[cell] = xb - l;
    lp_emit[cell] = y >= 0 ? xy - l : -INFINITY;
    xbe[cell] = make_float2(xb, xy);
  }
}

// lp_ext as a tiled fp32 GEMM per utterance: C[frame, slot] = hs[b, frame, :] . ctc_lo.weight[class(slot), :].
// One CTA = 32 frames x 128 slots (127 labels l0 .. l0 + 126 of the utterance and, in slot 127, the blank),
// 256 threads with 4 x 4 register tiles, operands staged k-major in shared memory in blocks of 16 k (two
// stages, global -> register prefetch of the next block under the multiplies of the current one).  The
// states the CTA owns, 2 l0 .. 2 l0 + 253 (blank before each of its labels, the labels, and the last blank),
// are contiguous in a row of lp_ext: the tile goes through shared memory and out in 128-byte lines.
// Before: one thread per dot product, two shared-memory reads per FMA: 717 us at cfg5 - the largest kernel
// of the fused CTC module, for 3.3 of its 1500 GFLOP.
constexpr int kLpFrames = 32;
constexpr int kLpSlots = 128;   // 127 labels + blank
constexpr int kLpBK = 16;
__global__ void __launch_bounds__(256)
ctc_lp_ext_kernel(const float* __restrict__ hs, const float* __restrict__ W, const float* __restrict__ bias,
                  const long long* __restrict__ ys_pad, const int32_t* __restrict__ hlens,
                  const int32_t* __restrict__ ys_lens, int Tm, int E, int V, int Lmax, int blank,
                  const float* __restrict__ lse, float* __restrict__ lp_ext) {
  __shared__ __align__(16) float As[2][kLpBK][kLpFrames + 4];
  __shared__ __align__(16) float Ws[2][kLpBK][kLpSlots + 4];
  __shared__ int s_cls[kLpSlots];
  const int tid = threadIdx.x;
  const int b = blockIdx.y;
  const int t0 = blockIdx.x * kLpFrames;
  const int Tb = hlens[b];
  const int L = ys_lens[b];
  const int l0 = blockIdx.z * (kLpSlots - 1);
  if (t0 >= Tb || l0 > L) return;  // l0 == L: only the final blank state 2 L is left for this CTA
  const long long* ys = ys_pad + (long long)b * Lmax;
  if (tid < kLpSlots) {  // class of every slot; -1: no such label (beyond L, or an index outside [0, V))
    int cls = -1;
    if (tid == kLpSlots - 1) {
      cls = blank;
    } else if (l0 + tid < L) {
      const long long y = ys[l0 + tid];
      if (y >= 0 && y < V) cls = (int)y;
    }
    s_cls[tid] = cls;
  }
  __syncthreads();
  const int nfr = min(kLpFrames, Tb - t0);
  // loaders: thread -> (row tid / 4 (+ 64), four consecutive k at (tid % 4) * 4)
  const int lk = (tid & 3) * 4;
  const int lr = tid >> 2;
  const float* a_src = lr < nfr ? hs + ((long long)b * Tm + t0 + lr) * E : nullptr;  // rows 0..31: threads 0..127
  const float* w_src[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int cls = s_cls[lr + 64 * i];
    w_src[i] = cls >= 0 ? W + (long long)cls * E : nullptr;
  }
  float4 ra, rw[2];
  auto load_regs = [&](int k0) {
    ra = make_float4(0.f, 0.f, 0.f, 0.f);
    if (tid < 4 * kLpFrames && a_src) ra = __ldg(reinterpret_cast<const float4*>(a_src + k0 + lk));
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      rw[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (w_src[i]) rw[i] = __ldg(reinterpret_cast<const float4*>(w_src[i] + k0 + lk));
    }
  };
  auto store_regs = [&](int st) {
    if (tid < 4 * kLpFrames) {
      As[st][lk + 0][lr] = ra.x;
      As[st][lk + 1][lr] = ra.y;
      As[st][lk + 2][lr] = ra.z;
      As[st][lk + 3][lr] = ra.w;
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      Ws[st][lk + 0][lr + 64 * i] = rw[i].x;
      Ws[st][lk + 1][lr + 64 * i] = rw[i].y;
      Ws[st][lk + 2][lr + 64 * i] = rw[i].z;
      Ws[st][lk + 3][lr + 64 * i] = rw[i].w;
    }
  };
  const int tx = tid & 31, ty = tid >> 5;  // slots tx*4..+3, frames ty*4..+3
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  load_regs(0);
  store_regs(0);
  __syncthreads();
  int st = 0;
  for (int k0 = 0; k0 < E; k0 += kLpBK) {  // E is a multiple of 16 (checked on the host)
    const bool more = k0 + kLpBK < E;
    if (more) load_regs(k0 + kLpBK);
#pragma unroll
    for (int k = 0; k < kLpBK; ++k) {
      float av[4], wv[4];
      *reinterpret_cast<float4*>(av) = *reinterpret_cast<const float4*>(&As[st][k][ty * 4]);
      *reinterpret_cast<float4*>(wv) = *reinterpret_cast<const float4*>(&Ws[st][k][tx * 4]);
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], wv[j], acc[i][j]);
    }
    if (more) store_regs(st ^ 1);
    __syncthreads();
    st ^= 1;
  }
  // epilogue: lp = dot + bias - lse into the tile [frame][slot] (aliases the W stages), then rows of lp_ext
  float* tile = &Ws[0][0][0];  // 32 x (128 + 4) floats = the two W stages
  constexpr int kPitch = kLpSlots + 4;
  static_assert(kLpFrames * kPitch <= 2 * kLpBK * (kLpSlots + 4), "tile must fit the W stages");
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int cls = s_cls[tx * 4 + j];
    const float bj = cls >= 0 ? __ldg(bias + cls) : 0.f;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int f = ty * 4 + i;
      float v = -INFINITY;
      if (cls >= 0 && f < nfr) v = acc[i][j] + bj - lse[(long long)b * Tm + t0 + f];
      tile[f * kPitch + tx * 4 + j] = v;
    }
  }
  __syncthreads();
  const int Smax = 2 * Lmax + 1;
  const int s_lo = 2 * l0;
  const int s_hi = min(2 * L + 1, s_lo + 2 * (kLpSlots - 1));  // states [s_lo, s_hi) belong to this CTA
  const int ns = s_hi - s_lo;
  for (int o = tid; o < nfr * ns; o += 256) {
    const int f = o / ns, c = o - f * ns;
    // even state: blank (slot 127); odd state 2 l + 1: label l = l0 + c / 2
    const float v = (c & 1) ? tile[f * kPitch + (c >> 1)] : tile[f * kPitch + kLpSlots - 1];
    lp_ext[((long long)b * Tm + t0 + f) * Smax + s_lo + c] = v;
  }
}

__device__ __forceinline__ float split_get(const __nv_bfloat16* hi, const __nv_bfloat16* lo, long long i) {
  return __bfloat162float(hi[i]) + (lo ? __bfloat162float(lo[i]) : 0.f);
}
__device__ __forceinline__ void split_put(__nv_bfloat16* hi, __nv_bfloat16* lo, long long i, float v) {
  __nv_bfloat16 h, l;
  split1(v, h, l);
  hi[i] = h;
  if (lo) lo[i] = l;
}

// one thread per lattice cell
__global__ void __launch_bounds__(256)
rnnt_fixup_kernel(const float4* __restrict__ coef, const float2* __restrict__ xbe, const int32_t* __restrict__ targets,
                  int T, int U1, int blank, long long ncells, __nv_bfloat16* __restrict__ Ghi,
                  __nv_bfloat16* __restrict__ Glo, long long ldg) {
  const long long cell = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (cell >= ncells) return;
  const float4 c = coef[cell];
  if (c.w == 0.f) return;
  const float2 x = xbe[cell];
  const long long bt = cell / U1;
  const int u = (int)(cell - bt * U1);
  const int b = (int)(bt / T);
  // same if / else-if order as torchaudio when a target equals blank: both terms land on one class
  float fb = c.y != -INFINITY ? c.w * expf(x.x + c.y) : 0.f;
  int y = -1;
  float fe = 0.f;
  if (c.z != -INFINITY) {
    y = targets[(long long)b * (U1 - 1) + u];
    fe = c.w * expf(x.y + c.z);
  }
  const long long row = cell * ldg;
  if (y == blank) {
    fb += fe;
    y = -1;
  }
  if (fb != 0.f) split_put(Ghi, Glo, row + blank, split_get(Ghi, Glo, row + blank) - fb);
  if (y >= 0) split_put(Ghi, Glo, row + y, split_get(Ghi, Glo, row + y) - fe);
}

// one thread per lattice cell: rowscale and the two patched entries of the split Pt (see joint_nr.cu)
__global__ void __launch_bounds__(256)
joint_patch_split_kernel(const float4* __restrict__ coef, const float2* __restrict__ xbe,
                         const int32_t* __restrict__ targets, int T, int U1, int blank, long long ncells,
                         __nv_bfloat16* __restrict__ Phi, __nv_bfloat16* __restrict__ Plo, long long ldg,
                         float* __restrict__ rowscale) {
  const long long cell = blockIdx.x * (long long)blockDim.x + threadIdx.x;
  if (cell >= ncells) return;
  const float4 c = coef[cell];
  const float rs = c.x == -INFINITY ? 0.f : c.w * expf(c.x);
  rowscale[cell] = rs;
  if (rs == 0.f) return;
  const float2 x = xbe[cell];
  const bool has_b = c.y != -INFINITY, has_e = c.z != -INFINITY;
  int y = -1;
  if (has_e) {
    const long long bt = cell / U1;
    const int u = (int)(cell - bt * U1);
    y = targets[(bt / T) * (U1 - 1) + u];
  }
  const long long row = cell * ldg;
  if (has_b) {
    float v = expf(x.x) - expf(x.x + (c.y - c.x));
    if (y == blank) v -= expf(x.x + (c.z - c.x));
    split_put(Phi, Plo, row + blank, v);
  }
  if (has_e && !(y == blank && has_b)) split_put(Phi, Plo, row + y, expf(x.y) - expf(x.y + (c.z - c.x)));
}

// Hs = rowscale * (H_hi + H_lo), split again; 8 elements per thread
__global__ void __launch_bounds__(256)
hs_split_kernel(const float* __restrict__ rowscale, const __nv_bfloat16* __restrict__ Hhi,
                const __nv_bfloat16* __restrict__ Hlo, int J, long long ncells, int ldh,
                __nv_bfloat16* __restrict__ Shi, __nv_bfloat16* __restrict__ Slo) {
  const int g8 = J >> 3;
  const long long total = ncells * g8;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long cell = i / g8;
    const int k0 = (int)(i - cell * g8) << 3;
    const float rs = __ldg(rowscale + cell);
    const uint4 a = __ldg(reinterpret_cast<const uint4*>(Hhi + cell * ldh + k0));
    const uint4 b = __ldg(reinterpret_cast<const uint4*>(Hlo + cell * ldh + k0));
    const uint32_t* pa = &a.x;
    const uint32_t* pb = &b.x;
    __align__(16) __nv_bfloat16 h[8], l[8];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float v0 = (__uint_as_float(pa[k] << 16) + __uint_as_float(pb[k] << 16)) * rs;
      const float v1 = (__uint_as_float(pa[k] & 0xffff0000u) + __uint_as_float(pb[k] & 0xffff0000u)) * rs;
      split1(v0, h[2 * k], l[2 * k]);
      split1(v1, h[2 * k + 1], l[2 * k + 1]);
    }
    *reinterpret_cast<uint4*>(Shi + cell * ldh + k0) = *reinterpret_cast<const uint4*>(h);
    *reinterpret_cast<uint4*>(Slo + cell * ldh + k0) = *reinterpret_cast<const uint4*>(l);
  }
}

// one warp per (b, t): posteriors of the extended states, merged per class through the leader table
__global__ void __launch_bounds__(256)
ctc_fixup_kernel(const float* __restrict__ lp_ext, const float* __restrict__ alpha, const float* __restrict__ beta,
                 const double* __restrict__ frame_off, const float* __restrict__ scale,
                 const long long* __restrict__ ys_pad, const int32_t* __restrict__ hlens,
                 const int32_t* __restrict__ ys_lens, const int32_t* __restrict__ leader, int B, int Tm, int V, int Lmax,
                 int blank, __nv_bfloat16* __restrict__ Ghi, __nv_bfloat16* __restrict__ Glo, long long ldg) {
  extern __shared__ float occ_all[];  // [8][Smax]
  const long long rowi = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (rowi >= (long long)B * Tm) return;
  const int Smax = 2 * Lmax + 1;
  float* occ = occ_all + (threadIdx.x >> 5) * Smax;
  const int b = (int)(rowi / Tm), t = (int)(rowi - (long long)b * Tm);
  const float sc = scale[b];
  if (t >= hlens[b] || sc == 0.f) return;
  const int S = 2 * ys_lens[b] + 1;
  const double* oa = frame_off + (long long)b * 2 * (Tm + 1);
  const float nl = (float)(oa[t] + oa[Tm + 1 + t] - oa[Tm]);
  const float* al = alpha + rowi * Smax;
  const float* be = beta + rowi * Smax;
  const float* lp = lp_ext + rowi * Smax;
  const int32_t* ld = leader + (long long)b * Smax;
  const long long* ys = ys_pad + (long long)b * Lmax;
  for (int s = lane; s < S; s += 32) occ[s] = 0.f;
  __syncwarp();
  float occ_blank = 0.f;
  for (int s = lane; s < S; s += 32) {
    const float post = expf(al[s] + be[s] + nl - lp[s]);  // alpha*beta counts the emission at t twice
    if (s & 1)
      atomicAdd(&occ[ld[s]], post);
    else
      occ_blank += post;
  }
  occ_blank = warp_sum(occ_blank);
  __syncwarp();
  const long long row = rowi * ldg;
  if (lane == 0) split_put(Ghi, Glo, row + blank, split_get(Ghi, Glo, row + blank) - sc * occ_blank);
  for (int s = 1 + 2 * lane; s < S; s += 64) {
    if (ld[s] != s) continue;  // the first state of each distinct label carries the class total
    const int c = (int)ys[s >> 1];
    if (c < 0 || c >= V || c == blank) continue;
    split_put(Ghi, Glo, row + c, split_get(Ghi, Glo, row + c) - sc * occ[s]);
  }
}

// The sparse part of the CTC gradient, in the form the EXP-mode GEMM epilogue subtracts on the fly
// (split_gemm.cu MODE_EXPSUB) instead of a read-modify-write pass over G (ctc_fixup_kernel above):
//   words [B][nwords] {bitmap over the classes: blank + the utterance's distinct valid labels,
//                      number of set bits in the words before this one}
//   val   [B][Lmax + 1][T]: val[b][rank(c)][t] = scale_b * posterior(class c at frame t), rank(c) = set bits below c
// (frame-minor, so that the 32 rows = frames of an epilogue warp read 128 contiguous bytes).
// One CTA = 32 frames of one utterance: warps form the state posteriors of four rows each (occupancies merged
// per class through the leader table in shared memory), then each (patching state, 32 frames) goes out as one
// 128-byte line.  Rows beyond hlens / with scale 0 get zeros.
constexpr int kPostFrames = 32;
__global__ void __launch_bounds__(256)
ctc_post_kernel(const float* __restrict__ lp_ext, const float* __restrict__ alpha, const float* __restrict__ beta,
                const double* __restrict__ frame_off, const float* __restrict__ scale,
                const long long* __restrict__ ys_pad, const int32_t* __restrict__ hlens,
                const int32_t* __restrict__ ys_lens, const int32_t* __restrict__ leader, int Tm, int V, int Lmax,
                int blank, int nwords, uint2* __restrict__ words, float* __restrict__ val) {
  extern __shared__ float post_sm[];  // occ [32][Smax] | bitmap [nwords] | prefix [nwords]
  const int Smax = 2 * Lmax + 1;
  float* occ_all = post_sm;
  uint32_t* bm = reinterpret_cast<uint32_t*>(post_sm + kPostFrames * Smax);
  uint32_t* pre = bm + nwords;
  const int b = blockIdx.y;
  const int t0 = blockIdx.x * kPostFrames;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const float sc = scale[b];
  const int hl = hlens[b];
  const int S = 2 * ys_lens[b] + 1;
  const int32_t* ld = leader + (long long)b * Smax;
  const long long* ys = ys_pad + (long long)b * Lmax;
  auto patch_class = [&](int s) -> int {  // class whose entry state s carries, -1: none
    if (s == 0) return blank;
    if (!(s & 1) || ld[s] != s) return -1;
    const long long c = ys[s >> 1];
    return (c >= 0 && c < V && c != blank) ? (int)c : -1;
  };
  // class bitmap and its running popcount
  for (int i = tid; i < nwords; i += 256) bm[i] = 0u;
  __syncthreads();
  for (int s = tid; s < S; s += 256) {
    const int c = patch_class(s);
    if (c >= 0) atomicOr(&bm[c >> 5], 1u << (c & 31));
  }
  __syncthreads();
  if (warp == 0) {
    const int chunk = (nwords + 31) / 32;
    int cnt = 0;
    for (int j = 0; j < chunk; ++j) {
      const int w = lane * chunk + j;
      if (w < nwords) cnt += __popc(bm[w]);
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
      if (w < nwords) {
        pre[w] = (uint32_t)run;
        run += __popc(bm[w]);
      }
    }
  }
  // state posteriors of this warp's rows
  const double* oa = frame_off + (long long)b * 2 * (Tm + 1);
  for (int r = warp * 4; r < warp * 4 + 4; ++r) {
    const int t = t0 + r;
    float* occ = occ_all + r * Smax;
    for (int s = lane; s < S; s += 32) occ[s] = 0.f;
    if (t >= Tm || t >= hl || sc == 0.f) continue;
    __syncwarp();
    const float nl = (float)(oa[t] + oa[Tm + 1 + t] - oa[Tm]);
    const long long rowi = (long long)b * Tm + t;
    const float* al = alpha + rowi * Smax;
    const float* be = beta + rowi * Smax;
    const float* lp = lp_ext + rowi * Smax;
    float occ_blank = 0.f;
    for (int s = lane; s < S; s += 32) {
      const float post = expf(al[s] + be[s] + nl - lp[s]);  // alpha*beta counts the emission at t twice
      if (s & 1)
        atomicAdd(&occ[ld[s]], post);
      else
        occ_blank += post;
    }
    occ_blank = warp_sum(occ_blank);
    __syncwarp();
    if (lane == 0) occ[0] = occ_blank;
  }
  __syncthreads();
  if (blockIdx.x == 0)
    for (int i = tid; i < nwords; i += 256) words[(long long)b * nwords + i] = make_uint2(bm[i], pre[i]);
  // lane = frame, warps walk the patching states
  const int t = t0 + lane;
  if (t < Tm) {
    float* vb = val + (long long)b * (Lmax + 1) * Tm + t;
    for (int s = warp; s < S; s += 8) {
      const int c = patch_class(s);
      if (c < 0) continue;
      const int rank = (int)pre[c >> 5] + __popc(bm[c >> 5] & ((1u << (c & 31)) - 1u));
      vb[(long long)rank * Tm] = sc * occ_all[lane * Smax + s];
    }
  }
}

// grid (column blocks of 256, row slabs); thread = one column pair over a slab of rows
__global__ void __launch_bounds__(128)
colsum_split_kernel(const __nv_bfloat16* __restrict__ Ghi, const __nv_bfloat16* __restrict__ Glo, long long rows, int N,
                    long long ldg, long long rows_per_slab, const float* __restrict__ w, float* __restrict__ out) {
  const int c = (blockIdx.x * 128 + threadIdx.x) * 2;
  if (c >= N) return;
  const long long r0 = blockIdx.y * rows_per_slab;
  const long long r1 = min(rows, r0 + rows_per_slab);
  float s0 = 0.f, s1 = 0.f;
  const bool pair_ok = c + 1 < N;
  for (long long r = r0; r < r1; ++r) {
    const __nv_bfloat162 h = *reinterpret_cast<const __nv_bfloat162*>(Ghi + r * ldg + c);
    float2 v = __bfloat1622float2(h);
    if (Glo) {
      const float2 lo2 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(Glo + r * ldg + c));
      v.x += lo2.x;
      v.y += lo2.y;
    }
    const float wr = w ? __ldg(w + r) : 1.f;
    s0 = fmaf(v.x, wr, s0);
    s1 = fmaf(v.y, wr, s1);
  }
  atomicAdd(out + c, s0);
  if (pair_ok) atomicAdd(out + c + 1, s1);
}

}  // namespace sa
}  // namespace b200

using namespace b200;

static int grid_for(long long items, int per_block, int cap) {
  long long g = (items + per_block - 1) / per_block;
  if (g > cap) g = cap;
  if (g < 1) g = 1;
  return (int)g;
}

extern "C" int b200_split_f32(const float* x, long long rows, int cols, long long ldx, void* hi, void* lo,
                              long long ldo, b200_stream_t stream) {
  B200_CHECK_ARG(x && hi && lo && rows > 0 && cols > 0 && ldx >= cols && ldo >= cols, "b200_split_f32: bad arguments");
  sa::split_f32_kernel<<<grid_for(rows * cols, 256, 148 * 16), 256, 0, (cudaStream_t)stream>>>(
      x, rows, cols, ldx, reinterpret_cast<__nv_bfloat16*>(hi), reinterpret_cast<__nv_bfloat16*>(lo), ldo);
  B200_CHECK_LAUNCH("split_f32_kernel");
  return B200_OK;
}

extern "C" int b200_joint_h_split(const float* e, const float* p, int B, int T, int U1, int J, void* H_hi, void* H_lo,
                                  int ldh, b200_stream_t stream) {
  B200_CHECK_ARG(e && p && H_hi && H_lo, "b200_joint_h_split: null pointer");
  B200_CHECK_ARG(B > 0 && T > 0 && U1 > 0 && J > 0 && J % 8 == 0 && ldh >= J && ldh % 8 == 0,
                 "b200_joint_h_split: bad shape (J and ldh multiples of 8)");
  B200_CHECK_ARG((((uintptr_t)e | (uintptr_t)p | (uintptr_t)H_hi | (uintptr_t)H_lo) & 15) == 0,
                 "b200_joint_h_split: pointers must be 16-byte aligned");
  const long long ncells = (long long)B * T * U1;
  sa::joint_h_split_kernel<<<grid_for(ncells * (J / 8), 256, 148 * 16), 256, 0, (cudaStream_t)stream>>>(
      e, p, T, U1, J, ncells, reinterpret_cast<__nv_bfloat16*>(H_hi), reinterpret_cast<__nv_bfloat16*>(H_lo), ldh);
  B200_CHECK_LAUNCH("joint_h_split_kernel");
  return B200_OK;
}

extern "C" int b200_joint_pick(const float* e, const float* p, const float* W, const float* bias,
                               const int32_t* targets, const int32_t* logit_len, const int32_t* target_len, int B,
                               int T, int U1, int V, int J, int blank, const float* lse, float* lp_blank,
                               float* lp_emit, float* xbe, b200_stream_t stream) {
  B200_CHECK_ARG(e && p && W && bias && logit_len && target_len && lse && lp_blank && lp_emit && xbe &&
                     (U1 == 1 || targets),
                 "b200_joint_pick: null pointer");
  B200_CHECK_ARG(J % 4 == 0 && blank >= 0 && blank < V, "b200_joint_pick: J %% 4, blank in range");
  B200_CHECK_ARG((((uintptr_t)e | (uintptr_t)p | (uintptr_t)W) & 15) == 0 && ((uintptr_t)xbe & 7) == 0,
                 "b200_joint_pick: e, p, W must be 16-byte aligned");
  const long long ncells = (long long)B * T * U1;
  sa::joint_pick_kernel<<<(unsigned)((ncells + 7) / 8), 256, 0, (cudaStream_t)stream>>>(
      e, p, W, bias, targets, logit_len, target_len, T, U1, J, blank, ncells, lse, lp_blank, lp_emit,
      reinterpret_cast<float2*>(xbe));
  B200_CHECK_LAUNCH("joint_pick_kernel");
  return B200_OK;
}

extern "C" int b200_ctc_lp_ext(const float* hs, const float* W, const float* bias, const int64_t* ys_pad,
                               const int32_t* hlens, const int32_t* ys_lens, int B, int T, int E, int V, int Lmax,
                               int blank, const float* lse, float* lp_ext, b200_stream_t stream) {
  B200_CHECK_ARG(hs && W && bias && hlens && ys_lens && lse && lp_ext && (Lmax == 0 || ys_pad),
                 "b200_ctc_lp_ext: null pointer");
  B200_CHECK_ARG(B > 0 && T > 0 && E > 0 && V > 0 && Lmax >= 0 && blank >= 0 && blank < V, "b200_ctc_lp_ext: bad shape");
  B200_CHECK_ARG(E % sa::kLpBK == 0, "b200_ctc_lp_ext: encoder dim E=%d must be a multiple of 16", E);
  B200_CHECK_ARG((((uintptr_t)hs | (uintptr_t)W) & 15) == 0, "b200_ctc_lp_ext: hs, W must be 16-byte aligned");
  // label blocks of 127 (+ blank) per CTA; the block that starts AT label L writes the final blank state
  const int nlb = Lmax / (sa::kLpSlots - 1) + 1;
  dim3 grid((T + sa::kLpFrames - 1) / sa::kLpFrames, B, nlb);
  sa::ctc_lp_ext_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(hs, W, bias, reinterpret_cast<const long long*>(ys_pad),
                                                                   hlens, ys_lens, T, E, V, Lmax, blank, lse, lp_ext);
  B200_CHECK_LAUNCH("ctc_lp_ext_kernel");
  return B200_OK;
}

extern "C" int b200_rnnt_fixup(const float* coef, const float* xbe, const int32_t* targets, int B, int T, int U1,
                               int V, int blank, void* G_hi, void* G_lo, long long ldg, b200_stream_t stream) {
  B200_CHECK_ARG(coef && xbe && G_hi && (U1 == 1 || targets), "b200_rnnt_fixup: null pointer");
  B200_CHECK_ARG(blank >= 0 && blank < V && ldg >= V, "b200_rnnt_fixup: bad arguments");
  const long long ncells = (long long)B * T * U1;
  sa::rnnt_fixup_kernel<<<(unsigned)((ncells + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const float4*>(coef), reinterpret_cast<const float2*>(xbe), targets, T, U1, blank, ncells,
      reinterpret_cast<__nv_bfloat16*>(G_hi), reinterpret_cast<__nv_bfloat16*>(G_lo), ldg);
  B200_CHECK_LAUNCH("rnnt_fixup_kernel");
  return B200_OK;
}

extern "C" int b200_ctc_fixup(const float* lp_ext, const float* alpha, const float* beta, const double* frame_off,
                              const float* scale, const int64_t* ys_pad, const int32_t* hlens,
                              const int32_t* ys_lens, const int32_t* leader, int B, int T, int V, int Lmax, int blank,
                              void* G_hi, void* G_lo, long long ldg, b200_stream_t stream) {
  B200_CHECK_ARG(lp_ext && alpha && beta && frame_off && scale && hlens && ys_lens && leader && G_hi,
                 "b200_ctc_fixup: null pointer");
  const int Smax = 2 * Lmax + 1;
  const long long rows = (long long)B * T;
  sa::ctc_fixup_kernel<<<(unsigned)((rows + 7) / 8), 256, 8 * Smax * sizeof(float), (cudaStream_t)stream>>>(
      lp_ext, alpha, beta, frame_off, scale, reinterpret_cast<const long long*>(ys_pad), hlens, ys_lens, leader, B, T, V,
      Lmax, blank, reinterpret_cast<__nv_bfloat16*>(G_hi), reinterpret_cast<__nv_bfloat16*>(G_lo), ldg);
  B200_CHECK_LAUNCH("ctc_fixup_kernel");
  return B200_OK;
}

extern "C" int b200_ctc_post(const float* lp_ext, const float* alpha, const float* beta, const double* frame_off,
                             const float* scale, const int64_t* ys_pad, const int32_t* hlens,
                             const int32_t* ys_lens, const int32_t* leader, int B, int T, int V, int Lmax, int blank,
                             uint32_t* words, float* val, b200_stream_t stream) {
  B200_CHECK_ARG(lp_ext && alpha && beta && frame_off && scale && hlens && ys_lens && leader && words && val &&
                     (Lmax == 0 || ys_pad),
                 "b200_ctc_post: null pointer");
  B200_CHECK_ARG(B > 0 && T > 0 && V > 0 && Lmax >= 0 && blank >= 0 && blank < V && 2 * Lmax + 1 <= 1024,
                 "b200_ctc_post: bad shape B=%d T=%d V=%d Lmax=%d blank=%d", B, T, V, Lmax, blank);
  B200_CHECK_ARG(((uintptr_t)words & 7) == 0, "b200_ctc_post: words must be 8-byte aligned");
  const int Smax = 2 * Lmax + 1;
  const int nwords = (V + 31) / 32;
  const size_t smem = ((size_t)sa::kPostFrames * Smax + 2 * (size_t)nwords) * sizeof(float);
  B200_CHECK_ARG(smem <= 200 * 1024, "b200_ctc_post: %zu bytes of shared memory needed (V=%d, Lmax=%d)", smem, V, Lmax);
  B200_CHECK_CUDA(cudaFuncSetAttribute(sa::ctc_post_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  dim3 grid((T + sa::kPostFrames - 1) / sa::kPostFrames, B);
  sa::ctc_post_kernel<<<grid, 256, smem, (cudaStream_t)stream>>>(
      lp_ext, alpha, beta, frame_off, scale, reinterpret_cast<const long long*>(ys_pad), hlens, ys_lens, leader, T, V,
      Lmax, blank, nwords, reinterpret_cast<uint2*>(words), val);
  B200_CHECK_LAUNCH("ctc_post_kernel");
  return B200_OK;
}

static int colsum_split_impl(const void* G_hi, const void* G_lo, long long rows, int N, long long ldg, const float* w,
                             float* out, b200_stream_t stream) {
  B200_CHECK_ARG(G_hi && out && rows > 0 && N > 0 && ldg >= N && ldg % 2 == 0, "b200_colsum_split: bad arguments");
  const int cblocks = (N + 255) / 256;
  int slabs = (148 * 8 + cblocks - 1) / cblocks;
  if (slabs > rows) slabs = (int)rows;
  const long long per = (rows + slabs - 1) / slabs;
  dim3 grid(cblocks, (unsigned)((rows + per - 1) / per));
  sa::colsum_split_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(reinterpret_cast<const __nv_bfloat16*>(G_hi),
                                                                  reinterpret_cast<const __nv_bfloat16*>(G_lo), rows, N,
                                                                  ldg, per, w, out);
  B200_CHECK_LAUNCH("colsum_split_kernel");
  return B200_OK;
}

extern "C" int b200_colsum_split(const void* G_hi, const void* G_lo, long long rows, int N, long long ldg, float* out,
                                 b200_stream_t stream) {
  return colsum_split_impl(G_hi, G_lo, rows, N, ldg, nullptr, out, stream);
}

// out[c] += sum_r w[r] (G_hi + G_lo)[r, c]: d bias of the no-recompute backward (w = rowscale, G = Pt')
extern "C" int b200_colsum_split_w(const void* G_hi, const void* G_lo, long long rows, int N, long long ldg,
                                   const float* w, float* out, b200_stream_t stream) {
  B200_CHECK_ARG(w, "b200_colsum_split_w: null weights");
  return colsum_split_impl(G_hi, G_lo, rows, N, ldg, w, out, stream);
}

// fp32-mode twin of b200_joint_rowscale (joint_nr.cu): rowscale, the patched entries of the split Pt from the
// exact logits xbe of b200_joint_pick, Hs = rowscale * H split again
extern "C" int b200_joint_rowscale_split(const float* coef, const float* xbe, const int32_t* targets, int B, int T,
                                         int U1, int V, int J, int blank, void* P_hi, void* P_lo, long long ldg,
                                         const void* H_hi, const void* H_lo, int ldh, void* Hs_hi, void* Hs_lo,
                                         float* rowscale, b200_stream_t stream) {
  B200_CHECK_ARG(coef && xbe && P_hi && P_lo && H_hi && H_lo && Hs_hi && Hs_lo && rowscale && (U1 == 1 || targets),
                 "b200_joint_rowscale_split: null pointer");
  B200_CHECK_ARG(B > 0 && T > 0 && U1 > 0 && blank >= 0 && blank < V && ldg >= V && J % 8 == 0 && ldh >= J && ldh % 8 == 0,
                 "b200_joint_rowscale_split: bad arguments");
  B200_CHECK_ARG((((uintptr_t)coef | (uintptr_t)H_hi | (uintptr_t)H_lo | (uintptr_t)Hs_hi | (uintptr_t)Hs_lo) & 15) == 0 &&
                     ((uintptr_t)xbe & 7) == 0,
                 "b200_joint_rowscale_split: misaligned pointer");
  const long long ncells = (long long)B * T * U1;
  sa::joint_patch_split_kernel<<<(unsigned)((ncells + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const float4*>(coef), reinterpret_cast<const float2*>(xbe), targets, T, U1, blank, ncells,
      reinterpret_cast<__nv_bfloat16*>(P_hi), reinterpret_cast<__nv_bfloat16*>(P_lo), ldg, rowscale);
  B200_CHECK_LAUNCH("joint_patch_split_kernel");
  sa::hs_split_kernel<<<grid_for(ncells * (J / 8), 256, 148 * 16), 256, 0, (cudaStream_t)stream>>>(
      rowscale, reinterpret_cast<const __nv_bfloat16*>(H_hi), reinterpret_cast<const __nv_bfloat16*>(H_lo), J, ncells, ldh,
      reinterpret_cast<__nv_bfloat16*>(Hs_hi), reinterpret_cast<__nv_bfloat16*>(Hs_lo));
  B200_CHECK_LAUNCH("hs_split_kernel");
  return B200_OK;
}
