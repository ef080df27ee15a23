// fp32-accurate tensor-core GEMM with a fused row-wise softmax epilogue, for the paths the reference
// computes in fp32 (rtol 1e-4): the ctc_lo projection + log_softmax of wenet/transformer/ctc.py:57-60,
// :66-74 and the fp32 joint of wenet/transducer/joint.py:58-69 (BASELINE configs[1]).
//
//   x[r, c] = sum_k A[r, k] * B[c, k] + bias[c]            r < rows, c < N
// Both operands are given as a SPLIT pair of bf16 arrays, X = X_hi + X_lo with X_hi = bf16(X) and
// X_lo = bf16(X - X_hi) (16 mantissa bits together), and the product is accumulated in fp32 as
//   A_hi.B_hi + A_lo.B_hi + A_hi.B_lo          (the dropped A_lo.B_lo term is 2^-18 relative)
// - three bf16 tcgen05.mma per K step, i.e. 1.5x the cost of one kind::tf32 pass for ~2^-17 instead of
// 2^-11 relative operand error.  With *_lo == NULL a single product is issued (plain bf16 operands).
//
// One CTA PAIR (cta_group::2) owns a [256 rows x N] stripe and walks its N tiles of 256 columns:
// M = 256 (128 rows per CTA), N = 256 (128 B rows per CTA), both operands K-major through TMA
// (SWIZZLE_128B boxes of 64 k), 3-stage ring of {A_hi, A_lo, B_hi, B_lo} = 64 KB, two TMEM accumulators
// of 256 columns so the epilogue of tile n overlaps the MMAs of tile n+1.  The logits never reach HBM:
//   MODE_LSE    lse[r] = log sum_c exp(x[r, c])                        (streaming, running maximum)
//   MODE_EXP    G[r, c] = scale_r * exp(x[r, c] + c0_r), written SPLIT (G_hi, G_lo: the operands of the
//               split backward GEMMs), coef = {c0, -, -, scale} per row; the sparse "- posterior"
//               terms of the RNN-T / CTC gradients are applied afterwards by a fix-up kernel
//   MODE_LSM    out[r, c] = x[r, c] - lse[r] in fp32                   (CTC.log_softmax)
//   MODE_EXPLSE the training forward of the NO-RECOMPUTE backward (wenet_celoss_b200/joint.py, fp32 mode):
//               Pt[r, c] = exp(x[r, c]) written SPLIT and lse[r] from the same pass.  exp without a
//               maximum needs bounded logits: the kernel reads the device scalar `logit_bound` and, when
//               bound * log2(e) >= 60, degrades to MODE_LSE (nothing stored) - the host applies the
//               same test to the same scalar when it picks the backward variant
// Roofline: tensor pipe; algorithmic flops per launch = 2 * rows * K * N (x3 issued in split mode).
#include <cuda.h>

#include "common.cuh"
#include "sm100_pair.cuh"
#include "sm100_ptx.cuh"

namespace b200 {
namespace sg {

using namespace pair;

constexpr int kBlockK = 64;
constexpr int kStages = 3;
constexpr int kTileN = 256;
constexpr int kOpBytes = 128 * kBlockK * 2;     // one operand block: [128 rows x 64 k] bf16 = 16 KB
constexpr int kStageBytes = 4 * kOpBytes;       // A_hi, A_lo, B_hi, B_lo
constexpr int kEpiWarp0 = 2;                    // warp 0 TMA, warp 1 MMA + TMEM, warps 2..9 epilogue
constexpr int kEpiWarps = 8;                    // 4 lane quarters x 2 column halves of 128
constexpr int kThreads = (kEpiWarp0 + kEpiWarps) * 32;
constexpr int kStagingBytes = kEpiWarps * 4096; // per warp: 32 rows x 128 B (fp32) or 2 x (32 x 64 B) (hi, lo)
constexpr int kBiasBytes = kEpiWarps * 32 * 4;
constexpr int kXchgBytes = 128 * 8;
constexpr int kSmemBytes = kStages * kStageBytes + kStagingBytes + kBiasBytes + kXchgBytes + 256;
static_assert(kSmemBytes <= 232448, "shared memory budget");

enum { MODE_LSE = 0, MODE_EXP = 1, MODE_LSM = 2, MODE_EXPLSE = 3, MODE_EXPSUB = 4 };
constexpr float kNrMaxLog2Bound = 60.f;  // = jf::kNrMaxLog2Bound (joint_fused.cu)

struct Params {
  long long rows;
  int K, N;
  const float* bias;   // [ceil(N/256)*256] natural units
  float* lse;          // MODE_LSE out / MODE_LSM in, natural log
  const float4* coef;  // MODE_EXP: {c0, -, -, scale} per row
  const float* logit_bound;  // MODE_EXPLSE: device scalar max_v(|W_v|_1 + |b_v|)
  // MODE_EXPSUB: G = scale exp(x + c0) - sub, a sparse term per (row, class) in the layout of b200_ctc_post:
  // rows come in groups of group_rows (utterances); per group a bitmap over the classes with running popcounts,
  // and the values [group][rank][row in group]
  const uint2* sub_words;    // [groups][nwords] {bitmap word, set bits before it}
  const float* sub_val;      // [groups][nranks][group_rows]
  int group_rows, nranks, nwords;
};

struct Bars {
  uint64_t full[kStages], empty[kStages], tmem_full[2], tmem_empty[2];
  uint32_t tmem_base;
};

__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

template <int MODE, int NPROD>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
    split_gemm_kernel(const __grid_constant__ CUtensorMap a_hi, const __grid_constant__ CUtensorMap a_lo,
                      const __grid_constant__ CUtensorMap b_hi, const __grid_constant__ CUtensorMap b_lo,
                      const __grid_constant__ CUtensorMap o_hi, const __grid_constant__ CUtensorMap o_lo,
                      const Params P) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t base = ptx::smem_u32(smem);
  if (base & 1023u) __trap();
  uint8_t* staging = smem + kStages * kStageBytes;
  float* bias_s = reinterpret_cast<float*>(staging + kStagingBytes);
  float2* xchg = reinterpret_cast<float2*>(staging + kStagingBytes + kBiasBytes);
  Bars* bars = reinterpret_cast<Bars*>(staging + kStagingBytes + kBiasBytes + kXchgBytes);

  const uint32_t rank = cluster_ctarank();
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int nkb = (P.K + kBlockK - 1) / kBlockK;
  const int NT = (P.N + kTileN - 1) / kTileN;
  const int mtiles = (int)((P.rows + 255) / 256);
  const int pair_id = (int)cluster_id_x();
  const int npairs = (int)nclusters_x();
  constexpr int kLoadBytes = (NPROD == 3 ? 4 : 2) * kOpBytes;

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      ptx::mbar_init(ptx::smem_u32(&bars->full[s]), 1);
      ptx::mbar_init(ptx::smem_u32(&bars->empty[s]), 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(ptx::smem_u32(&bars->tmem_full[i]), 1);
      ptx::mbar_init(ptx::smem_u32(&bars->tmem_empty[i]), 2 * kEpiWarps);
    }
    ptx::fence_barrier_init();
    ptx::prefetch_tensormap(&a_hi);
    ptx::prefetch_tensormap(&b_hi);
  }
  if (warp == 1) {
    tmem_alloc_pair(ptx::smem_u32(&bars->tmem_base), 512);
    tmem_relinquish_pair();
  }
  ptx::tc_fence_before();
  cluster_sync_all();
  ptx::tc_fence_after();
  const uint32_t tmem = bars->tmem_base;

  if (warp == 0) {
    // ---------------- TMA producer (both CTAs: own 128 rows of A, own 128 rows of B)
    const uint32_t leader_full0 = map_to_cta(ptx::smem_u32(&bars->full[0]), 0);
    const uint64_t pol_a = ptx::l2_policy_evict_first();  // re-read NT times, back to back, then never
    const uint64_t pol_b = ptx::l2_policy_evict_last();   // B (the weights) is shared by every stripe
    long long it = 0;
    for (int mt = pair_id; mt < mtiles; mt += npairs) {
      const int r0 = mt * 256 + (int)rank * 128;
      for (int n = 0; n < NT; ++n) {
        const int c0 = n * kTileN + (int)rank * 128;
        for (int kb = 0; kb < nkb; ++kb, ++it) {
          const int s = (int)(it % kStages);
          const uint32_t ph = (uint32_t)((it / kStages) & 1);
          ptx::mbar_wait(ptx::smem_u32(&bars->empty[s]), ph ^ 1u);
          if (ptx::elect_one()) {
            if (rank == 0) ptx::mbar_arrive_expect_tx(ptx::smem_u32(&bars->full[s]), 2u * kLoadBytes);
            const uint32_t bar = leader_full0 + 8u * s;
            const uint32_t dst = base + (uint32_t)s * kStageBytes;
            const int k0 = kb * kBlockK;
            tma_load_2d_pair(dst, &a_hi, bar, k0, r0, pol_a);
            tma_load_2d_pair(dst + 2 * kOpBytes, &b_hi, bar, k0, c0, pol_b);
            if (NPROD == 3) {
              tma_load_2d_pair(dst + kOpBytes, &a_lo, bar, k0, r0, pol_a);
              tma_load_2d_pair(dst + 3 * kOpBytes, &b_lo, bar, k0, c0, pol_b);
            }
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer (leader CTA only)
    if (rank == 0) {
      constexpr uint32_t idesc = ptx::make_idesc_bf16_f32(256, kTileN);
      long long it = 0;
      int acc = 0, acc_ph = 0;
      for (int mt = pair_id; mt < mtiles; mt += npairs) {
        for (int n = 0; n < NT; ++n) {
          ptx::mbar_wait(ptx::smem_u32(&bars->tmem_empty[acc]), (uint32_t)(acc_ph ^ 1));
          ptx::tc_fence_after();
          const uint32_t d_tmem = tmem + (uint32_t)acc * kTileN;
          for (int kb = 0; kb < nkb; ++kb, ++it) {
            const int s = (int)(it % kStages);
            const uint32_t ph = (uint32_t)((it / kStages) & 1);
            ptx::mbar_wait(ptx::smem_u32(&bars->full[s]), ph);
            ptx::tc_fence_after();
            if (ptx::elect_one()) {
              const uint32_t sa = base + (uint32_t)s * kStageBytes;
              const uint64_t dah = ptx::make_sw128_kmajor_desc(sa);
              const uint64_t dal = ptx::make_sw128_kmajor_desc(sa + kOpBytes);
              const uint64_t dbh = ptx::make_sw128_kmajor_desc(sa + 2 * kOpBytes);
              const uint64_t dbl = ptx::make_sw128_kmajor_desc(sa + 3 * kOpBytes);
#pragma unroll
              for (int k = 0; k < kBlockK / 16; ++k) {
                umma_bf16_pair(d_tmem, dah + 2 * k, dbh + 2 * k, idesc, (kb > 0 || k > 0) ? 1u : 0u);
                if (NPROD == 3) {
                  umma_bf16_pair(d_tmem, dal + 2 * k, dbh + 2 * k, idesc, 1u);
                  umma_bf16_pair(d_tmem, dah + 2 * k, dbl + 2 * k, idesc, 1u);
                }
              }
              umma_commit_pair(ptx::smem_u32(&bars->empty[s]));
              if (kb == nkb - 1) umma_commit_pair(ptx::smem_u32(&bars->tmem_full[acc]));
            }
            __syncwarp();
          }
          acc ^= 1;
          if (acc == 0) acc_ph ^= 1;
        }
      }
    }
  } else {
    // ---------------- epilogue (both CTAs): thread = one row (TMEM lane), 128 of the 256 columns
    const int ew = warp - kEpiWarp0;
    const int q = warp & 3;        // TMEM lane quarter this warp may access
    const int half = ew >> 2;      // column half of the N tile
    const int rrow = q * 32 + lane;
    const uint32_t lane_addr = tmem + ((uint32_t)(q * 32) << 16) + (uint32_t)half * 128;
    uint8_t* const stg = staging + ew * 4096;
    float* const bias_w = bias_s + ew * 32;
    const uint32_t tmem_empty0 = map_to_cta(ptx::smem_u32(&bars->tmem_empty[0]), 0);
    const uint64_t o_policy = ptx::l2_policy_evict_first();
    int acc = 0, acc_ph = 0;
    // raw value: the conversion to the log2 domain happens at the use, one chunk later (a multiply right
    // behind the load waited for it on the spot - a quarter of the epilogue's stall samples in ncu);
    // classes >= N are -inf
    auto bias_at = [&](int n, int c) {
      const int col = n * kTileN + half * 128 + c * 32 + lane;
      return col < P.N ? __ldg(P.bias + col) : -INFINITY;
    };
    float bnext = bias_at(0, 0);
    constexpr bool kStats = MODE == MODE_LSE || MODE == MODE_EXPLSE;
    const bool nr_ok = MODE == MODE_EXPLSE && P.logit_bound != nullptr &&
                       __ldg(P.logit_bound) * kLog2e < kNrMaxLog2Bound;
    constexpr bool kExp = MODE == MODE_EXP || MODE == MODE_EXPSUB;
    const bool stores = kExp || MODE == MODE_LSM || nr_ok;
    for (int mt = pair_id; mt < mtiles; mt += npairs) {
      const long long row = (long long)mt * 256 + rank * 128 + rrow;
      const bool in_range = row < P.rows;
      float m2 = nr_ok ? 0.f : kNegBig, ssum = 0.f;  // MODE_LSE / MODE_EXPLSE
      float c0 = -INFINITY, scale = 0.f;      // MODE_EXP (log2 domain) / MODE_LSM (c0 = -lse2)
      if (kExp && in_range) {
        const float4 c = __ldg(P.coef + row);
        scale = c.w;
        c0 = c.x * kLog2e;
      }
      if (MODE == MODE_LSM && in_range) c0 = -__ldg(P.lse + row) * kLog2e;
      // MODE_EXPSUB: this row's group (utterance): bitmap words and values; rows with scale 0 subtract nothing.
      // The bitmap word of a chunk is loaded one chunk ahead (like the bias), the first value it selects at the
      // top of its chunk, under the tcgen05.ld / store-drain waits.
      const uint2* sub_w = nullptr;
      const float* sub_v = nullptr;
      uint2 wnext = make_uint2(0u, 0u);
      auto word_at = [&](int n, int c) {
        const int wi = n * (kTileN / 32) + half * 4 + c;
        return (sub_w != nullptr && wi < P.nwords) ? __ldg(sub_w + wi) : make_uint2(0u, 0u);
      };
      if (MODE == MODE_EXPSUB && in_range && scale != 0.f) {
        const long long grp = row / P.group_rows;
        sub_w = P.sub_words + grp * P.nwords;
        sub_v = P.sub_val + grp * P.nranks * P.group_rows + (row - grp * P.group_rows);
        wnext = word_at(0, 0);
      }
      // scale >= 0 on every row of the warp: move it into the exponent (log2(0) = -inf: exact zeros)
      const bool pos = kExp && __all_sync(0xffffffffu, scale >= 0.f);
      const float c0s = pos ? c0 + log2f(scale) : c0;
      for (int n = 0; n < NT; ++n) {
        ptx::mbar_wait(ptx::smem_u32(&bars->tmem_full[acc]), (uint32_t)acc_ph);
        ptx::tc_fence_after();
#pragma unroll 1
        for (int c = 0; c < 4; ++c) {
          uint32_t r[32];
          ptx::tmem_ld32(lane_addr + (uint32_t)acc * kTileN + c * 32, r);
          // this chunk's bias into the warp's slot (loaded one chunk ahead), the next one into a register
          __syncwarp();
          bias_w[lane] = bnext * kLog2e;
          __syncwarp();
          {
            int nn = n, cc = c + 1;
            if (cc == 4) {
              cc = 0;
              nn = (n + 1 == NT) ? 0 : n + 1;
            }
            bnext = bias_at(nn, cc);
          }
          uint2 wp = make_uint2(0u, 0u);
          float d0 = 0.f;
          if (MODE == MODE_EXPSUB) {
            wp = wnext;
            if (c + 1 < 4 || n + 1 < NT) wnext = word_at(c + 1 < 4 ? n : n + 1, c + 1 < 4 ? c + 1 : 0);
            if (wp.x) d0 = __ldg(sub_v + (long long)wp.y * P.group_rows);
          }
          if (stores) {  // the previous chunk's store must have drained the staging tile
            if (lane == 0) ptx::bulk_wait_group_read0();
            __syncwarp();
          }
          ptx::tmem_ld_wait();
          if (c == 3) {  // the accumulator is in registers: hand it back to the MMA warp
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(tmem_empty0 + 8u * acc);
          }
          float t[32];
          const float4* b4 = reinterpret_cast<const float4*>(bias_w);
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const float4 bv = b4[i];
            t[4 * i + 0] = fmaf(__uint_as_float(r[4 * i + 0]), kLog2e, bv.x);
            t[4 * i + 1] = fmaf(__uint_as_float(r[4 * i + 1]), kLog2e, bv.y);
            t[4 * i + 2] = fmaf(__uint_as_float(r[4 * i + 2]), kLog2e, bv.z);
            t[4 * i + 3] = fmaf(__uint_as_float(r[4 * i + 3]), kLog2e, bv.w);
          }
          const int col0 = n * kTileN + half * 128 + c * 32;
          if (kStats && !nr_ok) {
            float cm[4] = {t[0], t[1], t[2], t[3]};
#pragma unroll
            for (int i = 4; i < 32; ++i) cm[i & 3] = fmaxf(cm[i & 3], t[i]);
            const float mn = fmaxf(m2, fmaxf(fmaxf(cm[0], cm[1]), fmaxf(cm[2], cm[3])));
            float a4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int i = 0; i < 32; ++i) a4[i & 3] += ex2_approx(t[i] - mn);
            ssum = ssum * ex2_approx(m2 - mn) + ((a4[0] + a4[1]) + (a4[2] + a4[3]));
            m2 = mn;
          } else if (kExp || MODE == MODE_EXPLSE) {
            if (MODE == MODE_EXPLSE) {  // Pt = 2^t (bounded logits), summed for lse on the way
              float a4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
              for (int i = 0; i < 32; ++i) {
                t[i] = ex2_approx(t[i]);
                a4[i & 3] += t[i];
              }
              ssum += (a4[0] + a4[1]) + (a4[2] + a4[3]);
            } else if (pos) {
#pragma unroll
              for (int i = 0; i < 32; ++i) t[i] = ex2_approx(t[i] + c0s);
            } else {
#pragma unroll
              for (int i = 0; i < 32; ++i) t[i] = scale * ex2_approx(t[i] + c0s);
            }
            if (MODE == MODE_EXPSUB && __any_sync(0xffffffffu, wp.x != 0u)) {
              // the selected classes of this chunk: bit k of the word = column col0 + k, values in rank order.
              // The register index is compile-time only, hence a compare-select sweep per set bit (1.4 bits on
              // average in a word that has any; the lanes of a warp share the word unless they straddle groups)
              uint32_t w = wp.x;
              int kk = w ? __ffs(w) - 1 : -1;
#pragma unroll
              for (int i = 0; i < 32; ++i) t[i] -= (i == kk) ? d0 : 0.f;
              w &= w - 1u;
              long long rk = (long long)wp.y + 1;
              while (__any_sync(0xffffffffu, w != 0u)) {
                kk = w ? __ffs(w) - 1 : -1;
                const float d = w ? __ldg(sub_v + rk * P.group_rows) : 0.f;
#pragma unroll
                for (int i = 0; i < 32; ++i) t[i] -= (i == kk) ? d : 0.f;
                w &= w - 1u;
                ++rk;
              }
            }
            // split: hi = bf16(v), lo = bf16(v - hi); two 32 x 64-byte tiles (SWIZZLE_64B: 16-byte
            // piece j of row r at j ^ ((r >> 1) & 3)), written by two TMA stores
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              uint4 vh, vl;
              uint32_t* ph_ = &vh.x;
              uint32_t* pl_ = &vl.x;
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const float x0 = t[8 * j + 2 * e], x1 = t[8 * j + 2 * e + 1];
                const uint32_t h = pack_bf16x2(x0, x1);
                ph_[e] = h;
                pl_[e] = pack_bf16x2(x0 - __uint_as_float(h << 16), x1 - __uint_as_float(h & 0xffff0000u));
              }
              const int off = lane * 64 + ((j ^ ((lane >> 1) & 3)) << 4);
              *reinterpret_cast<uint4*>(stg + off) = vh;
              *reinterpret_cast<uint4*>(stg + 2048 + off) = vl;
            }
            ptx::fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
              const int orow = mt * 256 + (int)rank * 128 + q * 32;
              ptx::tma_store_2d_hint(&o_hi, ptx::smem_u32(stg), col0, orow, o_policy);
              if (NPROD == 3) ptx::tma_store_2d_hint(&o_lo, ptx::smem_u32(stg + 2048), col0, orow, o_policy);
              ptx::bulk_commit_group();
            }
          } else {
            // log-softmax in fp32: 32 rows x 128 bytes (SWIZZLE_128B: 16-byte piece j of row r at j ^ (r & 7))
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              float4 v;
              v.x = (t[4 * j + 0] + c0s) * kLn2;
              v.y = (t[4 * j + 1] + c0s) * kLn2;
              v.z = (t[4 * j + 2] + c0s) * kLn2;
              v.w = (t[4 * j + 3] + c0s) * kLn2;
              *reinterpret_cast<float4*>(stg + lane * 128 + ((j ^ (lane & 7)) << 4)) = v;
            }
            ptx::fence_proxy_async();
            __syncwarp();
            if (lane == 0) {
              ptx::tma_store_2d_hint(&o_hi, ptx::smem_u32(stg), col0, mt * 256 + (int)rank * 128 + q * 32, o_policy);
              ptx::bulk_commit_group();
            }
          }
        }
        acc ^= 1;
        if (acc == 0) acc_ph ^= 1;
      }
      if (kStats) {
        // the two column halves of a row sit in two warps: half 1 hands (m2, ssum) to half 0
        // (named barriers 1 + q: data ready, 5 + q: slot free again; 64 threads each)
        if (half == 1) xchg[rrow] = make_float2(m2, ssum);
        asm volatile("bar.sync %0, 64;" ::"r"(1 + q) : "memory");
        if (half == 0) {
          const float2 o = xchg[rrow];
          const float mn = fmaxf(m2, o.x);
          const float s = ssum * ex2_approx(m2 - mn) + o.y * ex2_approx(o.x - mn);
          if (in_range) P.lse[row] = (mn + log2f(s)) * kLn2;
        }
        asm volatile("bar.sync %0, 64;" ::"r"(5 + q) : "memory");
      }
    }
    if (stores && lane == 0) ptx::bulk_wait_group0();
  }

  ptx::tc_fence_before();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc_pair(tmem, 512);
}

}  // namespace sg
}  // namespace b200

using namespace b200;

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn sg_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// 2-D map [rows, cols] with a row pitch in bytes; out-of-range parts read as zero / are not written
int sg_map(CUtensorMap* map, const void* ptr, CUtensorMapDataType dt, int esize, long long rows, long long cols,
           long long pitch_bytes, int box_cols, int box_rows, CUtensorMapSwizzle sw) {
  EncodeTiledFn fn = sg_encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled entry point not available");
    return B200_ERR_CUDA;
  }
  (void)esize;
  cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)pitch_bytes};
  cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, dt, 2, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return B200_ERR_CUDA;
  }
  return B200_OK;
}

template <int MODE, int NPROD>
int sg_launch(const CUtensorMap* m, const sg::Params& P, int max_ctas, cudaStream_t st) {
  B200_CHECK_CUDA(cudaFuncSetAttribute(sg::split_gemm_kernel<MODE, NPROD>,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, sg::kSmemBytes));
  int sms = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  int ctas = sms & ~1;
  if (max_ctas > 0 && max_ctas < ctas) ctas = max_ctas & ~1;
  if (ctas < 2) ctas = 2;
  const long long mtiles = (P.rows + 255) / 256;
  if (mtiles < ctas / 2) ctas = 2 * (int)mtiles;
  sg::split_gemm_kernel<MODE, NPROD><<<ctas, sg::kThreads, sg::kSmemBytes, st>>>(m[0], m[1], m[2], m[3], m[4], m[5], P);
  B200_CHECK_LAUNCH("split_gemm_kernel");
  return B200_OK;
}

int sg_run(int mode, const void* a_hi, const void* a_lo, long long rows, int K, int lda, const void* b_hi,
           const void* b_lo, int N, int ldb, const float* bias_pad, float* lse, const float* coef, void* o_hi,
           void* o_lo, long long ldo, int max_ctas, b200_stream_t stream, const sg::Params* sub = nullptr) {
  B200_CHECK_ARG(a_hi && b_hi && bias_pad, "b200_split_gemm: null pointer");
  B200_CHECK_ARG((a_lo == nullptr) == (b_lo == nullptr), "b200_split_gemm: a_lo and b_lo go together");
  B200_CHECK_ARG(rows > 0 && N > 0 && K > 0 && K % 8 == 0, "b200_split_gemm: bad shape rows=%lld N=%d K=%d (K %% 8)",
                 rows, N, K);
  B200_CHECK_ARG(rows < (1ll << 31) - 256, "b200_split_gemm: too many rows");
  B200_CHECK_ARG(lda >= K && lda % 8 == 0 && ldb >= K && ldb % 8 == 0, "b200_split_gemm: lda / ldb must be >= K, %% 8");
  B200_CHECK_ARG((((uintptr_t)a_hi | (uintptr_t)a_lo | (uintptr_t)b_hi | (uintptr_t)b_lo) & 15) == 0,
                 "b200_split_gemm: operands must be 16-byte aligned");
  const bool split = a_lo != nullptr;
  CUtensorMap m[6];
  const auto bf = CU_TENSOR_MAP_DATA_TYPE_BFLOAT16;
  if (int rc = sg_map(&m[0], a_hi, bf, 2, rows, K, (long long)lda * 2, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
  if (int rc = sg_map(&m[1], split ? a_lo : a_hi, bf, 2, rows, K, (long long)lda * 2, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
  if (int rc = sg_map(&m[2], b_hi, bf, 2, N, K, (long long)ldb * 2, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
  if (int rc = sg_map(&m[3], split ? b_lo : b_hi, bf, 2, N, K, (long long)ldb * 2, 64, 128, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
  m[4] = m[0];
  m[5] = m[0];
  sg::Params P{};
  P.rows = rows;
  P.K = K;
  P.N = N;
  P.bias = bias_pad;
  P.lse = lse;
  P.coef = reinterpret_cast<const float4*>(coef);
  cudaStream_t st = (cudaStream_t)stream;
  if (mode == sg::MODE_EXPLSE) {
    // `coef` carries the logit-bound scalar in this mode
    P.coef = nullptr;
    P.logit_bound = coef;
    B200_CHECK_ARG(lse && coef && o_hi && (!split || o_lo), "b200_split_gemm_explse: null pointer");
    B200_CHECK_ARG(ldo >= N && ldo % 8 == 0 && (((uintptr_t)o_hi | (uintptr_t)o_lo) & 15) == 0,
                   "b200_split_gemm_explse: Pt must be 16-byte aligned, ldo=%lld >= N and a multiple of 8", ldo);
    if (int rc = sg_map(&m[4], o_hi, bf, 2, rows, N, ldo * 2, 32, 32, CU_TENSOR_MAP_SWIZZLE_64B)) return rc;
    if (int rc = sg_map(&m[5], split ? o_lo : o_hi, bf, 2, rows, N, ldo * 2, 32, 32, CU_TENSOR_MAP_SWIZZLE_64B)) return rc;
    return split ? sg_launch<sg::MODE_EXPLSE, 3>(m, P, max_ctas, st) : sg_launch<sg::MODE_EXPLSE, 1>(m, P, max_ctas, st);
  }
  if (mode == sg::MODE_LSE) {
    B200_CHECK_ARG(lse, "b200_split_gemm_lse: null lse");
    return split ? sg_launch<sg::MODE_LSE, 3>(m, P, max_ctas, st) : sg_launch<sg::MODE_LSE, 1>(m, P, max_ctas, st);
  }
  if (mode == sg::MODE_EXP || mode == sg::MODE_EXPSUB) {
    B200_CHECK_ARG(coef && o_hi && (!split || o_lo), "b200_split_gemm_exp: null pointer");
    B200_CHECK_ARG(((uintptr_t)coef & 15) == 0 && ldo >= N && ldo % 8 == 0 &&
                       (((uintptr_t)o_hi | (uintptr_t)o_lo) & 15) == 0,
                   "b200_split_gemm_exp: coef / G must be 16-byte aligned, ldg=%lld >= N and a multiple of 8", ldo);
    if (int rc = sg_map(&m[4], o_hi, bf, 2, rows, N, ldo * 2, 32, 32, CU_TENSOR_MAP_SWIZZLE_64B)) return rc;
    if (int rc = sg_map(&m[5], split ? o_lo : o_hi, bf, 2, rows, N, ldo * 2, 32, 32, CU_TENSOR_MAP_SWIZZLE_64B)) return rc;
    if (mode == sg::MODE_EXPSUB) {
      B200_CHECK_ARG(sub && sub->sub_words && sub->sub_val && sub->group_rows > 0 && sub->nranks > 0 &&
                         sub->nwords == (N + 31) / 32 && rows % sub->group_rows == 0,
                     "b200_split_gemm_exp_sub: bad sparse term (rows must be whole groups, nwords = ceil(N / 32))");
      P.sub_words = sub->sub_words;
      P.sub_val = sub->sub_val;
      P.group_rows = sub->group_rows;
      P.nranks = sub->nranks;
      P.nwords = sub->nwords;
      return split ? sg_launch<sg::MODE_EXPSUB, 3>(m, P, max_ctas, st)
                   : sg_launch<sg::MODE_EXPSUB, 1>(m, P, max_ctas, st);
    }
    return split ? sg_launch<sg::MODE_EXP, 3>(m, P, max_ctas, st) : sg_launch<sg::MODE_EXP, 1>(m, P, max_ctas, st);
  }
  B200_CHECK_ARG(lse && o_hi, "b200_split_gemm_lsm: null pointer");
  B200_CHECK_ARG(ldo >= N && ldo % 4 == 0 && ((uintptr_t)o_hi & 15) == 0,
                 "b200_split_gemm_lsm: out must be 16-byte aligned, ldo=%lld >= N and a multiple of 4", ldo);
  if (int rc = sg_map(&m[4], o_hi, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, rows, N, ldo * 4, 32, 32, CU_TENSOR_MAP_SWIZZLE_128B))
    return rc;
  return split ? sg_launch<sg::MODE_LSM, 3>(m, P, max_ctas, st) : sg_launch<sg::MODE_LSM, 1>(m, P, max_ctas, st);
}

}  // namespace

extern "C" int b200_split_gemm_lse(const void* a_hi, const void* a_lo, long long rows, int K, int lda,
                                   const void* b_hi, const void* b_lo, int N, int ldb, const float* bias_pad,
                                   float* lse, int max_ctas, b200_stream_t stream) {
  return sg_run(sg::MODE_LSE, a_hi, a_lo, rows, K, lda, b_hi, b_lo, N, ldb, bias_pad, lse, nullptr, nullptr, nullptr,
                0, max_ctas, stream);
}

extern "C" int b200_split_gemm_exp(const void* a_hi, const void* a_lo, long long rows, int K, int lda,
                                   const void* b_hi, const void* b_lo, int N, int ldb, const float* bias_pad,
                                   const float* coef, void* g_hi, void* g_lo, long long ldg, int max_ctas,
                                   b200_stream_t stream) {
  return sg_run(sg::MODE_EXP, a_hi, a_lo, rows, K, lda, b_hi, b_lo, N, ldb, bias_pad, nullptr, coef, g_hi, g_lo, ldg,
                max_ctas, stream);
}

// G = scale exp(x + c0) - sub: the CTC gradient w.r.t. the ctc_lo logits in one pass (softmax part from the GEMM,
// state posteriors of the label classes from b200_ctc_post), no read-modify-write of G afterwards
extern "C" int b200_split_gemm_exp_sub(const void* a_hi, const void* a_lo, long long rows, int K, int lda,
                                       const void* b_hi, const void* b_lo, int N, int ldb, const float* bias_pad,
                                       const float* coef, const uint32_t* sub_words, const float* sub_val,
                                       int group_rows, int nranks, void* g_hi, void* g_lo, long long ldg,
                                       int max_ctas, b200_stream_t stream) {
  B200_CHECK_ARG(((uintptr_t)sub_words & 7) == 0, "b200_split_gemm_exp_sub: sub_words must be 8-byte aligned");
  sg::Params sub{};
  sub.sub_words = reinterpret_cast<const uint2*>(sub_words);
  sub.sub_val = sub_val;
  sub.group_rows = group_rows;
  sub.nranks = nranks;
  sub.nwords = (N + 31) / 32;
  return sg_run(sg::MODE_EXPSUB, a_hi, a_lo, rows, K, lda, b_hi, b_lo, N, ldb, bias_pad, nullptr, coef, g_hi, g_lo,
                ldg, max_ctas, stream, &sub);
}

// lse AND Pt = exp(x) (split) from one pass: the training forward of the no-recompute backward, fp32 mode
extern "C" int b200_split_gemm_explse(const void* a_hi, const void* a_lo, long long rows, int K, int lda,
                                      const void* b_hi, const void* b_lo, int N, int ldb, const float* bias_pad,
                                      const float* logit_bound, float* lse, void* p_hi, void* p_lo, long long ldp,
                                      int max_ctas, b200_stream_t stream) {
  return sg_run(sg::MODE_EXPLSE, a_hi, a_lo, rows, K, lda, b_hi, b_lo, N, ldb, bias_pad, lse, logit_bound, p_hi, p_lo,
                ldp, max_ctas, stream);
}

extern "C" int b200_split_gemm_lsm(const void* a_hi, const void* a_lo, long long rows, int K, int lda,
                                   const void* b_hi, const void* b_lo, int N, int ldb, const float* bias_pad,
                                   const float* lse, float* out, long long ldo, int max_ctas, b200_stream_t stream) {
  return sg_run(sg::MODE_LSM, a_hi, a_lo, rows, K, lda, b_hi, b_lo, N, ldb, bias_pad, const_cast<float*>(lse), nullptr,
                out, nullptr, ldo, max_ctas, stream);
}
