// Fused transducer joint: logits tile = tanh(e[b,t,:] + p[b,u,:]) . W_out^T + b_out on the
// 5th-gen tensor cores (tcgen05.mma, accumulators in TMEM), with the epilogue consuming the
// tile straight out of TMEM so the [B,T,U+1,V] logits never reach HBM.
//
// Replaces, at wenet/transducer/transducer.py:132-147, the chain
//   TransducerJoint.forward (wenet/transducer/joint.py:58-69: broadcast add, tanh, ffn_out)
//   + the log-softmax statistics / gradient of torchaudio.functional.rnnt_loss.
// enc_ffn / pred_ffn (joint.py:55-56) stay library GEMMs in the Python wrapper and arrive
// here as e [B*T, J] and p [B*U1, J] (fp32).
//
// One persistent CTA (640 threads) per SM walks M-tiles of 128 lattice cells
// (cell = (b*T + t)*U1 + u):
//   warp  0     TMA-loads W_out tiles [128 vocab rows x 2 x 64 k] (bf16) into a ring
//   warp  1     at the start of a tile copies the staged A into TENSOR MEMORY (tcgen05.cp,
//               128 lanes x 256 columns = 128 x 512 bf16) and then issues tcgen05.mma with A
//               read from tensor memory (M=128, N=128, K=16) into two 128-column accumulators.
//               A from TMEM matters: an M128 N128 MMA with both operands in shared memory
//               reads 8 KB per 64 tensor cycles = the whole shared-memory bandwidth of the SM
//               (measured: 94 instead of 64 cycles per MMA); with A in TMEM only W is read
//   warp  2     allocates / frees tensor memory
//   warp  3     MODE_GRAD: TMA-loads A tiles from the bf16 H scratch into a 2-slot ring
//   warps 4..   epilogue: 2 (MODE_FWD) or 4 (MODE_GRAD) warps per TMEM lane quarter (= per SM
//               sub-partition); each thread owns one cell row and 64 / 32 of the 128 columns of
//               every N tile; the accumulator is released as soon as it is in registers.
//     MODE_FWD   log2-domain sum-exp over V (no running maximum when the host-provided logit
//                bound allows), picks the blank and target logits
//                -> lse, lp_blank, lp_emit  (3 floats per cell leave the CTA).
//                Warps 12-15 build the NEXT tile's A = tanh(e+p) (bf16, K-major SWIZZLE_128B
//                staging, 128 KB) while the current tile is multiplied (folding this into the
//                epilogue warps was measured slower: 58 vs 54 us per tile)
//     MODE_GRAD  G = scale*(exp(x+c0) - [v==blank]exp(x+cb) - [v==y]exp(x+ce)) in bf16, staged
//                per warp in shared memory and written with TMA stores into a bounded scratch
//                chunk that feeds the two plain backward GEMMs
// Roofline: tensor pipe; algorithmic flops per launch = 2 * cells * J * V.
#include "common.cuh"
#include "sm100_ptx.cuh"

namespace b200 {
namespace jf {

constexpr int kTileM = 128;
constexpr int kTileN = 128;
constexpr int kBiasBlock = 256;  // bias is staged (and padded on the host) in blocks of 256 classes
constexpr int kBlockK = 64;      // bf16 elements per 128-byte swizzle row
constexpr int kMaxKB = 8;        // J <= 512
constexpr int kMaxStages = 4;    // W ring: each stage holds TWO k-blocks (128 x 128 k = 32 KB), so
                                 // one barrier round trip of the issuing warp feeds 8 MMAs (512
                                 // tensor cycles); with 4 MMAs per wait the issue latency showed
constexpr int kABlockBytes = kTileM * kBlockK * 2;      // 16 KB per k-block
constexpr int kEpiWarp0 = 4;  // warps 0..3: TMA (W), MMA, TMEM alloc, TMA (A, MODE_GRAD)
constexpr int kASlotBytes = 2 * kABlockBytes;           // one pair of k-blocks
constexpr int kOperandBytes = 224 * 1024;  // A staging + W ring (+ G staging), per mode:
//   MODE_FWD : A 4 slots (whole next tile, written by the builder warps) | W ring of 3
//   MODE_GRAD: A 2 slots (TMA ring fed from the bf16 H scratch) | W ring of 4 | G staging 8 x 4 KB
//              (the G stores and H loads lengthen the L2 round trip of the W stream: with 3
//              stages the MMA warp waited 40 % of the time for W)
//   MODE_FWD : 512 threads: 8 epilogue warps (64 columns each) + 4 A-builder warps
//   MODE_GRAD: 640 threads: 16 epilogue warps (32 columns each; its epilogue is the longer one:
//              FADD + bf16 packing + staging stores on top of FFMA / MUFU)
//   MODE_FWDP: the MODE_GRAD layout (A by TMA from a precomputed H, 16 epilogue warps, staging + TMA
//              stores) with the forward statistics: see below
template <int MODE> struct Cfg {
  static constexpr int kEpiWarps = MODE == 0 ? 8 : 16;
  static constexpr int kEpiThreads = kEpiWarps * 32;
  static constexpr int kBuilderWarp0 = kEpiWarp0 + kEpiWarps;  // MODE_FWD: 4 builder warps
  static constexpr int kThreads = MODE == 0 ? 512 : 640;
  static constexpr int kASlots = MODE == 0 ? 4 : 2;
// -DB200_GRAD_GBUFS=2: a second G staging tile per epilogue warp (W ring 3 stages to fit). Measured at
// cfg3, all cells: 8.69 ms against 8.66 ms with one tile -> the TMA-store drain does not bound GRAD.
// -DB200_GRAD_DBG=bits: timing knock-outs of the GRAD epilogue (results are wrong): 1 = no G store, 2 = no
// d b_out column sums, 4 = no exponentials (scripts/grad_variants.sh)
#ifndef B200_GRAD_DBG
#define B200_GRAD_DBG 0
#endif
#ifndef B200_GRAD_GBUFS
#define B200_GRAD_GBUFS 1
#endif
  static constexpr int kGBufs = MODE == 0 ? 0 : B200_GRAD_GBUFS;  // G staging tiles per epilogue warp
  static constexpr int kStages = MODE == 0 ? 3 : (B200_GRAD_GBUFS == 2 ? 3 : 4);
  static constexpr int kABytes = kASlots * kASlotBytes;
  static constexpr int kGStageBytes = kGBufs * 16 * 2048;
};
constexpr int kWBlockBytes = kTileN * kBlockK * 2;      // 16 KB per k-block
constexpr int kWStageBytes = 2 * kWBlockBytes;          // 32 KB
constexpr float kNrMaxLog2Bound = 60.f;  // MODE_FWDP keeps exp(logit) only below this bound * log2(e)
constexpr int kTmemCols = 512;
constexpr int kTmemA0 = 2 * kTileN;  // TMEM columns [256,512): A, 2 bf16 per 32-bit column
constexpr int kBarBytes = 512;
constexpr int kBiasBytes = 2 * kBiasBlock * 4;  // bias block in use and the next one
constexpr int kXchgBytes = kTileM * 4;          // forward statistics handed between column halves
// A + W ring leave 3 KB of the 227 KB: no slack for re-aligning the base (checked at run time)
constexpr int kSmemBytes = kOperandBytes + kBarBytes + kBiasBytes + kXchgBytes;
static_assert(kSmemBytes <= 232448, "shared memory budget");
static_assert(Cfg<0>::kABytes + Cfg<0>::kStages * (2 * kTileN * kBlockK * 2) + Cfg<0>::kGStageBytes <= kOperandBytes &&
              Cfg<1>::kABytes + Cfg<1>::kStages * (2 * kTileN * kBlockK * 2) + Cfg<1>::kGStageBytes <= kOperandBytes,
              "operand area");
static_assert(kTmemA0 + kMaxKB * kBlockK / 2 <= kTmemCols, "tensor memory budget");

// MODE_FWDP ("forward + probabilities"): the training forward of the no-recompute backward.  Besides
// lse / lp_blank / lp_emit its epilogue stores Pt = 2^(logit * log2 e) = exp(logit) in bf16 (the
// unnormalised softmax numerator; only taken when the logit bound keeps 2^(+-bound) inside fp32 / bf16
// range).  d loss / d logits is then rowscale[c] * Pt[c, v] with two patched entries per row
// (joint_nr.cu), so the backward pass runs dW and dH straight from Pt: the logits GEMM is NOT recomputed.
enum { MODE_FWD = 0, MODE_GRAD = 1, MODE_FWDP = 2 };

struct Params {
  const float* e;       // [B*T, J]
  const float* p;       // [B*U1, J]
  const float* bias;    // [V]
  const int32_t* targets;
  const int32_t* logit_len;
  const int32_t* target_len;
  int B, T, U1, V, J, blank;
  long long cell0, ncells, total_cells;
  // MODE_FWD
  float* lse;
  float* lp_blank;
  float* lp_emit;
  const float* logit_bound;  // optional device scalar, see b200_joint_logit_bound
  // MODE_GRAD
  const float4* coef;
  __nv_bfloat16* G;  // [ncells, ldg]
  int ldg;
  __nv_bfloat16* H;  // [ncells, ldh]: the bf16 activations (A operand, and for the dW GEMM)
  int ldh;
  float* db_part;    // optional [16][ld_db]: partial column sums of G (= d b_out), accumulated
  int ld_db;
};

struct Barriers {
  uint64_t full[kMaxStages];
  uint64_t empty[kMaxStages];
  uint64_t tmem_full[2];
  uint64_t tmem_empty[2];
  uint64_t a_ready[4];  // A slot filled (builder warps / TMA) -> MMA warp
  uint64_t a_free[4];   // A slot copied to TMEM -> builders / TMA producer
  uint64_t a_done;               // all MMAs of a tile complete: its A in TMEM may be overwritten
  uint32_t tmem_base;
};
static_assert(sizeof(Barriers) <= kBarBytes, "barrier block overlaps the bias staging area");

__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

template <int MODE>
__global__ void __launch_bounds__(Cfg<MODE>::kThreads, 1)
joint_kernel(const __grid_constant__ CUtensorMap w_map, const __grid_constant__ CUtensorMap h_map,
             const __grid_constant__ CUtensorMap g_map, const Params P) {
  constexpr int kASlots = Cfg<MODE>::kASlots;
  constexpr int kStages = Cfg<MODE>::kStages;
  constexpr int kEpiWarps = Cfg<MODE>::kEpiWarps;
  constexpr int kEpiThreads = Cfg<MODE>::kEpiThreads;
  constexpr int kBuilderWarp0 = Cfg<MODE>::kBuilderWarp0;
  constexpr int kABytes = Cfg<MODE>::kABytes;
  extern __shared__ __align__(1024) uint8_t smem[];
  const uint32_t base = ptx::smem_u32(smem);
  if (base & 1023u) __trap();  // SWIZZLE_128B operands need 1024-B alignment
  const uint32_t sA = base;
  const uint32_t sW = base + kABytes;
  uint8_t* g_stage = smem + kABytes + kStages * kWStageBytes;  // MODE_GRAD: 2 KB per epilogue warp
  Barriers* bars = reinterpret_cast<Barriers*>(smem + kOperandBytes);
  float* bias_s = reinterpret_cast<float*>(smem + kOperandBytes + kBarBytes);
  float* xchg = bias_s + 2 * kBiasBlock;
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int nkb = P.J / kBlockK;
  const int nkp = (nkb + 1) >> 1;  // pairs of k-blocks = W ring stages per N tile
  const int NT = (P.V + kTileN - 1) / kTileN;
  const int ntiles = (int)((P.ncells + kTileM - 1) / kTileM);

  if (threadIdx.x == 0) {
    for (int i = 0; i < kStages; ++i) {
      ptx::mbar_init(ptx::smem_u32(&bars->full[i]), 1);
      ptx::mbar_init(ptx::smem_u32(&bars->empty[i]), 1);
    }
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(ptx::smem_u32(&bars->tmem_full[i]), 1);
      ptx::mbar_init(ptx::smem_u32(&bars->tmem_empty[i]), kEpiThreads);
    }
    for (int i = 0; i < 4; ++i) {
      ptx::mbar_init(ptx::smem_u32(&bars->a_ready[i]), MODE == MODE_FWD ? 4 : 1);
      ptx::mbar_init(ptx::smem_u32(&bars->a_free[i]), 1);
    }
    ptx::mbar_init(ptx::smem_u32(&bars->a_done), 1);
    ptx::fence_barrier_init();
    ptx::prefetch_tensormap(&w_map);
    if (MODE != MODE_FWD) {
      ptx::prefetch_tensormap(&h_map);
      ptx::prefetch_tensormap(&g_map);
    }
  }
  if (warp == 2) {
    ptx::tmem_alloc(ptx::smem_u32(&bars->tmem_base), kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem = bars->tmem_base;

  if (warp == 0) {
    // ===================== TMA producer =====================
    // The whole warp walks the loop (warp-uniform control flow keeps every operand in
    // uniform registers); one elected lane issues.
    const bool leader = ptx::elect_one();
    // W_out (5 MB) is re-read by every tile of every SM: keep it in L2 while the G / H streams
    // of MODE_GRAD pass through (15 % of the W sectors missed L2 without the hint)
    const uint64_t w_policy = ptx::l2_policy_evict_last();
    int stage = 0, ph = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
      for (int n = 0; n < NT; ++n)
        for (int kp = 0; kp < nkp; ++kp) {
          ptx::mbar_wait(ptx::smem_u32(&bars->empty[stage]), ph ^ 1);
          if (leader) {
            const uint32_t fb = ptx::smem_u32(&bars->full[stage]);
            const bool two = 2 * kp + 1 < nkb;
            ptx::mbar_arrive_expect_tx(fb, two ? kWStageBytes : kWBlockBytes);
            const uint32_t dst = sW + stage * kWStageBytes;
            ptx::tma_load_2d_hint(dst, &w_map, fb, 2 * kp * kBlockK, n * kTileN, w_policy);
            if (two)
              ptx::tma_load_2d_hint(dst + kWBlockBytes, &w_map, fb, (2 * kp + 1) * kBlockK, n * kTileN,
                                    w_policy);
          }
          if (++stage == kStages) {
            stage = 0;
            ph ^= 1;
          }
        }
    __syncwarp();
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    // Warp-uniform loop, one elected lane issues: with the loop inside `if (lane == 0)` the
    // compiler could not prove the descriptors uniform and wrapped every tcgen05.mma in an
    // ELECT / BRA.U.ANY waterfall (~85 instructions per 4 MMAs: the issue thread, not the
    // tensor pipe, paced the kernel -- profiles/README.md).
    const bool leader = ptx::elect_one();
    constexpr uint32_t idesc = ptx::make_idesc_bf16_f32(kTileM, kTileN);
    const uint64_t desc_hi = ptx::make_sw128_kmajor_desc(0);
    int stage = 0, ph = 0, acc = 0, acc_ph = 0, it = 0, aslot = 0, aph = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      // the previous tile's MMAs still read A from tensor memory: wait until they are done
      if (it > 0) ptx::mbar_wait(ptx::smem_u32(&bars->a_done), (it - 1) & 1);
      for (int n = 0; n < NT; ++n) {
        ptx::mbar_wait(ptx::smem_u32(&bars->tmem_empty[acc]), acc_ph ^ 1);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem + acc * kTileN;
        for (int kp = 0; kp < nkp; ++kp) {
          const int nk = (2 * kp + 1 < nkb) ? 2 * (kBlockK / 16) : (kBlockK / 16);  // K=16 steps
          const uint32_t ta = tmem + kTmemA0 + kp * kBlockK;  // 2 k-blocks = 64 columns per pair
          if (n == 0) {
            // staged A (pair kp) -> tensor memory: 128 lanes x 32 bytes per tcgen05.cp
            // MODE_FWD: slot = pair index (the staging holds a whole tile); MODE_GRAD: ring
            const int as = MODE == MODE_FWD ? kp : aslot;
            ptx::mbar_wait(ptx::smem_u32(&bars->a_ready[as]), MODE == MODE_FWD ? (it & 1) : aph);
            ptx::tc_fence_after();
            if (leader) {
              const uint64_t da = desc_hi | (uint64_t)(((sA + as * kASlotBytes) & 0x3FFFF) >> 4);
#pragma unroll
              for (int k = 0; k < 2 * (kBlockK / 16); ++k)
                if (k < nk)
                  ptx::tmem_cp_128x256b(ta + 8 * k, da + (k >> 2) * (kABlockBytes >> 4) + 2 * (k & 3));
              ptx::umma_commit(ptx::smem_u32(&bars->a_free[as]));
            }
            if (++aslot == kASlots) {
              aslot = 0;
              aph ^= 1;
            }
          }
          ptx::mbar_wait(ptx::smem_u32(&bars->full[stage]), ph);
          ptx::tc_fence_after();
          if (leader) {
            // W: k-block j of the stage at +16 KB (>>4 = 1024), +32 bytes (>>4 = 2) per K=16 step
            const uint64_t db = desc_hi | (uint64_t)(((sW + stage * kWStageBytes) & 0x3FFFF) >> 4);
#pragma unroll
            for (int k = 0; k < 2 * (kBlockK / 16); ++k)  // A: +8 columns per K=16 step
              if (k < nk)
                ptx::umma_bf16_ts(d_tmem, ta + 8 * k, db + (k >> 2) * (kWBlockBytes >> 4) + 2 * (k & 3),
                                  idesc, (kp | k) != 0);
            ptx::umma_commit(ptx::smem_u32(&bars->empty[stage]));
          }
          if (++stage == kStages) {
            stage = 0;
            ph ^= 1;
          }
        }
        if (leader) {
          ptx::umma_commit(ptx::smem_u32(&bars->tmem_full[acc]));
          if (n == NT - 1) ptx::umma_commit(ptx::smem_u32(&bars->a_done));
        }
        acc ^= 1;
        if (acc == 0) acc_ph ^= 1;
      }
    }
    __syncwarp();
  } else if (MODE != MODE_FWD && warp == 3) {
    // ===================== A producer (MODE_GRAD / MODE_FWDP): TMA ring fed from the bf16 H scratch =====================
    const bool leader = ptx::elect_one();
    const uint64_t h_policy = ptx::l2_policy_evict_first();  // each H tile is read once here
    int aslot = 0, aph = 0;
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x)
      for (int kp = 0; kp < nkp; ++kp) {
        ptx::mbar_wait(ptx::smem_u32(&bars->a_free[aslot]), aph ^ 1);
        if (leader) {
          const uint32_t fb = ptx::smem_u32(&bars->a_ready[aslot]);
          const bool two = 2 * kp + 1 < nkb;
          ptx::mbar_arrive_expect_tx(fb, two ? kASlotBytes : kABlockBytes);
          const uint32_t dst = sA + aslot * kASlotBytes;
          ptx::tma_load_2d_hint(dst, &h_map, fb, 2 * kp * kBlockK, tile * kTileM, h_policy);
          if (two)
            ptx::tma_load_2d_hint(dst + kABlockBytes, &h_map, fb, (2 * kp + 1) * kBlockK, tile * kTileM,
                                  h_policy);
        }
        if (++aslot == kASlots) {
          aslot = 0;
          aph ^= 1;
        }
      }
    __syncwarp();
  } else if (MODE == MODE_FWD && warp >= kBuilderWarp0) {
    // ===================== A-tile builders: H = tanh(e + p) -> bf16, swizzled staging =====================
    const int bw = warp - kBuilderWarp0;  // rows bw*32 .. bw*32+31
    const int rr = lane >> 3;             // row within a group of 4
    const int cc = lane & 7;              // 16-byte chunk within the 128-byte swizzle row
    // With P.H set (training forward) every finished [32 rows x 64 k] piece also leaves as a TMA
    // store into the bf16 H scratch: the backward pass reads H instead of recomputing tanh(e + p)
    // in a kernel of its own (joint_h_kernel: 0.7 ms per cfg3 step).
    const bool save_h = P.H != nullptr;
    const uint64_t hs_policy = ptx::l2_policy_evict_first();
    int it = 0;  // slot = pair index kp (the staging holds a whole tile), phase = tile parity
    for (int tile = blockIdx.x; tile < ntiles; tile += gridDim.x, ++it) {
      if (save_h && it > 0) {  // the previous tile's H stores must have read the staging area
        if (lane == 0) ptx::bulk_wait_group_read0();
        __syncwarp();
      }
      int eoff[8], poff[8];
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const int r = bw * 32 + i * 4 + rr;
        long long cell = P.cell0 + (long long)tile * kTileM + r;
        if (cell >= P.total_cells) cell = P.total_cells - 1;
        const int bt = (int)(cell / P.U1);
        const int u = (int)(cell - (long long)bt * P.U1);
        const int b = bt / P.T;
        eoff[i] = bt * P.J;  // < 2^31: B*T*J and B*U1*J elements are checked on the host
        poff[i] = (b * P.U1 + u) * P.J;
      }
      for (int kp = 0; kp < nkp; ++kp) {
        ptx::mbar_wait(ptx::smem_u32(&bars->a_free[kp]), (it & 1) ^ 1);
        for (int kb = 2 * kp; kb < 2 * kp + 2 && kb < nkb; ++kb) {
          const int k0 = kb * kBlockK + cc * 8;
          uint8_t* ablk = smem + kb * kABlockBytes;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const int r = bw * 32 + i * 4 + rr;
            const float4* ep = reinterpret_cast<const float4*>(P.e + eoff[i] + k0);
            const float4* pp = reinterpret_cast<const float4*>(P.p + poff[i] + k0);
            const float4 e0 = __ldg(ep), e1 = __ldg(ep + 1);
            const float4 p0 = __ldg(pp), p1 = __ldg(pp + 1);
            uint4 h;
            h.x = pack_bf16x2(tanh_approx(e0.x + p0.x), tanh_approx(e0.y + p0.y));
            h.y = pack_bf16x2(tanh_approx(e0.z + p0.z), tanh_approx(e0.w + p0.w));
            h.z = pack_bf16x2(tanh_approx(e1.x + p1.x), tanh_approx(e1.y + p1.y));
            h.w = pack_bf16x2(tanh_approx(e1.z + p1.z), tanh_approx(e1.w + p1.w));
            *reinterpret_cast<uint4*>(ablk + r * 128 + ((cc ^ (r & 7)) << 4)) = h;
          }
        }
        ptx::fence_proxy_async();  // generic-proxy stores -> visible to the tensor-core proxy
        __syncwarp();
        if (lane == 0) {
          ptx::mbar_arrive(ptx::smem_u32(&bars->a_ready[kp]));
          if (save_h) {
            for (int kb = 2 * kp; kb < 2 * kp + 2 && kb < nkb; ++kb)
              ptx::tma_store_2d_hint(&h_map, base + kb * kABlockBytes + bw * 32 * 128, kb * kBlockK,
                                     tile * kTileM + bw * 32, hs_policy);
            ptx::bulk_commit_group();
          }
        }
      }
    }
    if (save_h && lane == 0) ptx::bulk_wait_group0();
  } else if (warp >= kEpiWarp0 && warp < kBuilderWarp0) {
    // ===================== epilogue =====================
#include "joint_fused_epilogue.inc"
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem, kTmemCols);
  }
}

// H[cell - cell0, 0:J] = bf16(tanh(e[b,t,:] + p[b,u,:])) for one chunk of cells: the A operand of
// the backward joint kernel (read back through TMA while still in L2) and of the dW GEMM.
// Same arithmetic as the forward builders (tanh.approx.f32, round-to-nearest bf16).
// Block = J/8 threads x (256 / (J/8)) frame slots: a thread owns 8 consecutive k of one (b, t) frame, keeps
// that frame's e values in registers and walks the U1 labels (the fp32 e row was re-read through L1 for
// every cell before: 0.42 of HBM peak, issue-bound).  Two labels per pass keep four 16-byte p loads in
// flight; 32-bit index arithmetic only (the host checks cells, B*T*J, B*U1*J < 2^31).  Chunks that do
// not start / end on a frame boundary are handled by clamping the label range of the edge frames.
__global__ void __launch_bounds__(256)
joint_h_kernel(const float* __restrict__ e, const float* __restrict__ p, int T, int U1, int J,
               long long cell0, long long ncells, __nv_bfloat16* __restrict__ H, int ldh) {
  const int g8 = J >> 3;                      // 16-byte groups per row
  const int rows = blockDim.x / g8;           // frames per block and pass
  const int r = threadIdx.x / g8;
  const int k0 = (threadIdx.x - r * g8) << 3;
  if (r >= rows) return;
  const unsigned c0 = (unsigned)cell0, c1 = (unsigned)(cell0 + ncells);  // cells [c0, c1)
  const unsigned bt0 = c0 / (unsigned)U1, bt1 = (c1 + U1 - 1) / (unsigned)U1;  // frames touched
  for (unsigned bt = bt0 + blockIdx.x * rows + r; bt < bt1; bt += gridDim.x * rows) {
    const unsigned b = bt / (unsigned)T;
    const float4* ep = reinterpret_cast<const float4*>(e + (size_t)bt * J + k0);
    const float4 e0 = __ldg(ep), e1 = __ldg(ep + 1);
    const unsigned first = bt * (unsigned)U1;
    const int ua = first >= c0 ? 0 : (int)(c0 - first);
    const int ub = first + U1 <= c1 ? U1 : (int)(c1 - first);
    const float* pb = p + (size_t)b * U1 * J + k0;
    __nv_bfloat16* hb = H + ((size_t)first - c0) * ldh + k0;
    auto one = [&](const float4 p0, const float4 p1) {
      uint4 h;
      h.x = pack_bf16x2(tanh_approx(e0.x + p0.x), tanh_approx(e0.y + p0.y));
      h.y = pack_bf16x2(tanh_approx(e0.z + p0.z), tanh_approx(e0.w + p0.w));
      h.z = pack_bf16x2(tanh_approx(e1.x + p1.x), tanh_approx(e1.y + p1.y));
      h.w = pack_bf16x2(tanh_approx(e1.z + p1.z), tanh_approx(e1.w + p1.w));
      return h;
    };
    int u = ua;
    for (; u + 1 < ub; u += 2) {
      const float4* pa = reinterpret_cast<const float4*>(pb + (size_t)u * J);
      const float4* pc = reinterpret_cast<const float4*>(pb + (size_t)(u + 1) * J);
      const float4 a0 = __ldg(pa), a1 = __ldg(pa + 1), b0 = __ldg(pc), b1 = __ldg(pc + 1);
      *reinterpret_cast<uint4*>(hb + (ptrdiff_t)u * ldh) = one(a0, a1);
      *reinterpret_cast<uint4*>(hb + (ptrdiff_t)(u + 1) * ldh) = one(b0, b1);
    }
    if (u < ub) {
      const float4* pa = reinterpret_cast<const float4*>(pb + (size_t)u * J);
      *reinterpret_cast<uint4*>(hb + (ptrdiff_t)u * ldh) = one(__ldg(pa), __ldg(pa + 1));
    }
  }
}

}  // namespace jf
}  // namespace b200

#include "joint_fused_host.inc"
