// d e, d p from G: the hidden-side gradient of the fused joint backward as ONE 2-CTA tcgen05 kernel
//   dH = G . W_out            (autograd of ffn_out, wenet/transducer/joint.py:68-69)
//   dZ = dH * (1 - H^2)       (tanh, joint.py:63-64)
//   d e[b,t,:] = sum_u dZ     d p[b,u,:] = sum_t dZ      (broadcast add, joint.py:58-60)
// reached from wenet/transducer/transducer.py:132-147.  Replaces a library GEMM that wrote a
// [cells, J] fp32 dH (3.3 GB at cfg3) plus a separate reduction kernel that read it back.
//
// The product is computed TRANSPOSED, dH^T [J, cells] = W^T [J, V] . G^T [V, cells], so that a TMEM
// lane is one j and the columns are lattice cells: both reductions then run over the registers of
// one thread (no shuffles, no shared memory), and every global access of the epilogue is
// coalesced over j.
//   A = W^T   M = j, K = v: W_out [V, J] row-major is MN-major; [64 v x 64 j] SWIZZLE_128B boxes
//   B = G     N = cells, K = v: G [cells, ldg] row-major is K-major; one box of 128 rows x 64 v
//   cta_group::2, M = 256 (128 j per CTA), N = 256 cells (128 rows from each CTA), one accumulator per
//   256 j: J = 512 fills all 512 TMEM columns of both SMs; 0.0059 B of shared-memory fill per flop.
// A tile is NOT 256 consecutive cells but a block of UB labels x 256/UB frames of one utterance
// (a 4-D TMA box over G viewed as [b, t, u, v]): the thread then holds UB running sums for d p
// (over the tile's frames) and one running sum per frame for d e, i.e. 256/UB + UB atomics per thread
// and tile instead of 256.  Frames / labels beyond T / U1 are zero-filled by TMA and add nothing.
// Roofline: tensor pipe; algorithmic flops per launch = 2 * cells * J * V.
#include <cuda.h>

#include "common.cuh"
#include "sm100_pair.cuh"
#include "sm100_ptx.cuh"

namespace b200 {
namespace dh {

using namespace pair;

constexpr int kBlockK = 64;   // vocabulary entries per pipeline stage
constexpr int kStages = 4;
constexpr int kTileN = 256;   // cells per tile (128 rows of G from each CTA of the pair)
constexpr int kBoxBytes = 64 * kBlockK * 2;       // [64 v x 64 j] bf16 = 8 KB
constexpr int kBBytes = 128 * kBlockK * 2;        // [128 cells x 64 v] bf16 = 16 KB
constexpr int kEpiWarp0 = 2;  // warp 0 TMA, warp 1 MMA + TMEM, warps 2.. epilogue (4 per accumulator)
constexpr int kThreads = (kEpiWarp0 + 8) * 32;
constexpr int kMaxJ = 512;

struct Params {
  int nb, T, U1, V, J;
  int ntb, nub;          // frame blocks / label blocks per utterance
  int kblocks;           // ceil(V / 64)
  const __nv_bfloat16* H;  // [nb*T*U1, ldh]
  const __nv_bfloat16* H_lo;  // optional: H = H + H_lo (split operands, fp32-accurate mode)
  int ldh;
  int npass;             // 1: G.W   3: G_hi.W_hi + G_lo.W_hi + G_hi.W_lo
  long long rows;        // PLAIN: rows of G (= nb, with T = U1 = 1 unused)
  float* out;            // PLAIN: [rows, J] fp32 = G.W, written (no tanh', no sums)
  const int32_t* logit_len;   // optional [nb]: tiles beyond T_b / U_b are skipped (their G is zero)
  const int32_t* target_len;
  const float* rowscale; // optional [nb*T*U1]: dH = rowscale[cell] * (G W) (no-recompute backward, joint_nr.cu)
  float* de;             // [nb, T, J]  accumulated (zero it first)
  float* dp;             // [nb, U1, J] accumulated
};

__host__ __device__ constexpr int stage_bytes(int J) { return (J / 64) / 2 * kBoxBytes + kBBytes; }

// MN-major SWIZZLE_128B operand made of 8 KB boxes [64 k rows x 128 bytes] (see joint_dw.cu)
__device__ __forceinline__ uint64_t make_sw128_mnmajor_desc(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFF) >> 4);
  d |= (uint64_t)(kBoxBytes >> 4) << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// kind::f16, bf16 x bf16 -> f32, A MN-major (bit 15), B K-major
__host__ __device__ constexpr uint32_t make_idesc_amn(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | ((uint32_t)(N >> 3) << 17) |
         ((uint32_t)(M >> 4) << 24);
}
// 4-D TMA load into THIS CTA's shared memory, completion on the leader CTA's mbarrier
__device__ __forceinline__ void tma_load_4d_pair(uint32_t dst, const CUtensorMap* m, uint32_t bar, int c0,
                                                 int c1, int c2, int c3, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.4d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes"
      ".L2::cache_hint [%0], [%1, {%3, %4, %5, %6}], [%2], %7;" ::"r"(dst),
      "l"(m), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "l"(policy)
      : "memory");
}
// volatile: keeps the H loads where the epilogue's software pipeline places them (ptxas otherwise hoists all
// 256 of a tile above the accumulator wait and spills)
__device__ __forceinline__ uint32_t ldg_u32_pinned(const void* p) {
  uint32_t v;
  asm volatile("ld.global.nc.u32 %0, [%1];" : "=r"(v) : "l"(p));
  return v;
}
__device__ __forceinline__ void red_add(float* p, float v) {
  asm volatile("red.global.add.f32 [%0], %1;" ::"l"(p), "f"(v) : "memory");
}

// (utterance, frame block, label block) of work unit u; label block fastest
struct Tile {
  int b, t0, u0;
};
__device__ __forceinline__ Tile decode(const Params& P, int u, int tb) {
  Tile t;
  const int ub_i = u % P.nub;
  const int r = u / P.nub;
  t.b = r / P.ntb;
  t.t0 = (r - t.b * P.ntb) * tb;
  t.u0 = ub_i;
  return t;
}

// PLAIN: the same product for a plain row-major G [rows, ldg] (2-D TMA map, tiles of 256 consecutive rows),
// epilogue out[row, j] = (G W)[row, j] - the d hs GEMM of the CTC projection (autograd of ctc.py:57)
// SPLIT: fp32-accurate mode (three passes over split operands, H = H_hi + H_lo read inside the epilogue).
// !SPLIT: the tile's 256 H values of a thread are loaded into registers (packed pairs) BEFORE the wait for
// the accumulator, i.e. under the tile's own MMA main loop: the epilogue that the tensor pipe has to wait
// for (one accumulator set fills all of TMEM, so there is no second buffer to compute into) is then pure
// register work - tcgen05.ld one chunk ahead, 5 instructions per cell, no memory latency on the chain.
template <int UB, bool PLAIN, bool SPLIT>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(kThreads, 1)
    joint_dh_kernel(const __grid_constant__ CUtensorMap wmap, const __grid_constant__ CUtensorMap gmap,
                    const __grid_constant__ CUtensorMap wmap_lo, const __grid_constant__ CUtensorMap gmap_lo,
                    const Params P) {
  constexpr int TBH = 128 / UB;  // frames per CTA half of a tile
  constexpr int TB = 2 * TBH;    // frames per tile
  extern __shared__ uint8_t smem_raw[];
  __shared__ __align__(8) uint64_t bars[2 * kStages + 2];
  __shared__ uint32_t tmem_holder;
  __shared__ __align__(16) float rs_s[2][kTileN];  // per-cell factors of the tile (1 without P.rowscale)

  const uint32_t rank = cluster_ctarank();
  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const int J = P.J;
  const int nacc = J / 256;
  const int sbytes = stage_bytes(J);
  const int abytes = sbytes - kBBytes;
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t bar0 = ptx::smem_u32(bars);
  auto full_bar = [&](int s) { return bar0 + 8u * s; };
  auto empty_bar = [&](int s) { return bar0 + 8u * (kStages + s); };
  const uint32_t acc_full = bar0 + 8u * (2 * kStages);
  const uint32_t acc_empty = bar0 + 8u * (2 * kStages + 1);

  if (threadIdx.x == 0) {
    for (int s = 0; s < kStages; ++s) {
      ptx::mbar_init(full_bar(s), 1);
      ptx::mbar_init(empty_bar(s), 1);
    }
    ptx::mbar_init(acc_full, 1);
    ptx::mbar_init(acc_empty, 8 * nacc);  // 4 epilogue warps per accumulator in each CTA of the pair
    ptx::fence_barrier_init();
    ptx::prefetch_tensormap(&wmap);
    ptx::prefetch_tensormap(&gmap);
  }
  if (warp == 1) {
    tmem_alloc_pair(ptx::smem_u32(&tmem_holder), 512);
    tmem_relinquish_pair();
  }
  ptx::tc_fence_before();
  cluster_sync_all();
  ptx::tc_fence_after();
  const uint32_t tmem_base = tmem_holder;

  const int nunits = PLAIN ? (int)((P.rows + 255) / 256) : P.nb * P.ntb * P.nub;
  const int pair_id = (int)cluster_id_x();
  const int npairs = (int)nclusters_x();
  // a tile whose frames / labels all lie beyond the utterance's lengths holds G == 0: every role skips it
  auto live = [&](const Tile& t) {
    if (PLAIN || P.logit_len == nullptr) return true;
    return t.t0 < __ldg(P.logit_len + t.b) && t.u0 * UB <= __ldg(P.target_len + t.b);
  };

  if (warp == 0) {
    // ---------------- TMA producer (both CTAs; each loads its own half of the operands)
    const uint32_t leader_full0 = map_to_cta(full_bar(0), 0);
    const uint64_t pol_stream = ptx::l2_policy_evict_first();
    const uint64_t pol_keep = ptx::l2_policy_evict_last();
    long long it = 0;
    for (int u = pair_id; u < nunits; u += npairs) {
      const Tile tl = decode(P, u, TB);
      if (!live(tl)) continue;
      for (int pass = 0; pass < P.npass; ++pass) {
        const CUtensorMap* wm = pass == 2 ? &wmap_lo : &wmap;
        const CUtensorMap* gm = pass == 1 ? &gmap_lo : &gmap;
        for (int kb = 0; kb < P.kblocks; ++kb, ++it) {
          const int s = (int)(it % kStages);
          const uint32_t ph = (uint32_t)((it / kStages) & 1);
          ptx::mbar_wait(empty_bar(s), ph ^ 1u);
          if (ptx::elect_one()) {
            if (rank == 0) ptx::mbar_arrive_expect_tx(full_bar(s), 2u * (uint32_t)sbytes);
            const uint32_t bar = leader_full0 + 8u * s;
            const uint32_t dst = smem_base + (uint32_t)s * sbytes;
            const int v0 = kb * kBlockK;
            // A: W^T, this CTA's 128 j of every accumulator (two boxes of 64 j); shared by all tiles
            uint32_t adst = dst;
            for (int a = 0; a < nacc; ++a) {
              const int j0 = 256 * a + (int)rank * 128;
              tma_load_2d_pair(adst, wm, bar, j0, v0, pol_keep);
              tma_load_2d_pair(adst + kBoxBytes, wm, bar, j0 + 64, v0, pol_keep);
              adst += 2 * kBoxBytes;
            }
            // B: G rows of UB labels x TBH frames (this CTA's half of the tile); read once per pass
            if (PLAIN)
              tma_load_2d_pair(dst + abytes, gm, bar, v0, u * 256 + (int)rank * 128, pol_stream);
            else
              tma_load_4d_pair(dst + abytes, gm, bar, v0, tl.u0 * UB, tl.t0 + (int)rank * TBH, tl.b, pol_stream);
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 1) {
    // ---------------- MMA issuer (leader CTA only)
    if (rank == 0) {
      constexpr uint32_t idesc = make_idesc_amn(256, kTileN);
      long long it = 0;
      int nu = 0;
      for (int u = pair_id; u < nunits; u += npairs) {
        const Tile tl = decode(P, u, TB);
        if (!live(tl)) continue;
        ptx::mbar_wait(acc_empty, (uint32_t)((nu & 1) ^ 1));  // drained by both CTAs' epilogues
        ++nu;
        ptx::tc_fence_after();
        for (int pass = 0; pass < P.npass; ++pass)
          for (int kb = 0; kb < P.kblocks; ++kb, ++it) {
            const int s = (int)(it % kStages);
            const uint32_t ph = (uint32_t)((it / kStages) & 1);
            ptx::mbar_wait(full_bar(s), ph);
            ptx::tc_fence_after();
            if (ptx::elect_one()) {
              const uint32_t sa = smem_base + (uint32_t)s * sbytes;
              const uint64_t bdesc0 = ptx::make_sw128_kmajor_desc(sa + abytes);
#pragma unroll
              for (int k = 0; k < kBlockK / 16; ++k) {
                for (int a = 0; a < nacc; ++a) {
                  const uint64_t adesc = make_sw128_mnmajor_desc(sa + a * 2 * kBoxBytes + k * 2048);
                  umma_bf16_pair(tmem_base + 256u * a, adesc, bdesc0 + 2 * k, idesc,
                                 (pass > 0 || kb > 0 || k > 0) ? 1u : 0u);
                }
              }
              umma_commit_pair(empty_bar(s));
              if (pass == P.npass - 1 && kb == P.kblocks - 1) umma_commit_pair(acc_full);
            }
            __syncwarp();
          }
      }
    }
  } else if (warp - kEpiWarp0 < 4 * nacc) {
    // ---------------- epilogue (both CTAs): thread = one j, columns = the tile's cells
    const int q = warp & 3;                  // TMEM lane quarter this warp may access
    const int a = (warp - kEpiWarp0) >> 2;   // accumulator
    const int j = 256 * a + (int)rank * 128 + q * 32 + lane;
    const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + 256u * a;
    int nu = 0;
    for (int u = pair_id; u < nunits; u += npairs) {
      const Tile tl = decode(P, u, TB);
      if (!live(tl)) continue;
      const int u0 = tl.u0 * UB;
      if (PLAIN) {
        // out[row, j] for the 256 rows of this tile (column c = row u * 256 + c): plain coalesced stores
        ptx::mbar_wait(acc_full, (uint32_t)(nu & 1));
        ++nu;
        ptx::tc_fence_after();
        const long long row0 = (long long)u * 256;
#pragma unroll 1
        for (int ch = 0; ch < 16; ++ch) {
          uint32_t r[16];
          ptx::tmem_ld16(taddr + ch * 16, r);
          ptx::tmem_ld_wait();
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            const long long row = row0 + ch * 16 + k;
            if (row < P.rows) P.out[row * J + j] = __uint_as_float(r[k]);
          }
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(map_to_cta(acc_empty, 0));
        continue;
      }
      // per-cell factors of the tile's 256 columns -> shared memory (double-buffered over tiles: the
      // barrier of the next tile orders this tile's reads before the buffer's reuse)
      auto stage_rowscale = [&](int tile_no) {
        float* const rs_t = rs_s[tile_no & 1];
        for (int c = (warp - kEpiWarp0) * 32 + lane; c < kTileN; c += 4 * nacc * 32) {
          float w = 1.f;
          if (P.rowscale != nullptr) {
            const int row = c & 127;
            const int t = min(tl.t0 + (c >> 7) * TBH + row / UB, P.T - 1);
            const long long cell = (long long)(tl.b * P.T + t) * P.U1 + min(u0 + row % UB, P.U1 - 1);
            w = __ldg(P.rowscale + cell);
          }
          rs_t[c] = w;
        }
        asm volatile("bar.sync 1, %0;" ::"r"(4 * nacc * 32) : "memory");
        return rs_t;
      };
      if constexpr (!SPLIT) {
        // column c = 128 * half + UB * f + i: frame f of CTA half `half`, label i of the tile (the order the
        // 4-D TMA box lays the G rows out); chunk ch = columns 16 ch .. 16 ch + 15.  Frames / labels beyond
        // T / U1 have a zero accumulator: their H index is clamped, the value is irrelevant.
        // The H values of chunks 0 .. kAhead-1 are loaded BEFORE the accumulator wait (under the tile's MMA
        // main loop), those of chunk ch + kAhead while chunk ch is processed (168 registers per thread at
        // 10 warps: all 256 values at once would spill)
        // Lanes 2m / 2m+1 (j, j+1) share their loads: each fetches the 32-bit word {H[.., j], H[.., j+1]} of ONE
        // cell of a column pair (even lane: the even column, odd lane: the odd one) and the pair is
        // recombined by one shuffle + one byte permute at consumption - 128 loads and 128 registers per
        // thread and tile instead of 256, and no pack instruction waiting on a load
        constexpr int kAhead = 8;
        const int par = lane & 1;
        const uint16_t* Hj = reinterpret_cast<const uint16_t*>(P.H) + (j & ~1);
        const uint32_t psel = par ? 0x3276u : 0x5410u;
        auto load_chunk = [&](int ch, uint32_t* dst) {
#pragma unroll
          for (int k = 0; k < 16; k += 2) {
            const int c = ch * 16 + k, row = c & 127;
            const int t = min(tl.t0 + (c >> 7) * TBH + row / UB, P.T - 1);
            const uint32_t cell = (uint32_t)((tl.b * P.T + t) * P.U1 + min(u0 + row % UB + par, P.U1 - 1));
            dst[k >> 1] = ldg_u32_pinned(Hj + cell * (uint32_t)P.ldh);  // < 2^31 elements (host check)
          }
        };
        uint32_t hreg[16][8];
#pragma unroll
        for (int ch = 0; ch < kAhead; ++ch) load_chunk(ch, hreg[ch]);
        float* const rs_t = stage_rowscale(nu);
        float dp_acc[UB];
#pragma unroll
        for (int i = 0; i < UB; ++i) dp_acc[i] = 0.f;
        float de_run = 0.f;
        ptx::mbar_wait(acc_full, (uint32_t)(nu & 1));
        ++nu;
        ptx::tc_fence_after();
        uint32_t r[2][16];
        uint32_t hpair = 0;
        ptx::tmem_ld16(taddr, r[0]);
#pragma unroll
        for (int ch = 0; ch < 16; ++ch) {
          if (ch + kAhead < 16) load_chunk(ch + kAhead, hreg[ch + kAhead]);
          ptx::tmem_ld_wait();
          if (ch + 1 < 16) ptx::tmem_ld16(taddr + (ch + 1) * 16, r[(ch + 1) & 1]);
          float4 w4[4];
#pragma unroll
          for (int k = 0; k < 4; ++k) w4[k] = reinterpret_cast<const float4*>(rs_t)[ch * 4 + k];
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            const int c = ch * 16 + k;
            const float4 wq = w4[k >> 2];
            const float wk = (k & 3) == 0 ? wq.x : ((k & 3) == 1 ? wq.y : ((k & 3) == 2 ? wq.z : wq.w));
            uint32_t hp = 0;
            if ((k & 1) == 0) {
              const uint32_t x = hreg[ch][k >> 1];
              hpair = __byte_perm(x, __shfl_xor_sync(0xffffffffu, x, 1), psel);
            }
            hp = hpair;
            const float hv = __uint_as_float((k & 1) ? (hp & 0xffff0000u) : (hp << 16));
            const float z = (__uint_as_float(r[ch & 1][k]) * wk) * fmaf(-hv, hv, 1.f);
            dp_acc[c % UB] += z;
            de_run += z;
            if (c % UB == UB - 1) {  // last label of a frame: flush d e
              const int t = tl.t0 + (c >> 7) * TBH + (c & 127) / UB;
              if (t < P.T) red_add(P.de + (long long)(tl.b * P.T + t) * J + j, de_run);
              de_run = 0.f;
            }
          }
        }
        ptx::tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(map_to_cta(acc_empty, 0));
#pragma unroll
        for (int i = 0; i < UB; ++i)
          if (u0 + i < P.U1) red_add(P.dp + (long long)(tl.b * P.U1 + u0 + i) * J + j, dp_acc[i]);
        continue;
      }
      // chunk ch = columns 16ch .. 16ch+15 = rows (ch & 7) * 16 + k of CTA half ch >> 3; row r of a half is
      // frame r / UB, label r % UB of the tile (the order the 4-D TMA box lays the G rows out).  Frames /
      // labels beyond T / U1 have a zero accumulator: their H index is clamped, the value is irrelevant
      const uint16_t* Hj = reinterpret_cast<const uint16_t*>(P.H) + j;
      const uint16_t* Hlj = P.H_lo ? reinterpret_cast<const uint16_t*>(P.H_lo) + j : nullptr;
      auto load_h = [&](int ch, uint16_t* h, uint16_t* hl) {
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          const int row = (ch & 7) * 16 + k;
          const int t = min(tl.t0 + (ch >> 3) * TBH + row / UB, P.T - 1);
          const int uu = min(u0 + row % UB, P.U1 - 1);
          const uint32_t off = (uint32_t)(((tl.b * P.T + t) * P.U1 + uu) * P.ldh);  // < 2^31 (host check)
          h[k] = __ldg(Hj + off);
          if (Hlj) hl[k] = __ldg(Hlj + off);
        }
      };
      float dp_acc[UB];
#pragma unroll
      for (int i = 0; i < UB; ++i) dp_acc[i] = 0.f;
      float de_run = 0.f;
      uint16_t hbuf[2][16], lbuf[2][16];
      load_h(0, hbuf[0], lbuf[0]);
      const float* const rs_t = stage_rowscale(nu);
      ptx::mbar_wait(acc_full, (uint32_t)(nu & 1));
      ++nu;
      ptx::tc_fence_after();
#pragma unroll 1
      for (int c2 = 0; c2 < 8; ++c2) {
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {  // chunk parity is static: hbuf / dp_acc indices stay in registers
          const int ch = 2 * c2 + hh;
          uint32_t r[16];
          ptx::tmem_ld16(taddr + ch * 16, r);
          if (ch + 1 < 16) load_h(ch + 1, hbuf[hh ^ 1], lbuf[hh ^ 1]);
          ptx::tmem_ld_wait();
          const uint16_t* h = hbuf[hh];
          const uint16_t* hl = lbuf[hh];
#pragma unroll
          for (int k = 0; k < 16; ++k) {
            const int rs = hh * 16 + k;  // row within a pair of chunks; UB divides 32
            float hv = __uint_as_float((uint32_t)h[k] << 16);
            if (Hlj) hv += __uint_as_float((uint32_t)hl[k] << 16);
            const float z = (__uint_as_float(r[k]) * rs_t[ch * 16 + k]) * fmaf(-hv, hv, 1.f);
            dp_acc[rs % UB] += z;
            de_run += z;
            if (rs % UB == UB - 1) {  // last label of a frame: flush d e
              const int t = tl.t0 + (ch >> 3) * TBH + ((ch & 7) * 16 + k) / UB;
              if (t < P.T) red_add(P.de + (long long)(tl.b * P.T + t) * J + j, de_run);
              de_run = 0.f;
            }
          }
        }
      }
      // accumulator is in registers / flushed: hand TMEM back to the MMA warp
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(map_to_cta(acc_empty, 0));
#pragma unroll
      for (int i = 0; i < UB; ++i)
        if (u0 + i < P.U1) red_add(P.dp + (long long)(tl.b * P.U1 + u0 + i) * J + j, dp_acc[i]);
    }
  }

  // teardown: both CTAs done with TMEM and with each other's shared memory / barriers
  ptx::tc_fence_before();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc_pair(tmem_base, 512);
}

}  // namespace dh
}  // namespace b200

using namespace b200;

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn dh_encode_fn() {
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

int encode(CUtensorMap* map, const void* ptr, int rank, const cuuint64_t* dims, const cuuint64_t* strides,
           const cuuint32_t* box) {
  EncodeTiledFn fn = dh_encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled entry point not available");
    return B200_ERR_CUDA;
  }
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, (cuuint32_t)rank, const_cast<void*>(ptr), dims,
                  strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r);
    return B200_ERR_CUDA;
  }
  return B200_OK;
}

template <int UB, bool PLAIN, bool SPLIT>
int launch_dh(const CUtensorMap& wmap, const CUtensorMap& wmap_lo, const void* G, const void* G_lo, int ldg,
              dh::Params P, int ctas, cudaStream_t st) {
  constexpr int TBH = 128 / UB;
  CUtensorMap gmap, gmap_lo;
  long long units;
  if (PLAIN) {
    // G [rows, ldg]: boxes of [128 rows x 64 v]; rows beyond `rows`, v beyond V read as zero
    cuuint64_t dims[2] = {(cuuint64_t)P.V, (cuuint64_t)P.rows};
    cuuint64_t strides[1] = {(cuuint64_t)ldg * 2};
    cuuint32_t box[2] = {(cuuint32_t)dh::kBlockK, 128};
    if (int rc = encode(&gmap, G, 2, dims, strides, box)) return rc;
    gmap_lo = gmap;
    if (G_lo)
      if (int rc = encode(&gmap_lo, G_lo, 2, dims, strides, box)) return rc;
    units = (P.rows + 255) / 256;
    P.nb = (int)units;
    P.ntb = P.nub = 1;
  } else {
    P.ntb = (P.T + 2 * TBH - 1) / (2 * TBH);
    P.nub = (P.U1 + UB - 1) / UB;
    // G [nb*T*U1, ldg] viewed as [b, t, u, v]; v beyond V, u beyond U1, t beyond T read as zero
    cuuint64_t dims[4] = {(cuuint64_t)P.V, (cuuint64_t)P.U1, (cuuint64_t)P.T, (cuuint64_t)P.nb};
    cuuint64_t strides[3] = {(cuuint64_t)ldg * 2, (cuuint64_t)P.U1 * ldg * 2, (cuuint64_t)P.T * P.U1 * ldg * 2};
    cuuint32_t box[4] = {(cuuint32_t)dh::kBlockK, (cuuint32_t)UB, (cuuint32_t)TBH, 1};
    if (int rc = encode(&gmap, G, 4, dims, strides, box)) return rc;
    gmap_lo = gmap;
    if (G_lo)
      if (int rc = encode(&gmap_lo, G_lo, 4, dims, strides, box)) return rc;
    units = (long long)P.nb * P.ntb * P.nub;
  }
  const int smem = dh::kStages * dh::stage_bytes(P.J) + 1024;
  B200_CHECK_CUDA(cudaFuncSetAttribute(dh::joint_dh_kernel<UB, PLAIN, SPLIT>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       dh::kStages * dh::stage_bytes(dh::kMaxJ) + 1024));
  int grid = ctas;
  if (units < grid / 2) grid = 2 * (int)units;
  dh::joint_dh_kernel<UB, PLAIN, SPLIT><<<grid, dh::kThreads, smem, st>>>(wmap, gmap, wmap_lo, gmap_lo, P);
  B200_CHECK_LAUNCH("joint_dh_kernel");
  return B200_OK;
}

int encode_w(CUtensorMap* wmap, const void* w, int V, int J) {
  // W [V, J] bf16: boxes of [64 v x 64 j] (MN-major A operand); rows beyond V read as zero
  cuuint64_t dims[2] = {(cuuint64_t)J, (cuuint64_t)V};
  cuuint64_t strides[1] = {(cuuint64_t)J * 2};
  cuuint32_t box[2] = {64, (cuuint32_t)dh::kBlockK};
  return encode(wmap, w, 2, dims, strides, box);
}

int dh_ctas(int max_ctas) {
  int sms = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  int ctas = sms & ~1;
  if (max_ctas > 0 && max_ctas < ctas) ctas = max_ctas & ~1;
  if (ctas < 2) ctas = 2;
  return ctas;
}

int joint_dh_impl(const void* G, const void* G_lo, int ldg, const void* w, const void* w_lo, const void* H,
                  const void* H_lo, int ldh, const float* rowscale, int nb, int T, int U1, int V, int J,
                  const int32_t* logit_len, const int32_t* target_len, float* de, float* dp, int max_ctas,
                  b200_stream_t stream) {

  B200_CHECK_ARG(G && w && H && de && dp, "b200_joint_dh: null pointer");
  B200_CHECK_ARG(nb > 0 && T > 0 && U1 > 0 && V > 0, "b200_joint_dh: bad shape nb=%d T=%d U1=%d V=%d", nb, T,
                 U1, V);
  B200_CHECK_ARG(J == 256 || J == 512, "b200_joint_dh: join dim J=%d must be 256 or 512", J);
  B200_CHECK_ARG(ldg >= V && ldg % 8 == 0 && ((uintptr_t)G & 15) == 0 && ((uintptr_t)G_lo & 15) == 0,
                 "b200_joint_dh: ldg=%d must be >= V and a multiple of 8, G 16-byte aligned", ldg);
  B200_CHECK_ARG(ldh >= J && ldh % 2 == 0 && ((uintptr_t)H & 3) == 0 && ((uintptr_t)H_lo & 1) == 0,
                 "b200_joint_dh: ldh=%d must be >= J and even, H 4-byte aligned", ldh);
  B200_CHECK_ARG(((uintptr_t)w & 15) == 0 && ((uintptr_t)w_lo & 15) == 0,
                 "b200_joint_dh: W must be 16-byte aligned");
  B200_CHECK_ARG((long long)nb * T * U1 * ldh < (1ll << 31),
                 "b200_joint_dh: chunk of %d utterances too large (H chunk must stay below 2^31 elements)", nb);
  B200_CHECK_ARG((logit_len == nullptr) == (target_len == nullptr),
                 "b200_joint_dh: logit_len and target_len go together");
  B200_CHECK_ARG((G_lo == nullptr) == (w_lo == nullptr), "b200_joint_dh: G_lo and W_lo go together");
  CUtensorMap wmap, wmap_lo;
  if (int rc = encode_w(&wmap, w, V, J)) return rc;
  wmap_lo = wmap;
  if (w_lo)
    if (int rc = encode_w(&wmap_lo, w_lo, V, J)) return rc;
  const int ctas = dh_ctas(max_ctas);
  dh::Params P{};
  P.nb = nb;
  P.T = T;
  P.U1 = U1;
  P.V = V;
  P.J = J;
  P.kblocks = (V + dh::kBlockK - 1) / dh::kBlockK;
  P.H = reinterpret_cast<const __nv_bfloat16*>(H);
  P.H_lo = reinterpret_cast<const __nv_bfloat16*>(H_lo);
  P.ldh = ldh;
  P.npass = G_lo ? 3 : 1;
  P.logit_len = logit_len;
  P.target_len = target_len;
  P.rowscale = rowscale;
  P.de = de;
  P.dp = dp;
  // label-block size: the power of two that wastes the fewest padded cells (ties: closest to 16,
  // which minimises the 256/UB + UB atomics per thread and tile)
  int best = 8;
  double best_cost = 1e30;
  const int cands[5] = {16, 8, 32, 4, 2};
  for (int i = 0; i < 5; ++i) {
    const int ub = cands[i], tb = 256 / ub;
    const double padded = (double)((U1 + ub - 1) / ub * ub) * (double)((T + tb - 1) / tb * tb);
    const double cost = padded / ((double)U1 * T) * (1.0 + (256.0 / ub + ub) / 1500.0);
    if (cost < best_cost - 1e-9) {
      best_cost = cost;
      best = ub;
    }
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (G_lo) switch (best) {
      case 2: return launch_dh<2, false, true>(wmap, wmap_lo, G, G_lo, ldg, P, ctas, st);
      case 4: return launch_dh<4, false, true>(wmap, wmap_lo, G, G_lo, ldg, P, ctas, st);
      case 8: return launch_dh<8, false, true>(wmap, wmap_lo, G, G_lo, ldg, P, ctas, st);
      case 16: return launch_dh<16, false, true>(wmap, wmap_lo, G, G_lo, ldg, P, ctas, st);
      default: return launch_dh<32, false, true>(wmap, wmap_lo, G, G_lo, ldg, P, ctas, st);
    }
  switch (best) {
    case 2: return launch_dh<2, false, false>(wmap, wmap_lo, G, G_lo, ldg, P, ctas, st);
    case 4: return launch_dh<4, false, false>(wmap, wmap_lo, G, G_lo, ldg, P, ctas, st);
    case 8: return launch_dh<8, false, false>(wmap, wmap_lo, G, G_lo, ldg, P, ctas, st);
    case 16: return launch_dh<16, false, false>(wmap, wmap_lo, G, G_lo, ldg, P, ctas, st);
    default: return launch_dh<32, false, false>(wmap, wmap_lo, G, G_lo, ldg, P, ctas, st);
  }
}

}  // namespace

extern "C" int b200_joint_dh(const void* G, int ldg, const void* w_bf16, const void* H, int ldh, int nb, int T,
                             int U1, int V, int J, const int32_t* logit_len, const int32_t* target_len,
                             float* de, float* dp, int max_ctas, b200_stream_t stream) {
  return joint_dh_impl(G, nullptr, ldg, w_bf16, nullptr, H, nullptr, ldh, nullptr, nb, T, U1, V, J, logit_len,
                       target_len, de, dp, max_ctas, stream);
}

// with a per-cell factor on dH (no-recompute backward: G = Pt', factor = rowscale of b200_joint_rowscale)
extern "C" int b200_joint_dh_scaled(const void* G, int ldg, const void* w_bf16, const void* H, int ldh,
                                    const float* rowscale, int nb, int T, int U1, int V, int J,
                                    const int32_t* logit_len, const int32_t* target_len, float* de, float* dp,
                                    int max_ctas, b200_stream_t stream) {
  B200_CHECK_ARG(rowscale, "b200_joint_dh_scaled: null pointer");
  return joint_dh_impl(G, nullptr, ldg, w_bf16, nullptr, H, nullptr, ldh, rowscale, nb, T, U1, V, J, logit_len,
                       target_len, de, dp, max_ctas, stream);
}

extern "C" int b200_joint_dh_split(const void* G_hi, const void* G_lo, int ldg, const void* w_hi,
                                   const void* w_lo, const void* H_hi, const void* H_lo, int ldh, int nb, int T,
                                   int U1, int V, int J, const int32_t* logit_len, const int32_t* target_len,
                                   float* de, float* dp, int max_ctas, b200_stream_t stream) {
  B200_CHECK_ARG(G_lo && w_lo && H_lo, "b200_joint_dh_split: null pointer");
  return joint_dh_impl(G_hi, G_lo, ldg, w_hi, w_lo, H_hi, H_lo, ldh, nullptr, nb, T, U1, V, J, logit_len,
                       target_len, de, dp, max_ctas, stream);
}

// split operands with a per-cell factor (no-recompute backward in fp32 mode)
extern "C" int b200_joint_dh_split_scaled(const void* G_hi, const void* G_lo, int ldg, const void* w_hi,
                                          const void* w_lo, const void* H_hi, const void* H_lo, int ldh,
                                          const float* rowscale, int nb, int T, int U1, int V, int J,
                                          const int32_t* logit_len, const int32_t* target_len, float* de, float* dp,
                                          int max_ctas, b200_stream_t stream) {
  B200_CHECK_ARG(G_lo && w_lo && H_lo && rowscale, "b200_joint_dh_split_scaled: null pointer");
  return joint_dh_impl(G_hi, G_lo, ldg, w_hi, w_lo, H_hi, H_lo, ldh, rowscale, nb, T, U1, V, J, logit_len,
                       target_len, de, dp, max_ctas, stream);
}

extern "C" int b200_split_gemm_nn(const void* G_hi, const void* G_lo, int ldg, const void* w_hi, const void* w_lo,
                                  long long rows, int V, int J, float* out, int max_ctas, b200_stream_t stream) {
  B200_CHECK_ARG(G_hi && w_hi && out, "b200_split_gemm_nn: null pointer");
  B200_CHECK_ARG(rows > 0 && V > 0, "b200_split_gemm_nn: bad shape rows=%lld V=%d", rows, V);
  B200_CHECK_ARG(J == 256 || J == 512, "b200_split_gemm_nn: output width J=%d must be 256 or 512", J);
  B200_CHECK_ARG(ldg >= V && ldg % 8 == 0 && ((uintptr_t)G_hi & 15) == 0 && ((uintptr_t)G_lo & 15) == 0,
                 "b200_split_gemm_nn: ldg=%d must be >= V and a multiple of 8, G 16-byte aligned", ldg);
  B200_CHECK_ARG(((uintptr_t)w_hi & 15) == 0 && ((uintptr_t)w_lo & 15) == 0,
                 "b200_split_gemm_nn: W must be 16-byte aligned");
  B200_CHECK_ARG((G_lo == nullptr) == (w_lo == nullptr), "b200_split_gemm_nn: G_lo and W_lo go together");
  CUtensorMap wmap, wmap_lo;
  if (int rc = encode_w(&wmap, w_hi, V, J)) return rc;
  wmap_lo = wmap;
  if (w_lo)
    if (int rc = encode_w(&wmap_lo, w_lo, V, J)) return rc;
  dh::Params P{};
  P.T = P.U1 = 1;
  P.V = V;
  P.J = J;
  P.kblocks = (V + dh::kBlockK - 1) / dh::kBlockK;
  P.npass = G_lo ? 3 : 1;
  P.rows = rows;
  P.out = out;
  return launch_dh<16, true, true>(wmap, wmap_lo, G_hi, G_lo, ldg, P, dh_ctas(max_ctas), (cudaStream_t)stream);
}
