// Batched RNN-T greedy search, all bookkeeping on the device.
//
// Replaces the per-utterance Python loop of
//   wenet/transducer/search/greedy_search.py:34-176 (fork; context_filter_state='off')
//   wenet/transducer/search/greedy_search copy.py:6-63 (upstream core algorithm)
// which issues ~10 tiny kernels and one device->host sync per lattice step for ONE
// utterance.  Here every utterance of the batch advances one lattice step per round:
//   state per utterance: t, token, per-frame emission count, LSTM (h,c) per layer,
//   "predictor must run" flag (set after a non-blank emission, as prev_out_nblk).
//   round:  [lstm layer 0..L-1]  (flagged utterances only; predictor.forward_step,
//                                 wenet/transducer/predictor.py:180-200)
//           [pred_z]   pp = pred_ffn(projection(h)), z = tanh(enc_ffn(enc)[t] + pp)   (joint.py:55-68)
//           [logits_argmax]  ffn_out(z) per vocab tile, partial argmax          (joint.py:69, :152-154)
//           [update]   final argmax (lowest index wins ties), token append, t / counters
// All arithmetic is fp32 FMA (no tensor cores, no bf16): the outputs are integers and
// must match the reference's argmax.  enc_ffn(enc_out) is one library GEMM done by the
// caller; projection and pred_ffn are pre-composed on the host (fp64) into W_c, b_c.
#include <cooperative_groups.h>

#include "common.cuh"

namespace b200 {
namespace gs {

constexpr int kMaxLayers = 4;

struct Model {
  const float* embed;                 // [V, Emb]
  const float* w_ih[kMaxLayers];      // [4H, in]
  const float* w_hh[kMaxLayers];      // [4H, H]
  const float* b_ih[kMaxLayers];      // [4H]
  const float* b_hh[kMaxLayers];      // [4H]
  const float* w_c;                   // [J, H]   pred_ffn.W @ projection.W
  const float* b_c;                   // [J]
  const float* w_out;                 // [V, J]
  const float* b_out;                 // [V]
  int V, Emb, H, L, J, blank;
  int has_bias;                       // pp comes from pred_bias_kernel (hot-word predictor bias)
};

struct State {
  const float* enc_proj;  // [B, T, J]
  const int32_t* lens;    // [B]
  int B, T, n_steps, max_tokens, n_tiles;
  int32_t* t;        // [B]
  int32_t* token;    // [B]
  int32_t* cnt;      // [B]
  int32_t* need;     // [B] predictor must run this round
  int32_t* par;      // [B] parity of the current LSTM state
  int32_t* n_tok;    // [B]
  int32_t* tokens;   // [B, max_tokens]
  float* h;          // [2][L][B][H]
  float* c;          // [2][L][B][H]
  float* pp;         // [B][J]
  float* z;          // [B][J]
  float* part_val;   // [B][n_tiles]
  int32_t* part_idx; // [B][n_tiles]
  int32_t* act_list;   // [2][B]
  int32_t* need_list;  // [2][B]
  int32_t* counters;   // [2][2]: {n_act, n_need} per parity; [4] = rounds run (persistent kernel)
  // ---- hot-word gate with "go-back" (greedy_search.py:100-150); null / 2 slots otherwise ----
  int nslot;                    // LSTM state slots: 2, or 3 = {before previous call, current, written by this call}
  const float* enc_proj_empty;  // [B, T, J] encoder stream biased with the EMPTY phrase list (decision 0)
  int32_t* last_t;     // [B] frame of the last decision-0 call
  int32_t* gb_flag;    // [B] going back: decisions forced to 1
  int32_t* gb_end;     // [B] frame at which the go-back started
  int32_t* last_res;   // [B] result[-1]
  int32_t* prev_tok;   // [B] predictor input of the previous call
  int32_t* skip;       // [B] went back this round: the joint / argmax of this round do not count
  int32_t* n_dec;      // [B] decisions kept (len(result))
  int32_t* dec;        // [B, max_dec]
  int max_dec;
};

__device__ __forceinline__ int next_slot(int p, int nslot) { return p + 1 == nslot ? 0 : p + 1; }

__global__ void init_kernel(State S, int blank, int L, int H) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b == 0) {
    S.counters[0] = 0;
    S.counters[1] = 0;
    S.counters[2] = 0;
    S.counters[3] = 0;
    S.counters[6] = 0;  // arrival counter of the persistent kernel's grid barrier
  }
  if (b >= S.B) return;
  S.t[b] = 0;
  S.token[b] = blank;
  S.cnt[b] = 0;
  S.par[b] = 0;
  S.n_tok[b] = 0;
  S.need[b] = 0;
  if (S.skip) {
    S.last_t[b] = 0;
    S.gb_flag[b] = 0;
    S.gb_end[b] = -1;
    S.last_res[b] = 1;
    S.prev_tok[b] = blank;
    S.skip[b] = 0;
    S.n_dec[b] = 0;
  }
}
// second init pass (after the counters are zero): build the round-0 lists
__global__ void init_lists_kernel(State S) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= S.B) return;
  if (S.lens[b] > 0) {
    S.need[b] = 1;
    S.act_list[atomicAdd(&S.counters[0], 1)] = b;
    S.need_list[atomicAdd(&S.counters[1], 1)] = b;
  }
}

__device__ __forceinline__ float sigmoidf_acc(float x) { return 1.f / (1.f + expf(-x)); }

// ---- LSTM layer: one warp per hidden unit, grid.y over groups of 8 flagged utterances ----
// Weight rows are re-read per utterance; the 8 units of a CTA keep their 4*(In+H) rows in L1.
__global__ void __launch_bounds__(256) lstm_layer_kernel(Model M, State S, int layer, int round) {
  const int set = round & 1;
  const int n_need = S.counters[set * 2 + 1];
  const int g0 = blockIdx.y * 8;
  if (g0 >= n_need) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int j = blockIdx.x * 8 + warp;  // hidden unit
  const int H = M.H;
  const int In = (layer == 0) ? M.Emb : H;
  if (j >= H) return;
  float bsum[4];
#pragma unroll
  for (int g = 0; g < 4; ++g) bsum[g] = M.b_ih[layer][g * H + j] + M.b_hh[layer][g * H + j];
  const float* wi = M.w_ih[layer] + (size_t)j * In;   // gate g at + g*H*In
  const float* wh = M.w_hh[layer] + (size_t)j * H;    // gate g at + g*H*H
  const int* list = S.need_list + set * S.B;
  const size_t lay = (size_t)S.B * H;
  for (int q = g0; q < min(n_need, g0 + 8); ++q) {
    const int b = list[q];
    const int p = S.par[b];
    const float* hin = S.h + ((size_t)p * M.L + layer) * lay + (size_t)b * H;
    const float* cin = S.c + ((size_t)p * M.L + layer) * lay + (size_t)b * H;
    const int pn = next_slot(p, S.nslot);
    const float* x = (layer == 0) ? M.embed + (size_t)S.token[b] * In
                                  : S.h + ((size_t)pn * M.L + layer - 1) * lay + (size_t)b * H;
    float a[4] = {0.f, 0.f, 0.f, 0.f};
    for (int i = lane; i < In; i += 32) {
      const float xv = x[i];
#pragma unroll
      for (int g = 0; g < 4; ++g) a[g] = fmaf(__ldg(wi + (size_t)g * H * In + i), xv, a[g]);
    }
    for (int i = lane; i < H; i += 32) {
      const float hv = hin[i];
#pragma unroll
      for (int g = 0; g < 4; ++g) a[g] = fmaf(__ldg(wh + (size_t)g * H * H + i), hv, a[g]);
    }
#pragma unroll
    for (int g = 0; g < 4; ++g) a[g] = warp_sum(a[g]) + bsum[g];
    if (lane == 0) {
      const float ig = sigmoidf_acc(a[0]), fg = sigmoidf_acc(a[1]);
      const float gg = tanhf(a[2]), og = sigmoidf_acc(a[3]);
      const float cn = fg * cin[j] + ig * gg;
      S.c[((size_t)pn * M.L + layer) * lay + (size_t)b * H + j] = cn;
      S.h[((size_t)pn * M.L + layer) * lay + (size_t)b * H + j] = og * tanhf(cn);
    }
  }
}

// ---- pp = W_c h_top + b_c (flagged only), z = tanh(enc_proj[b,t] + pp) (all active) ----
__global__ void __launch_bounds__(256) pred_z_kernel(Model M, State S, int round) {
  const int set = round & 1;
  const int n_act = S.counters[set * 2 + 0];
  const int g0 = blockIdx.y * 8;
  if (g0 >= n_act) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int j = blockIdx.x * 8 + warp;
  if (j >= M.J) return;
  const int H = M.H;
  const float* w = M.w_c + (size_t)j * H;
  const float bc = M.b_c[j];
  const int* list = S.act_list + set * S.B;
  const size_t lay = (size_t)S.B * H;
  for (int q = g0; q < min(n_act, g0 + 8); ++q) {
    const int b = list[q];
    float pp;
    if (S.need[b] && !M.has_bias) {
      const int p = next_slot(S.par[b], S.nslot);  // state written by the LSTM kernels of this round
      const float* h = S.h + ((size_t)p * M.L + M.L - 1) * lay + (size_t)b * H;
      float a = 0.f;
      for (int i = lane; i < H; i += 32) a = fmaf(__ldg(w + i), h[i], a);
      pp = warp_sum(a) + bc;
      if (lane == 0) S.pp[(size_t)b * M.J + j] = pp;
    } else {
      pp = S.pp[(size_t)b * M.J + j];
    }
    if (lane == 0) {
      // decision 0 pairs the predictor stream with the EMPTY-biased encoder stream (greedy_search.py:146-150)
      const float* ep = (S.enc_proj_empty && S.last_res[b] == 0) ? S.enc_proj_empty : S.enc_proj;
      const float e = ep[((size_t)b * S.T + S.t[b]) * M.J + j];
      S.z[(size_t)b * M.J + j] = tanhf(e + pp);
    }
  }
}

// ---- logits for a tile of kVT vocabulary rows x all active utterances, partial argmax ----
// 34 rows: V = 5000 is 148 tiles, one per SM (32 rows gave 157 CTAs: nine SMs ran two)
constexpr int kVT = 34;   // vocabulary rows per CTA (even)
constexpr int kLaThreads = 8 * (kVT / 2);  // 8 utterance pairs x kVT/2 row slots
constexpr int kUG = 16;   // utterances per shared-memory group
constexpr int kPad = 4;   // floats of row padding (bank spread)

__global__ void __launch_bounds__(kLaThreads) logits_argmax_kernel(Model M, State S, int round) {
  extern __shared__ float sm[];
  const int set = round & 1;
  const int n_act = S.counters[set * 2 + 0];
  if (n_act == 0) return;
  const int J = M.J, ld = J + kPad;
  float* sw = sm;                 // [kVT][ld]
  float* sz = sm + kVT * ld;      // [kUG][ld]
  float* sl = sz + kUG * ld;      // [kUG][kVT + 1] logits of the current group
  const int v0 = blockIdx.x * kVT;
  const int tid = threadIdx.x;
  for (int i = tid; i < kVT * (J / 4); i += blockDim.x) {
    const int r = i / (J / 4), k4 = i - r * (J / 4);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (v0 + r < M.V) v = *reinterpret_cast<const float4*>(M.w_out + (size_t)(v0 + r) * J + k4 * 4);
    *reinterpret_cast<float4*>(sw + r * ld + k4 * 4) = v;
  }
  const int u2 = tid & 7;    // utterance pair within the group
  const int rs = tid >> 3;   // rows rs and rs + kVT / 2
  const float b0 = (v0 + rs < M.V) ? M.b_out[v0 + rs] : 0.f;
  const float b1 = (v0 + rs + kVT / 2 < M.V) ? M.b_out[v0 + rs + kVT / 2] : 0.f;
  const int* list = S.act_list + set * S.B;
  for (int g0 = 0; g0 < n_act; g0 += kUG) {
    __syncthreads();
    for (int i = tid; i < kUG * (J / 4); i += blockDim.x) {
      const int uu = i / (J / 4), k4 = i - uu * (J / 4);
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (g0 + uu < n_act)
        v = *reinterpret_cast<const float4*>(S.z + (size_t)list[g0 + uu] * J + k4 * 4);
      *reinterpret_cast<float4*>(sz + uu * ld + k4 * 4) = v;
    }
    __syncthreads();
    float a00 = 0.f, a01 = 0.f, a10 = 0.f, a11 = 0.f;  // [row][utt]
    const float* w0 = sw + rs * ld;
    const float* w1 = sw + (rs + kVT / 2) * ld;
    const float* z0 = sz + (2 * u2) * ld;
    const float* z1 = sz + (2 * u2 + 1) * ld;
#pragma unroll 4
    for (int k = 0; k < J; k += 4) {
      const float4 wa = *reinterpret_cast<const float4*>(w0 + k);
      const float4 wb = *reinterpret_cast<const float4*>(w1 + k);
      const float4 za = *reinterpret_cast<const float4*>(z0 + k);
      const float4 zb = *reinterpret_cast<const float4*>(z1 + k);
      a00 = fmaf(wa.x, za.x, a00); a00 = fmaf(wa.y, za.y, a00); a00 = fmaf(wa.z, za.z, a00); a00 = fmaf(wa.w, za.w, a00);
      a01 = fmaf(wa.x, zb.x, a01); a01 = fmaf(wa.y, zb.y, a01); a01 = fmaf(wa.z, zb.z, a01); a01 = fmaf(wa.w, zb.w, a01);
      a10 = fmaf(wb.x, za.x, a10); a10 = fmaf(wb.y, za.y, a10); a10 = fmaf(wb.z, za.z, a10); a10 = fmaf(wb.w, za.w, a10);
      a11 = fmaf(wb.x, zb.x, a11); a11 = fmaf(wb.y, zb.y, a11); a11 = fmaf(wb.z, zb.z, a11); a11 = fmaf(wb.w, zb.w, a11);
    }
    sl[(2 * u2) * (kVT + 1) + rs] = a00 + b0;
    sl[(2 * u2 + 1) * (kVT + 1) + rs] = a01 + b0;
    sl[(2 * u2) * (kVT + 1) + rs + kVT / 2] = a10 + b1;
    sl[(2 * u2 + 1) * (kVT + 1) + rs + kVT / 2] = a11 + b1;
    __syncthreads();
    if (tid < kUG && g0 + tid < n_act) {
      float best = -INFINITY;
      int bi = 0x7fffffff;
      for (int r = 0; r < kVT; ++r) {
        const float v = sl[tid * (kVT + 1) + r];
        if (v0 + r < M.V && v > best) {  // strict '>' keeps the lowest index on ties
          best = v;
          bi = v0 + r;
        }
      }
      const int b = list[g0 + tid];
      S.part_val[(size_t)b * S.n_tiles + blockIdx.x] = best;
      S.part_idx[(size_t)b * S.n_tiles + blockIdx.x] = bi;
    }
  }
}

// ---- final argmax + the greedy state machine; builds the next round's lists ----
__global__ void __launch_bounds__(1024) update_kernel(Model M, State S, int round) {
  const int set = round & 1, nxt = set ^ 1;
  const int n_act = S.counters[set * 2 + 0];
  if (threadIdx.x == 0) {
    S.counters[nxt * 2 + 0] = 0;
    S.counters[nxt * 2 + 1] = 0;
  }
  __syncthreads();
  if (n_act == 0) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  const int* list = S.act_list + set * S.B;
  for (int q = warp; q < n_act; q += nw) {
    const int b = list[q];
    float best = -INFINITY;
    int bi = 0x7fffffff;
    for (int i = lane; i < S.n_tiles; i += 32) {
      const float v = S.part_val[(size_t)b * S.n_tiles + i];
      const int ix = S.part_idx[(size_t)b * S.n_tiles + i];
      if (v > best || (v == best && ix < bi)) {
        best = v;
        bi = ix;
      }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float ov = __shfl_xor_sync(0xffffffffu, best, o);
      const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
      if (ov > best || (ov == best && oi < bi)) {
        best = ov;
        bi = oi;
      }
    }
    if (lane == 0 && S.skip && S.skip[b]) {
      // went back this round (pred_gate_kernel rewound t / token / LSTM slot): re-run the previous call
      S.skip[b] = 0;
      S.act_list[nxt * S.B + atomicAdd(&S.counters[nxt * 2 + 0], 1)] = b;
      S.need_list[nxt * S.B + atomicAdd(&S.counters[nxt * 2 + 1], 1)] = b;
    } else if (lane == 0) {
      const int k = bi;
      int t = S.t[b], cnt = S.cnt[b];
      if (S.need[b]) S.par[b] = next_slot(S.par[b], S.nslot);  // the LSTM state written this round becomes current
      int need = 0;
      if (k != M.blank) {
        const int n = S.n_tok[b];
        if (n < S.max_tokens) S.tokens[(size_t)b * S.max_tokens + n] = k;
        S.n_tok[b] = n + 1;
        S.token[b] = k;
        cnt += 1;
        need = 1;
      }
      if (k == M.blank || cnt >= S.n_steps) {
        t += 1;
        cnt = 0;
      }
      S.t[b] = t;
      S.cnt[b] = cnt;
      S.need[b] = need;
      if (t < S.lens[b]) {
        S.act_list[nxt * S.B + atomicAdd(&S.counters[nxt * 2 + 0], 1)] = b;
        if (need) S.need_list[nxt * S.B + atomicAdd(&S.counters[nxt * 2 + 1], 1)] = b;
      }
    }
  }
}

#include "greedy_bias.inc"
#include "greedy_gate_cluster.inc"
#include "greedy_persistent.inc"
#include "greedy_replicated.inc"

}  // namespace gs
}  // namespace b200

#include "greedy_host.inc"
