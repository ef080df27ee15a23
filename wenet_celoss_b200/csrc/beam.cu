// Batched RNN-T prefix beam search, hypothesis bookkeeping on the device.
//
// Replaces PrefixBeamSearch.prefix_beam_search
//   (wenet/transducer/search/prefix_beam_search.py:42-148; one utterance, Python lists,
//    N*N .item() syncs per frame) for a whole batch.  Per frame i and utterance b with
//    N live hypotheses ("rows", row = b*beam + j):
//   lstm / pred_z   predictor.forward_step on every row from the row's cached state
//                   (prefix_beam_search.py:31-40, predictor.py:180-200) and the joint's
//                   hidden z = tanh(enc_ffn(enc)[i] + pred_ffn(pred))       (joint.py:55-68)
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
  float* new_h;        // [L][B*beam][H]     state AFTER consuming it (this frame)
  float* new_c;
  float* z;            // [B*beam][J]
  float* logits;       // [B*beam][V]
  float* topv;         // [B*beam][beam]
  int32_t* topi;       // [B*beam][beam]
  int32_t* row_list;   // [2][B*beam]
  int32_t* n_rows;     // [2]
};

__device__ __forceinline__ unsigned long long hash_push(unsigned long long h, int tok) {
  h ^= (unsigned long long)(unsigned)tok + 0x9e3779b97f4a7c15ull + (h << 6) + (h >> 2);
  return h * 0xff51afd7ed558ccdull;
}
__device__ __forceinline__ float sigm(float x) { return 1.f / (1.f + expf(-x)); }

__global__ void init_kernel(State S, int blank) {
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b == 0) {
    S.n_rows[0] = 0;
    S.n_rows[1] = 0;
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
  if (S.lens[b] > 0) S.row_list[atomicAdd(&S.n_rows[0], 1)] = b * S.beam;
}

// ---- LSTM layer over the live rows (one warp per hidden unit x 8 rows) ----
__global__ void __launch_bounds__(256) lstm_kernel(Model M, State S, int layer, int frame) {
  const int set = frame & 1;
  // the other parity's row counter was last read one frame ago: reset it for this frame's update
  if (layer == 0 && blockIdx.x == 0 && blockIdx.y == 0 && threadIdx.x == 0) S.n_rows[set ^ 1] = 0;
  const int n = S.n_rows[set];
  const int g0 = blockIdx.y * 8;
  if (g0 >= n) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int j = blockIdx.x * 8 + warp;
  const int H = M.H, In = layer == 0 ? M.Emb : H;
  if (j >= H) return;
  float bsum[4];
#pragma unroll
  for (int g = 0; g < 4; ++g) bsum[g] = M.b_ih[layer][g * H + j] + M.b_hh[layer][g * H + j];
  const float* wi = M.w_ih[layer] + (size_t)j * In;
  const float* wh = M.w_hh[layer] + (size_t)j * H;
  const int* list = S.row_list + set * S.n_rows_max;
  const size_t lay = (size_t)S.n_rows_max * H;
  for (int q = g0; q < min(n, g0 + 8); ++q) {
    const int row = list[q];
    const float* hin = S.cache_h + ((size_t)set * M.L + layer) * lay + (size_t)row * H;
    const float* cin = S.cache_c + ((size_t)set * M.L + layer) * lay + (size_t)row * H;
    const int len = S.hyp_len[set * S.n_rows_max + row];
    const int tok = S.hyp[((size_t)set * S.n_rows_max + row) * S.max_len + len - 1];
    const float* x = layer == 0 ? M.embed + (size_t)tok * In
                                : S.new_h + (size_t)(layer - 1) * lay + (size_t)row * H;
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
      const float ig = sigm(a[0]), fg = sigm(a[1]), gg = tanhf(a[2]), og = sigm(a[3]);
      const float cn = fg * cin[j] + ig * gg;
      S.new_c[(size_t)layer * lay + (size_t)row * H + j] = cn;
      S.new_h[(size_t)layer * lay + (size_t)row * H + j] = og * tanhf(cn);
    }
  }
}

__global__ void __launch_bounds__(256) pred_z_kernel(Model M, State S, int frame) {
  const int set = frame & 1;
  const int n = S.n_rows[set];
  const int g0 = blockIdx.y * 8;
  if (g0 >= n) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int j = blockIdx.x * 8 + warp;
  if (j >= M.J) return;
  const int H = M.H;
  const float* w = M.w_c + (size_t)j * H;
  const float bc = M.b_c[j];
  const int* list = S.row_list + set * S.n_rows_max;
  const size_t lay = (size_t)S.n_rows_max * H;
  for (int q = g0; q < min(n, g0 + 8); ++q) {
    const int row = list[q];
    const float* h = S.new_h + (size_t)(M.L - 1) * lay + (size_t)row * H;
    float a = 0.f;
    for (int i = lane; i < H; i += 32) a = fmaf(__ldg(w + i), h[i], a);
    a = warp_sum(a) + bc;
    if (lane == 0) {
      const int b = row / S.beam;
      const float e = S.enc_proj[((size_t)b * S.T + frame) * M.J + j];
      S.z[(size_t)row * M.J + j] = tanhf(e + a);
    }
  }
}

// ---- logits[row, v] for a tile of 32 vocabulary rows x all live rows ----
constexpr int kVT = 32, kUG = 16, kPad = 4;

__global__ void __launch_bounds__(128) logits_kernel(Model M, State S, int frame) {
  extern __shared__ float sm[];
  const int set = frame & 1;
  const int n = S.n_rows[set];
  if (n == 0) return;
  const int J = M.J, ld = J + kPad;
  float* sw = sm;
  float* sz = sm + kVT * ld;
  const int v0 = blockIdx.x * kVT;
  const int tid = threadIdx.x;
  for (int i = tid; i < kVT * (J / 4); i += blockDim.x) {
    const int r = i / (J / 4), k4 = i - r * (J / 4);
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (v0 + r < M.V) v = *reinterpret_cast<const float4*>(M.w_out + (size_t)(v0 + r) * J + k4 * 4);
    *reinterpret_cast<float4*>(sw + r * ld + k4 * 4) = v;
  }
  const int u2 = tid & 7, rs = tid >> 3;
  const float b0 = (v0 + rs < M.V) ? M.b_out[v0 + rs] : 0.f;
  const float b1 = (v0 + rs + 16 < M.V) ? M.b_out[v0 + rs + 16] : 0.f;
  const int* list = S.row_list + set * S.n_rows_max;
  for (int g0 = 0; g0 < n; g0 += kUG) {
    __syncthreads();
    for (int i = tid; i < kUG * (J / 4); i += blockDim.x) {
      const int uu = i / (J / 4), k4 = i - uu * (J / 4);
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (g0 + uu < n) v = *reinterpret_cast<const float4*>(S.z + (size_t)list[g0 + uu] * J + k4 * 4);
      *reinterpret_cast<float4*>(sz + uu * ld + k4 * 4) = v;
    }
    __syncthreads();
    float a00 = 0.f, a01 = 0.f, a10 = 0.f, a11 = 0.f;
    const float* w0 = sw + rs * ld;
    const float* w1 = sw + (rs + 16) * ld;
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
    const int ra = g0 + 2 * u2, rb = ra + 1;
    if (ra < n) {
      float* o = S.logits + (size_t)list[ra] * M.V + v0;
      if (v0 + rs < M.V) o[rs] = a00 + b0;
      if (v0 + rs + 16 < M.V) o[rs + 16] = a10 + b1;
    }
    if (rb < n) {
      float* o = S.logits + (size_t)list[rb] * M.V + v0;
      if (v0 + rs < M.V) o[rs] = a01 + b0;
      if (v0 + rs + 16 < M.V) o[rs + 16] = a11 + b1;
    }
  }
}

#include "beam_select.inc"

}  // namespace bs
}  // namespace b200

#include "beam_host.inc"
