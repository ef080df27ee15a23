// No-recompute backward of the fused joint (autograd of wenet/transducer/joint.py:58-69 under
// torchaudio.functional.rnnt_loss, reached from wenet/transducer/transducer.py:132-147).
//
// The training forward (joint_kernel<MODE_FWDP>, joint_fused.cu) left Pt[c, v] = exp(logit[c, v]) in bf16.
// With the coefficients of b200_rnnt_grad_coef, d loss / d logits of cell c is
//   G[c, v] = scale * (exp(x_v + c0) - [v == blank] exp(x_v + cb) - [v == y] exp(x_v + ce))
//           = rowscale[c] * Pt'[c, v],     rowscale = scale * exp(c0),
// where Pt' equals Pt except for the blank and target entries of the row,
//   Pt'[c, blank] = exp(x_b) - exp(x_b + cb - c0)      Pt'[c, y] = exp(x_y) - exp(x_y + ce - c0),
// which this kernel recomputes from the fp32 statistics of the forward pass (x_b = lp_blank + lse,
// x_y = lp_emit + lse) and stores over the two bf16 entries: one rounding, as in the recompute path.
// The kernel also writes rowscale[c] and Hs[c, :] = bf16(rowscale[c] * H[c, :]), so that
//   dW = Pt'^T . Hs   (joint_dw.cu)       d b = Pt'^T . rowscale   (column sums inside joint_dw.cu)
//   dH = rowscale * (Pt' . W)             (joint_dh.cu, factor applied in its epilogue)
// need no second logits GEMM.  HBM-bound: reads H, writes Hs (2 * cells * J * 2 bytes).
#include "common.cuh"

namespace b200 {
namespace nr {

// Hs = bf16(rowscale * H), streaming.  Block = (256 / (J/8)) cell rows x J/8 threads of 8 elements,
// 32-bit index arithmetic (cells < 2^31).
__global__ void __launch_bounds__(256)
joint_rowscale_kernel(const float4* __restrict__ coef, int J, long long cells,
                      const __nv_bfloat16* __restrict__ H, long long ldh, __nv_bfloat16* __restrict__ Hs) {
  const int g8 = J >> 3;  // 16-byte groups per row
  const int rows = blockDim.x / g8;
  const int r = threadIdx.x / g8;
  const int k0 = (threadIdx.x - r * g8) << 3;
  if (r >= rows) return;
  const unsigned n = (unsigned)cells;
  const unsigned stride = (unsigned)gridDim.x * rows;
  auto scale_row = [&](unsigned cell, float rs, uint4 h) {
    uint4 o;
    const uint32_t* hi = &h.x;
    uint32_t* oo = &o.x;
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const float lo = __uint_as_float(hi[k] << 16) * rs;
      const float up = __uint_as_float(hi[k] & 0xffff0000u) * rs;
      const __nv_bfloat162 v = __floats2bfloat162_rn(lo, up);
      oo[k] = *reinterpret_cast<const uint32_t*>(&v);
    }
    *reinterpret_cast<uint4*>(Hs + (size_t)cell * ldh + k0) = o;
  };
  unsigned cell = blockIdx.x * rows + r;
  for (; cell + stride < n; cell += 2 * stride) {  // two rows per pass: four 16-byte loads in flight
    const float4 c0 = __ldg(coef + cell), c1 = __ldg(coef + cell + stride);
    const uint4 h0 = __ldg(reinterpret_cast<const uint4*>(H + (size_t)cell * ldh + k0));
    const uint4 h1 = __ldg(reinterpret_cast<const uint4*>(H + (size_t)(cell + stride) * ldh + k0));
    const float r0 = c0.x == -INFINITY ? 0.f : c0.w * __expf(c0.x);
    const float r1 = c1.x == -INFINITY ? 0.f : c1.w * __expf(c1.x);
    scale_row(cell, r0, h0);
    scale_row(cell + stride, r1, h1);
  }
  if (cell < n) {
    const float4 c0 = __ldg(coef + cell);
    const uint4 h0 = __ldg(reinterpret_cast<const uint4*>(H + (size_t)cell * ldh + k0));
    const float r0 = c0.x == -INFINITY ? 0.f : c0.w * __expf(c0.x);
    scale_row(cell, r0, h0);
  }
}

// one thread per cell: rowscale[c] and the two patched entries of the row of Pt (kept apart from the
// streaming kernel above: one diverged lane per row cost it a fifth of its bandwidth)
__global__ void __launch_bounds__(256)
joint_patch_kernel(const float4* __restrict__ coef, const float* __restrict__ lse,
                   const float* __restrict__ lp_blank, const float* __restrict__ lp_emit,
                   const int32_t* __restrict__ targets, int T, int U1, int blank, long long cells,
                   __nv_bfloat16* __restrict__ Pt, long long ldg, float* __restrict__ rowscale) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= cells) return;
  const unsigned cell = (unsigned)i;
  const float4 c = __ldg(coef + cell);
  const float rs = c.x == -INFINITY ? 0.f : c.w * __expf(c.x);
  rowscale[cell] = rs;
  if (rs == 0.f) return;
  const float ls = lse[cell];
  const bool has_b = c.y != -INFINITY, has_e = c.z != -INFINITY;
  int y = -1;
  if (has_e) {
    const unsigned bt = cell / (unsigned)U1;
    const unsigned u = cell - bt * (unsigned)U1;
    const unsigned b = bt / (unsigned)T;
    y = targets[(size_t)b * (U1 - 1) + u];
  }
  // exp(x + (cb - c0)) <= exp(lse): lp + beta_next <= beta (the lattice recursion)
  if (has_b) {
    const float xb = lp_blank[cell] + ls;
    float v = __expf(xb) - __expf(xb + (c.y - c.x));
    if (y == blank) v -= __expf(xb + (c.z - c.x));
    Pt[(size_t)cell * ldg + blank] = __float2bfloat16_rn(v);
  }
  if (has_e && !(y == blank && has_b)) {
    const float xe = lp_emit[cell] + ls;
    Pt[(size_t)cell * ldg + y] = __float2bfloat16_rn(__expf(xe) - __expf(xe + (c.z - c.x)));
  }
}

}  // namespace nr
}  // namespace b200

using namespace b200;

extern "C" int b200_joint_rowscale(const float* coef, const float* lse, const float* lp_blank,
                                   const float* lp_emit, const int32_t* targets, int B, int T, int U1, int V,
                                   int J, int blank, void* Pt, int ldg, const void* H, int ldh, void* Hs,
                                   float* rowscale, b200_stream_t stream) {
  B200_CHECK_ARG(coef && lse && lp_blank && lp_emit && Pt && H && Hs && rowscale && (U1 == 1 || targets),
                 "b200_joint_rowscale: null pointer");
  B200_CHECK_ARG(B > 0 && T > 0 && U1 > 0 && V > 0, "b200_joint_rowscale: bad shape B=%d T=%d U1=%d V=%d", B,
                 T, U1, V);
  B200_CHECK_ARG(J > 0 && J % 8 == 0 && ldh >= J && ldh % 8 == 0 && ldg >= V,
                 "b200_joint_rowscale: J=%d must be a multiple of 8, ldh=%d >= J and a multiple of 8, ldg=%d >= V",
                 J, ldh, ldg);
  B200_CHECK_ARG(blank >= 0 && blank < V, "b200_joint_rowscale: blank %d out of range", blank);
  B200_CHECK_ARG((((uintptr_t)coef | (uintptr_t)H | (uintptr_t)Hs) & 15) == 0,
                 "b200_joint_rowscale: coef, H, Hs must be 16-byte aligned");
  const long long cells = (long long)B * T * U1;
  B200_CHECK_ARG(cells < (1ll << 31) && J <= 2048, "b200_joint_rowscale: more than 2^31 cells or J > 2048");
  int sms = 0, dev = 0;
  cudaGetDevice(&dev);
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  if (sms <= 0) sms = 148;
  const int g8 = J / 8, rows = 256 / g8;
  long long blocks = (cells + rows - 1) / rows;
  if (blocks > 16ll * sms) blocks = 16ll * sms;
  nr::joint_patch_kernel<<<(unsigned)((cells + 255) / 256), 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const float4*>(coef), lse, lp_blank, lp_emit, targets, T, U1, blank, cells,
      reinterpret_cast<__nv_bfloat16*>(Pt), ldg, rowscale);
  B200_CHECK_LAUNCH("joint_patch_kernel");
  nr::joint_rowscale_kernel<<<(int)blocks, rows * g8, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const float4*>(coef), J, cells, reinterpret_cast<const __nv_bfloat16*>(H), ldh,
      reinterpret_cast<__nv_bfloat16*>(Hs));
  B200_CHECK_LAUNCH("joint_rowscale_kernel");
  return B200_OK;
}
