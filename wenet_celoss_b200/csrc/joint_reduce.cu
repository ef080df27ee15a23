// Backward tail of the fused joint: dZ = dH * (1 - H^2) (tanh'), reduced over u into
// d e[b,t,:] and over t into d p[b,u,:].  Mirrors autograd through
//   out = enc_out.unsqueeze(2) + pred_out.unsqueeze(1); out = tanh(out)
// (wenet/transducer/joint.py:58-68).  HBM-bound: reads dH (fp32) and H (bf16) once.
//
// grid (k-slices of 64, t-blocks, utterances of the chunk); 256 threads = 64 k x 4 u-phases.
// d e rows are owned by one CTA (plain stores); d p is accumulated per CTA in shared memory
// and added to global memory with one atomic per (u,k) and t-block.
#include "common.cuh"

namespace b200 {

constexpr int kDzK = 64;
constexpr int kDzPh = 4;

__global__ void __launch_bounds__(kDzK* kDzPh)
joint_dz_reduce_kernel(const float* __restrict__ dH, const __nv_bfloat16* __restrict__ H, int ldh,
                       int T, int U1, int J, int t_per_block, float* __restrict__ de,
                       float* __restrict__ dp) {
  extern __shared__ float sm[];  // dp_s[U1][64], then e_s[4][64]
  float* dp_s = sm;
  float* e_s = sm + (size_t)U1 * kDzK;
  const int kk = threadIdx.x & (kDzK - 1);
  const int up = threadIdx.x / kDzK;
  const int k = blockIdx.x * kDzK + kk;
  const int b = blockIdx.z;
  const int t0 = blockIdx.y * t_per_block;
  const int t1 = min(T, t0 + t_per_block);
  for (int i = threadIdx.x; i < U1 * kDzK; i += blockDim.x) dp_s[i] = 0.f;
  __syncthreads();
  for (int t = t0; t < t1; ++t) {
    const long long row0 = ((long long)b * T + t) * U1;
    float acc = 0.f;
#pragma unroll 4
    for (int u = up; u < U1; u += kDzPh) {
      const long long row = row0 + u;
      const float x = dH[row * J + k];
      const float h = __bfloat162float(H[row * ldh + k]);
      const float z = x * (1.f - h * h);
      acc += z;
      dp_s[u * kDzK + kk] += z;  // (u,kk) is owned by exactly one thread
    }
    e_s[up * kDzK + kk] = acc;
    __syncthreads();
    if (up == 0)
      de[((long long)b * T + t) * J + k] = e_s[kk] + e_s[kDzK + kk] + e_s[2 * kDzK + kk] + e_s[3 * kDzK + kk];
    __syncthreads();
  }
  for (int i = threadIdx.x; i < U1 * kDzK; i += blockDim.x) {
    const int u = i / kDzK, c = i - u * kDzK;
    atomicAdd(&dp[((long long)b * U1 + u) * J + blockIdx.x * kDzK + c], dp_s[i]);
  }
}

// bound[0] = max_v (sum_j |W[v,j]| + |b[v]|) >= |logit| of every lattice cell (|tanh| <= 1):
// lets the forward joint kernel drop the running maximum of its log-sum-exp.  One warp per
// vocabulary row; non-negative floats order like their bit patterns.
__global__ void joint_logit_bound_kernel(const __nv_bfloat16* __restrict__ w,
                                         const float* __restrict__ bias, int V, int J,
                                         float* __restrict__ bound) {
  const int v = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (v >= V) return;
  const int lane = threadIdx.x & 31;
  const __nv_bfloat162* row = reinterpret_cast<const __nv_bfloat162*>(w + (long long)v * J);
  float s = 0.f;
  for (int j = lane; j < J / 2; j += 32) {
    const float2 f = __bfloat1622float2(row[j]);
    s += fabsf(f.x) + fabsf(f.y);
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) {
    s = (s + fabsf(bias[v])) * 1.001f;  // headroom for the summation order
    if (!(s == s)) s = INFINITY;         // NaN weights: no bound
    atomicMax(reinterpret_cast<int*>(bound), __float_as_int(s));
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_joint_logit_bound(const void* w_bf16, const float* bias, int V, int J,
                                      float* bound, b200_stream_t stream) {
  B200_CHECK_ARG(w_bf16 && bias && bound, "b200_joint_logit_bound: null pointer");
  B200_CHECK_ARG(V > 0 && J > 0 && J % 2 == 0, "b200_joint_logit_bound: bad shape V=%d J=%d", V, J);
  B200_CHECK_CUDA(cudaMemsetAsync(bound, 0, sizeof(float), (cudaStream_t)stream));
  joint_logit_bound_kernel<<<(V + 7) / 8, 256, 0, (cudaStream_t)stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(w_bf16), bias, V, J, bound);
  B200_CHECK_LAUNCH("joint_logit_bound_kernel");
  return B200_OK;
}

extern "C" int b200_joint_dz_reduce(const float* dH, const void* H, int ldh, int nb, int T, int U1,
                                    int J, float* de, float* dp, b200_stream_t stream) {
  B200_CHECK_ARG(dH && H && de && dp, "b200_joint_dz_reduce: null pointer");
  B200_CHECK_ARG(nb > 0 && T > 0 && U1 > 0 && J > 0 && J % kDzK == 0 && ldh >= J,
                 "b200_joint_dz_reduce: bad shape nb=%d T=%d U1=%d J=%d ldh=%d", nb, T, U1, J, ldh);
  const size_t smem = ((size_t)U1 * kDzK + kDzPh * kDzK) * sizeof(float);
  B200_CHECK_ARG(smem <= 200 * 1024, "b200_joint_dz_reduce: U+1 = %d too large for shared memory", U1);
  static bool attr_set = false;
  if (!attr_set) {
    B200_CHECK_CUDA(cudaFuncSetAttribute(joint_dz_reduce_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_set = true;
  }
  const int tpb = 25;
  dim3 grid(J / kDzK, (T + tpb - 1) / tpb, nb);
  joint_dz_reduce_kernel<<<grid, kDzK * kDzPh, smem, (cudaStream_t)stream>>>(
      dH, reinterpret_cast<const __nv_bfloat16*>(H), ldh, T, U1, J, tpb, de, dp);
  B200_CHECK_LAUNCH("joint_dz_reduce_kernel");
  return B200_OK;
}
