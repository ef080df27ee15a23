// Backward tail of the fused joint: dZ = dH * (1 - H^2) (tanh'), reduced over u into
// d e[b,t,:] and over t into d p[b,u,:].  Mirrors autograd through
//   out = enc_out.unsqueeze(2) + pred_out.unsqueeze(1); out = tanh(out)
// (wenet/transducer/joint.py:58-68).  HBM-bound: reads dH (fp32) and H (bf16) once.
//
// grid (k-slices of 128, t-blocks, utterances of the chunk); 256 threads = 32 lanes x float4
// (128 k) x 8 u-phases: 16-byte dH loads and 8-byte H loads, several in flight per thread.
// d e rows are owned by one CTA (plain stores); d p is accumulated per CTA in shared memory
// and added to global memory with one vector atomic per (u, 4 k) and t-block.
#include "common.cuh"

namespace b200 {

constexpr int kDzK = 128;
constexpr int kDzPh = 8;

__global__ void __launch_bounds__(32 * kDzPh)
joint_dz_reduce_kernel(const float* __restrict__ dH, const __nv_bfloat16* __restrict__ H, int ldh,
                       int T, int U1, int J, int t_per_block, float* __restrict__ de,
                       float* __restrict__ dp) {
  extern __shared__ float4 sm4[];  // dp_s[U1][32], then e_s[8][32]
  float4* dp_s = sm4;
  float4* e_s = sm4 + (size_t)U1 * 32;
  const int lane = threadIdx.x & 31;
  const int up = threadIdx.x >> 5;
  const int k = blockIdx.x * kDzK + lane * 4;
  const bool on = k < J;  // J % 128 == 64: the last slice is half empty
  const int b = blockIdx.z;
  const int t0 = blockIdx.y * t_per_block;
  const int t1 = min(T, t0 + t_per_block);
  for (int i = threadIdx.x; i < U1 * 32; i += blockDim.x) dp_s[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  __syncthreads();
  for (int t = t0; t < t1; ++t) {
    const long long row0 = ((long long)b * T + t) * U1;
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
    for (int u = up; u < U1 && on; u += kDzPh) {
      const long long row = row0 + u;
      const float4 x = __ldcs(reinterpret_cast<const float4*>(dH + row * J + k));
      const uint2 hb = __ldcs(reinterpret_cast<const uint2*>(H + row * ldh + k));
      const float2 h01 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&hb.x));
      const float2 h23 = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&hb.y));
      float4 z;
      z.x = x.x * (1.f - h01.x * h01.x);
      z.y = x.y * (1.f - h01.y * h01.y);
      z.z = x.z * (1.f - h23.x * h23.x);
      z.w = x.w * (1.f - h23.y * h23.y);
      acc.x += z.x; acc.y += z.y; acc.z += z.z; acc.w += z.w;
      float4 d = dp_s[u * 32 + lane];  // (u, lane) is owned by exactly one thread
      d.x += z.x; d.y += z.y; d.z += z.z; d.w += z.w;
      dp_s[u * 32 + lane] = d;
    }
    e_s[up * 32 + lane] = acc;
    __syncthreads();
    if (up == 0 && on) {
      float4 s = e_s[lane];
#pragma unroll
      for (int j = 1; j < kDzPh; ++j) {
        const float4 v = e_s[j * 32 + lane];
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      }
      *reinterpret_cast<float4*>(de + ((long long)b * T + t) * J + k) = s;
    }
    __syncthreads();
  }
  for (int i = threadIdx.x; i < U1 * 32; i += blockDim.x) {
    const int u = i >> 5, c = i & 31;
    if (blockIdx.x * kDzK + c * 4 < J)
      atomicAdd(reinterpret_cast<float4*>(dp + ((long long)b * U1 + u) * J + blockIdx.x * kDzK + c * 4),
              dp_s[i]);
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
  B200_CHECK_ARG(nb > 0 && T > 0 && U1 > 0 && J > 0 && J % 4 == 0 && ldh >= J && ldh % 4 == 0,
                 "b200_joint_dz_reduce: bad shape nb=%d T=%d U1=%d J=%d ldh=%d (J %% 4 == 0)", nb, T,
                 U1, J, ldh);
  B200_CHECK_ARG(((uintptr_t)dH & 15) == 0 && ((uintptr_t)H & 7) == 0 && ((uintptr_t)de & 15) == 0 &&
                     ((uintptr_t)dp & 15) == 0,
                 "b200_joint_dz_reduce: dH, de, dp must be 16-byte aligned, H 8-byte aligned");
  const size_t smem = ((size_t)U1 * 32 + kDzPh * 32) * sizeof(float4);
  B200_CHECK_ARG(smem <= 200 * 1024, "b200_joint_dz_reduce: U+1 = %d too large for shared memory", U1);
  // per device, and cheap: set on every call (a process may drive several GPUs)
  B200_CHECK_CUDA(cudaFuncSetAttribute(joint_dz_reduce_kernel,
                                       cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
  // frames per CTA: about 8 CTAs per SM, but at least 4 frames (bounds the d p atomics)
  int sms = 148;
  {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
  }
  const int kslices = (J + kDzK - 1) / kDzK;
  long long tpb = ((long long)T * kslices * nb + 8LL * sms - 1) / (8LL * sms);
  if (tpb < 4) tpb = 4;
  if (tpb > T) tpb = T;
  dim3 grid(kslices, (T + (int)tpb - 1) / (int)tpb, nb);
  joint_dz_reduce_kernel<<<grid, 32 * kDzPh, smem, (cudaStream_t)stream>>>(
      dH, reinterpret_cast<const __nv_bfloat16*>(H), ldh, T, U1, J, (int)tpb, de, dp);
  B200_CHECK_LAUNCH("joint_dz_reduce_kernel");
  return B200_OK;
}

