// Shared device/host helpers for libb200_transducer (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include <math.h>
#include <atomic>

#include "../../include/b200_transducer.h"

namespace b200 {

// ---- host side: error string + launch accounting ------------------------
void set_error(const char* fmt, ...);
extern std::atomic<uint64_t> g_launches;
inline void count_launch(int n = 1) { g_launches.fetch_add((uint64_t)n, std::memory_order_relaxed); }

#define B200_CHECK_ARG(cond, ...)                 \
  do {                                            \
    if (!(cond)) {                                \
      ::b200::set_error(__VA_ARGS__);             \
      return B200_ERR_ARG;                        \
    }                                             \
  } while (0)

#define B200_CHECK_LAUNCH(name)                                             \
  do {                                                                      \
    cudaError_t e__ = cudaGetLastError();                                   \
    if (e__ != cudaSuccess) {                                               \
      ::b200::set_error("%s: launch failed: %s", name, cudaGetErrorString(e__)); \
      return B200_ERR_CUDA;                                                 \
    }                                                                       \
    ::b200::count_launch();                                                 \
  } while (0)

#define B200_CHECK_CUDA(call)                                               \
  do {                                                                      \
    cudaError_t e__ = (call);                                               \
    if (e__ != cudaSuccess) {                                               \
      ::b200::set_error("%s failed: %s", #call, cudaGetErrorString(e__));   \
      return B200_ERR_CUDA;                                                 \
    }                                                                       \
  } while (0)

// ---- device helpers -------------------------------------------------------
constexpr float kLog2e = 1.4426950408889634f;
constexpr float kLn2 = 0.6931471805599453f;
constexpr float kNegBig = -1.0e30f;  // finite stand-in for -inf in running maxima (x log2e must not overflow)

__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float lg2_approx(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// log(exp(a)+exp(b)), exact-ish (libdevice expf/log1pf), -inf safe
__device__ __forceinline__ float log_add_exp(float a, float b) {
  float m = fmaxf(a, b);
  if (m == -INFINITY) return -INFINITY;
  float d = -fabsf(a - b);  // -inf if one side is -inf -> exp -> 0
  return m + log1pf(expf(d));
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

constexpr int kLag = 4;  // lattice steps between rescalings (power of two)

// ---- log2-domain helpers of the trellis / lattice chains (ctc_loss.cu, rnnt_loss.cu) ----
// log2(2^a + 2^b); -inf arguments allowed, both -inf -> -inf.  Maximum factored out: 1 ex2 + 1 lg2
__device__ __forceinline__ float lae2_log2(float a, float b) {
  const float m = fmaxf(a, b), mn = fminf(a, b);
  const float ms = (m == -INFINITY) ? 0.f : m;
  return m + lg2_approx(1.f + ex2_approx(mn - ms));
}
// log2(2^a + 2^b + 2^c): 2 ex2 + 1 lg2
__device__ __forceinline__ float lae3_log2(float a, float b, float c) {
  const float lo = fminf(a, b), hi = fmaxf(a, b);
  const float m = fmaxf(hi, c);
  const float mid = fmaxf(lo, fminf(hi, c));
  const float mn = fminf(lo, c);
  const float ms = (m == -INFINITY) ? 0.f : m;
  return m + lg2_approx(1.f + ex2_approx(mid - ms) + ex2_approx(mn - ms));
}
// warp maximum of floats (no NaNs) through the integer redux unit: one instruction instead of 5 shuffles
__device__ __forceinline__ float warp_max_redux(float v) {
  int k = __float_as_int(v);
  k = k >= 0 ? k : k ^ 0x7fffffff;
  k = __reduce_max_sync(0xffffffffu, k);
  k = k >= 0 ? k : k ^ 0x7fffffff;
  return __int_as_float(k);
}

// maximum over the warps' slots of one parity of red = [2][32]; all -inf -> 0 (nothing to subtract)
__device__ __forceinline__ float block_max_read(const float* red, int parity, int nw) {
  float m = red[parity * 32];
  for (int i = 1; i < nw; ++i) m = fmaxf(m, red[parity * 32 + i]);
  return m == -INFINITY ? 0.f : m;
}

// Lagged block maximum for the rescaled lattices: at step i every warp publishes the max
// of the values it just produced (before the step's __syncthreads); at step i+1 all
// threads read the other parity slot.  No extra barrier.  red = [2][32] floats.
__device__ __forceinline__ void lagged_max_write(float v, float* red, int parity) {
  v = warp_max(v);
  if ((threadIdx.x & 31) == 0) red[parity * 32 + (threadIdx.x >> 5)] = v;
}
__device__ __forceinline__ float lagged_max_read(const float* red, int parity) {
  const int nw = (blockDim.x + 31) >> 5;
  float m = red[parity * 32];
  for (int i = 1; i < nw; ++i) m = fmaxf(m, red[parity * 32 + i]);
  return m == -INFINITY ? 0.f : m;
}

// streaming 16-byte global load: read once, do not pollute L1
__device__ __forceinline__ uint4 ldg_stream16(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void stg_stream16(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// element <-> float conversion for the three logits dtypes.
template <typename T>
struct Elem;
template <>
struct Elem<float> {
  static constexpr int kPerVec = 4;
  __device__ static __forceinline__ void unpack(const uint4& v, float* f) {
    f[0] = __uint_as_float(v.x);
    f[1] = __uint_as_float(v.y);
    f[2] = __uint_as_float(v.z);
    f[3] = __uint_as_float(v.w);
  }
  __device__ static __forceinline__ uint4 pack(const float* f) {
    return make_uint4(__float_as_uint(f[0]), __float_as_uint(f[1]), __float_as_uint(f[2]),
                      __float_as_uint(f[3]));
  }
  __device__ static __forceinline__ float to_float(float x) { return x; }
  __device__ static __forceinline__ float from_float(float x) { return x; }
};
template <>
struct Elem<__half> {
  static constexpr int kPerVec = 8;
  __device__ static __forceinline__ void unpack(const uint4& v, float* f) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __half2 h = *reinterpret_cast<const __half2*>(&w[i]);
      float2 p = __half22float2(h);
      f[2 * i] = p.x;
      f[2 * i + 1] = p.y;
    }
  }
  __device__ static __forceinline__ uint4 pack(const float* f) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __half2 h = __floats2half2_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
  __device__ static __forceinline__ float to_float(__half x) { return __half2float(x); }
  __device__ static __forceinline__ __half from_float(float x) { return __float2half_rn(x); }
};
template <>
struct Elem<__nv_bfloat16> {
  static constexpr int kPerVec = 8;
  __device__ static __forceinline__ void unpack(const uint4& v, float* f) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {  // bf16 -> f32 is a 16-bit shift
      f[2 * i] = __uint_as_float(w[i] << 16);
      f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  __device__ static __forceinline__ uint4 pack(const float* f) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<uint32_t*>(&h);
    }
    return make_uint4(w[0], w[1], w[2], w[3]);
  }
  __device__ static __forceinline__ float to_float(__nv_bfloat16 x) { return __bfloat162float(x); }
  __device__ static __forceinline__ __nv_bfloat16 from_float(float x) {
    return __float2bfloat16_rn(x);
  }
};

// A row of V elements starting at an arbitrary element-aligned address, seen as
// the 16-byte segments that cover it.  Segment s holds row indices
// [s*kPerVec - head, s*kPerVec - head + kPerVec); indices outside [0,V) belong
// to the neighbouring rows (same allocation: the aligned segment containing a
// valid byte never leaves that byte's page, so reading it is safe).
template <typename T>
struct RowSegs {
  const uint4* seg0;  // first covering segment (16-B aligned)
  int head;           // elements of seg 0 that precede the row
  int nseg;           // covering segments
  __device__ __forceinline__ RowSegs(const T* row, int V) {
    uintptr_t a = reinterpret_cast<uintptr_t>(row);
    uintptr_t a0 = a & ~uintptr_t(15);
    head = int((a - a0) / sizeof(T));
    seg0 = reinterpret_cast<const uint4*>(a0);
    nseg = (head + V + Elem<T>::kPerVec - 1) / Elem<T>::kPerVec;
  }
};

}  // namespace b200
