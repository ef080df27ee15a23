// Warp-per-row streaming primitives shared by the RNN-T and CTC passes.
#pragma once
#include "common.cuh"

namespace b200 {

constexpr int kRowWarps = 8;  // warps (= rows) per CTA in the streaming passes
constexpr int kUnroll = 4;    // 16-byte loads in flight per lane

// log-sum-exp of one row of V elements, computed by one warp with an online
// (max, sum-exp) per lane and a shuffle reduction.  Every lane returns the result.
template <typename T>
__device__ __forceinline__ float warp_row_lse(const T* __restrict__ row, int V, int lane) {
  constexpr int K = Elem<T>::kPerVec;
  RowSegs<T> rs(row, V);
  float m = kNegBig, s = 0.f;
  // software pipeline: the loads of block i+1 are in flight while block i is reduced
  uint4 v[kUnroll], nv[kUnroll];
#pragma unroll
  for (int j = 0; j < kUnroll; ++j) {
    int idx = j * 32 + lane;
    if (idx < rs.nseg) nv[j] = ldg_stream16(rs.seg0 + idx);
  }
  for (int base = 0; base < rs.nseg; base += 32 * kUnroll) {
#pragma unroll
    for (int j = 0; j < kUnroll; ++j) v[j] = nv[j];
#pragma unroll
    for (int j = 0; j < kUnroll; ++j) {
      int idx = base + 32 * kUnroll + j * 32 + lane;
      if (idx < rs.nseg) nv[j] = ldg_stream16(rs.seg0 + idx);
    }
    float f[kUnroll][K];
    float lmax = kNegBig;
#pragma unroll
    for (int j = 0; j < kUnroll; ++j) {
      int idx = base + j * 32 + lane;
      if (idx < rs.nseg) {
        Elem<T>::unpack(v[j], f[j]);
        if (idx == 0 || idx == rs.nseg - 1) {  // edge segments may hold neighbours' elements
#pragma unroll
          for (int k = 0; k < K; ++k) {
            int e = idx * K - rs.head + k;
            if (e < 0 || e >= V) f[j][k] = -INFINITY;
          }
        }
      } else {
#pragma unroll
        for (int k = 0; k < K; ++k) f[j][k] = -INFINITY;
      }
#pragma unroll
      for (int k = 0; k < K; ++k) lmax = fmaxf(lmax, f[j][k]);
    }
    float mn = fmaxf(m, lmax);
    float mn2 = mn * kLog2e;
    float acc = 0.f;
#pragma unroll
    for (int j = 0; j < kUnroll; ++j)
#pragma unroll
      for (int k = 0; k < K; ++k) acc += ex2_approx(fmaf(f[j][k], kLog2e, -mn2));
    s = s * ex2_approx((m - mn) * kLog2e) + acc;
    m = mn;
  }
  float M = warp_max(m);
  s *= ex2_approx((m - M) * kLog2e);
  float S = warp_sum(s);
  return M + logf(S);
}

// First block of 16-byte loads of a row stream.  Split from the body so that a caller can put
// independent work (whose own loads are already issued) between the two: the stream's first
// round trip to HBM then overlaps that work instead of following it.
template <typename T>
__device__ __forceinline__ void warp_row_stream_begin(const RowSegs<T>& rs, int lane, uint4 (&nv)[kUnroll]) {
#pragma unroll
  for (int j = 0; j < kUnroll; ++j) {
    int idx = j * 32 + lane;
    if (idx < rs.nseg) nv[j] = ldg_stream16(rs.seg0 + idx);
  }
}

// Stream one row: out[e] = clamp(scale * exp(x[e] + c0)) (or zeros when dead), with
// 16-byte stores for whole segments and element stores on the two edge segments that
// may be shared with neighbouring rows.  c0l2 = c0 * log2(e).  nv: the loads of
// warp_row_stream_begin (unused when dead).
template <typename T>
__device__ __forceinline__ void warp_row_softmax_store_from(const RowSegs<T>& rs, uint4 (&nv)[kUnroll],
                                                            T* __restrict__ out, int V, int lane,
                                                            bool dead, float c0l2, float scale,
                                                            float clamp) {
  constexpr int K = Elem<T>::kPerVec;
  uint4* oseg0 = reinterpret_cast<uint4*>(reinterpret_cast<uintptr_t>(out) & ~uintptr_t(15));
  uint4 v[kUnroll];
  for (int base = 0; base < rs.nseg; base += 32 * kUnroll) {
    if (!dead) {  // software pipeline: next block's loads in flight while this one is written
#pragma unroll
      for (int j = 0; j < kUnroll; ++j) v[j] = nv[j];
#pragma unroll
      for (int j = 0; j < kUnroll; ++j) {
        int idx = base + 32 * kUnroll + j * 32 + lane;
        if (idx < rs.nseg) nv[j] = ldg_stream16(rs.seg0 + idx);
      }
    }
#pragma unroll
    for (int j = 0; j < kUnroll; ++j) {
      int idx = base + j * 32 + lane;
      if (idx >= rs.nseg) continue;
      float f[K];
      if (!dead) {
        Elem<T>::unpack(v[j], f);
#pragma unroll
        for (int k = 0; k < K; ++k) {
          float g = scale * ex2_approx(fmaf(f[k], kLog2e, c0l2));
          f[k] = clamp > 0.f ? fminf(fmaxf(g, -clamp), clamp) : g;
        }
      } else {
#pragma unroll
        for (int k = 0; k < K; ++k) f[k] = 0.f;
      }
      const int e0 = idx * K - rs.head;
      if (e0 >= 0 && e0 + K <= V) {
        stg_stream16(oseg0 + idx, Elem<T>::pack(f));
      } else {
#pragma unroll
        for (int k = 0; k < K; ++k) {
          int e = e0 + k;
          if (e >= 0 && e < V) out[e] = Elem<T>::from_float(f[k]);
        }
      }
    }
  }
}

template <typename T>
__device__ __forceinline__ void warp_row_softmax_store(const T* __restrict__ row,
                                                       T* __restrict__ out, int V, int lane,
                                                       bool dead, float c0l2, float scale,
                                                       float clamp) {
  RowSegs<T> rs(row, V);
  uint4 nv[kUnroll];
  if (!dead) warp_row_stream_begin<T>(rs, lane, nv);
  warp_row_softmax_store_from<T>(rs, nv, out, V, lane, dead, c0l2, scale, clamp);
}

}  // namespace b200
