// Micro-benchmark behind the layout of the row-streaming passes (ctc_grad / rnnt_grad / *_logprobs):
//   out[r, v] = s[r] * exp2(x[r, v] * log2e + c[r])   over [R, V] fp32, V not a multiple of the tile.
// Variants differ only in which thread touches which bytes and how many 16-byte loads are in flight.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/stream_variants scripts/stream_variants.cu
// Usage: stream_variants [R] [V]
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ uint4 ld16(const uint4* p) {
  uint4 v;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ void st16(uint4* p, uint4 v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint4 xform(uint4 v, float c, float s) {
  float* f = reinterpret_cast<float*>(&v);
#pragma unroll
  for (int k = 0; k < 4; ++k) f[k] = s * ex2(fmaf(f[k], 1.4426950408889634f, c));
  return v;
}

// one row = nseg 16-byte segments handled by NT threads, U loads in flight per thread
template <int NT, int U>
__device__ __forceinline__ void stream_row(const uint4* __restrict__ in, uint4* __restrict__ out, int nseg, int tid,
                                           float c, float s) {
  uint4 v[U], nv[U];
#pragma unroll
  for (int j = 0; j < U; ++j) {
    const int idx = j * NT + tid;
    if (idx < nseg) nv[j] = ld16(in + idx);
  }
  for (int base = 0; base < nseg; base += NT * U) {
#pragma unroll
    for (int j = 0; j < U; ++j) v[j] = nv[j];
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const int idx = base + NT * U + j * NT + tid;
      if (idx < nseg) nv[j] = ld16(in + idx);
    }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const int idx = base + j * NT + tid;
      if (idx < nseg) st16(out + idx, xform(v[j], c, s));
    }
  }
}

// V0 / V1: one warp per row, 8 rows per CTA
template <int U, int MINB>
__global__ void __launch_bounds__(256, MINB) warp_per_row(const float* x, const float* c, const float* s, float* y,
                                                          int R, int V) {
  const long long r = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (r >= R) return;
  stream_row<32, U>(reinterpret_cast<const uint4*>(x + r * V), reinterpret_cast<uint4*>(y + r * V), V / 4,
                    threadIdx.x & 31, c[r], s[r]);
}

// V2: the CTA walks its 8 rows one after the other, 256 threads x 16 bytes contiguous per instruction
template <int U, int MINB>
__global__ void __launch_bounds__(256, MINB) cta_per_rows(const float* x, const float* c, const float* s, float* y,
                                                          int R, int V) {
  for (int w = 0; w < 8; ++w) {
    const long long r = (long long)blockIdx.x * 8 + w;
    if (r >= R) return;
    stream_row<256, U>(reinterpret_cast<const uint4*>(x + r * V), reinterpret_cast<uint4*>(y + r * V), V / 4,
                       threadIdx.x, c[r], s[r]);
  }
}

// V3: flat, per-row factors looked up per segment (upper bound of any row mapping)
template <int U>
__global__ void __launch_bounds__(256, 4) flat(const float* x, const float* c, const float* s, float* y, long long nseg,
                                               int segs_per_row) {
  const long long stride = (long long)gridDim.x * 256 * U;
  for (long long base = (long long)blockIdx.x * 256 * U; base < nseg; base += stride) {
    uint4 v[U];
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const long long idx = base + j * 256 + threadIdx.x;
      if (idx < nseg) v[j] = ld16(reinterpret_cast<const uint4*>(x) + idx);
    }
#pragma unroll
    for (int j = 0; j < U; ++j) {
      const long long idx = base + j * 256 + threadIdx.x;
      if (idx < nseg) {
        const int r = (int)(idx / segs_per_row);
        st16(reinterpret_cast<uint4*>(y) + idx, xform(v[j], c[r], s[r]));
      }
    }
  }
}

// V4: persistent warps, each takes rows r = w, w + nwarps, ... (no CTA drain between rows)
template <int U, int MINB>
__global__ void __launch_bounds__(256, MINB) warp_persistent(const float* x, const float* c, const float* s, float* y,
                                                             int R, int V) {
  const int nw = gridDim.x * 8;
  for (long long r = (long long)blockIdx.x * 8 + (threadIdx.x >> 5); r < R; r += nw)
    stream_row<32, U>(reinterpret_cast<const uint4*>(x + r * V), reinterpret_cast<uint4*>(y + r * V), V / 4,
                      threadIdx.x & 31, c[r], s[r]);
}

template <typename F>
static void bench(const char* name, F launch, double bytes) {
  for (int i = 0; i < 3; ++i) launch();
  CK(cudaDeviceSynchronize());
  cudaEvent_t a, b;
  cudaEventCreate(&a);
  cudaEventCreate(&b);
  float best = 1e9f, tot = 0.f;
  const int reps = 10;
  for (int i = 0; i < reps; ++i) {
    cudaEventRecord(a);
    launch();
    cudaEventRecord(b);
    CK(cudaEventSynchronize(b));
    float ms;
    cudaEventElapsedTime(&ms, a, b);
    best = ms < best ? ms : best;
    tot += ms;
  }
  CK(cudaGetLastError());
  printf("%-34s avg %7.1f us  best %7.1f us  %6.0f GB/s (avg)\n", name, tot / reps * 1e3, best * 1e3,
         bytes / (tot / reps * 1e-3) / 1e9);
}

int main(int argc, char** argv) {
  const int R = argc > 1 ? atoi(argv[1]) : 64000;
  const int V = argc > 2 ? atoi(argv[2]) : 5000;
  const size_t n = (size_t)R * V;
  float *x, *y, *c, *s;
  CK(cudaMalloc(&x, n * 4));
  CK(cudaMalloc(&y, n * 4));
  CK(cudaMalloc(&c, R * 4));
  CK(cudaMalloc(&s, R * 4));
  CK(cudaMemset(x, 0, n * 4));
  CK(cudaMemset(c, 0, R * 4));
  CK(cudaMemset(s, 0, R * 4));
  const double bytes = 2.0 * n * 4;
  const int grid = (R + 7) / 8;
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  printf("R=%d V=%d  %.2f GB per pass, %d SMs\n", R, V, bytes / 1e9, sms);
  bench("copy (cudaMemcpy D2D)", [&] { cudaMemcpyAsync(y, x, n * 4, cudaMemcpyDeviceToDevice); }, bytes);
  bench("warp/row U=4 (256,4)", [&] { warp_per_row<4, 4><<<grid, 256>>>(x, c, s, y, R, V); }, bytes);
  bench("warp/row U=4 (256,3)", [&] { warp_per_row<4, 3><<<grid, 256>>>(x, c, s, y, R, V); }, bytes);
  bench("warp/row U=8 (256,2)", [&] { warp_per_row<8, 2><<<grid, 256>>>(x, c, s, y, R, V); }, bytes);
  bench("warp/row U=2 (256,8)", [&] { warp_per_row<2, 8><<<grid, 256>>>(x, c, s, y, R, V); }, bytes);
  bench("cta/8 rows U=1 (256,8)", [&] { cta_per_rows<1, 8><<<grid, 256>>>(x, c, s, y, R, V); }, bytes);
  bench("cta/8 rows U=2 (256,6)", [&] { cta_per_rows<2, 6><<<grid, 256>>>(x, c, s, y, R, V); }, bytes);
  bench("cta/8 rows U=4 (256,4)", [&] { cta_per_rows<4, 4><<<grid, 256>>>(x, c, s, y, R, V); }, bytes);
  bench("flat U=4, 4/SM", [&] { flat<4><<<sms * 4, 256>>>(x, c, s, y, (long long)n / 4, V / 4); }, bytes);
  bench("flat U=8, 4/SM", [&] { flat<8><<<sms * 4, 256>>>(x, c, s, y, (long long)n / 4, V / 4); }, bytes);
  bench("flat U=4, 8/SM", [&] { flat<4><<<sms * 8, 256>>>(x, c, s, y, (long long)n / 4, V / 4); }, bytes);
  bench("persistent warp/row U=4 4/SM", [&] { warp_persistent<4, 4><<<sms * 4, 256>>>(x, c, s, y, R, V); }, bytes);
  bench("persistent warp/row U=8 2/SM", [&] { warp_persistent<8, 2><<<sms * 2, 256>>>(x, c, s, y, R, V); }, bytes);
  return 0;
}
