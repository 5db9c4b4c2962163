// Microbenchmark: scalar FFMA/FADD vs packed FFMA2/FADD2 issue throughput on sm_100a,
// alone and interleaved with shared-memory loads.  Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, int iters, float c) {
  __shared__ float2 sm[1024];
  for (int i = threadIdx.x; i < 1024; i += 256) sm[i] = make_float2(i * 1e-3f, 1.f);
  __syncthreads();
  float2 a[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) a[i] = make_float2(threadIdx.x * 1e-3f + i, i * 0.5f);
  const float2 cc = make_float2(c, c), dd = make_float2(1e-3f, 1e-3f);
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int r = 0; r < 4; ++r) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (MODE == 0) { a[i].x = fmaf(a[i].x, c, 1e-3f); a[i].y = fmaf(a[i].y, c, 1e-3f); }
        if (MODE == 1) a[i] = __ffma2_rn(a[i], cc, dd);
        if (MODE == 2) { a[i].x = a[i].x + c; a[i].y = a[i].y + c; }
        if (MODE == 3) a[i] = __fadd2_rn(a[i], cc);
        if (MODE == 4) {  // scalar fma + one 64-bit LDS per 2 fma pairs
          a[i].x = fmaf(a[i].x, c, 1e-3f); a[i].y = fmaf(a[i].y, c, 1e-3f);
          if ((i & 1) == 0) { float2 v = sm[(threadIdx.x + 32 * (i + 8 * r) + it) & 1023]; a[i].x += v.x; a[i].y += v.y; }
        }
        if (MODE == 5) {
          a[i] = __ffma2_rn(a[i], cc, dd);
          if ((i & 1) == 0) { float2 v = sm[(threadIdx.x + 32 * (i + 8 * r) + it) & 1023]; a[i] = __fadd2_rn(a[i], v); }
        }
      }
    }
  }
  float s = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += a[i].x + a[i].y;
  out[blockIdx.x * 256 + threadIdx.x] = s;
}

template <int MODE>
void run(const char* name, float* d, int ctas_per_sm) {
  const int iters = 4096, grid = 148 * ctas_per_sm;
  k<MODE><<<grid, 256>>>(d, 16, 0.999f);
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  cudaDeviceSynchronize();
  cudaEventRecord(e0);
  k<MODE><<<grid, 256>>>(d, iters, 0.999f);
  cudaEventRecord(e1); cudaEventSynchronize(e1);
  float ms; cudaEventElapsedTime(&ms, e0, e1);
  const double lane_ops = (double)grid * 256 * iters * 4 * 8 * 2;  // fp32 lane-ops (fma = 1 op here)
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("%-28s ctas/SM=%d  %.3f ms  %.1f lane-ops/clk/SM (at %d MHz nominal)  %.2f Tops/s\n", name,
         ctas_per_sm, ms, lane_ops / (ms * 1e-3) / (clk * 1e3) / 148, clk / 1000, lane_ops / ms / 1e9);
}

int main() {
  float* d; cudaMalloc(&d, 148 * 8 * 256 * 4);
  for (int c : {2, 4, 8}) {
    run<0>("scalar FFMA", d, c);
    run<1>("packed FFMA2", d, c);
    run<2>("scalar FADD", d, c);
    run<3>("packed FADD2", d, c);
    run<4>("scalar FFMA + LDS.64", d, c);
    run<5>("packed FFMA2 + LDS.64", d, c);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
