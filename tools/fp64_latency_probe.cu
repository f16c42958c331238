// Dev probe: latency of a dependent FP64 FMA chain and the per-warp rate with C independent chains, one warp on one SM.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/fp64_latency_probe tools/fp64_latency_probe.cu && /tmp/fp64_latency_probe
#include <cstdio>
#include <cuda_runtime.h>
template <int C>
__global__ void probe(double* out, long long* cycles, double a, double b) {
  double v[C];
#pragma unroll
  for (int i = 0; i < C; ++i) v[i] = a + i + threadIdx.x;
  const long long t0 = clock64();
#pragma unroll 1
  for (int it = 0; it < 1024; ++it) {
#pragma unroll
    for (int r = 0; r < 8; ++r)
#pragma unroll
      for (int i = 0; i < C; ++i) v[i] = fma(v[i], b, a);
  }
  const long long t1 = clock64();
  double s = 0;
#pragma unroll
  for (int i = 0; i < C; ++i) s += v[i];
  out[threadIdx.x] = s;
  if (threadIdx.x == 0) cycles[0] = t1 - t0;
}
template <int C>
void run(double* out, long long* cyc) {
  probe<C><<<1, 32>>>(out, cyc, 1.0, 0.999999);
  cudaDeviceSynchronize();
  probe<C><<<1, 32>>>(out, cyc, 1.0, 0.999999);
  cudaDeviceSynchronize();
  long long h;
  cudaMemcpy(&h, cyc, sizeof h, cudaMemcpyDeviceToHost);
  printf("chains %2d: %.1f cycles per dependent step, %.2f cycles per warp-wide DFMA\n", C, (double)h / (1024.0 * 8), (double)h / (1024.0 * 8 * C));
}
int main() {
  double* out; long long* cyc;
  cudaMalloc(&out, 32 * sizeof(double)); cudaMalloc(&cyc, sizeof(long long));
  run<1>(out, cyc); run<2>(out, cyc); run<4>(out, cyc); run<8>(out, cyc); run<16>(out, cyc); run<32>(out, cyc);
  return 0;
}
