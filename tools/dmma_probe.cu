// Probe: throughput of mma.sync.aligned.m8n8k4.row.col.f64 (DMMA) against plain DFMA on this GPU.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/dmma_probe tools/dmma_probe.cu && /tmp/dmma_probe
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) k_dmma(double* out, int iters) {
  double c[16][2];
#pragma unroll
  for (int i = 0; i < 16; ++i) { c[i][0] = threadIdx.x * 1e-3 + i; c[i][1] = 1.0 + i; }
  double a[4], b[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) { a[i] = 1.0 + 1e-9 * (threadIdx.x + i); b[i] = 1.0 - 1e-9 * (threadIdx.x + 2 * i); }
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 4; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j)
        asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                     : "+d"(c[4 * i + j][0]), "+d"(c[4 * i + j][1]) : "d"(a[i]), "d"(b[j]));
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 16; ++i) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) k_dfma(double* out, int iters) {
  double c[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) c[i] = threadIdx.x * 1e-3 + i;
  const double a = 1.0000001, b = 1e-9;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 32; ++i) c[i] = fma(c[i], a, b);
  }
  double s = 0;
#pragma unroll
  for (int i = 0; i < 32; ++i) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  int sms = 0;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double* out;
  cudaMalloc(&out, sizeof(double) * sms * 8 * 256);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  for (int warps_per_sm : {4, 8, 16, 32}) {
    const int ctas = sms * warps_per_sm / 8;
    const int iters = 20000;
    float ms;
    k_dmma<<<ctas, 256>>>(out, 100); cudaDeviceSynchronize();
    cudaEventRecord(e0); k_dmma<<<ctas, 256>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    const double fl_mma = 2.0 * 8 * 8 * 4 * 16.0 * iters * (double)ctas * 8;   // per warp: 16 MMAs of 256 FMAs per iteration
    printf("DMMA m8n8k4: %2d warps/SM  %.2f TFLOP/s  (%.1f clk per warp-MMA per SMSP at 1.9 GHz)\n", warps_per_sm, fl_mma / ms * 1e-9,
           ms * 1e-3 * 1.9e9 / (16.0 * iters * warps_per_sm / 4.0));
    k_dfma<<<ctas, 256>>>(out, 100); cudaDeviceSynchronize();
    cudaEventRecord(e0); k_dfma<<<ctas, 256>>>(out, iters); cudaEventRecord(e1); cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    const double fl_fma = 2.0 * 32 * 32.0 * iters * (double)ctas * 8;
    printf("DFMA       : %2d warps/SM  %.2f TFLOP/s\n", warps_per_sm, fl_fma / ms * 1e-9);
  }
  printf("%s\n", cudaGetErrorString(cudaGetLastError()));
  return 0;
}
