// FP64 peak probe for the roofline denominators MEASURED_PEAKS.json does not carry:
// register-resident DMMA.8x8x4 and DFMA loops, 148 x k CTAs, timed with CUDA events.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/fp64_peak tools/fp64_peak.cu && tools/fp64_peak
#include <cstdio>
#include <cuda_runtime.h>

__global__ void __launch_bounds__(256) k_dmma(double* out, int iters) {
  double c[16][2];
#pragma unroll
  for (int i = 0; i < 16; i++) { c[i][0] = 0.0; c[i][1] = 0.0; }
  double a = 1.0 + threadIdx.x * 1e-9, b = 1.0 - threadIdx.x * 1e-9;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 16; i++)
      asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                   : "+d"(c[i][0]), "+d"(c[i][1]) : "d"(a), "d"(b));
  }
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += c[i][0] + c[i][1];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

__global__ void __launch_bounds__(256) k_dfma(double* out, int iters) {
  double c[16];
#pragma unroll
  for (int i = 0; i < 16; i++) c[i] = threadIdx.x * 1e-9 + i;
  double a = 1.0000001, b = 1e-9;
  for (int it = 0; it < iters; it++) {
#pragma unroll
    for (int i = 0; i < 16; i++) c[i] = fma(c[i], a, b);
  }
  double s = 0.0;
#pragma unroll
  for (int i = 0; i < 16; i++) s += c[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

int main() {
  int sms = 148;
  cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
  double* out;
  cudaMalloc(&out, sizeof(double) * sms * 8 * 256);
  cudaEvent_t e0, e1;
  cudaEventCreate(&e0); cudaEventCreate(&e1);
  double best_dmma = 0, best_dfma = 0;
  for (int occ = 1; occ <= 8; occ *= 2) {
    int grid = sms * occ, iters = 20000;
    k_dmma<<<grid, 256>>>(out, 100);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k_dmma<<<grid, 256>>>(out, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    // per warp instruction: 8x8x4 MACs = 256 FMA = 512 flop; 8 warps per CTA, 16 per iteration
    double tf = (double)grid * 8 * 16 * iters * 512.0 / (ms * 1e-3) / 1e12;
    if (tf > best_dmma) best_dmma = tf;
    k_dfma<<<grid, 256>>>(out, 100);
    cudaDeviceSynchronize();
    cudaEventRecord(e0);
    k_dfma<<<grid, 256>>>(out, iters);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    cudaEventElapsedTime(&ms, e0, e1);
    double tf2 = (double)grid * 256 * 16 * iters * 2.0 / (ms * 1e-3) / 1e12;
    if (tf2 > best_dfma) best_dfma = tf2;
    fprintf(stderr, "ctas/SM=%d dmma=%.2f TF/s dfma=%.2f TF/s\n", occ, tf, tf2);
  }
  printf("{\"dmma_tflops\": %.3f, \"dfma_tflops\": %.3f, \"sms\": %d, \"how\": \"tools/fp64_peak.cu: register-resident "
         "DMMA.8x8x4 / DFMA loops, best over 1-8 CTAs per SM, CUDA events\"}\n", best_dmma, best_dfma, sms);
  return 0;
}
