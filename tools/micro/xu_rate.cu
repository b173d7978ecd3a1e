// Micro-benchmark: per-SM throughput of MUFU.EX2, F2FP.BF16 pack, FFMA on this GPU.
#include <cstdio>
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
template <int MODE>
__global__ void k(float* out, int iters) {
  float a[8];
  for (int i = 0; i < 8; ++i) a[i] = 0.001f * (threadIdx.x + i);
  uint32_t acc = 0;
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (MODE == 0) { asm volatile("ex2.approx.ftz.f32 %0, %0;" : "+f"(a[i])); }
      if (MODE == 1) { uint32_t w; asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w) : "f"(a[i]), "f"(a[(i + 1) & 7])); acc ^= w; a[i] += 1.0f; }
      if (MODE == 2) { asm volatile("fma.rn.f32 %0, %0, %0, %0;" : "+f"(a[i])); }
      if (MODE == 3) { asm volatile("tanh.approx.f32 %0, %0;" : "+f"(a[i])); }
    }
  }
  float s = 0; for (int i = 0; i < 8; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + acc;
}
template <int MODE> void run(const char* name) {
  int sms = 148, tpb = 1024, bps = 2, iters = 4096;
  float* out; cudaMalloc(&out, sizeof(float) * sms * bps * tpb);
  k<MODE><<<sms * bps, tpb>>>(out, 16);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a); k<MODE><<<sms * bps, tpb>>>(out, iters); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  double ops = (double)sms * bps * tpb * iters * 8;
  int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  printf("%-10s %.3f ms  %.1f Gop/s  -> %.2f ops/clk/SM at %d MHz nominal (%.2f at 1.7 GHz)\n", name, ms, ops / ms / 1e6,
         ops / (ms * 1e-3) / sms / (clk * 1e3), clk / 1000, ops / (ms * 1e-3) / sms / 1.7e9);
  cudaFree(out);
}
int main() { run<0>("ex2"); run<1>("f2fp.bf16x2"); run<2>("ffma"); run<3>("tanh"); return 0; }
