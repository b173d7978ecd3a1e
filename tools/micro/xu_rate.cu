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
      if (MODE == 5) { asm volatile("fma.rn.f32x2 %0, %0, %0, %0;" : "+l"(*reinterpret_cast<unsigned long long*>(&a[i & 6]))); }
      if (MODE == 4) { uint32_t w = __float_as_uint(a[i]); asm volatile("ex2.approx.f16x2 %0, %0;" : "+r"(w)); a[i] = __uint_as_float(w); }
    }
  }
  float s = 0; for (int i = 0; i < 8; ++i) s += a[i];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s + acc;
}
// the softmax inner loop of the attention kernel: 64 scores -> scale, exp2 (a quarter as a cubic), row sum, bf16 pack
__device__ __forceinline__ float ex2_poly(float x) {
  x = fmaxf(x, -126.0f);
  const float t = x + 12582912.0f, n = t - 12582912.0f, f = x - n;
  float p = fmaf(f, 0.0555041f, 0.2402265f);
  p = fmaf(p, f, 0.6931472f);
  p = fmaf(p, f, 1.0f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}
template <int POLY_OF_16>
__global__ void softmax_loop(float* out, const float* in, int iters) {
  float v[64];
  for (int i = 0; i < 64; ++i) v[i] = in[(threadIdx.x * 64 + i) & 4095];
  float m = 0.5f, l = 0.f;
  uint32_t acc = 0;
  for (int it = 0; it < iters; ++it) {
    float s4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      float p0, p1;
      const float x0 = fmaf(v[2 * i], 1.44269504f, -m), x1 = fmaf(v[2 * i + 1], 1.44269504f, -m);
      asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(p0) : "f"(x0));
      if ((i & 15) < POLY_OF_16) p1 = ex2_poly(x1); else asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(p1) : "f"(x1));
      s4[i & 3] += p0 + p1;
      uint32_t w; asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w) : "f"(p1), "f"(p0));
      acc ^= w;
    }
    l += (s4[0] + s4[1]) + (s4[2] + s4[3]);
    m += 1e-6f;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = l + acc;
}
// the same loop on packed fp32x2 instructions (FFMA2 / FADD2, sm_100): scale and row sum cost half the issue slots
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  float2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(*reinterpret_cast<unsigned long long*>(&d))
      : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)), "l"(*reinterpret_cast<unsigned long long*>(&c)));
  return d;
}
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
  float2 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(*reinterpret_cast<unsigned long long*>(&d))
      : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)));
  return d;
}
// two cubic exp2 at once
__device__ __forceinline__ float2 ex2_poly2(float2 x) {
  x.x = fmaxf(x.x, -126.0f); x.y = fmaxf(x.y, -126.0f);
  const float2 magic = make_float2(12582912.0f, 12582912.0f), nmagic = make_float2(-12582912.0f, -12582912.0f);
  const float2 t = fadd2(x, magic);
  const float2 n = fadd2(t, nmagic);
  const float2 f = ffma2(n, make_float2(-1.f, -1.f), x);
  float2 p = ffma2(f, make_float2(0.0555041f, 0.0555041f), make_float2(0.2402265f, 0.2402265f));
  p = ffma2(p, f, make_float2(0.6931472f, 0.6931472f));
  p = ffma2(p, f, make_float2(1.0f, 1.0f));
  p.x = __int_as_float(__float_as_int(p.x) + (__float_as_int(t.x) << 23));
  p.y = __int_as_float(__float_as_int(p.y) + (__float_as_int(t.y) << 23));
  return p;
}
// POLY_PAIRS_OF_8: of every 8 element pairs, this many whole pairs go through the cubic
template <int POLY_PAIRS_OF_8>
__global__ void softmax_loop2(float* out, const float* in, int iters) {
  float2 v[32];
  for (int i = 0; i < 32; ++i) v[i] = make_float2(in[(threadIdx.x * 64 + 2 * i) & 4095], in[(threadIdx.x * 64 + 2 * i + 1) & 4095]);
  float m = 0.5f, l = 0.f;
  uint32_t acc = 0;
  for (int it = 0; it < iters; ++it) {
    float2 s2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
    const float2 sc = make_float2(1.44269504f, 1.44269504f), nm = make_float2(-m, -m);
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const float2 x = ffma2(v[i], sc, nm);
      float2 p;
      if ((i & 7) < POLY_PAIRS_OF_8) p = ex2_poly2(x);
      else {
        asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(p.x) : "f"(x.x));
        asm volatile("ex2.approx.ftz.f32 %0, %1;" : "=f"(p.y) : "f"(x.y));
      }
      s2[i & 1] = fadd2(s2[i & 1], p);
      uint32_t w; asm volatile("cvt.rn.bf16x2.f32 %0, %1, %2;" : "=r"(w) : "f"(p.y), "f"(p.x));
      acc ^= w;
    }
    l += (s2[0].x + s2[0].y) + (s2[1].x + s2[1].y);
    m += 1e-6f;
  }
  out[blockIdx.x * blockDim.x + threadIdx.x] = l + acc;
}
template <int PP> void run_softmax2(int warps_per_sm) {
  int sms = 148, iters = 2000;
  float *out, *in; cudaMalloc(&out, sizeof(float) * sms * warps_per_sm * 32); cudaMalloc(&in, 4096 * 4); cudaMemset(in, 0, 4096 * 4);
  softmax_loop2<PP><<<sms, warps_per_sm * 32>>>(out, in, 16);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a); softmax_loop2<PP><<<sms, warps_per_sm * 32>>>(out, in, iters); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  const double el = (double)warps_per_sm * 32 * 64 * iters;
  printf("packed softmax loop poly %4.1f%%  warps/SM %2d: %.3f ms -> %.2f elements/ns/SM\n", PP * 12.5, warps_per_sm, ms, el / (ms * 1e6));
  cudaFree(out); cudaFree(in);
}
template <int POLY_OF_16> void run_softmax(int warps_per_sm) {
  int sms = 148, iters = 2000;
  float *out, *in; cudaMalloc(&out, sizeof(float) * sms * warps_per_sm * 32); cudaMalloc(&in, 4096 * 4); cudaMemset(in, 0, 4096 * 4);
  softmax_loop<POLY_OF_16><<<sms, warps_per_sm * 32>>>(out, in, 16);
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  cudaEventRecord(a); softmax_loop<POLY_OF_16><<<sms, warps_per_sm * 32>>>(out, in, iters); cudaEventRecord(b); cudaEventSynchronize(b);
  float ms; cudaEventElapsedTime(&ms, a, b);
  // clk per 64-element row block per SM sub-partition (warps_per_sm / 4 warps share one)
  const double el = (double)warps_per_sm * 32 * 64 * iters;      // per SM
  printf("softmax loop poly %2d/32  warps/SM %2d: %.3f ms -> %.2f elements/ns/SM (MUFU-only bound at 1.9 GHz: %.1f)\n",
         POLY_OF_16 * 2, warps_per_sm, ms, el / (ms * 1e6), 16 * 1.9 / (1.0 - POLY_OF_16 / 32.0));
  cudaFree(out); cudaFree(in);
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
int main() { run<0>("ex2"); run<1>("f2fp.bf16x2"); run<2>("ffma"); run<3>("tanh"); run<4>("ex2.f16x2 (instr)"); run<5>("ffma2 (instr)");
  for (int w : {8, 16}) { run_softmax2<0>(w); run_softmax2<1>(w); run_softmax2<2>(w); run_softmax2<3>(w); run_softmax2<4>(w); }
  for (int w : {8, 16}) { run_softmax<0>(w); run_softmax<4>(w); run_softmax<8>(w); }
  return 0; }
