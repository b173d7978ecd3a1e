// Micro-benchmark: the attention kernel's tcgen05.mma mix (per 96-key block and key group: 4 x SS M128 N96 K16 for S = Q K^T,
// then 6 x TS M128 N64 K16 for O += P V) issued (a) by ONE thread for both key groups, round robin, (b) by TWO threads in
// different warps, one per group, as attention_sm100.cu does.  Cycles per block pair (both groups), issue + completion.
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../eventgpt_b200/csrc/sm100_ptx.cuh"
using namespace egpt;
__host__ __device__ constexpr uint32_t idesc_bf16(int m, int n, int b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)b_mn_major << 16) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(m >> 4) << 24);
}
__device__ __forceinline__ uint64_t desc_mn_sw128(uint32_t a) {
  uint64_t d = 0;
  d |= (uint64_t)((a & 0x3FFFFu) >> 4); d |= (uint64_t)1 << 16; d |= (uint64_t)(1024 >> 4) << 32; d |= (uint64_t)1 << 46; d |= (uint64_t)2 << 61;
  return d;
}
__device__ __forceinline__ void mma_ts(uint32_t d, uint32_t a, uint64_t db, uint32_t idesc, uint32_t acc) {
  asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
               ::"r"(d), "r"(a), "l"(db), "r"(idesc), "r"(acc) : "memory");
}
// one block of group g: S into tmem columns g*192 + (n&1)*96, P V from those columns into 384 + g*64
__device__ __forceinline__ void block(uint32_t tb, uint32_t sq, uint32_t sk, uint32_t sv, int g, int n, int what) {
  const uint64_t dq = ptx::make_smem_desc_sw128(sq), dk = ptx::make_smem_desc_sw128(sk + g * 36864 + (n % 3) * 12288);
  const uint64_t dv = desc_mn_sw128(sv + g * 36864 + (n % 3) * 12288);
  const uint32_t ts = tb + g * 192 + (n & 1) * 96, to = tb + 384 + g * 64;
  if (what & 1)
    for (int k = 0; k < 4; ++k) ptx::umma_bf16_ss(ts, dq + (uint64_t)(k * 2), dk + (uint64_t)(k * 2), idesc_bf16(128, 96, 0), k != 0);
  if (what & 2)
    for (int kk = 0; kk < 6; ++kk) mma_ts(to, ts + kk * 8, dv + (uint64_t)(kk * 128), idesc_bf16(128, 64, 1), 1u);
}
template <int MODE>   // 0: one issuer, both groups; 1: two issuers (warps 1 and 3); 2: one issuer, one group only
__global__ void k(long long* out, int blocks, int what) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar[2]; __shared__ uint32_t tptr;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 176 * 1024 / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { ptx::mbar_init(&bar[0], 1); ptx::mbar_init(&bar[1], 1); ptx::fence_barrier_init(); }
  if (warp == 0) { ptx::tmem_alloc(&tptr, 512); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async();
  ptx::tcgen05_fence_before(); __syncthreads(); ptx::tcgen05_fence_after();
  const uint32_t tb = tptr;
  const uint32_t sq = ptx::smem_u32(smem), sk = sq + 16384, sv = sk + 73728;
  const bool issuer = (threadIdx.x & 31) == 0 && (MODE == 1 ? (warp == 1 || warp == 3) : warp == 1);
  if (issuer) {
    const int g = MODE == 1 ? (warp == 3) : 0;
    long long t0 = clock64();
    for (int n = 0; n < blocks; ++n) {
      block(tb, sq, sk, sv, g, n, what);
      if (MODE == 0) block(tb, sq, sk, sv, 1, n, what);
    }
    long long t1 = clock64();
    ptx::umma_commit(&bar[g]);
    ptx::mbar_wait(&bar[g], 0);
    long long t2 = clock64();
    out[2 * g] = t1 - t0; out[2 * g + 1] = t2 - t0;
  }
  ptx::tcgen05_fence_before(); __syncthreads(); ptx::tcgen05_fence_after();
  if (warp == 0) ptx::tmem_dealloc(tb, 512);
}
template <int MODE> void run(const char* name, int what) {
  long long* out; cudaMalloc(&out, 32); cudaMemset(out, 0, 32);
  cudaFuncSetAttribute(k<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, 180 * 1024);
  const int blocks = 96;
  k<MODE><<<1, 128, 180 * 1024>>>(out, blocks, what);
  k<MODE><<<1, 128, 180 * 1024>>>(out, blocks, what);
  long long h[4]; cudaMemcpy(h, out, 32, cudaMemcpyDeviceToHost);
  const long long tot = h[1] > h[3] ? h[1] : h[3];
  printf("%-52s issue %7.1f  issue+complete %7.1f clk per block%s  (%s)\n", name, (double)h[0] / blocks, (double)tot / blocks,
         MODE == 2 ? " (one group)" : " pair", cudaGetErrorString(cudaGetLastError()));
  cudaFree(out);
}
int main() {
  run<2>("one issuer, one group: S only (4 x SS N96)", 1);
  run<2>("one issuer, one group: PV only (6 x TS N64)", 2);
  run<2>("one issuer, one group: S + PV", 3);
  run<0>("one issuer, both groups round robin: S + PV", 3);
  run<1>("two issuers (warps 1, 3), one group each: S + PV", 3);
  run<1>("two issuers: S only", 1);
  run<1>("two issuers: PV only", 2);
  return 0;
}
