// Micro-benchmark: tcgen05.mma issue/completion time for dependent vs independent accumulators.
#include <cstdio>
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../eventgpt_b200/csrc/sm100_ptx.cuh"
using namespace egpt;
template <int N, int NACC, int TS>
__global__ void k(long long* out, int iters) {
  extern __shared__ uint8_t raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)raw + 1023) & ~(uintptr_t)1023);
  __shared__ uint64_t bar; __shared__ uint32_t tptr;
  const int warp = threadIdx.x >> 5;
  for (int i = threadIdx.x; i < 64 * 1024 / 4; i += blockDim.x) ((uint32_t*)smem)[i] = 0x3c003c00u;
  if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_barrier_init(); }
  if (warp == 1) { ptx::tmem_alloc(&tptr, 512); ptx::tmem_relinquish(); }
  ptx::fence_proxy_async();
  ptx::tcgen05_fence_before(); __syncthreads(); ptx::tcgen05_fence_after();
  const uint32_t tb = tptr;
  if (threadIdx.x == 0) {
    const uint32_t sa = ptx::smem_u32(smem), sb = sa + 16384;
    const uint32_t idesc = ptx::make_idesc_bf16(128, N);
    const uint64_t da = ptx::make_smem_desc_sw128(sa), db = ptx::make_smem_desc_sw128(sb);
    long long t0 = clock64();
    for (int i = 0; i < iters; ++i) {
      const uint32_t d = tb + 256 + (uint32_t)((i % NACC) * 64);
      if (TS) {
        asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\ttcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
                     ::"r"(d), "r"(tb + (uint32_t)((i & 7) * 8)), "l"(db + (uint64_t)((i & 3) * 2)), "r"(idesc), "r"(1u) : "memory");
      } else {
        ptx::umma_bf16_ss(d, da + (uint64_t)((i & 3) * 2), db + (uint64_t)((i & 3) * 2), idesc, 1u);
      }
    }
    long long t1 = clock64();
    ptx::umma_commit(&bar);
    ptx::mbar_wait(&bar, 0);
    long long t2 = clock64();
    out[0] = t1 - t0; out[1] = t2 - t0;
  }
  ptx::tcgen05_fence_before(); __syncthreads(); ptx::tcgen05_fence_after();
  if (warp == 1) ptx::tmem_dealloc(tb, 512);
}
template <int N, int NACC, int TS> void run(const char* name) {
  long long* out; cudaMalloc(&out, 16);
  cudaFuncSetAttribute(k<N, NACC, TS>, cudaFuncAttributeMaxDynamicSharedMemorySize, 80 * 1024);
  const int iters = 256;
  k<N, NACC, TS><<<1, 128, 80 * 1024>>>(out, iters);
  k<N, NACC, TS><<<1, 128, 80 * 1024>>>(out, iters);
  long long h[2]; cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
  printf("%-34s issue %.1f clk/MMA, issue+complete %.1f clk/MMA  (%s)\n", name, (double)h[0] / iters, (double)h[1] / iters,
         cudaGetErrorString(cudaGetLastError()));
  cudaFree(out);
}
int main() {
  run<64, 1, 0>("SS N=64  1 accumulator");
  run<64, 2, 0>("SS N=64  2 accumulators");
  run<64, 4, 0>("SS N=64  4 accumulators");
  run<64, 1, 1>("TS N=64  1 accumulator");
  run<64, 2, 1>("TS N=64  2 accumulators");
  run<64, 4, 1>("TS N=64  4 accumulators");
  run<192, 1, 0>("SS N=192 1 accumulator");
  run<256, 1, 0>("SS N=256 1 accumulator");
  return 0;
}
