// Persistent warp-specialised bf16 GEMM for sm_100a:  C = epilogue(A[M,K] . B[N,K]^T + bias)
//
// Used for every linear layer of the visual-token stage (reference:
// transformers modeling_clip.py CLIPAttention / CLIPMLP called from
// model/EventChatModel.py:196, visual_projector :96-102,199 and feature_adaptor
// :84-85,338).  nn.Linear stores weight as [out, in] = [N, K] K-major, which is
// exactly the B operand layout tcgen05 wants, so weights are used as they are.
//
// Structure (one CTA per SM, 256 threads):
//   warp 0    TMA producer: A (128 x 64) and B (256 x 64) bf16 tiles, SWIZZLE_128B,
//             4-stage mbarrier ring
//   warp 1    MMA issuer: one thread issues tcgen05.mma (M128 N256 K16), fp32
//             accumulators in TMEM, double-buffered (2 x 256 columns)
//   warp 2    TMEM allocator
//   warps 4-7 epilogue: tcgen05.ld 32 lanes x 32 columns -> bias / activation -> 128-byte
//             swizzled staging rows in shared memory -> TMA store (bf16 / fp32), or TMA
//             reduce-add for the fp32 residual stream (h += acc + bias is done in L2;
//             the SM never reads h).  Rows past M are clipped by the TMA unit.
// Tiles are walked n-fastest so the CTAs running concurrently share A rows through L2.
#include "common.cuh"
#include "sm100_ptx.cuh"
#include <mutex>

namespace egpt {

constexpr int BM = 128, BN = 256, BK = 64, UMMA_K = 16;
constexpr int kStages = 4;
constexpr int kABytes = BM * BK * 2;               // 16 KB
constexpr int kBBytes = BN * BK * 2;               // 32 KB
constexpr int kStageBytes = kABytes + kBBytes;     // 48 KB
constexpr int kTmemCols = 2 * BN;                  // two accumulator stages
constexpr int kGemmThreads = 256;
constexpr int kStagingBytes = 4 * 2 * 4096;        // 4 epilogue warps x 2 buffers x (32 rows x 128 B)
constexpr int kGemmSmem = kStages * kStageBytes + kStagingBytes + 1024 /*align slack*/ + 256 /*barriers*/;

enum { EPI_BIAS_BF16 = 0, EPI_QUICKGELU_BF16 = 1, EPI_GELU_BF16 = 2, EPI_RESIDUAL_F32 = 3, EPI_STORE_F32 = 4 };

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

template <int EPI>
__device__ __forceinline__ float activate(float v) {
  if (EPI == EPI_QUICKGELU_BF16) {
    // x * sigmoid(1.702 x) with sigmoid(z) = 0.5 + 0.5 tanh(z/2): one MUFU op, no division
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.851f * v));
    return v * fmaf(0.5f, t, 0.5f);
  }
  if (EPI == EPI_GELU_BF16) return 0.5f * v * (1.0f + erff(v * 0.70710678118654752f));
  return v;
}

template <int EPI>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ CUtensorMap tmap_out, const float* __restrict__ bias, long long M, int N, int K,
                 long long* __restrict__ stats) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* staging = smem + kStages * kStageBytes;                 // 1024-byte aligned
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(staging + kStagingBytes);
  uint64_t* empty_bar = full_bar + kStages;
  uint64_t* tmem_full = empty_bar + kStages;
  uint64_t* tmem_empty = tmem_full + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_tiles = N / BN;
  const long long m_tiles = (M + BM - 1) / BM;
  const long long num_tiles = m_tiles * n_tiles;
  const int num_kb = K / BK;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tmap_a);
    ptx::prefetch_tensormap(&tmap_b);
    ptx::prefetch_tensormap(&tmap_out);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages; ++i) { ptx::mbar_init(full_bar + i, 1); ptx::mbar_init(empty_bar + i, 1); }
    for (int i = 0; i < 2; ++i) { ptx::mbar_init(tmem_full + i, 1); ptx::mbar_init(tmem_empty + i, 128); }
    ptx::fence_barrier_init();
  }
  if (warp == 2) {
    ptx::tmem_alloc(tmem_ptr, kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0 && lane == 0) {
    // ===== TMA producer =====
    int stage = 0; uint32_t phase = 0;
    long long stat_prod_wait = 0;
    for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int m0 = (int)(tile / n_tiles) * BM, n0 = (int)(tile % n_tiles) * BN;
      for (int kb = 0; kb < num_kb; ++kb) {
#ifdef EGPT_GEMM_STATS
        const long long w0 = clock64();
        ptx::mbar_wait(empty_bar + stage, phase ^ 1);
        if (stats) stat_prod_wait += clock64() - w0;
#else
        ptx::mbar_wait(empty_bar + stage, phase ^ 1);
#endif
        uint8_t* sa = smem + stage * kStageBytes;
        ptx::mbar_arrive_expect_tx(full_bar + stage, kStageBytes);
        ptx::tma_load_2d(sa, &tmap_a, full_bar + stage, kb * BK, m0);
        ptx::tma_load_2d(sa + kABytes, &tmap_b, full_bar + stage, kb * BK, n0);
        if (++stage == kStages) { stage = 0; phase ^= 1; }
      }
    }
#ifdef EGPT_GEMM_STATS
    if (stats) stats[blockIdx.x * 8 + 0] = stat_prod_wait;
#endif
  } else if (warp == 1 && lane == 0) {
    // ===== MMA issuer =====
    constexpr uint32_t idesc = ptx::make_idesc_bf16(BM, BN);
    int stage = 0; uint32_t phase = 0;
    int acc = 0; uint32_t acc_phase = 0;
    long long stat_full = 0, stat_tmem = 0; const long long stat_t0 = clock64();
    for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
#ifdef EGPT_GEMM_STATS
      { const long long w0 = clock64(); ptx::mbar_wait(tmem_empty + acc, acc_phase ^ 1); stat_tmem += clock64() - w0; }
#else
      ptx::mbar_wait(tmem_empty + acc, acc_phase ^ 1);
#endif
      ptx::tcgen05_fence_after();
      const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BN);
      for (int kb = 0; kb < num_kb; ++kb) {
#ifdef EGPT_GEMM_STATS
        { const long long w0 = clock64(); ptx::mbar_wait(full_bar + stage, phase); stat_full += clock64() - w0; }
#else
        ptx::mbar_wait(full_bar + stage, phase);
#endif
        ptx::tcgen05_fence_after();
        const uint32_t sa = ptx::smem_u32(smem + stage * kStageBytes);
        const uint64_t da = ptx::make_smem_desc_sw128(sa), db = ptx::make_smem_desc_sw128(sa + kABytes);
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; ++k) {
          // advance 32 B (16 bf16) inside the 128 B swizzle atom: +2 in 16-byte units
          ptx::umma_bf16_ss(tmem_d, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | k) != 0);
        }
        ptx::umma_commit(empty_bar + stage);          // smem slot free once these MMAs retire
        if (++stage == kStages) { stage = 0; phase ^= 1; }
      }
      ptx::umma_commit(tmem_full + acc);              // accumulator ready for the epilogue
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
#ifdef EGPT_GEMM_STATS
    if (stats) { stats[blockIdx.x * 8 + 1] = stat_full; stats[blockIdx.x * 8 + 2] = stat_tmem; stats[blockIdx.x * 8 + 3] = clock64() - stat_t0; }
#else
    (void)stat_full; (void)stat_tmem; (void)stat_t0;
#endif
  } else if (warp >= 4) {
    // ===== epilogue =====
    const int q = warp & 3;                            // TMEM lane quadrant this warp may read
    constexpr bool kF32Out = (EPI == EPI_RESIDUAL_F32 || EPI == EPI_STORE_F32);
    constexpr int CH = kF32Out ? 32 : 64;              // columns per staged chunk (128 B per row)
    uint8_t* my_stage = staging + q * 8192;
    const uint32_t sw = (uint32_t)(lane & 7);          // SWIZZLE_128B: 16-byte chunk index ^= row & 7
    int acc = 0; uint32_t acc_phase = 0; int buf = 0;
    for (long long tile = blockIdx.x; tile < num_tiles; tile += gridDim.x) {
      const int row0 = (int)(tile / n_tiles) * BM + q * 32;
      const int n0 = (int)(tile % n_tiles) * BN;
      ptx::mbar_wait(tmem_full + acc, acc_phase);
      ptx::tcgen05_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN);
#pragma unroll 1
      for (int c0 = 0; c0 < BN; c0 += CH) {
        uint32_t r[CH];
        ptx::tmem_ld_32x32b_x32(taddr + (uint32_t)c0, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
        if (CH == 64) ptx::tmem_ld_32x32b_x32(taddr + (uint32_t)c0 + 32u, *reinterpret_cast<uint32_t(*)[32]>(&r[CH - 32]));
        ptx::tmem_ld_wait();
        if (c0 + CH == BN) {                           // accumulator fully read: hand it back to the MMA warp
          ptx::tcgen05_fence_before();
          ptx::mbar_arrive(tmem_empty + acc);
        }
        // the TMA store that last used this staging buffer must have finished reading it
        if (lane == 0) ptx::bulk_wait_group_read<1>();
        __syncwarp();
        uint8_t* dst = my_stage + buf * 4096 + lane * 128;
        if (bias != nullptr) {
          const float4* bp = reinterpret_cast<const float4*>(bias + n0 + c0);
#pragma unroll
          for (int j = 0; j < CH / 4; ++j) {
            const float4 b = __ldg(bp + j);
            r[4 * j] = __float_as_uint(__uint_as_float(r[4 * j]) + b.x);
            r[4 * j + 1] = __float_as_uint(__uint_as_float(r[4 * j + 1]) + b.y);
            r[4 * j + 2] = __float_as_uint(__uint_as_float(r[4 * j + 2]) + b.z);
            r[4 * j + 3] = __float_as_uint(__uint_as_float(r[4 * j + 3]) + b.w);
          }
        }
        if (kF32Out) {
#pragma unroll
          for (int j = 0; j < 8; ++j)
            *reinterpret_cast<uint4*>(dst + (((uint32_t)j ^ sw) << 4)) = make_uint4(r[4 * j], r[4 * j + 1], r[4 * j + 2], r[4 * j + 3]);
        } else {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            uint4 o;
            o.x = pack_bf16(activate<EPI>(__uint_as_float(r[8 * j + 0])), activate<EPI>(__uint_as_float(r[8 * j + 1])));
            o.y = pack_bf16(activate<EPI>(__uint_as_float(r[8 * j + 2])), activate<EPI>(__uint_as_float(r[8 * j + 3])));
            o.z = pack_bf16(activate<EPI>(__uint_as_float(r[8 * j + 4])), activate<EPI>(__uint_as_float(r[8 * j + 5])));
            o.w = pack_bf16(activate<EPI>(__uint_as_float(r[8 * j + 6])), activate<EPI>(__uint_as_float(r[8 * j + 7])));
            *reinterpret_cast<uint4*>(dst + (((uint32_t)j ^ sw) << 4)) = o;
          }
        }
        ptx::fence_proxy_async();                      // generic-proxy smem writes -> visible to the TMA unit
        __syncwarp();
        if (lane == 0) {
          if (EPI == EPI_RESIDUAL_F32) ptx::tma_reduce_add_2d(&tmap_out, my_stage + buf * 4096, n0 + c0, row0);
          else ptx::tma_store_2d(&tmap_out, my_stage + buf * 4096, n0 + c0, row0);
          ptx::bulk_commit_group();
        }
        buf ^= 1;
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (lane == 0) ptx::bulk_wait_group<0>();          // all stores / reductions complete before exit
  }

  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  if (warp == 2) ptx::tmem_dealloc(tmem_base, kTmemCols);
}

// ---- host side -----------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// 2-D row-major [rows, cols] tensor, box = {box_cols, box_rows} with box_cols * elem = 128 B, 128-byte swizzle
static int make_tmap(CUtensorMap* map, const void* ptr, bool f32, uint64_t rows, uint64_t cols, uint32_t box_cols,
                     uint32_t box_rows) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled entry point not available"); return EGPT_ERR_CUDA; }
  const uint64_t esz = f32 ? 4 : 2;
  cuuint64_t dims[2] = {cols, rows};
  cuuint64_t strides[1] = {cols * esz};
  cuuint32_t box[2] = {box_cols, box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, f32 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32 : CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2,
                  const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("cuTensorMapEncodeTiled failed with CUresult %d", (int)r); return EGPT_ERR_CUDA; }
  return EGPT_OK;
}

template <int EPI>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to, const float* bias,
                       long long M, int N, int K, cudaStream_t stream, long long* stats = nullptr) {
  static std::once_flag once;   // attribute is per function; one process drives one device
  cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [&] {
    attr_err = cudaFuncSetAttribute(gemm_bf16_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmem);
  });
  EGPT_CUDA(attr_err);
  const long long tiles = ((M + BM - 1) / BM) * (N / BN);
  const int grid = (int)(tiles < kNumSMs ? tiles : kNumSMs);
  gemm_bf16_kernel<EPI><<<grid, kGemmThreads, kGemmSmem, stream>>>(ta, tb, to, bias, M, N, K, stats);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}

int gemm_bf16_stats(const void* A, const void* B, const float* bias, void* out, long long M, int N, int K, int epilogue,
                    cudaStream_t stream, long long* stats);
int gemm_bf16(const void* A, const void* B, const float* bias, void* out, long long M, int N, int K, int epilogue,
              cudaStream_t stream) {
  return gemm_bf16_stats(A, B, bias, out, M, N, K, epilogue, stream, nullptr);
}
int gemm_bf16_stats(const void* A, const void* B, const float* bias, void* out, long long M, int N, int K, int epilogue,
                    cudaStream_t stream, long long* stats) {
  EGPT_CHECK_ARG(A && B && out, "gemm: null pointer");
  EGPT_CHECK_ARG(M > 0 && N > 0 && K > 0 && N % BN == 0 && K % BK == 0, "gemm: unsupported shape M=%lld N=%d K=%d (N %% 256, K %% 64)",
                 M, N, K);
  EGPT_CHECK_ARG(((uintptr_t)A % 16 == 0) && ((uintptr_t)B % 16 == 0) && ((uintptr_t)out % 16 == 0) &&
                 (bias == nullptr || (uintptr_t)bias % 16 == 0), "gemm: pointers must be 16-byte aligned");
  CUtensorMap ta, tb;
  ProfScope prof(PROF_GEMM, 2.0 * (double)M * N * K, stream);
  CUtensorMap to;
  const bool f32_out = epilogue == EPI_RESIDUAL_F32 || epilogue == EPI_STORE_F32;
  EGPT_TRY(make_tmap(&ta, A, false, (uint64_t)M, (uint64_t)K, BK, BM));
  EGPT_TRY(make_tmap(&tb, B, false, (uint64_t)N, (uint64_t)K, BK, BN));
  EGPT_TRY(make_tmap(&to, out, f32_out, (uint64_t)M, (uint64_t)N, f32_out ? 32 : 64, 32));
  switch (epilogue) {
    case EPI_BIAS_BF16: return launch_gemm<EPI_BIAS_BF16>(ta, tb, to, bias, M, N, K, stream, stats);
    case EPI_QUICKGELU_BF16: return launch_gemm<EPI_QUICKGELU_BF16>(ta, tb, to, bias, M, N, K, stream, stats);
    case EPI_GELU_BF16: return launch_gemm<EPI_GELU_BF16>(ta, tb, to, bias, M, N, K, stream, stats);
    case EPI_RESIDUAL_F32: return launch_gemm<EPI_RESIDUAL_F32>(ta, tb, to, bias, M, N, K, stream, stats);
    case EPI_STORE_F32: return launch_gemm<EPI_STORE_F32>(ta, tb, to, bias, M, N, K, stream, stats);
  }
  set_error("gemm: unknown epilogue %d", epilogue);
  return EGPT_ERR_INVALID;
}

}  // namespace egpt

extern "C" int egpt_gemm_bf16(const uint16_t* A, const uint16_t* B, const float* bias, void* out, int64_t M, int N,
                              int K, int epilogue, void* stream) {
  return egpt::gemm_bf16(A, B, bias, out, M, N, K, epilogue, (cudaStream_t)stream);
}

#ifdef EGPT_GEMM_STATS
// developer aid: per-CTA cycles {producer waiting for a free stage, MMA waiting for data, MMA waiting for TMEM, MMA total}
extern "C" int egpt_gemm_stats(const uint16_t* A, const uint16_t* B, const float* bias, void* out, int64_t M, int N, int K,
                               int epilogue, long long* stats, void* stream) {
  return egpt::gemm_bf16_stats(A, B, bias, out, M, N, K, epilogue, (cudaStream_t)stream, stats);
}
#endif
