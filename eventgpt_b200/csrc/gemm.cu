// Persistent warp-specialised bf16 GEMM for sm_100a:  C = epilogue(A[M,K] . B[N,K]^T + bias)
//
// Used for every linear layer of the visual-token stage (reference:
// transformers modeling_clip.py CLIPAttention / CLIPMLP called from
// model/EventChatModel.py:196, visual_projector :96-102,199 and feature_adaptor
// :84-85,338).  nn.Linear stores weight as [out, in] = [N, K] K-major, which is
// exactly the B operand layout tcgen05 wants, so weights are used as they are.
//
// Structure: persistent, one CTA per SM, CTAs paired in clusters of 2 (tcgen05 cta_group::2), 256 or 384 threads:
//   warp 0    TMA producer (both CTAs): own 128 x 64 slice of A and HALF (128 rows) of the 256 x 64 B tile,
//             SWIZZLE_128B, 6-stage mbarrier ring (32 KB per stage per SM); completion bytes of both CTAs
//             are credited to the leader CTA's "full" barrier
//   warp 1    MMA issuer (leader CTA only): one thread issues tcgen05.mma.cta_group::2 (M256 N256 K16),
//             fp32 accumulators in both CTAs' TMEM, double-buffered (2 x 256 columns); commits are
//             multicast to both CTAs ("slot free", "accumulator ready")
//   warp 2    TMEM allocator
//   warps 4.. epilogue (both CTAs, own 128 rows): tcgen05.ld 32 lanes x 32 columns -> bias / activation / folded
//             LayerNorm -> 128-byte swizzled staging rows in shared memory -> TMA store.  Eight warps (two per TMEM
//             lane quarter, half the tile's columns each) for the bf16-output epilogues, four for the fp32-output ones:
//             plain store, TMA reduce-add (h += acc + bias done in L2), or the residual read-add-store that also emits
//             the bf16 copy and the LayerNorm partial sums (see the epilogue enum).  Rows past M are clipped by the TMA unit.
// Tiles are walked n-fastest so the CTA pairs running concurrently share A rows through L2.
// Measured (profiles/r01_bench_history.md): a single-CTA 128x256 version with 4 x 48 KB stages spent 17-33 % of
// the MMA thread's time waiting for operand data; pairing CTAs cut the L2 -> SM operand traffic by a third and
// made room for six stages.
#include "common.cuh"
#include "sm100_ptx.cuh"
#include <mutex>

namespace egpt {

constexpr int BM = 128, BN = 256, BK = 64, UMMA_K = 16;
constexpr int kABytes = BM * BK * 2;               // 16 KB
constexpr int kTmemCols = 2 * BN;                  // two accumulator stages

enum { EPI_BIAS_BF16 = 0, EPI_QUICKGELU_BF16 = 1, EPI_GELU_BF16 = 2, EPI_RESIDUAL_F32 = 3, EPI_STORE_F32 = 4,
       // LayerNorm folded into the two GEMMs around it (the separate LayerNorm pass over the residual stream disappears):
       // 5: producer of the residual stream: h = h + acc + bias read-add-stored through shared memory, a bf16 copy of the new
       //    h (the A operand of the next GEMM) and the per-row sum / sum of squares of this tile's 256 columns;
       // 6, 7: consumer of LayerNorm(h): A is the bf16 copy of h itself, B holds gamma (.) W, and the epilogue applies
       //    rstd * (acc - mean * colsum) + bias' per row (mean, rstd from the four column tiles' partial sums);  7 adds QuickGELU
       EPI_RESIDUAL_LN = 5, EPI_LN_BIAS_BF16 = 6, EPI_LN_QUICKGELU_BF16 = 7,
       // 5 for a short K (out-proj, K = 1024: the epilogue is the critical path there): four staging tiles per warp so that a
       // load never waits for the newest store, at the price of an operand stage; chosen by gemm_bf16, not by callers
       EPI_RESIDUAL_LN_SHORTK = 8 };
constexpr float kLnEpsG = 1e-5f;

__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

template <int EPI>
__device__ __forceinline__ float activate(float v) {
  if (EPI == EPI_QUICKGELU_BF16 || EPI == EPI_LN_QUICKGELU_BF16) {
    // x * sigmoid(1.702 x) with sigmoid(z) = 0.5 + 0.5 tanh(z/2): one MUFU op, no division
    float t;
    asm("tanh.approx.f32 %0, %1;" : "=f"(t) : "f"(0.851f * v));
    return v * fmaf(0.5f, t, 0.5f);
  }
  if (EPI == EPI_GELU_BF16) return 0.5f * v * (1.0f + erff(v * 0.70710678118654752f));
  return v;
}

// ---- 2-CTA variant ------------------------------------------------------------------------------
// A CTA pair (cluster of 2) computes a 256 x 256 tile with tcgen05.mma.cta_group::2: each CTA loads its own
// 128 x 64 slice of A and HALF (128 rows) of the 256 x 64 B tile, so a stage costs 32 KB per SM instead of
// 48 KB -> six stages fit where four did, and the L2 -> SM operand traffic drops by a third.  The leader
// CTA issues the MMAs; full barriers live in the leader (both CTAs' TMA bytes are credited to them), empty /
// accumulator-ready barriers are multicast to both CTAs; each CTA drains its own 128 accumulator rows.
constexpr int kBHalfBytes = (BN / 2) * BK * 2;          // 16 KB
constexpr int kStageBytes2 = kABytes + kBHalfBytes;     // 32 KB per CTA
// the residual + LayerNorm-statistics epilogue stages a bf16 copy next to the fp32 tile: one operand stage less
// the residual + LayerNorm-statistics epilogue keeps three fp32 staging tiles per warp in flight (the old residual tile is
// fetched two chunks ahead) and stages a bf16 copy beside them: one operand stage less
// The bf16-output epilogues (bias, activations, folded LayerNorm) run on EIGHT warps: two per TMEM lane quarter, each taking
// half of the tile's 256 columns.  One warp per scheduler left fc1's QuickGELU + LayerNorm epilogue latency-bound (tensor
// pipe 82 % active); the fp32-output epilogues are bound by memory or by the MMA and keep four warps.
constexpr int epi_warps_of(int epi) {
  return (epi == EPI_BIAS_BF16 || epi == EPI_QUICKGELU_BF16 || epi == EPI_GELU_BF16 || epi == EPI_LN_BIAS_BF16 ||
          epi == EPI_LN_QUICKGELU_BF16) ? 8 : 4;
}
constexpr int threads_of(int epi) { return 128 + 32 * epi_warps_of(epi); }
constexpr bool is_producer(int epi) { return epi == EPI_RESIDUAL_LN || epi == EPI_RESIDUAL_LN_SHORTK; }
constexpr int stages_of(int epi) { return epi == EPI_RESIDUAL_LN ? 5 : (epi == EPI_RESIDUAL_LN_SHORTK ? 4 : 6); }
constexpr int fbufs_of(int epi) {            // staging tiles (4 KB) per epilogue warp
  return epi == EPI_RESIDUAL_LN ? 3 : (epi == EPI_RESIDUAL_LN_SHORTK ? 4 : (epi_warps_of(epi) == 8 ? 1 : 2));
}
constexpr int bbufs_of(int epi) { return epi == EPI_RESIDUAL_LN ? 1 : (epi == EPI_RESIDUAL_LN_SHORTK ? 2 : 0); }   // bf16 staging tiles per warp
constexpr int staging_of(int epi) { return epi_warps_of(epi) * fbufs_of(epi) * 4096; }
constexpr int extra_staging_of(int epi) { return 4 * bbufs_of(epi) * 4096; }
constexpr int kVecBytes = 2 * BN * 4;                     // bias and column sums of the tile's 256 columns
constexpr int gemm_smem_of(int epi) { return stages_of(epi) * kStageBytes2 + staging_of(epi) + extra_staging_of(epi) + kVecBytes + 512; }

template <int EPI>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(threads_of(EPI), 1)
gemm2_bf16_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                  const __grid_constant__ CUtensorMap tmap_out, const __grid_constant__ CUtensorMap tmap_out2,
                  const float* __restrict__ bias, const float* __restrict__ colsum, float* __restrict__ stats,
                  long long M, int N, int K) {
  constexpr int kStages2 = stages_of(EPI);
  extern __shared__ __align__(1024) uint8_t smem[];                      // SWIZZLE_128B tiles want 1024-byte alignment
  constexpr int NBUF = fbufs_of(EPI);
  constexpr bool kProducer = is_producer(EPI);
  constexpr int kNewest = EPI == EPI_RESIDUAL_LN_SHORTK ? 1 : 0;        // stores that may still be reading when a load is issued
  uint8_t* staging = smem + kStages2 * kStageBytes2;
  uint8_t* staging2 = staging + staging_of(EPI);                         // bf16 copy (EPI_RESIDUAL_LN only)
  float* vec_s = reinterpret_cast<float*>(staging2 + extra_staging_of(EPI));   // [2][256]: bias, column sums of this tile
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(reinterpret_cast<uint8_t*>(vec_s) + kVecBytes);
  uint64_t* empty_bar = full_bar + kStages2;
  uint64_t* tmem_full = empty_bar + kStages2;
  uint64_t* tmem_empty = tmem_full + 2;
  uint64_t* ld_bar = tmem_empty + 2;                                     // [4 warps][4 buffers]: residual tile loads
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(ld_bar + 16);

  // warp index and CTA rank are made warp-uniform by construction: the producer and MMA warps then run their loops with
  // uniform control flow and operands (descriptors, coordinates and barrier addresses in uniform registers) and only the
  // TMA / MMA instructions themselves sit under elect.sync
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const uint32_t rank = __shfl_sync(0xffffffffu, ptx::cluster_ctarank(), 0);
  const bool leader = rank == 0;
  const int n_tiles = N / BN;
  const long long m_pairs = (M + 2 * BM - 1) / (2 * BM);
  const long long num_tiles = m_pairs * n_tiles;           // pair tiles
  const int num_kb = K / BK;
  const long long first = blockIdx.x >> 1, step = gridDim.x >> 1;

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tensormap(&tmap_a);
    ptx::prefetch_tensormap(&tmap_b);
    ptx::prefetch_tensormap(&tmap_out);
    if (kProducer) ptx::prefetch_tensormap(&tmap_out2);
  }
  if (warp == 1 && lane == 0) {
    for (int i = 0; i < kStages2; ++i) { ptx::mbar_init(full_bar + i, 1); ptx::mbar_init(empty_bar + i, 1); }
    for (int i = 0; i < 2; ++i) { ptx::mbar_init(tmem_full + i, 1); ptx::mbar_init(tmem_empty + i, 2 * 32 * epi_warps_of(EPI)); }
    for (int i = 0; i < 16; ++i) ptx::mbar_init(ld_bar + i, 1);
    ptx::fence_barrier_init();
  }
  if (warp == 2) {
    ptx::tmem_alloc_2sm(tmem_ptr, kTmemCols);
    ptx::tmem_relinquish_2sm();
  }
  ptx::tcgen05_fence_before();
  ptx::cluster_sync_all();                                  // barriers of both CTAs initialised before anyone signals them
  ptx::tcgen05_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, *tmem_ptr, 0);

  if (warp == 0) {
    // ===== TMA producer (both CTAs) =====
    int stage = 0; uint32_t phase = 0;
    for (long long tile = first; tile < num_tiles; tile += step) {
      const int m0 = (int)(tile / n_tiles) * (2 * BM) + (int)rank * BM;
      const int n0 = (int)(tile % n_tiles) * BN + (int)rank * (BN / 2);
      for (int kb = 0; kb < num_kb; ++kb) {
        ptx::mbar_wait(empty_bar + stage, phase ^ 1);
        uint8_t* sa = smem + stage * kStageBytes2;
        if (ptx::elect_one()) {
          if (leader) ptx::mbar_arrive_expect_tx(full_bar + stage, 2 * kStageBytes2);   // bytes of both CTAs
          ptx::tma_load_2d_2sm(sa, &tmap_a, full_bar + stage, kb * BK, m0);
          ptx::tma_load_2d_2sm(sa + kABytes, &tmap_b, full_bar + stage, kb * BK, n0);
        }
        __syncwarp();
        if (++stage == kStages2) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1 && leader) {
    // ===== MMA issuer (leader CTA only) =====
    constexpr uint32_t idesc = ptx::make_idesc_bf16(2 * BM, BN);
    int stage = 0; uint32_t phase = 0;
    int acc = 0; uint32_t acc_phase = 0;
    for (long long tile = first; tile < num_tiles; tile += step) {
      ptx::mbar_wait(tmem_empty + acc, acc_phase ^ 1);        // both CTAs' epilogues have drained this stage
      ptx::tcgen05_fence_after();
      const uint32_t tmem_d = tmem_base + (uint32_t)(acc * BN);
      for (int kb = 0; kb < num_kb; ++kb) {
        ptx::mbar_wait(full_bar + stage, phase);
        ptx::tcgen05_fence_after();
        const uint32_t sa = ptx::smem_u32(smem + stage * kStageBytes2);
        const uint64_t da = ptx::make_smem_desc_sw128(sa), db = ptx::make_smem_desc_sw128(sa + kABytes);
        if (ptx::elect_one()) {
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k)
            ptx::umma_bf16_ss_2sm(tmem_d, da + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, (kb | k) != 0);
          ptx::umma_commit_2sm(empty_bar + stage);            // both producers may refill this slot
          if (kb == num_kb - 1) ptx::umma_commit_2sm(tmem_full + acc);   // both epilogues may read their accumulator half
        }
        __syncwarp();
        if (++stage == kStages2) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else if (warp >= 4) {
    // ===== epilogue (both CTAs, own 128 rows) =====
    constexpr int kEpiWarps = epi_warps_of(EPI);
    constexpr int kSpan = BN / (kEpiWarps / 4);                 // columns of the tile this warp drains
    const int q = warp & 3;                                     // TMEM lane quarter (rows q*32 .. q*32+31 of the CTA's 128)
    const int ew = warp - 4;                                    // epilogue warp 0 .. kEpiWarps-1
    const int col_lo = (ew >> 2) * kSpan;
    constexpr bool kF32Out = (EPI == EPI_RESIDUAL_F32 || EPI == EPI_STORE_F32 || kProducer);
    constexpr bool kLnIn = (EPI == EPI_LN_BIAS_BF16 || EPI == EPI_LN_QUICKGELU_BF16);
    constexpr int CH = kF32Out ? 32 : 64;
    uint8_t* my_stage = staging + ew * (NBUF * 4096);
    uint8_t* my_stage2 = staging2 + q * (bbufs_of(EPI) * 4096);
    uint64_t* my_ld = ld_bar + q * 4;
    uint32_t ld_phase = 0;                                   // bit b: phase of buffer b's load barrier
    const uint32_t sw = (uint32_t)(lane & 7);
    const uint32_t leader_tmem_empty = ptx::mapa_u32(tmem_empty, 0);
    int acc = 0; uint32_t acc_phase = 0; int buf = 0;
    for (long long tile = first; tile < num_tiles; tile += step) {
      const int row0 = (int)(tile / n_tiles) * (2 * BM) + (int)rank * BM + q * 32;
      const int n0 = (int)(tile % n_tiles) * BN;
      float mean = 0.f, rstd = 0.f, s1 = 0.f, s2 = 0.f;
      if (kLnIn) {
        // LayerNorm statistics of this lane's row from the partial sums the producing GEMM left (K = row length)
        const long long r = (long long)row0 + lane;
        if (r < M) {
          const float4 a = __ldg(reinterpret_cast<const float4*>(stats + r * 8)), b = __ldg(reinterpret_cast<const float4*>(stats + r * 8) + 1);
          const float inv = 1.0f / (float)K;
          mean = ((a.x + a.z) + (b.x + b.z)) * inv;
          const float var = ((a.y + a.w) + (b.y + b.w)) * inv - mean * mean;
          rstd = rsqrtf(fmaxf(var, 0.f) + kLnEpsG);
        }
      }
      if (kProducer) {
        // the old residual tile travels into the staging buffers the new one will leave from, two chunks ahead
        if (lane == 0) {
          ptx::bulk_wait_group_read<kNewest>();
#pragma unroll
          for (int a = 0; a < 2; ++a) {
            const int b = (buf + a) % NBUF;
            ptx::mbar_arrive_expect_tx(my_ld + b, 4096);
            ptx::tma_load_2d(my_stage + b * 4096, &tmap_out, my_ld + b, n0 + a * CH, row0);
          }
        }
        __syncwarp();
      }
      if (bias != nullptr) {
        // bias (and column sums) of the tile's 256 columns: once into shared memory -- with the whole L1 carved out for
        // the operand ring, per-chunk global loads would each pay an L2 round trip
        asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpiWarps) : "memory");       // the previous tile's readers are done
        const int e = ew * 32 + lane;
        if (e < 64) reinterpret_cast<float4*>(vec_s)[e] = __ldg(reinterpret_cast<const float4*>(bias + n0) + e);
        else if (kLnIn && e < 128) reinterpret_cast<float4*>(vec_s)[e] = __ldg(reinterpret_cast<const float4*>(colsum + n0) + (e - 64));
        asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpiWarps) : "memory");
      }
      ptx::mbar_wait(tmem_full + acc, acc_phase);
      ptx::tcgen05_fence_after();
      const uint32_t taddr = tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(acc * BN);
#pragma unroll 1
      for (int c0 = col_lo; c0 < col_lo + kSpan; c0 += CH) {
        uint32_t r[CH];
        ptx::tmem_ld_32x32b_x32(taddr + (uint32_t)c0, *reinterpret_cast<uint32_t(*)[32]>(&r[0]));
        if (CH == 64) ptx::tmem_ld_32x32b_x32(taddr + (uint32_t)c0 + 32u, *reinterpret_cast<uint32_t(*)[32]>(&r[CH - 32]));
        ptx::tmem_ld_wait();
        if (c0 + CH == col_lo + kSpan) {
          ptx::tcgen05_fence_before();
          ptx::mbar_arrive_cluster(leader_tmem_empty + (uint32_t)(acc * 8));
        }
        if (kProducer) {
          // the residual tile two chunks ahead goes into a buffer whose store has been read out (ring of three: the previous
          // chunk's, so the newest store is waited for; ring of four: the one before, and the newest store stays in flight);
          // the same wait frees the bf16 tile about to be rewritten.  Then wait for this chunk's tile
          if (lane == 0) {
            ptx::bulk_wait_group_read<kNewest>();
            if (c0 + 2 * CH < BN) {
              const int b = (buf + 2) % NBUF;
              ptx::mbar_arrive_expect_tx(my_ld + b, 4096);
              ptx::tma_load_2d(my_stage + b * 4096, &tmap_out, my_ld + b, n0 + c0 + 2 * CH, row0);
            }
          }
          __syncwarp();
          ptx::mbar_wait(my_ld + buf, (ld_phase >> buf) & 1u);
          ld_phase ^= 1u << buf;
        } else {
          if (lane == 0) ptx::bulk_wait_group_read<NBUF - 1>();      // the store that last read this staging tile is done with it
          __syncwarp();
        }
        uint8_t* dst = my_stage + buf * 4096 + lane * 128;
        if (kLnIn) {
          const float4* cp = reinterpret_cast<const float4*>(vec_s + BN + c0);
          const float4* bp = reinterpret_cast<const float4*>(vec_s + c0);
          // rstd * acc + (bias' - rstd * mean * colsum)
          const float nmr = -mean * rstd;
#pragma unroll
          for (int j = 0; j < CH / 4; ++j) {
            const float4 c = cp[j], b = bp[j];
            r[4 * j] = __float_as_uint(fmaf(rstd, __uint_as_float(r[4 * j]), fmaf(nmr, c.x, b.x)));
            r[4 * j + 1] = __float_as_uint(fmaf(rstd, __uint_as_float(r[4 * j + 1]), fmaf(nmr, c.y, b.y)));
            r[4 * j + 2] = __float_as_uint(fmaf(rstd, __uint_as_float(r[4 * j + 2]), fmaf(nmr, c.z, b.z)));
            r[4 * j + 3] = __float_as_uint(fmaf(rstd, __uint_as_float(r[4 * j + 3]), fmaf(nmr, c.w, b.w)));
          }
        } else if (bias != nullptr) {
          const float4* bp = reinterpret_cast<const float4*>(vec_s + c0);
#pragma unroll
          for (int j = 0; j < CH / 4; ++j) {
            const float4 b = bp[j];
            r[4 * j] = __float_as_uint(__uint_as_float(r[4 * j]) + b.x);
            r[4 * j + 1] = __float_as_uint(__uint_as_float(r[4 * j + 1]) + b.y);
            r[4 * j + 2] = __float_as_uint(__uint_as_float(r[4 * j + 2]) + b.z);
            r[4 * j + 3] = __float_as_uint(__uint_as_float(r[4 * j + 3]) + b.w);
          }
        }
        if (kProducer) {
          // h_new = h_old + acc + bias in place in the staging tile; bf16 copy beside it; running row statistics
          uint8_t* dst2 = my_stage2 + (bbufs_of(EPI) == 2 ? ((c0 >> 6) & 1) * 4096 : 0) + lane * 128;
          const uint32_t half = (uint32_t)((c0 >> 5) & 1) * 4u;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            uint4* cell = reinterpret_cast<uint4*>(dst + (((uint32_t)j ^ sw) << 4));
            const uint4 o = *cell;
            const float v0 = __uint_as_float(o.x) + __uint_as_float(r[4 * j]), v1 = __uint_as_float(o.y) + __uint_as_float(r[4 * j + 1]);
            const float v2 = __uint_as_float(o.z) + __uint_as_float(r[4 * j + 2]), v3 = __uint_as_float(o.w) + __uint_as_float(r[4 * j + 3]);
            s1 += (v0 + v1) + (v2 + v3);
            s2 = fmaf(v0, v0, fmaf(v1, v1, fmaf(v2, v2, fmaf(v3, v3, s2))));
            *cell = make_uint4(__float_as_uint(v0), __float_as_uint(v1), __float_as_uint(v2), __float_as_uint(v3));
            r[4 * j] = __float_as_uint(v0); r[4 * j + 1] = __float_as_uint(v1); r[4 * j + 2] = __float_as_uint(v2); r[4 * j + 3] = __float_as_uint(v3);
          }
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint4 o;
            o.x = pack_bf16(__uint_as_float(r[8 * j + 0]), __uint_as_float(r[8 * j + 1]));
            o.y = pack_bf16(__uint_as_float(r[8 * j + 2]), __uint_as_float(r[8 * j + 3]));
            o.z = pack_bf16(__uint_as_float(r[8 * j + 4]), __uint_as_float(r[8 * j + 5]));
            o.w = pack_bf16(__uint_as_float(r[8 * j + 6]), __uint_as_float(r[8 * j + 7]));
            *reinterpret_cast<uint4*>(dst2 + (((half + (uint32_t)j) ^ sw) << 4)) = o;
          }
        } else if (kF32Out) {
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
        ptx::fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          if (EPI == EPI_RESIDUAL_F32) ptx::tma_reduce_add_2d(&tmap_out, my_stage + buf * 4096, n0 + c0, row0);
          else ptx::tma_store_2d(&tmap_out, my_stage + buf * 4096, n0 + c0, row0);
          if (kProducer && (c0 & 32))                        // two 32-column chunks make one 64-column bf16 tile
            ptx::tma_store_2d(&tmap_out2, my_stage2 + (bbufs_of(EPI) == 2 ? ((c0 >> 6) & 1) * 4096 : 0), n0 + c0 - 32, row0);
          ptx::bulk_commit_group();
        }
        buf = buf + 1 == NBUF ? 0 : buf + 1;
      }
      if (kProducer) {
        const long long r = (long long)row0 + lane;
        if (r < M) *reinterpret_cast<float2*>(stats + r * 8 + 2 * (n0 / BN)) = make_float2(s1, s2);
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (lane == 0) ptx::bulk_wait_group<0>();
  }

  ptx::tcgen05_fence_before();
  ptx::cluster_sync_all();                                   // nobody frees TMEM or exits while the peer still uses it
  ptx::tcgen05_fence_after();
  if (warp == 2) ptx::tmem_dealloc_2sm(tmem_base, kTmemCols);
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
static int launch_gemm2(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to, const CUtensorMap& to2,
                        const float* bias, const float* colsum, float* stats, long long M, int N, int K, cudaStream_t stream) {
  // per launch: the attribute is per device, and one process may drive several GPUs
  constexpr int smem = gemm_smem_of(EPI);
  EGPT_CUDA(cudaFuncSetAttribute(gemm2_bf16_kernel<EPI>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
  const long long pairs = ((M + 2 * BM - 1) / (2 * BM)) * (N / BN);
  const int grid = (int)(2 * (pairs < kNumSMs / 2 ? pairs : kNumSMs / 2));
  gemm2_bf16_kernel<EPI><<<grid, threads_of(EPI), smem, stream>>>(ta, tb, to, to2, bias, colsum, stats, M, N, K);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}

// out2 / colsum / stats: only for the LayerNorm-folding epilogues (see the enum)
int gemm_bf16(const void* A, const void* B, const float* bias, void* out, long long M, int N, int K, int epilogue,
              cudaStream_t stream, void* out2, const float* colsum, float* stats) {
  EGPT_CHECK_ARG(A && B && out, "gemm: null pointer");
  EGPT_CHECK_ARG(M > 0 && N > 0 && K > 0 && N % BN == 0 && K % BK == 0, "gemm: unsupported shape M=%lld N=%d K=%d (N %% 256, K %% 64)",
                 M, N, K);
  EGPT_CHECK_ARG(((uintptr_t)A % 16 == 0) && ((uintptr_t)B % 16 == 0) && ((uintptr_t)out % 16 == 0) &&
                 (bias == nullptr || (uintptr_t)bias % 16 == 0), "gemm: pointers must be 16-byte aligned");
  const bool producer = epilogue == EPI_RESIDUAL_LN, consumer = epilogue == EPI_LN_BIAS_BF16 || epilogue == EPI_LN_QUICKGELU_BF16;
  EGPT_CHECK_ARG(!producer || (N == 4 * BN && out2 && stats && bias && (uintptr_t)out2 % 16 == 0 && (uintptr_t)stats % 32 == 0),
                 "gemm: the residual + statistics epilogue wants N = 1024, a bf16 copy buffer, a statistics buffer and a bias");
  EGPT_CHECK_ARG(!consumer || (K == 4 * BN && colsum && stats && bias && (uintptr_t)colsum % 16 == 0 && (uintptr_t)stats % 32 == 0),
                 "gemm: the LayerNorm-folding epilogue wants K = 1024, column sums, a statistics buffer and a bias");
  CUtensorMap ta, tb, to, to2;
  ProfScope prof(PROF_GEMM, 2.0 * (double)M * N * K, stream);
  const bool f32_out = epilogue == EPI_RESIDUAL_F32 || epilogue == EPI_STORE_F32 || producer;
  EGPT_TRY(make_tmap(&ta, A, false, (uint64_t)M, (uint64_t)K, BK, BM));
  EGPT_TRY(make_tmap(&tb, B, false, (uint64_t)N, (uint64_t)K, BK, BN / 2));          // each CTA of a pair loads half of B
  EGPT_TRY(make_tmap(&to, out, f32_out, (uint64_t)M, (uint64_t)N, f32_out ? 32 : 64, 32));
  if (producer) EGPT_TRY(make_tmap(&to2, out2, false, (uint64_t)M, (uint64_t)N, 64, 32));
  else to2 = to;
  switch (epilogue) {
    case EPI_BIAS_BF16: return launch_gemm2<EPI_BIAS_BF16>(ta, tb, to, to2, bias, colsum, stats, M, N, K, stream);
    case EPI_QUICKGELU_BF16: return launch_gemm2<EPI_QUICKGELU_BF16>(ta, tb, to, to2, bias, colsum, stats, M, N, K, stream);
    case EPI_GELU_BF16: return launch_gemm2<EPI_GELU_BF16>(ta, tb, to, to2, bias, colsum, stats, M, N, K, stream);
    case EPI_RESIDUAL_F32: return launch_gemm2<EPI_RESIDUAL_F32>(ta, tb, to, to2, bias, colsum, stats, M, N, K, stream);
    case EPI_STORE_F32: return launch_gemm2<EPI_STORE_F32>(ta, tb, to, to2, bias, colsum, stats, M, N, K, stream);
    case EPI_RESIDUAL_LN:
      if (K <= 2048) return launch_gemm2<EPI_RESIDUAL_LN_SHORTK>(ta, tb, to, to2, bias, colsum, stats, M, N, K, stream);
      return launch_gemm2<EPI_RESIDUAL_LN>(ta, tb, to, to2, bias, colsum, stats, M, N, K, stream);
    case EPI_LN_BIAS_BF16: return launch_gemm2<EPI_LN_BIAS_BF16>(ta, tb, to, to2, bias, colsum, stats, M, N, K, stream);
    case EPI_LN_QUICKGELU_BF16: return launch_gemm2<EPI_LN_QUICKGELU_BF16>(ta, tb, to, to2, bias, colsum, stats, M, N, K, stream);
  }
  set_error("gemm: unknown epilogue %d", epilogue);
  return EGPT_ERR_INVALID;
}

}  // namespace egpt

extern "C" int egpt_gemm_bf16(const uint16_t* A, const uint16_t* B, const float* bias, void* out, int64_t M, int N,
                              int K, int epilogue, void* stream) {
  EGPT_CHECK_ARG(epilogue >= 0 && epilogue <= egpt::EPI_STORE_F32, "egpt_gemm_bf16: epilogue %d (0..4)", epilogue);
  return egpt::gemm_bf16(A, B, bias, out, M, N, K, epilogue, (cudaStream_t)stream, nullptr, nullptr, nullptr);
}

// the LayerNorm-folding epilogues, exposed for tests: 5 needs out2 (bf16 copy) + stats, 6 / 7 need colsum + stats
extern "C" int egpt_gemm_bf16_ln(const uint16_t* A, const uint16_t* B, const float* bias, void* out, int64_t M, int N,
                                 int K, int epilogue, void* out2, const float* colsum, float* stats, void* stream) {
  return egpt::gemm_bf16(A, B, bias, out, M, N, K, epilogue, (cudaStream_t)stream, out2, colsum, stats);
}
