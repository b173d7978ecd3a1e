// Multi-head self-attention of the CLIP tower on tcgen05 (sm_100a): softmax(Q K^T) V per
// (frame, head), S = 577 tokens, 16 heads x 64 (transformers modeling_clip.py CLIPAttention,
// called from model/EventChatModel.py:196).  Input: packed QKV projection [F*577, 3072] bf16
// (q pre-scaled by 1/8 in the weights); output bf16 [F*577, 1024].
//
// Persistent CTAs of 512 threads, one per SM, each working through a contiguous range of (frame, head, Q tile)
// units.  The 577 keys of a head are split into two key groups of 288 (three blocks of 96) plus one left-over
// key; each group is an independent flash-attention pipeline over the same 128-row Q tile, and the two partial
// results of a row are merged once per tile:
//   warps 0-3 / 4-7  softmax of key group 0 / 1: one thread per (row, group).  S_n is read from TMEM into
//              registers, P_n = exp2(S_n - m) goes back over it as packed bf16 (tcgen05.st), the group's partial
//              output accumulates in TMEM across the three blocks and is rescaled in place only when a block
//              raises the row maximum by more than 2^8 (otherwise the stale maximum is kept: p <= 256 is harmless
//              in fp32 / bf16).  A quarter of the exp2 runs as a cubic on the FMA pipe.  The two warps that share a
//              scheduler (same lane quarter, different group) take turns on its MUFU unit: a warp starts its exp2 stretch
//              only when its partner is a third into its own (named barriers, 64 threads each).  The softmax warps do
//              nothing else: at the end of a tile they publish (max, sum) and go straight on to the next tile.
//   warps 8-11 merge / output: one thread per row.  The left-over key 576 is handled here on the CUDA cores (a
//              64-long dot product per row, computed one tile ahead while the Q tile is still resident).  As soon as
//              group 0's last P V of tile t has retired the thread reads that partial output into registers and
//              releases the accumulator; when group 1's has, it combines the two with the left-over key's term,
//              normalises, and the warp stores its 32 rows x 64 dims with one TMA store.  This used to run inside the
//              softmax warps and cost them a third of every tile.
//   warp 12    TMA producer: K and V of a head (576 rows x 64 in blocks of 96, each block re-loaded for the next
//              head as soon as the current head's last product on it has retired; rows 576..591 -- key 576 and
//              zero fill from the 3-D tensor map -- in a small double buffer), 128-row Q tiles into a double buffer
//   warps 13/15 MMA issuer of group 0 / 1 (one thread each): S_n = Q K_n^T (M128 x N96 x K64, SS) into a double
//              buffer, O_g += P_n V_n (TS form: P read from TMEM, V from shared memory as an MN-major operand)
//   warp 14    TMEM allocator (2 groups x 2 x 96 columns for S / P, 2 x 64 for the partial outputs)
// The control warps sit in the highest warp ids: the warp arbiter favours them, and the single-thread issuers
// must not starve for issue slots behind the busy softmax warps.  Registers are handed from the control and merge
// warps to the softmax warps with setmaxnreg (56 / 104 / 176 per thread).
#include "common.cuh"
#include "sm100_ptx.cuh"

namespace egpt {

namespace at5 {
constexpr int kS = EGPT_TOKENS;            // 577
constexpr int kD = 64;
constexpr int kHeads = 16;
constexpr int kQTile = 128;
constexpr int kQTiles = (kS + kQTile - 1) / kQTile;      // 5
constexpr int kBlk = 96;                                  // keys per block
constexpr int kGroupKeys = 3 * kBlk;                      // 288: key group 0 = keys [0,288), group 1 = [288,576)
constexpr int kXKey = 2 * kGroupKeys;                     // 576: the left-over key, handled outside the MMAs
constexpr int kNBlk = 3;                                  // blocks per group and tile
constexpr int kThreads = 512;
// control warp ids: TMA producer, the two MMA issuers, TMEM allocator
constexpr int kWarpProd = 12, kWarpIss0 = 13;
constexpr int kWarpIss1 = 15, kWarpAlloc = 14;
constexpr int kRowBytes = 128;                            // 64 bf16
constexpr int kKVBytes = kXKey * kRowBytes;               // 73728: 576 keys
constexpr int kQBytes = kQTile * kRowBytes;               // 16384
constexpr int kXBytes = 16 * kRowBytes;                   // rows 576..591 of K or V (key 576 + zero fill)
constexpr int kSmemK = 0, kSmemV = kKVBytes, kSmemQ = 2 * kKVBytes;
constexpr int kSmemX = kSmemQ + 2 * kQBytes;              // [2 (head parity)][2 (K, V)][16 rows]
constexpr int kSmemStage = kSmemX + 4 * kXBytes;          // output staging: 4 merge warps x (32 rows x 128 B), SWIZZLE_128B,
constexpr int kStageBytes = 32 * kRowBytes;               // so each slab sits on a 1024-byte boundary
static_assert(kSmemStage % 1024 == 0, "swizzled staging slabs must be 1024-byte aligned");
constexpr int kSmemBar = kSmemStage + 4 * kStageBytes;
constexpr int kSmemXch = kSmemBar + 512;                  // float [2 (max, sum)][2 (tile parity)][2 (group)][128]
constexpr int kSmemTotal = kSmemXch + 4 * 1024 + 1024;    // + alignment slack
constexpr int kTmemCols = 512;
constexpr uint32_t kTmemS = 0, kTmemO = 4 * kBlk;         // columns: S/P [group][buffer][96], then O [group][64]
#ifndef EGPT_ATTN_POLY_EVERY
#define EGPT_ATTN_POLY_EVERY 4
#endif
#ifndef EGPT_ATTN_RELEASE_AT
#define EGPT_ATTN_RELEASE_AT 0
#endif
constexpr int kReleaseAt = EGPT_ATTN_RELEASE_AT;          // third of the exp2 stretch after which the partner warp may start its own
constexpr int kPolyEvery = EGPT_ATTN_POLY_EVERY;          // every kPolyEvery-th pair of exp2 runs as a cubic on the FMA pipe (0: none)

// MN-major B operand (V block: rows = keys (K), 64 contiguous dims (N) = 128 B per key, SWIZZLE_128B):
// 8-key groups are 1024 B apart (SBO); a single 64-wide N block, so LBO is unused.
__device__ __forceinline__ uint64_t make_desc_mn_sw128(uint32_t smem_addr) {
  uint64_t d = 0;
  d |= (uint64_t)((smem_addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// kind::f16 instruction descriptor with selectable B major-ness (bit 16: 1 = MN-major)
__host__ __device__ constexpr uint32_t idesc_bf16(int m, int n, int b_mn_major) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)b_mn_major << 16) | ((uint32_t)(n >> 3) << 17) |
         ((uint32_t)(m >> 4) << 24);
}

__device__ __forceinline__ void tma_load_3d(void* smem_dst, const CUtensorMap* m, uint64_t* bar, int c0, int c1, int c2) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4, %5}], [%2];"
      ::"r"(ptx::smem_u32(smem_dst)), "l"(reinterpret_cast<uint64_t>(m)), "r"(ptx::smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2)
      : "memory");
}
__device__ __forceinline__ void tma_store_3d(const CUtensorMap* m, const void* smem_src, int c0, int c1, int c2) {
  asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];"
               ::"l"(reinterpret_cast<uint64_t>(m)), "r"(ptx::smem_u32(smem_src)), "r"(c0), "r"(c1), "r"(c2) : "memory");
}
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// packed fp32x2 arithmetic (sm_100): one issue slot for two lanes of fma / add
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
  float2 d;
  asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(*reinterpret_cast<unsigned long long*>(&d))
      : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)),
        "l"(*reinterpret_cast<unsigned long long*>(&c)));
  return d;
}
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
  float2 d;
  asm("add.rn.f32x2 %0, %1, %2;" : "=l"(*reinterpret_cast<unsigned long long*>(&d))
      : "l"(*reinterpret_cast<unsigned long long*>(&a)), "l"(*reinterpret_cast<unsigned long long*>(&b)));
  return d;
}
// 2^x for x <= ~8 on the FMA pipe, two at once: Cody-Waite split (adding 1.5 * 2^23 rounds x to the nearest integer n) and a
// cubic in f = x - n in [-0.5, 0.5]; the integer part goes into the exponent field.  Relative error < 1e-4, far below the bf16
// rounding P undergoes anyway.
__device__ __forceinline__ float2 ex2_poly2(float2 x) {
  x.x = fmaxf(x.x, -126.0f); x.y = fmaxf(x.y, -126.0f);
  const float2 t = fadd2(x, make_float2(12582912.0f, 12582912.0f));
  const float2 n = fadd2(t, make_float2(-12582912.0f, -12582912.0f));
  const float2 f = ffma2(n, make_float2(-1.f, -1.f), x);
  float2 p = ffma2(f, make_float2(0.0555041f, 0.0555041f), make_float2(0.2402265f, 0.2402265f));
  p = ffma2(p, f, make_float2(0.6931472f, 0.6931472f));
  p = ffma2(p, f, make_float2(1.0f, 1.0f));
  p.x = __int_as_float(__float_as_int(p.x) + (__float_as_int(t.x) << 23));
  p.y = __int_as_float(__float_as_int(p.y) + (__float_as_int(t.y) << 23));
  return p;
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}
__device__ __forceinline__ void tmem_st_32x32b_x16(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
      "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
        "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
// D[tmem] (+)= A[tmem] * B[smem]: A is the bf16 P tile, row i in lane i, keys (2c, 2c+1) packed in column c
__device__ __forceinline__ void umma_bf16_ts(uint32_t tmem_d, uint32_t tmem_a, uint64_t desc_b, uint32_t idesc,
                                             uint32_t accumulate) {
  asm volatile(
      "{\n\t"
      ".reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t"
      "}"
      ::"r"(tmem_d), "r"(tmem_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}

struct Bars {
  uint64_t k_full[2][kNBlk], v_full[2][kNBlk];   // per head: a 96-key block of K / V landed
  uint64_t k_free[2][kNBlk], v_free[2][kNBlk];   // per head: the head's last product on that block retired (commit)
  uint64_t x_full[2], x_free[2];         // [head parity]: rows 576..591 of K and V landed / were last read (128 arrivals)
  uint64_t q_full[2], q_empty[2];        // Q tile double buffer; q_empty counts both issuers
  uint64_t s_full[2][2];                 // [group][buffer]: S block landed
  uint64_t p_full[2][2];                 // [group][buffer]: P stored over it by the group's 128 threads
  uint64_t o_done[2][2];                 // [group][block parity]: P V of a block retired
  uint64_t o_free[2];                    // [group]: the tile's partial output of that group was read out by the 128 merge threads
  uint64_t g_done[2];                    // [group], per tile: the group's (max, sum) are in shared memory (128 arrivals) and its
                                         // last P V has retired (1 commit)
  uint32_t tmem_ptr;
  float zero;                            // 0.0f, re-read after the turn-taking barrier (see the softmax warps)
};

// 8 bf16 (one 16-byte chunk of a SWIZZLE_128B row) -> fp32
__device__ __forceinline__ void unpack8(const uint4& u, float (&f)[8]) {
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    f[2 * i] = __uint_as_float(w[i] << 16);
    f[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
  }
}

// Persistent: each CTA works through a contiguous range of (head, Q tile) units of the n_heads = frames x 16 (frame, head)
// pairs.  Every barrier phase below is derived from running counters (it = head index inside this CTA's range, T = tile,
// N = block), so the pipelines run straight across head boundaries: the next head's K / V / Q stream in while the last
// tile of the current one is still in the softmax warps, and a head's last merge happens after the first block of the next.
__global__ void __launch_bounds__(kThreads, 1)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmap_q, const __grid_constant__ CUtensorMap tmap_kv,
                    const __grid_constant__ CUtensorMap tmap_out, int n_heads, long long* __restrict__ trace) {
  extern __shared__ uint8_t smem_raw[];
#ifdef EGPT_ATTN_TRACE   // developer aid: clock64 phase stamps of CTA 0's first head (tools/attn_trace.py)
#ifndef EGPT_ATTN_TRACE_IT
#define EGPT_ATTN_TRACE_IT 0
#endif
#define TR(slot) do { if (trace != nullptr && blockIdx.x == 0 && it == EGPT_ATTN_TRACE_IT) trace[(slot)] = clock64(); } while (0)
#else
#define TR(slot) do { } while (0)
#endif
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  Bars* bars = reinterpret_cast<Bars*>(smem + kSmemBar);
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0);    // warp-uniform by construction
  const int lane = threadIdx.x & 31;
  // This CTA's share: a contiguous range [g0, g1) of the n_heads * 5 (head, Q tile) units, so the load is even to within one
  // tile; a head that straddles two CTAs has its K / V loaded by both.
  const long long n_units = (long long)n_heads * kQTiles;
  const int g0 = (int)(n_units * blockIdx.x / gridDim.x), g1 = (int)(n_units * (blockIdx.x + 1) / gridDim.x);
  const int head0 = g0 / kQTiles, n_it = (g1 - 1) / kQTiles - head0 + 1;      // heads touched (>= 1)
  // tiles [tb, te) of the it-th head of this CTA
  auto tile_begin = [&](int it) { return it == 0 ? g0 - head0 * kQTiles : 0; };
  auto tile_end = [&](int it) { return it == n_it - 1 ? g1 - (head0 + it) * kQTiles : kQTiles; };

  if (warp == kWarpProd && lane == 0) { ptx::prefetch_tensormap(&tmap_q); ptx::prefetch_tensormap(&tmap_kv); ptx::prefetch_tensormap(&tmap_out); }
  if (warp == kWarpIss0 && lane == 0) {
    for (int g = 0; g < 2; ++g) {
      for (int i = 0; i < kNBlk; ++i) {
        ptx::mbar_init(&bars->k_full[g][i], 1); ptx::mbar_init(&bars->v_full[g][i], 1);
        ptx::mbar_init(&bars->k_free[g][i], 1); ptx::mbar_init(&bars->v_free[g][i], 1);
      }
      ptx::mbar_init(&bars->x_full[g], 1); ptx::mbar_init(&bars->x_free[g], 128);
      ptx::mbar_init(&bars->q_full[g], 1); ptx::mbar_init(&bars->q_empty[g], 2);
      for (int i = 0; i < 2; ++i) {
        ptx::mbar_init(&bars->s_full[g][i], 1); ptx::mbar_init(&bars->p_full[g][i], 128); ptx::mbar_init(&bars->o_done[g][i], 1);
      }
    }
    for (int g = 0; g < 2; ++g) { ptx::mbar_init(&bars->o_free[g], 128); ptx::mbar_init(&bars->g_done[g], 129); }
    bars->zero = 0.f;
    ptx::fence_barrier_init();
  }
  if (warp == kWarpAlloc) {
    ptx::tmem_alloc(&bars->tmem_ptr, kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, bars->tmem_ptr, 0);

  if (warp >= 12) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
    if (warp == kWarpProd && lane == 0) {
      // ===== TMA producer =====
      int T = 0;                                                       // tiles loaded so far
      for (int it = 0; it < n_it; ++it) {
        const int w = head0 + it, f = w / kHeads, h = w % kHeads;
        const int tb = tile_begin(it), te = tile_end(it);
        auto load_q = [&](int t) {
          if (T >= 2) ptx::mbar_wait(&bars->q_empty[T & 1], (uint32_t)(((T >> 1) - 1) & 1));   // tile T-2's last Q K^T retired
          uint8_t* dst = smem + kSmemQ + (T & 1) * kQBytes;
          ptx::mbar_arrive_expect_tx(&bars->q_full[T & 1], kQBytes);
          tma_load_3d(dst, &tmap_q, &bars->q_full[T & 1], h * kD, t * kQTile, f);
          tma_load_3d(dst + 64 * kRowBytes, &tmap_q, &bars->q_full[T & 1], h * kD, t * kQTile + 64, f);
          ++T;
        };
        load_q(tb);
        // K then V, block by block and alternating between the two key groups (16-row boxes), each block as soon as the
        // previous head has released it; the left-over key's rows right after the first K blocks (the softmax warps need
        // them during their first block)
        for (int kv = 0; kv < 2; ++kv) {
          uint8_t* dst = smem + (kv == 0 ? kSmemK : kSmemV);
          const int col = (kv + 1) * kHeads * kD + h * kD;
          for (int j = 0; j < kNBlk; ++j) {
            for (int g = 0; g < 2; ++g) {
              uint64_t* full = kv == 0 ? &bars->k_full[g][j] : &bars->v_full[g][j];
              if (it > 0) ptx::mbar_wait(kv == 0 ? &bars->k_free[g][j] : &bars->v_free[g][j], (uint32_t)((it - 1) & 1));
              const int row0 = g * kGroupKeys + j * kBlk;
              ptx::mbar_arrive_expect_tx(full, kBlk * kRowBytes);
              for (int r = 0; r < kBlk; r += 16) tma_load_3d(dst + (row0 + r) * kRowBytes, &tmap_kv, full, col, row0 + r, f);
            }
            if (kv == 0 && j == 0) {
              const int xs = it & 1;
              if (it >= 2) ptx::mbar_wait(&bars->x_free[xs], (uint32_t)(((it >> 1) - 1) & 1));
              uint8_t* xdst = smem + kSmemX + xs * 2 * kXBytes;
              ptx::mbar_arrive_expect_tx(&bars->x_full[xs], 2 * kXBytes);
              tma_load_3d(xdst, &tmap_kv, &bars->x_full[xs], kHeads * kD + h * kD, kXKey, f);
              tma_load_3d(xdst + kXBytes, &tmap_kv, &bars->x_full[xs], 2 * kHeads * kD + h * kD, kXKey, f);
            }
          }
          if (kv == 0 && tb + 1 < te) load_q(tb + 1);
        }
        for (int t = tb + 2; t < te; ++t) load_q(t);
      }
    } else if (warp == kWarpIss0 || warp == kWarpIss1) {
      // ===== MMA issuer of key group gi: an independent in-order stream of S and P V products.  The whole warp walks the
      // loop (uniform control flow and operands, so descriptors live in uniform registers); one elected lane issues =====
      const int gi = warp == kWarpIss0 ? 0 : 1;
      const uint64_t dK0 = ptx::make_smem_desc_sw128(ptx::smem_u32(smem + kSmemK + gi * kGroupKeys * kRowBytes));
      const uint64_t dV0 = make_desc_mn_sw128(ptx::smem_u32(smem + kSmemV + gi * kGroupKeys * kRowBytes));
      const uint64_t dQ0 = ptx::make_smem_desc_sw128(ptx::smem_u32(smem + kSmemQ));
      const uint32_t tmem_s0 = tmem_base + kTmemS + (uint32_t)(gi * 2 * kBlk);
      const uint32_t tmem_o = tmem_base + kTmemO + (uint32_t)(gi * kD);
      constexpr uint32_t idesc_pv = idesc_bf16(kQTile, kD, 1);
      constexpr uint32_t idesc_s = idesc_bf16(kQTile, kBlk, 0);
      // P V of block N: head it of this CTA, local tile counter T, block j; first / last = the head's first / last tile here
      auto issue_pv = [&](int N, int it, int T, int j, bool first, bool last) {
        const int b = N & 1;
        if (first) ptx::mbar_wait(&bars->v_full[gi][j], (uint32_t)(it & 1));
        TR(2048 + gi * 256 + (N % 16) * 8 + 2);
        ptx::mbar_wait(&bars->p_full[gi][b], (uint32_t)((N >> 1) & 1));
        TR(2048 + gi * 256 + (N % 16) * 8 + 3);
        // the first block of a tile overwrites the accumulator: the previous tile's output must have been read out
        if (j == 0 && T > 0) ptx::mbar_wait(&bars->o_free[gi], (uint32_t)((T - 1) & 1));
        ptx::tcgen05_fence_after();
        const uint32_t tmem_p = tmem_s0 + (uint32_t)(b * kBlk);
        // A = P from TMEM: 16 keys = 8 packed columns per k-step; B = V rows 16*kk .. +15 of the block (2048 B = 128 units)
        const uint64_t dv = dV0 + (uint64_t)(j * (kBlk * kRowBytes >> 4));
        if (ptx::elect_one()) {
#pragma unroll
          for (int kk = 0; kk < kBlk / 16; ++kk)
            umma_bf16_ts(tmem_o, tmem_p + (uint32_t)(kk * 8), dv + (uint64_t)(kk * 128), idesc_pv, (j != 0 || kk != 0) ? 1u : 0u);
          ptx::umma_commit(&bars->o_done[gi][b]);
          if (j == kNBlk - 1) ptx::umma_commit(&bars->g_done[gi]);
          if (last) ptx::umma_commit(&bars->v_free[gi][j]);                  // the head's last use of this V block
        }
        __syncwarp();
        TR(2048 + gi * 256 + (N % 16) * 8 + 4);
      };
      int N = 0, T = 0;
      int p_it = 0, p_T = 0, p_j = 0;                                  // block N-1, whose P V is still to be issued
      bool p_first = false, p_last = false;
      for (int it = 0; it < n_it; ++it) {
        const int tb = tile_begin(it), te = tile_end(it);
        for (int t = tb; t < te; ++t, ++T) {
          ptx::mbar_wait(&bars->q_full[T & 1], (uint32_t)((T >> 1) & 1));
          const uint64_t dq = dQ0 + (uint64_t)((T & 1) * (kQBytes >> 4));
          for (int j = 0; j < kNBlk; ++j, ++N) {
            if (t == tb) ptx::mbar_wait(&bars->k_full[gi][j], (uint32_t)(it & 1));
            TR(2048 + gi * 256 + (N % 16) * 8 + 0);
            // S_N overwrites the buffer that held P_{N-2}; P V of block N-2 was issued earlier by this same thread and
            // the tensor pipe runs one thread's MMAs in issue order, so that reuse needs no barrier.
            ptx::tcgen05_fence_after();
            const uint32_t tmem_s = tmem_s0 + (uint32_t)((N & 1) * kBlk);
            const uint64_t dk = dK0 + (uint64_t)(j * (kBlk * kRowBytes >> 4));
            if (ptx::elect_one()) {
#pragma unroll
              for (int k = 0; k < kD / 16; ++k) ptx::umma_bf16_ss(tmem_s, dq + (uint64_t)(k * 2), dk + (uint64_t)(k * 2), idesc_s, k != 0);
              ptx::umma_commit(&bars->s_full[gi][N & 1]);
              if (j == kNBlk - 1) ptx::umma_commit(&bars->q_empty[T & 1]);
              if (t == te - 1) ptx::umma_commit(&bars->k_free[gi][j]);       // the head's last use of this K block
            }
            __syncwarp();
            TR(2048 + gi * 256 + (N % 16) * 8 + 1);
            if (N >= 1) issue_pv(N - 1, p_it, p_T, p_j, p_first, p_last);
            p_it = it; p_T = T; p_j = j; p_first = (t == tb); p_last = (t == te - 1);
          }
        }
      }
      issue_pv(N - 1, p_it, p_T, p_j, p_first, p_last);
    }
  } else if (warp >= 8) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 104;");
    // ===== merge / output: thread = row r of the tile, all 64 dims.  Runs one tile behind the softmax warps. =====
    const int q = warp & 3;
    const int r = q * 32 + lane;                       // row inside the tile == TMEM lane
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    const uint32_t tmem_oa = tmem_base + lane_base + kTmemO, tmem_ob = tmem_oa + (uint32_t)kD;
    const float* xch = reinterpret_cast<const float*>(smem + kSmemXch);     // [2][2 (tile parity)][2 (group)][128]
    uint8_t* stage = smem + kSmemStage + q * kStageBytes;
    constexpr float kLog2e = 1.4426950408889634f;
    // q_r . k_576 of local tile T (head it of this CTA): the Q tile and the head's rows 576.. are TMA-written, the waits are
    // the acquire.  Called one tile AHEAD of the merge that uses it: the Q buffer is recycled when the tile's last S product
    // retires, which cannot happen before the previous tile's merge has released the accumulators (o_free, below).
    auto xdot = [&](int T, int it, bool head_first) {
      ptx::mbar_wait(&bars->q_full[T & 1], (uint32_t)((T >> 1) & 1));
      if (head_first) ptx::mbar_wait(&bars->x_full[it & 1], (uint32_t)((it >> 1) & 1));
      const uint8_t* qrow = smem + kSmemQ + (T & 1) * kQBytes + r * kRowBytes;
      const uint4* kx = reinterpret_cast<const uint4*>(smem + kSmemX + (it & 1) * 2 * kXBytes);   // K row 576
      float d4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int c = 0; c < 8; ++c) {
        float qf[8], kf[8];
        unpack8(*reinterpret_cast<const uint4*>(qrow + ((c ^ (r & 7)) << 4)), qf);
        unpack8(kx[c], kf);
#pragma unroll
        for (int i = 0; i < 8; ++i) d4[i & 3] = fmaf(qf[i], kf[i], d4[i & 3]);
      }
      return (d4[0] + d4[1]) + (d4[2] + d4[3]);
    };
    int T = 0;
    float dot_cur = xdot(0, 0, true);
    for (int it = 0; it < n_it; ++it) {
      const int w = head0 + it, f = w / kHeads, h = w % kHeads;
      const int tb = tile_begin(it), te = tile_end(it);
      for (int t = tb; t < te; ++t, ++T) {
        float dot_next = 0.f;
        if (q == 0 && lane == 0) TR(3072 + (T % 16) * 8 + 0);
        if (t + 1 < te) dot_next = xdot(T + 1, it, false);
        else if (it + 1 < n_it) dot_next = xdot(T + 1, it + 1, true);
        // The two key groups run on their own clocks (group 0 leads by about one block).  Each group's accumulator is read out and
        // released as soon as THAT group's last P V of the tile has retired, so neither group's next tile waits for the other:
        // group 0's partial output sits in registers, unscaled, until group 1's maximum is known.
        if (q == 0 && lane == 0) TR(3072 + (T % 16) * 8 + 1);
        const float* x_t = xch + (T & 1) * 256;
        ptx::mbar_wait(&bars->g_done[0], (uint32_t)(T & 1));
        if (q == 0 && lane == 0) TR(3072 + (T % 16) * 8 + 2);
        ptx::tcgen05_fence_after();
        uint32_t oa[4][16];
#pragma unroll
        for (int c = 0; c < 4; ++c) ptx::tmem_ld_32x32b_x16(tmem_oa + (uint32_t)(16 * c), oa[c]);
        const float m0 = x_t[r], l0 = x_t[512 + r];
        ptx::tmem_ld_wait();
        ptx::tcgen05_fence_before();
        ptx::mbar_arrive(&bars->o_free[0]);
        if (q == 0 && lane == 0) TR(3072 + (T % 16) * 8 + 3);
        ptx::mbar_wait(&bars->g_done[1], (uint32_t)(T & 1));
        if (q == 0 && lane == 0) TR(3072 + (T % 16) * 8 + 5);
        ptx::tcgen05_fence_after();
        const float m1 = x_t[128 + r], l1 = x_t[512 + 128 + r];
        const float sx = dot_cur * kLog2e;
        const float mt = fmaxf(fmaxf(m0, m1), sx);
        const float f0 = ex2(m0 - mt), f1 = ex2(m1 - mt), px = ex2(sx - mt);
        const float inv = __fdividef(1.0f, fmaf(l0, f0, l1 * f1) + px);
        const float c0 = f0 * inv, c1 = f1 * inv, cx = px * inv;
        const float2 ca2 = make_float2(c0, c0), cb2 = make_float2(c1, c1), cx2 = make_float2(cx, cx), z2 = make_float2(0.f, 0.f);
        const uint4* vx = reinterpret_cast<const uint4*>(smem + kSmemX + (it & 1) * 2 * kXBytes + kXBytes);   // V row 576
        if (lane == 0) ptx::bulk_wait_group_read<0>();                   // the previous tile's store has read the slab
        __syncwarp();
#pragma unroll
        for (int c = 0; c < 4; ++c) {                                    // 16 dims at a time
          uint32_t ob[16];
          ptx::tmem_ld_32x32b_x16(tmem_ob + (uint32_t)(16 * c), ob);
          ptx::tmem_ld_wait();
          if (c == 3) {
            ptx::tcgen05_fence_before();
            ptx::mbar_arrive(&bars->o_free[1]);
            if (q == 0 && lane == 0) TR(3072 + (T % 16) * 8 + 6);
          }
#pragma unroll
          for (int hh = 0; hh < 2; ++hh) {
            float vf[8];
            unpack8(vx[2 * c + hh], vf);
            uint32_t wd[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
              float2 e = ffma2(make_float2(vf[2 * i], vf[2 * i + 1]), cx2, z2);
              e = ffma2(make_float2(__uint_as_float(ob[8 * hh + 2 * i]), __uint_as_float(ob[8 * hh + 2 * i + 1])), cb2, e);
              e = ffma2(make_float2(__uint_as_float(oa[c][8 * hh + 2 * i]), __uint_as_float(oa[c][8 * hh + 2 * i + 1])), ca2, e);
              wd[i] = pack2(e.x, e.y);
            }
            // SWIZZLE_128B slab: 16-byte chunk index XOR (row & 7) -- the minimum of four wavefronts per warp store
            *reinterpret_cast<uint4*>(stage + lane * kRowBytes + (((2 * c + hh) ^ (lane & 7)) << 4)) = make_uint4(wd[0], wd[1], wd[2], wd[3]);
          }
        }
        if (t == te - 1) ptx::mbar_arrive(&bars->x_free[it & 1]);        // the head's rows 576.. are no longer needed
        ptx::fence_proxy_async();
        __syncwarp();
        if (lane == 0) {
          tma_store_3d(&tmap_out, stage, h * kD, t * kQTile + q * 32, f);   // rows past 577 are clipped by the tensor map
          ptx::bulk_commit_group();
        }
        if (q == 0 && lane == 0) TR(3072 + (T % 16) * 8 + 4);
        dot_cur = dot_next;
      }
    }
    if (lane == 0) ptx::bulk_wait_group_read<0>();
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 176;");
    // ===== softmax: thread (row r, key group gi) walks its group's key blocks on its own: own running maximum, own
    // row sum, own partial output in TMEM =====
    const int q = warp & 3, gi = warp >> 2;
    const int r = q * 32 + lane;                       // row inside the tile == TMEM lane
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    const uint32_t tmem_s0 = tmem_base + lane_base + kTmemS + (uint32_t)(gi * 2 * kBlk);
    const uint32_t tmem_o = tmem_base + lane_base + kTmemO + (uint32_t)(gi * kD);
    float* xch = reinterpret_cast<float*>(smem + kSmemXch);           // [2][2 (tile parity)][2 (group)][128]
    constexpr float kLog2e = 1.4426950408889634f;
    float m = -INFINITY, l = 0.f;
    const int n_blocks = (g1 - g0) * kNBlk;                            // blocks this thread will process
    if (gi == 1) asm volatile("bar.arrive %0, 64;" ::"r"(1 + q) : "memory");   // group 0 takes the first turn

    int N = 0, T = 0;
    for (int it = 0; it < n_it; ++it) {
      const bool trc = (lane == 0);
      const int tb = tile_begin(it), te = tile_end(it);
      for (int t = tb; t < te; ++t, ++T) {
        for (int j = 0; j < kNBlk; ++j, ++N) {
          const int b = N & 1;
          [[maybe_unused]] const int n = N % 16;                        // trace slot
          const uint32_t ts = tmem_s0 + (uint32_t)(b * kBlk);
          if (trc) TR(warp * 256 + n * 16 + 0);
          ptx::mbar_wait(&bars->s_full[gi][b], (uint32_t)((N >> 1) & 1));
          ptx::tcgen05_fence_after();
          if (trc) TR(warp * 256 + n * 16 + 1);
          uint32_t v[3][32];
          ptx::tmem_ld_32x32b_x32(ts, v[0]);
          ptx::tmem_ld_32x32b_x32(ts + 32u, v[1]);
          ptx::tmem_ld_32x32b_x32(ts + 64u, v[2]);
          ptx::tmem_ld_wait();
          if (trc) TR(warp * 256 + n * 16 + 2);
          float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
          for (int c = 0; c < 3; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i) m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(v[c][i]));
          const float m_blk = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3])) * kLog2e;   // exp2 domain; finite
          if (j == 0) {
            m = m_blk; l = 0.f;                                        // fresh accumulator, nothing to rescale
          } else {
            // Lazy rescale: only when this block raises the row maximum by more than 2^8 are the accumulator and the
            // running sum moved to the new frame; otherwise the stale maximum stays and p may reach 256.
            const bool need = m_blk - m > 8.0f;
            if (__any_sync(0xffffffffu, need)) {
              const float fac = need ? ex2(m - m_blk) : 1.0f;
              // P V of block N-1 must have retired (blocks <= N-2 have: S_N was committed after them, so this barrier is
              // at most one phase behind and the parity wait is exact)
              ptx::mbar_wait(&bars->o_done[gi][(N - 1) & 1], (uint32_t)(((N - 1) >> 1) & 1));
              ptx::tcgen05_fence_after();
#pragma unroll
              for (int c = 0; c < 2; ++c) {
                uint32_t a[32];
                ptx::tmem_ld_32x32b_x32(tmem_o + (uint32_t)(32 * c), a);
                ptx::tmem_ld_wait();
#pragma unroll
                for (int i = 0; i < 32; ++i) a[i] = __float_as_uint(__uint_as_float(a[i]) * fac);
                tmem_st_32x32b_x16(tmem_o + (uint32_t)(32 * c), a);
                tmem_st_32x32b_x16(tmem_o + (uint32_t)(32 * c + 16), a + 16);
              }
              l *= fac;
              if (need) m = m_blk;
            }
          }
          if (trc) TR(warp * 256 + n * 16 + 3);
          // The exp2 stretch is bound by the scheduler's MUFU unit, which this warp shares with the other key group's warp of
          // the same TMEM lane quarter (warp ^ 4).  The two take turns: one runs its exp2 stretch at the unit's full rate
          // while the other waits for S, loads it and finds the row maximum -- instead of both crawling through overlapping
          // stretches and then idling the unit together.  Two named barriers per warp pair (64 threads: one warp syncs, the
          // other arrives); group 0 goes first, and both groups have the same number of blocks.
          asm volatile("bar.sync %0, 64;" ::"r"(1 + gi * 4 + q) : "memory");
          // ptxas moves register-only math across a barrier (it hoisted 28 of the 72 MUFUs above this one): every exp2 below
          // takes -m as an input, and -m now depends on a shared-memory load that cannot move above the barrier.
          float zero;
          asm volatile("ld.shared.f32 %0, [%1];" : "=f"(zero) : "r"(ptx::smem_u32(&bars->zero)) : "memory");
          const float mz = m + zero;
          // exp2 on packed fp32x2 instructions (FFMA2 / FADD2): scale, cubic and row sum cost half the issue slots
          float2 s2[2] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
          const float2 sc = make_float2(kLog2e, kLog2e), nm = make_float2(-mz, -mz);
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            uint32_t wd[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float2 x = ffma2(make_float2(__uint_as_float(v[c][2 * i]), __uint_as_float(v[c][2 * i + 1])), sc, nm);
              float2 pp;
              if (kPolyEvery > 0 && (i % (kPolyEvery > 0 ? kPolyEvery : 1)) == kPolyEvery - 1) {
                pp = ex2_poly2(x);
              } else {
                pp.x = ex2(x.x); pp.y = ex2(x.y);
              }
              s2[i & 1] = fadd2(s2[i & 1], pp);
              wd[i] = pack2(pp.x, pp.y);
            }
            tmem_st_32x32b_x16(ts + (uint32_t)(16 * c), wd);          // P over S: column c <- keys 2c, 2c+1
            if (c == kReleaseAt && (gi == 0 || N + 1 < n_blocks))      // the partner warp may start its stretch
              asm volatile("bar.arrive %0, 64;" ::"r"(1 + (gi ^ 1) * 4 + q) : "memory");
          }
          l += (s2[0].x + s2[0].y) + (s2[1].x + s2[1].y);
          if (trc) TR(warp * 256 + n * 16 + 4);
          tmem_st_wait();
          ptx::tcgen05_fence_before();
          ptx::mbar_arrive(&bars->p_full[gi][b]);
          if (trc) TR(warp * 256 + n * 16 + 5);
          if (j == kNBlk - 1) {
            // end of the tile: hand (max, sum) to the merge warps and go straight on
            float* x_t = xch + (T & 1) * 256;
            x_t[gi * 128 + r] = m;
            x_t[512 + gi * 128 + r] = l;
            ptx::mbar_arrive(&bars->g_done[gi]);
          }
        }
      }
    }
  }

  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  if (warp == kWarpAlloc) ptx::tmem_dealloc(tmem_base, kTmemCols);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      return reinterpret_cast<EncodeTiledFn>(p);
    return (EncodeTiledFn) nullptr;
  }();
  return fn;
}
}  // namespace at5

int attention_impl(const void* qkv, int n_frames_total, void* out, cudaStream_t stream, long long* trace) {
  using namespace at5;
  EGPT_CHECK_ARG(qkv && out && n_frames_total > 0, "attention: bad arguments");
  EGPT_CHECK_ARG(((uintptr_t)qkv % 16 == 0) && ((uintptr_t)out % 16 == 0), "attention: pointers must be 16-byte aligned");
  EncodeTiledFn fn = encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled entry point not available"); return EGPT_ERR_CUDA; }
  // qkv viewed as [F][577][3072]; rows past 577 of a frame are out of bounds -> zero filled
  CUtensorMap tm_q, tm_kv;
  cuuint64_t dims[3] = {3 * kHeads * kD, (cuuint64_t)kS, (cuuint64_t)n_frames_total};
  cuuint64_t strides[2] = {(cuuint64_t)3 * kHeads * kD * 2, (cuuint64_t)kS * 3 * kHeads * kD * 2};
  cuuint32_t estr[3] = {1, 1, 1};
  cuuint32_t box_q[3] = {kD, 64, 1}, box_kv[3] = {kD, 16, 1};
  CUresult r = fn(&tm_q, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(qkv), dims, strides, box_q, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r == CUDA_SUCCESS)
    r = fn(&tm_kv, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(qkv), dims, strides, box_kv, estr,
           CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
           CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  // out viewed as [F][577][1024]: a merge warp stores 32 rows x 64 dims (one head) at a time from a SWIZZLE_128B slab,
  // rows past 577 are clipped
  CUtensorMap tm_out;
  cuuint64_t odims[3] = {kHeads * kD, (cuuint64_t)kS, (cuuint64_t)n_frames_total};
  cuuint64_t ostrides[2] = {(cuuint64_t)kHeads * kD * 2, (cuuint64_t)kS * kHeads * kD * 2};
  cuuint32_t box_o[3] = {kD, 32, 1};
  if (r == CUDA_SUCCESS)
    r = fn(&tm_out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, out, odims, ostrides, box_o, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("attention: cuTensorMapEncodeTiled failed with CUresult %d", (int)r); return EGPT_ERR_CUDA; }
  // per launch: the attribute is per device, and one process may drive several GPUs
  EGPT_CUDA(cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal));
  ProfScope prof(PROF_ATTN, 4.0 * kS * kS * kD * kHeads * (double)n_frames_total, stream);
  const int n_heads = n_frames_total * kHeads, n_units = n_heads * kQTiles;
  attention_tc_kernel<<<n_units < kNumSMs ? n_units : kNumSMs, kThreads, kSmemTotal, stream>>>(tm_q, tm_kv, tm_out, n_heads, trace);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}

int attention(const void* qkv, int n_frames_total, void* out, cudaStream_t stream) {
  return attention_impl(qkv, n_frames_total, out, stream, nullptr);
}

}  // namespace egpt

#ifdef EGPT_ATTN_TRACE
// developer aid (tools/attn_trace.py): clock64 stamps of CTA 0 -- lane 0 of softmax warp w at trace[w*256 + n*16 + k],
// the two MMA issuers at trace[2048 + group*256 + n*8 + k]
extern "C" int egpt_attention_trace(const uint16_t* qkv, int n_frames_total, uint16_t* out, long long* trace, void* stream) {
  return egpt::attention_impl(qkv, n_frames_total, out, (cudaStream_t)stream, trace);
}
#endif

extern "C" int egpt_attention(const uint16_t* qkv, int n_frames_total, uint16_t* out, void* stream) {
  return egpt::attention(qkv, n_frames_total, out, (cudaStream_t)stream);
}
