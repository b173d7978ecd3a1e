// Multi-head self-attention of the CLIP tower on tcgen05 (sm_100a): softmax(Q K^T) V per
// (frame, head), S = 577 tokens, 16 heads x 64 (transformers modeling_clip.py CLIPAttention,
// called from model/EventChatModel.py:196).  Input: packed QKV projection [F*577, 3072] bf16
// (q pre-scaled by 1/8 in the weights); output bf16 [F*577, 1024].
//
// One CTA per (frame, head), 384 threads:
//   warp 8     TMA producer: K and V of the head (592 padded rows x 64, zero filled past 577 by the
//              3-D tensor map) once, then one 128-row Q tile at a time
//   warp 9     MMA issuer (one thread): S_j = Q K_j^T (M128 x N192 x K64) into a double-buffered TMEM
//              region, then O_j = P_j V_j (M128 x N64 x K192; P read straight from TMEM where the
//              softmax warps wrote it over S_j, V from shared memory as an MN-major operand)
//   warp 10    TMEM allocator (512 columns: 2 x 192 for S, 2 x 64 for the output of even / odd tiles)
//   warps 0-7  softmax: two threads per query row, each owning half of the key columns of a block
//              and half of the 64 output dims.  S_j is read from TMEM once into registers, the
//              row maximum is exchanged with the partner thread through shared memory, P_j goes
//              back into the same TMEM columns as packed bf16 (tcgen05.st).  The output accumulates
//              in TMEM across the key blocks of a tile; it is rescaled in place only when a block
//              raises a row's maximum by more than 2^8 (otherwise the stale maximum is kept: p <= 256
//              is harmless in fp32 / bf16), so the common path is load S, max, exp2, store P.
// Keys are walked in four blocks of 192, 192, 192 and 16 (= 592 >= 577).
#include "common.cuh"
#include "sm100_ptx.cuh"
#include <mutex>

namespace egpt {

namespace at5 {
constexpr int kS = EGPT_TOKENS;            // 577
constexpr int kD = 64;
constexpr int kHeads = 16;
constexpr int kKeyPad = 592;
constexpr int kQTile = 128;
constexpr int kQTiles = (kS + kQTile - 1) / kQTile;      // 5
constexpr int kBlocks = 4;                                // key blocks per q tile: 192,192,192,16
constexpr int kBlk = 192;
constexpr int kTail = kKeyPad - 3 * kBlk;                 // 16
constexpr int kThreads = 384;
constexpr int kRowBytes = 128;                            // 64 bf16
constexpr int kKVBytes = kKeyPad * kRowBytes;             // 75776
constexpr int kQBytes = kQTile * kRowBytes;               // 16384
constexpr int kSmemK = 0, kSmemV = kKVBytes, kSmemQ = 2 * kKVBytes;
constexpr int kSmemBar = kSmemQ + kQBytes;                // 167936
constexpr int kSmemXch = kSmemBar + 256;                  // float [2][2][128] row maxima + [2][128] row sums
constexpr int kSmemTotal = kSmemXch + 3 * 1024 + 1024;    // + alignment slack
constexpr int kTmemCols = 512;
constexpr uint32_t kTmemS = 0, kTmemO = 2 * kBlk;         // columns

__device__ __forceinline__ int block_cols(int j) { return j < 3 ? kBlk : kTail; }   // 192 / 16

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
__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// 2^x for x <= 0 on the FMA/ALU pipes (Cody-Waite split + cubic), used for a quarter of the
// elements so the MUFU pipe (16 exp2/clk/SM, the unit this kernel saturates) is not the only
// exponential engine.  Relative error < 1e-4, far below the bf16 rounding P undergoes anyway.
__device__ __forceinline__ float ex2_poly(float x) {
  x = fmaxf(x, -126.0f);
  const float t = x + 12582912.0f;                 // 1.5 * 2^23: rounds x to the nearest integer n
  const float n = t - 12582912.0f;
  const float f = x - n;                           // [-0.5, 0.5]
  float p = fmaf(f, 0.0555041f, 0.2402265f);
  p = fmaf(p, f, 0.6931472f);
  p = fmaf(p, f, 1.0f);
  return __int_as_float(__float_as_int(p) + (__float_as_int(t) << 23));
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

__device__ __forceinline__ void tmem_ld_32x32b_x16(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr)
      : "memory");
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
__device__ __forceinline__ void named_bar_sync(int id, int nthreads) {
  asm volatile("bar.sync %0, %1;" ::"r"(id), "r"(nthreads) : "memory");
}

struct Bars {
  uint64_t k_full[kBlocks], v_full[kBlocks];
  uint64_t q_full, q_empty;
  uint64_t s_full[2];
  uint64_t p_full[2];
  uint64_t o_done[2];                    // [block parity]: the P V product of a block has retired
  uint64_t o_free[2];                    // [tile parity]: the tile's output was read out of TMEM
  uint32_t tmem_ptr;
};

__global__ void __launch_bounds__(kThreads, 1)
attention_tc_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const __grid_constant__ CUtensorMap tmap_tail,
                    __nv_bfloat16* __restrict__ out, long long* __restrict__ trace) {
  extern __shared__ uint8_t smem_raw[];
#ifdef EGPT_ATTN_TRACE   // developer aid: clock64 phase stamps of CTA 0 (tools/attn_trace.py)
#define TR(slot) do { if (trace != nullptr && blockIdx.x == 0) trace[(slot)] = clock64(); } while (0)
#else
#define TR(slot) do { } while (0)
#endif
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  Bars* bars = reinterpret_cast<Bars*>(smem + kSmemBar);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int f = blockIdx.x / kHeads, h = blockIdx.x % kHeads;

  if (warp == 8 && lane == 0) { ptx::prefetch_tensormap(&tmap_qkv); ptx::prefetch_tensormap(&tmap_tail); }
  if (warp == 9 && lane == 0) {
    for (int i = 0; i < kBlocks; ++i) { ptx::mbar_init(&bars->k_full[i], 1); ptx::mbar_init(&bars->v_full[i], 1); }
    ptx::mbar_init(&bars->q_full, 1); ptx::mbar_init(&bars->q_empty, 1);
    for (int i = 0; i < 2; ++i) {
      ptx::mbar_init(&bars->s_full[i], 1); ptx::mbar_init(&bars->p_full[i], 256);
      ptx::mbar_init(&bars->o_done[i], 1); ptx::mbar_init(&bars->o_free[i], 256);
    }
    ptx::fence_barrier_init();
  }
  if (warp == 10) {
    ptx::tmem_alloc(&bars->tmem_ptr, kTmemCols);
    ptx::tmem_relinquish();
  }
  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  const uint32_t tmem_base = bars->tmem_ptr;
  // hand registers from the four control warps to the softmax warps (the row block lives in registers)
  // Softmax warps are 0-7, control warps 8-11: the warp arbiter favours higher warp ids and the single-thread
  // TMA / MMA issuers must not starve for issue slots behind eight busy softmax warps.
  if (warp >= 8) {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 56;");
  if (warp == 8 && lane == 0) {
    // ===== TMA producer =====
    // Q tile 0 first, then K and V block by block (64-row boxes) so the first MMAs can start early
    ptx::mbar_arrive_expect_tx(&bars->q_full, kQBytes);
    tma_load_3d(smem + kSmemQ, &tmap_qkv, &bars->q_full, h * kD, 0, f);
    tma_load_3d(smem + kSmemQ + 64 * kRowBytes, &tmap_qkv, &bars->q_full, h * kD, 64, f);
    for (int kv = 0; kv < 2; ++kv) {
      uint8_t* dst = smem + (kv == 0 ? kSmemK : kSmemV);
      uint64_t* full = kv == 0 ? bars->k_full : bars->v_full;
      const int col = (kv + 1) * kHeads * kD + h * kD;
      for (int j = 0; j < kBlocks; ++j) {
        const int rows = block_cols(j);
        ptx::mbar_arrive_expect_tx(&full[j], rows * kRowBytes);
        if (j < 3) {
          for (int r = 0; r < rows; r += 64)
            tma_load_3d(dst + (j * kBlk + r) * kRowBytes, &tmap_qkv, &full[j], col, j * kBlk + r, f);
        } else {
          tma_load_3d(dst + 3 * kBlk * kRowBytes, &tmap_tail, &full[j], col, 3 * kBlk, f);      // 16-row box
        }
      }
    }
    for (int t = 1; t < kQTiles; ++t) {
      ptx::mbar_wait(&bars->q_empty, (uint32_t)((t - 1) & 1));       // tile t-1's last Q K^T has retired
      ptx::mbar_arrive_expect_tx(&bars->q_full, kQBytes);
      // two 64-row boxes: the tensor map's box is 64 rows
      tma_load_3d(smem + kSmemQ, &tmap_qkv, &bars->q_full, h * kD, t * kQTile, f);
      tma_load_3d(smem + kSmemQ + 64 * kRowBytes, &tmap_qkv, &bars->q_full, h * kD, t * kQTile + 64, f);
    }
  } else if (warp == 9 && lane == 0) {
    // ===== MMA issuer =====
    const uint32_t sK = ptx::smem_u32(smem + kSmemK), sV = ptx::smem_u32(smem + kSmemV);
    const uint32_t sQ = ptx::smem_u32(smem + kSmemQ);
    const uint64_t dV0 = make_desc_mn_sw128(sV), dK0 = ptx::make_smem_desc_sw128(sK), dQ = ptx::make_smem_desc_sw128(sQ);
    auto issue_pv = [&](int g) {
      const int j = g & 3, b = g & 1, n = g >> 1;
      if (g < kBlocks) ptx::mbar_wait(&bars->v_full[j], 0);
      if (g < 12) TR(256 + g * 8 + 2);
      ptx::mbar_wait(&bars->p_full[b], (uint32_t)(n & 1));
      if (g < 12) TR(256 + g * 8 + 3);
      const int t = g >> 2;
      // the first block of a tile starts a fresh accumulator: tile t-2's output must have been read out
      if (j == 0) ptx::mbar_wait(&bars->o_free[t & 1], (uint32_t)(((t >> 1) & 1) ^ 1));
      ptx::tcgen05_fence_after();
      constexpr uint32_t idesc = idesc_bf16(kQTile, kD, 1);
      const uint32_t tmem_o = tmem_base + kTmemO + (uint32_t)((t & 1) * kD);
      const uint32_t tmem_p = tmem_base + kTmemS + (uint32_t)(b * kBlk);
      // A = P from TMEM: 16 keys = 8 packed columns per k-step; B = V rows j*192 + 16*kk .. +15 (2048 B = 128 desc units)
      const uint64_t db0 = dV0 + (uint64_t)(j * (kBlk * kRowBytes >> 4));
      if (j < 3) {
#pragma unroll
        for (int kk = 0; kk < kBlk / 16; ++kk)
          umma_bf16_ts(tmem_o, tmem_p + (uint32_t)(kk * 8), db0 + (uint64_t)(kk * 128), idesc, (j != 0 || kk != 0) ? 1u : 0u);
      } else {
        umma_bf16_ts(tmem_o, tmem_p, db0, idesc, 1u);              // tail block: never the first of a tile
      }
      ptx::umma_commit(&bars->o_done[b]);
      if (g < 12) TR(256 + g * 8 + 4);
    };
    for (int t = 0; t < kQTiles; ++t) {
      ptx::mbar_wait(&bars->q_full, (uint32_t)(t & 1));
      for (int j = 0; j < kBlocks; ++j) {
        const int g = t * kBlocks + j, b = g & 1;
        if (g < 12) TR(256 + g * 8 + 0);
        if (t == 0) ptx::mbar_wait(&bars->k_full[j], 0);
        // S_g overwrites the buffer that held S_{g-2} / P_{g-2}: the tensor pipe runs MMAs in issue order,
        // and P V of block g-2 was issued before this point, so no barrier is needed for that reuse.
        ptx::tcgen05_fence_after();
        const uint32_t idesc = j < 3 ? idesc_bf16(kQTile, kBlk, 0) : idesc_bf16(kQTile, kTail, 0);
        const uint32_t tmem_s = tmem_base + kTmemS + (uint32_t)(b * kBlk);
        const uint64_t db = dK0 + (uint64_t)(j * (kBlk * kRowBytes >> 4));
#pragma unroll
        for (int k = 0; k < kD / 16; ++k) ptx::umma_bf16_ss(tmem_s, dQ + (uint64_t)(k * 2), db + (uint64_t)(k * 2), idesc, k != 0);
        ptx::umma_commit(&bars->s_full[b]);
        if (g < 12) TR(256 + g * 8 + 1);
        if (j == kBlocks - 1) ptx::umma_commit(&bars->q_empty);
        if (g >= 1) issue_pv(g - 1);
      }
    }
    issue_pv(kQTiles * kBlocks - 1);
  }
  } else {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 224;");
    // ===== softmax: thread (row r, half hf) owns key columns [hf*96, hf*96+96) of a 192-key block
    // (the 16-key tail block belongs to half 0) and output dims [hf*32, hf*32+32) =====
    const int q = warp & 3, hf = warp >> 2;
    const int r = q * 32 + lane;                       // row inside the tile == TMEM lane
    const uint32_t lane_base = (uint32_t)(q * 32) << 16;
    float* xmax = reinterpret_cast<float*>(smem + kSmemXch);          // [2 (block parity)][2 (half)][128]
    float* xsum = xmax + 2 * 2 * 128;                                 // [2 (half)][128]
    constexpr float kLog2e = 1.4426950408889634f;
    float m = -INFINITY, l = 0.f;
    for (int t = 0; t < kQTiles; ++t) {
      const uint32_t to = tmem_base + lane_base + kTmemO + (uint32_t)((t & 1) * kD + hf * 32);
      m = -INFINITY; l = 0.f;
      for (int j = 0; j < kBlocks; ++j) {
        const int g = t * kBlocks + j, b = g & 1, n = g >> 1;
        const uint32_t ts = tmem_base + lane_base + kTmemS + (uint32_t)(b * kBlk);
        // number of valid (un-masked) columns among this thread's columns, and where they start
        const int col0 = j < 3 ? hf * 96 : 0;
        const int nvalid = j < 3 ? 96 : (hf == 0 ? kS - 3 * kBlk : 0);         // 96 / 1 / 0
        const bool trc = (warp == 0 && lane == 0 && g < 12);
        if (trc) TR(g * 16 + 0);
        ptx::mbar_wait(&bars->s_full[b], (uint32_t)(n & 1));
        ptx::tcgen05_fence_after();
        if (trc) TR(g * 16 + 1);
        uint32_t v[3][32];
        if (j < 3) {
          ptx::tmem_ld_32x32b_x32(ts + (uint32_t)col0, v[0]);
          ptx::tmem_ld_32x32b_x32(ts + (uint32_t)col0 + 32u, v[1]);
          ptx::tmem_ld_32x32b_x32(ts + (uint32_t)col0 + 64u, v[2]);
        } else if (hf == 0) {
          tmem_ld_32x32b_x16(ts, v[0]);
        }
        ptx::tmem_ld_wait();
        if (trc) TR(g * 16 + 2);
        float mx = -INFINITY;
        if (j < 3) {
          float m4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
          for (int c = 0; c < 3; ++c)
#pragma unroll
            for (int i = 0; i < 32; ++i) m4[i & 3] = fmaxf(m4[i & 3], __uint_as_float(v[c][i]));
          mx = fmaxf(fmaxf(m4[0], m4[1]), fmaxf(m4[2], m4[3]));
        } else if (hf == 0) {
          mx = __uint_as_float(v[0][0]);                               // key 576, the only valid one of the tail
        }
        xmax[(b * 2 + hf) * 128 + r] = mx;
        // Both warps of this row group have S_j in registers past this barrier, so P_j may overwrite it.
        named_bar_sync(1 + q, 64);
        mx = fmaxf(mx, xmax[(b * 2 + (hf ^ 1)) * 128 + r]);
        const float m_blk = mx * kLog2e;                               // exp2 domain; finite; same in both threads of a row
        if (j == 0) {
          m = m_blk;                                                   // fresh accumulator, nothing to rescale
        } else {
          // Lazy rescale: only when this block raises the row maximum by more than 2^8 are the accumulator and the
          // running sum moved to the new frame; otherwise the stale maximum stays and p may reach 256.
          const bool need = m_blk - m > 8.0f;
          if (__any_sync(0xffffffffu, need)) {
            const float fac = need ? ex2(m - m_blk) : 1.0f;
            // P V of block g-1 must have retired (blocks <= g-2 have: S_g was committed after them, so this barrier is
            // at most one phase behind and the parity wait is exact)
            ptx::mbar_wait(&bars->o_done[(g - 1) & 1], (uint32_t)(((g - 1) >> 1) & 1));
            ptx::tcgen05_fence_after();
            uint32_t a[32];
            ptx::tmem_ld_32x32b_x32(to, a);
            ptx::tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < 32; ++i) a[i] = __float_as_uint(__uint_as_float(a[i]) * fac);
            tmem_st_32x32b_x16(to, a);
            tmem_st_32x32b_x16(to + 16u, a + 16);
            l *= fac;
            if (need) m = m_blk;
          }
        }
        const float m_new = m;
        if (trc) TR(g * 16 + 3);
        float s4[4] = {0.f, 0.f, 0.f, 0.f};
        // P_j (bf16 pairs) goes to columns [hf*48, hf*48+48) of the S_j buffer
        const uint32_t tp = ts + (uint32_t)(hf * 48);
        if (j < 3) {
#pragma unroll
          for (int c = 0; c < 3; ++c) {
            uint32_t w[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) {
              const float x0 = fmaf(__uint_as_float(v[c][2 * i]), kLog2e, -m_new);
              const float x1 = fmaf(__uint_as_float(v[c][2 * i + 1]), kLog2e, -m_new);
              const float p0 = ex2(x0);
              const float p1 = (i & 1) ? ex2_poly(x1) : ex2(x1);
              s4[i & 3] += p0 + p1;
              w[i] = pack2(p0, p1);
            }
            tmem_st_32x32b_x16(tp + (uint32_t)(16 * c), w);
          }
        } else if (hf == 0) {
          uint32_t w[16];
          const float p0 = ex2(fmaf(__uint_as_float(v[0][0]), kLog2e, -m_new));
          s4[0] = p0;
          w[0] = pack2(p0, 0.f);
#pragma unroll
          for (int i = 1; i < 16; ++i) w[i] = 0u;
          tmem_st_32x32b_x16(ts, w);                                   // 16 tail keys -> 8 columns used by the MMA
        }
        const float sum = (s4[0] + s4[1]) + (s4[2] + s4[3]);
        (void)nvalid; (void)col0;
        if (trc) TR(g * 16 + 4);
        tmem_st_wait();
        ptx::tcgen05_fence_before();
        ptx::mbar_arrive(&bars->p_full[b]);
        if (trc) TR(g * 16 + 5);
        l += sum;
        if (trc) TR(g * 16 + 7);
      }
      // end of the tile: once the last P V has retired read the accumulator, combine the two halves' row sums,
      // normalise and store
      {
        const int g = t * kBlocks + kBlocks - 1;
        xsum[hf * 128 + r] = l;
        ptx::mbar_wait(&bars->o_done[g & 1], (uint32_t)((g >> 1) & 1));
        ptx::tcgen05_fence_after();
        uint32_t vo[32];
        ptx::tmem_ld_32x32b_x32(to, vo);
        ptx::tmem_ld_wait();
        ptx::tcgen05_fence_before();
        ptx::mbar_arrive(&bars->o_free[t & 1]);
        named_bar_sync(1 + q, 64);
        const float inv = 1.0f / (l + xsum[(hf ^ 1) * 128 + r]);
        const int row = t * kQTile + r;
        if (row < kS) {
          uint4* dst = reinterpret_cast<uint4*>(out + ((size_t)f * kS + row) * (kHeads * kD) + h * kD + hf * 32);
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            float e[8];
#pragma unroll
            for (int i = 0; i < 8; ++i) e[i] = __uint_as_float(vo[8 * c + i]) * inv;
            dst[c] = make_uint4(pack2(e[0], e[1]), pack2(e[2], e[3]), pack2(e[4], e[5]), pack2(e[6], e[7]));
          }
        }
        named_bar_sync(1 + q, 64);                                     // xsum is rewritten by the next tile
      }
    }
  }

  ptx::tcgen05_fence_before();
  __syncthreads();
  ptx::tcgen05_fence_after();
  if (warp == 10) ptx::tmem_dealloc(tmem_base, kTmemCols);
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static EncodeTiledFn encode_fn() {
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
}  // namespace at5

int attention_impl(const void* qkv, int n_frames_total, void* out, cudaStream_t stream, long long* trace);
int attention(const void* qkv, int n_frames_total, void* out, cudaStream_t stream) {
  return attention_impl(qkv, n_frames_total, out, stream, nullptr);
}
int attention_impl(const void* qkv, int n_frames_total, void* out, cudaStream_t stream, long long* trace) {
  using namespace at5;
  EGPT_CHECK_ARG(qkv && out && n_frames_total > 0, "attention: bad arguments");
  EGPT_CHECK_ARG(((uintptr_t)qkv % 16 == 0) && ((uintptr_t)out % 16 == 0), "attention: pointers must be 16-byte aligned");
  EncodeTiledFn fn = encode_fn();
  if (!fn) { set_error("cuTensorMapEncodeTiled entry point not available"); return EGPT_ERR_CUDA; }
  // qkv viewed as [F][577][3072]; rows past 577 of a frame are out of bounds -> zero filled
  CUtensorMap tm;
  cuuint64_t dims[3] = {3 * kHeads * kD, (cuuint64_t)kS, (cuuint64_t)n_frames_total};
  cuuint64_t strides[2] = {(cuuint64_t)3 * kHeads * kD * 2, (cuuint64_t)kS * 3 * kHeads * kD * 2};
  cuuint32_t box[3] = {kD, 64, 1};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = fn(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(qkv), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("attention: cuTensorMapEncodeTiled failed with CUresult %d", (int)r); return EGPT_ERR_CUDA; }
  CUtensorMap tm_tail;
  cuuint32_t box_tail[3] = {kD, kTail, 1};
  r = fn(&tm_tail, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(qkv), dims, strides, box_tail, estr,
         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { set_error("attention: cuTensorMapEncodeTiled (tail) failed with CUresult %d", (int)r); return EGPT_ERR_CUDA; }
  // per launch: the attribute is per device, and one process may drive several GPUs
  EGPT_CUDA(cudaFuncSetAttribute(attention_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemTotal));
  ProfScope prof(PROF_ATTN, 4.0 * kS * kS * kD * kHeads * (double)n_frames_total, stream);
  attention_tc_kernel<<<n_frames_total * kHeads, kThreads, kSmemTotal, stream>>>(tm, tm_tail, reinterpret_cast<__nv_bfloat16*>(out), trace);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}

}  // namespace egpt

// developer aid: clock64 timestamps of CTA 0's phases (softmax warp 4 lane 0: slots g*16+k; MMA thread: 256+g*8+k)
extern "C" int egpt_attention_trace(const uint16_t* qkv, int n_frames_total, uint16_t* out, long long* trace, void* stream) {
  return egpt::attention_impl(qkv, n_frames_total, out, (cudaStream_t)stream, trace);
}

extern "C" int egpt_attention(const uint16_t* qkv, int n_frames_total, uint16_t* out, void* stream) {
  return egpt::attention(qkv, n_frames_total, out, (cudaStream_t)stream);
}
