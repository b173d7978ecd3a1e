// Multi-head self-attention of the CLIP tower: softmax(Q K^T) V per (frame, head),
// S = 577 tokens, 16 heads x 64 (transformers modeling_clip.py CLIPAttention, called
// from model/EventChatModel.py:196).  Input is the packed QKV projection
// [F*577, 3072] (q pre-scaled by 1/8 in the weights); output bf16 [F*577, 1024].
//
// One CTA per (frame, head): the whole K and V of the head (2 x 577 x 64 bf16) are
// staged once in shared memory with cp.async; each of the 8 warps then walks 16-row
// query tiles with flash-style online softmax in fp32 (exp2 domain), the two matmuls
// on mma.sync m16n8k16 bf16 tensor-core tiles fed by ldmatrix.
#include "common.cuh"

namespace egpt {

constexpr int kS = EGPT_TOKENS;          // 577
constexpr int kD = 64;
constexpr int kHeads = 16;
constexpr int kQKVStride = 3 * kHeads * kD;     // 3072
constexpr int kKeyPad = 640;             // keys padded to a multiple of 64
constexpr int kPitch = 72;               // smem row pitch in elements (144 B): conflict-free ldmatrix
constexpr int kAttnThreads = 256;
constexpr int kAttnSmem = 2 * kKeyPad * kPitch * 2;   // 184320 B
constexpr int kQTiles = (kS + 15) / 16;  // 37

__device__ __forceinline__ void cp_async16(void* smem, const void* gmem) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem)), "l"(gmem));
}
__device__ __forceinline__ void ldmatrix_x4(uint32_t (&r)[4], const void* smem) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"((uint32_t)__cvta_generic_to_shared(smem)));
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t (&r)[4], const void* smem) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"((uint32_t)__cvta_generic_to_shared(smem)));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

__global__ void __launch_bounds__(kAttnThreads, 1)
attention_kernel(const __nv_bfloat16* __restrict__ qkv, __nv_bfloat16* __restrict__ out) {
  extern __shared__ __align__(16) uint8_t smem_attn[];
  __nv_bfloat16* sK = reinterpret_cast<__nv_bfloat16*>(smem_attn);
  __nv_bfloat16* sV = sK + kKeyPad * kPitch;
  const int f = blockIdx.x / kHeads, h = blockIdx.x % kHeads;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const __nv_bfloat16* base = qkv + (size_t)f * kS * kQKVStride + h * kD;

  // stage K and V (rows >= 577 zero-filled)
  for (int i = threadIdx.x; i < kKeyPad * 8; i += kAttnThreads) {
    const int row = i >> 3, ch = i & 7;
    __nv_bfloat16* dk = sK + row * kPitch + ch * 8;
    __nv_bfloat16* dv = sV + row * kPitch + ch * 8;
    if (row < kS) {
      const __nv_bfloat16* src = base + (size_t)row * kQKVStride + ch * 8;
      cp_async16(dk, src + kHeads * kD);
      cp_async16(dv, src + 2 * kHeads * kD);
    } else {
      *reinterpret_cast<uint4*>(dk) = make_uint4(0, 0, 0, 0);
      *reinterpret_cast<uint4*>(dv) = make_uint4(0, 0, 0, 0);
    }
  }
  asm volatile("cp.async.commit_group;\ncp.async.wait_group 0;" ::: "memory");
  __syncthreads();

  const int g = lane >> 2, tig = lane & 3;
  const int lm = lane >> 3, lr = lane & 7;          // ldmatrix: matrix index, row within matrix
  constexpr float kLog2e = 1.4426950408889634f;

  for (int qt = warp; qt < kQTiles; qt += kAttnThreads / 32) {
    const int r0 = qt * 16;
    // Q fragments straight from global (A operand, row-major 16 x 64)
    uint32_t qa[4][4];
    {
      const int ra = r0 + g, rb = r0 + g + 8;
      const uint32_t* pa = reinterpret_cast<const uint32_t*>(base + (size_t)ra * kQKVStride);
      const uint32_t* pb = reinterpret_cast<const uint32_t*>(base + (size_t)rb * kQKVStride);
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        qa[ks][0] = ra < kS ? __ldg(pa + ks * 8 + tig) : 0u;
        qa[ks][1] = rb < kS ? __ldg(pb + ks * 8 + tig) : 0u;
        qa[ks][2] = ra < kS ? __ldg(pa + ks * 8 + 4 + tig) : 0u;
        qa[ks][3] = rb < kS ? __ldg(pb + ks * 8 + 4 + tig) : 0u;
      }
    }
    float o[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
    float m0 = -INFINITY, m1 = -INFINITY, l0 = 0.f, l1 = 0.f;

#pragma unroll 1
    for (int kb = 0; kb < kKeyPad / 64; ++kb) {
      const int key0 = kb * 64;
      float s[8][4];
#pragma unroll
      for (int j = 0; j < 8; ++j) s[j][0] = s[j][1] = s[j][2] = s[j][3] = 0.f;
      // S = Q K^T : n-tile j = keys key0 + 8j .. +7
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        uint32_t kf[4];
        const __nv_bfloat16* kp = sK + (key0 + j * 8 + lr) * kPitch + lm * 8;
        ldmatrix_x4(kf, kp);                 // dims 0..31  -> (b0,b1) for ks = 0, 1
        mma_bf16(s[j], qa[0], kf[0], kf[1]);
        mma_bf16(s[j], qa[1], kf[2], kf[3]);
        ldmatrix_x4(kf, kp + 32);            // dims 32..63 -> ks = 2, 3
        mma_bf16(s[j], qa[2], kf[0], kf[1]);
        mma_bf16(s[j], qa[3], kf[2], kf[3]);
      }
      // to the exp2 domain, mask the padded keys
      float bm0 = -INFINITY, bm1 = -INFINITY;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const int kc = key0 + j * 8 + tig * 2;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const bool valid = (kc + (e & 1)) < kS;
          s[j][e] = valid ? s[j][e] * kLog2e : -INFINITY;
        }
        bm0 = fmaxf(bm0, fmaxf(s[j][0], s[j][1]));
        bm1 = fmaxf(bm1, fmaxf(s[j][2], s[j][3]));
      }
      bm0 = fmaxf(bm0, __shfl_xor_sync(0xffffffffu, bm0, 1));
      bm0 = fmaxf(bm0, __shfl_xor_sync(0xffffffffu, bm0, 2));
      bm1 = fmaxf(bm1, __shfl_xor_sync(0xffffffffu, bm1, 1));
      bm1 = fmaxf(bm1, __shfl_xor_sync(0xffffffffu, bm1, 2));
      const float n0 = fmaxf(m0, bm0), n1 = fmaxf(m1, bm1);     // finite: every block holds >= 1 valid key
      const float sc0 = exp2f(m0 - n0), sc1 = exp2f(m1 - n1);
      m0 = n0; m1 = n1;
      float rs0 = 0.f, rs1 = 0.f;
      uint32_t pa[4][4];                   // P as A fragments for the 4 k-steps of 16 keys
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float p0 = exp2f(s[j][0] - n0), p1 = exp2f(s[j][1] - n0);
        const float p2 = exp2f(s[j][2] - n1), p3 = exp2f(s[j][3] - n1);
        rs0 += p0 + p1; rs1 += p2 + p3;
        pa[j >> 1][(j & 1) * 2 + 0] = pack2(p0, p1);
        pa[j >> 1][(j & 1) * 2 + 1] = pack2(p2, p3);
      }
      l0 = l0 * sc0 + rs0; l1 = l1 * sc1 + rs1;
#pragma unroll
      for (int j = 0; j < 8; ++j) { o[j][0] *= sc0; o[j][1] *= sc0; o[j][2] *= sc1; o[j][3] *= sc1; }
      // O += P V : k-step kk = keys key0 + 16kk .. +15, n-tile jd = dims 8jd .. +7
#pragma unroll
      for (int kk = 0; kk < 4; ++kk) {
#pragma unroll
        for (int jd = 0; jd < 8; jd += 2) {
          uint32_t vf[4];
          const __nv_bfloat16* vp = sV + (key0 + kk * 16 + (lm & 1) * 8 + lr) * kPitch + (jd + (lm >> 1)) * 8;
          ldmatrix_x4_trans(vf, vp);
          mma_bf16(o[jd], pa[kk], vf[0], vf[1]);
          mma_bf16(o[jd + 1], pa[kk], vf[2], vf[3]);
        }
      }
    }
    // row sums live in 4 lanes (tig); finish and store
    l0 += __shfl_xor_sync(0xffffffffu, l0, 1); l0 += __shfl_xor_sync(0xffffffffu, l0, 2);
    l1 += __shfl_xor_sync(0xffffffffu, l1, 1); l1 += __shfl_xor_sync(0xffffffffu, l1, 2);
    const float i0 = 1.f / l0, i1 = 1.f / l1;
    const int ra = r0 + g, rb = r0 + g + 8;
    __nv_bfloat16* oa = out + ((size_t)f * kS + ra) * (kHeads * kD) + h * kD;
    __nv_bfloat16* ob = out + ((size_t)f * kS + rb) * (kHeads * kD) + h * kD;
#pragma unroll
    for (int j = 0; j < 8; ++j) {
      if (ra < kS) *reinterpret_cast<uint32_t*>(oa + j * 8 + tig * 2) = pack2(o[j][0] * i0, o[j][1] * i0);
      if (rb < kS) *reinterpret_cast<uint32_t*>(ob + j * 8 + tig * 2) = pack2(o[j][2] * i1, o[j][3] * i1);
    }
  }
}

int attention(const void* qkv, int n_frames_total, void* out, cudaStream_t stream) {
  EGPT_CHECK_ARG(qkv && out && n_frames_total > 0, "attention: bad arguments");
  static bool attr_set = false;
  if (!attr_set) {
    EGPT_CUDA(cudaFuncSetAttribute(attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAttnSmem));
    attr_set = true;
  }
  ProfScope prof(PROF_ATTN, 4.0 * kS * kS * kD * kHeads * (double)n_frames_total, stream);
  attention_kernel<<<n_frames_total * kHeads, kAttnThreads, kAttnSmem, stream>>>(
      reinterpret_cast<const __nv_bfloat16*>(qkv), reinterpret_cast<__nv_bfloat16*>(out));
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}

}  // namespace egpt

namespace egpt { int attention_tc(const void* qkv, int n_frames_total, void* out, cudaStream_t stream); }

extern "C" int egpt_attention(const uint16_t* qkv, int n_frames_total, uint16_t* out, void* stream) {
  return egpt::attention_tc(qkv, n_frames_total, out, (cudaStream_t)stream);
}
// bring-up comparator (tests only): the mma.sync kernel
extern "C" int egpt_attention_legacy(const uint16_t* qkv, int n_frames_total, uint16_t* out, void* stream) {
  return egpt::attention(qkv, n_frames_total, out, (cudaStream_t)stream);
}
