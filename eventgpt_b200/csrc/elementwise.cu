// Bandwidth-bound glue kernels of the visual-token stage: LayerNorm, embedding assembly,
// im2col, fp32 -> bf16, spatio-temporal pooling, weight packing.
#include "common.cuh"

namespace egpt {

constexpr int kHid = EGPT_HIDDEN;      // 1024
constexpr int kTok = EGPT_TOKENS;      // 577
constexpr float kLnEps = 1e-5f;

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float a, float b) {
  __nv_bfloat162 v = __floats2bfloat162_rn(a, b);
  return *reinterpret_cast<uint32_t*>(&v);
}

// One warp per row of 1024 fp32: two-pass mean / variance in registers (torch.layer_norm
// numerics), affine, output bf16 (OUT_BF16) or fp32 in place.
template <bool OUT_BF16>
__device__ __forceinline__ void ln_row(const float (&x)[32], const float* __restrict__ gamma,
                                       const float* __restrict__ beta, int lane, void* __restrict__ out_row) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 32; ++i) s += x[i];
  const float mean = warp_sum(s) * (1.0f / kHid);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 32; ++i) { const float d = x[i] - mean; q += d * d; }
  const float rstd = rsqrtf(warp_sum(q) * (1.0f / kHid) + kLnEps);
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const int col = c * 128 + lane * 4;
    const float4 gm = __ldg(reinterpret_cast<const float4*>(gamma + col));
    const float4 bt = __ldg(reinterpret_cast<const float4*>(beta + col));
    const float y0 = (x[4 * c + 0] - mean) * rstd * gm.x + bt.x;
    const float y1 = (x[4 * c + 1] - mean) * rstd * gm.y + bt.y;
    const float y2 = (x[4 * c + 2] - mean) * rstd * gm.z + bt.z;
    const float y3 = (x[4 * c + 3] - mean) * rstd * gm.w + bt.w;
    if (OUT_BF16) {
      uint2 o = make_uint2(pack_bf16x2(y0, y1), pack_bf16x2(y2, y3));
      *reinterpret_cast<uint2*>(reinterpret_cast<__nv_bfloat16*>(out_row) + col) = o;
    } else {
      *reinterpret_cast<float4*>(reinterpret_cast<float*>(out_row) + col) = make_float4(y0, y1, y2, y3);
    }
  }
}

// y = LayerNorm(h) as bf16 (input of the next GEMM); h fp32 [rows,1024]
__global__ void __launch_bounds__(256)
layernorm_bf16_kernel(const float* __restrict__ h, const float* __restrict__ gamma, const float* __restrict__ beta,
                      __nv_bfloat16* __restrict__ y, long long rows) {
  // Rows are walked from the last to the first: the GEMM that produced h wrote its rows in ascending order, so the tail of h
  // is what is still in L2 when this kernel starts; and the rows of y written last (the first ones) are the ones the next
  // GEMM reads first.
  const long long row = (long long)(gridDim.x - 1 - blockIdx.x) * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  float x[32];
  const float4* src = reinterpret_cast<const float4*>(h + row * kHid);
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const float4 v = src[c * 32 + lane];
    x[4 * c] = v.x; x[4 * c + 1] = v.y; x[4 * c + 2] = v.z; x[4 * c + 3] = v.w;
  }
  ln_row<true>(x, gamma, beta, lane, y + row * kHid);
}

// CLIPVisionEmbeddings + pre_layrnorm: h[f,0] = LN(cls + pos[0]); h[f,1+p] = LN(patch[f,p] + pos[1+p])
__global__ void __launch_bounds__(256)
embed_preln_kernel(const float* __restrict__ patch, const float* __restrict__ cls, const float* __restrict__ pos,
                   const float* __restrict__ gamma, const float* __restrict__ beta, float* __restrict__ h,
                   __nv_bfloat16* __restrict__ hb, float* __restrict__ stats, long long rows) {
  const long long row = (long long)blockIdx.x * 8 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const long long f = row / kTok;
  const int tok = (int)(row - f * kTok);
  const float4* src = tok == 0 ? reinterpret_cast<const float4*>(cls)
                               : reinterpret_cast<const float4*>(patch + (f * (kTok - 1) + tok - 1) * kHid);
  const float4* pp = reinterpret_cast<const float4*>(pos + (size_t)tok * kHid);
  float x[32];
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const float4 v = src[c * 32 + lane];
    const float4 p = __ldg(pp + c * 32 + lane);
    x[4 * c] = v.x + p.x; x[4 * c + 1] = v.y + p.y; x[4 * c + 2] = v.z + p.z; x[4 * c + 3] = v.w + p.w;
  }
  ln_row<false>(x, gamma, beta, lane, h + row * kHid);
  // what the first encoder layer's LayerNorm-folding GEMM needs: a bf16 copy of the row and its sum / sum of squares
  // (ln_row left the row in h; every lane re-reads the 32 values it wrote)
  __syncwarp();
  float s1 = 0.f, s2 = 0.f;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    const int col = c * 128 + lane * 4;
    const float4 v = *reinterpret_cast<const float4*>(h + row * kHid + col);
    s1 += (v.x + v.y) + (v.z + v.w);
    s2 = fmaf(v.x, v.x, fmaf(v.y, v.y, fmaf(v.z, v.z, fmaf(v.w, v.w, s2))));
    *reinterpret_cast<uint2*>(hb + row * kHid + col) = make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
  }
  s1 = warp_sum(s1);
  s2 = warp_sum(s2);
  if (lane < 4) *reinterpret_cast<float2*>(stats + row * 8 + 2 * lane) = lane == 0 ? make_float2(s1, s2) : make_float2(0.f, 0.f);
}

// LayerNorm folded into the linear layer that follows it (model set-up, once per layer):
//   LN(h) W^T + b = rstd * (h (gamma . W)^T - mean * colsum) + (beta W^T + b),  colsum_n = sum_k (gamma . W)_nk
// w [N, K] bf16 is overwritten with bf16(gamma_k * w_nk); colsum is taken over the ROUNDED products (what the tensor
// cores will multiply), bias_out over the original weights.  One warp per output row.
__global__ void __launch_bounds__(256)
ln_fold_kernel(__nv_bfloat16* __restrict__ w, const float* __restrict__ gamma, const float* __restrict__ beta,
               const float* __restrict__ bias, float* __restrict__ colsum, float* __restrict__ bias_out, int N, int K) {
  const int n = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (n >= N) return;
  __nv_bfloat16* row = w + (size_t)n * K;
  float c = 0.f, d = 0.f;
  for (int k = lane; k < K; k += 32) {
    const float wv = __bfloat162float(row[k]);
    const __nv_bfloat16 f = __float2bfloat16_rn(gamma[k] * wv);
    row[k] = f;
    c += __bfloat162float(f);
    d = fmaf(beta[k], wv, d);
  }
  c = warp_sum(c);
  d = warp_sum(d);
  if (lane == 0) { colsum[n] = c; bias_out[n] = d + bias[n]; }
}

// fp32 -> bf16, 8 elements per thread
__global__ void __launch_bounds__(256)
f32_to_bf16_kernel(const float* __restrict__ in, __nv_bfloat16* __restrict__ out, long long n8) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n8; i += (long long)gridDim.x * blockDim.x) {
    const float4 a = reinterpret_cast<const float4*>(in)[2 * i], b = reinterpret_cast<const float4*>(in)[2 * i + 1];
    reinterpret_cast<uint4*>(out)[i] =
        make_uint4(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w), pack_bf16x2(b.x, b.y), pack_bf16x2(b.z, b.w));
  }
}

// pixel_values bf16 [F,3,336,336] -> im2col rows [F*576, 640] (col = c*196 + dy*14 + dx, zero padded)
__global__ void __launch_bounds__(256)
im2col_kernel(const uint16_t* __restrict__ pix, uint16_t* __restrict__ patches, long long total) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const int col = (int)(i % 640);
    const long long r = i / 640;
    uint16_t v = 0;
    if (col < 588) {
      const long long f = r / 576;
      const int p = (int)(r - f * 576), py = p / 24, px = p - py * 24;
      const int c = col / 196, rem = col - c * 196, dy = rem / 14, dx = rem - dy * 14;
      v = pix[((f * 3 + c) * 336 + py * 14 + dy) * 336 + px * 14 + dx];
    }
    patches[i] = v;
  }
}

// ---- spatio-temporal pooling (model/EventChatModel.py:15-38) --------------------------------
// features [B,t,s,c]; grid (c/ (threads*VEC), s_chunks, B).  Each CTA covers `rows_per_chunk`
// spatial positions: writes the spatial rows (mean over t) directly and a partial temporal
// sum per frame into `partial` [B, chunks, t, c] fp32; pool_finish sums the chunks in order.
template <typename T> struct Vec8;
template <> struct Vec8<__nv_bfloat16> {
  typedef uint4 Raw;                                    // 8 values as loaded (kept packed while in flight)
  static __device__ __forceinline__ Raw load_raw(const __nv_bfloat16* p) { return *reinterpret_cast<const uint4*>(p); }
  static __device__ __forceinline__ void unpack(const Raw& u, float (&v)[8]) {
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __uint_as_float(w[i] << 16);
      v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  static __device__ __forceinline__ void load(const __nv_bfloat16* p, float (&v)[8]) {
    const uint4 u = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      v[2 * i] = __uint_as_float(w[i] << 16);
      v[2 * i + 1] = __uint_as_float(w[i] & 0xffff0000u);
    }
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, const float (&v)[8]) {
    *reinterpret_cast<uint4*>(p) = make_uint4(pack_bf16x2(v[0], v[1]), pack_bf16x2(v[2], v[3]),
                                              pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
  }
};
template <> struct Vec8<float> {
  struct Raw { float4 a, b; };
  static __device__ __forceinline__ Raw load_raw(const float* p) {
    return Raw{reinterpret_cast<const float4*>(p)[0], reinterpret_cast<const float4*>(p)[1]};
  }
  static __device__ __forceinline__ void unpack(const Raw& r, float (&v)[8]) {
    v[0] = r.a.x; v[1] = r.a.y; v[2] = r.a.z; v[3] = r.a.w; v[4] = r.b.x; v[5] = r.b.y; v[6] = r.b.z; v[7] = r.b.w;
  }
  static __device__ __forceinline__ void load(const float* p, float (&v)[8]) {
    const float4 a = reinterpret_cast<const float4*>(p)[0], b = reinterpret_cast<const float4*>(p)[1];
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
  }
  static __device__ __forceinline__ void store(float* p, const float (&v)[8]) {
    reinterpret_cast<float4*>(p)[0] = make_float4(v[0], v[1], v[2], v[3]);
    reinterpret_cast<float4*>(p)[1] = make_float4(v[4], v[5], v[6], v[7]);
  }
};

constexpr int kPoolMaxT = 8;
constexpr int kPoolRows = 32;     // spatial positions per CTA

template <typename T, int TMAX>
__global__ void __launch_bounds__(128)
pool_kernel(const T* __restrict__ feat, int t, int s, int c, int n_temporal, T* __restrict__ out,
            float* __restrict__ partial) {
  const int b = blockIdx.z, chunk = blockIdx.y, chunks = gridDim.y;
  const int col = (blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (col >= c) return;
  const int s0 = chunk * kPoolRows, s1 = min(s, s0 + kPoolRows);
  float tacc[TMAX][8];
#pragma unroll
  for (int i = 0; i < TMAX; ++i)
#pragma unroll
    for (int j = 0; j < 8; ++j) tacc[i][j] = 0.f;
  const float inv_t = 1.0f / (float)t;
  // two spatial positions per iteration: 2 t independent 16-byte loads in flight per thread, kept packed until used
  // (the kernel is a pure stream; the additions happen in the same order as a one-row-at-a-time loop)
  for (int si = s0; si < s1; si += 2) {
    const bool two = si + 1 < s1;
    typename Vec8<T>::Raw raw[2][TMAX];
#pragma unroll
    for (int ti = 0; ti < TMAX; ++ti) {
      if (ti < t) {
        const T* src = feat + (((size_t)b * t + ti) * s + si) * c + col;
        raw[0][ti] = Vec8<T>::load_raw(src);
        if (two) raw[1][ti] = Vec8<T>::load_raw(src + c);
      }
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      if (r == 1 && !two) break;
      float sp[8] = {0, 0, 0, 0, 0, 0, 0, 0};
#pragma unroll
      for (int ti = 0; ti < TMAX; ++ti) {
        if (ti < t) {
          float v[8];
          Vec8<T>::unpack(raw[r][ti], v);
#pragma unroll
          for (int j = 0; j < 8; ++j) { sp[j] += v[j]; tacc[ti][j] += v[j]; }
        }
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) sp[j] *= inv_t;
      Vec8<T>::store(out + ((size_t)b * (n_temporal + s) + n_temporal + si + r) * c + col, sp);
    }
  }
#pragma unroll
  for (int ti = 0; ti < TMAX; ++ti)
    if (ti < t) {
      float* p = partial + (((size_t)b * chunks + chunk) * t + ti) * c + col;
      reinterpret_cast<float4*>(p)[0] = make_float4(tacc[ti][0], tacc[ti][1], tacc[ti][2], tacc[ti][3]);
      reinterpret_cast<float4*>(p)[1] = make_float4(tacc[ti][4], tacc[ti][5], tacc[ti][6], tacc[ti][7]);
    }
}

// Any number of frames (t > 8: the reference takes any n, common/common.py:17-37).  Same grid, same partial-sum layout and the same
// order of additions as pool_kernel, but nothing is kept per frame in registers: the CTA's slab (32 positions x t frames) is walked
// twice, once position-major for the spatial means and once frame-major for the temporal partial sums (the second walk hits L2).
template <typename T>
__global__ void __launch_bounds__(128)
pool_any_kernel(const T* __restrict__ feat, int t, int s, int c, int n_temporal, T* __restrict__ out, float* __restrict__ partial) {
  const int b = blockIdx.z, chunk = blockIdx.y, chunks = gridDim.y;
  const int col = (blockIdx.x * blockDim.x + threadIdx.x) * 8;
  if (col >= c) return;
  const int s0 = chunk * kPoolRows, s1 = min(s, s0 + kPoolRows);
  const float inv_t = 1.0f / (float)t;
  for (int si = s0; si < s1; ++si) {
    float sp[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int ti = 0; ti < t; ++ti) {
      float v[8];
      Vec8<T>::load(feat + (((size_t)b * t + ti) * s + si) * c + col, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) sp[j] += v[j];
    }
#pragma unroll
    for (int j = 0; j < 8; ++j) sp[j] *= inv_t;
    Vec8<T>::store(out + ((size_t)b * (n_temporal + s) + n_temporal + si) * c + col, sp);
  }
  for (int ti = 0; ti < t; ++ti) {
    float acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    for (int si = s0; si < s1; ++si) {
      float v[8];
      Vec8<T>::load(feat + (((size_t)b * t + ti) * s + si) * c + col, v);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += v[j];
    }
    float* p = partial + (((size_t)b * chunks + chunk) * t + ti) * c + col;
    reinterpret_cast<float4*>(p)[0] = make_float4(acc[0], acc[1], acc[2], acc[3]);
    reinterpret_cast<float4*>(p)[1] = make_float4(acc[4], acc[5], acc[6], acc[7]);
  }
}

template <typename T>
__global__ void pool_finish_kernel(const float* __restrict__ partial, int chunks, int t, int s, int c, int n_temporal,
                                   T* __restrict__ out, long long total) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;   // over B * n_temporal * c
  if (i >= total) return;
  const int col = (int)(i % c);
  const int ti = (int)((i / c) % n_temporal);
  const long long b = i / ((long long)c * n_temporal);
  float acc = 0.f;
  if (ti < t)
    for (int k = 0; k < chunks; ++k) acc += partial[((b * chunks + k) * t + ti) * c + col];
  const float v = ti < t ? acc / (float)s : 0.f;      // zero rows when num_temporal_tokens > t (:29-31)
  T* o = out + (b * (n_temporal + s) + ti) * c + col;
  if (sizeof(T) == 2) *reinterpret_cast<__nv_bfloat16*>(o) = __float2bfloat16_rn(v);
  else *reinterpret_cast<float*>(o) = v;
}

// ---- multimodal splice gather (model/EventChatModel.py:355-465) -----------------------------------
// out[row, :] = embed_table[src[row]] (src >= 0), zero (src == -1, right padding), or
// event_features row -(src + 2) (src <= -2; rows of the [NF*T, C] feature matrix).  16-byte vectors.
__global__ void __launch_bounds__(256)
splice_gather_kernel(const uint4* __restrict__ table, const uint4* __restrict__ feats, const int32_t* __restrict__ src,
                     uint4* __restrict__ out, long long rows, int vec_per_row) {
  const long long total = rows * vec_per_row;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long row = i / vec_per_row;
    const int v = (int)(i - row * vec_per_row);
    const int s = __ldg(src + row);
    uint4 val = make_uint4(0u, 0u, 0u, 0u);
    if (s >= 0) val = __ldg(table + (size_t)s * vec_per_row + v);
    else if (s <= -2) val = __ldg(feats + (size_t)(-(long long)s - 2) * vec_per_row + v);
    out[i] = val;
  }
}

// ---- weight packing ------------------------------------------------------------------------------
__global__ void scale_bf16_kernel(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out, float scale,
                                  long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[i] = __float2bfloat16_rn(__bfloat162float(in[i]) * scale);
}
__global__ void scale_f32_kernel(const float* __restrict__ in, float* __restrict__ out, float scale, long long n) {
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x)
    out[i] = in[i] * scale;
}
// [1024, 588] -> [1024, 640] zero padded
__global__ void pad_patch_weight_kernel(const __nv_bfloat16* __restrict__ in, __nv_bfloat16* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= kHid * 640) return;
  const int r = i / 640, c = i - r * 640;
  out[i] = c < 588 ? in[r * 588 + c] : __float2bfloat16_rn(0.f);
}

// ---- launchers -----------------------------------------------------------------------------------
int layernorm_bf16(const float* h, const float* gamma, const float* beta, void* y, long long rows, cudaStream_t st) {
  ProfScope prof(PROF_ELEMENTWISE, (double)rows * kHid * 6, st);
  layernorm_bf16_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(h, gamma, beta, (__nv_bfloat16*)y, rows);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}
int embed_preln(const float* patch, const float* cls, const float* pos, const float* gamma, const float* beta, float* h,
                void* hb, float* stats, long long rows, cudaStream_t st) {
  ProfScope prof(PROF_ELEMENTWISE, (double)rows * kHid * 10, st);
  embed_preln_kernel<<<(unsigned)((rows + 7) / 8), 256, 0, st>>>(patch, cls, pos, gamma, beta, h, (__nv_bfloat16*)hb, stats, rows);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}
int ln_fold(void* w, const float* gamma, const float* beta, const float* bias, float* colsum, float* bias_out, int N, int K,
            cudaStream_t st) {
  ln_fold_kernel<<<(N + 7) / 8, 256, 0, st>>>((__nv_bfloat16*)w, gamma, beta, bias, colsum, bias_out, N, K);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}
int f32_to_bf16(const float* in, void* out, long long n, cudaStream_t st) {
  const long long n8 = n / 8;
  ProfScope prof(PROF_ELEMENTWISE, (double)n * 6, st);
  long long blocks = (n8 + 255) / 256;
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  f32_to_bf16_kernel<<<(unsigned)blocks, 256, 0, st>>>(in, (__nv_bfloat16*)out, n8);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}
int im2col(const uint16_t* pix, uint16_t* patches, int F, cudaStream_t st) {
  const long long total = (long long)F * 576 * 640;
  ProfScope prof(PROF_ELEMENTWISE, (double)F * 3 * 336 * 336 * 2 + (double)total * 2, st);
  long long blocks = (total + 255) / 256;
  if (blocks > kNumSMs * 32) blocks = kNumSMs * 32;
  im2col_kernel<<<(unsigned)blocks, 256, 0, st>>>(pix, patches, total);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}
int scale_bf16(const void* in, void* out, float scale, long long n, cudaStream_t st) {
  scale_bf16_kernel<<<(unsigned)((n + 255) / 256 > 4096 ? 4096 : (n + 255) / 256), 256, 0, st>>>(
      (const __nv_bfloat16*)in, (__nv_bfloat16*)out, scale, n);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}
int scale_f32(const float* in, float* out, float scale, long long n, cudaStream_t st) {
  scale_f32_kernel<<<(unsigned)((n + 255) / 256 > 4096 ? 4096 : (n + 255) / 256), 256, 0, st>>>(in, out, scale, n);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}
int pad_patch_weight(const void* in, void* out, cudaStream_t st) {
  pad_patch_weight_kernel<<<(kHid * 640 + 255) / 256, 256, 0, st>>>((const __nv_bfloat16*)in, (__nv_bfloat16*)out);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}

static int pool_chunks(int s) { return (s + kPoolRows - 1) / kPoolRows; }

}  // namespace egpt

using namespace egpt;

extern "C" int egpt_splice_embeddings(const uint16_t* embed_table, const uint16_t* event_features, const int32_t* src,
                                      int64_t rows, int c, uint16_t* out, void* stream_) {
  cudaStream_t st = (cudaStream_t)stream_;
  EGPT_CHECK_ARG(embed_table && event_features && src && out, "egpt_splice_embeddings: null pointer argument");
  EGPT_CHECK_ARG(rows > 0 && c > 0 && c % 8 == 0, "egpt_splice_embeddings: rows=%lld c=%d (c must be a multiple of 8)", (long long)rows, c);
  EGPT_CHECK_ARG(((uintptr_t)embed_table % 16 == 0) && ((uintptr_t)event_features % 16 == 0) && ((uintptr_t)out % 16 == 0),
                 "egpt_splice_embeddings: pointers must be 16-byte aligned");
  ProfScope prof(PROF_ELEMENTWISE, 2.0 * (double)rows * c * 2, st);
  long long blocks = (rows * (c / 8) + 255) / 256;
  if (blocks > kNumSMs * 16) blocks = kNumSMs * 16;
  splice_gather_kernel<<<(unsigned)blocks, 256, 0, st>>>((const uint4*)embed_table, (const uint4*)event_features, src,
                                                         (uint4*)out, rows, c / 8);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}

extern "C" size_t egpt_pool_workspace_bytes(int batch, int t, int c) {
  if (batch <= 0 || t <= 0 || c <= 0) return 0;
  // sized for s up to 1024 spatial positions
  return align_up((size_t)batch * 32 * t * c * sizeof(float), 1024);
}

extern "C" int egpt_spatio_temporal_pool(const void* features, int is_f32, int batch, int t, int s, int c,
                                         int n_temporal, void* out, void* workspace, size_t workspace_bytes,
                                         void* stream_) {
  cudaStream_t st = (cudaStream_t)stream_;
  EGPT_CHECK_ARG(features && out && workspace, "egpt_spatio_temporal_pool: null pointer argument");
  EGPT_CHECK_ARG(batch > 0 && t > 0 && s > 0 && c > 0 && n_temporal >= 0, "egpt_spatio_temporal_pool: bad shape");
  EGPT_CHECK_ARG(c % 8 == 0, "egpt_spatio_temporal_pool: c=%d must be a multiple of 8", c);
  const int chunks = pool_chunks(s);
  const size_t need = (size_t)batch * chunks * t * c * sizeof(float);
  if (workspace_bytes < need) { set_error("egpt_spatio_temporal_pool: workspace %zu < %zu", workspace_bytes, need); return EGPT_ERR_WORKSPACE; }
  float* partial = (float*)workspace;
  const double esz = is_f32 ? 4.0 : 2.0;
  ProfScope prof(PROF_POOL, ((double)batch * t * s * c + (double)batch * (n_temporal + s) * c) * esz, st);
  dim3 grid((c / 8 + 127) / 128, chunks, batch);
  const long long total = (long long)batch * n_temporal * c;
  if (is_f32) {
    if (t <= 5) pool_kernel<float, 5><<<grid, 128, 0, st>>>((const float*)features, t, s, c, n_temporal, (float*)out, partial);
    else if (t > kPoolMaxT) pool_any_kernel<float><<<grid, 128, 0, st>>>((const float*)features, t, s, c, n_temporal, (float*)out, partial);
    else pool_kernel<float, kPoolMaxT><<<grid, 128, 0, st>>>((const float*)features, t, s, c, n_temporal, (float*)out, partial);
    EGPT_LAUNCH_CHECK();
    if (total > 0) {
      pool_finish_kernel<float><<<(unsigned)((total + 255) / 256), 256, 0, st>>>(partial, chunks, t, s, c, n_temporal, (float*)out, total);
      EGPT_LAUNCH_CHECK();
    }
  } else {
    if (t <= 5) pool_kernel<__nv_bfloat16, 5><<<grid, 128, 0, st>>>((const __nv_bfloat16*)features, t, s, c, n_temporal, (__nv_bfloat16*)out, partial);
    else if (t > kPoolMaxT) pool_any_kernel<__nv_bfloat16><<<grid, 128, 0, st>>>((const __nv_bfloat16*)features, t, s, c, n_temporal, (__nv_bfloat16*)out, partial);
    else pool_kernel<__nv_bfloat16, kPoolMaxT><<<grid, 128, 0, st>>>((const __nv_bfloat16*)features, t, s, c, n_temporal, (__nv_bfloat16*)out, partial);
    EGPT_LAUNCH_CHECK();
    if (total > 0) {
      pool_finish_kernel<__nv_bfloat16><<<(unsigned)((total + 255) / 256), 256, 0, st>>>(partial, chunks, t, s, c, n_temporal, (__nv_bfloat16*)out, total);
      EGPT_LAUNCH_CHECK();
    }
  }
  return EGPT_OK;
}
