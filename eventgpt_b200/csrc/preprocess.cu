// CLIP preprocessing on the device, bit-exact with the reference's CPU path.
//
// Replaces `event_processor(img, return_tensors='pt')['pixel_values'][0].to(bf16)`
// (common/common.py:124-127): transformers' CLIPImageProcessor = resize shortest
// edge -> 336 with Pillow's antialiased bicubic, centre crop 336x336, x 1/255,
// (v - mean) / std, HWC -> CHW, bf16.  Pillow's resampler (Resample.c) works in
// 8-bit fixed point: coefficients are doubles normalised to sum 1 and rounded to
// 22 fractional bits, each pass accumulates in int32 from 2^21 and clips to uint8;
// the horizontal pass runs first and its uint8 output feeds the vertical pass.
// Canvas sizes are data dependent (hard part 2), so the
// coefficient tables are computed on the device from the per-frame {h, w}; all
// double arithmetic uses the *_rn intrinsics so nvcc cannot contract it into FMAs
// and the tables equal the CPU's IEEE results bit for bit.
#include "common.cuh"

namespace egpt {

constexpr int kCrop = EGPT_CROP;
constexpr int kTaps = EGPT_MAX_TAPS;
constexpr int kPrecisionBits = 32 - 8 - 2;
constexpr int kTapInts = kCrop * (2 + kTaps);         // xmin[336], cnt[336], then the taps: horizontal axis k[taps][336] (lanes of a warp
                                                      // = neighbouring output columns read neighbouring words), vertical axis k[336][taps]
// ... then, for the byte-parallel fast path (at most 12 taps): the first 12 coefficients of every output index split into byte
// planes and packed four taps to a word, planes[336][12] = {plane 0 (bits 0-7, unsigned): 3 words, plane 1 (bits 8-15, unsigned):
// 3 words, plane 2 (bits 16-23, signed): 3 words, 3 words unused}; and the largest tap count of the axis.
constexpr int kPlaneOff = kTapInts, kPlaneInts = 12;
constexpr int kMaxCntOff = kPlaneOff + kCrop * kPlaneInts;
constexpr int kAxisInts = kMaxCntOff + 16;
constexpr int kPatchK = 640;                           // 3*14*14 = 588 padded to a multiple of 64

struct Geometry { int new_h, new_w, top, left; };

__device__ __forceinline__ Geometry resize_geometry(int h, int w) {
  // transformers get_resize_output_image_size(shortest_edge=336) + center_crop
  const int shrt = w <= h ? w : h, lng = w <= h ? h : w;
  const int new_long = (int)__ddiv_rn((double)(kCrop * (long long)lng), (double)shrt);
  Geometry g;
  g.new_h = w <= h ? new_long : kCrop;
  g.new_w = w <= h ? kCrop : new_long;
  g.top = (g.new_h - kCrop) / 2;
  g.left = (g.new_w - kCrop) / 2;
  return g;
}

__device__ __forceinline__ double bicubic_filter(double x) {
  // Resample.c bicubic_filter, a = -0.5
  if (x < 0.0) x = -x;
  if (x < 1.0) {
    double t = __dsub_rn(__dmul_rn(1.5, x), 2.5);
    t = __dmul_rn(__dmul_rn(t, x), x);
    return __dadd_rn(t, 1.0);
  }
  if (x < 2.0) {
    double t = __dmul_rn(__dsub_rn(x, 5.0), x);
    t = __dmul_rn(__dadd_rn(t, 8.0), x);
    return __dmul_rn(__dsub_rn(t, 4.0), -0.5);
  }
  return 0.0;
}

// grid (F, 2): axis 0 = horizontal (columns left..left+335 of the resized image),
// axis 1 = vertical (rows top..top+335).  One thread per output index.
__global__ void __launch_bounds__(352)
resize_coeffs_kernel(const int32_t* __restrict__ hw, int32_t* __restrict__ tables, int32_t* __restrict__ status) {
  __shared__ int smax;
  const int f = blockIdx.x, axis = blockIdx.y, o = threadIdx.x;
  const int h = hw[2 * f], w = hw[2 * f + 1];
  int32_t* tab = tables + ((size_t)f * 2 + axis) * kAxisInts;
  if (o == 0) smax = 0;
  __syncthreads();
  if (o < kCrop) {
    if (h <= 0 || w <= 0) {
      tab[o] = 0; tab[kCrop + o] = 0;
    } else {
      const Geometry g = resize_geometry(h, w);
      const int in_size = axis == 0 ? w : h;
      const int out_size = axis == 0 ? g.new_w : g.new_h;
      const int xx = (axis == 0 ? g.left : g.top) + o;

      // Resample.c precompute_coeffs
      const double scale = __ddiv_rn((double)in_size, (double)out_size);
      const double filterscale = scale < 1.0 ? 1.0 : scale;
      const double support = __dmul_rn(2.0, filterscale);
      const double ss = __ddiv_rn(1.0, filterscale);
      const double center = __dmul_rn(__dadd_rn((double)xx, 0.5), scale);
      int xmin = (int)__dadd_rn(__dsub_rn(center, support), 0.5);
      if (xmin < 0) xmin = 0;
      int xmax = (int)__dadd_rn(__dadd_rn(center, support), 0.5);
      if (xmax > in_size) xmax = in_size;
      int cnt = xmax - xmin;
      if (cnt > kTaps) { atomicOr(status, EGPT_FLAG_FILTER_TOO_WIDE); cnt = kTaps; }
      double wts[kTaps];
      double ww = 0.0;
      for (int x = 0; x < cnt; ++x) {
        const double arg = __dmul_rn(__dadd_rn(__dsub_rn((double)(x + xmin), center), 0.5), ss);
        const double v = bicubic_filter(arg);
        wts[x] = v;
        ww = __dadd_rn(ww, v);
      }
      // normalize_coeffs_8bpc
      uint32_t pl[3][3] = {{0u, 0u, 0u}, {0u, 0u, 0u}, {0u, 0u, 0u}};
      for (int x = 0; x < cnt; ++x) {
        double v = wts[x];
        if (ww != 0.0) v = __ddiv_rn(v, ww);
        const double sc = __dmul_rn(v, (double)(1 << kPrecisionBits));
        const int k = v < 0 ? (int)__dadd_rn(-0.5, sc) : (int)__dadd_rn(0.5, sc);
        tab[2 * kCrop + (axis == 0 ? x * kCrop + o : o * kTaps + x)] = k;
        if (x < 12) {
          // k = b0 + 2^8 b1 + 2^16 b2 with b0, b1 in [0, 255] and b2 = k >> 16 signed (|k| < 2^23)
          const uint32_t sh = (uint32_t)(x & 3) * 8u;
          pl[0][x >> 2] |= ((uint32_t)k & 0xffu) << sh;
          pl[1][x >> 2] |= (((uint32_t)k >> 8) & 0xffu) << sh;
          pl[2][x >> 2] |= (((uint32_t)(k >> 16)) & 0xffu) << sh;
        }
      }
      int32_t* pp = tab + kPlaneOff + o * kPlaneInts;
#pragma unroll
      for (int pi = 0; pi < 3; ++pi)
#pragma unroll
        for (int wi = 0; wi < 3; ++wi) pp[pi * 3 + wi] = (int32_t)pl[pi][wi];
      tab[o] = xmin;
      tab[kCrop + o] = cnt;
      atomicMax(&smax, cnt);
    }
  }
  __syncthreads();
  if (o == 0) tab[kMaxCntOff] = smax;
}

__device__ __forceinline__ int clip8(int v) {
  v >>= kPrecisionBits;
  return v < 0 ? 0 : (v > 255 ? 255 : v);
}

// ---- byte-parallel fast path (both axes need at most 12 taps: down-scale factors up to 2.75, i.e. every event camera) ----
// Four taps at a time with IDP.4A: the coefficient k (22 fractional bits, signed) is split into three byte planes and the
// exact integer sum rebuilt as d0 + 2^8 d1 + 2^16 d2.  That needs four consecutive taps of ONE channel in one register, so the
// horizontal pass first de-interleaves its staged rows into planar R / G / B rows, and hands its result to the vertical pass
// with four consecutive rows packed into each word (tmp[f][y / 4][o][c] as uint32: byte y % 4), which is exactly the operand
// the vertical taps want.  A tap window starts at an arbitrary byte: two or three funnel shifts align it.
__device__ __forceinline__ int dp4a_us(uint32_t a, uint32_t b, int c) {      // unsigned pixels x signed coefficient bytes
  int d;
  asm("dp4a.u32.s32 %0, %1, %2, %3;" : "=r"(d) : "r"(a), "r"(b), "r"(c));
  return d;
}
// NW packed words of taps starting at byte offset `sh8 / 8` of w[0]; planes pk[plane * 3 + word]
template <int NW>
__device__ __forceinline__ int tap_sum(const uint32_t (&w)[4], uint32_t sh8, const uint32_t (&pk)[9]) {
  uint32_t a[3];
#pragma unroll
  for (int i = 0; i < NW; ++i) a[i] = __funnelshift_r(w[i], w[i + 1], sh8);
  uint32_t d0 = 1u << (kPrecisionBits - 1), d1 = 0u;
  int d2 = 0;
#pragma unroll
  for (int i = 0; i < NW; ++i) {
    d0 = __dp4a(a[i], pk[i], d0);
    d1 = __dp4a(a[i], pk[3 + i], d1);
    d2 = dp4a_us(a[i], pk[6 + i], d2);
  }
  return (int)(d0 + (d1 << 8)) + (d2 << 16);
}

// Horizontal pass: tmp[f][y][o][c] for every input row y < h and the 336 cropped columns.
// grid (ceil(H/8), F), 352 threads.  The CTA first stages its 8 input rows in shared memory with 16-byte
// loads (rows are packed at an arbitrary byte offset, so each is copied from the enclosing aligned span);
// thread o then owns output column o and reads its taps as bytes from shared memory.
constexpr int kHRows = 8;
__host__ __device__ constexpr int planar_words(int pitch) { return pitch / 12 + 4; }   // words per planar row (>= w / 4 + 6)

// 0: generic path; 2 / 3: fast path with 8 / 12 taps.  Both passes of a frame must agree (the intermediate's layout differs).
__device__ __forceinline__ int frame_mode(const int32_t* __restrict__ tables, int f, int w, int pitch) {
  const bool staged = pitch > 0 && 3 * w + 32 <= pitch;
  const int mc = max(tables[((size_t)f * 2 + 0) * kAxisInts + kMaxCntOff], tables[((size_t)f * 2 + 1) * kAxisInts + kMaxCntOff]);
  return !staged ? 0 : (mc <= 8 ? 2 : (mc <= 12 ? 3 : 0));
}

template <int NW>
__device__ __forceinline__ void horizontal_fast(const uint8_t* __restrict__ srows, uint32_t* __restrict__ planar, int pitch, int w,
                                                int h, int y0, int y1, size_t frame_off, const int32_t* __restrict__ tab,
                                                uint32_t* __restrict__ dstq) {
  const int PW = planar_words(pitch);
  // de-interleave the staged rows: 4 pixels = 12 bytes -> one word of R, G and B each
  const int nq = (w + 3) >> 2;
  const int r = threadIdx.x / 44;                                      // 44 threads per row
  if (r < y1 - y0) {
    const uint32_t off0 = (uint32_t)((frame_off + (size_t)(y0 + r) * w * 3) & 15);
    const uint32_t sh8 = (off0 & 3u) * 8u;
    const uint32_t* wrow = reinterpret_cast<const uint32_t*>(srows + r * pitch + (off0 & ~3u));
    uint32_t* pr = planar + r * PW;
    for (int xq = threadIdx.x - r * 44; xq < nq; xq += 44) {
      const uint32_t* wp = wrow + 3 * xq;
      const uint32_t w0 = wp[0], w1 = wp[1], w2 = wp[2], w3 = wp[3];
      const uint32_t A = __funnelshift_r(w0, w1, sh8), B = __funnelshift_r(w1, w2, sh8), C = __funnelshift_r(w2, w3, sh8);
      // A = R0 G0 B0 R1, B = G1 B1 R2 G2, C = B2 R3 G3 B3
      pr[xq] = __byte_perm(__byte_perm(A, B, 0x0630), C, 0x5210);
      pr[kHRows * PW + xq] = __byte_perm(__byte_perm(A, B, 0x0741), C, 0x6210);
      pr[2 * kHRows * PW + xq] = __byte_perm(__byte_perm(A, B, 0x0052), C, 0x7410);
    }
  }
  __syncthreads();
  const int o = threadIdx.x;
  if (o >= kCrop) return;
  const int xmin = tab[o];
  const uint32_t sh8 = (uint32_t)(xmin & 3) * 8u;
  uint32_t pk[9];
  {
    const uint4* pp = reinterpret_cast<const uint4*>(tab + kPlaneOff + o * kPlaneInts);
    const uint4 a = __ldg(pp), b = __ldg(pp + 1);
    const uint32_t c = __ldg(reinterpret_cast<const uint32_t*>(pp + 2));
    pk[0] = a.x; pk[1] = a.y; pk[2] = a.z; pk[3] = a.w; pk[4] = b.x; pk[5] = b.y; pk[6] = b.z; pk[7] = b.w; pk[8] = c;
  }
  const uint32_t* pr = planar + (xmin >> 2);                            // planar row (c * 8 + qd * 4 + r): walked in order
  uint32_t* dq = dstq + (size_t)(y0 >> 2) * 3 * kCrop + o;              // tmp[y / 4][c][o], byte y % 4
  const bool second = y0 + 4 < h;
#pragma unroll
  for (int c = 0; c < 3; ++c) {
#pragma unroll
    for (int qd = 0; qd < kHRows / 4; ++qd) {
      uint32_t packed = 0u;
#pragma unroll
      for (int r = 0; r < 4; ++r) {
        uint32_t wv[4];
#pragma unroll
        for (int i = 0; i <= NW; ++i) wv[i] = pr[i];
        pr += PW;
        packed |= (uint32_t)clip8(tap_sum<NW>(wv, sh8, pk)) << (8 * r);
      }
      if (qd == 0 || second) dq[(qd * 3 + c) * kCrop] = packed;
    }
  }
}

__global__ void __launch_bounds__(352)
resample_horizontal_kernel(const uint8_t* __restrict__ frames, size_t frame_stride, const int32_t* __restrict__ hw,
                           const int32_t* __restrict__ tables, uint8_t* __restrict__ tmp, size_t tmp_stride, int pitch) {
  extern __shared__ __align__(16) uint8_t srows[];          // [kHRows][pitch], pitch % 16 == 0 (0: no staging), then the planar rows
  const int f = blockIdx.y, o = threadIdx.x;
  const int h = hw[2 * f], w = hw[2 * f + 1];
  const int y0 = blockIdx.x * kHRows;
  if (y0 >= h) return;
  const int y1 = min(y0 + kHRows, h);
  const uint8_t* src = frames + (size_t)f * frame_stride;
  const bool staged = pitch > 0 && 3 * w + 32 <= pitch;
  if (staged) {
    // 44 threads per row (8 x 44 = 352); every thread has up to three 16-byte loads in flight before it stores the first, so the
    // CTA pays one global-memory latency per 132 vectors of a row instead of one per row
    const int r = threadIdx.x / 44, l44 = threadIdx.x - r * 44;
    if (r < y1 - y0) {
      const size_t gs = (size_t)(y0 + r) * w * 3, a0 = gs & ~(size_t)15;
      const int nvec = (int)((gs + (size_t)w * 3 - a0 + 15) >> 4);
      const uint4* g = reinterpret_cast<const uint4*>(src + a0);
      uint4* d = reinterpret_cast<uint4*>(srows + r * pitch);
      for (int i0 = l44; i0 < nvec; i0 += 132) {
        uint4 v[3];
#pragma unroll
        for (int k = 0; k < 3; ++k) if (i0 + 44 * k < nvec) v[k] = __ldg(g + i0 + 44 * k);
#pragma unroll
        for (int k = 0; k < 3; ++k) if (i0 + 44 * k < nvec) d[i0 + 44 * k] = v[k];
      }
    }
    __syncthreads();
  }
  const int32_t* tab = tables + ((size_t)f * 2 + 0) * kAxisInts;
  const int mode = frame_mode(tables, f, w, pitch);
  if (mode != 0) {
    uint32_t* planar = reinterpret_cast<uint32_t*>(srows + (size_t)pitch * kHRows);
    uint32_t* dstq = reinterpret_cast<uint32_t*>(tmp + (size_t)f * tmp_stride);
    if (mode == 2) horizontal_fast<2>(srows, planar, pitch, w, h, y0, y1, 0, tab, dstq);
    else horizontal_fast<3>(srows, planar, pitch, w, h, y0, y1, 0, tab, dstq);
    return;
  }
  if (o >= kCrop) return;
  const int xmin = tab[o], cnt = tab[kCrop + o];
  const int32_t* k = tab + 2 * kCrop + o;                   // tap j of this column at k[j * kCrop]
  int kr[12];
#pragma unroll
  for (int j = 0; j < 12; ++j) kr[j] = j < cnt ? __ldg(k + j * kCrop) : 0;
  uint8_t* dst = tmp + (size_t)f * tmp_stride;
  for (int y = y0; y < y1; ++y) {
    const size_t gs = (size_t)y * w * 3;
    const uint8_t* row = staged ? srows + (y - y0) * pitch + (gs & 15) + xmin * 3 : src + gs + (size_t)xmin * 3;
    int a0 = 1 << (kPrecisionBits - 1), a1 = a0, a2 = a0;
    if (cnt <= 12) {
#pragma unroll
      for (int j = 0; j < 12; ++j)
        if (j < cnt) { a0 += row[3 * j + 0] * kr[j]; a1 += row[3 * j + 1] * kr[j]; a2 += row[3 * j + 2] * kr[j]; }
    } else {
      for (int j = 0; j < cnt; ++j) {
        const int kj = __ldg(k + j * kCrop);
        a0 += row[3 * j + 0] * kj; a1 += row[3 * j + 1] * kj; a2 += row[3 * j + 2] * kj;
      }
    }
    uint8_t* d = dst + ((size_t)y * kCrop + o) * 3;
    d[0] = (uint8_t)clip8(a0); d[1] = (uint8_t)clip8(a1); d[2] = (uint8_t)clip8(a2);
  }
}

struct NormLut { uint16_t v[256 * 3]; };   // bf16 bits of ((u/255) - mean_c) / std_c, [u][c]

// Vertical pass + crop + normalise + layout.  grid (336/8, F), 352 threads; thread o owns output column o for
// 8 output rows.  The rows of the horizontal intermediate that those 8 outputs need are staged in shared
// memory first (up to kVRows of them; a frame so tall that more are needed reads global memory instead).
constexpr int kVRows = 40;
constexpr int kTmpRowBytes = kCrop * 3;                    // 1008 = 63 x 16
constexpr int kVQuads = 10;                                // fast path: quad rows (4 rows x 336 x 3 bytes = 4032 B) staged per CTA
constexpr int kQuadWords = kCrop * 3;
constexpr int kVSmemBytes = kVQuads * kQuadWords * 4;      // 40 320 = kVRows * kTmpRowBytes

__device__ __forceinline__ void store_pixel(uint16_t* __restrict__ out_pix, uint16_t* __restrict__ out_patches, int f, int yo, int o,
                                            uint16_t v0, uint16_t v1, uint16_t v2) {
  if (out_pix) {
    uint16_t* p = out_pix + (size_t)f * 3 * kCrop * kCrop + (size_t)yo * kCrop + o;
    p[0] = v0; p[(size_t)kCrop * kCrop] = v1; p[(size_t)2 * kCrop * kCrop] = v2;
  }
  if (out_patches) {
    // im2col row = patch (py*24+px), column = c*196 + dy*14 + dx  (conv weight [1024,3,14,14] flattened)
    const int py = yo / 14, dy = yo - py * 14, px = o / 14, dx = o - px * 14;
    uint16_t* p = out_patches + ((size_t)f * 576 + py * 24 + px) * kPatchK + dy * 14 + dx;
    p[0] = v0; p[196] = v1; p[392] = v2;
  }
}

// Fast vertical pass: the intermediate holds four consecutive rows per word, tmp[y / 4][c][o]; the taps of output row yo start
// at row ymin, i.e. at byte ymin % 4 of quad ymin / 4.  Quads past the frame's last one are never copied: their bytes meet
// zero coefficients only.
template <int NW>
__device__ __forceinline__ void vertical_fast(const uint32_t* __restrict__ srcq, uint32_t* __restrict__ sq, const uint16_t* slut,
                                              int h, int yo0, int f, const int32_t* __restrict__ tab,
                                              uint16_t* __restrict__ out_pix, uint16_t* __restrict__ out_patches) {
  __shared__ __align__(16) uint32_t sk[8][kPlaneInts];
  const int o = threadIdx.x;
  const int q_lo = tab[yo0] >> 2, q_need = (tab[yo0 + 7] >> 2) + NW + 1 - q_lo, q_valid = ((h + 3) >> 2) - q_lo;
  const bool staged = q_need <= kVQuads;
  if (staged) {
    const uint4* g = reinterpret_cast<const uint4*>(srcq + (size_t)q_lo * kQuadWords);
    uint4* d = reinterpret_cast<uint4*>(sq);
    const int nvec = min(q_need, q_valid) * (kQuadWords / 4);
    for (int i0 = threadIdx.x; i0 < nvec; i0 += 4 * 352) {               // four loads in flight per thread
      uint4 v[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) if (i0 + 352 * k < nvec) v[k] = __ldg(g + i0 + 352 * k);
#pragma unroll
      for (int k = 0; k < 4; ++k) if (i0 + 352 * k < nvec) d[i0 + 352 * k] = v[k];
    }
  }
  if (o < 8 * kPlaneInts) {
    const int i8 = o / kPlaneInts, i = o - i8 * kPlaneInts;              // words 0..8: planes, word 9: the row's first tap
    sk[i8][i] = i == 9 ? (uint32_t)tab[yo0 + i8] : (uint32_t)tab[kPlaneOff + (yo0 + i8) * kPlaneInts + i];
  }
  __syncthreads();
  if (o >= kCrop) return;
  // per-thread output addressing: column o of pixel_values, patch column px / offset dx of the im2col rows
  const int px = o / 14, dx = o - px * 14;
  uint16_t* ppix = out_pix ? out_pix + (size_t)f * 3 * kCrop * kCrop + o : nullptr;
  uint16_t* ppat = out_patches ? out_patches + ((size_t)f * 576 + px) * kPatchK + dx : nullptr;
  const int q_last = ((h + 3) >> 2) - 1;
#pragma unroll 1
  for (int i8 = 0; i8 < 8; ++i8) {
    const int yo = yo0 + i8;
    uint32_t pk[9];
    int ymin;
    {
      const uint4* s4 = reinterpret_cast<const uint4*>(sk[i8]);
      const uint4 a = s4[0], b = s4[1];
      const uint2 c2 = *reinterpret_cast<const uint2*>(&sk[i8][8]);
      pk[0] = a.x; pk[1] = a.y; pk[2] = a.z; pk[3] = a.w; pk[4] = b.x; pk[5] = b.y; pk[6] = b.z; pk[7] = b.w; pk[8] = c2.x;
      ymin = (int)c2.y;
    }
    const uint32_t sh8 = (uint32_t)(ymin & 3) * 8u;
    uint16_t v[3];
    if (staged) {
      const uint32_t* pc = sq + ((ymin >> 2) - q_lo) * kQuadWords + o;
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        uint32_t wv[4];
#pragma unroll
        for (int i = 0; i <= NW; ++i) wv[i] = pc[i * kQuadWords + c * kCrop];
        v[c] = slut[clip8(tap_sum<NW>(wv, sh8, pk)) * 3 + c];
      }
    } else {
#pragma unroll
      for (int c = 0; c < 3; ++c) {
        uint32_t wv[4];
#pragma unroll
        for (int i = 0; i <= NW; ++i) wv[i] = __ldg(srcq + ((size_t)min((ymin >> 2) + i, q_last) * 3 + c) * kCrop + o);
        v[c] = slut[clip8(tap_sum<NW>(wv, sh8, pk)) * 3 + c];
      }
    }
    if (ppix) {
      uint16_t* p = ppix + (size_t)yo * kCrop;
      p[0] = v[0]; p[(size_t)kCrop * kCrop] = v[1]; p[(size_t)2 * kCrop * kCrop] = v[2];
    }
    if (ppat) {
      // im2col row = patch (py*24+px), column = c*196 + dy*14 + dx  (conv weight [1024,3,14,14] flattened)
      const int py = yo / 14, dy = yo - py * 14;
      uint16_t* p = ppat + (size_t)py * 24 * kPatchK + dy * 14;
      p[0] = v[0]; p[196] = v[1]; p[392] = v[2];
    }
  }
}

__global__ void __launch_bounds__(352)
resample_vertical_kernel(const uint8_t* __restrict__ tmp, size_t tmp_stride, const int32_t* __restrict__ hw,
                         const int32_t* __restrict__ tables, const NormLut lut, int pitch,
                         uint16_t* __restrict__ out_pix, uint16_t* __restrict__ out_patches) {
  __shared__ uint16_t slut[256 * 3];
  extern __shared__ __align__(16) uint8_t srows[];          // kVSmemBytes
  for (int i = threadIdx.x; i < 256 * 3; i += blockDim.x) slut[i] = lut.v[i];
  const int f = blockIdx.y, o = threadIdx.x;
  const int h = hw[2 * f];
  const int32_t* tab = tables + ((size_t)f * 2 + 1) * kAxisInts;
  const uint8_t* src = tmp + (size_t)f * tmp_stride;
  const int yo0 = blockIdx.x * 8;
  const int mode = h > 0 ? frame_mode(tables, f, hw[2 * f + 1], pitch) : 0;
  if (mode != 0) {
    if (mode == 2) vertical_fast<2>(reinterpret_cast<const uint32_t*>(src), reinterpret_cast<uint32_t*>(srows), slut, h, yo0, f, tab, out_pix, out_patches);
    else vertical_fast<3>(reinterpret_cast<const uint32_t*>(src), reinterpret_cast<uint32_t*>(srows), slut, h, yo0, f, tab, out_pix, out_patches);
  }
  int rbase = 0, rend = 0;
  bool staged = false;
  if (h > 0 && mode == 0) {
    rbase = tab[yo0];
    rend = tab[yo0 + 7] + tab[kCrop + yo0 + 7];
    staged = rend - rbase <= kVRows;
    if (staged) {
      const uint4* g = reinterpret_cast<const uint4*>(src + (size_t)rbase * kTmpRowBytes);
      uint4* d = reinterpret_cast<uint4*>(srows);
      const int nvec = (rend - rbase) * (kTmpRowBytes / 16);
      for (int i = threadIdx.x; i < nvec; i += blockDim.x) d[i] = __ldg(g + i);
    }
  }
  __syncthreads();
  if (o < kCrop && h > 0 && mode == 0) {
    for (int yo = yo0; yo < yo0 + 8; ++yo) {
      const int ymin = tab[yo], cnt = tab[kCrop + yo];
      const int32_t* k = tab + 2 * kCrop + yo * kTaps;
      const uint8_t* col = staged ? srows + (size_t)(ymin - rbase) * kTmpRowBytes + o * 3
                                  : src + ((size_t)ymin * kCrop + o) * 3;
      int a0 = 1 << (kPrecisionBits - 1), a1 = a0, a2 = a0;
      for (int j = 0; j < cnt; ++j) {
        const int kj = __ldg(k + j);
        const uint8_t* q = col + (size_t)j * kTmpRowBytes;
        a0 += q[0] * kj; a1 += q[1] * kj; a2 += q[2] * kj;
      }
      store_pixel(out_pix, out_patches, f, yo, o, slut[clip8(a0) * 3 + 0], slut[clip8(a1) * 3 + 1], slut[clip8(a2) * 3 + 2]);
    }
  }
  // zero the K padding (columns 588..639) of this frame's patch rows: 576 rows x 52 columns,
  // spread over the 42 CTAs of the frame
  if (out_patches) {
    const int per_cta = (576 + gridDim.x - 1) / gridDim.x;
    const int r0 = blockIdx.x * per_cta, r1 = min(576, r0 + per_cta);
    for (int i = threadIdx.x; i < (r1 - r0) * 52; i += blockDim.x) {
      const int r = r0 + i / 52, c = 588 + i % 52;
      out_patches[((size_t)f * 576 + r) * kPatchK + c] = 0;
    }
  }
}

static uint16_t f32_to_bf16_rne(float f) {
  uint32_t b;
  memcpy(&b, &f, 4);
  b += 0x7fffu + ((b >> 16) & 1u);
  return (uint16_t)(b >> 16);
}

static NormLut make_norm_lut() {
  // transformers rescale(): image * (1/255) in float64, cast to float32; normalize(): (image - mean) / std in float32
  static const double mean_d[3] = {0.48145466, 0.4578275, 0.40821073};
  static const double std_d[3] = {0.26862954, 0.26130258, 0.27577711};
  NormLut l;
  for (int u = 0; u < 256; ++u) {
    volatile float v = (float)((double)u * (1.0 / 255.0));
    for (int c = 0; c < 3; ++c) {
      volatile float m = (float)mean_d[c], s = (float)std_d[c];
      volatile float d = v - m;
      volatile float r = d / s;
      l.v[u * 3 + c] = f32_to_bf16_rne(r);
    }
  }
  return l;
}

}  // namespace egpt

using namespace egpt;

static size_t tables_bytes(int F) { return align_up((size_t)F * 2 * kAxisInts * sizeof(int32_t), 1024); }
// horizontal-pass intermediate of one frame: max_h rows of 336 x 3 bytes, or (fast path) ceil(max_h / 4) quad rows of the same size
static size_t tmp_stride_bytes(int max_h) { return align_up((size_t)(max_h + 4) * kCrop * 3, 256); }

extern "C" size_t egpt_preprocess_workspace_bytes(int n_frames_total, int max_h) {
  if (n_frames_total <= 0 || max_h <= 0) return 0;
  // coefficient tables + the horizontal-pass intermediate [F][max_h][336][3]
  return tables_bytes(n_frames_total) + align_up(tmp_stride_bytes(max_h) * (size_t)n_frames_total, 1024);
}

extern "C" int egpt_preprocess(const uint8_t* frames, size_t frame_stride, const int32_t* hw, int n_frames_total,
                               int max_h, uint16_t* out_pixel_values, uint16_t* out_patches, int32_t* status,
                               void* workspace, size_t workspace_bytes, void* stream_) {
  static const NormLut lut = make_norm_lut();
  cudaStream_t stream = (cudaStream_t)stream_;
  EGPT_CHECK_ARG(frames && hw && status && workspace, "egpt_preprocess: null pointer argument");
  EGPT_CHECK_ARG(out_pixel_values || out_patches, "egpt_preprocess: no output requested");
  EGPT_CHECK_ARG(n_frames_total > 0 && max_h > 0, "egpt_preprocess: n_frames_total=%d max_h=%d", n_frames_total, max_h);
  EGPT_CHECK_ARG(((uintptr_t)frames & 15) == 0 && frame_stride % 16 == 0 && ((uintptr_t)workspace & 15) == 0 &&
                 ((uintptr_t)out_pixel_values & 15) == 0 && ((uintptr_t)out_patches & 15) == 0,
                 "egpt_preprocess: frames, frame_stride, outputs and workspace must be 16-byte aligned");
  if (workspace_bytes < egpt_preprocess_workspace_bytes(n_frames_total, max_h)) {
    set_error("egpt_preprocess: workspace %zu < %zu bytes", workspace_bytes,
              egpt_preprocess_workspace_bytes(n_frames_total, max_h));
    return EGPT_ERR_WORKSPACE;
  }
  const int F = n_frames_total;
  ProfScope prof(PROF_PREPROCESS, 0.0, stream);
  int32_t* tables = (int32_t*)workspace;
  uint8_t* tmp = (uint8_t*)workspace + tables_bytes(F);
  const size_t tmp_stride = tmp_stride_bytes(max_h);
  resize_coeffs_kernel<<<dim3(F, 2), 352, 0, stream>>>(hw, tables, status);
  EGPT_LAUNCH_CHECK();
  // staging pitch: frames written with capacity (max_h, W) have frame_stride >= 3 * max_h * W
  int pitch = (int)align_up(frame_stride / (size_t)max_h + 3 + 32, 16);
  if ((size_t)pitch * kHRows > 40 * 1024) pitch = 0;       // very wide frames: no staging
  // raw rows + (fast path) planar R / G / B rows of the same eight input rows
  const size_t h_smem = (size_t)pitch * kHRows + (pitch > 0 ? (size_t)3 * kHRows * planar_words(pitch) * 4 : 0);
  EGPT_CUDA(cudaFuncSetAttribute(resample_horizontal_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 96 * 1024));
  EGPT_CUDA(cudaFuncSetAttribute(resample_vertical_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kVSmemBytes));
  resample_horizontal_kernel<<<dim3((max_h + kHRows - 1) / kHRows, F), 352, h_smem, stream>>>(
      frames, frame_stride, hw, tables, tmp, tmp_stride, pitch);
  EGPT_LAUNCH_CHECK();
  resample_vertical_kernel<<<dim3(kCrop / 8, F), 352, kVSmemBytes, stream>>>(tmp, tmp_stride, hw, tables, lut, pitch, out_pixel_values,
                                                                            out_patches);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}
