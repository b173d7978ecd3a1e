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
constexpr int kAxisInts = kCrop * (2 + kTaps);        // xmin[336], cnt[336], then the taps: horizontal axis k[taps][336] (lanes of a warp
                                                      // = neighbouring output columns read neighbouring words), vertical axis k[336][taps]
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
  const int f = blockIdx.x, axis = blockIdx.y, o = threadIdx.x;
  if (o >= kCrop) return;
  const int h = hw[2 * f], w = hw[2 * f + 1];
  int32_t* tab = tables + ((size_t)f * 2 + axis) * kAxisInts;
  if (h <= 0 || w <= 0) { tab[o] = 0; tab[kCrop + o] = 0; return; }
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
  for (int x = 0; x < cnt; ++x) {
    double v = wts[x];
    if (ww != 0.0) v = __ddiv_rn(v, ww);
    const double s = __dmul_rn(v, (double)(1 << kPrecisionBits));
    tab[2 * kCrop + (axis == 0 ? x * kCrop + o : o * kTaps + x)] = v < 0 ? (int)__dadd_rn(-0.5, s) : (int)__dadd_rn(0.5, s);
  }
  tab[o] = xmin;
  tab[kCrop + o] = cnt;
}

__device__ __forceinline__ int clip8(int v) {
  v >>= kPrecisionBits;
  return v < 0 ? 0 : (v > 255 ? 255 : v);
}

// Horizontal pass: tmp[f][y][o][c] for every input row y < h and the 336 cropped columns.
// grid (ceil(H/8), F), 352 threads.  The CTA first stages its 8 input rows in shared memory with 16-byte
// loads (rows are packed at an arbitrary byte offset, so each is copied from the enclosing aligned span);
// thread o then owns output column o and reads its taps as bytes from shared memory.
constexpr int kHRows = 8;
__global__ void __launch_bounds__(352)
resample_horizontal_kernel(const uint8_t* __restrict__ frames, size_t frame_stride, const int32_t* __restrict__ hw,
                           const int32_t* __restrict__ tables, uint8_t* __restrict__ tmp, size_t tmp_stride, int pitch) {
  extern __shared__ __align__(16) uint8_t srows[];          // [kHRows][pitch], pitch % 16 == 0 (0: no staging)
  const int f = blockIdx.y, o = threadIdx.x;
  const int h = hw[2 * f], w = hw[2 * f + 1];
  const int y0 = blockIdx.x * kHRows;
  if (y0 >= h) return;
  const int y1 = min(y0 + kHRows, h);
  const uint8_t* src = frames + (size_t)f * frame_stride;
  const bool staged = pitch > 0 && 3 * w + 32 <= pitch;
  if (staged) {
    for (int r = 0; r < y1 - y0; ++r) {
      const size_t gs = (size_t)(y0 + r) * w * 3, a0 = gs & ~(size_t)15;
      const int nvec = (int)((gs + (size_t)w * 3 - a0 + 15) >> 4);
      const uint4* g = reinterpret_cast<const uint4*>(src + a0);
      uint4* d = reinterpret_cast<uint4*>(srows + r * pitch);
      for (int i = threadIdx.x; i < nvec; i += blockDim.x) d[i] = __ldg(g + i);
    }
    __syncthreads();
  }
  if (o >= kCrop) return;
  const int32_t* tab = tables + ((size_t)f * 2 + 0) * kAxisInts;
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
__global__ void __launch_bounds__(352)
resample_vertical_kernel(const uint8_t* __restrict__ tmp, size_t tmp_stride, const int32_t* __restrict__ hw,
                         const int32_t* __restrict__ tables, const NormLut lut,
                         uint16_t* __restrict__ out_pix, uint16_t* __restrict__ out_patches) {
  __shared__ uint16_t slut[256 * 3];
  __shared__ __align__(16) uint8_t srows[kVRows * kTmpRowBytes];
  for (int i = threadIdx.x; i < 256 * 3; i += blockDim.x) slut[i] = lut.v[i];
  const int f = blockIdx.y, o = threadIdx.x;
  const int h = hw[2 * f];
  const int32_t* tab = tables + ((size_t)f * 2 + 1) * kAxisInts;
  const uint8_t* src = tmp + (size_t)f * tmp_stride;
  const int yo0 = blockIdx.x * 8;
  int rbase = 0, rend = 0;
  bool staged = false;
  if (h > 0) {
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
  if (o < kCrop && h > 0) {
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
      const uint16_t v0 = slut[clip8(a0) * 3 + 0], v1 = slut[clip8(a1) * 3 + 1], v2 = slut[clip8(a2) * 3 + 2];
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
static size_t tmp_stride_bytes(int max_h) { return align_up((size_t)max_h * kCrop * 3, 256); }

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
  resample_horizontal_kernel<<<dim3((max_h + kHRows - 1) / kHRows, F), 352, (size_t)pitch * kHRows, stream>>>(
      frames, frame_stride, hw, tables, tmp, tmp_stride, pitch);
  EGPT_LAUNCH_CHECK();
  resample_vertical_kernel<<<dim3(kCrop / 8, F), 352, 0, stream>>>(tmp, tmp_stride, hw, tables, lut, out_pixel_values, out_patches);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}
