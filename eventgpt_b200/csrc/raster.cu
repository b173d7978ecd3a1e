// Event-stream splitting and order-exact rasterisation (sm_100a).
//
// Replaces, bit for bit, common/common.py:17-37 (get_event_images_list),
// :64-74 (generate_event_image) and :76-107 (split_event_by_time) of the
// reference, plus the "negative wins" raster of
// feasible/my_egpt_dsec_dataset/preprocess_event_images.py:58-90.
//
// Rasterisation is "last event at a pixel wins", which is not commutative.  The per-event
// operation is a PLAIN one-byte shared-memory store inside a CTA that owns a band of rows of
// one frame (no atomic per event):
//   * a frame is cut into R bands of rows so that one byte per pixel of a band fits in the
//     shared memory of one SM (640x480 -> 2 bands of 240 rows, 1280x720 -> 5 bands of 144);
//     a long slice is also cut into up to K consecutive event ranges.  One CTA per
//     (frame, band, event range) streams its range with 128-bit loads (8 events per load),
//     keeps the events of its band and drops the others after two instructions;
//   * events are consumed in batches of 16 384 with a CTA barrier between batches, so order
//     ACROSS batches is free.  INSIDE a batch every warp owns consecutive 256-event chunks
//     and stores the byte  (chunk_rank+1)<<1 | (p != 0).  After the barrier each thread reads
//     its cells back: a cell holding a lower rank (an earlier chunk overwrote a later one) or
//     the same rank with the other polarity (two events of one chunk on one pixel) is
//     repaired -- the warp finds the latest event of its chunk on that pixel with one
//     shuffle/reduce step -- and re-checked until no thread objects.  With uniformly spread
//     events a few pixels per batch need this; a hot pixel costs one step per chunk;
//   * the finished band leaves as 2 bits per pixel (0 untouched, 2 blue, 3 red) into the
//     caller's scratch; the second kernel overlays the K partial planes of a frame in range
//     order and writes packed RGB at the data-dependent pitch.
// DRAM traffic is the algorithmic 5 B/event + 3 B/pixel plus 2 x 0.25 B/pixel/range of scratch.
#include "common.cuh"
#include "sm100_ptx.cuh"
#include <mutex>

namespace egpt {

constexpr int kTileThreads = 512;
constexpr int kTileWarps = kTileThreads / 32;
constexpr int kGroupsPerBatch = 4;                              // 8-event groups per thread per batch
constexpr int kBatchGroups = kTileThreads * kGroupsPerBatch;    // 2048 groups = 16 384 events between barriers
constexpr int kMaxPlaneBytes = 200 * 1024;                      // band plane held in shared memory
constexpr int kMaxRanges = 12;                                  // K: event ranges per slice
constexpr long long kMinRangeGroups = 4096;                     // a range is at least 32 768 events

struct SliceInfo {
  long long start, end;
};

// slice i of a window [s, e): k = (e-s)/n, [s+i*k, s+(i+1)*k), last slice takes the remainder
// (window w = events [offsets[w*stride], offsets[w*stride + 1]): stride 1 = CSR offsets, stride 2 = (start, end) pairs,
// which may overlap)
__device__ __forceinline__ SliceInfo slice_bounds(const int64_t* __restrict__ offsets, int stride, int window, int i, int n) {
  long long s = offsets[(size_t)window * stride], e = offsets[(size_t)window * stride + 1];
  if (e < s) e = s;
  long long k = (e - s) / n;
  SliceInfo r;
  r.start = s + (long long)i * k;
  r.end = (i < n - 1) ? s + (long long)(i + 1) * k : e;
  return r;
}

// how a frame is cut into row bands, how the frames of a call are cut into event ranges, and how much scratch one
// partial plane takes.  The first `head_frames` frames take one CTA per band (whole waves of the machine); the rest
// (what would be a thin last wave, or a batch smaller than the machine) are cut into K event ranges per slice.
struct TilePlan {
  int band_rows, R, K;
  int head_frames;
  unsigned plane_smem;      // bytes of shared memory per CTA (one byte per pixel of a band)
  unsigned plane_words;     // 32-bit words of one 2-bit-per-pixel partial plane (whole frame), multiple of 4
};

static bool make_tile_plan(int H, int W, long long frames, TilePlan* pl) {
  int rows_fit = kMaxPlaneBytes / W;
  if (rows_fit >= H) {
    pl->band_rows = H;
    pl->R = 1;
  } else {
    rows_fit &= ~15;                                  // bands start on a 16-pixel boundary of the 2-bit plane
    if (rows_fit < 16) return false;
    int R = (H + rows_fit - 1) / rows_fit;
    int rows = ((H + R - 1) / R + 15) & ~15;
    pl->band_rows = rows;
    pl->R = (H + rows - 1) / rows;
  }
  pl->plane_smem = (unsigned)(((size_t)pl->band_rows * W + 15) & ~(size_t)15);
  pl->plane_words = (unsigned)((((size_t)H * W + 15) / 16 + 3) & ~(size_t)3);
  const long long items = frames * pl->R;             // CTAs at one range per slice
  long long head = (items / kNumSMs) * kNumSMs / pl->R;          // frames that fill whole waves
  long long tail = frames - head;
  long long K = 1;
  if (tail > 0) {
    // a small batch gets two waves of CTAs, the left-over of a large one a single wave of short CTAs
    K = ((head ? 1LL : 2LL) * kNumSMs) / (tail * pl->R);
    if (K < 1) K = 1;
    if (K > kMaxRanges) K = kMaxRanges;
  }
  if (K == 1) { head = frames; }
  pl->head_frames = (int)head;
  pl->K = (int)K;
  return true;
}

// partial planes in the scratch: one per head frame, K per tail frame
__host__ __device__ __forceinline__ size_t plane_slot(const TilePlan& pl, long long frame, int range) {
  return frame < pl.head_frames ? (size_t)frame : (size_t)pl.head_frames + (size_t)(frame - pl.head_frames) * pl.K + range;
}
static size_t plane_slots(const TilePlan& pl, long long frames) {
  return (size_t)pl.head_frames + (size_t)(frames - pl.head_frames) * pl.K;
}

// number of event ranges a slice of `ng` 8-event groups is really cut into (same rule in both kernels)
__device__ __forceinline__ int ranges_of(long long ng, int K) {
  if (ng <= 0) return 0;
  long long k = (ng + kMinRangeGroups - 1) / kMinRangeGroups;
  return (int)(k < K ? k : K);
}

struct RasterArgs {
  const void *x, *y, *p;
  int xdt, ydt, pdt;
  const int64_t* offsets;
  int off_stride;
  int n_frames, H, W, fixed_canvas;
  TilePlan plan;
  uint32_t* scratch;        // [frame][K][plane_words]
  int32_t* out_hw;
  int32_t* out_counts;
  int32_t* status;
};

// eight consecutive events as loaded: x, y as 8 x u16, p as 8 x u8.  An event that is not part of the slice (or lies
// outside the sensor) is stored as y = 0xffff, which no band accepts.
struct EventGroup {
  uint4 x, y;
  uint2 p;
};

struct ScanState {
  unsigned px, py;          // packed u16x2 running maxima of the raw coordinates of whole groups
  unsigned mx, my;          // running maxima of the coordinates of element-loaded groups
  unsigned flags;
  unsigned whole;           // bit g: group g of the batch came straight from memory (not yet range-checked)
};

template <bool FAST>
__device__ __forceinline__ void load_group(const RasterArgs& a, long long g, long long g_end, const SliceInfo& sl,
                                           EventGroup& e, ScanState& st, int slot) {
  const long long base = g << 3;
  st.whole &= ~(1u << slot);
  if (FAST && g < g_end && base >= sl.start && base + 8 <= sl.end) {
    e.x = __ldg(reinterpret_cast<const uint4*>(a.x) + g);
    e.y = __ldg(reinterpret_cast<const uint4*>(a.y) + g);
    e.p = __ldg(reinterpret_cast<const uint2*>(a.p) + g);
    st.whole |= 1u << slot;
    return;
  }
  // slice boundary (rare) or generic element types: element loads, never outside [start, end); range-checked here
  unsigned long long xlo = 0, xhi = 0, ylo = ~0ull, yhi = ~0ull, pq = 0;
  if (g < g_end) {
#pragma unroll 1
    for (int j = 0; j < 8; ++j) {
      const long long i = base + j;
      if (i < sl.start || i >= sl.end) continue;
      long long xl, yl, pl;
      if (FAST) {
        xl = ((const uint16_t*)a.x)[i]; yl = ((const uint16_t*)a.y)[i]; pl = ((const uint8_t*)a.p)[i];
      } else {
        xl = load_int(a.x, a.xdt, i); yl = load_int(a.y, a.ydt, i); pl = load_int(a.p, a.pdt, i);
      }
      if (xl < 0 || yl < 0 || xl >= a.W || yl >= a.H) { st.flags |= EGPT_FLAG_OUT_OF_RANGE; continue; }
      const int sh = (j & 3) * 16;
      if (j < 4) { xlo |= (unsigned long long)xl << sh; ylo = (ylo & ~(0xffffull << sh)) | ((unsigned long long)yl << sh); }
      else { xhi |= (unsigned long long)xl << sh; yhi = (yhi & ~(0xffffull << sh)) | ((unsigned long long)yl << sh); }
      pq |= (unsigned long long)(pl == 0 ? 0u : (pl == 1 ? 1u : 2u)) << (j * 8);
      st.mx = max(st.mx, (unsigned)xl);
      st.my = max(st.my, (unsigned)yl);
    }
  }
  e.x = make_uint4((unsigned)xlo, (unsigned)(xlo >> 32), (unsigned)xhi, (unsigned)(xhi >> 32));
  e.y = make_uint4((unsigned)ylo, (unsigned)(ylo >> 32), (unsigned)yhi, (unsigned)(yhi >> 32));
  e.p = make_uint2((unsigned)pq, (unsigned)(pq >> 32));
}

// 4 one-byte cells -> 4 two-bit codes: 0 untouched, 2 | (cell & 1) otherwise
__device__ __forceinline__ unsigned pack_cells4(unsigned w) {
  const unsigned nz = (((w & 0x7f7f7f7fu) + 0x7f7f7f7fu) | w) & 0x80808080u;
  const unsigned c = (nz >> 6) | (w & 0x01010101u);
  return (c * 0x01041040u) >> 24;
}

// one-byte shared-memory accesses at a 32-bit shared address (keeps the address arithmetic out of the unrolled loops)
__device__ __forceinline__ void sts_u8(unsigned addr, unsigned v) {
  asm volatile("st.shared.u8 [%0], %1;" ::"r"(addr), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned lds_u8(unsigned addr) {
  unsigned v;
  asm volatile("ld.shared.u8 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
  return v;
}

// byte j of the eight packed cell values of a group
__device__ __forceinline__ unsigned cell_value(const unsigned (&vw)[2], int j) {
  return __dp4a(vw[j >> 2], 1u << (8 * (j & 3)), 0u);
}

// bytes != 0 -> 1, per byte
__device__ __forceinline__ unsigned nonzero_bytes(unsigned w) {
  return ((((w & 0x7f7f7f7fu) + 0x7f7f7f7fu) | w) >> 7) & 0x01010101u;
}

// One CTA per (frame, band of rows, event range).  MODE 0: last event wins; MODE 1: negative wins.
template <bool FAST, int MODE>
__global__ void __launch_bounds__(kTileThreads, 1)
raster_tile_kernel(const __grid_constant__ RasterArgs a) {
  extern __shared__ __align__(16) uint8_t plane[];
  const int tid = threadIdx.x, lane = tid & 31;
  // chunk order within a batch: the warp scheduler favours high warp ids, so they take the EARLY chunks -- stores then tend
  // to land in chunk order and few cells need the repair step (a matter of speed only)
  const int wrank = kTileWarps - 1 - (tid >> 5);
  const int R = a.plan.R, H = a.H, W = a.W;
  const long long head_items = (long long)a.plan.head_frames * R;
  long long item = blockIdx.x;
  int K = 1;
  long long frame = item / R;
  const int band = (int)(item % R);
  int range = 0;
  if (item >= head_items) {
    item -= head_items;
    K = a.plan.K;
    range = (int)((item / R) % K);
    frame = a.plan.head_frames + item / ((long long)R * K);
  }
  const int window = (int)(frame / a.n_frames), fi = (int)(frame % a.n_frames);
  const SliceInfo sl = slice_bounds(a.offsets, a.off_stride, window, fi, a.n_frames);
  if (sl.end <= sl.start) {
    if (band == 0 && range == 0 && tid == 0) atomicOr(a.status, EGPT_FLAG_EMPTY_SLICE);
    return;
  }
  const long long g0 = sl.start >> 3, ng = ((sl.end + 7) >> 3) - g0;
  const int n_ranges = ranges_of(ng, K);
  if (range >= n_ranges) return;
  const long long g_begin = g0 + ng * range / n_ranges, g_end = g0 + ng * (range + 1) / n_ranges;
  const unsigned y0 = (unsigned)band * a.plan.band_rows;
  const unsigned bh = min((unsigned)a.plan.band_rows, (unsigned)H - y0);

  for (unsigned i = tid; i < a.plan.plane_smem / 16; i += kTileThreads) reinterpret_cast<uint4*>(plane)[i] = make_uint4(0, 0, 0, 0);

  ScanState st = {0, 0, 0, 0, 0, 0};
  EventGroup ev[kGroupsPerBatch];
#pragma unroll
  for (int g = 0; g < kGroupsPerBatch; ++g)
    load_group<FAST>(a, g_begin + (long long)(g * kTileWarps + wrank) * 32 + lane, g_end, sl, ev[g], st, g);
  __syncthreads();

  unsigned plane_s = (unsigned)__cvta_generic_to_shared(plane);
  asm volatile("mov.u32 %0, %0;" : "+r"(plane_s));          // opaque: kept in a register, not recomputed per access
  const unsigned limit = bh * (unsigned)W;                  // cells of this band; a larger index is not this band's event
  const unsigned band_off = y0 * (unsigned)W;
  for (long long gb = g_begin; gb < g_end; gb += kBatchGroups) {
    // ---- store: cell = rank of the 256-event chunk within the batch, and the polarity ------------------
    unsigned pix[kGroupsPerBatch][8];           // cell index within the band (>= limit: none)
    unsigned valw[kGroupsPerBatch][2];          // the eight cell values of a group, one byte each
#pragma unroll
    for (int g = 0; g < kGroupsPerBatch; ++g) {
      const unsigned rank2 = (unsigned)(g * kTileWarps + wrank + 1) << 1;
      const unsigned xw[4] = {ev[g].x.x, ev[g].x.y, ev[g].x.z, ev[g].x.w};
      const unsigned yw[4] = {ev[g].y.x, ev[g].y.y, ev[g].y.z, ev[g].y.w};
      if ((st.whole >> g) & 1u) {               // whole group straight from memory: its coordinates count for the
        st.px = __vmaxu2(__vmaxu2(st.px, xw[0]), __vmaxu2(xw[1], __vmaxu2(xw[2], xw[3])));   // canvas and are range-checked
        st.py = __vmaxu2(__vmaxu2(st.py, yw[0]), __vmaxu2(yw[1], __vmaxu2(yw[2], yw[3])));   // once, at the end
      }
      if (MODE == EGPT_RASTER_LAST_WINS) {
        valw[g][0] = nonzero_bytes(ev[g].p.x) | (rank2 * 0x01010101u);
        valw[g][1] = nonzero_bytes(ev[g].p.y) | (rank2 * 0x01010101u);
      } else {                                  // p == 0: negative (0xfe, wins), p == 1: positive (0x01), other p: ignored (0)
        valw[g][0] = (nonzero_bytes(ev[g].p.x) ^ 0x01010101u) * 0xfeu | (nonzero_bytes(ev[g].p.x ^ 0x01010101u) ^ 0x01010101u);
        valw[g][1] = (nonzero_bytes(ev[g].p.y) ^ 0x01010101u) * 0xfeu | (nonzero_bytes(ev[g].p.y ^ 0x01010101u) ^ 0x01010101u);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        // field extraction on the multiply-add pipe (dot products with a one-hot selector): the integer-logic pipe is
        // the busier one in this kernel
        const unsigned yv = __dp2a_lo(yw[j >> 1], (j & 1) ? 0x0100u : 0x0001u, 0u);
        const unsigned val = cell_value(valw[g], j);
        // rows below the band wrap to a huge index, rows above it (and y = 0xffff: not an event) exceed the limit; an x
        // beyond the sensor can land on a wrong pixel of the band, but never outside it (the call then reports
        // EGPT_FLAG_OUT_OF_RANGE and the frames are not used)
        const unsigned q = __dp2a_lo(xw[j >> 1], (j & 1) ? 0x0100u : 0x0001u, yv * (unsigned)W - band_off);
        pix[g][j] = q;
        if (MODE == EGPT_RASTER_LAST_WINS) {
          if (q < limit) sts_u8(plane_s + q, val);
        } else {
          if (q < limit && lds_u8(plane_s + q) < val) sts_u8(plane_s + q, val);
        }
      }
    }
    if (a.out_counts) {                                            // auxiliary per-pixel counts (no reference counterpart)
      int32_t* counts = a.out_counts + (size_t)frame * 2 * H * W + band_off;
#pragma unroll 1
      for (int g = 0; g < kGroupsPerBatch; ++g)
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          unsigned q = 0, v = 0;
#pragma unroll
          for (int gg = 0; gg < kGroupsPerBatch; ++gg)
            if (gg == g) { q = pix[gg][j]; v = cell_value(valw[gg], j); }
          if (q < limit) atomicAdd(counts + (size_t)(v & 1u) * H * W + q, 1);
        }
    }
    // the next batch's events travel while this one is checked
#pragma unroll
    for (int g = 0; g < kGroupsPerBatch; ++g)
      load_group<FAST>(a, gb + kBatchGroups + (long long)(g * kTileWarps + wrank) * 32 + lane, g_end, sl, ev[g], st, g);
    __syncthreads();

    // ---- check: does every cell hold the latest event of the batch on its pixel? --------------------------
    unsigned bad = 0;
    unsigned rq = 0, rv = 0, ridx = 0;          // the (last) event found wrong: cell address, value, index in the batch
#pragma unroll
    for (int g = 0; g < kGroupsPerBatch; ++g) {
      const unsigned hi = ((unsigned)(g * kTileWarps + wrank + 1) << 1) | 1u;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const unsigned val = cell_value(valw[g], j);
        unsigned c = 0xffu;
        if (pix[g][j] < limit) c = lds_u8(plane_s + pix[g][j]);
        const bool wrong = MODE == EGPT_RASTER_LAST_WINS ? (c != val && c <= hi) : c < val;
        if (wrong) { bad |= 1u << (g * 8 + j); rq = plane_s + pix[g][j]; rv = val; ridx = g * 8 + j; }
      }
    }
    while (__syncthreads_or(bad != 0)) {
      unsigned recheck = 0;
      bool single = false;
      if (MODE == EGPT_RASTER_LAST_WINS) {
        // the common dispute: ONE event of this thread found an earlier chunk's value in its cell (that store landed
        // after ours).  The thread stores again, alone; everything else goes to the warp-wide step below.
        const unsigned rhi = ((unsigned)((ridx >> 3) * kTileWarps + wrank + 1) << 1) | 1u;
        unsigned coop = bad;
        if (bad != 0 && bad == (1u << ridx)) {
          const unsigned c = lds_u8(rq);
          if (c > rhi || c == rv) { bad = 0; coop = 0; }                    // a later chunk (or our value) holds the pixel by now
          else if (c < (rhi ^ 1u)) { sts_u8(rq, rv); single = true; coop = 0; }
        }
        if (__ballot_sync(0xffffffffu, coop != 0)) {
#pragma unroll
          for (int g = 0; g < kGroupsPerBatch; ++g) {
            const unsigned hi = ((unsigned)(g * kTileWarps + wrank + 1) << 1) | 1u;
            unsigned gm = (coop >> (g * 8)) & 0xffu;
            // one step per disputed pixel of this warp's chunk: the latest event of the chunk on it writes
            for (;;) {
              const unsigned ballot = __ballot_sync(0xffffffffu, gm != 0);
              if (!ballot) break;
              unsigned mine = 0;
#pragma unroll
              for (int j = 7; j >= 0; --j) if ((gm >> j) & 1u) mine = pix[g][j];
              const unsigned P = __shfl_sync(0xffffffffu, mine, __ffs(ballot) - 1);
              int best = -1;
              unsigned val = 0;
#pragma unroll
              for (int j = 0; j < 8; ++j)
                if (pix[g][j] == P) { best = j; val = cell_value(valw[g], j); gm &= ~(1u << j); }
              const int key = best >= 0 ? lane * 8 + best : -1;
              const int top = __reduce_max_sync(0xffffffffu, key);
              if (best >= 0 && key == top) {
                if (lds_u8(plane_s + P) <= hi) sts_u8(plane_s + P, val);
                recheck |= 1u << (g * 8 + best);
              }
            }
          }
        }
      } else {
#pragma unroll
        for (int g = 0; g < kGroupsPerBatch; ++g)
#pragma unroll
          for (int j = 0; j < 8; ++j)
            if ((bad >> (g * 8 + j)) & 1u) {
              const unsigned val = cell_value(valw[g], j);
              if (lds_u8(plane_s + pix[g][j]) < val) sts_u8(plane_s + pix[g][j], val);
              recheck |= 1u << (g * 8 + j);
            }
      }
      __syncthreads();
      unsigned still = 0;
      if (single) {
        const unsigned rhi = ((unsigned)((ridx >> 3) * kTileWarps + wrank + 1) << 1) | 1u;
        const unsigned c = lds_u8(rq);
        if (c != rv && c <= rhi) still = bad;
      }
      if (recheck) {
#pragma unroll
        for (int g = 0; g < kGroupsPerBatch; ++g) {
          const unsigned hi = ((unsigned)(g * kTileWarps + wrank + 1) << 1) | 1u;
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            if (!((recheck >> (g * 8 + j)) & 1u)) continue;
            const unsigned val = cell_value(valw[g], j);
            const unsigned c = lds_u8(plane_s + pix[g][j]);
            const bool wrong = MODE == EGPT_RASTER_LAST_WINS ? (c != val && c <= hi) : c < val;
            if (wrong) still |= 1u << (g * 8 + j);
          }
        }
      }
      bad = still;
    }
  }

  // ---- the band leaves as 2 bits per pixel ------------------------------------------------------------
  {
    uint32_t* dst = a.scratch + plane_slot(a.plan, frame, range) * a.plan.plane_words + ((size_t)y0 * W >> 4);
    const unsigned words = (unsigned)(((size_t)bh * W + 15) >> 4);
    for (unsigned i = tid; i < words; i += kTileThreads) {
      const uint4 c = reinterpret_cast<const uint4*>(plane)[i];
      dst[i] = pack_cells4(c.x) | (pack_cells4(c.y) << 8) | (pack_cells4(c.z) << 16) | (pack_cells4(c.w) << 24);
    }
  }

  // whole groups are range-checked here, through their running maxima
  unsigned mx = max(st.px & 0xffffu, st.px >> 16), my = max(st.py & 0xffffu, st.py >> 16);
  if (mx >= (unsigned)W || my >= (unsigned)H) st.flags |= EGPT_FLAG_OUT_OF_RANGE;
  mx = min(max(mx, st.mx), (unsigned)W - 1);
  my = min(max(my, st.my), (unsigned)H - 1);
  // canvas extent = (max y + 1, max x + 1) of the slice (common/common.py:65); a non-empty range holds an event
  // (or the out-of-range flag is up, and the frame is not used)
  if (!a.fixed_canvas) {
    const unsigned vx = __reduce_max_sync(0xffffffffu, mx), vy = __reduce_max_sync(0xffffffffu, my);
    if (lane == 0) {
      atomicMax(a.out_hw + 2 * frame + 0, (int)vy + 1);
      atomicMax(a.out_hw + 2 * frame + 1, (int)vx + 1);
    }
  }
  const unsigned flags = __reduce_or_sync(0xffffffffu, st.flags);
  if (lane == 0 && flags) atomicOr(a.status, (int)flags);
}

__global__ void set_fixed_hw_kernel(int32_t* __restrict__ hw, int n, int H, int W) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) { hw[2 * i] = H; hw[2 * i + 1] = W; }
}

__device__ __forceinline__ unsigned rgb_of_code(unsigned code) {
  // little-endian byte triple: white 255,255,255 (0); blue 0,0,255 (2); red 255,0,0 (3)
  return code == 0 ? 0x00ffffffu : ((code & 1u) ? 0x000000ffu : 0x00ff0000u);
}

// a later range's pixels lie over an earlier range's
__device__ __forceinline__ unsigned overlay_codes(unsigned under, unsigned over) {
  const unsigned m = ((over >> 1) & 0x55555555u) * 3u;
  return (under & ~m) | (over & m);
}

// partial planes -> packed RGB.  A thread turns 16 pixels (one word of codes per range) into 48 B of RGB; a CTA stages
// 256 x 48 B = 12 KB in shared memory and hands it to the bulk-copy (TMA) engine as ONE contiguous store, so the frame
// leaves in large pieces without tying up the load/store path; two staging buffers alternate.
constexpr int kColThreads = 256;
constexpr int kColChunk = kColThreads * 48;

__device__ __forceinline__ void bulk_store_1d(void* gdst, const void* ssrc, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
               ::"l"(gdst), "r"((unsigned)__cvta_generic_to_shared(ssrc)), "r"(bytes) : "memory");
}

__global__ void __launch_bounds__(kColThreads)
raster_colourise_kernel(const uint32_t* __restrict__ scratch, const int32_t* __restrict__ out_hw,
                        const int64_t* __restrict__ offsets, int off_stride, int n_frames,
                        const __grid_constant__ TilePlan plan, int W,
                        uint8_t* __restrict__ out_frames, size_t frame_stride) {
  __shared__ __align__(128) uint8_t stage[2][kColChunk];
  const long long frame = blockIdx.y;
  const int tid = threadIdx.x;
  const int h = out_hw[2 * frame], w = out_hw[2 * frame + 1];
  const long long npix = (long long)h * w;
  if (npix == 0) return;
  const SliceInfo sl = slice_bounds(offsets, off_stride, (int)(frame / n_frames), (int)(frame % n_frames), n_frames);
  const int K = frame < plan.head_frames ? 1 : plan.K;
  const int n_ranges = sl.end > sl.start ? ranges_of(((sl.end + 7) >> 3) - (sl.start >> 3), K) : 0;
  const unsigned plane_words = plan.plane_words;
  const uint32_t* planes = scratch + plane_slot(plan, frame, 0) * plane_words;
  uint8_t* dst = out_frames + (size_t)frame * frame_stride;
  const long long ngroups = (npix + 15) >> 4, total_bytes = npix * 3;
  const long long nchunks = (ngroups + kColThreads - 1) / kColThreads;
  int buf = 0;
  for (long long chunk = blockIdx.x; chunk < nchunks; chunk += gridDim.x, buf ^= 1) {
    const long long g = chunk * kColThreads + tid;
    if (tid == 0) ptx::bulk_wait_group_read<1>();       // the copy that last read this buffer has drained it
    __syncthreads();
    if (g < ngroups) {
      unsigned m = 0;
      if (w == W) {
        for (int k = 0; k < n_ranges; ++k) m = overlay_codes(m, __ldg(planes + (size_t)k * plane_words + g));
      } else {
        const long long p0 = g << 4;
#pragma unroll
        for (int j = 0; j < 16; ++j) {
          const long long p = p0 + j;
          unsigned code = 0;
          if (p < npix) {
            const int yy = (int)(p / w), xx = (int)(p - (long long)yy * w);
            const size_t q = (size_t)yy * W + xx;
            for (int k = 0; k < n_ranges; ++k) {
              const unsigned v = (planes[(size_t)k * plane_words + (q >> 4)] >> ((q & 15) * 2)) & 3u;
              if (v) code = v;
            }
          }
          m |= code << (2 * j);
        }
      }
      // 4 pixels (12 B) -> 3 words; 16 pixels -> 12 words -> 3 x uint4
      unsigned wds[12];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const unsigned a = rgb_of_code((m >> (8 * q)) & 3u), b = rgb_of_code((m >> (8 * q + 2)) & 3u),
                       d = rgb_of_code((m >> (8 * q + 4)) & 3u), e = rgb_of_code((m >> (8 * q + 6)) & 3u);
        wds[3 * q + 0] = a | (b << 24);
        wds[3 * q + 1] = (b >> 8) | (d << 16);
        wds[3 * q + 2] = (d >> 16) | (e << 8);
      }
      uint4* o = reinterpret_cast<uint4*>(stage[buf] + tid * 48);
      o[0] = make_uint4(wds[0], wds[1], wds[2], wds[3]);
      o[1] = make_uint4(wds[4], wds[5], wds[6], wds[7]);
      o[2] = make_uint4(wds[8], wds[9], wds[10], wds[11]);
    }
    ptx::fence_proxy_async();
    __syncthreads();
    const long long off = chunk * kColChunk;
    const long long left = total_bytes - off;
    const unsigned bytes = (unsigned)(left < kColChunk ? left : kColChunk), bulk = bytes & ~15u;
    if (tid == 0) {
      if (bulk) bulk_store_1d(dst + off, stage[buf], bulk);
      ptx::bulk_commit_group();
    }
    if (tid < (int)(bytes - bulk)) dst[off + bulk + tid] = stage[buf][bulk + tid];      // the last few bytes of a frame
  }
  if (tid == 0) ptx::bulk_wait_group<0>();
}

// ------------------------------------------------------------------------------------------
// split_event_by_time: stable counting sort of events by time bin.
// ------------------------------------------------------------------------------------------
constexpr int kSplitMaxChunks = 4096;

__device__ __forceinline__ long long floor_div(long long a, long long b) {
  long long q = a / b;
  if ((a % b != 0) && ((a < 0) != (b < 0))) --q;
  return q;
}

struct SplitPlan {
  long long n;
  long long chunk;     // events per chunk (multiple of 32)
  int n_chunks;
  int span_cap;        // capacity of the dense bin table (= max_windows)
};

// minimum and maximum TIMESTAMP (the bin is monotone in t, so the division happens once per kernel, not once per event)
template <bool T64>
__global__ void split_minmax_kernel(const void* __restrict__ t, int tdt, long long n, long long* __restrict__ minmax) {
  long long lo = 0x7fffffffffffffffLL, hi = -0x7fffffffffffffffLL - 1;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += 4 * stride) {
    long long v[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long j = (i + u * stride) < n ? i + u * stride : i;       // past the end: element i again
      v[u] = T64 ? ((const long long*)t)[j] : load_int(t, tdt, j);
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) { lo = min(lo, v[u]); hi = max(hi, v[u]); }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    lo = min(lo, __shfl_xor_sync(0xffffffffu, lo, o));
    hi = max(hi, __shfl_xor_sync(0xffffffffu, hi, o));
  }
  if ((threadIdx.x & 31) == 0) { atomicMin(minmax, lo); atomicMax(minmax + 1, hi); }
}

// bin of a timestamp relative to the first bin of the stream.  When the stream spans less than 2^32 us past the start of
// its first bin (any real recording), the 64-bit floor division becomes one 32-bit division.
struct BinOf {
  long long lo, span, t0, interval;
  unsigned iv32;
  bool small;
  __device__ BinOf(const long long* __restrict__ minmax, long long interval_) : interval(interval_) {
    const long long tmin = minmax[0], tmax = minmax[1];
    lo = floor_div(tmin, interval);
    span = floor_div(tmax, interval) - lo + 1;
    t0 = lo * interval;
    small = interval < (1LL << 31) && (tmax - t0) < (1LL << 32);
    iv32 = (unsigned)interval;
  }
  __device__ __forceinline__ int operator()(long long tv) const {
    return small ? (int)((unsigned)(tv - t0) / iv32) : (int)(floor_div(tv, interval) - lo);
  }
};

__global__ void split_init_kernel(long long* minmax) {
  minmax[0] = 0x7fffffffffffffffLL;
  minmax[1] = -0x7fffffffffffffffLL - 1;
}

// one warp per chunk: zeroes its row of the table (the bin span is known by now), then counts
template <bool T64>
__global__ void split_count_kernel(const void* __restrict__ t, int tdt, SplitPlan plan, long long interval,
                                   const long long* __restrict__ minmax, int32_t* __restrict__ table,
                                   int32_t* __restrict__ status) {
  const int chunk = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (chunk >= plan.n_chunks) return;
  const BinOf bin(minmax, interval);
  const long long span = bin.span;
  if (span > plan.span_cap) { if (lane == 0 && chunk == 0) atomicOr(status, EGPT_FLAG_TOO_MANY_BINS); return; }
  const long long s = (long long)chunk * plan.chunk, e = min(plan.n, s + plan.chunk);
  int32_t* row = table + (size_t)chunk * plan.span_cap;
  for (long long i = lane; i < span; i += 32) row[i] = 0;
  __syncwarp();
  for (long long b0 = s; b0 < e; b0 += 128) {
    int rb[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {                            // four rows of 32 timestamps in flight
      const long long i = b0 + u * 32 + lane;
      rb[u] = i < e ? bin(T64 ? ((const long long*)t)[i] : load_int(t, tdt, i)) : -1;
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const unsigned m = __match_any_sync(0xffffffffu, rb[u]);
      if (rb[u] >= 0 && lane == __ffs(m) - 1) row[rb[u]] += __popc(m);
      __syncwarp();
    }
  }
}

// one CTA per time bin: exclusive prefix of the bin's counts over the chunks, in place; the bin's total on the side
__global__ void __launch_bounds__(256)
split_column_scan_kernel(SplitPlan plan, long long interval, const long long* __restrict__ minmax, int32_t* __restrict__ table,
                         long long* __restrict__ bin_total) {
  const long long span = BinOf(minmax, interval).span;
  if (span > plan.span_cap) return;
  __shared__ int warp_sum[8];
  __shared__ int carry_s;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  for (long long rb = blockIdx.x; rb < span; rb += gridDim.x) {
    if (tid == 0) carry_s = 0;
    __syncthreads();
    for (int c0 = 0; c0 < plan.n_chunks; c0 += 256) {
      const int c = c0 + tid;
      int32_t* cell = table + (size_t)c * plan.span_cap + rb;
      const int v = c < plan.n_chunks ? *cell : 0;
      int incl = v;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const int u = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += u; }
      if (lane == 31) warp_sum[wid] = incl;
      __syncthreads();
      int before = carry_s;
      for (int w = 0; w < wid; ++w) before += warp_sum[w];
      if (c < plan.n_chunks) *cell = before + incl - v;
      __syncthreads();
      if (tid == 255) carry_s = before + incl;
      __syncthreads();
    }
    if (tid == 0) bin_total[rb] = carry_s;
    __syncthreads();
  }
}

// single CTA: exclusive prefix of the bin totals (first destination of every bin) and compaction of the non-empty bins
// into window offsets
__global__ void __launch_bounds__(1024)
split_bin_scan_kernel(SplitPlan plan, long long interval, const long long* __restrict__ minmax, long long* __restrict__ bin_start,
                      int64_t* __restrict__ out_offsets, int32_t* __restrict__ out_nwindows, int max_windows) {
  const long long span = BinOf(minmax, interval).span;
  if (span > plan.span_cap) { if (threadIdx.x == 0) { *out_nwindows = 0; out_offsets[0] = 0; } return; }
  __shared__ long long wsum[32];
  __shared__ int wcnt[32];
  __shared__ long long carry_s;
  __shared__ int carry_n;
  const int tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
  if (tid == 0) { carry_s = 0; carry_n = 0; out_offsets[0] = 0; }
  __syncthreads();
  for (long long r0 = 0; r0 < span; r0 += 1024) {
    const long long rb = r0 + tid;
    const long long v = rb < span ? bin_start[rb] : 0;        // holds the bin's total on entry
    const int ne = v > 0;
    long long incl = v;
    int ninc = ne;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const long long u = __shfl_up_sync(0xffffffffu, incl, o);
      const int un = __shfl_up_sync(0xffffffffu, ninc, o);
      if (lane >= o) { incl += u; ninc += un; }
    }
    if (lane == 31) { wsum[wid] = incl; wcnt[wid] = ninc; }
    __syncthreads();
    long long before = carry_s;
    int nbefore = carry_n;
    for (int w = 0; w < wid; ++w) { before += wsum[w]; nbefore += wcnt[w]; }
    if (rb < span) {
      bin_start[rb] = before + incl - v;
      if (ne && nbefore + ninc <= max_windows) out_offsets[nbefore + ninc] = before + incl;
    }
    __syncthreads();
    if (tid == 1023) { carry_s = before + incl; carry_n = nbefore + ninc; }
    __syncthreads();
  }
  if (tid == 0) *out_nwindows = carry_n;
}

__device__ __forceinline__ void copy_elem(void* __restrict__ dst, const void* __restrict__ src, int size,
                                          long long d, long long s) {
  switch (size) {
    case 1: ((uint8_t*)dst)[d] = ((const uint8_t*)src)[s]; break;
    case 2: ((uint16_t*)dst)[d] = ((const uint16_t*)src)[s]; break;
    case 4: ((uint32_t*)dst)[d] = ((const uint32_t*)src)[s]; break;
    default: ((uint64_t*)dst)[d] = ((const uint64_t*)src)[s]; break;
  }
}

// FAST: t 64-bit, x and y 16-bit, p 8-bit (the layout of the reference's DSEC windows): typed copies
template <bool FAST>
__global__ void split_scatter_kernel(const void* __restrict__ x, const void* __restrict__ y,
                                     const void* __restrict__ t, const void* __restrict__ p,
                                     int xs, int ys, int tdt, int ps, SplitPlan plan, long long interval,
                                     const long long* __restrict__ minmax, int32_t* __restrict__ table,
                                     const long long* __restrict__ bin_start,
                                     void* __restrict__ ox, void* __restrict__ oy, void* __restrict__ ot,
                                     void* __restrict__ op) {
  const int chunk = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (chunk >= plan.n_chunks) return;
  const BinOf bin(minmax, interval);
  if (bin.span > plan.span_cap) return;
  const long long s = (long long)chunk * plan.chunk, e = min(plan.n, s + plan.chunk);
  int32_t* row = table + (size_t)chunk * plan.span_cap;
  const int ts = dtype_size(tdt);
  for (long long b0 = s; b0 < e; b0 += 128) {
    int rb[4];
    long long tv[4];
    uint16_t xv[4], yv[4];
    uint8_t pv[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long i = b0 + u * 32 + lane;
      if (FAST) {
        if (i < e) { tv[u] = ((const long long*)t)[i]; xv[u] = ((const uint16_t*)x)[i]; yv[u] = ((const uint16_t*)y)[i]; pv[u] = ((const uint8_t*)p)[i]; }
        else tv[u] = 0;
        rb[u] = i < e ? bin(tv[u]) : -1;
      } else {
        rb[u] = i < e ? bin(load_int(t, tdt, i)) : -1;
      }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const long long i = b0 + u * 32 + lane;
      const unsigned m = __match_any_sync(0xffffffffu, rb[u]);
      const int leader = __ffs(m) - 1;
      int base = 0;
      if (rb[u] >= 0 && lane == leader) { base = row[rb[u]]; row[rb[u]] = base + __popc(m); }
      base = __shfl_sync(0xffffffffu, base, leader);
      if (rb[u] >= 0) {
        const long long d = bin_start[rb[u]] + base + __popc(m & ((1u << lane) - 1u));
        if (FAST) {
          ((long long*)ot)[d] = tv[u]; ((uint16_t*)ox)[d] = xv[u]; ((uint16_t*)oy)[d] = yv[u]; ((uint8_t*)op)[d] = pv[u];
        } else {
          copy_elem(ox, x, xs, d, i);
          copy_elem(oy, y, ys, d, i);
          copy_elem(ot, t, ts, d, i);
          copy_elem(op, p, ps, d, i);
        }
      }
      __syncwarp();
    }
  }
}

static SplitPlan make_split_plan(long long n, int max_windows) {
  SplitPlan pl;
  pl.n = n;
  long long chunks = (n + 1023) / 1024;
  if (chunks > kSplitMaxChunks) chunks = kSplitMaxChunks;
  if (chunks < 1) chunks = 1;
  long long chunk = (n + chunks - 1) / chunks;
  chunk = (chunk + 31) / 32 * 32;
  if (chunk < 32) chunk = 32;
  pl.chunk = chunk;
  pl.n_chunks = (int)((n + chunk - 1) / chunk);
  if (pl.n_chunks < 1) pl.n_chunks = 1;
  pl.span_cap = max_windows;
  return pl;
}

}  // namespace egpt

using namespace egpt;

extern "C" size_t egpt_raster_frame_stride(int H, int W) { return align_up((size_t)3 * H * W, 256); }

extern "C" size_t egpt_raster_workspace_bytes(int n_windows, int n_frames, int H, int W) {
  if (n_windows <= 0 || n_frames <= 0 || H <= 0 || W <= 0) return 0;
  TilePlan pl;
  if (!make_tile_plan(H, W, (long long)n_windows * n_frames, &pl)) return 0;
  return plane_slots(pl, (long long)n_windows * n_frames) * pl.plane_words * sizeof(uint32_t) + 1024;
}

namespace egpt {
template <bool FAST, int MODE>
static int launch_tiles(const RasterArgs& a, long long items, cudaStream_t stream) {
  EGPT_CUDA(cudaFuncSetAttribute(raster_tile_kernel<FAST, MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)a.plan.plane_smem));
  raster_tile_kernel<FAST, MODE><<<(unsigned)items, kTileThreads, a.plan.plane_smem, stream>>>(a);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}

// shrink_to_fit: use fewer event ranges per slice when the scratch is smaller than the query asks for
int raster_impl(const void* x, const void* y, const void* p, int x_dtype, int y_dtype, int p_dtype,
                       const int64_t* window_offsets, int off_stride, int n_windows, int n_frames, int mode, int H, int W,
                       int fixed_canvas, uint8_t* out_frames, int32_t* out_hw, int32_t* out_counts, int32_t* status,
                       void* workspace, size_t workspace_bytes, bool shrink_to_fit, cudaStream_t stream) {
  EGPT_CHECK_ARG(x && y && p && window_offsets && out_frames && out_hw && status && workspace,
                 "egpt_raster: null pointer argument");
  EGPT_CHECK_ARG(n_windows > 0 && n_frames > 0, "egpt_raster: n_windows=%d n_frames=%d must be positive", n_windows, n_frames);
  EGPT_CHECK_ARG(H > 0 && W > 0 && H < 65535 && W < 65535 && (long long)H * W < (1LL << 31) / 4,
                 "egpt_raster: bad canvas capacity %dx%d", H, W);
  EGPT_CHECK_ARG(mode == EGPT_RASTER_LAST_WINS || mode == EGPT_RASTER_NEG_WINS, "egpt_raster: unknown mode %d", mode);
  EGPT_CHECK_ARG(dtype_size(x_dtype) && dtype_size(y_dtype) && dtype_size(p_dtype), "egpt_raster: unknown dtype");
  EGPT_CHECK_ARG(out_counts == nullptr || mode == EGPT_RASTER_LAST_WINS, "egpt_raster: counts only in LAST_WINS mode");
  EGPT_CHECK_ARG(((uintptr_t)out_frames & 15) == 0 && ((uintptr_t)workspace & 15) == 0,
                 "egpt_raster: out_frames and workspace must be 16-byte aligned");
  const long long F = (long long)n_windows * n_frames;
  EGPT_CHECK_ARG(F <= 65535, "egpt_raster: more than 65535 frames in one call");
  RasterArgs a;
  if (!make_tile_plan(H, W, F, &a.plan)) {
    set_error("egpt_raster: a %d-pixel-wide sensor does not fit the shared-memory band (16 rows x W <= %d bytes)", W, kMaxPlaneBytes);
    return EGPT_ERR_UNSUPPORTED;
  }
  const size_t plane_bytes = (size_t)a.plan.plane_words * sizeof(uint32_t);
  if (shrink_to_fit)
    while (a.plan.K > 1 && plane_slots(a.plan, F) * plane_bytes > workspace_bytes) --a.plan.K;
  if (plane_slots(a.plan, F) * plane_bytes > workspace_bytes) {
    set_error("egpt_raster: workspace %zu < %zu bytes", workspace_bytes, egpt_raster_workspace_bytes(n_windows, n_frames, H, W));
    return EGPT_ERR_WORKSPACE;
  }
  const long long items = ((long long)a.plan.head_frames + (F - a.plan.head_frames) * a.plan.K) * a.plan.R;
  const size_t frame_stride = egpt_raster_frame_stride(H, W);
  // algorithmic bytes are data dependent (5 B/event + 3 B/pixel written); the caller of
  // egpt_profile_collect supplies them, so only time is recorded here
  ProfScope prof(PROF_RASTER, 0.0, stream);
  a.x = x; a.y = y; a.p = p; a.xdt = x_dtype; a.ydt = y_dtype; a.pdt = p_dtype;
  a.offsets = window_offsets; a.off_stride = off_stride; a.n_frames = n_frames; a.H = H; a.W = W; a.fixed_canvas = fixed_canvas;
  a.scratch = (uint32_t*)workspace; a.out_hw = out_hw; a.out_counts = out_counts; a.status = status;
  const bool fast = x_dtype == EGPT_U16 && y_dtype == EGPT_U16 && p_dtype == EGPT_U8 &&
                    ((uintptr_t)x % 16 == 0) && ((uintptr_t)y % 16 == 0) && ((uintptr_t)p % 8 == 0);

  if (fixed_canvas) {
    set_fixed_hw_kernel<<<(unsigned)((F + 255) / 256), 256, 0, stream>>>(out_hw, (int)F, H, W);
    EGPT_LAUNCH_CHECK();
  } else {
    EGPT_CUDA(cudaMemsetAsync(out_hw, 0, sizeof(int32_t) * 2 * F, stream));
  }
  if (out_counts) EGPT_CUDA(cudaMemsetAsync(out_counts, 0, sizeof(int32_t) * 2 * (size_t)F * H * W, stream));

  if (mode == EGPT_RASTER_LAST_WINS) {
    if (fast) EGPT_TRY((launch_tiles<true, EGPT_RASTER_LAST_WINS>(a, items, stream)));
    else EGPT_TRY((launch_tiles<false, EGPT_RASTER_LAST_WINS>(a, items, stream)));
  } else {
    if (fast) EGPT_TRY((launch_tiles<true, EGPT_RASTER_NEG_WINS>(a, items, stream)));
    else EGPT_TRY((launch_tiles<false, EGPT_RASTER_NEG_WINS>(a, items, stream)));
  }
  int cgrid = (int)(((size_t)H * W / 16 + kColThreads - 1) / kColThreads);
  const int want = (int)((kNumSMs * 6 + F - 1) / F);
  if (cgrid > want) cgrid = want;
  if (cgrid < 1) cgrid = 1;
  raster_colourise_kernel<<<dim3(cgrid, (unsigned)F), kColThreads, 0, stream>>>(a.scratch, out_hw, window_offsets, off_stride,
                                                                                n_frames, a.plan, W, out_frames, frame_stride);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}
}  // namespace egpt

extern "C" int egpt_raster(const void* x, const void* y, const void* p, int x_dtype, int y_dtype, int p_dtype,
                           const int64_t* window_offsets, int n_windows, int n_frames, int mode, int H, int W,
                           int fixed_canvas, uint8_t* out_frames, int32_t* out_hw, int32_t* out_counts,
                           int32_t* status, void* workspace, size_t workspace_bytes, void* stream) {
  return raster_impl(x, y, p, x_dtype, y_dtype, p_dtype, window_offsets, 1, n_windows, n_frames, mode, H, W, fixed_canvas,
                     out_frames, out_hw, out_counts, status, workspace, workspace_bytes, false, (cudaStream_t)stream);
}

extern "C" int egpt_raster_ranges(const void* x, const void* y, const void* p, int x_dtype, int y_dtype, int p_dtype,
                                  const int64_t* window_bounds, int n_windows, int n_frames, int mode, int H, int W,
                                  int fixed_canvas, uint8_t* out_frames, int32_t* out_hw, int32_t* out_counts,
                                  int32_t* status, void* workspace, size_t workspace_bytes, void* stream) {
  return raster_impl(x, y, p, x_dtype, y_dtype, p_dtype, window_bounds, 2, n_windows, n_frames, mode, H, W, fixed_canvas,
                     out_frames, out_hw, out_counts, status, workspace, workspace_bytes, false, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------
// Window / clip bounds of a long time-sorted recording, on the device.
// ------------------------------------------------------------------------------------------
namespace egpt {
// first index i in [0, n) with t[i] >= v (n if none)
__device__ __forceinline__ long long lower_bound_t(const void* __restrict__ t, int tdt, long long n, long long v) {
  long long lo = 0, hi = n;
  while (lo < hi) {
    const long long mid = (lo + hi) >> 1;
    if (load_int(t, tdt, mid) < v) lo = mid + 1; else hi = mid;
  }
  return lo;
}

__global__ void clip_bounds_kernel(const void* __restrict__ t, int tdt, long long n, long long first_us, long long interval_us,
                                   long long step_us, int n_clips, int64_t* __restrict__ out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_clips) return;
  const long long s = first_us + (long long)c * step_us;
  out[2 * c] = lower_bound_t(t, tdt, n, s);
  out[2 * c + 1] = lower_bound_t(t, tdt, n, s + interval_us);
}

__global__ void ms_bounds_kernel(const void* __restrict__ ms_to_idx, int dt, long long n_ms, long long first_ms, long long window_ms,
                                 long long step_ms, int n_windows, int64_t* __restrict__ out, int32_t* __restrict__ status) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= n_windows) return;
  const long long a = first_ms + (long long)c * step_ms, b = a + window_ms;
  if (a < 0 || b >= n_ms) { atomicOr(status, EGPT_FLAG_OUT_OF_RANGE); out[2 * c] = out[2 * c + 1] = 0; return; }
  out[2 * c] = load_int(ms_to_idx, dt, a);
  out[2 * c + 1] = load_int(ms_to_idx, dt, b);
}
}  // namespace egpt

extern "C" int egpt_clip_bounds(const void* t, int t_dtype, int64_t n_events, int64_t first_us, int64_t interval_us,
                                int64_t step_us, int n_clips, int64_t* out_bounds, void* stream) {
  EGPT_CHECK_ARG(t && out_bounds && dtype_size(t_dtype), "egpt_clip_bounds: bad pointer or dtype");
  EGPT_CHECK_ARG(n_events >= 0 && interval_us > 0 && step_us > 0 && n_clips > 0, "egpt_clip_bounds: bad sizes");
  clip_bounds_kernel<<<(n_clips + 127) / 128, 128, 0, (cudaStream_t)stream>>>(t, t_dtype, n_events, first_us, interval_us,
                                                                             step_us, n_clips, out_bounds);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}

extern "C" int egpt_ms_window_bounds(const void* ms_to_idx, int dtype, int64_t n_ms, int64_t first_ms, int64_t window_ms,
                                     int64_t step_ms, int n_windows, int64_t* out_bounds, int32_t* status, void* stream) {
  EGPT_CHECK_ARG(ms_to_idx && out_bounds && status && dtype_size(dtype), "egpt_ms_window_bounds: bad pointer or dtype");
  EGPT_CHECK_ARG(n_ms > 0 && window_ms >= 0 && step_ms > 0 && n_windows > 0, "egpt_ms_window_bounds: bad sizes");
  ms_bounds_kernel<<<(n_windows + 127) / 128, 128, 0, (cudaStream_t)stream>>>(ms_to_idx, dtype, n_ms, first_ms, window_ms, step_ms,
                                                                             n_windows, out_bounds, status);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}

extern "C" size_t egpt_split_time_workspace_bytes(int64_t n_events, int max_windows) {
  if (n_events < 0 || max_windows <= 0) return 0;
  SplitPlan pl = make_split_plan(n_events, max_windows);
  return 1024 + align_up((size_t)pl.n_chunks * pl.span_cap * sizeof(int32_t), 1024) +
         align_up((size_t)pl.span_cap * sizeof(long long), 1024);
}

extern "C" int egpt_split_time(const void* x, const void* y, const void* t, const void* p, int x_dtype, int y_dtype,
                               int t_dtype, int p_dtype, int64_t n_events, int64_t interval, void* out_x, void* out_y,
                               void* out_t, void* out_p, int64_t* out_offsets, int32_t* out_nwindows, int max_windows,
                               int32_t* status, void* workspace, size_t workspace_bytes, void* stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  EGPT_CHECK_ARG(x && y && t && p && out_x && out_y && out_t && out_p && out_offsets && out_nwindows && status && workspace,
                 "egpt_split_time: null pointer argument");
  EGPT_CHECK_ARG(n_events > 0 && n_events < (1LL << 31), "egpt_split_time: n_events=%lld out of range", (long long)n_events);
  EGPT_CHECK_ARG(interval > 0, "egpt_split_time: interval must be positive");
  EGPT_CHECK_ARG(max_windows > 0 && max_windows <= (1 << 20), "egpt_split_time: max_windows=%d out of range", max_windows);
  EGPT_CHECK_ARG(dtype_size(x_dtype) && dtype_size(y_dtype) && dtype_size(t_dtype) && dtype_size(p_dtype),
                 "egpt_split_time: unknown dtype");
  if (workspace_bytes < egpt_split_time_workspace_bytes(n_events, max_windows)) {
    set_error("egpt_split_time: workspace too small");
    return EGPT_ERR_WORKSPACE;
  }
  SplitPlan pl = make_split_plan(n_events, max_windows);
  ProfScope prof(PROF_SPLIT, 2.0 * (double)n_events * (dtype_size(x_dtype) + dtype_size(y_dtype) + dtype_size(t_dtype) +
                                                        dtype_size(p_dtype)) + 2.0 * (double)n_events * dtype_size(t_dtype), stream);
  char* ws = (char*)workspace;
  long long* minmax = (long long*)ws;
  int32_t* table = (int32_t*)(ws + 1024);
  const size_t table_bytes = align_up((size_t)pl.n_chunks * pl.span_cap * sizeof(int32_t), 1024);
  long long* bin_start = (long long*)(ws + 1024 + table_bytes);

  split_init_kernel<<<1, 1, 0, stream>>>(minmax);
  EGPT_LAUNCH_CHECK();
  int blocks = (int)((n_events + 1023) / 1024);
  if (blocks > kNumSMs * 8) blocks = kNumSMs * 8;
  const bool t64 = dtype_size(t_dtype) == 8;
  const bool fast = t64 && dtype_size(x_dtype) == 2 && dtype_size(y_dtype) == 2 && dtype_size(p_dtype) == 1;
  if (t64) split_minmax_kernel<true><<<blocks, 256, 0, stream>>>(t, t_dtype, n_events, minmax);
  else split_minmax_kernel<false><<<blocks, 256, 0, stream>>>(t, t_dtype, n_events, minmax);
  EGPT_LAUNCH_CHECK();
  const int warps_per_cta = 4;
  const int cblocks = (pl.n_chunks + warps_per_cta - 1) / warps_per_cta;
  if (t64) split_count_kernel<true><<<cblocks, warps_per_cta * 32, 0, stream>>>(t, t_dtype, pl, interval, minmax, table, status);
  else split_count_kernel<false><<<cblocks, warps_per_cta * 32, 0, stream>>>(t, t_dtype, pl, interval, minmax, table, status);
  EGPT_LAUNCH_CHECK();
  split_column_scan_kernel<<<max_windows < 4 * kNumSMs ? max_windows : 4 * kNumSMs, 256, 0, stream>>>(pl, interval, minmax, table, bin_start);
  EGPT_LAUNCH_CHECK();
  split_bin_scan_kernel<<<1, 1024, 0, stream>>>(pl, interval, minmax, bin_start, out_offsets, out_nwindows, max_windows);
  EGPT_LAUNCH_CHECK();
  if (fast)
    split_scatter_kernel<true><<<cblocks, warps_per_cta * 32, 0, stream>>>(x, y, t, p, 2, 2, t_dtype, 1, pl, interval, minmax, table,
                                                                          bin_start, out_x, out_y, out_t, out_p);
  else
    split_scatter_kernel<false><<<cblocks, warps_per_cta * 32, 0, stream>>>(x, y, t, p, dtype_size(x_dtype), dtype_size(y_dtype),
        t_dtype, dtype_size(p_dtype), pl, interval, minmax, table, bin_start, out_x, out_y, out_t, out_p);
  EGPT_LAUNCH_CHECK();
  return EGPT_OK;
}


// ------------------------------------------------------------------------------------------
// egpt_raster_static: rasterisation for callers that cannot allocate device memory.
// ------------------------------------------------------------------------------------------
// Same result as egpt_raster, but the partial planes (2 bits per pixel per event range) live in a
// fixed array that is part of the library image instead of a caller-supplied workspace.
// Capacity: n_windows * n_frames * H * W <= kStaticMaxKeys pixel cells, i.e. one 5-frame window of a
// 1280x720 sensor, or three 5-frame windows of a 640x480 one.  Every GPU that loads the library
// carries these 17.6 MiB until the process exits.
namespace egpt {
constexpr long long kStaticMaxKeys = 5LL * 720 * 1280;
constexpr long long kStaticSlackWords = 1024;
static __device__ __align__(256) uint32_t g_static_key_planes[kStaticMaxKeys + kStaticSlackWords];

// The array's name is only a handle on the host; the device address has to be asked for, once per
// device, while that device is current.
static std::mutex g_static_mu;
static void* g_static_ptr[64] = {nullptr};
static int static_scratch_address(void** out) {
  int dev = -1;
  EGPT_CUDA(cudaGetDevice(&dev));
  EGPT_CHECK_ARG(dev >= 0 && dev < 64, "egpt_raster_static: device ordinal %d not supported", dev);
  std::lock_guard<std::mutex> lock(g_static_mu);
  if (g_static_ptr[dev] == nullptr) {
    void* ptr = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&ptr, g_static_key_planes);
    if (e != cudaSuccess || ptr == nullptr) {
      set_error("egpt_raster_static: cudaGetSymbolAddress failed on device %d: %s", dev, cudaGetErrorString(e));
      return EGPT_ERR_CUDA;
    }
    g_static_ptr[dev] = ptr;
  }
  *out = g_static_ptr[dev];
  return EGPT_OK;
}
}  // namespace egpt

extern "C" int64_t egpt_raster_static_max_keys(void) { return egpt::kStaticMaxKeys; }

extern "C" int egpt_raster_static(const void* x, const void* y, const void* p, int x_dtype, int y_dtype, int p_dtype,
                                  const int64_t* window_offsets, int n_windows, int n_frames, int mode, int H, int W,
                                  int fixed_canvas, uint8_t* out_frames, int32_t* out_hw, int32_t* status,
                                  void* stream) {
  EGPT_CHECK_ARG(n_windows > 0 && n_frames > 0 && H > 0 && W > 0, "egpt_raster_static: bad shape %d x %d x %d x %d",
                 n_windows, n_frames, H, W);
  // compare by division: the product of four ints may not fit 64 bits
  const long long frames = (long long)n_windows * n_frames, pixels = (long long)H * W;
  if (frames > kStaticMaxKeys || pixels > kStaticMaxKeys / frames) {
    set_error("egpt_raster_static: %d windows x %d frames x %dx%d exceeds the %lld pixel cells of the built-in scratch "
              "(use egpt_raster with a caller workspace for larger requests)", n_windows, n_frames, H, W, kStaticMaxKeys);
    return EGPT_ERR_TOO_LARGE;
  }
  void* scratch = nullptr;
  EGPT_TRY(static_scratch_address(&scratch));
  // every partial plane that is read is written in full by its band CTAs first, so whatever the
  // previous call left in the array is harmless; the number of event ranges shrinks to what fits
  return raster_impl(x, y, p, x_dtype, y_dtype, p_dtype, window_offsets, 1, n_windows, n_frames, mode, H, W, fixed_canvas,
                     out_frames, out_hw, nullptr, status, scratch,
                     (size_t)(kStaticMaxKeys + kStaticSlackWords) * sizeof(uint32_t), true, (cudaStream_t)stream);
}
