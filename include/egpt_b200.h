/*
 * egpt_b200.h -- C ABI of libegpt_b200.so: the B200 (sm_100a) implementation of
 * EventGPT's event -> visual-token front end.
 *
 * The reference (xifan-zhang/EventGPT) is pure Python and has no FFI of its
 * own; each entry point below replaces one reference function (cited as
 * file:line relative to the reference tree) and is what a binding for that
 * function would call.  INTEGRATION.md shows the ctypes stubs.
 *
 * Conventions
 *   - every function returns 0 (EGPT_OK) or a negative egpt_status; the text of
 *     the last error on the calling thread is egpt_last_error().
 *   - all data pointers are DEVICE pointers owned by the caller unless the
 *     parameter name ends in _host.  The library allocates nothing on the data
 *     path: scratch comes from the caller (`workspace`), sized by the matching
 *     *_workspace_bytes() query.  egpt_create() is the only allocating call (it
 *     packs the weights once).
 *   - calls are asynchronous on `stream` (a cudaStream_t passed as void*), never
 *     synchronise, install no host callbacks, and keep no global mutable state:
 *     concurrent calls on different streams with different workspaces are safe.
 *     The one exception to "scratch comes from the caller" and to "no shared state"
 *     is egpt_raster_static (its comment below says what it shares and when calls race).
 *   - conditions the reference turns into Python exceptions (an empty slice ->
 *     ValueError from y.max(), common/common.py:65) are data dependent; they are
 *     written to a caller-provided device `status` word that the host checks
 *     after it synchronises (see egpt_status bits).
 */
#ifndef EGPT_B200_H
#define EGPT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  EGPT_OK = 0,
  EGPT_ERR_INVALID = -1,      /* bad argument (null pointer, size, dtype, alignment)            */
  EGPT_ERR_CUDA = -2,         /* a CUDA runtime / driver call failed                             */
  EGPT_ERR_WORKSPACE = -3,    /* workspace too small for the request                             */
  EGPT_ERR_UNSUPPORTED = -4,  /* valid upstream, outside the limits of this build                */
  EGPT_ERR_TOO_LARGE = -5     /* request exceeds a fixed capacity                                */
} egpt_status;

/* bits of the device status word written by the raster / split / preprocess kernels */
enum {
  EGPT_FLAG_EMPTY_SLICE = 1,   /* a frame slice has no events: the reference raises ValueError    */
  EGPT_FLAG_OUT_OF_RANGE = 2,  /* x >= W, y >= H or a negative coordinate                         */
  EGPT_FLAG_TOO_MANY_BINS = 4, /* split: the bins between the first and last event exceed max_windows */
  EGPT_FLAG_FILTER_TOO_WIDE = 8 /* preprocess: down-scale needs more filter taps than EGPT_MAX_TAPS */
};

/* element types of the event arrays (numpy dtypes found in the reference's .npy files) */
typedef enum {
  EGPT_U8 = 0, EGPT_I8 = 1, EGPT_U16 = 2, EGPT_I16 = 3,
  EGPT_U32 = 4, EGPT_I32 = 5, EGPT_U64 = 6, EGPT_I64 = 7
} egpt_dtype;

typedef enum {
  EGPT_RASTER_LAST_WINS = 0,  /* common/common.py:64-74 (inference path)                         */
  EGPT_RASTER_NEG_WINS = 1    /* feasible/my_egpt_dsec_dataset/preprocess_event_images.py:58-90  */
} egpt_raster_mode;

#define EGPT_CROP 336          /* CLIP ViT-L/14-336 input edge                                    */
#define EGPT_TOKENS 577        /* 24*24 patches + class token                                     */
#define EGPT_HIDDEN 1024
#define EGPT_LLM_HIDDEN 4096
#define EGPT_MAX_TAPS 32       /* resize filter taps supported (down-scale factor <= 7.5)         */

const char* egpt_last_error(void);
int egpt_version(void);

/* ------------------------------------------------------------------------------------------ */
/* a-2  split_event_by_time(event_npy, time_interval)            common/common.py:76-107        */
/* Stable partition of a stream into time bins  bin = floor(t / interval);  bins ascending,     */
/* empty bins skipped, event order inside a bin preserved, t need not be sorted.  Outputs are   */
/* the permuted x,y,t,p (same dtypes, n elements), out_offsets[0..n_windows] and *out_nwindows. */
/* max_windows bounds the SPAN of bins, floor(t_max/interval) - floor(t_min/interval) + 1, empty  */
/* bins included (the scratch table is chunks x span); beyond it EGPT_FLAG_TOO_MANY_BINS is set   */
/* and no window is returned -- call again with max_windows >= span (the Python mirror does).     */
size_t egpt_split_time_workspace_bytes(int64_t n_events, int max_windows);
int egpt_split_time(const void* x, const void* y, const void* t, const void* p,
                    int x_dtype, int y_dtype, int t_dtype, int p_dtype,
                    int64_t n_events, int64_t interval,
                    void* out_x, void* out_y, void* out_t, void* out_p,
                    int64_t* out_offsets, int32_t* out_nwindows, int max_windows,
                    int32_t* status, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------ */
/* a-3 + a-4  get_event_images_list(event_npy, n) / generate_event_image(x, y, p)               */
/*                                                   common/common.py:17-37 and 64-74           */
/* For each of n_windows windows (events [window_offsets[b], window_offsets[b+1]) ) cut         */
/* n_frames equal-count slices (the last takes the remainder) and paint one uint8 RGB frame per */
/* slice: white canvas, blue (0,0,255) where the last event has p == 0, red (255,0,0)           */
/* otherwise.  The canvas is (max y + 1, max x + 1) of the slice (fixed_canvas == 0) or H x W   */
/* (fixed_canvas != 0); H, W are the capacity.  Frame f = b*n_frames + i is stored tightly      */
/* packed, row pitch 3*w_f bytes, at out_frames + f*frame_stride; out_hw[f] = {h_f, w_f}.       */
/* out_counts (optional, may be NULL): int32 [F, 2, H, W] per-pixel event counts (p==0, p!=0),  */
/* an auxiliary product with no counterpart in the reference.                                   */
/* Preconditions (checked): out_frames and workspace 16-byte aligned; H, W < 65535; a band of   */
/* 16 rows of the sensor (16*W bytes) must fit in shared memory (W <= 12800); at most 65535     */
/* frames per call.  A slice may hold any number of events (no per-slice index is kept).        */
size_t egpt_raster_frame_stride(int H, int W);   /* bytes between frames: 3*H*W rounded up to 256 */
size_t egpt_raster_workspace_bytes(int n_windows, int n_frames, int H, int W);
int egpt_raster(const void* x, const void* y, const void* p, int x_dtype, int y_dtype, int p_dtype,
                const int64_t* window_offsets, int n_windows, int n_frames,
                int mode, int H, int W, int fixed_canvas,
                uint8_t* out_frames, int32_t* out_hw, int32_t* out_counts,
                int32_t* status, void* workspace, size_t workspace_bytes, void* stream);

/* f-2  the same rasterisation over (start, end) pairs -- windows that may overlap -- instead of CSR */
/* offsets: window b = events [window_bounds[2b], window_bounds[2b+1]) of ONE time-ordered recording.   */
/* This is how the 50 %-overlap clips of build_my_egpt_dsec_seq.py:137-202 and the DSEC windows of      */
/* dataset/io.py:68-87 are rasterised without copying the shared events (egpt_clip_bounds /             */
/* egpt_ms_window_bounds below produce the pairs on the device).                                        */
int egpt_raster_ranges(const void* x, const void* y, const void* p, int x_dtype, int y_dtype, int p_dtype,
                       const int64_t* window_bounds, int n_windows, int n_frames,
                       int mode, int H, int W, int fixed_canvas,
                       uint8_t* out_frames, int32_t* out_hw, int32_t* out_counts,
                       int32_t* status, void* workspace, size_t workspace_bytes, void* stream);

/* f-2  clip bounds of a time-SORTED recording   feasible/my_egpt_dsec_dataset/build_my_egpt_dsec_seq.py:137-202 */
/* clip c covers times [first_us + c*step_us, first_us + c*step_us + interval_us): out_bounds[c] = {first event  */
/* index with t >= start, first index with t >= end} (two binary searches per clip on the device; t any integer  */
/* dtype).  step_us = interval_us/2 gives the reference's 50 %-overlap clips, step_us = interval_us consecutive   */
/* windows.  An empty clip has start == end (the reference drops it; egpt_raster flags it as an empty slice).      */
int egpt_clip_bounds(const void* t, int t_dtype, int64_t n_events, int64_t first_us, int64_t interval_us,
                     int64_t step_us, int n_clips, int64_t* out_bounds, void* stream);

/* f-2  window bounds from DSEC's per-millisecond index                          dataset/io.py:68-87             */
/* window c = events [ms_to_idx[first_ms + c*step_ms], ms_to_idx[first_ms + c*step_ms + window_ms]); a window     */
/* beyond the index sets EGPT_FLAG_OUT_OF_RANGE (the reference's assert on t_max).                                 */
int egpt_ms_window_bounds(const void* ms_to_idx, int dtype, int64_t n_ms, int64_t first_ms, int64_t window_ms,
                          int64_t step_ms, int n_windows, int64_t* out_bounds, int32_t* status, void* stream);

/* egpt_raster_static -- no counterpart in the reference.  The same rasterisation for callers   */
/* that are not allowed to create device memory: a sensor-acquisition or plug-in thread that    */
/* was handed its input and output buffers by the host application and must not call cudaMalloc */
/* (it would stall the device the application is sharing, and it can fail at a moment the       */
/* thread has no way to recover from).  The scratch planes that egpt_raster keeps in a caller    */
/* workspace live here in a fixed array inside the library, so the call allocates nothing and   */
/* takes no workspace argument.  Limits and rules:                                              */
/*   - n_windows * n_frames * H * W <= egpt_raster_static_max_keys() (= 5*720*1280: one 5-frame */
/*     window at 1280x720, or three at 640x480); larger requests return EGPT_ERR_TOO_LARGE.     */
/*     Every GPU that loads the library holds this 17.6 MiB array until the process exits.      */
/*   - the first call on a device looks up the array's device address (with that device         */
/*     current); make that first call before starting any stream capture.                       */
/*   - all calls on a device share the one array: two calls that may overlap on the GPU         */
/*     (different streams, different threads) race -- serialise them.  The array keeps the last */
/*     call's planes; a call writes every plane it later reads in full, so nothing is inherited.*/
/*   - per-pixel counts are not offered here; everything else is as in egpt_raster.             */
int64_t egpt_raster_static_max_keys(void);
int egpt_raster_static(const void* x, const void* y, const void* p, int x_dtype, int y_dtype, int p_dtype,
                       const int64_t* window_offsets, int n_windows, int n_frames,
                       int mode, int H, int W, int fixed_canvas,
                       uint8_t* out_frames, int32_t* out_hw, int32_t* status, void* stream);

/* ------------------------------------------------------------------------------------------ */
/* a-5  event_processor(img)['pixel_values'][0].to(bf16)         common/common.py:124-127       */
/* CLIPImageProcessor: resize shortest edge -> 336 with Pillow's antialiased bicubic (8-bit     */
/* fixed point, horizontal pass then vertical pass), centre crop 336x336, 1/255, (v-mean)/std,  */
/* CHW, bf16.  Input frames as written by egpt_raster (packed, per-frame {h,w} in `hw`).        */
/* out_pixel_values: bf16 [F,3,336,336] (may be NULL); out_patches: bf16 [F*576, 640] im2col    */
/* rows (c*196+dy*14+dx, zero padded from 588) for the patch-embedding GEMM (may be NULL).      */
/* max_h bounds every frame's height (the raster capacity H).                                   */
size_t egpt_preprocess_workspace_bytes(int n_frames_total, int max_h);
int egpt_preprocess(const uint8_t* frames, size_t frame_stride, const int32_t* hw, int n_frames_total,
                    int max_h, uint16_t* out_pixel_values, uint16_t* out_patches,
                    int32_t* status, void* workspace, size_t workspace_bytes, void* stream);

/* ------------------------------------------------------------------------------------------ */
/* Weights: device pointers in the checkpoint's own layouts (nn.Linear weight = [out, in]).     */
/* Matrices are bf16; vectors (biases, LayerNorm affine, class embedding) are fp32.             */
typedef struct {
  const void *ln1_w, *ln1_b;            /* layer_norm1                                  fp32 [1024]      */
  const void *q_w, *k_w, *v_w, *o_w;    /* self_attn.{q,k,v,out}_proj.weight            bf16 [1024,1024] */
  const void *q_b, *k_b, *v_b, *o_b;    /*                          .bias               fp32 [1024]      */
  const void *ln2_w, *ln2_b;            /* layer_norm2                                  fp32 [1024]      */
  const void *fc1_w, *fc1_b;            /* mlp.fc1   bf16 [4096,1024], fp32 [4096]                       */
  const void *fc2_w, *fc2_b;            /* mlp.fc2   bf16 [1024,4096], fp32 [1024]                       */
} egpt_layer_weights;

typedef struct {
  int num_layers;                       /* 24 for CLIP ViT-L/14-336                                      */
  const void* class_embedding;          /* fp32 [1024]                                                   */
  const void* patch_weight;             /* bf16 [1024, 3*14*14]                                          */
  const void* position_embedding;       /* fp32 [577,1024]                                               */
  const void *pre_ln_w, *pre_ln_b;      /* pre_layrnorm  fp32 [1024]                                     */
  const egpt_layer_weights* layers;     /* host array [num_layers]                                       */
  const void *proj0_w, *proj0_b;        /* visual_projector.0   bf16 [4096,1024], fp32 [4096]            */
  const void *proj2_w, *proj2_b;        /* visual_projector.2   bf16 [4096,4096], fp32 [4096]            */
  const void *adaptor_w, *adaptor_b;    /* feature_adaptor      bf16 [4096,4096], fp32 [4096]            */
} egpt_weights;

typedef struct egpt_handle egpt_handle;

/* Model construction: EventChatLlamaModel.__init__ (model/EventChatModel.py:73-90).  Packs     */
/* q/k/v into one [3072,1024] matrix (q pre-scaled by 0.125, exact in bf16) and pads the patch  */
/* kernel to K = 640.  Reads the weights on `device`; they may be freed afterwards.             */
int egpt_create(const egpt_weights* weights_host, int device, egpt_handle** out);
int egpt_destroy(egpt_handle* h);

/* workspace for n_frames_total frames through tower + projector + pooling */
size_t egpt_workspace_bytes(int n_frames_total);

/* a-6 (tower half)  CLIPVisionModel(x).last_hidden_state        model/EventChatModel.py:196    */
/* pixel_values bf16 [F,3,336,336] -> out bf16 [F,577,1024] (post_layernorm is NOT applied).    */
int egpt_tower(egpt_handle* h, const uint16_t* pixel_values, int n_frames_total,
               uint16_t* out_hidden, void* workspace, size_t workspace_bytes, void* stream);

/* a-6  EventChatModel.visval_encode(x)                          model/EventChatModel.py:194-200 */
/* pixel_values bf16 [F,3,336,336] -> out bf16 [F,577,4096] = visual_projector(tower(x)).       */
int egpt_visval_encode(egpt_handle* h, const uint16_t* pixel_values, int n_frames_total,
                       uint16_t* out_features, void* workspace, size_t workspace_bytes, void* stream);

/* a-7  feature_adaptor(x) = Linear(4096,4096)                   model/EventChatModel.py:84-85,338 */
/* in bf16 [rows,4096] -> out bf16 [rows,4096].                                                  */
int egpt_feature_adaptor(egpt_handle* h, const uint16_t* in, int64_t rows, uint16_t* out, void* stream);

/* a-8  get_spatio_temporal_features(features, num_temporal_tokens) model/EventChatModel.py:15-38 */
/* features [B,t,s,c] (bf16 if is_f32 == 0, else fp32) -> out [B, n_temporal + s, c] same dtype: */
/* rows 0..n_temporal-1 = mean over s of frame i (zero rows beyond t), then s rows = mean over t. */
size_t egpt_pool_workspace_bytes(int batch, int t, int c);
int egpt_spatio_temporal_pool(const void* features, int is_f32, int batch, int t, int s, int c,
                              int n_temporal, void* out, void* workspace, size_t workspace_bytes,
                              void* stream);

/* a-9  list branch of prepare_inputs_labels_for_multimodal      model/EventChatModel.py:326-343 */
/* pixel_values bf16 [B*n,3,336,336] -> tokens bf16 [B, n+577, 4096]: tower, projector, adaptor, */
/* pooling.  The pooling is a mean and the projector's second linear and the adaptor are affine, */
/* so both run AFTER the pooling, on n+577 instead of n*577 rows per window (the result is the   */
/* same up to rounding).                                                           */
int egpt_encode_pool(egpt_handle* h, const uint16_t* pixel_values, int n_windows, int n_frames,
                     uint16_t* out_tokens, void* workspace, size_t workspace_bytes, void* stream);

/* a-6 (second half) .. a-8 on cached tower outputs              model/EventChatModel.py:199, 335-342 */
/* tower_out bf16 [B*n,577,1024] (what egpt_tower returned) -> tokens bf16 [B, n+577, 4096]:       */
/* visual_projector, feature_adaptor and the pooling without running the tower again; same op      */
/* order as egpt_encode_pool (pooled before the second projector linear).  Any n_frames >= 1, like */
/* egpt_encode_pool and egpt_visual_tokens (the reference's callers use 5, common.py:119; up to 8   */
/* frames the pooling keeps a frame per register set, beyond that it walks the features twice).     */
int egpt_project_pool(egpt_handle* h, const uint16_t* tower_out, int n_windows, int n_frames,
                      uint16_t* out_tokens, void* workspace, size_t workspace_bytes, void* stream);

/* f-1  multimodal splice of prepare_inputs_labels_for_multimodal  model/EventChatModel.py:355-465 */
/* One gather builds inputs_embeds [rows = B*L_out, c] (bf16): row i copies embed_table[src[i]]    */
/* when src[i] >= 0, event_features row -(src[i]+2) of the flattened [NF*T, c] matrix when         */
/* src[i] <= -2, and is zero (right padding) when src[i] == -1.  The host computes `src` (token    */
/* masking, <event> positions, truncation to 2048, padding) exactly as the reference's Python does. */
int egpt_splice_embeddings(const uint16_t* embed_table, const uint16_t* event_features, const int32_t* src,
                           int64_t rows, int c, uint16_t* out, void* stream);

/* a-3..a-9 fused: events -> tokens with no host round trip      common/common.py:110-129 +      */
/*                                                               model/EventChatModel.py:326-343 */
size_t egpt_visual_tokens_workspace_bytes(int n_windows, int n_frames, int H, int W);
int egpt_visual_tokens(egpt_handle* h, const void* x, const void* y, const void* p,
                       int x_dtype, int y_dtype, int p_dtype,
                       const int64_t* window_offsets, int n_windows, int n_frames, int H, int W,
                       uint16_t* out_tokens, int32_t* out_hw, int32_t* status,
                       void* workspace, size_t workspace_bytes, void* stream);

/* the same from (start, end) pairs of one recording (see egpt_raster_ranges) */
int egpt_visual_tokens_ranges(egpt_handle* h, const void* x, const void* y, const void* p,
                              int x_dtype, int y_dtype, int p_dtype,
                              const int64_t* window_bounds, int n_windows, int n_frames, int H, int W,
                              uint16_t* out_tokens, int32_t* out_hw, int32_t* status,
                              void* workspace, size_t workspace_bytes, void* stream);

/* Building block exposed for tests and profiling: C = epilogue(A[M,K] . B[N,K]^T + bias).       */
/* epilogue: 0 bias -> bf16, 1 bias+QuickGELU -> bf16, 2 bias+GELU(erf) -> bf16,                  */
/*           3 out_f32[M,N] += acc + bias (fp32 residual), 4 plain fp32 store.                    */
int egpt_gemm_bf16(const uint16_t* A, const uint16_t* B, const float* bias, void* out,
                   int64_t M, int N, int K, int epilogue, void* stream);

/* Building block: softmax(Q K^T) V per (frame, head) on the packed qkv buffer [F*577, 3072]     */
/* (q pre-scaled) -> out bf16 [F*577, 1024].                                                     */
int egpt_attention(const uint16_t* qkv, int n_frames_total, uint16_t* out, void* stream);

/* number of kernel launches issued by this library on the calling thread since the last reset   */
int64_t egpt_launch_count(int reset);

/* Measurement aid (no counterpart in the reference): while enabled on the calling thread, every  */
/* kernel family launch is bracketed by CUDA events on its own stream.  egpt_profile_collect      */
/* synchronises those events, adds up per category {0 gemm, 1 attention, 2 layernorm/elementwise, */
/* 3 raster, 4 preprocess, 5 pooling, 6 split} the device milliseconds, the launches and the      */
/* algorithmic work (FLOPs for 0-1, bytes for 2-6), writes three arrays of EGPT_PROFILE_CATEGORIES */
/* entries and clears the record.                                                                 */
#define EGPT_PROFILE_CATEGORIES 8
int egpt_profile_enable(int on);
int egpt_profile_collect(double* ms, int64_t* launches, double* work);

#ifdef __cplusplus
}
#endif
#endif /* EGPT_B200_H */
