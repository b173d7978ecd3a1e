// C ABI: model handle, tower / projector / pooling orchestration, fused events -> tokens path.
#include "common.cuh"
#include <string.h>
#include <vector>

namespace egpt {

// ---- error text + launch counter (thread local; no shared mutable state) ------------------------
static thread_local char g_err[512] = "";
static thread_local long long g_launches = 0;
void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
void count_launch(int n) { g_launches += n; }

struct ProfRec { int cat; double work; cudaEvent_t a, b; };
static thread_local bool g_prof_on = false;
static thread_local std::vector<ProfRec>* g_prof = nullptr;
void prof_begin(int category, double work, cudaStream_t stream) {
  if (!g_prof_on) return;
  if (!g_prof) g_prof = new std::vector<ProfRec>();
  ProfRec r{category, work, nullptr, nullptr};
  cudaEventCreate(&r.a);
  cudaEventCreate(&r.b);
  cudaEventRecord(r.a, stream);
  g_prof->push_back(r);
}
void prof_end(cudaStream_t stream) {
  if (!g_prof_on || !g_prof || g_prof->empty()) return;
  cudaEventRecord(g_prof->back().b, stream);
}

// kernels implemented in the other translation units
int gemm_bf16(const void* A, const void* B, const float* bias, void* out, long long M, int N, int K, int epilogue, cudaStream_t,
              void* out2 = nullptr, const float* colsum = nullptr, float* stats = nullptr);
int ln_fold(void* w, const float* gamma, const float* beta, const float* bias, float* colsum, float* bias_out, int N, int K, cudaStream_t);
int attention(const void* qkv, int n_frames_total, void* out, cudaStream_t);
int layernorm_bf16(const float* h, const float* gamma, const float* beta, void* y, long long rows, cudaStream_t);
int embed_preln(const float* patch, const float* cls, const float* pos, const float* gamma, const float* beta, float* h, void* hb,
                float* stats, long long rows, cudaStream_t);
int f32_to_bf16(const float* in, void* out, long long n, cudaStream_t);
int im2col(const uint16_t* pix, uint16_t* patches, int F, cudaStream_t);
int scale_bf16(const void* in, void* out, float scale, long long n, cudaStream_t);
int scale_f32(const float* in, float* out, float scale, long long n, cudaStream_t);
int pad_patch_weight(const void* in, void* out, cudaStream_t);

enum { EPI_BIAS_BF16 = 0, EPI_QUICKGELU_BF16 = 1, EPI_GELU_BF16 = 2, EPI_RESIDUAL_F32 = 3, EPI_STORE_F32 = 4,
       EPI_RESIDUAL_LN = 5, EPI_LN_BIAS_BF16 = 6, EPI_LN_QUICKGELU_BF16 = 7 };

constexpr int kHid = EGPT_HIDDEN, kTok = EGPT_TOKENS, kLlm = EGPT_LLM_HIDDEN, kMlp = 4096, kPatchK = 640;

// layer_norm1 / layer_norm2 are folded into qkv / fc1 at set-up: qkv_w, fc1_w hold gamma . W, qkv_b, fc1_b hold
// beta W^T + b, qkv_c, fc1_c the column sums of the folded weights (see ln_fold_kernel)
struct Layer {
  float *ln1_w, *ln1_b, *ln2_w, *ln2_b;
  uint16_t *qkv_w, *o_w, *fc1_w, *fc2_w;
  float *qkv_b, *o_b, *fc1_b, *fc2_b;
  float *qkv_c, *fc1_c;
};

}  // namespace egpt

struct egpt_handle {
  int device;
  int num_layers;
  char* slab;          // one allocation holding every packed weight
  size_t slab_bytes;
  float *cls, *pos, *pre_w, *pre_b;
  uint16_t* patch_w;   // [1024, 640]
  std::vector<egpt::Layer> layers;
  uint16_t *proj0_w, *proj2_w, *adaptor_w;
  float *proj0_b, *proj2_b, *adaptor_b;
};

namespace egpt {

struct SlabPlan {
  size_t off = 0;
  size_t take(size_t bytes) { size_t o = off; off = align_up(off + bytes, 256); return o; }
};

// workspace layout of the tower / projector stage for F frames
struct TowerWs {
  float* h;          // [M,1024] fp32 residual stream
  uint16_t* y;       // [M,1024] bf16 copy of the residual stream (A operand of the LayerNorm-folding GEMMs)
  float* stats;      // [M,4,2] per-row sum / sum of squares of h, one pair per 256-column tile
  uint16_t* qkv;     // [M,3072] bf16        (aliases: im2col patches [Mp,640])
  uint16_t* attn;    // [M,1024] bf16
  uint16_t* mlp;     // [M,4096] bf16        (aliases: patch GEMM output fp32 [Mp,1024])
  uint16_t* z;       // [M,4096] bf16 projector output
  uint16_t* pooled;  // [B,582,4096] bf16 (pre-adaptor), sized for F rows of windows
  float* pool_partial;
  size_t pool_partial_bytes;
};

static size_t tower_ws_bytes(int F) {
  const size_t M = (size_t)F * kTok;
  size_t b = 0;
  b += align_up(M * kHid * 4, 1024);
  b += align_up(M * kHid * 2, 1024);
  b += align_up(M * 3 * kHid * 2, 1024);
  b += align_up(M * kHid * 2, 1024);
  b += align_up(M * kMlp * 2, 1024);
  b += align_up(M * kLlm * 2, 1024);
  b += align_up((size_t)F * (8 + kTok) * kLlm * 2, 1024);          // pooled (n_frames >= 1 per window)
  b += align_up((size_t)F * 32 * kLlm * 4, 1024);                  // pooling partial sums
  b += align_up(M * 8 * 4, 1024);                                  // LayerNorm statistics
  return b + 16 * 1024;
}

static bool carve_tower_ws(void* ws, size_t bytes, int F, TowerWs* o) {
  Arena a(ws, bytes);
  const size_t M = (size_t)F * kTok;
  o->h = (float*)a.take(M * kHid * 4);
  o->y = (uint16_t*)a.take(M * kHid * 2);
  o->qkv = (uint16_t*)a.take(M * 3 * kHid * 2);
  o->attn = (uint16_t*)a.take(M * kHid * 2);
  o->mlp = (uint16_t*)a.take(M * kMlp * 2);
  o->z = (uint16_t*)a.take(M * kLlm * 2);
  o->pooled = (uint16_t*)a.take((size_t)F * (8 + kTok) * kLlm * 2);
  o->pool_partial_bytes = (size_t)F * 32 * kLlm * 4;
  o->pool_partial = (float*)a.take(o->pool_partial_bytes);
  o->stats = (float*)a.take(M * 8 * 4);
  return a.ok();
}

// patches [F*576, 640] bf16 already in ws.qkv -> residual stream after the last layer in ws.h
static int run_tower_from_patches(egpt_handle* hd, int F, TowerWs& w, cudaStream_t st) {
  const long long M = (long long)F * kTok, Mp = (long long)F * (kTok - 1);
  float* patch_out = reinterpret_cast<float*>(w.mlp);
  EGPT_TRY(gemm_bf16(w.qkv, hd->patch_w, nullptr, patch_out, Mp, kHid, kPatchK, EPI_STORE_F32, st));
  EGPT_TRY(embed_preln(patch_out, hd->cls, hd->pos, hd->pre_w, hd->pre_b, w.h, w.y, w.stats, M, st));
  // Per layer (transformers CLIPEncoderLayer): the two LayerNorms never run as passes of their own.  Each residual GEMM
  // (out-proj, fc2) leaves the new h in fp32, a bf16 copy of it in y and the row sums in stats; the GEMM that consumes
  // LayerNorm(h) (qkv, fc1) multiplies the bf16 copy by gamma . W and normalises in its epilogue.
  for (int l = 0; l < hd->num_layers; ++l) {
    const Layer& L = hd->layers[l];
    EGPT_TRY(gemm_bf16(w.y, L.qkv_w, L.qkv_b, w.qkv, M, 3 * kHid, kHid, EPI_LN_BIAS_BF16, st, nullptr, L.qkv_c, w.stats));
    EGPT_TRY(attention(w.qkv, F, w.attn, st));
    EGPT_TRY(gemm_bf16(w.attn, L.o_w, L.o_b, w.h, M, kHid, kHid, EPI_RESIDUAL_LN, st, w.y, nullptr, w.stats));
    EGPT_TRY(gemm_bf16(w.y, L.fc1_w, L.fc1_b, w.mlp, M, kMlp, kHid, EPI_LN_QUICKGELU_BF16, st, nullptr, L.fc1_c, w.stats));
    EGPT_TRY(gemm_bf16(w.mlp, L.fc2_w, L.fc2_b, w.h, M, kHid, kMlp, EPI_RESIDUAL_LN, st, w.y, nullptr, w.stats));
  }
  return EGPT_OK;
}

// ws.h (fp32 last_hidden_state) -> visual_projector -> dst bf16 [M,4096]
static int run_projector(egpt_handle* hd, int F, TowerWs& w, uint16_t* dst, cudaStream_t st) {
  const long long M = (long long)F * kTok;
  // ws.y already holds the bf16 copy of the last hidden state (written by the last residual GEMM)
  EGPT_TRY(gemm_bf16(w.y, hd->proj0_w, hd->proj0_b, w.mlp, M, kLlm, kHid, EPI_GELU_BF16, st));
  EGPT_TRY(gemm_bf16(w.mlp, hd->proj2_w, hd->proj2_b, dst, M, kLlm, kLlm, EPI_BIAS_BF16, st));
  return EGPT_OK;
}

// ws.h (fp32 last_hidden_state of B windows x n frames) -> out_tokens [B, n+577, 4096]:
//   feature_adaptor(get_spatio_temporal_features(visual_projector(h)))          (model/EventChatModel.py:194-200, 326-343)
// The projector is Linear -> GELU -> Linear and the pooling is a mean, so the second Linear (affine) commutes with it:
//   pool(W2 g + b2) = W2 pool(g) + b2.  Pooling the GELU output first shrinks the second projector GEMM from
// n*577 to n+577 rows per window (5x fewer at n = 5); the adaptor then runs on the same pooled rows.
// hidden == nullptr: the last hidden state is what the tower left in ws.y; otherwise bf16 [M,1024] from the caller
static int run_projector_pool_adaptor(egpt_handle* hd, int B, int n, TowerWs& w, uint16_t* out_tokens, cudaStream_t st,
                                      const uint16_t* hidden = nullptr) {
  const long long M = (long long)B * n * kTok, Mo = (long long)B * (n + kTok);
  if (hidden == nullptr) hidden = w.y;       // the bf16 copy of the last hidden state, written by the last residual GEMM
  EGPT_TRY(gemm_bf16(hidden, hd->proj0_w, hd->proj0_b, w.mlp, M, kLlm, kHid, EPI_GELU_BF16, st));
  EGPT_TRY(egpt_spatio_temporal_pool(w.mlp, 0, B, n, kTok, kLlm, n, w.pooled, w.pool_partial, w.pool_partial_bytes, st));
  EGPT_TRY(gemm_bf16(w.pooled, hd->proj2_w, hd->proj2_b, w.z, Mo, kLlm, kLlm, EPI_BIAS_BF16, st));
  EGPT_TRY(gemm_bf16(w.z, hd->adaptor_w, hd->adaptor_b, out_tokens, Mo, kLlm, kLlm, EPI_BIAS_BF16, st));
  return EGPT_OK;
}

}  // namespace egpt

using namespace egpt;

extern "C" const char* egpt_last_error(void) { return g_err; }
extern "C" int egpt_version(void) { return 100; }
extern "C" int64_t egpt_launch_count(int reset) {
  const long long v = g_launches;
  if (reset) g_launches = 0;
  return v;
}

extern "C" int egpt_profile_enable(int on) {
  g_prof_on = on != 0;
  return EGPT_OK;
}
extern "C" int egpt_profile_collect(double* ms, int64_t* launches, double* work) {
  EGPT_CHECK_ARG(ms && launches && work, "egpt_profile_collect: null argument");
  for (int i = 0; i < EGPT_PROFILE_CATEGORIES; ++i) { ms[i] = 0; launches[i] = 0; work[i] = 0; }
  if (!g_prof) return EGPT_OK;
  for (ProfRec& r : *g_prof) {
    float t = 0.f;
    if (cudaEventSynchronize(r.b) == cudaSuccess && cudaEventElapsedTime(&t, r.a, r.b) == cudaSuccess &&
        r.cat >= 0 && r.cat < EGPT_PROFILE_CATEGORIES) {
      ms[r.cat] += t; launches[r.cat] += 1; work[r.cat] += r.work;
    }
    cudaEventDestroy(r.a);
    cudaEventDestroy(r.b);
  }
  g_prof->clear();
  return EGPT_OK;
}

extern "C" int egpt_create(const egpt_weights* wh, int device, egpt_handle** out) {
  EGPT_CHECK_ARG(wh && out, "egpt_create: null argument");
  EGPT_CHECK_ARG(wh->num_layers >= 0 && wh->num_layers <= 64 && (wh->num_layers == 0 || wh->layers),
                 "egpt_create: num_layers=%d", wh->num_layers);
  EGPT_CHECK_ARG(wh->class_embedding && wh->patch_weight && wh->position_embedding && wh->pre_ln_w && wh->pre_ln_b &&
                 wh->proj0_w && wh->proj0_b && wh->proj2_w && wh->proj2_b && wh->adaptor_w && wh->adaptor_b,
                 "egpt_create: a weight pointer is null");
  EGPT_CUDA(cudaSetDevice(device));
  egpt_handle* h = new egpt_handle();
  h->device = device;
  h->num_layers = wh->num_layers;
  const int L = wh->num_layers;
  SlabPlan p;
  const size_t o_cls = p.take(kHid * 4), o_pos = p.take((size_t)kTok * kHid * 4), o_prew = p.take(kHid * 4),
               o_preb = p.take(kHid * 4), o_patch = p.take((size_t)kHid * kPatchK * 2);
  struct LOff { size_t ln1w, ln1b, ln2w, ln2b, qkvw, ow, fc1w, fc2w, qkvb, ob, fc1b, fc2b, qkvc, fc1c; };
  std::vector<LOff> lo(L);
  for (int l = 0; l < L; ++l) {
    lo[l].ln1w = p.take(kHid * 4); lo[l].ln1b = p.take(kHid * 4); lo[l].ln2w = p.take(kHid * 4); lo[l].ln2b = p.take(kHid * 4);
    lo[l].qkvw = p.take((size_t)3 * kHid * kHid * 2); lo[l].ow = p.take((size_t)kHid * kHid * 2);
    lo[l].fc1w = p.take((size_t)kMlp * kHid * 2); lo[l].fc2w = p.take((size_t)kHid * kMlp * 2);
    lo[l].qkvb = p.take(3 * kHid * 4); lo[l].ob = p.take(kHid * 4); lo[l].fc1b = p.take(kMlp * 4); lo[l].fc2b = p.take(kHid * 4);
    lo[l].qkvc = p.take(3 * kHid * 4); lo[l].fc1c = p.take(kMlp * 4);
  }
  const size_t o_p0w = p.take((size_t)kLlm * kHid * 2), o_p0b = p.take(kLlm * 4), o_p2w = p.take((size_t)kLlm * kLlm * 2),
               o_p2b = p.take(kLlm * 4), o_aw = p.take((size_t)kLlm * kLlm * 2), o_ab = p.take(kLlm * 4);
  h->slab_bytes = p.off;
  cudaError_t e = cudaMalloc((void**)&h->slab, h->slab_bytes);
  if (e != cudaSuccess) { set_error("egpt_create: cudaMalloc(%zu) failed: %s", h->slab_bytes, cudaGetErrorString(e)); delete h; return EGPT_ERR_CUDA; }
  char* s = h->slab;
  cudaStream_t st = 0;
  int rc = EGPT_OK;
  auto cp = [&](size_t off, const void* src, size_t bytes) {
    if (rc == EGPT_OK && cudaMemcpyAsync(s + off, src, bytes, cudaMemcpyDeviceToDevice, st) != cudaSuccess) {
      set_error("egpt_create: weight copy failed: %s", cudaGetErrorString(cudaGetLastError()));
      rc = EGPT_ERR_CUDA;
    }
  };
  cp(o_cls, wh->class_embedding, kHid * 4); cp(o_pos, wh->position_embedding, (size_t)kTok * kHid * 4);
  cp(o_prew, wh->pre_ln_w, kHid * 4); cp(o_preb, wh->pre_ln_b, kHid * 4);
  if (rc == EGPT_OK) rc = pad_patch_weight(wh->patch_weight, s + o_patch, st);
  h->cls = (float*)(s + o_cls); h->pos = (float*)(s + o_pos); h->pre_w = (float*)(s + o_prew); h->pre_b = (float*)(s + o_preb);
  h->patch_w = (uint16_t*)(s + o_patch);
  h->layers.resize(L);
  for (int l = 0; l < L && rc == EGPT_OK; ++l) {
    const egpt_layer_weights& w = wh->layers[l];
    if (!(w.ln1_w && w.ln1_b && w.q_w && w.k_w && w.v_w && w.o_w && w.q_b && w.k_b && w.v_b && w.o_b && w.ln2_w && w.ln2_b &&
          w.fc1_w && w.fc1_b && w.fc2_w && w.fc2_b)) { set_error("egpt_create: layer %d has a null weight", l); rc = EGPT_ERR_INVALID; break; }
    Layer& Lr = h->layers[l];
    cp(lo[l].ln1w, w.ln1_w, kHid * 4); cp(lo[l].ln1b, w.ln1_b, kHid * 4); cp(lo[l].ln2w, w.ln2_w, kHid * 4); cp(lo[l].ln2b, w.ln2_b, kHid * 4);
    // q is pre-scaled by head_dim^-0.5 = 0.125 (exact in bf16/fp32): CLIPAttention scales q after the projection
    if (rc == EGPT_OK) rc = scale_bf16(w.q_w, s + lo[l].qkvw, 0.125f, (long long)kHid * kHid, st);
    cp(lo[l].qkvw + (size_t)kHid * kHid * 2, w.k_w, (size_t)kHid * kHid * 2);
    cp(lo[l].qkvw + (size_t)2 * kHid * kHid * 2, w.v_w, (size_t)kHid * kHid * 2);
    if (rc == EGPT_OK) rc = scale_f32((const float*)w.q_b, (float*)(s + lo[l].qkvb), 0.125f, kHid, st);
    cp(lo[l].qkvb + kHid * 4, w.k_b, kHid * 4); cp(lo[l].qkvb + 2 * kHid * 4, w.v_b, kHid * 4);
    cp(lo[l].ow, w.o_w, (size_t)kHid * kHid * 2); cp(lo[l].ob, w.o_b, kHid * 4);
    cp(lo[l].fc1w, w.fc1_w, (size_t)kMlp * kHid * 2); cp(lo[l].fc1b, w.fc1_b, kMlp * 4);
    cp(lo[l].fc2w, w.fc2_w, (size_t)kHid * kMlp * 2); cp(lo[l].fc2b, w.fc2_b, kHid * 4);
    Lr.ln1_w = (float*)(s + lo[l].ln1w); Lr.ln1_b = (float*)(s + lo[l].ln1b); Lr.ln2_w = (float*)(s + lo[l].ln2w); Lr.ln2_b = (float*)(s + lo[l].ln2b);
    Lr.qkv_w = (uint16_t*)(s + lo[l].qkvw); Lr.o_w = (uint16_t*)(s + lo[l].ow); Lr.fc1_w = (uint16_t*)(s + lo[l].fc1w); Lr.fc2_w = (uint16_t*)(s + lo[l].fc2w);
    Lr.qkv_b = (float*)(s + lo[l].qkvb); Lr.o_b = (float*)(s + lo[l].ob); Lr.fc1_b = (float*)(s + lo[l].fc1b); Lr.fc2_b = (float*)(s + lo[l].fc2b);
    Lr.qkv_c = (float*)(s + lo[l].qkvc); Lr.fc1_c = (float*)(s + lo[l].fc1c);
    // fold layer_norm1 into the packed qkv projection and layer_norm2 into fc1 (in place, after the copies above)
    if (rc == EGPT_OK) rc = ln_fold(Lr.qkv_w, Lr.ln1_w, Lr.ln1_b, Lr.qkv_b, Lr.qkv_c, Lr.qkv_b, 3 * kHid, kHid, st);
    if (rc == EGPT_OK) rc = ln_fold(Lr.fc1_w, Lr.ln2_w, Lr.ln2_b, Lr.fc1_b, Lr.fc1_c, Lr.fc1_b, kMlp, kHid, st);
  }
  cp(o_p0w, wh->proj0_w, (size_t)kLlm * kHid * 2); cp(o_p0b, wh->proj0_b, kLlm * 4);
  cp(o_p2w, wh->proj2_w, (size_t)kLlm * kLlm * 2); cp(o_p2b, wh->proj2_b, kLlm * 4);
  cp(o_aw, wh->adaptor_w, (size_t)kLlm * kLlm * 2); cp(o_ab, wh->adaptor_b, kLlm * 4);
  h->proj0_w = (uint16_t*)(s + o_p0w); h->proj0_b = (float*)(s + o_p0b); h->proj2_w = (uint16_t*)(s + o_p2w); h->proj2_b = (float*)(s + o_p2b);
  h->adaptor_w = (uint16_t*)(s + o_aw); h->adaptor_b = (float*)(s + o_ab);
  if (rc == EGPT_OK && cudaStreamSynchronize(st) != cudaSuccess) { set_error("egpt_create: sync failed: %s", cudaGetErrorString(cudaGetLastError())); rc = EGPT_ERR_CUDA; }
  if (rc != EGPT_OK) { cudaFree(h->slab); delete h; return rc; }
  *out = h;
  return EGPT_OK;
}

extern "C" int egpt_destroy(egpt_handle* h) {
  if (!h) return EGPT_OK;
  cudaFree(h->slab);
  delete h;
  return EGPT_OK;
}

extern "C" size_t egpt_workspace_bytes(int n_frames_total) {
  return n_frames_total > 0 ? tower_ws_bytes(n_frames_total) : 0;
}

static int check_device(const egpt_handle* h, const char* what) {
  int dev = -1;
  EGPT_CUDA(cudaGetDevice(&dev));
  if (dev != h->device) {
    set_error("%s: the handle was created on device %d but device %d is current", what, h->device, dev);
    return EGPT_ERR_INVALID;
  }
  return EGPT_OK;
}

#define EGPT_WS_OR_FAIL(F)                                                                       \
  TowerWs w;                                                                                     \
  if (!carve_tower_ws(workspace, workspace_bytes, (F), &w)) {                                    \
    set_error("workspace %zu < %zu bytes for %d frames", workspace_bytes, tower_ws_bytes(F), (F)); \
    return EGPT_ERR_WORKSPACE;                                                                   \
  }

extern "C" int egpt_tower(egpt_handle* h, const uint16_t* pixel_values, int F, uint16_t* out_hidden, void* workspace,
                          size_t workspace_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  EGPT_CHECK_ARG(h && pixel_values && out_hidden && workspace && F > 0, "egpt_tower: bad arguments");
  EGPT_TRY(check_device(h, "egpt_tower"));
  EGPT_WS_OR_FAIL(F);
  EGPT_TRY(im2col(pixel_values, w.qkv, F, st));
  EGPT_TRY(run_tower_from_patches(h, F, w, st));
  EGPT_CUDA(cudaMemcpyAsync(out_hidden, w.y, (size_t)F * kTok * kHid * 2, cudaMemcpyDeviceToDevice, st));
  return EGPT_OK;
}

extern "C" int egpt_visval_encode(egpt_handle* h, const uint16_t* pixel_values, int F, uint16_t* out_features,
                                  void* workspace, size_t workspace_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  EGPT_CHECK_ARG(h && pixel_values && out_features && workspace && F > 0, "egpt_visval_encode: bad arguments");
  EGPT_TRY(check_device(h, "egpt_visval_encode"));
  EGPT_WS_OR_FAIL(F);
  EGPT_TRY(im2col(pixel_values, w.qkv, F, st));
  EGPT_TRY(run_tower_from_patches(h, F, w, st));
  EGPT_TRY(run_projector(h, F, w, out_features, st));
  return EGPT_OK;
}

extern "C" int egpt_feature_adaptor(egpt_handle* h, const uint16_t* in, int64_t rows, uint16_t* out, void* stream) {
  EGPT_CHECK_ARG(h && in && out && rows > 0, "egpt_feature_adaptor: bad arguments");
  EGPT_TRY(check_device(h, "egpt_feature_adaptor"));
  return gemm_bf16(in, h->adaptor_w, h->adaptor_b, out, rows, kLlm, kLlm, EPI_BIAS_BF16, (cudaStream_t)stream);
}

extern "C" int egpt_encode_pool(egpt_handle* h, const uint16_t* pixel_values, int n_windows, int n_frames,
                                uint16_t* out_tokens, void* workspace, size_t workspace_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  EGPT_CHECK_ARG(h && pixel_values && out_tokens && workspace && n_windows > 0 && n_frames > 0,
                 "egpt_encode_pool: bad arguments");
  EGPT_TRY(check_device(h, "egpt_encode_pool"));
  const int F = n_windows * n_frames;
  EGPT_WS_OR_FAIL(F);
  EGPT_TRY(im2col(pixel_values, w.qkv, F, st));
  EGPT_TRY(run_tower_from_patches(h, F, w, st));
  EGPT_TRY(run_projector_pool_adaptor(h, n_windows, n_frames, w, out_tokens, st));
  return EGPT_OK;
}

extern "C" int egpt_project_pool(egpt_handle* h, const uint16_t* tower_out, int n_windows, int n_frames,
                                 uint16_t* out_tokens, void* workspace, size_t workspace_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  EGPT_CHECK_ARG(h && tower_out && out_tokens && workspace && n_windows > 0 && n_frames > 0,
                 "egpt_project_pool: bad arguments");
  EGPT_CHECK_ARG(((uintptr_t)tower_out & 15) == 0 && ((uintptr_t)out_tokens & 15) == 0, "egpt_project_pool: pointers must be 16-byte aligned");
  EGPT_TRY(check_device(h, "egpt_project_pool"));
  const int F = n_windows * n_frames;
  EGPT_WS_OR_FAIL(F);
  EGPT_TRY(run_projector_pool_adaptor(h, n_windows, n_frames, w, out_tokens, st, tower_out));
  return EGPT_OK;
}

// ---- fused events -> tokens ---------------------------------------------------------------------
extern "C" size_t egpt_visual_tokens_workspace_bytes(int n_windows, int n_frames, int H, int W) {
  if (n_windows <= 0 || n_frames <= 0 || H <= 0 || W <= 0) return 0;
  const int F = n_windows * n_frames;
  return align_up(egpt_raster_frame_stride(H, W) * (size_t)F, 1024) +
         align_up(egpt_raster_workspace_bytes(n_windows, n_frames, H, W), 1024) +
         align_up(egpt_preprocess_workspace_bytes(F, H), 1024) + tower_ws_bytes(F) + 8 * 1024;
}

namespace egpt {
int raster_impl(const void* x, const void* y, const void* p, int x_dtype, int y_dtype, int p_dtype, const int64_t* window_offsets,
                int off_stride, int n_windows, int n_frames, int mode, int H, int W, int fixed_canvas, uint8_t* out_frames,
                int32_t* out_hw, int32_t* out_counts, int32_t* status, void* workspace, size_t workspace_bytes, bool shrink_to_fit,
                cudaStream_t stream);

static int visual_tokens_impl(egpt_handle* h, const void* x, const void* y, const void* p, int x_dtype, int y_dtype, int p_dtype,
                              const int64_t* windows, int off_stride, int n_windows, int n_frames, int H, int W,
                              uint16_t* out_tokens, int32_t* out_hw, int32_t* status, void* workspace, size_t workspace_bytes,
                              cudaStream_t st) {
  EGPT_CHECK_ARG(h && x && y && p && windows && out_tokens && out_hw && status && workspace,
                 "egpt_visual_tokens: null pointer argument");
  EGPT_CHECK_ARG(n_windows > 0 && n_frames > 0 && H > 0 && W > 0, "egpt_visual_tokens: bad shape");
  EGPT_TRY(check_device(h, "egpt_visual_tokens"));
  if (workspace_bytes < egpt_visual_tokens_workspace_bytes(n_windows, n_frames, H, W)) {
    set_error("egpt_visual_tokens: workspace %zu < %zu bytes", workspace_bytes,
              egpt_visual_tokens_workspace_bytes(n_windows, n_frames, H, W));
    return EGPT_ERR_WORKSPACE;
  }
  const int F = n_windows * n_frames;
  Arena a(workspace, workspace_bytes);
  const size_t frame_stride = egpt_raster_frame_stride(H, W);
  uint8_t* frames = (uint8_t*)a.take(frame_stride * (size_t)F);
  const size_t rws = egpt_raster_workspace_bytes(n_windows, n_frames, H, W);
  void* raster_ws = a.take(rws);
  const size_t pws = egpt_preprocess_workspace_bytes(F, H);
  void* pre_ws = a.take(pws);
  const size_t tws = tower_ws_bytes(F);
  void* tower_ws = a.take(tws);
  if (!a.ok()) { set_error("egpt_visual_tokens: workspace carve failed"); return EGPT_ERR_WORKSPACE; }
  TowerWs w;
  if (!carve_tower_ws(tower_ws, tws, F, &w)) { set_error("egpt_visual_tokens: tower workspace carve failed"); return EGPT_ERR_WORKSPACE; }
  EGPT_TRY(raster_impl(x, y, p, x_dtype, y_dtype, p_dtype, windows, off_stride, n_windows, n_frames, EGPT_RASTER_LAST_WINS, H, W,
                       0, frames, out_hw, nullptr, status, raster_ws, rws, false, st));
  // resized + normalised pixels go straight into the im2col layout of the patch GEMM
  EGPT_TRY(egpt_preprocess(frames, frame_stride, out_hw, F, H, nullptr, w.qkv, status, pre_ws, pws, st));
  EGPT_TRY(run_tower_from_patches(h, F, w, st));
  EGPT_TRY(run_projector_pool_adaptor(h, n_windows, n_frames, w, out_tokens, st));
  return EGPT_OK;
}
}  // namespace egpt

extern "C" int egpt_visual_tokens(egpt_handle* h, const void* x, const void* y, const void* p, int x_dtype, int y_dtype,
                                  int p_dtype, const int64_t* window_offsets, int n_windows, int n_frames, int H, int W,
                                  uint16_t* out_tokens, int32_t* out_hw, int32_t* status, void* workspace,
                                  size_t workspace_bytes, void* stream) {
  return visual_tokens_impl(h, x, y, p, x_dtype, y_dtype, p_dtype, window_offsets, 1, n_windows, n_frames, H, W, out_tokens, out_hw,
                            status, workspace, workspace_bytes, (cudaStream_t)stream);
}

extern "C" int egpt_visual_tokens_ranges(egpt_handle* h, const void* x, const void* y, const void* p, int x_dtype, int y_dtype,
                                         int p_dtype, const int64_t* window_bounds, int n_windows, int n_frames, int H, int W,
                                         uint16_t* out_tokens, int32_t* out_hw, int32_t* status, void* workspace,
                                         size_t workspace_bytes, void* stream) {
  return visual_tokens_impl(h, x, y, p, x_dtype, y_dtype, p_dtype, window_bounds, 2, n_windows, n_frames, H, W, out_tokens, out_hw,
                            status, workspace, workspace_bytes, (cudaStream_t)stream);
}
