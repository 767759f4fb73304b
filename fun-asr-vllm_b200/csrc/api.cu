// C-ABI orchestration: error state, device check, the encoder/adaptor plan (workspace carve-up +
// TMA descriptors built once) and the layer schedule.  No device allocation, no synchronisation.
#include <cstring>
#include <new>
#include <vector>

#include "common.cuh"

namespace fa {

static thread_local char g_err[1024] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
  }
  return n;
}

// ---- profiling: event pairs around every launch, aggregated per kernel class on collect ----
bool g_prof_on = false;
namespace {
struct ProfRec { int kind; cudaEvent_t a, b; };
std::vector<ProfRec> g_prof_recs;
std::vector<cudaEvent_t> g_prof_pool;
cudaEvent_t prof_event() {
  cudaEvent_t e = nullptr;
  if (!g_prof_pool.empty()) { e = g_prof_pool.back(); g_prof_pool.pop_back(); }
  else cudaEventCreate(&e);
  return e;
}
}  // namespace
void prof_mark(int kind, cudaStream_t s, bool begin) {
  if (begin) {
    ProfRec r{kind, prof_event(), nullptr};
    cudaEventRecord(r.a, s);
    g_prof_recs.push_back(r);
  } else if (!g_prof_recs.empty()) {
    ProfRec& r = g_prof_recs.back();
    r.b = prof_event();
    cudaEventRecord(r.b, s);
  }
}

}  // namespace fa

using namespace fa;

extern "C" int fa_profile_enable(int32_t on) {
  g_prof_on = on != 0;
  return 0;
}
extern "C" int fa_profile_collect(float* ms_per_kind, int32_t* launches_per_kind, int32_t n_kinds) {
  FA_REQUIRE(ms_per_kind && launches_per_kind && n_kinds >= K_NUM, "profile_collect: need %d slots", (int)K_NUM);
  for (int i = 0; i < n_kinds; ++i) { ms_per_kind[i] = 0.f; launches_per_kind[i] = 0; }
  for (ProfRec& r : g_prof_recs) {
    if (r.b) {
      FA_CHECK_CUDA(cudaEventSynchronize(r.b));
      float ms = 0.f;
      FA_CHECK_CUDA(cudaEventElapsedTime(&ms, r.a, r.b));
      ms_per_kind[r.kind] += ms;
      launches_per_kind[r.kind] += 1;
      g_prof_pool.push_back(r.b);
    }
    g_prof_pool.push_back(r.a);
  }
  g_prof_recs.clear();
  return 0;
}

extern "C" int fa_abi_version(void) { return FA_ABI_VERSION; }
extern "C" const char* fa_last_error(void) { return g_err; }

extern "C" int fa_check_device(int32_t* sm, int32_t* major, int32_t* minor) {
  int dev = 0, mj = 0, mn = 0, n = 0;
  FA_CHECK_CUDA(cudaGetDevice(&dev));
  FA_CHECK_CUDA(cudaDeviceGetAttribute(&mj, cudaDevAttrComputeCapabilityMajor, dev));
  FA_CHECK_CUDA(cudaDeviceGetAttribute(&mn, cudaDevAttrComputeCapabilityMinor, dev));
  FA_CHECK_CUDA(cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev));
  if (sm) *sm = n;
  if (major) *major = mj;
  if (minor) *minor = mn;
  FA_REQUIRE(mj == 10, "libfunasr_b200 is built for sm_100a only; device reports sm_%d%d", mj, mn);
  return 0;
}

// ------------------------------------------------------------------------------------------ plan
namespace {

struct GemmOp {
  TensorMap2D ta;     // activations [cap, K], box 128 x 64
  TensorMap2D tb[3];  // weights [N, K], box {64,128,256} x 64 (one per tile width)
  bool tb_ok[3] = {false, false, false};
  int N = 0, K = 0;
  const float* bias = nullptr;
};

struct SanmLayerPlan {
  fa_sanm_layer_w w;
  GemmOp qkv, out, ffn1, ffn2;
  int d_in;
};
struct AdaptorLayerPlan {
  fa_adaptor_layer_w w;
  GemmOp qkv, out, ffn1, ffn2;
};

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct Workspace {
  // all row counts are cap = max_tokens rounded up to 128
  float* x;              // [cap, d_model]   residual stream (f32)
  float* x2;             // [cap, d_model]   ping-pong for the residual stream
  __nv_bfloat16* xn0;    // [cap, d_in]      LN'd input of layer 0
  __nv_bfloat16* xn;     // [cap, d_model]   LN output (GEMM A operand)
  __nv_bfloat16* qkv;    // [cap, 3*d_model]
  float* mem;            // [cap, d_model]   FSMN memory
  __nv_bfloat16* ctx;    // [cap, d_model]
  __nv_bfloat16* h;      // [cap, max(d_ff, ad_ffn)]
  __nv_bfloat16* enc_bf; // [cap, d_model]   encoder output, bf16 (adaptor input)
  float* enc_f32;        // [cap, d_model]   encoder output, f32
  float* ax;             // [cap, ad_dim]
  float* ax2;            // [cap, ad_dim]
  __nv_bfloat16* axn;    // [cap, ad_dim]
  __nv_bfloat16* aqkv;   // [cap, 3*ad_dim]
  __nv_bfloat16* actx;   // [cap, ad_dim]
  __nv_bfloat16* ah;     // [cap, ad_ff]
  size_t bytes;
};

Workspace carve(const fa_model_cfg& c, int cap, unsigned char* base) {
  Workspace w;
  size_t off = 0;
  auto take = [&](size_t bytes) {
    unsigned char* p = base ? base + off : nullptr;
    off = align_up(off + bytes, 1024);
    return p;
  };
  const size_t n = static_cast<size_t>(cap);
  const int hmax = c.d_ff > c.ad_ffn ? c.d_ff : c.ad_ffn;
  w.x = reinterpret_cast<float*>(take(n * c.d_model * 4));
  w.x2 = reinterpret_cast<float*>(take(n * c.d_model * 4));
  w.xn0 = reinterpret_cast<__nv_bfloat16*>(take(n * c.d_in * 2));
  w.xn = reinterpret_cast<__nv_bfloat16*>(take(n * c.d_model * 2));
  w.qkv = reinterpret_cast<__nv_bfloat16*>(take(n * 3 * c.d_model * 2));
  w.mem = reinterpret_cast<float*>(take(n * c.d_model * 4));
  w.ctx = reinterpret_cast<__nv_bfloat16*>(take(n * c.d_model * 2));
  w.h = reinterpret_cast<__nv_bfloat16*>(take(n * hmax * 2));
  w.enc_bf = reinterpret_cast<__nv_bfloat16*>(take(n * c.d_model * 2));
  w.enc_f32 = reinterpret_cast<float*>(take(n * c.d_model * 4));
  w.ax = reinterpret_cast<float*>(take(n * c.ad_dim * 4));
  w.ax2 = reinterpret_cast<float*>(take(n * c.ad_dim * 4));
  w.axn = reinterpret_cast<__nv_bfloat16*>(take(n * c.ad_dim * 2));
  w.aqkv = reinterpret_cast<__nv_bfloat16*>(take(n * 3 * c.ad_dim * 2));
  w.actx = reinterpret_cast<__nv_bfloat16*>(take(n * c.ad_dim * 2));
  w.ah = reinterpret_cast<__nv_bfloat16*>(take(n * c.ad_ff * 2));
  w.bytes = off;
  return w;
}

int check_model_cfg(const fa_model_cfg* c) {
  FA_REQUIRE(c != nullptr, "model cfg is NULL");
  FA_REQUIRE(c->d_model % 128 == 0 && c->d_model / c->n_heads == 128, "encoder head dim must be 128 (d_model %d, heads %d)",
             c->d_model, c->n_heads);
  FA_REQUIRE(c->ad_dim % 128 == 0 && c->ad_dim / c->ad_heads == 128, "adaptor head dim must be 128 (dim %d, heads %d)",
             c->ad_dim, c->ad_heads);
  FA_REQUIRE(c->d_in % 8 == 0 && c->d_in <= 1024 && c->d_model <= 1024 && c->ad_dim <= 1024, "d_in/d_model/ad_dim must be <= 1024, d_in %% 8 == 0");
  FA_REQUIRE(c->d_ff % 64 == 0 && c->ad_ffn % 64 == 0 && c->ad_ff % 64 == 0, "FFN widths must be multiples of 64");
  FA_REQUIRE(c->n_blocks >= 0 && c->n_tp_blocks >= 0 && c->ad_layers >= 0, "bad block counts");
  FA_REQUIRE(c->n_blocks >= 1 || c->n_tp_blocks == 0, "tp blocks need at least one main block");
  FA_REQUIRE(c->ad_downsample == 1, "adaptor downsample_rate %d: only 1 is compiled", c->ad_downsample);
  FA_REQUIRE(c->fsmn_kernel >= 1 && c->fsmn_kernel <= 16 && c->sanm_shift >= 0, "bad FSMN kernel/shift");
  return 0;
}

}  // namespace

struct fa_plan {
  fa_model_cfg cfg;
  fa_model_w w;
  int cap;
  Workspace ws;
  std::vector<SanmLayerPlan> layers;
  std::vector<AdaptorLayerPlan> ad_layers;
  GemmOp lin1, lin2;
  int fsmn_left;
};

namespace {

int make_op(GemmOp* op, const void* a, int lda, const void* w, int N, int K, const float* bias, int cap) {
  op->N = N;
  op->K = K;
  op->bias = bias;
  FA_REQUIRE(w != nullptr, "NULL weight pointer (N=%d K=%d)", N, K);
  if (int r = make_tensor_map_bf16(&op->ta, a, cap, K, lda, 128)) return r;
  for (int i = 0; i < 3; ++i) {
    const int bn = 64 << i;
    if (N % bn != 0) continue;
    if (int r = make_tensor_map_bf16(&op->tb[i], w, N, K, K, bn)) return r;
    op->tb_ok[i] = true;
  }
  return 0;
}

}  // namespace

static int run_gemm(GemmOp& op, int M, const GemmEpilogue& e_in, void* c, int64_t ldc, cudaStream_t s) {
  const int bn = gemm_block_n(M, op.N);
  const int idx = bn == 64 ? 0 : (bn == 128 ? 1 : 2);
  FA_REQUIRE(bn != 0 && op.tb_ok[idx], "gemm N=%d unsupported", op.N);
  GemmEpilogue e = e_in;
  e.bias = op.bias;
  return launch_gemm(op.ta, op.tb[idx], bn, M, op.N, op.K, e, c, ldc, s);
}

extern "C" size_t fa_plan_workspace_bytes(const fa_model_cfg* cfg, int32_t max_tokens) {
  if (check_model_cfg(cfg) != 0 || max_tokens <= 0) return 0;
  const int cap = (max_tokens + 127) / 128 * 128;
  return carve(*cfg, cap, nullptr).bytes;
}

extern "C" int fa_plan_create(const fa_model_cfg* cfg, const fa_model_w* w, void* workspace_dev, size_t workspace_bytes,
                              int32_t max_tokens, fa_plan** plan_out) {
  if (int r = check_model_cfg(cfg)) return r;
  FA_REQUIRE(w && workspace_dev && plan_out && max_tokens > 0, "plan_create: bad arguments");
  FA_REQUIRE((reinterpret_cast<uintptr_t>(workspace_dev) & 1023) == 0, "workspace must be 1024-byte aligned");
  const int cap = (max_tokens + 127) / 128 * 128;
  Workspace ws = carve(*cfg, cap, static_cast<unsigned char*>(workspace_dev));
  FA_REQUIRE(ws.bytes <= workspace_bytes, "workspace too small: need %zu bytes, got %zu", ws.bytes, workspace_bytes);
  FA_REQUIRE(cfg->n_blocks == 0 ||
                 (w->layers && w->after_norm_g && w->after_norm_b && w->tp_norm_g && w->tp_norm_b && w->pe_inv_timescales),
             "plan_create: NULL encoder weights");
  FA_REQUIRE(cfg->n_blocks > 0 || w->lin1_w, "plan_create: neither encoder blocks nor adaptor weights given");
  fa_plan* p = new (std::nothrow) fa_plan();
  FA_REQUIRE(p != nullptr, "out of host memory");
  p->cfg = *cfg;
  p->w = *w;
  p->cap = cap;
  p->ws = ws;
  const int left = (cfg->fsmn_kernel - 1) / 2 + (cfg->sanm_shift > 0 ? cfg->sanm_shift : 0);  // VM:175-179
  p->fsmn_left = left;
  const int nl = cfg->n_blocks + cfg->n_tp_blocks;
  p->layers.resize(nl);
  int rc = 0;
  for (int i = 0; i < nl && rc == 0; ++i) {
    SanmLayerPlan& L = p->layers[i];
    L.w = w->layers[i];
    L.d_in = (i == 0) ? cfg->d_in : cfg->d_model;
    const void* a_in = (i == 0) ? static_cast<const void*>(ws.xn0) : static_cast<const void*>(ws.xn);
    rc = make_op(&L.qkv, a_in, L.d_in, L.w.qkv_w, 3 * cfg->d_model, L.d_in, L.w.qkv_b, cap);
    if (!rc) rc = make_op(&L.out, ws.ctx, cfg->d_model, L.w.out_w, cfg->d_model, cfg->d_model, L.w.out_b, cap);
    if (!rc) rc = make_op(&L.ffn1, ws.xn, cfg->d_model, L.w.ffn1_w, cfg->d_ff, cfg->d_model, L.w.ffn1_b, cap);
    if (!rc) rc = make_op(&L.ffn2, ws.h, cfg->d_ff, L.w.ffn2_w, cfg->d_model, cfg->d_ff, L.w.ffn2_b, cap);
  }
  if (!rc && cfg->ad_layers >= 0 && w->lin1_w) {
    rc = make_op(&p->lin1, ws.enc_bf, cfg->d_model, w->lin1_w, cfg->ad_ffn, cfg->d_model, w->lin1_b, cap);
    if (!rc) rc = make_op(&p->lin2, ws.h, cfg->ad_ffn, w->lin2_w, cfg->ad_dim, cfg->ad_ffn, w->lin2_b, cap);
    p->ad_layers.resize(cfg->ad_layers);
    for (int i = 0; i < cfg->ad_layers && rc == 0; ++i) {
      AdaptorLayerPlan& L = p->ad_layers[i];
      if (!w->ad_layers) { set_error("plan_create: NULL adaptor layers"); rc = 2; break; }
      L.w = w->ad_layers[i];
      rc = make_op(&L.qkv, ws.axn, cfg->ad_dim, L.w.qkv_w, 3 * cfg->ad_dim, cfg->ad_dim, L.w.qkv_b, cap);
      if (!rc) rc = make_op(&L.out, ws.actx, cfg->ad_dim, L.w.out_w, cfg->ad_dim, cfg->ad_dim, L.w.out_b, cap);
      if (!rc) rc = make_op(&L.ffn1, ws.axn, cfg->ad_dim, L.w.ffn1_w, cfg->ad_ff, cfg->ad_dim, L.w.ffn1_b, cap);
      if (!rc) rc = make_op(&L.ffn2, ws.ah, cfg->ad_ff, L.w.ffn2_w, cfg->ad_dim, cfg->ad_ff, L.w.ffn2_b, cap);
    }
  }
  if (rc) {
    delete p;
    return rc;
  }
  *plan_out = p;
  return 0;
}

extern "C" void fa_plan_destroy(fa_plan* plan) { delete plan; }

static int check_batch(const fa_plan* p, const int32_t* cu, int n_utts, int n_tokens, int max_len) {
  FA_REQUIRE(p != nullptr, "NULL plan");
  FA_REQUIRE(cu != nullptr && n_utts > 0, "empty batch");
  FA_REQUIRE(n_tokens > 0 && n_tokens <= p->cap, "n_tokens %d exceeds the plan capacity %d", n_tokens, p->cap);
  FA_REQUIRE(max_len > 0 && max_len <= n_tokens, "max_len %d out of range", max_len);
  return 0;
}

static int check_tmax(int tmax, int max_len, const char* what) {
  FA_REQUIRE(tmax == 0 || tmax >= max_len, "%s: padded row count %d is smaller than the longest utterance %d", what, tmax, max_len);
  return 0;
}

extern "C" int fa_encoder_forward(fa_plan* p, const float* feats, int32_t feats_tmax, const int32_t* cu, int32_t n_utts,
                                  int32_t M, int32_t max_len, float* enc_out, int32_t enc_out_tmax, fa_stream_t stream) {
  if (int r = check_batch(p, cu, n_utts, M, max_len)) return r;
  FA_REQUIRE(feats != nullptr, "NULL feats");
  FA_REQUIRE(p->cfg.n_blocks > 0, "plan was created without encoder weights");
  if (int r = check_tmax(feats_tmax, max_len, "feats")) return r;
  if (int r = check_tmax(enc_out_tmax, max_len, "enc_out")) return r;
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const fa_model_cfg& c = p->cfg;
  Workspace& ws = p->ws;
  const int d = c.d_model;
  const int nl = c.n_blocks + c.n_tp_blocks;
  float* x = ws.x;    // current residual stream
  float* xo = ws.x2;  // next
  int rc = 0;

  // input stage: xs*sqrt(d) + PE, then norm1 of encoders0 (VM:388-390, :121)
  rc = launch_embed_layernorm(feats, feats_tmax, p->w.pe_inv_timescales, cu, n_utts, sqrtf(static_cast<float>(d)), p->layers[0].w.ln1_g,
                              p->layers[0].w.ln1_b, ws.xn0, M, c.d_in, s);
  if (rc) return rc;
  for (int i = 0; i < nl; ++i) {
    SanmLayerPlan& L = p->layers[i];
    GemmEpilogue e;
    // q|k|v = LN1(x) W^T + b                                              VM:206
    if ((rc = run_gemm(L.qkv, M, e, ws.qkv, 3 * d, s))) return rc;
    // FSMN memory on v                                                     VM:249,181-201
    if ((rc = launch_fsmn(ws.qkv, 3 * d, 2 * d, L.w.fsmn_w, cu, n_utts, ws.mem, M, d, c.fsmn_kernel, p->fsmn_left, s))) return rc;
    // masked MHA                                                           VM:250-252,222-236
    if ((rc = launch_attention(ws.qkv, 3 * d, cu, n_utts, max_len, c.n_heads, ws.ctx, d, s))) return rc;
    // x' = [x +] ctx Wo^T + b + mem    (no residual when in_size != size)   VM:238,253,123-136
    e = GemmEpilogue();
    e.out_f32 = 1;
    e.res1 = ws.mem;
    e.res2 = (i == 0) ? nullptr : x;
    if ((rc = run_gemm(L.out, M, e, xo, d, s))) return rc;
    // LN2, FFN                                                              VM:138-140,412-431
    if ((rc = launch_layernorm(xo, L.w.ln2_g, L.w.ln2_b, nullptr, nullptr, ws.xn, nullptr, M, d, s))) return rc;
    e = GemmEpilogue();
    e.relu = 1;
    if ((rc = run_gemm(L.ffn1, M, e, ws.h, c.d_ff, s))) return rc;
    e = GemmEpilogue();
    e.out_f32 = 1;
    e.res1 = xo;
    if ((rc = run_gemm(L.ffn2, M, e, x, d, s))) return rc;  // residual stream back in x
    // next layer's norm1, with after_norm / tp_norm folded in                VM:400,408
    const bool last_main = (i == c.n_blocks - 1);
    const bool last = (i == nl - 1);
    if (last) {
      if (c.n_tp_blocks == 0) {
        // after_norm then tp_norm back to back (VM:400,408 with an empty tp_encoders list)
        if ((rc = launch_layernorm(x, p->w.after_norm_g, p->w.after_norm_b, nullptr, nullptr, nullptr, xo, M, d, s))) return rc;
        if ((rc = launch_layernorm(xo, p->w.tp_norm_g, p->w.tp_norm_b, nullptr, nullptr, ws.enc_bf, ws.enc_f32, M, d, s))) return rc;
      } else {
        if ((rc = launch_layernorm(x, p->w.tp_norm_g, p->w.tp_norm_b, nullptr, nullptr, ws.enc_bf, ws.enc_f32, M, d, s))) return rc;
      }
    } else if (last_main) {
      // x <- after_norm(x) becomes the residual stream of the tp blocks; xn <- norm1_next(x)
      const SanmLayerPlan& Ln = p->layers[i + 1];
      if ((rc = launch_layernorm(x, p->w.after_norm_g, p->w.after_norm_b, Ln.w.ln1_g, Ln.w.ln1_b, ws.xn, xo, M, d, s))) return rc;
      float* t = x; x = xo; xo = t;
    } else {
      const SanmLayerPlan& Ln = p->layers[i + 1];
      if ((rc = launch_layernorm(x, Ln.w.ln1_g, Ln.w.ln1_b, nullptr, nullptr, ws.xn, nullptr, M, d, s))) return rc;
    }
  }
  if (enc_out) rc = launch_repack_rows(ws.enc_f32, 0, enc_out, enc_out_tmax, 0, cu, n_utts, M, d, s);
  return rc;
}

extern "C" int fa_adaptor_forward(fa_plan* p, const float* enc, int32_t enc_tmax, const int32_t* cu, int32_t n_utts,
                                  int32_t M, int32_t max_len, void* out, int32_t out_tmax, int32_t out_bf16,
                                  fa_stream_t stream) {
  if (int r = check_batch(p, cu, n_utts, M, max_len)) return r;
  if (int r = check_tmax(enc_tmax, max_len, "enc")) return r;
  if (int r = check_tmax(out_tmax, max_len, "out")) return r;
  FA_REQUIRE(out != nullptr, "NULL out");
  FA_REQUIRE(p->w.lin1_w != nullptr, "plan was created without adaptor weights");
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const fa_model_cfg& c = p->cfg;
  Workspace& ws = p->ws;
  const int D = c.ad_dim;
  int rc = 0;
  if (enc) {
    if ((rc = launch_repack_rows(enc, enc_tmax, ws.enc_bf, 0, 1, cu, n_utts, M, c.d_model, s))) return rc;
  }
  // linear1 + ReLU, linear2                                                 VM:585-587
  GemmEpilogue e;
  e.relu = 1;
  if ((rc = run_gemm(p->lin1, M, e, ws.h, c.ad_ffn, s))) return rc;
  e = GemmEpilogue();
  e.out_f32 = 1;
  float* x = ws.ax;
  float* xo = ws.ax2;
  if ((rc = run_gemm(p->lin2, M, e, x, D, s))) return rc;
  for (int i = 0; i < c.ad_layers; ++i) {
    AdaptorLayerPlan& L = p->ad_layers[i];
    // pre-LN MHA block                                                       VM:434-455
    if ((rc = launch_layernorm(x, L.w.ln1_g, L.w.ln1_b, nullptr, nullptr, ws.axn, nullptr, M, D, s))) return rc;
    e = GemmEpilogue();
    if ((rc = run_gemm(L.qkv, M, e, ws.aqkv, 3 * D, s))) return rc;
    if ((rc = launch_attention(ws.aqkv, 3 * D, cu, n_utts, max_len, c.ad_heads, ws.actx, D, s))) return rc;
    e = GemmEpilogue();
    e.out_f32 = 1;
    e.res1 = x;
    if ((rc = run_gemm(L.out, M, e, xo, D, s))) return rc;
    if ((rc = launch_layernorm(xo, L.w.ln2_g, L.w.ln2_b, nullptr, nullptr, ws.axn, nullptr, M, D, s))) return rc;
    e = GemmEpilogue();
    e.relu = 1;
    if ((rc = run_gemm(L.ffn1, M, e, ws.ah, c.ad_ff, s))) return rc;
    e = GemmEpilogue();
    e.out_f32 = 1;
    e.res1 = xo;
    if ((rc = run_gemm(L.ffn2, M, e, x, D, s))) return rc;
  }
  return launch_repack_rows(x, 0, out, out_tmax, out_bf16, cu, n_utts, M, D, s);
}

extern "C" int fa_encode_adapt(fa_plan* p, const float* feats, int32_t feats_tmax, const int32_t* cu, int32_t n_utts,
                               int32_t M, int32_t max_len, float* enc_out, int32_t enc_out_tmax, void* out,
                               int32_t out_tmax, int32_t out_bf16, fa_stream_t stream) {
  if (int r = fa_encoder_forward(p, feats, feats_tmax, cu, n_utts, M, max_len, enc_out, enc_out_tmax, stream)) return r;
  return fa_adaptor_forward(p, nullptr, 0, cu, n_utts, M, max_len, out, out_tmax, out_bf16, stream);
}
