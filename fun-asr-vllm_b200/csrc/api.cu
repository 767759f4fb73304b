// C-ABI orchestration: error state, device check, the encoder/adaptor plan (workspace carve-up +
// TMA descriptors built once) and the layer schedule.  No device allocation, no synchronisation.
#include <cmath>
#include <cstdlib>
#include <algorithm>
#include <cstring>
#include <functional>
#include <list>
#include <mutex>
#include <new>
#include <vector>

#include "common.cuh"

namespace fa {

static thread_local char g_err[1024] = "";

bool pdl_enabled() {  // FA_PDL=0 turns programmatic dependent launch off (A/B measurements)
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("FA_PDL");
    on = (e && atoi(e) == 0) ? 0 : 1;
  }
  return on == 1;
}

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int sm_count() {  // of the current device; cached per device (a process may drive several GPUs)
  static std::atomic<int> cache[64];
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev >= 0 && dev < 64) {
    const int c = cache[dev].load(std::memory_order_relaxed);
    if (c > 0) return c;
  }
  int n = 0;
  if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0) n = 148;
  if (dev >= 0 && dev < 64) cache[dev].store(n, std::memory_order_relaxed);
  return n;
}

// ---- profiling: event pairs around every launch, aggregated per kernel class on collect ----
bool g_prof_on = false;
namespace {
struct ProfRec { int kind; cudaEvent_t a, b; };
std::mutex g_prof_mu;  // the records are process-wide; launches may come from several host threads
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
  std::lock_guard<std::mutex> lk(g_prof_mu);
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
  std::lock_guard<std::mutex> lk(g_prof_mu);
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
// A plan is a "program": the exact kernel sequence of the encoder and of the adaptor, built once with every
// pointer, TMA descriptor and tile choice resolved; the forward calls just replay it on the caller's stream.
namespace {

size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

struct Workspace {
  // all row counts are cap = max_tokens rounded up to 128
  float* x;              // [cap, d_model]   residual stream (f32)
  float* x2;             // [cap, d_model]   ping-pong for the residual stream
  __nv_bfloat16* xn0;    // [cap, d_in]      LN'd input of layer 0
  __nv_bfloat16* xn;     // [cap, d_model]   LN output, or (folded LN) the raw bf16 copy of the residual stream
  float2* stats;         // [cap, 16]        folded LN: per-row partial (sum, sum of squares), <= 16 slots
  __nv_bfloat16* qkv;    // [cap, 3*d_model]
  __nv_bfloat16* ctx;    // [cap, 2*d_model] [attention context | bf16 FSMN memory]: A operand and bf16 residual of the out-projection
  __nv_bfloat16* h;      // [cap, max(d_ff, ad_ffn)]
  __nv_bfloat16* enc_bf; // [cap, d_model]   encoder output, bf16 (adaptor input)
  float* enc_f32;        // [cap, d_model]   encoder output, f32
  float* ax;             // [cap, ad_dim]
  __nv_bfloat16* axn;    // [cap, ad_dim]
  __nv_bfloat16* aqkv;   // [cap, 3*ad_dim]
  __nv_bfloat16* actx;   // [cap, ad_dim]
  __nv_bfloat16* ah;     // [cap, ad_ff]
  int32_t* cu;           // [cap + 8]        cu_seqlens at a FIXED address, padded to the utterance bucket (graph replay)
  int32_t* chain_flags;  // [2][chain_ints]  dependency counters of the GEMM chain, one set per half-batch stream
  int chain_ints;
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
  w.stats = reinterpret_cast<float2*>(take(n * 16 * sizeof(float2)));
  w.qkv = reinterpret_cast<__nv_bfloat16*>(take(n * 3 * c.d_model * 2));
  w.ctx = reinterpret_cast<__nv_bfloat16*>(take(n * 2 * c.d_model * 2));
  w.h = reinterpret_cast<__nv_bfloat16*>(take(n * hmax * 2));
  w.enc_bf = reinterpret_cast<__nv_bfloat16*>(take(n * c.d_model * 2));
  w.enc_f32 = reinterpret_cast<float*>(take(n * c.d_model * 4));
  w.ax = reinterpret_cast<float*>(take(n * c.ad_dim * 4));
  w.axn = reinterpret_cast<__nv_bfloat16*>(take(n * c.ad_dim * 2));
  w.aqkv = reinterpret_cast<__nv_bfloat16*>(take(n * 3 * c.ad_dim * 2));
  w.actx = reinterpret_cast<__nv_bfloat16*>(take(n * c.ad_dim * 2));
  w.ah = reinterpret_cast<__nv_bfloat16*>(take(n * c.ad_ff * 2));
  w.cu = reinterpret_cast<int32_t*>(take((n + 8) * 4));
  w.chain_ints = gemm_chain_flag_ints(FA_CHAIN_MAX_PROBS, cap);
  w.chain_flags = reinterpret_cast<int32_t*>(take(static_cast<size_t>(2) * w.chain_ints * 4));
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

// per-call arguments of a program: one SLAB = rows [row0, row0 + M) of every workspace matrix, holding the n_utts
// utterances cu[0..n_utts] (cu already points at the slab's first utterance and keeps absolute row numbers).
struct RunArgs {
  int M, n_utts, max_len, row0;
  const int32_t* cu;
  cudaStream_t s;
  int lane = 0;  // which half-batch stream of the dual-stream schedule (own set of chain counters)
};
using Step = std::function<int(const RunArgs&)>;

// One nn.Linear resolved against fixed workspace buffers: A map (activations), one B map per tile width, C map.
struct GemmOp {
  TensorMap2D ta, tc, tr1, tr2;
  TensorMap2D txb;  // folded-LN producer: the bf16 copy as 32 x 32 boxes (the chain kernel stores it by TMA)
  TensorMap2D tb[3];
  bool tb_ok[3] = {false, false, false};
  int N = 0, K = 0;
  GemmEpilogue epi;
  void* c = nullptr;
  int64_t ldc = 0;
  int prod_N = 0;  // folded-LN consumer: width of the GEMM that produced the statistics (its tile choice fixes ln_slots)
};

int make_gemm_op(GemmOp* op, int cap, const void* a, int lda, const void* w, int N, int K, const float* bias, void* c,
                 bool out_f32, bool relu, const float* res1, const float* res2, bool accumulate) {
  op->N = N;
  op->K = K;
  op->c = c;
  op->ldc = N;
  op->epi.bias = bias;
  op->epi.res1 = res1;
  op->epi.res2 = res2;
  op->epi.relu = relu ? 1 : 0;
  op->epi.out_f32 = out_f32 ? 1 : 0;
  op->epi.accumulate = accumulate ? 1 : 0;
  FA_REQUIRE(w != nullptr, "NULL weight pointer (N=%d K=%d)", N, K);
  if (int r = make_tensor_map_bf16(&op->ta, a, cap, K, lda, 128)) return r;
  if (int r = make_tensor_map(&op->tc, c, out_f32 ? 4 : 2, cap, N, N, 32)) return r;
  FA_REQUIRE(out_f32 || (!res1 && !res2), "residuals need the f32 output");
  if (res1) if (int r = make_tensor_map(&op->tr1, res1, 4, cap, N, N, 32)) return r;
  if (res2) if (int r = make_tensor_map(&op->tr2, res2, 4, cap, N, N, 32)) return r;
  for (int i = 0; i < 3; ++i) {
    const int bn = 64 << i;
    if (N % bn != 0) continue;
    if (int r = make_tensor_map_bf16(&op->tb[i], w, N, K, K, bn)) return r;
    op->tb_ok[i] = true;
  }
  return 0;
}

int run_gemm_op(const GemmOp& op, const RunArgs& a) {
  const int bn = gemm_block_n(a.M, op.N, (op.epi.res1 ? 1 : 0) + (op.epi.res2 ? 1 : 0));
  const int idx = bn == 64 ? 0 : (bn == 128 ? 1 : 2);
  FA_REQUIRE(bn != 0 && op.tb_ok[idx], "gemm N=%d unsupported", op.N);
  const bool pair = gemm_use_pair(a.M, bn) && op.tb_ok[1];
  GemmEpilogue epi = op.epi;
  if (epi.ln_mode == 2) epi.txb = &op.txb;
  if (epi.ln_mode == 1) epi.ln_slots = op.prod_N / 32;  // the producer leaves one partial per 32 columns of its output
  return launch_gemm(op.ta, op.tb[pair ? 1 : idx], op.tc, epi.res1 ? &op.tr1 : nullptr, epi.res2 ? &op.tr2 : nullptr, bn, a.M,
                     op.N, op.K, epi, op.c, op.ldc, a.row0, pair, a.s);
}

}  // namespace

// One instantiated CUDA graph of a program for one (rows, utterances, longest utterance) bucket.
struct GraphEntry {
  int prog, rows, utts, max_len;
  cudaGraphExec_t exec;
};

struct fa_plan {
  fa_model_cfg cfg;
  fa_model_w w;
  int cap;
  int n_streams;        // 1: everything on the caller's stream; 2: two half-batches, the second on `aux` (fork/join by events)
  int graph_max_rows;   // batches of up to this many rows replay a captured CUDA graph of the program (0: never)
  int chain = 0;        // 1: the row-local GEMMs between two attention kernels run as ONE launch (gemm_chain.cu); measured slower
                        // than one launch per GEMM so far (profiles/r02/chain_r2ae.md), so it is opt-in
  cudaStream_t aux = nullptr;
  cudaStream_t cap_stream = nullptr;  // only ever used in capture mode: nothing executes on it
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr;
  Workspace ws;
  std::vector<Step> enc_prog, ad_prog;
  std::list<GraphEntry> graphs;  // most recently used first
  ~fa_plan() {
    for (GraphEntry& g : graphs) cudaGraphExecDestroy(g.exec);
    if (aux) cudaStreamDestroy(aux);
    if (cap_stream) cudaStreamDestroy(cap_stream);
    if (ev_fork) cudaEventDestroy(ev_fork);
    if (ev_join) cudaEventDestroy(ev_join);
  }
};

namespace {

struct Builder {
  fa_plan* p;
  int rc = 0;
  void gemm(std::vector<Step>& prog, const void* a, int lda, const void* w, int N, int K, const float* bias, void* c,
            bool out_f32, bool relu = false, const float* res1 = nullptr, const float* res2 = nullptr, bool accumulate = false,
            int ln_mode = 0, const float* colsum = nullptr, int ln_dim = 0, int prod_N = 0, const __nv_bfloat16* res_bf16 = nullptr,
            int ld_res_bf16 = 0) {
    if (rc) return;
    GemmOp op;
    rc = make_gemm_op(&op, p->cap, a, lda, w, N, K, bias, c, out_f32, relu, res1, res2, accumulate);
    op.epi.res_bf16 = res_bf16;
    op.epi.ld_res_bf16 = ld_res_bf16;
    op.epi.ln_mode = ln_mode;
    if (ln_mode != 0) op.epi.ln_stats = p->ws.stats;
    if (ln_mode == 1) { op.epi.colsum = colsum; op.epi.ln_dim = ln_dim; op.prod_N = prod_N; }
    if (ln_mode == 2) {
      op.epi.xb = p->ws.xn;
      op.epi.ld_xb = N;
      if (!rc) rc = make_tensor_map_w(&op.txb, p->ws.xn, 2, p->cap, N, N, 32, 64);
    }
    if (!rc) prog.push_back([op](const RunArgs& r) { return run_gemm_op(op, r); });
  }
  // Layer i's out-projection, ffn1, ffn2 and layer i+1's qkv as one step: the chain launch when the plan's switch is on,
  // the four GEMMs one after the other otherwise (same arithmetic, bit-identical results).
  void chain(std::vector<Step>& prog, const GemmOp (&ops)[4]) {
    if (rc) return;
    fa_plan* pl = p;
    struct Pack { GemmOp op[4]; };
    Pack pk;
    for (int i = 0; i < 4; ++i) pk.op[i] = ops[i];
    prog.push_back([pk, pl](const RunArgs& r) {
      if (!pl->chain) {
        for (int i = 0; i < 4; ++i)
          if (int e = run_gemm_op(pk.op[i], r)) return e;
        return 0;
      }
      ChainDesc d;
      d.n_probs = 4;
      for (int i = 0; i < 4; ++i) {
        const GemmOp& o = pk.op[i];
        ChainProb& c = d.p[i];
        c.ta = &o.ta; c.tb = &o.tb[1]; c.tc = &o.tc; c.txb = &o.txb;
        c.N = o.N; c.K = o.K;
        c.kind = o.epi.ln_mode == 2 ? 0 : 1;
        c.relu = o.epi.relu;
        c.bias = o.epi.bias; c.colsum = o.epi.colsum; c.stats = o.epi.ln_stats;
        c.xb = o.epi.xb; c.ld_xb = o.epi.ld_xb;
        c.res_bf16 = o.epi.res_bf16; c.ld_res_bf16 = o.epi.ld_res_bf16;
        c.c = o.c;
      }
      return launch_gemm_chain(d, r.M, r.row0, pl->ws.chain_flags + r.lane * pl->ws.chain_ints, r.s);
    });
  }
  int make(GemmOp* op, const void* a, int lda, const void* w, int N, int K, const float* bias, void* c, bool out_f32, bool relu,
           const float* res1, int ln_mode, const float* colsum, int ln_dim, int prod_N, const __nv_bfloat16* res_bf16, int ld_res_bf16) {
    if (int e = make_gemm_op(op, p->cap, a, lda, w, N, K, bias, c, out_f32, relu, res1, nullptr, false)) return e;
    op->epi.res_bf16 = res_bf16;
    op->epi.ld_res_bf16 = ld_res_bf16;
    op->epi.ln_mode = ln_mode;
    op->epi.ln_stats = p->ws.stats;
    if (ln_mode == 1) { op->epi.colsum = colsum; op->epi.ln_dim = ln_dim; op->prod_N = prod_N; }
    if (ln_mode == 2) {
      op->epi.xb = p->ws.xn;
      op->epi.ld_xb = N;
      if (int e = make_tensor_map_w(&op->txb, p->ws.xn, 2, p->cap, N, N, 32, 64)) return e;
    }
    return 0;
  }
  void ln(std::vector<Step>& prog, const float* x, const float* g, const float* b, const float* g2, const float* b2,
          __nv_bfloat16* y_bf16, float* y_f32, int D) {
    prog.push_back([=](const RunArgs& r) {
      const int64_t o = static_cast<int64_t>(r.row0) * D;
      return launch_layernorm(x + o, g, b, g2, b2, y_bf16 ? y_bf16 + o : nullptr, y_f32 ? y_f32 + o : nullptr, r.M, D, r.s);
    });
  }
  // fsmn_w != nullptr: the FSMN memory is produced by the attention kernel itself (bf16, beside the context: column D + c)
  void attention(std::vector<Step>& prog, const __nv_bfloat16* qkv, __nv_bfloat16* ctx, int H, int D, int ld_ctx,
                 const float* fsmn_w = nullptr, int fsmn_left = 0, bool zero_chain = false, float* fsmn_acc = nullptr) {
    if (rc) return;
    // one tensor map over the whole qkv buffer, encoded once per plan: utterances are addressed by their ABSOLUTE packed rows
    // (base 0).  Key / value rows past the end of a batch are workspace rows, which hold finite values (the caller hands in a
    // zero-filled workspace and every kernel keeps them finite), so 0 * garbage stays 0 in P V.
    TensorMap2D tm;
    const int cap = p->cap;
    rc = make_tensor_map_bf16(&tm, qkv, cap, 3 * D, 3 * D, attention_kv_box());
    if (rc) return;
    TensorMap2D tmc;  // the context rows as 32 x 64 boxes for the epilogue's TMA stores
    rc = make_tensor_map(&tmc, ctx, 2, cap, D, ld_ctx, 32);
    if (rc) return;
    fa_plan* pl = p;
    prog.push_back([tm, tmc, qkv, ctx, H, D, ld_ctx, cap, fsmn_w, fsmn_left, zero_chain, pl, fsmn_acc](const RunArgs& r) {
      // the chain that follows this attention wants its dependency counters zero: this kernel clears them on its way
      const bool z = zero_chain && pl->chain;
      const int nz = gemm_chain_flag_ints(FA_CHAIN_MAX_PROBS, r.M);
      return launch_attention(&tm, qkv, 3 * D, r.cu, r.n_utts, 0, r.max_len, H, ctx, ld_ctx, cap, fsmn_w, fsmn_left,
                              fsmn_w ? ctx + D : nullptr, ld_ctx, r.s, z ? pl->ws.chain_flags + r.lane * pl->ws.chain_ints : nullptr, z ? nz : 0,
                              fsmn_acc, D, &tmc);
    });
  }
};

int build_programs(fa_plan* p) {
  const fa_model_cfg& c = p->cfg;
  const fa_model_w& w = p->w;
  Workspace& ws = p->ws;
  Builder B{p};
  const int d = c.d_model;
  const int nl = c.n_blocks + c.n_tp_blocks;
  const int left = (c.fsmn_kernel - 1) / 2 + (c.sanm_shift > 0 ? c.sanm_shift : 0);  // VM:175-179
  // ---------------- encoder: the embed+LN input stage is issued by run_encoder (it needs the caller's feats)
  float* x = ws.x;    // current residual stream
  float* xo = ws.x2;  // next
  std::vector<Step>& E = p->enc_prog;
  bool qkv_in_chain = false;  // this layer's qkv GEMM is the last problem of the previous layer's chain step
  for (int i = 0; i < nl; ++i) {
    const fa_sanm_layer_w& L = w.layers[i];
    const int din = (i == 0) ? c.d_in : d;
    const void* a_in = (i == 0) ? static_cast<const void*>(ws.xn0) : static_cast<const void*>(ws.xn);
    // Folded LayerNorm (fa_sanm_layer_w.qkv_colsum / ffn1_colsum): the GEMM that updates the residual stream also emits
    // its bf16 copy and row statistics, the next GEMM normalises in its epilogue -> no LN kernel, x is not re-read.
    const bool fold_qkv = L.qkv_colsum != nullptr, fold_ffn1 = L.ffn1_colsum != nullptr;
    const bool fold_next = (i + 1 < nl) && w.layers[i + 1].qkv_colsum != nullptr;
    if (fold_qkv && (i == 0 || i == c.n_blocks)) { set_error("plan_create: layer %d takes its input from a LayerNorm kernel; its qkv weights must not be LN-folded", i); return 2; }
    if (fold_ffn1 && i == 0) { set_error("plan_create: layer 0 has no residual out-projection; its ffn1 weights must not be LN-folded"); return 2; }
    // Chain step (gemm_chain.cu): out-projection, ffn1, ffn2 of this layer and the qkv GEMM of the next one are row-local and
    // LN-folded into each other -> one launch.  Needs the folded forms and 256-wide tiles everywhere.
    const bool chain_ok = fold_ffn1 && fold_next && d % 256 == 0 && d <= 512 && c.d_ff % 256 == 0 &&
                          2 * (c.d_ff + 3 * d) * 4 <= 28 * 1024;
    // q|k|v = LN1(x) W^T + b                                              VM:206
    if (qkv_in_chain) {}
    else if (fold_qkv) B.gemm(E, ws.xn, d, L.qkv_w, 3 * d, d, L.qkv_b, ws.qkv, false, false, nullptr, nullptr, false, 1, L.qkv_colsum, d, d);
    else B.gemm(E, a_in, din, L.qkv_w, 3 * d, din, L.qkv_b, ws.qkv, false);
    qkv_in_chain = false;
    // FSMN memory on v                                                     VM:249,181-201
    // The memory is kept in bf16 beside the attention context (row = [ctx | mem]); the out-projection's epilogue adds it to
    // the f32 result (GemmEpilogue::res_bf16).  With the model's 11-tap kernel the FSMN runs inside the attention kernel, on
    // the V blocks that are in shared memory for P V anyway: no FSMN launch, V is read once.
    __nv_bfloat16* mem = ws.ctx + d;
    const bool fuse_fsmn = attention_can_fuse_fsmn(c.fsmn_kernel, 2 * d, mem);
    if (!fuse_fsmn) {
      const __nv_bfloat16* qkv = ws.qkv;
      const float* fw = L.fsmn_w;
      __nv_bfloat16* cm = ws.ctx;
      const int K = c.fsmn_kernel;
      E.push_back([=](const RunArgs& r) {
        return launch_fsmn(qkv + static_cast<int64_t>(r.row0) * 3 * d, 3 * d, 2 * d, fw, r.cu, r.n_utts, r.row0,
                           cm + static_cast<int64_t>(r.row0) * 2 * d + d, 1, 2 * d, r.M, d, K, left, r.s);
      });
    }
    // masked MHA                                                           VM:250-252,222-236
    // Layers with a residual out-projection (all but layer 0): the fused FSMN adds its memory straight into the f32 residual
    // stream (x += mem in L2; x then += ctx Wo^T + b), so the out-projection has no memory operand to read.
    // (measured against the bf16 memory + residual operand form inside the step: 43.9 / 44.0 against 44.2 / 44.2 ms at cfg5)
    const bool mem_into_x = fuse_fsmn && i != 0;
    if (mem_into_x) mem = nullptr;
    B.attention(E, ws.qkv, ws.ctx, c.n_heads, d, 2 * d, fuse_fsmn ? L.fsmn_w : nullptr, left, chain_ok, mem_into_x ? x : nullptr);
    if (chain_ok) {
      const fa_sanm_layer_w& Ln = w.layers[i + 1];
      GemmOp ops[4];
      if (!B.rc) B.rc = B.make(&ops[0], ws.ctx, 2 * d, L.out_w, d, d, L.out_b, x, true, false, x, 2, nullptr, 0, 0, mem, 2 * d);
      if (!B.rc) B.rc = B.make(&ops[1], ws.xn, d, L.ffn1_w, c.d_ff, d, L.ffn1_b, ws.h, false, true, nullptr, 1, L.ffn1_colsum, d, d, nullptr, 0);
      if (!B.rc) B.rc = B.make(&ops[2], ws.h, c.d_ff, L.ffn2_w, d, c.d_ff, L.ffn2_b, x, true, false, x, 2, nullptr, 0, 0, nullptr, 0);
      if (!B.rc) B.rc = B.make(&ops[3], ws.xn, d, Ln.qkv_w, 3 * d, d, Ln.qkv_b, ws.qkv, false, false, nullptr, 1, Ln.qkv_colsum, d, d, nullptr, 0);
      B.chain(E, ops);
      qkv_in_chain = true;
      continue;  // (a chained layer is neither the last main block nor the last block: nothing else follows it)
    }
    // x (+)= ctx Wo^T + b + mem   (no residual when in_size != size: layer 0 overwrites x)   VM:238,253,123-136
    // Folded-LN form: the epilogue reads the old x through its TMA ring and writes x_new, its bf16 copy and the row
    // statistics.  Otherwise the residual stream is updated IN PLACE by a TMA reduce-add of acc + bias + mem.
    if (fold_ffn1) B.gemm(E, ws.ctx, 2 * d, L.out_w, d, d, L.out_b, x, true, false, x, nullptr, false, 2, nullptr, 0, 0, mem, 2 * d);
    else B.gemm(E, ws.ctx, 2 * d, L.out_w, d, d, L.out_b, x, true, false, nullptr, nullptr, /*accumulate=*/i != 0, 0, nullptr, 0, 0, mem, 2 * d);
    // LN2, FFN                                                              VM:138-140,412-431
    if (fold_ffn1) {
      B.gemm(E, ws.xn, d, L.ffn1_w, c.d_ff, d, L.ffn1_b, ws.h, false, true, nullptr, nullptr, false, 1, L.ffn1_colsum, d, d);
    } else {
      B.ln(E, x, L.ln2_g, L.ln2_b, nullptr, nullptr, ws.xn, nullptr, d);
      B.gemm(E, ws.xn, d, L.ffn1_w, c.d_ff, d, L.ffn1_b, ws.h, false, true);
    }
    if (fold_next) B.gemm(E, ws.h, c.d_ff, L.ffn2_w, d, c.d_ff, L.ffn2_b, x, true, false, x, nullptr, false, 2);
    else B.gemm(E, ws.h, c.d_ff, L.ffn2_w, d, c.d_ff, L.ffn2_b, x, true, false, nullptr, nullptr, /*accumulate=*/true);
    // next layer's norm1, with after_norm / tp_norm folded in                VM:400,408
    const bool last_main = (i == c.n_blocks - 1);
    const bool last = (i == nl - 1);
    if (last) {
      if (c.n_tp_blocks == 0) {  // after_norm then tp_norm back to back (VM:400,408 with an empty tp_encoders list)
        B.ln(E, x, w.after_norm_g, w.after_norm_b, nullptr, nullptr, nullptr, xo, d);
        B.ln(E, xo, w.tp_norm_g, w.tp_norm_b, nullptr, nullptr, ws.enc_bf, ws.enc_f32, d);
      } else {
        B.ln(E, x, w.tp_norm_g, w.tp_norm_b, nullptr, nullptr, ws.enc_bf, ws.enc_f32, d);
      }
    } else if (last_main) {
      // x <- after_norm(x) becomes the residual stream of the tp blocks; xn <- norm1_next(x)
      B.ln(E, x, w.after_norm_g, w.after_norm_b, w.layers[i + 1].ln1_g, w.layers[i + 1].ln1_b, ws.xn, xo, d);
      float* t = x; x = xo; xo = t;
    } else if (!fold_next) {
      B.ln(E, x, w.layers[i + 1].ln1_g, w.layers[i + 1].ln1_b, nullptr, nullptr, ws.xn, nullptr, d);
    }
  }
  // ---------------- adaptor (input: ws.enc_bf)
  if (w.lin1_w) {
    std::vector<Step>& A = p->ad_prog;
    const int D = c.ad_dim;
    float* ax = ws.ax;
    // linear1 + ReLU, linear2                                                 VM:585-587
    B.gemm(A, ws.enc_bf, c.d_model, w.lin1_w, c.ad_ffn, c.d_model, w.lin1_b, ws.h, false, true);
    B.gemm(A, ws.h, c.ad_ffn, w.lin2_w, D, c.ad_ffn, w.lin2_b, ax, true);
    for (int i = 0; i < c.ad_layers; ++i) {
      if (!w.ad_layers) { set_error("plan_create: NULL adaptor layers"); return 2; }
      const fa_adaptor_layer_w& L = w.ad_layers[i];
      // pre-LN MHA block                                                       VM:434-455
      B.ln(A, ax, L.ln1_g, L.ln1_b, nullptr, nullptr, ws.axn, nullptr, D);
      B.gemm(A, ws.axn, D, L.qkv_w, 3 * D, D, L.qkv_b, ws.aqkv, false);
      B.attention(A, ws.aqkv, ws.actx, c.ad_heads, D, D);
      B.gemm(A, ws.actx, D, L.out_w, D, D, L.out_b, ax, true, false, nullptr, nullptr, /*accumulate=*/true);
      B.ln(A, ax, L.ln2_g, L.ln2_b, nullptr, nullptr, ws.axn, nullptr, D);
      B.gemm(A, ws.axn, D, L.ffn1_w, c.ad_ff, D, L.ffn1_b, ws.ah, false, true);
      B.gemm(A, ws.ah, c.ad_ff, L.ffn2_w, D, c.ad_ff, L.ffn2_b, ax, true, false, nullptr, nullptr, /*accumulate=*/true);
    }
  }
  return B.rc;
}

int default_streams() {
  // Measured on B200 (profiles/r01_d_dual_stream.md): cfg3 (16 k rows) 16.9 -> 15.2 ms.  On by default; FA_STREAMS=1 or
  // fa_plan_set_streams(plan, 1) turn it off.
  if (const char* e = getenv("FA_STREAMS")) return atoi(e) >= 2 ? 2 : 1;
  return 2;
}

}  // namespace

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
  p->n_streams = default_streams();
  p->graph_max_rows = FA_DEFAULT_GRAPH_ROWS;
  if (const char* e = getenv("FA_CHAIN")) p->chain = atoi(e) != 0;  // A/B measurements
  p->ws = ws;
  if (int rc = build_programs(p)) {
    delete p;
    return rc;
  }
  *plan_out = p;
  return 0;
}

extern "C" void fa_plan_destroy(fa_plan* plan) { delete plan; }

extern "C" int fa_plan_set_streams(fa_plan* plan, int32_t n_streams) {
  FA_REQUIRE(plan != nullptr && (n_streams == 1 || n_streams == 2), "set_streams: 1 or 2 streams");
  plan->n_streams = n_streams;
  return 0;
}

extern "C" int fa_plan_set_chain(fa_plan* plan, int32_t on) {
  FA_REQUIRE(plan != nullptr, "set_chain: NULL plan");
  if ((on != 0) != (plan->chain != 0)) {  // captured graphs hold the other schedule
    for (GraphEntry& g : plan->graphs) cudaGraphExecDestroy(g.exec);
    plan->graphs.clear();
  }
  plan->chain = on != 0;
  return 0;
}

extern "C" int fa_plan_set_graph_rows(fa_plan* plan, int32_t max_rows) {
  FA_REQUIRE(plan != nullptr && max_rows >= 0, "set_graph_rows: bad arguments");
  plan->graph_max_rows = max_rows;
  return 0;
}

namespace {

int check_batch(const fa_plan* p, const int32_t* cu, int n_utts, int n_tokens, int max_len) {
  FA_REQUIRE(p != nullptr, "NULL plan");
  FA_REQUIRE(cu != nullptr && n_utts > 0, "empty batch");
  FA_REQUIRE(n_tokens > 0 && n_tokens <= p->cap, "n_tokens %d exceeds the plan capacity %d", n_tokens, p->cap);
  FA_REQUIRE(max_len > 0 && max_len <= n_tokens, "max_len %d out of range", max_len);
  FA_REQUIRE(n_utts <= n_tokens, "%d utterances in %d rows", n_utts, n_tokens);
  return 0;
}

int check_tmax(int tmax, int max_len, const char* what) {
  FA_REQUIRE(tmax == 0 || tmax >= max_len, "%s: padded row count %d is smaller than the longest utterance %d", what, tmax, max_len);
  return 0;
}

// A run of consecutive utterances [b0, b1) = packed rows [row0, row0 + M).
struct Slab { int b0, b1, row0, M, max_len; };

RunArgs args_of(const Slab& sl, const int32_t* cu, cudaStream_t s) {
  return RunArgs{sl.M, sl.b1 - sl.b0, sl.max_len, sl.row0, cu + sl.b0, s};
}

// input stage: xs*sqrt(d) + PE, then norm1 of encoders0 (VM:388-390, :121)
int encoder_pre(fa_plan* p, const Slab& sl, const RunArgs& a, const float* feats, int feats_tmax) {
  const fa_model_cfg& c = p->cfg;
  const float* f = feats + (feats_tmax > 0 ? static_cast<int64_t>(sl.b0) * feats_tmax : static_cast<int64_t>(sl.row0)) * c.d_in;
  return launch_embed_layernorm(f, feats_tmax, p->w.pe_inv_timescales, a.cu, a.n_utts, a.row0, sqrtf(static_cast<float>(c.d_model)),
                                p->w.layers[0].ln1_g, p->w.layers[0].ln1_b, p->ws.xn0 + static_cast<int64_t>(a.row0) * c.d_in,
                                a.M, c.d_in, a.s);
}

int encoder_post(fa_plan* p, const Slab& sl, const RunArgs& a, float* enc_out, int enc_out_tmax) {
  const fa_model_cfg& c = p->cfg;
  if (!enc_out) return 0;
  float* dst = enc_out + (enc_out_tmax > 0 ? static_cast<int64_t>(sl.b0) * enc_out_tmax : static_cast<int64_t>(sl.row0)) * c.d_model;
  return launch_repack_rows(p->ws.enc_f32 + static_cast<int64_t>(a.row0) * c.d_model, 0, dst, enc_out_tmax, 0, a.cu, a.n_utts,
                            a.row0, a.M, c.d_model, a.s);
}

int adaptor_pre(fa_plan* p, const Slab& sl, const RunArgs& a, const float* enc, int enc_tmax) {
  const fa_model_cfg& c = p->cfg;
  if (!enc) return 0;
  const float* src = enc + (enc_tmax > 0 ? static_cast<int64_t>(sl.b0) * enc_tmax : static_cast<int64_t>(sl.row0)) * c.d_model;
  return launch_repack_rows(src, enc_tmax, p->ws.enc_bf + static_cast<int64_t>(a.row0) * c.d_model, 0, 1, a.cu, a.n_utts, a.row0,
                            a.M, c.d_model, a.s);
}

int adaptor_post(fa_plan* p, const Slab& sl, const RunArgs& a, void* out, int out_tmax, int out_bf16) {
  const fa_model_cfg& c = p->cfg;
  const int64_t dst_row = out_tmax > 0 ? static_cast<int64_t>(sl.b0) * out_tmax : static_cast<int64_t>(sl.row0);
  void* dst = static_cast<unsigned char*>(out) + dst_row * c.ad_dim * (out_bf16 ? 2 : 4);
  return launch_repack_rows(p->ws.ax + static_cast<int64_t>(a.row0) * c.ad_dim, 0, dst, out_tmax, out_bf16, a.cu, a.n_utts, a.row0,
                            a.M, c.ad_dim, a.s);
}

int run_prog(const std::vector<Step>& prog, const RunArgs& a) {
  for (const Step& st : prog)
    if (int rc = st(a)) return rc;
  return 0;
}

// ---- CUDA-graph replay ------------------------------------------------------------------------------------------------
// A program is ~440 dependent launches; at the reference's own batch sizes (16 x 10 s = 2 672 rows, infer.py:415) a launch
// runs for 5-10 us, about what the host needs to issue it, so the step is launch-bound.  The program of a plan is static
// (fixed workspace pointers, prebuilt tensor maps), and everything that varies per batch is either read from cu_seqlens ON
// THE DEVICE (utterance boundaries, the true row count) or only sizes a grid.  So the program is captured once per BUCKET
// (rows rounded up to ~1/8 octave, utterances to a multiple of 8, longest utterance to a multiple of 128 rows) with
// cu_seqlens at a fixed workspace address, and replayed with one cudaGraphLaunch.  Per call a tiny kernel copies the caller's
// cu_seqlens there, padded with the total row count: padded utterances are empty (their CTAs exit), rows between the true
// count and the bucket are computed on finite workspace contents and never read.  The programmatic-dependent-launch edges
// survive capture.  (The reference's vLLM side captures graphs the same way: infer_kaldidata.py:655-656.)
int bucket_rows(int M, int cap) {
  int gran = 128;
  while (gran < 4096 && gran * 16 <= M) gran *= 2;  // 128 below 2048 rows, then ~1/8 .. 1/16 of M
  const int b = (M + gran - 1) / gran * gran;
  return b < cap ? b : cap;
}
int bucket_utts(int n, int rows) {
  const int g = n <= 64 ? 8 : 32;
  const int b = (n + g - 1) / g * g;
  return b < rows ? b : rows;
}

int ensure_capture_stream(fa_plan* p) {
  if (p->cap_stream) return 0;
  FA_CHECK_CUDA(cudaStreamCreateWithFlags(&p->cap_stream, cudaStreamNonBlocking));
  return 0;
}

constexpr size_t MAX_GRAPHS = 48;

int graph_for(fa_plan* p, int prog_id, const std::vector<Step>& prog, int rows, int utts, int max_len, cudaGraphExec_t* out) {
  for (auto it = p->graphs.begin(); it != p->graphs.end(); ++it) {
    if (it->prog == prog_id && it->rows == rows && it->utts == utts && it->max_len == max_len) {
      p->graphs.splice(p->graphs.begin(), p->graphs, it);
      *out = p->graphs.front().exec;
      return 0;
    }
  }
  if (int rc = ensure_capture_stream(p)) return rc;
  FA_CHECK_CUDA(cudaStreamBeginCapture(p->cap_stream, cudaStreamCaptureModeThreadLocal));
  const RunArgs a{rows, utts, max_len, 0, p->ws.cu, p->cap_stream};
  const int rc = run_prog(prog, a);
  cudaGraph_t g = nullptr;
  const cudaError_t e = cudaStreamEndCapture(p->cap_stream, &g);
  if (rc) {
    if (g) cudaGraphDestroy(g);
    return rc;
  }
  if (e != cudaSuccess) {
    set_error("cudaStreamEndCapture failed: %s", cudaGetErrorString(e));
    return 1;
  }
  cudaGraphExec_t exec = nullptr;
  const cudaError_t ei = cudaGraphInstantiate(&exec, g, 0);
  cudaGraphDestroy(g);
  if (ei != cudaSuccess) {
    set_error("cudaGraphInstantiate failed: %s", cudaGetErrorString(ei));
    return 1;
  }
  if (p->graphs.size() >= MAX_GRAPHS) {  // the in-flight replays of an evicted graph keep it alive until they finish
    cudaGraphExecDestroy(p->graphs.back().exec);
    p->graphs.pop_back();
  }
  p->graphs.push_front(GraphEntry{prog_id, rows, utts, max_len, exec});
  *out = exec;
  return 0;
}

// true when this call replays graphs; then a.cu of the returned bucket args points at the plan's padded copy
bool use_graph(const fa_plan* p, int M) { return p->graph_max_rows > 0 && M <= p->graph_max_rows && !g_prof_on; }

int run_graphed(fa_plan* p, int n_utts, int M, int max_len, const int32_t* cu, cudaStream_t s, const std::function<int(const RunArgs&)>& pre,
                const std::vector<std::pair<int, const std::vector<Step>*>>& progs, const std::function<int(const RunArgs&)>& mid,
                const std::function<int(const RunArgs&)>& post) {
  const int rows = bucket_rows(M, p->cap), utts = bucket_utts(n_utts, rows), ml = std::min(rows, (max_len + 127) / 128 * 128);
  const RunArgs a{M, n_utts, max_len, 0, cu, s};
  if (int rc = launch_pad_cu(cu, n_utts, p->ws.cu, utts, s)) return rc;
  if (int rc = pre(a)) return rc;
  for (size_t i = 0; i < progs.size(); ++i) {
    if (i > 0) if (int rc = mid(a)) return rc;
    cudaGraphExec_t exec = nullptr;
    if (int rc = graph_for(p, progs[i].first, *progs[i].second, rows, utts, ml, &exec)) return rc;
    FA_CHECK_CUDA(cudaGraphLaunch(exec, s));
  }
  return post(a);
}

// ---- dual-stream schedule -------------------------------------------------------------------------------------------
// The batch is cut on an utterance boundary into two halves of about equal cost; half 0 runs on the caller's stream, half 1
// on the plan's auxiliary stream (forked and joined with events, so for the caller everything is still "enqueued on
// `stream`").  Utterances are independent, so the result is bit-identical to the single-stream schedule.  The point: the
// HBM-bound kernels of one half (LayerNorm, FSMN, attention softmax phases) co-reside on the SMs with the tensor-bound GEMM
// of the other half, and the ramp-up / drain of every dependent launch is covered by the other stream's work.
bool split_two(const fa_plan* p, const int32_t* cu_host, int n_utts, int M, Slab (&out)[2]) {
  if (p->n_streams < 2 || cu_host == nullptr || n_utts < 2 || M < 2048) return false;
  if (cu_host[0] != 0 || cu_host[n_utts] != M) return false;
  // cost ~ 0.4657 T + 1.5155e-4 T^2 GFLOP per utterance
  double total = 0.0;
  for (int b = 0; b < n_utts; ++b) {
    const double T = cu_host[b + 1] - cu_host[b];
    if (T <= 0) return false;
    total += 0.4657152 * T + 1.51552e-4 * T * T;
  }
  double acc = 0.0;
  int cut = 1;
  for (int b = 0; b < n_utts - 1; ++b) {
    const double T = cu_host[b + 1] - cu_host[b];
    const double c = 0.4657152 * T + 1.51552e-4 * T * T;
    if (acc + 0.5 * c >= 0.5 * total) break;
    acc += c;
    cut = b + 1;
  }
  auto mk = [&](int b0, int b1) {
    int ml = 0;
    for (int u = b0; u < b1; ++u) ml = std::max(ml, cu_host[u + 1] - cu_host[u]);
    return Slab{b0, b1, cu_host[b0], cu_host[b1] - cu_host[b0], ml};
  };
  out[0] = mk(0, cut);
  out[1] = mk(cut, n_utts);
  return true;
}

int ensure_streams(fa_plan* p) {
  if (p->aux) return 0;
  FA_CHECK_CUDA(cudaStreamCreateWithFlags(&p->aux, cudaStreamNonBlocking));
  FA_CHECK_CUDA(cudaEventCreateWithFlags(&p->ev_fork, cudaEventDisableTiming));
  FA_CHECK_CUDA(cudaEventCreateWithFlags(&p->ev_join, cudaEventDisableTiming));
  return 0;
}

// Issue pre, the programs (mid between them) and post for both halves, alternating step by step.
int run_dual(fa_plan* p, const Slab (&sl)[2], const int32_t* cu, cudaStream_t s, const std::function<int(int, const RunArgs&)>& pre,
             const std::vector<const std::vector<Step>*>& progs, const std::function<int(int, const RunArgs&)>& mid,
             const std::function<int(int, const RunArgs&)>& post) {
  if (int rc = ensure_streams(p)) return rc;
  FA_CHECK_CUDA(cudaEventRecord(p->ev_fork, s));
  FA_CHECK_CUDA(cudaStreamWaitEvent(p->aux, p->ev_fork, 0));
  RunArgs a[2] = {args_of(sl[0], cu, s), args_of(sl[1], cu, p->aux)};
  a[1].lane = 1;
  int rc = 0;
  auto both = [&](const std::function<int(int)>& f) {
    for (int k = 0; k < 2 && !rc; ++k) rc = f(k);
  };
  both([&](int k) { return pre(k, a[k]); });
  for (size_t pi = 0; pi < progs.size() && !rc; ++pi) {
    if (pi > 0) both([&](int k) { return mid(k, a[k]); });
    for (const Step& st : *progs[pi]) {
      if (rc) break;
      both([&](int k) { return st(a[k]); });
    }
  }
  if (!rc) both([&](int k) { return post(k, a[k]); });
  // join ALSO on the error path: work already enqueued on `aux` must stay ordered before whatever the caller does next on
  // `s` (it may free or reuse the buffers once this call has returned an error)
  const cudaError_t e1 = cudaEventRecord(p->ev_join, p->aux);
  const cudaError_t e2 = e1 == cudaSuccess ? cudaStreamWaitEvent(s, p->ev_join, 0) : e1;
  if (rc) return rc;
  FA_CHECK_CUDA(e2);
  return 0;
}

const Slab whole(int n_utts, int M, int max_len) { return Slab{0, n_utts, 0, M, max_len}; }

}  // namespace

extern "C" int fa_encoder_forward(fa_plan* p, const float* feats, int32_t feats_tmax, const int32_t* cu,
                                  const int32_t* cu_host, int32_t n_utts, int32_t M, int32_t max_len, float* enc_out,
                                  int32_t enc_out_tmax, fa_stream_t stream) {
  if (int r = check_batch(p, cu, n_utts, M, max_len)) return r;
  FA_REQUIRE(feats != nullptr, "NULL feats");
  FA_REQUIRE(p->cfg.n_blocks > 0, "plan was created without encoder weights");
  if (int r = check_tmax(feats_tmax, max_len, "feats")) return r;
  if (int r = check_tmax(enc_out_tmax, max_len, "enc_out")) return r;
  const cudaStream_t s = static_cast<cudaStream_t>(stream);
  const Slab all = whole(n_utts, M, max_len);
  if (use_graph(p, M))
    return run_graphed(p, n_utts, M, max_len, cu, s, [&](const RunArgs& a) { return encoder_pre(p, all, a, feats, feats_tmax); },
                       {{0, &p->enc_prog}}, [](const RunArgs&) { return 0; },
                       [&](const RunArgs& a) { return encoder_post(p, all, a, enc_out, enc_out_tmax); });
  Slab two[2];
  if (split_two(p, cu_host, n_utts, M, two)) {
    const auto none = [](int, const RunArgs&) { return 0; };
    return run_dual(p, two, cu, s, [&](int k, const RunArgs& a) { return encoder_pre(p, two[k], a, feats, feats_tmax); },
                    {&p->enc_prog}, none, [&](int k, const RunArgs& a) { return encoder_post(p, two[k], a, enc_out, enc_out_tmax); });
  }
  const RunArgs a = args_of(all, cu, s);
  if (int rc = encoder_pre(p, all, a, feats, feats_tmax)) return rc;
  if (int rc = run_prog(p->enc_prog, a)) return rc;
  return encoder_post(p, all, a, enc_out, enc_out_tmax);
}

extern "C" int fa_adaptor_forward(fa_plan* p, const float* enc, int32_t enc_tmax, const int32_t* cu, const int32_t* cu_host,
                                  int32_t n_utts, int32_t M, int32_t max_len, void* out, int32_t out_tmax, int32_t out_bf16,
                                  fa_stream_t stream) {
  if (int r = check_batch(p, cu, n_utts, M, max_len)) return r;
  if (int r = check_tmax(enc_tmax, max_len, "enc")) return r;
  if (int r = check_tmax(out_tmax, max_len, "out")) return r;
  FA_REQUIRE(out != nullptr, "NULL out");
  FA_REQUIRE(p->w.lin1_w != nullptr, "plan was created without adaptor weights");
  const cudaStream_t s = static_cast<cudaStream_t>(stream);
  const Slab all = whole(n_utts, M, max_len);
  if (use_graph(p, M))
    return run_graphed(p, n_utts, M, max_len, cu, s, [&](const RunArgs& a) { return adaptor_pre(p, all, a, enc, enc_tmax); },
                       {{1, &p->ad_prog}}, [](const RunArgs&) { return 0; },
                       [&](const RunArgs& a) { return adaptor_post(p, all, a, out, out_tmax, out_bf16); });
  Slab two[2];
  if (split_two(p, cu_host, n_utts, M, two)) {
    const auto none = [](int, const RunArgs&) { return 0; };
    return run_dual(p, two, cu, s, [&](int k, const RunArgs& a) { return adaptor_pre(p, two[k], a, enc, enc_tmax); }, {&p->ad_prog},
                    none, [&](int k, const RunArgs& a) { return adaptor_post(p, two[k], a, out, out_tmax, out_bf16); });
  }
  const RunArgs a = args_of(all, cu, s);
  if (int rc = adaptor_pre(p, all, a, enc, enc_tmax)) return rc;
  if (int rc = run_prog(p->ad_prog, a)) return rc;
  return adaptor_post(p, all, a, out, out_tmax, out_bf16);
}

extern "C" int fa_encode_adapt(fa_plan* p, const float* feats, int32_t feats_tmax, const int32_t* cu, const int32_t* cu_host,
                               int32_t n_utts, int32_t M, int32_t max_len, float* enc_out, int32_t enc_out_tmax, void* out,
                               int32_t out_tmax, int32_t out_bf16, fa_stream_t stream) {
  if (int r = check_batch(p, cu, n_utts, M, max_len)) return r;
  FA_REQUIRE(feats != nullptr && out != nullptr, "NULL feats / out");
  FA_REQUIRE(p->cfg.n_blocks > 0 && p->w.lin1_w != nullptr, "plan needs both encoder and adaptor weights");
  if (int r = check_tmax(feats_tmax, max_len, "feats")) return r;
  if (int r = check_tmax(enc_out_tmax, max_len, "enc_out")) return r;
  if (int r = check_tmax(out_tmax, max_len, "out")) return r;
  const cudaStream_t s = static_cast<cudaStream_t>(stream);
  const Slab all = whole(n_utts, M, max_len);
  if (use_graph(p, M))
    return run_graphed(p, n_utts, M, max_len, cu, s, [&](const RunArgs& a) { return encoder_pre(p, all, a, feats, feats_tmax); },
                       {{0, &p->enc_prog}, {1, &p->ad_prog}},
                       [&](const RunArgs& a) { return encoder_post(p, all, a, enc_out, enc_out_tmax); },
                       [&](const RunArgs& a) { return adaptor_post(p, all, a, out, out_tmax, out_bf16); });
  Slab two[2];
  if (split_two(p, cu_host, n_utts, M, two)) {
    return run_dual(p, two, cu, s, [&](int k, const RunArgs& a) { return encoder_pre(p, two[k], a, feats, feats_tmax); },
                    {&p->enc_prog, &p->ad_prog}, [&](int k, const RunArgs& a) { return encoder_post(p, two[k], a, enc_out, enc_out_tmax); },
                    [&](int k, const RunArgs& a) { return adaptor_post(p, two[k], a, out, out_tmax, out_bf16); });
  }
  const RunArgs a = args_of(all, cu, s);
  if (int rc = encoder_pre(p, all, a, feats, feats_tmax)) return rc;
  if (int rc = run_prog(p->enc_prog, a)) return rc;
  if (int rc = encoder_post(p, all, a, enc_out, enc_out_tmax)) return rc;
  if (int rc = run_prog(p->ad_prog, a)) return rc;
  return adaptor_post(p, all, a, out, out_tmax, out_bf16);
}
