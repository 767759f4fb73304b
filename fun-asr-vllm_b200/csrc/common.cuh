// Shared host/device helpers for libfunasr_b200 (sm_100a only).
#pragma once
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <atomic>
#include <cstdarg>
#include <cstdio>

#include "../../include/funasr_b200.h"

namespace fa {

// ---- host-side error plumbing --------------------------------------------------------------
void set_error(const char* fmt, ...);

#define FA_CHECK_CUDA(expr)                                                                 \
  do {                                                                                      \
    cudaError_t _e = (expr);                                                                \
    if (_e != cudaSuccess) {                                                                \
      fa::set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__, cudaGetErrorString(_e)); \
      return 1;                                                                             \
    }                                                                                       \
  } while (0)

#define FA_REQUIRE(cond, ...)          \
  do {                                 \
    if (!(cond)) {                     \
      fa::set_error(__VA_ARGS__);      \
      return 2;                        \
    }                                  \
  } while (0)

#define FA_CHECK_LAUNCH() FA_CHECK_CUDA(cudaGetLastError())

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
int sm_count();  // of the CURRENT device (cached per device)

// cudaFuncAttributeMaxDynamicSharedMemorySize is a PER-DEVICE attribute of a kernel: remember, per kernel, the devices it has
// been opted in on (one process may drive several GPUs; host threads may race here - setting it twice is harmless).
struct SmemOptIn {
  std::atomic<unsigned long long> done{0};
};
template <typename Kernel>
inline int opt_in_smem(Kernel kernel, int bytes, SmemOptIn& st) {
  int dev = 0;
  FA_CHECK_CUDA(cudaGetDevice(&dev));
  const unsigned long long bit = dev < 64 ? (1ull << dev) : 0ull;
  if (bit && (st.done.load(std::memory_order_acquire) & bit)) return 0;
  FA_CHECK_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  if (bit) st.done.fetch_or(bit, std::memory_order_release);
  return 0;
}

// ---- per-kernel-class device timing (bench / profiling only; off by default, zero cost when off) ----
enum Kind { K_FRONTEND = 0, K_GEMM, K_ATTENTION, K_LAYERNORM, K_FSMN, K_REPACK, K_SPLICE, K_NUM };
extern bool g_prof_on;
void prof_mark(int kind, cudaStream_t s, bool begin);
struct ProfScope {  // brackets the launches issued during its lifetime with two events on `s`
  int kind;
  cudaStream_t s;
  ProfScope(int k, cudaStream_t st) : kind(k), s(st) { if (g_prof_on) prof_mark(kind, s, true); }
  ~ProfScope() { if (g_prof_on) prof_mark(kind, s, false); }
};

// ---- internal launchers shared between api.cu and the op entry points -----------------------
struct TensorMap2D {  // opaque 128-byte CUtensorMap
  alignas(64) unsigned char bytes[128];
};

struct GemmEpilogue {
  const float* bias = nullptr;  // [N]
  const float* res1 = nullptr;  // f32 [M, N] (row stride ldc)
  const float* res2 = nullptr;  // f32 [M, N]
  // f32 output: one more residual kept in bf16, [M, N] with row stride ld_res_bf16 (the FSMN memory the attention kernel
  // leaves beside its context): read straight from global by the epilogue lane that owns the row, one unit ahead
  const __nv_bfloat16* res_bf16 = nullptr;
  int64_t ld_res_bf16 = 0;
  int relu = 0;
  int out_f32 = 0;
  int accumulate = 0;  // f32 output only: C += acc + bias (+ res*) instead of C = ...; the residual stream is updated in place
  // LayerNorm folded into the GEMMs around it (no LN kernel, no second pass over the f32 residual stream):
  //  ln_mode 2 (producer; f32 output with res1 = the old residual stream): besides C = x_new the epilogue writes
  //    xb = bf16(x_new) (the consumer's A operand) and per-row partial sums (sum, sum of squares of the bf16 values) into
  //    ln_stats[row * ln_slots + slot], slot = (n tile, epilogue half): plain stores, no atomics, no zeroing.
  //  ln_mode 1 (consumer; bf16 output): A is the raw xb, the weight is W * gamma, bias is b + W beta, and the epilogue
  //    applies  rstd_r * (acc - mean_r * colsum[n]) + bias[n]  with mean/rstd from the ln_slots partials (eps 1e-12).
  int ln_mode = 0;
  float2* ln_stats = nullptr;
  int ln_slots = 0;             // consumer: partials per row (= 2 * N_producer / BN_producer); producer: set by the launcher
  int ln_dim = 0;               // consumer: row length the statistics are over
  const float* colsum = nullptr;  // consumer: [N] row sums of the bf16 weight
  __nv_bfloat16* xb = nullptr;  // producer: bf16 copy of C, row stride ld_xb
  int64_t ld_xb = 0;
  const TensorMap2D* txb = nullptr;  // producer: prebuilt map of xb (32 x 32 boxes, 64B swizzle); NULL: encoded per launch
};

// rows x cols bf16 matrix, row stride ld (elements); box = box_rows x 64 cols, 128B swizzle.
int make_tensor_map_bf16(TensorMap2D* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld,
                         uint32_t box_rows);
// elem_bytes 2 (bf16) or 4 (f32); box = box_rows x (128 / elem_bytes) columns, 128B swizzle.
int make_tensor_map(TensorMap2D* out, const void* base, int elem_bytes, uint64_t rows, uint64_t cols, uint64_t ld,
                    uint32_t box_rows);

// box = box_rows x (box_bytes / elem_bytes) columns; box_bytes 128 (128B swizzle) or 64 (64B swizzle)
int make_tensor_map_w(TensorMap2D* out, const void* base, int elem_bytes, uint64_t rows, uint64_t cols, uint64_t ld,
                      uint32_t box_rows, uint32_t box_bytes);

int gemm_block_n(int M, int N, int n_res);  // tile-N heuristic (64 / 128 / 256); n_res = f32 residuals in the epilogue
// Rows [row0, row0 + M) of the A / C / residual matrices the maps describe (tr1 / tr2: f32 maps like tc, NULL when
// epi.res1 / epi.res2 are; the epi.res* pointers only select the epilogue variant).
bool gemm_use_pair(int M, int block_n);  // CTA-pair (cta_group::2) kernel for this launch? then tb must box block_n / 2 rows
int launch_gemm(const TensorMap2D& ta, const TensorMap2D& tb, const TensorMap2D& tc, const TensorMap2D* tr1,
                const TensorMap2D* tr2, int block_n, int M, int N, int K, const GemmEpilogue& epi, void* c, int64_t ldc,
                int row0, bool pair, cudaStream_t stream);

// ---- the row-local GEMM chain of a SAN-M block in one launch (gemm_chain.cu) ----
#define FA_CHAIN_MAX_PROBS 4
#define FA_CHAIN_MAX_SLOTS 32
struct ChainProb {
  const TensorMap2D* ta;  // A [rows, K] bf16, box 128 rows x 64 columns
  const TensorMap2D* tb;  // W [N, K] bf16, box 128 rows x 64 columns (each CTA of a pair stages half of a 256-wide tile)
  const TensorMap2D* tc;  // C [rows, N]: f32 box 32 x 32 (kind 0, also the map the old C rows are read through) or bf16 box 32 x 64
  const TensorMap2D* txb = nullptr;  // kind 0: the bf16 copy xb [rows, N], box 32 rows x 32 columns (64 bytes, 64B swizzle)
  int N = 0, K = 0;
  int kind = 0;  // 0: f32 C = acc + bias + C (+ res_bf16), xb = bf16(C), row partials -> stats (folded-LN producer)
                 // 1: bf16 C = rstd * (acc - mean * colsum) + bias (, relu) from the partials of the problem before it
  int relu = 0;
  const float* bias = nullptr;
  const float* colsum = nullptr;
  float2* stats = nullptr;
  __nv_bfloat16* xb = nullptr;
  int64_t ld_xb = 0;
  const __nv_bfloat16* res_bf16 = nullptr;
  int64_t ld_res_bf16 = 0;
  void* c = nullptr;
};
struct ChainDesc {
  int n_probs = 0;
  ChainProb p[FA_CHAIN_MAX_PROBS];  // problem i + 1 reads what problem i writes for the same rows
};
// rows [row0, row0 + M) of every matrix; flags: n_probs * ceil(M / 256) ints, zero on entry (dirty afterwards)
int launch_gemm_chain(const ChainDesc& d, int M, int row0, int* flags, cudaStream_t stream);
int gemm_chain_flag_ints(int n_probs, int max_rows);

int launch_layernorm(const float* x, const float* g, const float* b, const float* g2, const float* b2,
                     __nv_bfloat16* y_bf16, float* y_f32, int M, int D, cudaStream_t stream);
// Row-range convention of the kernels below: every data pointer addresses LOCAL row 0 of a slab of M rows holding the
// n_utts utterances cu[0..n_utts]; cu keeps ABSOLUTE packed row numbers and `base` (= cu[0] of the slab) is subtracted
// inside the kernel.  base == 0 for a whole batch; the chunked schedule in api.cu passes slabs.
int launch_embed_layernorm(const float* x, int in_tmax, const float* inv_ts, const int32_t* cu, int n_utts, int base,
                           float scale, const float* g, const float* b, __nv_bfloat16* y, int M, int D, cudaStream_t stream);
int launch_fsmn(const __nv_bfloat16* qkv, int64_t ld, int v_col, const float* w, const int32_t* cu, int n_utts, int base,
                void* mem, int out_bf16, int64_t ld_out, int M, int D, int kernel, int left_pad, cudaStream_t stream);
// tm: prebuilt tensor map over qkv (box = attention_kv_box() rows x 64 columns), or NULL to encode one over n_rows rows here.
// fsmn_w != NULL: the FSMN memory of the same value rows (f32 taps [11, n_heads * 128] tap-major, `fsmn_left` rows of left
// context) is computed inside the attention kernel and written as bf16 to mem[row, h * 128 + c] (row stride ld_mem).
int attention_kv_box();
bool attention_can_fuse_fsmn(int kernel, int64_t ld_mem, const void* mem);
int launch_attention(const TensorMap2D* tm, const __nv_bfloat16* qkv, int64_t ld, const int32_t* cu_seqlens, int n_utts, int base,
                     int max_len, int n_heads, __nv_bfloat16* ctx, int64_t ld_ctx, int n_rows, const float* fsmn_w, int fsmn_left,
                     __nv_bfloat16* mem, int64_t ld_mem, cudaStream_t stream, int* zero_ints = nullptr, int n_zero = 0,
                     float* acc = nullptr, int64_t ld_acc = 0,  // acc != NULL: acc[row, h * 128 + c] += memory instead of the bf16 store
                     const TensorMap2D* tm_ctx = nullptr);      // prebuilt map of ctx (32 x 64 boxes), or NULL to encode one here
// dst[i] = cu[min(i, n_utts)], i = 0..n_pad: cu_seqlens copied to a fixed address, padded with the total row count
int launch_pad_cu(const int32_t* cu, int n_utts, int32_t* dst, int n_pad, cudaStream_t stream);
int launch_repack_rows(const float* src, int src_tmax, void* dst, int dst_tmax, int dst_bf16, const int32_t* cu,
                       int n_utts, int base, int M, int D, cudaStream_t stream);

}  // namespace fa

// ---- device helpers --------------------------------------------------------------------------
#ifdef __CUDACC__
#include "ptx.cuh"
namespace fa {

// Programmatic dependent launch: every kernel of the path lets its successor in the stream start early
// (pdl_launch_dependents at entry) and blocks, after its own prologue (barrier init, TMEM allocation, descriptor
// prefetch, constant staging), until the predecessor grid has completed and flushed (pdl_wait).  Nothing produced by
// another kernel may be read, and nothing written, before pdl_wait().
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// Launch with the programmatic-stream-serialization attribute (and an optional cluster width).
bool pdl_enabled();
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, int cluster_x,
                              Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  int n = 0;
  if (cluster_x > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = cluster_x;
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = 1;
    ++n;
  }
  if (pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ uint32_t pack_bf16x2(float lo, float hi) {
  __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&t);
}
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

}  // namespace fa
#endif
