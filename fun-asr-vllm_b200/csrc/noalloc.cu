// fa_op_ln_linear_static: LayerNorm -> Linear with the intermediate rows in a file-scope device array.
//
// For callers that cannot allocate: code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc.
// The op is the pre-LN projection of a block done the plain way (LayerNorm kernel, then the tcgen05 GEMM): the bf16
// normalized rows between the two launches are scratch.  Every other entry point takes that scratch from the caller
// (fa_op_layernorm's y_bf16_dev handed to fa_op_gemm as a_dev, or a slice of the plan workspace); this one keeps it in
// `g_ln_rows` below, so the call allocates nothing - no cudaMalloc / cudaFree, no allocating library routine.
//
// The array is part of the module image: every GPU that loads the library holds FA_STATIC_LN_ELEMS * 2 bytes (4 MiB)
// until the process exits.  In host code the name `g_ln_rows` is only a host-side stand-in for the symbol, not a device
// address: the address is looked up with cudaGetSymbolAddress, once per device with that device current, and THAT pointer
// is what the kernels receive.
//
// Sharing: all calls on a device use the same array and it keeps the last call's contents.  Nothing in it has to start at
// zero: the LayerNorm launch writes rows [0, M) completely and the GEMM's tensor map covers exactly those rows, so a call
// never reads what an earlier call left.  Calls that overlap on the GPU (two streams, two host threads on one device) RACE
// on the array: issue them on one stream, or order them with events.
#include <mutex>

#include "common.cuh"

namespace fa {
namespace noalloc {

static __device__ __align__(256) __nv_bfloat16 g_ln_rows[FA_STATIC_LN_ELEMS];

constexpr int MAX_DEVICES = 64;
std::mutex g_mu;
__nv_bfloat16* g_addr[MAX_DEVICES] = {};

// Device address of the scratch array on the CURRENT device (looked up once per device, then cached).
int scratch_address(__nv_bfloat16** out) {
  int dev = -1;
  FA_CHECK_CUDA(cudaGetDevice(&dev));
  FA_REQUIRE(dev >= 0 && dev < MAX_DEVICES, "device index %d outside the %d the static scratch table tracks", dev, MAX_DEVICES);
  std::lock_guard<std::mutex> lk(g_mu);
  if (g_addr[dev] == nullptr) {
    void* p = nullptr;
    FA_CHECK_CUDA(cudaGetSymbolAddress(&p, g_ln_rows));
    FA_REQUIRE(p != nullptr, "cudaGetSymbolAddress returned a NULL address for the static scratch");
    g_addr[dev] = static_cast<__nv_bfloat16*>(p);
  }
  *out = g_addr[dev];
  return 0;
}

}  // namespace noalloc
}  // namespace fa

using namespace fa;

extern "C" size_t fa_static_ln_capacity(void) { return static_cast<size_t>(FA_STATIC_LN_ELEMS); }

extern "C" int fa_op_ln_linear_static(const float* x_dev, const float* g_dev, const float* b_dev, const void* w_dev, int64_t ldw,
                                      const float* bias_dev, void* c_dev, int64_t ldc, int32_t M, int32_t N, int32_t D,
                                      int32_t relu, int32_t out_f32, fa_stream_t stream) {
  FA_REQUIRE(x_dev && g_dev && b_dev && w_dev && c_dev, "ln_linear_static: NULL pointer");
  FA_REQUIRE(M > 0 && N > 0 && D > 0, "ln_linear_static: M, N, D must be positive (got %d, %d, %d)", M, N, D);
  FA_REQUIRE(D % 8 == 0, "ln_linear_static: D = %d must be a multiple of 8 (16-byte bf16 rows for the GEMM's tensor map)", D);
  FA_REQUIRE(static_cast<int64_t>(M) * D <= static_cast<int64_t>(FA_STATIC_LN_ELEMS),
             "ln_linear_static: M * D = %lld exceeds the static scratch of %d bf16 elements (e.g. %d rows at D = %d); use "
             "fa_op_layernorm + fa_op_gemm with a caller-owned buffer for larger batches",
             static_cast<long long>(M) * D, FA_STATIC_LN_ELEMS, FA_STATIC_LN_ELEMS / D, D);
  __nv_bfloat16* rows = nullptr;
  if (int r = noalloc::scratch_address(&rows)) return r;
  // LayerNorm (eps 1e-12) -> bf16 rows [M, D] in the static array, then C = rows W^T + bias on the same stream
  if (int r = fa_op_layernorm(x_dev, g_dev, b_dev, rows, nullptr, M, D, stream)) return r;
  return fa_op_gemm(rows, D, w_dev, ldw, bias_dev, nullptr, nullptr, nullptr, 0, c_dev, ldc, M, N, D, relu, out_f32, stream);
}
