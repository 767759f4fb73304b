// HBM/L2-bound row kernels of the SAN-M encoder: LayerNorm (+ fused input scaling / positional
// encoding, + fused second LayerNorm), FSMN depthwise memory block, row metadata, casts, splice.
// All work on PACKED rows; utterance boundaries come from cu_seqlens / row_meta.
#include <math.h>

#include "common.cuh"

namespace fa {
namespace ew {

constexpr int LN_MAX_D = 1024;
constexpr int LN_VEC = LN_MAX_D / 128;  // float4 per lane

// ---------------------------------------------------------------- utterance lookup
// (first row, one-past-last row) of the utterance owning packed row r; cu_seqlens is tiny and
// L1-resident, so a binary search per row is cheaper than materialising a per-row table.
__device__ __forceinline__ int utt_index(const int32_t* __restrict__ cu, int n_utts, int r) {
  int lo = 0, hi = n_utts;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(cu + mid) <= r) lo = mid; else hi = mid;
  }
  return lo;
}
__device__ __forceinline__ int2 utt_bounds(const int32_t* __restrict__ cu, int n_utts, int r) {
  const int lo = utt_index(cu, n_utts, r);
  return make_int2(__ldg(cu + lo), __ldg(cu + lo + 1));
}

// ---------------------------------------------------------------- LayerNorm (eps 1e-12)
// One warp per row, the row lives in registers.  Optional fused pre-op (x*scale + PE(pos)) and an
// optional second LayerNorm applied to the first one's output (after_norm -> next norm1).
struct LnArgs {
  const float* x;
  const float* g;
  const float* b;
  const float* g2;  // second LN (nullable)
  const float* b2;
  __nv_bfloat16* y_bf16;  // output of the LAST LN (nullable)
  float* y_f32;           // output of the FIRST LN (nullable)
  const float* inv_ts;    // positional encoding (nullable)
  const int32_t* cu;
  int n_utts;
  int base;  // absolute packed row of local row 0
  float scale;
  int M, D;
  int in_tmax;  // EMBED only: 0 = x is packed [M, D]; else x is padded [n_utts, in_tmax, D]
};

__device__ __forceinline__ void ln_stats(const float4 (&v)[LN_VEC], int nvec, int lane, int D, float& mean, float& rstd) {
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < LN_VEC; ++i)
    if (lane + 32 * i < nvec) s += v[i].x + v[i].y + v[i].z + v[i].w;
  mean = warp_sum(s) / static_cast<float>(D);
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < LN_VEC; ++i)
    if (lane + 32 * i < nvec) {
      const float a = v[i].x - mean, b = v[i].y - mean, c = v[i].z - mean, d = v[i].w - mean;
      q += a * a + b * b + c * c + d * d;
    }
  const float var = warp_sum(q) / static_cast<float>(D);
  rstd = rsqrtf(var + 1e-12f);  // VLLM/model_executor/models/funasr.py:85 eps=1e-12
}

template <bool EMBED>
__global__ void __launch_bounds__(256) layernorm_kernel(const LnArgs a) {
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (row >= a.M) return;
  const int D = a.D, nvec = D >> 2;
  int64_t src_row = row;
  int pos_i = 0;
  if (EMBED) {
    const int b = utt_index(a.cu, a.n_utts, row + a.base);
    pos_i = row + a.base - __ldg(a.cu + b);
    if (a.in_tmax > 0) src_row = static_cast<int64_t>(b) * a.in_tmax + pos_i;
  }
  const float4* xr = reinterpret_cast<const float4*>(a.x + src_row * D);
  float4 v[LN_VEC];
#pragma unroll
  for (int i = 0; i < LN_VEC; ++i) {
    const int c = lane + 32 * i;
    v[i] = (c < nvec) ? xr[c] : make_float4(0, 0, 0, 0);
  }
  if (EMBED) {
    // xs * sqrt(d_model) + [sin(p*inv) | cos(p*inv)],  p = 1..T within the utterance
    // (VLLM/model_executor/models/funasr.py:388-390, 256-290); fp32 like the reference.
    const float pos = static_cast<float>(pos_i + 1);
    const int half = D >> 1;
#pragma unroll
    for (int i = 0; i < LN_VEC; ++i) {
      const int c = lane + 32 * i;
      if (c < nvec) {
        float* e = reinterpret_cast<float*>(&v[i]);
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          const int col = 4 * c + k;
          const float ang = pos * __ldg(a.inv_ts + (col < half ? col : col - half));
          const float pe = col < half ? sinf(ang) : cosf(ang);
          e[k] = e[k] * a.scale + pe;
        }
      }
    }
  }
  float mean, rstd;
  ln_stats(v, nvec, lane, D, mean, rstd);
#pragma unroll
  for (int i = 0; i < LN_VEC; ++i) {
    const int c = lane + 32 * i;
    if (c < nvec) {
      const float4 g = __ldg(reinterpret_cast<const float4*>(a.g) + c);
      const float4 b = __ldg(reinterpret_cast<const float4*>(a.b) + c);
      v[i].x = (v[i].x - mean) * rstd * g.x + b.x;
      v[i].y = (v[i].y - mean) * rstd * g.y + b.y;
      v[i].z = (v[i].z - mean) * rstd * g.z + b.z;
      v[i].w = (v[i].w - mean) * rstd * g.w + b.w;
      if (a.y_f32) reinterpret_cast<float4*>(a.y_f32 + static_cast<int64_t>(row) * D)[c] = v[i];
    }
  }
  if (a.g2) {
    ln_stats(v, nvec, lane, D, mean, rstd);
#pragma unroll
    for (int i = 0; i < LN_VEC; ++i) {
      const int c = lane + 32 * i;
      if (c < nvec) {
        const float4 g = __ldg(reinterpret_cast<const float4*>(a.g2) + c);
        const float4 b = __ldg(reinterpret_cast<const float4*>(a.b2) + c);
        v[i].x = (v[i].x - mean) * rstd * g.x + b.x;
        v[i].y = (v[i].y - mean) * rstd * g.y + b.y;
        v[i].z = (v[i].z - mean) * rstd * g.z + b.z;
        v[i].w = (v[i].w - mean) * rstd * g.w + b.w;
      }
    }
  }
  if (a.y_bf16) {
    uint2* yr = reinterpret_cast<uint2*>(a.y_bf16 + static_cast<int64_t>(row) * D);
#pragma unroll
    for (int i = 0; i < LN_VEC; ++i) {
      const int c = lane + 32 * i;
      if (c < nvec) yr[c] = make_uint2(pack_bf16x2(v[i].x, v[i].y), pack_bf16x2(v[i].z, v[i].w));
    }
  }
}

// ---------------------------------------------------------------- FSMN memory block
// mem[t, c] = v[t, c] + sum_k w[k, c] * v[t + k - left, c]   (zero outside the utterance; w is TAP-MAJOR [K, D])
// (VLLM/model_executor/models/funasr.py:181-201: mask, ConstantPad1d, depthwise Conv1d, +inputs, mask).
// Output: f32 [M, ld_out] or bf16 (the plan writes bf16 straight into the right half of the out-projection's A operand).
// The row count of a launch is min(M, cu[n_utts] - base): a launch captured in a CUDA graph for a row BUCKET (api.cu) still
// stops at the batch's true last row, which it reads from cu_seqlens on the device.
constexpr int FSMN_MAXK = 16;
constexpr int FSMN_THREADS = 128;

__device__ __forceinline__ float4 bf16x4_to_f32(uint2 r) {
  return make_float4(__uint_as_float(r.x << 16), __uint_as_float(r.x & 0xffff0000u), __uint_as_float(r.y << 16),
                     __uint_as_float(r.y & 0xffff0000u));
}

// Generic fallback (any tap count <= 16, rows only 8-byte aligned): one thread = one output row x 4 channels, taps read
// straight from global memory.  The model's own geometry (11 taps, 16-byte aligned rows) takes fsmn_staged_kernel below.
template <bool OUT_BF16>
__global__ void __launch_bounds__(FSMN_THREADS)
fsmn_generic_kernel(const __nv_bfloat16* __restrict__ qkv, int64_t ld, int v_col, const float* __restrict__ w,
                    const int32_t* __restrict__ cu, int n_utts, int base, void* __restrict__ mem, int64_t ld_out, int M, int D,
                    int K, int left) {
  pdl_launch_dependents();
  const int c4 = blockIdx.y * FSMN_THREADS + threadIdx.x;  // channel quad
  const int r = blockIdx.x;
  if (4 * c4 >= D) return;
  M = min(M, __ldg(cu + n_utts) - base);
  if (r >= M) return;
  pdl_wait();
  const int2 ub = utt_bounds(cu, n_utts, r + base);
  const int lo = max(ub.x - base, 0), hi = min(ub.y - base, M);
  const __nv_bfloat16* vcol = qkv + v_col + 4 * c4;
  float4 a = bf16x4_to_f32(__ldg(reinterpret_cast<const uint2*>(vcol + static_cast<int64_t>(r) * ld)));  // "+ inputs"
  for (int k = 0; k < K; ++k) {
    const int row = r + k - left;
    if (row < lo || row >= hi) continue;
    const float4 x = bf16x4_to_f32(__ldg(reinterpret_cast<const uint2*>(vcol + static_cast<int64_t>(row) * ld)));
    const float4 wt = __ldg(reinterpret_cast<const float4*>(w + static_cast<int64_t>(k) * D) + c4);
    a.x = fmaf(wt.x, x.x, a.x); a.y = fmaf(wt.y, x.y, a.y); a.z = fmaf(wt.z, x.z, a.z); a.w = fmaf(wt.w, x.w, a.w);
  }
  const int64_t o = static_cast<int64_t>(r) * ld_out + 4 * c4;
  if (OUT_BF16) *reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(mem) + o) = make_uint2(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w));
  else *reinterpret_cast<float4*>(static_cast<float*>(mem) + o) = a;
}

// ---------------------------------------------------------------- FSMN, shared-memory staged (the default path)
// A CTA owns R consecutive rows x 512 channels and pulls ALL the input rows it needs (R + K - 1, one 1 KB bulk copy per
// row: cp.async.bulk global -> shared, completion on one mbarrier) in a single round trip.  One thread owns 4 adjacent
// channels and walks the rows with a K-row sliding window held in registers (the row loop is unrolled K times so the window
// rotates through compile-time register names), fed by one conflict-free LDS.64 per row; a CTA's rows are cut into
// per-utterance segments and rows outside the segment's utterance enter the window as zeros, so the inner loop carries no
// per-tap predicates.  The "+ inputs" term is tap `left` with weight + 1.
template <int K, int R, bool OUT_BF16>
__global__ void __launch_bounds__(FSMN_THREADS)
fsmn_staged_kernel(const __nv_bfloat16* __restrict__ qkv, int64_t ld, int v_col, const float* __restrict__ w,
                   const int32_t* __restrict__ cu, int n_utts, int base, void* __restrict__ mem, int64_t ld_out, int M, int D,
                   int left) {
  extern __shared__ __align__(128) unsigned char fs_smem[];
  pdl_launch_dependents();
  const int ch0 = blockIdx.y * (4 * FSMN_THREADS);  // first channel of this CTA's 512-channel chunk
  const int wch = min(4 * FSMN_THREADS, D - ch0);   // channels in the chunk (multiple of 8: 16-byte bulk copies)
  M = min(M, __ldg(cu + n_utts) - base);  // cu_seqlens is an input of the forward call, not a predecessor's output
  const int r0 = blockIdx.x * R, r_end = min(M, r0 + R);
  if (r0 >= M) return;  // uniform per CTA, before any barrier use
  const int ahead = K - 1 - left;  // newest input row of output row t is t + ahead
  const int in0 = max(r0 - left, 0), in1 = min(r_end + ahead, M);  // input rows staged: [in0, in1), at most R + K - 1
  const uint32_t tile = smem_u32(fs_smem), bar = tile + (R + K - 1) * 1024;
  if (threadIdx.x == 0) {
    ptx::mbar_init(bar, 1);
    ptx::fence_mbar_init();
    ptx::mbar_expect_tx(bar, static_cast<uint32_t>(in1 - in0) * wch * 2);  // the only arrival: the phase cannot end before it
  }
  const bool active = 4 * static_cast<int>(threadIdx.x) < wch;
  float4 wt[K];
#pragma unroll
  for (int k = 0; k < K; ++k) {
    wt[k] = active ? __ldg(reinterpret_cast<const float4*>(w + static_cast<int64_t>(k) * D + ch0) + threadIdx.x)
                   : make_float4(0.f, 0.f, 0.f, 0.f);
    const float one = (k == left) ? 1.f : 0.f;  // "+ inputs" = tap `left` with weight + 1
    wt[k].x += one; wt[k].y += one; wt[k].z += one; wt[k].w += one;
  }
  __syncthreads();  // barrier initialised and armed
  pdl_wait();       // the taps are constants; v comes from the previous kernel
  if (static_cast<int>(threadIdx.x) < in1 - in0)
    ptx::bulk_g2s(tile + threadIdx.x * 1024, qkv + static_cast<int64_t>(in0 + threadIdx.x) * ld + v_col + ch0, wch * 2, bar);
  if (!active) return;
  int2 ub = utt_bounds(cu, n_utts, r0 + base);  // overlaps the copies
  ptx::mbar_wait(bar, 0);
  const uint32_t col = tile + threadIdx.x * 8;
  int r = r0;
  while (r < r_end) {
    if (r != r0) ub = utt_bounds(cu, n_utts, r + base);
    const int lo = max(ub.x - base, in0), hi = min(ub.y - base, in1);  // rows of this utterance that exist in the tile
    const int seg_end = min(r_end, ub.y - base);
    auto load = [&](int row) -> float4 {
      uint2 v = make_uint2(0u, 0u);
      if (row >= lo && row < hi) v = ptx::lds64(col + static_cast<uint32_t>(row - in0) * 1024u);
      return bf16x4_to_f32(v);
    };
    float4 win[K];
#pragma unroll
    for (int i = 0; i < K - 1; ++i) win[i] = load(r - left + i);
    for (int tb = r; tb < seg_end; tb += K) {
#pragma unroll
      for (int j = 0; j < K; ++j) {
        if (tb + j < seg_end) {
          win[(j + K - 1) % K] = load(tb + j + ahead);
          float2 a01 = make_float2(0.f, 0.f), a23 = make_float2(0.f, 0.f);
#pragma unroll
          for (int k = 0; k < K; ++k) {  // packed f32x2 FMAs (Blackwell FFMA2): same IEEE fma per channel, half the issue slots
            const float4 x = win[(j + k) % K];
            a01 = __ffma2_rn(make_float2(wt[k].x, wt[k].y), make_float2(x.x, x.y), a01);
            a23 = __ffma2_rn(make_float2(wt[k].z, wt[k].w), make_float2(x.z, x.w), a23);
          }
          const float4 a = make_float4(a01.x, a01.y, a23.x, a23.y);
          const int64_t o = static_cast<int64_t>(tb + j) * ld_out + ch0 + 4 * threadIdx.x;
          if (OUT_BF16) *reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(mem) + o) = make_uint2(pack_bf16x2(a.x, a.y), pack_bf16x2(a.z, a.w));
          else *reinterpret_cast<float4*>(static_cast<float*>(mem) + o) = a;
        }
      }
    }
    r = seg_end;
  }
}

// ---------------------------------------------------------------- row re-layout / splice
// Row re-layout + cast: dst row r (packed [M,D] or padded [n_utts, dst_tmax, D]) <- the matching
// src row (packed or padded [n_utts, src_tmax, D]), f32 -> f32 | bf16; padding rows are zero-filled.
// This is the only place the padded [B, Tmax, .] layout of the reference API exists.
__global__ void __launch_bounds__(256)
repack_rows_kernel(const float* __restrict__ src, int src_tmax, void* __restrict__ dst, int dst_tmax, int dst_bf16,
                   const int32_t* __restrict__ cu, int n_utts, int base, int dst_rows, int D) {
  pdl_launch_dependents();
  pdl_wait();
  const int lane = threadIdx.x & 31;
  const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  if (r >= dst_rows) return;
  int b, t;
  bool valid = true;
  if (dst_tmax > 0) {
    b = r / dst_tmax;
    t = r - b * dst_tmax;
    valid = t < __ldg(cu + b + 1) - __ldg(cu + b);
  } else {
    b = utt_index(cu, n_utts, r + base);
    t = r + base - __ldg(cu + b);
  }
  const int64_t sr = src_tmax > 0 ? static_cast<int64_t>(b) * src_tmax + t : static_cast<int64_t>(__ldg(cu + b) - base) + t;
  const float4* s4 = reinterpret_cast<const float4*>(src + sr * D);
  const int nvec = D >> 2;
  for (int c = lane; c < nvec; c += 32) {
    const float4 v = valid ? __ldg(s4 + c) : make_float4(0.f, 0.f, 0.f, 0.f);
    if (dst_bf16)
      reinterpret_cast<uint2*>(static_cast<__nv_bfloat16*>(dst) + static_cast<int64_t>(r) * D)[c] =
          make_uint2(pack_bf16x2(v.x, v.y), pack_bf16x2(v.z, v.w));
    else
      reinterpret_cast<float4*>(static_cast<float*>(dst) + static_cast<int64_t>(r) * D)[c] = v;
  }
}

template <typename T>
__device__ __forceinline__ float ld_as_f32(const void* p, int64_t i);
template <>
__device__ __forceinline__ float ld_as_f32<float>(const void* p, int64_t i) { return static_cast<const float*>(p)[i]; }
template <>
__device__ __forceinline__ float ld_as_f32<__nv_bfloat16>(const void* p, int64_t i) {
  return __bfloat162float(static_cast<const __nv_bfloat16*>(p)[i]);
}

// grid: (row within utterance output, utterance).  Row j of utterance b is prefix[j], speech
// [cu[b] + j - P] or suffix[j - P - take]; torch.cat of infer.py:462-466 as one gather.
__global__ void __launch_bounds__(256)
splice_kernel(const void* __restrict__ speech, int speech_bf16, const int32_t* __restrict__ cu,
              const int32_t* __restrict__ take, const void* __restrict__ prefix, int P, const void* __restrict__ suffix,
              int S, const int64_t* __restrict__ out_off, int dim, void* __restrict__ out, int out_bf16) {
  const int b = blockIdx.y, j = blockIdx.x;
  const int len = cu[b + 1] - cu[b];
  const int tk = take ? min(max(take[b], 0), len) : len;
  if (j >= P + tk + S) return;
  const void* src;
  int64_t src_row;
  int src_bf16;
  if (j < P) { src = prefix; src_row = j; src_bf16 = out_bf16; }
  else if (j < P + tk) { src = speech; src_row = cu[b] + (j - P); src_bf16 = speech_bf16; }
  else { src = suffix; src_row = j - P - tk; src_bf16 = out_bf16; }
  const int64_t orow = out_off[b] + j;
  for (int c = threadIdx.x; c < dim; c += blockDim.x) {
    const float v = src_bf16 ? ld_as_f32<__nv_bfloat16>(src, src_row * dim + c) : ld_as_f32<float>(src, src_row * dim + c);
    if (out_bf16) static_cast<__nv_bfloat16*>(out)[orow * dim + c] = __float2bfloat16_rn(v);
    else static_cast<float*>(out)[orow * dim + c] = v;
  }
}

__global__ void pad_cu_kernel(const int32_t* __restrict__ cu, int n_utts, int32_t* __restrict__ dst, int n_pad) {
  pdl_launch_dependents();
  pdl_wait();
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i <= n_pad; i += gridDim.x * blockDim.x) dst[i] = __ldg(cu + min(i, n_utts));
}

}  // namespace ew

int launch_pad_cu(const int32_t* cu, int n_utts, int32_t* dst, int n_pad, cudaStream_t stream) {
  FA_REQUIRE(cu && dst && n_utts > 0 && n_pad >= n_utts, "pad_cu: bad arguments");
  ew::pad_cu_kernel<<<ceil_div(n_pad + 1, 256), 256, 0, stream>>>(cu, n_utts, dst, n_pad);
  FA_CHECK_LAUNCH();
  return 0;
}

static int ln_check(int D, const void* x) {
  FA_REQUIRE(D % 4 == 0 && D <= ew::LN_MAX_D, "layernorm: D=%d must be a multiple of 4 and <= %d", D, ew::LN_MAX_D);
  FA_REQUIRE((reinterpret_cast<uintptr_t>(x) & 15) == 0, "layernorm: x must be 16-byte aligned");
  return 0;
}

int launch_layernorm(const float* x, const float* g, const float* b, const float* g2, const float* b2,
                     __nv_bfloat16* y_bf16, float* y_f32, int M, int D, cudaStream_t stream) {
  if (int r = ln_check(D, x)) return r;
  ew::LnArgs a{x, g, b, g2, b2, y_bf16, y_f32, nullptr, nullptr, 0, 0, 1.f, M, D, 0};
  ProfScope _ps(K_LAYERNORM, stream);
  FA_CHECK_CUDA(launch_pdl(ew::layernorm_kernel<false>, dim3(ceil_div(M, 8)), dim3(256), 0, stream, 1, a));
  return 0;
}

int launch_embed_layernorm(const float* x, int in_tmax, const float* inv_ts, const int32_t* cu, int n_utts, int base,
                           float scale, const float* g, const float* b, __nv_bfloat16* y, int M, int D, cudaStream_t stream) {
  if (int r = ln_check(D, x)) return r;
  ew::LnArgs a{x, g, b, nullptr, nullptr, y, nullptr, inv_ts, cu, n_utts, base, scale, M, D, in_tmax};
  ProfScope _ps(K_LAYERNORM, stream);
  FA_CHECK_CUDA(launch_pdl(ew::layernorm_kernel<true>, dim3(ceil_div(M, 8)), dim3(256), 0, stream, 1, a));
  return 0;
}

int launch_fsmn(const __nv_bfloat16* qkv, int64_t ld, int v_col, const float* w, const int32_t* cu, int n_utts, int base,
                void* mem, int out_bf16, int64_t ld_out, int M, int D, int kernel, int left_pad, cudaStream_t stream) {
  FA_REQUIRE(kernel >= 1 && kernel <= ew::FSMN_MAXK, "fsmn: kernel size %d not in [1,%d]", kernel, ew::FSMN_MAXK);
  FA_REQUIRE(left_pad >= 0 && left_pad < kernel, "fsmn: left pad %d outside the %d-tap window", left_pad, kernel);
  FA_REQUIRE(D % 4 == 0 && v_col % 4 == 0 && ld % 4 == 0 && ld_out % 4 == 0 && ld_out >= D, "fsmn: D, v_col, ld and ld_out must be multiples of 4");
  FA_REQUIRE((reinterpret_cast<uintptr_t>(qkv) & 7) == 0 && (reinterpret_cast<uintptr_t>(mem) & 15) == 0 &&
                 (reinterpret_cast<uintptr_t>(w) & 15) == 0,
             "fsmn: qkv must be 8-byte, mem and w 16-byte aligned");
  const int threads = ew::FSMN_THREADS;
  ProfScope _ps(K_FSMN, stream);
  if (kernel == 11 && D % 8 == 0 && v_col % 8 == 0 && ld % 8 == 0 && (reinterpret_cast<uintptr_t>(qkv) & 15) == 0) {
    // rows per CTA: the largest of 32 / 16 / 8 that still gives every SM a CTA.  For small batches the serial walk over a
    // CTA's rows is the kernel's whole duration (cfg2 half-batch: 1 336 rows = 42 CTAs of 32 rows, 10.6 us per launch; 167
    // CTAs of 8 rows, 8.6 us); once every SM has work, more and smaller CTAs only add halo reads (profiles/r01_s_fsmn_rows.md).
    const int sms = sm_count();
    const int R = M >= 32 * sms ? 32 : M >= 16 * sms ? 16 : 8;
    const size_t smem = static_cast<size_t>(R + kernel - 1) * 1024 + 16;
    dim3 sgrid(ceil_div(M, R), ceil_div(D, 4 * threads));
#define FA_FSMN_STAGED_LAUNCH(RR, BF)                                                                                              \
  do {                                                                                                                             \
    static SmemOptIn opt;                                                                                                          \
    if (int r_ = opt_in_smem(ew::fsmn_staged_kernel<11, RR, BF>, static_cast<int>(smem), opt)) return r_;                          \
    FA_CHECK_CUDA(launch_pdl(ew::fsmn_staged_kernel<11, RR, BF>, sgrid, dim3(threads), smem, stream, 1, qkv, ld, v_col, w, cu,     \
                             n_utts, base, mem, ld_out, M, D, left_pad));                                                          \
  } while (0)
#define FA_FSMN_STAGED_ROWS(RR) \
  do { if (out_bf16) FA_FSMN_STAGED_LAUNCH(RR, true); else FA_FSMN_STAGED_LAUNCH(RR, false); } while (0)
    if (R == 32) FA_FSMN_STAGED_ROWS(32);
    else if (R == 16) FA_FSMN_STAGED_ROWS(16);
    else FA_FSMN_STAGED_ROWS(8);
#undef FA_FSMN_STAGED_ROWS
#undef FA_FSMN_STAGED_LAUNCH
    return 0;
  }
  dim3 grid(M, ceil_div(D / 4, threads));
  if (out_bf16)
    FA_CHECK_CUDA(launch_pdl(ew::fsmn_generic_kernel<true>, grid, dim3(threads), 0, stream, 1, qkv, ld, v_col, w, cu, n_utts, base, mem,
                             ld_out, M, D, kernel, left_pad));
  else
    FA_CHECK_CUDA(launch_pdl(ew::fsmn_generic_kernel<false>, grid, dim3(threads), 0, stream, 1, qkv, ld, v_col, w, cu, n_utts, base, mem,
                             ld_out, M, D, kernel, left_pad));
  return 0;
}

int launch_repack_rows(const float* src, int src_tmax, void* dst, int dst_tmax, int dst_bf16, const int32_t* cu,
                       int n_utts, int base, int M, int D, cudaStream_t stream) {
  FA_REQUIRE(D % 4 == 0, "repack: D=%d must be a multiple of 4", D);
  FA_REQUIRE((reinterpret_cast<uintptr_t>(src) & 15) == 0 && (reinterpret_cast<uintptr_t>(dst) & 15) == 0,
             "repack: buffers must be 16-byte aligned");
  const int rows = dst_tmax > 0 ? n_utts * dst_tmax : M;
  ProfScope _ps(K_REPACK, stream);
  FA_CHECK_CUDA(launch_pdl(ew::repack_rows_kernel, dim3(ceil_div(rows, 8)), dim3(256), 0, stream, 1, src, src_tmax, dst, dst_tmax,
                           dst_bf16, cu, n_utts, base, rows, D));
  return 0;
}

}  // namespace fa

using namespace fa;

extern "C" int fa_op_layernorm(const float* x_dev, const float* g_dev, const float* b_dev, void* y_bf16_dev,
                               float* y_f32_dev, int32_t M, int32_t D, fa_stream_t stream) {
  FA_REQUIRE(x_dev && g_dev && b_dev && M > 0, "layernorm: bad arguments");
  return launch_layernorm(x_dev, g_dev, b_dev, nullptr, nullptr, static_cast<__nv_bfloat16*>(y_bf16_dev), y_f32_dev, M,
                          D, static_cast<cudaStream_t>(stream));
}

extern "C" int fa_op_embed_layernorm(const float* x_dev, const float* inv_ts_dev, const int32_t* cu_seqlens_dev,
                                     int32_t n_utts, float scale, const float* g_dev, const float* b_dev,
                                     void* y_bf16_dev, int32_t M, int32_t D, fa_stream_t stream) {
  FA_REQUIRE(x_dev && inv_ts_dev && cu_seqlens_dev && g_dev && b_dev && y_bf16_dev && M > 0 && n_utts > 0,
             "embed_layernorm: bad arguments");
  return launch_embed_layernorm(x_dev, 0, inv_ts_dev, cu_seqlens_dev, n_utts, 0, scale, g_dev, b_dev,
                                static_cast<__nv_bfloat16*>(y_bf16_dev), M, D, static_cast<cudaStream_t>(stream));
}

extern "C" int fa_op_fsmn(const void* qkv_dev, int64_t ld_qkv, int32_t v_col, const float* w_dev,
                          const int32_t* cu_seqlens_dev, int32_t n_utts, float* mem_dev, int32_t M, int32_t D,
                          int32_t kernel, int32_t left_pad, fa_stream_t stream) {
  FA_REQUIRE(qkv_dev && w_dev && cu_seqlens_dev && mem_dev && M > 0 && n_utts > 0, "fsmn: bad arguments");
  return launch_fsmn(static_cast<const __nv_bfloat16*>(qkv_dev), ld_qkv, v_col, w_dev, cu_seqlens_dev, n_utts, 0, mem_dev, 0, D,
                     M, D, kernel, left_pad, static_cast<cudaStream_t>(stream));
}

extern "C" int fa_op_fsmn_bf16(const void* qkv_dev, int64_t ld_qkv, int32_t v_col, const float* w_dev,
                               const int32_t* cu_seqlens_dev, int32_t n_utts, void* mem_bf16_dev, int64_t ld_mem, int32_t M,
                               int32_t D, int32_t kernel, int32_t left_pad, fa_stream_t stream) {
  FA_REQUIRE(qkv_dev && w_dev && cu_seqlens_dev && mem_bf16_dev && M > 0 && n_utts > 0, "fsmn: bad arguments");
  return launch_fsmn(static_cast<const __nv_bfloat16*>(qkv_dev), ld_qkv, v_col, w_dev, cu_seqlens_dev, n_utts, 0, mem_bf16_dev, 1,
                     ld_mem, M, D, kernel, left_pad, static_cast<cudaStream_t>(stream));
}

extern "C" int fa_splice_embeds(const void* speech_dev, int32_t speech_bf16, const int32_t* cu_seqlens_dev,
                                const int32_t* take_dev, const void* prefix_dev, int32_t n_prefix,
                                const void* suffix_dev, int32_t n_suffix, const int64_t* out_off_dev, int32_t n_utts,
                                int32_t dim, void* out_dev, int32_t out_bf16, int32_t max_rows_per_utt,
                                fa_stream_t stream) {
  FA_REQUIRE(speech_dev && cu_seqlens_dev && out_off_dev && out_dev && n_utts > 0 && dim > 0, "splice: bad arguments");
  FA_REQUIRE((n_prefix == 0 || prefix_dev) && (n_suffix == 0 || suffix_dev), "splice: NULL prefix/suffix");
  FA_REQUIRE(max_rows_per_utt > 0, "splice: max_rows_per_utt must be positive");
  dim3 grid(max_rows_per_utt, n_utts);
  ProfScope _ps(K_SPLICE, static_cast<cudaStream_t>(stream));
  ew::splice_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(speech_dev, speech_bf16, cu_seqlens_dev,
                                                                         take_dev, prefix_dev, n_prefix, suffix_dev,
                                                                         n_suffix, out_off_dev, dim, out_dev, out_bf16);
  FA_CHECK_LAUNCH();
  return 0;
}
