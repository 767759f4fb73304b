// Variable-length multi-head self-attention over packed utterances, head dim 128.
//   ctx = softmax(q k^T / sqrt(128)) v   per (utterance, head); keys never cross utterances, which is
//   exactly the reference's key-masked padded attention on valid rows
//   (VLLM/model_executor/models/funasr.py:222-235,250-252; adaptor :506-512).
//
// Round-1 kernel: flash-style online softmax, 64 queries per CTA (4 warps x 16 rows), 64-key tiles
// double-buffered with cp.async, bf16 mma.sync.m16n8k16 with fp32 accumulation, XOR-swizzled smem
// read by ldmatrix.  (The tcgen05/TMEM version is the next step)
#include "common.cuh"

namespace fa {
namespace attn {

constexpr int HD = 128;  // head dim
constexpr int BQ = 64;
constexpr int BKV = 64;
constexpr int THREADS = 128;
constexpr int TILE_BYTES = BKV * HD * 2;  // 16 KB
constexpr int SMEM_BYTES = BQ * HD * 2 + 4 * TILE_BYTES;

__device__ __forceinline__ uint32_t swz(int row, int chunk) {  // byte offset of a 16B chunk
  return static_cast<uint32_t>(row * (HD * 2) + ((chunk ^ (row & 7)) << 4));
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, bool valid) {
  const int bytes = valid ? 16 : 0;
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void ldsm_x4_t(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
__device__ __forceinline__ void mma_bf16(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

// rows [row0, row0+64) of one operand (q, k or v slice of head h) -> swizzled smem tile; rows >= row_end zero-filled
__device__ __forceinline__ void load_tile(uint32_t smem_tile, const __nv_bfloat16* __restrict__ base, int64_t ld,
                                          int row0, int row_end, int tid) {
#pragma unroll
  for (int i = 0; i < (BKV * 16) / THREADS; ++i) {
    const int idx = tid + i * THREADS;
    const int r = idx >> 4, c = idx & 15;
    const int gr = row0 + r;
    const bool ok = gr < row_end;
    const __nv_bfloat16* src = base + static_cast<int64_t>(ok ? gr : row0) * ld + c * 8;
    cp_async16(smem_tile + swz(r, c), src, ok);
  }
}

__global__ void __launch_bounds__(THREADS)
attention_kernel(const __nv_bfloat16* __restrict__ qkv, int64_t ld, const int32_t* __restrict__ cu, int n_heads,
                 __nv_bfloat16* __restrict__ ctx, int64_t ld_ctx, float scale_log2e) {
  extern __shared__ __align__(128) unsigned char smem_raw[];
  const int b = blockIdx.z, h = blockIdx.y;
  const int start = cu[b], len = cu[b + 1] - start;
  const int q0 = blockIdx.x * BQ;
  if (q0 >= len) return;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int dm = n_heads * HD;

  const uint32_t sQ = smem_u32(smem_raw);
  const uint32_t sK = sQ + BQ * HD * 2;
  const uint32_t sV = sK + 2 * TILE_BYTES;
  const __nv_bfloat16* qbase = qkv + static_cast<int64_t>(start) * ld + h * HD;
  const __nv_bfloat16* kbase = qbase + dm;
  const __nv_bfloat16* vbase = qbase + 2 * dm;

  const int n_tiles = (len + BKV - 1) / BKV;
  load_tile(sQ, qbase, ld, q0, len, tid);
  load_tile(sK, kbase, ld, 0, len, tid);
  load_tile(sV, vbase, ld, 0, len, tid);
  cp_async_commit();

  uint32_t qf[8][4];
  float o[16][4];
#pragma unroll
  for (int i = 0; i < 16; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};

  for (int t = 0; t < n_tiles; ++t) {
    const int buf = t & 1;
    if (t + 1 < n_tiles) {
      load_tile(sK + (buf ^ 1) * TILE_BYTES, kbase, ld, (t + 1) * BKV, len, tid);
      load_tile(sV + (buf ^ 1) * TILE_BYTES, vbase, ld, (t + 1) * BKV, len, tid);
      cp_async_commit();
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    if (t == 0) {
#pragma unroll
      for (int kk = 0; kk < 8; ++kk)
        ldsm_x4(sQ + swz(warp * 16 + (lane & 15), 2 * kk + (lane >> 4)), qf[kk][0], qf[kk][1], qf[kk][2], qf[kk][3]);
    }
    const uint32_t sKt = sK + buf * TILE_BYTES, sVt = sV + buf * TILE_BYTES;

    // ---- S = Q K^T  (16 x 64 per warp)
    float s[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i) s[i][0] = s[i][1] = s[i][2] = s[i][3] = 0.f;
#pragma unroll
    for (int kk = 0; kk < 8; ++kk) {
#pragma unroll
      for (int nb = 0; nb < 8; nb += 2) {
        uint32_t b0, b1, b2, b3;
        const int mi = lane >> 3, r = lane & 7;
        ldsm_x4(sKt + swz((nb + (mi >> 1)) * 8 + r, 2 * kk + (mi & 1)), b0, b1, b2, b3);
        mma_bf16(s[nb], qf[kk], b0, b1);
        mma_bf16(s[nb + 1], qf[kk], b2, b3);
      }
    }

    // ---- online softmax (rows g = lane/4 and g+8), keys beyond the utterance masked to -inf
    const int kv0 = t * BKV;
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int nb = 0; nb < 8; ++nb) {
      const int key = kv0 + nb * 8 + 2 * (lane & 3);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const bool ok = (key + (e & 1)) < len;
        s[nb][e] = ok ? s[nb][e] * scale_log2e : -INFINITY;
        mx[e >> 1] = fmaxf(mx[e >> 1], s[nb][e]);
      }
    }
    float alpha[2];
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      mx[i] = fmaxf(mx[i], __shfl_xor_sync(0xffffffffu, mx[i], 1));
      mx[i] = fmaxf(mx[i], __shfl_xor_sync(0xffffffffu, mx[i], 2));
      const float m_new = fmaxf(m_run[i], mx[i]);  // finite: every tile holds >= 1 valid key
      alpha[i] = exp2f(m_run[i] - m_new);
      m_run[i] = m_new;
    }
    float rs[2] = {0.f, 0.f};
    uint32_t pf[4][4];
#pragma unroll
    for (int nb = 0; nb < 8; ++nb) {
      const float p0 = exp2f(s[nb][0] - m_run[0]), p1 = exp2f(s[nb][1] - m_run[0]);
      const float p2 = exp2f(s[nb][2] - m_run[1]), p3 = exp2f(s[nb][3] - m_run[1]);
      rs[0] += p0 + p1;
      rs[1] += p2 + p3;
      pf[nb >> 1][(nb & 1) * 2 + 0] = pack_bf16x2(p0, p1);
      pf[nb >> 1][(nb & 1) * 2 + 1] = pack_bf16x2(p2, p3);
    }
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      rs[i] += __shfl_xor_sync(0xffffffffu, rs[i], 1);
      rs[i] += __shfl_xor_sync(0xffffffffu, rs[i], 2);
      l_run[i] = l_run[i] * alpha[i] + rs[i];
    }
#pragma unroll
    for (int dn = 0; dn < 16; ++dn) {
      o[dn][0] *= alpha[0]; o[dn][1] *= alpha[0];
      o[dn][2] *= alpha[1]; o[dn][3] *= alpha[1];
    }

    // ---- O += P V
#pragma unroll
    for (int kk2 = 0; kk2 < 4; ++kk2) {
#pragma unroll
      for (int dn = 0; dn < 16; dn += 2) {
        uint32_t b0, b1, b2, b3;
        const int mi = lane >> 3, r = lane & 7;
        ldsm_x4_t(sVt + swz(kk2 * 16 + (mi & 1) * 8 + r, dn + (mi >> 1)), b0, b1, b2, b3);
        mma_bf16(o[dn], pf[kk2], b0, b1);
        mma_bf16(o[dn + 1], pf[kk2], b2, b3);
      }
    }
    __syncthreads();  // everyone done with buf before the next prefetch overwrites it
  }

  // ---- normalise and store (head merge = column offset h*128)
  const int g = lane >> 2;
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    const int row = q0 + warp * 16 + g + 8 * i;
    if (row < len) {
      const float inv = 1.f / l_run[i];
      __nv_bfloat16* dst = ctx + static_cast<int64_t>(start + row) * ld_ctx + h * HD + 2 * (lane & 3);
#pragma unroll
      for (int dn = 0; dn < 16; ++dn)
        *reinterpret_cast<uint32_t*>(dst + dn * 8) = pack_bf16x2(o[dn][2 * i] * inv, o[dn][2 * i + 1] * inv);
    }
  }
}

}  // namespace attn

int launch_attention(const __nv_bfloat16* qkv, int64_t ld, const int32_t* cu_seqlens, int n_utts, int max_len,
                     int n_heads, __nv_bfloat16* ctx, int64_t ld_ctx, cudaStream_t stream) {
  FA_REQUIRE(ld % 8 == 0 && ld_ctx % 2 == 0, "attention: qkv row stride must be a multiple of 8 elements");
  FA_REQUIRE((reinterpret_cast<uintptr_t>(qkv) & 15) == 0, "attention: qkv must be 16-byte aligned");
  FA_REQUIRE(max_len > 0 && n_utts > 0 && n_heads > 0, "attention: empty problem");
  static bool attr_done = false;
  if (!attr_done) {
    FA_CHECK_CUDA(cudaFuncSetAttribute(attn::attention_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, attn::SMEM_BYTES));
    attr_done = true;
  }
  dim3 grid(ceil_div(max_len, attn::BQ), n_heads, n_utts);
  const float scale_log2e = 1.4426950408889634f / sqrtf(static_cast<float>(attn::HD));
  ProfScope _ps(K_ATTENTION, stream);
  attn::attention_kernel<<<grid, attn::THREADS, attn::SMEM_BYTES, stream>>>(qkv, ld, cu_seqlens, n_heads, ctx, ld_ctx,
                                                                           scale_log2e);
  FA_CHECK_LAUNCH();
  return 0;
}

}  // namespace fa

extern "C" int fa_op_attention(const void* qkv_dev, int64_t ld_qkv, const int32_t* cu_seqlens_dev, int32_t n_utts,
                               int32_t max_len, int32_t n_heads, void* ctx_dev, int64_t ld_ctx, fa_stream_t stream) {
  using namespace fa;
  FA_REQUIRE(qkv_dev && cu_seqlens_dev && ctx_dev, "attention: NULL pointer");
  return launch_attention(static_cast<const __nv_bfloat16*>(qkv_dev), ld_qkv, cu_seqlens_dev, n_utts, max_len, n_heads,
                          static_cast<__nv_bfloat16*>(ctx_dev), ld_ctx, static_cast<cudaStream_t>(stream));
}
