// Variable-length multi-head self-attention over packed utterances, head dim 128, on tcgen05 / TMEM.
//   ctx = softmax(q k^T / sqrt(128)) v   per (utterance, head); keys never cross utterances, which is exactly the
//   reference's key-masked padded attention on valid rows
//   (VLLM/model_executor/models/funasr.py:222-235,250-252; adaptor :506-512).
//
// One CTA = (utterance b, head h, 128 query rows).  Keys/values stream through a 2-stage shared-memory ring in blocks of
// 64 keys (TMA, 128B swizzle); per block j
//     S_j = Q K_j^T         tcgen05.mma, A = Q (smem, K-major), B = K_j (smem, K-major), D in TMEM (two S buffers)
//     P_j = exp2(S c - m)   softmax warps: thread = row, TMEM -> registers (one pass) -> bf16 -> swizzled smem in the
//                           A-operand layout, written over K_j (dead once S_j is complete)
//     O  += P_j V_j         tcgen05.mma, A = P (smem, K-major), B = V_j (smem, MN-major: V rows are keys), D in TMEM
// software-pipelined: S_{j+1} is issued before the softmax of block j finishes, so the tensor pipe, the MUFU/ALU softmax
// and the TMA of block j+2 overlap inside one CTA; the online softmax only raises its running maximum when it grows by
// more than 2^8 (the O accumulator in TMEM is then rescaled once), so most blocks never touch O.  The CTA is lean —
// 97 KB of smem and 256 TMEM columns — so that TWO CTAs share an SM (utterances are short: T = 167 for 10 s = three
// key blocks).  Warp roles: 0 TMA producer, 1 MMA issuer (+ TMEM alloc), 2..5 softmax / epilogue (one TMEM lane quarter
// each).
// Measured alternative (kept out): one CTA per SM owning BOTH 128-row query tiles of an utterance with a 3-stage ring and
// two softmax warpgroups (K/V loaded once, all 512 TMEM columns): 138 us instead of 121 us at T = 167 -- with a single
// CTA per SM the load phase of one work item no longer overlaps the compute phase of another.
// Second measured alternative (kept out, profiles/r01_m_attention_short.md): a single-pass kernel for utterances of <= 192 keys
// (S for all keys in one MMA series, exact softmax, V loaded over Q, O over S's TMEM columns; still two CTAs per SM):
// 115.7 us instead of 118.8 us at T = 167 -- the three block rounds are not what bounds the kernel -- and it would break the
// bit-identity of the slab / whole-batch schedules (two kernels, two roundings), so the block loop stays the only path.
#include <cuda.h>

#include <cstring>

#include "common.cuh"

namespace fa {
namespace attn {

constexpr int HD = 128;   // head dim
constexpr int BQ = 128;   // query rows per CTA = UMMA M
constexpr int BKV = 64;   // keys per block
constexpr int STAGES = 2;
constexpr int THREADS = 192;
constexpr int QPANEL = BQ * 128;           // [128 rows x 64 bf16] 128B-swizzled panel = 16 KB
constexpr int KPANEL = BKV * 128;          // [64 rows x 64 bf16] panel = 8 KB
constexpr int Q_BYTES = 2 * QPANEL;        // d = 128 -> two panels
constexpr int K_BYTES = 2 * KPANEL;        // one K block; P_j (128 rows x 64 keys = one 16 KB panel) later overwrites it
constexpr int V_BYTES = 2 * KPANEL;
constexpr int SMEM_BYTES = Q_BYTES + STAGES * (K_BYTES + V_BYTES) + 1024 + 256;
constexpr int TMEM_COLS = 256;             // S0 [0,64)  S1 [64,128)  O [128,256)
constexpr float RESCALE_THRESHOLD = 8.f;   // log2 units

using namespace fa::ptx;  // shared inline-PTX wrappers (ptx.cuh)

__device__ __forceinline__ unsigned long long gtime() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// Per-CTA timeline for profiles/scripts/trace_attention.py: compiled in only with -DFA_ATTN_TRACE (make TRACE=1).
#ifdef FA_ATTN_TRACE
#define TRACE(slot)                                                                              \
  do {                                                                                           \
    if (trace && cta_lin < 512) trace[cta_lin * 16 + (slot)] = gtime();                          \
  } while (0)
#else
#define TRACE(slot) do { } while (0)
#endif

// ---------------------------------------------------------------- kernel
__global__ void __launch_bounds__(THREADS, 2)
attention_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const int32_t* __restrict__ cu, int base, int n_heads,
                 __nv_bfloat16* __restrict__ ctx, int64_t ld_ctx, float scale_log2e, unsigned long long* trace) {
  extern __shared__ unsigned char smem_raw[];
  const int cta_lin = blockIdx.x + gridDim.x * (blockIdx.y + gridDim.y * blockIdx.z);
  if (threadIdx.x == 0) TRACE(0);
  pdl_launch_dependents();
  const int b = blockIdx.z, h = blockIdx.y;  // (cu_seqlens is an input of the whole forward call, not a kernel's output)
  const int start = cu[b] - base, len = cu[b + 1] - cu[b];
  const int q0 = blockIdx.x * BQ;
  if (q0 >= len) return;  // uniform per CTA, before any barrier / TMEM use
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int dm = n_heads * HD;
  const int n_blk = (len + BKV - 1) / BKV;

  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ = sbase, sK0 = sQ + Q_BYTES, sV0 = sK0 + STAGES * K_BYTES, bars = sV0 + STAGES * V_BYTES;
  auto sK = [&](int st) { return sK0 + st * K_BYTES; };
  auto sV = [&](int st) { return sV0 + st * V_BYTES; };
  const uint32_t q_full = bars, o_full = bars + 8, tmem_slot = bars + 16;
  auto k_full = [&](int st) { return bars + 24 + 8 * st; };
  auto v_full = [&](int st) { return bars + 40 + 8 * st; };
  auto kv_empty = [&](int st) { return bars + 56 + 8 * st; };
  auto s_full = [&](int st) { return bars + 72 + 8 * st; };
  auto p_full = [&](int st) { return bars + 88 + 8 * st; };
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  auto issue_kv = [&](int j) {  // K_j and V_j into stage j & 1 (the caller has made sure the stage is free)
    const int st = j & 1;
    const int xk = dm + h * HD, xv = 2 * dm + h * HD, row = start + j * BKV;
    mbar_expect_tx(k_full(st), K_BYTES);
    tma_load_2d(sK(st), &tmap_qkv, k_full(st), xk, row);
    tma_load_2d(sK(st) + KPANEL, &tmap_qkv, k_full(st), xk + 64, row);
    mbar_expect_tx(v_full(st), V_BYTES);
    tma_load_2d(sV(st), &tmap_qkv, v_full(st), xv, row);
    tma_load_2d(sV(st) + KPANEL, &tmap_qkv, v_full(st), xv + 64, row);
  };
  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_qkv);
    mbar_init(q_full, 1);
    mbar_init(o_full, 1);
    for (int st = 0; st < STAGES; ++st) {
      mbar_init(k_full(st), 1);
      mbar_init(v_full(st), 1);
      mbar_init(kv_empty(st), 1);
      mbar_init(s_full(st), 1);
      mbar_init(p_full(st), 128);  // every softmax thread arrives
    }
    fence_mbar_init();
    // Q and the first two K/V blocks are requested right here, before the TMEM allocation and the CTA-wide sync
    // (the producer only needs its own barriers): ~0.4 us of the 1.7 us Q+K0 wait leaves the critical path
    pdl_wait();
    mbar_expect_tx(q_full, Q_BYTES);
#pragma unroll
    for (int i = 0; i < 4; ++i)  // (panel, row half): box = 64 rows x 64 columns
      tma_load_2d(sQ + (i >> 1) * QPANEL + (i & 1) * KPANEL, &tmap_qkv, q_full, h * HD + (i >> 1) * 64, start + q0 + (i & 1) * 64);
    for (int j = 0; j < min(n_blk, STAGES); ++j) issue_kv(j);
  }
  if (warp == 1) {
    tmem_alloc<TMEM_COLS>(tmem_slot);
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (threadIdx.x == 0) TRACE(1);
  const uint32_t tmem = *tmem_slot_ptr;
  pdl_wait();
  const uint32_t tO = tmem + 2 * BKV;
  auto tS = [&](int st) { return tmem + st * BKV; };

  if (warp == 0) {
    // ===================== TMA producer (blocks 0 and 1 were requested during set-up) =====================
    if (lane == 0) {
      for (int j = STAGES; j < n_blk; ++j) {
        mbar_wait(kv_empty(j & 1), ((j >> 1) & 1) ^ 1);  // P_{j-2} V_{j-2} retired: the K/P and V slots of this stage are free
        issue_kv(j);
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      auto issue_s = [&](int j) {  // S_j = Q K_j^T into S buffer j & 1
        const int st = j & 1;
        const int nkeys = min(BKV, len - j * BKV);
        tc_fence_after();
        const uint32_t idesc = make_idesc(BQ, (nkeys + 15) & ~15, 0);
#pragma unroll
        for (int k = 0; k < HD / 16; ++k)
          umma_f16(tS(st), desc_kmajor(sQ + (k >> 2) * QPANEL + (k & 3) * 32), desc_kmajor(sK(st) + (k >> 2) * KPANEL + (k & 3) * 32),
                   idesc, k != 0);
        umma_commit(s_full(st));
      };
      mbar_wait(q_full, 0);
      mbar_wait(k_full(0), 0);
      TRACE(2);
      issue_s(0);
      for (int j = 0; j < n_blk; ++j) {
        const int st = j & 1, nx = (j + 1) & 1;
        const uint32_t nx_phase = ((j + 1) >> 1) & 1;
        // S_{j+1} first if its K block has landed (it then overlaps the softmax of block j); otherwise after P_j V_j
        bool s_next_done = (j + 1 >= n_blk);
        if (!s_next_done && mbar_test(k_full(nx), nx_phase)) {
          issue_s(j + 1);
          s_next_done = true;
        }
        // O (+)= P_j V_j
        mbar_wait(p_full(st), (j >> 1) & 1);
        mbar_wait(v_full(st), (j >> 1) & 1);
        tc_fence_after();
        {
          const int nkeys = min(BKV, len - j * BKV);
          const int ksteps = (nkeys + 15) >> 4;
          const uint32_t idesc = make_idesc(BQ, HD, 1);
          for (int k = 0; k < ksteps; ++k)
            umma_f16(tO, desc_kmajor(sK(st) + k * 32), desc_mnmajor(sV(st) + k * 16 * 128, KPANEL, 1024), idesc, (j | k) != 0);
          umma_commit(kv_empty(st));
          if (j == n_blk - 1) umma_commit(o_full);
        }
        if (!s_next_done) {
          mbar_wait(k_full(nx), nx_phase);
          issue_s(j + 1);
        }
      }
    }
  } else {
    // ===================== softmax / epilogue warps: thread = query row =====================
    const int q = warp & 3;
    const int row = q * 32 + lane;                 // row within the tile == TMEM lane
    const uint32_t lane_sel = static_cast<uint32_t>(q * 32) << 16;
    const uint32_t rsw = static_cast<uint32_t>(row & 7);
    const bool row_ok = q0 + row < len;
    float m_used = -INFINITY, l_run = 0.f;
    for (int j = 0; j < n_blk; ++j) {
      const int st = j & 1;
      const int nkeys = min(BKV, len - j * BKV);
      mbar_wait(s_full(st), (j >> 1) & 1);
      tc_fence_after();
      if (threadIdx.x == 64 && j < 3) TRACE(3 + j);
      // ---- the whole S row of this block into registers (one TMEM pass); keys past the end of the utterance become
      // -inf once, so the arithmetic below is predicate-free; row maximum with four independent chains
      float r[2][32];
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        if (c * 32 < nkeys) {
          uint32_t t[32];
          tmem_ld32(tS(st) + lane_sel + c * 32, t);
#pragma unroll
          for (int i = 0; i < 32; ++i) r[c][i] = __uint_as_float(t[i]);
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i) r[c][i] = -INFINITY;
        }
      }
      if (nkeys < BKV) {
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (c * 32 + i >= nkeys) r[c][i] = -INFINITY;
      }
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
      for (int c = 0; c < 2; ++c)
#pragma unroll
        for (int i = 0; i < 32; ++i) mx4[i & 3] = fmaxf(mx4[i & 3], r[c][i]);
      const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
      const float m_blk = mx * scale_log2e;
      // ---- running maximum: only raised when it would grow by more than 2^8 (then O in TMEM is rescaled once)
      float alpha = 1.f;
      if (j == 0) {
        m_used = m_blk;
      } else if (row_ok && m_blk - m_used > RESCALE_THRESHOLD) {
        alpha = ex2(m_used - m_blk);
        m_used = m_blk;
      }
      if (__any_sync(0xffffffffu, alpha != 1.f)) {
        // O holds blocks 0..j-1: P_{j-1} V_{j-1} must have retired (its commit is this stage ring's "empty" barrier)
        mbar_wait(kv_empty((j - 1) & 1), ((j - 1) >> 1) & 1);
        tc_fence_after();
        l_run *= alpha;
#pragma unroll 1
        for (int c = 0; c < HD / 32; ++c) {
          uint32_t o[32];
          tmem_ld32(tO + lane_sel + c * 32, o);
#pragma unroll
          for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
          tmem_st32(tO + lane_sel + c * 32, o);
        }
      }
      // ---- p = exp2(s c - m), row sum, bf16 -> swizzled smem (A operand of the P V MMA), over the dead K block.
      // Rows past the end of the utterance skip the arithmetic: their P rows only feed O rows that are never stored.
      if (row_ok) {
        float rs4[4] = {0.f, 0.f, 0.f, 0.f};
        const int nch16 = (nkeys + 15) >> 4;  // 16-key MMA steps that will be issued
        const uint32_t prow = sK(st) + static_cast<uint32_t>(row) * 128u;
#pragma unroll
        for (int c = 0; c < 2; ++c) {
          if (c * 2 < nch16) {  // chunk read by at least one MMA step
            uint32_t pk[16];
#pragma unroll
            for (int i = 0; i < 32; i += 2) {
              const float p0 = ex2(fmaf(r[c][i], scale_log2e, -m_used));      // masked keys: 2^-inf = 0
              const float p1 = ex2(fmaf(r[c][i + 1], scale_log2e, -m_used));
              rs4[(i >> 1) & 3] += p0 + p1;
              pk[i >> 1] = pack_bf16x2(p0, p1);
            }
#pragma unroll
            for (int g = 0; g < 4; ++g)  // 32 keys = 64 B = four 16-byte chunks of this row
              sts128(prow + ((static_cast<uint32_t>(c * 4 + g) ^ rsw) << 4), pk[4 * g], pk[4 * g + 1], pk[4 * g + 2], pk[4 * g + 3]);
          }
        }
        l_run += (rs4[0] + rs4[1]) + (rs4[2] + rs4[3]);
      }
      fence_async_smem();
      tc_fence_before();
      mbar_arrive(p_full(st));
      if (threadIdx.x == 64 && j < 3) TRACE(6 + j);
    }
    // ---- epilogue: O / l -> bf16 -> ctx (head merge = column offset h*128)
    mbar_wait(o_full, 0);
    tc_fence_after();
    if (threadIdx.x == 64) TRACE(9);
    const float inv = 1.f / l_run;
    __nv_bfloat16* dst = ctx + static_cast<int64_t>(start + q0 + row) * ld_ctx + h * HD;
#pragma unroll 1
    for (int c = 0; c < HD / 32; ++c) {
      uint32_t o[32];
      tmem_ld32(tO + lane_sel + c * 32, o);
      if (row_ok) {
#pragma unroll
        for (int i = 0; i < 32; i += 8) {
          uint4 v;
          v.x = pack_bf16x2(__uint_as_float(o[i]) * inv, __uint_as_float(o[i + 1]) * inv);
          v.y = pack_bf16x2(__uint_as_float(o[i + 2]) * inv, __uint_as_float(o[i + 3]) * inv);
          v.z = pack_bf16x2(__uint_as_float(o[i + 4]) * inv, __uint_as_float(o[i + 5]) * inv);
          v.w = pack_bf16x2(__uint_as_float(o[i + 6]) * inv, __uint_as_float(o[i + 7]) * inv);
          *reinterpret_cast<uint4*>(dst + c * 32 + i) = v;
        }
      }
    }
  }

  if (threadIdx.x == 64) TRACE(10);
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc<TMEM_COLS>(tmem);
  }
  if (threadIdx.x == 0) {
    TRACE(11);
#ifdef FA_ATTN_TRACE
    if (trace && cta_lin < 512) {
      unsigned smid;
      asm volatile("mov.u32 %0, %smid;" : "=r"(smid));
      trace[cta_lin * 16 + 12] = smid;
    }
#endif
  }
}
unsigned long long* g_trace = nullptr;

}  // namespace attn

int attention_kv_box() { return attn::BKV; }

int launch_attention(const TensorMap2D* tm_in, const __nv_bfloat16* qkv, int64_t ld, const int32_t* cu_seqlens, int n_utts, int base,
                     int max_len, int n_heads, __nv_bfloat16* ctx, int64_t ld_ctx, int n_rows, cudaStream_t stream) {
  FA_REQUIRE(ld % 8 == 0 && ld_ctx % 8 == 0, "attention: row strides must be multiples of 8 elements");
  FA_REQUIRE((reinterpret_cast<uintptr_t>(qkv) & 15) == 0 && (reinterpret_cast<uintptr_t>(ctx) & 15) == 0,
             "attention: qkv and ctx must be 16-byte aligned");
  FA_REQUIRE(max_len > 0 && n_utts > 0 && n_heads > 0 && n_rows > 0, "attention: empty problem");
  FA_REQUIRE(ld >= 3 * n_heads * attn::HD, "attention: qkv rows hold q|k|v of %d heads x 128", n_heads);
  static SmemOptIn opt;
  if (int r = opt_in_smem(attn::attention_kernel, attn::SMEM_BYTES, opt)) return r;
  // Without a prebuilt map: rows = the valid packed rows only, so key/value rows past the end are zero-filled by TMA instead
  // of reading memory whose contents are undefined (NaN there would leak into O through 0 * NaN).
  TensorMap2D tm;
  if (tm_in == nullptr) {
    if (int r = make_tensor_map_bf16(&tm, qkv, n_rows, 3 * n_heads * attn::HD, ld, attn::BKV)) return r;
    tm_in = &tm;
  }
  CUtensorMap m;
  memcpy(&m, tm_in->bytes, sizeof(m));
  dim3 grid(ceil_div(max_len, attn::BQ), n_heads, n_utts);
  const float scale_log2e = 1.4426950408889634f / sqrtf(static_cast<float>(attn::HD));
  ProfScope _ps(K_ATTENTION, stream);
  FA_CHECK_CUDA(launch_pdl(attn::attention_kernel, grid, dim3(attn::THREADS), attn::SMEM_BYTES, stream, 1, m, cu_seqlens, base, n_heads,
                           ctx, ld_ctx, scale_log2e, attn::g_trace));
  return 0;
}

}  // namespace fa

extern "C" int fa_op_attention(const void* qkv_dev, int64_t ld_qkv, const int32_t* cu_seqlens_dev, int32_t n_utts,
                               int32_t max_len, int32_t n_heads, int32_t n_rows, void* ctx_dev, int64_t ld_ctx,
                               fa_stream_t stream) {
  using namespace fa;
  FA_REQUIRE(qkv_dev && cu_seqlens_dev && ctx_dev, "attention: NULL pointer");
  return launch_attention(nullptr, static_cast<const __nv_bfloat16*>(qkv_dev), ld_qkv, cu_seqlens_dev, n_utts, 0, max_len, n_heads,
                          static_cast<__nv_bfloat16*>(ctx_dev), ld_ctx, n_rows, static_cast<cudaStream_t>(stream));
}

#ifdef FA_ATTN_TRACE
// debug only (not in the public header): device buffer of 512 x 16 u64 receiving per-CTA %globaltimer stamps
extern "C" __attribute__((visibility("default"))) void fa_debug_attention_trace(void* buf) {
  fa::attn::g_trace = static_cast<unsigned long long*>(buf);
}
#endif
