// Variable-length multi-head self-attention over packed utterances, head dim 128, fused with the FSMN memory block of
// SAN-M, on tcgen05 / TMEM.  One persistent kernel per layer:
//   ctx = softmax(q k^T / sqrt(128)) v   per (utterance, head); keys never cross utterances, which is exactly the
//         reference's key-masked padded attention on valid rows (VLLM/model_executor/models/funasr.py:222-235,250-252;
//         adaptor :506-512)
//   mem = v + depthwise_conv_time(v)     the FSMN memory on the SAME value rows (VLLM/...:181-201), computed by two filler
//         warps of the same CTAs while the attention pipeline pulls those rows through L2, so V crosses HBM once per layer
//         and the separate FSMN pass (one launch and one more read of V per layer) is gone.
//
// Persistent CTAs (one per SM, static round-robin over work items).  A work item = (utterance b, head h, group of GQ <= 2
// query tiles of 128 rows): the K/V blocks (64 keys each) of the utterance stream ONCE through a 3-stage shared-memory ring
// (TMA, 128B swizzle) and serve both query tiles; per block j and query tile q
//     S_qj = Q_q K_j^T       tcgen05.mma, A = Q_q (smem, K-major), B = K_j (smem, K-major), D in TMEM (2 buffers per tile)
//     P_qj = exp2(S c - m)   softmax warpgroup q: thread = row, TMEM -> registers (one pass) -> bf16 -> swizzled smem in the
//                            A-operand layout (2 buffers per tile)
//     O_q += P_qj V_j        tcgen05.mma, A = P (smem, K-major), B = V_j (smem, MN-major: V rows are keys), D in TMEM
// S and P are both double-buffered, so a softmax warpgroup never waits for the P V product of the block before it: the only
// MMA results it waits for are S of its own block (issued two blocks ahead) and, on the rare block that raises the running
// maximum by more than 2^8, the P V products so far (O is then rescaled once in TMEM).  All 512 TMEM columns are used:
// S[2 tiles][2 buffers] x 64 + O[2 tiles] x 128.
// Warp roles (16 warps, four warpgroups): 0..3 FSMN, 4..7 softmax / epilogue of query tile 0, 8..11 of query tile 1, 12 TMA
// producer, 13 MMA issuer (+ TMEM alloc), 14..15 idle.  The producer runs ahead across work items (next item's Q as soon as the last S of the current one
// has been issued, K/V as stages free up), every role has the next item's cu_seqlens loads in flight while it works on the
// current one, and the epilogue of one item (O -> ctx) overlaps the first S MMAs of the next: nothing is set up or torn
// down per work item.
#include <cuda.h>

#include <cstring>

#include "common.cuh"

namespace fa {
namespace attn {

using namespace fa::ptx;

constexpr int HD = 128;   // head dim
constexpr int BQ = 128;   // query rows per tile = UMMA M
constexpr int BKV = 64;   // keys per block
constexpr int NST = 3;    // K/V stages
constexpr int NQ = 2;     // query tiles per work item (at most)
// Four warpgroups (setmaxnreg works on aligned groups of four warps): 0 = FSMN filler warps, 1 and 2 = softmax / epilogue of
// query tiles 0 and 1 (TMEM lane quarter = warp id mod 4), 3 = TMA producer, MMA issuer and two idle warps.  The kernel is
// compiled for 128 registers per thread (512 threads); the control group hands most of its registers back (56 each), the softmax
// groups take 152 (a 64-key S row lives in registers) and the FSMN group 152 (taps, accumulators and an 11-row ring): no spills.  The FSMN walk is bound by the bytes one warp can
// keep in flight (11 rows x 256 B): two warps per CTA took as long as the whole attention pipeline (fused 113 us against 72 us
// for attention alone at 1 h of 10 s utterances), hence four.
constexpr int FSMN_WARPS = 4;   // each takes every FSMN_WARPS-th FSMN unit of its CTA
constexpr int THREADS = 16 * 32;
#ifndef FA_ATTN_RS
#define FA_ATTN_RS 152
#define FA_ATTN_RC 56
#define FA_ATTN_RF 152
#endif
constexpr int REGS_SOFTMAX = FA_ATTN_RS, REGS_CONTROL = FA_ATTN_RC, REGS_FSMN = FA_ATTN_RF;
static_assert(128 * REGS_FSMN + 256 * REGS_SOFTMAX + 128 * REGS_CONTROL <= 65536, "attention: register split exceeds the register file");  // 128 x 136 + 256 x 168 + 128 x 40 = the whole register file
// Warp ids by role.  The issue arbiter of a sub-partition favours its higher warp ids: the two latency-critical single-lane
// roles sit on top, the FSMN filler warps at the bottom (they only take slots nobody else wants); the softmax warps keep
// ids 4..11 (TMEM lane quarter = warp id mod 4).
constexpr int W_FSMN = 0, W_SOFTMAX = 4, W_PRODUCER = 12, W_MMA = 13;
constexpr int FSMN_K = 11;                 // taps compiled into the fused path (the model's kernel_size)
constexpr int QPANEL = BQ * 128;           // [128 rows x 64 bf16] 128B-swizzled panel = 16 KB
constexpr int KPANEL = BKV * 128;          // [64 rows x 64 bf16] panel = 8 KB
constexpr int QT_BYTES = 2 * QPANEL;       // one query tile: d = 128 -> two panels
constexpr int K_BYTES = 2 * KPANEL;
constexpr int V_BYTES = 2 * KPANEL;
constexpr int O_BYTES = 2 * QPANEL;        // epilogue staging of one query tile: 128 rows x 128 bf16 = two swizzled panels
constexpr int BAR_BYTES = 512;
constexpr int SMEM_BYTES = NQ * QT_BYTES + NST * (K_BYTES + V_BYTES) + NQ * O_BYTES + 1024 + BAR_BYTES;
constexpr int TMEM_COLS = 512;             // S[q][buf] at (2q + buf) * 64, O[q] at 256 + 128 q
constexpr float RESCALE_THRESHOLD = 8.f;   // log2 units
static_assert(SMEM_BYTES <= 227 * 1024, "attention: shared memory budget");

struct Params {
  const int32_t* cu;
  int base;       // absolute packed row of local row 0 of qkv / ctx
  int n_utts, n_heads;
  int groups;     // query-tile groups per utterance (ceil(max_len / (128 GQ)))
  int gq;         // query tiles per group: 1 or 2
  __nv_bfloat16* ctx;
  long long ld_ctx;
  float scale_log2e;
  // fused FSMN (nullptr: plain attention, e.g. the adaptor)
  const __nv_bfloat16* qkv;  // local row 0 (the FSMN warps read the value rows directly)
  long long ld_qkv;
  const float* fsmn_w;  // f32 [FSMN_K, dm] tap-major
  int fsmn_left;
  __nv_bfloat16* mem;   // bf16 [rows, ld_mem], column h * 128 + c
  long long ld_mem;
  float* acc;           // alternative destination: the f32 residual stream, acc[row, h * 128 + c] += memory (one vector red per
  long long ld_acc;     // lane and row, added in L2; the out-projection then needs no memory operand at all)
  int fsmn_rows, fsmn_chunks;  // FSMN work unit = fsmn_rows output rows of one (utterance, head); chunks = ceil(max_len / rows)
  // housekeeping for the launch that follows: the dependency counters of the GEMM chain (gemm_chain.cu) must be zero on its entry
  int* zero_ints;
  int n_zero;
};

struct Item {  // decoded identically by every warp role
  int h, start, len, q_row0, nq, nb;
};

// cu_seqlens[b], cu_seqlens[b + 1] of work item i (an input of the forward call, not a predecessor's output).  Every role
// issues these loads one item early, so their latency is spent under the current item.
__device__ __forceinline__ void fetch_bounds(const Params& p, int i, int n_items, int& c0, int& c1) {
  if (i < n_items) {
    const int b = i / (p.n_heads * p.groups);
    c0 = __ldg(p.cu + b);
    c1 = __ldg(p.cu + b + 1);
  }
}

__device__ __forceinline__ bool decode_item(const Params& p, int i, int c0, int c1, Item& it) {
  it.h = i % p.n_heads;
  const int g = (i / p.n_heads) % p.groups;
  it.start = c0 - p.base;
  it.len = c1 - c0;
  it.q_row0 = g * p.gq * BQ;
  if (it.q_row0 >= it.len) return false;
  it.nq = min(p.gq, (it.len - it.q_row0 + BQ - 1) / BQ);
  it.nb = (it.len + BKV - 1) / BKV;
  return true;
}

// Runs body(it) for every work item of this CTA (static round-robin); items whose query group lies past the end of the
// utterance are skipped by every role alike.
template <typename Body>
__device__ __forceinline__ void for_items(const Params& p, int n_items, Body&& body) {
  int c0n = 0, c1n = 0;
  fetch_bounds(p, blockIdx.x, n_items, c0n, c1n);
  for (int i = blockIdx.x; i < n_items; i += gridDim.x) {
    Item it;
    const bool ok = decode_item(p, i, c0n, c1n, it);
    fetch_bounds(p, i + gridDim.x, n_items, c0n, c1n);
    if (ok) body(it);
  }
}

// ---------------------------------------------------------------- pipeline trace (make TRACE=1: profiling builds only)
#ifdef FA_ATTN_TRACE
// SM clock | event | item | block/tile, CTA 0 only, lane 0 of a warp.  One region per warp and a counter in a register: plain
// stores, no atomics (an atomicAdd per event put a ~700-cycle L2 round trip into the role that was being timed).
constexpr int TRACE_PER_WARP = 4096;
__device__ unsigned long long g_trace[16 * TRACE_PER_WARP];
__device__ __forceinline__ void trace(int& tn, int ev, int a, int b) {
  if (blockIdx.x == 0 && (threadIdx.x & 31) == 0 && tn < TRACE_PER_WARP) {
    g_trace[(threadIdx.x >> 5) * TRACE_PER_WARP + tn] =
        (static_cast<unsigned long long>(clock64()) << 24) | ((ev & 0xff) << 16) | ((a & 0xff) << 8) | (b & 0xff);
    ++tn;
  }
}
#define FA_TRACE(ev, a, b) trace(tn, ev, a, b)
#else
#define FA_TRACE(ev, a, b)
#endif

// ---------------------------------------------------------------- kernel
// ACC: the FSMN memory is added into the f32 residual stream (p.acc) instead of being stored as bf16 (p.mem).  Compile-time:
// with both stores in one instruction stream the walk over the value rows lost its load pipelining (141 against 108 us per
// layer at 1 h of 10 s utterances, profiles/r02/attn_r2af.txt).
template <bool ACC>
__global__ void __launch_bounds__(THREADS, 1)
attention_kernel(const __grid_constant__ CUtensorMap tmap_qkv, const __grid_constant__ CUtensorMap tmap_ctx, const Params p) {
  extern __shared__ unsigned char smem_raw[];
  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
#ifdef FA_ATTN_TRACE
  int tn = 0;
#endif
  const int dm = p.n_heads * HD;
  const int n_items = p.n_utts * p.groups * p.n_heads;

  const uint32_t sbase = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sQ0 = sbase, sK0 = sQ0 + NQ * QT_BYTES, sV0 = sK0 + NST * K_BYTES, sO0 = sV0 + NST * V_BYTES;
  const uint32_t bars = sO0 + NQ * O_BYTES;
  auto sQ = [&](int q) { return sQ0 + q * QT_BYTES; };
  auto sK = [&](int st) { return sK0 + st * K_BYTES; };
  auto sV = [&](int st) { return sV0 + st * V_BYTES; };
  auto sO = [&](int q) { return sO0 + q * O_BYTES; };
  const uint32_t q_full = bars, q_empty = bars + 8, tmem_slot = bars + 16;
  auto k_full = [&](int st) { return bars + 32 + 8 * st; };
  auto v_full = [&](int st) { return bars + 64 + 8 * st; };
  auto kv_empty = [&](int st) { return bars + 96 + 8 * st; };
  auto s_full = [&](int q, int buf) { return bars + 128 + 8 * (2 * q + buf); };   // S_q[buf] written by the MMA
  auto p_full = [&](int q, int buf) { return bars + 160 + 8 * (2 * q + buf); };   // P_q[buf] written by the softmax (and S_q[buf] read)
  auto pv_done = [&](int q, int buf) { return bars + 192 + 8 * (2 * q + buf); };  // the P V product that read P_q[buf] retired
  static_assert(224 <= BAR_BYTES, "barrier region");
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));
  const bool fsmn_on = p.fsmn_w != nullptr;

  if (warp == W_PRODUCER && lane == 0) {
    prefetch_tmap(&tmap_qkv);
    prefetch_tmap(&tmap_ctx);
    mbar_init(q_full, 1);
    mbar_init(q_empty, 1);
    for (int st = 0; st < NST; ++st) {
      mbar_init(k_full(st), 1);
      mbar_init(v_full(st), 1);
      mbar_init(kv_empty(st), 1);  // the last P V of the block retired
    }
    for (int q = 0; q < NQ; ++q) {
      for (int buf = 0; buf < 2; ++buf) {
        mbar_init(s_full(q, buf), 1);
        mbar_init(p_full(q, buf), 128);  // every softmax thread of the tile arrives
        mbar_init(pv_done(q, buf), 1);
      }
    }
    fence_mbar_init();
  }
  if (warp == W_MMA) tmem_alloc<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = *tmem_slot_ptr;
  auto tS = [&](int q, int buf) { return tmem + (2 * q + buf) * BKV; };
  auto tO = [&](int q) { return tmem + 4 * BKV + q * HD; };
  pdl_wait();  // qkv comes from the previous kernel
  for (int i = blockIdx.x * THREADS + threadIdx.x; i < p.n_zero; i += gridDim.x * THREADS) p.zero_ints[i] = 0;

  if (warp >= W_PRODUCER) {
  asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(REGS_CONTROL));  // the whole control warpgroup, idle warps included
  if (warp == W_PRODUCER) {
    // ===================== TMA producer =====================
    // (producer and MMA warps run their loops with all 32 lanes - waits and counters are warp-uniform - and elect one lane
    // only around the TMA / MMA instructions: inside a lane-0-only region the compiler has to wrap every such instruction
    // in an elect loop and feed its uniform-register operands through R2UR, ~12 instructions per MMA)
    int n_it = 0, gk = 0;
    for_items(p, n_items, [&](const Item& it) {
      if (n_it > 0) mbar_wait(q_empty, (n_it - 1) & 1);  // every S MMA of the previous item has read Q
      FA_TRACE(1, n_it, 0);
      if (elect_one()) {
        mbar_expect_tx(q_full, it.nq * QT_BYTES);
        for (int q = 0; q < it.nq; ++q)
#pragma unroll
          for (int x = 0; x < 4; ++x)  // (panel, row half): box = 64 rows x 64 columns
            tma_load_2d(sQ(q) + (x >> 1) * QPANEL + (x & 1) * KPANEL, &tmap_qkv, q_full, it.h * HD + (x >> 1) * 64,
                        it.start + it.q_row0 + q * BQ + (x & 1) * 64);
      }
      __syncwarp();
      const int xk = dm + it.h * HD, xv = 2 * dm + it.h * HD;
      for (int j = 0; j < it.nb; ++j, ++gk) {
        const int st = gk % NST;
        if (gk >= NST) mbar_wait(kv_empty(st), ((gk / NST) - 1) & 1);
        const int row = it.start + j * BKV;
        FA_TRACE(2, n_it, j);
        if (elect_one()) {
          mbar_expect_tx(k_full(st), K_BYTES);
          tma_load_2d(sK(st), &tmap_qkv, k_full(st), xk, row);
          tma_load_2d(sK(st) + KPANEL, &tmap_qkv, k_full(st), xk + 64, row);
          mbar_expect_tx(v_full(st), V_BYTES);
          tma_load_2d(sV(st), &tmap_qkv, v_full(st), xv, row);
          tma_load_2d(sV(st) + KPANEL, &tmap_qkv, v_full(st), xv + 64, row);
        }
        __syncwarp();
      }
      ++n_it;
    });
  } else if (warp == W_MMA) {
    // ===================== MMA issuer =====================
    int n_it = 0, gk0 = 0;
    // low words of the shared-memory descriptors of Q tile 0, K / V stage 0 and P buffer (0, 0); the high word is the same for
    // all of them (SBO 1024 B, version 1, 128B swizzle)
    const uint32_t q_lo = static_cast<uint32_t>(desc_kmajor(sQ(0))), k_lo = static_cast<uint32_t>(desc_kmajor(sK(0)));
    const uint32_t v_lo = static_cast<uint32_t>(desc_mnmajor(sV(0), KPANEL, 1024));
    constexpr uint32_t desc_hi = static_cast<uint32_t>((static_cast<uint64_t>(1024 >> 4) << 32 | (1ull << 46) | (2ull << 61)) >> 32);
    auto mk_desc = [&](uint32_t lo) { return (static_cast<uint64_t>(desc_hi) << 32) | lo; };
    int cs0 = 0, cs1 = 0;  // S blocks issued per query tile
    int cp0 = 0, cp1 = 0;  // P V blocks issued per query tile
    for_items(p, n_items, [&](const Item& it) {
      auto stage_of = [&](int j) { return (gk0 + j) % NST; };
      auto phase_of = [&](int j) { return static_cast<uint32_t>(((gk0 + j) / NST) & 1); };
      auto issue_s = [&](int q, int j) {  // S_qj = Q_q K_j^T into S buffer cs & 1
        int& cs = q ? cs1 : cs0;
        if (elect_one()) {
          const int st = stage_of(j);
          const int nkeys = min(BKV, it.len - j * BKV);
          const uint32_t idesc = make_idesc(BQ, (nkeys + 15) & ~15, 0);
          const uint32_t d = tS(q, cs & 1);
          // descriptors: only the 14-bit address field (bytes >> 4) changes from one MMA to the next, so each one is a 32-bit
          // add on a precomputed low word (the timeline showed the issuing thread, not the tensor pipe, bounding the kernel:
          // ~85 cycles per S MMA and ~185 per P V MMA went into assembling 64-bit descriptors in a dependent chain)
          const uint32_t aq = q_lo + static_cast<uint32_t>(q) * (QT_BYTES >> 4), bk = k_lo + static_cast<uint32_t>(st) * (K_BYTES >> 4);
#pragma unroll
          for (int k = 0; k < HD / 16; ++k)
            umma_f16(d, mk_desc(aq + (k >> 2) * (QPANEL >> 4) + (k & 3) * 2), mk_desc(bk + (k >> 2) * (KPANEL >> 4) + (k & 3) * 2), idesc, k != 0);
          umma_commit(s_full(q, cs & 1));
          if (j == it.nb - 1 && q == it.nq - 1) umma_commit(q_empty);  // the producer may overwrite Q
        }
        __syncwarp();
        FA_TRACE(3, n_it, 2 * j + q);
        ++cs;
      };
      FA_TRACE(11, n_it, 0);
      mbar_wait(q_full, n_it & 1);
      FA_TRACE(12, n_it, 0);
      for (int j = 0; j < min(2, it.nb); ++j) {
        mbar_wait(k_full(stage_of(j)), phase_of(j));
        tc_fence_after();
        for (int q = 0; q < it.nq; ++q) issue_s(q, j);
      }
      for (int j = 0; j < it.nb; ++j) {
        const int st = stage_of(j);
        mbar_wait(v_full(st), phase_of(j));
        const int nkeys = min(BKV, it.len - j * BKV);
        const int ksteps = (nkeys + 15) >> 4;
        for (int q = 0; q < it.nq; ++q) {
          int& cp = q ? cp1 : cp0;
          mbar_wait(p_full(q, cp & 1), (cp >> 1) & 1);
          // (j == 0 overwrites O_q: the previous item's epilogue has read it, because the same threads read O, waited for
          // the TMEM loads and only then produced the P this MMA waits for)
          tc_fence_after();
          if (elect_one()) {
            constexpr uint32_t idesc = make_idesc(BQ, HD, 1);
            const uint32_t bv = v_lo + static_cast<uint32_t>(st) * (V_BYTES >> 4);
            const uint32_t d = tO(q);
#pragma unroll
            for (int k = 0; k < BKV / 16; ++k)  // O_q (+)= P_qj V_j
              if (k < ksteps) {
                // A = P: 16 keys = 8 columns of packed bf16 pairs, over the first half of the S buffer the softmax read
                umma_f16_ts(d, tS(q, cp & 1) + k * 8, mk_desc(bv + k * (16 * 128 >> 4)), idesc, (j | k) != 0);
              }
            umma_commit(pv_done(q, cp & 1));
            if (q == it.nq - 1) umma_commit(kv_empty(st));  // K_j / V_j are free once everything issued so far has retired
          }
          __syncwarp();
          FA_TRACE(4, n_it, 2 * j + q);
          ++cp;
        }
        if (j + 2 < it.nb) {        // S two blocks ahead: its buffer was read by the softmax of block j (p_full above)
          mbar_wait(k_full(stage_of(j + 2)), phase_of(j + 2));
          tc_fence_after();
          for (int q = 0; q < it.nq; ++q) issue_s(q, j + 2);
        }
      }
      FA_TRACE(16, n_it, 0);
      gk0 += it.nb;
      ++n_it;
    });
  }
  } else if (warp >= W_SOFTMAX) {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(REGS_SOFTMAX));
    // ===================== softmax / epilogue warpgroups: thread = query row of tile q =====================
    const int q = (warp - W_SOFTMAX) >> 2;
    const int quarter = warp & 3;                   // TMEM lane quarter this warp may access (hardware: warp id mod 4)
    const int row = quarter * 32 + lane;            // row within the tile == TMEM lane
    const uint32_t lane_sel = static_cast<uint32_t>(quarter * 32) << 16;
    const uint32_t rsw = static_cast<uint32_t>(row & 7);
    int c = 0;     // blocks this warpgroup has processed (buffer / phase counter)
    for_items(p, n_items, [&](const Item& it) {
      if (q >= it.nq) return;
      const int q0 = it.q_row0 + q * BQ;
      const bool row_ok = q0 + row < it.len;
      float m_used = -INFINITY, l_run = 0.f;
      for (int j = 0; j < it.nb; ++j, ++c) {
        const int buf = c & 1;
        const uint32_t par = static_cast<uint32_t>(c >> 1) & 1u;
        const int nkeys = min(BKV, it.len - j * BKV);
        if ((threadIdx.x & 127) == 64) FA_TRACE(5, c, q);
        mbar_wait(s_full(q, buf), par);
        if ((threadIdx.x & 127) == 64) FA_TRACE(6, c, q);
        tc_fence_after();
        // ---- the whole S row of this block into registers (one TMEM pass); keys past the end of the utterance become
        // -inf once, so the arithmetic below is predicate-free
        float r[2][32];
        {
          uint32_t t0[32], t1[32];
          tmem_ld32_nowait(tS(q, buf) + lane_sel, t0);
          if (nkeys > 32) tmem_ld32_nowait(tS(q, buf) + lane_sel + 32, t1);
          tmem_ld_wait();
#pragma unroll
          for (int x = 0; x < 32; ++x) {
            r[0][x] = __uint_as_float(t0[x]);
            r[1][x] = nkeys > 32 ? __uint_as_float(t1[x]) : -INFINITY;
          }
        }
        if (nkeys < BKV) {
#pragma unroll
          for (int h2 = 0; h2 < 2; ++h2)
#pragma unroll
            for (int x = 0; x < 32; ++x)
              if (h2 * 32 + x >= nkeys) r[h2][x] = -INFINITY;
        }
        float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
#pragma unroll
        for (int h2 = 0; h2 < 2; ++h2)
#pragma unroll
          for (int x = 0; x < 32; ++x) mx4[x & 3] = fmaxf(mx4[x & 3], r[h2][x]);
        const float mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        const float m_blk = mx * p.scale_log2e;
        // ---- running maximum: only raised when it would grow by more than 2^8 (then O in TMEM is rescaled once)
        float alpha = 1.f;
        if (j == 0) {
          m_used = m_blk;
        } else if (row_ok && m_blk - m_used > RESCALE_THRESHOLD) {
          alpha = ex2(m_used - m_blk);
          m_used = m_blk;
        }
        if (__any_sync(0xffffffffu, alpha != 1.f)) {
          // O_q holds blocks 0..j-1: the P V product of the block before this one (and with it every earlier one) must have
          // retired before O is read, scaled and written back; the MMA thread only issues this block's P V after p_full.
          mbar_wait(pv_done(q, buf ^ 1), static_cast<uint32_t>((c - 1) >> 1) & 1u);
          tc_fence_after();
          l_run *= alpha;
#pragma unroll 1
          for (int x = 0; x < HD / 32; ++x) {
            uint32_t o[32];
            tmem_ld32(tO(q) + lane_sel + x * 32, o);
#pragma unroll
            for (int y = 0; y < 32; ++y) o[y] = __float_as_uint(__uint_as_float(o[y]) * alpha);
            tmem_st32(tO(q) + lane_sel + x * 32, o);
          }
        }
        // ---- p = exp2(s c - m), row sum, bf16 pairs -> tensor memory, over the S columns this thread has just read: the A
        // operand of the P V MMA (M = 128 lanes = query rows, 2 keys per 32-bit column).  The tensor pipe runs its MMAs in order,
        // so the S MMA that next writes these columns (block c + 2) comes after the P V product that reads them.  No smem P
        // buffer, no proxy fence; the P V MMA reads only V from shared memory (its operand traffic was what paced it).
        {
          uint32_t pk[32];
          float rs4[4] = {0.f, 0.f, 0.f, 0.f};
          const int nch16 = (nkeys + 15) >> 4;  // 16-key MMA steps that will be issued
#pragma unroll
          for (int h2 = 0; h2 < 2; ++h2) {
            if (row_ok && h2 * 2 < nch16) {
#pragma unroll
              for (int x = 0; x < 32; x += 2) {
                const float p0 = ex2(fmaf(r[h2][x], p.scale_log2e, -m_used));      // masked keys: 2^-inf = 0
                const float p1 = ex2(fmaf(r[h2][x + 1], p.scale_log2e, -m_used));
                rs4[(x >> 1) & 3] += p0 + p1;
                pk[h2 * 16 + (x >> 1)] = pack_bf16x2(p0, p1);
              }
            } else {
#pragma unroll
              for (int x = 0; x < 16; ++x) pk[h2 * 16 + x] = 0u;
            }
          }
          l_run += (rs4[0] + rs4[1]) + (rs4[2] + rs4[3]);
          tmem_st32(tS(q, buf) + lane_sel, pk);  // (warp-wide: rows past the end of the utterance store zeros)
        }
        tc_fence_before();
        mbar_arrive(p_full(q, buf));
        if ((threadIdx.x & 127) == 64) FA_TRACE(7, c, q);
      }
      // ---- epilogue: O / l -> bf16 -> ctx (head merge = column offset h*128)
      mbar_wait(pv_done(q, (c - 1) & 1), static_cast<uint32_t>((c - 1) >> 1) & 1u);
      tc_fence_after();
      if ((threadIdx.x & 127) == 64) FA_TRACE(8, c, q);
      const float inv = 1.f / l_run;
      // A warp whose 32 rows all belong to the utterance stages them in shared memory (its 32-row slice of the tile's two
      // swizzled panels) and stores them with two TMA boxes; a 16-byte store per lane with a 2 KB row pitch is 32 LSU
      // wavefronts per instruction, and the timeline showed this epilogue taking 3.3 k cycles per tile.  The ragged last
      // rows of an utterance keep the per-lane stores (a box must not cover the next utterance's rows).
      const bool box_rows = q0 + quarter * 32 + 32 <= it.len;  // warp-uniform
      const uint32_t so = sO(q) + static_cast<uint32_t>(row) * 128u;
      __nv_bfloat16* dst = p.ctx + static_cast<long long>(it.start + q0 + row) * p.ld_ctx + it.h * HD;
      if (lane == 0) bulk_wait_read<0>();  // the boxes of the previous item have left this warp's staging rows
      __syncwarp();
#pragma unroll 1
      for (int x = 0; x < 2; ++x) {
        uint32_t o0[32], o1[32];
        tmem_ld32_nowait(tO(q) + lane_sel + x * 64, o0);
        tmem_ld32_nowait(tO(q) + lane_sel + x * 64 + 32, o1);
        tmem_ld_wait();
#pragma unroll
        for (int hh = 0; hh < 2; ++hh) {
          const uint32_t(&o)[32] = hh ? o1 : o0;
#pragma unroll
          for (int y = 0; y < 32; y += 8) {
            uint4 v;
            v.x = pack_bf16x2(__uint_as_float(o[y]) * inv, __uint_as_float(o[y + 1]) * inv);
            v.y = pack_bf16x2(__uint_as_float(o[y + 2]) * inv, __uint_as_float(o[y + 3]) * inv);
            v.z = pack_bf16x2(__uint_as_float(o[y + 4]) * inv, __uint_as_float(o[y + 5]) * inv);
            v.w = pack_bf16x2(__uint_as_float(o[y + 6]) * inv, __uint_as_float(o[y + 7]) * inv);
            if (box_rows) sts128(so + x * QPANEL + ((static_cast<uint32_t>(hh * 4 + (y >> 3)) ^ rsw) << 4), v.x, v.y, v.z, v.w);
            else if (row_ok) *reinterpret_cast<uint4*>(dst + x * 64 + hh * 32 + y) = v;
          }
        }
      }
      if (box_rows) {
        fence_async_smem();
        __syncwarp();
        if (lane == 0) {
          const uint32_t sw = sO(q) + static_cast<uint32_t>(quarter * 32) * 128u;
          tma_store_2d(&tmap_ctx, sw, it.h * HD, it.start + q0 + quarter * 32);
          tma_store_2d(&tmap_ctx, sw + QPANEL, it.h * HD + 64, it.start + q0 + quarter * 32);
          bulk_commit();
        }
      }
      if ((threadIdx.x & 127) == 64) FA_TRACE(15, c, q);
    });
    if (lane == 0) bulk_wait<0>();  // this warp's last context boxes are in global memory before the CTA retires
  } else if (fsmn_on) {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(REGS_FSMN));
    // ===================== FSMN warps: mem[t, c] = v[t, c] + sum_k w[k, c] v[t + k - left, c] ==============================
    // Four filler warps per CTA that share NOTHING with the attention pipeline (no barrier, no shared memory): they stream
    // value rows straight from global memory - the same rows the TMA producers pull for P V at about the same time, so one
    // of the two reads hits L2 and V still crosses HBM once per layer - using issue slots the softmax warps leave idle.
    // lane = 4 adjacent channels of the head's 128 (one 8-byte load per row); a unit's rows are walked ONCE, in order: an
    // input row is scattered into the K outputs it feeds (K independent packed FMAs per channel pair into a ring of K
    // accumulators, slot = (walk index - k) mod K, so the unrolled-by-K loop indexes registers statically); the output that
    // has then seen its last tap is stored.  Every output sums its taps in the order k = 0 .. K-1 with fma, the same
    // arithmetic as the stand-alone kernel.  Rows outside the utterance enter as zeros (VLLM/...:181-201: mask,
    // ConstantPad1d, depthwise Conv1d, + inputs, mask).
    constexpr int K = FSMN_K;
    const int f = warp - W_FSMN;
    const int left = p.fsmn_left;
    // FSMN work units are independent of the attention work items: (utterance, head, chunk of fsmn_rows output rows), dealt
    // round-robin to every FSMN warp of the grid, so that a small batch still spreads over all of them.
    const int n_units = p.n_utts * p.n_heads * p.fsmn_chunks;
    // Per unit the timeline showed 4-7 k cycles before the first row (utterance bounds, then the taps, then the ring's first
    // loads: three dependent round trips).  The bounds of the next unit are requested a unit ahead, and the taps are only
    // reloaded when the head changes (with a unit stride that is a multiple of the head count it never does).
    float4 w4[K];
    int w_head = -1;
    const int u_first = blockIdx.x * FSMN_WARPS + f, u_step = gridDim.x * FSMN_WARPS;
    int c0n = 0, c1n = 0;
    if (u_first < n_units) {
      const int b = u_first / (p.n_heads * p.fsmn_chunks);
      c0n = __ldg(p.cu + b);
      c1n = __ldg(p.cu + b + 1);
    }
    for (int u = u_first; u < n_units; u += u_step) {
      const int h = u % p.n_heads, c = (u / p.n_heads) % p.fsmn_chunks;
      const int c0 = c0n, c1 = c1n;
      if (u + u_step < n_units) {
        const int b = (u + u_step) / (p.n_heads * p.fsmn_chunks);
        c0n = __ldg(p.cu + b);
        c1n = __ldg(p.cu + b + 1);
      }
      const int start = c0 - p.base, len = c1 - c0;
      const int t0 = c * p.fsmn_rows;  // output rows [t0, t1) of the utterance
      if (t0 >= len) continue;
      const int t1 = min(len, t0 + p.fsmn_rows);
      float2 a01[K], a23[K];
      if (h != w_head) {
        w_head = h;
#pragma unroll
        for (int k = 0; k < K; ++k) {
          w4[k] = __ldg(reinterpret_cast<const float4*>(p.fsmn_w + static_cast<long long>(k) * dm + h * HD) + lane);
          const float one = (k == left) ? 1.f : 0.f;  // "+ inputs" = tap `left` with weight + 1
          w4[k].x += one; w4[k].y += one; w4[k].z += one; w4[k].w += one;
        }
      }
#pragma unroll
      for (int k = 0; k < K; ++k) {
        a01[k] = make_float2(0.f, 0.f);
        a23[k] = make_float2(0.f, 0.f);
      }
      // walk i = 0 .. total-1 reads input row t0 - left + i and completes output row t0 + i - (K - 1)
      const int r0 = t0 - left, total = (t1 - t0) + K - 1;
      const char* vnext = reinterpret_cast<const char*>(p.qkv + static_cast<long long>(start + r0) * p.ld_qkv + 2 * dm + h * HD + 4 * lane);
      constexpr bool to_acc = ACC;
      char* mrow = to_acc ? reinterpret_cast<char*>(p.acc + static_cast<long long>(start + t0 - (K - 1)) * p.ld_acc + h * HD + 4 * lane)
                          : reinterpret_cast<char*>(p.mem + static_cast<long long>(start + t0 - (K - 1)) * p.ld_mem + h * HD + 4 * lane);
      const long long vstep = p.ld_qkv * 2, mstep = to_acc ? p.ld_acc * 4 : p.ld_mem * 2;
      auto load_row = [&](int i) -> uint2 {  // next value row in walking order (zeros outside the utterance)
        uint2 v = make_uint2(0u, 0u);
        if (static_cast<unsigned>(r0 + i) < static_cast<unsigned>(len)) v = __ldg(reinterpret_cast<const uint2*>(vnext));
        vnext += vstep;
        return v;
      };
      auto row_step = [&](const uint2 raw, int x, bool store) {  // x = walk index mod K (compile-time after unrolling)
        const float2 v01 = make_float2(__uint_as_float(raw.x << 16), __uint_as_float(raw.x & 0xffff0000u));
        const float2 v23 = make_float2(__uint_as_float(raw.y << 16), __uint_as_float(raw.y & 0xffff0000u));
        // tap 0 opens a new output in the slot the previous row has just retired: fma(w, v, 0), no reset needed.
        // Packed f32x2 FMAs: measured faster here than scalar FFMA (106.6 vs 114.8 us per layer at 1 h of 10 s utterances).
        a01[x] = __ffma2_rn(make_float2(w4[0].x, w4[0].y), v01, make_float2(0.f, 0.f));
        a23[x] = __ffma2_rn(make_float2(w4[0].z, w4[0].w), v23, make_float2(0.f, 0.f));
#pragma unroll
        for (int k = 1; k < K; ++k) {
          const int sl = (x - k + K) % K;
          a01[sl] = __ffma2_rn(make_float2(w4[k].x, w4[k].y), v01, a01[sl]);
          a23[sl] = __ffma2_rn(make_float2(w4[k].z, w4[k].w), v23, a23[sl]);
        }
        const int sd = (x + 1) % K;  // this output has now seen tap K - 1
        if (store) {
          if (to_acc) red_add_v4_f32(reinterpret_cast<float*>(mrow), a01[sd].x, a01[sd].y, a23[sd].x, a23[sd].y);
          else *reinterpret_cast<uint2*>(mrow) = make_uint2(pack_bf16x2(a01[sd].x, a01[sd].y), pack_bf16x2(a23[sd].x, a23[sd].y));
        }
        mrow += mstep;
      };
      // a ring of K rows in registers: row i + K is requested as soon as row i has been consumed, so K - 1 rows of
      // arithmetic cover each load's latency
      uint2 ring[K];
#pragma unroll
      for (int x = 0; x < K; ++x) ring[x] = load_row(x);
      for (int ib = 0; ib < total; ib += K) {
        if (ib >= K - 1 && ib + K <= total) {  // interior: every row is walked and every output exists
#pragma unroll
          for (int x = 0; x < K; ++x) {
            const uint2 raw = ring[x];
            ring[x] = load_row(ib + K + x);
            row_step(raw, x, true);
          }
        } else {
#pragma unroll
          for (int x = 0; x < K; ++x) {
            const uint2 raw = ring[x];
            ring[x] = load_row(ib + K + x);
            if (ib + x < total) row_step(raw, x, ib + x >= K - 1);
          }
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == W_MMA) {
    tc_fence_after();
    tmem_dealloc<TMEM_COLS>(tmem);
  }
}

}  // namespace attn

int attention_kv_box() { return attn::BKV; }

// Fused FSMN (fsmn_w != NULL): the compiled tap count, 8-byte aligned memory rows.
bool attention_can_fuse_fsmn(int kernel, int64_t ld_mem, const void* mem) {
  return kernel == attn::FSMN_K && ld_mem % 4 == 0 && (reinterpret_cast<uintptr_t>(mem) & 7) == 0;
}

int launch_attention(const TensorMap2D* tm_in, const __nv_bfloat16* qkv, int64_t ld, const int32_t* cu_seqlens, int n_utts, int base,
                     int max_len, int n_heads, __nv_bfloat16* ctx, int64_t ld_ctx, int n_rows, const float* fsmn_w, int fsmn_left,
                     __nv_bfloat16* mem, int64_t ld_mem, cudaStream_t stream, int* zero_ints, int n_zero, float* acc, int64_t ld_acc,
                     const TensorMap2D* tm_ctx) {
  FA_REQUIRE(ld % 8 == 0 && ld_ctx % 8 == 0, "attention: row strides must be multiples of 8 elements");
  FA_REQUIRE((reinterpret_cast<uintptr_t>(qkv) & 15) == 0 && (reinterpret_cast<uintptr_t>(ctx) & 15) == 0,
             "attention: qkv and ctx must be 16-byte aligned");
  FA_REQUIRE(max_len > 0 && n_utts > 0 && n_heads > 0 && n_rows > 0, "attention: empty problem");
  FA_REQUIRE(ld >= 3 * n_heads * attn::HD, "attention: qkv rows hold q|k|v of %d heads x 128", n_heads);
  if (fsmn_w != nullptr) {
    FA_REQUIRE((reinterpret_cast<uintptr_t>(fsmn_w) & 15) == 0 &&
                   (acc != nullptr ? (ld_acc % 4 == 0 && (reinterpret_cast<uintptr_t>(acc) & 15) == 0)
                                   : (mem != nullptr && attention_can_fuse_fsmn(attn::FSMN_K, ld_mem, mem))),
               "attention: fused FSMN needs an 8-byte aligned bf16 memory buffer (or a 16-byte aligned f32 accumulator) and 16-byte aligned taps");
    FA_REQUIRE(fsmn_left >= 0 && fsmn_left < attn::FSMN_K, "attention: FSMN left pad %d outside the %d-tap window", fsmn_left, attn::FSMN_K);
  }
  static SmemOptIn opt;
  static SmemOptIn opt_acc;
  const bool use_acc = fsmn_w != nullptr && acc != nullptr;
  if (int r = use_acc ? opt_in_smem(attn::attention_kernel<true>, attn::SMEM_BYTES, opt_acc)
                      : opt_in_smem(attn::attention_kernel<false>, attn::SMEM_BYTES, opt))
    return r;
  // Without a prebuilt map: rows = the valid packed rows only, so key/value rows past the end are zero-filled by TMA instead
  // of reading memory whose contents are undefined (NaN there would leak into O through 0 * NaN).
  TensorMap2D tm;
  if (tm_in == nullptr) {
    if (int r = make_tensor_map_bf16(&tm, qkv, n_rows, 3 * n_heads * attn::HD, ld, attn::BKV)) return r;
    tm_in = &tm;
  }
  CUtensorMap m, mc;
  memcpy(&m, tm_in->bytes, sizeof(m));
  TensorMap2D tmc;  // the context rows as 32 x 64 boxes (the epilogue's TMA stores)
  if (tm_ctx == nullptr) {
    if (int r = make_tensor_map(&tmc, ctx, 2, n_rows, static_cast<uint64_t>(n_heads) * attn::HD, ld_ctx, 32)) return r;
    tm_ctx = &tmc;
  }
  memcpy(&mc, tm_ctx->bytes, sizeof(mc));
  const int sms = sm_count();
  // Query tiles per work item: two (K/V streamed once for both) unless that leaves SMs without an item.
  const int tiles = ceil_div(max_len, attn::BQ);
  int gq = tiles >= 2 ? 2 : 1;
  if (gq == 2 && n_utts * n_heads * ceil_div(tiles, 2) < sms) gq = 1;
  attn::Params p;
  p.cu = cu_seqlens; p.base = base; p.n_utts = n_utts; p.n_heads = n_heads;
  p.gq = gq; p.groups = ceil_div(tiles, gq);
  p.ctx = ctx; p.ld_ctx = ld_ctx;
  p.scale_log2e = 1.4426950408889634f / sqrtf(static_cast<float>(attn::HD));
  p.qkv = qkv; p.ld_qkv = ld;
  p.fsmn_w = fsmn_w; p.fsmn_left = fsmn_left; p.mem = mem; p.ld_mem = ld_mem;
  p.zero_ints = zero_ints; p.n_zero = zero_ints ? n_zero : 0;
  p.acc = fsmn_w ? acc : nullptr; p.ld_acc = ld_acc;
  const int n_items = n_utts * p.groups * n_heads;
  const int grid = n_items < sms ? n_items : sms;
  // FSMN units (each chunk re-reads K - 1 halo rows, so no smaller than 32 rows): whole utterances when every FSMN warp of
  // the grid gets several; one chunk per warp when there are at most half as many (utterance, head) pairs as warps;
  // otherwise about four chunks per warp, which evens out utterances of different lengths.
  const int fs_warps = grid * attn::FSMN_WARPS, fs_items = n_utts * n_heads;
  int fs_chunks = 1;
  if (2 * fs_items <= fs_warps) fs_chunks = fs_warps / fs_items;
  else if (fs_items < 4 * fs_warps) fs_chunks = ceil_div(4 * fs_warps, fs_items);
  p.fsmn_rows = ceil_div(max_len, fs_chunks) < 32 ? 32 : ceil_div(max_len, fs_chunks);
  p.fsmn_chunks = ceil_div(max_len, p.fsmn_rows);
  ProfScope _ps(K_ATTENTION, stream);
  if (use_acc) FA_CHECK_CUDA(launch_pdl(attn::attention_kernel<true>, dim3(grid), dim3(attn::THREADS), attn::SMEM_BYTES, stream, 1, m, mc, p));
  else FA_CHECK_CUDA(launch_pdl(attn::attention_kernel<false>, dim3(grid), dim3(attn::THREADS), attn::SMEM_BYTES, stream, 1, m, mc, p));
  return 0;
}

}  // namespace fa

extern "C" int fa_op_attention(const void* qkv_dev, int64_t ld_qkv, const int32_t* cu_seqlens_dev, int32_t n_utts,
                               int32_t max_len, int32_t n_heads, int32_t n_rows, void* ctx_dev, int64_t ld_ctx,
                               fa_stream_t stream) {
  using namespace fa;
  FA_REQUIRE(qkv_dev && cu_seqlens_dev && ctx_dev, "attention: NULL pointer");
  return launch_attention(nullptr, static_cast<const __nv_bfloat16*>(qkv_dev), ld_qkv, cu_seqlens_dev, n_utts, 0, max_len, n_heads,
                          static_cast<__nv_bfloat16*>(ctx_dev), ld_ctx, n_rows, nullptr, 0, nullptr, 0, static_cast<cudaStream_t>(stream));
}

extern "C" int fa_op_attention_fsmn(const void* qkv_dev, int64_t ld_qkv, const int32_t* cu_seqlens_dev, int32_t n_utts,
                                    int32_t max_len, int32_t n_heads, int32_t n_rows, void* ctx_dev, int64_t ld_ctx,
                                    const float* fsmn_w_dev, int32_t fsmn_kernel, int32_t fsmn_left, void* mem_bf16_dev, int64_t ld_mem,
                                    float* acc_f32_dev, int64_t ld_acc, fa_stream_t stream) {
  using namespace fa;
  FA_REQUIRE(qkv_dev && cu_seqlens_dev && ctx_dev && fsmn_w_dev && (mem_bf16_dev || acc_f32_dev), "attention_fsmn: NULL pointer");
  FA_REQUIRE(fsmn_kernel == attn::FSMN_K && (acc_f32_dev || attention_can_fuse_fsmn(fsmn_kernel, ld_mem, mem_bf16_dev)),
             "attention_fsmn: only the %d-tap kernel with 8-byte aligned memory rows is fused", attn::FSMN_K);
  return launch_attention(nullptr, static_cast<const __nv_bfloat16*>(qkv_dev), ld_qkv, cu_seqlens_dev, n_utts, 0, max_len, n_heads,
                          static_cast<__nv_bfloat16*>(ctx_dev), ld_ctx, n_rows, fsmn_w_dev, fsmn_left, static_cast<__nv_bfloat16*>(mem_bf16_dev),
                          ld_mem, static_cast<cudaStream_t>(stream), nullptr, 0, acc_f32_dev, ld_acc);
}

#ifdef FA_ATTN_TRACE
// profiling builds only: copies the trace of CTA 0 to the host and clears it; returns the number of events
extern "C" __attribute__((visibility("default"))) int fa_debug_attention_trace(unsigned long long* dst_host, int cap) {
  int n = 16 * fa::attn::TRACE_PER_WARP;
  if (n > cap) n = cap;
  cudaDeviceSynchronize();
  cudaMemcpyFromSymbol(dst_host, fa::attn::g_trace, n * sizeof(unsigned long long));
  static unsigned long long zeros[16 * fa::attn::TRACE_PER_WARP];
  cudaMemcpyToSymbol(fa::attn::g_trace, zeros, sizeof(zeros));
  return n;
}
#endif
