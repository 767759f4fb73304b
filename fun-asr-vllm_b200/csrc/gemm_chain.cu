// Row-local GEMM chain of one SAN-M block in ONE persistent launch:
//     x += ctx Wo^T + b + mem            (out-projection, LN producer)       VLLM/model_executor/models/funasr.py:238,253,123-136
//     h  = relu(LN2(x) W1^T + b1)        (ffn1, LN consumer)                  :138-140,415-431
//     x += h W2^T + b2                   (ffn2, LN producer)                  :415-431,140
//     qkv' = LN1'(x) Wqkv'^T + b'        (next block's qkv, LN consumer)      :120,206
// Everything between two attention kernels touches a token row on its own, so the four GEMMs need no grid-wide barrier:
// a tile of GEMM s for row block rb (256 rows) only needs the tiles of GEMM s-1 of the SAME row block.  The launch walks a
// static list of tiles ordered in macro steps t = 0, 1, ...: step t holds the tiles of (GEMM s, row block t - delay_s), dealt
// round-robin to the CTA pairs; a tile's producer warp waits on a per-(GEMM, row block) counter in global memory that the
// epilogue warps of the tiles it depends on have bumped (release) after their stores completed.  Dependencies always point
// to earlier list positions and every CTA is resident (cooperative launch), so the walk cannot deadlock.
// What it buys: three of the five kernel boundaries of a block disappear (fill / drain / tail wave / grid hand-over, the
// whole cost of a small batch), the HBM-bound tiles of the out-projection run next to the tensor-bound ones of the FFN on
// other SMs, and h [256, d_ff], the bf16 residual copy and the row statistics of a row block are consumed a few macro
// steps after they were written, i.e. out of L2.
//
// The tile arithmetic is that of gemm_tcgen05.cu's CTA-pair kernel (256 x 256 tiles, tcgen05.mma cta_group::2, same k
// order, same epilogue formulas), so the results are bit-identical to the launch-by-launch schedule.
#include <cuda.h>
#include <cooperative_groups.h>

#include <cstdlib>
#include <cstring>

#include "common.cuh"

namespace fa {
namespace chain {

using namespace fa::ptx;

constexpr int BM = 128;            // rows per CTA; a pair owns 256
constexpr int BN = 256;
constexpr int BK = 64;
constexpr int UMMA_K = 16;
constexpr int EPI_WARPS = 8;
constexpr int NUM_THREADS = 64 + 32 * EPI_WARPS + 32;  // + the publisher warp
constexpr int W_PUBLISH = 2 + EPI_WARPS;
constexpr int SLAB_BYTES = 32 * 128;
constexpr int A_BYTES = BM * BK * 2;
constexpr int B_BYTES = BN / 2 * BK * 2;
constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
constexpr int STAGES = 4;
constexpr int SLABS_BYTES = EPI_WARPS * 2 * SLAB_BYTES;  // per warp: 2 ring slots (f32 tiles) or 2 output slabs (bf16 tiles)
constexpr int XSLAB_BYTES = 32 * 64;                     // bf16 copy of a 32 x 32 unit: 64-byte rows, 64B swizzle
constexpr int XSLABS_BYTES = EPI_WARPS * XSLAB_BYTES;
constexpr int CVEC_BYTES = 2 * 2 * BN * 4;               // [bias | colsum] of a bf16 tile's 256 columns, two tiles deep
constexpr int BAR_BYTES = 1024;
constexpr int PUB_RING = 8;
constexpr int SMEM_BYTES = 1024 + STAGES * STAGE_BYTES + SLABS_BYTES + XSLABS_BYTES + CVEC_BYTES + BAR_BYTES;
static_assert(SMEM_BYTES <= 227 * 1024, "chain: shared memory budget");
constexpr int TMEM_COLS = 2 * BN;
constexpr int MAX_PROBS = FA_CHAIN_MAX_PROBS;
constexpr int MAX_SLOTS = FA_CHAIN_MAX_SLOTS;
#ifndef FA_CHAIN_DIST
#define FA_CHAIN_DIST 24
#endif
constexpr int PUBS_PER_TILE = 2;  // the publisher warp of each CTA of the pair bumps a tile's counter once

struct Maps {
  CUtensorMap a[MAX_PROBS], b[MAX_PROBS], c[MAX_PROBS], xb[MAX_PROBS];
};

struct Prob {
  int N, K, tiles_n, num_kb;
  int kind;      // 0: f32 C = acc + bias + C_old (+ rb), bf16 copy + row partials (LN producer); 1: bf16 C = LN-folded epilogue
  int relu;
  int dep;       // problem whose tiles of the same row block must be complete (-1: only the previous kernel)
  int dep_need;  // PUBS_PER_TILE * tiles_n of that problem
  int delay;     // macro steps this problem runs behind problem 0
  int slots;     // row partials per row (producer: N / 32 written; consumer: read)
  float inv_dim;
  const float* bias;
  const float* colsum;
  float2* stats;
  __nv_bfloat16* xb;
  long long ld_xb;
  const __nv_bfloat16* rb;
  long long ld_rb;
  void* c;
  long long ldc;
};

struct Params {
  int M, row0, n_rb, n_prob, n_slots, n_steps;
  int defer;   // > 0: a publication may wait for the pair's next tile when that tile lies at most this many list positions ahead
  int* flags;  // [n_prob][n_rb], zero on entry
  unsigned char slot_prob[MAX_SLOTS], slot_tn[MAX_SLOTS];
  Prob pr[MAX_PROBS];
};

__device__ __forceinline__ int ld_acquire(const int* p) {
  int v;
  asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_add(int* p, int v) {
  asm volatile("red.release.gpu.global.add.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void fence_proxy_async_all() { asm volatile("fence.proxy.async;" ::: "memory"); }
__device__ __forceinline__ float4 ldg_cg_f4(const float4* p) {  // L2 load: the line may have been rewritten by another SM in this launch
  float4 v;
  asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
  return v;
}
__device__ __forceinline__ float2 ldg_cg_f2(const float2* p) {
  float2 v;
  asm volatile("ld.global.cg.v2.f32 {%0, %1}, [%2];" : "=f"(v.x), "=f"(v.y) : "l"(p));
  return v;
}
__device__ __forceinline__ uint4 ldg_cg_v4(const uint4* p) {
  uint4 v;
  asm volatile("ld.global.cg.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
// Waits until *f >= need.  A wait that lasts seconds can only be a scheduling bug: trap instead of hanging the GPU.
__device__ __forceinline__ void wait_counter(const int* f, int need) {
  if (ld_acquire(f) >= need) return;
  const long long t0 = clock64();
  while (ld_acquire(f) < need) {
    __nanosleep(40);
    if (clock64() - t0 > 6000000000ll) __trap();
  }
}

// ---- per-tile timeline of one CTA (make variant NAME=ctrace DEFS=-DFA_CHAIN_TRACE: profiling builds only) ----
#ifdef FA_CHAIN_TRACE
constexpr int TR_TILES = 128, TR_EV = 8, TR_ROLES = 4;  // roles: 0 producer, 1 MMA, 2 epilogue warp 2, 3 epilogue warp 6
__device__ long long g_ctrace[TR_ROLES * TR_TILES * TR_EV];
__device__ int g_ctrace_block = 0;
#define FA_CTRACE(role, it, ev, val)                                                                             \
  do {                                                                                                           \
    if (static_cast<int>(blockIdx.x) == g_ctrace_block && (it) < TR_TILES) g_ctrace[((role) * TR_TILES + (it)) * TR_EV + (ev)] = (val); \
  } while (0)
#else
#define FA_CTRACE(role, it, ev, val)
#endif

// mbarrier wait with the same watchdog (shadows ptx::mbar_wait inside this file): a tile-list bug must fail, not hang
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_test(bar, parity)) return;
  const long long t0 = clock64();
  int n = 0;
  while (!mbar_test(bar, parity)) {
    if ((++n & 1023) == 0 && clock64() - t0 > 6000000000ll) __trap();
  }
}

__global__ void __launch_bounds__(NUM_THREADS, 1)
chain_kernel(const __grid_constant__ Maps maps, const __grid_constant__ Params p) {
  extern __shared__ unsigned char smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t slab_base = smem_base + STAGES * STAGE_BYTES;
  const uint32_t xslab_base = slab_base + SLABS_BYTES;
  const uint32_t cvec_base = xslab_base + XSLABS_BYTES;
  const uint32_t bar_base = cvec_base + CVEC_BYTES;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (STAGES + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + 2 + a); };
  // "the dependency of this CTA's tile number i has been observed": a ring of four (index i & 3).  The producer is never more
  // than two tiles ahead of the epilogue (operand ring + accumulator hand-over), so a barrier is not completed twice before
  // the epilogue has waited on it - with a ring of two it could be (arrive for tile i + 2 while tile i - 1 still drains).
  auto dep_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + 4 + a); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * STAGES + 8);
  auto res_bar = [&](int w, int slot) { return bar_base + 8u * (2 * STAGES + 9 + w * 2 + slot); };
  // "every epilogue warp of this CTA has its part of tile number i in global memory" (ring of PUB_RING, index i % PUB_RING):
  // the publisher warp waits here and does the gpu-scope release + counter bump, so the ~4 k cycles of that fence (measured:
  // a fifth of the epilogue warps' time when they published themselves) are spent by a warp that has nothing else to do
  auto pub_bar = [&](int a) { return bar_base + 8u * (2 * STAGES + 9 + EPI_WARPS * 2 + a); };
  static_assert(8 * (2 * STAGES + 9 + EPI_WARPS * 2 + PUB_RING) <= BAR_BYTES, "barrier region too small");
  volatile uint32_t* tmem_slot_ptr = reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int cl = static_cast<int>(blockIdx.x >> 1), ncl = static_cast<int>(gridDim.x >> 1);
  const int total = p.n_steps * p.n_slots;

  // tile g of the list -> (problem, row block, n tile); false for the empty slots at the ends of the pipeline
  auto decode = [&](int g, int& pi, int& rbk, int& tn) {
    const int t = g / p.n_slots, s = g - t * p.n_slots;
    pi = p.slot_prob[s];
    tn = p.slot_tn[s];
    rbk = t - p.pr[pi].delay;
    return rbk >= 0 && rbk < p.n_rb;
  };

  if (warp == 0 && lane == 0) {
    for (int i = 0; i < p.n_prob; ++i) {
      prefetch_tmap(&maps.a[i]);
      prefetch_tmap(&maps.b[i]);
      prefetch_tmap(&maps.c[i]);
      if (p.pr[i].kind == 0) prefetch_tmap(&maps.xb[i]);
    }
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), EPI_WARPS * 2);
    }
    for (int a = 0; a < 4; ++a) mbar_init(dep_bar(a), 1);
    for (int a = 0; a < PUB_RING; ++a) mbar_init(pub_bar(a), EPI_WARPS);
    for (int w = 0; w < EPI_WARPS; ++w)
      for (int sl = 0; sl < 2; ++sl) mbar_init(res_bar(w, sl), 1);
    fence_mbar_init();
  }
  if (warp == 1) tmem_alloc_pair<TMEM_COLS>(tmem_slot);
  tc_fence_before();
  cluster_sync();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  pdl_wait();

  if (warp == 0) {
    // ===================== TMA producer (both CTAs: own 128 rows of A, own half of the B tile) =====================
    if (lane == 0) {
      int stage = 0, it = 0;
      uint32_t phase = 0;
      for (int g = cl; g < total; g += ncl) {
        int pi, rbk, tn;
        if (!decode(g, pi, rbk, tn)) continue;
        const Prob& P = p.pr[pi];
        FA_CTRACE(0, it, 0, (static_cast<long long>(pi) << 40) | (static_cast<long long>(rbk) << 8) | tn);
        FA_CTRACE(0, it, 1, clock64());
        if (P.dep >= 0) {
          wait_counter(p.flags + P.dep * p.n_rb + rbk, P.dep_need);
          fence_proxy_async_all();  // the TMA reads below are ordered after the acquire
        }
        mbar_arrive(dep_bar(it & 3));
        FA_CTRACE(0, it, 2, clock64());
        const int arow = p.row0 + (rbk * 2 + static_cast<int>(rank)) * BM;
        if (P.kind == 0) {
          // The f32 epilogue is a chain of dependent round trips (old C rows in, new rows out, per 32-column unit): ask L2
          // for the tile's old rows now, a main loop ahead, so those round trips end in L2 instead of HBM.
          for (int r4 = 0; r4 < BM / 32; ++r4)
#pragma unroll
            for (int c8 = 0; c8 < BN / 32; ++c8) tma_prefetch_l2_2d(&maps.c[pi], tn * BN + c8 * 32, arow + r4 * 32);
        }
        const int brow = tn * BN + static_cast<int>(rank) * (BN / 2);
        for (int kb = 0; kb < P.num_kb; ++kb) {
          mbar_wait(empty_bar(stage), phase ^ 1);
          const uint32_t sa = smem_base + stage * STAGE_BYTES;
          const uint32_t lead_bar = mapa(full_bar(stage), 0);
          if (rank == 0) mbar_expect_tx(full_bar(stage), 2 * STAGE_BYTES);
          tma_load_2d_pair(sa, &maps.a[pi], lead_bar, kb * BK, arow);
          tma_load_2d_pair(sa + A_BYTES, &maps.b[pi], lead_bar, kb * BK, brow);
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        FA_CTRACE(0, it, 3, clock64());
        ++it;
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (the pair's leader) =====================
    constexpr uint32_t idesc = make_idesc(2 * BM, BN, 0);
    int stage = 0, it = 0;
    uint32_t phase = 0;
    if (rank == 0) {
      for (int g = cl; g < total; g += ncl) {
        int pi, rbk, tn;
        if (!decode(g, pi, rbk, tn)) continue;
        const int num_kb = p.pr[pi].num_kb;
        const int acc = it & 1;
        if (lane == 0) FA_CTRACE(1, it, 1, clock64());
        mbar_wait(tempty_bar(acc), ((it >> 1) & 1) ^ 1);
        if (lane == 0) FA_CTRACE(1, it, 2, clock64());
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + acc * BN;
        for (int kb = 0; kb < num_kb; ++kb) {
          mbar_wait(full_bar(stage), phase);
          if (lane == 0 && kb == 0) FA_CTRACE(1, it, 3, clock64());
          tc_fence_after();
          if (elect_one()) {
            const uint32_t sa = smem_base + stage * STAGE_BYTES;
            const uint32_t sb = sa + A_BYTES;
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k)
              umma_f16_pair(tmem_d, desc_kmajor(sa + k * UMMA_K * 2), desc_kmajor(sb + k * UMMA_K * 2), idesc, (kb | k) != 0);
            umma_commit_pair(empty_bar(stage));
            if (kb == num_kb - 1) umma_commit_pair(tfull_bar(acc));
          }
          __syncwarp();
          if (++stage == STAGES) { stage = 0; phase ^= 1; }
        }
        if (lane == 0) FA_CTRACE(1, it, 4, clock64());
        ++it;
      }
    }
  } else if (warp == W_PUBLISH) {
    // ===================== publisher: tile complete in this CTA -> release at gpu scope, bump the tile's counter =====================
    if (lane == 0) {
      int it = 0;
      for (int g = cl; g < total; g += ncl) {
        int pi, rbk, tn;
        if (!decode(g, pi, rbk, tn)) continue;
        mbar_wait(pub_bar(it % PUB_RING), (it / PUB_RING) & 1);  // all eight epilogue warps: stores performed / ordered before their arrive
        fence_proxy_async_all();
        red_release_add(p.flags + pi * p.n_rb + rbk, 1);   // release.gpu: cumulative over what the arrivals ordered before it
        ++it;
      }
    }
  } else {
    // ===================== epilogue warps =====================
    const int q = warp & 3;
    const int ew = warp - 2;
    const int hf = ew >> 2;
    const uint32_t rsw = static_cast<uint32_t>(lane & 7);
    const uint32_t row_off = static_cast<uint32_t>(lane) * 128u;
    const uint32_t my_slabs = slab_base + ew * 2 * SLAB_BYTES;
    const uint32_t my_xslab = xslab_base + ew * XSLAB_BYTES;
    const uint32_t xrow_off = static_cast<uint32_t>(lane) * 64u;
    const uint32_t xsw = static_cast<uint32_t>(lane >> 1) & 3u;  // 64B swizzle: 16-byte chunk index ^ bits [8:7] of the byte offset
    int it = 0;
    int nres = 0;  // residual slabs this warp has requested so far (ring slot = nres & 1)
    int pend = -1;  // number of the previous tile when its publication was deferred
    int nbf = 0;    // bf16 tiles this warp has seen (parity = constants buffer)
    for (int g = cl; g < total; g += ncl) {
      int pi, rbk, tn;
      if (!decode(g, pi, rbk, tn)) continue;
      const Prob& P = p.pr[pi];
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      const int row0 = (rbk * 2 + static_cast<int>(rank)) * BM + q * 32;  // first row of this warp, local to the launch
      const bool rows_full = row0 + 32 <= p.M;
      const bool my_row = row0 + lane < p.M;
      const uint32_t tmem_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
      const CUtensorMap* mc = &maps.c[pi];
      [[maybe_unused]] const int trole = 2 + (ew >> 2);
      [[maybe_unused]] const bool tr = lane == 0 && (ew & 3) == 0;
      if (tr) FA_CTRACE(trole, it, 1, clock64());
#ifdef FA_CHAIN_TRACE
      long long tr_res = 0, tr_rd = 0;
#endif
      mbar_wait(dep_bar(it & 3), (it >> 2) & 1);  // whatever this tile reads from earlier tiles is complete and visible
      if (P.kind == 0) {
        // ---- f32 tile: C = acc + bias + C_old (+ bf16 residual); bf16 copy and row partials of the new rows.  Units of 32
        // columns; this warp takes units hf, hf + 2, ...; C_old arrives by TMA in the slab the result leaves from.
        constexpr int UH = BN / 32 / 2;
        auto coords = [&](int i, int& col0, int& u) {
          u = hf + 2 * i;
          col0 = tn * BN + u * 32;
        };
        const int nres0 = nres;  // unit i of this tile is residual slab nres0 + i of the warp's sequence (warp-uniform)
        auto issue_load = [&](int i) {  // lane 0
          if (i >= UH) return;
          int col0, u;
          coords(i, col0, u);
          const int slot = (nres0 + i) & 1;
          mbar_expect_tx(res_bar(ew, slot), SLAB_BYTES);
          tma_load_2d(my_slabs + slot * SLAB_BYTES, mc, res_bar(ew, slot), col0, p.row0 + row0);
        };
        auto ld_rb = [&](int i, uint4 (&v)[4]) {
#pragma unroll
          for (int k = 0; k < 4; ++k) v[k] = make_uint4(0u, 0u, 0u, 0u);
          if (P.rb == nullptr || i >= UH || !my_row) return;
          int col0, u;
          coords(i, col0, u);
          const uint4* gp = reinterpret_cast<const uint4*>(P.rb + static_cast<long long>(p.row0 + row0 + lane) * P.ld_rb + col0);
#pragma unroll
          for (int k = 0; k < 4; ++k) v[k] = ldg_cg_v4(gp + k);
        };
        if (lane == 0) {
          fence_proxy_async_all();
          issue_load(0);
          issue_load(1);
        }
        uint4 rbv[4];
        ld_rb(0, rbv);
        if (tr) FA_CTRACE(trole, it, 2, clock64());
        mbar_wait(tfull_bar(acc), acc_phase);
        if (tr) FA_CTRACE(trole, it, 3, clock64());
        tc_fence_after();
        for (int i = 0; i < UH; ++i) {
          float ln_s = 0.f, ln_ss = 0.f;
          int col0, u;
          coords(i, col0, u);
          const int rn = nres0 + i;
          const uint32_t slab = my_slabs + (rn & 1) * SLAB_BYTES;
          uint4 rbn[4];
          ld_rb(i + 1, rbn);
#ifdef FA_CHAIN_TRACE
          const long long tq0 = clock64();
#endif
          mbar_wait(res_bar(ew, rn & 1), (rn >> 1) & 1);
#ifdef FA_CHAIN_TRACE
          tr_res += clock64() - tq0;
#endif
          uint32_t r[32];
          tmem_ld32_nowait(tmem_row + u * 32, r);
          tmem_ld_wait();
          const float4* bp = reinterpret_cast<const float4*>(P.bias + col0);
          uint32_t xbp[4];
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            const uint32_t addr = slab + row_off + ((static_cast<uint32_t>(j) ^ rsw) << 4);
            const float4 bv = __ldg(bp + j);
            float4 v = lds128(addr);
            v.x += __uint_as_float(r[4 * j]) + bv.x;
            v.y += __uint_as_float(r[4 * j + 1]) + bv.y;
            v.z += __uint_as_float(r[4 * j + 2]) + bv.z;
            v.w += __uint_as_float(r[4 * j + 3]) + bv.w;
            {
              const uint32_t m0 = (j & 1) ? rbv[j >> 1].z : rbv[j >> 1].x, m1 = (j & 1) ? rbv[j >> 1].w : rbv[j >> 1].y;
              v.x += __uint_as_float(m0 << 16); v.y += __uint_as_float(m0 & 0xffff0000u);
              v.z += __uint_as_float(m1 << 16); v.w += __uint_as_float(m1 & 0xffff0000u);
            }
            sts128(addr, __float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z), __float_as_uint(v.w));
            const uint32_t lo = pack_bf16x2(v.x, v.y), hi = pack_bf16x2(v.z, v.w);
            xbp[2 * (j & 1)] = lo;
            xbp[2 * (j & 1) + 1] = hi;
            const float b0 = __uint_as_float(lo << 16), b1 = __uint_as_float(lo & 0xffff0000u);
            const float b2 = __uint_as_float(hi << 16), b3 = __uint_as_float(hi & 0xffff0000u);
            ln_s += (b0 + b1) + (b2 + b3);
            ln_ss = fmaf(b0, b0, fmaf(b1, b1, fmaf(b2, b2, fmaf(b3, b3, ln_ss))));
            if (j & 1) {
              // bf16 copy: through its own swizzled slab and a TMA store (a 16-byte store per lane with a 1 KB row pitch is 32
              // LSU wavefronts per instruction - the f32 epilogue was LSU-bound on these); ragged last rows: straight to global
              if (rows_full) sts128(my_xslab + xrow_off + ((static_cast<uint32_t>(j >> 1) ^ xsw) << 4), xbp[0], xbp[1], xbp[2], xbp[3]);
              else if (my_row)
                *reinterpret_cast<uint4*>(P.xb + static_cast<long long>(p.row0 + row0 + lane) * P.ld_xb + col0 + 4 * (j - 1)) =
                    make_uint4(xbp[0], xbp[1], xbp[2], xbp[3]);
            }
          }
          if (my_row) P.stats[static_cast<long long>(p.row0 + row0 + lane) * P.slots + (col0 >> 5)] = make_float2(ln_s, ln_ss);
          if (i == UH - 1) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(mapa(tempty_bar(acc), 0));
          }
          if (rows_full) {
            fence_async_smem();
            __syncwarp();
#ifdef FA_CHAIN_TRACE
            const long long tq1 = clock64();
#endif
            if (lane == 0) {
              tma_store_2d(mc, slab, col0, p.row0 + row0);
              tma_store_2d(&maps.xb[pi], my_xslab, col0, p.row0 + row0);
              bulk_commit();
              bulk_wait_read<0>();
            }
#ifdef FA_CHAIN_TRACE
            tr_rd += clock64() - tq1;
#endif
          } else {
            __syncwarp();
            if (my_row) {
              float* gp = static_cast<float*>(P.c) + static_cast<long long>(p.row0 + row0 + lane) * P.ldc + col0;
#pragma unroll
              for (int j = 0; j < 8; ++j)
                *reinterpret_cast<float4*>(gp + 4 * j) = lds128(slab + row_off + ((static_cast<uint32_t>(j) ^ rsw) << 4));
            }
            fence_async_smem();
            __syncwarp();
          }
          if (lane == 0) issue_load(i + 2);
#pragma unroll
          for (int k = 0; k < 4; ++k) rbv[k] = rbn[k];
          __syncwarp();
        }
        nres = nres0 + UH;
      } else {
        // ---- bf16 tile: rstd * (acc - mean * colsum) + bias (, relu) with the row's statistics from the producer's partials
        // The tile's 256 bias / colsum values go to shared memory first (one of each per epilogue thread; buffer = parity of
        // the warp's bf16-tile count, so a warp that runs ahead never overwrites what a slower one still reads): broadcast
        // LDS in the inner loop instead of L1 round trips (4.7 k against 3.1 k cycles per tile measured).
        const uint32_t cv = cvec_base + static_cast<uint32_t>(nbf & 1) * (2 * BN * 4);
        {
          const int c = (ew << 5) | lane;
          asm volatile("st.shared.f32 [%0], %1;" ::"r"(cv + 4u * c), "f"(__ldg(P.bias + tn * BN + c)) : "memory");
          asm volatile("st.shared.f32 [%0], %1;" ::"r"(cv + 4u * (BN + c)), "f"(__ldg(P.colsum + tn * BN + c)) : "memory");
        }
        ++nbf;
        float ln_rs, ln_nm;
        {
          float s = 0.f, ss = 0.f;
          if (my_row) {
            const float2* sp = P.stats + static_cast<long long>(p.row0 + row0 + lane) * P.slots;
            if (P.slots == 16) {
              float4 st[8];
#pragma unroll
              for (int k = 0; k < 8; ++k) st[k] = ldg_cg_f4(reinterpret_cast<const float4*>(sp + 2 * k));
#pragma unroll
              for (int k = 0; k < 8; ++k) {
                s += st[k].x; ss += st[k].y;
                s += st[k].z; ss += st[k].w;
              }
            } else {
              for (int k = 0; k < P.slots; ++k) {
                const float2 t = ldg_cg_f2(sp + k);
                s += t.x;
                ss += t.y;
              }
            }
          }
          const float mu = s * P.inv_dim;
          ln_rs = rsqrtf(fmaxf(fmaf(-mu, mu, ss * P.inv_dim), 0.f) + 1e-12f);
          ln_nm = -ln_rs * mu;
        }
        asm volatile("bar.sync 1, %0;" ::"n"(32 * EPI_WARPS) : "memory");  // the eight epilogue warps: constants of this tile staged
        if (tr) FA_CTRACE(trole, it, 2, clock64());
        mbar_wait(tfull_bar(acc), acc_phase);
        if (tr) FA_CTRACE(trole, it, 3, clock64());
        tc_fence_after();
        constexpr int UNITS = BN / 64;
#pragma unroll 1
        for (int u = hf; u < UNITS; u += 2) {
          const int col0 = tn * BN + u * 64;
          const uint32_t slab = my_slabs + (u >> 1) * SLAB_BYTES;  // two units per warp and tile: one slab each, free at tile start
          uint32_t r0[32], r1[32];
          tmem_ld32_nowait(tmem_row + u * 64, r0);
          tmem_ld32_nowait(tmem_row + u * 64 + 32, r1);
          tmem_ld_wait();
#pragma unroll
          for (int half = 0; half < 2; ++half) {
            const uint32_t(&r)[32] = half ? r1 : r0;
            const uint32_t bp = cv + 4u * (u * 64 + half * 32);
            const uint32_t cp = bp + 4u * BN;
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              const float4 b0 = lds128(bp + 32 * j), b1 = lds128(bp + 32 * j + 16);
              const float4 c0 = lds128(cp + 32 * j), c1 = lds128(cp + 32 * j + 16);
              const float cs[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
              const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
              float v[8];
#pragma unroll
              for (int e = 0; e < 8; e += 2)
                ffma2(v[e], v[e + 1], ln_rs, __uint_as_float(r[8 * j + e]), __uint_as_float(r[8 * j + e + 1]), ln_nm, cs[e], cs[e + 1],
                      bb[e], bb[e + 1]);
              if (P.relu) {
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] = fmaxf(v[e], 0.f);
              }
              sts128(slab + row_off + ((static_cast<uint32_t>(half * 4 + j) ^ rsw) << 4), pack_bf16x2(v[0], v[1]),
                     pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
            }
          }
          if (u + 2 >= UNITS) {
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive_cluster(mapa(tempty_bar(acc), 0));
          }
          if (rows_full) {
            fence_async_smem();
            __syncwarp();
            if (lane == 0) {
              tma_store_2d(mc, slab, col0, p.row0 + row0);
              bulk_commit();
            }
          } else {
            __syncwarp();
            if (my_row) {
              unsigned char* gp = static_cast<unsigned char*>(P.c) + (static_cast<long long>(p.row0 + row0 + lane) * P.ldc + col0) * 2;
#pragma unroll
              for (int j = 0; j < 8; ++j)
                *reinterpret_cast<float4*>(gp + 16 * j) = lds128(slab + row_off + ((static_cast<uint32_t>(j) ^ rsw) << 4));
            }
            fence_async_smem();
            __syncwarp();
          }
        }
      }
      // ---- publish this warp's part of the tile: its TMA stores must have been PERFORMED (not only have read their slabs),
      // which takes microseconds.  When the list guarantees that nothing this pair does next can depend on the tile (p.defer,
      // see the launcher) the wait is deferred: the counter is bumped after the NEXT tile's epilogue, by which time the stores
      // are long complete (wait_group N with N = the groups committed since), so the epilogue never stalls on a store.
      __syncwarp();  // the other lanes' plain stores (bf16 copy, partials, ragged rows) are ordered before lane 0's release
      if (tr) FA_CTRACE(trole, it, 4, clock64());
#ifdef FA_CHAIN_TRACE
      if (tr) { FA_CTRACE(trole, it, 6, tr_res); FA_CTRACE(trole, it, 7, tr_rd); }
#endif
      const int my_groups = rows_full ? (P.kind == 0 ? BN / 32 / 2 : BN / 64 / 2) : 0;
      if (pend >= 0) {
        if (lane == 0) {
          if (my_groups == 4) bulk_wait<4>();
          else if (my_groups == 2) bulk_wait<2>();
          else bulk_wait<0>();
          mbar_arrive(pub_bar(pend % PUB_RING));
        }
        pend = -1;
      }
      bool defer = false;
      if (p.defer > 0) {
        // this pair's next tile: deferring needs one within p.defer list positions (see the launcher), on another row block
        for (int g2 = g + ncl; g2 < total && g2 - g <= p.defer; g2 += ncl) {
          int pi2, rb2, tn2;
          if (decode(g2, pi2, rb2, tn2)) {
            defer = rb2 != rbk;
            break;
          }
        }
      }
      if (defer) {
        pend = it;
      } else if (lane == 0) {
        bulk_wait<0>();
        mbar_arrive(pub_bar(it % PUB_RING));
      }
      if (tr) FA_CTRACE(trole, it, 5, clock64());
      ++it;
    }
    if (pend >= 0 && lane == 0) {
      bulk_wait<0>();
      mbar_arrive(pub_bar(pend % PUB_RING));
    }
  }

  tc_fence_before();
  cluster_sync();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc_pair<TMEM_COLS>(tmem_base);
  }
}

}  // namespace chain

#ifdef FA_CHAIN_TRACE
// profiling builds only: choose the traced CTA (block index) / copy the trace of the last launch to the host
extern "C" __attribute__((visibility("default"))) int fa_debug_chain_trace(long long* dst_host, int n, int next_block) {
  cudaDeviceSynchronize();
  const int cap = chain::TR_ROLES * chain::TR_TILES * chain::TR_EV;
  if (n > cap) n = cap;
  cudaMemcpyFromSymbol(dst_host, chain::g_ctrace, sizeof(long long) * n);
  cudaMemcpyToSymbol(chain::g_ctrace_block, &next_block, sizeof(int));
  return n;
}
#endif

int gemm_chain_flag_ints(int n_probs, int max_rows) { return n_probs * ceil_div(max_rows, 2 * chain::BM); }

// Runs the chain over rows [row0, row0 + M).  flags: n_probs * ceil(M / 256) zeroed ints.
int launch_gemm_chain(const ChainDesc& d, int M, int row0, int* flags, cudaStream_t stream) {
  using namespace chain;
  FA_REQUIRE(d.n_probs >= 2 && d.n_probs <= MAX_PROBS, "chain: 2..%d problems", MAX_PROBS);
  FA_REQUIRE(M > 0 && row0 >= 0 && flags != nullptr, "chain: bad arguments");
  Maps maps;
  Params p;
  memset(&p, 0, sizeof(p));
  p.M = M;
  p.row0 = row0;
  p.n_rb = ceil_div(M, 2 * BM);
  p.n_prob = d.n_probs;
  p.flags = flags;
  // Distance (in macro steps) between a tile and the tiles it depends on: far enough that they were dealt a wave or two
  // earlier when there are many row blocks, 1 when the launch is a handful of row blocks and latency is everything.
  const int sms = sm_count();
  const int pairs = sms / 2;
  int n_slots = 0;
  for (int i = 0; i < d.n_probs; ++i) n_slots += d.p[i].N / BN;
  FA_REQUIRE(n_slots <= MAX_SLOTS, "chain: %d tiles per row block exceed %d", n_slots, MAX_SLOTS);
  // A tile is published ~15 us after it was picked up (32 k-blocks of ffn2, epilogue, store completion, counter) and a macro
  // step is picked up every ~0.8 us by 74 pairs: its dependencies must lie ~20 steps back or the pairs wait in list order
  // (measured with 2 steps at 11 row blocks: 118 us per chain against 45 us for four launches).  Few row blocks: problem-major
  // order (dist = n_rb), i.e. four launches' worth of tiles without the three grid hand-overs.
  int dist = FA_CHAIN_DIST;
  if (dist > p.n_rb) dist = p.n_rb;
  for (int i = 0; i < d.n_probs; ++i) {
    const ChainProb& s = d.p[i];
    Prob& P = p.pr[i];
    FA_REQUIRE(s.N % BN == 0 && s.K % BK == 0 && s.K >= BK, "chain: problem %d: N %% 256 == 0 and K %% 64 == 0 required (N %d K %d)", i, s.N, s.K);
    FA_REQUIRE(s.bias != nullptr && s.c != nullptr, "chain: problem %d: NULL bias / output", i);
    P.N = s.N; P.K = s.K; P.tiles_n = s.N / BN; P.num_kb = s.K / BK;
    P.kind = s.kind; P.relu = s.relu;
    P.dep = i - 1;
    P.dep_need = i > 0 ? PUBS_PER_TILE * (d.p[i - 1].N / BN) : 0;
    P.delay = i * dist;
    P.bias = s.bias; P.colsum = s.colsum; P.stats = s.stats;
    P.xb = s.xb; P.ld_xb = s.ld_xb; P.rb = s.res_bf16; P.ld_rb = s.ld_res_bf16;
    P.c = s.c; P.ldc = s.N;
    if (s.kind == 0) {
      FA_REQUIRE(s.stats && s.xb && s.ld_xb % 8 == 0 && s.N / 32 <= 16, "chain: producer %d needs ln_stats, xb and N <= 512", i);
      FA_REQUIRE(!s.res_bf16 || (s.ld_res_bf16 % 8 == 0 && (reinterpret_cast<uintptr_t>(s.res_bf16) & 15) == 0), "chain: bf16 residual alignment");
      P.slots = s.N / 32;
    } else {
      FA_REQUIRE(i > 0 && d.p[i - 1].kind == 0 && s.colsum && s.stats && s.K == d.p[i - 1].N, "chain: consumer %d must follow its LN producer", i);
      P.slots = d.p[i - 1].N / 32;
      P.inv_dim = 1.f / static_cast<float>(s.K);
    }
    memcpy(&maps.a[i], s.ta->bytes, sizeof(CUtensorMap));
    memcpy(&maps.b[i], s.tb->bytes, sizeof(CUtensorMap));
    memcpy(&maps.c[i], s.tc->bytes, sizeof(CUtensorMap));
    if (s.kind == 0) {
      FA_REQUIRE(s.txb != nullptr, "chain: producer %d needs the tensor map of its bf16 copy", i);
      memcpy(&maps.xb[i], s.txb->bytes, sizeof(CUtensorMap));
    } else {
      maps.xb[i] = maps.c[i];
    }
  }
  for (int i = d.n_probs; i < MAX_PROBS; ++i) { maps.a[i] = maps.a[0]; maps.b[i] = maps.b[0]; maps.c[i] = maps.c[0]; maps.xb[i] = maps.c[0]; }
  // CTA pairs the device can hold at once (GPCs with an odd SM count leave a few SMs without a partner): cached per device
  static SmemOptIn opt;
  if (int r = opt_in_smem(chain_kernel, SMEM_BYTES, opt)) return r;
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(NUM_THREADS);
  cfg.dynamicSmemBytes = SMEM_BYTES;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = 2;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  const long long tiles = static_cast<long long>(p.n_rb) * n_slots;
  int use_pairs = pairs;
  {
    static std::atomic<int> max_pairs[64];
    int dev = 0;
    FA_CHECK_CUDA(cudaGetDevice(&dev));
    int mp = dev < 64 ? max_pairs[dev].load(std::memory_order_relaxed) : 0;
    if (mp <= 0) {
      cfg.gridDim = dim3(2 * pairs);
      FA_CHECK_CUDA(cudaOccupancyMaxActiveClusters(&mp, chain_kernel, &cfg));
      FA_REQUIRE(mp > 0, "chain: the device cannot hold a single CTA pair of the chain kernel");
      if (dev < 64) max_pairs[dev].store(mp, std::memory_order_relaxed);
    }
    if (use_pairs > mp) use_pairs = mp;
  }
  if (tiles < use_pairs) use_pairs = static_cast<int>(tiles);
  // Slots of a macro step.  List position g goes to pair g % use_pairs, so the k-th tile of a pair sits in slot
  // (c + k * stride) % n_slots with stride = use_pairs % n_slots: a pair walks one of gcd(stride, n_slots) cycles through the
  // slots.  Every cycle gets its share of each problem, in the order that keeps the tensor pipe busy while an f32 epilogue
  // (HBM round trips, ~4x its main loop) drains: f32 tiles back to back by growing K - the long main loop of ffn2 covers
  // the epilogue of the out-projection tile before it - then the bf16 problems interleaved.
  {
    auto gcd = [](int a, int b) { while (b) { const int t = a % b; a = b; b = t; } return a; };
    while (use_pairs > 1 && gcd(use_pairs % n_slots, n_slots) > 2) --use_pairs;
    const int stride = use_pairs % n_slots;
    const int cycles = use_pairs > 1 ? gcd(stride, n_slots) : 1;
    bool even = true;
    for (int i = 0; i < d.n_probs; ++i) even = even && (d.p[i].N / BN) % cycles == 0;
    int next_tn[MAX_PROBS] = {0, 0, 0, 0};
    if (even && use_pairs > 1) {
      const int len = n_slots / cycles;
      for (int c0 = 0; c0 < cycles; ++c0) {
        int left[MAX_PROBS], seq[MAX_SLOTS], n = 0;
        for (int i = 0; i < d.n_probs; ++i) left[i] = d.p[i].N / BN / cycles;
        for (;;) {  // f32 problems first, by growing K
          int best = -1;
          for (int i = 0; i < d.n_probs; ++i)
            if (left[i] > 0 && d.p[i].kind == 0 && (best < 0 || d.p[i].K < d.p[best].K)) best = i;
          if (best < 0) break;
          while (left[best] > 0) { seq[n++] = best; --left[best]; }
        }
        while (n < len) {  // bf16 problems: the one furthest behind its share
          int best = -1;
          double best_w = -1.0;
          for (int i = 0; i < d.n_probs; ++i) {
            if (left[i] == 0) continue;
            const double w = static_cast<double>(left[i]) / (d.p[i].N / BN / cycles) + 1e-6 * i;
            if (w > best_w) { best_w = w; best = i; }
          }
          seq[n++] = best;
          --left[best];
        }
        for (int j = 0; j < len; ++j) {
          const int slot = (c0 + j * stride) % n_slots;
          p.slot_prob[slot] = static_cast<unsigned char>(seq[j]);
          p.slot_tn[slot] = static_cast<unsigned char>(next_tn[seq[j]]++);
        }
      }
    } else {
      int n = 0;
      for (int i = 0; i < d.n_probs; ++i)
        for (int t = 0; t < d.p[i].N / BN; ++t) { p.slot_prob[n] = static_cast<unsigned char>(i); p.slot_tn[n] = static_cast<unsigned char>(t); ++n; }
    }
  }
  p.n_slots = n_slots;
  p.n_steps = p.n_rb + p.pr[d.n_probs - 1].delay;
  // Deferred publication makes "tile at list position g is published" wait for the epilogue of the pair's next tile at
  // g2 <= g + R.  That tile's dependencies sit in macro step t2 - dist, i.e. at positions <= g2 - (dist - 1) n_slots - 1
  // <= g - n_slots - 1 for R = (dist - 2) n_slots: every wait points to a strictly earlier list position than g, whatever
  // is deferred there in turn, so the walk stays deadlock-free.  (Empty slots make "the next tile" arbitrarily far in the
  // problem-major order of small launches: there the publication is immediate.)
  p.defer = (dist - 2) * n_slots;
  if (p.defer < pairs) p.defer = 0;
  cfg.gridDim = dim3(2 * use_pairs);
  // every CTA spins on counters other CTAs of the same grid bump: all of them must be resident, whatever else is running
  // on the device (another stream's chain, the other plan of the overlapped pipeline) -> cooperative launch
  attr[1].id = cudaLaunchAttributeCooperative;
  attr[1].val.cooperative = 1;
  cfg.numAttrs = 2;
  ProfScope _ps(K_GEMM, stream);
  FA_CHECK_CUDA(cudaLaunchKernelEx(&cfg, chain_kernel, maps, p));
  return 0;
}

}  // namespace fa
