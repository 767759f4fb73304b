// bf16 GEMM  C[M,N] = A[M,K] * W[N,K]^T  (+bias, ReLU, residuals) on sm_100a tensor cores.
//
// Persistent, warp-specialised kernel:
//   warp 0      TMA producer   (cp.async.bulk.tensor 2D, 128B swizzle, mbarrier complete_tx)
//   warp 1      MMA issuer     (tcgen05.mma cta_group::1 kind::f16, M=128, N=BN, K=16 per instruction,
//                               fp32 accumulators in TMEM, double buffered: 2 x BN columns)
//   warps 2..9  epilogue       (tcgen05.ld 32x32b -> registers -> bias/ReLU -> 128B-swizzled smem slab -> TMA store).
//                               f32 residuals are PREFETCHED by TMA into the very slab the result leaves from (a 2-slot
//                               ring per warp, loads issued two units ahead), so the epilogue is lds + add + sts in place
//                               and no warp ever waits on a global load it has just issued;
//                               two warps per TMEM lane quarter (= per SM sub-partition), each taking every other
//                               128-byte column unit: ncu showed the 4-warp epilogue (1500 instr/tile/warp, one warp
//                               per scheduler) pacing the tensor pipe at 51 % active (profiles/r01_b_gemm.md)
// Three pipelines: smem full/empty (TMA<->MMA), TMEM full/empty (MMA<->epilogue), static tile
// scheduler (tile = blockIdx.x + i * gridDim.x).  Both operands are K-major (nn.Linear layout),
// so no transposes are ever materialised.
//
// Replaces the cuBLAS calls behind nn.Linear in VLLM/model_executor/models/funasr.py:164-168 (qkv),
// :159-163,238 (out), :415-431 (FFN), :543-552,585-587 (adaptor linears).
#include <cuda.h>

#include <cstdlib>

#include "common.cuh"

namespace fa {
namespace gemm {

constexpr int BM = 128;
constexpr int BK = 64;  // 64 bf16 = 128 bytes = one swizzle span
constexpr int UMMA_K = 16;
constexpr int EPI_WARPS = 8;                    // two per TMEM lane quarter
constexpr int NUM_THREADS = 64 + 32 * EPI_WARPS;
constexpr int SLAB_BYTES = 32 * 128;            // 32 rows x 128 B, SWIZZLE_128B (64 bf16 or 32 f32 columns)
constexpr int SMEM_LIMIT = 227 * 1024;

// residual / output slots per epilogue warp (f32 epilogue with residuals): loads run RING - 1 units ahead; a slot is
// refilled only after the store that left from it has been read.  One residual: 3 slots (no exposed store drain).
constexpr int res_ring(int) { return 2; }  // (3 slots, i.e. no exposed store drain, measured slower: one operand stage less)
constexpr int RES_RING_MAX = 3;
constexpr int BAR_BYTES = 512;

// CTAS = 2: a CTA pair (one cluster, the two SMs of a TPC) computes a 256 x BN tile with tcgen05.mma.cta_group::2.  Each
// CTA stages its own 128 rows of A and only HALF of the B tile, so the operand bytes an SM pulls through its 64 B/clk
// L2 port drop from 48 KB to 32 KB per 128x256x64 of MMA work (ncu, profiles/r01_e_campaign.md: the 1-CTA kernel moves
// 54 B/clk/SM into the SMs at 63 % tensor-pipe activity, i.e. it is L2->SM ingress bound).
template <int BN, int NRES, int CTAS, bool OUT_F32, bool XBS = false>
struct Cfg {
  static constexpr int A_BYTES = BM * BK * 2;
  static constexpr int B_BYTES = BN / CTAS * BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  // staging slabs per epilogue warp: one (no residual) or RES_RING slots of NRES slabs (slab 0 of a slot receives
  // residual 1 and then holds the result; slab 1 receives residual 2)
  // no residual: bf16 output (K = d_model GEMMs, little main-loop time per tile) gets two slabs so a store drains while the
  // next unit is produced; the f32 accumulate output (K = d_ff) keeps one slab and one more operand stage
  static constexpr int RING = res_ring(NRES);
  static constexpr int SLABS_PER_WARP = NRES == 0 ? (OUT_F32 ? 1 : 2) : RING * NRES;
  // XBS (folded-LN producer with a short k loop): one more slab per warp, 32 rows x 64 B (64B swizzle), for the bf16 copy of
  // a unit, which then leaves by TMA with the f32 rows instead of as 16-byte stores per lane with a 1 KB row pitch (32 LSU
  // wavefronts per instruction).  It costs the ring its fifth stage, which the K = 2048 producer (ffn2) needs more than the
  // stores hurt it (127 -> 136 us with the slab); the K = 512 out-projection gains 90 -> 85 us.
  static constexpr int XSLAB_BYTES = 32 * 64;
  static constexpr int SLABS_BYTES = EPI_WARPS * (SLABS_PER_WARP * SLAB_BYTES + (XBS ? XSLAB_BYTES : 0));
  // bf16 output: bias (and the folded-LN column sums) of all N columns are staged in smem once per CTA, so the epilogue's
  // per-column constants are short-latency broadcast LDS instead of L1 round trips (ncu: a fifth of its stall samples)
  static constexpr int CVEC_BYTES = OUT_F32 ? 0 : 16384;
  static constexpr int MAX_STAGES = (SMEM_LIMIT - 1024 - BAR_BYTES - SLABS_BYTES - CVEC_BYTES) / STAGE_BYTES;
  static constexpr int STAGES = MAX_STAGES > 8 ? 8 : MAX_STAGES;
  static_assert(STAGES >= 2, "not enough shared memory left for the operand ring");
  static constexpr int TMEM_COLS = 2 * BN;  // double-buffered accumulator (power of two >= 32)
  static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + SLABS_BYTES + CVEC_BYTES + 1024 /*align slack*/ + BAR_BYTES;
};

struct Params {
  int M, N, K;
  int row0;  // first row of A / C inside the matrices the tensor maps describe
  int tiles_m, tiles_n;
  const float* bias;
  const float* res1;
  const float* res2;
  void* c;
  long long ldc;
  int relu;
  int out_f32;
  int accumulate;  // C += result (TMA reduce-add in L2) instead of C = result
  float2* ln_stats;  // see GemmEpilogue
  int ln_slots;
  float ln_inv_dim;
  const float* colsum;
  __nv_bfloat16* xb;
  long long ld_xb;
  const __nv_bfloat16* rb;  // f32 output: optional bf16 residual (the FSMN memory beside the attention context), row stride ld_rb
  long long ld_rb;
};

using namespace fa::ptx;  // shared inline-PTX wrappers (ptx.cuh)
__device__ __forceinline__ uint64_t make_smem_desc(uint32_t a) { return desc_kmajor(a); }

// ---------------------------------------------------------------- kernel
// Epilogue variants are compile-time: OUT_F32 (f32 C, optional NRES f32 residuals) or bf16 C (optional RELU).
// LN: 0 none, 1 consumer (bf16 output), 2 producer (f32 output, NRES == 1) of a folded LayerNorm.
template <int BN, bool OUT_F32, bool RELU, int NRES, int CTAS, int LN, bool XBS = false>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmap_a, const __grid_constant__ CUtensorMap tmap_b,
                 const __grid_constant__ CUtensorMap tmap_c, const __grid_constant__ CUtensorMap tmap_r1,
                 const __grid_constant__ CUtensorMap tmap_r2, const Params p) {
  // (XBS: tmap_r2 is not a residual - NRES is 1 - but the 32 x 32 map of the bf16 copy xb)
  using C = Cfg<BN, NRES, CTAS, OUT_F32, XBS>;
  extern __shared__ unsigned char smem_raw[];
  const uint32_t smem_base = (smem_u32(smem_raw) + 1023u) & ~1023u;  // SWIZZLE_128B wants 1024B alignment
  const uint32_t slab_base = smem_base + C::STAGES * C::STAGE_BYTES;  // 1024-aligned: STAGE_BYTES is a multiple of 1024
  const uint32_t cvec_base = slab_base + C::SLABS_BYTES;  // [bias[N] | colsum[N]] (bf16 output only)
  const uint32_t bar_base = cvec_base + C::CVEC_BYTES;
  auto full_bar = [&](int s) { return bar_base + 8u * s; };
  auto empty_bar = [&](int s) { return bar_base + 8u * (C::STAGES + s); };
  auto tfull_bar = [&](int a) { return bar_base + 8u * (2 * C::STAGES + a); };
  auto tempty_bar = [&](int a) { return bar_base + 8u * (2 * C::STAGES + 2 + a); };
  const uint32_t tmem_slot = bar_base + 8u * (2 * C::STAGES + 4);
  auto res_bar = [&](int w, int slot) { return bar_base + 8u * (2 * C::STAGES + 5 + w * RES_RING_MAX + slot); };
  static_assert(8 * (2 * C::STAGES + 5 + EPI_WARPS * RES_RING_MAX) <= BAR_BYTES, "barrier region too small");
  volatile uint32_t* tmem_slot_ptr =
      reinterpret_cast<volatile uint32_t*>(smem_raw + (tmem_slot - smem_u32(smem_raw)));

  pdl_launch_dependents();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int num_kb = (p.K + BK - 1) / BK;
  const int num_tiles = p.tiles_m * p.tiles_n;  // tiles of (BM * CTAS) x BN
  const uint32_t rank = CTAS == 2 ? cluster_ctarank() : 0u;  // position in the CTA pair; rank 0 issues the MMAs
  const int cl = CTAS == 2 ? static_cast<int>(blockIdx.x >> 1) : static_cast<int>(blockIdx.x);    // tile-scheduler slot
  const int ncl = CTAS == 2 ? static_cast<int>(gridDim.x >> 1) : static_cast<int>(gridDim.x);
  auto tile_row = [&](int tm) { return (tm * CTAS + static_cast<int>(rank)) * BM; };  // first row this CTA owns in tile row tm
  auto tm_of = [&](int tile) { return tile / p.tiles_n; };

  if (warp == 0 && lane == 0) {
    prefetch_tmap(&tmap_a);
    prefetch_tmap(&tmap_b);
    prefetch_tmap(&tmap_c);
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(full_bar(s), 1);
      mbar_init(empty_bar(s), 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(tfull_bar(a), 1);
      mbar_init(tempty_bar(a), EPI_WARPS * CTAS);  // one arrive per epilogue warp (of both CTAs of a pair)
    }
    if (NRES > 0) {
      prefetch_tmap(&tmap_r1);
      if (NRES > 1) prefetch_tmap(&tmap_r2);
      for (int w = 0; w < EPI_WARPS; ++w)
        for (int sl = 0; sl < RES_RING_MAX; ++sl) mbar_init(res_bar(w, sl), 1);
    }
    fence_mbar_init();
  }
  if constexpr (!OUT_F32) {
    for (int i = threadIdx.x; i < p.N; i += NUM_THREADS) {
      asm volatile("st.shared.f32 [%0], %1;" ::"r"(cvec_base + 4u * i), "f"(__ldg(p.bias + i)) : "memory");
      if constexpr (LN == 1) asm volatile("st.shared.f32 [%0], %1;" ::"r"(cvec_base + 4u * (p.N + i)), "f"(__ldg(p.colsum + i)) : "memory");
    }
  }
  if (warp == 1) {
    if constexpr (CTAS == 2) tmem_alloc_pair<C::TMEM_COLS>(tmem_slot);
    else tmem_alloc<C::TMEM_COLS>(tmem_slot);
  }
  tc_fence_before();
  if constexpr (CTAS == 2) cluster_sync(); else __syncthreads();  // barriers of BOTH CTAs are initialised past this point
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;
  // The weight tiles of this CTA's first k-blocks do not depend on the previous kernel: their TMA loads go out BEFORE the
  // grid dependency resolves, so in a PDL chain (and in a replayed graph) the operand ring is already half full when the
  // activations become visible.  Small batches are a chain of ~10 us kernels; this takes a load round trip off each.
  const int pre_kb = (cl < num_tiles) ? (num_kb < C::STAGES ? num_kb : C::STAGES) : 0;
  if (warp == 0 && lane == 0) {
    const int tn = cl % p.tiles_n;
    for (int kb = 0; kb < pre_kb; ++kb) {
      const uint32_t sb = smem_base + kb * C::STAGE_BYTES + C::A_BYTES;
      if constexpr (CTAS == 2) {
        if (rank == 0) mbar_expect_tx(full_bar(kb), 2 * C::STAGE_BYTES);
        tma_load_2d_pair(sb, &tmap_b, mapa(full_bar(kb), 0), kb * BK, tn * BN + static_cast<int>(rank) * (BN / 2));
      } else {
        mbar_expect_tx(full_bar(kb), C::STAGE_BYTES);
        tma_load_2d(sb, &tmap_b, full_bar(kb), kb * BK, tn * BN);
      }
    }
  }
  pdl_wait();  // prologue done (weights / bias are constants); activations of the previous kernel are visible from here

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int tile = cl; tile < num_tiles; tile += ncl) {
        const int tm = tm_of(tile), tn = tile % p.tiles_n;
        for (int kb = 0; kb < num_kb; ++kb) {
          // (an L2 prefetch of the A stream 8-32 k-blocks ahead was measured 8-12 % SLOWER on every shape: the ring is not
          // waiting on HBM latency)
          mbar_wait(empty_bar(stage), phase ^ 1);
          const uint32_t sa = smem_base + stage * C::STAGE_BYTES;
          const uint32_t sb = sa + C::A_BYTES;
          if (tile == cl && kb < pre_kb) {  // weights already under way (above): only the activations are missing
            if constexpr (CTAS == 2) tma_load_2d_pair(sa, &tmap_a, mapa(full_bar(stage), 0), kb * BK, p.row0 + tile_row(tm));
            else tma_load_2d(sa, &tmap_a, full_bar(stage), kb * BK, p.row0 + tm * BM);
          } else if constexpr (CTAS == 2) {
            // both CTAs' bytes are counted on the LEADER's full barrier (the only one the MMA warp waits on)
            const uint32_t lead_bar = mapa(full_bar(stage), 0);
            if (rank == 0) mbar_expect_tx(full_bar(stage), 2 * C::STAGE_BYTES);
            tma_load_2d_pair(sa, &tmap_a, lead_bar, kb * BK, p.row0 + tile_row(tm));
            tma_load_2d_pair(sb, &tmap_b, lead_bar, kb * BK, tn * BN + static_cast<int>(rank) * (BN / 2));
          } else {
            mbar_expect_tx(full_bar(stage), C::STAGE_BYTES);
            tma_load_2d(sa, &tmap_a, full_bar(stage), kb * BK, p.row0 + tm * BM);
            tma_load_2d(sb, &tmap_b, full_bar(stage), kb * BK, tn * BN);
          }
          if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer (the pair's leader only) =====================
    constexpr uint32_t idesc = make_idesc(BM * CTAS, BN, 0);
    int stage = 0;
    uint32_t phase = 0;
    int it = 0;
    for (int tile = (rank == 0 ? cl : num_tiles); tile < num_tiles; tile += ncl, ++it) {
      const int acc = it & 1;
      const uint32_t acc_phase = (it >> 1) & 1;
      mbar_wait(tempty_bar(acc), acc_phase ^ 1);  // epilogue has drained this accumulator
      tc_fence_after();
      const uint32_t tmem_d = tmem_base + acc * BN;
      for (int kb = 0; kb < num_kb; ++kb) {
        mbar_wait(full_bar(stage), phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = smem_base + stage * C::STAGE_BYTES;
          const uint32_t sb = sa + C::A_BYTES;
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t adesc = make_smem_desc(sa + k * UMMA_K * 2);
            const uint64_t bdesc = make_smem_desc(sb + k * UMMA_K * 2);
            if constexpr (CTAS == 2) umma_f16_pair(tmem_d, adesc, bdesc, idesc, (kb | k) != 0);
            else umma_f16(tmem_d, adesc, bdesc, idesc, (kb | k) != 0);
          }
          if constexpr (CTAS == 2) {
            umma_commit_pair(empty_bar(stage));                       // frees the slot in BOTH CTAs
            if (kb == num_kb - 1) umma_commit_pair(tfull_bar(acc));   // both epilogues may drain their half
          } else {
            umma_commit(empty_bar(stage));                       // frees the smem slot when the MMAs retire
            if (kb == num_kb - 1) umma_commit(tfull_bar(acc));   // accumulator complete
          }
        }
        __syncwarp();
        if (++stage == C::STAGES) { stage = 0; phase ^= 1; }
      }
    }
  } else {
    // ===================== epilogue warps =====================
    // Warp w owns TMEM lanes [32 (w & 3), +32) = 32 rows of the tile (hardware rule: lane quarter = warp id mod 4) and
    // every other "unit" (64 bf16 or 32 f32 output columns = one 128-byte row segment) of those rows.  Per unit it fills
    // a 32x128B slab in the SWIZZLE_128B layout and hands it to TMA, so every global access of the epilogue is a
    // full-line bulk access.
    const int q = warp & 3;
    const int ew = warp - 2;
    const int hf = ew >> 2;  // which half of the units
    constexpr int COLS = OUT_F32 ? 32 : 64;
    constexpr int UNITS = BN / COLS;
    const uint32_t rsw = static_cast<uint32_t>(lane & 7);  // swizzle phase of this lane's row
    const uint32_t row_off = static_cast<uint32_t>(lane) * 128u;
    // a 32-row box that straddles the end of this launch's rows (the last rows of a slab) must not be written by TMA:
    // rows >= M belong to the next slab, which may be running concurrently on another stream.  Lane = row.
    auto store_rows_guarded = [&](uint32_t slab, int row0, int col0) {
      __syncwarp();
      if (row0 + lane < p.M) {
        unsigned char* g = static_cast<unsigned char*>(p.c) +
                           (static_cast<long long>(p.row0 + row0 + lane) * p.ldc + col0) * (OUT_F32 ? 4 : 2);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float4 v = lds128(slab + row_off + ((static_cast<uint32_t>(j) ^ rsw) << 4));
          float4* gp = reinterpret_cast<float4*>(g + 16 * j);
          if (OUT_F32 && p.accumulate) {  // these elements have no other writer in this launch
            const float4 o = *gp;
            v.x += o.x; v.y += o.y; v.z += o.z; v.w += o.w;
          }
          *gp = v;
        }
      }
      fence_async_smem();
      __syncwarp();
    };
    if constexpr (OUT_F32 && NRES >= 1) {
      // ---- f32 result with TMA-prefetched residuals.  This warp's units form one flat sequence n = it * UH + i over its
      // tiles; unit n lives in ring slot n % RES_RING.
      constexpr int UH = UNITS / 2;
      constexpr int RES_RING = C::RING;
      constexpr int LAG = RES_RING >= 3 ? 1 : 0;  // stores allowed to be still reading their slab when the next load is issued
      constexpr uint32_t SLOT_BYTES = NRES * SLAB_BYTES;
      const uint32_t ring = slab_base + ew * RES_RING * SLOT_BYTES;
      [[maybe_unused]] const uint32_t xslab = slab_base + EPI_WARPS * RES_RING * SLOT_BYTES + ew * C::XSLAB_BYTES;
      [[maybe_unused]] const uint32_t xrow_off = static_cast<uint32_t>(lane) * 64u;
      [[maybe_unused]] const uint32_t xsw = static_cast<uint32_t>(lane >> 1) & 3u;  // 64B swizzle: chunk ^ bits [8:7] of the offset
      const int my_tiles = (num_tiles - cl + ncl - 1) / ncl;
      const int total = my_tiles * UH;
      auto coords = [&](int n, int& row0, int& col0, int& u) {
        const int it = n / UH, i = n - it * UH;
        const int tile = cl + it * ncl;
        const int tm = tm_of(tile), tn = tile % p.tiles_n;
        u = hf + 2 * i;
        row0 = tile_row(tm) + q * 32;
        col0 = tn * BN + u * COLS;
      };
      auto issue_load = [&](int n) {  // lane 0 only
        if (n >= total) return;
        int row0, col0, u;
        coords(n, row0, col0, u);
        const int slot = n % RES_RING;
        const uint32_t dst = ring + slot * SLOT_BYTES;
        mbar_expect_tx(res_bar(ew, slot), SLOT_BYTES);
        tma_load_2d(dst, &tmap_r1, res_bar(ew, slot), col0, p.row0 + row0);
        if constexpr (NRES >= 2) tma_load_2d(dst + SLAB_BYTES, &tmap_r2, res_bar(ew, slot), col0, p.row0 + row0);
      };
      if (lane == 0) {
#pragma unroll
        for (int n = 0; n < RES_RING - LAG; ++n) issue_load(n);
      }
      // bf16 residual (p.rb): this lane's row, the 32 columns of a unit = four 16-byte loads, requested one unit ahead
      auto ld_rb = [&](int n_, uint4 (&v)[4]) {
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = make_uint4(0u, 0u, 0u, 0u);
        if (p.rb == nullptr || n_ >= total) return;
        int row0_, col0_, u_;
        coords(n_, row0_, col0_, u_);
        if (row0_ + lane < p.M) {
          const uint4* g = reinterpret_cast<const uint4*>(p.rb + static_cast<long long>(p.row0 + row0_ + lane) * p.ld_rb + col0_);
#pragma unroll
          for (int k = 0; k < 4; ++k) v[k] = ldg_nc_v4(g + k);
        }
      };
      uint4 rbv[4];
      ld_rb(0, rbv);
      for (int n = 0; n < total; ++n) {
        [[maybe_unused]] float ln_s = 0.f, ln_ss = 0.f;  // LN == 2: this lane's row, sums over the 32 columns of this unit
        uint4 rbn[4];
        ld_rb(n + 1, rbn);
        const int it = n / UH, i = n - it * UH;
        const int acc = it & 1;
        int row0, col0, u;
        coords(n, row0, col0, u);
        const int slot = n % RES_RING;
        const uint32_t slab = ring + slot * SLOT_BYTES;
        if (i == 0) {
          mbar_wait(tfull_bar(acc), (it >> 1) & 1);
          tc_fence_after();
        }
        mbar_wait(res_bar(ew, slot), (n / RES_RING) & 1);  // residual slab(s) of this unit have landed
        const uint32_t tmem_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
        uint32_t r[32];
        tmem_ld32_nowait(tmem_row + u * 32, r);
        tmem_ld_wait();
        const float4* bp = reinterpret_cast<const float4*>(p.bias + col0);
        [[maybe_unused]] uint32_t xbp[4];
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const uint32_t addr = slab + row_off + ((static_cast<uint32_t>(j) ^ rsw) << 4);
          const float4 bv = __ldg(bp + j);
          float4 v = lds128(addr);
          v.x += __uint_as_float(r[4 * j]) + bv.x;
          v.y += __uint_as_float(r[4 * j + 1]) + bv.y;
          v.z += __uint_as_float(r[4 * j + 2]) + bv.z;
          v.w += __uint_as_float(r[4 * j + 3]) + bv.w;
          if constexpr (NRES >= 2) {
            const float4 w2 = lds128(addr + SLAB_BYTES);
            v.x += w2.x; v.y += w2.y; v.z += w2.z; v.w += w2.w;
          }
          {  // bf16 residual: columns 4j .. 4j+3 = half of rbv[j / 2] (zeros when there is none)
            const uint32_t m0 = (j & 1) ? rbv[j >> 1].z : rbv[j >> 1].x, m1 = (j & 1) ? rbv[j >> 1].w : rbv[j >> 1].y;
            v.x += __uint_as_float(m0 << 16); v.y += __uint_as_float(m0 & 0xffff0000u);
            v.z += __uint_as_float(m1 << 16); v.w += __uint_as_float(m1 & 0xffff0000u);
          }
          sts128(addr, __float_as_uint(v.x), __float_as_uint(v.y), __float_as_uint(v.z), __float_as_uint(v.w));
          if constexpr (LN == 2) {  // bf16 copy for the next GEMM + the statistics of exactly those bf16 values
            const uint32_t lo = pack_bf16x2(v.x, v.y), hi = pack_bf16x2(v.z, v.w);
            xbp[2 * (j & 1)] = lo;
            xbp[2 * (j & 1) + 1] = hi;
            const float b0 = __uint_as_float(lo << 16), b1 = __uint_as_float(lo & 0xffff0000u);
            const float b2 = __uint_as_float(hi << 16), b3 = __uint_as_float(hi & 0xffff0000u);
            ln_s += (b0 + b1) + (b2 + b3);
            ln_ss = fmaf(b0, b0, fmaf(b1, b1, fmaf(b2, b2, fmaf(b3, b3, ln_ss))));
            if (j & 1) {
              if (XBS && row0 + 32 <= p.M) sts128(xslab + xrow_off + ((static_cast<uint32_t>(j >> 1) ^ xsw) << 4), xbp[0], xbp[1], xbp[2], xbp[3]);
              else if (row0 + lane < p.M)  // (ragged last rows of the launch always go straight to global memory)
                *reinterpret_cast<uint4*>(p.xb + static_cast<long long>(p.row0 + row0 + lane) * p.ld_xb + col0 + 4 * (j - 1)) =
                    make_uint4(xbp[0], xbp[1], xbp[2], xbp[3]);
            }
          }
        }
        if constexpr (LN == 2) {
          // one partial per 32-column unit, whatever the tile width: the consumer adds the N / 32 partials of a row in a
          // fixed order, so the statistics (and everything downstream) do not depend on the tile choice of this launch
          if (row0 + lane < p.M)
            p.ln_stats[static_cast<long long>(p.row0 + row0 + lane) * p.ln_slots + (col0 >> 5)] = make_float2(ln_s, ln_ss);
        }
        if (i == UH - 1) {  // last TMEM read of this tile: hand the accumulator back to the MMA warp
          tc_fence_before();
          __syncwarp();
          if (lane == 0) {
            if constexpr (CTAS == 2) mbar_arrive_cluster(mapa(tempty_bar(acc), 0));
            else mbar_arrive(tempty_bar(acc));
          }
        }
        if (row0 + 32 <= p.M) {
          fence_async_smem();  // generic-proxy smem writes -> visible to the TMA (async proxy)
          __syncwarp();
          if (lane == 0) {
            if (p.accumulate) tma_reduce_add_2d(&tmap_c, slab, col0, p.row0 + row0);
            else tma_store_2d(&tmap_c, slab, col0, p.row0 + row0);
            if constexpr (XBS) tma_store_2d(&tmap_r2, xslab, col0, p.row0 + row0);
            bulk_commit();
            bulk_wait_read<LAG>();  // the store that left slot (n + RES_RING - LAG) % RES_RING has read its slab
          }
        } else {
          store_rows_guarded(slab, row0, col0);
          if (lane == 0) bulk_wait_read<0>();
        }
        if (lane == 0) issue_load(n + RES_RING - LAG);
#pragma unroll
        for (int k = 0; k < 4; ++k) rbv[k] = rbn[k];
        __syncwarp();
      }
    } else {
      // two slabs per warp: unit n fills slab n & 1 while the TMA store of unit n - 1 still reads the other one
      // folded-LN consumer: the row partials (sum, sum of squares) of a tile are fetched one tile ahead
      // folded-LN consumer: the row partials (sum, sum of squares per 32-column unit of the producer's output) of a tile are
      // fetched one tile ahead and added in slot order
      [[maybe_unused]] float4 st[8];
      [[maybe_unused]] auto ld_stats = [&](int tile_) {
#pragma unroll
        for (int k = 0; k < 8; ++k) st[k] = make_float4(0.f, 0.f, 0.f, 0.f);
        const int r_ = tile_row(tm_of(tile_)) + q * 32 + lane;
        if (r_ >= p.M) return;
        const float2* sp = p.ln_stats + static_cast<long long>(p.row0 + r_) * p.ln_slots;  // written by the previous kernel
        if (p.ln_slots == 16) {  // 512 columns: eight 16-byte loads
#pragma unroll
          for (int k = 0; k < 8; ++k) st[k] = *reinterpret_cast<const float4*>(sp + 2 * k);
        } else {
          float s_ = 0.f, ss_ = 0.f;
          for (int k = 0; k < p.ln_slots; ++k) {
            const float2 t = sp[k];
            s_ += t.x;
            ss_ += t.y;
          }
          st[0] = make_float4(s_, ss_, 0.f, 0.f);
        }
      };
      if constexpr (LN == 1) {
        if (cl < num_tiles) ld_stats(cl);
      }
      constexpr int NSLAB = C::SLABS_PER_WARP;
      const uint32_t slab2 = slab_base + ew * NSLAB * SLAB_BYTES;
      int pending = 0;  // bulk groups (TMA stores) this warp's lane 0 has committed and not yet waited for
      int n = 0;        // units this warp has produced
      int it = 0;
      for (int tile = cl; tile < num_tiles; tile += ncl, ++it) {
        const int tm = tm_of(tile), tn = tile % p.tiles_n;
        const int acc = it & 1;
        const uint32_t acc_phase = (it >> 1) & 1;
        const int row0 = tile_row(tm) + q * 32;
        const uint32_t tmem_row = tmem_base + (static_cast<uint32_t>(q * 32) << 16) + acc * BN;
        [[maybe_unused]] float ln_rs = 1.f, ln_nm = 0.f;  // LN == 1: rstd and -rstd * mean of this lane's row
        if constexpr (LN == 1) {
          // partials of THIS tile's row were requested one tile ago (or before the loop); request the next tile's now
          float s = 0.f, ss = 0.f;
#pragma unroll
          for (int k = 0; k < 8; ++k) {
            s += st[k].x; ss += st[k].y;
            s += st[k].z; ss += st[k].w;
          }
          if (tile + ncl < num_tiles) ld_stats(tile + ncl);
          const float mu = s * p.ln_inv_dim;
          ln_rs = rsqrtf(fmaxf(fmaf(-mu, mu, ss * p.ln_inv_dim), 0.f) + 1e-12f);  // VLLM/model_executor/models/funasr.py:85 eps
          ln_nm = -ln_rs * mu;
        }
        mbar_wait(tfull_bar(acc), acc_phase);
        tc_fence_after();
#pragma unroll 1
        for (int u = hf; u < UNITS; u += 2, ++n) {
          const int col0 = tn * BN + u * COLS;
          const uint32_t slab = slab2 + (n % NSLAB) * SLAB_BYTES;
          if constexpr (OUT_F32) {
            uint32_t r[32];
            tmem_ld32_nowait(tmem_row + u * 32, r);
            float4 bv[8];
            const float4* bp = reinterpret_cast<const float4*>(p.bias + col0);
#pragma unroll
            for (int j = 0; j < 8; ++j) bv[j] = __ldg(bp + j);
            uint4 rbv[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) rbv[k] = make_uint4(0u, 0u, 0u, 0u);
            if (p.rb != nullptr && row0 + lane < p.M) {  // bf16 residual of this lane's row (see the residual path above)
              const uint4* g = reinterpret_cast<const uint4*>(p.rb + static_cast<long long>(p.row0 + row0 + lane) * p.ld_rb + col0);
#pragma unroll
              for (int k = 0; k < 4; ++k) rbv[k] = ldg_nc_v4(g + k);
            }
            if (pending >= NSLAB) {  // the store that read this slab NSLAB units ago must have drained it
              if (lane == 0) bulk_wait_read<NSLAB - 1>();
              pending = NSLAB - 1;
            }
            __syncwarp();
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 8; ++j) {
              const uint32_t m0 = (j & 1) ? rbv[j >> 1].z : rbv[j >> 1].x, m1 = (j & 1) ? rbv[j >> 1].w : rbv[j >> 1].y;
              sts128(slab + row_off + ((static_cast<uint32_t>(j) ^ rsw) << 4),
                     __float_as_uint(__uint_as_float(r[4 * j]) + bv[j].x + __uint_as_float(m0 << 16)),
                     __float_as_uint(__uint_as_float(r[4 * j + 1]) + bv[j].y + __uint_as_float(m0 & 0xffff0000u)),
                     __float_as_uint(__uint_as_float(r[4 * j + 2]) + bv[j].z + __uint_as_float(m1 << 16)),
                     __float_as_uint(__uint_as_float(r[4 * j + 3]) + bv[j].w + __uint_as_float(m1 & 0xffff0000u)));
            }
          } else {
            uint32_t r0[32], r1[32];
            tmem_ld32_nowait(tmem_row + u * 64, r0);
            tmem_ld32_nowait(tmem_row + u * 64 + 32, r1);
            if (pending >= NSLAB) {
              if (lane == 0) bulk_wait_read<NSLAB - 1>();
              pending = NSLAB - 1;
            }
            __syncwarp();
            tmem_ld_wait();
#pragma unroll
            for (int half = 0; half < 2; ++half) {
              const uint32_t(&r)[32] = half ? r1 : r0;
              const uint32_t bp = cvec_base + 4u * (col0 + half * 32);
              [[maybe_unused]] const uint32_t cp = bp + 4u * p.N;
#pragma unroll
              for (int j = 0; j < 4; ++j) {
                const float4 b0 = lds128(bp + 32 * j), b1 = lds128(bp + 32 * j + 16);
                float v[8];
                if constexpr (LN == 1) {  // rstd * (acc - mean * colsum) + bias
                  const float4 c0 = lds128(cp + 32 * j), c1 = lds128(cp + 32 * j + 16);
                  const float cs[8] = {c0.x, c0.y, c0.z, c0.w, c1.x, c1.y, c1.z, c1.w};
                  const float bb[8] = {b0.x, b0.y, b0.z, b0.w, b1.x, b1.y, b1.z, b1.w};
#pragma unroll
                  for (int e = 0; e < 8; e += 2)  // two packed f32x2 FMAs per column pair (Blackwell FFMA2)
                    ffma2(v[e], v[e + 1], ln_rs, __uint_as_float(r[8 * j + e]), __uint_as_float(r[8 * j + e + 1]), ln_nm, cs[e], cs[e + 1],
                          bb[e], bb[e + 1]);
                } else {
                  v[0] = __uint_as_float(r[8 * j]) + b0.x;     v[1] = __uint_as_float(r[8 * j + 1]) + b0.y;
                  v[2] = __uint_as_float(r[8 * j + 2]) + b0.z; v[3] = __uint_as_float(r[8 * j + 3]) + b0.w;
                  v[4] = __uint_as_float(r[8 * j + 4]) + b1.x; v[5] = __uint_as_float(r[8 * j + 5]) + b1.y;
                  v[6] = __uint_as_float(r[8 * j + 6]) + b1.z; v[7] = __uint_as_float(r[8 * j + 7]) + b1.w;
                }
                if constexpr (RELU) {
#pragma unroll
                  for (int e = 0; e < 8; ++e) v[e] = fmaxf(v[e], 0.f);
                }
                sts128(slab + row_off + ((static_cast<uint32_t>(half * 4 + j) ^ rsw) << 4), pack_bf16x2(v[0], v[1]),
                       pack_bf16x2(v[2], v[3]), pack_bf16x2(v[4], v[5]), pack_bf16x2(v[6], v[7]));
              }
            }
          }
          if (u + 2 >= UNITS) {  // last TMEM read of this tile: hand the accumulator back before the stores go out
            tc_fence_before();
            __syncwarp();
            if (lane == 0) {
              if constexpr (CTAS == 2) mbar_arrive_cluster(mapa(tempty_bar(acc), 0));
              else mbar_arrive(tempty_bar(acc));
            }
          }
          if (row0 + 32 <= p.M) {
            fence_async_smem();  // generic-proxy smem writes -> visible to the TMA (async proxy)
            __syncwarp();
            if (lane == 0) {
              if (OUT_F32 && p.accumulate) tma_reduce_add_2d(&tmap_c, slab, col0, p.row0 + row0);
              else tma_store_2d(&tmap_c, slab, col0, p.row0 + row0);
              bulk_commit();
            }
            ++pending;
          } else {
            store_rows_guarded(slab, row0, col0);
            if (lane == 0) bulk_wait_read<0>();  // rare (ragged last row tile): restart the two-slab bookkeeping
            pending = 0;
            __syncwarp();
          }
        }
      }
    }
    if (lane == 0) bulk_wait<0>();
    __syncwarp();
  }

  tc_fence_before();
  if constexpr (CTAS == 2) cluster_sync(); else __syncthreads();  // the leader's MMAs read the peer's smem until here
  if (warp == 1) {
    tc_fence_after();
    if constexpr (CTAS == 2) tmem_dealloc_pair<C::TMEM_COLS>(tmem_base);
    else tmem_dealloc<C::TMEM_COLS>(tmem_base);
  }
}

// ---------------------------------------------------------------- host
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* ptr = nullptr;
  cudaDriverEntryPointQueryResult qres;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) != cudaSuccess ||
      qres != cudaDriverEntryPointSuccess || ptr == nullptr)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(ptr);
  return fn;
}

template <int BN, bool OUT_F32, bool RELU, int NRES, int CTAS, int LN, bool XBS = false>
static int launch_v(const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mc, const CUtensorMap& mr1,
                    const CUtensorMap& mr2, const Params& p, int grid, cudaStream_t stream) {
  using C = Cfg<BN, NRES, CTAS, OUT_F32, XBS>;
  auto kern = gemm_bf16_kernel<BN, OUT_F32, RELU, NRES, CTAS, LN, XBS>;
  static SmemOptIn opt;  // one per kernel instantiation
  if (int r = opt_in_smem(kern, C::SMEM_BYTES, opt)) return r;
  // the CTA pair of cta_group::2 is one cluster
  FA_CHECK_CUDA(launch_pdl(kern, dim3(grid), dim3(NUM_THREADS), C::SMEM_BYTES, stream, CTAS, ma, mb, mc, mr1, mr2, p));
  return 0;
}

// pair == true: BN is 256 and `mb` boxes BN/2 rows (each CTA of a pair stages half of the B tile)
template <int BN, bool OUT_F32, bool RELU, int NRES, int LN = 0, bool XBS = false>
static int launch_p(bool pair, const CUtensorMap& ma, const CUtensorMap& mb, const CUtensorMap& mc, const CUtensorMap& mr1,
                    const CUtensorMap& mr2, const Params& p, int grid, cudaStream_t stream) {
  if constexpr (BN == 256) {
    if (pair) return launch_v<BN, OUT_F32, RELU, NRES, 2, LN, XBS>(ma, mb, mc, mr1, mr2, p, grid, stream);
  }
  return launch_v<BN, OUT_F32, RELU, NRES, 1, LN, XBS>(ma, mb, mc, mr1, mr2, p, grid, stream);
}

template <int BN>
static int launch_t(const TensorMap2D& ta, const TensorMap2D& tb, const TensorMap2D& tc, const TensorMap2D* tr1,
                    const TensorMap2D* tr2, int M, int N, int K, const GemmEpilogue& epi, void* c, int64_t ldc, int row0,
                    bool pair, cudaStream_t stream) {
  Params p;
  p.M = M; p.N = N; p.K = K;
  p.row0 = row0;
  FA_REQUIRE(!pair || BN == 256, "gemm: the CTA-pair kernel is compiled for tile N 256 only");
  p.tiles_m = ceil_div(M, pair ? 2 * BM : BM);
  p.tiles_n = N / BN;
  p.bias = epi.bias; p.res1 = epi.res1; p.res2 = epi.res2;
  p.c = c; p.ldc = ldc; p.relu = epi.relu; p.out_f32 = epi.out_f32; p.accumulate = epi.accumulate;
  p.ln_stats = epi.ln_stats; p.colsum = epi.colsum; p.xb = epi.xb; p.ld_xb = epi.ld_xb;
  p.rb = epi.res_bf16; p.ld_rb = epi.ld_res_bf16;
  FA_REQUIRE(!epi.res_bf16 || (epi.out_f32 && !epi.res2 && epi.ld_res_bf16 % 8 == 0 && (reinterpret_cast<uintptr_t>(epi.res_bf16) & 15) == 0),
             "gemm: the bf16 residual needs the f32 output, at most one f32 residual, 16-byte aligned rows");
  p.ln_slots = epi.ln_mode == 2 ? N / 32 : epi.ln_slots;
  p.ln_inv_dim = epi.ln_dim > 0 ? 1.f / static_cast<float>(epi.ln_dim) : 0.f;
  FA_REQUIRE(epi.out_f32 || N * 4 * (epi.ln_mode == 1 ? 2 : 1) <= 16384, "gemm: N=%d too wide for the staged epilogue constants", N);
  if (epi.ln_mode == 1)
    FA_REQUIRE(!epi.out_f32 && epi.ln_stats && epi.colsum && epi.ln_slots > 0 && epi.ln_dim > 0, "gemm: folded-LN consumer needs the bf16 output, ln_stats, colsum, ln_slots and ln_dim");
  if (epi.ln_mode == 2)
    FA_REQUIRE(epi.out_f32 && epi.res1 && !epi.res2 && !epi.accumulate && epi.ln_stats && epi.xb && epi.ld_xb % 8 == 0 && N / 32 <= 16,
               "gemm: folded-LN producer needs the f32 output with exactly one staged residual, ln_stats and xb");
  FA_REQUIRE(!epi.accumulate || epi.out_f32, "gemm: accumulate needs the f32 output");
  const int tiles = p.tiles_m * p.tiles_n;
  int grid = tiles < sm_count() ? tiles : sm_count();
  if (pair) grid = 2 * (tiles < sm_count() / 2 ? tiles : sm_count() / 2);
  CUtensorMap ma, mb, mc, mr1, mr2;
  static_assert(sizeof(CUtensorMap) == sizeof(TensorMap2D::bytes), "CUtensorMap size");
  memcpy(&ma, ta.bytes, sizeof(ma));
  memcpy(&mb, tb.bytes, sizeof(mb));
  memcpy(&mc, tc.bytes, sizeof(mc));
  memcpy(&mr1, (tr1 ? tr1 : &tc)->bytes, sizeof(mr1));
  memcpy(&mr2, (tr2 ? tr2 : &tc)->bytes, sizeof(mr2));
  // (inside the cfg5 step: 43.9 / 44.2 ms with it, 44.6 / 44.7 ms without)
  const bool xbs = epi.ln_mode == 2 && K <= 1024;  // short k loop: the bf16 copy leaves by TMA (see Cfg::XSLAB_BYTES)
  if (xbs) {  // its 32 x 32 map travels in the (unused) second-residual slot
    TensorMap2D txb_local;
    const TensorMap2D* txb = epi.txb;
    if (txb == nullptr) {
      if (int r = make_tensor_map_w(&txb_local, epi.xb, 2, static_cast<uint64_t>(row0) + M, N, epi.ld_xb, 32, 64)) return r;
      txb = &txb_local;
    }
    memcpy(&mr2, txb->bytes, sizeof(mr2));
  }
  FA_REQUIRE((epi.res1 == nullptr) == (tr1 == nullptr) && (epi.res2 == nullptr) == (tr2 == nullptr),
             "gemm: every residual needs its tensor map");
  ProfScope _ps(K_GEMM, stream);
  if (epi.out_f32) {
    FA_REQUIRE(!epi.relu, "gemm: ReLU is only compiled for the bf16 output");
    FA_REQUIRE(epi.res1 || !epi.res2, "gemm: res2 without res1");
    if (epi.res2) {
      if constexpr (BN <= 128) return launch_p<BN, true, false, 2>(pair, ma, mb, mc, mr1, mr2, p, grid, stream);
      else FA_REQUIRE(false, "gemm: two residuals are compiled for tile N <= 128 only");
    }
    if (epi.res1 && epi.ln_mode == 2)
      return xbs ? launch_p<BN, true, false, 1, 2, true>(pair, ma, mb, mc, mr1, mr2, p, grid, stream)
                 : launch_p<BN, true, false, 1, 2>(pair, ma, mb, mc, mr1, mr2, p, grid, stream);
    if (epi.res1) return launch_p<BN, true, false, 1>(pair, ma, mb, mc, mr1, mr2, p, grid, stream);
    return launch_p<BN, true, false, 0>(pair, ma, mb, mc, mr1, mr2, p, grid, stream);
  }
  FA_REQUIRE(!epi.res1 && !epi.res2, "gemm: residuals are only compiled for the f32 output");
  if (epi.ln_mode == 1) {
    if (epi.relu) return launch_p<BN, false, true, 0, 1>(pair, ma, mb, mc, mr1, mr2, p, grid, stream);
    return launch_p<BN, false, false, 0, 1>(pair, ma, mb, mc, mr1, mr2, p, grid, stream);
  }
  if (epi.relu) return launch_p<BN, false, true, 0>(pair, ma, mb, mc, mr1, mr2, p, grid, stream);
  return launch_p<BN, false, false, 0>(pair, ma, mb, mc, mr1, mr2, p, grid, stream);
}

}  // namespace gemm

// rows x cols matrix of bf16 (elem_bytes 2) or f32 (4), row stride ld elements; box = box_rows x box_cols with
// box_cols * elem_bytes == 128 (one SWIZZLE_128B span).
int make_tensor_map(TensorMap2D* out, const void* base, int elem_bytes, uint64_t rows, uint64_t cols, uint64_t ld,
                    uint32_t box_rows) {
  return make_tensor_map_w(out, base, elem_bytes, rows, cols, ld, box_rows, 128);
}

int make_tensor_map_w(TensorMap2D* out, const void* base, int elem_bytes, uint64_t rows, uint64_t cols, uint64_t ld,
                      uint32_t box_rows, uint32_t box_bytes) {
  gemm::EncodeTiledFn fn = gemm::get_encode_fn();
  FA_REQUIRE(fn != nullptr, "cuTensorMapEncodeTiled not available from the driver");
  FA_REQUIRE(elem_bytes == 2 || elem_bytes == 4, "TMA element size must be 2 or 4");
  FA_REQUIRE((reinterpret_cast<uintptr_t>(base) & 15) == 0, "TMA base must be 16-byte aligned");
  FA_REQUIRE((ld * elem_bytes) % 16 == 0, "TMA row pitch (%llu bytes) must be a multiple of 16", (unsigned long long)(ld * elem_bytes));
  FA_REQUIRE(box_rows >= 1 && box_rows <= 256, "TMA box rows out of range");
  CUtensorMap m;
  cuuint64_t gdim[2] = {cols, rows};
  cuuint64_t gstride[1] = {ld * elem_bytes};
  FA_REQUIRE(box_bytes == 128 || box_bytes == 64, "TMA box width must be 128 or 64 bytes");
  cuuint32_t box[2] = {static_cast<cuuint32_t>(box_bytes / elem_bytes), box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(&m, elem_bytes == 2 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2,
                  const_cast<void*>(base), gdim, gstride, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  box_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                  CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  FA_REQUIRE(r == CUDA_SUCCESS, "cuTensorMapEncodeTiled failed with CUresult %d (rows %llu cols %llu ld %llu)", (int)r,
             (unsigned long long)rows, (unsigned long long)cols, (unsigned long long)ld);
  memcpy(out->bytes, &m, sizeof(m));
  return 0;
}
int make_tensor_map_bf16(TensorMap2D* out, const void* base, uint64_t rows, uint64_t cols, uint64_t ld, uint32_t box_rows) {
  return make_tensor_map(out, base, 2, rows, cols, ld, box_rows);
}

// CTA pairs (cta_group::2, 256 x 256 tiles) for the wide tile once there are enough rows.
bool gemm_use_pair(int M, int block_n) { return block_n == 256 && M >= 1024; }

int gemm_block_n(int M, int N, int n_res) {
  // Every SM works on 128 rows x BN columns per round (a CTA pair: 256 x 256 over two SMs), tiles are dealt round-robin, so
  // a launch takes ceil(tiles / slots) rounds of a duration that grows with BN plus a fixed per-tile part (pipeline fill,
  // epilogue drain: about 48 columns' worth).  Pick the width with the least rounds x duration; ties go to the wider tile
  // (fewer B re-reads).  Two prefetched f32 residuals leave room for the 256-wide operand ring only twice: they use <= 128.
  // The result of a GEMM does not depend on the choice (same k order per element, LN partials per 32-column unit).
  const int sms = sm_count();
  int best = 0;
  long best_cost = 0;
  for (int bn = 256; bn >= 64; bn >>= 1) {
    if (N % bn != 0 || (bn == 256 && n_res >= 2)) continue;
    const bool pair = gemm_use_pair(M, bn);
    const long tiles = static_cast<long>(ceil_div(M, pair ? 2 * gemm::BM : gemm::BM)) * (N / bn);
    const long slots = pair ? sms / 2 : sms;
    const long cost = ((tiles + slots - 1) / slots) * (bn + 48);  // (a fixed part of 96 or 160 columns: within 2 % at 2.7 k and 16 k rows)
    if (best == 0 || cost < best_cost) { best = bn; best_cost = cost; }
  }
  return best;
}

int launch_gemm(const TensorMap2D& ta, const TensorMap2D& tb, const TensorMap2D& tc, const TensorMap2D* tr1,
                const TensorMap2D* tr2, int block_n, int M, int N, int K, const GemmEpilogue& epi, void* c, int64_t ldc,
                int row0, bool pair, cudaStream_t stream) {
  FA_REQUIRE(M > 0 && N > 0 && K > 0 && row0 >= 0, "gemm: empty problem %dx%dx%d", M, N, K);
  FA_REQUIRE(block_n == 64 || block_n == 128 || block_n == 256, "gemm: unsupported tile N %d", block_n);
  FA_REQUIRE(N % block_n == 0, "gemm: N=%d is not a multiple of the tile %d", N, block_n);
  FA_REQUIRE(K % 8 == 0, "gemm: K=%d must be a multiple of 8", K);
  FA_REQUIRE(epi.bias != nullptr, "gemm: bias is required (nn.Linear of the reference always has one)");
  FA_REQUIRE(ldc % 8 == 0 && (reinterpret_cast<uintptr_t>(c) & 15) == 0, "gemm: C must be 16-byte aligned, ldc %% 8 == 0");
  switch (block_n) {
    case 64: return gemm::launch_t<64>(ta, tb, tc, tr1, tr2, M, N, K, epi, c, ldc, row0, false, stream);
    case 128: return gemm::launch_t<128>(ta, tb, tc, tr1, tr2, M, N, K, epi, c, ldc, row0, false, stream);
    default: return gemm::launch_t<256>(ta, tb, tc, tr1, tr2, M, N, K, epi, c, ldc, row0, pair, stream);
  }
}

}  // namespace fa

// One GEMM with the folded-LayerNorm epilogues (GemmEpilogue::ln_mode).  ln_mode 2: res1_dev is the old residual stream,
// c_dev (f32, may alias res1_dev) receives x_new, xb_dev its bf16 copy and ln_stats_dev [M, *n_slots] the row partials
// (*n_slots is written).  ln_mode 1: a_dev is the raw bf16 stream, w/bias are LN-folded, *n_slots partials are read.
extern "C" int fa_op_gemm_ln(const void* a_dev, int64_t lda, const void* w_dev, int64_t ldw, const float* bias_dev,
                             const float* res1_dev, void* c_dev, int64_t ldc, int32_t M, int32_t N, int32_t K, int32_t relu,
                             int32_t out_f32, int32_t ln_mode, void* ln_stats_dev, int32_t* n_slots, int32_t ln_dim,
                             const float* colsum_dev, void* xb_dev, int64_t ld_xb, const void* res_bf16_dev, int64_t ld_res_bf16,
                             fa_stream_t stream) {
  using namespace fa;
  FA_REQUIRE(a_dev && w_dev && c_dev && n_slots && (ln_mode == 1 || ln_mode == 2), "gemm_ln: bad arguments");
  const int bn = gemm_block_n(M, N, res1_dev ? 1 : 0);
  FA_REQUIRE(bn != 0, "gemm: N=%d must be a multiple of 64", N);
  TensorMap2D ta, tb, tc, tr1;
  if (res1_dev) if (int r = make_tensor_map(&tr1, res1_dev, 4, M, N, ldc, 32)) return r;
  if (int r = make_tensor_map_bf16(&ta, a_dev, M, K, lda, gemm::BM)) return r;
  const bool pair = gemm_use_pair(M, bn);
  if (int r = make_tensor_map_bf16(&tb, w_dev, N, K, ldw, pair ? bn / 2 : bn)) return r;
  if (int r = make_tensor_map(&tc, c_dev, out_f32 ? 4 : 2, M, N, ldc, 32)) return r;
  GemmEpilogue e;
  e.bias = bias_dev; e.res1 = res1_dev; e.relu = relu; e.out_f32 = out_f32;
  e.ln_mode = ln_mode; e.ln_stats = static_cast<float2*>(ln_stats_dev); e.ln_dim = ln_dim; e.colsum = colsum_dev;
  e.xb = static_cast<__nv_bfloat16*>(xb_dev); e.ld_xb = ld_xb;
  e.res_bf16 = static_cast<const __nv_bfloat16*>(res_bf16_dev); e.ld_res_bf16 = ld_res_bf16;
  if (ln_mode == 2) *n_slots = N / 32;
  e.ln_slots = *n_slots;
  return launch_gemm(ta, tb, tc, res1_dev ? &tr1 : nullptr, nullptr, bn, M, N, K, e, c_dev, ldc, 0, pair,
                     static_cast<cudaStream_t>(stream));
}

extern "C" int fa_op_gemm(const void* a_dev, int64_t lda, const void* w_dev, int64_t ldw, const float* bias_dev,
                          const float* res1_dev, const float* res2_dev, const void* res_bf16_dev, int64_t ld_res_bf16,
                          void* c_dev, int64_t ldc, int32_t M, int32_t N, int32_t K, int32_t relu, int32_t out_f32,
                          fa_stream_t stream) {
  using namespace fa;
  FA_REQUIRE(a_dev && w_dev && c_dev, "gemm: NULL pointer");
  // a residual that IS the output buffer is applied as an in-place accumulate (C += ...)
  int accumulate = 0;
  if (res2_dev && res2_dev == c_dev) { res2_dev = nullptr; accumulate = 1; }
  if (res1_dev && res1_dev == c_dev) { res1_dev = res2_dev; res2_dev = nullptr; accumulate = 1; }
  const int n_res = (res1_dev ? 1 : 0) + (res2_dev ? 1 : 0);
  const int bn = gemm_block_n(M, N, n_res);
  FA_REQUIRE(bn != 0, "gemm: N=%d must be a multiple of 64", N);
  FA_REQUIRE(n_res == 0 || out_f32, "gemm: residuals need the f32 output");
  TensorMap2D ta, tb, tc, tr1, tr2;
  if (res1_dev) if (int r = make_tensor_map(&tr1, res1_dev, 4, M, N, ldc, 32)) return r;
  if (res2_dev) if (int r = make_tensor_map(&tr2, res2_dev, 4, M, N, ldc, 32)) return r;
  if (int r = make_tensor_map_bf16(&ta, a_dev, M, K, lda, gemm::BM)) return r;
  const bool pair = gemm_use_pair(M, bn);
  if (int r = make_tensor_map_bf16(&tb, w_dev, N, K, ldw, pair ? bn / 2 : bn)) return r;
  if (int r = make_tensor_map(&tc, c_dev, out_f32 ? 4 : 2, M, N, ldc, 32)) return r;
  GemmEpilogue e;
  e.bias = bias_dev; e.res1 = res1_dev; e.res2 = res2_dev; e.relu = relu; e.out_f32 = out_f32; e.accumulate = accumulate;
  e.res_bf16 = static_cast<const __nv_bfloat16*>(res_bf16_dev); e.ld_res_bf16 = ld_res_bf16;
  return launch_gemm(ta, tb, tc, res1_dev ? &tr1 : nullptr, res2_dev ? &tr2 : nullptr, bn, M, N, K, e, c_dev, ldc, 0, pair,
                     static_cast<cudaStream_t>(stream));
}
