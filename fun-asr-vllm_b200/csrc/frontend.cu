// Fused Kaldi fbank -> LFR -> CMVN frontend for sm_100a.
//
// One CTA produces FA_FE_TILE_ROWS low-frame-rate rows of one utterance:
//   PCM (global, packed offsets) -> smem -> per frame {DC removal, pre-emphasis, window,
//   512-pt real FFT as a 256-pt complex FFT (two radix-16 passes in registers, one smem
//   transpose, conjugate-pair split through warp shuffles), power, sparse mel, log}
//   -> smem log-mel tile -> LFR gather + CMVN -> global.
// PCM is read once and the [T,560] features written once: no intermediate touches HBM.
//
// Precision.  Pre-emphasis leaves the lowest FFT bins ~30x (amplitude) below the loud ones, so fp32 round-off made
// inside an FFT (relative to the LOUD bins) shows up as up to ~2e-3 in the log-mel of the lowest mel bins on long
// full-scale inputs (profiles/scripts/fft_noise_study.py) — over the 1e-3 budget.  Therefore the pre-FFT steps follow
// the reference's fp32 operation order exactly (same roundings as torchaudio: no FMA contraction), and the FFT, the
// split and |X|^2 run in fp64 (B200 issues FP64 at half the FP32 rate); mel, log, LFR and CMVN are fp32 again.
//
// Arithmetic follows torchaudio.compliance.kaldi.fbank (TA:514-645, _get_window TA:154-217) and
// funasr's apply_lfr / apply_cmvn (VLLM/transformers_utils/processors/funasr.py:40-60,23-37).
#include <math.h>

#include <cstring>

#include "common.cuh"

namespace fa {
namespace fe {

#ifndef FA_FE_MEL_BATCH
#define FA_FE_MEL_BATCH 1  // 0: one mel group at a time for every bank (measurement builds)
#endif
constexpr int FRAME_LEN = 400;
constexpr int SHIFT = 160;
constexpr int NFFT = 512;
constexpr int NBIN = NFFT / 2 + 1;
constexpr int NMEL = 80;
constexpr int LFR_M = 7;
constexpr int LFR_N = 6;
constexpr int LFR_LEFT = (LFR_M - 1) / 2;
constexpr int FEAT = LFR_M * NMEL;  // 560
constexpr int TILE_ROWS = FA_FE_TILE_ROWS;
constexpr int MAX_FRAMES = LFR_N * (TILE_ROWS - 1) + LFR_M;        // 31
constexpr int RAW_TILE_FRAMES = 30;                                  // raw log-mel mode
constexpr int MAX_SAMPLES = (MAX_FRAMES - 1) * SHIFT + FRAME_LEN;    // 5200
constexpr int THREADS = 256;
constexpr int SLOTS = THREADS / 16;  // frames in flight per CTA (one per half-warp)
constexpr int MEL_GROUPS = NMEL / 16;  // a lane of a half-warp owns mel bins hl, hl + 16, ..., one per group
constexpr int MEL_MAXJ = 5;             // four-tap steps per band (bands start on a multiple of four FFT bins): <= 20 taps
constexpr int MEL_W = MEL_GROUPS * MEL_MAXJ * 16 * 4;  // weights as [group][step][lane] float4: conflict-free 16-byte loads

struct __align__(16) Tables {
  float window[FRAME_LEN];
  double2 tw256[16];  // [t] = W256^t = (cos, -sin)(2 pi t / 256): lane t builds its pass-1 -> pass-2 twiddles W256^(t q) from it
  double2 sc512[16];  // [t] = (sin, cos)(2 pi t / 512): lane t builds its split constants for k = t + 16 p from it
  int mel_start[NMEL];       // first FFT bin of the band, rounded down to a multiple of four (16-byte power loads)
  int mel_steps[8];          // per group: four-tap steps of its widest band (uniform over the lanes; shorter bands carry zeros);
                             // [7]: 1 when every group has exactly mel_cap<true>() steps (the batched mel path), else 0
  float mel_w[MEL_W];
};

// Everything a tile needs, resolved once: by every warp for a CTA's first tile, by warp 1 one tile ahead afterwards.
struct TileDesc {
  int valid;  // 0: nothing to do (utterance shorter than one frame, or a tile past its end)
  int m;      // frames of the utterance (TA:62-66 snip_edges)
  int i0, i1; // LFR rows [i0, i1) (MODE 0)
  int f_lo;   // first frame staged
  int nF;     // frames staged
  int orow;   // output row of the utterance
  int src_lo, src_hi;  // element offset of the tile's first sample in `wav` (two 32-bit halves)
};

struct __align__(16) Smem {
  double2 tw256[16];
  double2 sc512[16];
  // 16x17 padded complex transpose buffer; the slot's power spectrum (264 floats) reuses it.  Slots are 64 bytes (16 banks) out of
  // step with each other, so that the two half-warps of a warp (two slots) store their power values into different banks.
  double2 work[SLOTS][16 * 17 + 4];
  float samples[MAX_SAMPLES + 8];
  float logmel[MAX_FRAMES + 1][NMEL];
  float window[FRAME_LEN];
  float mel_w[MEL_W];
  TileDesc desc[2];  // this tile and the CTA's next one (resolved by warp 1 during the frame phase), by tile parity
};

typedef double2 cplx;
__device__ __forceinline__ cplx mk(double re, double im) { return make_double2(re, im); }
__device__ __forceinline__ cplx cmul(cplx a, cplx b) { return mk(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }

// cos / sin (p pi / 16), p = 0..8
__device__ __host__ constexpr double cos_p16(int p) {
  constexpr double t[9] = {1.0, 0.98078528040323044913, 0.92387953251128675613, 0.83146961230254523708, 0.70710678118654752440,
                           0.55557023301960222474, 0.38268343236508977173, 0.19509032201612826785, 0.0};
  return t[p];
}
__device__ __host__ constexpr double sin_p16(int p) { return cos_p16(8 - p); }

// 4-point DFT in place: (a,b,c,d) -> (X0,X1,X2,X3), W4 = -i.
__device__ __forceinline__ void dft4(cplx& a, cplx& b, cplx& c, cplx& d) {
  const cplx s0 = mk(a.x + c.x, a.y + c.y), s1 = mk(a.x - c.x, a.y - c.y);
  const cplx s2 = mk(b.x + d.x, b.y + d.y), s3 = mk(b.x - d.x, b.y - d.y);
  a = mk(s0.x + s2.x, s0.y + s2.y);
  c = mk(s0.x - s2.x, s0.y - s2.y);
  b = mk(s1.x + s3.y, s1.y - s3.x);  // s1 - i*s3
  d = mk(s1.x - s3.y, s1.y + s3.x);  // s1 + i*s3
}

// The same with d == 0 (the zero padding 400 -> 512 of the frame): four additions fewer.
__device__ __forceinline__ void dft4_d0(cplx& a, cplx& b, cplx& c, cplx& d) {
  const cplx s0 = mk(a.x + c.x, a.y + c.y), s1 = mk(a.x - c.x, a.y - c.y), s2 = b;
  a = mk(s0.x + s2.x, s0.y + s2.y);
  c = mk(s0.x - s2.x, s0.y - s2.y);
  b = mk(s1.x + s2.y, s1.y - s2.x);
  d = mk(s1.x - s2.y, s1.y + s2.x);
}

// 16-point DFT, 4x4 Cooley-Tukey in registers.  Input x[j] natural order; output X[q] is left at
// position POS16(q) = 4*(q&3) + (q>>2).  ZPAD: x[13], x[14], x[15] are known to be zero.
#define POS16(q) (4 * ((q)&3) + ((q) >> 2))
template <bool ZPAD>
__device__ __forceinline__ void fft16(cplx (&x)[16]) {
  constexpr double C1 = 0.92387953251128675613, S1 = 0.38268343236508977173, R = 0.70710678118654752440;
  dft4(x[0], x[4], x[8], x[12]);
#pragma unroll
  for (int j0 = 1; j0 < 4; ++j0) {
    if (ZPAD) dft4_d0(x[j0], x[4 + j0], x[8 + j0], x[12 + j0]);
    else dft4(x[j0], x[4 + j0], x[8 + j0], x[12 + j0]);
  }
  // A[j0][q0] now sits at x[4*q0 + j0]; multiply by W16^(j0*q0)
  x[4 * 1 + 1] = cmul(x[4 * 1 + 1], mk(C1, -S1));  // W^1
  x[4 * 2 + 1] = mk(R * (x[4 * 2 + 1].x + x[4 * 2 + 1].y), R * (x[4 * 2 + 1].y - x[4 * 2 + 1].x));  // W^2 = R(1 - i)
  x[4 * 3 + 1] = cmul(x[4 * 3 + 1], mk(S1, -C1));  // W^3
  x[4 * 1 + 2] = mk(R * (x[4 * 1 + 2].x + x[4 * 1 + 2].y), R * (x[4 * 1 + 2].y - x[4 * 1 + 2].x));  // W^2
  x[4 * 2 + 2] = mk(x[4 * 2 + 2].y, -x[4 * 2 + 2].x);                                               // W^4 = -i
  x[4 * 3 + 2] = mk(R * (x[4 * 3 + 2].y - x[4 * 3 + 2].x), -R * (x[4 * 3 + 2].x + x[4 * 3 + 2].y));  // W^6 = -R(1 + i)
  x[4 * 1 + 3] = cmul(x[4 * 1 + 3], mk(S1, -C1));  // W^3
  x[4 * 2 + 3] = mk(R * (x[4 * 2 + 3].y - x[4 * 2 + 3].x), -R * (x[4 * 2 + 3].x + x[4 * 2 + 3].y));  // W^6
  x[4 * 3 + 3] = cmul(x[4 * 3 + 3], mk(-C1, S1));  // W^9
#pragma unroll
  for (int q0 = 0; q0 < 4; ++q0) dft4(x[4 * q0 + 0], x[4 * q0 + 1], x[4 * q0 + 2], x[4 * q0 + 3]);
}

// Utterance of tile `id`: the last b with tile_off[b] <= id.  A 32-way search (every lane probes one pivot, a ballot picks the
// interval): two dependent load rounds for up to 1024 utterances instead of ten.  Called by whole, converged warps.
__device__ __forceinline__ int utterance_of_tile(const int32_t* __restrict__ tile_off, int n_utts, int id, int lane) {
  int lo = 0, hi = n_utts;
  while (hi - lo > 1) {
    const int step = (hi - lo + 31) >> 5;
    const int pos = lo + step * lane;
    const bool le = pos < hi && __ldg(tile_off + pos) <= id;
    const unsigned m = __ballot_sync(0xffffffffu, le) | 1u;  // tile_off[lo] <= id holds by construction
    const int k = 31 - __clz(m);
    lo += step * k;
    hi = min(lo + step, hi);
  }
  return lo;
}

template <int MODE>
__device__ __forceinline__ TileDesc resolve_tile(int id, const int32_t* __restrict__ tile_off, int n_utts,
                                                 const int64_t* __restrict__ wav_start, const int32_t* __restrict__ wav_len,
                                                 const int32_t* __restrict__ out_row, int lane) {
  TileDesc d;
  d.valid = 0; d.m = d.i0 = d.i1 = d.f_lo = d.nF = d.orow = d.src_lo = d.src_hi = 0;
  int b, ti;
  if (MODE == 0) {
    b = utterance_of_tile(tile_off, n_utts, id, lane);
    ti = id - __ldg(tile_off + b);
  } else {
    b = blockIdx.y;
    ti = blockIdx.x;
  }
  const int n = __ldg(wav_len + b);
  if (n < FRAME_LEN) return d;
  const int m = 1 + (n - FRAME_LEN) / SHIFT;
  int f_hi;
  if (MODE == 0) {
    const int T = (m + LFR_N - 1) / LFR_N;  // VLLM:43
    d.i0 = ti * TILE_ROWS;
    if (d.i0 >= T) return d;
    d.i1 = min(d.i0 + TILE_ROWS, T);
    d.f_lo = max(LFR_N * d.i0 - LFR_LEFT, 0);
    f_hi = min(LFR_N * (d.i1 - 1) + LFR_M - 1 - LFR_LEFT, m - 1);
  } else {
    d.f_lo = ti * RAW_TILE_FRAMES;
    if (d.f_lo >= m) return d;
    f_hi = min(d.f_lo + RAW_TILE_FRAMES - 1, m - 1);
  }
  d.m = m;
  d.nF = f_hi - d.f_lo + 1;
  d.orow = __ldg(out_row + b);
  const int64_t src = __ldg(wav_start + b) + static_cast<int64_t>(d.f_lo) * SHIFT;
  d.src_lo = static_cast<int>(static_cast<uint32_t>(src));
  d.src_hi = static_cast<int>(src >> 32);
  d.valid = 1;
  return d;
}
__device__ __forceinline__ int64_t tile_src(const TileDesc& d) {
  return (static_cast<int64_t>(d.src_hi) << 32) | static_cast<uint32_t>(d.src_lo);
}

// Compile-time bound on the four-tap steps of a group: the Kaldi mel banks of the compiled geometry (80 bands, 20 Hz .. Nyquist
// at 16 kHz) need exactly {2, 2, 3, 3, 5}; any other bank takes the generic bound, one group at a time (Tables::mel_steps[7]).
template <bool STD>
__device__ __host__ constexpr int mel_cap(int g) { return !STD ? MEL_MAXJ : g < 2 ? 2 : g < 4 ? 3 : 5; }

// Mel bands hl + 16 i, i in [I0, I1): loads first, multiplies after.  Accumulation order as in a plain loop over the steps (four
// independent partial sums per band), so the result does not depend on the batching.
template <bool STD>
__device__ __host__ constexpr int mel_off(int g0, int g) {  // steps of groups [g0, g)
  int n = 0;
  for (int i = g0; i < g; ++i) n += mel_cap<STD>(i);
  return n;
}
template <int I0, int I1, bool STD>
__device__ __forceinline__ void mel_groups(const float* power, const float* mel_w, const int (&mel_st)[MEL_GROUPS],
                                           const int (&mel_nj)[MEL_GROUPS], int hl, bool active, float* logrow, float log_floor,
                                           float log_of_floor) {
  constexpr int NS = mel_off<STD>(I0, I1);
  float4 w[NS], pv[NS];
#pragma unroll
  for (int i = I0; i < I1; ++i) {
    const float4* pw = reinterpret_cast<const float4*>(power + mel_st[i]);
    const float4* wv = reinterpret_cast<const float4*>(mel_w) + i * MEL_MAXJ * 16 + hl;
#pragma unroll
    for (int j = 0; j < mel_cap<STD>(i); ++j) {
      // STD: the group has exactly mel_cap steps (Tables::mel_steps[7]); otherwise a uniform predicate, zeros beyond the last step
      const bool on = STD || j < mel_nj[i];
      w[mel_off<STD>(I0, i) + j] = on ? wv[j * 16] : make_float4(0.f, 0.f, 0.f, 0.f);
      pv[mel_off<STD>(I0, i) + j] = on ? pw[j] : make_float4(0.f, 0.f, 0.f, 0.f);
    }
  }
#pragma unroll
  for (int i = I0; i < I1; ++i) {
    float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
#pragma unroll
    for (int j = 0; j < mel_cap<STD>(i); ++j) {
      if (STD || j < mel_nj[i]) {
        a0 = fmaf(w[mel_off<STD>(I0, i) + j].x, pv[mel_off<STD>(I0, i) + j].x, a0);
        a1 = fmaf(w[mel_off<STD>(I0, i) + j].y, pv[mel_off<STD>(I0, i) + j].y, a1);
        a2 = fmaf(w[mel_off<STD>(I0, i) + j].z, pv[mel_off<STD>(I0, i) + j].z, a2);
        a3 = fmaf(w[mel_off<STD>(I0, i) + j].w, pv[mel_off<STD>(I0, i) + j].w, a3);
      }
    }
    const float acc = (a0 + a1) + (a2 + a3);
    // fast log: |err| < 4e-6 at these magnitudes; the floor itself (silence) maps to the exactly rounded log(eps) of the reference
    if (active) logrow[hl + 16 * i] = acc > log_floor ? __logf(acc) : log_of_floor;
  }
}

template <int MODE>  // 0: LFR + CMVN rows, 1: raw log-mel rows
__global__ void __launch_bounds__(THREADS, 2)
fbank_kernel(const Tables* __restrict__ tab, const float* __restrict__ wav, const int64_t* __restrict__ wav_start,
             const int32_t* __restrict__ wav_len, const int32_t* __restrict__ out_row, const int32_t* __restrict__ tile_off, int n_utts,
             int n_tiles, const float* __restrict__ cmvn, float* __restrict__ out, int64_t out_ld, float win_scale, float smp_scale,
             float preemph, float log_floor) {
  const float log_of_floor = logf(log_floor);
  extern __shared__ __align__(16) unsigned char smem_raw[];
  Smem& s = *reinterpret_cast<Smem*>(smem_raw);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5, hl = lane & 15;
  pdl_launch_dependents();  // the first kernel of the path: its successor's prologue may overlap it

  // ---- tables into shared memory, once per CTA (MODE 0 CTAs are persistent over tiles).  The sample scale (x 32768,
  // VLLM:140-141) is folded into the window when it is a power of two: scaling by 2^k commutes with every fp32 rounding of the
  // DC removal, the pre-emphasis and the window product, so the frames entering the FFT carry the same bits either way.
  for (int i = tid; i < FRAME_LEN; i += THREADS) s.window[i] = __ldg(&tab->window[i]) * win_scale;
  if (tid < 16) {
    s.tw256[tid] = tab->tw256[tid];
    s.sc512[tid] = tab->sc512[tid];
  }
  for (int i = tid; i < MEL_W / 4; i += THREADS)
    reinterpret_cast<float4*>(s.mel_w)[i] = __ldg(reinterpret_cast<const float4*>(tab->mel_w) + i);
  // this lane's five mel bins (hl, hl + 16, ..., hl + 64): first FFT bin (a multiple of four), four-tap steps of the group
  int mel_st[MEL_GROUPS], mel_nj[MEL_GROUPS];
#pragma unroll
  for (int i = 0; i < MEL_GROUPS; ++i) {
    mel_st[i] = __ldg(&tab->mel_start[hl + 16 * i]);
    mel_nj[i] = __ldg(&tab->mel_steps[i]);
  }
  const bool mel_std = FA_FE_MEL_BATCH && __ldg(&tab->mel_steps[7]) != 0;

  // PCM of a tile -> shared memory, asynchronously (16-byte cp.async; nS is a multiple of 8); the caller waits before use
  auto stage = [&](const TileDesc& d) {
    const float* src = wav + tile_src(d);
    const int nS = (d.nF - 1) * SHIFT + FRAME_LEN;
    if ((reinterpret_cast<uintptr_t>(src) & 15) == 0) {
      const uint32_t dst = static_cast<uint32_t>(__cvta_generic_to_shared(s.samples));
      for (int i = tid; i < nS / 4; i += THREADS)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(dst + 16u * i), "l"(src + 4 * i) : "memory");
    } else {
      for (int i = tid; i < nS; i += THREADS) s.samples[i] = __ldg(src + i);
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
  };

  const int first_tile = MODE == 0 ? static_cast<int>(blockIdx.x) : 0;
  const int tile_step = MODE == 0 ? static_cast<int>(gridDim.x) : 1;
  const int tile_end = MODE == 0 ? n_tiles : 1;
  {
    const TileDesc first = resolve_tile<MODE>(first_tile, tile_off, n_utts, wav_start, wav_len, out_row, lane);
    if (tid == 0) s.desc[0] = first;
    if (first_tile < tile_end && first.valid) stage(first);
  }
  int par = 0;  // descriptors live in shared memory (only the parity is carried in a register across the frame loop)
  for (int id = first_tile; id < tile_end; id += tile_step, par ^= 1) {
  asm volatile("cp.async.wait_all;" ::: "memory");
  __syncthreads();  // this tile's samples have landed; the previous tile's rows have left the log-mel buffer
  const TileDesc& cur = s.desc[par];
  const bool has_next = MODE == 0 && id + tile_step < tile_end;
  if (has_next && warp == 1) {
    // one warp resolves the CTA's next tile while the others start on the frames; its PCM is requested into L2 now, a whole
    // tile of arithmetic ahead of the cp.async that stages it
    const TileDesc nx = resolve_tile<MODE>(id + tile_step, tile_off, n_utts, wav_start, wav_len, out_row, lane);
    if (lane == 0) {
      s.desc[par ^ 1] = nx;
      const float* nsrc = wav + tile_src(nx);
      if (nx.valid && (reinterpret_cast<uintptr_t>(nsrc) & 15) == 0)
        asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" ::"l"(nsrc),
                     "r"(static_cast<uint32_t>(((nx.nF - 1) * SHIFT + FRAME_LEN) * 4))
                     : "memory");
    }
  }
  const int nF = cur.valid ? cur.nF : 0;

  const int slot = warp * 2 + (lane >> 4);
  double2* work = s.work[slot];
  float* power = reinterpret_cast<float*>(s.work[slot]);  // written after the transpose has been read back
  const int partner = (16 - hl) & 15;  // lane holding Z[256 - k] for this lane's k = hl + 16 p

  for (int fb = warp * 2; fb < nF; fb += SLOTS) {
    const int f_raw = fb + (lane >> 4);
    const bool active = f_raw < nF;
    const int f = active ? f_raw : nF - 1;
    const float* xs = s.samples + f * SHIFT;

    // ---- load 13 sample pairs (complex point n = hl + 16 j  <->  samples 2n, 2n+1), frame mean
    float2 xf[13];
    float ps[13];
#pragma unroll
    for (int j = 0; j < 13; ++j) {
      const int sidx = 2 * hl + 32 * j;
      float2 v = make_float2(0.f, 0.f);
      if (j < 12 || hl < 8) v = *reinterpret_cast<const float2*>(xs + sidx);
      if (smp_scale != 1.f) { v.x *= smp_scale; v.y *= smp_scale; }  // a scale that is not a power of two (uniform branch)
      xf[j] = v;
      ps[j] = v.x + v.y;
    }
    // frame mean (TA:183-186): the lane's 26 samples by an fp32 pairwise tree, the 16 lane sums in fp64 (the reference sums in
    // fp32 as well; both are within a few 1e-8 relative of the exact mean, far inside the log-mel budget)
    const float lane_sum = (((ps[0] + ps[1]) + (ps[2] + ps[3])) + ((ps[4] + ps[5]) + (ps[6] + ps[7]))) +
                           (((ps[8] + ps[9]) + (ps[10] + ps[11])) + ps[12]);
    double part = static_cast<double>(lane_sum);
#pragma unroll
    for (int o = 8; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    const float mean = static_cast<float>(part * (1.0 / static_cast<double>(FRAME_LEN)));

    // ---- DC removal, pre-emphasis (replicate-left, TA:193-198), window (TA:201-204): the reference's fp32 operation
    // sequence, one rounding per torch op (no FMA contraction), so the frames entering the FFT carry the same bits.
    cplx x[16];
#pragma unroll
    for (int j = 0; j < 13; ++j) {
      const int sidx = 2 * hl + 32 * j;
      // the sample before the pair (replicate-left at the frame start, TA:193-198) is the second sample of lane hl - 1's pair j;
      // lane 0 takes lane 15's pair j - 1.  A shuffle instead of a second pass over the staged samples.
      const float snd = hl == 15 ? xf[j > 0 ? j - 1 : 0].y : xf[j].y;
      float rp = __shfl_sync(0xffffffffu, snd, (hl + 15) & 15, 16);
      if (j == 0 && hl == 0) rp = xf[0].x;
      if (j < 12 || hl < 8) {
        const float a0 = __fsub_rn(xf[j].x, mean), a1 = __fsub_rn(xf[j].y, mean), ap = __fsub_rn(rp, mean);
        const float y0 = __fsub_rn(a0, __fmul_rn(preemph, ap));
        const float y1 = __fsub_rn(a1, __fmul_rn(preemph, a0));
        x[j] = mk(static_cast<double>(__fmul_rn(y0, s.window[sidx])), static_cast<double>(__fmul_rn(y1, s.window[sidx + 1])));
      } else {
        x[j] = mk(0.0, 0.0);
      }
    }
    x[13] = x[14] = x[15] = mk(0.0, 0.0);  // zero padding 400 -> 512 (TA:207-211)

    // ---- pass 1: DFT16 over j, twiddle W256^(t q), transpose through smem
    fft16<true>(x);
#pragma unroll
    {  // twiddles W256^(hl q), q = 1..15, as two interleaved product chains (odd and even q) from the lane's W256^hl: fourteen
       // complex products instead of fifteen 16-byte table reads per lane (shared-memory wavefronts are what bounds this kernel);
       // the chains add a few 1e-16 of relative error
      const cplx w1 = s.tw256[hl];
      const cplx w2 = cmul(w1, w1);
      cplx po = w1, pe = w2;
      x[POS16(1)] = cmul(x[POS16(1)], po);
      x[POS16(2)] = cmul(x[POS16(2)], pe);
#pragma unroll
      for (int q = 3; q < 16; q += 2) {
        po = cmul(po, w2);
        x[POS16(q)] = cmul(x[POS16(q)], po);
        if (q + 1 < 16) {
          pe = cmul(pe, w2);
          x[POS16(q + 1)] = cmul(x[POS16(q + 1)], pe);
        }
      }
    }
    {  // 16-byte accesses: rows of 17 complex values keep both directions at the two-wavefront minimum
#pragma unroll
      for (int q = 0; q < 16; ++q) work[q * 17 + hl] = x[POS16(q)];
      __syncwarp();
#pragma unroll
      for (int t = 0; t < 16; ++t) x[t] = work[hl * 17 + t];
      __syncwarp();
    }
    // ---- pass 2: DFT16 over t -> Z[hl + 16 p] at x[POS16(p)]
    fft16<false>(x);

    // ---- split the packed real FFT and take |.|^2 in one step.  With zk = Z[k], zm = Z[256 - k], A = zk + conj(zm),
    // B = zk - conj(zm), w' = -i W512^k:  2 X[k] = A + w' B,  2 conj(X[256-k]) = A - w' B, hence (|w'| = 1)
    //   |X[k]|^2, |X[256-k]|^2 = (|zk|^2 + |zm|^2) / 2  +-  (Re(w') / 2 * (|zk|^2 - |zm|^2) - Im(w') * Im(zk zm))
    // (12 fp64 operations per bin pair instead of 20 through Xe and W512^k Xo).  The difference form loses
    // eps64 * (|X[k]| / |X[256-k]|)^2 relative precision on the weaker bin of a pair: 1e-8 at 80 dB between the two, 1e-4 at 120 dB,
    // still under the 1e-3 log-mel budget where the reference's own fp32 FFT is off by tens of per cent.
    // k = hl + 16 p; Z[256 - k] = Z[(16 - hl) + 16 (15 - p)] lives in lane (16 - hl) & 15 at p' = 15 - p; for hl == 0 it is
    // this lane's own Z[16 ((16 - p) & 15)].
    const cplx sc = s.sc512[hl];  // (sin, cos)(2 pi hl / 512); the constants of bin k = hl + 16 p follow by a rotation by p pi / 16
#pragma unroll
    for (int p = 0; p < 9; ++p) {
      const cplx own = x[POS16((16 - p) & 15)];
      const cplx snd = x[POS16(15 - (p & 7))];  // p == 8 is only used by hl == 0 (k = 128), which reads `own`
      cplx zm = mk(__shfl_sync(0xffffffffu, snd.x, partner, 16), __shfl_sync(0xffffffffu, snd.y, partner, 16));
      if (hl == 0) zm = own;
      const int k = hl + 16 * p;
      if (k <= 128) {
        const cplx zk = x[POS16(p & 15)];
        // u = (-sin(theta) / 2, cos(theta)), theta = 2 pi k / 512 = theta_hl + p pi / 16
        cplx u;
        if (p == 0) u = mk(-0.5 * sc.x, sc.y);
        else if (p == 8) u = mk(-0.5 * sc.y, -sc.x);
        else u = mk(fma(sc.x, -0.5 * cos_p16(p), sc.y * (-0.5 * sin_p16(p))), fma(sc.y, cos_p16(p), -(sc.x * sin_p16(p))));
        const double nk = fma(zk.x, zk.x, zk.y * zk.y), nm = fma(zm.x, zm.x, zm.y * zm.y);
        const double e = fma(zk.x, zm.y, zk.y * zm.x);
        const double sum = nk + nm, c = fma(u.y, e, u.x * (nk - nm));
        power[k] = static_cast<float>(fma(0.5, sum, c));  // TA:616-618 |rfft|^2
        power[256 - k] = static_cast<float>(fma(0.5, sum, -c));
      }
    }
    if (hl < 7) power[257 + hl] = 0.f;  // read (x 0) by the padded mel taps
    __syncwarp();

    // ---- sparse mel projection + log (TA:621-633); bins hl, hl+16, ..., hl+64.  Every band is a whole number of aligned
    // four-tap steps (zero weights in the padding); the step count of a group is uniform over the lanes.  The loads of groups
    // 0-1, 2-3 and 4 are all issued before the first multiply of their batch (the FFT's registers are free by now), so a lane
    // waits for shared memory three times per frame instead of once per group and step.
    if (mel_std) {
      mel_groups<0, 2, true>(power, s.mel_w, mel_st, mel_nj, hl, active, &s.logmel[f][0], log_floor, log_of_floor);
      mel_groups<2, 4, true>(power, s.mel_w, mel_st, mel_nj, hl, active, &s.logmel[f][0], log_floor, log_of_floor);
      mel_groups<4, MEL_GROUPS, true>(power, s.mel_w, mel_st, mel_nj, hl, active, &s.logmel[f][0], log_floor, log_of_floor);
    } else {
      mel_groups<0, 1, false>(power, s.mel_w, mel_st, mel_nj, hl, active, &s.logmel[f][0], log_floor, log_of_floor);
      mel_groups<1, 2, false>(power, s.mel_w, mel_st, mel_nj, hl, active, &s.logmel[f][0], log_floor, log_of_floor);
      mel_groups<2, 3, false>(power, s.mel_w, mel_st, mel_nj, hl, active, &s.logmel[f][0], log_floor, log_of_floor);
      mel_groups<3, 4, false>(power, s.mel_w, mel_st, mel_nj, hl, active, &s.logmel[f][0], log_floor, log_of_floor);
      mel_groups<4, 5, false>(power, s.mel_w, mel_st, mel_nj, hl, active, &s.logmel[f][0], log_floor, log_of_floor);
    }
    __syncwarp();
  }
  // this thread's CMVN column (MODE 0: threads 0..139 own one float4 column of the 560 for every row of the tile), requested
  // before the barrier so that its latency overlaps the wait
  constexpr int V = FEAT / 4;  // 140 float4 per row
  float4 mu = make_float4(0.f, 0.f, 0.f, 0.f), sc = make_float4(1.f, 1.f, 1.f, 1.f);
  if (MODE == 0 && cmvn != nullptr && tid < V) {
    mu = __ldg(reinterpret_cast<const float4*>(cmvn) + tid);
    sc = __ldg(reinterpret_cast<const float4*>(cmvn + FEAT) + tid);
  }
  __syncthreads();  // log-mel rows complete, samples free, s.next visible

  if (has_next && s.desc[par ^ 1].valid) stage(s.desc[par ^ 1]);  // lands while this tile's rows are written out
  const int m = cur.m, i0 = cur.i0, i1 = cur.i1, f_lo = cur.f_lo;

  if (MODE == 0) {
    // ---- LFR gather (VLLM:40-60 closed form) + CMVN (VLLM:23-37), coalesced float4 rows
    if (cur.valid && tid < V) {
      const int col = tid * 4, j = col / NMEL, c = col - j * NMEL;
      float* obase = out + static_cast<int64_t>(cur.orow + i0) * out_ld + col;
      for (int r = 0; r < i1 - i0; ++r) {
        int fr = LFR_N * (i0 + r) + j - LFR_LEFT;
        fr = min(max(fr, 0), m - 1) - f_lo;
        float4 v = *reinterpret_cast<const float4*>(&s.logmel[fr][c]);
        if (cmvn != nullptr) {
          v.x = (v.x + mu.x) * sc.x; v.y = (v.y + mu.y) * sc.y;
          v.z = (v.z + mu.z) * sc.z; v.w = (v.w + mu.w) * sc.w;
        }
        *reinterpret_cast<float4*>(obase + static_cast<int64_t>(r) * out_ld) = v;
      }
    }
  } else if (cur.valid) {
    constexpr int VM = NMEL / 4;
    float* obase = out + static_cast<int64_t>(cur.orow + f_lo) * out_ld;
    for (int idx = tid; idx < nF * VM; idx += THREADS) {
      const int r = idx / VM, c4 = idx - r * VM;
      *reinterpret_cast<float4*>(obase + static_cast<int64_t>(r) * out_ld + c4 * 4) =
          *reinterpret_cast<const float4*>(&s.logmel[r][c4 * 4]);
    }
  }
  }
  asm volatile("cp.async.wait_all;" ::: "memory");
}

// A power-of-two sample scale goes into the window (exact); anything else is applied to the samples as they are read.
struct Scales { float win, smp; };
static Scales split_scale(float upscale) {
  int e = 0;
  const bool pow2 = upscale > 0.f && frexpf(upscale, &e) == 0.5f;
  return pow2 ? Scales{upscale, 1.f} : Scales{1.f, upscale};
}

static int check_cfg(const fa_frontend_cfg* c) {
  FA_REQUIRE(c != nullptr, "frontend cfg is NULL");
  FA_REQUIRE(c->frame_len == FRAME_LEN && c->frame_shift == SHIFT && c->n_fft == NFFT && c->n_mels == NMEL &&
                 c->lfr_m == LFR_M && c->lfr_n == LFR_N,
             "frontend geometry %d/%d/%d/%d lfr %d/%d is not the compiled one (400/160/512/80 lfr 7/6)", c->frame_len,
             c->frame_shift, c->n_fft, c->n_mels, c->lfr_m, c->lfr_n);
  return 0;
}

}  // namespace fe
}  // namespace fa

using namespace fa;
using namespace fa::fe;

extern "C" size_t fa_frontend_tables_bytes(void) { return sizeof(Tables); }

extern "C" int fa_frontend_tables_build(const fa_frontend_cfg* cfg, const float* window, const float* mel_banks,
                                        void* blob_host) {
  if (int r = check_cfg(cfg)) return r;
  FA_REQUIRE(window && mel_banks && blob_host, "NULL table pointer");
  Tables* t = static_cast<Tables*>(blob_host);
  std::memset(t, 0, sizeof(Tables));
  for (int i = 0; i < FRAME_LEN; ++i) t->window[i] = window[i];
  const double two_pi = 6.283185307179586476925286766559;
  for (int l = 0; l < 16; ++l) {
    t->tw256[l] = make_double2(cos(two_pi * l / 256.0), -sin(two_pi * l / 256.0));
    t->sc512[l] = make_double2(sin(two_pi * l / 512.0), cos(two_pi * l / 512.0));
  }
  for (int b = 0; b < NMEL; ++b) {
    const float* row = mel_banks + static_cast<size_t>(b) * NBIN;
    int first = -1, last = -1;
    for (int k = 0; k < NBIN; ++k)
      if (row[k] != 0.f) {
        if (first < 0) first = k;
        last = k;
      }
    const int start = first < 0 ? 0 : (first & ~3);
    const int steps = first < 0 ? 0 : (last - start) / 4 + 1;
    FA_REQUIRE(steps <= MEL_MAXJ, "mel band %d spans %d FFT bins: more than the %d aligned taps the kernel is compiled for", b,
               last - first + 1, 4 * MEL_MAXJ);
    const int g = b / 16, l = b % 16;
    t->mel_start[b] = start;
    if (steps > t->mel_steps[g]) t->mel_steps[g] = steps;
    for (int k = first; first >= 0 && k <= last; ++k)
      t->mel_w[((g * MEL_MAXJ + (k - start) / 4) * 16 + l) * 4 + ((k - start) & 3)] = row[k];
  }
  t->mel_steps[7] = 1;
  for (int g = 0; g < MEL_GROUPS; ++g)
    if (t->mel_steps[g] != mel_cap<true>(g)) t->mel_steps[7] = 0;
  for (int b = 0; b < NMEL; ++b)  // the kernel reads whole steps: they must stay inside the zero-padded power row (264 floats)
    FA_REQUIRE(t->mel_start[b] + 4 * t->mel_steps[b / 16] <= 264, "mel band %d: padded taps run past the spectrum", b);
  return 0;
}

extern "C" int fa_frontend_forward(const fa_frontend_cfg* cfg, const void* tables_dev, const float* wav_dev,
                                   const int64_t* wav_start_dev, const int32_t* wav_len_dev,
                                   const int32_t* out_row_dev, const int32_t* tile_off_dev,
                                   int32_t n_utts, int32_t n_tiles, const float* cmvn_dev, float* out_dev,
                                   int64_t out_ld, fa_stream_t stream) {
  if (int r = check_cfg(cfg)) return r;
  FA_REQUIRE(n_utts > 0, "empty batch");
  FA_REQUIRE(tables_dev && wav_dev && wav_start_dev && wav_len_dev && out_row_dev && tile_off_dev && out_dev,
             "NULL device pointer");
  FA_REQUIRE(out_ld >= FEAT && out_ld % 4 == 0, "out_ld must be >= %d and a multiple of 4", FEAT);
  FA_REQUIRE((reinterpret_cast<uintptr_t>(out_dev) & 15) == 0, "out_dev must be 16-byte aligned");
  if (n_tiles <= 0) return 0;
  static SmemOptIn opt0;
  if (int r = opt_in_smem(fbank_kernel<0>, (int)sizeof(Smem), opt0)) return r;
  const Scales sc = split_scale(cfg->upscale);
  ProfScope _ps(K_FRONTEND, static_cast<cudaStream_t>(stream));
  // persistent CTAs (two per SM: registers and shared memory), tiles dealt round-robin: the tables are staged once per CTA
  const int grid = n_tiles < 2 * sm_count() ? n_tiles : 2 * sm_count();
  fbank_kernel<0><<<grid, THREADS, sizeof(Smem), static_cast<cudaStream_t>(stream)>>>(
      static_cast<const Tables*>(tables_dev), wav_dev, wav_start_dev, wav_len_dev, out_row_dev, tile_off_dev, n_utts, n_tiles,
      cmvn_dev,
      out_dev, out_ld, sc.win, sc.smp, cfg->preemph, cfg->log_floor);
  FA_CHECK_LAUNCH();
  return 0;
}

extern "C" int fa_frontend_logmel(const fa_frontend_cfg* cfg, const void* tables_dev, const float* wav_dev,
                                  const int64_t* wav_start_dev, const int32_t* wav_len_dev,
                                  const int32_t* out_row_dev, int32_t n_utts,
                                  int32_t max_frames, float* out_dev, int64_t out_ld, fa_stream_t stream) {
  if (int r = check_cfg(cfg)) return r;
  FA_REQUIRE(n_utts > 0 && max_frames > 0, "empty batch");
  FA_REQUIRE(tables_dev && wav_dev && wav_start_dev && wav_len_dev && out_row_dev && out_dev, "NULL device pointer");
  FA_REQUIRE(out_ld >= NMEL && out_ld % 4 == 0, "out_ld must be >= %d and a multiple of 4", NMEL);
  static SmemOptIn opt1;
  if (int r = opt_in_smem(fbank_kernel<1>, (int)sizeof(Smem), opt1)) return r;
  dim3 grid(ceil_div(max_frames, RAW_TILE_FRAMES), n_utts);
  const Scales sc = split_scale(cfg->upscale);
  ProfScope _ps(K_FRONTEND, static_cast<cudaStream_t>(stream));
  fbank_kernel<1><<<grid, THREADS, sizeof(Smem), static_cast<cudaStream_t>(stream)>>>(
      static_cast<const Tables*>(tables_dev), wav_dev, wav_start_dev, wav_len_dev, out_row_dev, nullptr, n_utts, 1, nullptr,
      out_dev,
      out_ld, sc.win, sc.smp, cfg->preemph, cfg->log_floor);
  FA_CHECK_LAUNCH();
  return 0;
}

extern "C" int32_t fa_num_frames(int32_t n_samples, int32_t frame_len, int32_t frame_shift) {
  return n_samples < frame_len ? 0 : 1 + (n_samples - frame_len) / frame_shift;
}
extern "C" int32_t fa_num_lfr(int32_t n_frames, int32_t lfr_n) { return (n_frames + lfr_n - 1) / lfr_n; }
extern "C" int32_t fa_fake_token_len(int32_t n_lfr, int32_t use_low_frame_rate) {
  if (!use_low_frame_rate) return n_lfr;
  // /root/reference/model.py:391-394 (python floor division)
  auto fdiv = [](int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); };
  int o = 1 + fdiv(n_lfr - 3 + 2, 2);
  o = 1 + fdiv(o - 3 + 2, 2);
  return fdiv(o - 1, 2) + 1;
}
