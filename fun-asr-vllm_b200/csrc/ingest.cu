// Audio ingest on the GPU: the reference resamples every utterance to 16 kHz on the host with
// torchaudio.transforms.Resample(sr, 16000) (/root/reference/infer.py:328-329, infer_kaldidata.py:454-455), i.e. the
// polyphase windowed-sinc filter bank of TA/functional/functional.py:1405-1432:
//     out[i*new + j] = sum_k kernel[j][k] * xpad[i*orig + k],  k < 2*width + orig,  xpad = (width zeros | x | zeros)
// cropped to ceil(new*L/orig) samples.  Here: one kernel over a packed batch (utterance b at in_start[b], length in_len[b]).
// HBM-bound (read 4 B/sample, write 4 B/sample); the FMA work (K per output sample) runs out of shared memory.
//   CTA  = `fw*32` consecutive frames of one utterance; the input span is staged once in shared memory (zero padded);
//   warp = one (block of J phases, group of 32*F frames): taps are broadcast 16-byte loads, the lanes read the staged span at
//          stride `orig`; results go through a shared-memory tile so that the global stores are coalesced.
#include <cuda_runtime.h>

#include "common.cuh"

namespace fa {
namespace ingest {

constexpr int THREADS = 256;

// One mono f32 sample from interleaved raw PCM frame `frame` (FA_PCM_*): every channel is converted to f32 (int16: exact,
// f64: round to nearest = torch's .float(), /root/reference/infer.py:321-324), scaled (1/32768 reproduces what soundfile
// hands librosa.load for 16-bit files, infer_kaldidata.py:101), the channels are summed in order and divided by their count
// (librosa.to_mono = np.mean over the channel axis).
__device__ __forceinline__ float load_pcm(const void* __restrict__ in, int fmt, int ch, float scale, int64_t frame) {
  float s = 0.f;
  for (int c = 0; c < ch; ++c) {
    const int64_t i = frame * ch + c;
    float v;
    if (fmt == FA_PCM_S16) v = static_cast<float>(__ldg(static_cast<const int16_t*>(in) + i));
    else if (fmt == FA_PCM_F64) v = __double2float_rn(__ldg(static_cast<const double*>(in) + i));
    else v = __ldg(static_cast<const float*>(in) + i);
    v = __fmul_rn(v, scale);
    s = c == 0 ? v : __fadd_rn(s, v);
  }
  return ch > 1 ? __fdiv_rn(s, static_cast<float>(ch)) : s;
}

// out[out_start[b] + s] = mono f32 of frame in_start[b] + s, s < in_len[b]   (grid: x = blocks of frames, y = utterance)
__global__ void __launch_bounds__(THREADS)
pcm_to_mono_kernel(const void* __restrict__ in, int fmt, int ch, float scale, const int64_t* __restrict__ in_start,
                   const int32_t* __restrict__ in_len, float* __restrict__ out, const int64_t* __restrict__ out_start) {
  const int b = blockIdx.y;
  const int L = in_len[b];
  const int64_t i0 = in_start[b];
  float* dst = out + out_start[b];
  for (int s = blockIdx.x * THREADS + threadIdx.x; s < L; s += gridDim.x * THREADS) dst[s] = load_pcm(in, fmt, ch, scale, i0 + s);
}

// Register blocking: a lane owns J phases x F frames (frames 32 apart), so every 4 taps cost J 16-byte broadcast loads of the
// (row-padded) table and 4*F shared-memory loads for 16*J*F/4 FMAs.  J*F = 8; J follows the number of phases (new_freq).
template <int J, int F>
__global__ void __launch_bounds__(THREADS)
resample_kernel(const void* __restrict__ in, int fmt, int ch, float scale, const int64_t* __restrict__ in_start,
                const int32_t* __restrict__ in_len, const float* __restrict__ kern, int orig, int nw, int width, int Kp, int fw, float* __restrict__ out,
                const int64_t* __restrict__ out_start, const int32_t* __restrict__ tile_off, int n_utts) {
  extern __shared__ float rs_smem[];
  // utterance of this tile: last b with tile_off[b] <= blockIdx.x
  int lo = 0, hi = n_utts;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(tile_off + mid) <= static_cast<int>(blockIdx.x)) lo = mid; else hi = mid;
  }
  const int b = lo;
  const int FI = fw * 32;
  const int64_t L = in_len[b];
  const int64_t target = (static_cast<int64_t>(nw) * L + orig - 1) / orig;  // ceil(new * L / orig)
  const int64_t i0 = static_cast<int64_t>(blockIdx.x - tile_off[b]) * FI;   // first frame of the tile
  const int n_in = (FI - 1) * orig + Kp;  // the table rows are zero-padded to Kp taps: the extra samples only need to be finite
  float* xs = rs_smem;
  float* ys = rs_smem + ((n_in + 3) & ~3);
  const int64_t src0 = in_start[b];
  const int64_t x0 = i0 * orig - width;  // sample index of xs[0]
  for (int q = threadIdx.x; q < n_in; q += THREADS) {  // conversion and down-mix happen here: raw PCM is read once
    const int64_t s = x0 + q;
    xs[q] = (s >= 0 && s < L) ? load_pcm(in, fmt, ch, scale, src0 + s) : 0.f;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_jb = (nw + J - 1) / J, n_fg = fw / F;
  for (int it = warp; it < n_jb * n_fg; it += THREADS / 32) {
    const int j0 = (it % n_jb) * J, f0 = (it / n_jb) * 32 * F + lane;
    const float4* kj[J];
#pragma unroll
    for (int a = 0; a < J; ++a) kj[a] = reinterpret_cast<const float4*>(kern + static_cast<int64_t>(min(j0 + a, nw - 1)) * Kp);
    const float* xf[F];
#pragma unroll
    for (int f = 0; f < F; ++f) xf[f] = xs + (f0 + 32 * f) * orig;
    float acc[J][F];
#pragma unroll
    for (int a = 0; a < J; ++a)
#pragma unroll
      for (int f = 0; f < F; ++f) acc[a][f] = 0.f;
    float4 wn[J];  // taps of the next 4-tap step, requested one step ahead (the table is L2-resident, not L1-resident)
#pragma unroll
    for (int a = 0; a < J; ++a) wn[a] = __ldg(kj[a]);
    for (int k = 0; k < Kp; k += 4) {
      float4 w[J];
#pragma unroll
      for (int a = 0; a < J; ++a) w[a] = wn[a];
      if (k + 4 < Kp) {
#pragma unroll
        for (int a = 0; a < J; ++a) wn[a] = __ldg(kj[a] + ((k + 4) >> 2));
      }
#pragma unroll
      for (int f = 0; f < F; ++f) {
        const float x0v = xf[f][k], x1v = xf[f][k + 1], x2v = xf[f][k + 2], x3v = xf[f][k + 3];
#pragma unroll
        for (int a = 0; a < J; ++a) {  // taps in ascending order, like the scalar sum
          acc[a][f] = fmaf(w[a].x, x0v, acc[a][f]);
          acc[a][f] = fmaf(w[a].y, x1v, acc[a][f]);
          acc[a][f] = fmaf(w[a].z, x2v, acc[a][f]);
          acc[a][f] = fmaf(w[a].w, x3v, acc[a][f]);
        }
      }
    }
#pragma unroll
    for (int a = 0; a < J; ++a)
      if (j0 + a < nw)
#pragma unroll
        for (int f = 0; f < F; ++f) ys[(f0 + 32 * f) * nw + j0 + a] = acc[a][f];
  }
  __syncthreads();
  float* dst = out + out_start[b];
  const int64_t o0 = i0 * nw;
  for (int o = threadIdx.x; o < FI * nw; o += THREADS)
    if (o0 + o < target) dst[o0 + o] = ys[o];
}

}  // namespace ingest
}  // namespace fa

static int check_pcm(int32_t format, int32_t channels, float scale) {
  FA_REQUIRE(format == FA_PCM_F32 || format == FA_PCM_S16 || format == FA_PCM_F64, "ingest: unknown PCM format %d", format);
  FA_REQUIRE(channels >= 1 && channels <= 64, "ingest: %d channels", channels);
  FA_REQUIRE(scale == scale && scale != 0.f, "ingest: scale must be a non-zero number");
  return 0;
}

// Replaces `torch.from_numpy(audio).float()` (/root/reference/infer.py:320-324) and librosa.load's mono mix
// (/root/reference/infer_kaldidata.py:101) for a packed batch of raw interleaved PCM.
extern "C" int fa_pcm_to_mono_f32(const void* in_dev, int32_t format, int32_t channels, float scale, const int64_t* in_start_dev,
                                  const int32_t* in_len_dev, float* out_dev, const int64_t* out_start_dev, int32_t n_utts,
                                  int32_t max_len, fa_stream_t stream) {
  using namespace fa;
  FA_REQUIRE(in_dev && in_start_dev && in_len_dev && out_dev && out_start_dev, "pcm_to_mono: NULL pointer");
  FA_REQUIRE(n_utts >= 1 && max_len >= 1, "pcm_to_mono: empty problem");
  if (int r = check_pcm(format, channels, scale)) return r;
  const int blocks = ceil_div(max_len, ingest::THREADS * 8);
  ingest::pcm_to_mono_kernel<<<dim3(blocks, n_utts), ingest::THREADS, 0, static_cast<cudaStream_t>(stream)>>>(
      in_dev, format, channels, scale, in_start_dev, in_len_dev, out_dev, out_start_dev);
  FA_CHECK_LAUNCH();
  return 0;
}

// Replaces `torchaudio.transforms.Resample(sr, 16000)(audio_tensor)` (/root/reference/infer.py:328-329) for a packed batch,
// reading raw PCM (format / channels / scale as above; in_start / in_len count FRAMES).
extern "C" int fa_resample_sinc_pcm(const void* in_dev, int32_t format, int32_t channels, float scale, const int64_t* in_start_dev,
                                    const int32_t* in_len_dev, const float* kernel_dev, int32_t orig, int32_t new_freq, int32_t width,
                                    int32_t taps_padded, int32_t frames_per_tile, float* out_dev, const int64_t* out_start_dev,
                                    const int32_t* tile_off_dev, int32_t n_utts, int32_t n_tiles, fa_stream_t stream) {
  using namespace fa;
  if (int r = check_pcm(format, channels, scale)) return r;
  FA_REQUIRE(in_dev && in_start_dev && in_len_dev && kernel_dev && out_dev && out_start_dev && tile_off_dev, "resample: NULL pointer");
  FA_REQUIRE(orig >= 1 && new_freq >= 1 && width >= 1 && n_utts >= 1 && n_tiles >= 1, "resample: empty problem");
  FA_REQUIRE(taps_padded % 4 == 0 && taps_padded >= 2 * width + orig && (reinterpret_cast<uintptr_t>(kernel_dev) & 15) == 0,
             "resample: table rows must be padded to a multiple of 4 taps (>= %d) and 16-byte aligned", 2 * width + orig);
  const int J = new_freq >= 8 ? 8 : new_freq >= 4 ? 4 : new_freq >= 2 ? 2 : 1, F = 8 / J;
  FA_REQUIRE(frames_per_tile >= 32 * F && frames_per_tile % (32 * F) == 0, "resample: frames_per_tile must be a multiple of %d for new_freq %d",
             32 * F, new_freq);
  const int fw = frames_per_tile / 32;
  const size_t n_in = static_cast<size_t>(frames_per_tile - 1) * orig + taps_padded;
  const size_t smem = (((n_in + 3) & ~static_cast<size_t>(3)) + static_cast<size_t>(frames_per_tile) * new_freq) * sizeof(float);
  FA_REQUIRE(smem <= 200 * 1024, "resample: tile of %d frames needs %zu bytes of shared memory (orig %d, new %d)", frames_per_tile,
             smem, orig, new_freq);
  const cudaStream_t st = static_cast<cudaStream_t>(stream);
#define FA_RS_LAUNCH(JJ, FF)                                                                                                        \
  do {                                                                                                                              \
    static SmemOptIn opt;                                                                                                           \
    if (int r_ = opt_in_smem(ingest::resample_kernel<JJ, FF>, 200 * 1024, opt)) return r_;                                          \
    ingest::resample_kernel<JJ, FF><<<n_tiles, ingest::THREADS, smem, st>>>(in_dev, format, channels, scale, in_start_dev, in_len_dev, kernel_dev, orig, \
                                                                             new_freq, width, taps_padded, fw, out_dev, out_start_dev, \
                                                                             tile_off_dev, n_utts);                                  \
  } while (0)
  if (J == 8) FA_RS_LAUNCH(8, 1);
  else if (J == 4) FA_RS_LAUNCH(4, 2);
  else if (J == 2) FA_RS_LAUNCH(2, 4);
  else FA_RS_LAUNCH(1, 8);
#undef FA_RS_LAUNCH
  FA_CHECK_LAUNCH();
  return 0;
}

// f32 mono input: `torchaudio.transforms.Resample(sr, 16000)(audio_tensor)` on what the reference's collator hands it
extern "C" int fa_resample_sinc(const float* in_dev, const int64_t* in_start_dev, const int32_t* in_len_dev, const float* kernel_dev,
                                int32_t orig, int32_t new_freq, int32_t width, int32_t taps_padded, int32_t frames_per_tile,
                                float* out_dev, const int64_t* out_start_dev, const int32_t* tile_off_dev, int32_t n_utts,
                                int32_t n_tiles, fa_stream_t stream) {
  return fa_resample_sinc_pcm(in_dev, FA_PCM_F32, 1, 1.f, in_start_dev, in_len_dev, kernel_dev, orig, new_freq, width, taps_padded,
                              frames_per_tile, out_dev, out_start_dev, tile_off_dev, n_utts, n_tiles, stream);
}
