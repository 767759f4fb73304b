/*
 * funasr_b200.h — C ABI of libfunasr_b200.so: the B200 (sm_100a) audio hot path of Fun-ASR-Nano.
 *
 * The reference (duj12/Fun-ASR-vllm) is pure Python and has no FFI of its own; the "binding" it
 * has for this path is the Python call boundary into funasr/torchaudio.  Each entry point below
 * names the reference call it replaces (paths under /root/reference; VLLM/ and TA/ are the on-box
 * copies of the third-party arithmetic, see SURVEY.md "Citation conventions"):
 *
 *   fa_frontend_forward   <- extract_fbank(...)            infer.py:446, infer_kaldidata.py:708,897, model.py:372
 *                            = WavFrontend.forward          VLLM/transformers_utils/processors/funasr.py:128-168
 *                            = kaldi.fbank + apply_lfr + apply_cmvn   TA/compliance/kaldi.py:514-645, VLLM/...:40-60,23-37
 *   fa_encoder_forward    <- model.audio_encoder(speech, lens)        infer.py:455, model.py:252,259-263
 *                            = SenseVoiceEncoderSmall.forward          VLLM/model_executor/models/funasr.py:378-409
 *   fa_adaptor_forward    <- model.audio_adaptor(enc, lens)           infer.py:458, model.py:253,491
 *                            = adaptor Transformer.forward             VLLM/model_executor/models/funasr.py:574-595
 *   fa_encode_adapt       <- FunASRNano.forward_export(speech, lens)  model.py:251-254
 *   fa_splice_embeds      <- torch.cat([prefix, speech[:len], suffix]) infer.py:462-466 (model.py:513-541)
 *   fa_fake_token_len     <- fake_token_len arithmetic                model.py:391-396
 *   fa_resample_sinc      <- torchaudio.transforms.Resample(sr, 16000)(audio)   infer.py:328-329, infer_kaldidata.py:454-455
 *                            = _apply_sinc_resample_kernel             TA/functional/functional.py:1405-1432   
 *   fa_pcm_to_mono_f32    <- torch.from_numpy(audio).float()           infer.py:320-324; librosa.load mono mix, infer_kaldidata.py:101
 *   fa_resample_sinc_pcm  <- both of the above in one pass
 *
 * Conventions
 *   - plain pointers and sizes only; every *_dev pointer is device memory owned by the caller;
 *   - no allocation and no synchronisation inside: every call only enqueues kernels on `stream`
 *     (a cudaStream_t / CUstream); the caller owns buffers, workspace and the stream;
 *   - return 0 on success, non-zero on error, message in fa_last_error() (thread local);
 *   - variable-length batches are PACKED: utterance b owns rows [cu_seqlens[b], cu_seqlens[b+1])
 *     of every [tokens, dim] matrix; no padding rows exist anywhere inside the library.
 */
#ifndef FUNASR_B200_H_
#define FUNASR_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif
#if defined(__GNUC__)
#pragma GCC visibility push(default) /* the library is built with -fvisibility=hidden */
#endif

typedef void* fa_stream_t; /* cudaStream_t */

#define FA_ABI_VERSION 10

int fa_abi_version(void);
const char* fa_last_error(void);
/* SM count / compute capability of the current device: fails unless it is sm_100 (B200). */
int fa_check_device(int32_t* sm_count, int32_t* cc_major, int32_t* cc_minor);

/* ------------------------------------------------------------------ frontend ---------------- */
typedef struct {
  int32_t sample_rate; /* 16000 */
  int32_t frame_len;   /* 400  (25 ms)   TA:135 */
  int32_t frame_shift; /* 160  (10 ms)   TA:134 */
  int32_t n_fft;       /* 512            TA:136 */
  int32_t n_mels;      /* 80 */
  int32_t lfr_m;       /* 7 */
  int32_t lfr_n;       /* 6 */
  float preemph;       /* 0.97           TA:532 */
  float upscale;       /* 32768 when WavFrontend.upsacle_samples (VLLM:140) else 1 */
  float log_floor;     /* FLT_EPSILON    TA:22,633 */
} fa_frontend_cfg;

/* Only this geometry is compiled (the one Fun-ASR-Nano uses); anything else returns an error. */
#define FA_FE_TILE_ROWS 5 /* LFR rows produced per CTA tile; host builds tile_off with it */

/* Size of the packed table blob (window, FFT twiddles, sparse mel filterbank). */
size_t fa_frontend_tables_bytes(void);
/* Build the blob on the HOST into `blob_host` from the reference-side tables:
 *   window[frame_len]            = _feature_window_function(...)       TA:86-113
 *   mel_banks[n_mels * (n_fft/2+1)] = get_mel_banks(...) zero Nyquist column appended  TA:436-511,627
 * The caller uploads the blob to the device once. */
int fa_frontend_tables_build(const fa_frontend_cfg* cfg, const float* window, const float* mel_banks,
                             void* blob_host);

/* Fused framing + DC removal + pre-emphasis + window + 512-pt rFFT + power + mel + log + LFR + CMVN.
 *   wav_dev       : packed f32 PCM in [-1,1)
 *   wav_start_dev : [n_utts] int64 first sample of each utterance in wav_dev
 *   wav_len_dev   : [n_utts] int32 samples per utterance (>= frame_len)
 *   out_row_dev  : [n_utts]   int32 first output row of each utterance (packed: cu_seqlens[b];
 *                  padded [B,Tmax,.]: b*Tmax)
 *   tile_off_dev : [n_utts+1] int32 prefix sum of ceil(T_b / FA_FE_TILE_ROWS)
 *   cmvn_dev     : [2, lfr_m*n_mels] (means; scales) or NULL            VLLM:23-37
 *   out_dev      : f32, row stride out_ld (>= lfr_m*n_mels) */
int fa_frontend_forward(const fa_frontend_cfg* cfg, const void* tables_dev, const float* wav_dev,
                        const int64_t* wav_start_dev, const int32_t* wav_len_dev, const int32_t* out_row_dev,
                        const int32_t* tile_off_dev, int32_t n_utts, int32_t n_tiles,
                        const float* cmvn_dev, float* out_dev, int64_t out_ld, fa_stream_t stream);

/* Raw 80-bin log-mel (no LFR/CMVN), [m_b, n_mels] rows at out_row_dev[b]; used by parity tests. */
int fa_frontend_logmel(const fa_frontend_cfg* cfg, const void* tables_dev, const float* wav_dev,
                       const int64_t* wav_start_dev, const int32_t* wav_len_dev, const int32_t* out_row_dev, int32_t n_utts,
                       int32_t max_frames, float* out_dev, int64_t out_ld, fa_stream_t stream);

int32_t fa_num_frames(int32_t n_samples, int32_t frame_len, int32_t frame_shift); /* TA:62-66 */
int32_t fa_num_lfr(int32_t n_frames, int32_t lfr_n);                              /* VLLM:43 */
int32_t fa_fake_token_len(int32_t n_lfr, int32_t use_low_frame_rate);             /* model.py:391-396 */

/* ------------------------------------------------------------------ encoder / adaptor ------- */
typedef struct {
  /* SenseVoiceEncoderSmall(input_size, **audio_encoder_conf)   VLLM/model_executor/models/funasr.py:293-376 */
  int32_t d_in;        /* 560 */
  int32_t d_model;     /* 512  output_size */
  int32_t n_heads;     /* 4 */
  int32_t d_ff;        /* 2048 linear_units */
  int32_t fsmn_kernel; /* 11 */
  int32_t sanm_shift;  /* 0 */
  int32_t n_blocks;    /* 50  (encoders0 + encoders) */
  int32_t n_tp_blocks; /* 20 */
  /* adaptor Transformer(**audio_adaptor_conf)                  VLLM/...:528-572, 625-637 */
  int32_t ad_downsample; /* 1 (only 1 is compiled) */
  int32_t ad_ffn;        /* 2048 */
  int32_t ad_dim;        /* 1024 llm_dim */
  int32_t ad_heads;      /* 8 */
  int32_t ad_layers;     /* 2 */
  int32_t ad_ff;         /* 256 = llm_dim // 4 */
} fa_model_cfg;

/* One SAN-M block (EncoderLayerSANM).  Matrices bf16 row-major [out, in] (nn.Linear layout);
 * vectors and the FSMN taps f32. */
typedef struct {
  const float* ln1_g;
  const float* ln1_b;
  const void* qkv_w;    /* bf16 [3*d_model, d_in|d_model]   self_attn.linear_q_k_v.weight */
  const float* qkv_b;   /* [3*d_model] */
  const float* fsmn_w;  /* f32 [fsmn_kernel, d_model] TAP-MAJOR = self_attn.fsmn_block.weight[d_model,1,K] transposed */
  const void* out_w;    /* bf16 [d_model, d_model]           self_attn.linear_out.weight.  The FSMN memory
                           (VLLM/model_executor/models/funasr.py:253 `att_outs + fsmn_memory`) is added by the attention kernel
                           straight into the f32 residual stream; layer 0 (no residual) keeps it in bf16 beside the attention
                           context and this GEMM's epilogue adds it. */
  const float* out_b;
  const float* ln2_g;
  const float* ln2_b;
  const void* ffn1_w;   /* bf16 [d_ff, d_model]              feed_forward.w_1.weight */
  const float* ffn1_b;
  const void* ffn2_w;   /* bf16 [d_model, d_ff]              feed_forward.w_2.weight */
  const float* ffn2_b;
  /* optional (ABI 4): LayerNorm folded into the GEMMs (norm1 -> qkv, norm2 -> ffn1; VLLM/...:83-91,120,138).  When
     qkv_colsum is non-NULL, qkv_w holds bf16(W * gamma1) (column k scaled by gamma1[k]), qkv_b holds b + W beta1 and
     qkv_colsum[n] = sum_k float(qkv_w[n, k]); the plan then feeds the GEMM the raw bf16 residual stream and applies
     rstd * (acc - mean * colsum) + bias in the epilogue.  Same for ffn1_colsum / ffn1_w / ffn1_b with norm2.  Not allowed
     on layers whose input comes from a LayerNorm kernel (layer 0, and qkv of the first tp block). */
  const float* qkv_colsum;
  const float* ffn1_colsum;
} fa_sanm_layer_w;

typedef struct {
  const float* ln1_g;
  const float* ln1_b;
  const void* qkv_w;    /* bf16 [3*ad_dim, ad_dim] = cat(linear_q, linear_k, linear_v) */
  const float* qkv_b;
  const void* out_w;    /* bf16 [ad_dim, ad_dim] */
  const float* out_b;
  const float* ln2_g;
  const float* ln2_b;
  const void* ffn1_w;   /* bf16 [ad_ff, ad_dim] */
  const float* ffn1_b;
  const void* ffn2_w;   /* bf16 [ad_dim, ad_ff] */
  const float* ffn2_b;
} fa_adaptor_layer_w;

typedef struct {
  const fa_sanm_layer_w* layers;       /* HOST array [n_blocks + n_tp_blocks] of device pointers */
  const float* after_norm_g;
  const float* after_norm_b;
  const float* tp_norm_g;
  const float* tp_norm_b;
  const float* pe_inv_timescales;      /* f32 [d_in/2]  exp(-k*ln(1e4)/(d_in/2-1))  VLLM/...:268-276 */
  const void* lin1_w;                  /* bf16 [ad_ffn, d_model] */
  const float* lin1_b;
  const void* lin2_w;                  /* bf16 [ad_dim, ad_ffn] */
  const float* lin2_b;
  const fa_adaptor_layer_w* ad_layers; /* HOST array [ad_layers] */
} fa_model_w;

/* host-side plan: TMA descriptors + workspace carve-up. No device allocation.  A plan may hold the
 * encoder only (lin1_w == NULL), the adaptor only (n_blocks == 0, layers == NULL) or both.
 * The weights are read through the pointers in `w` at every forward call (nothing is copied), so they must stay alive
 * and unchanged for the life of the plan; the workspace must be zero-filled by the caller before fa_plan_create.
 * A plan is not thread safe: one host thread (or external locking) per plan. */
typedef struct fa_plan fa_plan;

size_t fa_plan_workspace_bytes(const fa_model_cfg* cfg, int32_t max_tokens);
int fa_plan_create(const fa_model_cfg* cfg, const fa_model_w* w, void* workspace_dev, size_t workspace_bytes,
                   int32_t max_tokens, fa_plan** plan_out);
void fa_plan_destroy(fa_plan* plan);
/* Dual-stream schedule (new).  n_streams == 2: a batch of >= 2 utterances and >= 2048 rows is cut on an utterance boundary
 * into two halves of equal cost; half 0 runs on the caller's stream, half 1 on a plan-owned auxiliary stream that is forked
 * from and joined back into the caller's stream with events inside the same call, so the call still only ENQUEUES work
 * ordered on `stream`.  The HBM-bound kernels of one half then share the SMs with the tensor-bound GEMMs of the other.
 * Results are bit-identical to n_streams == 1.  Needs cu_seqlens_host.  Environment override at plan creation: FA_STREAMS. */
int fa_plan_set_streams(fa_plan* plan, int32_t n_streams);
/* CUDA-graph replay (new; the reference's vLLM side captures graphs for its decode batches, infer_kaldidata.py:655-656).
 * A batch of at most max_rows packed rows does not issue the ~440 launches of a forward pass one by one: the plan's launch
 * program is captured once per BUCKET (rows rounded up to 1/8..1/16 of their size, utterances to a multiple of 8, the longest
 * utterance to a multiple of 128 rows) and replayed with one cudaGraphLaunch; utterance boundaries and the true row count
 * are read from cu_seqlens on the device, so one captured graph serves every batch of its bucket.  Results are bit-identical
 * to the launch-by-launch schedule.  Capturing instantiates a CUDA graph (the one place the library lets the driver allocate);
 * up to 48 graphs are kept per plan, least recently used first out.  max_rows == 0 turns replay off; larger batches use the
 * dual-stream schedule above.  Needs a workspace that was zero-filled before fa_plan_create (rows past the end of a batch
 * are computed on whatever the workspace holds and must stay finite). */
/* Chain schedule (new, opt-in).  The GEMMs between two attention kernels of the encoder - out-projection, ffn1, ffn2
 * of block i and the qkv projection of block i + 1 - touch every token row on its own, so they run as ONE cooperative launch
 * whose tiles wait on per-row-block counters instead of on grid boundaries (csrc/gemm_chain.cu): three launch boundaries
 * per block disappear and the FFN hidden rows are consumed out of L2.  Same tile arithmetic: results are bit-identical to
 * on == 0 (one launch per GEMM, the default: on B200 the chain's pipeline fill and drain - three dependent stages of ~20 us
 * each per launch - and its f32 epilogues still cost more than the launch boundaries they replace, profiles/r02/chain_r2ae.md).
 * Environment override at plan creation: FA_CHAIN=1. */
int fa_plan_set_chain(fa_plan* plan, int32_t on);
#define FA_DEFAULT_GRAPH_ROWS 8192
int fa_plan_set_graph_rows(fa_plan* plan, int32_t max_rows);

/* Layouts.  Every [rows, dim] activation argument comes with a `*_tmax`:
 *   tmax == 0 : PACKED  [n_tokens, dim], utterance b at rows [cu_seqlens[b], cu_seqlens[b+1])
 *   tmax  > 0 : PADDED  [n_utts, tmax, dim] exactly as the reference passes it (xs_pad, infer.py:455);
 *               on output the padding rows are zero-filled.
 * feats_dev is NOT modified (the reference scales it in place, VLLM/...:388).
 * enc_out_dev f32 [.., d_model] may be NULL: the result stays in the plan's workspace for
 * fa_adaptor_forward(enc_dev = NULL).
 * cu_seqlens_host: HOST copy of cu_seqlens_dev ([n_utts + 1] int32), or NULL.  It is only used to cut the batch into the two
 * halves of the dual-stream schedule (fa_plan_set_streams); with NULL the batch runs on the caller's stream alone. */
int fa_encoder_forward(fa_plan* plan, const float* feats_dev, int32_t feats_tmax, const int32_t* cu_seqlens_dev,
                       const int32_t* cu_seqlens_host, int32_t n_utts, int32_t n_tokens, int32_t max_len,
                       float* enc_out_dev, int32_t enc_out_tmax, fa_stream_t stream);
/* enc_dev f32 [.., d_model], or NULL to consume the plan-resident encoder output.
 * out_dev [.., ad_dim]: f32 when out_bf16 == 0, bf16 otherwise. */
int fa_adaptor_forward(fa_plan* plan, const float* enc_dev, int32_t enc_tmax, const int32_t* cu_seqlens_dev,
                       const int32_t* cu_seqlens_host, int32_t n_utts, int32_t n_tokens, int32_t max_len, void* out_dev,
                       int32_t out_tmax, int32_t out_bf16, fa_stream_t stream);
/* FunASRNano.forward_export: encoder then adaptor, the intermediate never leaves the workspace. */
int fa_encode_adapt(fa_plan* plan, const float* feats_dev, int32_t feats_tmax, const int32_t* cu_seqlens_dev,
                    const int32_t* cu_seqlens_host, int32_t n_utts, int32_t n_tokens, int32_t max_len, float* enc_out_dev,
                    int32_t enc_out_tmax, void* out_dev, int32_t out_tmax, int32_t out_bf16, fa_stream_t stream);

/* out rows: utterance b -> [prefix (P rows) | speech rows [cu[b], cu[b]+take_b) | suffix (S rows)] at
 * out_off[b].  take_dev == NULL: take all rows.  All matrices [rows, dim]; speech f32 or bf16,
 * prefix/suffix/out bf16 when *_bf16 is set else f32. */
int fa_splice_embeds(const void* speech_dev, int32_t speech_bf16, const int32_t* cu_seqlens_dev,
                     const int32_t* take_dev, const void* prefix_dev, int32_t n_prefix, const void* suffix_dev,
                     int32_t n_suffix, const int64_t* out_off_dev, int32_t n_utts, int32_t dim, void* out_dev,
                     int32_t out_bf16, int32_t max_rows_per_utt, fa_stream_t stream);

/* Audio ingest: polyphase windowed-sinc resampling of a packed batch (utterance b = in_len[b] samples at in_start[b]) to
 * out_start[b] .. + ceil(new*len/orig).  `kernel_dev` is the [new_freq, taps_padded] f32 table (rows zero-padded from 2*width+orig taps to a multiple of 4) of
 * TA/functional/functional.py:1305-1402 (orig/new already divided by their gcd), computed once on the host.  One CTA
 * produces `frames_per_tile` (multiple of 32 * 8/min(8, pow2 <= new_freq)) output frames of `new_freq` samples; tile_off_dev [n_utts+1] is the prefix sum
 * of ceil((len/orig + 1) / frames_per_tile) per utterance and n_tiles its last entry. */
int fa_resample_sinc(const float* in_dev, const int64_t* in_start_dev, const int32_t* in_len_dev, const float* kernel_dev,
                     int32_t orig, int32_t new_freq, int32_t width, int32_t taps_padded, int32_t frames_per_tile,
                     float* out_dev, const int64_t* out_start_dev, const int32_t* tile_off_dev, int32_t n_utts,
                     int32_t n_tiles, fa_stream_t stream);

/* Raw PCM in: interleaved frames of `channels` samples (in_start / in_len count FRAMES), one of */
#define FA_PCM_F32 0
#define FA_PCM_S16 1
#define FA_PCM_F64 2
/* mono[s] = (sum over channels, in order, of f32(x[s][c]) * scale) / channels  -  `torch.from_numpy(audio).float()`
 * (infer.py:320-324; scale 1) and librosa.load's soundfile scaling (1/32768 for 16-bit files) + to_mono mean
 * (infer_kaldidata.py:101).  Output utterance b starts at out_start[b] (f32 samples). */
int fa_pcm_to_mono_f32(const void* in_dev, int32_t format, int32_t channels, float scale, const int64_t* in_start_dev,
                       const int32_t* in_len_dev, float* out_dev, const int64_t* out_start_dev, int32_t n_utts,
                       int32_t max_len, fa_stream_t stream);
/* fa_resample_sinc reading raw PCM: conversion and down-mix happen while the input span is staged, so the raw samples
 * cross HBM once and no f32 mono copy at the source rate is ever written. */
int fa_resample_sinc_pcm(const void* in_dev, int32_t format, int32_t channels, float scale, const int64_t* in_start_dev,
                         const int32_t* in_len_dev, const float* kernel_dev, int32_t orig, int32_t new_freq, int32_t width,
                         int32_t taps_padded, int32_t frames_per_tile, float* out_dev, const int64_t* out_start_dev,
                         const int32_t* tile_off_dev, int32_t n_utts, int32_t n_tiles, fa_stream_t stream);

/* ------------------------------------------------------------------ device timing (bench.py) */
/* When enabled, every kernel launch the library makes is bracketed by a CUDA event pair on the
 * caller's stream.  fa_profile_collect synchronises on them and returns, per kernel class
 * (0 frontend, 1 gemm, 2 attention, 3 layernorm, 4 fsmn, 5 repack, 6 splice), the summed device time
 * in ms and the launch count since the last collect.  Not thread safe; off by default.
 * (New: the reference has no device timing, only time.perf_counter around whole calls, model.py:361-386.) */
#define FA_PROFILE_KINDS 7
int fa_profile_enable(int32_t on);
int fa_profile_collect(float* ms_per_kind, int32_t* launches_per_kind, int32_t n_kinds);

/* ------------------------------------------------------------------ single ops (parity tests) */
/* C[M,N] = A[M,K] (bf16) * W[N,K]^T (bf16) + bias; relu optional; then
 *   out_f32 == 0: bf16 C;   out_f32 == 1: f32 C = acc + bias (+ res1) (+ res2) (+ res_bf16)
 *   (res1/res2 f32 [M,N] with row stride ldc, or NULL; res_bf16 bf16 [M,N] with row stride ld_res_bf16, or NULL, not
 *   together with res2) */
int fa_op_gemm(const void* a_dev, int64_t lda, const void* w_dev, int64_t ldw, const float* bias_dev,
               const float* res1_dev, const float* res2_dev, const void* res_bf16_dev, int64_t ld_res_bf16, void* c_dev,
               int64_t ldc, int32_t M, int32_t N, int32_t K, int32_t relu, int32_t out_f32, fa_stream_t stream);
/* One GEMM with a folded-LayerNorm epilogue (how the plan runs norm1 -> qkv and norm2 -> ffn1 without an LN kernel).
   ln_mode 2 (producer): c (f32, may alias res1) = a w^T + bias + res1 (+ res_bf16, bf16 [M,N] or NULL); xb = bf16(c); ln_stats[M, *n_slots] (float2 = sum, sum
   of squares of the bf16 row pieces; *n_slots is written).  ln_mode 1 (consumer): a = raw bf16 rows, w = bf16(W * gamma),
   bias = b + W beta, colsum[n] = sum_k w[n, k]; c = rstd * (a w^T - mean * colsum) + bias over ln_dim columns, eps 1e-12. */
int fa_op_gemm_ln(const void* a_dev, int64_t lda, const void* w_dev, int64_t ldw, const float* bias_dev, const float* res1_dev,
                  void* c_dev, int64_t ldc, int32_t M, int32_t N, int32_t K, int32_t relu, int32_t out_f32, int32_t ln_mode,
                  void* ln_stats_dev, int32_t* n_slots, int32_t ln_dim, const float* colsum_dev, void* xb_dev, int64_t ld_xb,
                  const void* res_bf16_dev, int64_t ld_res_bf16, fa_stream_t stream);
/* y = LN(x) (eps 1e-12); y_bf16 and/or y_f32 may be NULL */
int fa_op_layernorm(const float* x_dev, const float* g_dev, const float* b_dev, void* y_bf16_dev, float* y_f32_dev,
                    int32_t M, int32_t D, fa_stream_t stream);
/* y_bf16 = LN(x*scale + PE(pos))  (encoder input stage)   VLLM/...:388-390 + norm1 of encoders0 */
int fa_op_embed_layernorm(const float* x_dev, const float* inv_ts_dev, const int32_t* cu_seqlens_dev, int32_t n_utts,
                          float scale, const float* g_dev, const float* b_dev, void* y_bf16_dev, int32_t M, int32_t D,
                          fa_stream_t stream);
/* mem[M,D] f32 = v + depthwise_conv_time(v)   v = qkv[:, v_col : v_col+D] bf16; w_dev f32 [kernel, D] tap-major
 * VLLM/...:181-201 */
int fa_op_fsmn(const void* qkv_dev, int64_t ld_qkv, int32_t v_col, const float* w_dev, const int32_t* cu_seqlens_dev,
               int32_t n_utts, float* mem_dev, int32_t M, int32_t D, int32_t kernel, int32_t left_pad,
               fa_stream_t stream);
/* same, bf16 output with row stride ld_mem (the form the plan uses: memory written beside the attention context) */
int fa_op_fsmn_bf16(const void* qkv_dev, int64_t ld_qkv, int32_t v_col, const float* w_dev, const int32_t* cu_seqlens_dev,
                    int32_t n_utts, void* mem_bf16_dev, int64_t ld_mem, int32_t M, int32_t D, int32_t kernel,
                    int32_t left_pad, fa_stream_t stream);
/* ctx[M, H*128] bf16 = softmax(q k^T / sqrt(128)) v per utterance and head; q|k|v packed in the n_rows (= M) qkv rows */
int fa_op_attention(const void* qkv_dev, int64_t ld_qkv, const int32_t* cu_seqlens_dev, int32_t n_utts,
                    int32_t max_len, int32_t n_heads, int32_t n_rows, void* ctx_dev, int64_t ld_ctx, fa_stream_t stream);
/* attention and the FSMN memory of the same value rows in ONE kernel (the plan's form): ctx as above, and
 * mem[r, h*128 + c] bf16 (row stride ld_mem) = v + depthwise_conv_time(v), w f32 [11, H*128] tap-major, `fsmn_left` rows of
 * left context.  Only the model's 11-tap kernel is fused (FA_ERR_INVALID otherwise).  VLLM/...:181-201,222-252
 * acc_f32_dev != NULL (the plan's form for every block with a residual out-projection): the memory is not stored but ADDED,
 * in f32, to acc[r, h*128 + c] (row stride ld_acc) - the residual stream - and mem_bf16_dev may be NULL. */
int fa_op_attention_fsmn(const void* qkv_dev, int64_t ld_qkv, const int32_t* cu_seqlens_dev, int32_t n_utts,
                         int32_t max_len, int32_t n_heads, int32_t n_rows, void* ctx_dev, int64_t ld_ctx,
                         const float* fsmn_w_dev, int32_t fsmn_kernel, int32_t fsmn_left, void* mem_bf16_dev,
                         int64_t ld_mem, float* acc_f32_dev, int64_t ld_acc, fa_stream_t stream);

/* ------------------------------------------------------------------ allocation-free variant
 * No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
 *   C[M,N] = LayerNorm(x[M,D]; g, b, eps 1e-12) (rounded to bf16) * W[N,D]^T (bf16) + bias; bf16 C (relu optional), or f32
 *   C with out_f32 == 1.  Same two kernels as fa_op_layernorm(y_bf16) followed by fa_op_gemm, bit for bit; the
 *   normalized bf16 rows between them live in a file-scope `static __device__` array of FA_STATIC_LN_ELEMS elements
 *   (4 MiB, held by every GPU that loads the library until exit) instead of a caller buffer, and the call makes no
 *   cudaMalloc / cudaFree / allocating library call.
 *   Limit: M * D <= FA_STATIC_LN_ELEMS (4096 rows at D = 512), D % 8 == 0; anything larger is refused with an error.
 *   All calls on a device share the array (it keeps the last call's contents; nothing in it needs clearing, a call
 *   writes every row it reads): calls that overlap on the GPU - two streams, two host threads - race. */
#define FA_STATIC_LN_ELEMS (2 * 1024 * 1024)
size_t fa_static_ln_capacity(void); /* = FA_STATIC_LN_ELEMS */
int fa_op_ln_linear_static(const float* x_dev, const float* g_dev, const float* b_dev, const void* w_dev, int64_t ldw,
                           const float* bias_dev, void* c_dev, int64_t ldc, int32_t M, int32_t N, int32_t D, int32_t relu,
                           int32_t out_f32, fa_stream_t stream);

#if defined(__GNUC__)
#pragma GCC visibility pop
#endif
#ifdef __cplusplus
}
#endif
#endif /* FUNASR_B200_H_ */
