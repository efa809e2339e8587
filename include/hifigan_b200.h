/*
 * hifigan_b200.h -- C ABI of libhifigan_b200.so
 *
 * Drop-in boundary for ONE path of dmmagdal/NeuralTextToSpeech: the HiFi-GAN
 * generator forward pass, `Generator.call(mel) -> audio`
 * (reference HiFiGAN_TF/model.py:115-173).  The reference has no FFI of its own
 * (the boundary there is the Python class `Generator(keras.Model)`); these entry
 * points are what a ctypes binding of that class needs -- see INTEGRATION.md for
 * the reference-side stub.  Plain pointers and sizes only, no torch/TF types.
 *
 * All functions return 0 on success or a negative hfg_status; the message is
 * available from hfg_last_error().  Nothing here prints or exits.
 * A handle is bound to one CUDA device and is not thread-safe (one handle per
 * GPU / stream).  hfg_forward is asynchronous on `stream` and performs no
 * allocation; it fails (never falls back to a CPU path) if no CUDA device exists.
 */
#ifndef HIFIGAN_B200_H_
#define HIFIGAN_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define HFG_MAX_UPSAMPLES 8
#define HFG_MAX_KERNELS 8
#define HFG_MAX_DILATIONS 4

typedef enum {
  HFG_OK = 0,
  HFG_ERR_INVALID = -1,   /* bad argument / config / shape  -> ValueError in the shim   */
  HFG_ERR_STATE = -2,     /* call order (weights missing, not finalised ...)           */
  HFG_ERR_CUDA = -3,      /* CUDA runtime error                     -> RuntimeError     */
  HFG_ERR_NOMEM = -4,     /* workspace too small / allocation failed                   */
  HFG_ERR_UNSUPPORTED = -5
} hfg_status;

/* Arithmetic of the forward pass (hfg_finalize_weights). */
typedef enum {
  HFG_PRECISION_FP32 = 0, /* fp32 FFMA validation mode: <=1e-5 of the fp32 reference    */
  HFG_PRECISION_FP16 = 1, /* tcgen05 kind::f16, fp16 operands+storage, fp32 accumulate  */
  HFG_PRECISION_BF16 = 2, /* tcgen05 kind::f16, bf16 operands+storage, fp32 accumulate  */
  HFG_PRECISION_TF32 = 3  /* tcgen05 kind::tf32, fp32 storage (operands rounded to TF32), fp32 accumulate: fp32 range,
                             meets the tensor-core tolerance (<= 2e-3 max-abs, >= 40 dB) like fp16      */
} hfg_precision;

/* POD mirror of the hparams fields Generator.__init__ reads
 * (reference HiFiGAN_TF/model.py:119-143, HiFiGAN_TF/hparams.py:9-41) + num_mels
 * (the reference lets Keras infer it at first call, model.py:121-125). */
typedef struct {
  int32_t num_mels;                                   /* config_v*.json "num_mels" (80)            */
  int32_t upsample_initial_channel;                   /* 512 / 128 / 256                           */
  int32_t resblock;                                   /* 1 -> ResBlock1, 2 -> ResBlock2 (model.py:126) */
  int32_t num_upsamples;
  int32_t upsample_rates[HFG_MAX_UPSAMPLES];
  int32_t upsample_kernel_sizes[HFG_MAX_UPSAMPLES];
  int32_t num_kernels;
  int32_t resblock_kernel_sizes[HFG_MAX_KERNELS];
  int32_t num_dilations[HFG_MAX_KERNELS];
  int32_t resblock_dilation_sizes[HFG_MAX_KERNELS][HFG_MAX_DILATIONS];
  int32_t compat;  /* 0 = reference semantics (causal k=3,d=1 convs via get_padding,
                        HiFiGAN_TF/utils.py:29-37; final LeakyReLU slope 0.3, model.py:152);
                      1 = upstream jik876/hifi-gan semantics (symmetric pads, slope 0.01)  */
} hfg_config;

typedef struct hfg_ctx* hfg_handle;

/* Replaces Generator.__init__ (model.py:116-152).  `device` = CUDA ordinal. */
int hfg_create(const hfg_config* cfg, int device, hfg_handle* out);

/* Replaces the Keras variables of each tfa.layers.WeightNormalization wrapper
 * (model.py:25-66,88-101,121,131,145): name is "<layer>.v" | "<layer>.g" | "<layer>.b"
 * with <layer> in {conv_pre, ups.<i>, resblocks.<n>.convs1.<m>, resblocks.<n>.convs2.<m>,
 * resblocks.<n>.convs.<m>, conv_post}; `host` is fp32 in Keras layout
 * (v: Conv1D [k,Cin,Cout], Conv1DTranspose [k,Cout,Cin]; g: last axis of v; b: [Cout]). */
int hfg_set_weight(hfg_handle h, const char* name, const float* host, const int64_t* shape, int ndim);

/* Folds weight norm (kernel = g * v / max(||v||, 1e-6), norm over all axes but the last),
 * repacks for the kernels and uploads.  precision: hfg_precision. */
int hfg_finalize_weights(hfg_handle h, int precision);

/* Output samples per mel frame (prod(upsample_rates)). */
int hfg_total_upsample(hfg_handle h, int* out);

/* Device scratch needed by hfg_forward for a [B,T,num_mels] batch. */
int hfg_workspace_bytes(hfg_handle h, int B, int T, size_t* out);

/* Replaces Generator.call (model.py:155-173).
 *   mel_dev   : device, fp32, [B, T, num_mels] C-contiguous (channels-last, like the reference)
 *   audio_dev : device, fp32, [B, T*prod(rates), 1]
 *   workspace : device, >= hfg_workspace_bytes, 256-byte aligned; caller-owned
 *   stream    : cudaStream_t (NULL = legacy default stream)
 * Asynchronous on `stream`.  Launch-bound forwards (B*T <= 8192 frames, option "cuda_graphs" on, `stream` not being
 * captured by the caller) are replayed from a CUDA graph captured on the second call with the same
 * (mel_dev, audio_dev, workspace_dev, B, T): the graph is captured on an internal stream and launched into `stream`. */
int hfg_forward(hfg_handle h, const float* mel_dev, int B, int T, float* audio_dev,
                void* workspace_dev, size_t workspace_bytes, void* stream);

/* Same call with HOST buffers (what a numpy caller of Generator.call uses): copies mel to the
 * device, runs hfg_forward in micro-batches on an internal stream and workspace (grown on demand,
 * kept in the handle), copies audio back and synchronises.  Pinned host memory makes the copies
 * asynchronous; pageable memory works too. */
int hfg_forward_host(hfg_handle h, const float* mel_host, int B, int T, float* audio_host);

/* 16-bit PCM output, fused into the conv_post epilogue: sample = (int16)trunc(clamp(audio, -1, 1)
 * * 32768), +1.0 saturating to 32767 -- the step the reference intends after the vocoder
 * (Grad-TTS_TF/inference.py:89; FastSpeech2_TF/FastSpeech2/utils/model.py:82-85).  pcm: [B, T*prod(rates)] int16.
 * Halves the device->host bytes of the host variant. */
int hfg_forward_pcm16(hfg_handle h, const float* mel_dev, int B, int T, int16_t* pcm_dev,
                      void* workspace_dev, size_t workspace_bytes, void* stream);
int hfg_forward_host_pcm16(hfg_handle h, const float* mel_host, int B, int T, int16_t* pcm_host);

/* Options: "micro_batch" = utterances per forward of hfg_forward_host (0 = auto, the default: hfg_auto_micro_batch);
 * "stage_timing" 0/1; "fuse_resblocks" 0/1 (default 1: every
 * ResBlock1 iteration of a stage with C <= 64 is one launch, needs a new hfg_finalize_weights); "cuda_graphs" 0/1 (default 1: launch-bound micro-batches of hfg_forward_host, <= 8192 frames,
 * are replayed from a CUDA graph captured on their second use); "concurrent_resblocks" = largest B*T in frames
 * (default 2048, 0 = never) for which the resblocks of a stage -- independent up to their last conv, reference
 * HiFiGAN_TF/model.py:161-166 -- run as concurrent chains on internal streams forked from and joined to the caller's
 * stream (same kernels, same bits; the workspace holds two more P/Q/T sets per extra chain, so query
 * hfg_workspace_bytes after changing it); "force_ffma" 0/1 and "auto_fold" 0/1 (both need a new
 * hfg_finalize_weights). */
int hfg_set_option(hfg_handle h, const char* key, long long value);

/* Which kernel family each layer was bound to by hfg_finalize_weights: lines "<layer>:<tc|ffma|post>\n"
 * (tc = persistent tcgen05 kernel, ffma = direct fp32-accumulate kernel, post = conv_post+tanh kernel). */
int hfg_layer_engines(hfg_handle h, char* buf, size_t cap);

/* Utterances per forward the library would pick for a [B, T] batch (one activation buffer <= 1 GiB; the host path keeps
 * two micro-batches so copies overlap compute).  Callers of hfg_forward that split large batches can use the same rule. */
int hfg_auto_micro_batch(hfg_handle h, int B, int T, int host_path, int* out);

/* Time-fold factors chosen per layer on the tcgen05 path: lines "<layer>:<fold_in>:<fold_out>:<kernel-view taps>\n"
 * (a [L, C] tensor is bit-identical to [L/f, C*f]). */
int hfg_layer_folds(hfg_handle h, char* buf, size_t cap);

/* Kernels launched by the last hfg_forward / hfg_forward_host on this handle. */
int hfg_last_launch_count(hfg_handle h, long long* out);

/* Per-stage device time (CUDA events on the forward's stream) while option "stage_timing" is 1, averaged over
 * the forwards run since the previous call (the 128 most recent at most), then reset:
 * out[0]=conv_pre, out[1+i]=stage i (ups[i] + its resblocks), out[1+n]=conv_post. Returns count via n. */
int hfg_stage_times_ms(hfg_handle h, float* out, int* n);

/* Kernels launched per measurement stage by the last forward (same indexing as hfg_stage_times_ms; the mel cast of
 * the 16-bit path counts towards conv_pre): what one "launch" of a stage's roofline row means. */
int hfg_stage_launches(hfg_handle h, int* out, int* n);

/* Operator-level entry: ONE tapped convolution (the building block of every layer above); synchronous, for
 * parity tests.  A dilated Conv1D (reference model.py:26-66) is out_stride=1, tap_off[j] = j*d - pad_left; one
 * phase p of a Conv1DTranspose (model.py:131-136) is out_stride=u, out_off=p.  act: 0 fp32, 1 fp16, 2 bf16
 * activations; w_host fp32 [ntaps][c_in][c_out]; res / res2 / acc may be NULL.
 *   engine 0, direct FFMA kernel:
 *     v = bias[co] + sum_j sum_ci lrelu(x[b, q + tap_off[j], ci], in_slope) * w[j][ci][co] + res[...]
 *     mrf_mode 0: out = lrelu(v, out_slope); 1: acc = v; 2: acc += v; 3: out = lrelu((acc + v) / mrf_div, out_slope)
 *   engine 1, persistent tcgen05 kernel (16-bit only; x is stored already activated, in_slope must be 1):
 *     v = bias[co] + sum_j sum_ci x[b, q + tap_off[j], ci] * w[j][ci][co] + unact(res[...]) + res2[...]
 *     out = lrelu(v * out_scale, out_slope),   unact(y) = y >= 0 ? y : y * res_inv     (0 < out_slope <= 1 <= res_inv)
 *   out is indexed [b, q*out_stride + out_off, co]; res, res2, acc like out. */
int hfg_op_tapped_conv(int act, int engine, const void* x_dev, int B, int x_rows, int c_in,
                       const float* w_host, const float* bias_host, int c_out, int ntaps, const int* tap_off,
                       int out_stride, int out_off, float in_slope, const void* res_dev, float res_inv,
                       const void* res2_dev, float out_scale, float out_slope, float* acc_dev,
                       int mrf_mode, float mrf_div, void* out_dev, void* stream);

/* The same direct (engine 0) tapped convolution for callers that cannot allocate device memory (CUDA-graph capture, tight
 * serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * The device copy of the weights and the bias that hfg_op_tapped_conv cudaMallocs per call lives here in one
 * static device array of HFG_NOALLOC_SCRATCH_BYTES, which every GPU that loads the library holds until the process exits:
 * a layer needs round_up(c_out * 4, 256) + ntaps * c_in * c_out * 4 bytes of it and larger ones are rejected with
 * HFG_ERR_NOMEM.  No cudaMalloc / cudaFree / allocating routine is called; two host-to-device copies, one clear and the
 * kernel are enqueued on `stream` and the call returns WITHOUT synchronising (w_host / bias_host may be reused on return
 * when pageable; page-locked ones must stay unchanged until the stream reaches the copies).  All calls share the scratch,
 * which keeps the previous call's contents: calls on one stream are ordered, calls that overlap on the GPU (two streams
 * without an ordering, two host threads) race.  Runs on the current device.
 *   v = bias[co] + sum_j sum_ci lrelu(x[b, q + tap_off[j], ci], in_slope) * w[j][ci][co] + res[...];  out = lrelu(v, out_slope)
 * c_in % 4 == 0, c_out % 8 == 0; res may be NULL; err (may be NULL) receives the message of a failure. */
#define HFG_NOALLOC_SCRATCH_BYTES (4u << 20)
int hfg_op_tapped_conv_noalloc(int act, const void* x_dev, int B, int x_rows, int c_in, const float* w_host,
                               const float* bias_host, int c_out, int ntaps, const int* tap_off, int out_stride,
                               int out_off, float in_slope, const void* res_dev, float out_slope, void* out_dev,
                               void* stream, char* err, size_t err_cap);

/* Operator-level entry: ONE fused ResBlock1 iteration, xt = convs1[m](lrelu(x)); xt = convs2[m](lrelu(xt)); x = xt + x
 * (reference model.py:71-80), as a single launch whose intermediate stays in shared memory; synchronous, for parity
 * tests.  Kernel view of 64 channels per row (a [L, C] tensor with C < 64 is passed as [L/f, 64], f = 64/C, with
 * block-Toeplitz weights); act: 1 fp16, 2 bf16; w1/w2 fp32 [ntaps][64][64], b1/b2 [64]; x is stored activated:
 *   xt[q]  = lrelu(b1 + sum_j x[q + off1[j]] . W1_j, mid_slope), taken as 0 outside [0, rows)
 *   v[q]   = b2 + sum_j xt[q + off2[j]] . W2_j + unact(x[q]) + res2[q]      unact(y) = y >= 0 ? y : y * res_inv
 *   out[q] = lrelu(v * out_scale, out_slope)            res2 (running MRF sum, model.py:161-167) may be NULL
 * mode 1 is that iteration; mode 2 is a whole ResBlock2 (model.py:106-112), where both convs carry their own skip:
 *   x1[q]  = b1 + sum_j x[q + off1[j]] . W1_j + unact(x[q]);  xt = lrelu(x1, mid_slope) (16-bit), 0 outside [0, rows)
 *   v[q]   = b2 + sum_j xt[q + off2[j]] . W2_j + unact_mid(xt[q]) + res2[q] */
int hfg_op_resblock_fused(int act, int mode, const void* x_dev, int B, int rows, const float* w1_host, const float* b1_host, int ntaps1,
                          const int* off1, const float* w2_host, const float* b2_host, int ntaps2, const int* off2,
                          float mid_slope, float res_inv, const void* res2_dev, float out_scale, float out_slope,
                          void* out_dev, void* stream);

/* Pinned host memory helpers for callers without a CUDA runtime binding of their own. */
int hfg_host_alloc(void** out, size_t bytes);
int hfg_host_free(void* p);

/* Message of the last error on this handle (or of the last failed hfg_create when h is NULL). */
const char* hfg_last_error(hfg_handle h);

/* Library/ABI version: major*1000 + minor. */
int hfg_version(void);

void hfg_destroy(hfg_handle h);

/* ---------------------------------------------------------------------------------------------------------
 * Mel front-end: the step BEFORE the generator, `STFT.mel_spectrogram`
 * (reference HiFiGAN_TF/common/audio_processing_tf.py:41-104; callers data.py:142,210, gan.py:48-52,147):
 * tf.signal.stft (periodic Hann, no centre padding) -> power -> HTK mel filterbank -> log(clip(., 1e-5)).
 * The output layout is the generator's input layout. */
typedef struct {                 /* mirrors STFT.__init__ (audio_processing_tf.py:47-58) */
  int32_t fft_length;            /* filter_length (1024); 256, 1024 or 4096                          */
  int32_t frame_length;          /* 1024; <= fft_length (frames are zero-padded at the end)          */
  int32_t frame_step;            /* 256                                                              */
  int32_t n_mel_channels;        /* 80                                                               */
  int32_t sampling_rate;         /* 22050                                                            */
  float mel_fmin, mel_fmax;      /* 0, 8000                                                          */
} hfg_mel_config;

typedef struct hfg_mel_ctx* hfg_mel_handle;

int hfg_mel_create(const hfg_mel_config* cfg, int device, hfg_mel_handle* out);

/* Frames of an n_samples utterance: 1 + (n_samples - frame_length) / frame_step, 0 if shorter than one frame. */
int hfg_mel_num_frames(hfg_mel_handle h, long long n_samples, int* out);

/* wav_dev: device fp32 [B, n_samples]; mel_dev: device fp32 [B, T, n_mel_channels]; asynchronous on `stream`. */
int hfg_mel_forward(hfg_mel_handle h, const float* wav_dev, int B, long long n_samples, float* mel_dev, void* stream);

/* Same with HOST buffers (H2D, kernel, D2H, synchronise). */
int hfg_mel_forward_host(hfg_mel_handle h, const float* wav_host, int B, long long n_samples, float* mel_host);

const char* hfg_mel_last_error(hfg_mel_handle h);
void hfg_mel_destroy(hfg_mel_handle h);

/* ---------------------------------------------------------------------------------------------------------
 * Discriminator forward passes: `MultiPeriodDiscriminator` (reference HiFiGAN_TF/model.py:246-274, five
 * `DiscriminatorP` model.py:176-243) and `MultiScaleDiscriminator` (model.py:349-389, three `DiscriminatorS`
 * model.py:277-346) on ONE waveform batch; the reference's `call((y, y_hat))` is two such passes (or one pass over the
 * concatenated batch: no op couples utterances).  fp32 arithmetic.  Callers: gan.py:32-131 (training step).
 *   compat 0 = the reference's Keras semantics (TensorFlow SAME padding with strides, "valid" first DiscriminatorS conv,
 *              average pooling that ignores padding); 1 = upstream torch padding (what the vendored torch copies compute).
 * Weight names: "discriminators.<i>.convs.<l>.{v,g,b}" and "discriminators.<i>.conv_post.{v,g,b}" for
 * WeightNormalization-wrapped layers (kernel [k, c_in / groups, c_out]; Conv2D((k, 1)) kernels [k, 1, c_in, c_out] have
 * the same element order), or ".w" + ".b" for a kernel used as stored (SpectralNormalization at inference, model.py:353). */
typedef struct hfg_disc_ctx* hfg_disc_handle;

/* kind 0 = MultiPeriodDiscriminator (model.py:247-256), 1 = MultiScaleDiscriminator (model.py:350-364). */
int hfg_disc_create(int kind, int compat, int device, hfg_disc_handle* out);
int hfg_disc_set_weight(hfg_disc_handle h, const char* name, const float* host, const int64_t* shape, int ndim);
/* Folds weight norm (w = v * g / ||v||, per output channel) and uploads; needed before a forward. */
int hfg_disc_finalize_weights(hfg_disc_handle h);

/* Number of feature maps a forward writes: sub-discriminators x (convs + conv_post) = 30 (MPD) / 24 (MSD), in the order of
 * the reference's fmap lists (model.py:236-241, 337-342); the last map of each sub-discriminator is its score. */
int hfg_disc_num_maps(hfg_disc_handle h, int* out);
/* Where each map lives in the output buffer of a [B, T] forward: offsets[i] in floats, dims[4 i ..] = {B, period, len, C}
 * in MEMORY order (period = 1 for DiscriminatorS).  The reference's DiscriminatorP maps are [B, len, period, C]: swap
 * axes 1 and 2 of the view (model.py:235 reshapes to [b, t / period, period, c]).  offsets / dims may be NULL. */
int hfg_disc_layout(hfg_disc_handle h, int B, long long T, long long* offsets, long long* dims, long long* total_floats);

/* audio_dev: device fp32 [B, T] (the reference's [B, T, 1]); maps_dev: device fp32 [total_floats]; asynchronous on `stream`.
 * Arithmetic: fp32 results; every conv but the first (C_in = 1) and conv_post (C_out = 1) of a sub-discriminator runs on the
 * tensor cores as three TF32 products per fp32 product (within 2e-4 of the fp32 oracle per map, measured <= 6.2e-5;
 * environment HFG_DISC_TC=0 at hfg_disc_create keeps everything on fp32 FFMA, <= 2e-5).  The first forward of a new
 * (B, T) may allocate scratch and upload weight images (synchronously); forwards on one handle must not overlap. */
int hfg_disc_forward(hfg_disc_handle h, const float* audio_dev, int B, long long T, float* maps_dev, void* stream);
/* Same with HOST buffers (H2D, kernels, D2H, synchronise). */
int hfg_disc_forward_host(hfg_disc_handle h, const float* audio_host, int B, long long T, float* maps_host);

/* The GAN loss terms of the reference (HiFiGAN_TF/losses.py:8-43, called from gan.py:58-92) for ONE discriminator, from the
 * maps of a forward over the concatenated batch [y; y_hat] (2 B waveforms, real half first: what `call((y, y_hat))` runs).
 * out_host[3 + 3 n], n = sub-discriminators: [0] feature_loss (2 x sum over every map of mean |r - g|), [1] discriminator_loss
 * (sum of mean (1 - D(y))^2 + mean D(y_hat)^2), [2] generator_loss (sum of mean (1 - D(y_hat))^2), then r_losses[n],
 * g_losses[n], gen_losses[n].  Deterministic (fixed summation order); synchronises `stream`.  B = waveforms per half. */
int hfg_disc_pair_losses(hfg_disc_handle h, const float* maps_dev, int B, long long T, float* out_host, void* stream);

const char* hfg_disc_last_error(hfg_disc_handle h);
void hfg_disc_destroy(hfg_disc_handle h);

#ifdef __cplusplus
}
#endif
#endif /* HIFIGAN_B200_H_ */
