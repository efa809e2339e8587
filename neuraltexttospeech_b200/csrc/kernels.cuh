// Internal interface between the engine (engine.cu) and the kernel translation units.
#pragma once
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace hfg {

constexpr int kMaxTaps = 24;      // a k=11 dilated conv folded by 2 has 17 kernel-view taps

enum ActType : int { ACT_F32 = 0, ACT_F16 = 1, ACT_BF16 = 2 };

// MRF fusion in the epilogue of the last conv of a resblock
// (reference HiFiGAN_TF/model.py:161-167: xs = rb0(x)+rb1(x)+rb2(x); x = xs / num_kernels).
enum MrfMode : int {
  MRF_NONE = 0,    // out = val
  MRF_FIRST = 1,   // acc = val
  MRF_ADD = 2,     // acc += val
  MRF_FINAL = 3    // out = (acc + val) / num_kernels
};

// One "tapped GEMM": out[b, q*out_stride + out_off, co] =
//     epilogue( bias[co] + sum_j sum_ci act(x[b, q + tap_off[j], ci]) * w[j][ci][co] )
// A stride-1 (dilated) Conv1D is out_stride=1, tap_off[j] = j*d - pad_left.
// One output phase p of a Conv1DTranspose (stride u) is out_stride=u, out_off=p with the
// k/u taps j == (p+P) mod u at input offsets (p+P-j)/u   (polyphase, no zero insertion).
struct ConvArgs {
  const void* x;        // [B, x_rows, c_in] activations (ActType)
  const void* w;        // fp32 [ntaps][c_in][c_out]
  const float* bias;    // [c_out] fp32
  void* out;            // [B, out_rows, c_out] (ActType)
  const void* res;      // residual, indexed like out (ActType) or nullptr
  float* acc;           // MRF accumulator fp32, indexed like out, or nullptr
  int B;
  int x_rows;           // rows per batch item in x
  int n_q;              // q range per batch item
  int out_rows;         // rows per batch item in out
  int c_in, c_out;
  int ntaps;
  int tap_off[kMaxTaps];
  int out_stride, out_off;
  float in_slope;       // leaky-relu slope applied to x while staging (1 = identity)
  int mrf_mode;
  float mrf_div;        // num_kernels
  float out_slope;      // leaky-relu applied to the stored output (MRF_NONE / MRF_FINAL), 1 = none: used when the
                        // consumer is the tcgen05 kernel, which expects its input already activated
};

// conv_post + tanh (reference model.py:169-171): C_out = 1
struct PostArgs {
  const void* x;        // [B, rows, c_in]
  const float* w;       // fp32 [ntaps][c_in]
  float bias;
  float* out;           // fp32 [B, rows]
  int16_t* out_pcm;     // when non-NULL: 16-bit PCM [B, rows] instead of `out` (see pcm16_of)
  int B, rows, c_in, ntaps, pad_left;
  float in_slope;
};

// fp32 waveform sample -> 16-bit PCM, fused into the conv_post epilogue: the step the reference intends after the
// vocoder, `(audio.clamp(-1, 1) * 32768).astype(int16)` (Grad-TTS_TF/inference.py:89, same pattern in
// FastSpeech2_TF/FastSpeech2/utils/model.py:82-85): clamp, scale, truncate toward zero; +1.0 saturates to 32767
// (the numpy cast would wrap it to -32768).
__host__ __device__ inline int16_t pcm16_of(float v) {
  v = v < -1.f ? -1.f : (v > 1.f ? 1.f : v);
  const int i = (int)(v * 32768.f);
  return (int16_t)(i > 32767 ? 32767 : i);
}

// fp32 / generic direct kernels (conv_ffma.cu)
cudaError_t launch_conv_ffma(const ConvArgs& a, ActType act, cudaStream_t s);
cudaError_t launch_conv_post(const PostArgs& a, ActType act, cudaStream_t s);
cudaError_t launch_cast_pad(const float* src, void* dst, long long rows, int c_src, int c_dst,
                            ActType act, cudaStream_t s);

// ---------------------------------------------------------------- persistent tcgen05 kernel (conv_tc.cu)
// One launch = one layer: every output phase ("variant") of a Conv1DTranspose, or the single
// variant of a Conv1D.  x is stored already activated for this layer (see conv_tc.cu):
//   v = bias[co] + sum_j sum_ci x[b, q + var.tap_off[j], ci] * W_var[j][ci][co]
//       + unact(res[...]) + res2[...]                      unact(y) = y >= 0 ? y : y * res_inv
//   out[b, q*out_stride + var.out_off, co] = lrelu(v * out_scale, out_slope)
// res / res2 are indexed like out (16-bit, may be NULL): res is the ResBlock skip connection
// (reference model.py:79,111) stored activated with slope 1/res_inv, res2 the running raw MRF sum xs
// (model.py:161-166); out_scale = 1/num_kernels on the last resblock (model.py:167); out_slope is the
// activation the consumer of `out` would apply (model.py:73,76,108,159,169), 1 = store raw.
constexpr int kMaxVariants = 8;
struct TcVariant {
  const void* w;        // packed weights of this variant (see tc_pack_weights)
  int ntaps;
  int out_off;
  int tap_off[kMaxTaps];
  unsigned kz[kMaxTaps];   // 16-bit kernel only: bit kk of kz[j] = rows 16 kk .. 16 kk + 15 of tap j's weight block are all zero
                           // (time folding, channel padding): that K-step's MMA is not issued.  0 = issue everything
};
struct TcArgs {
  const void* x;
  void* out;
  const void* res;
  const void* res2;
  const float* bias;
  int B, x_rows, n_q, out_rows, out_stride;
  int c_in, c_out;
  float res_inv, out_scale, out_slope;
  int nvar;
  TcVariant var[kMaxVariants];
  int post_f;           // conv_post mode: samples per row (time folding): audio[b, q*post_f + s] = tanh(output channel s)
  float* post_out;      // conv_post mode (may be NULL): fp32 [B, n_q*post_f]; `out` is then only a scratch
                        // buffer of >= B*n_q*c_out 16-bit elements for the (unused) output tensor map
  int16_t* post_pcm;    // conv_post mode, when non-NULL: 16-bit PCM [B, n_q*post_f] instead of post_out
  int raw_out;          // TF32 kernel only: 1 = store the result unrounded (the running MRF sum is never an MMA operand)
  const void* w_base;   // allocation holding every variant's packed weights (256-byte aligned; the CTA-pair
  size_t w_bytes;       // configurations read it through one TMA map), and its size (multiple of 256)  // TF32 kernel only, extended form (discriminator layers, disc.cu): a strided and / or grouped Conv1D with up to
  // kTfMaxExtTaps taps.  ext_ntaps > 0 selects it: nvar must be 1 and var[0].tap_off / ntaps are ignored.
  //   out[b, q, co] = sum_j sum_ci x[b, q * in_stride + ext_tap_off[j], g(co) * cin_g + ci] * w[j][ci][co]
  const int* ext_tap_off;   // host array [ext_ntaps]: input-row offset of tap j (j - pad_left; taps may repeat an offset)
  int ext_ntaps;
  int in_stride;            // 1..4
  int cin_g;                // input channels per group (= c_in for a dense conv); the N tile is c_out / groups when grouped
  int groups;
};
struct TcConfig {       // tile configuration chosen for (c_in, c_out); see conv_tc.cu
  int n, mb, ks, kbw, tps, nst, nsl, nstg, cg;   // cg = 2: CTA pairs (cta_group::2)
};
bool tc_config(int c_in, int c_out, int span, TcConfig* cfg);   // span = max_off - min_off over ALL variants of the layer
bool tc_supported(int c_in, int c_out, int ntaps, int span);
size_t tc_packed_weight_bytes(int c_in, int c_out, int ntaps);
// host-side repack of fp32 [ntaps][c_in][c_out] into the kernel's smem-image weight stages (16-bit)
void tc_pack_weights(const float* w, int ntaps, int c_in, int c_out, int span, ActType act, void* dst_host);
// returns cudaSuccess or an error; *msg (optional) receives a static description on failure
cudaError_t launch_conv_tc(const TcArgs& a, ActType act, cudaStream_t s, int* grid_out);
void tc_set_trace(long long* dev);   // debug builds (-DHFG_TC_TRACE): device buffer of [6][64][8] clock stamps, see conv_tc.cu

// ---------------------------------------------------------------- TF32 tensor-core kernel (conv_tf32.cu)
// Same contract as launch_conv_tc (TcArgs; x / out / res / res2 are fp32, stored activated like the 16-bit path; no
// conv_post mode): tcgen05.mma.kind::tf32, fp32 storage with stored activations rounded to TF32.
bool tf32_supported(int c_in, int c_out, int ntaps, int span);
size_t tf32_packed_weight_bytes(int c_in, int c_out, int ntaps);
void tf32_pack_weights(const float* w, int ntaps, int c_in, int c_out, void* dst_host);
constexpr int kTfMaxExtTaps = 96;
// extended form: w fp32 [ntaps][cin_g][c_out] (Keras grouped-kernel layout), tap offsets as in TcArgs::ext_tap_off; the
// packed image has ntaps * cin_g * c_out floats.  tf32_ext_supported also needs x_rows >= in_stride at launch time.
bool tf32_ext_supported(int c_in, int c_out, int groups, int in_stride, int ntaps, const int* tap_off);
void tf32_pack_weights_ext(const float* w, int ntaps, const int* tap_off, int in_stride, int cin_g, int c_out, int groups, void* dst_host);
cudaError_t launch_conv_tf32(const TcArgs& a, cudaStream_t s);

// ---------------------------------------------------------------- fused ResBlock1 iteration (resblock_tc.cu)
// One launch = conv1 (dilated) -> lrelu -> conv2 -> + x (reference model.py:71-80), 64 kernel-view channels per row
// (C*fold = 64), both convs given as tap lists over kernel-view rows with 64x64 (block-Toeplitz when folded) weights:
//   xt[q]  = lrelu(b1 + sum_j x[q + off1[j]] . W1_j, mid_slope), 0 outside [0, n_q)      x stored activated
//   v[q]   = b2 + sum_j xt[q + off2[j]] . W2_j + unact(x[q]) + res2[q]                    unact(y) = y >= 0 ? y : y * res_inv
//   out[q] = lrelu(v * out_scale, out_slope)
struct RbArgs {
  const void* x;        // [B, n_q, 64] 16-bit, stored activated (slope 1/res_inv)
  void* out;            // [B, n_q, 64]
  const void* res2;     // running raw MRF sum, indexed like out, or nullptr
  const float* bias1; const float* bias2;   // [64] each (kernel view: the layer's bias replicated per fold slot)
  const void* w1; const void* w2;           // rb_pack_weights images (256-byte aligned)
  int cg;               // CTA grouping the weights were packed for (rb_supported)
  int mode;             // 1: ResBlock1 iteration (above); 2: whole ResBlock2 (model.py:106-112), both convs with their own skip:
                        //    x1 = b1 + conv1(x) + unact(x); xt = lrelu(x1); v = b2 + conv2(xt) + x1 (as stored in xt) + res2
  int B, n_q;
  int ntaps1, ntaps2;
  int off1[kMaxTaps], off2[kMaxTaps];
  float mid_slope, res_inv, out_scale, out_slope;
  // bit kk of kz1[j] / kz2[j]: rows 16 kk .. 16 kk + 15 of tap j's 64 x 64 weight block are all zero (rb_zero_ksteps), so
  // that K-step's MMA is not issued.  Time folding makes such blocks: a folded dilated tap reads one fold slot only.
  unsigned char kz1[kMaxTaps], kz2[kMaxTaps];
};
void rb_zero_ksteps(const float* w, int ntaps, unsigned char* kz);   // w: fp32 [ntaps][64][64]
bool rb_supported(const RbArgs& a, int* cg_out);          // needs ntaps / offsets only
size_t rb_packed_weight_bytes(int ntaps);
void rb_pack_weights(const float* w, int ntaps, int cg, ActType act, void* dst_host);   // w: fp32 [ntaps][64][64]
cudaError_t launch_resblock_tc(const RbArgs& a, ActType act, cudaStream_t s);
void rb_set_trace(long long* dev);     // debug builds (-DHFG_TC_TRACE): [7 roles][64 tiles][8 events] clock stamps

}  // namespace hfg
