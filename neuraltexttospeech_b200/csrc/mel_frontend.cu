// Mel front-end on the GPU: the step BEFORE the generator.
//
// Replaces `STFT.mel_spectrogram` of the reference (HiFiGAN_TF/common/audio_processing_tf.py:41-104; callers
// data.py:142,210 and gan.py:48-52,147):
//     tf.signal.stft(x, frame_length, frame_step, fft_length)   periodic Hann window, no centre padding, pad_end=False
//     -> tf.math.square(tf.abs(.))                              POWER spectrogram (reference quirk)
//     -> matmul with tf.signal.linear_to_mel_weight_matrix      HTK mel triangles, DC bin zeroed, no area normalisation
//     -> log(clip(., 1e-5))                                     dynamic_range_compression, audio_processing_tf.py:27-32
// Output layout = the generator's input layout, [B, T, n_mel] fp32 channels-last, so wav -> mel -> wav runs on device.
//
// One CTA = two consecutive frames of one utterance: the two real frames ride one complex FFT (z = a + i b;
// A[k] = (Z[k] + conj Z[N-k]) / 2, B[k] = (Z[k] - conj Z[N-k]) / 2i).  The FFT is a radix-4 Stockham autosort in
// shared memory (N/4 threads, one butterfly per thread per pass, twiddles from a host-computed fp64-accurate table);
// the power spectra stay in shared memory and each (frame, mel bin) is one thread's dot product over the triangle's
// non-zero span.  HBM traffic is the compulsory one: the waveform once (frame overlap is served by L2) and the mels.
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/hifigan_b200.h"

namespace {

struct MelKArgs {
  const float* wav;      // [B, n_samples]
  float* mel;            // [B, T, n_mel]
  const float2* tw;      // [N] (cos, -sin)(2 pi k / N)
  const float* window;   // [frame_length] periodic Hann
  const float* fb_w;     // [n_mel][fb_stride] weights over each triangle's span
  const int* fb_start;   // [n_mel] first spectrogram bin of the span
  const int* fb_len;     // [n_mel]
  int fb_stride;
  int n_samples, T, n_mel, frame_length, frame_step;
  float floor_;          // 1e-5
};

__device__ __forceinline__ float2 cmul(float2 a, float2 b) { return make_float2(a.x * b.x - a.y * b.y, a.x * b.y + a.y * b.x); }

// N = 4^LOG4N points, N/4 threads
template <int N>
__global__ void __launch_bounds__(N / 4) mel_frontend_kernel(const MelKArgs a) {
  extern __shared__ __align__(16) float2 msm[];
  float2* buf0 = msm;
  float2* buf1 = msm + N;
  const int tid = threadIdx.x;
  const int b = blockIdx.y, f0 = blockIdx.x * 2;
  const bool two = f0 + 1 < a.T;
  const float* xa = a.wav + (size_t)b * a.n_samples + (size_t)f0 * a.frame_step;
  // windowed frames, zero-padded to the FFT length (tf.signal.stft pads frames at the end)
  for (int i = tid; i < N; i += N / 4) {
    float2 z = make_float2(0.f, 0.f);
    if (i < a.frame_length) {
      const float w = a.window[i];
      z.x = xa[i] * w;
      if (two) z.y = xa[i + a.frame_step] * w;
    }
    buf0[i] = z;
  }
  __syncthreads();
  float2* in = buf0;
  float2* out = buf1;
#pragma unroll 1
  for (int ns = 1; ns < N; ns *= 4) {
    const int j = tid;
    const int k = j & (ns - 1);
    const int tstep = N / (4 * ns);                       // table stride: angle = -2 pi k / (4 ns)
    float2 u0 = in[j], u1 = in[j + N / 4], u2 = in[j + N / 2], u3 = in[j + 3 * N / 4];
    if (ns > 1) {
      u1 = cmul(u1, a.tw[k * tstep]);
      u2 = cmul(u2, a.tw[2 * k * tstep]);
      u3 = cmul(u3, a.tw[3 * k * tstep]);
    }
    const float2 s02 = make_float2(u0.x + u2.x, u0.y + u2.y), d02 = make_float2(u0.x - u2.x, u0.y - u2.y);
    const float2 s13 = make_float2(u1.x + u3.x, u1.y + u3.y), d13 = make_float2(u1.x - u3.x, u1.y - u3.y);
    const int o = ((j - k) << 2) + k;                     // (j / ns) * 4 ns + k
    out[o] = make_float2(s02.x + s13.x, s02.y + s13.y);
    out[o + ns] = make_float2(d02.x + d13.y, d02.y - d13.x);        // d02 - i d13
    out[o + 2 * ns] = make_float2(s02.x - s13.x, s02.y - s13.y);
    out[o + 3 * ns] = make_float2(d02.x - d13.y, d02.y + d13.x);    // d02 + i d13
    __syncthreads();
    float2* t = in; in = out; out = t;
  }
  // `in` holds Z; split into the two power spectra (bins 0..N/2), stored in `out` as (|A|^2, |B|^2)
  for (int k = tid; k <= N / 2; k += N / 4) {
    const float2 z = in[k], zc = in[(N - k) & (N - 1)];
    const float ar = 0.5f * (z.x + zc.x), ai = 0.5f * (z.y - zc.y);
    const float br = 0.5f * (z.y + zc.y), bi = 0.5f * (zc.x - z.x);
    out[k] = make_float2(ar * ar + ai * ai, br * br + bi * bi);
  }
  __syncthreads();
  for (int i = tid; i < 2 * a.n_mel; i += N / 4) {
    const int fr = i / a.n_mel, m = i - fr * a.n_mel;
    if (fr == 1 && !two) break;
    const int s = a.fb_start[m], len = a.fb_len[m];
    const float* w = a.fb_w + (size_t)m * a.fb_stride;
    float acc = 0.f;
    for (int q = 0; q < len; ++q) {
      const float2 p = out[s + q];
      acc = fmaf(fr ? p.y : p.x, w[q], acc);
    }
    a.mel[((size_t)b * a.T + f0 + fr) * a.n_mel + m] = logf(fmaxf(acc, a.floor_));
  }
}

thread_local std::string g_mel_create_error;

}  // namespace

struct hfg_mel_ctx {
  hfg_mel_config cfg{};
  int device = 0;
  int fb_stride = 0;
  float2* d_tw = nullptr;
  float* d_window = nullptr;
  float* d_fbw = nullptr;
  int* d_fbstart = nullptr;
  int* d_fblen = nullptr;
  cudaStream_t stream = nullptr;
  float* d_wav = nullptr; size_t d_wav_bytes = 0;
  float* d_mel = nullptr; size_t d_mel_bytes = 0;
  std::string err;
};

namespace {

int mfail(hfg_mel_ctx* h, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
  if (h) h->err = buf; else g_mel_create_error = buf;
  return code;
}

#define MEL_CUDA_TRY(h, expr)                                                                 \
  do { cudaError_t e__ = (expr);                                                              \
       if (e__ != cudaSuccess) return mfail(h, HFG_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e__)); } while (0)

double hz_to_mel(double f) { return 1127.0 * std::log1p(f / 700.0); }   // HTK scale (tf.signal.linear_to_mel_weight_matrix)

int frames_of(const hfg_mel_config& c, long long n) {
  return n < c.frame_length ? 0 : (int)(1 + (n - c.frame_length) / c.frame_step);
}

}  // namespace

extern "C" {

int hfg_mel_create(const hfg_mel_config* cfg, int device, hfg_mel_handle* out) {
  if (!cfg || !out) return mfail(nullptr, HFG_ERR_INVALID, "hfg_mel_create: null argument");
  *out = nullptr;
  const hfg_mel_config& c = *cfg;
  if (c.fft_length != 256 && c.fft_length != 1024 && c.fft_length != 4096)
    return mfail(nullptr, HFG_ERR_UNSUPPORTED, "fft_length must be 256, 1024 or 4096 (radix-4 kernel), got %d", c.fft_length);
  if (c.frame_length < 1 || c.frame_length > c.fft_length) return mfail(nullptr, HFG_ERR_INVALID, "frame_length must be in [1, fft_length]");
  if (c.frame_step < 1) return mfail(nullptr, HFG_ERR_INVALID, "frame_step must be >= 1");
  if (c.n_mel_channels < 1 || c.sampling_rate < 1 || c.mel_fmin < 0 || c.mel_fmax <= c.mel_fmin || c.mel_fmax > c.sampling_rate / 2.0f)
    return mfail(nullptr, HFG_ERR_INVALID, "bad mel filterbank parameters (need 0 <= fmin < fmax <= sampling_rate/2)");
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return mfail(nullptr, HFG_ERR_CUDA, "no CUDA device available (%s); this library has no CPU fallback", cudaGetErrorString(e));
  if (device < 0 || device >= ndev) return mfail(nullptr, HFG_ERR_INVALID, "device %d out of range [0,%d)", device, ndev);
  hfg_mel_ctx* h = new hfg_mel_ctx();
  h->cfg = c; h->device = device;
  const int N = c.fft_length, nbins = N / 2 + 1, M = c.n_mel_channels;
  // twiddles and the periodic Hann window, in double on the host
  std::vector<float2> tw(N);
  for (int k = 0; k < N; ++k) { const double ang = -2.0 * M_PI * k / N; tw[k] = make_float2((float)std::cos(ang), (float)std::sin(ang)); }
  std::vector<float> win(c.frame_length);
  for (int i = 0; i < c.frame_length; ++i) win[i] = (float)(0.5 - 0.5 * std::cos(2.0 * M_PI * i / c.frame_length));
  // tf.signal.linear_to_mel_weight_matrix restated (see oracle/adjacent_oracle.py): dense [nbins][M] in double, then
  // each triangle's non-zero span
  std::vector<double> edges(M + 2);
  const double mlo = hz_to_mel(c.mel_fmin), mhi = hz_to_mel(c.mel_fmax);
  for (int i = 0; i < M + 2; ++i) edges[i] = mlo + (mhi - mlo) * i / (M + 1);
  std::vector<std::vector<double>> col(M, std::vector<double>(nbins, 0.0));
  std::vector<int> start(M, 0), len(M, 0);
  int stride = 1;
  for (int m = 0; m < M; ++m) {
    int first = -1, last = -1;
    for (int k = 1; k < nbins; ++k) {                     // bin 0 (DC) is zeroed
      const double sm = hz_to_mel((c.sampling_rate / 2.0) * k / (nbins - 1));
      const double lo = (sm - edges[m]) / (edges[m + 1] - edges[m]), up = (edges[m + 2] - sm) / (edges[m + 2] - edges[m + 1]);
      const double w = std::max(0.0, std::min(lo, up));
      col[m][k] = w;
      if (w > 0.0) { if (first < 0) first = k; last = k; }
    }
    if (first >= 0) { start[m] = first; len[m] = last - first + 1; }
    stride = std::max(stride, len[m]);
  }
  h->fb_stride = stride;
  std::vector<float> fbw((size_t)M * stride, 0.f);
  for (int m = 0; m < M; ++m)
    for (int q = 0; q < len[m]; ++q) fbw[(size_t)m * stride + q] = (float)col[m][start[m] + q];
  auto up = [&](void** d, const void* src, size_t bytes) -> cudaError_t {
    cudaError_t e2 = cudaMalloc(d, bytes);
    return e2 != cudaSuccess ? e2 : cudaMemcpy(*d, src, bytes, cudaMemcpyHostToDevice);
  };
  e = cudaSetDevice(device);
  if (e == cudaSuccess) e = up((void**)&h->d_tw, tw.data(), tw.size() * sizeof(float2));
  if (e == cudaSuccess) e = up((void**)&h->d_window, win.data(), win.size() * sizeof(float));
  if (e == cudaSuccess) e = up((void**)&h->d_fbw, fbw.data(), fbw.size() * sizeof(float));
  if (e == cudaSuccess) e = up((void**)&h->d_fbstart, start.data(), start.size() * sizeof(int));
  if (e == cudaSuccess) e = up((void**)&h->d_fblen, len.data(), len.size() * sizeof(int));
  if (e != cudaSuccess) { hfg_mel_destroy(h); return mfail(nullptr, HFG_ERR_CUDA, "mel front-end setup failed: %s", cudaGetErrorString(e)); }
  *out = h;
  return HFG_OK;
}

int hfg_mel_num_frames(hfg_mel_handle h, long long n_samples, int* out) {
  if (!h || !out) return HFG_ERR_INVALID;
  if (n_samples < 0) return mfail(h, HFG_ERR_INVALID, "negative sample count");
  *out = frames_of(h->cfg, n_samples);
  return HFG_OK;
}

int hfg_mel_forward(hfg_mel_handle h, const float* wav_dev, int B, long long n_samples, float* mel_dev, void* stream) {
  if (!h) return HFG_ERR_INVALID;
  if (B < 0 || n_samples < 0) return mfail(h, HFG_ERR_INVALID, "negative batch or sample count");
  const int T = frames_of(h->cfg, n_samples);
  if (B == 0 || T == 0) return HFG_OK;
  if (!wav_dev || !mel_dev) return mfail(h, HFG_ERR_INVALID, "hfg_mel_forward: null device pointer");
  if (n_samples > 0x7fffffffLL || B > 65535) return mfail(h, HFG_ERR_INVALID, "utterance too long or batch above 65535");
  MEL_CUDA_TRY(h, cudaSetDevice(h->device));
  MelKArgs a{};
  a.wav = wav_dev; a.mel = mel_dev; a.tw = h->d_tw; a.window = h->d_window; a.fb_w = h->d_fbw; a.fb_start = h->d_fbstart;
  a.fb_len = h->d_fblen; a.fb_stride = h->fb_stride; a.n_samples = (int)n_samples; a.T = T; a.n_mel = h->cfg.n_mel_channels;
  a.frame_length = h->cfg.frame_length; a.frame_step = h->cfg.frame_step; a.floor_ = 1e-5f;
  const dim3 grid((unsigned)((T + 1) / 2), (unsigned)B);
  cudaStream_t s = (cudaStream_t)stream;
  const int N = h->cfg.fft_length;
  const size_t smem = (size_t)2 * N * sizeof(float2);
  cudaError_t e = cudaSuccess;
  if (N == 256) mel_frontend_kernel<256><<<grid, 64, smem, s>>>(a);
  else if (N == 1024) mel_frontend_kernel<1024><<<grid, 256, smem, s>>>(a);
  else {
    e = cudaFuncSetAttribute(mel_frontend_kernel<4096>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e == cudaSuccess) mel_frontend_kernel<4096><<<grid, 1024, smem, s>>>(a);
  }
  if (e == cudaSuccess) e = cudaGetLastError();
  if (e != cudaSuccess) return mfail(h, HFG_ERR_CUDA, "mel_frontend launch failed: %s", cudaGetErrorString(e));
  return HFG_OK;
}

int hfg_mel_forward_host(hfg_mel_handle h, const float* wav_host, int B, long long n_samples, float* mel_host) {
  if (!h) return HFG_ERR_INVALID;
  if (B < 0 || n_samples < 0) return mfail(h, HFG_ERR_INVALID, "negative batch or sample count");
  const int T = frames_of(h->cfg, n_samples);
  if (B == 0 || T == 0) return HFG_OK;
  if (!wav_host || !mel_host) return mfail(h, HFG_ERR_INVALID, "hfg_mel_forward_host: null host pointer");
  MEL_CUDA_TRY(h, cudaSetDevice(h->device));
  if (!h->stream) MEL_CUDA_TRY(h, cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  const size_t wb = (size_t)B * n_samples * 4, mb = (size_t)B * T * h->cfg.n_mel_channels * 4;
  if (wb > h->d_wav_bytes) {
    if (h->d_wav) cudaFree(h->d_wav);
    h->d_wav = nullptr; h->d_wav_bytes = 0;
    MEL_CUDA_TRY(h, cudaMalloc((void**)&h->d_wav, wb));
    h->d_wav_bytes = wb;
  }
  if (mb > h->d_mel_bytes) {
    if (h->d_mel) cudaFree(h->d_mel);
    h->d_mel = nullptr; h->d_mel_bytes = 0;
    MEL_CUDA_TRY(h, cudaMalloc((void**)&h->d_mel, mb));
    h->d_mel_bytes = mb;
  }
  MEL_CUDA_TRY(h, cudaMemcpyAsync(h->d_wav, wav_host, wb, cudaMemcpyHostToDevice, h->stream));
  int rc = hfg_mel_forward(h, h->d_wav, B, n_samples, h->d_mel, h->stream);
  if (rc) return rc;
  MEL_CUDA_TRY(h, cudaMemcpyAsync(mel_host, h->d_mel, mb, cudaMemcpyDeviceToHost, h->stream));
  MEL_CUDA_TRY(h, cudaStreamSynchronize(h->stream));
  return HFG_OK;
}

const char* hfg_mel_last_error(hfg_mel_handle h) { return h ? h->err.c_str() : g_mel_create_error.c_str(); }

void hfg_mel_destroy(hfg_mel_handle h) {
  if (!h) return;
  cudaSetDevice(h->device);
  if (h->d_tw) cudaFree(h->d_tw);
  if (h->d_window) cudaFree(h->d_window);
  if (h->d_fbw) cudaFree(h->d_fbw);
  if (h->d_fbstart) cudaFree(h->d_fbstart);
  if (h->d_fblen) cudaFree(h->d_fblen);
  if (h->d_wav) cudaFree(h->d_wav);
  if (h->d_mel) cudaFree(h->d_mel);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

}  // extern "C"
