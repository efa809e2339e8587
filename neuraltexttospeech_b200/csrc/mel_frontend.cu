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
// Two consecutive real frames of one utterance ride one complex FFT (z = a + i b;
// A[k] = (Z[k] + conj Z[N-k]) / 2, B[k] = (Z[k] - conj Z[N-k]) / 2i).  HBM traffic is the compulsory one: the waveform
// once (frame overlap is served by L1/L2) and the mels.  Two kernels:
//   * fft_length 1024 (the reference's configuration): mel_frontend_r16_kernel.  64 threads per FFT, 16 points per
//     thread in registers: Stockham passes of radix 16, 16, 4 (two shared-memory exchanges instead of five, padded
//     rows so every exchange is conflict-free, 64-thread named barriers), first pass straight from global memory
//     with the second frame's samples taken from the first frame's registers (hop 256 = 4 x 64).  Four FFTs (eight
//     frames) per CTA.  The filterbank is a host-built list of (mel bin, span chunk) items of bounded length so that
//     the 64 threads get equal work; partial sums are combined in a fixed order (deterministic).
//   * fft_length 256 / 4096: mel_frontend_kernel, a radix-4 Stockham autosort in shared memory (N/4 threads, one
//     butterfly per thread per pass), one thread per (frame, mel bin) dot product.
// Twiddles come from a host-computed fp64-accurate table in both.
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
  const float2* tw16;    // r16 kernel: [16][16] w^(4 r k) (pass 2) then [3][256] w^(r j), r = 1..3 (pass 3): lane-contiguous
  const float* fb_wt;    // r16 kernel: item weights transposed, [chunk][item_stride]: lanes (items) read consecutive floats
  int item_stride;
  const int4* items;     // r16 kernel: (mel bin, first spectrogram bin, length, -) chunks of the triangles
  const int2* item_of;   // r16 kernel: [n_mel] (first item, item count) of each mel bin
  int n_items;
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

// ---------------------------------------------------------------- fft_length 1024: radix 16 x 16 x 4 in registers
__device__ __forceinline__ float2 cadd(float2 a, float2 b) { return make_float2(a.x + b.x, a.y + b.y); }
__device__ __forceinline__ float2 csub(float2 a, float2 b) { return make_float2(a.x - b.x, a.y - b.y); }

// forward 4-point DFT in place: (a, b, c, d) -> (Y0, Y1, Y2, Y3)
__device__ __forceinline__ void dft4(float2& a, float2& b, float2& c, float2& d) {
  const float2 s02 = cadd(a, c), d02 = csub(a, c), s13 = cadd(b, d), d13 = csub(b, d);
  a = cadd(s02, s13);
  b = make_float2(d02.x + d13.y, d02.y - d13.x);       // d02 - i d13
  c = csub(s02, s13);
  d = make_float2(d02.x - d13.y, d02.y + d13.x);       // d02 + i d13
}

// forward 16-point DFT in place; on return u[4 c + d] holds Y[c + 4 d]
__device__ __forceinline__ void dft16(float2 (&u)[16]) {
  constexpr float C1 = 0.92387953251128674f, S1 = 0.38268343236508977f, H = 0.70710678118654752f;
#pragma unroll
  for (int b = 0; b < 4; ++b) dft4(u[b], u[4 + b], u[8 + b], u[12 + b]);      // u[4 c + b] = t[b][c]
  // t[b][c] *= W16^(b c)
  u[4 * 1 + 1] = cmul(u[4 * 1 + 1], make_float2(C1, -S1));                      // e = 1
  u[4 * 1 + 2] = cmul(u[4 * 1 + 2], make_float2(H, -H));                        // e = 2
  u[4 * 1 + 3] = cmul(u[4 * 1 + 3], make_float2(S1, -C1));                      // e = 3
  u[4 * 2 + 1] = cmul(u[4 * 2 + 1], make_float2(H, -H));                        // e = 2
  u[4 * 2 + 2] = make_float2(u[4 * 2 + 2].y, -u[4 * 2 + 2].x);                  // e = 4: * -i
  u[4 * 2 + 3] = cmul(u[4 * 2 + 3], make_float2(-H, -H));                       // e = 6
  u[4 * 3 + 1] = cmul(u[4 * 3 + 1], make_float2(S1, -C1));                      // e = 3
  u[4 * 3 + 2] = cmul(u[4 * 3 + 2], make_float2(-H, -H));                       // e = 6
  u[4 * 3 + 3] = cmul(u[4 * 3 + 3], make_float2(-C1, S1));                      // e = 9
#pragma unroll
  for (int c = 0; c < 4; ++c) dft4(u[4 * c], u[4 * c + 1], u[4 * c + 2], u[4 * c + 3]);
}

constexpr int kR16Fft = 4;                 // FFTs (frame pairs) per CTA
constexpr int kR16Row = 1024 + 128;        // float2 per FFT buffer: 16-element rows padded by 2 (see phys)
constexpr int kR16Partial = 640;           // partial sums of the filterbank items live above the 513 power bins
constexpr int kR16MaxItems = 448;

__device__ __forceinline__ int phys(int i) { return i + 2 * (i >> 4); }
__device__ __forceinline__ void group_sync(int g) {        // immediate ids: a register operand would reserve all 16 barriers
  switch (g) {
    case 0: asm volatile("bar.sync 1, 64;" ::: "memory"); break;
    case 1: asm volatile("bar.sync 2, 64;" ::: "memory"); break;
    case 2: asm volatile("bar.sync 3, 64;" ::: "memory"); break;
    default: asm volatile("bar.sync 4, 64;" ::: "memory"); break;
  }
}

// HOP4: frame_length == 1024 and frame_step == 256, so the second frame's samples are the first frame's, 4 registers on
template <bool HOP4>
__global__ void __launch_bounds__(64 * kR16Fft, 5) mel_frontend_r16_kernel(const MelKArgs a) {
  constexpr int N = 1024;
  __shared__ __align__(16) float2 r16sm[kR16Fft * kR16Row];
  const int g = threadIdx.x >> 6, j = threadIdx.x & 63;
  float2* sm = r16sm + g * kR16Row;
  const int b = blockIdx.y, f0 = (blockIdx.x * kR16Fft + g) * 2;
  if (f0 >= a.T) return;                                   // whole 64-thread group: its barrier is never used
  const bool two = f0 + 1 < a.T;
  const float* xa = a.wav + (size_t)b * a.n_samples + (size_t)f0 * a.frame_step;
  float2 u[16];
  // ---- pass 1 (ns = 1): u_r = z[j + 64 r], no twiddles; out[16 j + q]
  if (HOP4) {
    float x[20];
#pragma unroll
    for (int r = 0; r < 16; ++r) x[r] = __ldg(xa + j + 64 * r);
#pragma unroll
    for (int r = 16; r < 20; ++r) x[r] = two ? __ldg(xa + j + 64 * r) : 0.f;
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const float w = __ldg(a.window + j + 64 * r);
      u[r] = make_float2(x[r] * w, two ? x[r + 4] * w : 0.f);
    }
  } else {
#pragma unroll
    for (int r = 0; r < 16; ++r) {
      const int i = j + 64 * r;
      float2 z = make_float2(0.f, 0.f);
      if (i < a.frame_length) {
        const float w = __ldg(a.window + i);
        z.x = __ldg(xa + i) * w;
        if (two) z.y = __ldg(xa + i + a.frame_step) * w;
      }
      u[r] = z;
    }
  }
  dft16(u);
  {
    float4* row = reinterpret_cast<float4*>(sm + 18 * j);  // phys(16 j) = 18 j: 16-byte aligned, lanes 144 B apart
#pragma unroll
    for (int p = 0; p < 8; ++p) {                          // Y[2p], Y[2p+1]; Y[c + 4 d] sits in u[4 c + d]
      const float2 y0 = u[4 * ((2 * p) & 3) + ((2 * p) >> 2)], y1 = u[4 * ((2 * p + 1) & 3) + ((2 * p + 1) >> 2)];
      row[p] = make_float4(y0.x, y0.y, y1.x, y1.y);
    }
  }
  group_sync(g);
  // ---- pass 2 (ns = 16): u_r = in[j + 64 r] * w^(4 r k), k = j mod 16; out[(j / 16) 256 + k + 16 q]
  const int k2 = j & 15;
  {
    const int base = j + 2 * (j >> 4);                     // phys(j + 64 r) = base + 72 r
    u[0] = sm[base];
#pragma unroll
    for (int r = 1; r < 16; ++r) u[r] = cmul(sm[base + 72 * r], __ldg(a.tw16 + 16 * r + k2));
  }
  dft16(u);
  group_sync(g);                                           // everyone has read: the exchange is in place
  {
    float2* o = sm + (j >> 4) * 288 + k2;                  // phys((j / 16) 256 + k + 16 q) = (j / 16) 288 + k + 18 q
#pragma unroll
    for (int q = 0; q < 16; ++q) o[18 * q] = u[4 * (q & 3) + (q >> 2)];
  }
  group_sync(g);
  // ---- pass 3 (ns = 256, radix 4): butterfly jj = j + 64 i reads and writes the same four slots jj + 256 r
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int jj = j + 64 * i;
    float2* p = sm + jj + 2 * (jj >> 4);                   // phys(jj + 256 r) = phys(jj) + 288 r
    float2 v0 = p[0];
    float2 v1 = cmul(p[288], __ldg(a.tw16 + 256 + jj));
    float2 v2 = cmul(p[576], __ldg(a.tw16 + 512 + jj));
    float2 v3 = cmul(p[864], __ldg(a.tw16 + 768 + jj));
    dft4(v0, v1, v2, v3);
    p[0] = v0; p[288] = v1; p[576] = v2; p[864] = v3;
  }
  group_sync(g);
  // ---- the two power spectra, in place: slot k <- (|A[k]|^2, |B[k]|^2); slot k and N - k are touched by this k only
  for (int k = j; k <= N / 2; k += 64) {
    const float2 z = sm[phys(k)], zc = sm[phys((N - k) & (N - 1))];
    const float ar = 0.5f * (z.x + zc.x), ai = 0.5f * (z.y - zc.y);
    const float br = 0.5f * (z.y + zc.y), bi = 0.5f * (zc.x - z.x);
    sm[phys(k)] = make_float2(ar * ar + ai * ai, br * br + bi * bi);
  }
  group_sync(g);
  // ---- filterbank: bounded-length chunks of the triangles, then a fixed-order sum per mel bin
  for (int it = j; it < a.n_items; it += 64) {
    const int4 d = __ldg(a.items + it);                    // (mel bin, first bin, length, -)
    const float* w = a.fb_wt + it;
    float pa = 0.f, pb = 0.f;
    for (int q = 0; q < d.z; ++q) {
      const float2 p = sm[phys(d.y + q)];
      const float wq = __ldg(w + q * a.item_stride);
      pa = fmaf(p.x, wq, pa);
      pb = fmaf(p.y, wq, pb);
    }
    sm[kR16Partial + it] = make_float2(pa, pb);
  }
  group_sync(g);
  for (int m = j; m < a.n_mel; m += 64) {
    const int2 io = __ldg(a.item_of + m);
    float pa = 0.f, pb = 0.f;
    for (int q = 0; q < io.y; ++q) { const float2 p = sm[kR16Partial + io.x + q]; pa += p.x; pb += p.y; }
    float* dst = a.mel + ((size_t)b * a.T + f0) * a.n_mel + m;
    dst[0] = logf(fmaxf(pa, a.floor_));
    if (two) dst[a.n_mel] = logf(fmaxf(pb, a.floor_));
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
  int4* d_items = nullptr;
  float2* d_tw16 = nullptr;
  float* d_fbwt = nullptr;
  int item_stride = 0;
  int2* d_item_of = nullptr;
  int n_items = 0;
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
  // The reference sizes its filterbank with frame_length // 2 + 1 spectrogram bins while tf.signal.stft returns
  // fft_length // 2 + 1 (HiFiGAN_TF/common/audio_processing_tf.py:53-59, 85-97): with frame_length != filter_length its matmul
  // has mismatched inner dimensions and raises.  The drop-in refuses the same constructor arguments instead of inventing a result.
  if (c.frame_length != c.fft_length)
    return mfail(nullptr, HFG_ERR_INVALID, "frame_length (%d) != filter_length (%d): the reference's mel filterbank has frame_length//2+1 bins "
                 "and its matmul with the fft_length//2+1-bin spectrogram fails (audio_processing_tf.py:55,85-97)", c.frame_length, c.fft_length);
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
  if (N == 1024) {
    // work items of the r16 kernel: the smallest chunk length that fits the triangles into two rounds of 64 threads
    std::vector<int4> items;
    std::vector<int2> item_of(M);
    for (int chunk = 4;; ++chunk) {
      items.clear();
      for (int m = 0; m < M; ++m) {
        item_of[m] = make_int2((int)items.size(), 0);
        for (int q0 = 0; q0 < len[m]; q0 += chunk) {
          items.push_back(make_int4(m, start[m] + q0, std::min(chunk, len[m] - q0), m * stride + q0));
          ++item_of[m].y;
        }
      }
      if ((int)items.size() <= 128 || chunk >= nbins) break;
    }
    if ((int)items.size() > kR16MaxItems) { hfg_mel_destroy(h); return mfail(nullptr, HFG_ERR_UNSUPPORTED, "too many mel channels for the fft_length 1024 kernel"); }
    h->n_items = (int)items.size();
    if (items.empty()) items.push_back(make_int4(0, 0, 0, 0));
    int chunk_len = 1;
    for (const int4& it : items) chunk_len = std::max(chunk_len, it.z);
    h->item_stride = ((int)items.size() + 31) / 32 * 32;
    std::vector<float> wt((size_t)chunk_len * h->item_stride, 0.f);
    for (size_t i = 0; i < items.size(); ++i)
      for (int q = 0; q < items[i].z; ++q) wt[(size_t)q * h->item_stride + i] = fbw[(size_t)items[i].w + q];
    std::vector<float2> tw16(1024);
    for (int r = 0; r < 16; ++r)
      for (int k = 0; k < 16; ++k) tw16[16 * r + k] = tw[(4 * r * k) % N];
    for (int r = 1; r < 4; ++r)
      for (int jj = 0; jj < 256; ++jj) tw16[256 * r + jj] = tw[(r * jj) % N];
    if (e == cudaSuccess) e = up((void**)&h->d_tw16, tw16.data(), tw16.size() * sizeof(float2));
    if (e == cudaSuccess) e = up((void**)&h->d_fbwt, wt.data(), wt.size() * sizeof(float));
    if (e == cudaSuccess) e = up((void**)&h->d_items, items.data(), items.size() * sizeof(int4));
    if (e == cudaSuccess) e = up((void**)&h->d_item_of, item_of.data(), item_of.size() * sizeof(int2));
  }
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
  else if (N == 1024) {
    a.items = h->d_items; a.item_of = h->d_item_of; a.n_items = h->n_items;
    a.tw16 = h->d_tw16; a.fb_wt = h->d_fbwt; a.item_stride = h->item_stride;
    const dim3 g16((unsigned)((T + 2 * kR16Fft - 1) / (2 * kR16Fft)), (unsigned)B);
    if (a.frame_length == 1024 && a.frame_step == 256) mel_frontend_r16_kernel<true><<<g16, 64 * kR16Fft, 0, s>>>(a);
    else mel_frontend_r16_kernel<false><<<g16, 64 * kR16Fft, 0, s>>>(a);
  }
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
  if (h->d_items) cudaFree(h->d_items);
  if (h->d_tw16) cudaFree(h->d_tw16);
  if (h->d_fbwt) cudaFree(h->d_fbwt);
  if (h->d_item_of) cudaFree(h->d_item_of);
  if (h->d_wav) cudaFree(h->d_wav);
  if (h->d_mel) cudaFree(h->d_mel);
  if (h->stream) cudaStreamDestroy(h->stream);
  delete h;
}

}  // extern "C"
