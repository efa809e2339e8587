// HiFi-GAN discriminator forward passes: MultiPeriodDiscriminator and MultiScaleDiscriminator of the
// reference (HiFiGAN_TF/model.py:176-243 DiscriminatorP, 246-274 MPD, 277-346 DiscriminatorS, 349-389 MSD) behind the
// hfg_disc_* C ABI.  fp32 storage; convs on the tensor cores at fp32 accuracy (3xTF32, below) with an fp32 FFMA kernel for the
// first / last conv of each sub-discriminator; hfg_disc_pair_losses reduces the reference's GAN loss terms (losses.py:8-43).
//
// Data layout.  Every sub-discriminator works on independent SEQUENCES of channels-last rows [n_seq, len, C]:
//   DiscriminatorS   n_seq = B, the (pooled) waveform;
//   DiscriminatorP   n_seq = B * period: the reference reshapes [B, T'] to [B, T'/period, period, 1] and convolves with
//                    (5, 1) kernels, i.e. column w of the period axis never mixes with another column -- sequence
//                    (b, w) holds samples w, w + period, w + 2 period, ... of utterance b.
// Every feature map is written ONCE, straight into the caller's output buffer, as [B][period][len][C] (period = 1 for
// DiscriminatorS); it is the next layer's input as well (maps are stored after the LeakyReLU, like the reference's fmap
// list).  The reference's DiscriminatorP maps are [B][len][period][C]: the host shim returns a transposed view.
//
// Kernels: disc_split_kernel (reflect pad + period split), disc_pool_kernel (AveragePooling1D(4, 2, "same")),
// disc_conv_kernel (strided, grouped Conv1D with TensorFlow SAME / VALID padding, bias and LeakyReLU fused),
// disc_post_kernel (C_out = 1).
// Every conv with at least 8 input channels per group and more than one output channel (all but the first conv and
// conv_post of each sub-discriminator: > 99 % of the FLOPs) runs on the tensor cores at fp32 accuracy ("3xTF32"): tcgen05.mma.kind::tf32 reads only the top 10 mantissa bits of an fp32 operand, so with
// x = x_hi + x_lo (x_hi = what the tensor core sees, x_lo = the exact remainder) and w = w_hi + w_lo likewise,
//     x . w  ~=  x_hi . w_hi + x_hi . w_lo + x_lo . w_hi          (dropped: x_lo . w_lo, ~2^-22 relative)
// is three TF32 products accumulated in fp32 in TMEM: conv_tf32.cu's kernel run twice (x_lo with w_hi taps into a scratch
// map, then x with [w_hi | w_lo] as 2k taps + that scratch + bias + LeakyReLU).  Its extended form takes the stride as
// `stride` undilated convs over the phase sub-sequences of the input (one TMA map per phase) and a grouped conv as N tiles
// that read only their group's K range.  HFG_DISC_TC=0 keeps every layer on the FFMA kernel above.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include <cuda_runtime.h>

#include "../../include/hifigan_b200.h"
#include "kernels.cuh"

namespace {

constexpr float kSlope = 0.1f;            // LRELU_SLOPE, HiFiGAN_TF/model.py:14
constexpr int kPeriods[5] = {2, 3, 5, 7, 11};   // model.py:250-256

struct ConvSpec { int cin, cout, k, stride, groups; bool valid; };
// DiscriminatorP: model.py:186-216; DiscriminatorS: model.py:284-332 (the first conv is padding="valid")
const ConvSpec kPConvs[6] = {{1, 32, 5, 3, 1, false}, {32, 128, 5, 3, 1, false}, {128, 512, 5, 3, 1, false},
                             {512, 1024, 5, 3, 1, false}, {1024, 1024, 5, 1, 1, false}, {1024, 1, 3, 1, 1, false}};
const ConvSpec kSConvs[8] = {{1, 128, 15, 1, 1, true}, {128, 128, 41, 2, 4, false}, {128, 256, 41, 2, 16, false},
                             {256, 512, 41, 4, 16, false}, {512, 1024, 41, 4, 16, false}, {1024, 1024, 41, 1, 16, false},
                             {1024, 1024, 5, 1, 1, false}, {1024, 1, 3, 1, 1, false}};

// ------------------------------------------------------------------ kernels
// DiscriminatorP input (model.py:226-235): reflect-pad the END of the waveform to a multiple of the period, then
// sequence (b, w) = samples w, w + period, ...   out [B * period, H]
__global__ void disc_split_kernel(const float* __restrict__ audio, float* __restrict__ out, int B, long long T, int period, long long H) {
  const long long total = (long long)B * period * H;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long h = i % H;
    const long long bw = i / H;
    const int w = (int)(bw % period);
    const long long b = bw / period;
    long long t = h * period + w;
    if (t >= T) t = 2 * (T - 1) - t;                       // REFLECT: the edge sample is not repeated
    out[i] = audio[b * T + t];
  }
}

// AveragePooling1D(4, 2, "same") (model.py:357-364): window [2o - left, 2o - left + 4); TensorFlow divides by the number
// of in-range samples, the upstream torch AvgPool1d(4, 2, padding=1) by 4.
__global__ void disc_pool_kernel(const float* __restrict__ x, float* __restrict__ out, int B, long long n, long long n_out, int left, int count_pad) {
  const long long total = (long long)B * n_out;
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < total; i += (long long)gridDim.x * blockDim.x) {
    const long long o = i % n_out, b = i / n_out;
    const long long lo = 2 * o - left;
    float s = 0.f;
    int cnt = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const long long t = lo + j;
      if (t >= 0 && t < n) { s += x[b * n + t]; ++cnt; }
    }
    out[i] = s / (float)(count_pad ? 4 : cnt);
  }
}

// x_lo = x - (x with the low 13 mantissa bits cleared): the part of an fp32 operand the TF32 tensor core does not see
__global__ void disc_lo_kernel(const float4* __restrict__ x, float4* __restrict__ lo, long long n4) {
  for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 v = x[i];
    float4 o;
    o.x = v.x - __uint_as_float(__float_as_uint(v.x) & 0xFFFFE000u); o.y = v.y - __uint_as_float(__float_as_uint(v.y) & 0xFFFFE000u);
    o.z = v.z - __uint_as_float(__float_as_uint(v.z) & 0xFFFFE000u); o.w = v.w - __uint_as_float(__float_as_uint(v.w) & 0xFFFFE000u);
    lo[i] = o;
  }
}

// GAN loss terms of one (real, generated) pair forward (reference HiFiGAN_TF/losses.py:8-43): the maps of a forward over the
// concatenated batch [y; y_hat] hold the real half first.  Block (x, m) sums its strided share of map m:
//   [0] |r - g| (feature_loss, losses.py:8-17)   [1] (1 - r)^2, [2] g^2 (discriminator_loss, :20-33)   [3] (1 - g)^2 (generator_loss, :36-46)
struct LossSeg { long long off; long long half; };   // floats: start of the map, size of one half
constexpr int kLossBlocks = 64;
__global__ void __launch_bounds__(256) disc_loss_kernel(const float* __restrict__ maps, const LossSeg* __restrict__ segs, float* __restrict__ partial) {
  const LossSeg sg = segs[blockIdx.y];
  const float* __restrict__ r = maps + sg.off;
  const float* __restrict__ g = r + sg.half;
  float a0 = 0.f, a1 = 0.f, a2 = 0.f, a3 = 0.f;
  for (long long i = blockIdx.x * 256ll + threadIdx.x; i < sg.half; i += 256ll * kLossBlocks) {
    const float rv = r[i], gv = g[i];
    a0 += fabsf(rv - gv); a1 += (1.f - rv) * (1.f - rv); a2 += gv * gv; a3 += (1.f - gv) * (1.f - gv);
  }
  __shared__ float sh[4][8];
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a0 += __shfl_xor_sync(0xffffffffu, a0, o); a1 += __shfl_xor_sync(0xffffffffu, a1, o);
    a2 += __shfl_xor_sync(0xffffffffu, a2, o); a3 += __shfl_xor_sync(0xffffffffu, a3, o);
  }
  const int w = threadIdx.x >> 5;
  if ((threadIdx.x & 31) == 0) { sh[0][w] = a0; sh[1][w] = a1; sh[2][w] = a2; sh[3][w] = a3; }
  __syncthreads();
  if (threadIdx.x < 4) {
    float t = 0.f;
    for (int k = 0; k < 8; ++k) t += sh[threadIdx.x][k];        // fixed order: the result does not depend on scheduling
    partial[((size_t)blockIdx.y * kLossBlocks + blockIdx.x) * 4 + threadIdx.x] = t;
  }
}
// per map: the kLossBlocks partial sums in a fixed order, in double -> means [n_maps][4]
__global__ void disc_loss_final_kernel(const float* __restrict__ partial, const LossSeg* __restrict__ segs, float* __restrict__ out, int n_maps) {
  const int m = blockIdx.x * blockDim.x + threadIdx.x;
  if (m >= n_maps * 4) return;
  const int map = m >> 2, k = m & 3;
  double t = 0.0;
  for (int b = 0; b < kLossBlocks; ++b) t += (double)partial[((size_t)map * kLossBlocks + b) * 4 + k];
  out[m] = (float)(t / (double)segs[map].half);
}

struct DiscConvArgs {
  const float* x;       // [n_seq, n_in, cin]
  const float* w;       // [k][cin_g][cout]
  const float* bias;    // [cout]
  float* out;           // [n_seq, n_out, cout]
  int n_seq;
  long long n_in, n_out;
  int cin, cout, cin_g, cout_g, k, stride, pad_left;
  int tqs, spt;         // rows per sequence and sequences per block tile (set by launch_conv)
  float out_slope;
};

__device__ __forceinline__ float lrelu(float v, float slope) { return v >= 0.f ? v : v * slope; }

constexpr int kCiBlk = 32;                // input channels of a group staged per pass

template <int CO_TILE, int TT> struct DiscCfg {
  static constexpr int CGT = CO_TILE / 4;                      // threads along c_out (4 channels each)
  static constexpr int THREADS = 256;
  static constexpr int TG = THREADS / CGT;                     // threads along the output rows
  static constexpr int TQ = TG * TT;                           // output rows per block
};

// out[n, q, co] = lrelu(bias[co] + sum_j sum_ci x[n, q * stride + j - pad_left, g(co) * cin_g + ci] * w[j][ci][co]); rows outside
// [0, n_in) are zero (TensorFlow pads with zeros AFTER the previous layer's LeakyReLU, and lrelu(0) = 0).
// A block = TQ output rows x CO_TILE channels of one group; a thread = TT rows x 4 channels.  The deep DiscriminatorP
// layers have hundreds of SHORT sequences (15-50 rows), so a block's TQ rows are `spt` sequences x `tqs` rows each
// (tqs * spt <= TQ, chosen on the host to waste the fewest rows): tile row r -> sequence z * spt + r / tqs, row
// x * tqs + r % tqs; each sequence of the tile has its own (tqs - 1) * stride + k slab rows.
template <int CO_TILE, int TT, bool CI4, bool PACK>
__global__ void __launch_bounds__(256) disc_conv_kernel(const DiscConvArgs a) {
  using Cfg = DiscCfg<CO_TILE, TT>;
  extern __shared__ __align__(16) float slab[];
  const int tid = threadIdx.x;
  const int cgt = tid % Cfg::CGT, tg = tid / Cfg::CGT;
  const long long q0 = (long long)blockIdx.x * a.tqs;
  const int co = blockIdx.y * CO_TILE + cgt * 4;
  const int grp = (blockIdx.y * CO_TILE) / a.cout_g;
  const int seq0 = blockIdx.z * a.spt;
  const int cblk = a.cin_g < kCiBlk ? a.cin_g : kCiBlk;
  const int pitch = CI4 ? cblk + 4 : cblk;                     // floats; +4 keeps float4 alignment and shifts the banks
  const int rows_s = (a.tqs - 1) * a.stride + a.k;             // slab rows per sequence of the tile
  const int rows = rows_s * a.spt;
  const long long row0 = q0 * a.stride - a.pad_left;
  const float* __restrict__ x = a.x + (size_t)grp * a.cin_g;

  float acc[TT][4];
  int srow[TT];                                                // first slab row (tap 0) of each of this thread's rows, -1 = unused
#pragma unroll
  for (int t = 0; t < TT; ++t) {
    acc[t][0] = acc[t][1] = acc[t][2] = acc[t][3] = 0.f;
    const int r = t * Cfg::TG + tg;
    if (PACK) {
      const int sl = r / a.tqs, ql = r - sl * a.tqs;
      srow[t] = (sl < a.spt && seq0 + sl < a.n_seq && q0 + ql < a.n_out) ? sl * rows_s + ql * a.stride : -1;
    } else {
      srow[t] = r * a.stride;                                  // one sequence per tile (tqs = TQ): rows past n_out are computed, not stored
    }
  }

  for (int ci0 = 0; ci0 < a.cin_g; ci0 += kCiBlk) {
    const int cib = min(kCiBlk, a.cin_g - ci0);
    if (CI4) {
      const int c4n = cblk / 4;
      for (int idx = tid; idx < rows * c4n; idx += Cfg::THREADS) {
        const int r = idx / c4n, c4 = idx - r * c4n;
        const int sl = r / rows_s;
        const long long row = row0 + (r - sl * rows_s);
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (row >= 0 && row < a.n_in && seq0 + sl < a.n_seq && c4 * 4 < cib)
          v = __ldg(reinterpret_cast<const float4*>(x + ((size_t)(seq0 + sl) * a.n_in + row) * a.cin + ci0 + c4 * 4));
        *reinterpret_cast<float4*>(&slab[r * pitch + c4 * 4]) = v;
      }
    } else {
      for (int idx = tid; idx < rows * cblk; idx += Cfg::THREADS) {
        const int r = idx / cblk, c = idx - r * cblk;
        const int sl = r / rows_s;
        const long long row = row0 + (r - sl * rows_s);
        slab[r * pitch + c] = (row >= 0 && row < a.n_in && seq0 + sl < a.n_seq && c < cib)
                                  ? __ldg(x + ((size_t)(seq0 + sl) * a.n_in + row) * a.cin + ci0 + c) : 0.f;
      }
    }
    __syncthreads();
    if (CI4) {
      // one flat loop over (tap, 4-channel group) with the NEXT step's weights already in flight (they come from L2)
      const int c4n = cib / 4, nit = a.k * c4n;
      const float* __restrict__ wp = a.w + (size_t)ci0 * a.cout + co;
      const size_t tap_step = (size_t)a.cin_g * a.cout;
      float4 w0 = __ldg(reinterpret_cast<const float4*>(wp)), w1 = __ldg(reinterpret_cast<const float4*>(wp + a.cout));
      float4 w2 = __ldg(reinterpret_cast<const float4*>(wp + 2 * (size_t)a.cout)), w3 = __ldg(reinterpret_cast<const float4*>(wp + 3 * (size_t)a.cout));
      int j = 0, c4 = 0;
      for (int it = 0; it < nit; ++it) {
        int jn = j, cn = c4 + 1;
        if (cn == c4n) { cn = 0; ++jn; }
        float4 n0 = w0, n1 = w1, n2 = w2, n3 = w3;
        if (it + 1 < nit) {
          const float* __restrict__ np = wp + (size_t)jn * tap_step + (size_t)(cn * 4) * a.cout;
          n0 = __ldg(reinterpret_cast<const float4*>(np)); n1 = __ldg(reinterpret_cast<const float4*>(np + a.cout));
          n2 = __ldg(reinterpret_cast<const float4*>(np + 2 * (size_t)a.cout)); n3 = __ldg(reinterpret_cast<const float4*>(np + 3 * (size_t)a.cout));
        }
        const float* sj = &slab[j * pitch + c4 * 4];
#pragma unroll
        for (int t = 0; t < TT; ++t) {
          if (PACK && srow[t] < 0) continue;
          const float4 xv = *reinterpret_cast<const float4*>(&sj[srow[t] * pitch]);
          acc[t][0] = fmaf(xv.x, w0.x, acc[t][0]); acc[t][1] = fmaf(xv.x, w0.y, acc[t][1]);
          acc[t][2] = fmaf(xv.x, w0.z, acc[t][2]); acc[t][3] = fmaf(xv.x, w0.w, acc[t][3]);
          acc[t][0] = fmaf(xv.y, w1.x, acc[t][0]); acc[t][1] = fmaf(xv.y, w1.y, acc[t][1]);
          acc[t][2] = fmaf(xv.y, w1.z, acc[t][2]); acc[t][3] = fmaf(xv.y, w1.w, acc[t][3]);
          acc[t][0] = fmaf(xv.z, w2.x, acc[t][0]); acc[t][1] = fmaf(xv.z, w2.y, acc[t][1]);
          acc[t][2] = fmaf(xv.z, w2.z, acc[t][2]); acc[t][3] = fmaf(xv.z, w2.w, acc[t][3]);
          acc[t][0] = fmaf(xv.w, w3.x, acc[t][0]); acc[t][1] = fmaf(xv.w, w3.y, acc[t][1]);
          acc[t][2] = fmaf(xv.w, w3.z, acc[t][2]); acc[t][3] = fmaf(xv.w, w3.w, acc[t][3]);
        }
        w0 = n0; w1 = n1; w2 = n2; w3 = n3; j = jn; c4 = cn;
      }
    } else {
      for (int j = 0; j < a.k; ++j) {
        const float* __restrict__ wj = a.w + ((size_t)j * a.cin_g + ci0) * a.cout + co;
        const float* sj = &slab[j * pitch];
        for (int c = 0; c < cib; ++c) {
          const float4 w0 = __ldg(reinterpret_cast<const float4*>(wj + (size_t)c * a.cout));
#pragma unroll
          for (int t = 0; t < TT; ++t) {
            if (PACK && srow[t] < 0) continue;
            const float xv = sj[srow[t] * pitch + c];
            acc[t][0] = fmaf(xv, w0.x, acc[t][0]); acc[t][1] = fmaf(xv, w0.y, acc[t][1]);
            acc[t][2] = fmaf(xv, w0.z, acc[t][2]); acc[t][3] = fmaf(xv, w0.w, acc[t][3]);
          }
        }
      }
    }
    __syncthreads();
  }

  const float4 bias = __ldg(reinterpret_cast<const float4*>(a.bias + co));
#pragma unroll
  for (int t = 0; t < TT; ++t) {
    const int r = t * Cfg::TG + tg;
    const int sl = PACK ? r / a.tqs : 0;
    const long long q = q0 + (r - sl * a.tqs);
    if (PACK ? srow[t] < 0 : q >= a.n_out) continue;
    float4 v = make_float4(acc[t][0] + bias.x, acc[t][1] + bias.y, acc[t][2] + bias.z, acc[t][3] + bias.w);
    v.x = lrelu(v.x, a.out_slope); v.y = lrelu(v.y, a.out_slope); v.z = lrelu(v.z, a.out_slope); v.w = lrelu(v.w, a.out_slope);
    *reinterpret_cast<float4*>(a.out + ((size_t)(seq0 + sl) * a.n_out + q) * a.cout + co) = v;
  }
}

// conv_post (C_out = 1, model.py:213-216 / 328-332): one warp per output sample, the k * cin products reduced by shuffles.
__global__ void __launch_bounds__(256) disc_post_kernel(const DiscConvArgs a) {
  const int lane = threadIdx.x & 31;
  const long long warp = ((long long)blockIdx.x * blockDim.x + threadIdx.x) >> 5;
  const long long total = (long long)a.n_seq * a.n_out;
  if (warp >= total) return;
  const long long q = warp % a.n_out, seq = warp / a.n_out;
  float s = 0.f;
  for (int j = 0; j < a.k; ++j) {
    const long long row = q * a.stride + j - a.pad_left;
    if (row < 0 || row >= a.n_in) continue;
    const float* __restrict__ xr = a.x + ((size_t)seq * a.n_in + row) * a.cin;
    const float* __restrict__ wr = a.w + (size_t)j * a.cin;
    for (int c = lane * 4; c < a.cin; c += 128) {
      const float4 xv = __ldg(reinterpret_cast<const float4*>(xr + c));
      const float4 wv = __ldg(reinterpret_cast<const float4*>(wr + c));
      s = fmaf(xv.x, wv.x, s); s = fmaf(xv.y, wv.y, s); s = fmaf(xv.z, wv.z, s); s = fmaf(xv.w, wv.w, s);
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) s += __shfl_xor_sync(0xffffffffu, s, o);
  if (lane == 0) a.out[warp] = lrelu(s + a.bias[0], a.out_slope);
}

template <int CO_TILE, int TT, bool CI4>
cudaError_t launch_conv_t(const DiscConvArgs& a0, cudaStream_t s) {
  using Cfg = DiscCfg<CO_TILE, TT>;
  DiscConvArgs a = a0;
  // rows per sequence in a tile: the split of TQ into spt x tqs that computes the fewest padded rows overall
  int best_tqs = Cfg::TQ;
  double best_eff = 0.0;
  for (int tqs = Cfg::TQ; tqs >= 1; --tqs) {
    if (a.n_out > Cfg::TQ / 2 && tqs != Cfg::TQ) break;        // sequences that do not fit twice: one sequence per tile
    const int spt = Cfg::TQ / tqs;
    if ((size_t)((tqs - 1) * a.stride + a.k) * spt * (CI4 ? (a.cin_g < kCiBlk ? a.cin_g : kCiBlk) + 4 : 1) * sizeof(float) > 96 * 1024 && tqs != Cfg::TQ) continue;
    const long long qt = (a.n_out + tqs - 1) / tqs, zt = (a.n_seq + spt - 1) / spt;
    const double eff = (double)a.n_out * a.n_seq / ((double)qt * zt * Cfg::TQ);
    if (eff > best_eff + 1e-9) { best_eff = eff; best_tqs = tqs; }
  }
  a.tqs = best_tqs; a.spt = Cfg::TQ / best_tqs;
  const int cblk = a.cin_g < kCiBlk ? a.cin_g : kCiBlk;
  const int pitch = CI4 ? cblk + 4 : cblk;
  const size_t smem = (size_t)((a.tqs - 1) * a.stride + a.k) * a.spt * pitch * sizeof(float);
  if (smem > 200 * 1024) return cudaErrorInvalidValue;
  auto kern = a.spt > 1 ? disc_conv_kernel<CO_TILE, TT, CI4, true> : disc_conv_kernel<CO_TILE, TT, CI4, false>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  dim3 grid((unsigned)((a.n_out + a.tqs - 1) / a.tqs), a.cout / CO_TILE, (unsigned)((a.n_seq + a.spt - 1) / a.spt));
  kern<<<grid, Cfg::THREADS, smem, s>>>(a);
  return cudaGetLastError();
}

cudaError_t launch_conv(const DiscConvArgs& a, cudaStream_t s) {
  if (a.n_seq <= 0 || a.n_out <= 0) return cudaSuccess;
  if (a.cout == 1) {
    if (a.cin % 4) return cudaErrorInvalidValue;
    const long long warps = (long long)a.n_seq * a.n_out;
    disc_post_kernel<<<(unsigned)((warps * 32 + 255) / 256), 256, 0, s>>>(a);
    return cudaGetLastError();
  }
  const bool ci4 = a.cin_g % 4 == 0;
  if (!ci4) {                                            // the first conv of a sub-discriminator (C_in = 1)
    if (a.cout_g % 32 == 0) return launch_conv_t<32, 4, false>(a, s);
    return cudaErrorInvalidValue;
  }
  if (a.cout_g % 128 == 0) return launch_conv_t<128, 8, true>(a, s);
  if (a.cout_g % 64 == 0) return launch_conv_t<64, 8, true>(a, s);
  if (a.cout_g % 32 == 0) return launch_conv_t<32, 4, true>(a, s);
  if (a.cout_g % 16 == 0) return launch_conv_t<16, 4, true>(a, s);
  return cudaErrorInvalidValue;
}

}  // namespace

// ------------------------------------------------------------------ host side
namespace {

struct DiscLayer {
  ConvSpec spec;
  std::string name;                     // discriminators.<i>.convs.<l> / discriminators.<i>.conv_post
  std::vector<float> v, g, b, w;        // as set by the caller (w: stored kernel, used as is)
  float* d_w = nullptr;                 // folded [k][cin_g][cout]
  float* d_b = nullptr;
  bool tc = false;                      // on the tensor cores (3xTF32): every conv with c_in / groups >= 8 and c_out > 1
  std::vector<float> hi_lo;             // host [w_hi taps | w_lo taps], 2 * k * cin_g * cout floats
  struct Packed { void* main = nullptr; void* lo = nullptr; };
  // packed weight images, keyed by the left pad: with a stride, TensorFlow's SAME pad depends on the input length, the
  // pad decides which input phase each tap reads, and the kernel wants its taps phase by phase.  Built on first use.
  std::map<int, Packed> packed;
  float* d_zero = nullptr;              // zero bias of the x_lo pass
};

struct SubDisc {
  int period = 1;                       // DiscriminatorP; 1 for DiscriminatorS
  int pools = 0;                        // DiscriminatorS: number of mean pools in front of it (model.py:376-379)
  std::vector<DiscLayer> layers;
};

struct MapDesc { long long offset; int B, period; long long len; int C; };

}  // namespace

struct hfg_disc_ctx {
  int kind = 0, compat = 0, device = 0;
  bool finalized = false;
  bool use_tc = true;                                  // HFG_DISC_TC=0: every layer on the FFMA kernel
  std::vector<SubDisc> subs;
  std::string err;
  float* ws = nullptr; size_t ws_floats = 0;           // split / pooled waveforms
  float* d_audio = nullptr; size_t audio_floats = 0;   // host path staging
  float* d_maps = nullptr; size_t maps_floats = 0;
  void* d_loss = nullptr;                              // segs | partial sums | per-map means (hfg_disc_pair_losses)
};

namespace {

thread_local std::string g_disc_create_error;

int dfail(hfg_disc_ctx* h, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
  if (h) h->err = buf; else g_disc_create_error = buf;
  return code;
}

#define DISC_CUDA_TRY(h, expr)                                                               \
  do { cudaError_t e__ = (expr);                                                             \
       if (e__ != cudaSuccess) return dfail(h, HFG_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e__)); } while (0)

// TensorFlow SAME / VALID (compat 0) or torch padding (k - 1) / 2 (compat 1): -> left pad and output length
void conv_pads(long long n, const ConvSpec& c, int compat, int* left, long long* n_out) {
  if (compat == 1) {
    const int p = (c.k - 1) / 2;
    *left = p; *n_out = n + 2 * p - c.k >= 0 ? (n + 2 * p - c.k) / c.stride + 1 : 0;
    return;
  }
  if (c.valid) { *left = 0; *n_out = n >= c.k ? (n - c.k) / c.stride + 1 : 0; return; }
  const long long out = (n + c.stride - 1) / c.stride;
  const long long total = std::max<long long>((out - 1) * c.stride + c.k - n, 0);
  *left = (int)(total / 2); *n_out = out;
}

void pool_pads(long long n, int compat, int* left, long long* n_out) {
  if (compat == 1) { *left = 1; *n_out = n >= 2 ? (n + 2 - 4) / 2 + 1 : 0; return; }
  const long long out = (n + 1) / 2;
  const long long total = std::max<long long>((out - 1) * 2 + 4 - n, 0);
  *left = (int)(total / 2); *n_out = out;
}

// Length of the sequences the first conv of sub-discriminator `s` sees, and the scratch floats it needs.
long long first_len(const hfg_disc_ctx* h, const SubDisc& s, long long T) {
  if (h->kind == 0) return (T + s.period - 1) / s.period;
  long long n = T;
  for (int i = 0; i < s.pools; ++i) { int l; pool_pads(n, h->compat, &l, &n); }
  return n;
}

int build_layout(hfg_disc_ctx* h, int B, long long T, std::vector<MapDesc>* maps, long long* total) {
  long long off = 0;
  for (const SubDisc& s : h->subs) {
    long long n = first_len(h, s, T);
    for (const DiscLayer& L : s.layers) {
      int left; long long n_out;
      conv_pads(n, L.spec, h->compat, &left, &n_out);
      if (n_out < 1) return dfail(h, HFG_ERR_INVALID, "T=%lld is too short for %s (it would see %lld samples)", T, L.name.c_str(), n);
      maps->push_back({off, B, s.period, n_out, L.spec.cout});
      off += (long long)B * s.period * n_out * L.spec.cout;
      off = (off + 3) / 4 * 4;                               // 16-byte aligned maps (float4 loads in the kernels)
      n = n_out;
    }
  }
  *total = off;
  return HFG_OK;
}

int check_shape(hfg_disc_ctx* h, const char* name, const int64_t* shape, int ndim, size_t want, int last) {
  size_t n = 1;
  for (int i = 0; i < ndim; ++i) n *= (size_t)shape[i];
  if (ndim < 1 || n != want || (last > 0 && shape[ndim - 1] != last))
    return dfail(h, HFG_ERR_INVALID, "%s: expected %zu elements with last axis %d", name, want, last);
  return HFG_OK;
}

int run_disc_forward(hfg_disc_ctx* h, const float* audio, int B, long long T, float* maps_dev, cudaStream_t s) {
  std::vector<MapDesc> maps;
  long long total = 0;
  int rc = build_layout(h, B, T, &maps, &total);
  if (rc != HFG_OK) return rc;
  // scratch: DiscriminatorP's split waveform (one at a time, stream-ordered) / DiscriminatorS's pooled waveforms
  size_t need = 0;
  if (h->kind == 0) { for (const SubDisc& sd : h->subs) need = std::max(need, (size_t)B * sd.period * (size_t)first_len(h, sd, T)); }
  else { for (const SubDisc& sd : h->subs) if (sd.pools) need += (size_t)B * (size_t)first_len(h, sd, T); }
  need = (need + 63) / 64 * 64;                              // the tensor-core scratch below starts 256-byte aligned
  size_t tc_lo = 0, tc_part = 0;                             // x_lo and the x_lo . w_hi partial map of a 3xTF32 layer (largest)
  {
    size_t k = 0;
    for (const SubDisc& sd : h->subs)
      for (size_t l = 0; l < sd.layers.size(); ++l, ++k)
        if (sd.layers[l].tc) {
          tc_lo = std::max(tc_lo, (size_t)B * sd.period * (size_t)maps[k - 1].len * sd.layers[l].spec.cin);
          tc_part = std::max(tc_part, (size_t)B * sd.period * (size_t)maps[k].len * sd.layers[l].spec.cout);
        }
    tc_lo = (tc_lo + 63) / 64 * 64;
  }
  if (need + tc_lo + tc_part > h->ws_floats) {
    if (h->ws) { DISC_CUDA_TRY(h, cudaStreamSynchronize(s)); cudaFree(h->ws); h->ws = nullptr; h->ws_floats = 0; }
    DISC_CUDA_TRY(h, cudaMalloc((void**)&h->ws, (need + tc_lo + tc_part) * sizeof(float)));
    h->ws_floats = need + tc_lo + tc_part;
  }
  float* const ws_lo = h->ws + need;
  float* const ws_part = ws_lo + tc_lo;
  size_t mi = 0;
  const float* pooled = audio; long long pooled_n = T; float* ws_next = h->ws;
  for (const SubDisc& sd : h->subs) {
    const float* x = nullptr;
    long long n = first_len(h, sd, T);
    const int n_seq = B * sd.period;
    if (h->kind == 0) {
      const long long tot = (long long)n_seq * n;
      disc_split_kernel<<<(unsigned)std::min<long long>((tot + 255) / 256, 148 * 16), 256, 0, s>>>(audio, h->ws, B, T, sd.period, n);
      DISC_CUDA_TRY(h, cudaGetLastError());
      x = h->ws;
    } else if (sd.pools == 0) {
      x = audio;
    } else {                                                 // scale i pools scale i-1's input once more (model.py:376-379)
      int left; long long n_out;
      pool_pads(pooled_n, h->compat, &left, &n_out);
      const long long tot = (long long)B * n_out;
      disc_pool_kernel<<<(unsigned)std::min<long long>((tot + 255) / 256, 148 * 16), 256, 0, s>>>(pooled, ws_next, B, pooled_n, n_out, left, h->compat == 1);
      DISC_CUDA_TRY(h, cudaGetLastError());
      pooled = ws_next; pooled_n = n_out; ws_next += (size_t)B * n_out;
      x = pooled;
    }
    for (size_t l = 0; l < sd.layers.size(); ++l, ++mi) {
      const DiscLayer& L = sd.layers[l];
      DiscConvArgs a{};
      int left; long long n_out;
      conv_pads(n, L.spec, h->compat, &left, &n_out);
      a.x = x; a.w = L.d_w; a.bias = L.d_b; a.out = maps_dev + maps[mi].offset;
      a.n_seq = n_seq; a.n_in = n; a.n_out = n_out;
      a.cin = L.spec.cin; a.cout = L.spec.cout; a.cin_g = L.spec.cin / L.spec.groups; a.cout_g = L.spec.cout / L.spec.groups;
      a.k = L.spec.k; a.stride = L.spec.stride; a.pad_left = left;
      a.out_slope = l + 1 == sd.layers.size() ? 1.0f : kSlope;        // conv_post has no activation (model.py:241 / 342)
      cudaError_t e;
      if (L.tc && n >= L.spec.stride) {
        const int k = L.spec.k;
        int off[2 * 41];
        for (int j = 0; j < k; ++j) off[j] = off[j + k] = j - left;
        auto it = L.packed.find(left);
        if (it == L.packed.end()) {                              // first use of this pad: pack and upload (synchronous)
          const size_t kel = (size_t)k * a.cin_g * a.cout;
          std::vector<float> pk(2 * kel);
          DiscLayer::Packed pw;
          hfg::tf32_pack_weights_ext(L.hi_lo.data(), 2 * k, off, L.spec.stride, a.cin_g, a.cout, L.spec.groups, pk.data());
          DISC_CUDA_TRY(h, cudaMalloc(&pw.main, 2 * kel * sizeof(float)));
          DISC_CUDA_TRY(h, cudaMemcpy(pw.main, pk.data(), 2 * kel * sizeof(float), cudaMemcpyHostToDevice));
          hfg::tf32_pack_weights_ext(L.hi_lo.data(), k, off, L.spec.stride, a.cin_g, a.cout, L.spec.groups, pk.data());
          DISC_CUDA_TRY(h, cudaMalloc(&pw.lo, kel * sizeof(float)));
          DISC_CUDA_TRY(h, cudaMemcpy(pw.lo, pk.data(), kel * sizeof(float), cudaMemcpyHostToDevice));
          it = const_cast<DiscLayer&>(L).packed.emplace(left, pw).first;
        }
        const long long n4 = (long long)n_seq * n * L.spec.cin / 4;
        disc_lo_kernel<<<(unsigned)std::min<long long>((n4 + 255) / 256, 148 * 8), 256, 0, s>>>(
            reinterpret_cast<const float4*>(x), reinterpret_cast<float4*>(ws_lo), n4);
        e = cudaGetLastError();
        hfg::TcArgs t{};
        t.B = n_seq; t.x_rows = (int)n; t.n_q = (int)n_out; t.out_rows = (int)n_out; t.out_stride = 1;
        t.c_in = L.spec.cin; t.c_out = L.spec.cout; t.res_inv = 1.0f; t.out_scale = 1.0f; t.raw_out = 1; t.nvar = 1;
        t.in_stride = L.spec.stride; t.cin_g = a.cin_g; t.groups = L.spec.groups; t.ext_tap_off = off;
        // pass 1: partial = x_lo . w_hi
        t.x = ws_lo; t.out = ws_part; t.bias = L.d_zero; t.out_slope = 1.0f; t.var[0].w = it->second.lo; t.ext_ntaps = k;
        if (e == cudaSuccess) e = hfg::launch_conv_tf32(t, s);
        // pass 2: out = lrelu(bias + x_hi . w_hi + x_hi . w_lo + partial)
        t.x = x; t.out = a.out; t.bias = L.d_b; t.res2 = ws_part; t.out_slope = a.out_slope; t.var[0].w = it->second.main; t.ext_ntaps = 2 * k;
        if (e == cudaSuccess) e = hfg::launch_conv_tf32(t, s);
      } else {
        e = launch_conv(a, s);
      }
      if (e != cudaSuccess) return dfail(h, HFG_ERR_CUDA, "%s launch failed: %s", L.name.c_str(), cudaGetErrorString(e));
      x = a.out; n = n_out;
    }
  }
  return HFG_OK;
}

}  // namespace

extern "C" {

int hfg_disc_create(int kind, int compat, int device, hfg_disc_handle* out) {
  if (!out) return HFG_ERR_INVALID;
  *out = nullptr;
  if (kind != 0 && kind != 1) return dfail(nullptr, HFG_ERR_INVALID, "kind must be 0 (MultiPeriodDiscriminator) or 1 (MultiScaleDiscriminator)");
  if (compat != 0 && compat != 1) return dfail(nullptr, HFG_ERR_INVALID, "compat must be 0 (reference) or 1 (upstream)");
  int ndev = 0;
  if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev)
    return dfail(nullptr, HFG_ERR_CUDA, "no CUDA device %d (this library has no CPU path)", device);
  hfg_disc_ctx* h = new hfg_disc_ctx();
  h->kind = kind; h->compat = compat; h->device = device;
  { const char* e = getenv("HFG_DISC_TC"); if (e && atoi(e) == 0) h->use_tc = false; }
  const int nsub = kind == 0 ? 5 : 3;
  for (int i = 0; i < nsub; ++i) {
    SubDisc sd;
    sd.period = kind == 0 ? kPeriods[i] : 1;
    sd.pools = kind == 0 ? 0 : i;
    const ConvSpec* specs = kind == 0 ? kPConvs : kSConvs;
    const int nl = kind == 0 ? 6 : 8;
    for (int l = 0; l < nl; ++l) {
      DiscLayer L;
      L.spec = specs[l];
      char nm[96];
      if (l + 1 < nl) snprintf(nm, sizeof nm, "discriminators.%d.convs.%d", i, l); else snprintf(nm, sizeof nm, "discriminators.%d.conv_post", i);
      L.name = nm;
      sd.layers.push_back(L);
    }
    h->subs.push_back(sd);
  }
  *out = h;
  return HFG_OK;
}

int hfg_disc_set_weight(hfg_disc_handle h, const char* name, const float* host, const int64_t* shape, int ndim) {
  if (!h || !name || !host || !shape) return HFG_ERR_INVALID;
  const std::string full(name);
  const size_t dot = full.rfind('.');
  if (dot == std::string::npos) return dfail(h, HFG_ERR_INVALID, "bad weight name '%s'", name);
  const std::string layer = full.substr(0, dot), part = full.substr(dot + 1);
  for (SubDisc& sd : h->subs)
    for (DiscLayer& L : sd.layers) {
      if (L.name != layer) continue;
      const size_t kel = (size_t)L.spec.k * (L.spec.cin / L.spec.groups) * L.spec.cout;
      std::vector<float>* dst = nullptr; size_t want = 0;
      if (part == "v") { dst = &L.v; want = kel; } else if (part == "w") { dst = &L.w; want = kel; }
      else if (part == "g") { dst = &L.g; want = (size_t)L.spec.cout; } else if (part == "b") { dst = &L.b; want = (size_t)L.spec.cout; }
      else return dfail(h, HFG_ERR_INVALID, "unknown weight part '%s' (v, g, b or w)", part.c_str());
      const int rc = check_shape(h, name, shape, ndim, want, L.spec.cout);
      if (rc != HFG_OK) return rc;
      dst->assign(host, host + want);
      h->finalized = false;
      return HFG_OK;
    }
  return dfail(h, HFG_ERR_INVALID, "unknown layer '%s'", layer.c_str());
}

int hfg_disc_finalize_weights(hfg_disc_handle h) {
  if (!h) return HFG_ERR_INVALID;
  DISC_CUDA_TRY(h, cudaSetDevice(h->device));
  for (SubDisc& sd : h->subs)
    for (DiscLayer& L : sd.layers) {
      const int cout = L.spec.cout;
      const size_t kel = (size_t)L.spec.k * (L.spec.cin / L.spec.groups) * cout;
      std::vector<float> w(kel);
      if (L.b.empty()) return dfail(h, HFG_ERR_STATE, "%s.b was not set", L.name.c_str());
      if (!L.w.empty()) {
        w = L.w;                                             // SpectralNormalization at inference / an already folded kernel
      } else {
        if (L.v.empty() || L.g.empty()) return dfail(h, HFG_ERR_STATE, "%s needs (v, g) or w", L.name.c_str());
        // tfa WeightNormalization(data_init=False): w = v * g / ||v||, norm over every axis but the last, in fp64
        std::vector<double> nrm(cout, 0.0);
        for (size_t i = 0; i < kel; ++i) nrm[i % cout] += (double)L.v[i] * L.v[i];
        for (int c = 0; c < cout; ++c) nrm[c] = std::sqrt(nrm[c]);
        for (size_t i = 0; i < kel; ++i) w[i] = (float)((double)L.v[i] * ((double)L.g[i % cout] / nrm[i % cout]));
      }
      if (L.d_w) cudaFree(L.d_w);
      if (L.d_b) cudaFree(L.d_b);
      L.d_w = nullptr; L.d_b = nullptr;
      std::vector<float> bias((cout + 3) / 4 * 4, 0.f);
      memcpy(bias.data(), L.b.data(), sizeof(float) * cout);
      DISC_CUDA_TRY(h, cudaMalloc((void**)&L.d_w, kel * sizeof(float)));
      DISC_CUDA_TRY(h, cudaMemcpy(L.d_w, w.data(), kel * sizeof(float), cudaMemcpyHostToDevice));
      DISC_CUDA_TRY(h, cudaMalloc((void**)&L.d_b, bias.size() * sizeof(float)));
      DISC_CUDA_TRY(h, cudaMemcpy(L.d_b, bias.data(), bias.size() * sizeof(float), cudaMemcpyHostToDevice));
      // tensor cores at fp32 accuracy (3xTF32, see the header of this file); weight images are packed on first use
      for (auto& kv : L.packed) { if (kv.second.main) cudaFree(kv.second.main); if (kv.second.lo) cudaFree(kv.second.lo); }
      L.packed.clear();
      if (L.d_zero) cudaFree(L.d_zero);
      L.d_zero = nullptr;
      L.hi_lo.clear();
      {
        int off[2 * 41];
        for (int j = 0; j < L.spec.k; ++j) off[j] = off[j + L.spec.k] = j - (L.spec.k - 1) / 2;
        L.tc = h->use_tc && cout > 1 && L.spec.k <= 41 && L.spec.cin % 4 == 0 &&
               hfg::tf32_ext_supported(L.spec.cin, cout, L.spec.groups, L.spec.stride, 2 * L.spec.k, off);
      }
      if (L.tc) {
        L.hi_lo.resize(2 * kel);
        for (size_t i = 0; i < kel; ++i) {
          uint32_t u; memcpy(&u, &w[i], 4);
          u = (u + 0x0FFFu + ((u >> 13) & 1u)) & ~0x1FFFu;      // w_hi: round to nearest even at 10 mantissa bits
          float hi; memcpy(&hi, &u, 4);
          L.hi_lo[i] = hi; L.hi_lo[kel + i] = w[i] - hi;        // w_lo: exact in fp32
        }
        DISC_CUDA_TRY(h, cudaMalloc((void**)&L.d_zero, cout * sizeof(float)));
        DISC_CUDA_TRY(h, cudaMemset(L.d_zero, 0, cout * sizeof(float)));
      }
    }
  h->finalized = true;
  return HFG_OK;
}

int hfg_disc_num_maps(hfg_disc_handle h, int* out) {
  if (!h || !out) return HFG_ERR_INVALID;
  int n = 0;
  for (const SubDisc& sd : h->subs) n += (int)sd.layers.size();
  *out = n;
  return HFG_OK;
}

int hfg_disc_layout(hfg_disc_handle h, int B, long long T, long long* offsets, long long* dims, long long* total_floats) {
  if (!h || B < 1 || T < 1 || !total_floats) return h ? dfail(h, HFG_ERR_INVALID, "hfg_disc_layout: bad argument") : HFG_ERR_INVALID;
  std::vector<MapDesc> maps;
  const int rc = build_layout(h, B, T, &maps, total_floats);
  if (rc != HFG_OK) return rc;
  for (size_t i = 0; i < maps.size(); ++i) {
    if (offsets) offsets[i] = maps[i].offset;
    if (dims) { dims[4 * i] = maps[i].B; dims[4 * i + 1] = maps[i].period; dims[4 * i + 2] = maps[i].len; dims[4 * i + 3] = maps[i].C; }
  }
  return HFG_OK;
}

int hfg_disc_forward(hfg_disc_handle h, const float* audio_dev, int B, long long T, float* maps_dev, void* stream) {
  if (!h || !audio_dev || !maps_dev || B < 1 || T < 1) return h ? dfail(h, HFG_ERR_INVALID, "hfg_disc_forward: bad argument") : HFG_ERR_INVALID;
  if (!h->finalized) return dfail(h, HFG_ERR_STATE, "weights not finalised");
  if (h->kind == 0 && T < 12) return dfail(h, HFG_ERR_INVALID, "T=%lld: reflect padding to a multiple of period 11 needs at least 12 samples", T);
  DISC_CUDA_TRY(h, cudaSetDevice(h->device));
  return run_disc_forward(h, audio_dev, B, T, maps_dev, (cudaStream_t)stream);
}

int hfg_disc_forward_host(hfg_disc_handle h, const float* audio_host, int B, long long T, float* maps_host) {
  if (!h || !audio_host || !maps_host || B < 1 || T < 1) return h ? dfail(h, HFG_ERR_INVALID, "hfg_disc_forward_host: bad argument") : HFG_ERR_INVALID;
  if (!h->finalized) return dfail(h, HFG_ERR_STATE, "weights not finalised");
  std::vector<MapDesc> maps;
  long long total = 0;
  int rc = build_layout(h, B, T, &maps, &total);
  if (rc != HFG_OK) return rc;
  DISC_CUDA_TRY(h, cudaSetDevice(h->device));
  const size_t na = (size_t)B * (size_t)T;
  if (na > h->audio_floats) {
    if (h->d_audio) cudaFree(h->d_audio);
    h->d_audio = nullptr; h->audio_floats = 0;
    DISC_CUDA_TRY(h, cudaMalloc((void**)&h->d_audio, na * sizeof(float)));
    h->audio_floats = na;
  }
  if ((size_t)total > h->maps_floats) {
    if (h->d_maps) cudaFree(h->d_maps);
    h->d_maps = nullptr; h->maps_floats = 0;
    DISC_CUDA_TRY(h, cudaMalloc((void**)&h->d_maps, (size_t)total * sizeof(float)));
    h->maps_floats = (size_t)total;
  }
  DISC_CUDA_TRY(h, cudaMemcpyAsync(h->d_audio, audio_host, na * sizeof(float), cudaMemcpyHostToDevice, 0));
  rc = hfg_disc_forward(h, h->d_audio, B, T, h->d_maps, nullptr);
  if (rc != HFG_OK) return rc;
  DISC_CUDA_TRY(h, cudaMemcpyAsync(maps_host, h->d_maps, (size_t)total * sizeof(float), cudaMemcpyDeviceToHost, 0));
  DISC_CUDA_TRY(h, cudaStreamSynchronize(0));
  return HFG_OK;
}

int hfg_disc_pair_losses(hfg_disc_handle h, const float* maps_dev, int B, long long T, float* out_host, void* stream) {
  if (!h || !maps_dev || !out_host || B < 1 || T < 1) return h ? dfail(h, HFG_ERR_INVALID, "hfg_disc_pair_losses: bad argument") : HFG_ERR_INVALID;
  std::vector<MapDesc> maps;
  long long total = 0;
  int rc = build_layout(h, 2 * B, T, &maps, &total);
  if (rc != HFG_OK) return rc;
  DISC_CUDA_TRY(h, cudaSetDevice(h->device));
  cudaStream_t s = (cudaStream_t)stream;
  const int nm = (int)maps.size();
  const size_t seg_bytes = (nm * sizeof(LossSeg) + 255) / 256 * 256, part_bytes = (size_t)nm * kLossBlocks * 4 * sizeof(float);
  if (!h->d_loss) DISC_CUDA_TRY(h, cudaMalloc(&h->d_loss, seg_bytes + part_bytes + (size_t)nm * 4 * sizeof(float)));
  LossSeg* d_segs = reinterpret_cast<LossSeg*>(h->d_loss);
  float* d_part = reinterpret_cast<float*>(reinterpret_cast<char*>(h->d_loss) + seg_bytes);
  float* d_mean = d_part + (size_t)nm * kLossBlocks * 4;
  std::vector<LossSeg> segs(nm);
  for (int i = 0; i < nm; ++i) { segs[i].off = maps[i].offset; segs[i].half = (long long)B * maps[i].period * maps[i].len * maps[i].C; }
  DISC_CUDA_TRY(h, cudaMemcpyAsync(d_segs, segs.data(), nm * sizeof(LossSeg), cudaMemcpyHostToDevice, s));
  disc_loss_kernel<<<dim3(kLossBlocks, nm), 256, 0, s>>>(maps_dev, d_segs, d_part);
  DISC_CUDA_TRY(h, cudaGetLastError());
  disc_loss_final_kernel<<<(nm * 4 + 127) / 128, 128, 0, s>>>(d_part, d_segs, d_mean, nm);
  DISC_CUDA_TRY(h, cudaGetLastError());
  std::vector<float> mean((size_t)nm * 4);
  DISC_CUDA_TRY(h, cudaMemcpyAsync(mean.data(), d_mean, mean.size() * sizeof(float), cudaMemcpyDeviceToHost, s));
  DISC_CUDA_TRY(h, cudaStreamSynchronize(s));
  // losses.py: feature_loss = 2 * sum over every map of mean|r - g|; discriminator_loss = sum over sub-discriminators of
  // mean((1 - D(y))^2) + mean(D(y_hat)^2); generator_loss = sum of mean((1 - D(y_hat))^2); scores = each sub's last map
  const int nsub = (int)h->subs.size();
  double fm = 0.0, dl = 0.0, gl = 0.0;
  int mi = 0;
  for (int sd = 0; sd < nsub; ++sd) {
    const int nl = (int)h->subs[sd].layers.size();
    for (int l = 0; l < nl; ++l, ++mi) fm += mean[(size_t)mi * 4];
    const float* sc = &mean[(size_t)(mi - 1) * 4];
    out_host[3 + sd] = sc[1]; out_host[3 + nsub + sd] = sc[2]; out_host[3 + 2 * nsub + sd] = sc[3];
    dl += (double)sc[1] + (double)sc[2]; gl += sc[3];
  }
  out_host[0] = (float)(2.0 * fm); out_host[1] = (float)dl; out_host[2] = (float)gl;
  return HFG_OK;
}

const char* hfg_disc_last_error(hfg_disc_handle h) { return h ? h->err.c_str() : g_disc_create_error.c_str(); }

void hfg_disc_destroy(hfg_disc_handle h) {
  if (!h) return;
  cudaSetDevice(h->device);
  for (SubDisc& sd : h->subs)
    for (DiscLayer& L : sd.layers) {
      if (L.d_w) cudaFree(L.d_w);
      if (L.d_b) cudaFree(L.d_b);
      for (auto& kv : L.packed) { if (kv.second.main) cudaFree(kv.second.main); if (kv.second.lo) cudaFree(kv.second.lo); }
      if (L.d_zero) cudaFree(L.d_zero);
    }
  if (h->ws) cudaFree(h->ws);
  if (h->d_audio) cudaFree(h->d_audio);
  if (h->d_maps) cudaFree(h->d_maps);
  if (h->d_loss) cudaFree(h->d_loss);
  delete h;
}

}  // extern "C"
