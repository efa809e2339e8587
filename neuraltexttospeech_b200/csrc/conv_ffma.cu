// Direct (CUDA-core FFMA) kernels of the HiFi-GAN generator.
//
//  * conv_ffma_kernel : generic tapped conv, fp32 accumulate.  It is the whole forward pass in
//    the fp32 validation mode (north star: <=1e-5 of the fp32 reference) and the fallback for
//    configs outside the tcgen05 tile table (conv_tc.cu); no conv of the stock V1 / V2 / V3 configs
//    runs here in a 16-bit or TF32 mode (time folding covers the 32 / 16 / 8-channel stages).
//  * conv_post_kernel : lrelu(0.3) -> Conv1D(1, 7) -> tanh  (reference HiFiGAN_TF/model.py:169-171),
//    a pure bandwidth kernel (C_out = 1), shared by all precisions.
//  * cast_pad_kernel  : fp32 mel -> 16-bit, channel-padded, for the tensor-core conv_pre.
//
// Semantics restated from reference HiFiGAN_TF/model.py:71-80,106-112,155-173 (see kernels.cuh).
#include <type_traits>

#include "kernels.cuh"

namespace hfg {

// ------------------------------------------------------------------ 4-wide activation IO
template <typename T> struct Act4;
template <> struct Act4<float> {
  static __device__ __forceinline__ float4 load(const float* p) { return *reinterpret_cast<const float4*>(p); }
  static __device__ __forceinline__ void store(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
};
template <> struct Act4<__half> {
  static __device__ __forceinline__ float4 load(const __half* p) {
    uint2 u = *reinterpret_cast<const uint2*>(p);
    float2 a = __half22float2(*reinterpret_cast<__half2*>(&u.x));
    float2 b = __half22float2(*reinterpret_cast<__half2*>(&u.y));
    return make_float4(a.x, a.y, b.x, b.y);
  }
  static __device__ __forceinline__ void store(__half* p, float4 v) {
    __half2 a = __floats2half2_rn(v.x, v.y), b = __floats2half2_rn(v.z, v.w);
    uint2 u; u.x = *reinterpret_cast<uint32_t*>(&a); u.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(p) = u;
  }
};
template <> struct Act4<__nv_bfloat16> {
  static __device__ __forceinline__ float4 load(const __nv_bfloat16* p) {
    uint2 u = *reinterpret_cast<const uint2*>(p);
    float2 a = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u.x));
    float2 b = __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u.y));
    return make_float4(a.x, a.y, b.x, b.y);
  }
  static __device__ __forceinline__ void store(__nv_bfloat16* p, float4 v) {
    __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
    uint2 u; u.x = *reinterpret_cast<uint32_t*>(&a); u.y = *reinterpret_cast<uint32_t*>(&b);
    *reinterpret_cast<uint2*>(p) = u;
  }
};

__device__ __forceinline__ float lrelu(float v, float slope) { return v >= 0.f ? v : v * slope; }

// ------------------------------------------------------------------ generic tapped conv
constexpr int kCiBlk = 32;            // input channels staged per pass
constexpr int kPitch = kCiBlk + 4;    // smem row pitch in floats (16B aligned, bank-shifted)
constexpr int kTT = 8;                // q rows per thread

template <int CO_TILE> struct FfmaCfg {
  static constexpr int CG = CO_TILE / 4;                       // thread groups along c_out
  static constexpr int THREADS = (CG * 32 < 256) ? CG * 32 : 256;
  static constexpr int TG = THREADS / CG;                      // thread groups along q
  static constexpr int TQ = TG * kTT;                          // q rows per block
};

template <typename T, int CO_TILE>
__global__ void __launch_bounds__(FfmaCfg<CO_TILE>::THREADS)
conv_ffma_kernel(const ConvArgs a) {
  using Cfg = FfmaCfg<CO_TILE>;
  extern __shared__ __align__(16) float slab[];
  const int tid = threadIdx.x;
  const int cg = tid % Cfg::CG, tg = tid / Cfg::CG;
  const int q0 = blockIdx.x * Cfg::TQ;
  const int co = blockIdx.y * CO_TILE + cg * 4;
  const int b = blockIdx.z;

  int min_off = a.tap_off[0], max_off = a.tap_off[0];
  for (int j = 1; j < a.ntaps; ++j) { min_off = min(min_off, a.tap_off[j]); max_off = max(max_off, a.tap_off[j]); }
  const int rows = Cfg::TQ + (max_off - min_off);

  const T* __restrict__ x = reinterpret_cast<const T*>(a.x) + (size_t)b * a.x_rows * a.c_in;
  const float* __restrict__ w = reinterpret_cast<const float*>(a.w);

  float acc[kTT][4];
#pragma unroll
  for (int t = 0; t < kTT; ++t) { acc[t][0] = acc[t][1] = acc[t][2] = acc[t][3] = 0.f; }

  for (int ci0 = 0; ci0 < a.c_in; ci0 += kCiBlk) {
    const int cib = min(kCiBlk, a.c_in - ci0);
    // stage the pre-activated input slab (zero outside [0, x_rows): equals TF zero padding because
    // lrelu(0) = 0, reference pads after the activation)
    for (int idx = tid; idx < rows * (kCiBlk / 4); idx += Cfg::THREADS) {
      const int r = idx / (kCiBlk / 4), c4 = idx % (kCiBlk / 4);
      const int row = q0 + min_off + r;
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (row >= 0 && row < a.x_rows && c4 * 4 < cib) {
        v = Act4<T>::load(x + (size_t)row * a.c_in + ci0 + c4 * 4);
        v.x = lrelu(v.x, a.in_slope); v.y = lrelu(v.y, a.in_slope);
        v.z = lrelu(v.z, a.in_slope); v.w = lrelu(v.w, a.in_slope);
      }
      *reinterpret_cast<float4*>(&slab[r * kPitch + c4 * 4]) = v;
    }
    __syncthreads();
    for (int j = 0; j < a.ntaps; ++j) {
      const float* __restrict__ wj = w + ((size_t)j * a.c_in + ci0) * a.c_out + co;
      const float* sj = &slab[(a.tap_off[j] - min_off + tg) * kPitch];
      for (int c4 = 0; c4 < cib / 4; ++c4) {
        const float4 w0 = __ldg(reinterpret_cast<const float4*>(wj + (size_t)(c4 * 4 + 0) * a.c_out));
        const float4 w1 = __ldg(reinterpret_cast<const float4*>(wj + (size_t)(c4 * 4 + 1) * a.c_out));
        const float4 w2 = __ldg(reinterpret_cast<const float4*>(wj + (size_t)(c4 * 4 + 2) * a.c_out));
        const float4 w3 = __ldg(reinterpret_cast<const float4*>(wj + (size_t)(c4 * 4 + 3) * a.c_out));
#pragma unroll
        for (int t = 0; t < kTT; ++t) {
          const float4 xv = *reinterpret_cast<const float4*>(&sj[t * Cfg::TG * kPitch + c4 * 4]);
          acc[t][0] = fmaf(xv.x, w0.x, acc[t][0]); acc[t][1] = fmaf(xv.x, w0.y, acc[t][1]);
          acc[t][2] = fmaf(xv.x, w0.z, acc[t][2]); acc[t][3] = fmaf(xv.x, w0.w, acc[t][3]);
          acc[t][0] = fmaf(xv.y, w1.x, acc[t][0]); acc[t][1] = fmaf(xv.y, w1.y, acc[t][1]);
          acc[t][2] = fmaf(xv.y, w1.z, acc[t][2]); acc[t][3] = fmaf(xv.y, w1.w, acc[t][3]);
          acc[t][0] = fmaf(xv.z, w2.x, acc[t][0]); acc[t][1] = fmaf(xv.z, w2.y, acc[t][1]);
          acc[t][2] = fmaf(xv.z, w2.z, acc[t][2]); acc[t][3] = fmaf(xv.z, w2.w, acc[t][3]);
          acc[t][0] = fmaf(xv.w, w3.x, acc[t][0]); acc[t][1] = fmaf(xv.w, w3.y, acc[t][1]);
          acc[t][2] = fmaf(xv.w, w3.z, acc[t][2]); acc[t][3] = fmaf(xv.w, w3.w, acc[t][3]);
        }
      }
    }
    __syncthreads();
  }

  // epilogue: bias, residual, MRF sum / average (reference model.py:79,111,161-167)
  const float4 bias = *reinterpret_cast<const float4*>(a.bias + co);
#pragma unroll
  for (int t = 0; t < kTT; ++t) {
    const int q = q0 + t * Cfg::TG + tg;
    if (q >= a.n_q) continue;
    const size_t idx = ((size_t)b * a.out_rows + (size_t)q * a.out_stride + a.out_off) * a.c_out + co;
    float4 v = make_float4(acc[t][0] + bias.x, acc[t][1] + bias.y, acc[t][2] + bias.z, acc[t][3] + bias.w);
    if (a.res) {
      const float4 r = Act4<T>::load(reinterpret_cast<const T*>(a.res) + idx);
      v.x += r.x; v.y += r.y; v.z += r.z; v.w += r.w;
    }
    if (a.mrf_mode == MRF_FIRST) {
      *reinterpret_cast<float4*>(a.acc + idx) = v;
    } else if (a.mrf_mode == MRF_ADD) {
      float4 s = *reinterpret_cast<const float4*>(a.acc + idx);
      s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      *reinterpret_cast<float4*>(a.acc + idx) = s;
    } else {
      if (a.mrf_mode == MRF_FINAL) {
        const float4 s = *reinterpret_cast<const float4*>(a.acc + idx);
        v.x = (s.x + v.x) / a.mrf_div; v.y = (s.y + v.y) / a.mrf_div;
        v.z = (s.z + v.z) / a.mrf_div; v.w = (s.w + v.w) / a.mrf_div;
      }
      if (a.out_slope != 1.0f) {
        v.x = lrelu(v.x, a.out_slope); v.y = lrelu(v.y, a.out_slope);
        v.z = lrelu(v.z, a.out_slope); v.w = lrelu(v.w, a.out_slope);
      }
      Act4<T>::store(reinterpret_cast<T*>(a.out) + idx, v);
    }
  }
}

template <typename T, int CO_TILE>
static cudaError_t launch_ffma_t(const ConvArgs& a, cudaStream_t s) {
  using Cfg = FfmaCfg<CO_TILE>;
  int min_off = a.tap_off[0], max_off = a.tap_off[0];
  for (int j = 1; j < a.ntaps; ++j) { min_off = min_off < a.tap_off[j] ? min_off : a.tap_off[j]; max_off = max_off > a.tap_off[j] ? max_off : a.tap_off[j]; }
  const size_t smem = (size_t)(Cfg::TQ + (max_off - min_off)) * kPitch * sizeof(float);
  auto kern = conv_ffma_kernel<T, CO_TILE>;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
  }
  dim3 grid((a.n_q + Cfg::TQ - 1) / Cfg::TQ, a.c_out / CO_TILE, a.B);
  kern<<<grid, Cfg::THREADS, smem, s>>>(a);
  return cudaGetLastError();
}

template <typename T>
static cudaError_t launch_ffma_co(const ConvArgs& a, cudaStream_t s) {
  if (a.c_out % 128 == 0) return launch_ffma_t<T, 128>(a, s);
  if (a.c_out % 64 == 0) return launch_ffma_t<T, 64>(a, s);
  if (a.c_out % 32 == 0) return launch_ffma_t<T, 32>(a, s);
  if (a.c_out % 16 == 0) return launch_ffma_t<T, 16>(a, s);
  if (a.c_out % 8 == 0) return launch_ffma_t<T, 8>(a, s);
  return cudaErrorInvalidValue;
}

cudaError_t launch_conv_ffma(const ConvArgs& a, ActType act, cudaStream_t s) {
  if (a.ntaps < 1 || a.ntaps > kMaxTaps || a.c_in % 4 != 0) return cudaErrorInvalidValue;
  if (a.B <= 0 || a.n_q <= 0) return cudaSuccess;
  switch (act) {
    case ACT_F32: return launch_ffma_co<float>(a, s);
    case ACT_F16: return launch_ffma_co<__half>(a, s);
    case ACT_BF16: return launch_ffma_co<__nv_bfloat16>(a, s);
  }
  return cudaErrorInvalidValue;
}

// ------------------------------------------------------------------ conv_post + tanh
// Pure bandwidth kernel (C_out = 1, ~7 FLOP/B).  A block stages kPostRows + ntaps - 1 activation rows in
// shared memory with coalesced 16-byte loads (rows outside the utterance = TF's zero padding), then each
// thread produces kPostRows / kPostThreads output samples from the tile; a 16-byte pad per row keeps the
// row-strided 16-byte shared loads of a quarter-warp on distinct banks (row pitch 80 B for 32 x fp16).
constexpr int kPostThreads = 256;
constexpr int kPostRows = 1024;

template <typename T>
__global__ void __launch_bounds__(kPostThreads) conv_post_kernel(const PostArgs a) {
  constexpr int EPC = 16 / (int)sizeof(T);             // elements per 16-byte chunk
  extern __shared__ __align__(16) uint8_t psm[];
  const int chunks = a.c_in / EPC;                     // 16-byte chunks per row
  const int pitch = (chunks + 1) * 16;                 // bytes
  const int tile_rows = kPostRows + a.ntaps - 1;
  float* wsm = reinterpret_cast<float*>(psm + (size_t)tile_rows * pitch);   // [ntaps][c_in]
  for (int i = threadIdx.x; i < a.ntaps * a.c_in; i += kPostThreads) wsm[i] = a.w[i];
  const int b = blockIdx.y;
  const int t0 = blockIdx.x * kPostRows;
  const uint8_t* __restrict__ x = reinterpret_cast<const uint8_t*>(a.x) + (size_t)b * a.rows * a.c_in * sizeof(T);
  const int row_bytes = chunks * 16;
  for (int i = threadIdx.x; i < tile_rows * chunks; i += kPostThreads) {
    const int r = i / chunks, c = i - r * chunks;
    const int row = t0 + r - a.pad_left;
    uint4 v = make_uint4(0u, 0u, 0u, 0u);
    if (row >= 0 && row < a.rows) v = __ldg(reinterpret_cast<const uint4*>(x + (size_t)row * row_bytes + c * 16));
    *reinterpret_cast<uint4*>(psm + (size_t)r * pitch + c * 16) = v;
  }
  __syncthreads();
  const bool act = a.in_slope != 1.0f;
#pragma unroll 1
  for (int o = threadIdx.x; o < kPostRows; o += kPostThreads) {
    const int t = t0 + o;
    if (t >= a.rows) break;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    for (int j = 0; j < a.ntaps; ++j) {
      const uint8_t* xr = psm + (size_t)(o + j) * pitch;
      const float* wj = wsm + j * a.c_in;
      for (int c = 0; c < chunks; ++c) {
        const uint4 u = *reinterpret_cast<const uint4*>(xr + c * 16);
        float f[EPC];
        if constexpr (sizeof(T) == 4) {
          f[0] = __uint_as_float(u.x); f[1] = __uint_as_float(u.y); f[2] = __uint_as_float(u.z); f[3] = __uint_as_float(u.w);
        } else {
          const uint32_t uu[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            float2 p;
            if constexpr (std::is_same<T, __half>::value) p = __half22float2(*reinterpret_cast<const __half2*>(&uu[e]));
            else p = __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(&uu[e]));
            f[2 * e] = p.x; f[2 * e + 1] = p.y;
          }
        }
        if (act) {
#pragma unroll
          for (int e = 0; e < EPC; ++e) f[e] = lrelu(f[e], a.in_slope);
        }
        const float* wc = wj + c * EPC;
#pragma unroll
        for (int e = 0; e < EPC; e += 4) {
          const float4 w4 = *reinterpret_cast<const float4*>(wc + e);
          s0 = fmaf(f[e + 0], w4.x, s0); s1 = fmaf(f[e + 1], w4.y, s1);
          s2 = fmaf(f[e + 2], w4.z, s2); s3 = fmaf(f[e + 3], w4.w, s3);
        }
      }
    }
    const float y = tanhf((s0 + s1) + (s2 + s3) + a.bias);
    if (a.out_pcm) a.out_pcm[(size_t)b * a.rows + t] = pcm16_of(y);
    else a.out[(size_t)b * a.rows + t] = y;
  }
}

cudaError_t launch_conv_post(const PostArgs& a, ActType act, cudaStream_t s) {
  const int epc = act == ACT_F32 ? 4 : 8;
  if (a.c_in % epc != 0 || a.ntaps < 1 || a.ntaps > kMaxTaps) return cudaErrorInvalidValue;
  if (a.B <= 0 || a.rows <= 0) return cudaSuccess;
  const dim3 grid((unsigned)((a.rows + kPostRows - 1) / kPostRows), (unsigned)a.B);
  const size_t smem = (size_t)(kPostRows + a.ntaps - 1) * (a.c_in / epc + 1) * 16 + (size_t)a.ntaps * a.c_in * sizeof(float);
  if (smem > 200 * 1024) return cudaErrorInvalidConfiguration;
  cudaError_t e = cudaSuccess;
  switch (act) {
    case ACT_F32:
      e = cudaFuncSetAttribute(conv_post_kernel<float>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e == cudaSuccess) conv_post_kernel<float><<<grid, kPostThreads, smem, s>>>(a);
      break;
    case ACT_F16:
      e = cudaFuncSetAttribute(conv_post_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e == cudaSuccess) conv_post_kernel<__half><<<grid, kPostThreads, smem, s>>>(a);
      break;
    case ACT_BF16:
      e = cudaFuncSetAttribute(conv_post_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e == cudaSuccess) conv_post_kernel<__nv_bfloat16><<<grid, kPostThreads, smem, s>>>(a);
      break;
  }
  if (e != cudaSuccess) return e;
  return cudaGetLastError();
}

// ------------------------------------------------------------------ mel cast + channel pad
template <typename T>
__global__ void cast_pad_kernel(const float* __restrict__ src, T* __restrict__ dst, long long rows, int c_src, int c_dst) {
  const long long n = rows * (c_dst / 4);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const long long r = i / (c_dst / 4);
    const int c = (int)(i % (c_dst / 4)) * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c < c_src) v = *reinterpret_cast<const float4*>(src + r * c_src + c);
    Act4<T>::store(dst + r * c_dst + c, v);
  }
}

cudaError_t launch_cast_pad(const float* src, void* dst, long long rows, int c_src, int c_dst, ActType act, cudaStream_t s) {
  if (c_src % 4 || c_dst % 4 || c_dst < c_src) return cudaErrorInvalidValue;
  if (rows <= 0) return cudaSuccess;
  const long long n = rows * (c_dst / 4);
  const unsigned blocks = (unsigned)((n + 255) / 256 > 148 * 16 ? 148 * 16 : (n + 255) / 256);
  switch (act) {
    case ACT_F32: cast_pad_kernel<float><<<blocks, 256, 0, s>>>(src, (float*)dst, rows, c_src, c_dst); break;
    case ACT_F16: cast_pad_kernel<__half><<<blocks, 256, 0, s>>>(src, (__half*)dst, rows, c_src, c_dst); break;
    case ACT_BF16: cast_pad_kernel<__nv_bfloat16><<<blocks, 256, 0, s>>>(src, (__nv_bfloat16*)dst, rows, c_src, c_dst); break;
  }
  return cudaGetLastError();
}

}  // namespace hfg
