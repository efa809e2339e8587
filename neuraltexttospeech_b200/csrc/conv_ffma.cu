// Direct (CUDA-core FFMA) kernels of the HiFi-GAN generator.
//
//  * conv_ffma_kernel : generic tapped conv, fp32 accumulate.  It is the whole forward pass in
//    the fp32 validation mode (north star: <=1e-5 of the fp32 reference) and the fallback for
//    shapes the tcgen05 kernel (conv_tc.cu) does not cover (V2's 16/8-channel stages).
//  * conv_post_kernel : lrelu(0.3) -> Conv1D(1, 7) -> tanh  (reference HiFiGAN_TF/model.py:169-171),
//    a pure bandwidth kernel (C_out = 1), shared by all precisions.
//  * cast_pad_kernel  : fp32 mel -> 16-bit, channel-padded, for the tensor-core conv_pre.
//
// Semantics restated from reference HiFiGAN_TF/model.py:71-80,106-112,155-173 (see kernels.cuh).
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
template <typename T>
__global__ void __launch_bounds__(256) conv_post_kernel(const PostArgs a) {
  extern __shared__ float wsm[];                       // [ntaps][c_in]
  for (int i = threadIdx.x; i < a.ntaps * a.c_in; i += blockDim.x) wsm[i] = a.w[i];
  __syncthreads();
  const long long total = (long long)a.B * a.rows;
  const long long gid = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (gid >= total) return;
  const int b = (int)(gid / a.rows), t = (int)(gid % a.rows);
  const T* __restrict__ x = reinterpret_cast<const T*>(a.x) + (size_t)b * a.rows * a.c_in;
  float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
  for (int j = 0; j < a.ntaps; ++j) {
    const int row = t + j - a.pad_left;
    if (row < 0 || row >= a.rows) continue;
    const T* xr = x + (size_t)row * a.c_in;
    const float* wj = wsm + j * a.c_in;
    for (int c = 0; c < a.c_in; c += 4) {
      const float4 v = Act4<T>::load(xr + c);
      s0 = fmaf(lrelu(v.x, a.in_slope), wj[c + 0], s0);
      s1 = fmaf(lrelu(v.y, a.in_slope), wj[c + 1], s1);
      s2 = fmaf(lrelu(v.z, a.in_slope), wj[c + 2], s2);
      s3 = fmaf(lrelu(v.w, a.in_slope), wj[c + 3], s3);
    }
  }
  a.out[gid] = tanhf((s0 + s1) + (s2 + s3) + a.bias);
}

cudaError_t launch_conv_post(const PostArgs& a, ActType act, cudaStream_t s) {
  if (a.c_in % 4 != 0 || a.ntaps < 1) return cudaErrorInvalidValue;
  const long long total = (long long)a.B * a.rows;
  if (total <= 0) return cudaSuccess;
  const unsigned blocks = (unsigned)((total + 255) / 256);
  const size_t smem = (size_t)a.ntaps * a.c_in * sizeof(float);
  switch (act) {
    case ACT_F32: conv_post_kernel<float><<<blocks, 256, smem, s>>>(a); break;
    case ACT_F16: conv_post_kernel<__half><<<blocks, 256, smem, s>>>(a); break;
    case ACT_BF16: conv_post_kernel<__nv_bfloat16><<<blocks, 256, smem, s>>>(a); break;
  }
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
