// hfg_op_tapped_conv_noalloc: the direct (FFMA) tapped convolution of conv_ffma.cu for callers that cannot allocate
// device memory (code under CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc).  It has no
// counterpart in the reference.
//
// hfg_op_tapped_conv (engine.cu) cudaMallocs a device copy of the weights and the bias on every call and frees it after a
// synchronise.  Here that scratch is ONE file-scope `static __device__` array of kScratchBytes (4 MiB, held by every GPU
// that loads the library until the process exits):
//     [0, bias_bytes)                       bias, fp32 [c_out]            bias_bytes = c_out * 4 rounded up to 256
//     [bias_bytes, bias_bytes + w_bytes)    weights, fp32 [ntaps][c_in][c_out]
// Host code may not use the array's name as a pointer (it names a host-side stand-in): its device address is looked up
// once per device with cudaGetSymbolAddress while that device is current, and that address is what the kernel receives.
// Every call shares the array and it keeps the previous call's contents: the bytes between the bias and the weights
// are cleared on each call, everything the kernel reads is rewritten on each call, and two calls that overlap on the GPU
// (different streams without an ordering between them, or two host threads) race on it.
// The entry point calls no cudaMalloc / cudaFree / cudaHostAlloc and no allocating library routine; it enqueues two
// host-to-device copies, one clear and the kernel on the caller's stream and returns without synchronising.
#include <cuda_runtime.h>
#include <cstdarg>
#include <cstdio>
#include <mutex>

#include "../../include/hifigan_b200.h"
#include "kernels.cuh"

namespace {

constexpr size_t kScratchBytes = HFG_NOALLOC_SCRATCH_BYTES;
constexpr int kMaxDevices = 64;

static __device__ __align__(256) unsigned char g_noalloc_scratch[kScratchBytes];

// Device code may name the array directly: clears [off, off + n) (both multiples of 16) of the scratch.
__global__ void noalloc_clear_kernel(unsigned off, unsigned n) {
  uint4* p = reinterpret_cast<uint4*>(g_noalloc_scratch + off);
  for (unsigned i = blockIdx.x * blockDim.x + threadIdx.x; i < n / 16; i += gridDim.x * blockDim.x) p[i] = make_uint4(0u, 0u, 0u, 0u);
}

std::mutex g_mu;
unsigned char* g_scratch_addr[kMaxDevices];       // device address of g_noalloc_scratch per device, nullptr = not looked up yet

int report(char* err, size_t cap, int code, const char* fmt, ...) {
  if (err && cap) {
    va_list ap; va_start(ap, fmt); vsnprintf(err, cap, fmt, ap); va_end(ap);
  }
  return code;
}

// set-up, once per device, with that device current
cudaError_t scratch_address(unsigned char** out, int* dev_out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  if (dev < 0 || dev >= kMaxDevices) return cudaErrorInvalidDevice;
  *dev_out = dev;
  std::lock_guard<std::mutex> lk(g_mu);
  if (!g_scratch_addr[dev]) {
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    g_scratch_addr[dev] = static_cast<unsigned char*>(p);
  }
  *out = g_scratch_addr[dev];
  return cudaSuccess;
}

}  // namespace

extern "C" int hfg_op_tapped_conv_noalloc(int act, const void* x_dev, int B, int x_rows, int c_in, const float* w_host,
                                          const float* bias_host, int c_out, int ntaps, const int* tap_off, int out_stride,
                                          int out_off, float in_slope, const void* res_dev, float out_slope, void* out_dev,
                                          void* stream, char* err, size_t err_cap) {
  using namespace hfg;
  if (err && err_cap) err[0] = 0;
  if (act < 0 || act > 2 || !x_dev || !out_dev || !w_host || !bias_host || !tap_off || B < 0 || x_rows < 0 || out_stride < 1 ||
      out_off < 0 || out_off >= out_stride)
    return report(err, err_cap, HFG_ERR_INVALID, "hfg_op_tapped_conv_noalloc: bad argument");
  if (ntaps < 1 || ntaps > kMaxTaps)
    return report(err, err_cap, HFG_ERR_INVALID, "hfg_op_tapped_conv_noalloc: ntaps=%d outside [1, %d]", ntaps, kMaxTaps);
  if (c_in < 4 || c_in % 4 || c_out < 8 || c_out % 8)
    return report(err, err_cap, HFG_ERR_UNSUPPORTED,
                  "hfg_op_tapped_conv_noalloc: the direct kernel needs c_in %% 4 == 0 and c_out %% 8 == 0 (got %d, %d)", c_in, c_out);
  const size_t bias_bytes = ((size_t)c_out * 4 + 255) / 256 * 256;
  const size_t w_bytes = (size_t)ntaps * (size_t)c_in * (size_t)c_out * 4;
  if (bias_bytes + w_bytes > kScratchBytes)
    return report(err, err_cap, HFG_ERR_NOMEM,
                  "hfg_op_tapped_conv_noalloc: weights + bias need %zu bytes of scratch (ntaps=%d c_in=%d c_out=%d), the static "
                  "scratch holds HFG_NOALLOC_SCRATCH_BYTES = %zu; use hfg_op_tapped_conv for larger layers",
                  bias_bytes + w_bytes, ntaps, c_in, c_out, kScratchBytes);
  unsigned char* scratch = nullptr;
  int dev = -1;
  cudaError_t e = scratch_address(&scratch, &dev);
  if (e != cudaSuccess)
    return report(err, err_cap, HFG_ERR_CUDA, "hfg_op_tapped_conv_noalloc: cudaGetSymbolAddress failed: %s", cudaGetErrorString(e));
  cudaStream_t s = (cudaStream_t)stream;
  float* d_bias = reinterpret_cast<float*>(scratch);
  float* d_w = reinterpret_cast<float*>(scratch + bias_bytes);
  const size_t gap = bias_bytes - (size_t)c_out * 4;          // multiple of 16 (c_out % 8 == 0)
  e = cudaMemcpyAsync(d_bias, bias_host, (size_t)c_out * 4, cudaMemcpyHostToDevice, s);
  if (e == cudaSuccess && gap) {
    noalloc_clear_kernel<<<1, 64, 0, s>>>((unsigned)(c_out * 4), (unsigned)gap);
    e = cudaGetLastError();
  }
  if (e == cudaSuccess) e = cudaMemcpyAsync(d_w, w_host, w_bytes, cudaMemcpyHostToDevice, s);
  if (e == cudaSuccess) {
    ConvArgs a{};
    a.x = x_dev; a.out = out_dev; a.res = res_dev; a.acc = nullptr;
    a.w = d_w; a.bias = d_bias;
    a.B = B; a.x_rows = x_rows; a.n_q = x_rows; a.out_rows = x_rows * out_stride;
    a.c_in = c_in; a.c_out = c_out; a.ntaps = ntaps;
    for (int j = 0; j < ntaps; ++j) a.tap_off[j] = tap_off[j];
    a.out_stride = out_stride; a.out_off = out_off; a.in_slope = in_slope;
    a.mrf_mode = MRF_NONE; a.mrf_div = 1.0f; a.out_slope = out_slope;
    e = launch_conv_ffma(a, (ActType)act, s);
  }
  if (e != cudaSuccess)
    return report(err, err_cap, HFG_ERR_CUDA, "hfg_op_tapped_conv_noalloc failed on device %d: %s", dev, cudaGetErrorString(e));
  return HFG_OK;
}
