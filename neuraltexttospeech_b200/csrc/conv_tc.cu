// Persistent, warp-specialised tcgen05 / TMEM implicit-GEMM kernel for every 16-bit conv layer of
// the HiFi-GAN generator (ResBlock1/2 dilated Conv1D chains, polyphase Conv1DTranspose, conv_pre):
// reference HiFiGAN_TF/model.py:71-80, 106-112, 156-167.  sm_100a only.
//
// One launch = one layer.  The layer is cut into tiles (utterance b, MB*128 output rows, output
// phase "variant", N output channels); CTAs are persistent and walk the tile list with a static
// stride.  Per tile
//     D[q, co] = sum_j sum_ci lrelu(x[b, q + tap_off[j], ci]) * W[j][ci][co]
// runs on the tensor cores as sum over (K-slice, tap) of [128 x KS] x [KS x N] MMAs:
//
//   loader warps 0-3   global -> regs -> leaky-relu -> smem, one K-slice (KS input channels) of the
//                      activation slab [MB*128 + span rows] at a time into an NSL-deep ring, in the
//                      no-swizzle K-major core-matrix layout  smem[(chunk * R_pad + row) * 16 B]
//                      (chunk = 8 channels).  Tap j of the dilated conv is the SAME slab addressed at
//                      row offset tap_off[j]-min_off: no im2col, no per-tap reload; rows outside the
//                      utterance are zero = TF's zero padding (applied after the activation, lrelu(0)=0).
//   producer warp 4    streams host-prepacked weight stages [KBW/8 chunks][N rows][16 B] through an
//                      NST-deep ring with cp.async.bulk + mbarrier complete_tx.
//   MMA warp 5         one thread issues tcgen05.mma (cta_group::1, kind::f16, M=128, N, K=16) into
//                      TMEM accumulators (NACC buffers of MB x N columns); tcgen05.commit releases
//                      weight stages / slab slices and hands the accumulator to the epilogue.
//   epilogue warps 6-9 tcgen05.ld (one TMEM lane quarter each, thread = output row) -> +bias
//                      -> +residual (model.py:79,111) -> +MRF partial sum, *1/num_kernels
//                      (model.py:161-167) -> 16-bit -> the thread's private staging row in smem
//                      -> cp.async.bulk smem->global of that row (full-line writes, no cross-thread sync).
//
// Loads of tile i+1, MMAs of tile i and the epilogue of tile i-1 overlap (slab ring + TMEM double
// buffering where 2*MB*N <= 512 columns).
#include <cstdio>
#include <cstring>
#include <type_traits>
#include <vector>

#include "kernels.cuh"

namespace hfg {

namespace {

constexpr int kTcThreads = 320;
constexpr int kTcMaxSpan = 96;       // largest max_off - min_off supported (V1: 50, V3: 72)
constexpr size_t kTcMaxSmem = 227 * 1024;

__host__ __device__ constexpr uint32_t tc_tmem_cols(int need) { return need <= 32 ? 32 : need <= 64 ? 64 : need <= 128 ? 128 : need <= 256 ? 256 : 512; }

__host__ __device__ inline int tc_rows_padded(int rows, int chunks) {
  // conflict-free 16-byte staging stores: R_pad mod 8 == 1 when a quarter-warp spans 8 chunks of one
  // row, 2 / 4 when a row has only 4 / 2 chunks
  const int want = chunks >= 8 ? 1 : (chunks >= 2 ? 8 / chunks : 0);
  int r = rows;
  while (r % 8 != want) ++r;
  return r;
}

// ------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("{ .reg .b64 st; mbarrier.arrive.shared::cta.b64 st, [%0]; }" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("{ .reg .b64 st; mbarrier.arrive.expect_tx.shared::cta.b64 st, [%0], %1; }" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// Bounded wait: a protocol bug traps (launch failure) instead of hanging the GPU.
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t done;
  asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
               : "=r"(done) : "r"(addr), "r"(parity) : "memory");
  if (done) return;
  const long long t0 = clock64();
  for (;;) {
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(done) : "r"(addr), "r"(parity) : "memory");
    if (done) return;
    if (clock64() - t0 > 2000000000LL) {
      printf("hfg conv_tc: mbarrier timeout (block %d thread %d bar %u parity %u)\n", blockIdx.x, threadIdx.x, addr, parity);
      __trap();
    }
  }
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_s2g(void* dst_gmem, const void* src_smem, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(dst_gmem), "r"(smem_u32(src_smem)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group 0;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void fence_mbar_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc(uint32_t* slot_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(slot_smem)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{ .reg .pred p; setp.ne.b32 p, %4, 0; tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p; }"
               ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
        "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
        "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&r)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
        "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
      : "r"(taddr) : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// smem matrix descriptor, no swizzle, K-major (cute::UMMA::SmemDescriptor bit layout):
//  [0,14) start>>4 | [16,30) LBO>>4 (stride between the two 16-byte K chunks of one MMA)
//  [32,46) SBO>>4 (stride between 8-row groups) | [46,48) version=1 | [61,64) layout=0 (none)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) | (1ull << 46);
}
// instruction descriptor, kind::f16 (cute::UMMA::InstrDescriptor): D=f32, A/B = f16 or bf16, both K-major, M=128
__host__ __device__ constexpr uint32_t make_idesc(int n, bool bf16) {
  return (1u << 4) | ((bf16 ? 1u : 0u) << 7) | ((bf16 ? 1u : 0u) << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}

template <typename T> struct H2;
template <> struct H2<__half> {
  static __device__ __forceinline__ float2 up(uint32_t u) { return __half22float2(*reinterpret_cast<__half2*>(&u)); }
  static __device__ __forceinline__ uint32_t down(float a, float b) { __half2 h = __floats2half2_rn(a, b); return *reinterpret_cast<uint32_t*>(&h); }
  // leaky-relu on a packed pair: max(v, v*s) for 0 < s <= 1
  static __device__ __forceinline__ uint32_t lrelu(uint32_t u, float s) {
    __half2 v = *reinterpret_cast<__half2*>(&u);
    __half2 r = __hmax2(v, __hmul2(v, __float2half2_rn(s)));
    return *reinterpret_cast<uint32_t*>(&r);
  }
};
template <> struct H2<__nv_bfloat16> {
  static __device__ __forceinline__ float2 up(uint32_t u) { return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u)); }
  static __device__ __forceinline__ uint32_t down(float a, float b) { __nv_bfloat162 h = __floats2bfloat162_rn(a, b); return *reinterpret_cast<uint32_t*>(&h); }
  static __device__ __forceinline__ uint32_t lrelu(uint32_t u, float s) {
    __nv_bfloat162 v = *reinterpret_cast<__nv_bfloat162*>(&u);
    __nv_bfloat162 r = __hmax2(v, __hmul2(v, __float2bfloat162_rn(s)));
    return *reinterpret_cast<uint32_t*>(&r);
  }
};

__device__ __forceinline__ float lrelu_f(float v, float s) { return v >= 0.f ? v : v * s; }

// ------------------------------------------------------------------ shared-memory plan
struct TcSmemPlan {
  int r_pad;            // padded slab rows
  int slice_bytes;      // one K-slice of the slab
  int stage_bytes;      // one weight stage
  int stg_pitch;        // staging row pitch in bytes
  int off_w, off_stg, off_bias, off_bar;
  int total;
};

__host__ __device__ inline TcSmemPlan tc_plan(int n, int mb, int ks, int kbw, int nst, int nsl, int pw, int span) {
  TcSmemPlan p;
  p.r_pad = tc_rows_padded(mb * 128 + span, ks / 8);
  p.slice_bytes = (ks / 8) * p.r_pad * 16;
  p.stage_bytes = n * kbw * 2;
  p.stg_pitch = pw * 2 + 16;
  int off = nsl * p.slice_bytes;
  off = (off + 127) & ~127;
  p.off_w = off; off += nst * p.stage_bytes;
  p.off_stg = off; off += 2 * 128 * p.stg_pitch;
  p.off_bias = off; off += n * 4;
  off = (off + 15) & ~15;
  p.off_bar = off; off += (2 * nsl + 2 * nst + 4) * 8 + 16;
  p.total = off;
  return p;
}

struct TcKArgs {
  TcArgs c;
  int min_off, span;
  int tiles_q, n_tiles, total_tiles;
};

// ------------------------------------------------------------------ the kernel
template <typename T, int C_IN, int N, int MB, int KS, int KBW, int NST, int NSL, int PW>
__global__ void __launch_bounds__(kTcThreads, 1) conv_tc_kernel(const __grid_constant__ TcKArgs ka) {
  constexpr int NKS = C_IN / KS;                  // K-slices per tile
  constexpr int NKB = KS / KBW;                   // weight stages per (slice, tap)
  constexpr int CHS = KS / 8;                     // 16-byte chunks per slab row and slice
  constexpr int NACC = (2 * MB * N <= 512) ? 2 : 1;
  constexpr int ACC_COLS = MB * N;
  constexpr uint32_t TMEM_COLS = tc_tmem_cols(NACC * ACC_COLS);
  constexpr uint32_t IDESC = make_idesc(N, std::is_same<T, __nv_bfloat16>::value);
  static_assert(C_IN % KS == 0 && KS % KBW == 0 && KBW % 16 == 0 && N % 16 == 0 && N <= 256, "tile config");
  static_assert(PW % 16 == 0 && N % PW == 0 && PW <= 128, "staging pass width");
  const TcArgs& a = ka.c;
  const TcSmemPlan pl = tc_plan(N, MB, KS, KBW, NST, NSL, PW, ka.span);

  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* slab = smem;
  uint8_t* wst = smem + pl.off_w;
  uint8_t* stg = smem + pl.off_stg;
  float* bias_s = reinterpret_cast<float*>(smem + pl.off_bias);
  uint64_t* slice_full = reinterpret_cast<uint64_t*>(smem + pl.off_bar);
  uint64_t* slice_empty = slice_full + NSL;
  uint64_t* w_full = slice_empty + NSL;
  uint64_t* w_empty = w_full + NST;
  uint64_t* acc_full = w_empty + NST;
  uint64_t* acc_empty = acc_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    for (int s = 0; s < NSL; ++s) { mbar_init(&slice_full[s], 128); mbar_init(&slice_empty[s], 1); }
    for (int s = 0; s < NST; ++s) { mbar_init(&w_full[s], 1); mbar_init(&w_empty[s], 1); }
    for (int s = 0; s < 2; ++s) { mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], 128); }
    fence_mbar_init();
  }
  if (warp == 5) tmem_alloc(tmem_slot, TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // tile index -> (utterance b, row tile qt, variant v, channel tile nt); nt fastest so that CTAs that
  // run concurrently share the activation rows in L2
  auto decode = [&](int t, int& b, int& qt, int& v, int& nt) {
    nt = t % ka.n_tiles; t /= ka.n_tiles;
    v = t % a.nvar; t /= a.nvar;
    qt = t % ka.tiles_q; b = t / ka.tiles_q;
  };

  if (warp < 4) {
    // ===================== loaders: activation slab slices =====================
    const float slope = a.in_slope;
    const int rows = MB * 128 + ka.span;
    const int pieces = rows * CHS;
    uint32_t scnt = 0;
    for (int t = blockIdx.x; t < ka.total_tiles; t += gridDim.x) {
      int b, qt, v, nt; decode(t, b, qt, v, nt);
      const int row_base = qt * (MB * 128) + ka.min_off;
      const T* __restrict__ xb = reinterpret_cast<const T*>(a.x) + (size_t)b * a.x_rows * C_IN;
#pragma unroll 1
      for (int s = 0; s < NKS; ++s, ++scnt) {
        const int pos = scnt % NSL;
        mbar_wait(&slice_empty[pos], ((scnt / NSL) & 1) ^ 1);
        uint8_t* dst = slab + pos * pl.slice_bytes;
        const T* __restrict__ xs = xb + s * KS;
        constexpr int U = 8;
#pragma unroll 1
        for (int i0 = tid; i0 < pieces; i0 += 128 * U) {
          uint4 v4[U];
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int idx = i0 + u * 128;
            const int c = idx % CHS, r = idx / CHS;
            const int row = row_base + r;
            v4[u] = make_uint4(0u, 0u, 0u, 0u);
            if (idx < pieces && row >= 0 && row < a.x_rows)
              v4[u] = __ldg(reinterpret_cast<const uint4*>(xs + (size_t)row * C_IN + c * 8));
          }
#pragma unroll
          for (int u = 0; u < U; ++u) {
            const int idx = i0 + u * 128;
            if (idx < pieces) {
              const int c = idx % CHS, r = idx / CHS;
              uint4 w4 = v4[u];
              if (slope != 1.0f) {
                float2 f;
                f = H2<T>::up(w4.x); w4.x = H2<T>::down(lrelu_f(f.x, slope), lrelu_f(f.y, slope));
                f = H2<T>::up(w4.y); w4.y = H2<T>::down(lrelu_f(f.x, slope), lrelu_f(f.y, slope));
                f = H2<T>::up(w4.z); w4.z = H2<T>::down(lrelu_f(f.x, slope), lrelu_f(f.y, slope));
                f = H2<T>::up(w4.w); w4.w = H2<T>::down(lrelu_f(f.x, slope), lrelu_f(f.y, slope));
              }
              *reinterpret_cast<uint4*>(dst + ((size_t)c * pl.r_pad + r) * 16) = w4;
            }
          }
        }
        fence_proxy_async();               // generic-proxy smem writes -> visible to the tensor core (async proxy)
        mbar_arrive(&slice_full[pos]);
      }
    }
  } else if (warp == 4) {
    // ===================== weight producer =====================
    if (lane == 0) {
      uint32_t wcnt = 0;
      for (int t = blockIdx.x; t < ka.total_tiles; t += gridDim.x) {
        int b, qt, v, nt; decode(t, b, qt, v, nt);
        const int n_it = NKS * a.var[v].ntaps * NKB;
        const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(a.var[v].w) + (size_t)nt * n_it * pl.stage_bytes;
        for (int it = 0; it < n_it; ++it, ++wcnt) {
          const int st = wcnt % NST;
          mbar_wait(&w_empty[st], ((wcnt / NST) & 1) ^ 1);
          mbar_expect_tx(&w_full[st], pl.stage_bytes);
          bulk_g2s(wst + st * pl.stage_bytes, wsrc + (size_t)it * pl.stage_bytes, pl.stage_bytes, &w_full[st]);
        }
      }
    }
  } else if (warp == 5) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      const uint32_t slab_addr = smem_u32(slab), wst_addr = smem_u32(wst);
      const uint32_t a_lbo = (uint32_t)pl.r_pad * 16u;
      uint32_t scnt = 0, wcnt = 0, tcnt = 0;
      for (int t = blockIdx.x; t < ka.total_tiles; t += gridDim.x, ++tcnt) {
        int b, qt, v, nt; decode(t, b, qt, v, nt);
        const TcVariant& var = a.var[v];
        const int abuf = tcnt % NACC;
        mbar_wait(&acc_empty[abuf], ((tcnt / NACC) & 1) ^ 1);
        tc_fence_after();
        const uint32_t tacc = tmem_base + abuf * ACC_COLS;
        uint32_t first = 1;
#pragma unroll 1
        for (int s = 0; s < NKS; ++s, ++scnt) {
          const int pos = scnt % NSL;
          mbar_wait(&slice_full[pos], (scnt / NSL) & 1);
          tc_fence_after();
          const uint32_t sl = slab_addr + pos * pl.slice_bytes;
#pragma unroll 1
          for (int j = 0; j < var.ntaps; ++j) {
            const uint32_t row_off = (uint32_t)(var.tap_off[j] - ka.min_off);
#pragma unroll 1
            for (int kb = 0; kb < NKB; ++kb, ++wcnt) {
              const int st = wcnt % NST;
              mbar_wait(&w_full[st], (wcnt / NST) & 1);
              tc_fence_after();
              const uint32_t wb = wst_addr + st * pl.stage_bytes;
#pragma unroll
              for (int mb = 0; mb < MB; ++mb) {
#pragma unroll
                for (int kk = 0; kk < KBW / 16; ++kk) {
                  const uint32_t chunk = (uint32_t)(kb * (KBW / 8) + kk * 2);
                  const uint64_t adesc = make_desc(sl + (chunk * pl.r_pad + mb * 128 + row_off) * 16u, a_lbo, 128u);
                  const uint64_t bdesc = make_desc(wb + (kk * 2) * N * 16, N * 16, 128u);
                  umma_f16(tacc + mb * N, adesc, bdesc, IDESC, (first && kk == 0) ? 0u : 1u);
                }
              }
              first = 0;
              umma_commit(&w_empty[st]);          // frees the weight stage when these MMAs retire
            }
          }
          umma_commit(&slice_empty[pos]);         // slab slice may be overwritten
        }
        umma_commit(&acc_full[abuf]);             // accumulators of this tile complete
      }
    }
  } else {
    // ===================== epilogue warps 6-9 =====================
    const int quarter = warp & 3;                       // TMEM lane quarter this warp may access
    const int er = quarter * 32 + lane;                 // row of the 128-row block owned by this thread
    const uint32_t lane_base = ((uint32_t)(quarter * 32)) << 16;
    constexpr int CW = PW >= 32 ? 32 : 16;
    uint32_t tcnt = 0, pcnt = 0;
    int bias_nt = -1;
    for (int t = blockIdx.x; t < ka.total_tiles; t += gridDim.x, ++tcnt) {
      int b, qt, v, nt; decode(t, b, qt, v, nt);
      const int n0 = nt * N;
      if (nt != bias_nt) {                              // (re)load the bias slice; epilogue-warps-only barrier
        asm volatile("bar.sync 1, 128;" ::: "memory");
        for (int i = tid - 192; i < N; i += 128) bias_s[i] = a.bias[n0 + i];
        asm volatile("bar.sync 1, 128;" ::: "memory");
        bias_nt = nt;
      }
      const int abuf = tcnt % NACC;
      mbar_wait(&acc_full[abuf], (tcnt / NACC) & 1);
      tc_fence_after();
      const uint32_t tacc = tmem_base + lane_base + abuf * ACC_COLS;
      const int out_off = a.var[v].out_off;
#pragma unroll 1
      for (int mb = 0; mb < MB; ++mb) {
        const int q = qt * (MB * 128) + mb * 128 + er;
        const bool valid = q < a.n_q;
        const size_t row_idx = ((size_t)b * a.out_rows + (size_t)(valid ? q : 0) * a.out_stride + out_off) * a.c_out + n0;
#pragma unroll 1
        for (int p0 = 0; p0 < N; p0 += PW, ++pcnt) {
          uint8_t* my = stg + ((pcnt & 1) * 128 + er) * pl.stg_pitch;
          bulk_wait_read<1>();                          // the store that last used this staging row has read it
#pragma unroll 1
          for (int c0 = 0; c0 < PW; c0 += CW) {
            uint32_t r[32];
            if (CW == 32) tmem_ld32(tacc + mb * N + p0 + c0, r);
            else tmem_ld16(tacc + mb * N + p0 + c0, r);
            float vv[CW];
            const size_t idx = row_idx + p0 + c0;
            uint4 rr[CW / 8], rr2[CW / 8];
            const bool has_res = a.res != nullptr && valid, has_res2 = a.res2 != nullptr && valid;
            if (has_res) {
              const uint4* rp = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(a.res) + idx);
#pragma unroll
              for (int i = 0; i < CW / 8; ++i) rr[i] = __ldg(rp + i);
            }
            if (has_res2) {
              const uint4* rp = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(a.res2) + idx);
#pragma unroll
              for (int i = 0; i < CW / 8; ++i) rr2[i] = __ldg(rp + i);
            }
            tmem_ld_wait();
#pragma unroll
            for (int i = 0; i < CW; ++i) vv[i] = __uint_as_float(r[i]) + bias_s[p0 + c0 + i];
            if (has_res) {
#pragma unroll
              for (int i = 0; i < CW / 8; ++i) {
                float2 f;
                f = H2<T>::up(rr[i].x); vv[i * 8 + 0] += f.x; vv[i * 8 + 1] += f.y;
                f = H2<T>::up(rr[i].y); vv[i * 8 + 2] += f.x; vv[i * 8 + 3] += f.y;
                f = H2<T>::up(rr[i].z); vv[i * 8 + 4] += f.x; vv[i * 8 + 5] += f.y;
                f = H2<T>::up(rr[i].w); vv[i * 8 + 6] += f.x; vv[i * 8 + 7] += f.y;
              }
            }
            if (has_res2) {
#pragma unroll
              for (int i = 0; i < CW / 8; ++i) {
                float2 f;
                f = H2<T>::up(rr2[i].x); vv[i * 8 + 0] += f.x; vv[i * 8 + 1] += f.y;
                f = H2<T>::up(rr2[i].y); vv[i * 8 + 2] += f.x; vv[i * 8 + 3] += f.y;
                f = H2<T>::up(rr2[i].z); vv[i * 8 + 4] += f.x; vv[i * 8 + 5] += f.y;
                f = H2<T>::up(rr2[i].w); vv[i * 8 + 6] += f.x; vv[i * 8 + 7] += f.y;
              }
            }
            if (a.out_scale != 1.0f) {
#pragma unroll
              for (int i = 0; i < CW; ++i) vv[i] *= a.out_scale;
            }
            if (a.out_slope != 1.0f) {
#pragma unroll
              for (int i = 0; i < CW; ++i) vv[i] = lrelu_f(vv[i], a.out_slope);
            }
#pragma unroll
            for (int i = 0; i < CW / 8; ++i) {
              uint4 u;
              u.x = H2<T>::down(vv[i * 8 + 0], vv[i * 8 + 1]); u.y = H2<T>::down(vv[i * 8 + 2], vv[i * 8 + 3]);
              u.z = H2<T>::down(vv[i * 8 + 4], vv[i * 8 + 5]); u.w = H2<T>::down(vv[i * 8 + 6], vv[i * 8 + 7]);
              *reinterpret_cast<uint4*>(my + (c0 + i * 8) * 2) = u;
            }
          }
          fence_proxy_async();                          // my staging row -> visible to the bulk-copy engine
          if (valid) bulk_s2g(reinterpret_cast<T*>(a.out) + row_idx + p0, my, PW * 2);
          bulk_commit();
        }
      }
      tc_fence_before();
      mbar_arrive(&acc_empty[abuf]);                    // accumulator buffer may be overwritten
    }
    bulk_wait_all();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) { __syncwarp(); tc_fence_after(); tmem_dealloc(tmem_base, TMEM_COLS); }
}

// ------------------------------------------------------------------ host side
int g_num_sms = 0;

template <typename T, int C_IN, int N, int MB, int KS, int KBW, int NST, int NSL, int PW>
cudaError_t launch_t(const TcArgs& a, cudaStream_t s, int* grid_out) {
  TcKArgs ka;
  ka.c = a;
  int mn = a.var[0].tap_off[0], mx = mn;
  for (int v = 0; v < a.nvar; ++v)
    for (int j = 0; j < a.var[v].ntaps; ++j) { mn = a.var[v].tap_off[j] < mn ? a.var[v].tap_off[j] : mn; mx = a.var[v].tap_off[j] > mx ? a.var[v].tap_off[j] : mx; }
  ka.min_off = mn; ka.span = mx - mn;
  if (ka.span > kTcMaxSpan) return cudaErrorInvalidValue;
  const TcSmemPlan pl = tc_plan(N, MB, KS, KBW, NST, NSL, PW, ka.span);
  if ((size_t)pl.total > kTcMaxSmem) return cudaErrorInvalidConfiguration;
  ka.tiles_q = (a.n_q + MB * 128 - 1) / (MB * 128);
  ka.n_tiles = a.c_out / N;
  ka.total_tiles = a.B * ka.tiles_q * a.nvar * ka.n_tiles;
  auto kern = conv_tc_kernel<T, C_IN, N, MB, KS, KBW, NST, NSL, PW>;
  static int configured = 0, occ = 1;      // per instantiation
  if (pl.total > configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, pl.total);
    if (e != cudaSuccess) return e;
    configured = pl.total;
    if (!g_num_sms) {
      int dev = 0; cudaGetDevice(&dev);
      cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
      if (g_num_sms <= 0) g_num_sms = 148;
    }
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kTcThreads, pl.total);
    if (e != cudaSuccess) return e;
    constexpr int NACC = (2 * MB * N <= 512) ? 2 : 1;
    const int tmem_occ = 512 / (int)tc_tmem_cols(NACC * MB * N);
    if (occ > tmem_occ) occ = tmem_occ;
    if (occ < 1) occ = 1;
  }
  int grid = g_num_sms * occ;
  if (grid > ka.total_tiles) grid = ka.total_tiles;
  if (grid_out) *grid_out = grid;
  kern<<<grid, kTcThreads, pl.total, s>>>(ka);
  return cudaGetLastError();
}

struct TcRow { int c_in, c_out_tile; TcConfig cfg; };
//                 c_in  n     {n,  mb, ks, kbw, nst, nsl, pw}
const TcRow kTcTable[] = {
    {512, 256, {256, 1, 64, 32, 4, 3, 64}},
    {256, 256, {256, 2, 64, 32, 4, 3, 64}},
    {128, 256, {256, 1, 64, 32, 4, 2, 64}},
    {256, 128, {128, 2, 64, 32, 6, 3, 64}},
    {128, 128, {128, 2, 64, 32, 6, 3, 64}},
    {128, 64, {64, 4, 64, 64, 4, 2, 64}},
    {64, 64, {64, 4, 64, 64, 4, 2, 64}},
    {64, 32, {32, 4, 64, 64, 4, 2, 32}},
    {32, 32, {32, 4, 32, 32, 4, 2, 32}},
    {32, 16, {16, 4, 32, 32, 4, 2, 16}},
    {16, 16, {16, 4, 16, 16, 4, 2, 16}},
};

template <typename T>
cudaError_t launch_shape(const TcArgs& a, cudaStream_t s, int* grid_out) {
  const int ci = a.c_in, n = a.c_out > 256 ? 256 : a.c_out;
#define HFG_TC(CI, NN, MB, KS, KBW, NST, NSL, PW) if (ci == CI && n == NN) return launch_t<T, CI, NN, MB, KS, KBW, NST, NSL, PW>(a, s, grid_out);
  HFG_TC(512, 256, 1, 64, 32, 4, 3, 64)
  HFG_TC(256, 256, 2, 64, 32, 4, 3, 64)
  HFG_TC(128, 256, 1, 64, 32, 4, 2, 64)
  HFG_TC(256, 128, 2, 64, 32, 6, 3, 64)
  HFG_TC(128, 128, 2, 64, 32, 6, 3, 64)
  HFG_TC(128, 64, 4, 64, 64, 4, 2, 64)
  HFG_TC(64, 64, 4, 64, 64, 4, 2, 64)
  HFG_TC(64, 32, 4, 64, 64, 4, 2, 32)
  HFG_TC(32, 32, 4, 32, 32, 4, 2, 32)
  HFG_TC(32, 16, 4, 32, 32, 4, 2, 16)
  HFG_TC(16, 16, 4, 16, 16, 4, 2, 16)
#undef HFG_TC
  return cudaErrorInvalidValue;
}

}  // namespace

bool tc_config(int c_in, int c_out, TcConfig* cfg) {
  const int n = c_out > 256 ? 256 : c_out;
  if (c_out % n) return false;
  for (const TcRow& r : kTcTable)
    if (r.c_in == c_in && r.c_out_tile == n) { if (cfg) *cfg = r.cfg; return true; }
  return false;
}

bool tc_supported(int c_in, int c_out, int ntaps, int span) {
  if (ntaps < 1 || ntaps > kMaxTaps || span < 0 || span > kTcMaxSpan) return false;
  TcConfig c;
  if (!tc_config(c_in, c_out, &c)) return false;
  return (size_t)tc_plan(c.n, c.mb, c.ks, c.kbw, c.nst, c.nsl, c.pw, span).total <= kTcMaxSmem;
}

size_t tc_packed_weight_bytes(int c_in, int c_out, int ntaps) { return (size_t)ntaps * c_in * c_out * 2; }

// fp32 [ntaps][c_in][c_out] -> [n-tile][K-slice][tap][k-block] stages of [KBW/8 chunks][N rows][8 x 16-bit]
void tc_pack_weights(const float* w, int ntaps, int c_in, int c_out, ActType act, void* dst_host) {
  TcConfig c;
  if (!tc_config(c_in, c_out, &c)) return;
  uint16_t* dst = reinterpret_cast<uint16_t*>(dst_host);
  size_t o = 0;
  for (int nt = 0; nt < c_out / c.n; ++nt)
    for (int s = 0; s < c_in / c.ks; ++s)
      for (int j = 0; j < ntaps; ++j)
        for (int kb = 0; kb < c.ks / c.kbw; ++kb)
          for (int chunk = 0; chunk < c.kbw / 8; ++chunk)
            for (int row = 0; row < c.n; ++row)
              for (int e = 0; e < 8; ++e) {
                const int ci = s * c.ks + kb * c.kbw + chunk * 8 + e, co = nt * c.n + row;
                const float f = w[((size_t)j * c_in + ci) * c_out + co];
                uint16_t bits;
                if (act == ACT_BF16) { __nv_bfloat16 h = __float2bfloat16_rn(f); memcpy(&bits, &h, 2); }
                else { __half h = __float2half_rn(f); memcpy(&bits, &h, 2); }
                dst[o++] = bits;
              }
}

cudaError_t launch_conv_tc(const TcArgs& a, ActType act, cudaStream_t s, int* grid_out) {
  if (a.B <= 0 || a.n_q <= 0 || a.nvar < 1 || a.nvar > kMaxVariants) return a.nvar < 1 || a.nvar > kMaxVariants ? cudaErrorInvalidValue : cudaSuccess;
  if (act == ACT_F16) return launch_shape<__half>(a, s, grid_out);
  if (act == ACT_BF16) return launch_shape<__nv_bfloat16>(a, s, grid_out);
  return cudaErrorInvalidValue;
}

}  // namespace hfg
