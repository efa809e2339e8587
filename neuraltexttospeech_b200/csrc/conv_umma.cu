// tcgen05 / TMEM implicit-GEMM kernel for the tapped convolutions of the HiFi-GAN generator
// (ResBlock1/2 dilated Conv1D chains, polyphase ConvTranspose1d, conv_pre) -- sm_100a only.
//
// Computes, per (batch item b, time tile), for one conv
//     D[q, co] = sum_j sum_ci act(x[b, q + tap_off[j], ci]) * W[j][ci][co]          (kernels.cuh)
// as a sum over taps of [128 x C_in] x [C_in x N] GEMMs on the 5th-gen tensor cores:
//   * A operand: ONE activation slab per CTA, rows [q0+min_off, q0+MBLK*128+max_off), staged once
//     (global -> regs -> leaky-relu -> smem) in the no-swizzle K-major core-matrix layout
//         smem[(chunk * R_pad + row) * 16 B],   chunk = 8 consecutive channels
//     so that tap j of the dilated conv is the SAME slab addressed at a row offset:
//     descriptor start += (tap_off[j]-min_off)*16 B, SBO = 128 B, LBO = R_pad*16 B.  No im2col,
//     no per-tap reload; zero rows outside the utterance reproduce TF's zero padding.
//   * B operand: weights pre-packed on the host into smem-image tiles [tap][kblock] of
//     [KB/8 chunks][N rows][16 B]; streamed through an NSTAGE ring with 1-D bulk async copies
//     (cp.async.bulk + mbarrier complete_tx) by one producer thread.
//   * D: fp32 accumulators in TMEM, MBLK blocks of 128 lanes x N columns; one thread issues
//     tcgen05.mma (cta_group::1, kind::f16, M=128, N, K=16); tcgen05.commit releases weight
//     stages and finally signals the epilogue.
//   * Epilogue (4 warps, one TMEM lane quarter each): tcgen05.ld -> +bias -> +residual ->
//     MRF sum / average -> 16-bit store.  Fuses reference HiFiGAN_TF/model.py:79 (x = xt + x),
//     :111, :161-167 (xs += ...; x = xs / num_kernels); the leaky-relu of model.py:73,76,108,159
//     is fused into the consumer's slab staging.
#include <cstdio>
#include <cstring>
#include <type_traits>
#include <vector>

#include "kernels.cuh"

namespace hfg {

namespace {

// ------------------------------------------------------------------ tile configuration
__host__ __device__ constexpr int kKBlk(int c_in) { return c_in >= 64 ? 64 : c_in; }           // channels per weight stage
__host__ __device__ constexpr int kNTile(int c_out) { return c_out > 256 ? 256 : c_out; }
__host__ __device__ constexpr int kMBlk(int c_in, int n) {
  return c_in >= 512 ? 1 : (n >= 256 ? 2 : (n >= 128 ? 2 : 4));
}
__host__ __device__ constexpr int kStages(int c_in, int n) { return n >= 128 ? 2 : 4; }
constexpr int kThreads = 192;     // warps 0-3: slab staging + epilogue, warp 4: weight producer, warp 5: MMA + TMEM alloc
constexpr int kMaxSpan = 96;      // largest max_off - min_off supported (V1: 50, V3: 72)
constexpr size_t kMaxSmem = 227 * 1024;

__host__ __device__ constexpr uint32_t tmem_cols(int need) { return need <= 32 ? 32 : need <= 64 ? 64 : need <= 128 ? 128 : need <= 256 ? 256 : 512; }

__host__ __device__ inline int slab_rows_padded(int rows, int chunks) {
  // bank-conflict-free staging stores: R_pad mod 8 == 1 when a quarter-warp spans 8 chunks,
  // 2 / 4 when a row has only 4 / 2 chunks
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
  const long long t0 = clock64();
  for (;;) {
    uint32_t done;
    asm volatile("{ .reg .pred p; mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2; selp.u32 %0, 1, 0, p; }"
                 : "=r"(done) : "r"(addr), "r"(parity) : "memory");
    if (done) return;
    if (clock64() - t0 > 4000000000LL) { printf("hfg conv_umma: mbarrier timeout (block %d,%d,%d thread %d)\n", blockIdx.x, blockIdx.y, blockIdx.z, threadIdx.x); __trap(); }
  }
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
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

// smem matrix descriptor, no swizzle, K-major (cute::UMMA::SmemDescriptor layout):
//  [0,14) start>>4 | [16,30) LBO>>4 (stride between the two 16-byte K chunks of one MMA)
//  [32,46) SBO>>4 (stride between 8-row groups) | [46,48) version=1 | [61,64) layout=0 (none)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) | (1ull << 46);
}
// instruction descriptor, kind::f16 (cute::UMMA::InstrDescriptor): D=f32, A/B = f16 or bf16, both K-major
__host__ __device__ constexpr uint32_t make_idesc(int n, bool bf16) {
  return (1u << 4) | ((bf16 ? 1u : 0u) << 7) | ((bf16 ? 1u : 0u) << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}

template <typename T> struct H2;
template <> struct H2<__half> {
  static __device__ __forceinline__ float2 up(uint32_t u) { return __half22float2(*reinterpret_cast<__half2*>(&u)); }
  static __device__ __forceinline__ uint32_t down(float a, float b) { __half2 h = __floats2half2_rn(a, b); return *reinterpret_cast<uint32_t*>(&h); }
};
template <> struct H2<__nv_bfloat16> {
  static __device__ __forceinline__ float2 up(uint32_t u) { return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u)); }
  static __device__ __forceinline__ uint32_t down(float a, float b) { __nv_bfloat162 h = __floats2bfloat162_rn(a, b); return *reinterpret_cast<uint32_t*>(&h); }
};

__device__ __forceinline__ float lrelu_f(float v, float s) { return v >= 0.f ? v : v * s; }

struct UmmaArgs {
  ConvArgs c;
  int min_off, span, r_pad, n_iters;
};

// ------------------------------------------------------------------ the kernel
template <typename T, int C_IN, int N, int MBLK, int NSTAGE>
__global__ void __launch_bounds__(kThreads) conv_umma_kernel(const UmmaArgs ua) {
  constexpr int KB = kKBlk(C_IN);                 // channels per weight stage
  constexpr int NKB = C_IN / KB;                  // weight stages per tap
  constexpr int CH = C_IN / 8;                    // 16-byte chunks per slab row
  constexpr int STAGE_BYTES = N * KB * 2;
  constexpr uint32_t TMEM_COLS = tmem_cols(MBLK * N);
  constexpr uint32_t IDESC = make_idesc(N, sizeof(T) == 2 && !std::is_same<T, __half>::value);
  const ConvArgs& a = ua.c;

  extern __shared__ __align__(1024) uint8_t smem[];
  // layout: [weight stages][slab][bias][barriers][tmem slot]
  uint8_t* wst = smem;
  uint8_t* slab = smem + NSTAGE * STAGE_BYTES;
  const int slab_bytes = CH * ua.r_pad * 16;
  float* bias_s = reinterpret_cast<float*>(slab + slab_bytes);
  uint64_t* full = reinterpret_cast<uint64_t*>(bias_s + N);
  uint64_t* empty = full + NSTAGE;
  uint64_t* slab_ready = empty + NSTAGE;
  uint64_t* acc_ready = slab_ready + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_ready + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int q0 = blockIdx.x * (MBLK * 128);
  const int co0 = blockIdx.y * N;
  const int b = blockIdx.z;

  if (tid == 0) {
    for (int s = 0; s < NSTAGE; ++s) { mbar_init(&full[s], 1); mbar_init(&empty[s], 1); }
    mbar_init(slab_ready, 128);
    mbar_init(acc_ready, 1);
    fence_mbar_init();
  }
  if (warp == 5) tmem_alloc(tmem_slot, TMEM_COLS);
  if (tid < N) bias_s[tid] = a.bias[co0 + tid];
  if (N > kThreads && tid + kThreads < N) bias_s[tid + kThreads] = a.bias[co0 + tid + kThreads];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 4) {
    // ===== weight producer: one thread streams the packed tiles through the ring =====
    if (lane == 0) {
      const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(a.w) + (size_t)blockIdx.y * ua.n_iters * STAGE_BYTES;
      for (int it = 0; it < ua.n_iters; ++it) {
        const int s = it % NSTAGE;
        if (it >= NSTAGE) mbar_wait(&empty[s], ((it / NSTAGE) - 1) & 1);
        mbar_expect_tx(&full[s], STAGE_BYTES);
        bulk_g2s(wst + s * STAGE_BYTES, wsrc + (size_t)it * STAGE_BYTES, STAGE_BYTES, &full[s]);
      }
    }
  } else if (warp == 5) {
    // ===== MMA issuer =====
    if (lane == 0) {
      mbar_wait(slab_ready, 0);
      tc_fence_after();
      const uint32_t slab_addr = smem_u32(slab);
      const uint32_t wst_addr = smem_u32(wst);
      const uint32_t a_lbo = (uint32_t)ua.r_pad * 16u;
      for (int it = 0; it < ua.n_iters; ++it) {
        const int s = it % NSTAGE;
        const int j = it / NKB, kb = it % NKB;
        mbar_wait(&full[s], (it / NSTAGE) & 1);
        tc_fence_after();
        const uint32_t row_off = (uint32_t)(a.tap_off[j] - ua.min_off);
#pragma unroll
        for (int mb = 0; mb < MBLK; ++mb) {
#pragma unroll
          for (int kk = 0; kk < KB / 16; ++kk) {
            const uint32_t chunk = (uint32_t)(kb * (KB / 8) + kk * 2);
            const uint64_t adesc = make_desc(slab_addr + (chunk * ua.r_pad + mb * 128 + row_off) * 16u, a_lbo, 128u);
            const uint64_t bdesc = make_desc(wst_addr + s * STAGE_BYTES + (kk * 2) * N * 16, N * 16, 128u);
            umma_f16(tmem_base + mb * N, adesc, bdesc, IDESC, (it > 0 || kk > 0) ? 1u : 0u);
          }
        }
        umma_commit(&empty[s]);            // frees the weight stage when these MMAs retire
      }
      umma_commit(acc_ready);              // all MMAs retired -> accumulators complete
    }
  } else {
    // ===== warps 0-3: stage the activation slab, then run the epilogue =====
    {
      const T* __restrict__ x = reinterpret_cast<const T*>(a.x) + (size_t)b * a.x_rows * C_IN;
      const int rows = MBLK * 128 + ua.span;
      const int total = rows * CH;
      const int row_base = q0 + ua.min_off;
      const float slope = a.in_slope;
      for (int idx = tid; idx < total; idx += 128) {
        const int c = idx % CH, r = idx / CH;
        const int row = row_base + r;
        uint4 v = make_uint4(0u, 0u, 0u, 0u);
        if (row >= 0 && row < a.x_rows) {
          v = __ldg(reinterpret_cast<const uint4*>(x + (size_t)row * C_IN + c * 8));
          if (slope != 1.0f) {
            float2 f;
            f = H2<T>::up(v.x); v.x = H2<T>::down(lrelu_f(f.x, slope), lrelu_f(f.y, slope));
            f = H2<T>::up(v.y); v.y = H2<T>::down(lrelu_f(f.x, slope), lrelu_f(f.y, slope));
            f = H2<T>::up(v.z); v.z = H2<T>::down(lrelu_f(f.x, slope), lrelu_f(f.y, slope));
            f = H2<T>::up(v.w); v.w = H2<T>::down(lrelu_f(f.x, slope), lrelu_f(f.y, slope));
          }
        }
        *reinterpret_cast<uint4*>(slab + ((size_t)c * ua.r_pad + r) * 16) = v;
      }
      fence_proxy_async();                 // generic-proxy smem writes -> visible to the tensor core (async proxy)
      mbar_arrive(slab_ready);
    }
    mbar_wait(acc_ready, 0);
    __syncwarp();
    tc_fence_after();
    const uint32_t lane_base = ((uint32_t)(warp * 32)) << 16;
#pragma unroll 1
    for (int mb = 0; mb < MBLK; ++mb) {
      const int q = q0 + mb * 128 + warp * 32 + lane;
      const bool valid = q < a.n_q;
      const size_t row_idx = ((size_t)b * a.out_rows + (size_t)q * a.out_stride + a.out_off) * a.c_out + co0;
      constexpr int CW = N >= 32 ? 32 : 16;
#pragma unroll 1
      for (int c0 = 0; c0 < N; c0 += CW) {
        uint32_t r[32];
        __syncwarp();                      // tcgen05.ld is warp-collective (.sync.aligned)
        if (CW == 32) tmem_ld32(tmem_base + lane_base + mb * N + c0, r);
        else tmem_ld16(tmem_base + lane_base + mb * N + c0, r);
        tmem_ld_wait();
        if (!valid) continue;
        float v[CW];
#pragma unroll
        for (int i = 0; i < CW; ++i) v[i] = __uint_as_float(r[i]) + bias_s[c0 + i];
        const size_t idx = row_idx + c0;
        if (a.res) {
          const uint4* rp = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(a.res) + idx);
#pragma unroll
          for (int i = 0; i < CW / 8; ++i) {
            const uint4 u = rp[i];
            float2 f;
            f = H2<T>::up(u.x); v[i * 8 + 0] += f.x; v[i * 8 + 1] += f.y;
            f = H2<T>::up(u.y); v[i * 8 + 2] += f.x; v[i * 8 + 3] += f.y;
            f = H2<T>::up(u.z); v[i * 8 + 4] += f.x; v[i * 8 + 5] += f.y;
            f = H2<T>::up(u.w); v[i * 8 + 6] += f.x; v[i * 8 + 7] += f.y;
          }
        }
        if (a.mrf_mode == MRF_FIRST) {
          float4* ap = reinterpret_cast<float4*>(a.acc + idx);
#pragma unroll
          for (int i = 0; i < CW / 4; ++i) ap[i] = make_float4(v[i * 4], v[i * 4 + 1], v[i * 4 + 2], v[i * 4 + 3]);
        } else if (a.mrf_mode == MRF_ADD) {
          float4* ap = reinterpret_cast<float4*>(a.acc + idx);
#pragma unroll
          for (int i = 0; i < CW / 4; ++i) {
            float4 s = ap[i];
            s.x += v[i * 4]; s.y += v[i * 4 + 1]; s.z += v[i * 4 + 2]; s.w += v[i * 4 + 3];
            ap[i] = s;
          }
        } else {
          if (a.mrf_mode == MRF_FINAL) {
            const float4* ap = reinterpret_cast<const float4*>(a.acc + idx);
#pragma unroll
            for (int i = 0; i < CW / 4; ++i) {
              const float4 s = ap[i];
              v[i * 4] = (s.x + v[i * 4]) / a.mrf_div; v[i * 4 + 1] = (s.y + v[i * 4 + 1]) / a.mrf_div;
              v[i * 4 + 2] = (s.z + v[i * 4 + 2]) / a.mrf_div; v[i * 4 + 3] = (s.w + v[i * 4 + 3]) / a.mrf_div;
            }
          }
          uint4* op = reinterpret_cast<uint4*>(reinterpret_cast<T*>(a.out) + idx);
#pragma unroll
          for (int i = 0; i < CW / 8; ++i) {
            uint4 u;
            u.x = H2<T>::down(v[i * 8 + 0], v[i * 8 + 1]); u.y = H2<T>::down(v[i * 8 + 2], v[i * 8 + 3]);
            u.z = H2<T>::down(v[i * 8 + 4], v[i * 8 + 5]); u.w = H2<T>::down(v[i * 8 + 6], v[i * 8 + 7]);
            op[i] = u;
          }
        }
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) { __syncwarp(); tc_fence_after(); tmem_dealloc(tmem_base, TMEM_COLS); }
}

// ------------------------------------------------------------------ host side
template <int C_IN, int N> struct Tile {
  static constexpr int MBLK = kMBlk(C_IN, N);
  static constexpr int NSTAGE = kStages(C_IN, N);
  static constexpr int KB = kKBlk(C_IN);
  static constexpr int STAGE_BYTES = N * KB * 2;
};

template <typename T, int C_IN, int N>
cudaError_t launch_t(const ConvArgs& a, cudaStream_t s) {
  using Tl = Tile<C_IN, N>;
  UmmaArgs ua;
  ua.c = a;
  int mn = a.tap_off[0], mx = a.tap_off[0];
  for (int j = 1; j < a.ntaps; ++j) { mn = a.tap_off[j] < mn ? a.tap_off[j] : mn; mx = a.tap_off[j] > mx ? a.tap_off[j] : mx; }
  ua.min_off = mn; ua.span = mx - mn;
  if (ua.span > kMaxSpan) return cudaErrorInvalidValue;
  ua.r_pad = slab_rows_padded(Tl::MBLK * 128 + ua.span, C_IN / 8);
  ua.n_iters = a.ntaps * (C_IN / Tl::KB);
  const size_t smem = (size_t)Tl::NSTAGE * Tl::STAGE_BYTES + (size_t)(C_IN / 8) * ua.r_pad * 16 + N * sizeof(float) +
                      (2 * Tl::NSTAGE + 2) * sizeof(uint64_t) + 16;
  if (smem > kMaxSmem) return cudaErrorInvalidConfiguration;
  auto kern = conv_umma_kernel<T, C_IN, N, Tl::MBLK, Tl::NSTAGE>;
  static size_t configured = 0;      // per instantiation
  if (smem > configured) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    configured = smem;
  }
  dim3 grid((a.n_q + Tl::MBLK * 128 - 1) / (Tl::MBLK * 128), a.c_out / N, a.B);
  kern<<<grid, kThreads, smem, s>>>(ua);
  return cudaGetLastError();
}

template <typename T>
cudaError_t launch_shape(const ConvArgs& a, cudaStream_t s) {
  const int ci = a.c_in, n = kNTile(a.c_out);
#define HFG_CASE(CI, NN) if (ci == CI && n == NN) return launch_t<T, CI, NN>(a, s);
  HFG_CASE(512, 256) HFG_CASE(256, 256) HFG_CASE(128, 256) HFG_CASE(256, 128) HFG_CASE(128, 128)
  HFG_CASE(128, 64) HFG_CASE(64, 64) HFG_CASE(64, 32) HFG_CASE(32, 32) HFG_CASE(32, 16) HFG_CASE(16, 16)
#undef HFG_CASE
  return cudaErrorInvalidValue;
}

}  // namespace

bool umma_supported(int c_in, int c_out, int ntaps, int span) {
  if (ntaps < 1 || ntaps > kMaxTaps || span < 0 || span > kMaxSpan) return false;
  const int n = kNTile(c_out);
  if (c_out % n) return false;
  static const int combos[][2] = {{512, 256}, {256, 256}, {128, 256}, {256, 128}, {128, 128}, {128, 64},
                                  {64, 64}, {64, 32}, {32, 32}, {32, 16}, {16, 16}};
  bool listed = false;
  for (auto& c : combos) if (c[0] == c_in && c[1] == n) listed = true;
  if (!listed) return false;          // (also keeps slab_rows_padded away from c_in < 16)
  const int mblk = kMBlk(c_in, n), nst = kStages(c_in, n), kb = kKBlk(c_in);
  const size_t smem = (size_t)nst * n * kb * 2 + (size_t)(c_in / 8) * slab_rows_padded(mblk * 128 + span, c_in / 8) * 16 +
                      n * sizeof(float) + (2 * nst + 2) * sizeof(uint64_t) + 16;
  return smem <= kMaxSmem;
}

size_t umma_packed_weight_bytes(int c_in, int c_out, int ntaps) { return (size_t)ntaps * c_in * c_out * 2; }

// fp32 [ntaps][c_in][c_out] -> [n-split y][tap j][kblock kb] tiles of [KB/8 chunks][N rows][8 x 16-bit]
void umma_pack_weights(const float* w, int ntaps, int c_in, int c_out, ActType act, void* dst_host) {
  const int n = kNTile(c_out), kb_sz = kKBlk(c_in), nkb = c_in / kb_sz;
  uint16_t* dst = reinterpret_cast<uint16_t*>(dst_host);
  size_t o = 0;
  for (int y = 0; y < c_out / n; ++y)
    for (int j = 0; j < ntaps; ++j)
      for (int kb = 0; kb < nkb; ++kb)
        for (int chunk = 0; chunk < kb_sz / 8; ++chunk)
          for (int row = 0; row < n; ++row)
            for (int e = 0; e < 8; ++e) {
              const int ci = kb * kb_sz + chunk * 8 + e, co = y * n + row;
              const float f = w[((size_t)j * c_in + ci) * c_out + co];
              uint16_t bits;
              if (act == ACT_BF16) { __nv_bfloat16 h = __float2bfloat16_rn(f); memcpy(&bits, &h, 2); }
              else { __half h = __float2half_rn(f); memcpy(&bits, &h, 2); }
              dst[o++] = bits;
            }
}

cudaError_t launch_conv_umma(const ConvArgs& a, ActType act, cudaStream_t s) {
  if (a.B <= 0 || a.n_q <= 0) return cudaSuccess;
  if (act == ACT_F16) return launch_shape<__half>(a, s);
  if (act == ACT_BF16) return launch_shape<__nv_bfloat16>(a, s);
  return cudaErrorInvalidValue;
}

}  // namespace hfg
