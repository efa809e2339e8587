// Persistent, warp-specialised, TMA-fed tcgen05 / TMEM implicit-GEMM kernel for every 16-bit conv layer
// of the HiFi-GAN generator (ResBlock1/2 dilated Conv1D chains, polyphase Conv1DTranspose, conv_pre):
// reference HiFiGAN_TF/model.py:71-80, 106-112, 156-167.  sm_100a only.
//
// Storage convention of the 16-bit path: every activation tensor in HBM is stored ALREADY ACTIVATED for
// its consumer, y = leaky_relu(x, slope) (the reference applies the activation on the consumer side,
// model.py:73,76,108,159,169).  Where the raw value is needed again -- the ResBlock skip connection
// x = xt + x, model.py:79,111 -- it is recovered exactly as x = y >= 0 ? y : y / slope.  This is what
// lets the activation slab go from HBM to the tensor core without passing through registers.
//
// One launch = one layer.  The layer is cut into tiles (utterance b, MB*128 output rows, output phase
// "variant", N output channels); CTAs are persistent and walk the tile list with a static stride.
//     D[q, co] = sum_j sum_ci y[b, q + tap_off[j], ci] * W[j][ci][co]
// runs as sum over (K-slice, tap) of [128 x KS] x [KS x N] MMAs:
//
//   warp 0  slab producer: TMA tensor loads (cp.async.bulk.tensor.3d, SWIZZLE_128B/64B/32B, out-of-bounds
//           rows zero-filled = TF's zero padding, utterances never bleed into each other) of one K-slice
//           (KS input channels) of the halo tile [MB*128 + span rows] into an NSL-deep ring.  Tap j of the
//           dilated conv is the SAME slab addressed at row offset tap_off[j]-min_off (the swizzle is a
//           function of the absolute smem address, so an unaligned descriptor start is legal): no im2col.
//   warp 1  weight producer: host-prepacked weight stages [KBW/8 chunks][N rows][16 B] through an NST-deep
//           ring with cp.async.bulk + mbarrier complete_tx.
//   warp 2  MMA issuer: one thread issues tcgen05.mma (cta_group::1, kind::f16, M=128, N, K=16) into TMEM
//           accumulators (NACC buffers of MB x N columns); tcgen05.commit releases weight stages / slab
//           slices and hands the accumulator to the epilogue.
//   warp 3  staging producer: TMA-loads the residual tile of each epilogue pass into the staging ring.
//   warps 4-7 epilogue: tcgen05.ld (one TMEM lane quarter each, thread = output row) -> +bias -> +residual
//           (model.py:79,111) -> +MRF partial sum, *1/num_kernels (model.py:161-167) -> activation for the
//           consumer -> 16-bit, in place in the swizzled staging tile -> TMA tensor store.
//
// Loads of tile i+1, MMAs of tile i and the epilogue of tile i-1 overlap (slab ring, TMEM double
// buffering where 2*MB*N <= 512 columns, NSTG-deep staging ring).
#include <cuda.h>
#include <cstdio>
#include <cstring>
#include <type_traits>
#include <vector>

#include "kernels.cuh"

namespace hfg {

namespace {

constexpr int kTcThreads = 416;       // 13 warps: 4 producers/MMA, 2 x 4 epilogue, 1 store issuer
constexpr int kTcMaxCout = 512;
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

// one lane of a converged warp (elect.sync): the pattern the compiler turns into a single predicated issue
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{ .reg .pred P; elect.sync _|P, 0xffffffff; selp.u32 %0, 1, 0, P; }" : "=r"(pred));
  return pred != 0;
}

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

__device__ __forceinline__ void tma_load_3d(void* dst_smem, const CUtensorMap* tm, int c0, int c1, int c2, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4}], [%5];"
               ::"r"(smem_u32(dst_smem)), "l"(tm), "r"(c0), "r"(c1), "r"(c2), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_load_4d(void* dst_smem, const CUtensorMap* tm, int c0, int c1, int c2, int c3, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.4d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3, %4, %5}], [%6];"
               ::"r"(smem_u32(dst_smem)), "l"(tm), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_store_4d(const CUtensorMap* tm, const void* src_smem, int c0, int c1, int c2, int c3) {
  asm volatile("cp.async.bulk.tensor.4d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4, %5}], [%1];"
               ::"l"(tm), "r"(smem_u32(src_smem)), "r"(c0), "r"(c1), "r"(c2), "r"(c3) : "memory");
}
__device__ __forceinline__ void tma_prefetch_desc(const CUtensorMap* tm) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tm) : "memory");
}

// smem matrix descriptor, no swizzle, K-major (cute::UMMA::SmemDescriptor bit layout):
//  [0,14) start>>4 | [16,30) LBO>>4 (stride between the two 16-byte K chunks of one MMA)
//  [32,46) SBO>>4 (stride between 8-row groups) | [46,48) version=1 | [61,64) layout=0 (none)
__device__ __forceinline__ uint64_t make_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((saddr & 0x3FFFF) >> 4) | ((uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32) | (1ull << 46);
}
// K-major operand in a TMA-swizzled tile whose rows are ROWB = 128 / 64 / 32 bytes: 8-row atoms of 8*ROWB bytes
// (SBO), layout type 2 / 4 / 6, base_offset 0 (verified on B200: the XOR pattern follows the absolute smem
// address, so `saddr` may sit at any row of the tile).
template <int ROWB>
__device__ __forceinline__ uint64_t make_desc_sw(uint32_t saddr) {
  constexpr uint64_t LT = ROWB == 128 ? 2 : (ROWB == 64 ? 4 : 6);
  return (uint64_t)((saddr & 0x3FFFF) >> 4) | (1ull << 16) | ((uint64_t)((8 * ROWB) >> 4) << 32) | (1ull << 46) | (LT << 61);
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
  int slice_rows;       // rows of one K-slice buffer (nbox * box_rows)
  int slice_bytes;
  int stage_bytes;      // one weight stage
  int stg_bytes;        // one staging tile [128 rows][PW * 2 bytes]
  int off_stg, off_w, off_bias, off_bar;
  int total;
};

__host__ __device__ inline TcSmemPlan tc_plan(int n, int ks, int kbw, int nst, int nsl, int nstg, int slice_rows) {
  TcSmemPlan p;
  const int pw = n < 64 ? n : 64;
  p.slice_rows = slice_rows;
  p.slice_bytes = slice_rows * ks * 2;            // multiple of the swizzle atom (slice_rows % 8 == 0)
  p.stage_bytes = n * kbw * 2;
  p.stg_bytes = 128 * pw * 2;
  int off = nsl * p.slice_bytes;
  off = (off + 1023) & ~1023;
  p.off_stg = off; off += nstg * p.stg_bytes;
  off = (off + 127) & ~127;
  p.off_w = off; off += nst * p.stage_bytes;
  p.off_bias = off; off += kTcMaxCout * 4;     // bias of every output channel of the layer
  off = (off + 15) & ~15;
  p.off_bar = off; off += (2 * nsl + 2 * nst + 3 * nstg + 4) * 8 + 16;
  p.total = off;
  return p;
}

struct TcKArgs {
  const float* bias;
  const void* res2;
  int n_q, c_out, has_res;
  float res_inv, out_scale, out_slope;
  int nvar;
  TcVariant var[kMaxVariants];
  int min_off, span;
  int box_rows, nbox;   // slab slice = nbox TMA boxes of box_rows rows
  int tiles_q, n_tiles, total_tiles;
};

// ------------------------------------------------------------------ the kernel
template <typename T, int C_IN, int N, int MB, int KS, int KBW, int TPS, int NST, int NSL, int NSTG>
__global__ void __launch_bounds__(kTcThreads, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_res,
               const __grid_constant__ CUtensorMap tm_out, const __grid_constant__ TcKArgs ka) {
  constexpr int NKS = C_IN / KS;                  // K-slices per tile
  constexpr int NKB = KS / KBW;                   // weight k-blocks per (slice, tap)
  constexpr int BLK_BYTES = N * KBW * 2;          // one (tap, k-block) weight block; a stage holds TPS of them
  static_assert(TPS == 1 || NKB == 1, "multi-tap stages need KBW == KS");
  constexpr int ROWB = KS * 2;                    // bytes per slab row within a slice (= swizzle width)
  constexpr int PW = N < 64 ? N : 64;             // output columns per epilogue pass
  constexpr int SROWB = PW * 2;                   // bytes per staging row (= swizzle width)
  constexpr int NACC = (2 * MB * N <= 512) ? 2 : 1;
  constexpr int ACC_COLS = MB * N;
  constexpr uint32_t TMEM_COLS = tc_tmem_cols(NACC * ACC_COLS);
  constexpr uint32_t IDESC = make_idesc(N, std::is_same<T, __nv_bfloat16>::value);
  static_assert(C_IN % KS == 0 && KS % KBW == 0 && KBW % 16 == 0 && N % 16 == 0 && N <= 256, "tile config");
  static_assert(KS == 64 || KS == 32 || KS == 16, "K-slice = one swizzle row");
  static_assert(N % PW == 0, "pass width");
  const TcSmemPlan pl = tc_plan(N, KS, KBW * TPS, NST, NSL, NSTG, ka.nbox * ka.box_rows);

  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* slab = smem;
  uint8_t* stg = smem + pl.off_stg;
  uint8_t* wst = smem + pl.off_w;
  float* bias_s = reinterpret_cast<float*>(smem + pl.off_bias);
  uint64_t* slice_full = reinterpret_cast<uint64_t*>(smem + pl.off_bar);
  uint64_t* slice_empty = slice_full + NSL;
  uint64_t* w_full = slice_empty + NSL;
  uint64_t* w_empty = w_full + NST;
  uint64_t* stg_full = w_empty + NST;
  uint64_t* stg_empty = stg_full + NSTG;
  uint64_t* stg_done = stg_empty + NSTG;
  uint64_t* acc_full = stg_done + NSTG;
  uint64_t* acc_empty = acc_full + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;

  if (tid == 0) {
    for (int s = 0; s < NSL; ++s) { mbar_init(&slice_full[s], 1); mbar_init(&slice_empty[s], 1); }
    for (int s = 0; s < NST; ++s) { mbar_init(&w_full[s], 1); mbar_init(&w_empty[s], 1); }
    for (int s = 0; s < NSTG; ++s) { mbar_init(&stg_full[s], 1); mbar_init(&stg_empty[s], 1); mbar_init(&stg_done[s], 128); }
    for (int s = 0; s < 2; ++s) { mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], 256); }
    fence_mbar_init();
    tma_prefetch_desc(&tm_x); tma_prefetch_desc(&tm_res); tma_prefetch_desc(&tm_out);
  }
  if (warp == 2) tmem_alloc(tmem_slot, TMEM_COLS);
  for (int i = tid; i < ka.c_out; i += kTcThreads) bias_s[i] = ka.bias[i];
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  // tile index -> (utterance b, row tile qt, variant v, channel tile nt)
  auto decode = [&](int t, int& b, int& qt, int& v, int& nt) {
    nt = t % ka.n_tiles; t /= ka.n_tiles;
    v = t % ka.nvar; t /= ka.nvar;
    qt = t % ka.tiles_q; b = t / ka.tiles_q;
  };

  // Producer / MMA warps run their loops with ALL lanes (warp-uniform control flow keeps barrier addresses and
  // descriptors in uniform registers) and issue the asynchronous instructions from one elected lane.
  if (warp == 0) {
    // ===================== slab producer (TMA) =====================
    uint32_t scnt = 0;
    const uint32_t box_bytes = (uint32_t)ka.box_rows * ROWB;
    for (int t = blockIdx.x; t < ka.total_tiles; t += gridDim.x) {
      int b, qt, v, nt; decode(t, b, qt, v, nt);
      const int row_base = qt * (MB * 128) + ka.min_off;
      for (int s = 0; s < NKS; ++s, ++scnt) {
        const int pos = scnt % NSL;
        mbar_wait(&slice_empty[pos], ((scnt / NSL) & 1) ^ 1);
        if (elect_one()) {
          mbar_expect_tx(&slice_full[pos], box_bytes * ka.nbox);
          uint8_t* dst = slab + pos * pl.slice_bytes;
          for (int bx = 0; bx < ka.nbox; ++bx)
            tma_load_3d(dst + bx * box_bytes, &tm_x, s * KS, row_base + bx * ka.box_rows, b, &slice_full[pos]);
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // ===================== weight producer =====================
    uint32_t wcnt = 0;
    for (int t = blockIdx.x; t < ka.total_tiles; t += gridDim.x) {
      int b, qt, v, nt; decode(t, b, qt, v, nt);
      const int ntaps = ka.var[v].ntaps;
      const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(ka.var[v].w) + (size_t)nt * (NKS * ntaps * NKB) * BLK_BYTES;
      for (int s = 0; s < NKS; ++s) {
        for (int j0 = 0; j0 < ntaps; j0 += TPS) {
          const int nj = (ntaps - j0) < TPS ? (ntaps - j0) : TPS;
          for (int kb = 0; kb < NKB; ++kb, ++wcnt) {
            const int st = wcnt % NST;
            const uint32_t bytes = (uint32_t)nj * BLK_BYTES;
            mbar_wait(&w_empty[st], ((wcnt / NST) & 1) ^ 1);
            if (elect_one()) {
              mbar_expect_tx(&w_full[st], bytes);
              bulk_g2s(wst + st * pl.stage_bytes, wsrc, bytes, &w_full[st]);
            }
            __syncwarp();
            wsrc += bytes;
          }
        }
      }
    }
  } else if (warp == 2) {
    // ===================== MMA issuer =====================
    const uint32_t slab_addr = smem_u32(slab), wst_addr = smem_u32(wst);
    uint32_t scnt = 0, wcnt = 0, tcnt = 0;
    for (int t = blockIdx.x; t < ka.total_tiles; t += gridDim.x, ++tcnt) {
      int b, qt, v, nt; decode(t, b, qt, v, nt);
      const int ntaps = ka.var[v].ntaps;
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
        for (int j0 = 0; j0 < ntaps; j0 += TPS) {
          const int nj = (ntaps - j0) < TPS ? (ntaps - j0) : TPS;
#pragma unroll 1
          for (int kb = 0; kb < NKB; ++kb, ++wcnt) {
            const int st = wcnt % NST;
            mbar_wait(&w_full[st], (wcnt / NST) & 1);
            tc_fence_after();
            if (elect_one()) {
              const uint32_t wb = wst_addr + st * pl.stage_bytes;
#pragma unroll 1
              for (int jj = 0; jj < nj; ++jj) {
                const uint32_t row_off = (uint32_t)(ka.var[v].tap_off[j0 + jj] - ka.min_off);
                const uint64_t a0 = make_desc_sw<ROWB>(sl + row_off * ROWB + kb * KBW * 2);
                const uint64_t b0 = make_desc(wb + jj * BLK_BYTES, N * 16, 128u);
#pragma unroll
                for (int mb = 0; mb < MB; ++mb) {
#pragma unroll
                  for (int kk = 0; kk < KBW / 16; ++kk) {
                    // descriptor start-address field counts 16-byte units: advance with plain adds
                    const uint64_t adesc = a0 + (uint64_t)((mb * 128 * ROWB + kk * 32) >> 4);
                    const uint64_t bdesc = b0 + (uint64_t)((kk * 2 * N * 16) >> 4);
                    umma_f16(tacc + mb * N, adesc, bdesc, IDESC, (first && kk == 0) ? 0u : 1u);
                  }
                }
                first = 0;
              }
              umma_commit(&w_empty[st]);          // frees the weight stage when these MMAs retire
              if (kb == NKB - 1 && j0 + TPS >= ntaps) {
                umma_commit(&slice_empty[pos]);   // slab slice may be overwritten
                if (s == NKS - 1) umma_commit(&acc_full[abuf]);   // accumulators of this tile complete
              }
            }
            first = 0;
            __syncwarp();
          }
        }
      }
    }
  } else if (warp == 3) {
    // ===================== staging producer: residual tiles (or just buffer grants) =====================
    uint32_t pcnt = 0;
    for (int t = blockIdx.x; t < ka.total_tiles; t += gridDim.x) {
      int b, qt, v, nt; decode(t, b, qt, v, nt);
      for (int mb = 0; mb < MB; ++mb) {
        const int q0 = qt * (MB * 128) + mb * 128;
        for (int p0 = 0; p0 < N; p0 += PW, ++pcnt) {
          const int buf = pcnt % NSTG;
          mbar_wait(&stg_empty[buf], ((pcnt / NSTG) & 1) ^ 1);
          if (elect_one()) {
            if (ka.has_res && q0 < ka.n_q) {
              mbar_expect_tx(&stg_full[buf], pl.stg_bytes);
              tma_load_4d(stg + buf * pl.stg_bytes, &tm_res, nt * N + p0, 0, q0, b, &stg_full[buf]);
            } else {
              mbar_arrive(&stg_full[buf]);
            }
          }
          __syncwarp();
        }
      }
    }
  } else if (warp == 12) {
    // ===================== store issuer: TMA tensor store of every finished staging tile =====================
    uint32_t pcnt = 0;
    for (int t = blockIdx.x; t < ka.total_tiles; t += gridDim.x) {
      int b, qt, v, nt; decode(t, b, qt, v, nt);
      const int out_off = ka.var[v].out_off;
      for (int mb = 0; mb < MB; ++mb) {
        const int q0 = qt * (MB * 128) + mb * 128;
        for (int p0 = 0; p0 < N; p0 += PW, ++pcnt) {
          const int buf = pcnt % NSTG;
          mbar_wait(&stg_done[buf], (pcnt / NSTG) & 1);
          if (elect_one()) {
            if (q0 < ka.n_q) {
              tma_store_4d(&tm_out, stg + buf * pl.stg_bytes, nt * N + p0, out_off, q0, b);
              bulk_commit();
              bulk_wait_read<0>();                      // the store has read the tile: the buffer may be refilled
            }
            mbar_arrive(&stg_empty[buf]);
          }
          __syncwarp();
        }
      }
    }
    if (elect_one()) bulk_wait_all();                   // same lane every time (elect.sync picks the lowest)
    __syncwarp();
  } else {
    // ===================== epilogue: two groups of 4 warps (4-7, 8-11) take alternate passes =====================
    const int grp = (warp - 4) >> 2;
    const int quarter = warp & 3;                       // TMEM lane quarter this warp may access
    const int er = quarter * 32 + lane;                 // row of the 128-row block owned by this thread
    const uint32_t lane_base = ((uint32_t)(quarter * 32)) << 16;
    constexpr int CW = PW >= 32 ? 32 : 16;
    // XOR term of the TMA swizzle for this thread's staging row (row pitch SROWB, 1024-byte aligned tile)
    const int sx = SROWB == 128 ? (er & 7) : (SROWB == 64 ? ((er >> 1) & 3) : ((er >> 2) & 1));
    uint32_t tcnt = 0, pcnt = 0;
    for (int t = blockIdx.x; t < ka.total_tiles; t += gridDim.x, ++tcnt) {
      int b, qt, v, nt; decode(t, b, qt, v, nt);
      const int n0 = nt * N;
      const int abuf = tcnt % NACC;
      mbar_wait(&acc_full[abuf], (tcnt / NACC) & 1);
      tc_fence_after();
      const uint32_t tacc = tmem_base + lane_base + abuf * ACC_COLS;
#pragma unroll 1
      for (int mb = 0; mb < MB; ++mb) {
        const int q0 = qt * (MB * 128) + mb * 128;
        const int q = q0 + er;
        const bool valid = q < ka.n_q;
        const bool has_res = ka.has_res != 0 && q0 < ka.n_q;
#pragma unroll 1
        for (int p0 = 0; p0 < N; p0 += PW, ++pcnt) {
          if ((int)(pcnt & 1) != grp) continue;
          const int buf = pcnt % NSTG;
          uint8_t* my = stg + buf * pl.stg_bytes + er * SROWB;
          mbar_wait(&stg_full[buf], (pcnt / NSTG) & 1);   // residual tile landed / buffer granted
#pragma unroll
          for (int c0 = 0; c0 < PW; c0 += CW) {
            uint32_t r[32];
            if (CW == 32) tmem_ld32(tacc + mb * N + p0 + c0, r);
            else tmem_ld16(tacc + mb * N + p0 + c0, r);
            uint4 rr[CW / 8], rr2[CW / 8];
            if (has_res) {
#pragma unroll
              for (int i = 0; i < CW / 8; ++i) rr[i] = *reinterpret_cast<const uint4*>(my + (((c0 / 8 + i) ^ sx) * 16));
            }
            const bool has_res2 = ka.res2 != nullptr && valid;
            if (has_res2) {
              const uint4* rp = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(ka.res2) +
                                                               ((size_t)b * ka.n_q + q) * ka.c_out + n0 + p0 + c0);
#pragma unroll
              for (int i = 0; i < CW / 8; ++i) rr2[i] = rp[i];
            }
            tmem_ld_wait();
            float vv[CW];
#pragma unroll
            for (int i = 0; i < CW; ++i) vv[i] = __uint_as_float(r[i]) + bias_s[n0 + p0 + c0 + i];
            if (has_res) {
              const float ri = ka.res_inv;
#pragma unroll
              for (int i = 0; i < CW / 8; ++i) {
                float2 f;
                f = H2<T>::up(rr[i].x); vv[i * 8 + 0] += lrelu_f(f.x, ri); vv[i * 8 + 1] += lrelu_f(f.y, ri);
                f = H2<T>::up(rr[i].y); vv[i * 8 + 2] += lrelu_f(f.x, ri); vv[i * 8 + 3] += lrelu_f(f.y, ri);
                f = H2<T>::up(rr[i].z); vv[i * 8 + 4] += lrelu_f(f.x, ri); vv[i * 8 + 5] += lrelu_f(f.y, ri);
                f = H2<T>::up(rr[i].w); vv[i * 8 + 6] += lrelu_f(f.x, ri); vv[i * 8 + 7] += lrelu_f(f.y, ri);
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
            const float osc = ka.out_scale, osl = ka.out_slope;
#pragma unroll
            for (int i = 0; i < CW; ++i) vv[i] = lrelu_f(vv[i] * osc, osl);
#pragma unroll
            for (int i = 0; i < CW / 8; ++i) {
              uint4 u;
              u.x = H2<T>::down(vv[i * 8 + 0], vv[i * 8 + 1]); u.y = H2<T>::down(vv[i * 8 + 2], vv[i * 8 + 3]);
              u.z = H2<T>::down(vv[i * 8 + 4], vv[i * 8 + 5]); u.w = H2<T>::down(vv[i * 8 + 6], vv[i * 8 + 7]);
              *reinterpret_cast<uint4*>(my + (((c0 / 8 + i) ^ sx) * 16)) = u;
            }
          }
          fence_proxy_async();                          // staging writes -> visible to the TMA engine
          mbar_arrive(&stg_done[buf]);                  // hand the tile to the store warp
        }
      }
      tc_fence_before();
      mbar_arrive(&acc_empty[abuf]);                    // accumulator buffer may be overwritten
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) { __syncwarp(); tc_fence_after(); tmem_dealloc(tmem_base, TMEM_COLS); }
}

// ------------------------------------------------------------------ host side
int g_num_sms = 0;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

CUtensorMapSwizzle swizzle_for(int inner_bytes) {
  return inner_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B : (inner_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
}

// rank-3 map over activations [B][rows][C] (16-bit), box = [box_c channels][box_rows rows][1]
bool make_map3(CUtensorMap* m, ActType act, const void* base, int C, int rows, int B, int box_c, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  cuuint64_t gdim[3] = {(cuuint64_t)C, (cuuint64_t)rows, (cuuint64_t)B};
  cuuint64_t gstr[2] = {(cuuint64_t)C * 2, (cuuint64_t)C * 2 * (cuuint64_t)rows};
  cuuint32_t box[3] = {(cuuint32_t)box_c, (cuuint32_t)box_rows, 1};
  cuuint32_t est[3] = {1, 1, 1};
  return fn(m, act == ACT_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(base), gdim, gstr,
            box, est, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(box_c * 2), CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// rank-4 map over outputs viewed as [B][q][u][C] (row q*u + phase), box = [box_c][1][128][1]
bool make_map4(CUtensorMap* m, ActType act, const void* base, int C, int u, int nq, int B, int box_c) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)u, (cuuint64_t)nq, (cuuint64_t)B};
  cuuint64_t gstr[3] = {(cuuint64_t)C * 2, (cuuint64_t)C * 2 * u, (cuuint64_t)C * 2 * u * (cuuint64_t)nq};
  cuuint32_t box[4] = {(cuuint32_t)box_c, 1, 128, 1};
  cuuint32_t est[4] = {1, 1, 1, 1};
  return fn(m, act == ACT_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, const_cast<void*>(base), gdim, gstr,
            box, est, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(box_c * 2), CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// slab slice = nbox boxes of box_rows rows covering MB*128 + span rows with little waste
void tc_boxes(int rows, int* box_rows, int* nbox) {
  int best_br = 256, best_pad = 1 << 30;
  for (int br = 256; br >= 8; br >>= 1) {            // fewest padded rows with at most 24 boxes; ties -> larger box
    const int nb = (rows + br - 1) / br;
    if (nb > 24) break;
    if (nb * br < best_pad) { best_pad = nb * br; best_br = br; }
  }
  *box_rows = best_br; *nbox = (rows + best_br - 1) / best_br;
}

template <typename T, int C_IN, int N, int MB, int KS, int KBW, int TPS, int NST, int NSL, int NSTG>
cudaError_t launch_t(const TcArgs& a, ActType act, cudaStream_t s, int* grid_out) {
  constexpr int PW = N < 64 ? N : 64;
  TcKArgs ka{};
  ka.bias = a.bias; ka.res2 = a.res2; ka.n_q = a.n_q; ka.c_out = a.c_out; ka.has_res = a.res != nullptr;
  ka.res_inv = a.res_inv; ka.out_scale = a.out_scale; ka.out_slope = a.out_slope;
  ka.nvar = a.nvar;
  for (int v = 0; v < a.nvar; ++v) ka.var[v] = a.var[v];
  int mn = a.var[0].tap_off[0], mx = mn;
  for (int v = 0; v < a.nvar; ++v)
    for (int j = 0; j < a.var[v].ntaps; ++j) { mn = a.var[v].tap_off[j] < mn ? a.var[v].tap_off[j] : mn; mx = a.var[v].tap_off[j] > mx ? a.var[v].tap_off[j] : mx; }
  ka.min_off = mn; ka.span = mx - mn;
  if (ka.span > kTcMaxSpan) return cudaErrorInvalidValue;
  if ((a.res || a.res2) && a.out_stride != 1) return cudaErrorInvalidValue;
  tc_boxes(MB * 128 + ka.span, &ka.box_rows, &ka.nbox);
  const TcSmemPlan pl = tc_plan(N, KS, KBW * TPS, NST, NSL, NSTG, ka.nbox * ka.box_rows);
  if ((size_t)pl.total > kTcMaxSmem) return cudaErrorInvalidConfiguration;
  ka.tiles_q = (a.n_q + MB * 128 - 1) / (MB * 128);
  ka.n_tiles = a.c_out / N;
  ka.total_tiles = a.B * ka.tiles_q * a.nvar * ka.n_tiles;
  CUtensorMap tm_x, tm_res, tm_out;
  if (!make_map3(&tm_x, act, a.x, C_IN, a.x_rows, a.B, KS, ka.box_rows)) return cudaErrorNotSupported;
  if (!make_map4(&tm_out, act, a.out, a.c_out, a.out_stride, a.n_q, a.B, PW)) return cudaErrorNotSupported;
  if (a.res) { if (!make_map4(&tm_res, act, a.res, a.c_out, 1, a.n_q, a.B, PW)) return cudaErrorNotSupported; }
  else tm_res = tm_out;
  auto kern = conv_tc_kernel<T, C_IN, N, MB, KS, KBW, TPS, NST, NSL, NSTG>;
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
  kern<<<grid, kTcThreads, pl.total, s>>>(tm_x, tm_res, tm_out, ka);
  return cudaGetLastError();
}

// output channels per tile: at most 128 so that two accumulator buffers fit TMEM (epilogue overlaps the MMAs)
inline int tc_ntile(int c_out) { return c_out > 128 ? 128 : c_out; }

struct TcRow { int c_in, c_out_tile; TcConfig cfg; };
// Stage work at full tensor rate = MB * TPS * (KBW/16) * N/2 cycles: kept >= 512 so that the per-stage
// barrier round trip of the single issuing thread is amortised.
//   c_in  n    mb  ks kbw tps nst nsl nstg
#define HFG_TC_TABLE(X)               \
  X(512, 128, 1, 64, 64, 1, 3, 3, 4)  \
  X(256, 128, 2, 64, 64, 1, 3, 2, 5)  \
  X(128, 128, 2, 64, 64, 1, 3, 2, 5)  \
  X(128, 64, 2, 64, 64, 2, 3, 2, 5)   \
  X(64, 64, 2, 64, 64, 2, 3, 2, 5)    \
  X(64, 32, 4, 64, 64, 2, 4, 2, 5)    \
  X(32, 32, 4, 32, 32, 4, 4, 3, 8)    \
  X(32, 16, 4, 32, 32, 4, 4, 3, 8)    \
  X(16, 16, 4, 16, 16, 4, 4, 3, 8)

const TcRow kTcTable[] = {
#define X(CI, NN, MB, KS, KBW, TPS, NST, NSL, NSTG) {CI, NN, {NN, MB, KS, KBW, TPS, NST, NSL, NSTG}},
    HFG_TC_TABLE(X)
#undef X
};

template <typename T>
cudaError_t launch_shape(const TcArgs& a, ActType act, cudaStream_t s, int* grid_out) {
  const int ci = a.c_in, n = tc_ntile(a.c_out);
#define X(CI, NN, MB, KS, KBW, TPS, NST, NSL, NSTG) if (ci == CI && n == NN) return launch_t<T, CI, NN, MB, KS, KBW, TPS, NST, NSL, NSTG>(a, act, s, grid_out);
  HFG_TC_TABLE(X)
#undef X
  return cudaErrorInvalidValue;
}

}  // namespace

bool tc_config(int c_in, int c_out, TcConfig* cfg) {
  const int n = tc_ntile(c_out);
  if (c_out % n || c_out > kTcMaxCout) return false;
  for (const TcRow& r : kTcTable)
    if (r.c_in == c_in && r.c_out_tile == n) { if (cfg) *cfg = r.cfg; return true; }
  return false;
}

bool tc_supported(int c_in, int c_out, int ntaps, int span) {
  if (ntaps < 1 || ntaps > kMaxTaps || span < 0 || span > kTcMaxSpan) return false;
  TcConfig c;
  if (!tc_config(c_in, c_out, &c)) return false;
  int br, nb;
  tc_boxes(c.mb * 128 + span, &br, &nb);
  return (size_t)tc_plan(c.n, c.ks, c.kbw * c.tps, c.nst, c.nsl, c.nstg, br * nb).total <= kTcMaxSmem;
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
  if (a.nvar < 1 || a.nvar > kMaxVariants) return cudaErrorInvalidValue;
  if (a.B <= 0 || a.n_q <= 0) return cudaSuccess;
  if (act == ACT_F16) return launch_shape<__half>(a, act, s, grid_out);
  if (act == ACT_BF16) return launch_shape<__nv_bfloat16>(a, act, s, grid_out);
  return cudaErrorInvalidValue;
}

}  // namespace hfg
