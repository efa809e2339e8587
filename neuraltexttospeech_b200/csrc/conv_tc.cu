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
//           ring with cp.async.bulk + mbarrier complete_tx (pair mode: TMA 2-D loads of each CTA's half);
//           weights that fit the ring and are shared by every tile are loaded once (resident mode).
//   warp 2  MMA issuer: one thread issues tcgen05.mma (kind::f16, M=128 or 256 for a CTA pair, N, K=16) into
//           TMEM accumulators (NACC buffers of MB x N columns); tcgen05.commit releases weight stages / slab
//           slices and hands the accumulator to the epilogue.
//   warp 3  staging producer: TMA-loads the residual tile of each epilogue pass into the staging ring.
//   warps 4-15 epilogue, three groups of four taking passes round-robin (two on the 128-wide pair kernels, the third
//           then only gives its registers away): tcgen05.ld (one TMEM lane quarter per warp, thread = output row) -> +bias -> +residual (model.py:79,111) -> +MRF partial sum,
//           *1/num_kernels (model.py:161-167) -> activation for the consumer -> 16-bit, in place in the
//           swizzled staging tile; conv_post mode: tanh -> fp32 / 16-bit PCM samples (model.py:170-171).
//   warp 16 store issuer: TMA tensor store of every finished staging tile, two stores in flight (warps 17-19 idle:
//           they complete the warpgroup so that registers can be re-allocated between roles with setmaxnreg).
//
// Loads of tile i+1, MMAs of tile i and the epilogue of tile i-1 overlap (slab ring, TMEM double
// buffering where 2*MB*N <= 512 columns, NSTG-deep staging ring).  CG = 2 runs the kernel as CTA pairs
// (cta_group::2).  Every launch is a programmatic dependent launch: the prologue overlaps the previous
// layer's drain (griddepcontrol).
#include <cuda.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <type_traits>
#include <vector>

#include "kernels.cuh"
#include "tc_ptx.cuh"

namespace hfg {

namespace {

#ifndef HFG_TC_EPI_GROUPS
#define HFG_TC_EPI_GROUPS 3
#endif
constexpr int kEpiGroups = HFG_TC_EPI_GROUPS;                 // groups of 4 epilogue warps taking passes round-robin
constexpr int kStoreWarp = 4 + 4 * kEpiGroups;                // warp index of the store issuer
// Register re-allocation between roles (setmaxnreg, per warpgroup of 4 warps).  Each SM sub-partition holds one warp of
// every warpgroup and 512 registers per lane, so 13 warps cap every role at 128 registers and a third epilogue group
// would cap them at 96 (measured -6 %: the epilogue spills).  With HFG_TC_REGS_EPI > 0 the store issuer's warpgroup is
// padded to 4 warps (3 idle), the kernel starts at the launch bound and the producer / MMA / store warpgroups hand
// registers to the epilogue groups: 2 x kRegsIo + kEpiGroups x kRegsEpi <= warps-per-partition x launch registers.
#ifndef HFG_TC_REGS_EPI
#define HFG_TC_REGS_EPI (HFG_TC_EPI_GROUPS >= 3 ? 112 : 0)
#endif
#ifndef HFG_TC_REGS_IO
#define HFG_TC_REGS_IO 64
#endif
constexpr int kRegsEpi = HFG_TC_REGS_EPI, kRegsIo = HFG_TC_REGS_IO;
constexpr int kTcThreads = (kStoreWarp + (kRegsEpi > 0 ? 4 : 1)) * 32;   // 4 producer/MMA warps, kEpiGroups x 4 epilogue warps, store issuer
// Active epilogue groups of one instantiation (<= kEpiGroups; the others idle at 24 registers and the active ones take
// theirs).  HFG_TC_EG_RULE: 0 = every kernel uses all groups, 1 = the 128-wide CTA-pair kernels (operand-bound wide
// stages) keep two, 2 = every CTA-pair kernel keeps two.
#ifndef HFG_TC_EG_RULE
#define HFG_TC_EG_RULE 1
#endif
__host__ __device__ constexpr int tc_epi_groups(int n, int cg) {
  if (kRegsEpi == 0 || kEpiGroups <= 2) return kEpiGroups;
  if (HFG_TC_EG_RULE == 1 && cg == 2 && n >= 128) return 2;
  if (HFG_TC_EG_RULE == 2 && cg == 2) return 2;
  return kEpiGroups;
}
// registers of an active epilogue warp: what the idle groups give up is shared out (launch bound 96 x 5 warps per partition)
__host__ __device__ constexpr int tc_epi_regs(int eg) { return eg == kEpiGroups ? kRegsEpi : ((kEpiGroups * kRegsEpi - (kEpiGroups - eg) * 24) / eg) / 8 * 8; }
constexpr int kTcMaxCout = 512;
constexpr int kTcMaxSpan = 96;       // largest max_off - min_off supported (V1: 50, V3: 72)
constexpr size_t kTcMaxSmem = 227 * 1024;

constexpr int kTcMaxAcc = 4;          // accumulator buffers in TMEM (acc_full / acc_empty barrier pairs)
__host__ __device__ constexpr int tc_nacc(int mb, int n) { return 2 * mb * n <= 512 ? 2 : 1; }   // 4 buffers were tried: the epilogue, not the accumulator ring, paces the narrow layers

// 128-row blocks handled by one epilogue pass: narrow tiles (N < 64) take several, so that a pass always moves ~64
// values per thread (N = 32) and the per-pass fixed costs (barrier hand-offs, proxy fence, TMA issue) are paid half as often
// staging ring depth when a pass covers rb row blocks: the table's NSTG counts 128-row buffers
__host__ __device__ constexpr int tc_nstg(int nstg, int rb) { return nstg / rb < 3 ? 3 : nstg / rb; }
__host__ __device__ constexpr int tc_rb(int n, int mb) { return (n >= 64 || mb < 2) ? 1 : 2; }   // a TMA box holds at most 256 rows

__host__ __device__ constexpr uint32_t tc_tmem_cols(int need) { return need <= 32 ? 32 : need <= 64 ? 64 : need <= 128 ? 128 : need <= 256 ? 256 : 512; }

// ------------------------------------------------------------------ shared-memory plan
struct TcSmemPlan {
  int slice_rows;       // rows of one K-slice buffer (nbox * box_rows)
  int slice_bytes;
  int stage_bytes;      // one weight stage
  int stg_bytes;        // one staging tile [RB * 128 rows][PW * 2 bytes]
  int off_stg, off_w, off_bias, off_bar;
  int total;
};

__host__ __device__ inline TcSmemPlan tc_plan(int n, int ks, int kbw, int nst, int nsl, int nstg, int slice_rows, int cg = 1, int rb = 1) {
  TcSmemPlan p;
  const int pw = n < 64 ? n : 64;
  p.slice_rows = slice_rows;
  p.slice_bytes = slice_rows * ks * 2;            // multiple of the swizzle atom (slice_rows % 8 == 0)
  p.stage_bytes = (n / cg) * kbw * 2;     // a CTA of a pair stages half of the weight rows
  p.stg_bytes = rb * 128 * pw * 2;
  int off = nsl * p.slice_bytes;
  off = (off + 1023) & ~1023;
  p.off_stg = off; off += nstg * p.stg_bytes;
  off = (off + 127) & ~127;
  p.off_w = off; off += nst * p.stage_bytes;
  p.off_bias = off; off += kTcMaxCout * 4;     // bias of every output channel of the layer
  off = (off + 15) & ~15;
  p.off_bar = off; off += (2 * nsl + 2 * nst + 3 * nstg + 2 * kTcMaxAcc) * 8 + 16;
  p.total = off;
  return p;
}

// Debug build only (-DHFG_TC_TRACE, tools/trace_layer.py): CTA 0 stamps clock64() per role / tile / event.
#ifdef HFG_TC_TRACE
#define TC_TRACE(role, tile_i, ev)                                                                                   \
  do { if (ka.trace && blockIdx.x == 0 && (tile_i) < 64u) ka.trace[((role) * 64u + (tile_i)) * 8u + (ev)] = clock64(); } while (0)
#else
#define TC_TRACE(role, tile_i, ev) do { } while (0)
#endif

struct TcKArgs {
  long long* trace;                    // HFG_TC_TRACE builds: [6 roles][64 tiles][8 events] clock64 stamps of CTA 0, else unused
  const float* bias;
  const void* res2;
  int n_q, c_out, has_res;
  float res_inv, out_scale, out_slope;
  int nvar;
  TcVariant var[kMaxVariants];
  int min_off, span;
  int box_rows, nbox;   // slab slice = nbox TMA boxes of box_rows rows
  int tiles_q, n_tiles, total_tiles;   // CG == 2: tiles_q / total_tiles count PAIR tiles of 2*MB*128 rows
  const void* w_base;                  // CG == 2: base of the layer's packed weights (tm_w), 256-byte rows
  int resident;                        // every tile uses the same weights and they fit the NST stages: load them once
  float* post_out;                     // conv_post mode: audio[b, q*post_f + s] = tanh(channel s) written as fp32, no 16-bit tile
  int post_f;
  int16_t* post_pcm;                   // conv_post mode: 16-bit PCM output instead of post_out (pcm16_of)
  int store_in_flight;                 // debug knob: TMA stores kept in flight by the store warp (-1 = default rule)
  int nsl, nstg;                       // ring depths actually used (>= the table's NSL / NSTG when shared memory allows)
};

// ------------------------------------------------------------------ the kernel
// CG == 2: the kernel runs as CTA pairs (cluster of 2, cta_group::2): the pair computes 2*MB*128 rows with M=256
// MMAs issued by the leader (rank 0); each CTA stages its own rows of A and only HALF of the weight rows (B), which
// cuts the shared-memory operand reads per MMA from 4+N*32 to 4+N*16 bytes per N/2 cycles and halves weight staging.
template <typename T, int C_IN, int N, int MB, int KS, int KBW, int TPS, int NST, int NSL, int NSTG, int CG>
__global__ void __launch_bounds__(kTcThreads, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_res,
               const __grid_constant__ CUtensorMap tm_out, const __grid_constant__ CUtensorMap tm_w,
               const __grid_constant__ TcKArgs ka) {
  static_assert(CG == 1 || CG == 2, "cta group");
  constexpr int NB = N / CG;                      // weight rows (B operand) staged by this CTA
  constexpr int EG = tc_epi_groups(N, CG);        // epilogue groups at work in this instantiation
  constexpr int NKS = C_IN / KS;                  // K-slices per tile
  constexpr int NKB = KS / KBW;                   // weight k-blocks per (slice, tap)
  constexpr int BLK_BYTES = NB * KBW * 2;         // this CTA's part of one (tap, k-block) weight block; a stage holds TPS
  static_assert(TPS == 1 || NKB == 1, "multi-tap stages need KBW == KS");
  constexpr int ROWB = KS * 2;                    // bytes per slab row within a slice (= swizzle width)
  constexpr int PW = N < 64 ? N : 64;             // output columns per epilogue pass
  constexpr int SROWB = PW * 2;                   // bytes per staging row (= swizzle width)
  constexpr int RB = tc_rb(N, MB);                // 128-row blocks per epilogue pass
  constexpr int NACC = tc_nacc(MB, N);
  constexpr int ACC_COLS = MB * N;
  constexpr uint32_t TMEM_COLS = tc_tmem_cols(NACC * ACC_COLS);
  constexpr uint32_t IDESC = make_idesc(N, std::is_same<T, __nv_bfloat16>::value, 128 * CG);
  static_assert(C_IN % KS == 0 && KS % KBW == 0 && KBW % 16 == 0 && N % 16 == 0 && N <= 256 && NB % 8 == 0, "tile config");
  static_assert(CG == 1 || BLK_BYTES % 256 == 0, "pair weight blocks are whole 256-byte rows of tm_w");
  static_assert(KS == 64 || KS == 32 || KS == 16, "K-slice = one swizzle row");
  static_assert(N % PW == 0, "pass width");
  const int nsl = ka.nsl, nstg = ka.nstg;          // runtime ring depths (the template values are the table defaults)
  const TcSmemPlan pl = tc_plan(N, KS, KBW * TPS, NST, nsl, nstg, ka.nbox * ka.box_rows, CG, RB);

  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* slab = smem;
  uint8_t* stg = smem + pl.off_stg;
  uint8_t* wst = smem + pl.off_w;
  float* bias_s = reinterpret_cast<float*>(smem + pl.off_bias);
  uint64_t* slice_full = reinterpret_cast<uint64_t*>(smem + pl.off_bar);
  uint64_t* slice_empty = slice_full + nsl;
  uint64_t* w_full = slice_empty + nsl;
  uint64_t* w_empty = w_full + NST;
  uint64_t* stg_full = w_empty + NST;
  uint64_t* stg_empty = stg_full + nstg;
  uint64_t* stg_done = stg_empty + nstg;
  uint64_t* acc_full = stg_done + nstg;
  uint64_t* acc_empty = acc_full + kTcMaxAcc;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + kTcMaxAcc);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = CG == 2 ? cluster_ctarank() : 0u;       // 0 = leader of the pair
  pdl_launch_dependents();            // the next layer's kernel may begin its prologue on SMs this grid leaves free
  const int tile0 = blockIdx.x / CG, tile_step = gridDim.x / CG;

  if (tid == 0) {
    for (int s = 0; s < nsl; ++s) { mbar_init(&slice_full[s], 1); mbar_init(&slice_empty[s], 1); }
    for (int s = 0; s < NST; ++s) { mbar_init(&w_full[s], 1); mbar_init(&w_empty[s], 1); }
    for (int s = 0; s < nstg; ++s) { mbar_init(&stg_full[s], 1); mbar_init(&stg_empty[s], 1); mbar_init(&stg_done[s], 4); }
    for (int s = 0; s < kTcMaxAcc; ++s) { mbar_init(&acc_full[s], 1); mbar_init(&acc_empty[s], 4 * EG * CG); }
    fence_mbar_init();
    tma_prefetch_desc(&tm_x); tma_prefetch_desc(&tm_res); tma_prefetch_desc(&tm_out);
    if (CG == 2) tma_prefetch_desc(&tm_w);
  }
  if (warp == 2) { if (CG == 2) tmem_alloc2(tmem_slot, TMEM_COLS); else tmem_alloc(tmem_slot, TMEM_COLS); }
  for (int i = tid; i < ka.c_out; i += kTcThreads) bias_s[i] = ka.bias[i];
  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();   // pair: the peer's barriers are initialised before any remote signal
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  // Everything above (barriers, TMEM, bias, descriptor prefetch) touched only constants and ran while the previous
  // layer's kernel was still draining; activations are read / written only past this point.
  if (warp != 1) pdl_wait();          // the weight producer reads constants only: it starts staging right away

  // tile index -> (utterance b, row tile qt, variant v, channel tile nt)
  auto decode = [&](int t, int& b, int& qt, int& v, int& nt) {
    nt = t % ka.n_tiles; t /= ka.n_tiles;
    v = t % ka.nvar; t /= ka.nvar;
    qt = t % ka.tiles_q; b = t / ka.tiles_q;
    if (CG == 2) qt = qt * 2 + (int)rank;           // the pair's 2*MB*128 rows: leader takes the first half
  };

  // Producer / MMA warps run their loops with ALL lanes (warp-uniform control flow keeps barrier addresses and
  // descriptors in uniform registers) and issue the asynchronous instructions from one elected lane.
  if (warp < 4) {
  if constexpr (kRegsEpi > 0) regs_release<kRegsIo>();
  if (warp == 0) {
    // ===================== slab producer (TMA) =====================
    uint32_t scnt = 0;
    const uint32_t box_bytes = (uint32_t)ka.box_rows * ROWB;
    for (int t = tile0; t < ka.total_tiles; t += tile_step) {
      int b, qt, v, nt; decode(t, b, qt, v, nt);
      const int row_base = qt * (MB * 128) + ka.min_off;
      for (int s = 0; s < NKS; ++s, ++scnt) {
        const int pos = scnt % (uint32_t)nsl;
        mbar_wait(&slice_empty[pos], ((scnt / (uint32_t)nsl) & 1) ^ 1);
        if (elect_one()) {
          if (s == 0) TC_TRACE(0u, (uint32_t)((t - tile0) / tile_step), 0u);      // slot free, load issued
          uint8_t* dst = slab + pos * pl.slice_bytes;
          if (CG == 1) {
            mbar_expect_tx(&slice_full[pos], box_bytes * ka.nbox);
            for (int bx = 0; bx < ka.nbox; ++bx)
              tma_load_3d(dst + bx * box_bytes, &tm_x, s * KS, row_base + bx * ka.box_rows, b, &slice_full[pos]);
          } else {
            // both CTAs' slabs complete on the leader's barrier (the leader's MMAs read both)
            if (rank == 0) mbar_expect_tx(&slice_full[pos], 2 * box_bytes * ka.nbox);
            const uint32_t bar_cl = mapa_u32(smem_u32(&slice_full[pos]), 0);
            for (int bx = 0; bx < ka.nbox; ++bx)
              tma_load_3d_2cta(dst + bx * box_bytes, &tm_x, s * KS, row_base + bx * ka.box_rows, b, bar_cl);
          }
        }
        __syncwarp();
      }
    }
  } else if (warp == 1) {
    // ===================== weight producer =====================
    uint32_t wcnt = 0;
    for (int t = tile0; t < ka.total_tiles; t += tile_step) {
      int b, qt, v, nt; decode(t, b, qt, v, nt);
      const int ntaps = ka.var[v].ntaps;
      const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(ka.var[v].w) + (size_t)nt * (NKS * ntaps * NKB) * BLK_BYTES;
      // pair: block hb of the variant lives at 256-byte row (var_row + (2*hb + rank) * BLK_BYTES/256) of tm_w
      int wrow = CG == 2 ? (int)((reinterpret_cast<const uint8_t*>(ka.var[v].w) - reinterpret_cast<const uint8_t*>(ka.w_base)) >> 8) +
                               (2 * nt * (NKS * ntaps * NKB) + (int)rank) * (BLK_BYTES >> 8) : 0;
      for (int s = 0; s < NKS; ++s) {
        for (int j0 = 0; j0 < ntaps; j0 += TPS) {
          const int nj = (ntaps - j0) < TPS ? (ntaps - j0) : TPS;
          for (int kb = 0; kb < NKB; ++kb, ++wcnt) {
            const int st = wcnt % NST;
            const uint32_t bytes = (uint32_t)nj * BLK_BYTES;
            mbar_wait(&w_empty[st], ((wcnt / NST) & 1) ^ 1);
            if (elect_one()) {
              if (CG == 1) {
                mbar_expect_tx(&w_full[st], bytes);
                bulk_g2s(wst + st * pl.stage_bytes, wsrc, bytes, &w_full[st]);
              } else {
                if (rank == 0) mbar_expect_tx(&w_full[st], 2 * bytes);
                const uint32_t bar_cl = mapa_u32(smem_u32(&w_full[st]), 0);
                for (int jj = 0; jj < nj; ++jj)        // NKB == 1 when TPS > 1: consecutive taps are consecutive blocks
                  tma_load_2d_2cta(wst + st * pl.stage_bytes + jj * BLK_BYTES, &tm_w, 0, wrow + 2 * jj * (BLK_BYTES >> 8), bar_cl);
              }
            }
            __syncwarp();
            wsrc += bytes;
            wrow += 2 * nj * (BLK_BYTES >> 8);
          }
        }
      }
      if (ka.resident) break;                         // the stages stay in shared memory for every later tile
    }
  } else if (warp == 2) {
    // ===================== MMA issuer (pair: leader CTA only) =====================
    if (rank == 0) {
    const uint32_t slab_addr = smem_u32(slab), wst_addr = smem_u32(wst);
    uint32_t scnt = 0, wcnt = 0, tcnt = 0;
    for (int t = tile0; t < ka.total_tiles; t += tile_step, ++tcnt) {
      int b, qt, v, nt; decode(t, b, qt, v, nt);
      const int ntaps = ka.var[v].ntaps;
      const int abuf = tcnt % NACC;
      if (lane == 0) TC_TRACE(2u, tcnt, 0u);                      // MMA warp arrives at the tile
      mbar_wait(&acc_empty[abuf], ((tcnt / NACC) & 1) ^ 1);
      if (lane == 0) TC_TRACE(2u, tcnt, 1u);                      // accumulator free
      tc_fence_after();
      const uint32_t tacc = tmem_base + abuf * ACC_COLS;
      uint32_t first = 1;
      if (ka.resident) wcnt = 0;
#pragma unroll 1
      for (int s = 0; s < NKS; ++s, ++scnt) {
        const int pos = scnt % (uint32_t)nsl;
        mbar_wait(&slice_full[pos], (scnt / (uint32_t)nsl) & 1);
        if (lane == 0 && s == 0) TC_TRACE(2u, tcnt, 2u);          // first slab slice landed
        tc_fence_after();
        const uint32_t sl = slab_addr + pos * pl.slice_bytes;
#pragma unroll 1
        for (int j0 = 0; j0 < ntaps; j0 += TPS) {
          const int nj = (ntaps - j0) < TPS ? (ntaps - j0) : TPS;
#pragma unroll 1
          for (int kb = 0; kb < NKB; ++kb, ++wcnt) {
            const int st = wcnt % NST;                // resident: wcnt restarts per tile, stage = position in the tile
            if (!ka.resident || tcnt == 0) mbar_wait(&w_full[st], (wcnt / NST) & 1);
            tc_fence_after();
            if (elect_one()) {
              const uint32_t wb = wst_addr + st * pl.stage_bytes;
#pragma unroll
              uint32_t fst = first;
              for (int jj = 0; jj < TPS; ++jj) {
                if (jj >= nj) break;
                const uint32_t row_off = (uint32_t)(ka.var[v].tap_off[j0 + jj] - ka.min_off);
                // descriptor low words (start address / 16, LBO); the high words are compile-time constants, so that the issuing
                // thread does 32-bit adds per MMA instead of rebuilding two 64-bit values (its issue cost caps small MMAs)
                constexpr uint32_t A_HI = ((8u * ROWB) >> 4) | (1u << 14) | ((ROWB == 128 ? 2u : (ROWB == 64 ? 4u : 6u)) << 29);
                constexpr uint32_t B_HI = (128u >> 4) | (1u << 14);
                const uint32_t a0 = (uint32_t)make_desc_sw<ROWB>(sl + row_off * ROWB + kb * KBW * 2);
                const uint32_t b0 = (uint32_t)make_desc(wb + jj * BLK_BYTES, NB * 16, 128u);
                // K-steps of this tap whose weight rows are all zero (fold slots a shifted tap does not read, padded channels)
                const uint32_t zero = (ka.var[v].kz[j0 + jj] >> ((s * KS + kb * KBW) / 16)) & ((1u << (KBW / 16)) - 1u);
                if (zero == (1u << (KBW / 16)) - 1u) continue;
#pragma unroll
                for (int mb = 0; mb < MB; ++mb) {
                  uint32_t f = fst;
#pragma unroll
                  for (int kk = 0; kk < KBW / 16; ++kk) {
                    if ((zero >> kk) & 1u) continue;
                    // descriptor start-address field counts 16-byte units: advance with plain adds
                    const uint32_t adesc = a0 + (uint32_t)((mb * 128 * ROWB + kk * 32) >> 4);
                    const uint32_t bdesc = b0 + (uint32_t)((kk * 2 * NB * 16) >> 4);
                    if (CG == 2) umma_f16_2cta_lo<A_HI, B_HI>(tacc + mb * N, adesc, bdesc, IDESC, f ? 0u : 1u);
                    else umma_f16_lo<A_HI, B_HI>(tacc + mb * N, adesc, bdesc, IDESC, f ? 0u : 1u);
                    f = 0;
                  }
                }
                fst = 0;
              }
              // pair: the commits arrive on the same barrier of BOTH CTAs
              if (!ka.resident) { if (CG == 2) umma_commit_2cta(&w_empty[st]); else umma_commit(&w_empty[st]); }   // frees the weight stage
              if (kb == NKB - 1 && j0 + TPS >= ntaps) {
                if (CG == 2) umma_commit_2cta(&slice_empty[pos]); else umma_commit(&slice_empty[pos]);   // slab slice may be overwritten
                if (s == NKS - 1) { if (CG == 2) umma_commit_2cta(&acc_full[abuf]); else umma_commit(&acc_full[abuf]); TC_TRACE(2u, tcnt, 3u); }   // tile complete (all MMAs issued)
              }
            }
            for (int jj = 0; jj < nj; ++jj)                       // every lane tracks whether the accumulators have been started
              if (((ka.var[v].kz[j0 + jj] >> ((s * KS + kb * KBW) / 16)) & ((1u << (KBW / 16)) - 1u)) != (1u << (KBW / 16)) - 1u) first = 0;
            __syncwarp();
          }
        }
      }
    }
    }
  } else if (warp == 3) {
    // ===================== staging producer: residual tiles (or just buffer grants) =====================
    uint32_t pcnt = 0;
    for (int t = tile0; t < ka.total_tiles; t += tile_step) {
      int b, qt, v, nt; decode(t, b, qt, v, nt);
      for (int mb = 0; mb < MB; mb += RB) {
        const int q0 = qt * (MB * 128) + mb * 128;
        for (int p0 = 0; p0 < N; p0 += PW, ++pcnt) {
          const int buf = pcnt % (uint32_t)nstg;
          mbar_wait(&stg_empty[buf], ((pcnt / (uint32_t)nstg) & 1) ^ 1);
          if (elect_one()) {
            if (mb == 0 && p0 == 0) TC_TRACE(1u, (uint32_t)((t - tile0) / tile_step), 0u);   // first staging buffer of the tile granted
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
  }
  } else if (warp >= kStoreWarp) {
  if constexpr (kRegsEpi > 0) regs_release<kRegsIo>();
  if (warp == kStoreWarp) {
    // ===================== store issuer: TMA tensor store of every finished staging tile =====================
    // Up to two stores stay in flight: the buffer of pass p is handed back only once the store of pass p has finished
    // READING shared memory, which is checked two passes later (bulk groups retire in order).  Waiting for each store
    // right after issuing it serialised ~0.5 us per 8-16 KB tile and bounded the short-k layers of the narrow stages.
    // Holding buffers longer shortens the residual prefetch distance, so small rings keep fewer stores in flight.
    const int infl_env = ka.store_in_flight;            // debug knob (HFG_TC_STORES): -1 = default rule
    const uint32_t in_flight = infl_env >= 0 ? (uint32_t)(infl_env < nstg - 1 ? infl_env : nstg - 2)
                                             : (nstg >= 5 ? 2u : (nstg == 4 ? 1u : 0u));
    uint32_t pcnt = 0;
    for (int t = tile0; t < ka.total_tiles; t += tile_step) {
      int b, qt, v, nt; decode(t, b, qt, v, nt);
      const int out_off = ka.var[v].out_off;
      for (int mb = 0; mb < MB; mb += RB) {
        const int q0 = qt * (MB * 128) + mb * 128;
        for (int p0 = 0; p0 < N; p0 += PW, ++pcnt) {
          const int buf = pcnt % (uint32_t)nstg;
          mbar_wait(&stg_done[buf], (pcnt / (uint32_t)nstg) & 1);
          if (elect_one()) {
            if (mb == 0 && p0 == 0) TC_TRACE(3u, (uint32_t)((t - tile0) / tile_step), 0u);   // first finished pass of the tile seen by the store warp
            if (mb + RB >= MB && p0 + PW >= N) TC_TRACE(3u, (uint32_t)((t - tile0) / tile_step), 1u);   // last one
            if (q0 < ka.n_q && ka.post_out == nullptr && ka.post_pcm == nullptr)
              tma_store_4d(&tm_out, stg + buf * pl.stg_bytes, nt * N + p0, out_off, q0, b);
            bulk_commit();                              // one (possibly empty) group per pass keeps the group count = pass count
            switch (in_flight) {                        // every group but the in_flight newest has been read out
              case 0: bulk_wait_read<0>(); break;
              case 1: bulk_wait_read<1>(); break;
              case 2: bulk_wait_read<2>(); break;
              case 3: bulk_wait_read<3>(); break;
              case 4: bulk_wait_read<4>(); break;
              case 5: bulk_wait_read<5>(); break;
              default: bulk_wait_read<6>(); break;
            }
            if (pcnt >= in_flight) mbar_arrive(&stg_empty[(pcnt - in_flight) % (uint32_t)nstg]);
          }
          __syncwarp();
        }
      }
    }
    // the last in_flight buffers are not handed back: no pass is left to use them
    if (elect_one()) bulk_wait_all();                   // same lane every time (elect.sync picks the lowest)
    __syncwarp();
  }
  } else {
    // ===================== epilogue: kEpiGroups groups of 4 warps (4-7, 8-11, ...) take passes round-robin =====================
    const bool idle = EG < kEpiGroups && warp >= 4 + 4 * EG;       // a whole warpgroup: it only gives its registers away
    if constexpr (EG < kEpiGroups) { if (idle) regs_release<24>(); }
    if (!idle) {
    if constexpr (kRegsEpi > 0) regs_acquire<tc_epi_regs(EG)>();
    const int grp = (warp - 4) >> 2;
    const int quarter = warp & 3;                       // TMEM lane quarter this warp may access
    const int er = quarter * 32 + lane;                 // row of the 128-row block owned by this thread
    const uint32_t lane_base = ((uint32_t)(quarter * 32)) << 16;
    constexpr int CW = PW >= 32 ? 32 : 16;
    // XOR term of the TMA swizzle for this thread's staging row (row pitch SROWB, 1024-byte aligned tile)
    const int sx = SROWB == 128 ? (er & 7) : (SROWB == 64 ? ((er >> 1) & 3) : ((er >> 2) & 1));
    uint32_t tcnt = 0, pcnt = 0;
    for (int t = tile0; t < ka.total_tiles; t += tile_step, ++tcnt) {
      int b, qt, v, nt; decode(t, b, qt, v, nt);
      const int n0 = nt * N;
      const int abuf = tcnt % NACC;
      if (quarter == 0 && lane == 0) TC_TRACE(4u + grp, tcnt, 0u);   // epilogue group arrives at the tile
      mbar_wait(&acc_full[abuf], (tcnt / NACC) & 1);
      if (quarter == 0 && lane == 0) TC_TRACE(4u + grp, tcnt, 1u);   // accumulator complete
      tc_fence_after();
      const uint32_t tacc = tmem_base + lane_base + abuf * ACC_COLS;
#pragma unroll 1
      for (int mb = 0; mb < MB; mb += RB) {
        const int q0 = qt * (MB * 128) + mb * 128;
        const bool has_res = ka.has_res != 0 && q0 < ka.n_q;
#pragma unroll 1
        for (int p0 = 0; p0 < N; p0 += PW, ++pcnt) {
          if ((int)(pcnt % (uint32_t)EG) != grp) continue;
          const int buf = pcnt % (uint32_t)nstg;
#pragma unroll 1
          for (int rb = 0; rb < RB; ++rb) {              // the pass's 128-row blocks (one unless the tile is narrow)
          const int q = q0 + rb * 128 + er;
          const bool valid = q < ka.n_q;
          uint8_t* my = stg + buf * pl.stg_bytes + (rb * 128 + er) * SROWB;   // 128 rows keep the swizzle phase: sx is unchanged
#pragma unroll
          for (int c0 = 0; c0 < PW; c0 += CW) {
            uint32_t r[32];
            if (CW == 32) tmem_ld32(tacc + (mb + rb) * N + p0 + c0, r);
            else tmem_ld16(tacc + (mb + rb) * N + p0 + c0, r);
            if (rb == 0 && c0 == 0) {                    // the accumulator load is in flight while the staging buffer is awaited
              mbar_wait(&stg_full[buf], (pcnt / (uint32_t)nstg) & 1);   // residual tile landed / buffer granted
              if (quarter == 0 && lane == 0 && mb == 0 && p0 < 2 * PW) TC_TRACE(4u + grp, tcnt, 2u);   // first pass of the group: staging buffer ready
            }
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
            if (quarter == 0 && lane == 0 && mb == 0 && p0 < 2 * PW && c0 == 0) TC_TRACE(4u + grp, tcnt, 4u);   // accumulator chunk in registers
            float vv[CW];
            const float4* b4 = reinterpret_cast<const float4*>(bias_s + n0 + p0 + c0);   // 16-byte broadcast loads
#pragma unroll
            for (int i = 0; i < CW / 4; ++i) {
              const float4 bb = b4[i];
              vv[4 * i + 0] = __uint_as_float(r[4 * i + 0]) + bb.x; vv[4 * i + 1] = __uint_as_float(r[4 * i + 1]) + bb.y;
              vv[4 * i + 2] = __uint_as_float(r[4 * i + 2]) + bb.z; vv[4 * i + 3] = __uint_as_float(r[4 * i + 3]) + bb.w;
            }
            if (has_res) {
              const float ri = ka.res_inv;
#pragma unroll
              for (int i = 0; i < CW / 8; ++i) {
                float2 f;
                f = H2<T>::up(rr[i].x); vv[i * 8 + 0] += lrelu_ge1(f.x, ri); vv[i * 8 + 1] += lrelu_ge1(f.y, ri);
                f = H2<T>::up(rr[i].y); vv[i * 8 + 2] += lrelu_ge1(f.x, ri); vv[i * 8 + 3] += lrelu_ge1(f.y, ri);
                f = H2<T>::up(rr[i].z); vv[i * 8 + 4] += lrelu_ge1(f.x, ri); vv[i * 8 + 5] += lrelu_ge1(f.y, ri);
                f = H2<T>::up(rr[i].w); vv[i * 8 + 6] += lrelu_ge1(f.x, ri); vv[i * 8 + 7] += lrelu_ge1(f.y, ri);
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
            if (ka.post_out != nullptr || ka.post_pcm != nullptr) {   // conv_post + tanh (reference model.py:170-171): C_out = 1 lives in column 0
              if (valid && p0 + c0 == 0 && n0 == 0 && ka.post_pcm != nullptr) {
                int16_t* op = ka.post_pcm + ((size_t)b * ka.n_q + q) * ka.post_f;
#pragma unroll
                for (int i = 0; i < 16; ++i) if (i < ka.post_f) op[i] = pcm16_of(tanhf(vv[i]));
              } else if (valid && p0 + c0 == 0 && n0 == 0) {
                float* op = ka.post_out + ((size_t)b * ka.n_q + q) * ka.post_f;
                if (ka.post_f == 1) op[0] = tanhf(vv[0]);
                else if (ka.post_f == 2) *reinterpret_cast<float2*>(op) = make_float2(tanhf(vv[0]), tanhf(vv[1]));
                else {
#pragma unroll
                  for (int i = 0; i < 16; ++i) if (i < ka.post_f) op[i] = tanhf(vv[i]);
                }
              }
              continue;
            }
            const float osc = ka.out_scale, osl = ka.out_slope;
            if (osc != 1.0f) {                            // only the last conv of a stage scales (1 / num_kernels)
#pragma unroll
              for (int i = 0; i < CW; ++i) vv[i] *= osc;
            }
#pragma unroll
            for (int i = 0; i < CW; ++i) vv[i] = lrelu_le1(vv[i], osl);
#pragma unroll
            for (int i = 0; i < CW / 8; ++i) {
              uint4 u;
              u.x = H2<T>::down(vv[i * 8 + 0], vv[i * 8 + 1]); u.y = H2<T>::down(vv[i * 8 + 2], vv[i * 8 + 3]);
              u.z = H2<T>::down(vv[i * 8 + 4], vv[i * 8 + 5]); u.w = H2<T>::down(vv[i * 8 + 6], vv[i * 8 + 7]);
              *reinterpret_cast<uint4*>(my + (((c0 / 8 + i) ^ sx) * 16)) = u;
            }
          }
          }
          if (quarter == 0 && lane == 0 && mb == 0 && p0 < 2 * PW) TC_TRACE(4u + grp, tcnt, 5u);   // pass computed and staged
          fence_proxy_async();                          // staging writes -> visible to the TMA engine
          __syncwarp();
          if (quarter == 0 && lane == 0 && mb == 0 && p0 < 2 * PW) TC_TRACE(4u + grp, tcnt, 6u);   // fenced
          if (lane == 0) mbar_arrive(&stg_done[buf]);   // one arrival per warp: hand the tile to the store warp
          if (quarter == 0 && lane == 0 && mb == 0 && p0 < 2 * PW) TC_TRACE(4u + grp, tcnt, 7u);   // handed over
        }
      }
      tc_fence_before();
      __syncwarp();
      if (quarter == 0 && lane == 0) TC_TRACE(4u + grp, tcnt, 3u);   // all passes of the tile done
      if (lane == 0) {                                  // one arrival per warp: accumulator buffer may be overwritten
        if (CG == 2) mbar_arrive_cluster(mapa_u32(smem_u32(&acc_empty[abuf]), 0));   // the leader's MMA warp waits for both CTAs
        else mbar_arrive(&acc_empty[abuf]);
      }
    }
    }
  }
  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();   // pair: nobody exits while the peer may still signal its barriers
  if (warp == 2) {
    __syncwarp(); tc_fence_after();
    if (CG == 2) tmem_dealloc2(tmem_base, TMEM_COLS); else tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------ host side
int g_num_sms = 0;
long long* g_tc_trace = nullptr;       // device buffer for HFG_TC_TRACE builds (hfgdbg_set_trace)

int tc_num_sms() {
  if (!g_num_sms) {
    int dev = 0; cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&g_num_sms, cudaDevAttrMultiProcessorCount, dev);
    if (g_num_sms <= 0) g_num_sms = 148;
  }
  return g_num_sms;
}

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
bool make_map4(CUtensorMap* m, ActType act, const void* base, int C, int u, int nq, int B, int box_c, int box_rows = 128) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  cuuint64_t gdim[4] = {(cuuint64_t)C, (cuuint64_t)u, (cuuint64_t)nq, (cuuint64_t)B};
  cuuint64_t gstr[3] = {(cuuint64_t)C * 2, (cuuint64_t)C * 2 * u, (cuuint64_t)C * 2 * u * (cuuint64_t)nq};
  cuuint32_t box[4] = {(cuuint32_t)box_c, 1, (cuuint32_t)box_rows, 1};
  cuuint32_t est[4] = {1, 1, 1, 1};
  return fn(m, act == ACT_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, const_cast<void*>(base), gdim, gstr,
            box, est, CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle_for(box_c * 2), CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// slab slice = nbox boxes of box_rows rows covering MB*128 + span rows with little waste
void tc_boxes(int rows, int* box_rows, int* nbox) {
  int best_br = 256, best_pad = 1 << 30;
  static const int min_br = []() { const char* e = getenv("HFG_TC_MINBOX"); const int v = e ? atoi(e) : 8; return v < 8 ? 8 : v; }();
  for (int br = 256; br >= min_br; br >>= 1) {       // fewest padded rows with at most 24 boxes; ties -> larger box
    const int nb = (rows + br - 1) / br;
    if (nb > 24) break;
    if (nb * br < best_pad) { best_pad = nb * br; best_br = br; }
  }
  *box_rows = best_br; *nbox = (rows + best_br - 1) / best_br;
}

// rank-2 map over a layer's packed weights viewed as 256-byte rows; box = one CTA's part of a weight block
bool make_map_w(CUtensorMap* m, const void* base, size_t bytes, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) return false;
  cuuint64_t gdim[2] = {128, (cuuint64_t)(bytes / 256)};
  cuuint64_t gstr[1] = {256};
  cuuint32_t box[2] = {128, (cuuint32_t)box_rows};
  cuuint32_t est[2] = {1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, const_cast<void*>(base), gdim, gstr, box, est, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

int tc_env_int(const char* name, int dflt) { const char* e = getenv(name); return e ? atoi(e) : dflt; }

// Ring depths of a launch.  The table's (NSL, NSTG) always fit.  Layers whose weights stay resident in shared memory
// (C <= 64: the HBM-bound layers) have per-tile MMA time far below the latency of one slab load, so with the table's
// two slab slots only ONE load is in flight per SM while the other slot is consumed: their depth is raised as far as
// shared memory allows (Little's law: ~80 KB in flight per SM saturate HBM), trading staging slots above 3 for slab slots.
void tc_ring_depths(int n, int ks, int kbw_tps, int nst, int nsl0, int nstg0, int slice_rows, int cg, int rb, bool resident, int* nsl, int* nstg) {
  *nsl = nsl0; *nstg = nstg0;
  static const int want_nsl = tc_env_int("HFG_TC_NSL", 0), min_nstg = tc_env_int("HFG_TC_NSTG_MIN", 3);
  if (!resident || want_nsl <= nsl0) return;
  auto fits = [&](int a, int b) { return (size_t)tc_plan(n, ks, kbw_tps, nst, a, b, slice_rows, cg, rb).total <= kTcMaxSmem; };
  int a = nsl0, b = nstg0;
  while (a < want_nsl) {
    if (fits(a + 1, b)) { ++a; continue; }
    if (b > min_nstg) { --b; continue; }
    break;
  }
  while (!fits(a, b) && a > nsl0) --a;        // (cannot happen: every accepted step fits)
  if (fits(a, b)) { *nsl = a; *nstg = b; }
}

template <typename T, int C_IN, int N, int MB, int KS, int KBW, int TPS, int NST, int NSL, int NSTG, int CG>
cudaError_t launch_t(const TcArgs& a, ActType act, cudaStream_t s, int* grid_out) {
  constexpr int PW = N < 64 ? N : 64;
  TcKArgs ka{};
  ka.bias = a.bias; ka.res2 = a.res2; ka.n_q = a.n_q; ka.c_out = a.c_out; ka.has_res = a.res != nullptr;
  ka.res_inv = a.res_inv; ka.out_scale = a.out_scale; ka.out_slope = a.out_slope;
  ka.nvar = a.nvar;
  static const bool skip_zero = tc_env_int("HFG_TC_SKIPZ", 1) != 0;       // A/B knob: 0 issues every K-step
  for (int v = 0; v < a.nvar; ++v) { ka.var[v] = a.var[v]; if (!skip_zero) memset(ka.var[v].kz, 0, sizeof ka.var[v].kz); }
  int mn = a.var[0].tap_off[0], mx = mn;
  for (int v = 0; v < a.nvar; ++v)
    for (int j = 0; j < a.var[v].ntaps; ++j) { mn = a.var[v].tap_off[j] < mn ? a.var[v].tap_off[j] : mn; mx = a.var[v].tap_off[j] > mx ? a.var[v].tap_off[j] : mx; }
  ka.min_off = mn; ka.span = mx - mn;
  if (ka.span > kTcMaxSpan) return cudaErrorInvalidValue;
  if ((a.res || a.res2) && a.out_stride != 1) return cudaErrorInvalidValue;
  // the epilogue uses max / min forms of the leaky relu and of its inverse: slopes must lie on the right side of 1
  if (!(a.out_slope > 0.f && a.out_slope <= 1.f) || (a.res && !(a.res_inv >= 1.f))) return cudaErrorInvalidValue;
  tc_boxes(MB * 128 + ka.span, &ka.box_rows, &ka.nbox);
  {
    int max_taps = 0;
    for (int v = 0; v < a.nvar; ++v) max_taps = a.var[v].ntaps > max_taps ? a.var[v].ntaps : max_taps;
    const int stages = (C_IN / KS) * ((max_taps + TPS - 1) / TPS) * (KS / KBW);
    ka.resident = (a.nvar == 1 && a.c_out == N && stages <= NST) ? 1 : 0;
  }
  constexpr int RB = tc_rb(N, MB);
  tc_ring_depths(N, KS, KBW * TPS, NST, NSL, tc_nstg(NSTG, RB), ka.nbox * ka.box_rows, CG, RB, ka.resident != 0, &ka.nsl, &ka.nstg);
  const TcSmemPlan pl = tc_plan(N, KS, KBW * TPS, NST, ka.nsl, ka.nstg, ka.nbox * ka.box_rows, CG, RB);
  if ((size_t)pl.total > kTcMaxSmem) return cudaErrorInvalidConfiguration;
  ka.tiles_q = (a.n_q + CG * MB * 128 - 1) / (CG * MB * 128);
  ka.w_base = a.w_base;
  ka.post_out = a.post_out; ka.post_f = a.post_f; ka.post_pcm = a.post_pcm;
  ka.trace = g_tc_trace;
  static const int stores_env = tc_env_int("HFG_TC_STORES", -1);
  ka.store_in_flight = stores_env > 6 ? 6 : stores_env;
  const bool post_mode = a.post_out || a.post_pcm;
  if (post_mode && (a.post_f < 1 || a.post_f > 16)) return cudaErrorInvalidValue;
  if (post_mode && (a.out_stride != 1 || a.res || a.res2)) return cudaErrorInvalidValue;
  ka.n_tiles = a.c_out / N;
  ka.total_tiles = a.B * ka.tiles_q * a.nvar * ka.n_tiles;
  CUtensorMap tm_x, tm_res, tm_out, tm_w;
  if (CG == 2) {
    if (!a.w_base || a.w_bytes % 256 || (uintptr_t)a.w_base % 256) return cudaErrorInvalidValue;
    if (!make_map_w(&tm_w, a.w_base, a.w_bytes, (N / CG) * KBW * 2 / 256)) return cudaErrorNotSupported;
  }
  if (!make_map3(&tm_x, act, a.x, C_IN, a.x_rows, a.B, KS, ka.box_rows)) return cudaErrorNotSupported;
  if (!make_map4(&tm_out, act, a.out, a.c_out, a.out_stride, a.n_q, a.B, PW, RB * 128)) return cudaErrorNotSupported;
  if (a.res) { if (!make_map4(&tm_res, act, a.res, a.c_out, 1, a.n_q, a.B, PW, RB * 128)) return cudaErrorNotSupported; }
  else tm_res = tm_out;
  if (CG == 1) tm_w = tm_out;
  auto kern = conv_tc_kernel<T, C_IN, N, MB, KS, KBW, TPS, NST, NSL, NSTG, CG>;
  // per instantiation AND per device: function attributes live in the device's context (a process may hold handles
  // on several GPUs even though the intended deployment is one process per GPU)
  constexpr int kMaxDev = 32;
  static int configured_dev[kMaxDev] = {0}, occ_dev[kMaxDev] = {0};
  int dev = 0; cudaGetDevice(&dev);
  if (dev < 0 || dev >= kMaxDev) return cudaErrorInvalidDevice;
  if (pl.total > configured_dev[dev]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, pl.total);
    if (e != cudaSuccess) return e;
    configured_dev[dev] = pl.total;
    int occ = 1;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kTcThreads, pl.total);
    if (e != cudaSuccess) return e;
    constexpr int NACC = tc_nacc(MB, N);
    const int tmem_occ = 512 / (int)tc_tmem_cols(NACC * MB * N);
    if (occ > tmem_occ) occ = tmem_occ;
    occ_dev[dev] = occ < 1 ? 1 : occ;
  }
  int grid = tc_num_sms() * occ_dev[dev];
  static const bool use_pdl = tc_env_int("HFG_PDL", 1) != 0;
  cudaLaunchAttribute at[2];
  int nat = 0;
  if (use_pdl) { at[nat].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[nat].val.programmaticStreamSerializationAllowed = 1; ++nat; }
  cudaLaunchConfig_t lc{};
  lc.blockDim = dim3(kTcThreads); lc.dynamicSmemBytes = pl.total; lc.stream = s; lc.attrs = at;
  if (CG == 1) {
    if (grid > ka.total_tiles) grid = ka.total_tiles;
    if (grid_out) *grid_out = grid;
    lc.gridDim = dim3(grid); lc.numAttrs = nat;
    return cudaLaunchKernelEx(&lc, kern, tm_x, tm_res, tm_out, tm_w, ka);
  }
  // CTA pairs: cluster of 2 (same TPC), one pair per pair-tile stride
  int pairs = tc_num_sms() / 2;
  if (pairs > ka.total_tiles) pairs = ka.total_tiles;
  grid = 2 * pairs;
  if (grid_out) *grid_out = grid;
  at[nat].id = cudaLaunchAttributeClusterDimension;
  at[nat].val.clusterDim.x = 2; at[nat].val.clusterDim.y = 1; at[nat].val.clusterDim.z = 1;
  ++nat;
  lc.gridDim = dim3(grid); lc.numAttrs = nat;
  return cudaLaunchKernelEx(&lc, kern, tm_x, tm_res, tm_out, tm_w, ka);
}

// output channels per tile: at most 128 so that two accumulator buffers fit TMEM (epilogue overlaps the MMAs)
inline int tc_ntile(int c_out) { return c_out > 128 ? 128 : c_out; }

struct TcRow { int c_in, c_out_tile; TcConfig cfg; };
// Stage work at full tensor rate = MB * TPS * (KBW/16) * N/2 cycles (x2 rows for a pair): kept >= 512 so that the
// per-stage barrier round trip of the single issuing thread is amortised.  Rows are tried in order; the first row
// matching (c_in, n) that fits shared memory and is not masked by the debug knob HFG_TC_SKIP (bit i = skip row i) is
// the layer's config.  Measured on B200 (V1, 32 x 10 s): pairs win for N >= 64 (stage0 -12 %, stage1 -14 %, stage2 -3 %)
// and lose for the 32-channel layers (stage3 +20 %: tiles too short to hide the cross-CTA barrier round trips).
//   c_in  n    mb  ks kbw tps nst nsl nstg cg     (cg 2 = CTA pairs, cta_group::2)
#define HFG_TC_TABLE(X)                  \
  X(512, 128, 1, 64, 64, 1, 8, 3, 4, 2)  \
  X(256, 128, 2, 64, 64, 1, 6, 2, 5, 2)  \
  X(128, 128, 2, 64, 64, 1, 6, 2, 5, 2)  \
  X(128, 64, 2, 64, 64, 2, 6, 2, 5, 2)   \
  X(64, 64, 2, 64, 64, 2, 6, 2, 5, 2)    \
  X(512, 128, 1, 64, 64, 1, 3, 3, 4, 1)  \
  X(256, 128, 2, 64, 64, 1, 3, 2, 5, 1)  \
  X(128, 128, 2, 64, 64, 1, 3, 2, 5, 1)  \
  X(128, 64, 2, 64, 64, 2, 3, 2, 5, 1)   \
  X(64, 64, 2, 64, 64, 2, 6, 2, 3, 1)    \
  X(64, 32, 4, 64, 64, 2, 4, 2, 5, 1)    \
  X(32, 32, 4, 32, 32, 4, 4, 3, 8, 1)    \
  X(32, 16, 4, 32, 32, 4, 4, 3, 8, 1)    \
  X(16, 16, 4, 16, 16, 4, 4, 3, 8, 1)    \
  X(64, 16, 4, 64, 64, 2, 4, 2, 8, 1)    \
  X(128, 16, 4, 64, 64, 1, 4, 2, 8, 1)

// Small-problem variants: the same operand staging (identical packed weights: n, ks, kbw, tps, cg) with half the rows
// per tile, used when a layer has fewer tiles than the GPU has SMs (single utterances, the short early stages): twice
// the CTAs at work and half the serial MMA chain per tile.  V1, 1 x 256 frames: 1.09 -> see DESIGN.md section 5.
#define HFG_TC_SMALL(X)                  \
  X(256, 128, 1, 64, 64, 1, 6, 2, 5, 2)  \
  X(128, 128, 1, 64, 64, 1, 6, 2, 5, 2)  \
  X(128, 64, 1, 64, 64, 2, 6, 2, 5, 2)   \
  X(64, 64, 1, 64, 64, 2, 6, 2, 5, 2)    \
  X(64, 32, 2, 64, 64, 2, 4, 2, 5, 1)    \
  X(32, 32, 2, 32, 32, 4, 4, 3, 8, 1)    \
  X(32, 16, 2, 32, 32, 4, 4, 3, 8, 1)    \
  X(16, 16, 2, 16, 16, 4, 4, 3, 8, 1)

const TcRow kTcTable[] = {
#define X(CI, NN, MB, KS, KBW, TPS, NST, NSL, NSTG, CG) {CI, NN, {NN, MB, KS, KBW, TPS, NST, NSL, NSTG, CG}},
    HFG_TC_TABLE(X)
#undef X
};

unsigned tc_skip_mask() {
  static int init = 0; static unsigned mask = 0;
  if (!init) { const char* e = getenv("HFG_TC_SKIP"); mask = e ? (unsigned)strtoul(e, nullptr, 0) : 0u; init = 1; }
  return mask;
}

template <typename T>
cudaError_t launch_shape(const TcArgs& a, ActType act, cudaStream_t s, int* grid_out) {
  const int ci = a.c_in, n = tc_ntile(a.c_out);
  TcConfig c;
  int mn = a.var[0].tap_off[0], mx = mn;
  for (int v = 0; v < a.nvar; ++v)
    for (int j = 0; j < a.var[v].ntaps; ++j) { mn = a.var[v].tap_off[j] < mn ? a.var[v].tap_off[j] : mn; mx = a.var[v].tap_off[j] > mx ? a.var[v].tap_off[j] : mx; }
  if (!tc_config(a.c_in, a.c_out, mx - mn, &c)) return cudaErrorInvalidValue;
  // fewer tiles than SMs -> the half-height variant of the same row, if there is one
  static const bool allow_small = tc_env_int("HFG_TC_SMALL", 1) != 0;
  const long long ctas = (long long)a.B * ((a.n_q + c.cg * c.mb * 128 - 1) / (c.cg * c.mb * 128)) * a.nvar * (a.c_out / c.n) * c.cg;
  if (allow_small && ctas < tc_num_sms()) {
#define X(CI, NN, MB, KS, KBW, TPS, NST, NSL, NSTG, CG) \
    if (ci == CI && n == NN && c.cg == CG && c.nst == NST && c.nstg == NSTG && c.mb == 2 * MB && c.tps == TPS) \
      return launch_t<T, CI, NN, MB, KS, KBW, TPS, NST, NSL, NSTG, CG>(a, act, s, grid_out);
    HFG_TC_SMALL(X)
#undef X
  }
#define X(CI, NN, MB, KS, KBW, TPS, NST, NSL, NSTG, CG) \
  if (ci == CI && n == NN && c.cg == CG && c.nst == NST && c.nstg == NSTG) return launch_t<T, CI, NN, MB, KS, KBW, TPS, NST, NSL, NSTG, CG>(a, act, s, grid_out);
  HFG_TC_TABLE(X)
#undef X
  return cudaErrorInvalidValue;
}

}  // namespace

// first unmasked row for (c_in, n) whose shared-memory plan fits a layer of this tap span (max_off - min_off over all variants)
bool tc_config(int c_in, int c_out, int span, TcConfig* cfg) {
  const int n = tc_ntile(c_out);
  if (c_out % n || c_out > kTcMaxCout || span < 0 || span > kTcMaxSpan) return false;
  const unsigned skip = tc_skip_mask();
  int i = 0;
  for (const TcRow& r : kTcTable) {
    if (r.c_in == c_in && r.c_out_tile == n && !((skip >> i) & 1u)) {
      const TcConfig& c = r.cfg;
      int br, nb;
      tc_boxes(c.mb * 128 + span, &br, &nb);
      const int rb = tc_rb(c.n, c.mb);
      if ((size_t)tc_plan(c.n, c.ks, c.kbw * c.tps, c.nst, c.nsl, tc_nstg(c.nstg, rb), br * nb, c.cg, rb).total <= kTcMaxSmem) {
        if (cfg) *cfg = c;
        return true;
      }
    }
    ++i;
  }
  return false;
}

bool tc_supported(int c_in, int c_out, int ntaps, int span) {
  if (ntaps < 1 || ntaps > kMaxTaps) return false;
  return tc_config(c_in, c_out, span, nullptr);
}

size_t tc_packed_weight_bytes(int c_in, int c_out, int ntaps) { return (size_t)ntaps * c_in * c_out * 2; }

// fp32 [ntaps][c_in][c_out] -> [n-tile][K-slice][tap][k-block][CTA of the pair] blocks of [KBW/8 chunks][N/cg rows][8 x 16-bit]
void tc_pack_weights(const float* w, int ntaps, int c_in, int c_out, int span, ActType act, void* dst_host) {
  TcConfig c;
  if (!tc_config(c_in, c_out, span, &c)) return;
  uint16_t* dst = reinterpret_cast<uint16_t*>(dst_host);
  const int nb = c.n / c.cg;
  size_t o = 0;
  for (int nt = 0; nt < c_out / c.n; ++nt)
    for (int s = 0; s < c_in / c.ks; ++s)
      for (int j = 0; j < ntaps; ++j)
        for (int kb = 0; kb < c.ks / c.kbw; ++kb)
          for (int r = 0; r < c.cg; ++r)
            for (int chunk = 0; chunk < c.kbw / 8; ++chunk)
              for (int row = 0; row < nb; ++row)
                for (int e = 0; e < 8; ++e) {
                  const int ci = s * c.ks + kb * c.kbw + chunk * 8 + e, co = nt * c.n + r * nb + row;
                  const float f = w[((size_t)j * c_in + ci) * c_out + co];
                  uint16_t bits;
                  if (act == ACT_BF16) { __nv_bfloat16 h = __float2bfloat16_rn(f); memcpy(&bits, &h, 2); }
                  else { __half h = __float2half_rn(f); memcpy(&bits, &h, 2); }
                  dst[o++] = bits;
                }
}

void tc_set_trace(long long* dev) { g_tc_trace = dev; }

cudaError_t launch_conv_tc(const TcArgs& a, ActType act, cudaStream_t s, int* grid_out) {
  if (a.nvar < 1 || a.nvar > kMaxVariants) return cudaErrorInvalidValue;
  if (a.B <= 0 || a.n_q <= 0) return cudaSuccess;
  if (act == ACT_F16) return launch_shape<__half>(a, act, s, grid_out);
  if (act == ACT_BF16) return launch_shape<__nv_bfloat16>(a, act, s, grid_out);
  return cudaErrorInvalidValue;
}

}  // namespace hfg
