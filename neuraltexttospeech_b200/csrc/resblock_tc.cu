// Fused ResBlock iteration for the narrow stages (C <= 64) of the 16-bit path: ONE launch computes
//     xt = conv1(lrelu(x))  [dilated]  ->  xt = conv2(lrelu(xt))  [dilation 1]  ->  x' = xt + x
// (reference HiFiGAN_TF/model.py:71-80, one pass of the `for c1, c2 in zip(convs1, convs2)` loop), with the
// intermediate `xt` never leaving the SM.  conv_tc.cu runs one launch per conv: every conv pays a full HBM write and
// re-read of its output and the skip connection is read a second time; here a tile's x slab is read from HBM once, is
// conv1's A operand AND the residual, and only x' is written: 2 instead of 5 layer-boundary tensors per iteration.
// sm_100a only (tcgen05 / TMEM / TMA).
//
// Kernel view: 64 channels per row.  A [L, C] tensor with C < 64 is bit-identical to [L/f, C*f] with f = 64/C (time
// folding); the host hands both convs over as tap lists over folded rows with block-Toeplitz 64x64
// weights, so this kernel only ever sees K = N = 64 and 128-byte (SWIZZLE_128B) rows.
//
// A tile = (utterance b, R2 = MB*128 - span2 output rows).  Per tile:
//   slab   x rows [q0 - h1 - h2, ...) (MB*128 + span1 rows; h = left halo, OOB rows zero = TF padding)  TMA -> smem ring
//   c1     acc1[mb] (TMEM) = sum_taps slab[mb*128 + off1_j ..] . W1_j           tcgen05.mma, M = 128 (256 for a CTA pair)
//   (both convs end with a "bias tap": one K = 16 MMA of a constant ones tile with the bias as hi + lo 16-bit rows, so the
//    epilogues below carry no bias adds)
//   mid    acc1 -> zero outside the utterance (conv2 pads xt with zeros) -> lrelu -> 16-bit -> xt tile in smem,
//          written in the swizzled K-major layout the MMA reads (it is conv2's A operand);
//          the skip connection: slab row -> unactivate -> fp32 -> tcgen05.st into acc2 (the accumulator of conv2
//          starts from the residual, so the slab slot is free again before conv2 even starts)
//   c2     acc2[mb] += sum_taps xt[mb*128 + off2_j ..] . W2_j
//   fin    acc2 -> (+ running MRF sum, * 1/num_kernels on the last resblock: model.py:161-167) -> lrelu for the consumer
//          -> 16-bit, in place in the (now dead) xt tile -> TMA store.  The last span2 rows of a tile are not valid
//          (their taps read past the xt rows computed here) and are not stored: the last block uses a shorter box.
// MODE 2 is a whole ResBlock2 (model.py:106-112: x1 = c_a(lrelu(x)) + x; x2 = c_b(lrelu(x1)) + x1): the mid epilogue adds
// the first skip connection before activating (xt = lrelu(x1)), conv2 starts from zero, and the fin epilogue recovers x1
// from the 16-bit xt tile (exactly what the unfused path reads back from HBM) for the second one.  The xt tile then has
// `lead` spare rows on top so that output row r and its x1 share one shared-memory row at a 1024-byte-aligned block start.
// Roles: warp 0 slab producer, warp 1 weight producer (both convs' weights stay resident), warp 2 MMA issuer (order
// c1(i+1), c2(i): the tensor core works on the next tile's conv1 while the mid epilogue of this one runs), warps 4-15
// three epilogue groups taking mid / fin passes round-robin, warp 16 store issuer.  CG = 2 runs CTA pairs
// (cta_group::2, each CTA holds half of the weight rows).
#include <cuda.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <type_traits>

#include "kernels.cuh"
#include "tc_ptx.cuh"

namespace hfg {

namespace {

constexpr int kRbEpiGroups = 3;
constexpr int kRbStoreWarp = 4 + 4 * kRbEpiGroups;
constexpr int kRbThreads = (kRbStoreWarp + 4) * 32;          // 5 warpgroups: producers + MMA, 3 epilogue groups, store issuer (+3 idle)
constexpr int kRbRegsIo = 64, kRbRegsEpi = 112;              // setmaxnreg split (2 x 64 + 3 x 112 <= 5 x 96)
constexpr int kRbC = 64;                                     // kernel-view channels (K = N)
constexpr int kRbRowB = kRbC * 2;                            // bytes per row = swizzle width
constexpr size_t kRbMaxSmem = 227 * 1024;
constexpr int kRbMaxXt = 3, kRbMaxSl = 3;

struct RbPlan {
  int slab_bytes, xt_rows, xt_bytes, w_bytes;
  int off_xt, off_w, off_bias, off_bar, total;
};

__host__ __device__ inline RbPlan rb_plan(int mb, int cg, int slab_rows, int span2, int ntaps, int nsl, int nxt, int lead = 0) {
  RbPlan p;
  p.slab_bytes = slab_rows * kRbRowB;                         // slab_rows % 8 == 0 -> every slot 1024-byte aligned
  p.xt_rows = mb * 128 + ((lead + span2 + 7) & ~7);
  p.xt_bytes = p.xt_rows * kRbRowB;
  p.w_bytes = ntaps * (kRbC / cg) * kRbRowB;                  // both convs; a CTA of a pair holds half of the weight rows
  int off = nsl * p.slab_bytes;
  p.off_xt = off; off += nxt * p.xt_bytes;
  p.off_w = off; off += p.w_bytes;
  p.off_bias = off; off += 4096 + 2 * (2 * (kRbC / cg) * 16);   // "ones" A tile [2 K-chunks][128 rows][16 B] + per conv a bias B tile [2][64/cg rows][16 B]
  p.off_bar = off; off += (2 * kRbMaxSl + 2 + 8 + kRbMaxXt + kRbMaxXt * 2) * 8 + 16;
  p.total = off;
  return p;
}

struct RbKArgs {
  const float* bias1; const float* bias2;
  const void* res2;                    // running raw MRF sum (16-bit, indexed like the output) or nullptr
  const void* w1; const void* w2;      // CG == 1: packed weights (rb_pack_weights); CG == 2: read through tm_w1 / tm_w2
  int n_q, B;                          // kernel-view rows per utterance, utterances
  int ntaps1, ntaps2;
  int off1[kMaxTaps], off2[kMaxTaps];  // kernel-view row offsets of the taps
  unsigned char kz1[kMaxTaps], kz2[kMaxTaps];   // per tap: bit kk = K-step kk multiplies an all-zero weight block (skipped)
  int min1, min2, span2;
  int r2;                              // valid output rows per CTA tile: MB*128 - span2
  int tiles_q, total_tiles;            // per utterance / in total (CG == 2: PAIR tiles of 2*r2 rows)
  int box_rows, nbox;                  // slab = nbox TMA boxes of box_rows rows
  int nsl, nxt;                        // ring depths (slab slots, xt tiles)
  int lead;                            // MODE 2: spare rows on top of the xt tile (0 in MODE 1)
  float mid_slope, xt_inv, res_inv, out_scale, out_slope;   // xt_inv = 1 / mid_slope
  int two_issuers;                     // 1: warp 2 issues the conv1 MMAs and warp 3 the conv2 MMAs; 0: warp 2 issues both, c1(i+1) then c2(i)
  long long* trace;                    // HFG_TC_TRACE builds: [7 roles][64 tiles][8 events] clock64 stamps of CTA 0, else unused
};

// Debug build only (-DHFG_TC_TRACE, tools/trace_resblock.py): CTA 0 stamps clock64() per role / tile / event.
#ifdef HFG_TC_TRACE
#define RB_TRACE(role, tile_i, ev) \
  do { if (ka.trace && blockIdx.x == 0 && (uint32_t)(tile_i) < 64u) ka.trace[((role) * 64u + (uint32_t)(tile_i)) * 8u + (ev)] = clock64(); } while (0)
#else
#define RB_TRACE(role, tile_i, ev) do { } while (0)
#endif

template <typename T, int MB, int CG, int MODE>
__global__ void __launch_bounds__(kRbThreads, 1)
resblock_tc_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_out,
                   const __grid_constant__ CUtensorMap tm_out_last, const __grid_constant__ CUtensorMap tm_w1,
                   const __grid_constant__ CUtensorMap tm_w2, const __grid_constant__ RbKArgs ka) {
  constexpr int NB = kRbC / CG;                               // weight rows staged by this CTA
  constexpr int TAP_BYTES = NB * kRbRowB;                     // one tap's 64 x NB block
  constexpr uint32_t TMEM_COLS = MB == 2 ? 512 : 256;         // acc1[2][MB] + acc2[2][MB], 64 columns each
  constexpr uint32_t IDESC = make_idesc(kRbC, std::is_same<T, __nv_bfloat16>::value, 128 * CG);
  constexpr int EG = kRbEpiGroups;
  const int nsl = ka.nsl, nxt = ka.nxt;
  const int ntaps = ka.ntaps1 + ka.ntaps2;
  const RbPlan pl = rb_plan(MB, CG, ka.nbox * ka.box_rows, ka.span2, ntaps, nsl, nxt, ka.lead);
  const int lead = MODE == 2 ? ka.lead : 0;

  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* slab = smem;
  uint8_t* xt = smem + pl.off_xt;
  uint8_t* wsm = smem + pl.off_w;
  // Biases ride the tensor core: one extra K = 16 MMA per conv and block multiplies a constant A tile whose first two K
  // columns are 1 with a B tile whose first two K rows hold the bias as hi + lo 16-bit parts (exact to ~2^-22): the
  // epilogues carry no bias adds (64 FADD + 16 shared loads per 32 columns before).
  uint8_t* ones_s = smem + pl.off_bias;
  uint8_t* biasb_s = ones_s + 4096;                                      // conv1 tile, then conv2 tile (2 * NB * 16 bytes each)
  uint64_t* slab_full = reinterpret_cast<uint64_t*>(smem + pl.off_bar);
  uint64_t* slab_empty = slab_full + kRbMaxSl;
  uint64_t* w_full = slab_empty + kRbMaxSl;            // [2]: conv1 / conv2 weights landed
  uint64_t* acc1_full = w_full + 2;                    // [2]
  uint64_t* mid_done = acc1_full + 2;                  // [2] (leader's copy is the one waited on)
  uint64_t* acc2_full = mid_done + 2;                  // [2]
  uint64_t* acc2_free = acc2_full + 2;                 // [2]
  uint64_t* xt_free = acc2_free + 2;                   // [kRbMaxXt]
  uint64_t* out_done = xt_free + kRbMaxXt;             // [kRbMaxXt][2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(out_done + kRbMaxXt * 2);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const uint32_t rank = CG == 2 ? cluster_ctarank() : 0u;
  pdl_launch_dependents();
  const int tile0 = blockIdx.x / CG, tile_step = gridDim.x / CG;
  const int my_tiles = tile0 < ka.total_tiles ? (ka.total_tiles - tile0 + tile_step - 1) / tile_step : 0;

  if (tid == 0) {
    for (int s = 0; s < kRbMaxSl; ++s) { mbar_init(&slab_full[s], 1); mbar_init(&slab_empty[s], 4 * MB); }
    for (int s = 0; s < 2; ++s) {
      mbar_init(&w_full[s], 1); mbar_init(&acc1_full[s], 1); mbar_init(&mid_done[s], 4 * MB * CG);
      mbar_init(&acc2_full[s], 1); mbar_init(&acc2_free[s], 4 * MB);
    }
    for (int s = 0; s < kRbMaxXt; ++s) { mbar_init(&xt_free[s], 1); mbar_init(&out_done[2 * s], 4); mbar_init(&out_done[2 * s + 1], 4); }
    fence_mbar_init();
    tma_prefetch_desc(&tm_x); tma_prefetch_desc(&tm_out); tma_prefetch_desc(&tm_out_last);
    if (CG == 2) { tma_prefetch_desc(&tm_w1); tma_prefetch_desc(&tm_w2); }
  }
  if (warp == 2) { if (CG == 2) tmem_alloc2(tmem_slot, TMEM_COLS); else tmem_alloc(tmem_slot, TMEM_COLS); }
  {
    // ones tile: element (row r, k) at chunk (k / 8) * 2048 + r * 16 + (k % 8) * 2; k = 0, 1 -> 1.0, everything else 0
    const uint32_t one2 = std::is_same<T, __nv_bfloat16>::value ? 0x3F803F80u : 0x3C003C00u;
    if (tid < 256) *reinterpret_cast<uint4*>(ones_s + tid * 16) = make_uint4(tid < 128 ? one2 : 0u, 0u, 0u, 0u);
    // bias tiles: this CTA's NB output channels; element (k, n) at chunk (k / 8) * NB * 16 + n * 16 + (k % 8) * 2
    if (tid >= 256 && tid < 256 + 2 * 2 * NB) {
      const int e = tid - 256, c = e / (2 * NB), rem = e % (2 * NB), chunk = rem / NB, n = rem % NB;
      uint4 u = make_uint4(0u, 0u, 0u, 0u);
      if (chunk == 0) {
        const float bv = (c == 0 ? ka.bias1 : ka.bias2)[(int)rank * NB + n];
        float hi, lo;
        if (std::is_same<T, __nv_bfloat16>::value) { hi = __bfloat162float(__float2bfloat16_rn(bv)); lo = bv - hi; }
        else { hi = __half2float(__float2half_rn(bv)); lo = bv - hi; }
        u.x = H2<T>::down(hi, lo);
      }
      *reinterpret_cast<uint4*>(biasb_s + c * (2 * NB * 16) + chunk * NB * 16 + n * 16) = u;
    }
    fence_proxy_async();                                                 // generic writes -> visible to the tensor core
  }
  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (warp != 1) pdl_wait();                            // the weight producer reads constants only

  auto tile_q0 = [&](int t, int& b) -> int {            // first output row of this CTA's part of tile t
    b = t / ka.tiles_q;
    const int qt = t % ka.tiles_q;
    return (qt * CG + (int)rank) * ka.r2;
  };
  const int halo = -(ka.min1 + ka.min2);                // slab row of output row 0 (left halo of both convs)

  if (warp < 4) {
    regs_release<kRbRegsIo>();
    if (warp == 0) {
      // ===================== slab producer =====================
      const uint32_t box_bytes = (uint32_t)ka.box_rows * kRbRowB;
      for (int i = 0; i < my_tiles; ++i) {
        int b; const int q0 = tile_q0(tile0 + i * tile_step, b);
        const int pos = i % nsl;
        mbar_wait(&slab_empty[pos], ((i / nsl) & 1) ^ 1);
        if (lane == 0) RB_TRACE(0u, i, 0u);                          // slot free, load requested
        if (elect_one()) {
          uint8_t* dst = slab + pos * pl.slab_bytes;
          if (CG == 1) {
            mbar_expect_tx(&slab_full[pos], box_bytes * ka.nbox);
            for (int bx = 0; bx < ka.nbox; ++bx)
              tma_load_3d(dst + bx * box_bytes, &tm_x, 0, q0 - halo + bx * ka.box_rows, b, &slab_full[pos]);
          } else {
            if (rank == 0) mbar_expect_tx(&slab_full[pos], 2 * box_bytes * ka.nbox);
            const uint32_t bar_cl = mapa_u32(smem_u32(&slab_full[pos]), 0);
            for (int bx = 0; bx < ka.nbox; ++bx)
              tma_load_3d_2cta(dst + bx * box_bytes, &tm_x, 0, q0 - halo + bx * ka.box_rows, b, bar_cl);
          }
        }
        __syncwarp();
      }
    } else if (warp == 1) {
      // ===================== weight producer: both convs, once =====================
      if (my_tiles > 0 && elect_one()) {
        for (int c = 0; c < 2; ++c) {
          const int nt = c == 0 ? ka.ntaps1 : ka.ntaps2;
          uint8_t* dst = wsm + (c == 0 ? 0 : ka.ntaps1 * TAP_BYTES);
          if (CG == 1) {
            mbar_expect_tx(&w_full[c], (uint32_t)nt * TAP_BYTES);
            const uint8_t* src = reinterpret_cast<const uint8_t*>(c == 0 ? ka.w1 : ka.w2);
            for (int j = 0; j < nt; ++j) bulk_g2s(dst + j * TAP_BYTES, src + (size_t)j * TAP_BYTES, TAP_BYTES, &w_full[c]);
          } else {
            if (rank == 0) mbar_expect_tx(&w_full[c], 2u * nt * TAP_BYTES);
            const uint32_t bar_cl = mapa_u32(smem_u32(&w_full[c]), 0);
            // packed as [tap][CTA of the pair] blocks of TAP_BYTES, viewed by the map as 256-byte rows
            for (int j = 0; j < nt; ++j)
              tma_load_2d_2cta(dst + j * TAP_BYTES, c == 0 ? &tm_w1 : &tm_w2, 0, (2 * j + (int)rank) * (TAP_BYTES >> 8), bar_cl);
          }
        }
      }
      __syncwarp();
    } else {
      // ===================== MMA issuers (pair: leader only): warp 2 issues every conv1, warp 3 every conv2 =====================
      // One thread issues an MMA every ~60 clocks (descriptor arithmetic + tcgen05.mma), twice the 32 clocks an M = 256,
      // N = 64, K = 16 MMA keeps the tensor pipe busy: a single issuer caps the pipe at ~50 % (role timeline,
      // profiles/r2e_fused_issue.md).  conv1(i+1) and conv2(i) are independent streams, so each gets its own warp.
      if (rank == 0 && my_tiles > 0) {
        const uint32_t slab_addr = smem_u32(slab), xt_addr = smem_u32(xt), w_addr = smem_u32(wsm);
        auto issue = [&](uint32_t a_base, uint32_t w_base, const int* offs, const unsigned char* kz, int mn, int nt, uint32_t tacc, bool fresh, int bias_idx) {   // mn: offset of smem row 0
          if (elect_one()) {
#pragma unroll 1
            for (int mb = 0; mb < MB; ++mb) {
              uint32_t first = fresh ? 1u : 0u;
#pragma unroll 1
              for (int j = 0; j < nt; ++j) {
                // descriptor low words (start address / 16, LBO); the high words are compile-time constants
                constexpr uint32_t A_HI = ((8u * kRbRowB) >> 4) | (1u << 14) | (2u << 29);   // make_desc_sw<128>: SBO, version, SWIZZLE_128B
                constexpr uint32_t B_HI = (128u >> 4) | (1u << 14);                          // make_desc(.., sbo = 128)
                const uint32_t a0 = (uint32_t)make_desc_sw<kRbRowB>(a_base + (uint32_t)(mb * 128 + offs[j] - mn) * kRbRowB);
                const uint32_t b0 = (uint32_t)make_desc(w_base + j * TAP_BYTES, NB * 16, 128u);
                const uint32_t zero = kz[j];
#pragma unroll
                for (int kk = 0; kk < kRbC / 16; ++kk) {
                  if ((zero >> kk) & 1u) continue;                   // an all-zero 16 x 64 weight block (time folding): nothing to add
                  const uint32_t adesc = a0 + (uint32_t)((kk * 32) >> 4);
                  const uint32_t bdesc = b0 + (uint32_t)((kk * 2 * NB * 16) >> 4);
                  if (CG == 2) umma_f16_2cta_lo<A_HI, B_HI>(tacc + mb * kRbC, adesc, bdesc, IDESC, first ? 0u : 1u);
                  else umma_f16_lo<A_HI, B_HI>(tacc + mb * kRbC, adesc, bdesc, IDESC, first ? 0u : 1u);
                  first = 0;
                }
              }
              // bias tap: acc += ones[128 x 16] . biasB[16 x 64]
              const uint64_t oa = make_desc(smem_u32(ones_s), 2048u, 128u);
              const uint64_t ob = make_desc(smem_u32(biasb_s) + (uint32_t)bias_idx * (2 * NB * 16), NB * 16, 128u);
              if (CG == 2) umma_f16_2cta(tacc + mb * kRbC, oa, ob, IDESC, first ? 0u : 1u);
              else umma_f16(tacc + mb * kRbC, oa, ob, IDESC, first ? 0u : 1u);
            }
          }
          __syncwarp();
        };
        auto commit = [&](uint64_t* bar) {
          if (elect_one()) { if (CG == 2) umma_commit_2cta(bar); else umma_commit(bar); }
          __syncwarp();
        };
        auto conv1 = [&](int i) {
          const int pos = i % nsl;
          if (lane == 0) RB_TRACE(1u, i, 0u);                        // conv1's turn
          mbar_wait(&slab_full[pos], (i / nsl) & 1);
          if (lane == 0) RB_TRACE(1u, i, 1u);                        // slab landed
          tc_fence_after();
          issue(slab_addr + pos * pl.slab_bytes, w_addr, ka.off1, ka.kz1, ka.min1, ka.ntaps1, tmem_base + (i & 1) * (MB * kRbC), true, 0);
          commit(&acc1_full[i & 1]);
          if (lane == 0) RB_TRACE(1u, i, 2u);                        // conv1 issued
        };
        auto conv2 = [&](int i) {
          if (lane == 0) RB_TRACE(1u, i, 3u);                        // conv2's turn
          mbar_wait(&mid_done[i & 1], (i >> 1) & 1);      // xt tile written, acc2 pre-loaded with the residual, acc1 drained
          if (lane == 0) RB_TRACE(1u, i, 4u);                        // mid pass(es) done
          tc_fence_after();
          issue(xt_addr + (i % nxt) * pl.xt_bytes, w_addr + ka.ntaps1 * TAP_BYTES, ka.off2, ka.kz2, ka.min2 - lead, ka.ntaps2,
                tmem_base + (2 * MB + (i & 1) * MB) * kRbC, MODE == 2, 1);
          commit(&acc2_full[i & 1]);
          if (lane == 0) RB_TRACE(1u, i, 5u);                        // conv2 issued
        };
        if (!ka.two_issuers) {
          if (warp == 2) {                                  // one issuer, order c1(i+1), c2(i)
            mbar_wait(&w_full[0], 0);
            tc_fence_after();
            conv1(0);
            mbar_wait(&w_full[1], 0);
            tc_fence_after();
            for (int i = 0; i < my_tiles; ++i) {
              if (i + 1 < my_tiles) conv1(i + 1);           // acc1[(i+1)&1] was drained by mid(i-1): implied by mid_done(i-1)
              conv2(i);
            }
          }
        } else if (warp == 2) {
          mbar_wait(&w_full[0], 0);
          tc_fence_after();
          for (int i = 0; i < my_tiles; ++i) {
            if (i >= 2) {                                   // acc1[i&1] must have been drained by mid(i-2)
              mbar_wait(&mid_done[i & 1], ((i - 2) >> 1) & 1);
              tc_fence_after();
            }
            conv1(i);
          }
        } else {
          mbar_wait(&w_full[1], 0);
          tc_fence_after();
          for (int i = 0; i < my_tiles; ++i) conv2(i);
        }
      }
    }
  } else if (warp >= kRbStoreWarp) {
    regs_release<kRbRegsIo>();
    if (warp == kRbStoreWarp) {
      // ===================== store issuer =====================
      // one bulk group per tile
      const int out_row0 = MODE == 2 ? lead - ka.min2 : 0;    // xt row holding output row 0 (multiple of 8)
      for (int i = 0; i < my_tiles; ++i) {
        int b; const int q0 = tile_q0(tile0 + i * tile_step, b);
        const int buf = i % nxt;
        for (int mb = 0; mb < MB; ++mb) {
          mbar_wait(&out_done[2 * buf + mb], (i / nxt) & 1);
          if (lane == 0) RB_TRACE(2u, i, (uint32_t)mb);               // output block ready, store issued
          if (elect_one()) {
            const int q = q0 + mb * 128;
            if (q < ka.n_q) tma_store_3d(mb == MB - 1 ? &tm_out_last : &tm_out, xt + buf * pl.xt_bytes + (mb * 128 + out_row0) * kRbRowB, 0, q, b);
          }
          __syncwarp();
        }
        // The tile is handed back as soon as ITS stores have been read out (~1 us; the store warp has nothing else to
        // do).  Deferring the hand-back to the next tile's stores would make mid(i+1) wait for fin(i), which the
        // epilogue groups run in the opposite order.
        if (elect_one()) {
          bulk_commit();
          bulk_wait_read<0>();
          mbar_arrive(&xt_free[buf]);
        }
        __syncwarp();
        if (lane == 0) RB_TRACE(2u, i, 2u);                          // stores read out, tile handed back
      }
      if (elect_one()) bulk_wait_all();
      __syncwarp();
    }
  } else {
    // ===================== epilogue groups: mid / fin passes round-robin =====================
    regs_acquire<kRbRegsEpi>();
    const int grp = (warp - 4) >> 2;
    const int quarter = warp & 3;
    const int er = quarter * 32 + lane;                  // row of the 128-row block owned by this thread (= TMEM lane)
    const uint32_t lane_base = ((uint32_t)(quarter * 32)) << 16;
    const int lag = nxt >= 2 ? 1 : 0;                    // fin(i) runs after mid(i + lag): see the ordering note in DESIGN.md
    const float mid_slope = ka.mid_slope, res_inv = ka.res_inv, osc = ka.out_scale, osl = ka.out_slope;
    uint32_t pcnt = 0;
    for (int s = 0; s < my_tiles + lag; ++s) {
      // ---------------- mid(i): acc1 -> xt tile; residual -> acc2
      if (s < my_tiles) {
        const int i = s;
        int b; const int q0 = tile_q0(tile0 + i * tile_step, b);
        const int xbuf = i % nxt, sbuf = i % nsl;
#pragma unroll 1
        for (int mb = 0; mb < MB; ++mb, ++pcnt) {
          if ((int)(pcnt % (uint32_t)EG) != grp) continue;
          if (quarter == 0 && lane == 0) { RB_TRACE(3u + mb, i, 0u); RB_TRACE(3u + mb, i, 4u + grp); }
          mbar_wait(&acc1_full[i & 1], (i >> 1) & 1);                 // conv1 of the tile complete (hence its slab landed)
          if (quarter == 0 && lane == 0) RB_TRACE(3u + mb, i, 1u);
          tc_fence_after();
          mbar_wait(&xt_free[xbuf], ((i / nxt) & 1) ^ 1);            // the tile's previous tenant has been stored
          mbar_wait(&acc2_free[i & 1], ((i >> 1) & 1) ^ 1);          // fin(i-2) has read acc2[i&1]
          if (quarter == 0 && lane == 0) RB_TRACE(3u + mb, i, 2u);
          tc_fence_after();
          const int r = mb * 128 + er;                                // row of the CTA tile
          const int g = q0 + ka.min2 + r;                             // kernel-view row of this xt row (xt row 0 = q0 - h2)
          const bool inside = g >= 0 && g < ka.n_q;
          const bool tile_inside = q0 + ka.min2 >= 0 && q0 + ka.min2 + MB * 128 <= ka.n_q;   // warp-uniform: no row of the tile needs zeroing
          // MODE 1: slab row of OUTPUT row r (skip connection of the iteration, goes into acc2);
          // MODE 2: slab row of XT row r (skip connection of the first conv, added before the activation)
          const int sr = MODE == 2 ? r - ka.min1 : r + halo;
          const int xr = r + lead;                                    // shared-memory row of xt row r
          uint8_t* xrow = xt + xbuf * pl.xt_bytes + xr * kRbRowB;
          const uint8_t* srow = slab + sbuf * pl.slab_bytes + sr * kRbRowB;
          const uint32_t t1 = tmem_base + lane_base + ((i & 1) * MB + mb) * kRbC;
          const uint32_t t2 = tmem_base + lane_base + (2 * MB + (i & 1) * MB + mb) * kRbC;
#pragma unroll
          for (int c0 = 0; c0 < kRbC; c0 += 32) {
            uint32_t acc[32];
            tmem_ld32(t1 + c0, acc);
            uint4 rr[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) rr[k] = *reinterpret_cast<const uint4*>(srow + (((c0 / 8 + k) ^ (sr & 7)) * 16));
            tmem_ld_wait();
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              float v[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(acc[8 * k + e]);
              if (MODE == 2) {                                        // x1 = c_a(lrelu(x)) + x   (model.py:108-111)
                float2 f;
                f = H2<T>::up(rr[k].x); v[0] += lrelu_ge1(f.x, res_inv); v[1] += lrelu_ge1(f.y, res_inv);
                f = H2<T>::up(rr[k].y); v[2] += lrelu_ge1(f.x, res_inv); v[3] += lrelu_ge1(f.y, res_inv);
                f = H2<T>::up(rr[k].z); v[4] += lrelu_ge1(f.x, res_inv); v[5] += lrelu_ge1(f.y, res_inv);
                f = H2<T>::up(rr[k].w); v[6] += lrelu_ge1(f.x, res_inv); v[7] += lrelu_ge1(f.y, res_inv);
              }
#pragma unroll
              for (int e = 0; e < 8; ++e) v[e] = lrelu_le1(v[e], mid_slope);
              uint4 u;
              u.x = H2<T>::down(v[0], v[1]); u.y = H2<T>::down(v[2], v[3]); u.z = H2<T>::down(v[4], v[5]); u.w = H2<T>::down(v[6], v[7]);
              if (!tile_inside) { if (!inside) u = make_uint4(0u, 0u, 0u, 0u); }   // conv2 sees zero padding outside the utterance (edge tiles only)
              *reinterpret_cast<uint4*>(xrow + (((c0 / 8 + k) ^ (xr & 7)) * 16)) = u;
            }
            if (MODE == 1) {
              // skip connection (model.py:79): x is stored activated; acc2 starts from unact(x) + b2
              uint32_t init[32];
#pragma unroll
              for (int k = 0; k < 4; ++k) {
                float2 f;
                f = H2<T>::up(rr[k].x); init[8 * k + 0] = __float_as_uint(lrelu_ge1(f.x, res_inv)); init[8 * k + 1] = __float_as_uint(lrelu_ge1(f.y, res_inv));
                f = H2<T>::up(rr[k].y); init[8 * k + 2] = __float_as_uint(lrelu_ge1(f.x, res_inv)); init[8 * k + 3] = __float_as_uint(lrelu_ge1(f.y, res_inv));
                f = H2<T>::up(rr[k].z); init[8 * k + 4] = __float_as_uint(lrelu_ge1(f.x, res_inv)); init[8 * k + 5] = __float_as_uint(lrelu_ge1(f.y, res_inv));
                f = H2<T>::up(rr[k].w); init[8 * k + 6] = __float_as_uint(lrelu_ge1(f.x, res_inv)); init[8 * k + 7] = __float_as_uint(lrelu_ge1(f.y, res_inv));
              }
              tmem_st32(t2 + c0, init);
            }
          }
          if (MODE == 1) tmem_st_wait();
          tc_fence_before();
          fence_proxy_async();                                        // xt rows -> visible to the tensor core (async proxy)
          __syncwarp();
          if (lane == 0) {
            mbar_arrive(&slab_empty[sbuf]);                           // slab: conv1 done (acc1_full) and residual rows read
            if (CG == 2) mbar_arrive_cluster(mapa_u32(smem_u32(&mid_done[i & 1]), 0));
            else mbar_arrive(&mid_done[i & 1]);
          }
          if (quarter == 0 && lane == 0) RB_TRACE(3u + mb, i, 3u);
        }
      }
      // ---------------- fin(i): acc2 -> output rows, in place in the xt tile
      if (s >= lag) {
        const int i = s - lag;
        int b; const int q0 = tile_q0(tile0 + i * tile_step, b);
        const int xbuf = i % nxt;
#pragma unroll 1
        for (int mb = 0; mb < MB; ++mb, ++pcnt) {
          if ((int)(pcnt % (uint32_t)EG) != grp) continue;
          const int r = mb * 128 + er;
          const int q = q0 + r;
          const bool has_res2 = ka.res2 != nullptr && r < ka.r2 && q < ka.n_q;
          // the running MRF sum comes from global memory: request it before waiting for the accumulator
          uint4 rr2[8];
          if (has_res2) {
            const uint4* rp = reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(ka.res2) + ((size_t)b * ka.n_q + q) * kRbC);
#pragma unroll
            for (int k = 0; k < 8; ++k) rr2[k] = rp[k];
          }
          if (quarter == 0 && lane == 0) { RB_TRACE(5u + mb, i, 0u); RB_TRACE(5u + mb, i, 4u + grp); }
          mbar_wait(&acc2_full[i & 1], (i >> 1) & 1);                 // conv2 of the whole tile complete: xt rows are dead
          if (quarter == 0 && lane == 0) RB_TRACE(5u + mb, i, 1u);
          tc_fence_after();
          const int orr = MODE == 2 ? r + lead - ka.min2 : r;        // MODE 2: the row of x1 (second skip connection), reused for the output
          uint8_t* orow = xt + xbuf * pl.xt_bytes + orr * kRbRowB;
          const uint32_t t2 = tmem_base + lane_base + (2 * MB + (i & 1) * MB + mb) * kRbC;
#pragma unroll
          for (int c0 = 0; c0 < kRbC; c0 += 32) {
            uint32_t acc[32];
            tmem_ld32(t2 + c0, acc);
            uint4 rx[4];
            if (MODE == 2) {
#pragma unroll
              for (int k = 0; k < 4; ++k) rx[k] = *reinterpret_cast<const uint4*>(orow + (((c0 / 8 + k) ^ (orr & 7)) * 16));
            }
            tmem_ld_wait();
#pragma unroll
            for (int k = 0; k < 4; ++k) {
              float v[8];
#pragma unroll
              for (int e = 0; e < 8; ++e) v[e] = __uint_as_float(acc[8 * k + e]);
              if (MODE == 2) {                                        // x2 = c_b(lrelu(x1)) + x1, x1 recovered from the 16-bit xt tile
                const float xi = ka.xt_inv;
                float2 f;
                f = H2<T>::up(rx[k].x); v[0] += lrelu_ge1(f.x, xi); v[1] += lrelu_ge1(f.y, xi);
                f = H2<T>::up(rx[k].y); v[2] += lrelu_ge1(f.x, xi); v[3] += lrelu_ge1(f.y, xi);
                f = H2<T>::up(rx[k].z); v[4] += lrelu_ge1(f.x, xi); v[5] += lrelu_ge1(f.y, xi);
                f = H2<T>::up(rx[k].w); v[6] += lrelu_ge1(f.x, xi); v[7] += lrelu_ge1(f.y, xi);
              }
              if (has_res2) {
                float2 f;
                const uint4 s4 = rr2[c0 / 8 + k];
                f = H2<T>::up(s4.x); v[0] += f.x; v[1] += f.y;
                f = H2<T>::up(s4.y); v[2] += f.x; v[3] += f.y;
                f = H2<T>::up(s4.z); v[4] += f.x; v[5] += f.y;
                f = H2<T>::up(s4.w); v[6] += f.x; v[7] += f.y;
              }
              if (osc != 1.0f) {
#pragma unroll
                for (int e = 0; e < 8; ++e) v[e] *= osc;
              }
#pragma unroll
              for (int e = 0; e < 8; ++e) v[e] = lrelu_le1(v[e], osl);
              uint4 u;
              u.x = H2<T>::down(v[0], v[1]); u.y = H2<T>::down(v[2], v[3]); u.z = H2<T>::down(v[4], v[5]); u.w = H2<T>::down(v[6], v[7]);
              *reinterpret_cast<uint4*>(orow + (((c0 / 8 + k) ^ (orr & 7)) * 16)) = u;
            }
          }
          tc_fence_before();
          fence_proxy_async();                                        // output rows -> visible to the TMA engine
          __syncwarp();
          if (lane == 0) {
            mbar_arrive(&out_done[2 * xbuf + mb]);
            mbar_arrive(&acc2_free[i & 1]);
          }
          if (quarter == 0 && lane == 0) RB_TRACE(5u + mb, i, 3u);
        }
      }
    }
  }
  tc_fence_before();
  if (CG == 2) cluster_sync_all(); else __syncthreads();
  if (warp == 2) {
    __syncwarp(); tc_fence_after();
    if (CG == 2) tmem_dealloc2(tmem_base, TMEM_COLS); else tmem_dealloc(tmem_base, TMEM_COLS);
  }
}

// ------------------------------------------------------------------ host side
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn rb_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// rank-3 map over a [B][rows][64] 16-bit tensor, box = [64][box_rows][1], 128-byte swizzle
bool rb_map3(CUtensorMap* m, ActType act, const void* base, int rows, int B, int box_rows) {
  EncodeTiledFn fn = rb_encode_fn();
  if (!fn) return false;
  cuuint64_t gdim[3] = {(cuuint64_t)kRbC, (cuuint64_t)rows, (cuuint64_t)B};
  cuuint64_t gstr[2] = {(cuuint64_t)kRbRowB, (cuuint64_t)kRbRowB * (cuuint64_t)rows};
  cuuint32_t box[3] = {(cuuint32_t)kRbC, (cuuint32_t)box_rows, 1};
  cuuint32_t est[3] = {1, 1, 1};
  return fn(m, act == ACT_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 3, const_cast<void*>(base), gdim, gstr,
            box, est, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
            CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

bool rb_map_w(CUtensorMap* m, const void* base, size_t bytes, int box_rows) {
  EncodeTiledFn fn = rb_encode_fn();
  if (!fn) return false;
  cuuint64_t gdim[2] = {128, (cuuint64_t)(bytes / 256)};
  cuuint64_t gstr[1] = {256};
  cuuint32_t box[2] = {128, (cuuint32_t)box_rows};
  cuuint32_t est[2] = {1, 1};
  return fn(m, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, const_cast<void*>(base), gdim, gstr, box, est, CU_TENSOR_MAP_INTERLEAVE_NONE,
            CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

long long* g_rb_trace = nullptr;       // device buffer for HFG_TC_TRACE builds (hfgdbg_set_trace)

int rb_env_int(const char* name, int dflt) { const char* e = getenv(name); return e ? atoi(e) : dflt; }

void rb_boxes(int rows, int* box_rows, int* nbox) {
  int best_br = 256, best_pad = 1 << 30;
  for (int br = 256; br >= 32; br >>= 1) {              // fewest padded rows with boxes of >= 4 KB; ties -> larger box
    const int nb = (rows + br - 1) / br;
    if (nb > 24) break;
    if (nb * br < best_pad) { best_pad = nb * br; best_br = br; }
  }
  *box_rows = best_br; *nbox = (rows + best_br - 1) / best_br;
}

struct RbShape { int mb, cg, nsl, nxt, box_rows, nbox, min1, min2, span1, span2, lead; };

void rb_minmax(const int* off, int n, int* mn, int* mx) {
  *mn = off[0]; *mx = off[0];
  for (int j = 0; j < n; ++j) { *mn = off[j] < *mn ? off[j] : *mn; *mx = off[j] > *mx ? off[j] : *mx; }
}

// Tile height and ring depths for a tap configuration.  Preference: two 128-row blocks per tile with the xt tile double
// buffered (mid epilogue of tile i+1 overlaps conv2 / fin of tile i); when the resident weights leave no room (k = 11),
// one block per tile with deep rings; as a last resort two blocks and a single xt tile.
bool rb_choose(const RbArgs& a, RbShape* out) {
  if (a.ntaps1 < 1 || a.ntaps2 < 1 || a.ntaps1 > kMaxTaps || a.ntaps2 > kMaxTaps) return false;
  RbShape s{};
  int mx1, mx2;
  rb_minmax(a.off1, a.ntaps1, &s.min1, &mx1);
  rb_minmax(a.off2, a.ntaps2, &s.min2, &mx2);
  if (s.min1 > 0 || s.min2 > 0 || mx1 < 0 || mx2 < 0) return false;      // both convs see their own row (zero offset inside the taps' hull)
  s.span1 = mx1 - s.min1; s.span2 = mx2 - s.min2;
  if (s.span2 >= 100) return false;                      // the last block of a tile must keep some valid rows
  if (a.mode != 1 && a.mode != 2) return false;
  s.lead = a.mode == 2 ? (8 - ((-s.min2) % 8)) % 8 : 0;     // MODE 2: output row 0 / its x1 sit on an 8-row boundary of the xt tile
  static const int force_mb = rb_env_int("HFG_RB_MB", 0), force_cg = rb_env_int("HFG_RB_CG", 0), force_nxt = rb_env_int("HFG_RB_NXT", 0);
  const int ntaps = a.ntaps1 + a.ntaps2;
  struct Try { int mb, nxt_min; };
  const Try order[] = {{2, 2}, {1, 2}, {2, 1}, {1, 1}};
  for (int cg = 2; cg >= 1; --cg) {
    if (force_cg && cg != force_cg) continue;
    for (const Try& t : order) {
      if (force_mb && t.mb != force_mb) continue;
      const int halo = -(s.min1 + s.min2);
      const int need = t.mb * 128 + (a.mode == 2 ? s.span1 : (s.span1 > halo ? s.span1 : halo));
      int br, nb; rb_boxes(need, &br, &nb);
      int nsl = 2, nxt = t.nxt_min;
      if (force_nxt) { if (force_nxt < t.nxt_min) continue; nxt = force_nxt; }
      if ((size_t)rb_plan(t.mb, cg, br * nb, s.span2, ntaps, nsl, nxt, s.lead).total > kRbMaxSmem) continue;
      // spend what is left on a third xt tile, then a third slab slot
      if (!force_nxt && nxt < kRbMaxXt && nxt >= 2 && (size_t)rb_plan(t.mb, cg, br * nb, s.span2, ntaps, nsl, nxt + 1, s.lead).total <= kRbMaxSmem) ++nxt;
      if (nsl < kRbMaxSl && (size_t)rb_plan(t.mb, cg, br * nb, s.span2, ntaps, nsl + 1, nxt, s.lead).total <= kRbMaxSmem) ++nsl;
      s.mb = t.mb; s.cg = cg; s.nsl = nsl; s.nxt = nxt; s.box_rows = br; s.nbox = nb;
      *out = s;
      return true;
    }
  }
  return false;
}

template <typename T, int MB, int CG, int MODE>
cudaError_t rb_launch_t(const RbArgs& a, const RbShape& sh, ActType act, cudaStream_t st) {
  RbKArgs ka{};
  ka.bias1 = a.bias1; ka.bias2 = a.bias2; ka.res2 = a.res2; ka.w1 = a.w1; ka.w2 = a.w2;
  ka.n_q = a.n_q; ka.B = a.B; ka.ntaps1 = a.ntaps1; ka.ntaps2 = a.ntaps2;
  for (int j = 0; j < a.ntaps1; ++j) ka.off1[j] = a.off1[j];
  for (int j = 0; j < a.ntaps2; ++j) ka.off2[j] = a.off2[j];
  static const bool skip_zero = rb_env_int("HFG_RB_SKIPZ", 1) != 0;     // A/B knob: 0 issues every K-step
  if (skip_zero) { memcpy(ka.kz1, a.kz1, sizeof ka.kz1); memcpy(ka.kz2, a.kz2, sizeof ka.kz2); }
  ka.min1 = sh.min1; ka.min2 = sh.min2; ka.span2 = sh.span2;
  ka.r2 = MB * 128 - sh.span2;
  ka.tiles_q = (a.n_q + CG * ka.r2 - 1) / (CG * ka.r2);
  ka.total_tiles = a.B * ka.tiles_q;
  ka.box_rows = sh.box_rows; ka.nbox = sh.nbox; ka.nsl = sh.nsl; ka.nxt = sh.nxt; ka.lead = sh.lead;
  ka.xt_inv = 1.0f / a.mid_slope;
  ka.trace = g_rb_trace;
  // Two issuing warps double the MMA issue rate (-22 % on the k = 11 launches).  Under ncu (cold, serialised) the MB = 2 launches
  // that also read the running MRF sum looked 18 % slower with two issuers; inside bench.py they are not (V1 equal, V2 +0.6 %,
  // V3 +2 % with two issuers everywhere), so every launch uses two.  HFG_RB_ISSUERS=1 keeps the single issuer for A/B runs.
  static const int force_issuers = rb_env_int("HFG_RB_ISSUERS", 0);
  ka.two_issuers = force_issuers == 1 ? 0 : 1;
  ka.mid_slope = a.mid_slope; ka.res_inv = a.res_inv; ka.out_scale = a.out_scale; ka.out_slope = a.out_slope;
  if (!(a.mid_slope > 0.f && a.mid_slope <= 1.f) || !(a.out_slope > 0.f && a.out_slope <= 1.f) || !(a.res_inv >= 1.f)) return cudaErrorInvalidValue;
  const RbPlan pl = rb_plan(MB, CG, sh.nbox * sh.box_rows, sh.span2, a.ntaps1 + a.ntaps2, sh.nsl, sh.nxt, sh.lead);
  if ((size_t)pl.total > kRbMaxSmem) return cudaErrorInvalidConfiguration;
  CUtensorMap tm_x, tm_out, tm_last, tm_w1, tm_w2;
  if (!rb_map3(&tm_x, act, a.x, a.n_q, a.B, sh.box_rows)) return cudaErrorNotSupported;
  if (!rb_map3(&tm_out, act, a.out, a.n_q, a.B, 128)) return cudaErrorNotSupported;
  if (!rb_map3(&tm_last, act, a.out, a.n_q, a.B, 128 - sh.span2)) return cudaErrorNotSupported;
  if (MB == 1) tm_out = tm_last;                       // the only block of a tile is its last
  if (CG == 2) {
    constexpr int box = (kRbC / 2) * kRbRowB / 256;
    if ((uintptr_t)a.w1 % 256 || (uintptr_t)a.w2 % 256) return cudaErrorInvalidValue;
    if (!rb_map_w(&tm_w1, a.w1, (size_t)a.ntaps1 * kRbC * kRbRowB, box)) return cudaErrorNotSupported;
    if (!rb_map_w(&tm_w2, a.w2, (size_t)a.ntaps2 * kRbC * kRbRowB, box)) return cudaErrorNotSupported;
  } else { tm_w1 = tm_out; tm_w2 = tm_out; }
  auto kern = resblock_tc_kernel<T, MB, CG, MODE>;
  constexpr int kMaxDev = 32;
  static int configured_dev[kMaxDev] = {0};
  int dev = 0; cudaGetDevice(&dev);
  if (dev < 0 || dev >= kMaxDev) return cudaErrorInvalidDevice;
  if (pl.total > configured_dev[dev]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, pl.total);
    if (e != cudaSuccess) return e;
    configured_dev[dev] = pl.total;
  }
  static int num_sms = 0;
  if (!num_sms) { cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev); if (num_sms <= 0) num_sms = 148; }
  static const bool use_pdl = rb_env_int("HFG_PDL", 1) != 0;
  cudaLaunchAttribute at[2];
  int nat = 0;
  if (use_pdl) { at[nat].id = cudaLaunchAttributeProgrammaticStreamSerialization; at[nat].val.programmaticStreamSerializationAllowed = 1; ++nat; }
  cudaLaunchConfig_t lc{};
  lc.blockDim = dim3(kRbThreads); lc.dynamicSmemBytes = pl.total; lc.stream = st; lc.attrs = at;
  int ctas = CG == 1 ? num_sms : num_sms / 2;
  if (ctas > ka.total_tiles) ctas = ka.total_tiles;
  if (CG == 2) {
    at[nat].id = cudaLaunchAttributeClusterDimension;
    at[nat].val.clusterDim.x = 2; at[nat].val.clusterDim.y = 1; at[nat].val.clusterDim.z = 1;
    ++nat;
  }
  lc.gridDim = dim3(ctas * CG); lc.numAttrs = nat;
  return cudaLaunchKernelEx(&lc, kern, tm_x, tm_out, tm_last, tm_w1, tm_w2, ka);
}

}  // namespace

void rb_set_trace(long long* dev) { g_rb_trace = dev; }

bool rb_supported(const RbArgs& a, int* cg_out) {
  RbShape s;
  if (!rb_choose(a, &s)) return false;
  if (cg_out) *cg_out = s.cg;
  return true;
}

size_t rb_packed_weight_bytes(int ntaps) { return (size_t)ntaps * kRbC * kRbRowB; }

// fp32 [ntaps][64][64] (kernel view: [tap][c_in][c_out]) -> [tap][CTA of the pair] blocks of [8 K-chunks][64/cg rows][8 x 16-bit]
// (K-major, no swizzle: LBO = rows * 16 bytes between K chunks, SBO = 128 bytes between 8-row groups)
void rb_zero_ksteps(const float* w, int ntaps, unsigned char* kz) {
  for (int j = 0; j < ntaps; ++j) {
    unsigned m = 0;
    for (int kk = 0; kk < kRbC / 16; ++kk) {
      bool zero = true;
      for (int r = kk * 16; r < kk * 16 + 16 && zero; ++r)
        for (int c = 0; c < kRbC; ++c) if (w[((size_t)j * kRbC + r) * kRbC + c] != 0.0f) { zero = false; break; }
      if (zero) m |= 1u << kk;
    }
    kz[j] = (unsigned char)m;
  }
}

void rb_pack_weights(const float* w, int ntaps, int cg, ActType act, void* dst_host) {
  uint16_t* dst = reinterpret_cast<uint16_t*>(dst_host);
  const int nb = kRbC / cg;
  size_t o = 0;
  for (int j = 0; j < ntaps; ++j)
    for (int r = 0; r < cg; ++r)
      for (int chunk = 0; chunk < kRbC / 8; ++chunk)
        for (int row = 0; row < nb; ++row)
          for (int e = 0; e < 8; ++e) {
            const int ci = chunk * 8 + e, co = r * nb + row;
            const float f = w[((size_t)j * kRbC + ci) * kRbC + co];
            uint16_t bits;
            if (act == ACT_BF16) { __nv_bfloat16 h = __float2bfloat16_rn(f); memcpy(&bits, &h, 2); }
            else { __half h = __float2half_rn(f); memcpy(&bits, &h, 2); }
            dst[o++] = bits;
          }
}

cudaError_t launch_resblock_tc(const RbArgs& a, ActType act, cudaStream_t s) {
  if (a.B <= 0 || a.n_q <= 0) return cudaSuccess;
  RbShape sh;
  if (!rb_choose(a, &sh)) return cudaErrorInvalidValue;
  if (sh.cg != a.cg) return cudaErrorInvalidValue;          // weights were packed for the other CTA grouping
#define RB_DISPATCH(TT, MODE)                                                              \
  if (sh.mb == 2 && sh.cg == 2) return rb_launch_t<TT, 2, 2, MODE>(a, sh, act, s);         \
  if (sh.mb == 1 && sh.cg == 2) return rb_launch_t<TT, 1, 2, MODE>(a, sh, act, s);         \
  if (sh.mb == 2 && sh.cg == 1) return rb_launch_t<TT, 2, 1, MODE>(a, sh, act, s);         \
  if (sh.mb == 1 && sh.cg == 1) return rb_launch_t<TT, 1, 1, MODE>(a, sh, act, s);
  if (act == ACT_F16 && a.mode == 1) { RB_DISPATCH(__half, 1) }
  if (act == ACT_F16 && a.mode == 2) { RB_DISPATCH(__half, 2) }
  if (act == ACT_BF16 && a.mode == 1) { RB_DISPATCH(__nv_bfloat16, 1) }
  if (act == ACT_BF16 && a.mode == 2) { RB_DISPATCH(__nv_bfloat16, 2) }
#undef RB_DISPATCH
  return cudaErrorInvalidValue;
}

}  // namespace hfg
