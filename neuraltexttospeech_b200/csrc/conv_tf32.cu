// TF32 tensor-core mode (precision "tf32"): tcgen05.mma.kind::tf32 with fp32 storage -- the second tensor-core
// precision BASELINE.json's north star names ("bf16/TF32 tensor-core path", <= 2e-3 max-abs / >= 40 dB).  bf16 operands
// cannot meet that tolerance on this network (8-bit mantissa: ~5e-3 even with an fp32 trunk, tests/test_oracle.py); TF32
// keeps fp32's exponent range (no 65504 ceiling like the fp16 default) with a 10-bit mantissa on the MMA operands.
//
// One launch = one layer, like conv_tc.cu, with the same storage convention (tensors stored already activated for their
// consumer, residual recovered as y >= 0 ? y : y * res_inv) and the same tapped-GEMM contract (TcArgs), but a simpler
// kernel: one CTA per tile (utterance, 128 rows, variant, N <= 128 output channels), two CTAs per SM so that one tile's
// epilogue overlaps the other's MMAs.  Stored activations are rounded to TF32 (cvt.rna) so that the operand the tensor
// core reads -- it ignores the low 13 mantissa bits -- is the round-to-nearest value, not a truncation.
//   warp 0: TMA producer (slab = [128 + span rows][KS channels] fp32 per K-slice, 128/64/32-byte swizzle; weight stages)
//   warp 1: TMEM allocation + MMA issuer (K = 8 per instruction)
//   warps 2-5: epilogue, thread = output row: tcgen05.ld -> +bias -> +skip -> +MRF sum, scale -> lrelu -> global
// Reference: HiFiGAN_TF/model.py:71-80, 106-112, 156-167.  sm_100a only.
#include <cuda.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "kernels.cuh"
#include "tc_ptx.cuh"

namespace hfg {

namespace {

constexpr int kTfThreads = 192;
constexpr int kTfMaxSpan = 96;
constexpr int kTfNsl = 2, kTfNst = 3;

// descriptors as 32-bit low words (start address / 16, LBO) + compile-time high words: 32-bit adds per MMA for the issuing thread
template <uint32_t A_HI, uint32_t B_HI>
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint32_t a_lo, uint32_t b_lo, uint32_t idesc, uint32_t accumulate) {
  asm volatile("{ .reg .pred p; .reg .b64 da, db; setp.ne.b32 p, %4, 0; mov.b64 da, {%1, %5}; mov.b64 db, {%2, %6};\n"
               "tcgen05.mma.cta_group::1.kind::tf32 [%0], da, db, %3, p; }"
               ::"r"(tmem_d), "r"(a_lo), "r"(b_lo), "r"(idesc), "r"(accumulate), "n"(A_HI), "n"(B_HI) : "memory");
}
__device__ __forceinline__ float round_tf32(float v) {
  uint32_t r; asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v)); return __uint_as_float(r);
}
// kind::tf32 instruction descriptor (cute::UMMA::InstrDescriptor): D = f32 (bit 4), A / B format 2 = TF32, K-major, M = 128
__host__ __device__ constexpr uint32_t tf32_idesc(int n) {
  return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t)(n >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);
}

struct TfKArgs {
  float* out; const float* res; const float* res2; const float* bias;
  int n_q, c_out, out_stride, out_rows;
  float res_inv, out_scale, out_slope;
  int round_out;                       // 1: the output feeds an MMA -> round to TF32; 0: raw fp32 (running MRF sum)
  int nvar;
  TcVariant var[kMaxVariants];
  int min_off, nks, n, n_tiles, tiles_q, slab_rows;
  uint32_t idesc, tmem_cols;
  // extended form (TcArgs::ext_ntaps > 0): a conv of input stride `nph` is nph undilated convs over the phase
  // sub-sequences x[nph * q + r] (one tensor map per phase r): K-slice s of phase r is one slab, and tap j of that phase
  // reads it at row offset tap_a[ph_first[r] + j].  A grouped conv reads only its group's channels: K starts at
  // nt * kbase_per_tile for N tile nt (the N tile is one group's output channels).
  int ext, nph, kbase_per_tile;
  // short sequences (extended form): a 128-row M tile = pack_s consecutive sequences x pack_r slab rows each (one TMA box
  // [KS, pack_r, pack_s], rows outside a sequence zero-filled per sequence); tile row m = sequence m / pack_r, row
  // m % pack_r, of which the first tq are outputs (the others read the next chunk through their taps and are dropped)
  int pack_r, pack_s, tq, n_seq, load_bytes;
  int ph_first[5];
  short tap_a[kTfMaxExtTaps];
};

__device__ __forceinline__ int tf_ntaps(const TfKArgs& ka, int v, int r) { return ka.ext ? ka.ph_first[r + 1] - ka.ph_first[r] : ka.var[v].ntaps; }
__device__ __forceinline__ int tf_tap(const TfKArgs& ka, int v, int r, int j) { return ka.ext ? (int)ka.tap_a[ka.ph_first[r] + j] : ka.var[v].tap_off[j]; }

template <int KS>
__global__ void __launch_bounds__(kTfThreads, 2)
conv_tf32_kernel(const __grid_constant__ CUtensorMap tm_x, const __grid_constant__ CUtensorMap tm_x1, const __grid_constant__ CUtensorMap tm_x2,
                 const __grid_constant__ CUtensorMap tm_x3, const __grid_constant__ TfKArgs ka) {
  constexpr int ROWB = KS * 4;
  const int n = ka.n;
  const int slab_bytes = ka.slab_rows * ROWB;
  const int stage_bytes = KS * n * 4;

  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* slab = smem;
  uint8_t* wst = smem + kTfNsl * slab_bytes;                      // slab_bytes is a multiple of 256 (slab_rows % 8 == 0)
  uint64_t* bars = reinterpret_cast<uint64_t*>(wst + kTfNst * stage_bytes);
  uint64_t* slab_full = bars; uint64_t* slab_empty = bars + kTfNsl;
  uint64_t* w_full = slab_empty + kTfNsl; uint64_t* w_empty = w_full + kTfNst;
  uint64_t* acc_full = w_empty + kTfNst;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_full + 1);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  int t = blockIdx.x;
  const int nt = t % ka.n_tiles; t /= ka.n_tiles;
  const int v = t % ka.nvar; t /= ka.nvar;
  const int qt = t % ka.tiles_q; const int b = t / ka.tiles_q;
  const int nph = ka.ext ? ka.nph : 1;
  int ntaps = 0, last_r = 0;                                       // taps per K-slice over all phases; last phase that has taps
  for (int r = 0; r < nph; ++r) { const int c = tf_ntaps(ka, v, r); ntaps += c; if (c) last_r = r; }
  const int kbase = nt * ka.kbase_per_tile;

  if (tid == 0) {
    for (int s = 0; s < kTfNsl; ++s) { mbar_init(&slab_full[s], 1); mbar_init(&slab_empty[s], 1); }
    for (int s = 0; s < kTfNst; ++s) { mbar_init(&w_full[s], 1); mbar_init(&w_empty[s], 1); }
    mbar_init(acc_full, 1);
    fence_mbar_init();
    tma_prefetch_desc(&tm_x);
  }
  if (warp == 1) tmem_alloc(tmem_slot, ka.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ===================== producer: slab slices and weight stages =====================
    const uint8_t* wsrc = reinterpret_cast<const uint8_t*>(ka.var[v].w) + (size_t)nt * ka.nks * ntaps * stage_bytes;
    uint32_t wcnt = 0, scnt = 0;
    for (int s = 0; s < ka.nks; ++s) {
      for (int r = 0; r < nph; ++r) {
        const int nt_r = tf_ntaps(ka, v, r);
        if (nt_r == 0) continue;
        const int pos = scnt % kTfNsl;
        mbar_wait(&slab_empty[pos], ((scnt / kTfNsl) & 1) ^ 1);
        if (elect_one()) {
          const CUtensorMap* tm = r == 0 ? &tm_x : (r == 1 ? &tm_x1 : (r == 2 ? &tm_x2 : &tm_x3));
          mbar_expect_tx(&slab_full[pos], (uint32_t)ka.load_bytes);
          tma_load_3d(slab + pos * slab_bytes, tm, kbase + s * KS, qt * ka.tq + ka.min_off, b * ka.pack_s, &slab_full[pos]);
        }
        __syncwarp();
        ++scnt;
        for (int j = 0; j < nt_r; ++j, ++wcnt) {
          const int st = wcnt % kTfNst;
          mbar_wait(&w_empty[st], ((wcnt / kTfNst) & 1) ^ 1);
          if (elect_one()) {
            mbar_expect_tx(&w_full[st], (uint32_t)stage_bytes);
            bulk_g2s(wst + st * stage_bytes, wsrc, (uint32_t)stage_bytes, &w_full[st]);
          }
          __syncwarp();
          wsrc += stage_bytes;
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    const uint32_t slab_addr = smem_u32(slab), wst_addr = smem_u32(wst);
    uint32_t wcnt = 0, scnt = 0, first = 1;
    for (int s = 0; s < ka.nks; ++s) {
      for (int r = 0; r < nph; ++r) {
        const int nt_r = tf_ntaps(ka, v, r);
        if (nt_r == 0) continue;
        const int pos = scnt % kTfNsl;
        mbar_wait(&slab_full[pos], (scnt / kTfNsl) & 1);
        tc_fence_after();
        ++scnt;
        for (int j = 0; j < nt_r; ++j, ++wcnt) {
          const int st = wcnt % kTfNst;
          mbar_wait(&w_full[st], (wcnt / kTfNst) & 1);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t row_off = (uint32_t)(tf_tap(ka, v, r, j) - ka.min_off);
            constexpr uint32_t A_HI = ((8u * ROWB) >> 4) | (1u << 14) | ((ROWB == 128 ? 2u : (ROWB == 64 ? 4u : 6u)) << 29);
            constexpr uint32_t B_HI = (128u >> 4) | (1u << 14);
            const uint32_t a0 = (uint32_t)make_desc_sw<ROWB>(slab_addr + pos * slab_bytes + row_off * ROWB);
            const uint32_t b0 = (uint32_t)make_desc(wst_addr + st * stage_bytes, (uint32_t)n * 16u, 128u);
#pragma unroll
            for (int kk = 0; kk < KS / 8; ++kk) {
              umma_tf32<A_HI, B_HI>(tmem_base, a0 + (uint32_t)((kk * 32) >> 4), b0 + (uint32_t)((kk * 2 * n * 16) >> 4), ka.idesc, (first && kk == 0) ? 0u : 1u);
            }
            umma_commit(&w_empty[st]);
            if (j == nt_r - 1) {
              umma_commit(&slab_empty[pos]);
              if (s == ka.nks - 1 && r == last_r) umma_commit(acc_full);
            }
          }
          first = 0;
          __syncwarp();
        }
      }
    }
  } else {
    // ===================== epilogue: thread = output row =====================
    const int quarter = warp & 3;
    const int er = quarter * 32 + lane;
    const int sl = er / ka.pack_r, ql = er - sl * ka.pack_r;         // unpacked: pack_r = 128, pack_s = 1, tq = 128
    const int q = qt * ka.tq + ql;
    const int bs = b * ka.pack_s + sl;
    const bool valid = q < ka.n_q && ql < ka.tq && sl < ka.pack_s && bs < ka.n_seq;
    const uint32_t taddr = tmem_base + (((uint32_t)(quarter * 32)) << 16);
    const int n0 = nt * n;
    const size_t orow = ((size_t)bs * ka.out_rows + (size_t)q * ka.out_stride + ka.var[v].out_off) * ka.c_out + n0;
    mbar_wait(acc_full, 0);
    tc_fence_after();
    const int cw = n >= 32 ? 32 : 16;
    for (int c0 = 0; c0 < n; c0 += cw) {
      uint32_t r[32];
      if (cw == 32) tmem_ld32(taddr + c0, r); else tmem_ld16(taddr + c0, r);
      tmem_ld_wait();
      if (!valid) continue;
      const float4* b4 = reinterpret_cast<const float4*>(ka.bias + n0 + c0);
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        if (4 * i >= cw) break;
        const float4 bb = __ldg(b4 + i);
        float vv[4] = {__uint_as_float(r[4 * i]) + bb.x, __uint_as_float(r[4 * i + 1]) + bb.y, __uint_as_float(r[4 * i + 2]) + bb.z,
                       __uint_as_float(r[4 * i + 3]) + bb.w};
        if (ka.res) {                                   // skip connection, stored activated (model.py:79,111)
          const float4 rr = *reinterpret_cast<const float4*>(ka.res + orow + c0 + 4 * i);
          vv[0] += fminf(rr.x, rr.x * ka.res_inv); vv[1] += fminf(rr.y, rr.y * ka.res_inv);
          vv[2] += fminf(rr.z, rr.z * ka.res_inv); vv[3] += fminf(rr.w, rr.w * ka.res_inv);
        }
        if (ka.res2) {                                  // running raw MRF sum (model.py:161-166)
          const float4 rr = *reinterpret_cast<const float4*>(ka.res2 + orow + c0 + 4 * i);
          vv[0] += rr.x; vv[1] += rr.y; vv[2] += rr.z; vv[3] += rr.w;
        }
        float4 o;
        o.x = vv[0] * ka.out_scale; o.y = vv[1] * ka.out_scale; o.z = vv[2] * ka.out_scale; o.w = vv[3] * ka.out_scale;
        o.x = fmaxf(o.x, o.x * ka.out_slope); o.y = fmaxf(o.y, o.y * ka.out_slope);
        o.z = fmaxf(o.z, o.z * ka.out_slope); o.w = fmaxf(o.w, o.w * ka.out_slope);
        if (ka.round_out) { o.x = round_tf32(o.x); o.y = round_tf32(o.y); o.z = round_tf32(o.z); o.w = round_tf32(o.w); }
        *reinterpret_cast<float4*>(ka.out + orow + c0 + 4 * i) = o;
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) { tc_fence_after(); tmem_dealloc(tmem_base, ka.tmem_cols); }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn tf_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

inline int tf_ks(int c_in) { return c_in >= 32 ? 32 : c_in; }

// Extended form: tap j at input-row offset off[j] of a stride-s conv belongs to phase r = off mod s and reads phase row
// q + a with a = (off - r) / s.  Taps are processed phase by phase, in their given order inside a phase.
struct TfPhases { int order[kTfMaxExtTaps]; int a[kTfMaxExtTaps]; int first[5]; int min_a, max_a; };
inline bool tf_phases(const int* off, int ntaps, int stride, TfPhases* p) {
  if (ntaps < 1 || ntaps > kTfMaxExtTaps || stride < 1 || stride > 4) return false;
  int n = 0;
  p->min_a = 1 << 30; p->max_a = -(1 << 30);
  for (int r = 0; r < 4; ++r) {
    p->first[r] = n;
    if (r >= stride) continue;
    for (int j = 0; j < ntaps; ++j) {
      const int rr = ((off[j] % stride) + stride) % stride;
      if (rr != r) continue;
      const int a = (off[j] - rr) / stride;
      p->order[n] = j; p->a[n] = a; ++n;
      p->min_a = a < p->min_a ? a : p->min_a; p->max_a = a > p->max_a ? a : p->max_a;
    }
  }
  p->first[4] = n;
  return n == ntaps;
}
inline int tf_n(int c_out) { return c_out > 128 ? 128 : c_out; }

template <int KS>
cudaError_t tf_launch(const TcArgs& a, cudaStream_t s) {
  TfKArgs ka{};
  ka.out = (float*)a.out; ka.res = (const float*)a.res; ka.res2 = (const float*)a.res2; ka.bias = a.bias;
  ka.n_q = a.n_q; ka.c_out = a.c_out; ka.out_stride = a.out_stride; ka.out_rows = a.out_rows;
  ka.res_inv = a.res_inv; ka.out_scale = a.out_scale; ka.out_slope = a.out_slope;
  ka.round_out = a.raw_out ? 0 : 1;
  ka.nvar = a.nvar;
  int mn = a.var[0].tap_off[0], mx = mn;
  for (int v = 0; v < a.nvar; ++v) {
    ka.var[v] = a.var[v];
    for (int j = 0; j < a.var[v].ntaps; ++j) { mn = a.var[v].tap_off[j] < mn ? a.var[v].tap_off[j] : mn; mx = a.var[v].tap_off[j] > mx ? a.var[v].tap_off[j] : mx; }
  }
  const bool ext = a.ext_ntaps > 0;
  const int stride = ext ? a.in_stride : 1;
  const int groups = ext && a.groups > 1 ? a.groups : 1;
  if (ext) {
    TfPhases ph;
    if (a.nvar != 1 || a.out_stride != 1 || !a.ext_tap_off || !tf_phases(a.ext_tap_off, a.ext_ntaps, stride, &ph)) return cudaErrorInvalidValue;
    if (a.x_rows < stride || a.c_in % groups || a.c_out % groups || a.cin_g != a.c_in / groups) return cudaErrorInvalidValue;
    ka.ext = 1; ka.nph = stride;
    for (int r = 0; r < 5; ++r) ka.ph_first[r] = ph.first[r];
    for (int j = 0; j < a.ext_ntaps; ++j) ka.tap_a[j] = (short)ph.a[j];
    ka.var[0].out_off = 0; ka.var[0].ntaps = a.ext_ntaps;
    mn = ph.min_a; mx = ph.max_a;
  }
  if (mx - mn > kTfMaxSpan) return cudaErrorInvalidValue;
  if (!(a.out_slope > 0.f && a.out_slope <= 1.f) || (a.res && !(a.res_inv >= 1.f))) return cudaErrorInvalidValue;
  if ((a.res || a.res2) && a.out_stride != 1) return cudaErrorInvalidValue;
  ka.min_off = mn;
  ka.n = groups > 1 ? a.c_out / groups : tf_n(a.c_out); ka.n_tiles = a.c_out / ka.n; ka.nks = (ext ? a.cin_g : a.c_in) / KS;
  ka.kbase_per_tile = groups > 1 ? a.cin_g : 0;
  if (!(ka.n == 16 || (ka.n % 32 == 0 && ka.n <= 128))) return cudaErrorInvalidValue;
  ka.pack_r = 128; ka.pack_s = 1; ka.tq = 128; ka.n_seq = a.B;
  if (ext && a.B > 1) {
    // many short sequences (deep DiscriminatorP layers: 15-50 rows): chunks of tq outputs + span halo rows, as many
    // chunks as fit 128 MMA rows; kept only when it fills the M tile better than one sequence per tile
    const int span = mx - mn;
    const int tq = a.n_q < 128 - span ? a.n_q : 128 - span;
    const int r = tq + span, sq = 128 / r;
    const long long plain = (long long)a.B * ((a.n_q + 127) / 128);
    const long long packed = (long long)((a.B + sq - 1) / sq) * ((a.n_q + tq - 1) / tq);
    if (sq >= 2 && packed * 10 <= plain * 8) { ka.pack_r = r; ka.pack_s = sq; ka.tq = tq; }
  }
  ka.tiles_q = (a.n_q + ka.tq - 1) / ka.tq;
  ka.slab_rows = (128 + (mx - mn) + 7) & ~7;
  ka.load_bytes = (ka.pack_s > 1 ? ka.pack_s * ka.pack_r : ka.slab_rows) * KS * 4;
  ka.idesc = tf32_idesc(ka.n);
  ka.tmem_cols = ka.n <= 32 ? 32 : (ka.n <= 64 ? 64 : 128);
  const size_t smem = (size_t)kTfNsl * ka.slab_rows * KS * 4 + (size_t)kTfNst * KS * ka.n * 4 + (2 * kTfNsl + 2 * kTfNst + 1) * 8 + 16;
  EncodeTiledFn fn = tf_encode_fn();
  if (!fn) return cudaErrorNotSupported;
  CUtensorMap tm[4];
  for (int r = 0; r < 4; ++r) {
    if (r >= stride) { tm[r] = tm[0]; continue; }
    // phase r of a stride-s conv: rows r, r + s, r + 2 s, ... of every sequence (row pitch s * c_in)
    cuuint64_t gdim[3] = {(cuuint64_t)a.c_in, (cuuint64_t)((a.x_rows - r + stride - 1) / stride), (cuuint64_t)a.B};
    cuuint64_t gstr[2] = {(cuuint64_t)a.c_in * 4 * (cuuint64_t)stride, (cuuint64_t)a.c_in * 4 * (cuuint64_t)a.x_rows};
    cuuint32_t box[3] = {(cuuint32_t)KS, (cuuint32_t)(ka.pack_s > 1 ? ka.pack_r : ka.slab_rows), (cuuint32_t)ka.pack_s};
    cuuint32_t est[3] = {1, 1, 1};
    const CUtensorMapSwizzle sw = KS == 32 ? CU_TENSOR_MAP_SWIZZLE_128B : (KS == 16 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
    void* base = const_cast<float*>(reinterpret_cast<const float*>(a.x) + (size_t)r * a.c_in);
    if (fn(&tm[r], CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, base, gdim, gstr, box, est, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
           CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) return cudaErrorNotSupported;
  }
  auto kern = conv_tf32_kernel<KS>;
  constexpr int kMaxDev = 32;
  static size_t configured[kMaxDev] = {0};
  int dev = 0; cudaGetDevice(&dev);
  if (dev < 0 || dev >= kMaxDev) return cudaErrorInvalidDevice;
  if (smem > configured[dev]) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return e;
    configured[dev] = smem;
  }
  const long long tiles = (long long)((a.B + ka.pack_s - 1) / ka.pack_s) * ka.tiles_q * a.nvar * ka.n_tiles;
  if (tiles > 0x7fffffffLL) return cudaErrorInvalidValue;
  kern<<<(unsigned)tiles, kTfThreads, smem, s>>>(tm[0], tm[1], tm[2], tm[3], ka);
  return cudaGetLastError();
}

}  // namespace

bool tf32_supported(int c_in, int c_out, int ntaps, int span) {
  if (ntaps < 1 || ntaps > kMaxTaps || span < 0 || span > kTfMaxSpan) return false;
  if (!(c_in == 8 || c_in == 16 || (c_in >= 32 && c_in % 32 == 0))) return false;
  const int n = tf_n(c_out);
  return c_out >= 16 && c_out % 16 == 0 && c_out % n == 0 && (n == 16 || n % 32 == 0);
}

size_t tf32_packed_weight_bytes(int c_in, int c_out, int ntaps) { return (size_t)ntaps * c_in * c_out * 4; }

// fp32 [ntaps][c_in][c_out] -> [n-tile][K-slice][tap] stages of [KS/4 chunks][N rows][4 floats] (K-major, no swizzle), TF32-rounded
void tf32_pack_weights(const float* w, int ntaps, int c_in, int c_out, void* dst_host) {
  float* dst = reinterpret_cast<float*>(dst_host);
  const int ks = tf_ks(c_in), n = tf_n(c_out);
  size_t o = 0;
  for (int nt = 0; nt < c_out / n; ++nt)
    for (int s = 0; s < c_in / ks; ++s)
      for (int j = 0; j < ntaps; ++j)
        for (int chunk = 0; chunk < ks / 4; ++chunk)
          for (int row = 0; row < n; ++row)
            for (int e = 0; e < 4; ++e) {
              const int ci = s * ks + chunk * 4 + e, co = nt * n + row;
              float f = w[((size_t)j * c_in + ci) * c_out + co];
              uint32_t u; memcpy(&u, &f, 4);
              u = (u + 0x0FFFu + ((u >> 13) & 1u)) & ~0x1FFFu;      // round to nearest even at 10 mantissa bits
              memcpy(&f, &u, 4);
              dst[o++] = f;
            }
}

bool tf32_ext_supported(int c_in, int c_out, int groups, int in_stride, int ntaps, const int* tap_off) {
  TfPhases ph;
  if (groups < 1 || c_in % groups || c_out % groups || !tap_off || !tf_phases(tap_off, ntaps, in_stride, &ph)) return false;
  if (ph.max_a - ph.min_a > kTfMaxSpan) return false;
  const int cin_g = c_in / groups, n = groups > 1 ? c_out / groups : tf_n(c_out);
  if (!(cin_g == 8 || cin_g == 16 || (cin_g >= 32 && cin_g % 32 == 0))) return false;
  if (groups == 1 && c_out % n) return false;
  return n == 16 || (n % 32 == 0 && n <= 128);
}

// w fp32 [ntaps][cin_g][c_out] -> [n-tile][K-slice][tap, phase-major][KS/4 chunks][N rows][4 floats], TF32-rounded
void tf32_pack_weights_ext(const float* w, int ntaps, const int* tap_off, int in_stride, int cin_g, int c_out, int groups, void* dst_host) {
  float* dst = reinterpret_cast<float*>(dst_host);
  TfPhases ph;
  if (!tf_phases(tap_off, ntaps, in_stride, &ph)) return;
  const int ks = tf_ks(cin_g), n = groups > 1 ? c_out / groups : tf_n(c_out);
  size_t o = 0;
  for (int nt = 0; nt < c_out / n; ++nt)
    for (int s = 0; s < cin_g / ks; ++s)
      for (int t = 0; t < ntaps; ++t)
        for (int chunk = 0; chunk < ks / 4; ++chunk)
          for (int row = 0; row < n; ++row)
            for (int e = 0; e < 4; ++e) {
              const int ci = s * ks + chunk * 4 + e, co = nt * n + row;
              float f = w[((size_t)ph.order[t] * cin_g + ci) * c_out + co];
              uint32_t u; memcpy(&u, &f, 4);
              u = (u + 0x0FFFu + ((u >> 13) & 1u)) & ~0x1FFFu;
              memcpy(&f, &u, 4);
              dst[o++] = f;
            }
}

cudaError_t launch_conv_tf32(const TcArgs& a, cudaStream_t s) {
  if (a.nvar < 1 || a.nvar > kMaxVariants) return cudaErrorInvalidValue;
  if (a.B <= 0 || a.n_q <= 0) return cudaSuccess;
  if (a.ext_ntaps > 0) {
    if (!tf32_ext_supported(a.c_in, a.c_out, a.groups > 1 ? a.groups : 1, a.in_stride, a.ext_ntaps, a.ext_tap_off)) return cudaErrorInvalidValue;
    const int ks = tf_ks(a.cin_g);
    if (ks == 32) return tf_launch<32>(a, s);
    if (ks == 16) return tf_launch<16>(a, s);
    return tf_launch<8>(a, s);
  }
  int mxt = 0;
  for (int v = 0; v < a.nvar; ++v) mxt = a.var[v].ntaps > mxt ? a.var[v].ntaps : mxt;
  if (!tf32_supported(a.c_in, a.c_out, mxt, 0)) return cudaErrorInvalidValue;
  const int ks = tf_ks(a.c_in);
  if (ks == 32) return tf_launch<32>(a, s);
  if (ks == 16) return tf_launch<16>(a, s);
  return tf_launch<8>(a, s);
}

}  // namespace hfg
