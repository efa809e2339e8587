// Host-side engine behind the C ABI (include/hifigan_b200.h).
//
// Mirrors, as plain C++ control flow, what the reference builds and runs with Keras layers:
//   Generator.__init__   reference HiFiGAN_TF/model.py:116-152  -> build_layers()
//   Generator.call       reference HiFiGAN_TF/model.py:155-173  -> run_forward()
//   ResBlock1/2.call     reference HiFiGAN_TF/model.py:71-80, 106-112
//   get_padding          reference HiFiGAN_TF/utils.py:29-37    -> conv_pad_left()
//   tfa WeightNormalization(data_init=False) (call sites model.py:25-66,88-101,121,131,145)
//                                                               -> fold_weight_norm()
// There is no CPU compute path here: every arithmetic op is a CUDA kernel launch.
#include <algorithm>
#include <cmath>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <map>
#include <string>
#include <thread>
#include <vector>

#include "../../include/hifigan_b200.h"
#include "kernels.cuh"

using namespace hfg;

namespace {

struct Phase {            // one output phase of a transposed conv (or the single "phase" of a conv)
  int ntaps = 0;
  int tap_off[kMaxTaps];
  unsigned kz[kMaxTaps] = {0};   // all-zero 16-row K-steps of each tap's weight block (TcVariant::kz)
  size_t w32_off = 0;     // element offset into d_w32
  size_t wpk_off = 0;     // byte offset into d_wpk
};

struct Layer {
  std::string name;
  bool transposed = false;
  int k = 0, dil = 1, stride = 1, cin = 0, cout = 0;
  int cin_pad = 0;        // channel count the kernels see (conv_pre pads num_mels up for the MMA path)
  int cout_k = 0;         // output channels the kernels see (conv_post on the tcgen05 path pads 1 -> 16)
  int in_rate = 1, out_rate = 1;   // rows per mel frame of the input / output tensor
  // Time folding (tcgen05 path, channel counts below the 16-wide MMA tile, e.g. V2's 8-channel stage): a [L, C] tensor
  // is bit-identical to [L/f, C*f]; a conv over it is a conv over the folded rows with block-Toeplitz weights.
  int fold_in = 1, fold_out = 1;
  int u_k = 1;            // output stride in kernel-view rows (stride * fold_in / fold_out)
  int post_f = 0;         // conv_post on the tcgen05 path: samples per kernel-view row
  std::vector<float> v, g, b;
  bool has_v = false, has_g = false, has_b = false;
  std::vector<Phase> phases;
  float* d_w32 = nullptr;
  float* d_bias = nullptr;
  void* d_wpk = nullptr;
  bool use_tc = false;      // persistent tcgen05 kernel (conv_tc.cu): one launch per layer, phases = variants
  size_t wpk_bytes = 0;     // size of d_wpk (multiple of 256)
  // convs1[m] of a ResBlock1 whose iteration (convs1[m] -> convs2[m] -> + x) runs as ONE launch of resblock_tc.cu
  struct Fused {
    bool on = false;
    int fold = 1, cg = 2, ntaps1 = 0, ntaps2 = 0;
    int off1[kMaxTaps], off2[kMaxTaps];
    unsigned char kz1[kMaxTaps] = {0}, kz2[kMaxTaps] = {0};   // all-zero K-steps of each tap (rb_zero_ksteps)
    void* d_w1 = nullptr; void* d_w2 = nullptr;
    float* d_b1 = nullptr; float* d_b2 = nullptr;
  } fused;
};

thread_local std::string g_create_error;
constexpr int kTimingSlots = 128;    // forwards whose stage stamps are kept between two hfg_stage_times_ms calls

}  // namespace

struct hfg_ctx {
  hfg_config cfg{};
  int device = 0;
  std::vector<Layer> layers;
  std::map<std::string, int> index;     // weight name -> layer (hfg_set_weight)
  // forward plan (indices into `layers`, fixed at create): run_forward does no string work
  int i_pre = -1, i_post = -1;
  std::vector<int> i_ups;               // [stage]
  std::vector<int> i_rb;                // [resblock n][dilation m][0 = convs1 / convs, 1 = convs2]: (n * HFG_MAX_DILATIONS + m) * 2 + which
  int total_up = 1;
  bool finalized = false;
  int precision = HFG_PRECISION_FP32;
  bool tf32 = false;            // precision "tf32": fp32 storage (act == ACT_F32), layers on the kind::tf32 tensor-core kernel
  ActType act = ACT_F32;
  std::string err;
  long long launches = 0;
  bool stage_timing = false;
  std::vector<cudaEvent_t> events;   // kTimingSlots forwards x marks_per_fwd stamps (ring)
  int marks_per_fwd = 0;
  long long timed_forwards = 0;      // forwards stamped since the last hfg_stage_times_ms
  int n_stage_marks = 0;
  std::vector<long long> stage_launches;   // kernels launched between consecutive stage marks of the last forward
  // host-buffer path
  int micro_batch = 0;         // host path: utterances per forward; 0 = auto (auto_micro_batch)
  cudaStream_t stream = nullptr;
  cudaStream_t cap_stream = nullptr;                          // device path: graphs are captured here, launched on the caller's stream
  cudaStream_t h2d_stream = nullptr, d2h_stream = nullptr;   // host path: copies overlap the forward of the neighbouring micro-batch
  std::vector<cudaEvent_t> copy_events;                      // 2 per micro-batch (mel landed, audio ready)
  // host path, small forwards (launch-bound): the launch sequence of one micro-batch is captured into a CUDA graph on
  // its second use and replayed afterwards (tensor maps and pointers are baked in: the cache is keyed on them)
  struct GraphEntry { const void* mel; const void* out; const void* ws; int B, T, sample_bytes; int uses; long long launches; cudaGraphExec_t exec; };
  std::vector<GraphEntry> graphs;
  bool use_graphs = true;
  // pageable caller buffers are staged through pinned memory owned by the handle (one memcpy per micro-batch, overlapped
  // with the neighbouring micro-batch's kernels); pinned / registered caller buffers are copied directly
  char* pin_in = nullptr; size_t pin_in_bytes = 0;
  char* pin_out = nullptr; size_t pin_out_bytes = 0;
  void* ws = nullptr; size_t ws_bytes = 0;
  float* d_mel = nullptr; size_t d_mel_bytes = 0;
  void* d_audio = nullptr; size_t d_audio_bytes = 0;   // fp32 audio or 16-bit PCM
  bool force_ffma = false;
  bool auto_fold = true;       // per-layer fold factor from fold_cost (option "auto_fold")
  bool fuse_resblocks = true;  // ResBlock1 iterations of the narrow stages (C <= 64) as one launch each (option "fuse_resblocks")
  bool need_acc = true, need_sum = false;
  // launch-bound forwards (B*T <= chain_frames): the resblocks of a stage are independent chains up to their last conv
  // (reference model.py:161-166 sums rb_j(x) over j), so chain j >= 1 runs on its own stream with its own scratch
  // buffers; only the running-sum convs are ordered (events).  0 disables (option "concurrent_resblocks").
  long long chain_frames = 2048;
  cudaStream_t side[HFG_MAX_KERNELS] = {};
  cudaEvent_t ev_fork = nullptr, ev_join = nullptr, ev_sum[HFG_MAX_KERNELS] = {};
};

namespace {

void clear_graphs(hfg_ctx* h) {
  for (auto& g : h->graphs) if (g.exec) cudaGraphExecDestroy(g.exec);
  h->graphs.clear();
}

int fail(hfg_ctx* h, int code, const char* fmt, ...) {
  char buf[512];
  va_list ap; va_start(ap, fmt); vsnprintf(buf, sizeof buf, fmt, ap); va_end(ap);
  if (h) h->err = buf; else g_create_error = buf;
  return code;
}

#define CUDA_TRY(h, expr)                                                                   \
  do { cudaError_t e__ = (expr);                                                            \
       if (e__ != cudaSuccess) return fail(h, HFG_ERR_CUDA, "%s failed: %s", #expr, cudaGetErrorString(e__)); } while (0)

size_t act_bytes(ActType a) { return a == ACT_F32 ? 4 : 2; }
size_t align_up(size_t x, size_t a) { return (x + a - 1) / a * a; }

// reference HiFiGAN_TF/utils.py:29-37: int pad -> "valid"/"causal"/"same"; returns left zero padding.
// compat=1 (upstream) pads symmetrically.
int conv_pad_left(int k, int d, int compat, bool* ok) {
  const int span = d * (k - 1);
  *ok = true;
  if (compat == 1) return span / 2;
  const int pad = (int)((k * d - d) / 2);
  if (pad == 0) { *ok = (span == 0); return 0; }   // "valid": only length-preserving for k == 1
  if (pad == 1) return span;                       // "causal": all padding on the left
  return span / 2;                                 // "same"
}

void build_layers(hfg_ctx* h) {
  const hfg_config& c = h->cfg;
  int rate = 1;
  auto add = [&](const std::string& name, bool tr, int k, int d, int s, int cin, int cout) {
    Layer L; L.name = name; L.transposed = tr; L.k = k; L.dil = d; L.stride = s; L.cin = cin; L.cout = cout;
    L.cin_pad = cin;
    L.in_rate = rate; L.out_rate = rate * (tr ? s : 1);
    rate = L.out_rate;
    h->index[name] = (int)h->layers.size();
    h->layers.push_back(std::move(L));
  };
  const int c0 = c.upsample_initial_channel;
  h->i_rb.assign((size_t)c.num_upsamples * c.num_kernels * HFG_MAX_DILATIONS * 2, -1);
  auto rb_slot = [&](int n, int m, int which) -> int& { return h->i_rb[((size_t)n * HFG_MAX_DILATIONS + m) * 2 + which]; };
  h->i_pre = (int)h->layers.size();
  add("conv_pre", false, 7, 1, 1, c.num_mels, c0);
  int cprev = c0;
  h->total_up = 1;
  for (int i = 0; i < c.num_upsamples; ++i) {
    const int ch = c0 >> (i + 1);
    h->i_ups.push_back((int)h->layers.size());
    add("ups." + std::to_string(i), true, c.upsample_kernel_sizes[i], 1, c.upsample_rates[i], cprev, ch);
    h->total_up *= c.upsample_rates[i];
    for (int j = 0; j < c.num_kernels; ++j) {
      const int n = i * c.num_kernels + j;
      for (int m = 0; m < c.num_dilations[j]; ++m) {
        const std::string base = "resblocks." + std::to_string(n);
        rb_slot(n, m, 0) = (int)h->layers.size();
        if (c.resblock == 1) {
          add(base + ".convs1." + std::to_string(m), false, c.resblock_kernel_sizes[j], c.resblock_dilation_sizes[j][m], 1, ch, ch);
          rb_slot(n, m, 1) = (int)h->layers.size();
          add(base + ".convs2." + std::to_string(m), false, c.resblock_kernel_sizes[j], 1, 1, ch, ch);
        } else {
          add(base + ".convs." + std::to_string(m), false, c.resblock_kernel_sizes[j], c.resblock_dilation_sizes[j][m], 1, ch, ch);
        }
      }
    }
    cprev = ch;
  }
  h->i_post = (int)h->layers.size();
  add("conv_post", false, 7, 1, 1, cprev, 1);
}

int validate_config(const hfg_config& c) {
  if (c.num_mels <= 0 || c.num_mels % 4) return fail(nullptr, HFG_ERR_INVALID, "num_mels must be a positive multiple of 4 (got %d)", c.num_mels);
  if (c.resblock != 1 && c.resblock != 2) return fail(nullptr, HFG_ERR_INVALID, "resblock must be 1 or 2 (got %d)", c.resblock);
  if (c.num_upsamples < 1 || c.num_upsamples > HFG_MAX_UPSAMPLES) return fail(nullptr, HFG_ERR_INVALID, "num_upsamples out of range: %d", c.num_upsamples);
  if (c.num_kernels < 1 || c.num_kernels > HFG_MAX_KERNELS) return fail(nullptr, HFG_ERR_INVALID, "num_kernels out of range: %d", c.num_kernels);
  if (c.compat != 0 && c.compat != 1) return fail(nullptr, HFG_ERR_INVALID, "compat must be 0 or 1");
  const int c0 = c.upsample_initial_channel;
  if (c0 <= 0 || (c0 >> c.num_upsamples) < 8 || ((c0 >> c.num_upsamples) << c.num_upsamples) != c0 || (c0 >> c.num_upsamples) % 8)
    return fail(nullptr, HFG_ERR_INVALID, "upsample_initial_channel=%d must be divisible by 2^num_upsamples with a multiple of 8 channels left", c0);
  for (int i = 0; i < c.num_upsamples; ++i) {
    const int u = c.upsample_rates[i], k = c.upsample_kernel_sizes[i];
    if (u < 1 || k < u || (k + u - 1) / u > kMaxTaps) return fail(nullptr, HFG_ERR_INVALID, "unsupported upsample rate/kernel %d/%d", u, k);
  }
  for (int j = 0; j < c.num_kernels; ++j) {
    const int k = c.resblock_kernel_sizes[j];
    if (k < 1 || k > kMaxTaps) return fail(nullptr, HFG_ERR_INVALID, "unsupported resblock kernel size %d", k);
    if (c.num_dilations[j] < 1 || c.num_dilations[j] > HFG_MAX_DILATIONS) return fail(nullptr, HFG_ERR_INVALID, "bad dilation count");
    for (int m = 0; m < c.num_dilations[j]; ++m) {
      bool ok;
      if (c.resblock_dilation_sizes[j][m] < 1) return fail(nullptr, HFG_ERR_INVALID, "dilation must be >= 1");
      conv_pad_left(k, c.resblock_dilation_sizes[j][m], c.compat, &ok);
      if (!ok) return fail(nullptr, HFG_ERR_INVALID, "kernel %d / dilation %d gives a 'valid' (length-changing) conv; the reference's residual add would fail too", k, c.resblock_dilation_sizes[j][m]);
      conv_pad_left(k, 1, c.compat, &ok);
      if (!ok) return fail(nullptr, HFG_ERR_INVALID, "kernel %d gives a 'valid' (length-changing) conv", k);
    }
  }
  return HFG_OK;
}

// tfa WeightNormalization: kernel = l2_normalize(v, axes != last) * g; l2_normalize uses
// v * rsqrt(max(sum v^2, 1e-12)).  Returns canonical [k][cin][cout] fp32.
std::vector<float> fold_weight_norm(const Layer& L) {
  const int k = L.k, last = L.transposed ? L.cin : L.cout, inner = L.transposed ? L.cout : L.cin;
  std::vector<double> sq(last, 0.0);
  for (int j = 0; j < k; ++j)
    for (int a = 0; a < inner; ++a)
      for (int z = 0; z < last; ++z) { const double t = L.v[((size_t)j * inner + a) * last + z]; sq[z] += t * t; }
  std::vector<float> w((size_t)k * L.cin * L.cout);
  for (int j = 0; j < k; ++j)
    for (int a = 0; a < inner; ++a)
      for (int z = 0; z < last; ++z) {
        const double val = (double)L.v[((size_t)j * inner + a) * last + z] / std::sqrt(std::max(sq[z], 1e-12)) * (double)L.g[z];
        const int ci = L.transposed ? z : a, co = L.transposed ? a : z;
        w[((size_t)j * L.cin + ci) * L.cout + co] = (float)val;
      }
  return w;
}

// Kernel-visible form of a layer: phases ("variants") of taps over kernel-view rows.
//   unfolded:  out[q*u + p] = sum_{(o, j) in taps_p} x[q + o] . W_j            (conv: u = 1; transposed conv: polyphase)
//   conv folded by f (input and output [L, C] viewed as [L/f, C*f], row r = Q*f + s):
//       out'[Q, (s, co)] = sum_O sum_(s', ci) x'[Q + O, (s', ci)] . W'_O,  O = floor((s + o)/f), s' = (s + o) mod f
//   transposed conv whose OUTPUT is folded by f (f | u): kernel-view stride u/f, variant p' holds phases p'*f + s as channel blocks
//   conv_post on the tcgen05 path: C_out = 1 -> column s of a 16-column tile holds sample Q*f + s
struct KView {
  int cin_k = 0, cout_k = 0, u_k = 1, post_f = 0;
  std::vector<std::vector<int>> off;                        // [phase][tap] kernel-view row offset
  std::vector<std::vector<std::vector<float>>> w;           // [phase][tap] -> [cin_k][cout_k]
  int max_taps() const { size_t m = 0; for (auto& o : off) m = std::max(m, o.size()); return (int)m; }
  int span() const {                                        // over ALL phases: they share one halo slab
    int mn = off[0][0], mx = mn;
    for (auto& o : off) for (int v : o) { mn = std::min(mn, v); mx = std::max(mx, v); }
    return mx - mn;
  }
};

inline int floordiv(int a, int b) { return (a >= 0) ? a / b : -((-a + b - 1) / b); }

KView build_view(const Layer& L, const std::vector<float>& w, int pad_left, int fi, int fo, bool post_tc) {
  KView V;
  const int u = L.transposed ? L.stride : 1, cin = L.cin, cout = L.cout;
  V.cin_k = cin * fi; V.cout_k = post_tc ? 16 : cout * fo; V.post_f = post_tc ? fo : 0;
  std::vector<std::vector<std::pair<int, int>>> ut(u);      // unfolded (row offset, kernel index j) per phase
  if (!L.transposed) {
    for (int j = 0; j < L.k; ++j) ut[0].push_back({j * L.dil - pad_left, j});
  } else {
    const int P = (L.k - u) / 2;
    for (int p = 0; p < u; ++p)
      for (int j = (p + P) % u; j < L.k; j += u) ut[p].push_back({(p + P - j) / u, j});   // exact: (p+P-j) % u == 0
  }
  V.u_k = L.transposed ? u / fo : 1;
  const int nph = L.transposed ? V.u_k : 1;
  for (int pp = 0; pp < nph; ++pp) {
    std::map<int, std::vector<float>> acc;
    const int f = L.transposed ? fo : fi;                  // conv: fi == fo (conv_post: fo == fi, one output sample per s)
    for (int sidx = 0; sidx < f; ++sidx) {
      const auto& taps = ut[L.transposed ? pp * fo + sidx : 0];
      for (const auto& tj : taps) {
        const int O = L.transposed ? tj.first : floordiv(sidx + tj.first, f);
        const int sp = L.transposed ? 0 : sidx + tj.first - O * f;
        std::vector<float>& m = acc[O];
        if (m.empty()) m.assign((size_t)V.cin_k * V.cout_k, 0.f);
        for (int ci = 0; ci < cin; ++ci)
          for (int co = 0; co < cout; ++co)
            m[(size_t)(sp * cin + ci) * V.cout_k + (post_tc ? sidx : sidx * cout + co)] += w[((size_t)tj.second * cin + ci) * cout + co];
      }
    }
    V.off.emplace_back(); V.w.emplace_back();
    for (auto& kv : acc) { V.off.back().push_back(kv.first); V.w.back().push_back(std::move(kv.second)); }
  }
  return V;
}

// Modelled cycles per (unfolded) output row of a ResBlock conv run with fold factor f on the tcgen05 kernel: every
// kernel-view tap costs C*f/16 MMAs of max(N/2 tensor cycles, operand bytes / 128 B per clk of shared memory); weights
// that do not stay resident are re-staged per tile.  Folding trades redundant MACs (block-Toeplitz zeros) for wider
// MMAs: it wins for the d = 1 convs of the narrow stages and loses for dilated ones (tap count grows with d).
double fold_cost(const Layer& L, int pad_left, int f, bool* ok) {
  std::map<int, int> taps;                                     // kernel-view offsets O = floor((s + o) / f)
  for (int sidx = 0; sidx < f; ++sidx)
    for (int j = 0; j < L.k; ++j) taps[floordiv(sidx + j * L.dil - pad_left, f)] = 1;
  const int nt = (int)taps.size(), span = taps.rbegin()->first - taps.begin()->first, cf = L.cin * f;
  TcConfig c;
  const bool sup = nt <= kMaxTaps && tc_supported(cf, cf, nt, span) && tc_config(cf, cf, span, &c);
  if (ok) *ok = sup;
  if (!sup) return 1e30;
  const double mma = std::max(c.n / 2.0, (4096.0 + c.n * 32.0 / c.cg) / 128.0);
  double clk = nt * (cf / 16.0) * mma / 128.0;                 // per kernel-view row
  const int stages = (cf / c.ks) * ((nt + c.tps - 1) / c.tps) * (c.ks / c.kbw);
  if (stages > c.nst) clk += (double)nt * cf * cf * 2.0 / c.cg / 128.0 / (c.mb * 128.0);   // weight re-staging per tile
  return clk / f;
}

void free_layer_device(Layer& L) {
  if (L.d_w32) cudaFree(L.d_w32);
  if (L.d_bias) cudaFree(L.d_bias);
  if (L.d_wpk) cudaFree(L.d_wpk);
  if (L.fused.d_w1) cudaFree(L.fused.d_w1);
  if (L.fused.d_w2) cudaFree(L.fused.d_w2);
  if (L.fused.d_b1) cudaFree(L.fused.d_b1);
  if (L.fused.d_b2) cudaFree(L.fused.d_b2);
  L.fused = Layer::Fused();
  L.d_w32 = nullptr; L.d_bias = nullptr; L.d_wpk = nullptr;
}

struct Sizes {
  size_t max_act_elems = 0;   // largest [B, L, C] activation
  size_t mel_pad_elems = 0;
};

Sizes compute_sizes(const hfg_ctx* h, int B, int T) {
  Sizes s;
  const hfg_config& c = h->cfg;
  size_t L = (size_t)T;
  s.max_act_elems = (size_t)B * L * c.upsample_initial_channel;
  for (int i = 0; i < c.num_upsamples; ++i) {
    L *= c.upsample_rates[i];
    const size_t e = (size_t)B * L * (c.upsample_initial_channel >> (i + 1));
    if (e > s.max_act_elems) s.max_act_elems = e;
  }
  s.mel_pad_elems = (size_t)B * T * h->layers[0].cin_pad;
  return s;
}

// Utterances per forward when the caller leaves the choice to the library: as many as keep one activation buffer
// below 1 GiB (V1 at 862 frames: 64; V2: 256+) -- larger launches amortise the persistent kernels' prologue/drain
// (measured V1 +2.5 %, V2 +14 % from 32 to the cap).  The host path splits the batch so that the copies of one
// micro-batch overlap the compute of its neighbours: in two when the copies are cheap next to the compute (V1), in four
// when the exposed last device->host copy would otherwise be a visible share of the call (V2, V3).
int auto_micro_batch(const hfg_ctx* h, int B, int T, bool host_path, int sample_bytes = 4) {
  const Sizes one = compute_sizes(h, 1, T);
  const size_t per_utt = std::max<size_t>(one.max_act_elems * act_bytes(h->act), 1);
  long long mb = (long long)((size_t(1) << 30) / per_utt);
  if (mb < 1) mb = 1;
  if (host_path && B >= 16) {
    double macs = 0, bytes = 0;                                   // per mel frame
    for (const Layer& L : h->layers) {
      macs += (double)L.k * L.cin * L.cout * (L.transposed ? L.in_rate : L.out_rate);
      bytes += ((double)L.cin * L.in_rate + 2.0 * L.cout * L.out_rate) * act_bytes(h->act);
    }
    const double compute_us = std::max(2.0 * macs * T / 1.0e9, bytes * T / 5.0e6);   // ~1 PFLOP/s, ~5 TB/s
    const double copy_us = (double)T * h->total_up * sample_bytes / 25.0e3;          // ~25 GB/s host link
    const int parts = (copy_us < 0.1 * compute_us || B < 32) ? 2 : 4;
    if (mb > (B + parts - 1) / parts) mb = (B + parts - 1) / parts;
  }
  if (mb > B) mb = B;
  return (int)mb;
}

struct Workspace { char* U; char* P; char* Q; char* T; char* S; float* ACC; char* MEL;
                   char* CP[HFG_MAX_KERNELS]; char* CQ[HFG_MAX_KERNELS]; char* CT[HFG_MAX_KERNELS]; };   // per-chain P/Q/T ([0] = P/Q/T)

bool use_chains(const hfg_ctx* h, int B, int T) {
  return h->chain_frames > 0 && h->cfg.num_kernels > 1 && h->need_sum && !h->need_acc && h->ev_fork &&
         (long long)B * T <= h->chain_frames;
}

size_t workspace_layout(const hfg_ctx* h, int B, int T, char* base, Workspace* w) {
  const Sizes s = compute_sizes(h, B, T);
  const size_t ab = align_up(s.max_act_elems * act_bytes(h->act), 256);
  const size_t accb = h->need_acc ? align_up(s.max_act_elems * sizeof(float), 256) : 0;
  const size_t sumb = h->need_sum ? ab : 0;
  const size_t melb = align_up(s.mel_pad_elems * act_bytes(h->act), 256);
  size_t off = 0;
  auto take = [&](size_t n) { char* p = base ? base + off : nullptr; off += n; return p; };
  char* U = take(ab); char* P = take(ab); char* Q = take(ab); char* Tt = take(ab);
  char* S = take(sumb);
  float* ACC = (float*)take(accb); char* MEL = take(melb);
  if (w) { w->U = U; w->P = P; w->Q = Q; w->T = Tt; w->S = S; w->ACC = ACC; w->MEL = MEL; }
  for (int j = 0; j < h->cfg.num_kernels; ++j) {
    const bool own = j > 0 && use_chains(h, B, T);
    char* cp = own ? take(ab) : P; char* cq = own ? take(ab) : Q; char* ct = own ? take(ab) : Tt;
    if (w) { w->CP[j] = cp; w->CQ[j] = cq; w->CT[j] = ct; }
  }
  return off;
}

struct Runner {
  hfg_ctx* h;
  cudaStream_t s;
  int B;
  cudaError_t status = cudaSuccess;

  // persistent tcgen05 kernel: the whole layer (all phases of a transposed conv) in one launch; x is
  // stored already activated (conv_tc.cu), res is stored activated with slope 1/res_inv
  void conv_tc(const Layer& L, const void* x, int x_rows, void* out, const void* res, float res_inv,
               const void* res2, float out_scale, float out_slope, float* post_out = nullptr, int16_t* post_pcm = nullptr, bool raw_out = false) {
    if (status != cudaSuccess) return;
    const int u = L.u_k, rows = x_rows / L.fold_in;      // kernel-view rows (time folding, see Layer)
    TcArgs a{};
    a.x = x; a.out = out; a.res = res; a.res2 = res2; a.bias = L.d_bias;
    a.B = B; a.x_rows = rows; a.n_q = rows; a.out_rows = rows * u; a.out_stride = u;
    a.c_in = L.cin_pad; a.c_out = L.cout_k; a.post_out = post_out; a.post_pcm = post_pcm; a.post_f = L.post_f;
    a.res_inv = res_inv; a.out_scale = out_scale; a.out_slope = out_slope;
    a.nvar = (int)L.phases.size();
    a.w_base = L.d_wpk; a.w_bytes = L.wpk_bytes;
    for (int p = 0; p < a.nvar; ++p) {
      const Phase& ph = L.phases[p];
      a.var[p].w = (const char*)L.d_wpk + ph.wpk_off;
      a.var[p].ntaps = ph.ntaps; a.var[p].out_off = p;
      for (int j = 0; j < ph.ntaps; ++j) { a.var[p].tap_off[j] = ph.tap_off[j]; a.var[p].kz[j] = ph.kz[j]; }
    }
    a.raw_out = raw_out ? 1 : 0;
    status = h->tf32 ? launch_conv_tf32(a, s) : launch_conv_tc(a, h->act, s, nullptr);
    h->launches++;
  }

  // direct kernel: one launch per output phase; x / res are raw, in_slope applied while staging
  void conv_ffma(const Layer& L, const void* x, int x_rows, void* out, float in_slope, const void* res,
                 float* acc, int mrf_mode, float out_slope) {
    if (status != cudaSuccess) return;
    const int u = L.transposed ? L.stride : 1;
    for (int p = 0; p < (int)L.phases.size(); ++p) {
      const Phase& ph = L.phases[p];
      ConvArgs a{};
      a.x = x; a.bias = L.d_bias; a.out = out; a.res = res; a.acc = acc;
      a.B = B; a.x_rows = x_rows; a.n_q = x_rows; a.out_rows = x_rows * u;
      a.c_in = L.cin_pad; a.c_out = L.cout; a.ntaps = ph.ntaps;
      for (int j = 0; j < ph.ntaps; ++j) a.tap_off[j] = ph.tap_off[j];
      a.out_stride = u; a.out_off = p; a.in_slope = in_slope; a.mrf_mode = mrf_mode;
      a.mrf_div = (float)h->cfg.num_kernels; a.out_slope = out_slope;
      a.w = L.d_w32 + ph.w32_off;
      status = launch_conv_ffma(a, h->act, s);
      h->launches++;
      if (status != cudaSuccess) return;
    }
  }

  // A conv whose input needs activation `slope` (1 = none) and whose output feeds a consumer that would
  // apply `next_slope`; `next_tc` says whether that consumer wants it stored already activated.
  void conv(const Layer& L, const void* x, int x_rows, void* out, float slope, const void* res, float next_slope, bool next_tc) {
    const float os = next_tc ? next_slope : 1.0f;
    if (L.use_tc) conv_tc(L, x, x_rows, out, res, res ? 1.0f / slope_of_res : 1.0f, nullptr, 1.0f, os);
    else conv_ffma(L, x, x_rows, out, slope, res, nullptr, MRF_NONE, os);
  }

  // last conv of resblock j of a stage: residual add + MRF sum / average (reference model.py:79,111,161-167).
  // tcgen05 path: the running raw sum xs lives in a 16-bit buffer `sum` (res2); direct path: fp32 `acc`.
  void conv_mrf(const Layer& L, const void* x, int x_rows, void* out, void* sum, float slope, const void* res,
                float* acc, int j, int nk, float next_slope, bool next_tc) {
    const float os = next_tc ? next_slope : 1.0f;
    if (L.use_tc) {
      const float ri = 1.0f / slope_of_res;
      if (nk == 1) conv_tc(L, x, x_rows, out, res, ri, nullptr, 1.0f, os);
      else if (j == 0) conv_tc(L, x, x_rows, sum, res, ri, nullptr, 1.0f, 1.0f, nullptr, nullptr, true);
      else if (j < nk - 1) conv_tc(L, x, x_rows, sum, res, ri, sum, 1.0f, 1.0f, nullptr, nullptr, true);
      else conv_tc(L, x, x_rows, out, res, ri, sum, 1.0f / (float)nk, os);
    } else {
      const int mode = nk == 1 ? MRF_NONE : (j == 0 ? MRF_FIRST : (j == nk - 1 ? MRF_FINAL : MRF_ADD));
      conv_ffma(L, x, x_rows, out, slope, res, acc, mode, os);
    }
  }

  // one ResBlock1 iteration (convs1[m] -> convs2[m] -> + x, reference model.py:71-80) as one launch of resblock_tc.cu
  void resblock_fused(const Layer& c1, const void* x, int x_rows, void* out, const void* res2, float mid_slope,
                      float out_scale, float out_slope) {
    if (status != cudaSuccess) return;
    const Layer::Fused& F = c1.fused;
    RbArgs a{};
    a.x = x; a.out = out; a.res2 = res2; a.bias1 = F.d_b1; a.bias2 = F.d_b2; a.w1 = F.d_w1; a.w2 = F.d_w2; a.cg = F.cg;
    a.mode = h->cfg.resblock == 1 ? 1 : 2;
    a.B = B; a.n_q = x_rows / F.fold; a.ntaps1 = F.ntaps1; a.ntaps2 = F.ntaps2;
    memcpy(a.off1, F.off1, sizeof(int) * F.ntaps1); memcpy(a.off2, F.off2, sizeof(int) * F.ntaps2);
    memcpy(a.kz1, F.kz1, sizeof a.kz1); memcpy(a.kz2, F.kz2, sizeof a.kz2);
    a.mid_slope = mid_slope; a.res_inv = 1.0f / slope_of_res; a.out_scale = out_scale; a.out_slope = out_slope;
    status = launch_resblock_tc(a, h->act, s);
    h->launches++;
  }

  float slope_of_res = 1.0f;   // slope the residual tensor of the current stage was stored with (1 = raw)
};

// audio_dev (fp32) or pcm_dev (16-bit PCM, see pcm16_of in kernels.cuh): exactly one is non-NULL
int run_forward(hfg_ctx* h, const float* mel_dev, int B, int T, float* audio_dev, int16_t* pcm_dev, char* ws, size_t ws_bytes, cudaStream_t s) {
  const hfg_config& c = h->cfg;
  Workspace w;
  // the layout depends on (B, T) (concurrent resblock chains take extra scratch): never trust the caller's sizing
  const size_t ws_need = workspace_layout(h, B, T, ws, &w);
  if (ws_need > ws_bytes) return fail(h, HFG_ERR_NOMEM, "workspace too small for %d x %d frames: need %zu bytes, got %zu", B, T, ws_need, ws_bytes);
  Runner r{h, s, B};
  const bool chains = use_chains(h, B, T);
  const int nk = c.num_kernels;
  auto cu = [&](cudaError_t e) { if (r.status == cudaSuccess) r.status = e; };
  const float kSlope = 0.1f;                                    // LRELU_SLOPE, reference model.py:14
  const float kFinalSlope = c.compat == 0 ? 0.3f : 0.01f;       // keras LeakyReLU() default, model.py:152
  int mark = 0;
  const int slot = (int)(h->timed_forwards % kTimingSlots);
  long long launches_at_mark = h->launches;
  h->stage_launches.clear();
  auto stamp = [&]() {
    if (mark > 0) h->stage_launches.push_back(h->launches - launches_at_mark);   // mark 0 opens conv_pre
    launches_at_mark = h->launches;
    if (h->stage_timing && mark < h->marks_per_fwd) cudaEventRecord(h->events[(size_t)slot * h->marks_per_fwd + mark], s);
    ++mark;
  };
  auto rb = [&](int n, int m, int which) -> const Layer& { return h->layers[h->i_rb[((size_t)n * HFG_MAX_DILATIONS + m) * 2 + which]]; };
  auto up_layer = [&](int i) -> const Layer& { return h->layers[h->i_ups[i]]; };
  stamp();

  // conv_pre (model.py:156); its consumer is ups[0] behind lrelu(0.1)
  const Layer& pre = h->layers[h->i_pre];
  const void* x0 = mel_dev;
  if (h->act != ACT_F32 || pre.cin_pad != pre.cin) {
    cudaError_t e = launch_cast_pad(mel_dev, w.MEL, (long long)B * T, pre.cin, pre.cin_pad, h->act, s);
    h->launches++;
    if (e != cudaSuccess) return fail(h, HFG_ERR_CUDA, "cast_pad launch failed: %s", cudaGetErrorString(e));
    x0 = w.MEL;
  }
  r.conv(pre, x0, T, w.P, 1.0f, nullptr, kSlope, up_layer(0).use_tc);
  stamp();

  char* x = w.P;          // current stage input
  int L = T;
  bool x_preact = up_layer(0).use_tc;      // is x stored already activated for its consumer?
  for (int i = 0; i < c.num_upsamples; ++i) {
    const Layer& up = up_layer(i);
    const bool rb_tc = rb(i * c.num_kernels, 0, 0).use_tc;       // a stage's resblock convs all share one shape
    const bool last_stage = (i == c.num_upsamples - 1);
    const bool next_tc = last_stage ? rb_tc : up_layer(i + 1).use_tc;   // conv_post reads either form
    const float next_slope = last_stage ? kFinalSlope : kSlope;
    r.conv(up, x, L, w.U, x_preact ? 1.0f : kSlope, nullptr, kSlope, rb_tc);   // model.py:159-160
    L *= up.stride;
    r.slope_of_res = rb_tc ? kSlope : 1.0f;
    char* next_x = nullptr;
    if (chains) cu(cudaEventRecord(h->ev_fork, s));             // ups[i] done: the resblock chains may start
    for (int j = 0; j < nk; ++j) {                              // model.py:162-166
      const int n = i * nk + j;
      const int nd = c.num_dilations[j];
      const char* rcur = w.U;
      char* const P = w.CP[j]; char* const Q = w.CQ[j]; char* const Tj = w.CT[j];
      if (chains && j > 0) { r.s = h->side[j]; cu(cudaStreamWaitEvent(r.s, h->ev_fork, 0)); }
      for (int m = 0; m < nd; ++m) {
        const bool last = (m == nd - 1);
        char* rnext = (rcur == P) ? Q : P;
        if (last && j == nk - 1) next_x = rnext;
        if (c.resblock == 1) {                                   // ResBlock1.call model.py:71-80
          const Layer& c1 = rb(n, m, 0);
          const Layer& c2 = rb(n, m, 1);
          if (c1.fused.on) {                                     // the whole iteration in one launch; xt stays on chip
            if (last && chains && j > 0) cu(cudaStreamWaitEvent(r.s, h->ev_sum[j - 1], 0));
            const float os = next_tc ? next_slope : 1.0f;
            if (!last) r.resblock_fused(c1, rcur, L, rnext, nullptr, kSlope, 1.0f, kSlope);
            else if (nk == 1) r.resblock_fused(c1, rcur, L, rnext, nullptr, kSlope, 1.0f, os);
            else if (j == 0) r.resblock_fused(c1, rcur, L, w.S, nullptr, kSlope, 1.0f, 1.0f);           // xs = rb_0(x)   (model.py:163-164)
            else if (j < nk - 1) r.resblock_fused(c1, rcur, L, w.S, w.S, kSlope, 1.0f, 1.0f);           // xs += rb_j(x)  (model.py:165-166)
            else r.resblock_fused(c1, rcur, L, rnext, w.S, kSlope, 1.0f / (float)nk, os);               // x = xs / num_kernels (model.py:167)
          } else {
          r.conv(c1, rcur, L, Tj, kSlope, nullptr, kSlope, rb_tc);
          if (last && chains && j > 0) cu(cudaStreamWaitEvent(r.s, h->ev_sum[j - 1], 0));   // the running sum of chain j-1
          if (last) r.conv_mrf(c2, Tj, L, rnext, w.S, kSlope, rcur, w.ACC, j, nk, next_slope, next_tc);
          else r.conv(c2, Tj, L, rnext, kSlope, rcur, kSlope, rb_tc);
          }
        } else {                                                 // ResBlock2.call model.py:106-112
          const Layer& c1 = rb(n, m, 0);
          if (m == 0 && c1.fused.on) {                           // the whole ResBlock2 (both convs, both skips) in one launch
            if (chains && j > 0) cu(cudaStreamWaitEvent(r.s, h->ev_sum[j - 1], 0));
            rnext = (rcur == P) ? Q : P;
            if (j == nk - 1) next_x = rnext;
            const float os = next_tc ? next_slope : 1.0f;
            if (nk == 1) r.resblock_fused(c1, rcur, L, rnext, nullptr, kSlope, 1.0f, os);
            else if (j == 0) r.resblock_fused(c1, rcur, L, w.S, nullptr, kSlope, 1.0f, 1.0f);
            else if (j < nk - 1) r.resblock_fused(c1, rcur, L, w.S, w.S, kSlope, 1.0f, 1.0f);
            else r.resblock_fused(c1, rcur, L, rnext, w.S, kSlope, 1.0f / (float)nk, os);
            if (chains && j < nk - 1) cu(cudaEventRecord(h->ev_sum[j], r.s));
            break;
          }
          if (last && chains && j > 0) cu(cudaStreamWaitEvent(r.s, h->ev_sum[j - 1], 0));
          if (last) r.conv_mrf(c1, rcur, L, rnext, w.S, kSlope, rcur, w.ACC, j, nk, next_slope, next_tc);
          else r.conv(c1, rcur, L, rnext, kSlope, rcur, kSlope, rb_tc);
        }
        if (last && chains && j < nk - 1) cu(cudaEventRecord(h->ev_sum[j], r.s));
        rcur = rnext;
      }
    }
    if (chains) {                                                // the last chain's final conv depends on every other chain
      cu(cudaEventRecord(h->ev_join, r.s));
      r.s = s;
      cu(cudaStreamWaitEvent(s, h->ev_join, 0));
    }
    x = next_x;                                                  // x = xs / num_kernels (model.py:167)
    x_preact = next_tc;
    stamp();
  }
  if (r.status != cudaSuccess) return fail(h, HFG_ERR_CUDA, "kernel launch failed: %s", cudaGetErrorString(r.status));

  // leakyReLU(0.3) -> conv_post -> tanh (model.py:169-171)
  const Layer& post = h->layers[h->i_post];
  if (post.use_tc) {            // x is stored activated with the final slope (see hfg_finalize_weights); w.T is free scratch here
    r.conv_tc(post, x, L, w.T, nullptr, 1.0f, nullptr, 1.0f, 1.0f, audio_dev, pcm_dev);
    if (r.status != cudaSuccess) return fail(h, HFG_ERR_CUDA, "conv_post launch failed: %s", cudaGetErrorString(r.status));
    stamp();
    h->n_stage_marks = std::min(mark, h->marks_per_fwd);
    if (h->stage_timing) h->timed_forwards++;
    return HFG_OK;
  }
  PostArgs pa{};
  pa.x = x; pa.w = post.d_w32; pa.out = audio_dev; pa.out_pcm = pcm_dev; pa.B = B; pa.rows = L; pa.c_in = post.cin; pa.ntaps = post.k;
  pa.pad_left = (post.k - 1) / 2; pa.in_slope = x_preact ? 1.0f : kFinalSlope;
  pa.bias = post.b.empty() ? 0.f : post.b[0];
  cudaError_t e = launch_conv_post(pa, h->act, s);
  h->launches++;
  if (e != cudaSuccess) return fail(h, HFG_ERR_CUDA, "conv_post launch failed: %s", cudaGetErrorString(e));
  stamp();
  h->n_stage_marks = std::min(mark, h->marks_per_fwd);
  if (h->stage_timing) h->timed_forwards++;
  return HFG_OK;
}

}  // namespace

// =============================================================================== C ABI
extern "C" {

int hfg_version(void) { return 1000; }

int hfg_create(const hfg_config* cfg, int device, hfg_handle* out) {
  if (!cfg || !out) return fail(nullptr, HFG_ERR_INVALID, "hfg_create: null argument");
  *out = nullptr;
  int rc = validate_config(*cfg);
  if (rc) return rc;
  int ndev = 0;
  cudaError_t e = cudaGetDeviceCount(&ndev);
  if (e != cudaSuccess || ndev == 0)
    return fail(nullptr, HFG_ERR_CUDA, "no CUDA device available (%s); this library has no CPU fallback", cudaGetErrorString(e));
  if (device < 0 || device >= ndev) return fail(nullptr, HFG_ERR_INVALID, "device %d out of range [0,%d)", device, ndev);
  hfg_ctx* h = new hfg_ctx();
  h->cfg = *cfg; h->device = device;
  if (const char* e = getenv("HFG_AUTO_FOLD")) h->auto_fold = atoi(e) != 0;   // A/B knob
  if (const char* e = getenv("HFG_CHAIN_FRAMES")) h->chain_frames = atoll(e);  // A/B knob (option "concurrent_resblocks")
  if (const char* e = getenv("HFG_FUSE")) h->fuse_resblocks = atoi(e) != 0;   // A/B knob (option "fuse_resblocks")
  build_layers(h);
  *out = h;
  return HFG_OK;
}

int hfg_set_weight(hfg_handle h, const char* name, const float* host, const int64_t* shape, int ndim) {
  if (!h) return HFG_ERR_INVALID;
  if (!name || !host || !shape) return fail(h, HFG_ERR_INVALID, "hfg_set_weight: null argument");
  std::string full(name);
  const size_t dot = full.rfind('.');
  if (dot == std::string::npos) return fail(h, HFG_ERR_INVALID, "weight name '%s' must end in .v/.g/.b", name);
  const std::string lname = full.substr(0, dot), kind = full.substr(dot + 1);
  auto it = h->index.find(lname);
  if (it == h->index.end()) return fail(h, HFG_ERR_INVALID, "unknown layer '%s'", lname.c_str());
  Layer& L = h->layers[it->second];
  size_t n = 1;
  for (int i = 0; i < ndim; ++i) n *= (size_t)shape[i];
  if (kind == "v") {
    const int64_t e1 = L.transposed ? L.cout : L.cin, e2 = L.transposed ? L.cin : L.cout;
    if (ndim != 3 || shape[0] != L.k || shape[1] != e1 || shape[2] != e2)
      return fail(h, HFG_ERR_INVALID, "%s: expected shape [%d,%lld,%lld] (Keras %s kernel layout)", name, L.k, (long long)e1, (long long)e2,
                  L.transposed ? "Conv1DTranspose [k,Cout,Cin]" : "Conv1D [k,Cin,Cout]");
    L.v.assign(host, host + n); L.has_v = true;
  } else if (kind == "g") {
    const int64_t e = L.transposed ? L.cin : L.cout;
    if (n != (size_t)e) return fail(h, HFG_ERR_INVALID, "%s: expected %lld elements (last axis of v)", name, (long long)e);
    L.g.assign(host, host + n); L.has_g = true;
  } else if (kind == "b") {
    if (n != (size_t)L.cout) return fail(h, HFG_ERR_INVALID, "%s: expected %d elements", name, L.cout);
    L.b.assign(host, host + n); L.has_b = true;
  } else {
    return fail(h, HFG_ERR_INVALID, "weight name '%s' must end in .v/.g/.b", name);
  }
  h->finalized = false;
  return HFG_OK;
}

int hfg_finalize_weights(hfg_handle h, int precision) {
  if (!h) return HFG_ERR_INVALID;
  if (precision < 0 || precision > 3) return fail(h, HFG_ERR_INVALID, "precision must be 0 (fp32), 1 (fp16), 2 (bf16) or 3 (tf32)");
  for (const Layer& L : h->layers)
    if (!L.has_v || !L.has_g || !L.has_b) return fail(h, HFG_ERR_STATE, "layer '%s' is missing %s", L.name.c_str(), !L.has_v ? "v" : (!L.has_g ? "g" : "b"));
  CUDA_TRY(h, cudaSetDevice(h->device));
  clear_graphs(h);                                     // weight pointers are baked into captured launches
  h->precision = precision;
  h->tf32 = precision == HFG_PRECISION_TF32;
  h->act = (precision == HFG_PRECISION_FP32 || h->tf32) ? ACT_F32 : (precision == HFG_PRECISION_FP16 ? ACT_F16 : ACT_BF16);
  const hfg_config& c = h->cfg;
  // run_forward treats the resblock convs of a stage as one engine (pre-activated 16-bit storage and the 16-bit running
  // MRF sum on the tcgen05 path, raw storage and the fp32 accumulator on the direct path).  A non-stock config can have
  // one conv of a stage fall outside the tcgen05 tile table (tap span > kTcMaxSpan): the whole stage is then demoted to
  // the direct kernels in a second pass.
  std::vector<char> demote(h->layers.size(), 0);
  for (int pass = 0; pass < 2; ++pass) {
  for (size_t li = 0; li < h->layers.size(); ++li) {
    Layer& L = h->layers[li];
    free_layer_device(L);
    const std::vector<float> w = fold_weight_norm(L);   // canonical [k][cin][cout]
    const bool is_post = (L.name == "conv_post"), is_pre = (L.name == "conv_pre");
    bool ok;
    const int pl = (is_pre || is_post) ? (L.dil * (L.k - 1)) / 2 : (L.transposed ? 0 : conv_pad_left(L.k, L.dil, c.compat, &ok));
    const bool tf32 = h->tf32;
    const bool tc_mode = (h->act != ACT_F32 || tf32) && !h->force_ffma && !demote[li];
    // fold factor of a [rows, C] tensor on the tcgen05 path: C*f = 16 when C < 16 and the row count is divisible
    auto fold_of = [&](int C, int rate) { return (tc_mode && C < 16 && 16 % C == 0 && rate % (16 / C) == 0) ? 16 / C : 1; };
    // conv_post rides the tcgen05 kernel (16-column tile, tanh/fp32 epilogue) when its input is stored pre-activated,
    // i.e. when the last stage's resblocks run on the tcgen05 kernel too (layers are finalised in call order)
    const bool post_tc = is_post && tc_mode && !tf32 && li > 0 && h->layers[li - 1].use_tc;   // TF32: conv_post stays on its direct kernel
    KView V;
    L.use_tc = false;
    if (tc_mode && (!is_post || post_tc)) {
      int fi = fold_of(L.cin, L.in_rate), fo = is_post ? fi : fold_of(L.cout, L.out_rate);
      if (L.transposed ? (fi != 1 || L.stride % fo != 0) : (!is_post && fi != fo)) fi = fo = 1;
      if (post_tc && h->auto_fold) {
        // conv_post (C_out = 1 in a 16-column tile): every MMA streams a 4 KB A tile for 8 cycles of math, so the
        // fewest MMAs per sample win: folding by f turns 7 taps x C/16 K-steps per row into about (6 + f) / f + 1
        // taps x C*f/16 K-steps per f samples
        double best = 1e30;
        int best_f = fi;
        for (int f = 1; f <= 16; f *= 2) {
          const int cf = L.cin * f;
          if (cf < 16 || cf > 128 || L.in_rate % f != 0) continue;
          std::map<int, int> taps;
          for (int sidx = 0; sidx < f; ++sidx)
            for (int j = 0; j < L.k; ++j) taps[floordiv(sidx + j - pl, f)] = 1;
          const int nt = (int)taps.size(), span = taps.rbegin()->first - taps.begin()->first;
          if (nt > kMaxTaps || !tc_supported(cf, 16, nt, span)) continue;
          // MMAs actually issued per f samples: a tap's K-step is skipped when none of the fold slots it covers is read
          std::map<int, unsigned> used;                         // kernel-view row offset -> bit mask of 16-channel K-steps read
          for (int sidx = 0; sidx < f; ++sidx)
            for (int j = 0; j < L.k; ++j) {
              const int a = floordiv(sidx + j - pl, f), slot = sidx + j - pl - a * f;
              for (int ch = slot * L.cin; ch < (slot + 1) * L.cin; ch += 1) used[a] |= 1u << (ch / 16);
            }
          int issued = 0;
          for (auto& kv : used) issued += __builtin_popcount(kv.second);
          const double cst = (double)issued / f;
          if (cst < best) { best = cst; best_f = f; }
        }
        fi = fo = best_f;
      }
      if (!L.transposed && !is_pre && !is_post && L.cin == L.cout && h->auto_fold && !tf32) {
        // ResBlock convs: folding is free in memory (bit-identical tensor), so pick the fold factor with the lowest
        // modelled MMA + operand-staging time per output row (see fold_cost)
        double best = fold_cost(L, pl, fi, nullptr);
        for (int f = 1; f <= 16; f *= 2) {
          if (L.cin * f < 16 || L.cin * f > 64 || L.in_rate % f != 0 || f == fi) continue;
          bool ok2 = false;
          const double cst = fold_cost(L, pl, f, &ok2);
          if (ok2 && cst < best * 0.97) { best = cst; fi = fo = f; }
        }
      }
      V = build_view(L, w, pl, fi, fo, post_tc);
      const int cin_mma = is_pre ? (int)align_up((size_t)V.cin_k, tf32 ? 32 : 64) : V.cin_k;
      L.use_tc = (int)V.off.size() <= kMaxVariants && V.max_taps() <= kMaxTaps &&
                 (tf32 ? tf32_supported(cin_mma, V.cout_k, V.max_taps(), V.span()) : tc_supported(cin_mma, V.cout_k, V.max_taps(), V.span()));
      if (L.use_tc) { L.cin_pad = cin_mma; L.fold_in = fi; L.fold_out = fo; }
    }
    if (!L.use_tc) {                                     // direct kernels: unfolded view
      V = build_view(L, w, pl, 1, 1, false);
      L.cin_pad = L.cin; L.fold_in = L.fold_out = 1;
    }
    L.cout_k = V.cout_k; L.u_k = V.u_k; L.post_f = V.post_f;
    L.phases.clear();
    std::vector<float> w32;                              // per phase [ntaps][cin_pad][cout_k]
    for (size_t p = 0; p < V.off.size(); ++p) {
      Phase ph; ph.ntaps = (int)V.off[p].size(); ph.w32_off = w32.size();
      w32.resize(w32.size() + (size_t)ph.ntaps * L.cin_pad * L.cout_k, 0.f);
      for (int t = 0; t < ph.ntaps; ++t) {
        ph.tap_off[t] = V.off[p][t];
        for (int ci = 0; ci < V.cin_k; ++ci)
          memcpy(&w32[ph.w32_off + ((size_t)t * L.cin_pad + ci) * L.cout_k], &V.w[p][t][(size_t)ci * L.cout_k], sizeof(float) * L.cout_k);
        for (int kk = 0; kk < L.cin_pad / 16 && kk < 32; ++kk) {    // zero blocks: unread fold slots of a shifted tap, padded channels
          bool zero = true;
          const float* blk = &w32[ph.w32_off + ((size_t)t * L.cin_pad + (size_t)kk * 16) * L.cout_k];
          for (size_t e = 0; e < (size_t)16 * L.cout_k && zero; ++e) zero = blk[e] == 0.0f;
          if (zero) ph.kz[t] |= 1u << kk;
        }
      }
      {   // a phase whose weights are ALL zero must still start its accumulator: issue everything
        bool any = false;
        const unsigned full = L.cin_pad / 16 >= 32 ? 0xFFFFFFFFu : (1u << (L.cin_pad / 16)) - 1u;
        for (int t = 0; t < ph.ntaps; ++t) any = any || (ph.kz[t] & full) != full;
        if (!any) for (int t = 0; t < ph.ntaps; ++t) ph.kz[t] = 0;
      }
      L.phases.push_back(ph);
    }
    if (L.use_tc) {
      const int span = V.span();
      size_t total = 0;
      for (Phase& ph : L.phases) {
        ph.wpk_off = total;
        total += align_up(tf32 ? tf32_packed_weight_bytes(L.cin_pad, L.cout_k, ph.ntaps) : tc_packed_weight_bytes(L.cin_pad, L.cout_k, ph.ntaps), 256);
      }
      std::vector<char> pk(total);
      for (Phase& ph : L.phases) {
        if (tf32) tf32_pack_weights(&w32[ph.w32_off], ph.ntaps, L.cin_pad, L.cout_k, pk.data() + ph.wpk_off);
        else tc_pack_weights(&w32[ph.w32_off], ph.ntaps, L.cin_pad, L.cout_k, span, h->act, pk.data() + ph.wpk_off);
      }
      L.wpk_bytes = total;
      CUDA_TRY(h, cudaMalloc(&L.d_wpk, total));
      CUDA_TRY(h, cudaMemcpy(L.d_wpk, pk.data(), total, cudaMemcpyHostToDevice));
    } else {
      CUDA_TRY(h, cudaMalloc((void**)&L.d_w32, w32.size() * sizeof(float)));
      CUDA_TRY(h, cudaMemcpy(L.d_w32, w32.data(), w32.size() * sizeof(float), cudaMemcpyHostToDevice));
    }
    std::vector<float> bias(align_up((size_t)L.cout_k, 4), 0.f);
    if (L.post_f > 0) { for (int sidx = 0; sidx < L.post_f; ++sidx) bias[sidx] = L.b[0]; }
    else { for (int sidx = 0; sidx < L.fold_out; ++sidx) memcpy(&bias[(size_t)sidx * L.cout], L.b.data(), sizeof(float) * L.cout); }
    CUDA_TRY(h, cudaMalloc((void**)&L.d_bias, bias.size() * sizeof(float)));
    CUDA_TRY(h, cudaMemcpy(L.d_bias, bias.data(), bias.size() * sizeof(float), cudaMemcpyHostToDevice));
  }
  if (pass == 1) break;
  bool any = false;
  for (int i = 0; i < c.num_upsamples; ++i) {
    bool all_tc = true, none_tc = true;
    std::vector<int> ids;
    for (int j = 0; j < c.num_kernels; ++j)
      for (int m = 0; m < c.num_dilations[j]; ++m)
        for (int which = 0; which < 2; ++which) {
          const int id = h->i_rb[((size_t)(i * c.num_kernels + j) * HFG_MAX_DILATIONS + m) * 2 + which];
          if (id < 0) continue;
          ids.push_back(id);
          if (h->layers[id].use_tc) none_tc = false; else all_tc = false;
        }
    if (!all_tc && !none_tc) { for (int id : ids) demote[id] = 1; any = true; }
  }
  if (!any) break;
  }
  // Fused ResBlock1 iterations (resblock_tc.cu): both convs of an iteration on the tcgen05 path, C*f = 64 with
  // f = 64/C dividing the stage's rows per frame, tap lists / spans that fit the fused kernel's shared-memory plan.
  if (h->fuse_resblocks && h->act != ACT_F32 && !h->force_ffma) {
    for (int n = 0; n < c.num_upsamples * c.num_kernels; ++n)
      for (int m = 0; m < c.num_dilations[n % c.num_kernels]; ++m) {
        // ResBlock1: (convs1[m], convs2[m]) of iteration m; ResBlock2: (convs[0], convs[1]) = the whole block, once
        int i1, i2;
        if (c.resblock == 1) { i1 = h->i_rb[((size_t)n * HFG_MAX_DILATIONS + m) * 2]; i2 = h->i_rb[((size_t)n * HFG_MAX_DILATIONS + m) * 2 + 1]; }
        else {
          if (m != 0 || c.num_dilations[n % c.num_kernels] != 2) continue;
          i1 = h->i_rb[((size_t)n * HFG_MAX_DILATIONS + 0) * 2]; i2 = h->i_rb[((size_t)n * HFG_MAX_DILATIONS + 1) * 2];
        }
        if (i1 < 0 || i2 < 0) continue;
        Layer& A = h->layers[i1]; Layer& Bc = h->layers[i2];
        if (!A.use_tc || !Bc.use_tc || A.cin != A.cout || A.cin > 64 || 64 % A.cin) continue;
        const int f = 64 / A.cin;
        if (A.in_rate % f) continue;
        bool ok1, ok2;
        const int pl1 = conv_pad_left(A.k, A.dil, c.compat, &ok1), pl2 = conv_pad_left(Bc.k, Bc.dil, c.compat, &ok2);
        const KView V1 = build_view(A, fold_weight_norm(A), pl1, f, f, false), V2 = build_view(Bc, fold_weight_norm(Bc), pl2, f, f, false);
        if (V1.off.size() != 1 || V2.off.size() != 1 || V1.max_taps() > kMaxTaps || V2.max_taps() > kMaxTaps) continue;
        RbArgs ra{};
        ra.mode = c.resblock == 1 ? 1 : 2;
        ra.ntaps1 = (int)V1.off[0].size(); ra.ntaps2 = (int)V2.off[0].size();
        for (int t = 0; t < ra.ntaps1; ++t) ra.off1[t] = V1.off[0][t];
        for (int t = 0; t < ra.ntaps2; ++t) ra.off2[t] = V2.off[0][t];
        int cg = 2;
        if (!rb_supported(ra, &cg)) continue;
        Layer::Fused& F = A.fused;
        F.fold = f; F.cg = cg; F.ntaps1 = ra.ntaps1; F.ntaps2 = ra.ntaps2;
        memcpy(F.off1, ra.off1, sizeof(int) * ra.ntaps1); memcpy(F.off2, ra.off2, sizeof(int) * ra.ntaps2);
        auto upload_w = [&](const KView& V, void** dptr, unsigned char* kz) -> int {
          const int nt = (int)V.off[0].size();
          std::vector<float> w32((size_t)nt * 64 * 64);
          for (int t = 0; t < nt; ++t) memcpy(&w32[(size_t)t * 4096], V.w[0][t].data(), sizeof(float) * 4096);
          rb_zero_ksteps(w32.data(), nt, kz);
          std::vector<char> pk(rb_packed_weight_bytes(nt));
          rb_pack_weights(w32.data(), nt, cg, h->act, pk.data());
          CUDA_TRY(h, cudaMalloc(dptr, pk.size()));
          CUDA_TRY(h, cudaMemcpy(*dptr, pk.data(), pk.size(), cudaMemcpyHostToDevice));
          return HFG_OK;
        };
        auto upload_b = [&](const Layer& L, float** dptr) -> int {
          std::vector<float> bias(64);
          for (int sidx = 0; sidx < f; ++sidx) memcpy(&bias[(size_t)sidx * L.cout], L.b.data(), sizeof(float) * L.cout);
          CUDA_TRY(h, cudaMalloc((void**)dptr, 64 * sizeof(float)));
          CUDA_TRY(h, cudaMemcpy(*dptr, bias.data(), 64 * sizeof(float), cudaMemcpyHostToDevice));
          return HFG_OK;
        };
        int rc = upload_w(V1, &F.d_w1, F.kz1); if (rc) return rc;
        rc = upload_w(V2, &F.d_w2, F.kz2); if (rc) return rc;
        rc = upload_b(A, &F.d_b1); if (rc) return rc;
        rc = upload_b(Bc, &F.d_b2); if (rc) return rc;
        F.on = true;
      }
  }
  h->need_acc = false; h->need_sum = false;
  for (const Layer& L : h->layers)
    if (!L.transposed && L.name != "conv_pre" && L.name != "conv_post") { if (L.use_tc) h->need_sum = true; else h->need_acc = true; }
  if (h->events.empty()) {
    h->marks_per_fwd = c.num_upsamples + 3;
    h->events.resize((size_t)h->marks_per_fwd * kTimingSlots);
    for (auto& ev : h->events) CUDA_TRY(h, cudaEventCreate(&ev));
  }
  if (!h->ev_fork && c.num_kernels > 1) {                        // resblock chains of launch-bound forwards (run_forward)
    for (int j = 1; j < c.num_kernels; ++j) CUDA_TRY(h, cudaStreamCreateWithFlags(&h->side[j], cudaStreamNonBlocking));
    for (int j = 0; j < c.num_kernels; ++j) CUDA_TRY(h, cudaEventCreateWithFlags(&h->ev_sum[j], cudaEventDisableTiming));
    CUDA_TRY(h, cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming));
    CUDA_TRY(h, cudaEventCreateWithFlags(&h->ev_fork, cudaEventDisableTiming));
  }
  h->finalized = true;
  return HFG_OK;
}

int hfg_total_upsample(hfg_handle h, int* out) {
  if (!h || !out) return HFG_ERR_INVALID;
  *out = h->total_up;
  return HFG_OK;
}

int hfg_workspace_bytes(hfg_handle h, int B, int T, size_t* out) {
  if (!h || !out) return HFG_ERR_INVALID;
  if (!h->finalized) return fail(h, HFG_ERR_STATE, "hfg_workspace_bytes: weights not finalised");
  if (B < 0 || T < 0) return fail(h, HFG_ERR_INVALID, "negative batch or frame count");
  *out = workspace_layout(h, B, T, nullptr, nullptr) + 256;
  return HFG_OK;
}

// One micro-batch of the host path on h->stream.  Launch-bound forwards (few frames) are replayed from a CUDA graph
// captured on the second use of the same (buffers, shape): 79 launches + 230 tensor-map encodes become one graph launch.
static int forward_maybe_graph(hfg_ctx* h, const float* mel, int nb, int T, char* dev_out, int sample_bytes, char* ws, size_t ws_bytes,
                               cudaStream_t launch, bool own_stream) {
  float* audio = sample_bytes == 4 ? (float*)dev_out : nullptr;
  int16_t* pcm = sample_bytes == 2 ? (int16_t*)dev_out : nullptr;
  bool eligible = h->use_graphs && !h->stage_timing && (long long)nb * T <= 8192;
  if (eligible && !own_stream) {                       // a caller's stream: leave it alone while the caller captures it
    cudaStreamCaptureStatus st = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(launch, &st) != cudaSuccess) { cudaGetLastError(); eligible = false; }
    else if (st != cudaStreamCaptureStatusNone) eligible = false;
  }
  if (!eligible) return run_forward(h, mel, nb, T, audio, pcm, ws, ws_bytes, launch);
  hfg_ctx::GraphEntry* e = nullptr;
  for (auto& g : h->graphs)
    if (g.mel == mel && g.out == dev_out && g.ws == ws && g.B == nb && g.T == T && g.sample_bytes == sample_bytes) { e = &g; break; }
  if (!e) {
    if (h->graphs.size() >= 16) { if (h->graphs.front().exec) cudaGraphExecDestroy(h->graphs.front().exec); h->graphs.erase(h->graphs.begin()); }
    h->graphs.push_back({mel, dev_out, ws, nb, T, sample_bytes, 0, 0, nullptr});
    e = &h->graphs.back();
  }
  if (e->exec) {
    h->launches = e->launches;
    CUDA_TRY(h, cudaGraphLaunch(e->exec, launch));
    return HFG_OK;
  }
  if (e->uses++ == 0) return run_forward(h, mel, nb, T, audio, pcm, ws, ws_bytes, launch);   // first use: plain (also warms per-kernel attributes)
  // capture on an internal stream (a caller's stream may be the legacy default stream, which cannot be captured);
  // the instantiated graph is then launched into the caller's stream like any other work
  cudaStream_t cap = launch;
  if (!own_stream) {
    if (!h->cap_stream) CUDA_TRY(h, cudaStreamCreateWithFlags(&h->cap_stream, cudaStreamNonBlocking));
    cap = h->cap_stream;
  }
  CUDA_TRY(h, cudaStreamBeginCapture(cap, cudaStreamCaptureModeThreadLocal));
  int rc = run_forward(h, mel, nb, T, audio, pcm, ws, ws_bytes, cap);
  cudaGraph_t graph = nullptr;
  cudaError_t ce = cudaStreamEndCapture(cap, &graph);
  if (rc || ce != cudaSuccess || !graph) {
    if (graph) cudaGraphDestroy(graph);
    cudaGetLastError();
    h->use_graphs = false;                               // capture unsupported here: stay on plain launches
    if (rc) return rc;
    return run_forward(h, mel, nb, T, audio, pcm, ws, ws_bytes, launch);
  }
  ce = cudaGraphInstantiate(&e->exec, graph, 0);
  cudaGraphDestroy(graph);
  if (ce != cudaSuccess) { e->exec = nullptr; cudaGetLastError(); h->use_graphs = false; return run_forward(h, mel, nb, T, audio, pcm, ws, ws_bytes, launch); }
  e->launches = h->launches;
  CUDA_TRY(h, cudaGraphLaunch(e->exec, launch));
  return HFG_OK;
}

static int forward_device(hfg_handle h, const float* mel_dev, int B, int T, float* audio_dev, int16_t* pcm_dev,
                          void* workspace_dev, size_t workspace_bytes, void* stream, const char* fn) {
  if (!h) return HFG_ERR_INVALID;
  if (!h->finalized) return fail(h, HFG_ERR_STATE, "%s: weights not finalised", fn);
  if (B < 0 || T < 0) return fail(h, HFG_ERR_INVALID, "negative batch or frame count");
  h->launches = 0;
  if (B == 0 || T == 0) return HFG_OK;                 // empty batch / empty utterances: nothing to synthesise
  if (!mel_dev || (!audio_dev && !pcm_dev) || !workspace_dev) return fail(h, HFG_ERR_INVALID, "%s: null device pointer", fn);
  if ((uintptr_t)mel_dev % 16 || (uintptr_t)audio_dev % 4 || (uintptr_t)pcm_dev % 2)
    return fail(h, HFG_ERR_INVALID, "mel must be 16-byte aligned, the output naturally aligned");
  char* ws = (char*)align_up((size_t)(uintptr_t)workspace_dev, 256);
  const size_t need = workspace_layout(h, B, T, nullptr, nullptr);
  if ((size_t)(ws - (char*)workspace_dev) + need > workspace_bytes)
    return fail(h, HFG_ERR_NOMEM, "workspace too small: need %zu bytes, got %zu", need + 256, workspace_bytes);
  CUDA_TRY(h, cudaSetDevice(h->device));
  // launch-bound forwards are replayed from a CUDA graph captured on the second call with the same buffers and shape
  // (tensor-map encodes and ~60-80 launches become one graph launch into the caller's stream)
  return forward_maybe_graph(h, mel_dev, B, T, audio_dev ? (char*)audio_dev : (char*)pcm_dev, audio_dev ? 4 : 2, ws,
                             workspace_bytes - (size_t)(ws - (char*)workspace_dev), (cudaStream_t)stream, false);
}

int hfg_forward(hfg_handle h, const float* mel_dev, int B, int T, float* audio_dev, void* workspace_dev,
                size_t workspace_bytes, void* stream) {
  return forward_device(h, mel_dev, B, T, audio_dev, nullptr, workspace_dev, workspace_bytes, stream, "hfg_forward");
}

int hfg_forward_pcm16(hfg_handle h, const float* mel_dev, int B, int T, int16_t* pcm_dev, void* workspace_dev,
                      size_t workspace_bytes, void* stream) {
  return forward_device(h, mel_dev, B, T, nullptr, pcm_dev, workspace_dev, workspace_bytes, stream, "hfg_forward_pcm16");
}

// true when the CUDA runtime knows `p` as page-locked host memory (cudaHostAlloc / cudaHostRegister / managed)
static bool host_is_pinned(const void* p) {
  cudaPointerAttributes at{};
  if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
  return at.type == cudaMemoryTypeHost || at.type == cudaMemoryTypeManaged;
}

// host memcpy between pageable and pinned memory, split over a few threads (one thread moves ~6-10 GB/s, far below
// the host link; the first upload and the last download of a call are exposed)
static void staged_memcpy(char* dst, const char* src, size_t bytes) {
  const size_t kChunk = size_t(2) << 20;
  unsigned nt = std::thread::hardware_concurrency();
  nt = nt == 0 ? 1 : (nt > 8 ? 8 : nt);
  if (bytes < 2 * kChunk || nt == 1) { memcpy(dst, src, bytes); return; }
  if ((size_t)nt > bytes / kChunk) nt = (unsigned)(bytes / kChunk);
  const size_t per = (bytes / nt + 4095) & ~size_t(4095);
  std::vector<std::thread> th;
  for (unsigned t = 1; t < nt; ++t) {
    const size_t off = (size_t)t * per;
    if (off >= bytes) break;
    th.emplace_back([=]() { memcpy(dst + off, src + off, std::min(per, bytes - off)); });
  }
  memcpy(dst, src, std::min(per, bytes));
  for (auto& x : th) x.join();
}

static int grow_pinned(hfg_ctx* h, char** buf, size_t* have, size_t need) {
  if (need <= *have) return HFG_OK;
  if (*buf) cudaFreeHost(*buf);
  *buf = nullptr; *have = 0;
  CUDA_TRY(h, cudaHostAlloc((void**)buf, need, cudaHostAllocDefault));
  *have = need;
  return HFG_OK;
}

// host-buffer forward; `sample_bytes` = 4 (fp32 audio) or 2 (16-bit PCM)
static int forward_host(hfg_handle h, const float* mel_host, int B, int T, void* out_host, int sample_bytes, const char* fn) {
  if (!h) return HFG_ERR_INVALID;
  if (!h->finalized) return fail(h, HFG_ERR_STATE, "%s: weights not finalised", fn);
  if (B < 0 || T < 0) return fail(h, HFG_ERR_INVALID, "negative batch or frame count");
  h->launches = 0;
  if (B == 0 || T == 0) return HFG_OK;
  if (!mel_host || !out_host) return fail(h, HFG_ERR_INVALID, "%s: null host pointer", fn);
  CUDA_TRY(h, cudaSetDevice(h->device));
  if (!h->stream) CUDA_TRY(h, cudaStreamCreateWithFlags(&h->stream, cudaStreamNonBlocking));
  int mb = h->micro_batch <= 0 ? auto_micro_batch(h, B, T, true, sample_bytes) : (h->micro_batch > B ? B : h->micro_batch);
  // Micro-batch schedule.  Only the first upload and the last download are exposed, and both scale with the size of
  // their micro-batch; the kernels prefer large batches.  Where the rule asks for four parts (copy-heavy: V2 / V3) the
  // batch is walked as quarter, half, quarter instead: the same exposed copies with a larger middle part.  Utterances
  // are independent, so the split does not change a single output bit; an explicit micro_batch option stays uniform.
  std::vector<int> parts;
  static const int taper = []() { const char* e = getenv("HFG_HOST_TAPER"); return e ? atoi(e) : 1; }();   // A/B knob
  if (taper && h->micro_batch <= 0 && mb >= 8 && mb == (B + 3) / 4 && B - 2 * mb <= auto_micro_batch(h, B, T, false)) {
    const int edge = mb;
    parts.push_back(edge);
    parts.push_back(B - 2 * edge);
    parts.push_back(edge);
    mb = B - 2 * edge;
  } else {
    for (int left = B; left > 0; left -= mb) parts.push_back(left < mb ? left : mb);
  }
  const size_t mel_elems = (size_t)T * h->cfg.num_mels, aud_elems = (size_t)T * h->total_up;
  // every part lays out its own workspace (a small part may run its resblocks as concurrent chains with extra scratch
  // while a larger one does not): size the allocation for the most demanding part
  size_t need_ws = 0;
  for (int nb : parts) need_ws = std::max(need_ws, workspace_layout(h, nb, T, nullptr, nullptr));
  if (need_ws > h->ws_bytes) {
    clear_graphs(h);
    if (h->ws) cudaFree(h->ws);
    h->ws = nullptr; h->ws_bytes = 0;
    CUDA_TRY(h, cudaMalloc(&h->ws, need_ws));
    h->ws_bytes = need_ws;
  }
  if ((size_t)B * mel_elems * 4 > h->d_mel_bytes) {
    clear_graphs(h);
    if (h->d_mel) cudaFree(h->d_mel);
    h->d_mel = nullptr; h->d_mel_bytes = 0;
    CUDA_TRY(h, cudaMalloc((void**)&h->d_mel, (size_t)B * mel_elems * 4));
    h->d_mel_bytes = (size_t)B * mel_elems * 4;
  }
  if ((size_t)B * aud_elems * sample_bytes > h->d_audio_bytes) {
    clear_graphs(h);
    if (h->d_audio) cudaFree(h->d_audio);
    h->d_audio = nullptr; h->d_audio_bytes = 0;
    CUDA_TRY(h, cudaMalloc((void**)&h->d_audio, (size_t)B * aud_elems * sample_bytes));
    h->d_audio_bytes = (size_t)B * aud_elems * sample_bytes;
  }
  // Three streams: the mels of micro-batch i+1 go up and the waveforms of micro-batch i-1 come down while micro-batch i
  // computes; only the first H2D and the last D2H are exposed.
  if (!h->h2d_stream) CUDA_TRY(h, cudaStreamCreateWithFlags(&h->h2d_stream, cudaStreamNonBlocking));
  if (!h->d2h_stream) CUDA_TRY(h, cudaStreamCreateWithFlags(&h->d2h_stream, cudaStreamNonBlocking));
  const int n_mb = (int)parts.size();
  while ((int)h->copy_events.size() < 3 * n_mb) {
    cudaEvent_t ev;
    CUDA_TRY(h, cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    h->copy_events.push_back(ev);
  }
  // Pageable caller memory: cudaMemcpyAsync would stage it synchronously inside the driver and serialise with the kernel
  // launches.  It is staged here instead, one micro-batch at a time: the memcpy of part i+1's mels and of part i-1's
  // waveforms run on the host while part i computes.
  const bool in_direct = host_is_pinned(mel_host), out_direct = host_is_pinned(out_host);
  if (!in_direct) { int rc = grow_pinned(h, &h->pin_in, &h->pin_in_bytes, (size_t)B * mel_elems * 4); if (rc) return rc; }
  if (!out_direct) { int rc = grow_pinned(h, &h->pin_out, &h->pin_out_bytes, (size_t)B * aud_elems * sample_bytes); if (rc) return rc; }
  const char* src_base = in_direct ? (const char*)mel_host : h->pin_in;
  char* dst_base = out_direct ? (char*)out_host : h->pin_out;
  auto drain = [&](int i, int b0) -> int {               // part i's waveforms: pinned staging -> the caller's pageable buffer
    if (out_direct) return HFG_OK;
    CUDA_TRY(h, cudaEventSynchronize(h->copy_events[3 * i + 2]));
    const size_t off = (size_t)b0 * aud_elems * sample_bytes;
    staged_memcpy((char*)out_host + off, h->pin_out + off, (size_t)parts[i] * aud_elems * sample_bytes);
    return HFG_OK;
  };
  auto upload = [&](int i, int b0) -> int {
    const size_t off = (size_t)b0 * mel_elems * 4, bytes = (size_t)parts[i] * mel_elems * 4;
    if (!in_direct) staged_memcpy(h->pin_in + off, (const char*)mel_host + off, bytes);
    CUDA_TRY(h, cudaMemcpyAsync((char*)h->d_mel + off, src_base + off, bytes, cudaMemcpyHostToDevice, h->h2d_stream));
    CUDA_TRY(h, cudaEventRecord(h->copy_events[3 * i], h->h2d_stream));
    return HFG_OK;
  };
  long long launches = 0;
  std::vector<int> starts(n_mb, 0);
  for (int i = 1; i < n_mb; ++i) starts[i] = starts[i - 1] + parts[i - 1];
  if (in_direct) { for (int i = 0; i < n_mb; ++i) { int rc = upload(i, starts[i]); if (rc) return rc; } }
  else { int rc = upload(0, 0); if (rc) return rc; }
  for (int i = 0; i < n_mb; ++i) {
    const int nb = parts[i], b0 = starts[i];
    char* dev_out = (char*)h->d_audio + (size_t)b0 * aud_elems * sample_bytes;
    CUDA_TRY(h, cudaStreamWaitEvent(h->stream, h->copy_events[3 * i], 0));
    int rc = forward_maybe_graph(h, h->d_mel + (size_t)b0 * mel_elems, nb, T, dev_out, sample_bytes, (char*)h->ws, h->ws_bytes, h->stream, true);
    launches += h->launches; h->launches = 0;
    if (rc) { cudaStreamSynchronize(h->h2d_stream); cudaStreamSynchronize(h->stream); cudaStreamSynchronize(h->d2h_stream); return rc; }
    CUDA_TRY(h, cudaEventRecord(h->copy_events[3 * i + 1], h->stream));
    CUDA_TRY(h, cudaStreamWaitEvent(h->d2h_stream, h->copy_events[3 * i + 1], 0));
    CUDA_TRY(h, cudaMemcpyAsync(dst_base + (size_t)b0 * aud_elems * sample_bytes, dev_out,
                                (size_t)nb * aud_elems * sample_bytes, cudaMemcpyDeviceToHost, h->d2h_stream));
    CUDA_TRY(h, cudaEventRecord(h->copy_events[3 * i + 2], h->d2h_stream));
    // while part i runs on the GPU: stage the next part's mels, hand the previous part's waveforms to the caller
    if (!in_direct && i + 1 < n_mb) { rc = upload(i + 1, starts[i + 1]); if (rc) return rc; }
    if (i > 0) { rc = drain(i - 1, starts[i - 1]); if (rc) return rc; }
  }
  h->launches = launches;
  { int rc = drain(n_mb - 1, starts[n_mb - 1]); if (rc) return rc; }
  CUDA_TRY(h, cudaStreamSynchronize(h->d2h_stream));
  CUDA_TRY(h, cudaStreamSynchronize(h->stream));
  return HFG_OK;
}

int hfg_forward_host(hfg_handle h, const float* mel_host, int B, int T, float* audio_host) {
  return forward_host(h, mel_host, B, T, audio_host, 4, "hfg_forward_host");
}

int hfg_forward_host_pcm16(hfg_handle h, const float* mel_host, int B, int T, int16_t* pcm_host) {
  return forward_host(h, mel_host, B, T, pcm_host, 2, "hfg_forward_host_pcm16");
}

int hfg_set_option(hfg_handle h, const char* key, long long value) {
  if (!h || !key) return HFG_ERR_INVALID;
  const std::string k(key);
  if (k == "micro_batch") { if (value < 0) return fail(h, HFG_ERR_INVALID, "micro_batch must be >= 0"); h->micro_batch = (int)value; return HFG_OK; }
  if (k == "stage_timing") { h->stage_timing = value != 0; h->timed_forwards = 0; return HFG_OK; }
  if (k == "force_ffma") { h->force_ffma = value != 0; h->finalized = false; return HFG_OK; }
  if (k == "cuda_graphs") { h->use_graphs = value != 0; if (!h->use_graphs) clear_graphs(h); return HFG_OK; }
  if (k == "auto_fold") { h->auto_fold = value != 0; h->finalized = false; return HFG_OK; }
  if (k == "fuse_resblocks") { h->fuse_resblocks = value != 0; h->finalized = false; return HFG_OK; }
  if (k == "concurrent_resblocks") {      // largest B*T (frames) whose resblocks run as concurrent chains; 0 = never
    if (value < 0) return fail(h, HFG_ERR_INVALID, "concurrent_resblocks must be >= 0 (frames)");
    h->chain_frames = value; clear_graphs(h); return HFG_OK;
  }
  return fail(h, HFG_ERR_INVALID, "unknown option '%s'", key);
}

int hfg_layer_engines(hfg_handle h, char* buf, size_t cap) {
  if (!h || !buf || cap == 0) return HFG_ERR_INVALID;
  if (!h->finalized) return fail(h, HFG_ERR_STATE, "hfg_layer_engines: weights not finalised");
  std::string out;
  for (const Layer& L : h->layers)
    out += L.name + ":" + (L.use_tc ? "tc" : (L.name == "conv_post" ? "post" : "ffma")) + "\n";
  if (out.size() + 1 > cap) return fail(h, HFG_ERR_NOMEM, "need %zu bytes", out.size() + 1);
  memcpy(buf, out.c_str(), out.size() + 1);
  return HFG_OK;
}

int hfg_auto_micro_batch(hfg_handle h, int B, int T, int host_path, int* out) {
  if (!h || !out) return HFG_ERR_INVALID;
  if (!h->finalized) return fail(h, HFG_ERR_STATE, "hfg_auto_micro_batch: weights not finalised");
  if (B < 0 || T < 0) return fail(h, HFG_ERR_INVALID, "negative batch or frame count");
  *out = (B == 0 || T == 0) ? 0 : auto_micro_batch(h, B, T, host_path != 0);
  return HFG_OK;
}

int hfg_layer_folds(hfg_handle h, char* buf, size_t cap) {
  if (!h || !buf || cap == 0) return HFG_ERR_INVALID;
  if (!h->finalized) return fail(h, HFG_ERR_STATE, "hfg_layer_folds: weights not finalised");
  std::string out;
  for (const Layer& L : h->layers) {
    int nt = 0;
    for (const Phase& ph : L.phases) nt = std::max(nt, ph.ntaps);
    out += L.name + ":" + std::to_string(L.fold_in) + ":" + std::to_string(L.fold_out) + ":" + std::to_string(nt) + "\n";
  }
  if (out.size() + 1 > cap) return fail(h, HFG_ERR_NOMEM, "need %zu bytes", out.size() + 1);
  memcpy(buf, out.c_str(), out.size() + 1);
  return HFG_OK;
}

int hfg_last_launch_count(hfg_handle h, long long* out) {
  if (!h || !out) return HFG_ERR_INVALID;
  *out = h->launches;
  return HFG_OK;
}

int hfg_stage_launches(hfg_handle h, int* out, int* n) {
  if (!h || !out || !n) return HFG_ERR_INVALID;
  const int cnt = (int)h->stage_launches.size();
  if (cnt == 0) return fail(h, HFG_ERR_STATE, "no forward ran on this handle");
  if (*n < cnt) return fail(h, HFG_ERR_INVALID, "need room for %d ints", cnt);
  for (int i = 0; i < cnt; ++i) out[i] = (int)h->stage_launches[i];
  *n = cnt;
  return HFG_OK;
}

int hfg_stage_times_ms(hfg_handle h, float* out, int* n) {
  if (!h || !out || !n) return HFG_ERR_INVALID;
  if (!h->stage_timing || h->n_stage_marks < 2 || h->timed_forwards < 1) return fail(h, HFG_ERR_STATE, "stage timing not enabled or no forward ran");
  const int cnt = h->n_stage_marks - 1;
  if (*n < cnt) return fail(h, HFG_ERR_INVALID, "need room for %d floats", cnt);
  // average over the forwards stamped since the last call (at most the kTimingSlots most recent ones)
  const long long nf = h->timed_forwards < kTimingSlots ? h->timed_forwards : kTimingSlots;
  for (int i = 0; i < cnt; ++i) out[i] = 0.f;
  for (long long f = 0; f < nf; ++f) {
    const size_t base = (size_t)((h->timed_forwards - 1 - f) % kTimingSlots) * h->marks_per_fwd;
    CUDA_TRY(h, cudaEventSynchronize(h->events[base + cnt]));
    for (int i = 0; i < cnt; ++i) {
      float ms = 0.f;
      CUDA_TRY(h, cudaEventElapsedTime(&ms, h->events[base + i], h->events[base + i + 1]));
      out[i] += ms / (float)nf;
    }
  }
  *n = cnt;
  h->timed_forwards = 0;
  return HFG_OK;
}

// Single tapped-conv operator on caller-provided device buffers; packs and uploads the weights on every
// call -- meant for operator-level parity tests, not for speed.  engine 0: direct FFMA kernel (ConvArgs
// semantics, in_slope applied to x); engine 1: persistent tcgen05 kernel (TcArgs semantics: x already
// activated, in_slope must be 1; res stored activated with slope 1/res_inv).
int hfg_op_tapped_conv(int act, int engine, const void* x_dev, int B, int x_rows, int c_in,
                       const float* w_host, const float* bias_host, int c_out, int ntaps, const int* tap_off,
                       int out_stride, int out_off, float in_slope, const void* res_dev, float res_inv,
                       const void* res2_dev, float out_scale, float out_slope, float* acc_dev,
                       int mrf_mode, float mrf_div, void* out_dev, void* stream) {
  if (act < 0 || act > 2 || !x_dev || !w_host || !bias_host || !tap_off || ntaps < 1 || ntaps > kMaxTaps)
    return fail(nullptr, HFG_ERR_INVALID, "hfg_op_tapped_conv: bad argument");
  int mn = tap_off[0], mx = tap_off[0];
  for (int j = 0; j < ntaps; ++j) { mn = std::min(mn, tap_off[j]); mx = std::max(mx, tap_off[j]); }
  cudaStream_t s = (cudaStream_t)stream;
  float* d_bias = nullptr; void* d_w = nullptr;
  std::vector<float> bias(align_up((size_t)c_out, 4), 0.f);
  memcpy(bias.data(), bias_host, sizeof(float) * c_out);
  CUDA_TRY(nullptr, cudaMalloc((void**)&d_bias, bias.size() * 4));
  CUDA_TRY(nullptr, cudaMemcpy(d_bias, bias.data(), bias.size() * 4, cudaMemcpyHostToDevice));
  cudaError_t e;
  if (engine == 2) {                                   // TF32 tensor-core kernel (fp32 storage)
    if (act != ACT_F32 || in_slope != 1.0f || mrf_mode != MRF_NONE || !tf32_supported(c_in, c_out, ntaps, mx - mn)) {
      cudaFree(d_bias);
      return fail(nullptr, HFG_ERR_UNSUPPORTED, "tf32 kernel does not cover c_in=%d c_out=%d ntaps=%d span=%d act=%d", c_in, c_out, ntaps, mx - mn, act);
    }
    const size_t nb = align_up(tf32_packed_weight_bytes(c_in, c_out, ntaps), 256);
    std::vector<char> pk(nb);
    tf32_pack_weights(w_host, ntaps, c_in, c_out, pk.data());
    CUDA_TRY(nullptr, cudaMalloc(&d_w, nb));
    CUDA_TRY(nullptr, cudaMemcpy(d_w, pk.data(), nb, cudaMemcpyHostToDevice));
    TcArgs t{};
    t.x = x_dev; t.out = out_dev; t.res = res_dev; t.res2 = res2_dev; t.bias = d_bias;
    t.B = B; t.x_rows = x_rows; t.n_q = x_rows; t.out_rows = x_rows * out_stride; t.out_stride = out_stride;
    t.c_in = c_in; t.c_out = c_out; t.res_inv = res_inv; t.out_scale = out_scale; t.out_slope = out_slope;
    t.nvar = 1; t.var[0].w = d_w; t.var[0].ntaps = ntaps; t.var[0].out_off = out_off;
    for (int j = 0; j < ntaps; ++j) t.var[0].tap_off[j] = tap_off[j];
    e = launch_conv_tf32(t, s);
  } else if (engine == 1) {
    if (act == ACT_F32 || in_slope != 1.0f || mrf_mode != MRF_NONE || !tc_supported(c_in, c_out, ntaps, mx - mn)) {
      cudaFree(d_bias);
      return fail(nullptr, HFG_ERR_UNSUPPORTED, "tcgen05 kernel does not cover c_in=%d c_out=%d ntaps=%d span=%d act=%d in_slope=%g mrf_mode=%d",
                  c_in, c_out, ntaps, mx - mn, act, in_slope, mrf_mode);
    }
    const size_t nb = align_up(tc_packed_weight_bytes(c_in, c_out, ntaps), 256);
    std::vector<char> pk(nb);
    tc_pack_weights(w_host, ntaps, c_in, c_out, mx - mn, (ActType)act, pk.data());
    CUDA_TRY(nullptr, cudaMalloc(&d_w, nb));
    CUDA_TRY(nullptr, cudaMemcpy(d_w, pk.data(), nb, cudaMemcpyHostToDevice));
    TcArgs t{};
    t.x = x_dev; t.out = out_dev; t.res = res_dev; t.res2 = res2_dev; t.bias = d_bias;
    t.B = B; t.x_rows = x_rows; t.n_q = x_rows; t.out_rows = x_rows * out_stride; t.out_stride = out_stride;
    t.c_in = c_in; t.c_out = c_out; t.res_inv = res_inv; t.out_scale = out_scale; t.out_slope = out_slope;
    t.nvar = 1; t.var[0].w = d_w; t.var[0].ntaps = ntaps; t.var[0].out_off = out_off;
    t.w_base = d_w; t.w_bytes = nb;
    for (int j = 0; j < ntaps; ++j) t.var[0].tap_off[j] = tap_off[j];
    e = launch_conv_tc(t, (ActType)act, s, nullptr);
  } else {
    ConvArgs a{};
    a.x = x_dev; a.out = out_dev; a.res = res_dev; a.acc = acc_dev;
    a.B = B; a.x_rows = x_rows; a.n_q = x_rows; a.out_rows = x_rows * out_stride;
    a.c_in = c_in; a.c_out = c_out; a.ntaps = ntaps;
    for (int j = 0; j < ntaps; ++j) a.tap_off[j] = tap_off[j];
    a.out_stride = out_stride; a.out_off = out_off; a.in_slope = in_slope; a.mrf_mode = mrf_mode; a.mrf_div = mrf_div;
    a.out_slope = out_slope;
    const size_t nb = (size_t)ntaps * c_in * c_out * 4;
    CUDA_TRY(nullptr, cudaMalloc(&d_w, nb));
    CUDA_TRY(nullptr, cudaMemcpy(d_w, w_host, nb, cudaMemcpyHostToDevice));
    a.w = d_w; a.bias = d_bias;
    e = launch_conv_ffma(a, (ActType)act, s);
  }
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  cudaFree(d_bias); cudaFree(d_w);
  if (e != cudaSuccess) return fail(nullptr, HFG_ERR_CUDA, "tapped conv failed: %s", cudaGetErrorString(e));
  return HFG_OK;
}

// One fused ResBlock1 iteration (resblock_tc.cu) on caller-provided device buffers; packs and uploads the weights on
// every call -- operator-level parity tests only.
int hfg_op_resblock_fused(int act, int mode, const void* x_dev, int B, int rows, const float* w1_host, const float* b1_host, int ntaps1,
                          const int* off1, const float* w2_host, const float* b2_host, int ntaps2, const int* off2,
                          float mid_slope, float res_inv, const void* res2_dev, float out_scale, float out_slope,
                          void* out_dev, void* stream) {
  if ((act != ACT_F16 && act != ACT_BF16) || !x_dev || !out_dev || !w1_host || !w2_host || !b1_host || !b2_host || !off1 || !off2 ||
      ntaps1 < 1 || ntaps2 < 1 || ntaps1 > kMaxTaps || ntaps2 > kMaxTaps)
    return fail(nullptr, HFG_ERR_INVALID, "hfg_op_resblock_fused: bad argument");
  RbArgs a{};
  a.mode = mode;
  a.x = x_dev; a.out = out_dev; a.res2 = res2_dev; a.B = B; a.n_q = rows; a.ntaps1 = ntaps1; a.ntaps2 = ntaps2;
  memcpy(a.off1, off1, sizeof(int) * ntaps1); memcpy(a.off2, off2, sizeof(int) * ntaps2);
  a.mid_slope = mid_slope; a.res_inv = res_inv; a.out_scale = out_scale; a.out_slope = out_slope;
  int cg = 2;
  if (!rb_supported(a, &cg)) return fail(nullptr, HFG_ERR_UNSUPPORTED, "fused resblock kernel does not cover these taps");
  a.cg = cg;
  cudaStream_t s = (cudaStream_t)stream;
  void* d_w[2] = {nullptr, nullptr}; float* d_b[2] = {nullptr, nullptr};
  const float* wh[2] = {w1_host, w2_host}; const float* bh[2] = {b1_host, b2_host}; const int nt[2] = {ntaps1, ntaps2};
  cudaError_t e = cudaSuccess;
  for (int c = 0; c < 2 && e == cudaSuccess; ++c) {
    std::vector<char> pk(rb_packed_weight_bytes(nt[c]));
    rb_pack_weights(wh[c], nt[c], cg, (ActType)act, pk.data());
    rb_zero_ksteps(wh[c], nt[c], c == 0 ? a.kz1 : a.kz2);
    e = cudaMalloc(&d_w[c], pk.size());
    if (e == cudaSuccess) e = cudaMemcpy(d_w[c], pk.data(), pk.size(), cudaMemcpyHostToDevice);
    if (e == cudaSuccess) e = cudaMalloc((void**)&d_b[c], 64 * sizeof(float));
    if (e == cudaSuccess) e = cudaMemcpy(d_b[c], bh[c], 64 * sizeof(float), cudaMemcpyHostToDevice);
  }
  a.w1 = d_w[0]; a.w2 = d_w[1]; a.bias1 = d_b[0]; a.bias2 = d_b[1];
  if (e == cudaSuccess) e = launch_resblock_tc(a, (ActType)act, s);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s);
  for (int c = 0; c < 2; ++c) { cudaFree(d_w[c]); cudaFree(d_b[c]); }
  if (e != cudaSuccess) return fail(nullptr, HFG_ERR_CUDA, "fused resblock failed: %s", cudaGetErrorString(e));
  return HFG_OK;
}

#ifdef HFG_TC_TRACE
// debug builds only (tools/trace_layer.py); not part of the C ABI
void hfgdbg_set_trace(void* dev) { tc_set_trace((long long*)dev); rb_set_trace((long long*)dev); }
#endif

int hfg_host_alloc(void** out, size_t bytes) {
  if (!out) return HFG_ERR_INVALID;
  cudaError_t e = cudaHostAlloc(out, bytes ? bytes : 1, cudaHostAllocDefault);
  if (e != cudaSuccess) return fail(nullptr, HFG_ERR_CUDA, "cudaHostAlloc(%zu) failed: %s", bytes, cudaGetErrorString(e));
  return HFG_OK;
}

int hfg_host_free(void* p) {
  if (!p) return HFG_OK;
  return cudaFreeHost(p) == cudaSuccess ? HFG_OK : HFG_ERR_CUDA;
}

const char* hfg_last_error(hfg_handle h) { return h ? h->err.c_str() : g_create_error.c_str(); }

void hfg_destroy(hfg_handle h) {
  if (!h) return;
  cudaSetDevice(h->device);
  for (Layer& L : h->layers) free_layer_device(L);
  for (auto& ev : h->events) cudaEventDestroy(ev);
  if (h->ws) cudaFree(h->ws);
  if (h->pin_in) cudaFreeHost(h->pin_in);
  if (h->pin_out) cudaFreeHost(h->pin_out);
  if (h->d_mel) cudaFree(h->d_mel);
  if (h->d_audio) cudaFree(h->d_audio);
  if (h->stream) cudaStreamDestroy(h->stream);
  if (h->cap_stream) cudaStreamDestroy(h->cap_stream);
  if (h->h2d_stream) cudaStreamDestroy(h->h2d_stream);
  if (h->d2h_stream) cudaStreamDestroy(h->d2h_stream);
  for (auto& ev : h->copy_events) cudaEventDestroy(ev);
  for (auto& st : h->side) if (st) cudaStreamDestroy(st);
  for (auto& ev : h->ev_sum) if (ev) cudaEventDestroy(ev);
  if (h->ev_fork) cudaEventDestroy(h->ev_fork);
  if (h->ev_join) cudaEventDestroy(h->ev_join);
  clear_graphs(h);
  delete h;
}

}  // extern "C"
