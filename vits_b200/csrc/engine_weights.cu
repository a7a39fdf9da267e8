// state_dict -> packed device weights.
//
// Reference layouts (paths relative to emotional-vits/):
//   Conv1d.weight           [Cout][Cin][k]                 (models.py:288,301; modules.py:115,126,229-241,335-337)
//   ConvTranspose1d.weight  [Cin][Cout][k], stride u, padding (k-u)//2   (models.py:289-293)
//   Linear.weight           [out][in]                      (modules.py:109,243-245)
//   weight_norm             w = g * v / ||v|| over all dims but 0       (models.py:467-474)
// Packed layout: one K-major matrix [N][k*Cin] per implicit-GEMM conv, tap-major /
// channel-minor, with
//   - Flip (modules.py:278-289) folded into the channel order of flow pre/post,
//   - gated convs' rows tiled so that the tanh half and the sigmoid half of the
//     same channels share one N tile,
//   - ConvTranspose1d rewritten as a stride-1 conv over the input with N = u*Cout
//     phase-stacked outputs (time-major output [L][u*Cout] == [L*u][Cout]),
//   - every conditioning Linear stacked into one [Ncond][gin] matrix in packed
//     column order, conv bias folded into its bias.
#include <cmath>
#include <cstdlib>
#include <cstring>

#include "engine.h"

namespace vitsb {

void Engine::drop_graphs() {
  for (GraphEntry& x : graphs) if (x.exec) cudaGraphExecDestroy(x.exec);
  graphs.clear();
}

void Engine::free_device() {
  if (d_w) { cudaFree(d_w); d_w = nullptr; }
  if (d_bias) { cudaFree(d_bias); d_bias = nullptr; }
  if (d_condW) { cudaFree(d_condW); d_condW = nullptr; }
  if (d_condB) { cudaFree(d_condB); d_condB = nullptr; }
  if (d_post) { cudaFree(d_post); d_post = nullptr; }
}

Engine::~Engine() {
  p1_destroy(p1);
  drop_graphs();
  for (int i = 0; i < 2; ++i) { if (side_stream[i]) cudaStreamDestroy(side_stream[i]); if (ev_join[i]) cudaEventDestroy(ev_join[i]); }
  if (ev_fork) cudaEventDestroy(ev_fork);
  if (cap_stream) cudaStreamDestroy(cap_stream);
  free_device();
}

// ---- options (replace the environment switches of round 1; per handle, visible through the C ABI) ----
namespace {
struct OptDesc { const char* key; int Engine::*field; int lo, hi; };
const OptDesc kOpts[] = {
    {"cuda_graph", &Engine::opt_graph, 0, 1},             // replay the workspace-only launches of a repeated shape as a CUDA graph
    {"pdl", &Engine::opt_pdl, 0, 1},                      // programmatic dependent launch between tensor-core kernels
    {"tc_pair", &Engine::opt_tc_pair, 0, 1},              // CTA-pair mode of conv_tc and of the coupling kernel
    {"rb_fused", &Engine::opt_rb_fused, -1, 1},           // fused ResBlock kernel (-1 auto); applies at the next load
    {"flow_fused", &Engine::opt_flow_fused, -1, 1},       // fused coupling-layer kernel (-1 auto); applies at the next load
    {"flow_in_split", &Engine::opt_flow_in_split, -1, 1}, // 3-term split WN in_layers (-1: bf16 yes, fp16 no); applies at the next load
    {"rb_nslab", &Engine::opt_rb_nslab, 0, 16},           // rb_fused slab count override (0 auto)
    {"ups_split_mask", &Engine::opt_ups_split_mask, -1, 255}, // bit i: upsampler i runs 3-term split precision (16-bit paths); applies at the next load
    {"branch_streams", &Engine::opt_branch_streams, 0, 1}, // the MRF branches of a stage run concurrently on three streams
    {"ws_cap_mb", &Engine::opt_ws_cap_mb, 64, 1 << 20},   // workspace cap: larger batches are processed in row chunks
};
}  // namespace

int Engine::set_option(const char* key, int value) {
  if (!key) return fail(VITSB_ERR_BAD_ARG, "null option key");
  for (const OptDesc& o : kOpts)
    if (std::strcmp(o.key, key) == 0) {
      if (value < o.lo || value > o.hi) return fail(VITSB_ERR_BAD_ARG, std::string("option value out of range: ") + key);
      this->*(o.field) = value;
      return VITSB_OK;
    }
  return fail(VITSB_ERR_BAD_ARG, std::string("unknown option: ") + key);
}

int Engine::get_option(const char* key, int* value) const {
  if (!key || !value) return VITSB_ERR_BAD_ARG;
  for (const OptDesc& o : kOpts)
    if (std::strcmp(o.key, key) == 0) { *value = this->*(o.field); return VITSB_OK; }
  // read-only: the plan actually in force
  if (std::strcmp(key, "plan.rb_fused") == 0) { *value = rb_fused; return VITSB_OK; }
  if (std::strcmp(key, "plan.flow_fused") == 0) { *value = flow_fused; return VITSB_OK; }
  if (std::strcmp(key, "plan.flow_in_split") == 0) { *value = flow_in_split; return VITSB_OK; }
  if (std::strcmp(key, "plan.p1") == 0) { *value = p1 != nullptr; return VITSB_OK; }          // infer_p1 weights are loaded
  return VITSB_ERR_BAD_ARG;
}

int Engine::validate_config() {
  const vitsb_config& c = cfg;
  auto bad = [&](const char* m) { return fail(VITSB_ERR_UNSUPPORTED, std::string("unsupported config: ") + m); };
  if (c.precision < 0 || c.precision > 2) return bad("precision");
  if (c.inter_channels <= 0 || c.inter_channels % 32 != 0) return bad("inter_channels must be a multiple of 32");
  if (c.hidden_channels % 64 != 0) return bad("hidden_channels must be a multiple of 64");
  if (c.gin_channels <= 0 || c.gin_channels % 4 != 0) return bad("gin_channels must be a positive multiple of 4 "
                                                                 "(the reference hot path is always speaker-conditioned)");
  if (c.n_flows < 1 || c.n_flows > VITSB_MAX_LIST) return bad("n_flows");
  // every coupling is preceded by a Flip (models.py:213-217): an odd count leaves the state channel-reversed, which the
  // packed pre / post column orders do not undo
  if (c.n_flows % 2 != 0) return bad("n_flows must be even (the Flips are folded into the weight packing)");
  if (c.flow_kernel_size % 2 != 1) return bad("flow kernel size must be odd (modules.py:96)");
  if (c.flow_n_layers < 1 || c.flow_n_layers > VITSB_MAX_LIST) return bad("flow_n_layers");
  if (c.n_upsamples < 1 || c.n_upsamples > VITSB_MAX_LIST) return bad("n_upsamples");
  if (c.n_resblock_kernels < 1 || c.n_resblock_kernels > VITSB_MAX_LIST) return bad("n_resblock_kernels");
  if (c.n_resblock_dilations < 1 || c.n_resblock_dilations > VITSB_MAX_LIST) return bad("n_resblock_dilations");
  int C = c.upsample_initial_channel;
  if (C % 64 != 0) return bad("upsample_initial_channel must be a multiple of 64");
  for (int i = 0; i < c.n_upsamples; ++i) {
    if (C % 2) return bad("channel ladder");
    C /= 2;
    if (C % 32 != 0) return bad("every stage needs C % 32 == 0 (gate halves of 16 channels)");
    int u = c.upsample_rates[i], k = c.upsample_kernel_sizes[i];
    if (u < 1 || k < u || (k - u) % 2 != 0) return bad("upsample kernel/rate: need k >= u and (k-u) even");
  }
  for (int j = 0; j < c.n_resblock_kernels; ++j) {
    if (c.resblock_kernel_sizes[j] % 2 != 1) return bad("resblock kernel sizes must be odd");
    for (int d = 0; d < c.n_resblock_dilations; ++d) {
      int pad = (c.resblock_kernel_sizes[j] - 1) / 2 * c.resblock_dilation_sizes[j][d];
      if (pad > kHaloRows) return bad("conv receptive field exceeds the masking halo");
    }
  }
  for (int i = 0; i < c.n_flows; ++i) {
    int dr = c.flow_dilation_rate[i];
    int pad = (c.flow_kernel_size - 1) / 2;
    for (int l = 1; l < c.flow_n_layers; ++l) pad *= dr > 1 ? dr : 1;
    if (dr < 1 || pad > kHaloRows) return bad("flow dilation");
  }
  {
    int Cl = c.upsample_initial_channel;
    for (int i = 0; i < c.n_upsamples; ++i) Cl /= 2;
    if (!conv_post_supported(Cl, 7)) return bad("conv_post input channels");
  }
  up_total = 1;
  for (int i = 0; i < c.n_upsamples; ++i) up_total *= c.upsample_rates[i];
  return resolve_plan();
}

// precision + options -> the launch plan (which fused kernels run); called at create and at the top of every load
int Engine::resolve_plan() {
  const vitsb_config& c = cfg;
  op = c.precision == VITSB_PREC_FP32 ? OP_F32 : (c.precision == VITSB_PREC_BF16 ? OP_BF16 : OP_F16);
  split16 = op != OP_F32;
  flow_in_split = split16 && (opt_flow_in_split >= 0 ? opt_flow_in_split == 1 : op == OP_BF16);   // bf16 operands (8x the rounding) keep the split
  // fused ResBlock iterations: every (C, k, dilation) must be a compiled shape, and the MRF average is the separate
  // pass over <= 3 branch streams
  rb_fused = op != OP_F32 && opt_rb_fused != 0 && c.n_resblock_kernels <= 3;
  int Cs = c.upsample_initial_channel;
  for (int i = 0; i < c.n_upsamples && rb_fused; ++i) {
    Cs /= 2;
    for (int j = 0; j < c.n_resblock_kernels; ++j)
      for (int d = 0; d < c.n_resblock_dilations; ++d)
        if (!rb_fused_supported(Cs, c.resblock_kernel_sizes[j], c.resblock_dilation_sizes[j][d], 1)) rb_fused = false;
  }
  flow_fused = op != OP_F32 && opt_flow_fused != 0 && !flow_in_split &&
               flow_fused_supported(c.inter_channels, c.hidden_channels, c.flow_kernel_size, c.flow_n_layers, c.flow_dilation_rate,
                                    c.n_flows, op);
  return VITSB_OK;
}

namespace {

struct HostTensor { std::vector<float> v; std::vector<int64_t> shape; };

float load_elem(const void* p, int dtype, size_t i) {
  if (dtype == VITSB_F32) return reinterpret_cast<const float*>(p)[i];
  if (dtype == VITSB_F16) return __half2float(reinterpret_cast<const __half*>(p)[i]);
  return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p)[i]);
}

struct Loader {
  std::map<std::string, HostTensor> t;
  std::string err;

  // folded weight `name`.weight (or name.weight_g/_v), checked against `shape`
  const HostTensor* weight(const std::string& base, std::vector<int64_t> shape) {
    auto it = t.find(base + ".weight");
    if (it == t.end()) {
      auto ig = t.find(base + ".weight_g"), iv = t.find(base + ".weight_v");
      if (ig == t.end() || iv == t.end()) { err = "missing tensor " + base + ".weight"; return nullptr; }
      const HostTensor& g = ig->second; const HostTensor& v = iv->second;
      if (v.shape.empty() || (int64_t)g.v.size() != v.shape[0]) { err = "bad weight_g shape for " + base; return nullptr; }
      HostTensor w; w.shape = v.shape; w.v.resize(v.v.size());
      size_t inner = v.v.size() / (size_t)v.shape[0];
      for (int64_t r = 0; r < v.shape[0]; ++r) {
        double ss = 0;
        for (size_t i = 0; i < inner; ++i) { double x = v.v[r * inner + i]; ss += x * x; }
        float scale = (float)((double)g.v[r] / std::sqrt(ss));
        for (size_t i = 0; i < inner; ++i) w.v[r * inner + i] = v.v[r * inner + i] * scale;
      }
      it = t.emplace(base + ".weight", std::move(w)).first;
    }
    if (it->second.shape != shape) {
      err = "shape mismatch for " + base + ".weight";
      return nullptr;
    }
    return &it->second;
  }
  const HostTensor* bias(const std::string& base, int64_t n) {
    auto it = t.find(base + ".bias");
    if (it == t.end()) { err = "missing tensor " + base + ".bias"; return nullptr; }
    if (it->second.v.size() != (size_t)n) { err = "shape mismatch for " + base + ".bias"; return nullptr; }
    return &it->second;
  }
};

int pick_nt(int N) {
  if (N <= kMaxTileN) return N;
  for (int nt = kMaxTileN; nt >= 16; nt -= 16)
    if (N % nt == 0) return nt;
  return 16;
}

// packed row -> source out-channel for a gated conv with N_total = 2*Hout
std::vector<int> gate_rows(int N_total, int NT) {
  std::vector<int> r(N_total);
  int Hout = N_total / 2, half = NT / 2;
  for (int n = 0; n < N_total; ++n) {
    int tile = n / NT, q = n % NT;
    r[n] = q < half ? tile * half + q : Hout + tile * half + (q - half);
  }
  return r;
}
std::vector<int> iota_rows(int n0, int n) { std::vector<int> r(n); for (int i = 0; i < n; ++i) r[i] = n0 + i; return r; }

}  // namespace

int p1_load(Engine& e, const vitsb_tensor_desc* t, int n);

int Engine::load(const vitsb_tensor_desc* tens, int n) {
  if (!tens || n <= 0) return fail(VITSB_ERR_BAD_ARG, "no tensors");
  Loader L;
  for (int i = 0; i < n; ++i) {
    const vitsb_tensor_desc& d = tens[i];
    if (!d.name || !d.data) return fail(VITSB_ERR_BAD_ARG, "null tensor name/data");
    if (std::strncmp(d.name, "dec.", 4) != 0 && std::strncmp(d.name, "flow.", 5) != 0) continue;
    if (d.ndim < 1 || d.ndim > 4 || d.dtype < 0 || d.dtype > 2) return fail(VITSB_ERR_BAD_ARG, std::string("bad desc for ") + d.name);
    HostTensor h; size_t cnt = 1;
    for (int k = 0; k < d.ndim; ++k) { h.shape.push_back(d.shape[k]); cnt *= (size_t)d.shape[k]; }
    h.v.resize(cnt);
    for (size_t k = 0; k < cnt; ++k) h.v[k] = load_elem(d.data, d.dtype, k);
    L.t[d.name] = std::move(h);
  }

  const vitsb_config& c = cfg;
  // A reload invalidates everything derived from the old pools: cached graph execs hold the old weight pointers and
  // tensor maps, and a failed reload must not leave a half-built plan marked as loaded.
  drop_graphs();
  loaded = false;
  resolve_plan();
  h_w.clear(); h_bias.clear(); h_condW.clear(); h_condB.clear(); h_post.clear();
  couplings.clear(); stages.clear();
  Ncond = 0; flop_per_frame = 0;
  const int gin = c.gin_channels;
  bool ok = true;

  // generic Conv1d packer.  rows[n] = source out-channel; cols[cp] = source in-channel or -1 (zero)
  // operand-type rounding on the host (for the split-precision hi/lo decomposition)
  auto op_round = [&](float v) -> float {
    return op == OP_BF16 ? __bfloat162float(__float2bfloat16_rn(v)) : __half2float(__float2half_rn(v));
  };
  // writes one packed weight value; split convs get [Wh | Wh | Wl] per tap
  auto put_w = [&](float* dst, const PackedConv& p, int nn, int j, int cp, int C, float v) {
    if (!p.split) { dst[((size_t)nn * p.k + j) * p.Cin + cp] = v; return; }
    const float hi = op_round(v), lo = op_round(v - hi);
    float* row = dst + ((size_t)nn * p.k + j) * p.Cin;
    row[cp] = hi; row[C + cp] = hi; row[2 * C + cp] = lo;
  };
  auto pack_conv = [&](const HostTensor* w, const HostTensor* b, const std::vector<int>& rows, const std::vector<int>& cols,
                       int k, int dil, bool gate, int NT, const HostTensor* condW, const HostTensor* condB, int cond_row0,
                       int true_cin, bool split) -> PackedConv {
    PackedConv p;
    const int C = (int)cols.size();
    p.split = split && split16;
    p.Cin = p.split ? 3 * C : C; p.in_cols = p.split ? 2 * C : C;
    p.N = (int)rows.size(); p.NT = NT; p.k = k; p.dil = dil; p.pad = (k - 1) / 2 * dil; p.gate = gate;
    p.flop_per_pos = 2.0 * true_cin * p.N * k;
    if (!w || !b) { ok = false; return p; }
    const int Cin_src = (int)w->shape[1];
    p.w_off = h_w.size();
    h_w.resize(h_w.size() + (size_t)p.N * k * p.Cin, 0.f);
    float* dst = h_w.data() + p.w_off;
    for (int nn = 0; nn < p.N; ++nn)
      for (int j = 0; j < k; ++j)
        for (int cp = 0; cp < C; ++cp)
          if (cols[cp] >= 0) put_w(dst, p, nn, j, cp, C, w->v[((size_t)rows[nn] * Cin_src + cols[cp]) * k + j]);
    if (condW) {
      p.cond_off = Ncond;
      h_condW.resize((size_t)(Ncond + p.N) * gin);
      h_condB.resize(Ncond + p.N);
      for (int nn = 0; nn < p.N; ++nn) {
        std::memcpy(h_condW.data() + (size_t)(Ncond + nn) * gin, condW->v.data() + (size_t)(cond_row0 + rows[nn]) * gin, sizeof(float) * gin);
        h_condB[Ncond + nn] = condB->v[cond_row0 + rows[nn]] + b->v[rows[nn]];
      }
      Ncond += p.N;
    } else {
      p.bias_off = h_bias.size();
      for (int nn = 0; nn < p.N; ++nn) h_bias.push_back(b->v[rows[nn]]);
      while (h_bias.size() % 4) h_bias.push_back(0.f);
    }
    return p;
  };

  // ---------------- flow (execution order: flows[2*(n_flows-1)] ... flows[0]) ----------------
  const int H = c.hidden_channels, IC = c.inter_channels, half = IC / 2, fk = c.flow_kernel_size, nl = c.flow_n_layers;
  for (int step = 0; step < c.n_flows && ok; ++step) {
    const int fi = c.n_flows - 1 - step;
    const bool flipped = ((step + 1) % 2) == 1;          // number of Flips applied before this coupling is step+1
    const std::string p = "flow.flows." + std::to_string(2 * fi) + ".";
    Coupling cp;
    if (flow_fused) {
      // ---- packing of the fused coupling kernel (flow_fused.cu) ----
      const HostTensor* wpre = L.weight(p + "pre", {H, half, 1});
      const HostTensor* bpre = L.bias(p + "pre", H);
      const HostTensor* wpost = L.weight(p + "post", {half, H, 1});
      const HostTensor* bpost = L.bias(p + "post", half);
      const HostTensor* cw = L.weight(p + "enc.cond_layer", {2 * H * nl, gin});
      const HostTensor* cb = L.bias(p + "enc.cond_layer", 2 * H * nl);
      if (!wpre || !bpre || !wpost || !bpost || !cw || !cb) { ok = false; break; }
      constexpr int XW = 128;                              // x0 window of the z row: [0, 128) or [inter - 128, inter)
      FlowFusedWeights& fw = cp.ff;
      fw.xwin = flipped ? IC - XW : 0;
      fw.coff = cp.coff = flipped ? 0 : half;
      // pre: [H][Wh(XW) | Wl(XW)], zero on the window columns that belong to x1
      fw.w_pre = h_w.size();
      h_w.resize(h_w.size() + (size_t)H * 2 * XW, 0.f);
      for (int nn = 0; nn < H; ++nn)
        for (int wc = 0; wc < XW; ++wc) {
          const int pc = fw.xwin + wc;                     // physical z column
          const int cl = flipped ? IC - 1 - pc : pc;       // logical channel (Flip folded, modules.py:278-289)
          if (cl < 0 || cl >= half) continue;
          const float v = wpre->v[(size_t)nn * half + cl], hi = op_round(v);
          h_w[fw.w_pre + (size_t)nn * 2 * XW + wc] = hi;
          h_w[fw.w_pre + (size_t)nn * 2 * XW + XW + wc] = op_round(v - hi);
        }
      // in_layers: gate rows in 128-row groups [tanh 64 | sigmoid 64], all layers contiguous (weights and cond columns)
      std::vector<const HostTensor*> rsw(nl), rsb(nl);
      for (int l = 0; l < nl && ok; ++l) {
        const std::string li = std::to_string(l);
        PackedConv pc = pack_conv(L.weight(p + "enc.in_layers." + li, {2 * H, H, fk}), L.bias(p + "enc.in_layers." + li, 2 * H),
                                  gate_rows(2 * H, 128), iota_rows(0, H), fk, 1, true, 128, cw, cb, l * 2 * H, H, false);
        if (l == 0) { fw.w_in = pc.w_off; fw.cond_off = pc.cond_off; }
        cp.flop_per_pos += pc.flop_per_pos;
        const int rs = l < nl - 1 ? 2 * H : H;
        rsw[l] = L.weight(p + "enc.res_skip_layers." + li, {rs, H, 1});
        rsb[l] = L.bias(p + "enc.res_skip_layers." + li, rs);
        if (!rsw[l] || !rsb[l]) ok = false;
        else cp.flop_per_pos += 2.0 * H * rs;
      }
      if (!ok) break;
      cp.flop_per_pos += 2.0 * half * H * 2;               // pre + post
      // res | (post o skip): rows [res H | m half] per layer, K = [Wh(H) | Wl(H)].  post is a 1x1 conv on the skip sum
      // (modules.py:366), so post(sum_l skip_l(acts_l) + b) = sum_l (Wpost Wskip_l) acts_l + Wpost sum_l b_l + b_post.
      const int RS = H + half;
      fw.w_rs = h_w.size();
      h_w.resize(h_w.size() + (size_t)nl * RS * 2 * H, 0.f);
      std::vector<double> bskip(H, 0.0);
      std::vector<double> comp((size_t)half * H);
      for (int l = 0; l < nl; ++l) {
        float* base = h_w.data() + fw.w_rs + (size_t)l * RS * 2 * H;
        const bool lastl = l == nl - 1;
        const int skip0 = lastl ? 0 : H;                   // last layer: all H outputs go to the skip sum (modules.py:180-181)
        if (!lastl)
          for (int nn = 0; nn < H; ++nn)
            for (int ci = 0; ci < H; ++ci) {
              const float v = rsw[l]->v[(size_t)nn * H + ci], hi = op_round(v);
              base[(size_t)nn * 2 * H + ci] = hi;
              base[(size_t)nn * 2 * H + H + ci] = op_round(v - hi);
            }
        for (int r = 0; r < half; ++r) {
          const int pr = flipped ? half - 1 - r : r;       // physical x1 offset r <-> logical post row
          for (int ci = 0; ci < H; ++ci) {
            double acc = 0;
            for (int k2 = 0; k2 < H; ++k2) acc += (double)wpost->v[(size_t)pr * H + k2] * (double)rsw[l]->v[(size_t)(skip0 + k2) * H + ci];
            comp[(size_t)r * H + ci] = acc;
          }
        }
        for (int r = 0; r < half; ++r)
          for (int ci = 0; ci < H; ++ci) {
            const float v = (float)comp[(size_t)r * H + ci], hi = op_round(v);
            base[(size_t)(H + r) * 2 * H + ci] = hi;
            base[(size_t)(H + r) * 2 * H + H + ci] = op_round(v - hi);
          }
        for (int k2 = 0; k2 < H; ++k2) bskip[k2] += rsb[l]->v[skip0 + k2];
      }
      // biases: cumulative h bias per layer, and the composed bias of m
      fw.b_cum = h_bias.size();
      {
        std::vector<float> cum(bpre->v.begin(), bpre->v.end());
        for (int l = 0; l < nl; ++l) {
          for (int nn = 0; nn < H; ++nn) h_bias.push_back(cum[nn]);
          if (l < nl - 1) for (int nn = 0; nn < H; ++nn) cum[nn] += rsb[l]->v[nn];
        }
      }
      fw.b_m = h_bias.size();
      for (int r = 0; r < half; ++r) {
        const int pr = flipped ? half - 1 - r : r;
        double acc = bpost->v[pr];
        for (int k2 = 0; k2 < H; ++k2) acc += (double)wpost->v[(size_t)pr * H + k2] * bskip[k2];
        h_bias.push_back((float)acc);
      }
      while (h_bias.size() % 4) h_bias.push_back(0.f);
      flop_per_frame += cp.flop_per_pos;
      couplings.push_back(std::move(cp));
      continue;
    }
    // pre: 1x1 conv over the logical first half; K is the full physical row (zeros on the other half)
    {
      std::vector<int> cols(IC, -1);
      for (int cl = 0; cl < half; ++cl) cols[flipped ? IC - 1 - cl : cl] = cl;
      cp.pre = pack_conv(L.weight(p + "pre", {H, half, 1}), L.bias(p + "pre", H), iota_rows(0, H), cols, 1, 1, false, pick_nt(H),
                         nullptr, nullptr, 0, half, true);
    }
    const HostTensor* cw = L.weight(p + "enc.cond_layer", {2 * H * nl, gin});
    const HostTensor* cb = L.bias(p + "enc.cond_layer", 2 * H * nl);
    if (!cw || !cb) { ok = false; break; }
    int dil = 1;
    for (int l = 0; l < nl && ok; ++l) {
      const std::string li = std::to_string(l);
      const int NT = pick_nt(2 * H);
      const int GG = (NT % 64 == 0) ? 64 : NT;   // gate pairs packed in 64-column groups: the launcher may run any NT that is a multiple of it
      std::vector<int> cols = iota_rows(0, H);
      cp.in_layers.push_back(pack_conv(L.weight(p + "enc.in_layers." + li, {2 * H, H, fk}), L.bias(p + "enc.in_layers." + li, 2 * H),
                                       gate_rows(2 * H, GG), cols, fk, dil, true, NT, cw, cb, l * 2 * H, H, flow_in_split));
      cp.in_layers.back().gate_group = GG;
      if (split16 && !flow_in_split) cp.in_layers.back().in_ld = 2 * H;     // h operand rows stay [hi | lo] (pitch 2H); hi alone is read
      const int rs = l < nl - 1 ? 2 * H : H;
      const HostTensor* rw = L.weight(p + "enc.res_skip_layers." + li, {rs, H, 1});
      const HostTensor* rb = L.bias(p + "enc.res_skip_layers." + li, rs);
      if (l < nl - 1) {
        cp.res.push_back(pack_conv(rw, rb, iota_rows(0, H), cols, 1, 1, false, pick_nt(H), nullptr, nullptr, 0, H, true));
        cp.skip.push_back(pack_conv(rw, rb, iota_rows(H, H), cols, 1, 1, false, pick_nt(H), nullptr, nullptr, 0, H, true));
      } else {
        cp.skip.push_back(pack_conv(rw, rb, iota_rows(0, H), cols, 1, 1, false, pick_nt(H), nullptr, nullptr, 0, H, true));
      }
      dil *= c.flow_dilation_rate[fi];
    }
    // post: logical x1 channel cl lives at physical IC-1-(half+cl) when flipped, else half+cl
    {
      std::vector<int> rows(half);
      for (int r = 0; r < half; ++r) rows[r] = flipped ? half - 1 - r : r;
      cp.coff = flipped ? 0 : half;
      cp.post = pack_conv(L.weight(p + "post", {half, H, 1}), L.bias(p + "post", half), rows, iota_rows(0, H), 1, 1, false,
                          pick_nt(half), nullptr, nullptr, 0, H, true);
    }
    if (!ok) break;
    cp.flop_per_pos = cp.pre.flop_per_pos + cp.post.flop_per_pos;
    for (auto& x : cp.in_layers) cp.flop_per_pos += x.flop_per_pos;
    for (auto& x : cp.res) cp.flop_per_pos += x.flop_per_pos;
    for (auto& x : cp.skip) cp.flop_per_pos += x.flop_per_pos;
    flop_per_frame += cp.flop_per_pos;
    couplings.push_back(std::move(cp));
  }

  // ---------------- decoder ----------------
  int C = c.upsample_initial_channel;
  if (ok) {
    conv_pre = pack_conv(L.weight("dec.conv_pre", {C, IC, 7}), L.bias("dec.conv_pre", C), iota_rows(0, C), iota_rows(0, IC), 7, 1,
                         false, pick_nt(C), nullptr, nullptr, 0, IC, true);
    flop_per_frame += conv_pre.flop_per_pos;
  }
  int pos_per_frame = 1;
  for (int i = 0; i < c.n_upsamples && ok; ++i) {
    const int Cin = C, Cout = C / 2, u = c.upsample_rates[i], k = c.upsample_kernel_sizes[i], pd = (k - u) / 2;
    Stage st; st.u = u; st.C = Cout;
    const std::string up = "dec.ups." + std::to_string(i);
    const HostTensor* w = L.weight(up, {Cin, Cout, k});
    const HostTensor* b = L.bias(up, Cout);
    if (!w || !b) { ok = false; break; }
    // out[u*t + ph] += W[:, :, j] x[t + dlt]  with  j = ph + pd - u*dlt
    int dmin = 0, dmax = 0;
    for (int ph = 0; ph < u; ++ph)
      for (int j = 0; j < k; ++j) {
        int num = ph + pd - j;
        if (num % u != 0) continue;
        int dlt = num / u;
        dmin = dlt < dmin ? dlt : dmin; dmax = dlt > dmax ? dlt : dmax;
      }
    PackedConv& pc = st.ups;
    // automatic: fp16 splits the first two upsamplers only - un-splitting the last two costs 2e-4 of C2 max-abs error
    // (1.43e-3 -> 1.63e-3 of the 2e-3 budget) for 0.25 ms of a 4.4 ms step; bf16 operands keep every split
    const int ups_mask = opt_ups_split_mask >= 0 ? opt_ups_split_mask : (op == OP_BF16 ? 255 : 3);
    pc.split = split16 && ((ups_mask >> i) & 1);       // per-stage choice: DESIGN.md section 3 (error budget of the trunk convs)
    pc.Cin = pc.split ? 3 * Cin : Cin; pc.in_cols = pc.split ? 2 * Cin : Cin;
    pc.N = u * Cout; pc.NT = pick_nt(pc.N); pc.k = dmax - dmin + 1; pc.dil = 1; pc.pad = -dmin; pc.gate = false;
    pc.flop_per_pos = 2.0 * Cin * Cout * k;          // per INPUT position
    if (pc.pad > kHaloRows || dmax > kHaloRows) { ok = false; L.err = "upsample taps exceed halo"; break; }
    pc.w_off = h_w.size();
    h_w.resize(h_w.size() + (size_t)pc.N * pc.k * pc.Cin, 0.f);
    float* dst = h_w.data() + pc.w_off;
    for (int ph = 0; ph < u; ++ph)
      for (int co = 0; co < Cout; ++co)
        for (int jj = 0; jj < pc.k; ++jj) {
          int dlt = jj + dmin, j = ph + pd - u * dlt;
          if (j < 0 || j >= k) continue;
          for (int ci = 0; ci < Cin; ++ci)
            put_w(dst, pc, ph * Cout + co, jj, ci, Cin, w->v[((size_t)ci * Cout + co) * k + j]);
        }
    // which of the packed taps each phase really uses: with 3 taps the early phases use taps {0, 1} and the late ones
    // {1, 2}; the launcher skips the structurally zero tap for N tiles on one side of that boundary
    if (pc.k == 3) {
      int ps = u;                                        // first phase that uses tap 2
      bool pattern = true;
      for (int ph = 0; ph < u; ++ph) {
        bool use[3] = {false, false, false};
        for (int jj = 0; jj < 3; ++jj) { const int j = ph + pd - u * (jj + dmin); use[jj] = j >= 0 && j < k; }
        if (use[2] && ps == u) ps = ph;
        if (ph < ps && use[2]) pattern = false;
        if (ph >= ps && use[0]) pattern = false;
      }
      if (pattern && ps > 0 && ps < u) pc.tap_split_n = ps * Cout;
    }
    pc.bias_off = h_bias.size();
    for (int ph = 0; ph < u; ++ph) for (int co = 0; co < Cout; ++co) h_bias.push_back(b->v[co]);
    while (h_bias.size() % 4) h_bias.push_back(0.f);
    flop_per_frame += pc.flop_per_pos * pos_per_frame;
    pos_per_frame *= u;

    st.rb.resize(c.n_resblock_kernels);
    for (int j = 0; j < c.n_resblock_kernels && ok; ++j) {
      const int kk = c.resblock_kernel_sizes[j];
      const std::string rp = "dec.resblocks." + std::to_string(i * c.n_resblock_kernels + j) + ".";
      for (int d = 0; d < c.n_resblock_dilations && ok; ++d) {
        const std::string ds = std::to_string(d);
        ResIter it;
        const int NT = pick_nt(Cout);
        const HostTensor* gw = L.weight(rp + "conds." + ds, {Cout, gin});
        const HostTensor* gb = L.bias(rp + "conds." + ds, Cout);
        if (!gw || !gb) { ok = false; break; }
        it.c1 = pack_conv(L.weight(rp + "convs1." + ds, {Cout, Cout, kk}), L.bias(rp + "convs1." + ds, Cout), gate_rows(Cout, NT),
                          iota_rows(0, Cout), kk, c.resblock_dilation_sizes[j][d], true, NT, gw, gb, 0, Cout, false);
        it.c2 = pack_conv(L.weight(rp + "convs2." + ds, {Cout, Cout / 2, kk}), L.bias(rp + "convs2." + ds, Cout), iota_rows(0, Cout),
                          iota_rows(0, Cout / 2), kk, 1, false, pick_nt(Cout), nullptr, nullptr, 0, Cout / 2, false);
        flop_per_frame += (it.c1.flop_per_pos + it.c2.flop_per_pos) * pos_per_frame;
        st.rb[j].push_back(it);
      }
    }
    stages.push_back(std::move(st));
    C = Cout;
  }
  if (ok) {
    const HostTensor* w = L.weight("dec.conv_post", {1, C, 7});
    if (!w) ok = false;
    else {
      post_k = 7; post_C = C;
      h_post.resize((size_t)post_k * C);
      for (int j = 0; j < post_k; ++j) for (int ci = 0; ci < C; ++ci) h_post[(size_t)j * C + ci] = w->v[(size_t)ci * post_k + j];
      flop_per_frame += 2.0 * C * post_k * pos_per_frame;
    }
  }
  if (!ok) return fail(VITSB_ERR_WEIGHTS, L.err.empty() ? "weight packing failed" : L.err);

  // every conv the 16-bit plan hands to the tensor-core kernel must be inside its envelope (otherwise the failure
  // would surface as an opaque launch error in the middle of a call)
  if (op != OP_F32) {
    std::vector<const PackedConv*> tc = {&conv_pre};
    for (const Coupling& cp : couplings) {
      if (flow_fused) break;
      tc.push_back(&cp.pre); tc.push_back(&cp.post);
      for (auto& x : cp.in_layers) tc.push_back(&x);
      for (auto& x : cp.res) tc.push_back(&x);
      for (auto& x : cp.skip) tc.push_back(&x);
    }
    for (const Stage& st : stages) {
      tc.push_back(&st.ups);
      if (!rb_fused) for (auto& br : st.rb) for (auto& it : br) { tc.push_back(&it.c1); tc.push_back(&it.c2); }
    }
    for (const PackedConv* pc : tc)
      if (!conv_tc_supported(pc->Cin, pc->in_ld > 0 ? pc->in_ld : pc->in_cols, pc->N, pc->NT, pc->k) || (pc->gate && pc->NT % 32 != 0))
        return fail(VITSB_ERR_UNSUPPORTED, "unsupported config: a conv shape is outside the tensor-core kernel's envelope "
                                           "(channels per tap must be 16, 32 or a multiple of 64)");
  }

  // ---------------- upload ----------------
  free_device();
  cudaError_t e = cudaSetDevice(device);
  auto up = [&](void** dptr, const void* src, size_t bytes) {
    if (e != cudaSuccess) return;
    e = cudaMalloc(dptr, bytes ? bytes : 16);
    if (e == cudaSuccess && bytes) e = cudaMemcpy(*dptr, src, bytes, cudaMemcpyHostToDevice);
  };
  if (op == OP_F32) {
    up(&d_w, h_w.data(), h_w.size() * 4);
  } else if (op == OP_F16) {
    std::vector<__half> tmp(h_w.size());
    for (size_t i = 0; i < h_w.size(); ++i) tmp[i] = __float2half_rn(h_w[i]);
    up(&d_w, tmp.data(), tmp.size() * 2);
  } else {
    std::vector<__nv_bfloat16> tmp(h_w.size());
    for (size_t i = 0; i < h_w.size(); ++i) tmp[i] = __float2bfloat16_rn(h_w[i]);
    up(&d_w, tmp.data(), tmp.size() * 2);
  }
  up((void**)&d_bias, h_bias.data(), h_bias.size() * 4);
  up((void**)&d_condW, h_condW.data(), h_condW.size() * 4);
  up((void**)&d_condB, h_condB.data(), h_condB.size() * 4);
  up((void**)&d_post, h_post.data(), h_post.size() * 4);
  if (e != cudaSuccess) { free_device(); return fail(VITSB_ERR_CUDA, std::string("weight upload: ") + cudaGetErrorString(e)); }
  if (cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess || num_sms <= 0) num_sms = 148;
  std::vector<float>().swap(h_w); std::vector<float>().swap(h_condW);
  loaded = true;
  // optional infer_p1 side (text encoder, duration predictor, speaker table): packed when the checkpoint carries it
  p1_destroy(p1); p1 = nullptr;
  if (p1_present(tens, n)) return p1_load(*this, tens, n);
  return VITSB_OK;
}

}  // namespace vitsb
