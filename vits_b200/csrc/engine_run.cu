// Per-call launch plan: workspace carving and the layer sequence.
//
// Reference control flow being replaced (paths relative to emotional-vits/):
//   flow  : models.py:228-235 -> modules.py:362-375 -> modules.py:157-182
//   dec   : models.py:306-318 -> modules.py:250-260
// Buffer roles (all time-major [B][L][C]):
//   zf/zop      flow state, fp32 master + operand copy        [T][inter]
//   hf/hop      WN residual stream                             [T][H]      (layer-at-a-time flow only)
//   acts        gated activations (operand)                    [T][H]      (layer-at-a-time flow only)
//   skipf/skipop WN skip accumulator                            [T][H]      (layer-at-a-time flow only)
//   xin_op        stage input leaky_relu(x) (after ups)         [L_s][C_s]
//   xb_op/xb2_op  per-branch ResBlock residual stream (ping / pong; xb_op also holds the last branch's result)
//   br0/br1       results of the first two MRF branches
//   nxt_op        leaky_relu(MRF average) = next stage's ups input (conv_pre's output before stage 0)
//   xin_f/xb_f/acc_f/g_op  fp32 residual copies, MRF accumulator and gate tile of the layer-at-a-time plan (fp32 build)
#include "engine.h"

namespace vitsb {

static size_t align_up(size_t x, size_t a = 1024) { return (x + a - 1) / a * a; }

WsLayout Engine::layout(int B, int T) const {
  WsLayout w;
  const size_t ob = op_bytes(op);
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off = align_up(off + bytes); return o; };
  const size_t BT = (size_t)B * T;
  const size_t H = cfg.hidden_channels, IC = cfg.inter_channels;
  const size_t sm = split16 ? 2 : 1;             // split-precision operand rows are [hi | lo]
  w.len = take(sizeof(int) * (size_t)B);
  w.cond = take(sizeof(float) * (size_t)B * (Ncond > 0 ? Ncond : 1));
  w.xlen = take(sizeof(int) * (size_t)B);
  w.cum = take(sizeof(int) * (size_t)B * 1024);
  w.zf = take(4 * BT * IC);  w.zop = take(ob * BT * IC * sm);
  if (!flow_fused) {                             // the fused coupling kernel keeps h / acts / skip on chip
    w.hf = take(4 * BT * H);   w.hop = take(ob * BT * H * sm);
    w.acts = take(ob * BT * H * sm);
    w.skipf = take(4 * BT * H); w.skipop = take(ob * BT * H * sm);
  }
  // decoder: elements per frame of the widest stage tensor
  size_t E = cfg.upsample_initial_channel;     // conv_pre output lives in nxt_op
  size_t pos = 1; int C = cfg.upsample_initial_channel;
  for (int i = 0; i < cfg.n_upsamples; ++i) {
    pos *= cfg.upsample_rates[i]; C /= 2;
    if (pos * C > E) E = pos * C;
  }
  if (rb_fused) {
    // 16-bit streams only: stage input, ping / pong, two branch results (the third lands in ping), next operand
    w.xin_op = take(ob * BT * E); w.xb_op = take(ob * BT * E); w.xb2_op = take(ob * BT * E);
    w.br0 = take(ob * BT * E); w.br1 = take(ob * BT * E);
    w.x3a = take(ob * BT * E); w.x3b = take(ob * BT * E);       // third branch's ping / pong (branches run concurrently)
    const bool need_f32_post = !conv_post_y_supported(post_C, post_k, T * up_total);
    if (need_f32_post) w.acc_f = take(4 * BT * E);
  } else {
    w.xin_f = take(4 * BT * E); w.xb_f = take(4 * BT * E); w.acc_f = take(4 * BT * E);
    w.xin_op = take(ob * BT * E); w.xb_op = take(ob * BT * E); w.g_op = take(ob * BT * E / 2 + 16);
  }
  w.nxt_op = take(ob * BT * E * sm);
  w.total = off;
  return w;
}

// Rows per chunk: the whole batch when its layout fits both the kernels' tile tables and the workspace on offer.
int Engine::chunk_rows(int B, int T, size_t ws_bytes) const {
  int n = B < kMaxBatchRows ? B : kMaxBatchRows;
  if (layout(n, T).total <= ws_bytes) return n;
  const size_t per_row = layout(2, T).total - layout(1, T).total + 2048;     // layout is affine in B up to the 1 KB alignment
  const size_t one = layout(1, T).total;
  if (one > ws_bytes) return 0;
  n = (int)((ws_bytes - one) / per_row) + 1;
  while (n > 1 && layout(n, T).total > ws_bytes) --n;
  return n < 1 ? 0 : n;
}

size_t Engine::workspace_bytes(int B, int T) const {
  const size_t cap = (size_t)opt_ws_cap_mb << 20;
  const int n = B < kMaxBatchRows ? B : kMaxBatchRows;
  const size_t full = layout(n, T).total;
  if (full <= cap) return full;
  const int rows = chunk_rows(B, T, cap);
  return layout(rows > 0 ? rows : 1, T).total;
}

namespace {

// launch tags for the profile table
enum Tag { T_PACK = 0, T_COND = 1, T_UNPACK = 2, T_POST = 3, T_MRF = 4, T_FLOW_PRE = 10, T_FLOW_IN = 11, T_FLOW_RES = 12,
           T_FLOW_SKIP = 13, T_FLOW_POST = 14, T_FLOW_FUSED = 15, T_CONV_PRE = 20, T_UPS = 30, T_RB_C1 = 40, T_RB_C2 = 50,
           T_RB_FUSED = 60 };

struct ProfEv { int tag; double flops; cudaEvent_t a, b; };

struct Ctx {
  Engine* e;
  cudaStream_t stream;
  bool dry;
  bool skip = false;     // current phase is not part of this issue() pass: launches are neither made nor counted
  int launches = 0;
  std::vector<ProfEv> evs;
  double valid_frames = 0;   // sum_b min(len[b], T) (host-side, for algorithmic FLOP attribution)
  void ev_begin(int tag, double flops) {
    if (dry || !e->profile) return;
    ProfEv p{tag, flops, nullptr, nullptr};
    cudaEventCreate(&p.a); cudaEventCreate(&p.b);
    cudaEventRecord(p.a, stream);
    evs.push_back(p);
  }
  void ev_end() {
    if (dry || !e->profile) return;
    cudaEventRecord(evs.back().b, stream);
  }
  cudaError_t err = cudaSuccess;
  int B, T;
  const int* len_flow;   // device lens used for masking (nullptr = all rows live)
  const int* len_dec;
  const float* cond;     // device [B][Ncond]
  LaunchOpts lo;

  const int32_t* h_len = nullptr;   // host copy of len (tile counts for the launcher's wave balancing)
  int m_tiles(int L, int len_mul, bool ragged) const { return count_tiles(h_len, B, L, len_mul, ragged, kTileM, kHaloRows); }
  static int count_tiles(const int32_t* h_len, int B, int L, int len_mul, bool ragged, int tile_rows, int extra) {
    if (!ragged || !h_len) return B * ((L + tile_rows - 1) / tile_rows);
    int q = 0;
    for (int b = 0; b < B; ++b) {
      long long rows = (long long)h_len[b] * len_mul + extra;
      if (rows > L) rows = L;
      q += (int)((rows + tile_rows - 1) / tile_rows);
    }
    return q;
  }

  bool ok() const { return err == cudaSuccess; }

  // p2 / ep2: a second conv over the same input whose packed weights and biases directly follow p's
  // (WN res_skip halves): one launch, N tiles >= p.N take ep2 (tensor-core path only)
  static bool mergeable(const PackedConv& p, const PackedConv& q) {
    return q.w_off == p.w_off + (size_t)p.N * p.k * p.Cin && q.bias_off == p.bias_off + (size_t)p.N && q.NT == p.NT &&
           q.Cin == p.Cin && q.k == p.k && q.dil == p.dil && q.pad == p.pad && q.in_cols == p.in_cols && !p.gate && !q.gate &&
           p.cond_off < 0 && q.cond_off < 0;
  }
  void conv(int tag, const PackedConv& p, const void* in, int L, int len_mul, const int* len, const Epilogue& ep_in,
            const PackedConv* p2 = nullptr, const Epilogue* ep2 = nullptr) {
    if (skip) return;
    ++launches;
    if (dry || !ok()) return;
    ev_begin(tag, (p.flop_per_pos + (p2 ? p2->flop_per_pos : 0.0)) * (len ? valid_frames : (double)B * T) * len_mul);
    ConvArgs a{};
    const int in_ld = p.in_ld > 0 ? p.in_ld : p.in_cols;
    a.in = in; a.in_ld = in_ld; a.in_cols = p.in_cols; a.in_bstride = (long long)L * in_ld; a.Cin = p.Cin;
    a.w = reinterpret_cast<const char*>(e->d_w) + p.w_off * op_bytes(e->op);
    a.N = p.N; a.NT = p.NT; a.k = p.k; a.dil = p.dil; a.pad = p.pad; a.L = L; a.B = B;
    a.gate_group = p.gate_group; a.m_tiles = m_tiles(L, len_mul, len != nullptr);
    a.tap_split_n = p.tap_split_n;
    a.len = len; a.len_mul = len_mul;
    a.epi = ep_in;
    a.epi.gate = p.gate ? 1 : 0;
    if (p.cond_off >= 0) { a.epi.bias = cond + p.cond_off; a.epi.bias_bstride = e->Ncond; }
    else { a.epi.bias = e->d_bias + p.bias_off; a.epi.bias_bstride = 0; }
    if (p2) {
      a.N = p.N + p2->N; a.n_split = p.N;
      a.epi_b = *ep2; a.epi_b.gate = 0;
      a.epi_b.bias = e->d_bias + p2->bias_off; a.epi_b.bias_bstride = 0;
    }
    err = e->op == OP_F32 ? launch_conv_f32(a, stream) : launch_conv_tc(a, e->op, lo, stream);
    ev_end();
  }
};

// dense [B][L][C] epilogue target helpers
Epilogue epi_base() {
  Epilogue ep{};
  ep.alpha = 1.f; ep.beta = 1.f; ep.out_slope = 1.f;
  return ep;
}
void set_res(Epilogue& ep, const float* p, int ld, long long L) { ep.res = p; ep.res_ld = ld; ep.res_bstride = L * ld; }
void set_accum(Epilogue& ep, const float* p, int ld, long long L) { ep.accum = p; ep.accum_ld = ld; ep.accum_bstride = L * ld; }
void set_out_f32(Epilogue& ep, float* p, int ld, long long L) { ep.out_f32 = p; ep.out_f32_ld = ld; ep.out_f32_bstride = L * ld; }
void set_out_op(Epilogue& ep, void* p, int ld, long long L, float slope) {
  ep.out_op = p; ep.out_op_ld = ld; ep.out_op_bstride = L * ld; ep.out_slope = slope;
}
// split-precision operand target: logical C columns, physical row [hi(C) | lo(C)]
void set_out_op_split(Epilogue& ep, void* p, int C, long long L, float slope, bool split) {
  const int ld = split ? 2 * C : C;
  ep.out_op = p; ep.out_op_ld = ld; ep.out_op_bstride = L * ld; ep.out_slope = slope; ep.out_op_split = split ? C : 0;
}
// same row pitch as the split layout, lo half written only when a split conv will read it
void set_out_op_hi(Epilogue& ep, void* p, int C, long long L, bool split_pitch, bool write_lo) {
  set_out_op_split(ep, p, C, L, 1.f, split_pitch);
  if (!write_lo) ep.out_op_split = 0;
}

}  // namespace

bool Engine::side_streams_ready() {
  if (side_stream[0]) return true;
  if (side_failed) return false;
  bool ok = true;
  for (int i = 0; i < 2 && ok; ++i) ok = cudaStreamCreateWithFlags(&side_stream[i], cudaStreamNonBlocking) == cudaSuccess;
  if (ok) ok = cudaEventCreateWithFlags(&ev_fork, cudaEventDisableTiming) == cudaSuccess;
  for (int i = 0; i < 2 && ok; ++i) ok = cudaEventCreateWithFlags(&ev_join[i], cudaEventDisableTiming) == cudaSuccess;
  if (!ok) { cudaGetLastError(); side_failed = true; side_stream[0] = nullptr; }
  return ok;
}

int Engine::check_call(const Call& c, bool dry) {
  if (!dry) {
    if (!loaded) return fail(VITSB_ERR_WEIGHTS, "weights not loaded");
    if (c.dur) {
      const DurFront& d = *c.dur;
      if (!d.m_p || !d.s_p || !d.logw || !d.noise || !c.g || !c.ws) return fail(VITSB_ERR_BAD_ARG, "null pointer");
      if (d.Tx < 1 || d.Tx > 1024) return fail(VITSB_ERR_BAD_ARG, "Tx_max must be in [1, 1024]");
      if (d.noise_stride != 0 && d.noise_stride < c.T) return fail(VITSB_ERR_BAD_ARG, "noise_stride must be 0 (compact) or >= T_cap");
      if (c.len) return fail(VITSB_ERR_BAD_ARG, "durations front end computes the lengths itself");
      if (d.x_len) for (int b = 0; b < c.B; ++b)
        if (d.x_len[b] < 0 || d.x_len[b] > d.Tx) return fail(VITSB_ERR_BAD_ARG, "x_len[b] out of [0, Tx_max]");
    } else if (!c.z_in) return fail(VITSB_ERR_BAD_ARG, "null pointer");
    if (!c.g || !c.ws) return fail(VITSB_ERR_BAD_ARG, "null pointer");
    if ((c.what & RUN_DEC) && !c.wav && !c.pcm) return fail(VITSB_ERR_BAD_ARG, "null wav");
  }
  if (c.B < 1 || c.T < 1) return fail(VITSB_ERR_BAD_ARG, "B and T_max must be >= 1");
  if (c.B > 65535) return fail(VITSB_ERR_BAD_ARG, "B too large (at most 65535 rows per call)");
  if ((long long)c.T * up_total > 0x7fffffffLL / 64) return fail(VITSB_ERR_BAD_ARG, "T_max too large");
  if (!dry && c.len) for (int b = 0; b < c.B; ++b)
    if (c.len[b] < 0 || c.len[b] > c.T) return fail(VITSB_ERR_BAD_ARG, "len[b] out of [0, T_max]");
  return VITSB_OK;
}

int Engine::issue(int phases, const Call& c, bool dry, int* launches) {
  const int B = c.B, T = c.T, what = c.what;
  const int32_t* len = c.len;
  const WsLayout W = layout(B, T);
  if (!dry && c.ws_bytes < W.total) return fail(VITSB_ERR_WORKSPACE, "workspace too small");
  cudaStream_t stream = c.stream;

  char* base = reinterpret_cast<char*>(c.ws);
  auto F = [&](size_t off) { return reinterpret_cast<float*>(base + off); };
  auto P = [&](size_t off) { return reinterpret_cast<void*>(base + off); };
  const size_t ob = op_bytes(op);
  const int IC = cfg.inter_channels, H = cfg.hidden_channels;

  Ctx cx{this, stream, dry};
  cx.B = B; cx.T = T;
  cx.cond = F(W.cond);
  cx.h_len = len;
  cx.lo = lopts();
  int* d_len = reinterpret_cast<int*>(base + W.len);
  const bool dev_len = c.dur != nullptr;               // lengths computed on the device: the host never sees them, grids use upper bounds
  cx.len_flow = (len || dev_len) ? d_len : nullptr;
  cx.len_dec = ((len || dev_len) && c.ragged) ? d_len : nullptr;

#define AUX_LAUNCH(tag, expr) do { if (cx.skip) break; ++cx.launches; if (!dry && cx.ok()) { cx.ev_begin(tag, 0.0); cx.err = (expr); cx.ev_end(); } } while (0)
  if (!dry) {
    if (len) for (int b = 0; b < B; ++b) cx.valid_frames += len[b];
    else cx.valid_frames = (double)B * T;
  }

  if (!dry) {
    cudaError_t e0 = cudaSetDevice(device);
    if (e0 != cudaSuccess) return fail(VITSB_ERR_CUDA, cudaGetErrorString(e0));
    if (len && (phases & PH_FRONT)) cx.err = cudaMemcpyAsync(d_len, len, sizeof(int) * B, cudaMemcpyHostToDevice, stream);
  }
  cx.skip = !(phases & PH_FRONT);
  const bool front_fork = opt_branch_streams && !profile && !dry && !cx.skip && side_streams_ready();
  if (front_fork && cx.ok()) {
    cx.err = cudaEventRecord(ev_fork, stream);               // after everything earlier on the caller's stream (the workspace is reused)
    if (cx.ok()) cx.err = cudaStreamWaitEvent(side_stream[0], ev_fork, 0);
  }
  if (dev_len) {
    // durations -> y_len, cumulative durations; gather m_p + noise * s_p straight into the time-major state   infer.py:169-176, models.py:570-572
    const DurFront& d = *c.dur;
    int* d_xlen = reinterpret_cast<int*>(base + W.xlen);
    if (!dry && !cx.skip && d.x_len && cx.ok()) cx.err = cudaMemcpyAsync(d_xlen, d.x_len, sizeof(int) * B, cudaMemcpyHostToDevice, stream);
    if (!cx.skip) ++cx.launches;                         // two kernels
    AUX_LAUNCH(T_PACK, launch_durations(d.logw, d.x_len ? d_xlen : nullptr, d.m_p, d.s_p, d.noise, d.noise_stride, B, IC, d.Tx, T, d.scale,
                                        reinterpret_cast<int*>(base + W.cum), d_len, d.y_len_out, F(W.zf), P(W.zop), op, split16 ? 1 : 0, stream));
  } else
  // seam: channel-major fp32 -> time-major (masked beyond len)
  AUX_LAUNCH(T_PACK, launch_pack_time_major(c.z_in, B, IC, T, cx.len_flow, (what & RUN_FLOW) ? F(W.zf) : nullptr, P(W.zop), op, split16 ? 1 : 0, stream));
  // all conditioning Linears of the call: flow cond_layers (modules.py:161) + ResBlock conds (modules.py:254).  The GEMV reads
  // only g, the seam kernel above only the latents: with the handle's side streams available the two run concurrently
  // (fork before the seam kernel, join before the first coupling layer).
  if (front_fork && cx.ok()) {
    cudaStream_t gs = side_stream[0];
    AUX_LAUNCH(T_COND, launch_cond_gemv(d_condW, d_condB, c.g, B, Ncond, cfg.gin_channels, F(W.cond), gs));
    if (cx.ok()) cx.err = cudaEventRecord(ev_join[0], gs);
    if (cx.ok()) cx.err = cudaStreamWaitEvent(stream, ev_join[0], 0);
  } else
  AUX_LAUNCH(T_COND, launch_cond_gemv(d_condW, d_condB, c.g, B, Ncond, cfg.gin_channels, F(W.cond), stream));

  cx.skip = !(phases & PH_MID);
  if ((what & RUN_FLOW) && flow_fused) {
    // one kernel per coupling layer: pre -> WN -> post -> x1 -= m, h / acts / m on chip (flow_fused.cu)
    const int t_out = flow_fused_tile_rows(cfg.flow_n_layers);
    for (const Coupling& cp : couplings) {
      if (cx.skip) continue;
      ++cx.launches;
      if (dry || !cx.ok()) continue;
      FlowFusedArgs fa{};
      fa.zf = F(W.zf); fa.zop = P(W.zop); fa.w = d_w; fa.bias = d_bias;
      fa.cond = cx.cond; fa.cond_bstride = Ncond;
      fa.fw = cp.ff; fa.n_layers = cfg.flow_n_layers;
      fa.T = T; fa.B = B; fa.len = cx.len_flow;
      fa.m_tiles = Ctx::count_tiles(len, B, T, 1, len != nullptr, t_out, 0);
      cx.ev_begin(T_FLOW_FUSED, cp.flop_per_pos * cx.valid_frames);
      cx.err = launch_flow_fused(fa, op, cx.lo, stream);
      cx.ev_end();
    }
    if (c.z_out) AUX_LAUNCH(T_UNPACK, launch_unpack_channel_major(F(W.zf), B, IC, T, c.z_out, stream));
  } else if (what & RUN_FLOW) {
    for (const Coupling& cp : couplings) {
      // h = pre(x0)                                                   modules.py:364
      { Epilogue ep = epi_base(); set_out_f32(ep, F(W.hf), H, T); set_out_op_hi(ep, P(W.hop), H, T, split16, flow_in_split);
        cx.conv(T_FLOW_PRE, cp.pre, P(W.zop), T, 1, cx.len_flow, ep); }
      const int nl = (int)cp.in_layers.size();
      for (int l = 0; l < nl; ++l) {
        // acts = tanh(a)*sigmoid(b), a,b = in_layer(h) + cond slice     modules.py:163-172
        { Epilogue ep = epi_base(); set_out_op_split(ep, P(W.acts), H, T, 1.f, split16);
          cx.conv(T_FLOW_IN, cp.in_layers[l], P(W.hop), T, 1, cx.len_flow, ep); }
        // h += res_skip[:H]                                             modules.py:176-178
        Epilogue epr = epi_base(); set_res(epr, F(W.hf), H, T); set_out_f32(epr, F(W.hf), H, T); set_out_op_hi(epr, P(W.hop), H, T, split16, flow_in_split);
        // output += res_skip[H:] (all of it on the last layer)          modules.py:179-181
        Epilogue eps = epi_base();
        if (l > 0) set_accum(eps, F(W.skipf), H, T);
        if (l < nl - 1) set_out_f32(eps, F(W.skipf), H, T); else set_out_op_split(eps, P(W.skipop), H, T, 1.f, split16);
        if (l < nl - 1 && op != OP_F32 && Ctx::mergeable(cp.res[l], cp.skip[l])) {
          cx.conv(T_FLOW_RES, cp.res[l], P(W.acts), T, 1, cx.len_flow, epr, &cp.skip[l], &eps);   // both halves, one launch
        } else {
          if (l < nl - 1) cx.conv(T_FLOW_RES, cp.res[l], P(W.acts), T, 1, cx.len_flow, epr);
          cx.conv(T_FLOW_SKIP, cp.skip[l], P(W.acts), T, 1, cx.len_flow, eps);
        }
      }
      // x1 = x1 - post(out)   (mean-only, logs == 0)                    modules.py:366-372
      { Epilogue ep = epi_base(); ep.alpha = -1.f; ep.beta = 1.f;
        ep.res = F(W.zf) + cp.coff; ep.res_ld = IC; ep.res_bstride = (long long)T * IC;
        ep.out_f32 = F(W.zf) + cp.coff; ep.out_f32_ld = IC; ep.out_f32_bstride = (long long)T * IC;
        const int zld = split16 ? 2 * IC : IC;
        ep.out_op = reinterpret_cast<char*>(P(W.zop)) + (size_t)cp.coff * ob; ep.out_op_ld = zld; ep.out_op_bstride = (long long)T * zld;
        ep.out_op_split = split16 ? IC : 0;
        cx.conv(T_FLOW_POST, cp.post, P(W.skipop), T, 1, cx.len_flow, ep); }
    }
    if (c.z_out) AUX_LAUNCH(T_UNPACK, launch_unpack_channel_major(F(W.zf), B, IC, T, c.z_out, stream));
  }

  if (what & RUN_DEC) {
    // x = conv_pre(z); the next op's leaky_relu(0.1) is folded into the operand copy   models.py:307-309
    { Epilogue ep = epi_base(); set_out_op_split(ep, P(W.nxt_op), conv_pre.N, T, 0.1f, !stages.empty() && stages[0].ups.split);
      cx.conv(T_CONV_PRE, conv_pre, P(W.zop), T, 1, cx.len_dec, ep); }
    int L = T;              // rows per batch item at the current rate
    int mul = 1;            // valid rows = len * mul
    const int ns = (int)stages.size();
    // 16-bit plan: conv_post forms the last stage's MRF average from the branches' y streams itself when the shape allows
    const bool post_from_y = rb_fused && conv_post_y_supported(post_C, post_k, T * up_total);
    for (int s = 0; s < ns; ++s) {
      const Stage& st = stages[s];
      const int C = st.C, Lo = L * st.u;
      // x = ups[i](leaky_relu(x)): phase-stacked conv, output row t holds out[u*t .. u*t+u-1]   models.py:310
      { Epilogue ep = epi_base();
        if (!rb_fused) set_out_f32(ep, F(W.xin_f), st.ups.N, L);
        set_out_op(ep, P(W.xin_op), st.ups.N, L, 0.1f);
        cx.conv(T_UPS + s, st.ups, P(W.nxt_op), L, mul, cx.len_dec, ep); }
      L = Lo; mul *= st.u;
      const int nk = (int)st.rb.size();
      const float inv = 1.0f / nk;
      const float next_slope = 0.1f;                           // models.py:309 (the 0.01 of models.py:315 is in conv_post)
      // 16-bit plan: every MRF branch (models.py:311-314) is an independent chain of fused launches with its own ping / pong
      // buffers, so the branches run CONCURRENTLY on three streams (fork after the upsampler, join before the MRF pass):
      // another branch's CTAs fill the partial last wave of each persistent launch (stage 0 at 16 utterances: 3.4 tiles
      // per CTA), and at one utterance the launches do not fill the GPU at all.  Works the same under graph capture.
      void* bufs[3][2] = {{P(W.xb_op), P(W.xb2_op)}, {P(W.br0), P(W.br1)}, {P(W.x3a), P(W.x3b)}};
      void* br[3] = {nullptr, nullptr, nullptr};             // result buffer of each branch = the one its last iteration writes
      for (int j = 0; j < nk && j < 3; ++j) br[j] = bufs[j][((int)st.rb[j].size() - 1) & 1];   // (known without issuing: the PH_BACK pass needs it)
      const bool fork = rb_fused && opt_branch_streams && !profile && !dry && !cx.skip && nk > 1 && nk <= 3 && side_streams_ready();
      if (fork && cx.ok()) {
        cx.err = cudaEventRecord(ev_fork, stream);
        for (int j = 1; j < nk && cx.ok(); ++j) cx.err = cudaStreamWaitEvent(side_stream[j - 1], ev_fork, 0);
      }
      for (int j = 0; j < nk; ++j) {
        const int ni = (int)st.rb[j].size();
        cudaStream_t bstream = (fork && j > 0) ? side_stream[j - 1] : stream;
        for (int d = 0; d < ni; ++d) {
          const ResIter& it = st.rb[j][d];
          if (rb_fused) {
            // whole iteration in one kernel; the residual stream is y = op(leaky_relu(x)) only (rb_fused.cu).  Every branch
            // ends in an ordinary launch into br[j]; one coalesced pass (mrf_sum_kernel, or conv_post's tile load on the
            // last stage) forms the MRF average (models.py:311-314).
            if (cx.skip) continue;
            ++cx.launches;
            if (dry || !cx.ok()) continue;
            RbLaunch rl{};
            void* ping = bufs[j][d & 1];                            // out of iteration d
            void* pong = bufs[j][(d & 1) ^ 1];                      // out of iteration d-1
            rl.y_in = d == 0 ? P(W.xin_op) : pong;
            rl.w1 = reinterpret_cast<const char*>(d_w) + it.c1.w_off * ob;
            rl.w2 = reinterpret_cast<const char*>(d_w) + it.c2.w_off * ob;
            rl.C = C; rl.k = it.c1.k; rl.dil = it.c1.dil; rl.B = B;
            RbArgs& ra = rl.a;
            ra.L = L; ra.B = B; ra.len = cx.len_dec; ra.len_mul = mul;
            ra.bias1 = cx.cond + it.c1.cond_off; ra.bias1_bstride = Ncond;
            ra.bias2 = d_bias + it.c2.bias_off;
            ra.y_out = ping;                                        // the last iteration's output is the branch result br[j]
            cx.ev_begin(T_RB_FUSED + s, (it.c1.flop_per_pos + it.c2.flop_per_pos) * (cx.len_dec ? cx.valid_frames : (double)B * T) * mul);
            cx.err = launch_rb_fused(rl, op, cx.lo, bstream);
            cx.ev_end();
            continue;
          }
          const void* xin_op = d == 0 ? P(W.xin_op) : P(W.xb_op);
          const float* xin_f = d == 0 ? F(W.xin_f) : F(W.xb_f);
          // xt = tanh(.)*sigmoid(.) of c1(leaky_relu(x)) + cond           modules.py:252-257
          { Epilogue ep = epi_base(); set_out_op(ep, P(W.g_op), C / 2, L, 1.f);
            cx.conv(T_RB_C1 + s, it.c1, xin_op, L, mul, cx.len_dec, ep); }
          // x = c2(xt) + x                                                modules.py:258-259
          Epilogue ep = epi_base();
          set_res(ep, xin_f, C, L);
          if (d < ni - 1) {
            set_out_f32(ep, F(W.xb_f), C, L); set_out_op(ep, P(W.xb_op), C, L, 0.1f);
          } else {
            // MRF: xs += branch; x = xs / num_kernels                      models.py:311-314
            ep.alpha = inv; ep.beta = inv;
            if (j > 0) set_accum(ep, F(W.acc_f), C, L);
            if (j < nk - 1 || s == ns - 1) set_out_f32(ep, F(W.acc_f), C, L);   // last stage: conv_post reads fp32
            else set_out_op_split(ep, P(W.nxt_op), C, L, next_slope, stages[s + 1].ups.split);
          }
          cx.conv(T_RB_C2 + s, it.c2, P(W.g_op), L, mul, cx.len_dec, ep);
        }
      }
      if (fork) {
        for (int j = 1; j < nk && cx.ok(); ++j) {
          cx.err = cudaEventRecord(ev_join[j - 1], side_stream[j - 1]);
          if (cx.ok()) cx.err = cudaStreamWaitEvent(stream, ev_join[j - 1], 0);
        }
      }
      if (rb_fused) {
        if (s < ns - 1) {
          // x = sum_j resblock_j(x) / num_kernels; next operand = leaky_relu(x, 0.1)      models.py:309-314
          AUX_LAUNCH(T_MRF, launch_mrf_sum(br[0], nk > 1 ? br[1] : nullptr, nk > 2 ? br[2] : nullptr, B, L, C, cx.len_dec, mul,
                                           inv, next_slope, P(W.nxt_op), stages[s + 1].ups.split ? 1 : 0, nullptr, op, stream));
        } else if (!post_from_y) {
          AUX_LAUNCH(T_MRF, launch_mrf_sum(br[0], nk > 1 ? br[1] : nullptr, nk > 2 ? br[2] : nullptr, B, L, C, cx.len_dec, mul,
                                           inv, 1.f, nullptr, 0, F(W.acc_f), op, stream));
        }
      }
      if (s == ns - 1) {
        // tanh(conv_post(leaky_relu(x, 0.01)))                                 models.py:315-317
        cx.skip = !(phases & PH_BACK);
        if (post_from_y) {
          AUX_LAUNCH(T_POST, launch_conv_post_y(br[0], nk > 1 ? br[1] : nullptr, nk > 2 ? br[2] : nullptr, inv, B, L, post_C,
                                                post_k, 0.01f, d_post, cx.len_dec, mul, c.wav, c.pcm, c.volume, op, stream));
        } else {
          AUX_LAUNCH(T_POST, launch_conv_post(F(W.acc_f), B, L, post_C, post_k, 0.01f, d_post, cx.len_dec, mul, c.wav, c.pcm, c.volume, stream));
        }
      }
    }
  }
#undef AUX_LAUNCH
  if (launches) *launches = cx.launches;
  if (!cx.evs.empty()) {
    cudaStreamSynchronize(stream);
    for (ProfEv& p : cx.evs) {
      float ms = 0.f;
      if (p.b && cudaEventElapsedTime(&ms, p.a, p.b) == cudaSuccess) {
        ProfRow& r = prof[p.tag];
        r.ms += ms; r.launches += 1; r.flops += p.flops;
      }
      cudaEventDestroy(p.a); if (p.b) cudaEventDestroy(p.b);
    }
  }
  if (!cx.ok()) return fail(VITSB_ERR_CUDA, std::string("launch failed: ") + cudaGetErrorString(cx.err));
  return VITSB_OK;
}

// Everything the captured launches depend on besides device memory contents: the workspace, the shape, and the
// per-rate tile counts the launchers size their grids with.  The lengths themselves are read from device memory.
std::vector<long long> Engine::graph_key(const Call& c) const {
  std::vector<long long> key = {c.what, (long long)(uintptr_t)c.ws, (long long)c.ws_bytes, c.B, c.T, c.ragged, c.len ? 1 : 0,
                                opt_pdl, opt_tc_pair, opt_rb_nslab, c.dur ? 1 : 0, opt_branch_streams};
  if (c.len) {
    int mul = 1;
    key.push_back(Ctx::count_tiles(c.len, c.B, c.T, 1, true, kTileM, kHaloRows));
    key.push_back(Ctx::count_tiles(c.len, c.B, c.T, 1, true, flow_fused_tile_rows(cfg.flow_n_layers), 0));
    for (int i = 0; i < cfg.n_upsamples; ++i) {
      mul *= cfg.upsample_rates[i];
      key.push_back(Ctx::count_tiles(c.len, c.B, c.T * mul, mul, true, kTileM, kHaloRows));
    }
  }
  return key;
}

// One chunk of rows: issue the plan directly, or - from the second call with the same launch geometry on - replay its
// middle part (every launch that touches only the workspace) as an instantiated CUDA graph.  The seam kernels that
// read or write caller pointers stay outside the graph, so fresh input / output buffers do not invalidate it.
int Engine::run_chunk(const Call& c, bool dry, int* launches) {
  if (dry || profile || c.z_out || !opt_graph)
    return issue(PH_ALL, c, dry, launches);

  std::vector<long long> key = graph_key(c);
  GraphEntry* ge = nullptr;
  for (GraphEntry& x : graphs) if (x.key == key) { ge = &x; break; }
  if (!ge) {                                     // first sighting: run directly, remember the shape
    constexpr size_t kMaxGraphs = 16;
    if (graphs.size() >= kMaxGraphs) {
      size_t old = 0;
      for (size_t i = 1; i < graphs.size(); ++i) if (graphs[i].stamp < graphs[old].stamp) old = i;
      if (graphs[old].exec) cudaGraphExecDestroy(graphs[old].exec);
      graphs.erase(graphs.begin() + old);
    }
    GraphEntry n; n.key = std::move(key); n.seen = 1; n.stamp = ++graph_clock;
    graphs.push_back(std::move(n));
    return issue(PH_ALL, c, dry, launches);
  }
  ge->stamp = ++graph_clock;
  int n_front = 0, n_back = 0;
  int rc = issue(PH_FRONT, c, false, &n_front);
  if (rc != VITSB_OK) return rc;
  if (!ge->exec && ge->seen >= 0) {              // second sighting: capture the middle on a private stream
    if (!cap_stream && cudaStreamCreateWithFlags(&cap_stream, cudaStreamNonBlocking) != cudaSuccess) cap_stream = nullptr;
    cudaGraph_t graph = nullptr;
    bool ok = cap_stream && cudaStreamBeginCapture(cap_stream, cudaStreamCaptureModeRelaxed) == cudaSuccess;
    if (ok) {
      int n_mid = 0;
      Call cc = c; cc.stream = cap_stream;
      const int rc_mid = issue(PH_MID, cc, false, &n_mid);
      const cudaError_t ee = cudaStreamEndCapture(cap_stream, &graph);
      ok = rc_mid == VITSB_OK && ee == cudaSuccess && graph != nullptr;
      if (ok) ok = cudaGraphInstantiate(&ge->exec, graph, 0) == cudaSuccess;
      if (graph) cudaGraphDestroy(graph);
      ge->launches = n_mid;
    }
    if (!ok) {                                   // capture unavailable: never retry this shape, issue directly
      cudaGetLastError();
      ge->exec = nullptr; ge->seen = -1;
    }
  }
  if (ge->exec) {
    const cudaError_t le = cudaGraphLaunch(ge->exec, c.stream);
    if (le != cudaSuccess) return fail(VITSB_ERR_CUDA, std::string("graph launch failed: ") + cudaGetErrorString(le));
  } else {
    int n_mid = 0;
    rc = issue(PH_MID, c, false, &n_mid);
    if (rc != VITSB_OK) return rc;
    ge->launches = n_mid;
  }
  rc = issue(PH_BACK, c, false, &n_back);
  if (launches) *launches = n_front + ge->launches + n_back;
  return rc;
}

// Public entry.  Rows are independent (every kernel masks by len[b] and no op crosses batch rows), so a batch whose
// layout exceeds the workspace on offer - or the kernels' 512-row tile tables - is processed in row chunks through the
// same workspace: that is the memory plan for the C4-sized batches (64-512 utterances).
int Engine::run(const Call& c, bool dry, int* launches) {
  int rc = check_call(c, dry);
  if (rc != VITSB_OK) return rc;
  if (dry) {
    Call cc = c;
    cc.B = c.B < kMaxBatchRows ? c.B : kMaxBatchRows;
    return run_chunk(cc, true, launches);
  }
  const int rows = chunk_rows(c.B, c.T, c.ws_bytes);
  if (rows < 1) return fail(VITSB_ERR_WORKSPACE, "workspace too small (needs at least the layout of one row: vitsb_workspace_bytes(h, 1, T_max))");
  if (rows >= c.B) return run_chunk(c, false, launches);
  if (c.dur) return fail(VITSB_ERR_WORKSPACE, "the durations front end does not chunk rows: pass a workspace of vitsb_workspace_bytes(h, B, T_cap)");
  const size_t IC = cfg.inter_channels, gin = cfg.gin_channels, LT = (size_t)c.T * up_total;
  int total = 0;
  for (int b0 = 0; b0 < c.B; b0 += rows) {
    Call cc = c;
    cc.B = c.B - b0 < rows ? c.B - b0 : rows;
    cc.z_in = c.z_in + (size_t)b0 * IC * c.T;
    cc.g = c.g + (size_t)b0 * gin;
    if (c.len) cc.len = c.len + b0;
    if (c.z_out) cc.z_out = c.z_out + (size_t)b0 * IC * c.T;
    if (c.wav) cc.wav = c.wav + (size_t)b0 * LT;
    if (c.pcm) cc.pcm = c.pcm + (size_t)b0 * LT;
    int n = 0;
    rc = run_chunk(cc, false, &n);
    if (rc != VITSB_OK) return rc;
    total += n;
  }
  if (launches) *launches = total;
  return VITSB_OK;
}

}  // namespace vitsb
