// Per-call launch plan: workspace carving and the layer sequence.
//
// Reference control flow being replaced (paths relative to emotional-vits/):
//   flow  : models.py:228-235 -> modules.py:362-375 -> modules.py:157-182
//   dec   : models.py:306-318 -> modules.py:250-260
// Buffer roles (all time-major [B][L][C]):
//   zf/zop      flow state, fp32 master + operand copy        [T][inter]
//   hf/hop      WN residual stream                             [T][H]
//   acts        gated activations (operand)                    [T][H]
//   skipf/skipop WN skip accumulator                            [T][H]
//   xin_f/xin_op  stage input x (after ups) and leaky_relu(x)   [L_s][C_s]
//   xb_f/xb_op    per-branch ResBlock residual stream
//   g_op          gated activations                             [L_s][C_s/2]
//   acc_f         MRF accumulator (sum of branches / num_kernels)
//   nxt_op        leaky_relu(acc) = next stage's ups / conv_post input
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
  w.zf = take(4 * BT * IC);  w.zop = take(ob * BT * IC * sm);
  w.hf = take(4 * BT * H);   w.hop = take(ob * BT * H * sm);
  w.acts = take(ob * BT * H * sm);
  w.skipf = take(4 * BT * H); w.skipop = take(ob * BT * H * sm);
  // decoder: elements per frame of the widest stage tensor
  size_t E = cfg.upsample_initial_channel;     // conv_pre output lives in nxt_op
  size_t pos = 1; int C = cfg.upsample_initial_channel;
  for (int i = 0; i < cfg.n_upsamples; ++i) {
    pos *= cfg.upsample_rates[i]; C /= 2;
    if (pos * C > E) E = pos * C;
  }
  w.xin_f = take(4 * BT * E); w.xb_f = take(4 * BT * E); w.acc_f = take(4 * BT * E);
  w.xin_op = take(ob * BT * E); w.xb_op = take(ob * BT * E); w.xb2_op = take(rb_fused ? ob * BT * E : 16); w.g_op = take(ob * BT * E / 2 + 16); w.nxt_op = take(ob * BT * E * sm);
  w.total = off;
  return w;
}

namespace {

// launch tags for the profile table
enum Tag { T_PACK = 0, T_COND = 1, T_UNPACK = 2, T_POST = 3, T_MRF = 4, T_FLOW_PRE = 10, T_FLOW_IN = 11, T_FLOW_RES = 12,
           T_FLOW_SKIP = 13, T_FLOW_POST = 14, T_CONV_PRE = 20, T_UPS = 30, T_RB_C1 = 40, T_RB_C2 = 50, T_RB_FUSED = 60 };

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

  const int32_t* h_len = nullptr;   // host copy of len (tile counts for the launcher's wave balancing)
  int m_tiles(int L, int len_mul, bool ragged) const {
    if (!ragged || !h_len) return B * ((L + kTileM - 1) / kTileM);
    int q = 0;
    for (int b = 0; b < B; ++b) {
      long long rows = (long long)h_len[b] * len_mul + kHaloRows;
      if (rows > L) rows = L;
      q += (int)((rows + kTileM - 1) / kTileM);
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
    err = e->op == OP_F32 ? launch_conv_f32(a, stream) : launch_conv_tc(a, e->op, stream);
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

int Engine::issue(int phases, int what, const float* z_in, const float* g, const int32_t* len, int B, int T, int ragged,
                  float* z_out, float* wav, void* ws, size_t ws_bytes, cudaStream_t stream, bool dry, int* launches) {
  if (!dry) {
    if (!loaded) return fail(VITSB_ERR_WEIGHTS, "weights not loaded");
    if (!z_in || !g || !ws) return fail(VITSB_ERR_BAD_ARG, "null pointer");
    if ((what & RUN_DEC) && !wav && !pcm_out) return fail(VITSB_ERR_BAD_ARG, "null wav");
  }
  if (B < 1 || T < 1) return fail(VITSB_ERR_BAD_ARG, "B and T_max must be >= 1");
  if ((long long)T * up_total > 0x7fffffffLL / 64) return fail(VITSB_ERR_BAD_ARG, "T_max too large");
  const WsLayout W = layout(B, T);
  if (!dry && ws_bytes < W.total) return fail(VITSB_ERR_WORKSPACE, "workspace too small");
  if (!dry && len) for (int b = 0; b < B; ++b)
    if (len[b] < 0 || len[b] > T) return fail(VITSB_ERR_BAD_ARG, "len[b] out of [0, T_max]");

  char* base = reinterpret_cast<char*>(ws);
  auto F = [&](size_t off) { return reinterpret_cast<float*>(base + off); };
  auto P = [&](size_t off) { return reinterpret_cast<void*>(base + off); };
  const size_t ob = op_bytes(op);
  const int IC = cfg.inter_channels, H = cfg.hidden_channels;

  Ctx cx{this, stream, dry};
  cx.B = B; cx.T = T;
  cx.cond = F(W.cond);
  cx.h_len = len;
  int* d_len = reinterpret_cast<int*>(base + W.len);
  cx.len_flow = len ? d_len : nullptr;
  cx.len_dec = (len && ragged) ? d_len : nullptr;

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
  // seam: channel-major fp32 -> time-major (masked beyond len)
  AUX_LAUNCH(T_PACK, launch_pack_time_major(z_in, B, IC, T, cx.len_flow, (what & RUN_FLOW) ? F(W.zf) : nullptr, P(W.zop), op, split16 ? 1 : 0, stream));
  // all conditioning Linears of the call: flow cond_layers (modules.py:161) + ResBlock conds (modules.py:254)
  AUX_LAUNCH(T_COND, launch_cond_gemv(d_condW, d_condB, g, B, Ncond, cfg.gin_channels, F(W.cond), stream));

  cx.skip = !(phases & PH_MID);
  if (what & RUN_FLOW) {
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
    if (z_out) AUX_LAUNCH(T_UNPACK, launch_unpack_channel_major(F(W.zf), B, IC, T, z_out, stream));
  }

  if (what & RUN_DEC) {
    // x = conv_pre(z); the next op's leaky_relu(0.1) is folded into the operand copy   models.py:307-309
    { Epilogue ep = epi_base(); set_out_op_split(ep, P(W.nxt_op), conv_pre.N, T, 0.1f, split16);
      cx.conv(T_CONV_PRE, conv_pre, P(W.zop), T, 1, cx.len_dec, ep); }
    int L = T;              // rows per batch item at the current rate
    int mul = 1;            // valid rows = len * mul
    const int ns = (int)stages.size();
    // conv_post forms the last stage's MRF average from the branches' y streams itself (decided up front: the
    // decoder loop below is skipped in the PH_BACK pass of a graph replay)
    const bool post_from_y = rb_fused && mrf_separate && ns > 0 && stages.back().rb.size() <= 3 &&
                             conv_post_y_supported(post_C, post_k, T * up_total);
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
      for (int j = 0; j < nk; ++j) {
        const int ni = (int)st.rb[j].size();
        for (int d = 0; d < ni; ++d) {
          const ResIter& it = st.rb[j][d];
          if (rb_fused) {
            // whole iteration in one kernel; the residual stream is y = op(leaky_relu(x)) only (rb_fused.cu)
            if (cx.skip) continue;
            ++cx.launches;
            if (dry || !cx.ok()) continue;
            RbLaunch rl{};
            void* ping = (d & 1) ? P(W.xb2_op) : P(W.xb_op);        // out of iteration d
            void* pong = (d & 1) ? P(W.xb_op) : P(W.xb2_op);        // out of iteration d-1
            rl.y_in = d == 0 ? P(W.xin_op) : pong;
            rl.w1 = reinterpret_cast<const char*>(d_w) + it.c1.w_off * ob;
            rl.w2 = reinterpret_cast<const char*>(d_w) + it.c2.w_off * ob;
            rl.C = C; rl.k = it.c1.k; rl.dil = it.c1.dil; rl.B = B;
            RbArgs& ra = rl.a;
            ra.L = L; ra.B = B; ra.len = cx.len_dec; ra.len_mul = mul;
            ra.bias1 = cx.cond + it.c1.cond_off; ra.bias1_bstride = Ncond;
            ra.bias2 = d_bias + it.c2.bias_off;
            ra.scale = 1.f;
            // MRF (models.py:311-314).  With <= 3 branches the first branches' results stay 16-bit y streams (their
            // last iteration is an ordinary mid-type launch into br[j]) and the final branch sums them in its epilogue:
            // 12 B of traffic per element instead of 20 B through an fp32 accumulator.  br[] reuses the fp32
            // buffers of the layer-at-a-time plan, which the fused plan does not touch.
            void* br[3] = {P(W.xin_f), P(W.xb_f), P(W.acc_f)};
            const bool stream_mrf = nk <= 3;
            // Separate MRF pass (default): EVERY branch ends in an ordinary mid-type launch into br[j] and one coalesced
            // pass (mrf_sum_kernel, or conv_post's tile load on the last stage) averages the streams.  The summing
            // epilogue variant costs 40-100 us more per stage than a mid launch (one exposed DRAM round trip per
            // 16-column chunk in a row-per-thread epilogue); the separate pass is HBM-bound at 10 B per element.
            const bool sep_mrf = stream_mrf && mrf_separate && (s < ns - 1 || post_from_y);
            if (d < ni - 1) {
              ra.y_out = ping;
            } else if (sep_mrf) {
              ra.y_out = br[j];
            } else if (stream_mrf && j < nk - 1) {
              ra.y_out = br[j];
            } else if (stream_mrf) {
              ra.scale = inv;
              ra.acc_mode = RB_EMIT_NEXT;
              ra.add0 = nk > 1 ? br[0] : nullptr;
              ra.add1 = nk > 2 ? br[1] : nullptr;
              if (s == ns - 1) ra.nxt_f32 = F(W.acc_f);              // conv_post reads fp32
              else { ra.nxt_op = P(W.nxt_op); ra.nxt_split = split16 ? 1 : 0; }
            } else {
              ra.scale = inv;
              ra.acc = F(W.acc_f);
              ra.acc_mode = (j > 0 ? RB_ACC_READ : 0) | (j < nk - 1 ? RB_ACC_STORE : RB_EMIT_NEXT);
              if (j == nk - 1) {
                if (s == ns - 1) ra.nxt_f32 = F(W.acc_f);            // conv_post reads fp32
                else { ra.nxt_op = P(W.nxt_op); ra.nxt_split = split16 ? 1 : 0; }
              }
            }
            cx.ev_begin(T_RB_FUSED + s, (it.c1.flop_per_pos + it.c2.flop_per_pos) * (cx.len_dec ? cx.valid_frames : (double)B * T) * mul);
            cx.err = launch_rb_fused(rl, op, stream);
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
            else set_out_op_split(ep, P(W.nxt_op), C, L, next_slope, split16);
          }
          cx.conv(T_RB_C2 + s, it.c2, P(W.g_op), L, mul, cx.len_dec, ep);
        }
      }
      if (rb_fused && mrf_separate && nk <= 3 && s < ns - 1) {
        // x = sum_j resblock_j(x) / num_kernels; next operand = leaky_relu(x, 0.1)      models.py:309-314
        AUX_LAUNCH(T_MRF, launch_mrf_sum(P(W.xin_f), nk > 1 ? P(W.xb_f) : nullptr, nk > 2 ? P(W.acc_f) : nullptr, B, L, C, cx.len_dec, mul,
                                         inv, next_slope, P(W.nxt_op), split16 ? 1 : 0, op, stream));
      }
    }
    // tanh(conv_post(leaky_relu(x, 0.01)))                                 models.py:315-317
    cx.skip = !(phases & PH_BACK);
    if (post_from_y) {
      const int nk = (int)stages.back().rb.size();
      AUX_LAUNCH(T_POST, launch_conv_post_y(P(W.xin_f), nk > 1 ? P(W.xb_f) : nullptr, nk > 2 ? P(W.acc_f) : nullptr, 1.0f / nk, B, L, post_C,
                                            post_k, 0.01f, d_post, cx.len_dec, mul, wav, pcm_out, pcm_volume, op, stream));
    } else {
      AUX_LAUNCH(T_POST, launch_conv_post(F(W.acc_f), B, L, post_C, post_k, 0.01f, d_post, cx.len_dec, mul, wav, pcm_out, pcm_volume, stream));
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

static bool graphs_enabled() {
  static int v = -1;
  if (v < 0) { const char* e = getenv("VITSB_GRAPH"); v = (e && e[0] == '0') ? 0 : 1; }
  return v == 1;
}

// Public entry: issue the plan directly, or - from the second identical call on - replay its middle part
// (every launch that touches only the workspace) as an instantiated CUDA graph.  The seam kernels that read or
// write caller pointers stay outside the graph, so fresh input / output buffers do not invalidate it.
int Engine::run(int what, const float* z_in, const float* g, const int32_t* len, int B, int T, int ragged,
                float* z_out, float* wav, void* ws, size_t ws_bytes, cudaStream_t stream, bool dry, int* launches) {
  if (dry || profile || z_out || !graphs_enabled() || B < 1 || B > 4096)
    return issue(PH_ALL, what, z_in, g, len, B, T, ragged, z_out, wav, ws, ws_bytes, stream, dry, launches);

  std::vector<long long> key = {what, (long long)(uintptr_t)ws, (long long)ws_bytes, B, T, ragged, len ? 1 : 0};
  if (len) key.insert(key.end(), len, len + B);
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
    return issue(PH_ALL, what, z_in, g, len, B, T, ragged, z_out, wav, ws, ws_bytes, stream, dry, launches);
  }
  ge->stamp = ++graph_clock;
  int n_front = 0, n_back = 0;
  int rc = issue(PH_FRONT, what, z_in, g, len, B, T, ragged, z_out, wav, ws, ws_bytes, stream, false, &n_front);
  if (rc != VITSB_OK) return rc;
  if (!ge->exec && ge->seen >= 0) {              // second sighting: capture the middle on a private stream
    if (!cap_stream && cudaStreamCreateWithFlags(&cap_stream, cudaStreamNonBlocking) != cudaSuccess) cap_stream = nullptr;
    cudaGraph_t graph = nullptr;
    bool ok = cap_stream && cudaStreamBeginCapture(cap_stream, cudaStreamCaptureModeRelaxed) == cudaSuccess;
    if (ok) {
      int n_mid = 0;
      const int rc_mid = issue(PH_MID, what, z_in, g, len, B, T, ragged, z_out, wav, ws, ws_bytes, cap_stream, false, &n_mid);
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
    const cudaError_t le = cudaGraphLaunch(ge->exec, stream);
    if (le != cudaSuccess) return fail(VITSB_ERR_CUDA, std::string("graph launch failed: ") + cudaGetErrorString(le));
  } else {
    int n_mid = 0;
    rc = issue(PH_MID, what, z_in, g, len, B, T, ragged, z_out, wav, ws, ws_bytes, stream, false, &n_mid);
    if (rc != VITSB_OK) return rc;
    ge->launches = n_mid;
  }
  rc = issue(PH_BACK, what, z_in, g, len, B, T, ragged, z_out, wav, ws, ws_bytes, stream, false, &n_back);
  if (launches) *launches = n_front + ge->launches + n_back;
  return rc;
}

}  // namespace vitsb
