// infer_p1 on the library: text encoder + duration predictor + speaker embedding.
//
// Reference (paths relative to emotional-vits/): SynthesizerTrn.infer_p1 models.py:558-566 -> TextEncoder.infer
// models.py:180-189 -> attentions.Encoder.infer attentions.py:48-54 (MultiHeadAttention :78-102, FFN2 :160-166,
// modules.LayerNorm modules.py:33-44) and DurationPredictor.infer models.py:59-67.
//
// At serving sizes (one utterance, ~60 phonemes) this stage is 0.36 GFLOP against 85 MB of fp32 weights: it is bound
// by launch count and weight streaming, not by arithmetic (the reference issues ~500 tiny kernels).  The durations
// feed ceil(exp(logw)) (infer.py:169-171), which amplifies any error into a different output LENGTH, so the stage
// computes in fp32 on the CUDA cores (conv_f32_kernel: FFMA implicit GEMM with the library's fused epilogue - bias /
// conditioning, residual, tanh*sigmoid gate, relu) plus three small kernels below: layer norm with fused post-ops,
// the two-head attention, and the output finish.  Activations are time-major fp32 [B][Tx][C]; `text` already is.
// Batches are ragged: row b equals the B = 1 call on its own x_len[b] phonemes (padding rows are zeros everywhere, which
// is the conv zero padding; attention keys beyond x_len[b] are excluded).
#include <cmath>
#include <cstring>
#include <map>
#include <string>
#include <vector>

#include "engine.h"

namespace vitsb {

// ---------------------------------------------------------------------------
// kernels
// ---------------------------------------------------------------------------
// out[b][t][:] = ((LN(in[b][t][:]) * gamma + beta) + add[b][:]) * scale + pe[t][:] * pe_alpha[0]; rows t >= len[b] -> 0.
// One warp per row; C <= 1024, C % 4 == 0.  Two-pass mean / variance in fp32 like ATen's layer_norm.
__global__ void __launch_bounds__(256)
p1_layernorm_kernel(const float* __restrict__ in, int C, int T, const int* __restrict__ len, const float* __restrict__ gamma,
                    const float* __restrict__ beta, const float* __restrict__ add, long long add_bstride, float scale,
                    const float* __restrict__ pe, const float* __restrict__ pe_alpha, float eps, float* __restrict__ out, int rows) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= rows) return;
  const int b = row / T, t = row - b * T;
  float* o = out + (long long)row * C;
  const bool live = !len || t < len[b];
  if (!live) {
    for (int c = lane * 4; c < C; c += 128) *reinterpret_cast<float4*>(o + c) = make_float4(0.f, 0.f, 0.f, 0.f);
    return;
  }
  const float* x = in + (long long)row * C;
  float4 v[8];
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int c = lane * 4 + i * 128;
    v[i] = c < C ? *reinterpret_cast<const float4*>(x + c) : make_float4(0.f, 0.f, 0.f, 0.f);
    s += (v[i].x + v[i].y) + (v[i].z + v[i].w);
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
  const float mean = s / C;
  float q = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    if (lane * 4 + i * 128 >= C) continue;
    const float a0 = v[i].x - mean, a1 = v[i].y - mean, a2 = v[i].z - mean, a3 = v[i].w - mean;
    q += (a0 * a0 + a1 * a1) + (a2 * a2 + a3 * a3);
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) q += __shfl_xor_sync(0xffffffffu, q, d);
  const float rstd = rsqrtf(q / C + eps);
  const float pa = pe ? pe_alpha[0] : 0.f;
#pragma unroll
  for (int n = 0; n < 8; ++n) {
    const int c = lane * 4 + n * 128;
    if (c >= C) continue;
    const float4 g4 = *reinterpret_cast<const float4*>(gamma + c), b4 = *reinterpret_cast<const float4*>(beta + c);
    float4 r;
    r.x = (v[n].x - mean) * rstd * g4.x + b4.x; r.y = (v[n].y - mean) * rstd * g4.y + b4.y;
    r.z = (v[n].z - mean) * rstd * g4.z + b4.z; r.w = (v[n].w - mean) * rstd * g4.w + b4.w;
    if (add) {
      const float4 a4 = *reinterpret_cast<const float4*>(add + (long long)b * add_bstride + c);
      r.x += a4.x; r.y += a4.y; r.z += a4.z; r.w += a4.w;
    }
    if (pe) {
      const float4 p4 = *reinterpret_cast<const float4*>(pe + (long long)t * C + c);
      r.x = r.x * scale + p4.x * pa; r.y = r.y * scale + p4.y * pa; r.z = r.z * scale + p4.z * pa; r.w = r.w * scale + p4.w * pa;
    }
    *reinterpret_cast<float4*>(o + c) = r;
  }
}

// Two-head (n_heads general) self-attention over qkv [B][T][3H] (q | k | v, head h at columns h * kc): one warp per
// (b, head, query); scores staged in shared memory; keys beyond len[b] are excluded (attention.py:92-97 masks them
// with -1e4, whose softmax weight underflows to exactly 0 in fp32).  out [B][T][H].
constexpr int kAttQ = 8;           // queries (warps) per block
__global__ void __launch_bounds__(32 * kAttQ)
p1_attention_kernel(const float* __restrict__ qkv, int H, int kc, int T, const int* __restrict__ len, float qscale,
                    float* __restrict__ out) {
  extern __shared__ float s_scores[];                       // [kAttQ][T]
  const int b = blockIdx.z, h = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tq = blockIdx.x * kAttQ + warp;
  const int Tk = len ? len[b] : T;
  if (tq >= T) return;
  float* orow = out + ((long long)b * T + tq) * H + h * kc;
  if (tq >= Tk) {
    for (int d = lane; d < kc; d += 32) orow[d] = 0.f;
    return;
  }
  const float* base = qkv + (long long)b * T * 3 * H;
  const float* q = base + (long long)tq * 3 * H + h * kc;
  float* sc = s_scores + warp * T;
  // pass 1: scores (lane = key), q / sqrt(kc) first as the reference does
  float mx = -3.4e38f;
  for (int j = lane; j < Tk; j += 32) {
    const float* k = base + (long long)j * 3 * H + H + h * kc;
    float acc = 0.f;
    for (int d = 0; d < kc; d += 4) {
      const float4 q4 = *reinterpret_cast<const float4*>(q + d), k4 = *reinterpret_cast<const float4*>(k + d);
      acc = fmaf(q4.x * qscale, k4.x, acc); acc = fmaf(q4.y * qscale, k4.y, acc);
      acc = fmaf(q4.z * qscale, k4.z, acc); acc = fmaf(q4.w * qscale, k4.w, acc);
    }
    sc[j] = acc;
    mx = fmaxf(mx, acc);
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, d));
  float sum = 0.f;
  for (int j = lane; j < Tk; j += 32) { const float e = expf(sc[j] - mx); sc[j] = e; sum += e; }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
  __syncwarp();
  const float inv = 1.f / sum;
  // pass 2: out[d] = sum_j p_j v[j][d]   (lane = 4 consecutive dims, coalesced rows of v)
  for (int d0 = lane * 4; d0 < kc; d0 += 128) {
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
    const float* v = base + 2 * H + h * kc + d0;
    for (int j = 0; j < Tk; ++j) {
      const float p = sc[j] * inv;
      const float4 v4 = *reinterpret_cast<const float4*>(v + (long long)j * 3 * H);
      acc.x = fmaf(p, v4.x, acc.x); acc.y = fmaf(p, v4.y, acc.y); acc.z = fmaf(p, v4.z, acc.z); acc.w = fmaf(p, v4.w, acc.w);
    }
    *reinterpret_cast<float4*>(orow + d0) = acc;
  }
}

// stats [B][T][2*IC] time-major -> m_p [B][IC][T], s_p = exp(logs) [B][IC][T] (reference layouts; models.py:186-188,563),
// zero beyond len[b]
__global__ void p1_stats_out_kernel(const float* __restrict__ stats, int IC, int T, const int* __restrict__ len,
                                    float* __restrict__ m_p, float* __restrict__ s_p) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z, t0 = blockIdx.x * 32, c0 = blockIdx.y * 32;      // c over 2*IC
  const int valid = len ? len[b] : T;
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int t = t0 + i, c = c0 + threadIdx.x;
    tile[i][threadIdx.x] = (t < T && c < 2 * IC) ? stats[((long long)b * T + t) * (2 * IC) + c] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int c = c0 + i, t = t0 + threadIdx.x;
    if (c >= 2 * IC || t >= T) continue;
    const float v = tile[threadIdx.x][i];
    const bool live = t < valid;
    if (c < IC) m_p[((long long)b * IC + c) * T + t] = live ? v : 0.f;
    else s_p[((long long)b * IC + (c - IC)) * T + t] = live ? expf(v) : 0.f;
  }
}

// logw[b][0][t] = dot(x[b][t][:], w) + bias   (dp.proj, models.py:66), zero beyond len[b]; one warp per row
__global__ void __launch_bounds__(256)
p1_rowdot_kernel(const float* __restrict__ x, int C, int T, const int* __restrict__ len, const float* __restrict__ w,
                 const float* __restrict__ bias, float* __restrict__ out, int rows) {
  const int row = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (row >= rows) return;
  const int b = row / T, t = row - b * T;
  float s = 0.f;
  for (int c = lane * 4; c < C; c += 128) {
    const float4 a = *reinterpret_cast<const float4*>(x + (long long)row * C + c), w4 = *reinterpret_cast<const float4*>(w + c);
    s = fmaf(a.x, w4.x, s); s = fmaf(a.y, w4.y, s); s = fmaf(a.z, w4.z, s); s = fmaf(a.w, w4.w, s);
  }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
  if (lane == 0) out[row] = (!len || t < len[b]) ? s + bias[0] : 0.f;
}

// g[b][:] = emb_g[sid[b]][:]   (models.py:561)
__global__ void p1_gather_kernel(const float* __restrict__ table, const int* __restrict__ sid, int gin, float* __restrict__ g) {
  const int b = blockIdx.x;
  const float* src = table + (long long)sid[b] * gin;
  for (int i = threadIdx.x * 4; i < gin; i += blockDim.x * 4)
    *reinterpret_cast<float4*>(g + (long long)b * gin + i) = *reinterpret_cast<const float4*>(src + i);
}

// ---------------------------------------------------------------------------
// packed weights
// ---------------------------------------------------------------------------
struct P1Conv { size_t w = 0, b = 0; int Cin = 0, N = 0, k = 1; int cond = -1; };   // offsets into the fp32 pool; cond >= 0: per-row bias column

struct P1State {
  int text_ch = 0, H = 0, F = 0, D = 0, IC = 0, gin = 0, emo_ch = 0, n_heads = 2, n_layers = 0, k = 5, kd = 5, n_spk = 0;
  int Ncond = 0, max_T = 1024;
  P1Conv emb, proj, dp_pre, dp_c1, dp_c2;
  struct Layer { P1Conv qkv, o, c1, c2; size_t g1, b1, g2, b2; };
  std::vector<Layer> layers;
  size_t emb_g = 0, emb_b = 0, alpha = 0, dn1g = 0, dn1b = 0, dn2g = 0, dn2b = 0, dproj_w = 0, dproj_b = 0, pe = 0;
  size_t condW = 0, condB = 0, emoW = 0, emoB = 0, spk = 0;
  int cond2_col = 0;
  float* d_pool = nullptr;
  size_t pool_elems = 0;
};

void p1_destroy(P1State* s) {
  if (!s) return;
  if (s->d_pool) cudaFree(s->d_pool);
  delete s;
}

namespace {

struct HT { const vitsb_tensor_desc* d = nullptr; };

float ld_elem(const vitsb_tensor_desc& d, size_t i) {
  if (d.dtype == VITSB_F32) return reinterpret_cast<const float*>(d.data)[i];
  if (d.dtype == VITSB_F16) return __half2float(reinterpret_cast<const __half*>(d.data)[i]);
  return __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(d.data)[i]);
}

}  // namespace

bool p1_present(const vitsb_tensor_desc* t, int n) {
  for (int i = 0; i < n; ++i)
    if (t[i].name && std::strncmp(t[i].name, "enc_p.", 6) == 0) return true;
  return false;
}

// Pack enc_p.* / dp.* / emb_g.* (reference layouts: Conv1d [Cout][Cin][k], Linear [out][in]) into one fp32 pool:
// convs as K-major [N][k*Cin] tap-major matrices, q | k | v stacked into one N = 3H matrix, every Linear that
// consumes g stacked into one [Ncond][gin] matrix (its conv's bias folded into the stacked bias).
int p1_load(Engine& e, const vitsb_tensor_desc* tens, int n) {
  std::map<std::string, const vitsb_tensor_desc*> T;
  for (int i = 0; i < n; ++i) {
    const vitsb_tensor_desc& d = tens[i];
    if (!d.name || !d.data) continue;
    if (std::strncmp(d.name, "enc_p.", 6) && std::strncmp(d.name, "dp.", 3) && std::strncmp(d.name, "emb_g.", 6)) continue;
    if (d.ndim < 0 || d.ndim > 4 || d.dtype < 0 || d.dtype > 2) return e.fail(VITSB_ERR_BAD_ARG, std::string("bad desc for ") + d.name);
    T[d.name] = &d;
  }
  std::string err;
  auto get = [&](const std::string& k, std::vector<int64_t> shape) -> const vitsb_tensor_desc* {
    auto it = T.find(k);
    if (it == T.end()) { if (err.empty()) err = "missing tensor " + k; return nullptr; }
    const vitsb_tensor_desc* d = it->second;
    bool ok = (size_t)d->ndim == shape.size();
    for (size_t i = 0; ok && i < shape.size(); ++i) ok = shape[i] < 0 || d->shape[i] == shape[i];
    if (!ok) { if (err.empty()) err = "shape mismatch for " + k; return nullptr; }
    return d;
  };
  auto dim = [&](const std::string& k, int i) -> int {
    auto it = T.find(k);
    return (it == T.end() || i >= it->second->ndim) ? 0 : (int)it->second->shape[i];
  };
  P1State* S = new P1State();
  std::vector<float> pool;
  auto take = [&](size_t cnt) { size_t o = pool.size(); pool.resize(o + ((cnt + 3) & ~(size_t)3), 0.f); return o; };
  auto vec = [&](const std::string& k, int cnt) -> size_t {
    const vitsb_tensor_desc* d = get(k, {cnt});
    size_t o = take(cnt);
    if (d) for (int i = 0; i < cnt; ++i) pool[o + i] = ld_elem(*d, i);
    return o;
  };
  // conv / linear weight -> [N][k][Cin] at row offset n0 of a matrix allocated by the caller
  auto put_conv = [&](const std::string& k, size_t dst, int N, int Cin, int kk) {
    const vitsb_tensor_desc* d = kk > 0 ? get(k, {N, Cin, kk}) : get(k, {N, Cin});
    const int K = kk > 0 ? kk : 1;
    if (!d) return;
    for (int nn = 0; nn < N; ++nn)
      for (int j = 0; j < K; ++j)
        for (int c = 0; c < Cin; ++c) pool[dst + ((size_t)nn * K + j) * Cin + c] = ld_elem(*d, ((size_t)nn * Cin + c) * K + j);
  };
  auto conv = [&](const std::string& base, int N, int Cin, int kk) -> P1Conv {
    P1Conv c; c.N = N; c.Cin = Cin; c.k = kk > 0 ? kk : 1;
    c.w = take((size_t)N * c.k * Cin);
    put_conv(base + ".weight", c.w, N, Cin, kk);
    c.b = vec(base + ".bias", N);
    return c;
  };

  S->text_ch = dim("enc_p.emb.0.weight", 1); S->H = dim("enc_p.emb.0.weight", 0);
  S->emo_ch = dim("enc_p.emo_proj.weight", 1);
  S->IC = dim("enc_p.proj.weight", 0) / 2;
  S->F = dim("enc_p.encoder.ffn_layers.0.conv_2.weight", 1); S->k = dim("enc_p.encoder.ffn_layers.0.conv_2.weight", 2);
  S->D = dim("dp.pre.weight", 0); S->kd = dim("dp.conv_1.weight", 2);
  S->gin = dim("emb_g.weight", 1); S->n_spk = dim("emb_g.weight", 0);
  S->n_heads = e.cfg.enc_n_heads > 0 ? e.cfg.enc_n_heads : 2;
  while (T.count("enc_p.encoder.attn_layers." + std::to_string(S->n_layers) + ".conv_q.weight")) ++S->n_layers;
  const int H = S->H, F = S->F, D = S->D, gin = S->gin;
  auto bad = [&](const char* m) { p1_destroy(S); return e.fail(VITSB_ERR_UNSUPPORTED, std::string("infer_p1 weights: ") + m); };
  if (H <= 0 || H > 1024 || H % 16 || S->text_ch <= 0 || S->text_ch % 16 || F <= 0 || F % 16 || D <= 0 || D > 1024 || D % 16)
    return bad("channel counts must be multiples of 16 (hidden / duration hidden <= 1024)");
  if (S->n_layers < 1 || S->IC != e.cfg.inter_channels || gin != e.cfg.gin_channels || H % S->n_heads || (H / S->n_heads) % 4 ||
      S->k % 2 != 1 || S->kd % 2 != 1 || S->emo_ch % 4 || gin % 4)
    return bad("shape set does not match the handle's config");

  S->emb = conv("enc_p.emb.0", H, S->text_ch, 0);
  S->emb_g = vec("enc_p.emb.1.weight", H); S->emb_b = vec("enc_p.emb.1.bias", H);
  { const vitsb_tensor_desc* d = get("enc_p.alpha", {}); S->alpha = take(4); if (d) pool[S->alpha] = ld_elem(*d, 0); }
  // conditioning stack: per layer FFN2.cond (2F rows, + conv_1.bias), then dp.cond1 (D rows, + dp.pre.bias), dp.cond2 (D)
  S->Ncond = S->n_layers * 2 * F + 2 * D;
  S->condW = take((size_t)S->Ncond * gin); S->condB = take(S->Ncond);
  int col = 0;
  for (int i = 0; i < S->n_layers; ++i) {
    const std::string a = "enc_p.encoder.attn_layers." + std::to_string(i) + ".", f = "enc_p.encoder.ffn_layers." + std::to_string(i) + ".";
    P1State::Layer L;
    L.qkv.N = 3 * H; L.qkv.Cin = H; L.qkv.k = 1;
    L.qkv.w = take((size_t)3 * H * H); L.qkv.b = take(3 * H);
    const char* nm[3] = {"conv_q", "conv_k", "conv_v"};
    for (int j = 0; j < 3; ++j) {
      put_conv(a + nm[j] + ".weight", L.qkv.w + (size_t)j * H * H, H, H, 1);
      const vitsb_tensor_desc* bd = get(a + nm[j] + ".bias", {H});
      if (bd) for (int c = 0; c < H; ++c) pool[L.qkv.b + (size_t)j * H + c] = ld_elem(*bd, c);
    }
    L.o = conv(a + "conv_o", H, H, 1);
    L.g1 = vec("enc_p.encoder.norm_layers_1." + std::to_string(i) + ".gamma", H);
    L.b1 = vec("enc_p.encoder.norm_layers_1." + std::to_string(i) + ".beta", H);
    L.c1 = conv(f + "conv_1", 2 * F, H, S->k);          // natural row order: [tanh F | sigmoid F] (torch.chunk, attentions.py:162)
    L.c1.cond = col;
    put_conv(f + "cond.weight", S->condW + (size_t)col * gin, 2 * F, gin, 0);
    { const vitsb_tensor_desc* cb = get(f + "cond.bias", {2 * F});
      if (cb) for (int c = 0; c < 2 * F; ++c) pool[S->condB + col + c] = ld_elem(*cb, c) + pool[L.c1.b + c]; }
    col += 2 * F;
    L.c2 = conv(f + "conv_2", H, F, S->k);
    L.g2 = vec("enc_p.encoder.norm_layers_2." + std::to_string(i) + ".gamma", H);
    L.b2 = vec("enc_p.encoder.norm_layers_2." + std::to_string(i) + ".beta", H);
    S->layers.push_back(L);
  }
  S->proj = conv("enc_p.proj", 2 * S->IC, H, 1);
  S->dp_pre = conv("dp.pre", D, H, 1);
  S->dp_pre.cond = col;
  put_conv("dp.cond1.weight", S->condW + (size_t)col * gin, D, gin, 0);
  { const vitsb_tensor_desc* cb = get("dp.cond1.bias", {D});
    if (cb) for (int c = 0; c < D; ++c) pool[S->condB + col + c] = ld_elem(*cb, c) + pool[S->dp_pre.b + c]; }
  col += D;
  S->cond2_col = col;
  put_conv("dp.cond2.weight", S->condW + (size_t)col * gin, D, gin, 0);
  { const vitsb_tensor_desc* cb = get("dp.cond2.bias", {D});
    if (cb) for (int c = 0; c < D; ++c) pool[S->condB + col + c] = ld_elem(*cb, c); }
  S->dp_c1 = conv("dp.conv_1", D, D, S->kd);
  S->dn1g = vec("dp.norm_1.gamma", D); S->dn1b = vec("dp.norm_1.beta", D);
  S->dp_c2 = conv("dp.conv_2", D, D, S->kd);
  S->dn2g = vec("dp.norm_2.gamma", D); S->dn2b = vec("dp.norm_2.beta", D);
  S->dproj_w = take(D); S->dproj_b = vec("dp.proj.bias", 1);
  { const vitsb_tensor_desc* d = get("dp.proj.weight", {1, D, 1}); if (d) for (int c = 0; c < D; ++c) pool[S->dproj_w + c] = ld_elem(*d, c); }
  S->emoW = take((size_t)H * S->emo_ch);
  put_conv("enc_p.emo_proj.weight", S->emoW, H, S->emo_ch, 0);
  S->emoB = vec("enc_p.emo_proj.bias", H);
  // sinusoid table (commons.py:176-190): pe[t][2i] = sin(t * w_i), pe[t][2i+1] = cos(t * w_i), w_i = exp(2i * -(ln 10000 / H)) in fp32
  S->pe = take((size_t)S->max_T * H);
  for (int i = 0; i < H; i += 2) {
    const float div = std::exp((float)i * -(float)(std::log(10000.0) / H));
    for (int t = 0; t < S->max_T; ++t) {
      const float arg = (float)t * div;
      pool[S->pe + (size_t)t * H + i] = (float)std::sin((double)arg);
      if (i + 1 < H) pool[S->pe + (size_t)t * H + i + 1] = (float)std::cos((double)arg);
    }
  }
  S->spk = take((size_t)S->n_spk * gin);
  { const vitsb_tensor_desc* d = get("emb_g.weight", {S->n_spk, gin});
    if (d) for (size_t i = 0; i < (size_t)S->n_spk * gin; ++i) pool[S->spk + i] = ld_elem(*d, i); }
  if (!err.empty()) { p1_destroy(S); return e.fail(VITSB_ERR_WEIGHTS, err); }

  cudaError_t ce = cudaSetDevice(e.device);
  if (ce == cudaSuccess) ce = cudaMalloc(&S->d_pool, pool.size() * sizeof(float));
  if (ce == cudaSuccess) ce = cudaMemcpy(S->d_pool, pool.data(), pool.size() * sizeof(float), cudaMemcpyHostToDevice);
  if (ce != cudaSuccess) { p1_destroy(S); return e.fail(VITSB_ERR_CUDA, std::string("infer_p1 weight upload: ") + cudaGetErrorString(ce)); }
  S->pool_elems = pool.size();
  p1_destroy(e.p1);
  e.p1 = S;
  return VITSB_OK;
}

// ---------------------------------------------------------------------------
// launch plan
// ---------------------------------------------------------------------------
namespace {

struct P1Ws { size_t len, sid, cond, emo, X, Y, QKV, A, G, total; };

size_t a256(size_t x) { return (x + 255) / 256 * 256; }

P1Ws p1_layout(const P1State& S, int B, int T) {
  P1Ws w{};
  size_t off = 0;
  auto take = [&](size_t bytes) { size_t o = off; off += a256(bytes); return o; };
  const size_t BT = (size_t)B * T;
  const size_t wide = (size_t)std::max(std::max(3 * S.H, S.F), std::max(2 * S.IC, S.D));
  w.len = take(4 * (size_t)B); w.sid = take(4 * (size_t)B);
  w.cond = take(4 * (size_t)B * S.Ncond); w.emo = take(4 * (size_t)B * S.H);
  w.X = take(4 * BT * std::max(S.H, S.D)); w.Y = take(4 * BT * std::max(S.H, S.D));
  w.QKV = take(4 * BT * wide);                 // q | k | v, later the projection's stats
  w.A = take(4 * BT * std::max(S.H, S.D));
  w.G = take(4 * BT * S.F);
  w.total = off;
  return w;
}

}  // namespace

size_t p1_workspace_bytes(const Engine& e, int B, int T) {
  if (!e.p1 || B < 1 || T < 1) return 0;
  return p1_layout(*e.p1, B, T).total;
}

int p1_run(Engine& e, const float* text, const float* emo, const int32_t* sid, const int32_t* x_len, int B, int T, float* m_p,
           float* s_p, float* logw, float* g_out, void* ws, size_t ws_bytes, cudaStream_t stream) {
  if (!e.p1) return e.fail(VITSB_ERR_WEIGHTS, "infer_p1 weights not loaded (the state_dict had no enc_p.* / dp.* / emb_g.* entries)");
  const P1State& S = *e.p1;
  if (!text || !emo || !sid || !m_p || !s_p || !logw || !g_out || !ws) return e.fail(VITSB_ERR_BAD_ARG, "null pointer");
  if (B < 1 || B > 65535 || T < 1) return e.fail(VITSB_ERR_BAD_ARG, "B and Tx_max must be >= 1");
  if (T > S.max_T) return e.fail(VITSB_ERR_BAD_ARG, "Tx_max too large (at most 1024 phonemes per row)");
  for (int b = 0; b < B; ++b) {
    if (sid[b] < 0 || sid[b] >= S.n_spk) return e.fail(VITSB_ERR_BAD_ARG, "sid out of range");       // infer.py:145
    if (x_len && (x_len[b] < 0 || x_len[b] > T)) return e.fail(VITSB_ERR_BAD_ARG, "x_len[b] out of [0, Tx_max]");
  }
  const P1Ws W = p1_layout(S, B, T);
  if (ws_bytes < W.total) return e.fail(VITSB_ERR_WORKSPACE, "workspace too small (vitsb_p1_workspace_bytes)");
  cudaError_t ce = cudaSetDevice(e.device);
  if (ce != cudaSuccess) return e.fail(VITSB_ERR_CUDA, cudaGetErrorString(ce));

  char* base = reinterpret_cast<char*>(ws);
  auto Fp = [&](size_t off) { return reinterpret_cast<float*>(base + off); };
  const float* P = S.d_pool;
  int* d_len = reinterpret_cast<int*>(base + W.len);
  int* d_sid = reinterpret_cast<int*>(base + W.sid);
  const int* dl = x_len ? d_len : nullptr;
  const int H = S.H, F = S.F, D = S.D, rows = B * T;
  if (x_len) ce = cudaMemcpyAsync(d_len, x_len, 4 * (size_t)B, cudaMemcpyHostToDevice, stream);
  if (ce == cudaSuccess) ce = cudaMemcpyAsync(d_sid, sid, 4 * (size_t)B, cudaMemcpyHostToDevice, stream);
  auto ok = [&]() { return ce == cudaSuccess; };
  auto chk = [&]() { if (ok()) ce = cudaGetLastError(); };

  // one fused conv launch: out = epilogue(W * in)
  auto conv = [&](const P1Conv& c, const float* in, float* out, const float* res, int gate, bool relu) {
    if (!ok()) return;
    ConvArgs a{};
    a.in = in; a.in_ld = c.Cin; a.in_cols = c.Cin; a.in_bstride = (long long)T * c.Cin; a.Cin = c.Cin;
    a.w = P + c.w; a.N = c.N; a.NT = c.N; a.gate_group = 0;
    a.k = c.k; a.dil = 1; a.pad = (c.k - 1) / 2; a.L = T; a.B = B; a.len = dl; a.len_mul = 1;
    Epilogue& ep = a.epi;
    ep.alpha = 1.f; ep.beta = 1.f; ep.out_slope = 1.f; ep.gate = gate;
    if (c.cond >= 0) { ep.bias = Fp(W.cond) + c.cond; ep.bias_bstride = S.Ncond; }
    else { ep.bias = P + c.b; ep.bias_bstride = 0; }
    const int No = gate ? c.N / 2 : c.N;
    if (res) { ep.res = res; ep.res_ld = No; ep.res_bstride = (long long)T * No; }
    if (relu) { ep.out_op = out; ep.out_op_ld = No; ep.out_op_bstride = (long long)T * No; ep.out_slope = 0.f; }
    else { ep.out_f32 = out; ep.out_f32_ld = No; ep.out_f32_bstride = (long long)T * No; }
    ce = launch_conv_f32(a, stream);
  };
  auto ln = [&](const float* in, float* out, int C, size_t gamma, size_t beta, const float* add, long long add_bs, bool pos) {
    if (!ok()) return;
    p1_layernorm_kernel<<<(rows + 7) / 8, 256, 0, stream>>>(in, C, T, dl, P + gamma, P + beta, add, add_bs, std::sqrt((float)H),
                                                           pos ? P + S.pe : nullptr, P + S.alpha, 1e-5f, out, rows);
    chk();
  };

  // g = emb_g(sid); every Linear of g in one pass; emo projection                      models.py:561, attentions.py:161, models.py:182
  p1_gather_kernel<<<B, 256, 0, stream>>>(P + S.spk, d_sid, S.gin, g_out);
  chk();
  if (ok()) ce = launch_cond_gemv(P + S.condW, P + S.condB, g_out, B, S.Ncond, S.gin, Fp(W.cond), stream);
  if (ok()) ce = launch_cond_gemv(P + S.emoW, P + S.emoB, emo, B, H, S.emo_ch, Fp(W.emo), stream);
  // x = (LayerNorm(Linear(text)) + emo_proj(emo)) * sqrt(H) + pe * alpha                models.py:181-183, 157-164
  conv(S.emb, text, Fp(W.Y), nullptr, 0, false);
  ln(Fp(W.Y), Fp(W.X), H, S.emb_g, S.emb_b, Fp(W.emo), H, true);
  const int kc = H / S.n_heads;
  for (const P1State::Layer& L : S.layers) {
    conv(L.qkv, Fp(W.X), Fp(W.QKV), nullptr, 0, false);                                  // attentions.py:79-81
    if (ok()) {
      dim3 grid((T + kAttQ - 1) / kAttQ, S.n_heads, B);
      p1_attention_kernel<<<grid, 32 * kAttQ, sizeof(float) * kAttQ * T, stream>>>(Fp(W.QKV), H, kc, T, dl, 1.0f / std::sqrt((float)kc), Fp(W.A));
      chk();
    }
    conv(L.o, Fp(W.A), Fp(W.Y), Fp(W.X), 0, false);                                       // x + conv_o(attn)        attentions.py:83, 50-51
    ln(Fp(W.Y), Fp(W.X), H, L.g1, L.b1, nullptr, 0, false);
    conv(L.c1, Fp(W.X), Fp(W.G), nullptr, 1, false);                                      // tanh * sigmoid gate     attentions.py:161-164
    conv(L.c2, Fp(W.G), Fp(W.Y), Fp(W.X), 0, false);                                      // x + conv_2(.)           attentions.py:165, 52-53
    ln(Fp(W.Y), Fp(W.X), H, L.g2, L.b2, nullptr, 0, false);
  }
  // stats = proj(x); m_p, s_p = exp(logs_p)                                             models.py:185-188, 563
  conv(S.proj, Fp(W.X), Fp(W.QKV), nullptr, 0, false);
  if (ok()) {
    dim3 grid((T + 31) / 32, (2 * S.IC + 31) / 32, B), block(32, 8);
    p1_stats_out_kernel<<<grid, block, 0, stream>>>(Fp(W.QKV), S.IC, T, dl, m_p, s_p);
    chk();
  }
  // duration predictor                                                                 models.py:59-67
  conv(S.dp_pre, Fp(W.X), Fp(W.Y), nullptr, 0, false);                                    // pre(x) + cond1(g)
  conv(S.dp_c1, Fp(W.Y), Fp(W.A), nullptr, 0, true);                                      // relu(conv_1)
  ln(Fp(W.A), Fp(W.Y), D, S.dn1g, S.dn1b, Fp(W.cond) + S.cond2_col, S.Ncond, false);      // norm_1(.) + cond2(g)
  conv(S.dp_c2, Fp(W.Y), Fp(W.A), nullptr, 0, true);
  ln(Fp(W.A), Fp(W.Y), D, S.dn2g, S.dn2b, nullptr, 0, false);
  if (ok()) {
    p1_rowdot_kernel<<<(rows + 7) / 8, 256, 0, stream>>>(Fp(W.Y), D, T, dl, P + S.dproj_w, P + S.dproj_b, logw, rows);
    chk();
  }
  if (!ok()) return e.fail(VITSB_ERR_CUDA, std::string("infer_p1: ") + cudaGetErrorString(ce));
  return VITSB_OK;
}

}  // namespace vitsb

extern "C" {

size_t vitsb_p1_workspace_bytes(const vitsb_handle* h, int B, int Tx_max) {
  return h ? vitsb::p1_workspace_bytes(h->eng, B, Tx_max) : 0;
}

int vitsb_infer_p1(vitsb_handle* h, const float* text, const float* emo, const int32_t* sid, const int32_t* x_len, int B, int Tx_max,
                   float* m_p, float* s_p, float* logw, float* g, void* ws, size_t ws_bytes, vitsb_stream stream) {
  if (!h) return VITSB_ERR_BAD_ARG;
  std::lock_guard<std::mutex> lk(h->eng.mu);
  return vitsb::p1_run(h->eng, text, emo, sid, x_len, B, Tx_max, m_p, s_p, logw, g, ws, ws_bytes, (cudaStream_t)stream);
}

}  // extern "C"
