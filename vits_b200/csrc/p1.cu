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
  // pass 1: scores.  lane = 4 consecutive dims (kc / 4 lanes active; kc <= 128), four keys per step: every K row is one
  // coalesced read, the partial dots of the four keys reduce through independent shuffle trees (ILP).
  const int d0 = lane * 4;
  const bool dl_ok = d0 < kc;
  float4 q4 = dl_ok ? *reinterpret_cast<const float4*>(q + d0) : make_float4(0.f, 0.f, 0.f, 0.f);
  q4.x *= qscale; q4.y *= qscale; q4.z *= qscale; q4.w *= qscale;      // query / sqrt(k_channels) first, as attentions.py:92 does
  float mx = -3.4e38f;
  for (int j0 = 0; j0 < Tk; j0 += 4) {
    float p[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int j = j0 + u < Tk ? j0 + u : Tk - 1;
      const float* k = base + (long long)j * 3 * H + H + h * kc;
      const float4 k4 = dl_ok ? *reinterpret_cast<const float4*>(k + d0) : make_float4(0.f, 0.f, 0.f, 0.f);
      p[u] = fmaf(q4.x, k4.x, fmaf(q4.y, k4.y, fmaf(q4.z, k4.z, q4.w * k4.w)));
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) {
#pragma unroll
      for (int u = 0; u < 4; ++u) p[u] += __shfl_xor_sync(0xffffffffu, p[u], d);
    }
    if (lane < 4 && j0 + lane < Tk) sc[j0 + lane] = lane == 0 ? p[0] : (lane == 1 ? p[1] : (lane == 2 ? p[2] : p[3]));
#pragma unroll
    for (int u = 0; u < 4; ++u) if (j0 + u < Tk) mx = fmaxf(mx, p[u]);
  }
  __syncwarp();
  float sum = 0.f;
  for (int j = lane; j < Tk; j += 32) { const float e = expf(sc[j] - mx); sc[j] = e; sum += e; }
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
  __syncwarp();
  const float inv = 1.f / sum;
  // pass 2: out[d] = sum_j p_j v[j][d]   (coalesced rows of v, four independent accumulators)
  if (dl_ok) {
    float4 acc[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
    const float* v = base + 2 * H + h * kc + d0;
    for (int j0 = 0; j0 < Tk; j0 += 4) {
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        if (j0 + u < Tk) {
          const float p = sc[j0 + u];
          const float4 v4 = *reinterpret_cast<const float4*>(v + (long long)(j0 + u) * 3 * H);
          acc[u].x = fmaf(p, v4.x, acc[u].x); acc[u].y = fmaf(p, v4.y, acc[u].y); acc[u].z = fmaf(p, v4.z, acc[u].z); acc[u].w = fmaf(p, v4.w, acc[u].w);
        }
      }
    }
    float4 o4;
    o4.x = ((acc[0].x + acc[1].x) + (acc[2].x + acc[3].x)) * inv; o4.y = ((acc[0].y + acc[1].y) + (acc[2].y + acc[3].y)) * inv;
    o4.z = ((acc[0].z + acc[1].z) + (acc[2].z + acc[3].z)) * inv; o4.w = ((acc[0].w + acc[1].w) + (acc[2].w + acc[3].w)) * inv;
    *reinterpret_cast<float4*>(orow + d0) = o4;
  }
}

// Same attention for short sequences: K and V of one (b, head) staged ONCE in shared memory (coalesced), 32 queries per
// block (4 per warp).  The global-memory form above is latency bound (one dependent round trip per four keys).
constexpr int kAttQS = 8;            // one query per warp: at ~100 phonemes this gives 2 x 11 blocks instead of 2 x 3
__global__ void __launch_bounds__(256)
p1_attention_smem_kernel(const float* __restrict__ qkv, int H, int kc, int T, const int* __restrict__ len, float qscale,
                         float* __restrict__ out) {
  extern __shared__ float s_att[];                           // K [Tk][kc + 4] | V [Tk][kc] | scores [8][T] | q [8][kc] (pre-scaled)
  const int b = blockIdx.z, h = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int Tk = len ? len[b] : T;
  const int ks = kc + 4;                                     // padded K row: conflict-free float4 reads with lane = key
  float* sK = s_att;
  float* sV = sK + (size_t)T * ks;
  float* sc = sV + (size_t)T * kc + warp * T;
  float* sQ = sV + (size_t)T * kc + 8 * (size_t)T + warp * kc;
  const float* base = qkv + (long long)b * T * 3 * H;
  const int c4 = kc >> 2;
  for (int i = threadIdx.x; i < Tk * c4; i += 256) {
    const int j = i / c4, d = (i - j * c4) * 4;
    const float* row = base + (long long)j * 3 * H + h * kc + d;
    *reinterpret_cast<float4*>(sK + (size_t)j * ks + d) = *reinterpret_cast<const float4*>(row + H);
    *reinterpret_cast<float4*>(sV + (size_t)j * kc + d) = *reinterpret_cast<const float4*>(row + 2 * H);
  }
  __syncthreads();
  for (int qi = 0; qi < kAttQS / 8; ++qi) {
    const int tq = blockIdx.x * kAttQS + qi * 8 + warp;
    if (tq >= T) break;
    float* orow = out + ((long long)b * T + tq) * H + h * kc;
    if (tq >= Tk) {
      for (int d = lane; d < kc; d += 32) orow[d] = 0.f;
      continue;
    }
    const float* qg = base + (long long)tq * 3 * H + h * kc;
    for (int d = lane; d < kc; d += 32) sQ[d] = qg[d] * qscale;          // query / sqrt(k_channels) first (attentions.py:92)
    __syncwarp();
    const float* q = sQ;
    // pass 1: lane = key, the (scaled) query row is broadcast from shared memory
    float mx = -3.4e38f;
    for (int j = lane; j < Tk; j += 32) {
      const float* k = sK + (size_t)j * ks;
      float a0 = 0.f, a1 = 0.f;
#pragma unroll 4
      for (int d = 0; d < kc; d += 8) {
        const float4 q0 = *reinterpret_cast<const float4*>(q + d), q1 = *reinterpret_cast<const float4*>(q + d + 4);
        const float4 k0 = *reinterpret_cast<const float4*>(k + d), k1 = *reinterpret_cast<const float4*>(k + d + 4);
        a0 = fmaf(q0.x, k0.x, a0); a0 = fmaf(q0.y, k0.y, a0); a0 = fmaf(q0.z, k0.z, a0); a0 = fmaf(q0.w, k0.w, a0);
        a1 = fmaf(q1.x, k1.x, a1); a1 = fmaf(q1.y, k1.y, a1); a1 = fmaf(q1.z, k1.z, a1); a1 = fmaf(q1.w, k1.w, a1);
      }
      const float a = a0 + a1;
      sc[j] = a;
      mx = fmaxf(mx, a);
    }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, d));
    float sum = 0.f;
    for (int j = lane; j < Tk; j += 32) { const float e = expf(sc[j] - mx); sc[j] = e; sum += e; }
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
    __syncwarp();
    const float inv = 1.f / sum;
    // pass 2: lane = 4 consecutive dims
    const int d0 = lane * 4;
    if (d0 < kc) {
      float4 acc[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) acc[u] = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int j0 = 0; j0 < Tk; j0 += 4) {
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (j0 + u < Tk) {
            const float p = sc[j0 + u];
            const float4 v4 = *reinterpret_cast<const float4*>(sV + (size_t)(j0 + u) * kc + d0);
            acc[u].x = fmaf(p, v4.x, acc[u].x); acc[u].y = fmaf(p, v4.y, acc[u].y); acc[u].z = fmaf(p, v4.z, acc[u].z); acc[u].w = fmaf(p, v4.w, acc[u].w);
          }
        }
      }
      *reinterpret_cast<float4*>(orow + d0) = make_float4(((acc[0].x + acc[1].x) + (acc[2].x + acc[3].x)) * inv, ((acc[0].y + acc[1].y) + (acc[2].y + acc[3].y)) * inv,
                                                         ((acc[0].z + acc[1].z) + (acc[2].z + acc[3].z)) * inv, ((acc[0].w + acc[1].w) + (acc[2].w + acc[3].w)) * inv);
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------
// Skinny fp32 conv / GEMM for the text encoder's shapes (tens of rows, K up to a few thousand): K ACROSS THE LANES of a
// warp.  A warp owns an 8-row x 4-column output block; lane l multiplies channels c = 4 l + 128 i of every tap (all
// global loads are coalesced 16-byte reads, no shared memory, no barriers, dozens of independent loads in flight per
// lane), and a 31-shuffle transpose-reduce leaves output (row l / 4, column l % 4) in lane l.  The tiled kernel of
// conv_f32.cu is built for thousands of rows: here it runs 24-96 CTAs with a barrier per 16-wide K slice.
//   out[b][t][n] = epi( sum_{j, c} w[n][j * Cin + c] * in[b][t + j - pad][c] + bias[b][n] )
// epilogue: + res (residual) | relu | gate (columns n and n + N/2 = tanh / sigmoid halves, attentions.py:162-164);
// rows t >= len[b] are stored as zeros.
// ---------------------------------------------------------------------------
struct SkinnyArgs {
  const float* in; int Cin; int T; int B;
  const float* w; int N; int k; int pad;
  const float* bias; long long bias_bstride;
  const float* res;                 // [B][T][N_out] or null
  float* out;                       // [B][T][N_out]
  const int* len;
  int gate, relu;
};

__global__ void __launch_bounds__(256)
p1_skinny_conv_kernel(const SkinnyArgs a) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int n_out = a.gate ? a.N / 2 : a.N;
  const int cgroups = a.gate ? n_out / 2 : a.N / 4;          // column groups: 4 linear columns, or 2 gate channels (tanh, sigmoid of each)
  const int cg = blockIdx.x * 8 + warp;
  if (cg >= cgroups) return;
  const int b = blockIdx.z, t0 = blockIdx.y * 8;
  const int valid = a.len ? a.len[b] : a.T;
  int wr[4];                                                 // weight rows of this warp's 4 columns
  if (a.gate) { wr[0] = cg * 2; wr[1] = cg * 2 + 1; wr[2] = n_out + cg * 2; wr[3] = n_out + cg * 2 + 1; }
  else { wr[0] = cg * 4; wr[1] = cg * 4 + 1; wr[2] = cg * 4 + 2; wr[3] = cg * 4 + 3; }
  float acc[32];
#pragma unroll
  for (int i = 0; i < 32; ++i) acc[i] = 0.f;
  const int Ktot = a.k * a.Cin;
  const float* in = a.in + (long long)b * a.T * a.Cin;
  if (t0 < valid) {
    for (int j = 0; j < a.k; ++j) {
      const int tr0 = t0 + j - a.pad;
      for (int c = lane * 4; c < a.Cin; c += 128) {
        float4 wv[4], av[8];
#pragma unroll
        for (int n = 0; n < 4; ++n) wv[n] = __ldg(reinterpret_cast<const float4*>(a.w + (long long)wr[n] * Ktot + j * a.Cin + c));
#pragma unroll
        for (int r = 0; r < 8; ++r) {
          const int tr = tr0 + r;
          av[r] = (tr >= 0 && tr < a.T) ? *reinterpret_cast<const float4*>(in + (long long)tr * a.Cin + c) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
#pragma unroll
        for (int r = 0; r < 8; ++r)
#pragma unroll
          for (int n = 0; n < 4; ++n) {
            float s = acc[r * 4 + n];
            s = fmaf(av[r].x, wv[n].x, s); s = fmaf(av[r].y, wv[n].y, s); s = fmaf(av[r].z, wv[n].z, s); s = fmaf(av[r].w, wv[n].w, s);
            acc[r * 4 + n] = s;
          }
      }
    }
  }
  // transpose-reduce: after the step with offset o each lane keeps the half of its values whose index bit matches its
  // lane bit; 16 + 8 + 4 + 2 + 1 shuffles, and lane l ends with the full sum of element l
#pragma unroll
  for (int o = 16; o >= 1; o >>= 1) {
    const bool up = (lane & o) != 0;
#pragma unroll
    for (int i = 0; i < o; ++i) {
      const float send = up ? acc[i] : acc[i + o];
      const float keep = up ? acc[i + o] : acc[i];
      acc[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
    }
  }
  const int r = lane >> 2, n = lane & 3, t = t0 + r;
  const bool live = t < valid;
  const float* bias = a.bias + (long long)b * a.bias_bstride;
  float v = acc[0] + bias[wr[n]];
  // gate: lanes n = 0, 1 hold the tanh pre-activations of gate channels 2 cg, 2 cg + 1; lanes n = 2, 3 their sigmoid halves
  const float other = __shfl_xor_sync(0xffffffffu, v, 2);
  if (t >= a.T) return;
  if (a.gate) {
    if (n < 2) {
      const float g = live ? gate_fn<true>(v, other) : 0.f;
      a.out[((long long)b * a.T + t) * n_out + cg * 2 + n] = g;
    }
    return;
  }
  const long long o = ((long long)b * a.T + t) * a.N + wr[n];
  if (a.res) v += a.res[o];
  if (a.relu) v = fmaxf(v, 0.f);
  a.out[o] = live ? v : 0.f;
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
  // CUDA-graph replay of the workspace-only middle of a call (48 of its 55 launches), keyed by what the captured launches
  // depend on: workspace, shape, masked or not.  Same scheme as the synthesis path (engine_run.cu).
  struct Graph { const void* ws; int B, T, masked; int seen; cudaGraphExec_t exec; };
  std::vector<Graph> graphs;
};

void p1_destroy(P1State* s) {
  if (!s) return;
  for (P1State::Graph& g : s->graphs) if (g.exec) cudaGraphExecDestroy(g.exec);
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
  if (S->n_layers < 1 || S->IC != e.cfg.inter_channels || gin != e.cfg.gin_channels || H % S->n_heads || (H / S->n_heads) % 4 || (H / S->n_heads) > 128 ||
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
  P1State& S = *e.p1;
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
  cudaStream_t st = stream;                             // the stream the launch helpers below issue on (the capture stream while capturing)

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
    if (rows <= 4096 && c.Cin % 4 == 0 && c.N % 4 == 0 && (!gate || c.N % 8 == 0)) {
      // the encoder's regime: few rows, long K - K across the lanes of a warp (p1_skinny_conv_kernel)
      SkinnyArgs sa{};
      sa.in = in; sa.Cin = c.Cin; sa.T = T; sa.B = B; sa.w = P + c.w; sa.N = c.N; sa.k = c.k; sa.pad = (c.k - 1) / 2;
      sa.bias = ep.bias; sa.bias_bstride = ep.bias_bstride; sa.res = res; sa.out = out; sa.len = dl; sa.gate = gate; sa.relu = relu ? 1 : 0;
      const int cgroups = gate ? c.N / 4 : c.N / 4;
      dim3 grid((cgroups + 7) / 8, (T + 7) / 8, B);
      p1_skinny_conv_kernel<<<grid, 256, 0, st>>>(sa);
      chk();
      return;
    }
    if (res) { ep.res = res; ep.res_ld = No; ep.res_bstride = (long long)T * No; }
    if (relu) { ep.out_op = out; ep.out_op_ld = No; ep.out_op_bstride = (long long)T * No; ep.out_slope = 0.f; }
    else { ep.out_f32 = out; ep.out_f32_ld = No; ep.out_f32_bstride = (long long)T * No; }
    ce = launch_conv_f32(a, st);
  };
  auto ln = [&](const float* in, float* out, int C, size_t gamma, size_t beta, const float* add, long long add_bs, bool pos) {
    if (!ok()) return;
    p1_layernorm_kernel<<<(rows + 7) / 8, 256, 0, st>>>(in, C, T, dl, P + gamma, P + beta, add, add_bs, std::sqrt((float)H),
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
  const int kc = H / S.n_heads;
  // ---- middle: workspace and weights only (captured / replayed as a CUDA graph from the second identical call on) ----
  auto mid = [&]() {
  ln(Fp(W.Y), Fp(W.X), H, S.emb_g, S.emb_b, Fp(W.emo), H, true);
  for (const P1State::Layer& L : S.layers) {
    conv(L.qkv, Fp(W.X), Fp(W.QKV), nullptr, 0, false);                                  // attentions.py:79-81
    if (ok()) {
      const size_t smem_s = sizeof(float) * ((size_t)T * (kc + 4) + (size_t)T * kc + 8 * (size_t)T + 8 * (size_t)kc);
      if (smem_s <= 200 * 1024) {                           // K, V of one (row, head) fit shared memory
        static PerDeviceOnce optin;
        ce = optin.run([] { return cudaFuncSetAttribute(p1_attention_smem_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024); });
        if (ok()) {
          dim3 grid((T + kAttQS - 1) / kAttQS, S.n_heads, B);
          p1_attention_smem_kernel<<<grid, 256, smem_s, st>>>(Fp(W.QKV), H, kc, T, dl, 1.0f / std::sqrt((float)kc), Fp(W.A));
          chk();
        }
      } else {
        dim3 grid((T + kAttQ - 1) / kAttQ, S.n_heads, B);
        p1_attention_kernel<<<grid, 32 * kAttQ, sizeof(float) * kAttQ * T, st>>>(Fp(W.QKV), H, kc, T, dl, 1.0f / std::sqrt((float)kc), Fp(W.A));
        chk();
      }
    }
    conv(L.o, Fp(W.A), Fp(W.Y), Fp(W.X), 0, false);                                       // x + conv_o(attn)        attentions.py:83, 50-51
    ln(Fp(W.Y), Fp(W.X), H, L.g1, L.b1, nullptr, 0, false);
    conv(L.c1, Fp(W.X), Fp(W.G), nullptr, 1, false);                                      // tanh * sigmoid gate     attentions.py:161-164
    conv(L.c2, Fp(W.G), Fp(W.Y), Fp(W.X), 0, false);                                      // x + conv_2(.)           attentions.py:165, 52-53
    ln(Fp(W.Y), Fp(W.X), H, L.g2, L.b2, nullptr, 0, false);
  }
  // stats = proj(x); m_p, s_p = exp(logs_p)                                             models.py:185-188, 563
  conv(S.proj, Fp(W.X), Fp(W.QKV), nullptr, 0, false);
  // duration predictor                                                                 models.py:59-67
  conv(S.dp_pre, Fp(W.X), Fp(W.Y), nullptr, 0, false);                                    // pre(x) + cond1(g)
  conv(S.dp_c1, Fp(W.Y), Fp(W.A), nullptr, 0, true);                                      // relu(conv_1)
  ln(Fp(W.A), Fp(W.Y), D, S.dn1g, S.dn1b, Fp(W.cond) + S.cond2_col, S.Ncond, false);      // norm_1(.) + cond2(g)
  conv(S.dp_c2, Fp(W.Y), Fp(W.A), nullptr, 0, true);
  ln(Fp(W.A), Fp(W.Y), D, S.dn2g, S.dn2b, nullptr, 0, false);
  };   // mid

  P1State::Graph* ge = nullptr;
  if (e.opt_graph) {
    for (P1State::Graph& g : S.graphs) if (g.ws == ws && g.B == B && g.T == T && g.masked == (dl != nullptr)) { ge = &g; break; }
    if (!ge) {
      if (S.graphs.size() >= 8) { if (S.graphs.front().exec) cudaGraphExecDestroy(S.graphs.front().exec); S.graphs.erase(S.graphs.begin()); }
      S.graphs.push_back(P1State::Graph{ws, B, T, dl != nullptr, 0, nullptr});      // first sighting: issue directly
    } else if (!ge->exec && ge->seen >= 0 && ok()) {                                  // second: capture on the private stream
      if (!e.cap_stream && cudaStreamCreateWithFlags(&e.cap_stream, cudaStreamNonBlocking) != cudaSuccess) e.cap_stream = nullptr;
      cudaGraph_t graph = nullptr;
      bool good = e.cap_stream && cudaStreamBeginCapture(e.cap_stream, cudaStreamCaptureModeRelaxed) == cudaSuccess;
      if (good) {
        st = e.cap_stream;
        mid();
        st = stream;
        const cudaError_t ee = cudaStreamEndCapture(e.cap_stream, &graph);
        good = ok() && ee == cudaSuccess && graph != nullptr;
        if (good) good = cudaGraphInstantiate(&ge->exec, graph, 0) == cudaSuccess;
        if (graph) cudaGraphDestroy(graph);
      }
      if (!good) { cudaGetLastError(); ce = cudaSuccess; ge->exec = nullptr; ge->seen = -1; }    // never retry this shape
    }
  }
  if (ge && ge->exec) { if (ok()) ce = cudaGraphLaunch(ge->exec, stream); }
  else mid();
  // ---- back: the caller's outputs ----
  if (ok()) {
    dim3 grid((T + 31) / 32, (2 * S.IC + 31) / 32, B), block(32, 8);
    p1_stats_out_kernel<<<grid, block, 0, stream>>>(Fp(W.QKV), S.IC, T, dl, m_p, s_p);
    chk();
  }
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
