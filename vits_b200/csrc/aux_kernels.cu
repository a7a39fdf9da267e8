// Memory-bound kernels around the implicit-GEMM convs: layout pack/unpack at
// the seam, the per-call conditioning GEMV, and conv_post+tanh (N=1, CUDA cores).
#include "common.h"
#include "kernels.h"
#include "tc_common.cuh"

namespace vitsb {

// ---------------------------------------------------------------------------
// [B][C][T] -> [B][T][C] (fp32 + operand copy), masked beyond len[b]
// ---------------------------------------------------------------------------
template <typename OpT>
__global__ void pack_time_major_kernel(const float* __restrict__ src, int C, int T, const int* __restrict__ len,
                                       float* __restrict__ dst_f32, OpT* __restrict__ dst_op, int split) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z, t0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int valid = len ? len[b] : T;
  const float* s = src + (long long)b * C * T;
  for (int i = threadIdx.y; i < 32; i += 8) {
    int c = c0 + i, t = t0 + threadIdx.x;
    tile[i][threadIdx.x] = (c < C && t < T && t < valid) ? s[(long long)c * T + t] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += 8) {
    int t = t0 + i, c = c0 + threadIdx.x;
    if (t < T && c < C) {
      float v = tile[threadIdx.x][i];
      long long o = ((long long)b * T + t) * C + c;
      if (dst_f32) dst_f32[o] = v;
      if (dst_op) {
        if (split) {                       // row = [hi(C) | lo(C)], see common.h
          long long os = ((long long)b * T + t) * (2 * C) + c;
          OpT hi = to_op<OpT>(v);
          dst_op[os] = hi;
          dst_op[os + C] = to_op<OpT>(v - from_op<OpT>(hi));
        } else {
          dst_op[o] = to_op<OpT>(v);
        }
      }
    }
  }
}

cudaError_t launch_pack_time_major(const float* src, int B, int C, int T, const int* len,
                                   float* dst_f32, void* dst_op, int op, int split, cudaStream_t stream) {
  dim3 grid(ceil_div(T, 32), ceil_div(C, 32), B), block(32, 8);
  if (op == OP_F16) pack_time_major_kernel<__half><<<grid, block, 0, stream>>>(src, C, T, len, dst_f32, (__half*)dst_op, split);
  else if (op == OP_BF16) pack_time_major_kernel<__nv_bfloat16><<<grid, block, 0, stream>>>(src, C, T, len, dst_f32, (__nv_bfloat16*)dst_op, split);
  else pack_time_major_kernel<float><<<grid, block, 0, stream>>>(src, C, T, len, dst_f32, (float*)dst_op, 0);
  return cudaGetLastError();
}

__global__ void unpack_channel_major_kernel(const float* __restrict__ src, int C, int T, float* __restrict__ dst) {
  __shared__ float tile[32][33];
  const int b = blockIdx.z, t0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += 8) {
    int t = t0 + i, c = c0 + threadIdx.x;
    tile[i][threadIdx.x] = (t < T && c < C) ? src[((long long)b * T + t) * C + c] : 0.f;
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += 8) {
    int c = c0 + i, t = t0 + threadIdx.x;
    if (c < C && t < T) dst[((long long)b * C + c) * T + t] = tile[threadIdx.x][i];
  }
}

cudaError_t launch_unpack_channel_major(const float* src, int B, int C, int T, float* dst, cudaStream_t stream) {
  dim3 grid(ceil_div(T, 32), ceil_div(C, 32), B), block(32, 8);
  unpack_channel_major_kernel<<<grid, block, 0, stream>>>(src, C, T, dst);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// Head of infer_p2 (models.py:570-572): z_p = attn @ m_p^T + noise * (attn @ s_p^T).
// attn is the duration path (0/1, one source position per frame, commons.py:143-155), so the two matmuls are a
// gather; the kernel keeps the general weighted sum but skips zero weights, which makes a 0/1 path exact
// (1.0 * x plus exact zeros) and bit-identical to the reference's fp32 matmuls.  Tile: 32 frames x 64 channels per
// block, source positions in chunks of 32 staged through shared memory (all global accesses coalesced).
//   attn [B][T][Tx], m_p / s_p [B][C][Tx], noise / z_p [B][C][T]   (reference layouts)
// ---------------------------------------------------------------------------
constexpr int kLrT = 32, kLrC = 64, kLrX = 32;
__global__ void __launch_bounds__(256)
length_regulate_kernel(const float* __restrict__ attn, const float* __restrict__ m_p, const float* __restrict__ s_p,
                       const float* __restrict__ noise, int C, int T, int Tx, float* __restrict__ z_p) {
  __shared__ float sa[kLrT][kLrX + 1];
  __shared__ float sm[kLrX][kLrC + 1];
  __shared__ float ss[kLrX][kLrC + 1];
  const int b = blockIdx.z, t0 = blockIdx.x * kLrT, c0 = blockIdx.y * kLrC;
  const int tl = threadIdx.x & 31, cg = threadIdx.x >> 5;          // frame of this thread, group of 8 channels
  float m[8], sv[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) m[i] = sv[i] = 0.f;
  const float* A = attn + (long long)b * T * Tx;
  const float* M = m_p + (long long)b * C * Tx;
  const float* S = s_p + (long long)b * C * Tx;
  for (int x0 = 0; x0 < Tx; x0 += kLrX) {
    for (int i = threadIdx.x; i < kLrT * kLrX; i += 256) {
      const int r = i / kLrX, x = i - r * kLrX;
      sa[r][x] = (t0 + r < T && x0 + x < Tx) ? A[(long long)(t0 + r) * Tx + x0 + x] : 0.f;
    }
    for (int i = threadIdx.x; i < kLrC * kLrX; i += 256) {
      const int c = i / kLrX, x = i - c * kLrX;
      const bool ok = c0 + c < C && x0 + x < Tx;
      sm[x][c] = ok ? M[(long long)(c0 + c) * Tx + x0 + x] : 0.f;
      ss[x][c] = ok ? S[(long long)(c0 + c) * Tx + x0 + x] : 0.f;
    }
    __syncthreads();
#pragma unroll 4
    for (int x = 0; x < kLrX; ++x) {
      const float a = sa[tl][x];
      if (a != 0.f) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          m[i] = fmaf(a, sm[x][cg * 8 + i], m[i]);
          sv[i] = fmaf(a, ss[x][cg * 8 + i], sv[i]);
        }
      }
    }
    __syncthreads();
  }
  const int t = t0 + tl;
  if (t < T) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int c = c0 + cg * 8 + i;
      if (c < C) {
        const long long o = ((long long)b * C + c) * T + t;
        z_p[o] = __fadd_rn(m[i], __fmul_rn(noise[o], sv[i]));     // m + noise * s, two roundings like the reference
      }
    }
  }
}

cudaError_t launch_length_regulate(const float* attn, const float* m_p, const float* s_p, const float* noise,
                                   int B, int C, int T, int Tx, float* z_p, cudaStream_t stream) {
  dim3 grid(ceil_div(T, kLrT), ceil_div(C, kLrC), B);
  length_regulate_kernel<<<grid, 256, 0, stream>>>(attn, m_p, s_p, noise, C, T, Tx, z_p);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// conditioning GEMV: one warp per output row n, all batch rows per pass of W
// (W is read once per call: 12.5k x 1024 fp32 at base config)
// ---------------------------------------------------------------------------
// Two output rows per warp: their W rows (2 x 1024 floats per pass) are loaded into registers with all loads in
// flight at once (the matrix is read from HBM exactly once), then every batch row's g is streamed through L1
// against both.  The previous one-row version re-read W per 8 batch rows and kept a single load in flight per
// lane (70 us for a 51 MB read); this one is bound by the L1 reads of g.
constexpr int kCondRows = 2, kCondK = 1024;      // rows per warp, k values per register pass (8 float4 per lane)
__global__ void __launch_bounds__(256)
cond_gemv_kernel(const float* __restrict__ W, const float* __restrict__ bias, const float* __restrict__ g,
                 int B, int Ncond, int gin, float* __restrict__ out) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int n0 = warp * kCondRows;
  if (n0 >= Ncond) return;
  for (int b0 = 0; b0 < B; b0 += 32) {             // lane (b - b0) keeps the running sums of batch row b
    float acc[kCondRows];
#pragma unroll
    for (int r = 0; r < kCondRows; ++r) acc[r] = 0.f;
    for (int k0 = 0; k0 < gin; k0 += kCondK) {
      float4 wv[kCondRows][kCondK / 128];
#pragma unroll
      for (int r = 0; r < kCondRows; ++r)
#pragma unroll
        for (int i = 0; i < kCondK / 128; ++i) {
          const int k = k0 + i * 128 + lane * 4;
          wv[r][i] = (n0 + r < Ncond && k < gin) ? __ldg(reinterpret_cast<const float4*>(W + (long long)(n0 + r) * gin + k))
                                                 : make_float4(0.f, 0.f, 0.f, 0.f);
        }
      const int bend = b0 + 32 < B ? b0 + 32 : B;
      for (int b = b0; b < bend; ++b) {
        float p[kCondRows];
#pragma unroll
        for (int r = 0; r < kCondRows; ++r) p[r] = 0.f;
#pragma unroll
        for (int i = 0; i < kCondK / 128; ++i) {
          const int k = k0 + i * 128 + lane * 4;
          if (k < gin) {
            const float4 gv = __ldg(reinterpret_cast<const float4*>(g + (long long)b * gin + k));
#pragma unroll
            for (int r = 0; r < kCondRows; ++r) {
              p[r] = fmaf(wv[r][i].x, gv.x, p[r]); p[r] = fmaf(wv[r][i].y, gv.y, p[r]);
              p[r] = fmaf(wv[r][i].z, gv.z, p[r]); p[r] = fmaf(wv[r][i].w, gv.w, p[r]);
            }
          }
        }
#pragma unroll
        for (int r = 0; r < kCondRows; ++r) {
          float v = p[r];
#pragma unroll
          for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
          if (lane == b - b0) acc[r] += v;
        }
      }
    }
    const int b = b0 + lane;
    if (b < B) {
#pragma unroll
      for (int r = 0; r < kCondRows; ++r)
        if (n0 + r < Ncond) out[(long long)b * Ncond + n0 + r] = acc[r] + bias[n0 + r];
    }
  }
}

cudaError_t launch_cond_gemv(const float* W, const float* bias, const float* g, int B, int Ncond, int gin,
                             float* out, cudaStream_t stream) {
  const int warps_per_block = 8;
  cond_gemv_kernel<<<ceil_div(ceil_div(Ncond, kCondRows), warps_per_block), warps_per_block * 32, 0, stream>>>(W, bias, g, B, Ncond, gin, out);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// MRF average (models.py:311-314) over the branches' final 16-bit y streams (y = leaky_relu(x_j, 0.1), so
// x_j = min(y, 10 y)): next = op(leaky_relu(scale * sum_j x_j, 0.1)), written as the next upsampler's operand
// rows ([hi | lo] when split).  HBM-bound and fully coalesced: 8 channels (one 16-byte vector per stream) per
// thread; 2 B x n_streams in + 2-4 B out per element.  Inside the fused ResBlock kernel the same sum cost more
// than this whole pass: its row-per-thread epilogue paid one exposed DRAM round trip per 16-column chunk.
// ---------------------------------------------------------------------------
template <typename OpT>
__device__ __forceinline__ void y_accum8(const uint4& y, float (&x)[8]) {
  const uint32_t w[4] = {y.x, y.y, y.z, y.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = Pack16<OpT>::unpack2(w[i]);
    x[2 * i] += fminf(f.x, 10.f * f.x);
    x[2 * i + 1] += fminf(f.y, 10.f * f.y);
  }
}

template <typename OpT>
__global__ void __launch_bounds__(256)
mrf_sum_kernel(const OpT* __restrict__ y0, const OpT* __restrict__ y1, const OpT* __restrict__ y2, int C, int L,
               const int* __restrict__ len, int len_mul, float scale, float slope, OpT* __restrict__ out, int split,
               float* __restrict__ out_f32) {
  const int b = blockIdx.y;
  const int valid = len ? len[b] * len_mul : L;
  const int rows = valid + kHaloRows < L ? valid + kHaloRows : L;      // rows nobody reads are skipped (as every producer does)
  const int C8 = C >> 3;
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int t = (int)(i / C8), c = (int)(i - (long long)t * C8) * 8;
  if (t >= rows) return;
  const long long e = ((long long)b * L + t) * C + c;
  float x[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  if (t < valid) {
    const uint4 a0 = __ldg(reinterpret_cast<const uint4*>(y0 + e));
    uint4 a1 = make_uint4(0u, 0u, 0u, 0u), a2 = a1;
    if (y1) a1 = __ldg(reinterpret_cast<const uint4*>(y1 + e));
    if (y2) a2 = __ldg(reinterpret_cast<const uint4*>(y2 + e));
    y_accum8<OpT>(a0, x); y_accum8<OpT>(a1, x); y_accum8<OpT>(a2, x);
  }
  if (out_f32) {                          // fp32 average itself (input of the generic conv_post)
    float4* o = reinterpret_cast<float4*>(out_f32 + e);
    o[0] = make_float4(x[0] * scale, x[1] * scale, x[2] * scale, x[3] * scale);
    o[1] = make_float4(x[4] * scale, x[5] * scale, x[6] * scale, x[7] * scale);
    return;
  }
  uint32_t p[4], q[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    float u = x[2 * k] * scale, v = x[2 * k + 1] * scale;
    u = u >= 0.f ? u : slope * u; v = v >= 0.f ? v : slope * v;
    p[k] = Pack16<OpT>::pack2(u, v);
    const float2 hi = Pack16<OpT>::unpack2(p[k]);
    q[k] = Pack16<OpT>::pack2(u - hi.x, v - hi.y);
  }
  OpT* o = out + ((long long)b * L + t) * (split ? 2 * C : C) + c;
  *reinterpret_cast<uint4*>(o) = make_uint4(p[0], p[1], p[2], p[3]);
  if (split) *reinterpret_cast<uint4*>(o + C) = make_uint4(q[0], q[1], q[2], q[3]);
}

cudaError_t launch_mrf_sum(const void* y0, const void* y1, const void* y2, int B, int L, int C, const int* len, int len_mul,
                           float scale, float slope, void* out, int split, float* out_f32, int op, cudaStream_t stream) {
  if (op == OP_F32 || C % 8 != 0 || !y0 || (!out && !out_f32)) return cudaErrorNotSupported;
  const long long per_b = (long long)L * (C / 8);
  dim3 grid((unsigned)((per_b + 255) / 256), B);
  if (op == OP_F16) mrf_sum_kernel<__half><<<grid, 256, 0, stream>>>((const __half*)y0, (const __half*)y1, (const __half*)y2, C, L, len, len_mul,
                                                                     scale, slope, (__half*)out, split, out_f32);
  else mrf_sum_kernel<__nv_bfloat16><<<grid, 256, 0, stream>>>((const __nv_bfloat16*)y0, (const __nv_bfloat16*)y1, (const __nv_bfloat16*)y2, C, L,
                                                               len, len_mul, scale, slope, (__nv_bfloat16*)out, split, out_f32);
  return cudaGetLastError();
}

// ---------------------------------------------------------------------------
// leaky_relu(slope) + conv_post (Conv1d C->1, k taps, no bias) + tanh.  HBM-bound:
// reads C fp32 values (the MRF accumulator of the last stage, never rounded to
// 16 bits: conv_post sits right before the output and its input rounding would
// be the single largest error term) and writes one fp32 sample per position.
// ---------------------------------------------------------------------------
// Output post-processing of the serving path fused into the last kernel (vits_wrap.py:195:
// np.clip(wav * volume * 32767, -32768, 32767).astype(np.int16) - float32 products, truncation toward zero)
__device__ __forceinline__ short to_pcm16(float w, float volume) {
  float v = __fmul_rn(__fmul_rn(w, volume), 32767.f);
  v = fminf(fmaxf(v, -32768.f), 32767.f);
  return (short)__float2int_rz(v);
}

constexpr int kPostTile = 256;
__global__ void __launch_bounds__(kPostTile)
conv_post_kernel(const float* __restrict__ in, int L, int C, int k, float slope, const float* __restrict__ w,
                 const int* __restrict__ len, int len_mul, float* __restrict__ wav, short* __restrict__ pcm, float volume) {
  extern __shared__ float sm[];
  const int pad = (k - 1) / 2;
  const int rows = kPostTile + k - 1;
  const int ldc = C + 1;                       // +1: conflict-free row-per-thread reads
  float* tile = sm;                            // [rows][C+1]
  float* ws = sm + rows * ldc;                 // [k][C]
  const int b = blockIdx.y, t0 = blockIdx.x * kPostTile;
  const int valid = len ? len[b] * len_mul : L;
  if (t0 >= valid) {                           // fully masked tile: write zeros
    int t = t0 + threadIdx.x;
    if (t < L) { if (pcm) pcm[(long long)b * L + t] = 0; else wav[(long long)b * L + t] = 0.f; }
    return;
  }
  const float* src = in + (long long)b * L * C;
  for (int i = threadIdx.x; i < rows * C; i += kPostTile) {
    int r = i / C, c = i - r * C, t = t0 + r - pad;
    float x = (t >= 0 && t < L) ? src[(long long)t * C + c] : 0.f;
    tile[r * ldc + c] = x >= 0.f ? x : slope * x;
  }
  for (int i = threadIdx.x; i < k * C; i += kPostTile) ws[i] = w[i];
  __syncthreads();
  const int t = t0 + threadIdx.x;
  if (t >= L) return;
  float acc = 0.f;
  for (int j = 0; j < k; ++j) {
    const float* row = tile + (threadIdx.x + j) * ldc;
    const float* wj = ws + j * C;
#pragma unroll 8
    for (int c = 0; c < C; ++c) acc = fmaf(row[c], wj[c], acc);
  }
  const float o = t < valid ? tanhf(acc) : 0.f;
  if (pcm) pcm[(long long)b * L + t] = to_pcm16(o, volume); else wav[(long long)b * L + t] = o;
}

// Fast path for the base shape (C = 32, k = 7, L % 4 == 0): 64 threads x 4 consecutive outputs.  The generic
// kernel above issues two LDS.32 per FMA and is shared-memory-issue bound (~2.4x the HBM time of its input);
// here a thread keeps the 10 input rows of its 4 outputs in registers per 4-channel group, so one LDS.128 feeds
// up to 16 FMAs.  Tile rows are stored as 8 float4 with the float4 index XOR-swizzled by (row / 4), which makes
// the row loads of threads that sit 4 rows apart conflict-free.
constexpr int kPost4Out = 4, kPost4Threads = 64, kPost4Tile = kPost4Out * kPost4Threads;
// YT = float: `in` is the fp32 MRF average.  YT = 16-bit operand type: the MRF average is formed here from the
// branches' final y streams (in, in1, in2; x_j = min(y, 10 y)) times `scale` - the last stage then needs neither
// an fp32 accumulator pass nor a summing ResBlock epilogue.
template <int C, int K, typename YT>
__global__ void __launch_bounds__(kPost4Threads)
conv_post4_kernel(const YT* __restrict__ in, const YT* __restrict__ in1, const YT* __restrict__ in2, float scale, int L, float slope,
                  const float* __restrict__ w, const int* __restrict__ len, int len_mul, float* __restrict__ wav,
                  short* __restrict__ pcm, float volume) {
  constexpr int C4 = C / 4, ROWS = kPost4Tile + K - 1, PAD = (K - 1) / 2;
  static_assert(C4 == 8, "swizzle assumes 8 float4 per row");
  __shared__ float4 tile[ROWS * C4];
  __shared__ float4 ws[K * C4];
  const int b = blockIdx.y, t0 = blockIdx.x * kPost4Tile;
  const int valid = len ? len[b] * len_mul : L;
  const int t = t0 + threadIdx.x * kPost4Out;
  if (t0 >= valid) {                           // fully masked tile: write zeros
    if (t < L) {
      if (pcm) *reinterpret_cast<short4*>(pcm + (long long)b * L + t) = make_short4(0, 0, 0, 0);
      else *reinterpret_cast<float4*>(wav + (long long)b * L + t) = make_float4(0.f, 0.f, 0.f, 0.f);
    }
    return;
  }
  if constexpr (sizeof(YT) == 4) {
    const float4* src = reinterpret_cast<const float4*>(reinterpret_cast<const float*>(in) + (long long)b * L * C);
    for (int i = threadIdx.x; i < ROWS * C4; i += kPost4Threads) {
      const int r = i / C4, c4 = i - r * C4, tt = t0 + r - PAD;
      float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
      if (tt >= 0 && tt < L) x = __ldg(src + (long long)tt * C4 + c4);
      x.x = x.x >= 0.f ? x.x : slope * x.x; x.y = x.y >= 0.f ? x.y : slope * x.y;
      x.z = x.z >= 0.f ? x.z : slope * x.z; x.w = x.w >= 0.f ? x.w : slope * x.w;
      tile[r * C4 + ((c4 ^ (r >> 2)) & 7)] = x;
    }
  } else {
    constexpr int C8 = C / 8, U = 4;              // U items per thread per pass: 3 U independent 16-byte loads in flight
    const long long boff = (long long)b * L * C;
    for (int i0 = threadIdx.x; i0 < ROWS * C8; i0 += U * kPost4Threads) {
      uint4 a[U][3];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int i = i0 + u * kPost4Threads, r = i / C8, c8 = i - r * C8, tt = t0 + r - PAD;
        a[u][0] = a[u][1] = a[u][2] = make_uint4(0u, 0u, 0u, 0u);
        if (i < ROWS * C8 && tt >= 0 && tt < valid) {          // rows beyond len[b] are zero in every stream
          const long long e = boff + (long long)tt * C + c8 * 8;
          a[u][0] = __ldg(reinterpret_cast<const uint4*>(in + e));
          if (in1) a[u][1] = __ldg(reinterpret_cast<const uint4*>(in1 + e));
          if (in2) a[u][2] = __ldg(reinterpret_cast<const uint4*>(in2 + e));
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const int i = i0 + u * kPost4Threads, r = i / C8, c8 = i - r * C8;
        if (i >= ROWS * C8) break;
        float x[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        y_accum8<YT>(a[u][0], x); y_accum8<YT>(a[u][1], x); y_accum8<YT>(a[u][2], x);
#pragma unroll
        for (int q = 0; q < 8; ++q) { const float v = x[q] * scale; x[q] = v >= 0.f ? v : slope * v; }
        tile[r * C4 + (((2 * c8) ^ (r >> 2)) & 7)] = make_float4(x[0], x[1], x[2], x[3]);
        tile[r * C4 + (((2 * c8 + 1) ^ (r >> 2)) & 7)] = make_float4(x[4], x[5], x[6], x[7]);
      }
    }
  }
  for (int i = threadIdx.x; i < K * C4; i += kPost4Threads) ws[i] = __ldg(reinterpret_cast<const float4*>(w) + i);
  __syncthreads();
  if (t >= L) return;
  float acc[kPost4Out] = {0.f, 0.f, 0.f, 0.f};
  const int r0 = threadIdx.x * kPost4Out;
#pragma unroll 2
  for (int c4 = 0; c4 < C4; ++c4) {
    float4 x[kPost4Out + K - 1];
#pragma unroll
    for (int jj = 0; jj < kPost4Out + K - 1; ++jj) {
      const int r = r0 + jj;
      x[jj] = tile[r * C4 + ((c4 ^ (r >> 2)) & 7)];
    }
#pragma unroll
    for (int j = 0; j < K; ++j) {
      const float4 wj = ws[j * C4 + c4];
#pragma unroll
      for (int o = 0; o < kPost4Out; ++o) {
        acc[o] = fmaf(x[o + j].x, wj.x, acc[o]); acc[o] = fmaf(x[o + j].y, wj.y, acc[o]);
        acc[o] = fmaf(x[o + j].z, wj.z, acc[o]); acc[o] = fmaf(x[o + j].w, wj.w, acc[o]);
      }
    }
  }
  float4 o4;
  o4.x = t < valid ? tanhf(acc[0]) : 0.f;     o4.y = t + 1 < valid ? tanhf(acc[1]) : 0.f;
  o4.z = t + 2 < valid ? tanhf(acc[2]) : 0.f; o4.w = t + 3 < valid ? tanhf(acc[3]) : 0.f;
  if (pcm) *reinterpret_cast<short4*>(pcm + (long long)b * L + t) =
      make_short4(to_pcm16(o4.x, volume), to_pcm16(o4.y, volume), to_pcm16(o4.z, volume), to_pcm16(o4.w, volume));
  else *reinterpret_cast<float4*>(wav + (long long)b * L + t) = o4;
}

cudaError_t launch_conv_post(const float* in, int B, int L, int C, int k, float slope, const float* w,
                             const int* len, int len_mul, float* wav, short* pcm, float volume, cudaStream_t stream) {
  if (C == 32 && k == 7 && L % 4 == 0) {
    dim3 grid4(ceil_div(L, kPost4Tile), B);
    conv_post4_kernel<32, 7, float><<<grid4, kPost4Threads, 0, stream>>>(in, nullptr, nullptr, 1.f, L, slope, w, len, len_mul, wav, pcm, volume);
    return cudaGetLastError();
  }
  if (!conv_post_supported(C, k)) return cudaErrorNotSupported;
  dim3 grid(ceil_div(L, kPostTile), B);
  size_t smem = (size_t)((kPostTile + k - 1) * (C + 1) + k * C) * sizeof(float);
  static PerDeviceOnce smem_optin;
  const cudaError_t ae = smem_optin.run([] {
    return cudaFuncSetAttribute(conv_post_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  });
  if (ae != cudaSuccess) return ae;
  conv_post_kernel<<<grid, kPostTile, smem, stream>>>(in, L, C, k, slope, w, len, len_mul, wav, pcm, volume);
  return cudaGetLastError();
}

// generic kernel: the [256 + k - 1][C + 1] fp32 tile and the weights must fit the opted-in shared memory
bool conv_post_supported(int C, int k) {
  return C >= 1 && k >= 1 && (size_t)((kPostTile + k - 1) * (C + 1) + k * C) * sizeof(float) <= 200 * 1024;
}

bool conv_post_y_supported(int C, int k, int L) { return C == 32 && k == 7 && L % 4 == 0; }

cudaError_t launch_conv_post_y(const void* y0, const void* y1, const void* y2, float scale, int B, int L, int C, int k, float slope,
                               const float* w, const int* len, int len_mul, float* wav, short* pcm, float volume, int op,
                               cudaStream_t stream) {
  if (op == OP_F32 || !conv_post_y_supported(C, k, L) || !y0) return cudaErrorNotSupported;
  dim3 grid4(ceil_div(L, kPost4Tile), B);
  if (op == OP_F16)
    conv_post4_kernel<32, 7, __half><<<grid4, kPost4Threads, 0, stream>>>((const __half*)y0, (const __half*)y1, (const __half*)y2, scale, L, slope,
                                                                          w, len, len_mul, wav, pcm, volume);
  else
    conv_post4_kernel<32, 7, __nv_bfloat16><<<grid4, kPost4Threads, 0, stream>>>((const __nv_bfloat16*)y0, (const __nv_bfloat16*)y1,
                                                                                 (const __nv_bfloat16*)y2, scale, L, slope, w, len, len_mul, wav, pcm, volume);
  return cudaGetLastError();
}

}  // namespace vitsb

namespace vitsb {

// ---------------------------------------------------------------------------
// Device-side durations.  Replaces the host glue between infer_p1 and infer_p2
//   w = exp(logw) * length_scale; w_ceil = ceil(w); y_length = clamp_min(sum(w_ceil), 1).item()      infer.py:169-171, models.py:544-547
//   attn = infer_path(w_ceil, x_length, y_length)                                                   commons.py:143-155
// and the head of infer_p2 (models.py:570-572) without the dense 0/1 attn matrix, without the .item() host sync and
// without a channel-major z_p: frame t of row b takes source position x(t) with cum[x-1] <= t < cum[x].
// ---------------------------------------------------------------------------
// One block per batch row: cum[b][x] = inclusive cumsum of ceil(exp(logw) * scale) over x < x_len[b] (integers, exact),
// y_len[b] = min(max(cum[last], 1), T_cap); noise_off[b] = C * sum_{b' < b} y_len[b'] is filled by a second tiny pass.
__global__ void __launch_bounds__(256)
durations_scan_kernel(const float* __restrict__ logw, const int* __restrict__ x_len, int Tx, float scale, int T_cap,
                      int* __restrict__ cum, int* __restrict__ y_len, int* __restrict__ y_len_out) {
  __shared__ int s_part[256];
  const int b = blockIdx.x, tid = threadIdx.x;
  const int n = x_len ? x_len[b] : Tx;
  const int per = (Tx + 255) / 256;                    // consecutive positions per thread
  int local[4];                                         // Tx <= 1024
  int sum = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int x = tid * per + i;
    int w = 0;
    if (i < per && x < n) {
      const float v = ceilf(__fmul_rn(expf(logw[(long long)b * Tx + x]), scale));       // torch.ceil(torch.exp(logw) * length_scale)
      w = v > 0.f ? (v < 1e9f ? (int)v : 1000000000) : 0;
    }
    sum += w;
    local[i] = sum;
  }
  s_part[tid] = sum;
  __syncthreads();
  for (int o = 1; o < 256; o <<= 1) {                   // Hillis-Steele inclusive scan of the per-thread sums
    const int v = tid >= o ? s_part[tid - o] : 0;
    __syncthreads();
    s_part[tid] += v;
    __syncthreads();
  }
  const int before = tid > 0 ? s_part[tid - 1] : 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int x = tid * per + i;
    if (i < per && x < Tx) {
      long long c = (long long)before + local[i];
      cum[(long long)b * Tx + x] = c > 2000000000LL ? 2000000000 : (int)c;
    }
  }
  if (tid == 255) {
    int y = s_part[255] < 1 ? 1 : s_part[255];          // clamp_min(sum, 1)
    if (n == 0) y = 0;                                  // an empty row stays empty
    y = y > T_cap ? T_cap : y;
    y_len[b] = y;
    if (y_len_out) y_len_out[b] = y;
  }
}

// z[b][t][c] = m_p[b][c][x(t)] + noise[b][c][t] * s_p[b][c][x(t)] for t < y_len[b], else 0; written time-major as the fp32
// flow state and its operand copy ([hi | lo] when split).  Tile 32 frames x 32 channels through shared memory: the
// noise reads are coalesced along t, the stores along c.  noise_stride > 0: noise is [B][C][noise_stride]; 0: the
// reference's compact form, row b = a contiguous [C][y_len[b]] block after the blocks of the rows before it
// (infer.py:172-174 views a flat pool as [1, C, y_length]).
template <typename OpT>
__global__ void __launch_bounds__(256)
durations_gather_kernel(const float* __restrict__ m_p, const float* __restrict__ s_p, const float* __restrict__ noise,
                        const int* __restrict__ cum, const int* __restrict__ y_len, const int* __restrict__ x_len, int B, int C, int Tx,
                        int T, long long noise_stride, float* __restrict__ zf, OpT* __restrict__ zop, int split) {
  __shared__ float tile[32][33];
  __shared__ int s_x[32];
  const int b = blockIdx.z, t0 = blockIdx.x * 32, c0 = blockIdx.y * 32;
  const int yl = y_len[b];
  if (t0 < yl) {
    const int n = x_len ? x_len[b] : Tx;
    if (threadIdx.y == 0) {                             // source position of each of the 32 frames: first x with cum[x] > t
      const int t = t0 + threadIdx.x;
      int lo = 0, hi = n > 0 ? n - 1 : 0;
      const int* cb = cum + (long long)b * Tx;
      while (lo < hi) { const int mid = (lo + hi) >> 1; if (cb[mid] > t) hi = mid; else lo = mid + 1; }
      s_x[threadIdx.x] = lo;
    }
    long long nbase, nstr;
    if (noise_stride > 0) { nstr = noise_stride; nbase = (long long)b * C * noise_stride; }
    else { nstr = yl; nbase = 0; for (int bb = 0; bb < b; ++bb) nbase += (long long)C * y_len[bb]; }
    __syncthreads();
    for (int i = threadIdx.y; i < 32; i += 8) {
      const int c = c0 + i, t = t0 + threadIdx.x;
      float v = 0.f;
      if (c < C && t < yl) {
        const long long so = ((long long)b * C + c) * Tx + s_x[threadIdx.x];
        v = __fadd_rn(m_p[so], __fmul_rn(noise[nbase + (long long)c * nstr + t], s_p[so]));    // m + noise * s, two roundings (models.py:572)
      }
      tile[i][threadIdx.x] = v;
    }
    __syncthreads();
  }
  for (int i = threadIdx.y; i < 32; i += 8) {
    const int t = t0 + i, c = c0 + threadIdx.x;
    if (t >= T || c >= C) continue;
    const float v = t0 < yl ? tile[threadIdx.x][i] : 0.f;
    const long long o = ((long long)b * T + t) * C + c;
    if (zf) zf[o] = v;
    if (split) {
      const long long os = ((long long)b * T + t) * (2 * C) + c;
      const OpT hi = to_op<OpT>(v);
      zop[os] = hi;
      zop[os + C] = to_op<OpT>(v - from_op<OpT>(hi));
    } else {
      zop[o] = to_op<OpT>(v);
    }
  }
}

cudaError_t launch_durations(const float* logw, const int* x_len, const float* m_p, const float* s_p, const float* noise,
                             long long noise_stride, int B, int C, int Tx, int T, float scale, int* cum, int* y_len, int* y_len_out,
                             float* zf, void* zop, int op, int split, cudaStream_t stream) {
  if (Tx < 1 || Tx > 1024 || B < 1 || B > 65535) return cudaErrorInvalidValue;
  durations_scan_kernel<<<B, 256, 0, stream>>>(logw, x_len, Tx, scale, T, cum, y_len, y_len_out);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) return e;
  dim3 grid(ceil_div(T, 32), ceil_div(C, 32), B), block(32, 8);
  if (op == OP_F16) durations_gather_kernel<__half><<<grid, block, 0, stream>>>(m_p, s_p, noise, cum, y_len, x_len, B, C, Tx, T, noise_stride, zf, (__half*)zop, split);
  else if (op == OP_BF16) durations_gather_kernel<__nv_bfloat16><<<grid, block, 0, stream>>>(m_p, s_p, noise, cum, y_len, x_len, B, C, Tx, T, noise_stride, zf, (__nv_bfloat16*)zop, split);
  else durations_gather_kernel<float><<<grid, block, 0, stream>>>(m_p, s_p, noise, cum, y_len, x_len, B, C, Tx, T, noise_stride, zf, (float*)zop, 0);
  return cudaGetLastError();
}

}  // namespace vitsb
