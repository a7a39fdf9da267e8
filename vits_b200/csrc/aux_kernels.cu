// Memory-bound kernels around the implicit-GEMM convs: layout pack/unpack at
// the seam, the per-call conditioning GEMV, and conv_post+tanh (N=1, CUDA cores).
#include <atomic>
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
// conditioning GEMV: out[b][n] = sum_k W[n][k] g[b][k] + bias[n] for all conditioning Linears of a call
// (W: 12.5k x 1024 fp32 = 51 MB at base config, read from HBM exactly once per call)
// ---------------------------------------------------------------------------
// Fast kernel (gin <= 1024): the BATCH side lives in registers.  Warp q of a CTA owns the K quarter [256 q, 256 q + 256);
// a lane keeps its 8 k-values of g for 16 batch rows as 64 (b, b + 1) pairs, so a row of W is streamed straight from
// global memory into registers, used once (2 x 16-byte loads feed 64 packed fma.rn.f32x2) and never touches shared
// memory or L1 again.  Two W rows per step: 32 partial sums per lane, one 31-shuffle transpose-reduce leaves output
// (row l / 16, batch row l % 16) in lane l, the four K quarters meet in 1 KB of shared memory.  CTAs are persistent,
// load the next step's rows under the current step's FMAs and prefetch four steps ahead into L2.
// History (C2, 51 MB): one row per warp, W re-read per 8 batch rows: 70 us; W tile in registers, g streamed through
// L1 (two designs, 2 and 4 rows per warp): 42 us both - ncu: 71 % of issue slots empty on the L1 round trip of each
// g load (`long_scoreboard`), 4 warps per scheduler, 1.2 TB/s.
__device__ __forceinline__ unsigned long long pack_f32x2(float lo, float hi) {
  unsigned long long r;
  asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
  return r;
}
__device__ __forceinline__ void ffma2(unsigned long long& acc, unsigned long long a, unsigned long long b) {
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(acc) : "l"(a), "l"(b));
}
constexpr int kGvB = 16, kGvThreads = 128, kGvK = 1024;      // batch rows per pass, threads (4 K quarters), largest gin
__global__ void __launch_bounds__(kGvThreads, 2)
cond_gemv_regs_kernel(const float* __restrict__ W, const float* __restrict__ bias, const float* __restrict__ g,
                      int B, int Ncond, int gin, float* __restrict__ out) {
  __shared__ float part[2][kGvThreads / 32][32];
  const int q = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int kA = q * 256 + lane * 4, kB = kA + 128;            // the lane's two 4-float runs of k
  const bool okA = kA < gin, okB = kB < gin;
  const int n_steps = (Ncond + 1) / 2;
  const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
  auto load_rows = [&](int s, float4 (&w)[2][2]) {
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      const int n = 2 * s + r;
      const float* row = W + (long long)n * gin;
      w[r][0] = (n < Ncond && okA) ? __ldg(reinterpret_cast<const float4*>(row + kA)) : z4;
      w[r][1] = (n < Ncond && okB) ? __ldg(reinterpret_cast<const float4*>(row + kB)) : z4;
    }
  };
  for (int b0 = 0; b0 < B; b0 += kGvB) {
    unsigned long long g2[kGvB / 2][8];                        // [batch pair][k index] = (g[b][k], g[b + 1][k])
#pragma unroll
    for (int bp = 0; bp < kGvB / 2; ++bp) {
      const int ba = b0 + 2 * bp, bb = ba + 1;
      const float* ra = g + (long long)ba * gin;
      const float* rb = g + (long long)bb * gin;
      const float4 a0 = (ba < B && okA) ? __ldg(reinterpret_cast<const float4*>(ra + kA)) : z4;
      const float4 a1 = (ba < B && okB) ? __ldg(reinterpret_cast<const float4*>(ra + kB)) : z4;
      const float4 c0 = (bb < B && okA) ? __ldg(reinterpret_cast<const float4*>(rb + kA)) : z4;
      const float4 c1 = (bb < B && okB) ? __ldg(reinterpret_cast<const float4*>(rb + kB)) : z4;
      g2[bp][0] = pack_f32x2(a0.x, c0.x); g2[bp][1] = pack_f32x2(a0.y, c0.y);
      g2[bp][2] = pack_f32x2(a0.z, c0.z); g2[bp][3] = pack_f32x2(a0.w, c0.w);
      g2[bp][4] = pack_f32x2(a1.x, c1.x); g2[bp][5] = pack_f32x2(a1.y, c1.y);
      g2[bp][6] = pack_f32x2(a1.z, c1.z); g2[bp][7] = pack_f32x2(a1.w, c1.w);
    }
    float4 wn[2][2];
    if ((int)blockIdx.x < n_steps) load_rows(blockIdx.x, wn);
    int par = 0;
    for (int s = blockIdx.x; s < n_steps; s += gridDim.x, par ^= 1) {
      float4 wc[2][2];
#pragma unroll
      for (int r = 0; r < 2; ++r) { wc[r][0] = wn[r][0]; wc[r][1] = wn[r][1]; }
      if (s + (int)gridDim.x < n_steps) load_rows(s + gridDim.x, wn);
      if (b0 == 0 && threadIdx.x < 64) {                      // the rows of four steps ahead -> L2 (first pass only)
        const long long e = ((long long)(s + 4 * (int)gridDim.x) * 2) * gin + threadIdx.x * 32;
        if (threadIdx.x * 32 < 2 * gin && e < (long long)Ncond * gin) asm volatile("prefetch.global.L2 [%0];" ::"l"(W + e));
      }
      const int comb = (s / (int)gridDim.x) & 3;             // warp that combines this step's K quarters (rotates: balanced)
      const int n_out = 2 * s + (lane >> 4);
      float bias_n = 0.f;
      if (q == comb && n_out < Ncond) bias_n = __ldg(bias + n_out);      // needed after the barrier: in flight under the FMAs
      unsigned long long acc[2][kGvB / 2];
#pragma unroll
      for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int bp = 0; bp < kGvB / 2; ++bp) acc[r][bp] = 0ull;
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        const float wk[8] = {wc[r][0].x, wc[r][0].y, wc[r][0].z, wc[r][0].w, wc[r][1].x, wc[r][1].y, wc[r][1].z, wc[r][1].w};
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const unsigned long long ww = pack_f32x2(wk[j], wk[j]);
#pragma unroll
          for (int bp = 0; bp < kGvB / 2; ++bp) ffma2(acc[r][bp], ww, g2[bp][j]);
        }
      }
      float v[32];                                             // index = r * 16 + batch row
#pragma unroll
      for (int r = 0; r < 2; ++r)
#pragma unroll
        for (int bp = 0; bp < kGvB / 2; ++bp)
          asm("mov.b64 {%0, %1}, %2;" : "=f"(v[r * 16 + 2 * bp]), "=f"(v[r * 16 + 2 * bp + 1]) : "l"(acc[r][bp]));
      // transpose-reduce: after the step with offset o a lane keeps the half of its values whose index bit matches its
      // lane bit; 16 + 8 + 4 + 2 + 1 shuffles, and lane l ends with the warp's sum of element l
#pragma unroll
      for (int o = 16; o >= 1; o >>= 1) {
        const bool up = (lane & o) != 0;
#pragma unroll
        for (int i = 0; i < o; ++i) {
          const float send = up ? v[i] : v[i + o];
          const float keep = up ? v[i + o] : v[i];
          v[i] = keep + __shfl_xor_sync(0xffffffffu, send, o);
        }
      }
      part[par][q][lane] = v[0];
      __syncthreads();                                         // one barrier per step: `part` is double-buffered by step parity
      if (q == comb) {
        const int b = b0 + (lane & 15);
        if (n_out < Ncond && b < B)
          out[(long long)b * Ncond + n_out] = ((part[par][0][lane] + part[par][1][lane]) + (part[par][2][lane] + part[par][3][lane])) + bias_n;
      }
    }
    __syncthreads();
  }
}

// Generic kernel (any gin that is a multiple of 4): the W tile lives in registers and g streams through L1.
// A warp holds a 4-row x 512-k tile of W (two warps share a row group, one per K half) and runs the batch rows
// against it 8 at a time; 31-shuffle transpose-reduce as above.
constexpr int kCgRows = 4, kCgKS = 512, kCgBt = 8, kCgThreads = 128;
constexpr int kCgTileRows = kCgRows * (kCgThreads / 64);      // W rows per CTA tile (one row group per warp pair)
__global__ void __launch_bounds__(kCgThreads, 4)
cond_gemv_kernel(const float* __restrict__ W, const float* __restrict__ bias, const float* __restrict__ g,
                 int B, int Ncond, int gin, float* __restrict__ out) {
  static_assert(kCgRows * kCgBt == 32, "the transpose-reduce leaves one output per lane");
  __shared__ float part[kCgThreads / 64][32];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int grp = warp >> 1, kh = warp & 1;
  const int n_tiles = (Ncond + kCgTileRows - 1) / kCgTileRows;
  const bool one_pass = gin <= 2 * kCgKS;            // the warp's K half fits its register tile: W is loaded once per tile
  for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
    const int n0 = tile * kCgTileRows + grp * kCgRows;
    {                                                 // next tile of this CTA -> L2 (whole rows: both K halves)
      const long long nn = (long long)(tile + gridDim.x) * kCgTileRows * gin;
      const long long end = (long long)Ncond * gin;
      for (long long e = nn + (long long)threadIdx.x * 32; e < nn + (long long)kCgTileRows * gin && e < end; e += kCgThreads * 32)
        asm volatile("prefetch.global.L2 [%0];" ::"l"(W + e));
    }
    float4 wv[kCgRows][kCgKS / 128];
    auto load_w = [&](int k0) {
#pragma unroll
      for (int r = 0; r < kCgRows; ++r)
#pragma unroll
        for (int i = 0; i < kCgKS / 128; ++i) {
          const int k = k0 + i * 128 + lane * 4;
          wv[r][i] = (n0 + r < Ncond && k < gin) ? __ldg(reinterpret_cast<const float4*>(W + (long long)(n0 + r) * gin + k))
                                                 : make_float4(0.f, 0.f, 0.f, 0.f);
        }
    };
    if (one_pass) load_w(kh * kCgKS);
    for (int b0 = 0; b0 < B; b0 += kCgBt) {
      float acc[32];                                  // [row r][batch row bb] at r * kCgBt + bb
#pragma unroll
      for (int i = 0; i < 32; ++i) acc[i] = 0.f;
      for (int k0 = kh * kCgKS; k0 < gin; k0 += 2 * kCgKS) {
        if (!one_pass) load_w(k0);
#pragma unroll
        for (int bb = 0; bb < kCgBt; ++bb) {
          if (b0 + bb >= B) break;
          const float* gb = g + (long long)(b0 + bb) * gin;
#pragma unroll
          for (int i = 0; i < kCgKS / 128; ++i) {
            const int k = k0 + i * 128 + lane * 4;
            if (k < gin) {
              const float4 gv = __ldg(reinterpret_cast<const float4*>(gb + k));
#pragma unroll
              for (int r = 0; r < kCgRows; ++r) {
                float v = acc[r * kCgBt + bb];
                v = fmaf(wv[r][i].x, gv.x, v); v = fmaf(wv[r][i].y, gv.y, v);
                v = fmaf(wv[r][i].z, gv.z, v); v = fmaf(wv[r][i].w, gv.w, v);
                acc[r * kCgBt + bb] = v;
              }
            }
          }
        }
      }
      // transpose-reduce: after the step with offset o a lane keeps the half of its values whose index bit matches its
      // lane bit; 16 + 8 + 4 + 2 + 1 shuffles, and lane l ends with the warp's sum of element l
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
      if (kh == 1) part[grp][lane] = acc[0];
      __syncthreads();
      if (kh == 0) {
        const int r = lane / kCgBt, b = b0 + lane % kCgBt;
        if (n0 + r < Ncond && b < B) out[(long long)b * Ncond + n0 + r] = acc[0] + part[grp][lane] + bias[n0 + r];
      }
      __syncthreads();
    }
  }
}

// SM count of the current device (persistent grids), queried once per device
static cudaError_t current_sm_count(int* out) {
  static std::atomic<int> cached[64];
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  int n = (dev >= 0 && dev < 64) ? cached[dev].load(std::memory_order_relaxed) : 0;
  if (n <= 0) {
    e = cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (e != cudaSuccess) return e;
    if (dev >= 0 && dev < 64) cached[dev].store(n, std::memory_order_relaxed);
  }
  *out = n;
  return cudaSuccess;
}

cudaError_t launch_cond_gemv(const float* W, const float* bias, const float* g, int B, int Ncond, int gin,
                             float* out, cudaStream_t stream) {
  if (gin % 4 != 0) return cudaErrorNotSupported;
  int num_sms = 0;
  const cudaError_t se = current_sm_count(&num_sms);
  if (se != cudaSuccess) return se;
  if (gin <= kGvK) {
    const int n_steps = (Ncond + 1) / 2;
    const int grid = n_steps < 2 * num_sms ? n_steps : 2 * num_sms;
    cond_gemv_regs_kernel<<<grid, kGvThreads, 0, stream>>>(W, bias, g, B, Ncond, gin, out);
    return cudaGetLastError();
  }
  const int n_tiles = ceil_div(Ncond, kCgTileRows);
  const int grid = n_tiles < 4 * num_sms ? n_tiles : 4 * num_sms;
  cond_gemv_kernel<<<grid, kCgThreads, 0, stream>>>(W, bias, g, B, Ncond, gin, out);
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
#ifdef VITSB_NO_MIX16
    const float2 f = Pack16<OpT>::unpack2(w[i]);
    x[2 * i] += fminf(f.x, 10.f * f.x);
    x[2 * i + 1] += fminf(f.y, 10.f * f.y);
#else
    add_inv_lrelu2<OpT>(w[i], x[2 * i], x[2 * i + 1]);      // x += min(y, 10 y), 2.5 instructions per element (tc_common.cuh)
#endif
  }
}

constexpr int kMrfItems = 4;       // 16-byte items per thread: 3 x kMrfItems independent loads in flight per thread
template <typename OpT>
__global__ void __launch_bounds__(256)
mrf_sum_kernel(const OpT* __restrict__ y0, const OpT* __restrict__ y1, const OpT* __restrict__ y2, int C, int L,
               const int* __restrict__ len, int len_mul, float scale, float slope, OpT* __restrict__ out, int split,
               float* __restrict__ out_f32) {
  const int b = blockIdx.y;
  const int valid = len ? len[b] * len_mul : L;
  const int rows = valid + kHaloRows < L ? valid + kHaloRows : L;      // rows nobody reads are skipped (as every producer does)
  const unsigned C8 = (unsigned)C >> 3;                                 // items of one batch row are indexed in 32 bits (checked by the launcher)
  const unsigned n_valid = (unsigned)valid * C8, n_rows = (unsigned)rows * C8;
  const unsigned i0 = blockIdx.x * (256u * kMrfItems) + threadIdx.x;
  if (i0 >= n_rows) return;
  const long long base = (long long)b * L * C;
  uint4 a[kMrfItems][3];
#pragma unroll
  for (int u = 0; u < kMrfItems; ++u) {
    const unsigned i = i0 + u * 256u;
    a[u][0] = a[u][1] = a[u][2] = make_uint4(0u, 0u, 0u, 0u);
    if (i < n_valid) {                    // item i = 8 consecutive channels of one row; rows are contiguous, so e = base + 8 i
      const long long e = base + (long long)i * 8;
      a[u][0] = __ldg(reinterpret_cast<const uint4*>(y0 + e));
      if (y1) a[u][1] = __ldg(reinterpret_cast<const uint4*>(y1 + e));
      if (y2) a[u][2] = __ldg(reinterpret_cast<const uint4*>(y2 + e));
    }
  }
#pragma unroll
  for (int u = 0; u < kMrfItems; ++u) {
    const unsigned i = i0 + u * 256u;
    if (i >= n_rows) break;
    float x[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    y_accum8<OpT>(a[u][0], x); y_accum8<OpT>(a[u][1], x); y_accum8<OpT>(a[u][2], x);
    if (out_f32) {                          // fp32 average itself (input of the generic conv_post)
      float4* o = reinterpret_cast<float4*>(out_f32 + base + (long long)i * 8);
      o[0] = make_float4(x[0] * scale, x[1] * scale, x[2] * scale, x[3] * scale);
      o[1] = make_float4(x[4] * scale, x[5] * scale, x[6] * scale, x[7] * scale);
      continue;
    }
    uint32_t p[4], q[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      float v0 = x[2 * k] * scale, v1 = x[2 * k + 1] * scale;
      v0 = v0 >= 0.f ? v0 : slope * v0; v1 = v1 >= 0.f ? v1 : slope * v1;
      p[k] = Pack16<OpT>::pack2(v0, v1);
      const float2 hi = Pack16<OpT>::unpack2(p[k]);
      q[k] = Pack16<OpT>::pack2(v0 - hi.x, v1 - hi.y);
    }
    const unsigned t = i / C8, c = (i - t * C8) * 8u;
    OpT* o = out + ((long long)b * L + t) * (split ? 2 * C : C) + c;
    *reinterpret_cast<uint4*>(o) = make_uint4(p[0], p[1], p[2], p[3]);
    if (split) *reinterpret_cast<uint4*>(o + C) = make_uint4(q[0], q[1], q[2], q[3]);
  }
}

cudaError_t launch_mrf_sum(const void* y0, const void* y1, const void* y2, int B, int L, int C, const int* len, int len_mul,
                           float scale, float slope, void* out, int split, float* out_f32, int op, cudaStream_t stream) {
  const long long per_b = (long long)L * (C / 8);
  if (op == OP_F32 || C % 8 != 0 || !y0 || (!out && !out_f32) || per_b + 256 * kMrfItems >= (1ll << 32)) return cudaErrorNotSupported;
  dim3 grid((unsigned)((per_b + 256 * kMrfItems - 1) / (256 * kMrfItems)), B);
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
// `in` is the fp32 MRF average (fp32 build and the layer-at-a-time plans; the 16-bit fused plan uses
// conv_post_rows_kernel below, which forms the average itself).
template <int C, int K>
__global__ void __launch_bounds__(kPost4Threads)
conv_post4_kernel(const float* __restrict__ in, int L, float slope,
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
  {
    const float4* src = reinterpret_cast<const float4*>(in + (long long)b * L * C);
    for (int i = threadIdx.x; i < ROWS * C4; i += kPost4Threads) {
      const int r = i / C4, c4 = i - r * C4, tt = t0 + r - PAD;
      float4 x = make_float4(0.f, 0.f, 0.f, 0.f);
      if (tt >= 0 && tt < L) x = __ldg(src + (long long)tt * C4 + c4);
      x.x = x.x >= 0.f ? x.x : slope * x.x; x.y = x.y >= 0.f ? x.y : slope * x.y;
      x.z = x.z >= 0.f ? x.z : slope * x.z; x.w = x.w >= 0.f ? x.w : slope * x.w;
      tile[r * C4 + ((c4 ^ (r >> 2)) & 7)] = x;
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
    conv_post4_kernel<32, 7><<<grid4, kPost4Threads, 0, stream>>>(in, L, slope, w, len, len_mul, wav, pcm, volume);
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

bool conv_post_y_supported(int C, int k, int L) { return C == 32 && k == 7 && L >= 1; }

// Last stage on the 16-bit paths (base shape: C = 32, k = 7): the MRF average is formed from the branches' final y
// streams and conv_post runs on it without shared memory.  A thread owns one ROW (time position): it sums the three
// streams' 32 channels (x_j = min(y, 10 y)), applies leaky_relu and multiplies the row into its 7 per-tap partial dot
// products d_j = sum_c w[j][c] x[c] (weights: broadcast 16-byte reads of a 896-byte shared-memory copy); output
// t = sum_j d_j(row t + j - 3) then takes 12 warp shuffles.  A warp walks a run of 32 * kPrChunks - 6 outputs in chunks of 32 rows
// (6 rows of overlap per run), carrying the previous chunk's partial dots, with the next chunk's 12 x 16-byte loads in
// flight under the current chunk's math.  The tiled kernel it replaces (conv_post4_kernel with 16-bit input: fp32 tile
// in shared memory, 64-thread blocks, 12 warps per SM, four dependent load passes per tile) reached 3.3 TB/s.
#ifndef VITSB_PR_CHUNKS
#define VITSB_PR_CHUNKS 6              // 4-6 chunks per run: finer blocks fill the last wave (8: +2.5 %, 12: +10 %)
#endif
constexpr int kPrChunks = VITSB_PR_CHUNKS, kPrRun = 32 * kPrChunks - 6, kPrWarps = 4;
struct PrCtx {
  int out0, out_end, valid, row_hi, lane, zmask;
  long long boff, ooff;
  float scale, slope, volume;
  float* wav; short* pcm;
  const float4* ws;
};
// rows of chunk c: lane i holds row out0 - 3 + 32 c + i of the three streams (zeros outside [0, row_hi))
template <typename YT>
__device__ __forceinline__ void pr_load(const PrCtx& x, const YT* y0, const YT* y1, const YT* y2, int c, uint4 (&a)[3][4]) {
  const int r = x.out0 - 3 + 32 * c + x.lane;
  const bool ok = r >= 0 && r < x.row_hi;
  const long long e = x.boff + (long long)r * 32;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    a[0][k] = ok ? __ldg(reinterpret_cast<const uint4*>(y0 + e) + k) : make_uint4(0u, 0u, 0u, 0u);
    a[1][k] = (ok && y1) ? __ldg(reinterpret_cast<const uint4*>(y1 + e) + k) : make_uint4(0u, 0u, 0u, 0u);
    a[2][k] = (ok && y2) ? __ldg(reinterpret_cast<const uint4*>(y2 + e) + k) : make_uint4(0u, 0u, 0u, 0u);
  }
}
template <typename YT>
__device__ __forceinline__ void pr_step(const PrCtx& cx, int c, const uint4 (&a)[3][4], float (&dprev)[7]) {
  float d[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  float chain = dprev[6];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    // The 14 weight reads of an 8-channel group take their address from the previous group's result (`zmask` is a
    // kernel argument that is always 0): left free, ptxas issues all 56 16-byte weight loads of a chunk up front,
    // which pins 224 registers next to the two 48-register row buffers and spills.
    const float4* wz = cx.ws + (__float_as_int(chain) & cx.zmask);
    float x[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    y_accum8<YT>(a[0][k], x); y_accum8<YT>(a[1][k], x); y_accum8<YT>(a[2][k], x);
#pragma unroll
    for (int q = 0; q < 8; ++q) { const float v = x[q] * cx.scale; x[q] = v >= 0.f ? v : cx.slope * v; }
#pragma unroll
    for (int j = 0; j < 7; ++j) {
      const float4 w0 = wz[j * 8 + 2 * k], w1 = wz[j * 8 + 2 * k + 1];
      float t = d[j];
      t = fmaf(w0.x, x[0], t); t = fmaf(w0.y, x[1], t); t = fmaf(w0.z, x[2], t); t = fmaf(w0.w, x[3], t);
      t = fmaf(w1.x, x[4], t); t = fmaf(w1.y, x[5], t); t = fmaf(w1.z, x[6], t); t = fmaf(w1.w, x[7], t);
      d[j] = t;
    }
    chain = d[6];
  }
  // output t = out0 - 6 + 32 c + lane = sum_j d_j of local row (lane - 6 + j): this chunk's lane, or the previous chunk's
  float acc = d[6];
#pragma unroll
  for (int j = 0; j < 6; ++j) {
    const int sh = 6 - j, src = (cx.lane - sh) & 31;
    const float cur = __shfl_sync(0xffffffffu, d[j], src), prv = __shfl_sync(0xffffffffu, dprev[j], src);
    acc += cx.lane >= sh ? cur : prv;
  }
#pragma unroll
  for (int j = 0; j < 7; ++j) dprev[j] = d[j];
  const int t = cx.out0 - 6 + 32 * c + cx.lane;
  if (t >= cx.out0 && t < cx.out_end) {
    const float o = t < cx.valid ? tanhf(acc) : 0.f;
    if (cx.pcm) cx.pcm[cx.ooff + t] = to_pcm16(o, cx.volume); else cx.wav[cx.ooff + t] = o;
  }
}

// 3 blocks per SM (168 registers, no spills); 4 (128 registers, 116 B of spills) measured 4 % slower
template <typename YT>
__global__ void __launch_bounds__(kPrWarps * 32, 3)
conv_post_rows_kernel(const YT* __restrict__ y0, const YT* __restrict__ y1, const YT* __restrict__ y2, float scale, int L, float slope,
                      const float* __restrict__ w, const int* __restrict__ len, int len_mul, float* __restrict__ wav,
                      short* __restrict__ pcm, float volume, int zmask) {
  __shared__ float4 ws[7 * 8];                           // [tap][channel / 4]
  if (threadIdx.x < 7 * 8) ws[threadIdx.x] = __ldg(reinterpret_cast<const float4*>(w) + threadIdx.x);
  __syncthreads();
  const int b = blockIdx.y, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  PrCtx cx;
  cx.valid = len ? len[b] * len_mul : L;
  cx.out0 = (blockIdx.x * kPrWarps + warp) * kPrRun;
  if (cx.out0 >= L) return;
  cx.out_end = cx.out0 + kPrRun < L ? cx.out0 + kPrRun : L;
  cx.ooff = (long long)b * L;
  if (cx.out0 >= cx.valid) {                             // fully masked run: zeros
    for (int t = cx.out0 + lane; t < cx.out_end; t += 32) { if (pcm) pcm[cx.ooff + t] = 0; else wav[cx.ooff + t] = 0.f; }
    return;
  }
  cx.boff = (long long)b * L * 32;
  cx.row_hi = cx.out_end + 3 < cx.valid ? cx.out_end + 3 : cx.valid;      // rows at or beyond it are zero or not needed
  cx.lane = lane; cx.scale = scale; cx.slope = slope; cx.volume = volume; cx.wav = wav; cx.pcm = pcm; cx.ws = ws; cx.zmask = zmask;
  float dprev[7] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
  uint4 bufA[3][4], bufB[3][4];
  pr_load<YT>(cx, y0, y1, y2, 0, bufA);
#pragma unroll 1
  for (int c = 0; c < kPrChunks; c += 2) {
    if (cx.out0 - 6 + 32 * c >= cx.out_end) break;       // nothing left to emit (short last run)
    pr_load<YT>(cx, y0, y1, y2, c + 1, bufB);
    pr_step<YT>(cx, c, bufA, dprev);
    if (c + 2 < kPrChunks) pr_load<YT>(cx, y0, y1, y2, c + 2, bufA);
    pr_step<YT>(cx, c + 1, bufB, dprev);
  }
}

cudaError_t launch_conv_post_y(const void* y0, const void* y1, const void* y2, float scale, int B, int L, int C, int k, float slope,
                               const float* w, const int* len, int len_mul, float* wav, short* pcm, float volume, int op,
                               cudaStream_t stream) {
  if (op == OP_F32 || !conv_post_y_supported(C, k, L) || !y0) return cudaErrorNotSupported;
  dim3 grid(ceil_div(L, kPrRun * kPrWarps), B);
  if (op == OP_F16)
    conv_post_rows_kernel<__half><<<grid, kPrWarps * 32, 0, stream>>>((const __half*)y0, (const __half*)y1, (const __half*)y2, scale, L, slope,
                                                                      w, len, len_mul, wav, pcm, volume, 0);
  else
    conv_post_rows_kernel<__nv_bfloat16><<<grid, kPrWarps * 32, 0, stream>>>((const __nv_bfloat16*)y0, (const __nv_bfloat16*)y1,
                                                                             (const __nv_bfloat16*)y2, scale, L, slope, w, len, len_mul, wav, pcm, volume, 0);
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
