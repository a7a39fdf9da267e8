// fp32-verification path: CUDA-core FFMA implicit-GEMM conv1d over time-major
// fp32 activations.  Same packed weight layout, same fused epilogue and same
// masking rules as the tcgen05 path (conv_tc.cu); only the arithmetic differs
// (fp32 operands, fp32 accumulate, libm tanh/exp).  tcgen05 has no fp32 MMA
// (kind::tf32 keeps 10 mantissa bits), hence a separate CUDA-core kernel for
// the <=1e-5 verification build.
//
// Tile: 64 time rows x 64 packed columns per CTA, 256 threads, 4x4 outputs
// per thread, K consumed in (tap, 16-channel) slices staged in shared memory.
#include "common.h"
#include "kernels.h"

namespace vitsb {

constexpr int FT_M = 64, FT_N = 64, FT_K = 16;

__device__ __forceinline__ int packed_col_for_gate(int gc, int NT, int which) {
  // gate channel gc -> packed column of its tanh (which=0) / sigmoid (which=1) pre-activation
  int half = NT >> 1;
  return (gc / half) * NT + (gc % half) + which * half;
}

__global__ void __launch_bounds__(256)
conv_f32_kernel(const ConvArgs a) {
  const Epilogue& ep = a.epi;
  const int b = blockIdx.z;
  const int t0 = blockIdx.x * FT_M;
  const int valid = a.len ? a.len[b] * a.len_mul : a.L;
  if (t0 >= valid + kHaloRows) return;   // nothing downstream reads these rows

  __shared__ float As[FT_K][FT_M + 4];
  __shared__ float Bs[FT_K][FT_N + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;

  // packed weight rows handled by this CTA's 64 smem columns
  // linear: n0 + c ; gate: first 32 = tanh rows of 32 gate channels, last 32 = their sigmoid rows
  const int nblk = blockIdx.y;
  const int n_out_total = ep.gate ? a.N / 2 : a.N;
  auto wrow = [&](int c) -> int {
    if (!ep.gate) { int n = nblk * FT_N + c; return n < a.N ? n : -1; }
    int gc = nblk * 32 + (c & 31);
    if (gc >= n_out_total) return -1;
    return packed_col_for_gate(gc, a.gate_group > 0 ? a.gate_group : a.NT, c >> 5);
  };

  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  const float* in = reinterpret_cast<const float*>(a.in) + (long long)b * a.in_bstride;
  const float* w = reinterpret_cast<const float*>(a.w);
  const int Ktot = a.k * a.Cin;
  // loader mapping: thread -> (row r = tid/4, 4 consecutive k = (tid%4)*4)
  const int lr = tid >> 2, lk = (tid & 3) * 4;
  const int my_wrow = wrow(lr);

  for (int j = 0; j < a.k; ++j) {
    const int trow = t0 + lr + j * a.dil - a.pad;
    const bool row_ok = (trow >= 0) && (trow < a.L);
    for (int c0 = 0; c0 < a.Cin; c0 += FT_K) {
      float4 av = make_float4(0.f, 0.f, 0.f, 0.f), bv = av;
      if (row_ok) av = *reinterpret_cast<const float4*>(in + (long long)trow * a.in_ld + (c0 % a.in_cols) + lk);
      if (my_wrow >= 0) bv = *reinterpret_cast<const float4*>(w + (long long)my_wrow * Ktot + j * a.Cin + c0 + lk);
      __syncthreads();
      As[lk + 0][lr] = av.x; As[lk + 1][lr] = av.y; As[lk + 2][lr] = av.z; As[lk + 3][lr] = av.w;
      Bs[lk + 0][lr] = bv.x; Bs[lk + 1][lr] = bv.y; Bs[lk + 2][lr] = bv.z; Bs[lk + 3][lr] = bv.w;
      __syncthreads();
#pragma unroll
      for (int kk = 0; kk < FT_K; ++kk) {
        float ar[4], br[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) { ar[i] = As[kk][ty + 16 * i]; br[i] = Bs[kk][tx + 16 * i]; }
#pragma unroll
        for (int i = 0; i < 4; ++i)
#pragma unroll
          for (int jn = 0; jn < 4; ++jn) acc[i][jn] = fmaf(ar[i], br[jn], acc[i][jn]);
      }
    }
  }

  // ---- fused epilogue ----
  const float* bias = ep.bias + (long long)b * ep.bias_bstride;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int t = t0 + ty + 16 * i;
    if (t >= a.L) continue;
    const bool live = t < valid;
    if (ep.gate) {
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int c = tx + 16 * h;                 // smem col of tanh part; sigmoid part at c+32
        const int gc = nblk * 32 + c;
        if (gc >= n_out_total) continue;
        const int GG = a.gate_group > 0 ? a.gate_group : a.NT;
        const int na = packed_col_for_gate(gc, GG, 0), nb = packed_col_for_gate(gc, GG, 1);
        float va = ep.alpha * (acc[i][h] + bias[na]);
        float vb = ep.alpha * (acc[i][h + 2] + bias[nb]);
        float v = live ? gate_fn<true>(va, vb) : 0.f;
        if (ep.out_f32) ep.out_f32[(long long)b * ep.out_f32_bstride + (long long)t * ep.out_f32_ld + gc] = v;
        if (ep.out_op) {
          float o = v >= 0.f ? v : ep.out_slope * v;
          reinterpret_cast<float*>(ep.out_op)[(long long)b * ep.out_op_bstride + (long long)t * ep.out_op_ld + gc] = o;
        }
      }
    } else {
#pragma unroll
      for (int jn = 0; jn < 4; ++jn) {
        const int n = nblk * FT_N + tx + 16 * jn;
        if (n >= a.N) continue;
        float v = ep.alpha * (acc[i][jn] + bias[n]);
        if (ep.res) v += ep.beta * ep.res[(long long)b * ep.res_bstride + (long long)t * ep.res_ld + n];
        if (ep.accum) v += ep.accum[(long long)b * ep.accum_bstride + (long long)t * ep.accum_ld + n];
        if (!live) v = 0.f;
        if (ep.out_f32) ep.out_f32[(long long)b * ep.out_f32_bstride + (long long)t * ep.out_f32_ld + n] = v;
        if (ep.out_op) {
          float o = v >= 0.f ? v : ep.out_slope * v;
          reinterpret_cast<float*>(ep.out_op)[(long long)b * ep.out_op_bstride + (long long)t * ep.out_op_ld + n] = o;
        }
      }
    }
  }
}

cudaError_t launch_conv_f32(const ConvArgs& a, cudaStream_t stream) {
  const int ncols = a.epi.gate ? ceil_div(a.N / 2, 32) : ceil_div(a.N, FT_N);
  dim3 grid(ceil_div(a.L, FT_M), ncols, a.B);
  conv_f32_kernel<<<grid, 256, 0, stream>>>(a);
  return cudaGetLastError();
}

}  // namespace vitsb
