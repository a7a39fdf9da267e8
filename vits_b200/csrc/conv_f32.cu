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

constexpr int FT_K = 16;

__device__ __forceinline__ int packed_col_for_gate(int gc, int NT, int which) {
  // gate channel gc -> packed column of its tanh (which=0) / sigmoid (which=1) pre-activation
  int half = NT >> 1;
  return (gc / half) * NT + (gc % half) + which * half;
}

// TM x TN output tile per CTA (64 x 64: 4 x 4 outputs per thread; 32 x 32: 2 x 2 - four times the CTAs, for the small
// problems of the text encoder where a 64 x 64 grid leaves most SMs idle).  K slices stream through a 4-stage cp.async
// ring, so the global-load latency is off the per-slice critical path (one barrier per slice).
template <int TM, int TN>
__global__ void __launch_bounds__(256)
conv_f32_kernel(const ConvArgs a) {
  constexpr int RM = TM / 16, RN = TN / 16;       // outputs per thread
  constexpr int GH = TN / 2;                      // gate channels per CTA (tanh half | sigmoid half of the smem columns)
  const Epilogue& ep = a.epi;
  const int b = blockIdx.z;
  const int t0 = blockIdx.x * TM;
  const int valid = a.len ? a.len[b] * a.len_mul : a.L;
  if (t0 >= valid + kHaloRows) return;   // nothing downstream reads these rows

  constexpr int NS = 4;                           // cp.async stages in flight
  __shared__ float As[NS][FT_K][TM + 4];
  __shared__ float Bs[NS][FT_K][TN + 4];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;

  // packed weight rows handled by this CTA's TN smem columns
  // linear: n0 + c ; gate: first half = tanh rows of GH gate channels, second half = their sigmoid rows
  const int nblk = blockIdx.y;
  const int n_out_total = ep.gate ? a.N / 2 : a.N;
  auto wrow = [&](int c) -> int {
    if (!ep.gate) { int n = nblk * TN + c; return n < a.N ? n : -1; }
    int gc = nblk * GH + (c % GH);
    if (gc >= n_out_total) return -1;
    return packed_col_for_gate(gc, a.gate_group > 0 ? a.gate_group : a.NT, c / GH);
  };

  float acc[RM][RN];
#pragma unroll
  for (int i = 0; i < RM; ++i)
#pragma unroll
    for (int j = 0; j < RN; ++j) acc[i][j] = 0.f;

  const float* in = reinterpret_cast<const float*>(a.in) + (long long)b * a.in_bstride;
  const float* w = reinterpret_cast<const float*>(a.w);
  const int Ktot = a.k * a.Cin;
  // loader mapping: thread -> (row r = tid/4, 4 consecutive k = (tid%4)*4); rows beyond the tile load nothing.
  // 4-byte cp.async with a transposing destination (smem is [k][row] so that the inner product reads are conflict
  // free); src-size 0 zero-fills (conv padding rows, weight rows beyond N).
  const int lr = tid >> 2, lk = (tid & 3) * 4;
  const bool a_thr = lr < TM, b_thr = lr < TN;
  const int my_wrow = b_thr ? wrow(lr) : -1;
  const int kslices = a.Cin / FT_K, nsl = a.k * kslices;
  auto cp4 = [](float* dst, const float* src, bool ok) {
    const unsigned d = (unsigned)__cvta_generic_to_shared(dst);
    asm volatile("cp.async.ca.shared.global [%0], [%1], 4, %2;" ::"r"(d), "l"(src), "r"(ok ? 4 : 0) : "memory");
  };
  auto issue = [&](int s) {
    if (s < nsl) {
      const int slot = s % NS;
      const int j = s / kslices, c0 = (s - j * kslices) * FT_K;
      if (a_thr) {
        const int trow = t0 + lr + j * a.dil - a.pad;
        const bool ok = trow >= 0 && trow < a.L;
        const float* src = ok ? in + (long long)trow * a.in_ld + (c0 % a.in_cols) + lk : in;
#pragma unroll
        for (int q = 0; q < 4; ++q) cp4(&As[slot][lk + q][lr], src + (ok ? q : 0), ok);
      }
      if (b_thr) {
        const bool ok = my_wrow >= 0;
        const float* src = ok ? w + (long long)my_wrow * Ktot + j * a.Cin + c0 + lk : w;
#pragma unroll
        for (int q = 0; q < 4; ++q) cp4(&Bs[slot][lk + q][lr], src + (ok ? q : 0), ok);
      }
    }
    asm volatile("cp.async.commit_group;" ::: "memory");       // one group per slice, empty past the end
  };
#pragma unroll
  for (int s = 0; s < NS - 1; ++s) issue(s);
  for (int s = 0; s < nsl; ++s) {
    asm volatile("cp.async.wait_group %0;" ::"n"(NS - 2) : "memory");   // slice s has landed
    __syncthreads();                                            // ... for every thread; slot (s - 1) % NS is free again
    issue(s + NS - 1);
    const int slot = s % NS;
#pragma unroll
    for (int kk = 0; kk < FT_K; ++kk) {
      float ar[RM], br[RN];
#pragma unroll
      for (int i = 0; i < RM; ++i) ar[i] = As[slot][kk][ty + 16 * i];
#pragma unroll
      for (int i = 0; i < RN; ++i) br[i] = Bs[slot][kk][tx + 16 * i];
#pragma unroll
      for (int i = 0; i < RM; ++i)
#pragma unroll
        for (int jn = 0; jn < RN; ++jn) acc[i][jn] = fmaf(ar[i], br[jn], acc[i][jn]);
    }
  }

  // ---- fused epilogue ----
  const float* bias = ep.bias + (long long)b * ep.bias_bstride;
#pragma unroll
  for (int i = 0; i < RM; ++i) {
    const int t = t0 + ty + 16 * i;
    if (t >= a.L) continue;
    const bool live = t < valid;
    if (ep.gate) {
#pragma unroll
      for (int h = 0; h < RN / 2; ++h) {
        const int c = tx + 16 * h;                 // smem col of tanh part; sigmoid part at c + GH
        const int gc = nblk * GH + c;
        if (gc >= n_out_total) continue;
        const int GG = a.gate_group > 0 ? a.gate_group : a.NT;
        const int na = packed_col_for_gate(gc, GG, 0), nb = packed_col_for_gate(gc, GG, 1);
        float va = ep.alpha * (acc[i][h] + bias[na]);
        float vb = ep.alpha * (acc[i][h + RN / 2] + bias[nb]);
        float v = live ? gate_fn<true>(va, vb) : 0.f;
        if (ep.out_f32) ep.out_f32[(long long)b * ep.out_f32_bstride + (long long)t * ep.out_f32_ld + gc] = v;
        if (ep.out_op) {
          float o = v >= 0.f ? v : ep.out_slope * v;
          reinterpret_cast<float*>(ep.out_op)[(long long)b * ep.out_op_bstride + (long long)t * ep.out_op_ld + gc] = o;
        }
      }
    } else {
#pragma unroll
      for (int jn = 0; jn < RN; ++jn) {
        const int n = nblk * TN + tx + 16 * jn;
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
  const long long ctas64 = (long long)ceil_div(a.L, 64) * (a.epi.gate ? ceil_div(a.N / 2, 32) : ceil_div(a.N, 64)) * a.B;
  if (ctas64 < 120) {                             // fewer CTAs than SMs: quarter-size tiles
    dim3 grid(ceil_div(a.L, 32), a.epi.gate ? ceil_div(a.N / 2, 16) : ceil_div(a.N, 32), a.B);
    conv_f32_kernel<32, 32><<<grid, 256, 0, stream>>>(a);
  } else {
    dim3 grid(ceil_div(a.L, 64), a.epi.gate ? ceil_div(a.N / 2, 32) : ceil_div(a.N, 64), a.B);
    conv_f32_kernel<64, 64><<<grid, 256, 0, stream>>>(a);
  }
  return cudaGetLastError();
}

}  // namespace vitsb
