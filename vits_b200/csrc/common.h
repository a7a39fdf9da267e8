// Shared host/device definitions for the vitsb200 kernels.
//
// Data layout in HBM: every activation is TIME-MAJOR [B][L][C] (C contiguous),
// so a conv tap is a row shift and the implicit GEMM is
//   M = time positions, N = output channels, K = (tap, Cin)   (tap-major, Cin-minor).
// "op" tensors hold the GEMM operand type (fp16 / bf16 on tcgen05, fp32 on the
// verification path); residual streams and accumulators are always fp32.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace vitsb {

constexpr int kTileM = 128;      // time rows per CTA tile (tcgen05 M)
constexpr int kMaxTileN = 256;   // accumulator columns per CTA tile (tcgen05 N)
constexpr int kHaloRows = 32;    // >= largest one-sided receptive field of any single conv (k=11,d=5 -> 25)

// Fused epilogue, applied per output element (b, t, n):
//   v = alpha * (acc + bias[b*bias_bstride + n]) + beta * res[b,t,n] + accum[b,t,n]
//   gate: v = tanh(v_a) * sigmoid(v_b), (a, b) = packed columns (c, c + NT/2) of one N tile
//   rows t >= len[b]*len_mul are stored as zero (masked / ragged semantics)
//   out_f32[b,t,n] = v ; out_op[b,t,n] = op_t( v >= 0 ? v : out_slope * v )
//
// Split precision (16-bit paths only).  The convs whose operand rounding dominates
// the waveform error (the whole flow, conv_pre, the upsamplers)
// run as x*W ~= xh*Wh + xl*Wh + xh*Wl with x = xh + xl, W = Wh + Wl in the operand
// type: the activation row is stored as [hi(C) | lo(C)], the packed weights per tap
// as [Wh | Wh | Wl], and the K loop reads activation columns modulo 2C, i.e.
// (hi, lo, hi) against (Wh, Wh, Wl).
struct Epilogue {
  const float* bias;     long long bias_bstride;      // per-batch stride (0 = shared)
  float alpha, beta;
  const float* res;      int res_ld;   long long res_bstride;
  const float* accum;    int accum_ld; long long accum_bstride;
  float* out_f32;        int out_f32_ld; long long out_f32_bstride;
  void*  out_op;         int out_op_ld;  long long out_op_bstride;
  float out_slope;       // leaky-relu slope folded into the operand copy (1.0 = identity)
  int gate;              // 1: gated (tanh*sigmoid) epilogue, output has N/2 channels
  int out_op_split;      // >0: also store lo = op_t(x - float(hi)) at column offset +out_op_split (2-term split operand)
};

struct ConvArgs {
  const void* in;        // op_t [B][L][in_ld]
  long long in_bstride;  // elements
  int in_ld;             // row pitch in elements
  int Cin;               // K per tap (multiple of 16); 3*in_cols/2 for split-precision convs
  int in_cols;           // physical columns of an input row; K block kc reads columns (kc*BK) % in_cols
  const void* w;         // op_t [N][k*Cin]  K-major
  int N;                 // GEMM N (packed output columns)
  int NT;                // N tile (<=256, multiple of 16); the tensor-core launcher may pick a smaller divisor (wave balance)
  int gate_group;        // gated convs: packed columns come in groups of gate_group = [tanh half | sigmoid half]; 0 -> NT
  int m_tiles;           // host-known number of 128-row tiles that will run (ragged lengths); 0 -> B * ceil(L / 128)
  int k, dil, pad;       // tap j reads input row t + j*dil - pad
  int L;                 // rows per batch item
  int B;
  const int* len;        // device [B] valid frames
  int len_mul;           // valid rows = len[b] * len_mul
  Epilogue epi;
  // Two convs that read the same input fused into one launch (WN res_skip, modules.py:174-181):
  // N tiles with first column >= n_split use epi_b, with columns counted from n_split.  0 = single epilogue.
  int n_split;
  Epilogue epi_b;
  // Structural zeros of the phase-stacked transposed conv (3 taps, each output phase uses 2 of them): packed columns
  // < tap_split_n use taps [0, k-1) only, columns >= tap_split_n taps [1, k).  An N tile on one side of the split skips
  // the unused tap (its weights are exact zeros, so the result is bit-identical); 0 = every tap for every column.
  int tap_split_n;
};

__host__ __device__ inline int ceil_div(int a, int b) { return (a + b - 1) / b; }

// ---- operand type helpers -------------------------------------------------
template <typename T> __device__ __forceinline__ T to_op(float v);
template <> __device__ __forceinline__ float to_op<float>(float v) { return v; }
template <> __device__ __forceinline__ __half to_op<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 to_op<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

template <typename T> __device__ __forceinline__ float from_op(T v);
template <> __device__ __forceinline__ float from_op<float>(float v) { return v; }
template <> __device__ __forceinline__ float from_op<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float from_op<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

// Gate nonlinearity.  kAccurate: libm-grade (fp32-verification build);
// otherwise ex2/rcp based, ~1e-6 relative (well below fp16 rounding of the result).
template <bool kAccurate>
__device__ __forceinline__ float gate_fn(float a, float b) {
  if (kAccurate) {
    return tanhf(a) * (1.0f / (1.0f + expf(-b)));
  } else {
    // tanh(a) = 1 - 2/(1+e^{2a});  sigmoid(b) = 1/(1+e^{-b})
    float ea = __expf(2.0f * a);
    float eb = __expf(-b);
    float th = 1.0f - __fdividef(2.0f, 1.0f + ea);
    return __fdividef(th, 1.0f + eb);
  }
}

}  // namespace vitsb
