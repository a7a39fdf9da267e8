// Launch entry points of every kernel in the library.
#pragma once
#include <cuda.h>

#include <mutex>

#include "common.h"

namespace vitsb {

enum OpType { OP_F16 = 0, OP_BF16 = 1, OP_F32 = 2 };
inline int op_bytes(int op) { return op == OP_F32 ? 4 : 2; }

// Per-handle launch options (vitsb_set_option) and device facts handed to the launchers; nothing here is process-global.
struct LaunchOpts {
  bool pdl = true;         // programmatic dependent launch between consecutive tensor-core kernels: their prologues (and the
                           // coupling kernel's L2 weight prefetch) overlap the previous kernel's tail; -1 % (C2) .. -5 % (B = 1)
  bool pair = true;        // conv_tc (when the N tile allows it) and flow_fused: CTA-pair (cta_group::2) mode
  int num_sms = 148;
  int rb_nslab = 0;        // rb_fused: activation slab count override (0 = automatic)
};

// cudaFuncSetAttribute(MaxDynamicSharedMemorySize) is per DEVICE: remember, per kernel, which devices have it.
struct PerDeviceOnce {
  std::mutex mu;
  unsigned long long done = 0;
  template <typename F>
  cudaError_t run(F f) {
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    std::lock_guard<std::mutex> lk(mu);
    if (dev >= 0 && dev < 64 && ((done >> dev) & 1ull)) return cudaSuccess;
    e = f();
    if (e == cudaSuccess && dev >= 0 && dev < 64) done |= 1ull << dev;
    return e;
  }
};

// conv_f32.cu — CUDA-core fp32 verification path
cudaError_t launch_conv_f32(const ConvArgs& a, cudaStream_t stream);

// conv_tc.cu — tcgen05/TMEM/TMA implicit-GEMM path (fp16 / bf16 operands)
// Returns cudaErrorNotSupported for shapes outside its envelope.
cudaError_t launch_conv_tc(const ConvArgs& a, int op, const LaunchOpts& lo, cudaStream_t stream);
bool conv_tc_supported(int Cin, int in_ld, int N, int NT, int k);
constexpr int kMaxBatchRows = 512;       // batch rows per launch (tile tables live in shared memory); larger batches are chunked by the engine

// rb_fused.cu — fused ResBlock2 iteration (c1 -> gate -> c2 -> residual) as one persistent kernel
struct RbArgs {
  void* y_out;             // op_t [B][L][C] = leaky_relu(x_new, 0.1)
  const float* bias1; long long bias1_bstride;   // per-batch conditioning + conv bias, gate-packed order
  const float* bias2;
  int L, B;
  const int* len; int len_mul;
  long long* trace;        // debug: clock64 timeline of CTA 0 (VITSB_RB_TRACE), normally null
  const void* y_res;       // = RbLaunch::y_in (set by launch_rb_fused): residual rows re-read from global when C >= 64
};
struct RbLaunch {
  const void* y_in;        // op_t [B][L][C] = leaky_relu(x, 0.1)
  const void* w1;          // op_t [C][k*C]    (gate-packed rows)
  const void* w2;          // op_t [C][k*C/2]
  int C, k, dil, B;
  RbArgs a;
};
bool rb_fused_supported(int C, int k, int dil, int B);
cudaError_t launch_rb_fused(const RbLaunch& L, int op, const LaunchOpts& lo, cudaStream_t stream);

// flow_fused.cu — one ResidualCouplingLayer (reverse, mean-only; modules.py:362-375 with WN.infer modules.py:157-182)
// as ONE kernel: pre -> n_layers x [gated k=5 conv -> res 1x1 | (post o skip) 1x1] -> x1 -= m, with the WN residual
// stream h and the coupling mean m resident in TMEM and the gate tile in shared memory.  All weights of a coupling in
// the kernel's own packing (element offsets into the operand-typed weight pool / the fp32 bias pool):
struct FlowFusedWeights {
  size_t w_pre = 0;        // [H][2*FF_XW]            : [Wh | Wl] over the 128-column window of the z row that holds x0
  size_t w_in = 0;         // [n_layers*2H][k*H]      : gate rows in 128-row groups = [tanh 64 | sigmoid 64]
  size_t w_rs = 0;         // [n_layers*(H+half)][2H] : rows [res H | (post o skip) half], K = [Wh | Wl]
  size_t b_cum = 0;        // fp32 [n_layers][H]      : pre.bias + sum_{l' < l} res_skip[l'].bias[:H]
  size_t b_m = 0;          // fp32 [half]             : post.bias + post.weight @ sum_l skip bias
  int cond_off = 0;        // column of in_layers[0]'s conditioning vector in the per-call cond buffer (layer l at + l*2H)
  int xwin = 0;            // first z column of the x0 window (0 or inter - FF_XW)
  int coff = 0;            // first z column of x1 (updated in place)
};
struct FlowFusedArgs {
  float* zf;               // fp32 [B][T][inter] flow state (x1 half updated in place)
  void* zop;               // op_t [B][T][2*inter] = [hi | lo] operand copy of zf
  const void* w;           // operand-typed weight pool
  const float* bias;       // fp32 bias pool
  const float* cond; long long cond_bstride;    // per-call conditioning vectors [B][Ncond]
  FlowFusedWeights fw;
  int n_layers;
  int T, B;
  const int* len;          // device [B] or null
  int m_tiles;             // host-known tile count (0 -> upper bound)
  long long* trace;        // debug build: clock64 timeline of CTA 0 (VITSB_FF_TRACE), normally null
};
bool flow_fused_supported(int inter, int hidden, int kernel, int n_layers, const int* dilation_rates, int n_flows, int op);
int flow_fused_tile_rows(int n_layers);          // output rows per 128-row tile
cudaError_t launch_flow_fused(const FlowFusedArgs& a, int op, const LaunchOpts& lo, cudaStream_t stream);

// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time libcuda dependency)
typedef CUresult (*TmaEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
TmaEncodeFn tma_encode_fn();

// aux_kernels.cu
// [B][C][T] fp32 channel-major -> time-major fp32 (+ operand copy), rows >= len[b] zeroed
cudaError_t launch_pack_time_major(const float* src, int B, int C, int T, const int* len,
                                   float* dst_f32, void* dst_op, int op, int split, cudaStream_t stream);
// time-major fp32 [B][T][C] -> channel-major [B][C][T]
cudaError_t launch_unpack_channel_major(const float* src, int B, int C, int T, float* dst, cudaStream_t stream);
// cond[b][n] = sum_k W[n][k] g[b][k] + bias[n]   (all conditioning Linears of one call, fp32)
cudaError_t launch_cond_gemv(const float* W, const float* bias, const float* g, int B, int Ncond, int gin,
                             float* out, cudaStream_t stream);
// leaky_relu + conv_post + tanh: in = fp32 [B][L][C], w fp32 [k][C], out fp32 [B][L]
// pcm != nullptr: write int16 clip(wav * volume * 32767) there instead of the fp32 waveform (vits_wrap.py:195)
cudaError_t launch_conv_post(const float* in, int B, int L, int C, int k, float slope, const float* w,
                             const int* len, int len_mul, float* wav, short* pcm, float volume, cudaStream_t stream);
bool conv_post_supported(int C, int k);

// resample.cu - polyphase FIR resampler of the output waveform (vits_wrap.py:186-194; scipy.signal.resample_poly's method)
cudaError_t launch_resample_poly(const float* x, int B, long long n_in, long long x_stride, int up, int down, const double* taps,
                                 int n_taps, long long n_pre_remove, long long n_out, float* y, short* pcm, float volume,
                                 long long y_stride, cudaStream_t stream);

// head of infer_p2 (models.py:570-572): z_p[B][C][T] = attn @ m_p^T + noise * (attn @ s_p^T), reference layouts, fp32
cudaError_t launch_length_regulate(const float* attn, const float* m_p, const float* s_p, const float* noise,
                                   int B, int C, int T, int Tx, float* z_p, cudaStream_t stream);
// device-side durations + length regulation straight into the time-major flow state (aux_kernels.cu):
// cum / y_len: device int [B][Tx] / [B] scratch; y_len_out: optional second copy for the caller
cudaError_t launch_durations(const float* logw, const int* x_len, const float* m_p, const float* s_p, const float* noise,
                             long long noise_stride, int B, int C, int Tx, int T, float scale, int* cum, int* y_len, int* y_len_out,
                             float* zf, void* zop, int op, int split, cudaStream_t stream);
// MRF average over the branches' final 16-bit y streams (16-bit paths): next upsampler operand rows `out` (16-bit,
// [hi | lo] when split, leaky_relu(slope) applied) or, when out_f32 is given instead, the fp32 average itself
// (conv_post input of shapes conv_post_y does not cover)
cudaError_t launch_mrf_sum(const void* y0, const void* y1, const void* y2, int B, int L, int C, const int* len, int len_mul,
                           float scale, float slope, void* out, int split, float* out_f32, int op, cudaStream_t stream);
// conv_post whose input is the MRF average of up to three y streams, formed in the row load (base shape only)
bool conv_post_y_supported(int C, int k, int L);
cudaError_t launch_conv_post_y(const void* y0, const void* y1, const void* y2, float scale, int B, int L, int C, int k, float slope,
                               const float* w, const int* len, int len_mul, float* wav, short* pcm, float volume, int op,
                               cudaStream_t stream);

}  // namespace vitsb
