// Launch entry points of every kernel in the library.
#pragma once
#include <cuda.h>

#include "common.h"

namespace vitsb {

enum OpType { OP_F16 = 0, OP_BF16 = 1, OP_F32 = 2 };
inline int op_bytes(int op) { return op == OP_F32 ? 4 : 2; }

// conv_f32.cu — CUDA-core fp32 verification path
cudaError_t launch_conv_f32(const ConvArgs& a, cudaStream_t stream);

// conv_tc.cu — tcgen05/TMEM/TMA implicit-GEMM path (fp16 / bf16 operands)
// Returns cudaErrorNotSupported for shapes outside its envelope.
cudaError_t launch_conv_tc(const ConvArgs& a, int op, cudaStream_t stream);
bool conv_tc_supported(int Cin, int in_ld, int N, int NT, int k);
// rb_fused.cu — fused ResBlock2 iteration (c1 -> gate -> c2 -> residual) as one persistent kernel
enum { RB_ACC_READ = 1, RB_ACC_STORE = 2, RB_EMIT_NEXT = 4 };
struct RbArgs {
  void* y_out;             // op_t [B][L][C] = leaky_relu(x_new, 0.1) (mid iterations) or null
  float* acc;              // fp32 [B][L][C] MRF accumulator (models.py:311-314)
  int acc_mode;            // RB_ACC_READ: v += acc; RB_ACC_STORE: acc = v; RB_EMIT_NEXT: write nxt_*
  void* nxt_op;            // op_t leaky_relu(v, 0.1): [B][L][2C] hi|lo if nxt_split else [B][L][C]
  int nxt_split;
  float* nxt_f32;          // fp32 [B][L][C] = v (conv_post input)
  float scale;             // v = scale * (c2 + bias2 + x + x_add0 + x_add1)
  const void* add0;        // op_t [B][L][C] or null: y = leaky_relu(x_j, 0.1) streams of the other MRF branches, summed in
  const void* add1;        //   (x_j = min(y, 10 y)): 2 B per element instead of a 4 B fp32 accumulator read-modify-write
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
cudaError_t launch_rb_fused(const RbLaunch& L, int op, cudaStream_t stream);
// cuTensorMapEncodeTiled through the runtime's driver entry point (no link-time libcuda dependency)
typedef CUresult (*TmaEncodeFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
TmaEncodeFn tma_encode_fn();
// programmatic dependent launch between consecutive tensor-core kernels (VITSB_PDL=0 disables)
bool pdl_enabled();

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

// head of infer_p2 (models.py:570-572): z_p[B][C][T] = attn @ m_p^T + noise * (attn @ s_p^T), reference layouts, fp32
cudaError_t launch_length_regulate(const float* attn, const float* m_p, const float* s_p, const float* noise,
                                   int B, int C, int T, int Tx, float* z_p, cudaStream_t stream);
// MRF average over the branches' final 16-bit y streams -> next upsampler operand rows (16-bit paths)
cudaError_t launch_mrf_sum(const void* y0, const void* y1, const void* y2, int B, int L, int C, const int* len, int len_mul,
                           float scale, float slope, void* out, int split, int op, cudaStream_t stream);
// conv_post whose input is the MRF average of up to three y streams, formed in the tile load (base shape only)
bool conv_post_y_supported(int C, int k, int L);
cudaError_t launch_conv_post_y(const void* y0, const void* y1, const void* y2, float scale, int B, int L, int C, int k, float slope,
                               const float* w, const int* len, int len_mul, float* wav, short* pcm, float volume, int op,
                               cudaStream_t stream);

}  // namespace vitsb
