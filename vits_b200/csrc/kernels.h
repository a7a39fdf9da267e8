// Launch entry points of every kernel in the library.
#pragma once
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
cudaError_t launch_conv_post(const float* in, int B, int L, int C, int k, float slope, const float* w,
                             const int* len, int len_mul, float* wav, cudaStream_t stream);

}  // namespace vitsb
