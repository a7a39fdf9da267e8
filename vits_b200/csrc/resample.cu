// Sample-rate conversion of the synthesized waveform on the GPU (the `pitch` / `sampling_rate` step of
// VITSWrap.speaking, vits_wrap.py:186-194).  The reference calls librosa.resample with its default resampler (soxr_hq in
// current librosa, kaiser_best before 0.10 - requirements.txt does not pin it, and neither library is in this image), so
// there is no reference output to match bit for bit; this is the published polyphase method of scipy.signal.resample_poly
// (upfirdn with a Kaiser-windowed sinc low-pass, beta 5, 20 * max(up, down) + 1 taps), against which it is tested.
//
//   y[n] = sum_k taps[(n + n_pre_remove) * down - k * up] * x[k]        0 <= index < n_taps, 0 <= k < n_in
//
// `taps` is the zero-padded, up-scaled filter the host plan computes once per (up, down) in float64; one thread per output
// sample accumulates its ~20 * max(up, down) / up products in float64 (HBM / L2 bound: a waveform is a few MB).
// pcm != nullptr: the serving path's int16 conversion (vits_wrap.py:195) is applied to the resampled value instead.
#include "common.h"
#include "kernels.h"

namespace vitsb {
namespace {

__device__ __forceinline__ short rs_to_pcm16(float w, float volume) {      // np.clip(w * volume * 32767, -32768, 32767).astype(int16)
  float v = __fmul_rn(__fmul_rn(w, volume), 32767.f);
  v = fminf(fmaxf(v, -32768.f), 32767.f);
  return (short)__float2int_rz(v);
}

__global__ void __launch_bounds__(256)
resample_poly_kernel(const float* __restrict__ x, long long n_in, long long x_stride, int up, int down,
                     const double* __restrict__ taps, int n_taps, long long n_pre_remove, long long n_out,
                     float* __restrict__ y, short* __restrict__ pcm, float volume, long long y_stride) {
  const long long n = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (n >= n_out) return;
  const int b = blockIdx.y;
  const long long t = (n + n_pre_remove) * down;               // position on the up-sampled grid
  long long k_hi = t / up;
  if (k_hi > n_in - 1) k_hi = n_in - 1;
  const long long first = t - (n_taps - 1);                    // taps index < n_taps  <=>  k * up >= first
  const long long k_lo = first <= 0 ? 0 : (first + up - 1) / up;
  const float* xb = x + (long long)b * x_stride;
  double acc = 0.0;
  for (long long k = k_lo; k <= k_hi; ++k) acc = fma(taps[t - k * up], (double)__ldg(xb + k), acc);
  const float o = (float)acc;
  if (pcm) pcm[(long long)b * y_stride + n] = rs_to_pcm16(o, volume);
  else y[(long long)b * y_stride + n] = o;
}

}  // namespace

cudaError_t launch_resample_poly(const float* x, int B, long long n_in, long long x_stride, int up, int down, const double* taps,
                                 int n_taps, long long n_pre_remove, long long n_out, float* y, short* pcm, float volume,
                                 long long y_stride, cudaStream_t stream) {
  if (n_out <= 0) return cudaSuccess;
  dim3 grid((unsigned)((n_out + 255) / 256), B);
  resample_poly_kernel<<<grid, 256, 0, stream>>>(x, n_in, x_stride, up, down, taps, n_taps, n_pre_remove, n_out, y, pcm, volume, y_stride);
  return cudaGetLastError();
}

}  // namespace vitsb
