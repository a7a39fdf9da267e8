// Shared device code of the tcgen05 conv kernels: operand packing helpers and
// the fused epilogue (TMEM -> registers -> bias/cond/residual/accumulate/gate/
// leaky-relu -> fp32 and operand-typed global stores, masked beyond len[b]).
#pragma once
#include "common.h"
#include "ptx.h"

namespace vitsb {

template <typename OpT> struct Pack16;   // 16 floats -> 16 OpT (32 bytes) as two uint4
template <> struct Pack16<__half> {
  __device__ static __forceinline__ uint32_t pack2(float a, float b) {
    __half2 h = __floats2half2_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
  __device__ static __forceinline__ float2 unpack2(uint32_t u) { return __half22float2(*reinterpret_cast<__half2*>(&u)); }
};
template <> struct Pack16<__nv_bfloat16> {
  __device__ static __forceinline__ uint32_t pack2(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
  }
  __device__ static __forceinline__ float2 unpack2(uint32_t u) { return __bfloat1622float2(*reinterpret_cast<__nv_bfloat162*>(&u)); }
};

template <typename OpT>
__device__ __forceinline__ void store_op16(OpT* dst, const float (&v)[16], float slope, int split) {
  uint32_t p[8], q[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    float x = v[2 * i], y = v[2 * i + 1];
    x = x >= 0.f ? x : slope * x;
    y = y >= 0.f ? y : slope * y;
    p[i] = Pack16<OpT>::pack2(x, y);
    const float2 hi = Pack16<OpT>::unpack2(p[i]);
    q[i] = Pack16<OpT>::pack2(x - hi.x, y - hi.y);
  }
  uint4* d = reinterpret_cast<uint4*>(dst);
  d[0] = make_uint4(p[0], p[1], p[2], p[3]);
  d[1] = make_uint4(p[4], p[5], p[6], p[7]);
  if (split > 0) {
    uint4* l = reinterpret_cast<uint4*>(dst + split);
    l[0] = make_uint4(q[0], q[1], q[2], q[3]);
    l[1] = make_uint4(q[4], q[5], q[6], q[7]);
  }
}

__device__ __forceinline__ void load16(const float* src, float (&v)[16]) {
  const float4* s = reinterpret_cast<const float4*>(src);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    float4 x = s[i];
    v[4 * i] = x.x; v[4 * i + 1] = x.y; v[4 * i + 2] = x.z; v[4 * i + 3] = x.w;
  }
}
__device__ __forceinline__ void store16(float* dst, const float (&v)[16]) {
  float4* d = reinterpret_cast<float4*>(dst);
#pragma unroll
  for (int i = 0; i < 4; ++i) d[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
}

// One epilogue warp: q = TMEM lane quarter (warp_id % 4), thread = one time row.
// Must be called by all 32 lanes (tcgen05.ld is .sync.aligned) after the
// accumulator-ready barrier and a tcgen05.fence::after_thread_sync.
template <typename OpT>
__device__ __forceinline__ void tc_epilogue(const ConvArgs& a, const Epilogue& ep, int b, int t0, int ntile, int valid,
                                            uint32_t tmem_acc, int q, int lane) {
    const int row = q * 32 + lane;
    const int t = t0 + row;
    const bool in_range = t < a.L;
    const bool live = t < valid;
    const uint32_t trow = tmem_acc + (static_cast<uint32_t>(q * 32) << 16);
    const float* bias = ep.bias + (long long)b * ep.bias_bstride + (long long)ntile * a.NT;
    if (ep.gate) {
      // packed columns: groups of GG = [tanh GG/2 | sigmoid GG/2]; output channel = packed gate index
      const int GG = a.gate_group > 0 ? a.gate_group : a.NT, gh = GG >> 1;
      const long long orow_op = (long long)b * ep.out_op_bstride + (long long)t * ep.out_op_ld + (long long)ntile * (a.NT >> 1);
      for (int g0 = 0; g0 < a.NT; g0 += GG) {
        for (int c0 = 0; c0 < gh; c0 += 16) {
          uint32_t ra[16], rb[16];
          ptx::tmem_ld16(trow + g0 + c0, ra);
          ptx::tmem_ld16(trow + g0 + gh + c0, rb);
          ptx::tmem_ld_wait();
          float ba[16], bb[16], v[16];
          load16(bias + g0 + c0, ba);
          load16(bias + g0 + gh + c0, bb);
#pragma unroll
          for (int i = 0; i < 16; ++i) {
            float xa = ep.alpha * (__uint_as_float(ra[i]) + ba[i]);
            float xb = ep.alpha * (__uint_as_float(rb[i]) + bb[i]);
            v[i] = live ? gate_fn<false>(xa, xb) : 0.f;
          }
          const int oc = (g0 >> 1) + c0;       // output column inside this tile's NT/2 gate channels
          if (in_range) {
            if (ep.out_f32)
              store16(ep.out_f32 + (long long)b * ep.out_f32_bstride + (long long)t * ep.out_f32_ld + (long long)ntile * (a.NT >> 1) + oc, v);
            if (ep.out_op) store_op16<OpT>(reinterpret_cast<OpT*>(ep.out_op) + orow_op + oc, v, ep.out_slope, ep.out_op_split);
          }
        }
      }
    } else {
      const long long n0 = (long long)ntile * a.NT;
      for (int c0 = 0; c0 < a.NT; c0 += 16) {
        uint32_t r[16];
        ptx::tmem_ld16(trow + c0, r);
        ptx::tmem_ld_wait();
        float bv[16], v[16];
        load16(bias + c0, bv);
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = ep.alpha * (__uint_as_float(r[i]) + bv[i]);
        if (in_range) {
          if (ep.res) {
            float x[16];
            load16(ep.res + (long long)b * ep.res_bstride + (long long)t * ep.res_ld + n0 + c0, x);
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = fmaf(ep.beta, x[i], v[i]);
          }
          if (ep.accum) {
            float x[16];
            load16(ep.accum + (long long)b * ep.accum_bstride + (long long)t * ep.accum_ld + n0 + c0, x);
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] += x[i];
          }
          if (!live) {
#pragma unroll
            for (int i = 0; i < 16; ++i) v[i] = 0.f;
          }
          if (ep.out_f32)
            store16(ep.out_f32 + (long long)b * ep.out_f32_bstride + (long long)t * ep.out_f32_ld + n0 + c0, v);
          if (ep.out_op)
            store_op16<OpT>(reinterpret_cast<OpT*>(ep.out_op) + (long long)b * ep.out_op_bstride +
                                (long long)t * ep.out_op_ld + n0 + c0, v, ep.out_slope, ep.out_op_split);
        }
      }
    }
}

}  // namespace vitsb
