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

// Mixed-precision helpers (PTX ISA 8.6, sm_100: SASS FHADD / FHFMA / HMNMX2): a 16-bit value is converted and added (or
// multiplied-and-added) into an fp32 accumulator by ONE instruction, its half selected in the operand (no unpack).  Used
// where a stored y = op(leaky_relu(x, 0.1)) is turned back into x and added to an fp32 value:
//   x = min(y, 10 y) = y + 9 * min(y, 0)      ->  acc = fma(min(y, 0), 9, acc + y)
// (2.5 instructions per element instead of 4: unpack, 10 y, min, add; both forms are exact up to the fp32 adds).
template <typename OpT> struct Mix16;
template <> struct Mix16<__half> {
  static constexpr uint16_t kNine = 0x4880;
  __device__ static __forceinline__ float add(uint16_t h, float c) { float d; asm("add.rn.f32.f16 %0, %1, %2;" : "=f"(d) : "h"(h), "f"(c)); return d; }
  __device__ static __forceinline__ float fma(uint16_t a, uint16_t b, float c) { float d; asm("fma.rn.f32.f16 %0, %1, %2, %3;" : "=f"(d) : "h"(a), "h"(b), "f"(c)); return d; }
  __device__ static __forceinline__ uint32_t min0(uint32_t y) { uint32_t d; asm("min.f16x2 %0, %1, %2;" : "=r"(d) : "r"(y), "r"(0u)); return d; }
};
template <> struct Mix16<__nv_bfloat16> {
  static constexpr uint16_t kNine = 0x4110;
  __device__ static __forceinline__ float add(uint16_t h, float c) { float d; asm("add.rn.f32.bf16 %0, %1, %2;" : "=f"(d) : "h"(h), "f"(c)); return d; }
  __device__ static __forceinline__ float fma(uint16_t a, uint16_t b, float c) { float d; asm("fma.rn.f32.bf16 %0, %1, %2, %3;" : "=f"(d) : "h"(a), "h"(b), "f"(c)); return d; }
  __device__ static __forceinline__ uint32_t min0(uint32_t y) { uint32_t d; asm("min.bf16x2 %0, %1, %2;" : "=r"(d) : "r"(y), "r"(0u)); return d; }
};
// acc0 += x(lo half of y2), acc1 += x(hi half), x = inverse leaky_relu(0.1) of the stored value
template <typename OpT>
__device__ __forceinline__ void add_inv_lrelu2(uint32_t y2, float& acc0, float& acc1) {
  const uint32_t m = Mix16<OpT>::min0(y2);
  acc0 = Mix16<OpT>::fma((uint16_t)(m & 0xffffu), Mix16<OpT>::kNine, Mix16<OpT>::add((uint16_t)(y2 & 0xffffu), acc0));
  acc1 = Mix16<OpT>::fma((uint16_t)(m >> 16), Mix16<OpT>::kNine, Mix16<OpT>::add((uint16_t)(y2 >> 16), acc1));
}

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

// Swizzle<B,4,3> on a byte address whose buffer base is 1024-aligned: XOR bits [4,4+B) with bits [7,7+B)
template <int ROWBYTES>
__device__ __forceinline__ uint32_t swz_addr(uint32_t addr) {
  constexpr uint32_t mask = ROWBYTES == 128 ? 7u : (ROWBYTES == 64 ? 3u : 1u);
  return addr ^ (((addr >> 7) & mask) << 4);
}
__device__ __forceinline__ void st_shared_v4(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint4 ld_shared_v4(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}
// Prefetching global load that the compiler may not sink to its first use (plain __ldg of read-only data is
// free to move across the memory-clobbering tcgen05 asm statements, which silently undoes a software prefetch).
__device__ __forceinline__ uint4 ldg_nc_v4_pinned(const void* p) {
  uint4 v;
  asm volatile("ld.global.nc.v4.b32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}
// 16 consecutive floats from shared memory (all lanes read the same address: broadcast, conflict-free)
__device__ __forceinline__ void lds16(uint32_t addr, float (&v)[16]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const uint4 x = ld_shared_v4(addr + 16 * i);
    v[4 * i] = __uint_as_float(x.x); v[4 * i + 1] = __uint_as_float(x.y); v[4 * i + 2] = __uint_as_float(x.z); v[4 * i + 3] = __uint_as_float(x.w);
  }
}
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

// ring / buffer cursor without integer division
struct Cursor {
  int idx; uint32_t phase;
  __device__ __forceinline__ void next(int n) { if (++idx == n) { idx = 0; phase ^= 1u; } }
};

// tanh(a) * sigmoid(b) for the G tile (rounded to the 16-bit operand type right after).
//   default: 2 MUFU.TANH, sigmoid(b) = 0.5 + 0.5 tanh(b/2); tanh.approx has ~2^-11 relative
//   error, the same size as the fp16 rounding of the result (parity measured in tests).
//   VITSB_GATE_EXACT: 2x ex2 + rcp, ~1e-6 relative: (1 - e^-2a) / ((1 + e^-2a)(1 + e^-b)),
//   a clamped at -20 and b at -40 so the denominator stays inside the __fdividef range.
__device__ __forceinline__ float tanh_approx(float x) {
  float y;
  asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
// tanh(a) * sigmoid(b) = (1 - e^-2a) / ((1 + e^-2a) (1 + e^-b)): 2 ex2 + 1 rcp (3 MUFU instead of the 4 of gate_fn<false>),
// ~1e-6 relative; a clamped at -20 and b at -40 so the denominator stays inside the __fdividef range
__device__ __forceinline__ float gate_exact3(float a, float b) {
  const float ea = ex2_approx(-2.885390081777927f * fmaxf(a, -20.f));
  const float eb = ex2_approx(-1.4426950408889634f * fmaxf(b, -40.f));
  return __fdividef(1.0f - ea, (1.0f + ea) * (1.0f + eb));
}
__device__ __forceinline__ float gate_fast(float a, float b) {
#ifdef VITSB_GATE_EXACT
  const float ea = ex2_approx(-2.885390081777927f * fmaxf(a, -20.f));
  const float eb = ex2_approx(-1.4426950408889634f * fmaxf(b, -40.f));
  return __fdividef(1.0f - ea, (1.0f + ea) * (1.0f + eb));
#else
  const float th = 0.5f * tanh_approx(a);
  return fmaf(th, tanh_approx(0.5f * b), th);
#endif
}

// Tile table of a ragged batch, built by ONE full warp: prefix[b] = number of `tile_rows`-row tiles before batch
// row b, prefix[B] = total.  Tiles of row b cover rows [0, min(L, valid_b + kHaloRows)).  Lane-parallel length
// loads + a shuffle scan: a single thread walking len[] paid one dependent global-load latency per few rows in
// the prologue of every launch.
__device__ __forceinline__ void build_tile_prefix(int* prefix, int B, const int* len, int len_mul, int L, int tile_rows, int lane,
                                                  int extra_rows = kHaloRows, int* valid_out = nullptr) {
  int carry = 0;
  for (int b0 = 0; b0 < B; b0 += 32) {
    const int b = b0 + lane;
    int n = 0;
    if (b < B) {
      const int valid = len ? len[b] * len_mul : L;
      if (valid_out) valid_out[b] = valid;                                // valid rows per batch row, for the epilogues
      const int rows = valid + extra_rows < L ? valid + extra_rows : L;   // rows nobody reads are skipped
      n = (rows + tile_rows - 1) / tile_rows;
    }
    int inc = n;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int y = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += y;
    }
    if (b < B) prefix[b] = carry + inc - n;
    carry += __shfl_sync(0xffffffffu, inc, 31);
  }
  if (lane == 0) prefix[B] = carry;
}

// One epilogue warp: q = TMEM lane quarter (warp_id % 4), thread = one time row.
// The tile's columns are cut into 16-output units dealt round-robin to the `ngrp` epilogue groups
// (unit u -> group u % ngrp), so one tile's epilogue runs on 4 * ngrp warps at once: these launches have
// one or two tiles per CTA and the last tile's epilogue is exposed, so its latency is what matters.
// Each unit issues its global loads (bias / conditioning, residual, accumulator) next to the TMEM load so
// the latencies overlap.  (Prefetching them ahead of `wait_acc()` - the accumulator-ready wait - needs
// ~170 registers per thread, more than 18 warps leave; more warps won.)
// Must be called by all 32 lanes (tcgen05.ld is .sync.aligned); wait_acc() must end with a
// tcgen05.fence::after_thread_sync.
// Staged stores (optional): when the epilogue only writes unsplit 16-bit operand rows (the upsamplers), each 16-column
// unit is packed into a [128 rows][32 B] block of shared memory and written by ONE TMA store instead of 32-byte pieces
// of 32 different rows per warp instruction.  Two blocks per epilogue group, alternating; thread 0 of the group issues
// the stores (bulk async-groups are per thread) and waits for the block's previous store to have read it.
struct TcStage {
  uint32_t smem = 0;               // this group's staging block: [128 rows][cols] 16-bit, TMA-swizzled (0 = direct global stores)
  const CUtensorMap* tm = nullptr; // (columns, rows, batch) map of the output, box cols x 128 x 1, SWIZZLE_64B (cols 32) / 128B (cols 64)
  int bar_id = 0;                  // named barrier of the group's 128 threads
  int cols = 0;                    // columns per staged block: 32 or 64
};

// One staged block: W output columns of the tile's 128 rows.  Every thread turns its row's W accumulator columns into
// packed 16-bit values first (TMEM loads, bias, alpha, leaky_relu - everything that does not touch the staging block),
// then the group waits until the TMA store that last used the block has read it, writes the block (16-byte chunks at
// the swizzled position the tensor map expects: conflict-free for 8 consecutive rows), and one thread stores it.
// One fence + two named barriers + one TMA store per W columns; the first version staged 16 columns at a time and its
// per-unit barriers / store round trips made the narrow upsamplers latency-bound (1.7 k cycles per unit, ncu).
template <typename OpT, int W, typename WaitFn>
__device__ __forceinline__ void tc_stage_block(const Epilogue& ep, const TcStage& stg, const float* bias, uint32_t trow, int col0,
                                               int row, bool live, bool issuer, bool store_ok, int gcol, int t0, int b,
                                               bool first, WaitFn wait_acc) {
  constexpr int SUB = W / 16;
  // bias values come from global memory (the L1 next to ~200 KB of shared memory is a few KB: an L2 round trip); the
  // first 16 are requested before the accumulator wait, the next 16 under each sub-unit's math
  float bv[2][16];
  load16(bias + col0, bv[0]);
  if (first) wait_acc();
  uint32_t raw[SUB][16];
#pragma unroll
  for (int s = 0; s < SUB; ++s) ptx::tmem_ld16(trow + col0 + 16 * s, raw[s]);
  uint32_t pk[W / 2];
  ptx::tmem_ld_wait();
#pragma unroll
  for (int s = 0; s < SUB; ++s) {
    if (s + 1 < SUB) load16(bias + col0 + 16 * (s + 1), bv[(s + 1) & 1]);
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float x = ep.alpha * (__uint_as_float(raw[s][2 * i]) + bv[s & 1][2 * i]), y = ep.alpha * (__uint_as_float(raw[s][2 * i + 1]) + bv[s & 1][2 * i + 1]);
      x = x >= 0.f ? x : ep.out_slope * x;
      y = y >= 0.f ? y : ep.out_slope * y;
      pk[8 * s + i] = live ? Pack16<OpT>::pack2(x, y) : 0u;
    }
  }
  if (issuer) ptx::bulk_wait_read0();              // the store that last used this block has read it
  asm volatile("bar.sync %0, 128;" ::"r"(stg.bar_id) : "memory");
  constexpr int RB = W * 2;                        // row bytes: 64 (SWIZZLE_64B) or 128 (SWIZZLE_128B)
  const int sw = RB == 128 ? (row & 7) : ((row >> 1) & 3);
#pragma unroll
  for (int c = 0; c < RB / 16; ++c)
    st_shared_v4(stg.smem + row * RB + ((c ^ sw) << 4), pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
  fence_proxy_async_smem();
  asm volatile("bar.sync %0, 128;" ::"r"(stg.bar_id) : "memory");
  if (issuer && store_ok) {                        // rows beyond L are clipped by the map
    ptx::tma_store_3d(stg.tm, stg.smem, gcol, t0, b);
    ptx::bulk_commit();
  }
}

template <typename OpT, typename WaitFn>
__device__ __forceinline__ void tc_epilogue(const ConvArgs& a, const Epilogue& ep, int b, int t0, int ntile, int valid,
                                            uint32_t tmem_acc, int q, int lane, int ugrp, int ngrp, WaitFn wait_acc, int dbg = 0,
                                            TcStage* stg = nullptr) {
#ifdef VITSB_RB_DEBUG
#define TCE_DBG(m) ((dbg & (m)) != 0)      // timing ablations: 8 no global stores, 16 no bias / residual / accumulator loads, 32 no gate math
#else
#define TCE_DBG(m) (false)
#endif
    const int row = q * 32 + lane;
    const int t = t0 + row;
    const bool in_range = t < a.L;
    const bool live = t < valid;
    const uint32_t trow = tmem_acc + (static_cast<uint32_t>(q * 32) << 16);
    const float* bias = ep.bias + (long long)b * ep.bias_bstride + (long long)ntile * a.NT;
    if (stg && stg->smem) {
      // staged output (no gate / residual / accumulate / fp32 copy by construction, make_geom): blocks of stg->cols columns
      // dealt to the groups; the accumulator wait sits inside the first block, after its first bias request
      const bool issuer = q == 0 && lane == 0, store_ok = t0 < a.L;
      const int nblk = a.NT / stg->cols, n0 = ntile * a.NT;
      if (ugrp >= nblk) { wait_acc(); return; }
      bool first = true;
#pragma unroll 1
      for (int k = ugrp; k < nblk; k += ngrp, first = false) {
        const int col0 = k * stg->cols;
        if (stg->cols == 64) tc_stage_block<OpT, 64>(ep, *stg, bias, trow, col0, row, live, issuer, store_ok, n0 + col0, t0, b, first, wait_acc);
        else tc_stage_block<OpT, 32>(ep, *stg, bias, trow, col0, row, live, issuer, store_ok, n0 + col0, t0, b, first, wait_acc);
      }
      return;
    }
    wait_acc();
    if (ep.gate) {
      // packed columns: groups of GG = [tanh GG/2 | sigmoid GG/2]; output channel = packed gate index.
      // unit u = 16 gates: gate group gi = u / upg, slice ci = u % upg (advanced without integer division)
      const int GG = a.gate_group > 0 ? a.gate_group : a.NT, gh = GG >> 1;
      const int upg = gh >> 4, nu = a.NT >> 5;
      const long long orow_op = (long long)b * ep.out_op_bstride + (long long)t * ep.out_op_ld + (long long)ntile * (a.NT >> 1);
      int gi = 0, ci = ugrp;
      while (ci >= upg) { ci -= upg; ++gi; }
#pragma unroll 1
      for (int u = ugrp; u < nu; u += ngrp) {
        const int g0 = gi * GG, c0 = ci << 4;
        uint32_t ra[16], rb[16];
        ptx::tmem_ld16(trow + g0 + c0, ra);
        ptx::tmem_ld16(trow + g0 + gh + c0, rb);
        float ba[16], bb[16], v[16];
        if (!TCE_DBG(16)) { load16(bias + g0 + c0, ba); load16(bias + g0 + gh + c0, bb); }
        else {
#pragma unroll
          for (int i = 0; i < 16; ++i) ba[i] = bb[i] = 0.f;
        }
        ptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) {
          float xa = ep.alpha * (__uint_as_float(ra[i]) + ba[i]);
          float xb = ep.alpha * (__uint_as_float(rb[i]) + bb[i]);
          v[i] = TCE_DBG(32) ? xa + xb : (live ? gate_fn<false>(xa, xb) : 0.f);
        }
        const int oc = (g0 >> 1) + c0;       // output column inside this tile's NT/2 gate channels
        if (in_range && !TCE_DBG(8)) {
          if (ep.out_f32)
            store16(ep.out_f32 + (long long)b * ep.out_f32_bstride + (long long)t * ep.out_f32_ld + (long long)ntile * (a.NT >> 1) + oc, v);
          if (ep.out_op) store_op16<OpT>(reinterpret_cast<OpT*>(ep.out_op) + orow_op + oc, v, ep.out_slope, ep.out_op_split);
        }
        ci += ngrp;
        while (ci >= upg) { ci -= upg; ++gi; }
      }
    } else {
      const long long n0 = (long long)ntile * a.NT;
      const int nu = a.NT >> 4;
      const bool has_res = in_range && ep.res != nullptr, has_acc = in_range && ep.accum != nullptr;
      const float* res = ep.res + (long long)b * ep.res_bstride + (long long)t * ep.res_ld + n0;
      const float* accum = ep.accum + (long long)b * ep.accum_bstride + (long long)t * ep.accum_ld + n0;
#pragma unroll 1
      for (int u = ugrp; u < nu; u += ngrp) {
        const int c0 = u << 4;
        uint32_t r[16];
        ptx::tmem_ld16(trow + c0, r);
        float bv[16], v[16], rx[16], ax[16];
        if (!TCE_DBG(16)) {
          if (has_res) load16(res + c0, rx);
          if (has_acc) load16(accum + c0, ax);
          load16(bias + c0, bv);
        } else {
#pragma unroll
          for (int i = 0; i < 16; ++i) rx[i] = ax[i] = bv[i] = 0.f;
        }
        ptx::tmem_ld_wait();
#pragma unroll
        for (int i = 0; i < 16; ++i) v[i] = ep.alpha * (__uint_as_float(r[i]) + bv[i]);
        if (has_res) {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = fmaf(ep.beta, rx[i], v[i]);
        }
        if (has_acc) {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] += ax[i];
        }
        if (!live) {
#pragma unroll
          for (int i = 0; i < 16; ++i) v[i] = 0.f;
        }
        if (in_range && !TCE_DBG(8)) {
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
