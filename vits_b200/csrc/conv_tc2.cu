// tcgen05 implicit-GEMM conv1d, "slab" variant: the activation rows a tile needs
// for ALL taps — [t0 - pad, t0 - pad + 128 + (k-1)*dil) — are staged in shared
// memory once per channel block; every tap's A operand is then a row-shifted
// VIEW of that slab, selected purely by the UMMA descriptor's start address.
// L2->SMEM traffic for A drops by ~k compared to one TMA box per tap (conv_tc.cu).
//
// Slab layouts (mode):
//   1  no-swizzle K-major "interleaved": [BK/8 chunks][R rows][16 B].  Core matrix =
//      8 rows x 16 B contiguous (SBO = 128 B), next 16-byte K chunk LBO = R*16 B away.
//      A row shift of s rows is a start-address shift of s*16 B: any s is legal.
//      Filled by BK/8 TMA boxes of 8 channels x R rows.
//   2  128B-swizzled rows [R][128 B] (one TMA box), tap shift = s*128 B with the
//      descriptor base-offset field = (start >> 7) & 7.
//   3  as 2 with base-offset 0.
#include "common.h"
#include "kernels.h"
#include "ptx.h"
#include "tc_common.cuh"

#include <mutex>

namespace vitsb {

struct Tc2Geom {
  int mode;
  int BK, nCB;       // channels per block, number of channel blocks (Cin / BK)
  int R;             // slab rows = 128 + (k-1)*dil
  int SB;            // B ring depth
  uint32_t a_bytes;  // R * BK * 2, rounded to 1024
  uint32_t a_tx;     // exact bytes TMA delivers per slab
  uint32_t b_bytes, b_stride;
  uint32_t tmem_cols, idesc;
};

constexpr int kTc2Threads = 192;
constexpr int kSlabStages = 2;

// no-swizzle K-major descriptor: LBO = byte distance between 16-byte K chunks, SBO = between 8-row groups
__device__ __forceinline__ uint64_t umma_desc_interleaved(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((saddr & 0x3FFFFu) >> 4);
  d |= static_cast<uint64_t>((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= static_cast<uint64_t>((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= 1ull << 46;
  return d;                                   // layout type 0 = SWIZZLE_NONE
}

template <typename OpT>
__global__ void __launch_bounds__(kTc2Threads)
conv_tc2_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                const ConvArgs a, const Tc2Geom g) {
  const int b = blockIdx.z;
  const int t0 = blockIdx.x * kTileM;
  const int ntile = blockIdx.y;
  const int valid = a.len ? a.len[b] * a.len_mul : a.L;
  if (t0 >= valid + kHaloRows) return;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sA = smem_base;
  const uint32_t sB = sA + kSlabStages * g.a_bytes;
  const uint32_t bars = sB + g.SB * g.b_stride;
  auto a_full = [&](int s) { return bars + 8u * s; };
  auto a_empty = [&](int s) { return bars + 8u * (kSlabStages + s); };
  auto b_full = [&](int s) { return bars + 8u * (2 * kSlabStages + s); };
  auto b_empty = [&](int s) { return bars + 8u * (2 * kSlabStages + g.SB + s); };
  const uint32_t accum_bar = bars + 8u * (2 * kSlabStages + 2 * g.SB);
  const uint32_t tmem_slot = accum_bar + 8u;
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - ptx::smem_u32(smem_raw)));

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmB);
    for (int s = 0; s < kSlabStages; ++s) { ptx::mbar_init(a_full(s), 1); ptx::mbar_init(a_empty(s), 1); }
    for (int s = 0; s < g.SB; ++s) { ptx::mbar_init(b_full(s), 1); ptx::mbar_init(b_empty(s), 1); }
    ptx::mbar_init(accum_bar, 1);
    ptx::mbar_fence_init();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, g.tmem_cols);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  if (warp == 0) {
    if (lane == 0) {
      // ===== TMA producer: slab(cb) then its k weight boxes =====
      int itb = 0;
      for (int cb = 0; cb < g.nCB; ++cb) {
        const int sa = cb % kSlabStages;
        ptx::mbar_wait(a_empty(sa), ((cb / kSlabStages) & 1) ^ 1u);
        ptx::mbar_expect_tx(a_full(sa), g.a_tx);
        const int c0 = (cb * g.BK) % a.in_cols;
        const uint32_t dst = sA + sa * g.a_bytes;
        if (g.mode == 1) {
          for (int q = 0; q < g.BK / 8; ++q)
            ptx::tma_load_3d(dst + q * (g.R * 16), &tmA, a_full(sa), c0 + q * 8, t0 - a.pad, b);
        } else {
          ptx::tma_load_3d(dst, &tmA, a_full(sa), c0, t0 - a.pad, b);
        }
        for (int j = 0; j < a.k; ++j, ++itb) {
          const int sb = itb % g.SB;
          ptx::mbar_wait(b_empty(sb), ((itb / g.SB) & 1) ^ 1u);
          ptx::mbar_expect_tx(b_full(sb), g.b_bytes);
          ptx::tma_load_2d(sB + sb * g.b_stride, &tmB, b_full(sb), j * a.Cin + cb * g.BK, ntile * a.NT);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    if (lane == 0) {
      // ===== MMA issuer =====
      const uint32_t swz = g.BK * 2;
      int itb = 0;
      for (int cb = 0; cb < g.nCB; ++cb) {
        const int sa = cb % kSlabStages;
        ptx::mbar_wait(a_full(sa), (cb / kSlabStages) & 1);
        ptx::tc_fence_after();
        const uint32_t abase = sA + sa * g.a_bytes;
        for (int j = 0; j < a.k; ++j, ++itb) {
          const int sb = itb % g.SB;
          ptx::mbar_wait(b_full(sb), (itb / g.SB) & 1);
          ptx::tc_fence_after();
          const uint32_t bbase = sB + sb * g.b_stride;
          const uint32_t shift = static_cast<uint32_t>(j * a.dil);       // tap's row shift inside the slab
          for (int kk = 0; kk < g.BK / 16; ++kk) {
            uint64_t ad;
            if (g.mode == 1) {
              ad = umma_desc_interleaved(abase + (2 * kk) * (g.R * 16) + shift * 16, g.R * 16, 128);
            } else {
              const uint32_t addr = abase + shift * swz + kk * 32;
              ad = ptx::umma_smem_desc(addr, swz, g.mode == 2 ? ((addr >> 7) & 7u) : 0u);
            }
            const uint64_t bd = ptx::umma_smem_desc(bbase + kk * 32, swz, 0);
            ptx::umma_f16(tmem_base, ad, bd, g.idesc, (cb | j | kk) != 0 ? 1u : 0u);
          }
          ptx::umma_commit(b_empty(sb));
        }
        ptx::umma_commit(a_empty(sa));
      }
      ptx::umma_commit(accum_bar);
    }
    __syncwarp();
  } else {
    ptx::mbar_wait(accum_bar, 0);
    ptx::tc_fence_after();
    tc_epilogue<OpT>(a, b, t0, ntile, valid, tmem_base, warp & 3, lane);
  }

  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, g.tmem_cols);
  }
}

cudaError_t launch_conv_tc2(const ConvArgs& a, int op, int mode, cudaStream_t stream) {
  if (op == OP_F32 || !conv_tc_supported(a.Cin, a.in_ld, a.N, a.NT, a.k)) return cudaErrorNotSupported;
  if (a.epi.gate && (a.NT % 32 != 0)) return cudaErrorNotSupported;
  Tc2Geom g{};
  g.mode = mode;
  g.BK = a.Cin >= 64 ? 64 : a.Cin;
  if (a.in_cols <= 0 || a.in_cols % g.BK != 0) return cudaErrorNotSupported;
  g.nCB = a.Cin / g.BK;
  g.R = (kTileM + (a.k - 1) * a.dil + 7) & ~7;   // multiple of 8: every 16-byte-chunk plane stays 128 B aligned for TMA
  if (g.R > 256) return cudaErrorNotSupported;
  g.a_tx = (uint32_t)g.R * g.BK * 2;
  g.a_bytes = (g.a_tx + 1023u) & ~1023u;
  g.b_bytes = a.NT * g.BK * 2;
  g.b_stride = (g.b_bytes + 1023u) & ~1023u;
  int sb = (int)((150u * 1024 - kSlabStages * g.a_bytes) / g.b_stride);
  if (sb > 6) sb = 6;
  if (sb > a.k * g.nCB) sb = a.k * g.nCB;
  if (sb < 1) sb = 1;
  g.SB = sb;
  uint32_t cols = 32;
  while (cols < (uint32_t)a.NT) cols <<= 1;
  g.tmem_cols = cols;
  g.idesc = ptx::umma_idesc_f16(op == OP_BF16, kTileM, a.NT);

  TmaEncodeFn enc = tma_encode_fn();
  if (!enc) return cudaErrorNotSupported;
  const CUtensorMapDataType dt = op == OP_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  const CUtensorMapSwizzle swB = g.BK == 64 ? CU_TENSOR_MAP_SWIZZLE_128B
                                : (g.BK == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
  CUtensorMap tmA, tmB;
  {
    cuuint64_t dims[3] = {(cuuint64_t)a.in_cols, (cuuint64_t)a.L, (cuuint64_t)a.B};
    cuuint64_t strides[2] = {(cuuint64_t)a.in_ld * 2, (cuuint64_t)a.in_bstride * 2};
    cuuint32_t box[3] = {(cuuint32_t)(mode == 1 ? 8 : g.BK), (cuuint32_t)g.R, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(&tmA, dt, 3, const_cast<void*>(a.in), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     mode == 1 ? CU_TENSOR_MAP_SWIZZLE_NONE : swB,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)a.k * a.Cin, (cuuint64_t)a.N};
    cuuint64_t strides[1] = {(cuuint64_t)a.k * a.Cin * 2};
    cuuint32_t box[2] = {(cuuint32_t)g.BK, (cuuint32_t)a.NT};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&tmB, dt, 2, const_cast<void*>(a.w), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swB,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
  }
  const size_t smem = 1024 + (size_t)kSlabStages * g.a_bytes + (size_t)g.SB * g.b_stride + 8 * (2 * kSlabStages + 2 * g.SB + 1) + 16;
  dim3 grid(ceil_div(a.L, kTileM), a.N / a.NT, a.B);
  static std::once_flag once;
  std::call_once(once, [] {
    cudaFuncSetAttribute(conv_tc2_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(conv_tc2_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  });
  if (op == OP_BF16) conv_tc2_kernel<__nv_bfloat16><<<grid, kTc2Threads, smem, stream>>>(tmA, tmB, a, g);
  else conv_tc2_kernel<__half><<<grid, kTc2Threads, smem, stream>>>(tmA, tmB, a, g);
  return cudaGetLastError();
}

}  // namespace vitsb
