// tcgen05 implicit-GEMM conv1d over time-major 16-bit activations (sm_100a).
//
//   out[b, t, n] = sum_{j<k} sum_{c<Cin} W[n][j*Cin + c] * in[b, t + j*dil - pad, c]
//
// One CTA computes a 128 (time) x NT (channels) accumulator tile in TMEM.
//   warp 0      : TMA producer.  For every (tap j, 64-channel block) it loads the
//                 A box [128 rows x BK ch] at row coordinate t0 + j*dil - pad of the
//                 3-D activation map (C, L, B) — negative / past-the-end rows are
//                 zero-filled by TMA, which IS the conv's zero padding — and the
//                 B box [NT x BK] of the K-major packed weights.
//   warp 1      : allocates TMEM, issues tcgen05.mma (M=128, N=NT, K=16) from smem
//                 descriptors, commits stage release / accumulator-ready mbarriers.
//   warps 2..5  : epilogue.  tcgen05.ld 16 columns at a time (thread = time row),
//                 fused bias + conditioning + residual + accumulate + gate + leaky-relu,
//                 fp32 and operand-typed stores, masked beyond len[b].
// Roofline: tensor (2*128*NT*K flop per tile); HBM traffic is one pass over the
// input rows and the output rows, weights stay in L2.
#include "common.h"
#include "kernels.h"
#include "ptx.h"
#include "tc_common.cuh"

#include <cstdlib>
#include <mutex>

namespace vitsb {

struct TcGeom {
  int BK;            // channels per K block (16 / 32 / 64) = swizzle span / 2
  int kchunks;       // Cin / BK
  int iters;         // k * kchunks
  int stages;
  uint32_t a_bytes;  // 128 * BK * 2
  uint32_t b_bytes;  // NT * BK * 2 (rounded to 1024 in smem)
  uint32_t b_stride; // smem bytes per B stage
  uint32_t tmem_cols;
  uint32_t idesc;
};

constexpr int kTcThreads = 192;

template <typename OpT>
__global__ void __launch_bounds__(kTcThreads)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const ConvArgs a, const TcGeom g) {
  const int b = blockIdx.z;
  const int t0 = blockIdx.x * kTileM;
  const int ntile = blockIdx.y;
  const int valid = a.len ? a.len[b] * a.len_mul : a.L;
  if (t0 >= valid + kHaloRows) return;            // CTA-uniform: rows nobody reads

  extern __shared__ uint8_t smem_raw[];
  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sA = smem_base;
  const uint32_t sB = sA + g.stages * g.a_bytes;
  const uint32_t bars = sB + g.stages * g.b_stride;          // 8-byte aligned (multiples of 1024 above)
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (g.stages + s); };
  const uint32_t accum_bar = bars + 8u * (2 * g.stages);
  const uint32_t tmem_slot = accum_bar + 8u;
  // generic pointer to the TMEM-address slot
  uint32_t* tmem_slot_ptr = reinterpret_cast<uint32_t*>(smem_raw + (tmem_slot - ptx::smem_u32(smem_raw)));

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // provably warp-uniform

  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmB);
    for (int s = 0; s < g.stages; ++s) { ptx::mbar_init(full_bar(s), 1); ptx::mbar_init(empty_bar(s), 1); }
    ptx::mbar_init(accum_bar, 1);
    ptx::mbar_fence_init();
  }
  if (warp == 1) ptx::tmem_alloc(tmem_slot, g.tmem_cols);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot_ptr;

  // Role loops are warp-uniform; only the single-thread instructions are predicated on one
  // elected lane (see ptx::elect_one) so that descriptors / coordinates stay in uniform registers.
  const int iters = g.iters, stages = g.stages, kchunks = g.kchunks, BK = g.BK;
  const uint32_t a_bytes = g.a_bytes, b_stride = g.b_stride;
  if (warp == 0) {
    // ===== TMA producer =====
    const bool leader = ptx::elect_one();
    const uint32_t tx_bytes = g.a_bytes + g.b_bytes;
    const int in_cols = a.in_cols, dil = a.dil, row0 = t0 - a.pad, Cin = a.Cin, n0 = ntile * a.NT;
    for (int it = 0; it < iters; ++it) {
      const int s = it % stages;
      const uint32_t ph = (it / stages) & 1;
      ptx::mbar_wait(empty_bar(s), ph ^ 1u);
      if (leader) {
        ptx::mbar_expect_tx(full_bar(s), tx_bytes);
        const int j = it / kchunks, kc = it - j * kchunks;
        ptx::tma_load_3d(sA + s * a_bytes, &tmA, full_bar(s), (kc * BK) % in_cols, row0 + j * dil, b);
        ptx::tma_load_2d(sB + s * b_stride, &tmB, full_bar(s), j * Cin + kc * BK, n0);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===== MMA issuer =====
    const bool leader = ptx::elect_one();
    const uint64_t dhi = ptx::umma_smem_desc(0, BK * 2, 0);
    const uint32_t idesc = g.idesc;
    const int nk = BK / 16;
    for (int it = 0; it < iters; ++it) {
      const int s = it % stages;
      const uint32_t ph = (it / stages) & 1;
      ptx::mbar_wait(full_bar(s), ph);
      ptx::tc_fence_after();
      const uint64_t ad0 = dhi | ((sA + s * a_bytes) >> 4), bd0 = dhi | ((sB + s * b_stride) >> 4);
      if (leader) {
        for (int kk = 0; kk < nk; ++kk)
          ptx::umma_f16(tmem_base, ad0 + 2 * kk, bd0 + 2 * kk, idesc, (it | kk) != 0 ? 1u : 0u);
        ptx::umma_commit(empty_bar(s));           // frees the smem stage when these MMAs retire
      }
    }
    if (leader) ptx::umma_commit(accum_bar);      // accumulator complete
    __syncwarp();
  } else {
    // ===== epilogue: TMEM -> registers -> fused math -> global =====
    ptx::mbar_wait(accum_bar, 0);
    ptx::tc_fence_after();
    tc_epilogue<OpT>(a, b, t0, ntile, valid, tmem_base, warp & 3, lane);
  }

  // ---- teardown: all TMEM reads done before dealloc ----
  ptx::tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    ptx::tmem_dealloc(tmem_base, g.tmem_cols);
  }
}

// ---------------------------------------------------------------------------
// host side
// ---------------------------------------------------------------------------
typedef TmaEncodeFn EncodeTiledFn;
static EncodeTiledFn get_encode_fn() { return tma_encode_fn(); }

TmaEncodeFn tma_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

bool conv_tc_supported(int Cin, int in_ld, int N, int NT, int k) {
  // (in_cols, when different from Cin, must be a multiple of the K block; checked in make_geom)
  if (!(Cin == 16 || Cin == 32 || (Cin % 64 == 0 && Cin > 0))) return false;
  if (in_ld % 8 != 0) return false;                       // 16-byte row pitch for TMA
  if (NT < 16 || NT > kMaxTileN || NT % 16 != 0 || N % NT != 0) return false;
  if (k < 1) return false;
  return true;
}

static bool make_geom(const ConvArgs& a, int op, TcGeom* g) {
  if (!conv_tc_supported(a.Cin, a.in_ld, a.N, a.NT, a.k)) return false;
  if (a.epi.gate && (a.NT % 32 != 0)) return false;
  g->BK = a.Cin >= 64 ? 64 : a.Cin;
  if (a.in_cols <= 0 || a.in_cols % g->BK != 0) return false;
  g->kchunks = a.Cin / g->BK;
  g->iters = a.k * g->kchunks;
  g->a_bytes = kTileM * g->BK * 2;
  g->b_bytes = a.NT * g->BK * 2;
  g->b_stride = (g->b_bytes + 1023u) & ~1023u;
  // Pipeline depth vs occupancy: one tile per CTA means a CTA's epilogue only overlaps
  // another CTA's main loop, so prefer the smallest smem footprint that still gives a
  // >=3-deep ring (several co-resident CTAs per SM) over one deep ring per SM.
  const uint32_t stage = g->a_bytes + g->b_stride;
  const uint32_t budgets[4] = {40 * 1024, 72 * 1024, 110 * 1024, 192 * 1024};
  int st = 1;
  for (int i = 0; i < 4; ++i) {
    st = (int)(budgets[i] / stage);
    if (st > 8) st = 8;
    if (st > g->iters) st = g->iters;
    if (st >= (g->iters < 3 ? g->iters : 3)) break;
  }
  if (st < 1) st = 1;
  g->stages = st;
  uint32_t cols = 32;
  while (cols < (uint32_t)a.NT) cols <<= 1;
  g->tmem_cols = cols;
  g->idesc = ptx::umma_idesc_f16(op == OP_BF16, kTileM, a.NT);
  return true;
}

static int tc_mode() {
  static int mode = -1;
  if (mode < 0) { const char* e = getenv("VITSB_TC_MODE"); mode = e ? atoi(e) : 0; }
  return mode;
}

cudaError_t launch_conv_tc(const ConvArgs& a, int op, cudaStream_t stream) {
  if (tc_mode() > 0) {
    cudaError_t e = launch_conv_tc2(a, op, tc_mode(), stream);
    if (e != cudaErrorNotSupported) return e;
  }
  TcGeom g;
  if (op == OP_F32 || !make_geom(a, op, &g)) return cudaErrorNotSupported;
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return cudaErrorNotSupported;

  const CUtensorMapDataType dt = op == OP_BF16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16;
  const CUtensorMapSwizzle sw = g.BK == 64 ? CU_TENSOR_MAP_SWIZZLE_128B
                               : (g.BK == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B);
  CUtensorMap tmA, tmB;
  {
    cuuint64_t dims[3] = {(cuuint64_t)a.in_cols, (cuuint64_t)a.L, (cuuint64_t)a.B};
    cuuint64_t strides[2] = {(cuuint64_t)a.in_ld * 2, (cuuint64_t)a.in_bstride * 2};
    cuuint32_t box[3] = {(cuuint32_t)g.BK, (cuuint32_t)kTileM, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(&tmA, dt, 3, const_cast<void*>(a.in), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
  }
  {
    cuuint64_t dims[2] = {(cuuint64_t)a.k * a.Cin, (cuuint64_t)a.N};
    cuuint64_t strides[1] = {(cuuint64_t)a.k * a.Cin * 2};
    cuuint32_t box[2] = {(cuuint32_t)g.BK, (cuuint32_t)a.NT};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&tmB, dt, 2, const_cast<void*>(a.w), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
  }
  const size_t smem = 1024 + (size_t)g.stages * (g.a_bytes + g.b_stride) + 8 * (2 * g.stages + 1) + 16;
  dim3 grid(ceil_div(a.L, kTileM), a.N / a.NT, a.B);
  static std::once_flag once;
  std::call_once(once, [] {
    cudaFuncSetAttribute(conv_tc_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(conv_tc_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
  });
  if (op == OP_BF16) conv_tc_kernel<__nv_bfloat16><<<grid, kTcThreads, smem, stream>>>(tmA, tmB, a, g);
  else conv_tc_kernel<__half><<<grid, kTcThreads, smem, stream>>>(tmA, tmB, a, g);
  return cudaGetLastError();
}

}  // namespace vitsb
