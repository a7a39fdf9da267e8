// tcgen05 implicit-GEMM conv1d over time-major 16-bit activations (sm_100a), persistent.
//
//   out[b, t, n] = sum_{j<k} sum_{c<Cin} W[n][j*Cin + c] * in[b, t + j*dil - pad, c]
//
// Used for the reverse flow (pre / in_layers / res_skip / post), conv_pre and the polyphase
// upsamplers.  One CTA per SM walks a static round-robin list of 128 (time) x NT (channel)
// output tiles; the accumulator is double-buffered in TMEM (2 x NT columns) so the fused
// epilogue of tile i runs under the main loop of tile i+1:
//   warp 0      : TMA producer.  For every (tap j, 64-channel block) it loads the A box
//                 [128 rows x BK ch] at row coordinate t0 + j*dil - pad of the 3-D activation
//                 map (C, L, B) - negative / past-the-end rows are zero-filled by TMA, which
//                 IS the conv's zero padding - and the B box [NT x BK] of the K-major weights.
//   warp 1      : issues tcgen05.mma (M=128, N=NT, K=16) from smem descriptors, commits
//                 stage-release and accumulator-ready mbarriers.
//   warps 2..5  : epilogue group 0 (accumulator 0, even tiles of this CTA)
//   warps 6..9  : epilogue group 1 (accumulator 1, odd tiles)
//                 tcgen05.ld 16 columns at a time (thread = time row), fused bias + conditioning
//                 + residual + accumulate + gate + leaky-relu, fp32 and operand-typed stores,
//                 masked beyond len[b] (tc_common.cuh).
// Roofline: tensor (2*128*NT*K flop per tile); HBM traffic is one pass over the input rows and
// the output rows; weights (and the k-fold re-read of A) come from L2.
#include "common.h"
#include "kernels.h"
#include "ptx.h"
#include "tc_common.cuh"

#include <cstdio>
#include <mutex>

namespace vitsb {

struct TcGeom {
  int BK;            // channels per K block (16 / 32 / 64) = swizzle span / 2
  int kchunks;       // Cin / BK
  int stages;        // smem ring depth (<= 8)
  int nNT;           // N tiles
  uint32_t a_bytes;  // 128 * BK * 2
  uint32_t b_bytes;  // NT * BK * 2
  uint32_t b_stride; // smem bytes per B stage (b_bytes rounded to 1024)
  uint32_t tmem_cols;
  uint32_t idesc;
};

constexpr int kTcThreads = 320;
constexpr int kTcMaxB = 512;

template <typename OpT>
__global__ void __launch_bounds__(kTcThreads, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const ConvArgs a, const TcGeom g) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ int s_prefix[kTcMaxB + 1];          // m-tiles before batch row b
  __shared__ uint32_t s_tmem;

  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sA = smem_base;
  const uint32_t sB = sA + g.stages * g.a_bytes;
  const uint32_t bars = sB + g.stages * g.b_stride;          // 8-byte aligned (multiples of 1024 above)
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (8 + s); };
  auto acc_full = [&](int i) { return bars + 8u * (16 + i); };
  auto acc_empty = [&](int i) { return bars + 8u * (18 + i); };

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // provably warp-uniform

  if (threadIdx.x == 0) {
    int acc = 0;
    for (int b = 0; b < a.B; ++b) {
      s_prefix[b] = acc;
      const int valid = a.len ? a.len[b] * a.len_mul : a.L;
      const int rows = valid + kHaloRows < a.L ? valid + kHaloRows : a.L;   // rows nobody reads are skipped
      acc += (rows + kTileM - 1) / kTileM;
    }
    s_prefix[a.B] = acc;
  }
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmB);
    for (int s = 0; s < 8; ++s) { ptx::mbar_init(full_bar(s), 1); ptx::mbar_init(empty_bar(s), 1); }
    for (int i = 0; i < 2; ++i) { ptx::mbar_init(acc_full(i), 1); ptx::mbar_init(acc_empty(i), 4); }
    ptx::mbar_fence_init();
  }
  if (warp == 1) ptx::tmem_alloc(ptx::smem_u32(&s_tmem), g.tmem_cols);
  ptx::tc_fence_before();
  __syncthreads();
  ptx::tc_fence_after();
  const uint32_t tmem_base = s_tmem;

  const int nNT = g.nNT;
  const int total_tiles = s_prefix[a.B] * nNT;
  const int my_tiles = total_tiles > (int)blockIdx.x ? (total_tiles - 1 - (int)blockIdx.x) / (int)gridDim.x + 1 : 0;

  // tile -> (b, t0, ntile): N tile fastest, then time tile, then batch row; b only grows along a CTA's sequence
  int cur_b = 0;
  auto coords = [&](int tile, int& b, int& t0, int& ntile) {
    const int q = tile / nNT;
    ntile = tile - q * nNT;
    while (s_prefix[cur_b + 1] <= q) ++cur_b;
    b = cur_b;
    t0 = (q - s_prefix[cur_b]) * kTileM;
  };

  // Role loops are warp-uniform; only the single-thread instructions are predicated on one
  // elected lane (see ptx::elect_one) so that descriptors / coordinates stay in uniform registers.
  const int stages = g.stages, kchunks = g.kchunks, BK = g.BK, ktaps = a.k;
  const uint32_t a_bytes = g.a_bytes, b_stride = g.b_stride;
  if (warp == 0) {
    // ===== TMA producer =====
    const bool leader = ptx::elect_one();
    const uint32_t tx_bytes = g.a_bytes + g.b_bytes;
    const int in_cols = a.in_cols, dil = a.dil, Cin = a.Cin;
    int s = 0; uint32_t ph = 0;
    for (int n = 0; n < my_tiles; ++n) {
      int b, t0, ntile;
      coords(blockIdx.x + n * gridDim.x, b, t0, ntile);
      const int row0 = t0 - a.pad, n0 = ntile * a.NT;
      for (int j = 0; j < ktaps; ++j) {
        int col = 0;                                  // (kc * BK) % in_cols without the division
        for (int kc = 0; kc < kchunks; ++kc) {
          ptx::mbar_wait(empty_bar(s), ph ^ 1u);
          if (leader) {
            ptx::mbar_expect_tx(full_bar(s), tx_bytes);
            ptx::tma_load_3d(sA + s * a_bytes, &tmA, full_bar(s), col, row0 + j * dil, b);
            ptx::tma_load_2d(sB + s * b_stride, &tmB, full_bar(s), j * Cin + kc * BK, n0);
          }
          col += BK; if (col >= in_cols) col -= in_cols;
          if (++s == stages) { s = 0; ph ^= 1u; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===== MMA issuer =====
    const bool leader = ptx::elect_one();
    const uint64_t dhi = ptx::umma_smem_desc(0, BK * 2, 0);
    const uint32_t idesc = g.idesc;
    const int nk = BK / 16, iters = ktaps * kchunks;
    const uint32_t a16 = a_bytes >> 4, b16 = b_stride >> 4;
    int s = 0; uint32_t ph = 0;
    for (int n = 0; n < my_tiles; ++n) {
      const int buf = n & 1;
      ptx::mbar_wait(acc_empty(buf), ((n >> 1) & 1) ^ 1u);
      ptx::tc_fence_after();
      const uint32_t d_tmem = tmem_base + buf * a.NT;
      for (int it = 0; it < iters; ++it) {
        ptx::mbar_wait(full_bar(s), ph);
        ptx::tc_fence_after();
        const uint64_t ad0 = dhi | (uint64_t)((sA >> 4) + s * a16), bd0 = dhi | (uint64_t)((sB >> 4) + s * b16);
        if (leader) {
          if (nk == 4) {
#pragma unroll
            for (int kk = 0; kk < 4; ++kk)
              ptx::umma_f16(d_tmem, ad0 + 2 * kk, bd0 + 2 * kk, idesc, (it | kk) != 0 ? 1u : 0u);
          } else {
            for (int kk = 0; kk < nk; ++kk)
              ptx::umma_f16(d_tmem, ad0 + 2 * kk, bd0 + 2 * kk, idesc, (it | kk) != 0 ? 1u : 0u);
          }
          ptx::umma_commit(empty_bar(s));           // frees the smem stage when these MMAs retire
        }
        if (++s == stages) { s = 0; ph ^= 1u; }
      }
      if (leader) ptx::umma_commit(acc_full(buf));  // accumulator complete
    }
    __syncwarp();
  } else {
    // ===== epilogue groups: TMEM -> registers -> fused math -> global =====
    const int grp = warp >= 6 ? 1 : 0;
    for (int n = grp; n < my_tiles; n += 2) {
      int b, t0, ntile;
      coords(blockIdx.x + n * gridDim.x, b, t0, ntile);
      const int valid = a.len ? a.len[b] * a.len_mul : a.L;
      ptx::mbar_wait(acc_full(grp), (n >> 1) & 1);
      ptx::tc_fence_after();
      const int nsplit_tile = a.n_split > 0 ? a.n_split / a.NT : nNT;
      if (ntile < nsplit_tile) tc_epilogue<OpT>(a, a.epi, b, t0, ntile, valid, tmem_base + grp * a.NT, warp & 3, lane);
      else tc_epilogue<OpT>(a, a.epi_b, b, t0, ntile - nsplit_tile, valid, tmem_base + grp * a.NT, warp & 3, lane);
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) ptx::mbar_arrive(acc_empty(grp));
    }
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
TmaEncodeFn tma_encode_fn() {
  static TmaEncodeFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<TmaEncodeFn>(p);
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
  if (a.B > kTcMaxB) return false;
  if (a.n_split < 0 || a.n_split % a.NT != 0 || a.n_split >= a.N) return false;
  g->BK = a.Cin >= 64 ? 64 : a.Cin;
  if (a.in_cols <= 0 || a.in_cols % g->BK != 0) return false;
  g->kchunks = a.Cin / g->BK;
  g->nNT = a.N / a.NT;
  g->a_bytes = kTileM * g->BK * 2;
  g->b_bytes = a.NT * g->BK * 2;
  g->b_stride = (g->b_bytes + 1023u) & ~1023u;
  // one persistent CTA per SM: as deep a ring as fits (<= 8 stages, no deeper than two tiles' worth)
  const uint32_t stage = g->a_bytes + g->b_stride;
  int st = (int)((200u * 1024u) / stage);
  if (st > 8) st = 8;
  const int iters2 = 2 * a.k * g->kchunks;
  if (st > iters2) st = iters2;
  if (st < 2) st = 2;
  g->stages = st;
  uint32_t cols = 32;
  while (cols < (uint32_t)(2 * a.NT)) cols <<= 1;
  g->tmem_cols = cols;
  g->idesc = ptx::umma_idesc_f16(op == OP_BF16, kTileM, a.NT);
  return true;
}

cudaError_t launch_conv_tc(const ConvArgs& a, int op, cudaStream_t stream) {
  TcGeom g;
  if (op == OP_F32 || !make_geom(a, op, &g)) return cudaErrorNotSupported;
  TmaEncodeFn enc = tma_encode_fn();
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
  const size_t smem = 1024 + (size_t)g.stages * (g.a_bytes + g.b_stride) + 8 * 20 + 16;
  static int num_sms = 0;
  static std::once_flag once;
  static cudaError_t attr_err = cudaSuccess;
  std::call_once(once, [] {
    cudaError_t e1 = cudaFuncSetAttribute(conv_tc_kernel<__nv_bfloat16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 208 * 1024);
    cudaError_t e2 = cudaFuncSetAttribute(conv_tc_kernel<__half>, cudaFuncAttributeMaxDynamicSharedMemorySize, 208 * 1024);
    attr_err = e1 != cudaSuccess ? e1 : e2;
    int dev = 0; cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&num_sms, cudaDevAttrMultiProcessorCount, dev);
  });
  if (attr_err != cudaSuccess) return attr_err;
  const long long max_tiles = (long long)a.B * ceil_div(a.L, kTileM) * g.nNT;
  int grid = (int)(max_tiles < num_sms ? max_tiles : num_sms);
  if (grid < 1) grid = 1;
  if (op == OP_BF16) conv_tc_kernel<__nv_bfloat16><<<grid, kTcThreads, smem, stream>>>(tmA, tmB, a, g);
  else conv_tc_kernel<__half><<<grid, kTcThreads, smem, stream>>>(tmA, tmB, a, g);
  cudaError_t le = cudaGetLastError();
  if (le != cudaSuccess)
    fprintf(stderr, "conv_tc launch failed (%s): Cin=%d N=%d NT=%d k=%d grid=%d smem=%zu stages=%d\n", cudaGetErrorString(le),
            a.Cin, a.N, a.NT, a.k, grid, smem, g.stages);
  return le;
}

}  // namespace vitsb
