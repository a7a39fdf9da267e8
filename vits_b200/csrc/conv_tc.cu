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
#include <cstdlib>
#include <mutex>

namespace vitsb {

struct TcGeom {
  int BK;            // channels per K block (16 / 32 / 64) = swizzle span / 2
  int kchunks;       // Cin / BK
  int stages;        // smem ring depth (<= 8)
  int nNT;           // N tiles
  uint32_t a_bytes;  // 128 * BK * 2
  uint32_t b_bytes;  // per CTA: (NT / CTAs per item) * BK * 2
  uint32_t b_stride; // smem bytes per B stage (b_bytes rounded to 1024)
  uint32_t tmem_cols;
  uint32_t idesc;
  int dbg;           // VITSB_TC_DBG timing-experiment mask (only with -DVITSB_RB_DEBUG; results are wrong when set)
  int staged;        // 16-bit output rows leave through shared-memory blocks + TMA stores (tc_common.cuh TcStage)
  int stage_cols;    // columns per staged block (32 or 64); one block per epilogue group
};

#ifdef VITSB_RB_DEBUG
#define TC_DBG(mask) ((g.dbg & (mask)) != 0)
#else
#define TC_DBG(mask) (false)
#endif

#ifndef VITSB_TC_EPI_GROUPS
#define VITSB_TC_EPI_GROUPS 2
#endif
constexpr int kTcEpiGroups = VITSB_TC_EPI_GROUPS;                         // 4-warp epilogue groups; every group works on every tile (column units dealt round-robin)
constexpr int kTcThreads = 64 + 128 * kTcEpiGroups;
constexpr int kTcMaxB = kMaxBatchRows;

// TWO = true: CTA-pair mode (cluster of 2, tcgen05 cta_group::2).  A work item is a PAIR of time tiles
// (pair p = m-tiles 2p and 2p+1 of the flattened (batch row, time tile) list; rank r of the pair owns
// m-tile 2p+r) times one N tile.  Each CTA loads its own A box and HALF of the B box (NT/2 weight rows),
// so the L2 -> smem weight traffic per SM halves - these convs are L2-bound on weights at M = 128.
template <typename OpT, bool TWO>
__global__ void __launch_bounds__(kTcThreads, 1)
conv_tc_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
               const __grid_constant__ CUtensorMap tmO, const ConvArgs a, const TcGeom g) {
  extern __shared__ uint8_t smem_raw[];
  __shared__ int s_prefix[kTcMaxB + 1];          // m-tiles before batch row b
  __shared__ uint32_t s_tmem;

  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sA = smem_base;
  const uint32_t sB = sA + g.stages * g.a_bytes;
  const uint32_t stg_bytes = g.staged ? (uint32_t)kTileM * g.stage_cols * 2u : 0u;  // staged stores: [epilogue group][128 rows x stage_cols], swizzled
  const uint32_t sStg = sB + g.stages * g.b_stride;
  const uint32_t bars = sStg + kTcEpiGroups * stg_bytes;                             // 8-byte aligned (multiples of 1024 above)
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (8 + s); };
  auto acc_full = [&](int i) { return bars + 8u * (16 + i); };
  auto acc_empty = [&](int i) { return bars + 8u * (18 + i); };

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // provably warp-uniform
  const int rank = TWO ? (int)ptx::cluster_ctarank() : 0;
  const int ncl = TWO ? 2 : 1;                                                                // CTAs per work item
  const int cl_id = blockIdx.x / ncl, n_cl = gridDim.x / ncl;

  if (warp == 2) build_tile_prefix(s_prefix, a.B, a.len, a.len_mul, a.L, kTileM, lane);
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmB);
    for (int s = 0; s < 8; ++s) { ptx::mbar_init(full_bar(s), 1); ptx::mbar_init(empty_bar(s), 1); }
    for (int i = 0; i < 2; ++i) { ptx::mbar_init(acc_full(i), 1); ptx::mbar_init(acc_empty(i), 4 * kTcEpiGroups * ncl); }
    ptx::mbar_fence_init();
  }
  if (warp == 1) { if (TWO) ptx::tmem_alloc_2sm(ptx::smem_u32(&s_tmem), g.tmem_cols); else ptx::tmem_alloc(ptx::smem_u32(&s_tmem), g.tmem_cols); }
  ptx::tc_fence_before();
  if (TWO) ptx::cluster_sync_all(); else __syncthreads();    // barrier inits visible to the peer before any remote signal
  ptx::tc_fence_after();
  ptx::pdl_launch_dependents();
  ptx::pdl_wait();                   // prologue above overlapped the previous kernel's tail
  const uint32_t tmem_base = s_tmem;

  const int nNT = g.nNT;
  const int n_mt = s_prefix[a.B];                            // real m-tiles
  const int n_q = (n_mt + ncl - 1) / ncl;                    // m-tile groups (pairs)
  const int total_items = n_q * nNT;
  const int my_items = total_items > cl_id ? (total_items - 1 - cl_id) / n_cl + 1 : 0;

  // item -> (b, t0, ntile) of THIS CTA: N tile fastest, then time tile, then batch row; b only grows.
  // An odd m-tile count leaves rank 1 of the last pair without a tile: live = false (it still runs the protocol).
  int cur_b = 0;
  auto coords = [&](int item, int& b, int& t0, int& ntile, bool& live) {
    const int q = item / nNT;
    ntile = item - q * nNT;
    int mt = q * ncl + rank;
    live = mt < n_mt;
    if (!live) mt = n_mt - 1;
    while (s_prefix[cur_b + 1] <= mt) ++cur_b;
    b = cur_b;
    t0 = (mt - s_prefix[cur_b]) * kTileM;
  };

  // taps an N tile needs (see ConvArgs::tap_split_n)
  auto tap_range = [&](int ntile, int& j0, int& j1) {
    j0 = 0; j1 = a.k;
    if (a.tap_split_n > 0) {
      const int n_lo = ntile * a.NT, n_hi = n_lo + a.NT;
      if (n_hi <= a.tap_split_n) j1 = a.k - 1;
      else if (n_lo >= a.tap_split_n) j0 = 1;
    }
  };

  // Role loops are warp-uniform; only the single-thread instructions are predicated on one
  // elected lane (see ptx::elect_one) so that descriptors / coordinates stay in uniform registers.
  const int stages = g.stages, kchunks = g.kchunks, BK = g.BK, ktaps = a.k;
  const uint32_t a_bytes = g.a_bytes, b_stride = g.b_stride;
  if (warp == 0) {
    // ===== TMA producer (both CTAs of a pair; the leader arms the barrier for both) =====
    const bool leader = ptx::elect_one();
    const uint32_t tx_bytes = (g.a_bytes + g.b_bytes) * ncl;
    const int in_cols = a.in_cols, dil = a.dil, Cin = a.Cin;
    int s = 0; uint32_t ph = 0;
    for (int n = 0; n < my_items; ++n) {
      int b, t0, ntile; bool live;
      coords(cl_id + n * n_cl, b, t0, ntile, live);
      const int row0 = (live ? t0 : a.L + 1024) - a.pad;      // dead tile: every row out of range -> TMA zero fill
      const int n0 = ntile * a.NT + rank * (a.NT / ncl);
      int j_lo, j_hi;
      tap_range(ntile, j_lo, j_hi);
      for (int j = j_lo; j < j_hi; ++j) {
        int col = 0;                                  // (kc * BK) % in_cols without the division
        for (int kc = 0; kc < kchunks; ++kc) {
          ptx::mbar_wait(empty_bar(s), ph ^ 1u);
          if (leader && (TC_DBG(1) || TC_DBG(2))) {
            // ablation: 1 = A loaded only where a slab would load it (tap 0, first pass over the columns); 2 = no B loads
            const bool ldA = !TC_DBG(1) || (j == 0 && kc * BK < in_cols), ldB = !TC_DBG(2);
            const uint32_t tx = ((ldA ? g.a_bytes : 0u) + (ldB ? g.b_bytes : 0u)) * ncl;
            if (tx == 0) { if (rank == 0) ptx::mbar_arrive(full_bar(s)); }
            else if (!TWO || rank == 0) ptx::mbar_expect_tx(full_bar(s), tx);
            if (TWO) {
              if (ldA) ptx::tma_load_3d_2sm(sA + s * a_bytes, &tmA, full_bar(s), col, row0 + j * dil, b);
              if (ldB) ptx::tma_load_2d_2sm(sB + s * b_stride, &tmB, full_bar(s), j * Cin + kc * BK, n0);
            } else {
              if (ldA) ptx::tma_load_3d(sA + s * a_bytes, &tmA, full_bar(s), col, row0 + j * dil, b);
              if (ldB) ptx::tma_load_2d(sB + s * b_stride, &tmB, full_bar(s), j * Cin + kc * BK, n0);
            }
          } else if (leader) {
            if (TWO) {
              if (rank == 0) ptx::mbar_expect_tx(full_bar(s), tx_bytes);
              ptx::tma_load_3d_2sm(sA + s * a_bytes, &tmA, full_bar(s), col, row0 + j * dil, b);
              ptx::tma_load_2d_2sm(sB + s * b_stride, &tmB, full_bar(s), j * Cin + kc * BK, n0);
            } else {
              ptx::mbar_expect_tx(full_bar(s), tx_bytes);
              ptx::tma_load_3d(sA + s * a_bytes, &tmA, full_bar(s), col, row0 + j * dil, b);
              ptx::tma_load_2d(sB + s * b_stride, &tmB, full_bar(s), j * Cin + kc * BK, n0);
            }
          }
          col += BK; if (col >= in_cols) col -= in_cols;
          if (++s == stages) { s = 0; ph ^= 1u; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ===== MMA issuer (leader CTA only in pair mode) =====
    if (rank == 0) {
      const bool leader = ptx::elect_one();
      const uint64_t dhi = ptx::umma_smem_desc(0, BK * 2, 0);
      const uint32_t idesc = g.idesc;
      const int nk = BK / 16;
      const uint32_t a16 = a_bytes >> 4, b16 = b_stride >> 4;
      const uint32_t sA16 = (sA & 0x3FFFFu) >> 4, sB16 = (sB & 0x3FFFFu) >> 4;
      int s = 0; uint32_t ph = 0;
      for (int n = 0; n < my_items; ++n) {
        const int buf = n & 1;
        ptx::mbar_wait(acc_empty(buf), ((n >> 1) & 1) ^ 1u);
        ptx::tc_fence_after();
        const uint32_t d_tmem = tmem_base + buf * a.NT;
        int j_lo, j_hi;
        { const int item = cl_id + n * n_cl; tap_range(item - (item / nNT) * nNT, j_lo, j_hi); }
        const int iters = (j_hi - j_lo) * kchunks;
        for (int it = 0; it < iters; ++it) {
          ptx::mbar_wait(full_bar(s), ph);
          ptx::tc_fence_after();
          const uint64_t ad0 = dhi | (uint64_t)(sA16 + s * a16), bd0 = dhi | (uint64_t)(sB16 + s * b16);
          if (leader) {
#pragma unroll 4
            for (int kk = 0; kk < nk; ++kk) {
              if (TWO) ptx::umma_f16_2sm(d_tmem, ad0 + 2 * kk, bd0 + 2 * kk, idesc, (it | kk) != 0 ? 1u : 0u);
              else ptx::umma_f16(d_tmem, ad0 + 2 * kk, bd0 + 2 * kk, idesc, (it | kk) != 0 ? 1u : 0u);
            }
            if (TWO) ptx::umma_commit_2sm(empty_bar(s)); else ptx::umma_commit(empty_bar(s));   // frees the smem stage (both CTAs)
          }
          if (++s == stages) { s = 0; ph ^= 1u; }
        }
        if (leader) { if (TWO) ptx::umma_commit_2sm(acc_full(buf)); else ptx::umma_commit(acc_full(buf)); }   // accumulator complete
      }
    }
    __syncwarp();
  } else {
    // ===== epilogue groups: TMEM -> registers -> fused math -> global =====
    // All groups work on every tile (tc_epilogue deals the column units out); accumulators alternate per item.
    const int grp = (warp - 2) >> 2;
    TcStage stg;
    if (g.staged) { stg.smem = sStg + grp * stg_bytes; stg.tm = &tmO; stg.bar_id = 1 + grp; stg.cols = g.stage_cols; }
    for (int n = 0; n < my_items; ++n) {
      int b, t0, ntile; bool live;
      coords(cl_id + n * n_cl, b, t0, ntile, live);
      const int valid = a.len ? a.len[b] * a.len_mul : a.L;
      const int buf = n & 1;
      auto wait_acc = [&]() { ptx::mbar_wait(acc_full(buf), (n >> 1) & 1); ptx::tc_fence_after(); };
      if (live && !TC_DBG(4)) {
        const int nsplit_tile = a.n_split > 0 ? a.n_split / a.NT : nNT;
        if (ntile < nsplit_tile) tc_epilogue<OpT>(a, a.epi, b, t0, ntile, valid, tmem_base + buf * a.NT, warp & 3, lane, grp, kTcEpiGroups, wait_acc, g.dbg, &stg);
        else tc_epilogue<OpT>(a, a.epi_b, b, t0, ntile - nsplit_tile, valid, tmem_base + buf * a.NT, warp & 3, lane, grp, kTcEpiGroups, wait_acc, g.dbg);
      } else {
        wait_acc();
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) { if (TWO) ptx::mbar_arrive_leader(acc_empty(buf)); else ptx::mbar_arrive(acc_empty(buf)); }
    }
    if (g.staged && (warp & 3) == 0 && lane == 0) ptx::bulk_wait0();     // staged rows are in global memory before the CTA retires
  }

  // ---- teardown: all TMEM reads (and, in pair mode, the peer's) done before dealloc ----
  ptx::tc_fence_before();
  if (TWO) ptx::cluster_sync_all(); else __syncthreads();
  if (warp == 1) {
    ptx::tc_fence_after();
    if (TWO) ptx::tmem_dealloc_2sm(tmem_base, g.tmem_cols); else ptx::tmem_dealloc(tmem_base, g.tmem_cols);
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

static bool make_geom(const ConvArgs& a, int op, int ncl, TcGeom* g) {
  if (!conv_tc_supported(a.Cin, a.in_ld, a.N, a.NT, a.k)) return false;
  if (a.epi.gate && (a.NT % 32 != 0)) return false;
  if (a.B > kTcMaxB) return false;
  if (a.n_split < 0 || a.n_split % a.NT != 0 || a.n_split >= a.N) return false;
  g->BK = a.Cin >= 64 ? 64 : a.Cin;
  if (a.in_cols <= 0 || a.in_cols % g->BK != 0) return false;
  g->kchunks = a.Cin / g->BK;
  g->nNT = a.N / a.NT;
  g->a_bytes = kTileM * g->BK * 2;
  g->b_bytes = (a.NT / ncl) * g->BK * 2;
  g->b_stride = (g->b_bytes + 1023u) & ~1023u;
  // one persistent CTA per SM: as deep a ring as fits (<= 8 stages, no deeper than two tiles' worth)
  const uint32_t stage = g->a_bytes + g->b_stride;
  int st = (int)((200u * 1024u) / stage);
  if (st > 8) st = 8;
  // staged 16-bit output: unsplit operand rows only, dense rows, nothing else written or read by the epilogue
  const Epilogue& ep = a.epi;
  g->staged = (!ep.gate && ep.out_op && !ep.out_f32 && !ep.res && !ep.accum && ep.out_op_split == 0 && a.n_split == 0 &&
               a.NT % 32 == 0 && (ep.out_op_ld * 2) % 16 == 0 && ep.out_op_bstride == (long long)a.L * ep.out_op_ld) ? 1 : 0;
  // block width: 64 columns when that still gives every epilogue group a block, else 32
  g->stage_cols = g->staged ? (a.NT % (64 * kTcEpiGroups) == 0 ? 64 : 32) : 0;
  if (g->staged) st = (int)((200u * 1024u - kTcEpiGroups * (uint32_t)kTileM * g->stage_cols * 2u) / stage);
  if (st > 8) st = 8;
  const int iters2 = 2 * a.k * g->kchunks;
  if (st > iters2) st = iters2;
  if (st < 2) st = 2;
  g->stages = st;
  uint32_t cols = 32;
  while (cols < (uint32_t)(2 * a.NT)) cols <<= 1;
  g->tmem_cols = cols;
  g->idesc = ptx::umma_idesc_f16(op == OP_BF16, kTileM * ncl, a.NT);
  return true;
}

// Wave balance: one output tile is 128 rows x NT columns and a CTA (pair) runs its tiles back to back,
// so the kernel lasts ceil(items / groups) tiles.  With ~75 time tiles per call (16 utterances) NT = 256
// gives 150 tiles on 148 SMs - two waves, the second almost empty.  Pick the N tile that minimises
// waves x (NT + fixed per-tile cost) among the divisors the packing allows.
static int pick_launch_nt(const ConvArgs& a, int num_sms, bool pair_ok) {
  if (a.epi.gate && a.gate_group <= 0) return a.NT;          // gate pairs are packed per NT: fixed
  const int unit = a.epi.gate ? a.gate_group : 32;
  const int q = a.m_tiles > 0 ? a.m_tiles : a.B * ceil_div(a.L, kTileM);
  int best = a.NT; double best_cost = 1e30;
  for (int nt = a.NT; nt >= unit; nt -= unit) {
    if (a.N % nt != 0 || nt % unit != 0 || (a.n_split > 0 && a.n_split % nt != 0)) continue;
    const int ncl = (pair_ok && nt % 32 == 0) ? 2 : 1;
    const long long items = (long long)ceil_div(q, ncl) * (a.N / nt);
    const int groups = num_sms / ncl;
    const double waves = (double)((items + groups - 1) / groups);
    // tiles that do not straddle the tap split run k - 1 of the k taps
    const bool aligned = a.tap_split_n > 0 && a.tap_split_n % nt == 0 && a.k > 1;
    const double cost = waves * (nt * (aligned ? (double)(a.k - 1) / a.k : 1.0) + 40.0);
    if (cost < best_cost - 1e-9) { best_cost = cost; best = nt; }
  }
  return best;
}

cudaError_t launch_conv_tc(const ConvArgs& a_in, int op, const LaunchOpts& lo, cudaStream_t stream) {
  if (op == OP_F32) return cudaErrorNotSupported;
  const int num_sms = lo.num_sms > 0 ? lo.num_sms : 148;
  ConvArgs a = a_in;
  a.NT = pick_launch_nt(a_in, num_sms, lo.pair);
  // CTA-pair mode whenever each CTA's half of the weight box is a whole number of 8-row swizzle atoms
  const int ncl = (lo.pair && a.NT % 32 == 0 && a.NT >= 32) ? 2 : 1;
  TcGeom g;
  if (!make_geom(a, op, ncl, &g)) return cudaErrorNotSupported;
  g.dbg = 0;
#ifdef VITSB_RB_DEBUG
  { static int dbg = -1; if (dbg < 0) { const char* e = getenv("VITSB_TC_DBG"); dbg = e ? atoi(e) : 0; } g.dbg = dbg; }
#endif
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
    cuuint32_t box[2] = {(cuuint32_t)g.BK, (cuuint32_t)(a.NT / ncl)};
    cuuint32_t estr[2] = {1, 1};
    CUresult r = enc(&tmB, dt, 2, const_cast<void*>(a.w), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, sw,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
  }
  CUtensorMap tmO = tmA;                           // placeholder when the output is not staged (never dereferenced)
  if (g.staged) {
    cuuint64_t dims[3] = {(cuuint64_t)a.epi.out_op_ld, (cuuint64_t)a.L, (cuuint64_t)a.B};
    cuuint64_t strides[2] = {(cuuint64_t)a.epi.out_op_ld * 2, (cuuint64_t)a.epi.out_op_bstride * 2};
    cuuint32_t box[3] = {(cuuint32_t)g.stage_cols, (cuuint32_t)kTileM, 1};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(&tmO, dt, 3, a.epi.out_op, dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     g.stage_cols == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) return cudaErrorInvalidValue;
  }
  const size_t smem = 1024 + (size_t)g.stages * (g.a_bytes + g.b_stride) + (size_t)kTcEpiGroups * kTileM * g.stage_cols * 2 + 8 * 20 + 16;
  static PerDeviceOnce smem_optin;             // the opt-in is per device: a second engine on another GPU needs its own
  const cudaError_t attr_err = smem_optin.run([] {
    cudaError_t e[4] = {
        cudaFuncSetAttribute(conv_tc_kernel<__nv_bfloat16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 208 * 1024),
        cudaFuncSetAttribute(conv_tc_kernel<__half, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 208 * 1024),
        cudaFuncSetAttribute(conv_tc_kernel<__nv_bfloat16, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 208 * 1024),
        cudaFuncSetAttribute(conv_tc_kernel<__half, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 208 * 1024)};
    for (cudaError_t x : e) if (x != cudaSuccess) return x;
    return cudaSuccess;
  });
  if (attr_err != cudaSuccess) return attr_err;
  const long long max_items = (long long)ceil_div(a.m_tiles > 0 ? a.m_tiles : a.B * ceil_div(a.L, kTileM), ncl) * g.nNT;
  const int max_groups = num_sms / ncl;
  int groups = (int)(max_items < max_groups ? max_items : max_groups);
  if (groups < 1) groups = 1;
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = dim3(groups * ncl, 1, 1);
  cfg.blockDim = dim3(kTcThreads, 1, 1);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = ncl; at[0].val.clusterDim.y = 1; at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = lo.pdl ? 2 : 1;
  cudaError_t le;
  if (ncl == 2) {
    le = op == OP_BF16 ? cudaLaunchKernelEx(&cfg, conv_tc_kernel<__nv_bfloat16, true>, tmA, tmB, tmO, a, g)
                       : cudaLaunchKernelEx(&cfg, conv_tc_kernel<__half, true>, tmA, tmB, tmO, a, g);
  } else {
    le = op == OP_BF16 ? cudaLaunchKernelEx(&cfg, conv_tc_kernel<__nv_bfloat16, false>, tmA, tmB, tmO, a, g)
                       : cudaLaunchKernelEx(&cfg, conv_tc_kernel<__half, false>, tmA, tmB, tmO, a, g);
  }
  if (le == cudaSuccess) le = cudaGetLastError();
  if (le != cudaSuccess)
    fprintf(stderr, "conv_tc launch failed (%s): Cin=%d N=%d NT=%d k=%d grid=%d x%d smem=%zu stages=%d\n", cudaGetErrorString(le),
            a.Cin, a.N, a.NT, a.k, groups, ncl, smem, g.stages);
  return le;
}

}  // namespace vitsb
