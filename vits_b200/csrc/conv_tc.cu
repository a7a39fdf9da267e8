// tcgen05 implicit-GEMM conv1d over time-major 16-bit activations (sm_100a), persistent.
//
//   out[b, t, n] = sum_{j<k} sum_{c<Cin} W[n][j*Cin + c] * in[b, t + j*dil - pad, c]
//
// Used for the reverse flow (pre / in_layers / res_skip / post), conv_pre and the polyphase
// upsamplers.  One CTA per SM walks a static round-robin list of 128 (time) x NT (channel)
// output tiles; the accumulator is double-buffered in TMEM (2 x NT columns) so the fused
// epilogue of tile i runs under the main loop of tile i+1:
//   warp 0      : TMA producer of the activation (A) boxes.  For every (tap j, 64-channel block) the box
//                 [128 rows x BK ch] at row coordinate t0 + j*dil - pad of the 3-D activation map (C, L, B) -
//                 negative / past-the-end rows are zero-filled by TMA, which IS the conv's zero padding.
//   warp 2      : TMA producer of the weight (B) boxes [NT x BK] of the K-major weights, into the same ring slots.
//   warp 1      : issues tcgen05.mma (M=128, N=NT, K=16) from smem descriptors, commits
//                 slot-release and accumulator-ready mbarriers.
//   warps 3..6  : epilogue group 0 } every group works on every tile (column blocks dealt round-robin);
//   warps 7..10 : epilogue group 1 } the two TMEM accumulators alternate per tile.
//                 tcgen05.ld 16 columns at a time (thread = time row), fused bias + conditioning
//                 + residual + accumulate + gate + leaky-relu, fp32 and operand-typed stores,
//                 masked beyond len[b] (tc_common.cuh); plain 16-bit outputs leave through swizzled
//                 shared-memory blocks and TMA stores.
// A ring slot holds `grp` boxes (1 for long K loops, 2-3 for the narrow upsamplers whose tiles have 3 or 6 boxes).
// Roofline: tensor (2*128*NT*K flop per tile); HBM traffic is one pass over the input rows and
// the output rows; weights (and the k-fold re-read of A) come from L2.
#include "common.h"
#include "kernels.h"
#include "ptx.h"
#include "tc_common.cuh"

#include <cstdio>
#include <cstdlib>
#include <mutex>
#include <type_traits>

namespace vitsb {

struct TcGeom {
  int BK;            // channels per K block (16 / 32 / 64) = swizzle span / 2
  int kchunks;       // Cin / BK
  int stages;        // smem ring depth in slots (<= 8)
  int grp;           // (tap, 64-channel block) boxes per ring slot: one full / empty barrier round trip per `grp` boxes
  int nNT;           // N tiles
  uint32_t a_bytes;  // 128 * BK * 2
  uint32_t b_bytes;  // per CTA: (NT / CTAs per item) * BK * 2
  uint32_t b_stride; // smem bytes per B stage (b_bytes rounded to 1024)
  uint32_t tmem_cols;
  uint32_t idesc;
  int dbg;           // VITSB_TC_DBG timing-experiment mask (only with -DVITSB_RB_DEBUG; results are wrong when set)
  int staged;        // 16-bit output rows leave through shared-memory blocks + TMA stores (tc_common.cuh TcStage)
  int stage_cols;    // columns per staged block (32 or 64); one block per epilogue group
  int epi_alt;       // 1: an N tile is one or two staged blocks, the epilogue groups take alternate tiles (group = accumulator index)
  long long* trace;  // debug build: clock64 timeline of CTA 0 (VITSB_TC_TRACE), normally null
  int trace_t0;      // first traced item of the CTA's sequence
};

#ifdef VITSB_RB_DEBUG
#define TC_DBG(mask) ((g.dbg & (mask)) != 0)
#define TC_TRACE_ON 1
#else
#define TC_DBG(mask) (false)
#define TC_TRACE_ON 0
#endif

#ifndef VITSB_TC_EPI_GROUPS
#define VITSB_TC_EPI_GROUPS 2
#endif
constexpr int kTcEpiGroups = VITSB_TC_EPI_GROUPS;                         // 4-warp epilogue groups; every group works on every tile (column units dealt round-robin)
constexpr int kTcThreads = 96 + 128 * kTcEpiGroups;                       // warps: A producer, MMA issuer, B producer, then the epilogue groups
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
  __shared__ int s_valid[kTcMaxB];               // valid rows of batch row b (a global load per tile sat on the epilogue's critical path)
  __shared__ uint32_t s_tmem;

  const uint32_t smem_base = (ptx::smem_u32(smem_raw) + 1023u) & ~1023u;
  const uint32_t sA = smem_base;                              // [slot][grp] A boxes
  const uint32_t sB = sA + g.stages * g.grp * g.a_bytes;      // [slot][grp] B boxes
  const uint32_t stg_bytes = g.staged ? (uint32_t)kTileM * g.stage_cols * 2u : 0u;  // staged stores: [epilogue group][128 rows x stage_cols], swizzled
  const uint32_t sStg = sB + g.stages * g.grp * g.b_stride;
  const uint32_t bars = sStg + kTcEpiGroups * stg_bytes;                             // 8-byte aligned (multiples of 1024 above)
  auto full_bar = [&](int s) { return bars + 8u * s; };
  auto empty_bar = [&](int s) { return bars + 8u * (8 + s); };
  auto acc_full = [&](int i) { return bars + 8u * (16 + i); };
  auto acc_empty = [&](int i) { return bars + 8u * (18 + i); };

  const int warp = __shfl_sync(0xffffffffu, threadIdx.x >> 5, 0), lane = threadIdx.x & 31;   // provably warp-uniform
  const int rank = TWO ? (int)ptx::cluster_ctarank() : 0;
  const int ncl = TWO ? 2 : 1;                                                                // CTAs per work item
  const int cl_id = blockIdx.x / ncl, n_cl = gridDim.x / ncl;

  if (warp == 3) build_tile_prefix(s_prefix, a.B, a.len, a.len_mul, a.L, kTileM, lane, kHaloRows, s_valid);
  if (warp == 0 && lane == 0) {
    ptx::prefetch_tmap(&tmA);
    ptx::prefetch_tmap(&tmB);
    for (int s = 0; s < 8; ++s) { ptx::mbar_init(full_bar(s), 1); ptx::mbar_init(empty_bar(s), 1); }
    for (int i = 0; i < 2; ++i) { ptx::mbar_init(acc_full(i), 1); ptx::mbar_init(acc_empty(i), 4 * (g.epi_alt ? 1 : kTcEpiGroups) * ncl); }
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
  // Every role walks its items cl_id, cl_id + n_cl, ... with an ItemIter (no integer division per item: with three
  // k-blocks per tile the bookkeeping of the producer was as long as its loads).
  struct ItemIter {
    int q, nt, dq, dn, nNT;
    __device__ void next() { q += dq; nt += dn; if (nt >= nNT) { nt -= nNT; ++q; } }
  };
  auto item_iter = [&](int first) { ItemIter it; it.nNT = nNT; it.q = first / nNT; it.nt = first - it.q * nNT; it.dq = n_cl / nNT; it.dn = n_cl - it.dq * nNT; return it; };
  int cur_b = 0;
  auto coords = [&](const ItemIter& it, int& b, int& t0, int& ntile, bool& live) {
    ntile = it.nt;
    int mt = it.q * ncl + rank;
    live = mt < n_mt;
    if (!live) mt = n_mt - 1;
    while (s_prefix[cur_b + 1] <= mt) ++cur_b;
    b = cur_b;
    t0 = (mt - s_prefix[cur_b]) * kTileM;
  };

  // timeline of CTA 0 (VITSB_TC_TRACE, debug build): trace[(role * 16 + item) * 2 + kind] = clock64()
  auto trace = [&](int role, int item, int kind) {
    item -= g.trace_t0;
    if (TC_TRACE_ON && g.trace && blockIdx.x == 0 && item >= 0 && item < 16 && lane == 0) g.trace[(role * 16 + item) * 2 + kind] = clock64();
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
  const int stages = g.stages, kchunks = g.kchunks, BK = g.BK;
  const uint32_t a_bytes = g.a_bytes, b_stride = g.b_stride;
  if (warp == 0 || warp == 2) {
    // ===== TMA producers (both CTAs of a pair; the leader CTA's A producer arms the barrier for both) =====
    // Two warps: warp 0 loads the activation (A) boxes, warp 2 the weight (B) boxes of the same ring slots.  One warp
    // doing both spent 420-590 cycles per box in serial uniform-datapath code (barrier poll, expect_tx, two descriptor /
    // coordinate set-ups, kernel-parameter reloads) - longer than the box's four MMAs for every N below 256.
    auto produce = [&](auto is_a_tag) {
      constexpr bool kA = decltype(is_a_tag)::value;
      const bool leader = ptx::elect_one();
      // loop invariants pinned in registers (see the MMA issuer)
      uint32_t in_cols = (uint32_t)a.in_cols, dil = (uint32_t)a.dil, Cin = (uint32_t)a.Cin, grp = (uint32_t)g.grp, n_slots = (uint32_t)stages,
               box_bytes = kA ? a_bytes : b_stride, bk = (uint32_t)BK, nkc = (uint32_t)kchunks;
      asm volatile("" : "+r"(in_cols), "+r"(dil), "+r"(Cin), "+r"(grp), "+r"(n_slots), "+r"(box_bytes), "+r"(bk), "+r"(nkc));
      const uint32_t slot_bytes = grp * box_bytes, base = kA ? sA : sB;
      const uint32_t tx_box = (g.a_bytes + g.b_bytes) * ncl;
      uint32_t s = 0, ph = 0, dst_slot = base, fbar = full_bar(0);          // running slot state
      ItemIter iti = item_iter(cl_id);
      for (int n = 0; n < my_items; ++n, iti.next()) {
        int b, t0, ntile; bool live;
        coords(iti, b, t0, ntile, live);
        const int row0 = (live ? t0 : a.L + 1024) - a.pad;    // dead tile: every row out of range -> TMA zero fill
        const int n0 = ntile * a.NT + rank * (a.NT / ncl);
        int j_lo, j_hi;
        tap_range(ntile, j_lo, j_hi);
        if (kA) trace(0, n, 0);
        const uint32_t iters = (uint32_t)(j_hi - j_lo) * nkc;
        uint32_t kc = 0, col = 0;                             // col = (kc * BK) % in_cols without the division
        int row = row0 + j_lo * (int)dil;                     // A: row coordinate of the current tap
        uint32_t wcol = (uint32_t)j_lo * Cin;                 // B: K coordinate j * Cin + kc * BK
        for (uint32_t it0 = 0; it0 < iters; it0 += grp) {
          const uint32_t gsz = iters - it0 < grp ? iters - it0 : grp;
          ptx::mbar_wait(fbar + 64u, ph ^ 1u);                // empty_bar(s)
          if (kA) {
            uint32_t tx = gsz * tx_box;
#ifdef VITSB_RB_DEBUG
            // ablation: 1 = A loaded only where a slab would load it (tap 0, first pass over the columns); 2 = no B loads
            if (TC_DBG(1) || TC_DBG(2)) {
              tx = 0;
              int jj = j_lo + (int)(it0 / nkc); uint32_t kk = it0 - (it0 / nkc) * nkc;
              for (uint32_t i = 0; i < gsz; ++i) {
                const bool ldA = !TC_DBG(1) || (jj == 0 && kk * bk < in_cols), ldB = !TC_DBG(2);
                tx += ((ldA ? g.a_bytes : 0u) + (ldB ? g.b_bytes : 0u)) * ncl;
                if (++kk == nkc) { kk = 0; ++jj; }
              }
            }
#endif
            if (leader) {
              if (tx == 0) { if (rank == 0) ptx::mbar_arrive(fbar); }
              else if (!TWO || rank == 0) ptx::mbar_expect_tx(fbar, tx);
            }
          }
          uint32_t dst = dst_slot;
          for (uint32_t i = 0; i < gsz; ++i) {
            if (leader) {
              if (kA) {
                if (!TC_DBG(1) || (row == row0 && kc * bk < in_cols)) {
                  if (TWO) ptx::tma_load_3d_2sm(dst, &tmA, fbar, (int)col, row, b);
                  else ptx::tma_load_3d(dst, &tmA, fbar, (int)col, row, b);
                }
              } else if (!TC_DBG(2)) {
                if (TWO) ptx::tma_load_2d_2sm(dst, &tmB, fbar, (int)wcol, n0);
                else ptx::tma_load_2d(dst, &tmB, fbar, (int)wcol, n0);
              }
            }
            dst += box_bytes; wcol += bk;
            col += bk; if (col >= in_cols) col -= in_cols;
            if (++kc == nkc) { kc = 0; col = 0; row += (int)dil; }
          }
          dst_slot += slot_bytes; fbar += 8u;
          if (++s == n_slots) { s = 0; dst_slot = base; fbar = full_bar(0); ph ^= 1u; }
        }
        if (kA) trace(0, n, 1);
      }
    };
    if (warp == 0) produce(std::true_type{}); else produce(std::false_type{});
    __syncwarp();
  } else if (warp == 1) {
    // ===== MMA issuer (leader CTA only in pair mode) =====
    if (rank == 0) {
      const bool leader = ptx::elect_one();
      const uint64_t dhi = ptx::umma_smem_desc(0, BK * 2, 0);
      // Loop invariants of the per-box issue path are pinned in registers (the empty asm hides that they are kernel
      // parameters): left alone, the compiler re-loads them from the constant bank inside the loop (LDCU + dependent
      // UIMADs, ~300 cycles per box) and the issuing thread falls behind the tensor pipe for every N below 256.
      uint32_t idesc = g.idesc, a16 = a_bytes >> 4, b16 = b_stride >> 4, grp = (uint32_t)g.grp, n_slots = (uint32_t)stages, nk = (uint32_t)BK / 16;
      asm volatile("" : "+r"(idesc), "+r"(a16), "+r"(b16), "+r"(grp), "+r"(n_slots), "+r"(nk));
      const uint32_t slotA16 = grp * a16, slotB16 = grp * b16;
      const uint32_t sA16 = (sA & 0x3FFFFu) >> 4, sB16 = (sB & 0x3FFFFu) >> 4;
      auto mma = [&](uint64_t ad, uint64_t bd, uint32_t d_tmem, uint32_t acc) {
        if (TWO) ptx::umma_f16_2sm(d_tmem, ad, bd, idesc, acc); else ptx::umma_f16(d_tmem, ad, bd, idesc, acc);
      };
      uint32_t s = 0, ph = 0, a_off = 0, b_off = 0, fbar = full_bar(0);     // running slot state (no multiplies in the loop)
      ItemIter iti = item_iter(cl_id);
      for (int n = 0; n < my_items; ++n, iti.next()) {
        const int buf = n & 1;
        trace(1, n, 0);
        ptx::mbar_wait(acc_empty(buf), ((n >> 1) & 1) ^ 1u);
        ptx::tc_fence_after();
        trace(1, n, 1);
        const uint32_t d_tmem = tmem_base + buf * a.NT;
        int j_lo, j_hi;
        tap_range(iti.nt, j_lo, j_hi);
        const uint32_t iters = (uint32_t)((j_hi - j_lo) * kchunks);
        for (uint32_t it0 = 0; it0 < iters; it0 += grp) {
          const uint32_t gsz = iters - it0 < grp ? iters - it0 : grp;
          ptx::mbar_wait(fbar, ph);
          ptx::tc_fence_after();
          if (it0 == 0) trace(2, n, 0);
          if (leader) {
            uint64_t ad = dhi | (uint64_t)(sA16 + a_off), bd = dhi | (uint64_t)(sB16 + b_off);
            for (uint32_t i = 0; i < gsz; ++i) {
              const uint32_t acc0 = (it0 | i) != 0 ? 1u : 0u;
              if (nk == 4) {                       // BK = 64: the usual case, straight-line
                mma(ad, bd, d_tmem, acc0); mma(ad + 2, bd + 2, d_tmem, 1u); mma(ad + 4, bd + 4, d_tmem, 1u); mma(ad + 6, bd + 6, d_tmem, 1u);
              } else {
                for (uint32_t kk = 0; kk < nk; ++kk) mma(ad + 2 * kk, bd + 2 * kk, d_tmem, (acc0 | kk) != 0 ? 1u : 0u);
              }
              ad += a16; bd += b16;
            }
            if (TWO) ptx::umma_commit_2sm(fbar + 64u); else ptx::umma_commit(fbar + 64u);   // empty_bar(s): frees the ring slot (both CTAs)
          }
          a_off += slotA16; b_off += slotB16; fbar += 8u;
          if (++s == n_slots) { s = 0; a_off = 0; b_off = 0; fbar = full_bar(0); ph ^= 1u; }
        }
        if (leader) { if (TWO) ptx::umma_commit_2sm(acc_full(buf)); else ptx::umma_commit(acc_full(buf)); }   // accumulator complete
        trace(2, n, 1);
      }
    }
    __syncwarp();
  } else {
    // ===== epilogue groups: TMEM -> registers -> fused math -> global =====
    // All groups work on every tile (tc_epilogue deals the column units out); accumulators alternate per item.
    // epi_alt (an N tile of one or two staged blocks, i.e. NT <= 128): group i owns accumulator i and takes every second tile,
    // so a tile's fixed costs (accumulator wait, fence, two named barriers, store issue) are paid by one group only.
    const int grp = (warp - 3) >> 2;
    const bool alt = g.epi_alt != 0;
    TcStage stg;
    if (g.staged) { stg.smem = sStg + grp * stg_bytes; stg.tm = &tmO; stg.bar_id = 1 + grp; stg.cols = g.stage_cols; }
    ItemIter iti = item_iter(cl_id);
    for (int n = 0; n < my_items; ++n, iti.next()) {
      if (alt && (n & 1) != grp) continue;
      int b, t0, ntile; bool live;
      if (warp == 4) trace(5, n, 0);
      coords(iti, b, t0, ntile, live);
      const int valid = s_valid[b];
      const int buf = n & 1;
      auto wait_acc = [&]() {
        if (warp == 4) trace(5, n, 1);
        ptx::mbar_wait(acc_full(buf), (n >> 1) & 1); ptx::tc_fence_after();
        if ((warp & 3) == 0) trace(3 + grp, n, 0);
      };
      if (live && !TC_DBG(4)) {
        const int nsplit_tile = a.n_split > 0 ? a.n_split / a.NT : nNT;
        if (ntile < nsplit_tile) tc_epilogue<OpT>(a, a.epi, b, t0, ntile, valid, tmem_base + buf * a.NT, warp & 3, lane, alt ? 0 : grp, alt ? 1 : kTcEpiGroups, wait_acc, g.dbg, &stg);
        else tc_epilogue<OpT>(a, a.epi_b, b, t0, ntile - nsplit_tile, valid, tmem_base + buf * a.NT, warp & 3, lane, grp, kTcEpiGroups, wait_acc, g.dbg);
      } else {
        wait_acc();
      }
      ptx::tc_fence_before();
      __syncwarp();
      if (lane == 0) { if (TWO) ptx::mbar_arrive_leader(acc_empty(buf)); else ptx::mbar_arrive(acc_empty(buf)); }
      if ((warp & 3) == 0) trace(3 + grp, n, 1);
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
  // staged 16-bit output: unsplit operand rows only, dense rows, nothing else written or read by the epilogue
  const Epilogue& ep = a.epi;
  g->staged = (!ep.gate && ep.out_op && !ep.out_f32 && !ep.res && !ep.accum && ep.out_op_split == 0 && a.n_split == 0 &&
               a.NT % 32 == 0 && (ep.out_op_ld * 2) % 16 == 0 && ep.out_op_bstride == (long long)a.L * ep.out_op_ld) ? 1 : 0;
  // block width: 64 columns when that still gives every epilogue group a block, else 32
  g->epi_alt = (g->staged && a.NT <= 128) ? 1 : 0;
  g->stage_cols = g->staged ? ((g->epi_alt ? a.NT % 64 == 0 : a.NT % (64 * kTcEpiGroups) == 0) ? 64 : 32) : 0;
  // One persistent CTA per SM.  A ring slot holds `grp` (tap, channel-block) boxes and costs one full / empty barrier
  // round trip: a tile with few k-blocks (the narrow upsamplers: 3 or 6) is otherwise bound by those round trips
  // (timeline: ~700 cycles per box against ~320 of MMA issue).  grp = the largest count <= 4 that divides a tile's
  // k-blocks (with and without the skipped tap) and still leaves three slots; then as deep a ring as fits (<= 8 slots,
  // no deeper than two tiles' worth).
  const uint32_t box = g->a_bytes + g->b_stride;
  const uint32_t budget = 200u * 1024u - (g->staged ? kTcEpiGroups * (uint32_t)kTileM * g->stage_cols * 2u : 0u);
  // k-block counts a tile of this launch can have (ConvArgs::tap_split_n: tiles on one side of the split skip a tap)
  const int it_full = a.k * g->kchunks, it_skip = (a.k - 1) * g->kchunks;
  const bool has_skip = a.tap_split_n > 0 && a.k > 1 && a.N > a.NT, has_full = !(a.tap_split_n > 0 && a.k > 1 && a.tap_split_n % a.NT == 0);
  int grp = 1;
  if (it_full <= 8)                      // long K loops keep one box per slot: finer-grained overlap matters more there
    for (int c = 2; c <= 4; ++c)
      if ((!has_full || it_full % c == 0) && (!has_skip || it_skip % c == 0) &&
          (uint32_t)c * box * (c == it_full ? 2u : 3u) <= budget)       // three slots, or two when a slot is a whole tile
        grp = c;
  g->grp = grp;
  int st = (int)(budget / (box * grp));
  if (st > 8) st = 8;
  const int slots2 = 2 * ((it_full + grp - 1) / grp);
  if (st > slots2) st = slots2;
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
    // an MMA (K = 16) costs max(~80, N / 2) cycles: below N = 160 a narrower tile is not cheaper per instruction
    const double cost = waves * ((nt > 160 ? nt : 160) * (aligned ? (double)(a.k - 1) / a.k : 1.0) + 40.0);
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
  g.trace = nullptr; g.trace_t0 = 0;
#ifdef VITSB_RB_DEBUG
  // VITSB_TC_TRACE="Cin,N,count[,first]": dump the CTA-0 timeline of the first `count` launches with that Cin and N
  static long long* d_trace = nullptr;
  static int tr_cin = -1, tr_n = 0, tr_left = 0, tr_first = 0;
  if (tr_cin < 0) {
    const char* e = getenv("VITSB_TC_TRACE");
    tr_cin = 0;
    if (e && sscanf(e, "%d,%d,%d,%d", &tr_cin, &tr_n, &tr_left, &tr_first) >= 3) cudaMalloc(&d_trace, 6 * 16 * 2 * sizeof(long long));
  }
  const bool tracing = d_trace && tr_cin == a.Cin && tr_n == a.N && tr_left > 0;
  if (tracing) { cudaMemsetAsync(d_trace, 0, 6 * 16 * 2 * sizeof(long long), stream); g.trace = d_trace; g.trace_t0 = tr_first; }
#else
  const bool tracing = false;
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
  const size_t smem = 1024 + (size_t)g.stages * g.grp * (g.a_bytes + g.b_stride) + (size_t)kTcEpiGroups * kTileM * g.stage_cols * 2 + 8 * 20 + 16;
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
#ifdef VITSB_RB_DEBUG
  if (tracing && le == cudaSuccess) {
    --tr_left;
    long long h[6 * 16 * 2];
    cudaStreamSynchronize(stream);
    cudaMemcpy(h, d_trace, sizeof(h), cudaMemcpyDeviceToHost);
    long long t0 = 0;
    for (long long v : h) if (v && (!t0 || v < t0)) t0 = v;
    const char* names[6] = {"tma(first,last issue)", "mma(wait acc_empty)", "mma(first full,commit)", "epi0(acc ok,released)", "epi1(acc ok,released)", "epi0(loop top,wait acc)"};
    fprintf(stderr, "conv_tc trace Cin=%d N=%d NT=%d k=%d pair=%d slots=%d x %d boxes stage_cols=%d (cycles since first event, CTA 0)\n", a.Cin, a.N, a.NT, a.k,
            ncl == 2, g.stages, g.grp, g.stage_cols);
    for (int r = 0; r < 6; ++r) {
      fprintf(stderr, "  %-24s", names[r]);
      for (int t = 0; t < 12; ++t) {
        const long long x = h[(r * 16 + t) * 2], y = h[(r * 16 + t) * 2 + 1];
        if (!x && !y) break;
        fprintf(stderr, " [%lld,%lld]", x ? x - t0 : -1, y ? y - t0 : -1);
      }
      fprintf(stderr, "\n");
    }
  }
#else
  (void)tracing;
#endif
  if (le != cudaSuccess)
    fprintf(stderr, "conv_tc launch failed (%s): Cin=%d N=%d NT=%d k=%d grid=%d x%d smem=%zu stages=%d\n", cudaGetErrorString(le),
            a.Cin, a.N, a.NT, a.k, groups, ncl, smem, g.stages);
  return le;
}

}  // namespace vitsb
